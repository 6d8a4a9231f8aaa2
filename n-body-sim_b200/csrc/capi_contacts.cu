// libnbody.so -- collision broad phase entry points (include/nbody.h: nbody_set_radius, nbody_contacts,
// nbody_contact_stats).  SURVEY.md 8f row 1; kernels in contacts.cuh.
#include "capi_ctx.cuh"

#include "contacts.cuh"

// ============================================================================
// collision broad phase (contacts.cuh)
// ============================================================================

extern "C" int nbody_set_radius(nbody_ctx* c, const void* radius)
{
    DeviceGuard device_guard_;
    int rc = check_ready(c, "set_radius");
    if (rc != NBODY_OK) return rc;
    if (!radius) {
        c->has_radius = false;
        c->near_valid = false;
        return NBODY_OK;
    }
    std::vector<double> r(c->n);
    double rmax = 0.0;
    for (size_t i = 0; i < c->n; ++i) {
        r[i] = c->prec == NBODY_F32 ? double(static_cast<const float*>(radius)[i]) : static_cast<const double*>(radius)[i];
        if (!(r[i] >= 0.0)) return fail(c, NBODY_EINVAL, "set_radius: radius[%zu] = %g", i, r[i]);
        rmax = std::max(rmax, r[i]);
    }
    // A few dominant bodies (a galaxy's central mass, idx:750) would make every body a
    // candidate of the nearest-neighbour test; split them off when they are few.
    std::vector<double> sorted(r);
    std::nth_element(sorted.begin(), sorted.begin() + sorted.size() / 2, sorted.end());
    const double split = 4.0 * sorted[sorted.size() / 2];
    size_t n_big = 0;
    double small_max = 0.0;
    for (double v : r) {
        if (v > split) n_big++;
        else small_max = std::max(small_max, v);
    }
    const bool use_split = n_big > 0 && n_big <= std::max<size_t>(64, c->n / 256);
    c->r_split = use_split ? split : rmax;
    c->r_small_max = use_split ? small_max : rmax;
    for (Shard& s : c->shards) {
        CU(c, cudaSetDevice(s.device));
        CU(c, cudaMemsetAsync(s.radius, 0, c->n_pad * c->esz, s.stream));
        CU(c, cudaMemcpyAsync(s.radius, radius, c->n * c->esz, cudaMemcpyHostToDevice, s.stream));
        std::vector<unsigned> big;
        if (use_split)
            for (size_t i = s.i_begin; i < std::min(s.i_begin + s.n_local, c->n); ++i)
                if (r[i] > split) big.push_back(unsigned(i));
        cudaFree(s.big);
        s.big = nullptr;
        s.n_big = unsigned(big.size());
        if (s.n_big) {
            CU(c, cudaMalloc(&s.big, big.size() * sizeof(unsigned)));
            CU(c, cudaMemcpy(s.big, big.data(), big.size() * sizeof(unsigned), cudaMemcpyHostToDevice));
        }
    }
    for (Shard& s : c->shards) {
        CU(c, cudaSetDevice(s.device));
        CU(c, cudaStreamSynchronize(s.stream));
    }
    c->r_max = rmax;
    c->has_radius = true;
    c->near_valid = false;
    return NBODY_OK;
}

namespace {

template <typename Real, bool BOTH>
int launch_pairs(nbody_ctx* c, Shard& s, const unsigned* cand, unsigned n_cand, double margin)
{
    if (n_cand == 0) return NBODY_OK;
    const unsigned per_cta = kContactThreads * kContactPerThread;
    const unsigned gx = (n_cand + per_cta - 1) / per_cta;
    // enough CTAs to fill the chip: split the j range when there are few candidate blocks
    size_t jch = (size_t(c->sm_count) * 8 + gx - 1) / gx;
    const size_t max_jch = (c->n + kContactThreads - 1) / kContactThreads;
    jch = std::max<size_t>(1, std::min(jch, std::min<size_t>(max_jch, 65535)));
    const size_t chunk_len = round_up((c->n + jch - 1) / jch, size_t(kContactThreads));
    jch = (c->n + chunk_len - 1) / chunk_len;
    const dim3 grid(gx, unsigned(jch));
    contact_pairs_kernel<Real, BOTH><<<grid, kContactThreads, 0, s.stream>>>(static_cast<const Real*>(s.pos[c->cur]), static_cast<const Real*>(s.radius),
                                                                             cand, n_cand, s.i_begin, c->n, chunk_len, margin, s.pairs, s.pair_cap,
                                                                             s.pair_count);
    c->launches++;
    CU(c, cudaGetLastError());
    return NBODY_OK;
}

// Stage 1 on one shard: the nearest-neighbour values of the last force sweep -> candidate list
// (device, s.cand) and its length.
int shard_flag(nbody_ctx* c, Shard& s, double margin, unsigned* n_cand)
{
    *n_cand = 0;
    CU(c, cudaSetDevice(s.device));
    if (s.n_local_real == 0) return NBODY_OK;
    CU(c, cudaMemsetAsync(s.cand_count, 0, sizeof(unsigned), s.stream));
    const unsigned blocks = unsigned((s.n_local_real + 255) / 256);
    flag_near_kernel<<<blocks, 256, 0, s.stream>>>(s.near, static_cast<const float*>(s.radius), s.i_begin, s.n_local_real,
                                                   float(c->r_small_max + margin) * 1.000001f, float(c->r_split), float(c->near_eps2), s.cand,
                                                   s.cand_count);
    c->launches++;
    CU(c, cudaGetLastError());
    CU(c, cudaMemcpyAsync(n_cand, s.cand_count, sizeof *n_cand, cudaMemcpyDeviceToHost, s.stream));
    CU(c, cudaStreamSynchronize(s.stream));
    return NBODY_OK;
}

// Candidate x candidate (contacts.cuh): every device gets the candidates of all devices, concatenated in
// shard order (offsets[k] = where shard k's own begin), and dense (x, y, z, r) records of them.
int build_candidate_records(nbody_ctx* c, const std::vector<unsigned>& n_cand, std::vector<unsigned>& offsets, unsigned* total)
{
    offsets.assign(c->shards.size(), 0);
    unsigned F = 0;
    for (size_t k = 0; k < c->shards.size(); ++k) {
        offsets[k] = F;
        F += n_cand[k];
    }
    *total = F;
    std::vector<unsigned> all;
    if (c->shards.size() > 1) {
        all.resize(F);
        for (size_t k = 0; k < c->shards.size(); ++k) {
            Shard& s = c->shards[k];
            if (n_cand[k] == 0) continue;
            CU(c, cudaSetDevice(s.device));
            CU(c, cudaMemcpy(all.data() + offsets[k], s.cand, n_cand[k] * sizeof(unsigned), cudaMemcpyDeviceToHost));
        }
    }
    for (Shard& s : c->shards) {
        CU(c, cudaSetDevice(s.device));
        if (c->shards.size() > 1) {
            if (!s.cand_all) CU(c, cudaMalloc(&s.cand_all, c->cap_pad * sizeof(unsigned)));
            CU(c, cudaMemcpyAsync(s.cand_all, all.data(), F * sizeof(unsigned), cudaMemcpyHostToDevice, s.stream));
        }
        if (F > s.cand_rec_cap) {
            CU(c, cudaStreamSynchronize(s.stream));
            cudaFree(s.cand_rec);
            s.cand_rec = nullptr;
            s.cand_rec_cap = 0;
            const size_t want = std::min<size_t>(c->cap_pad, size_t(F) + F / 4 + 1024);
            CU(c, cudaMalloc(&s.cand_rec, want * sizeof(float4)));
            s.cand_rec_cap = want;
        }
        const unsigned* list = c->shards.size() > 1 ? s.cand_all : s.cand;
        gather_candidates_kernel<float><<<(F + 255) / 256, 256, 0, s.stream>>>(static_cast<const float*>(s.pos[c->cur]), static_cast<const float*>(s.radius),
                                                                               list, F, s.cand_rec);
        c->launches++;
        CU(c, cudaGetLastError());
    }
    if (c->shards.size() > 1)  // `all` is borrowed by the copies above
        for (Shard& s : c->shards) {
            CU(c, cudaSetDevice(s.device));
            CU(c, cudaStreamSynchronize(s.stream));
        }
    return NBODY_OK;
}

int launch_pairs_compact(nbody_ctx* c, Shard& s, unsigned i_off, unsigned n_i, unsigned n_all, double margin)
{
    if (n_i == 0) return NBODY_OK;
    const unsigned per_cta = kContactThreads * kContactPerThread;
    const unsigned gx = (n_i + per_cta - 1) / per_cta;
    size_t jch = (size_t(c->sm_count) * 8 + gx - 1) / gx;  // enough CTAs to fill the chip
    const size_t max_jch = (n_all + kContactThreads - 1) / kContactThreads;
    jch = std::max<size_t>(1, std::min(jch, std::min<size_t>(max_jch, 65535)));
    const unsigned chunk_len = unsigned(round_up((n_all + jch - 1) / jch, size_t(kContactThreads)));
    jch = (n_all + chunk_len - 1) / chunk_len;
    const unsigned* list = c->shards.size() > 1 ? s.cand_all : s.cand;
    contact_pairs_compact_kernel<float><<<dim3(gx, unsigned(jch)), kContactThreads, 0, s.stream>>>(s.cand_rec, list, i_off, n_i, n_all, chunk_len, margin,
                                                                                                    s.pairs, s.pair_cap, s.pair_count);
    c->launches++;
    CU(c, cudaGetLastError());
    return NBODY_OK;
}

// Stage 2 on one shard: `launch` queues its pair kernels; they run again with a larger buffer until every
// pair fits, then the pairs join `keys` as (i << 32 | j).
template <typename L> int shard_collect(nbody_ctx* c, Shard& s, L&& launch, std::vector<unsigned long long>& keys)
{
    CU(c, cudaSetDevice(s.device));
    if (s.pair_cap == 0) {
        s.pair_cap = 1 << 16;
        CU(c, cudaMalloc(&s.pairs, s.pair_cap * sizeof(uint2)));
    }
    for (;;) {
        CU(c, cudaMemsetAsync(s.pair_count, 0, sizeof(unsigned long long), s.stream));
        const int rc = launch();
        if (rc != NBODY_OK) return rc;
        unsigned long long found = 0;
        CU(c, cudaMemcpyAsync(&found, s.pair_count, sizeof found, cudaMemcpyDeviceToHost, s.stream));
        CU(c, cudaStreamSynchronize(s.stream));
        if (found > s.pair_cap) {  // grow and search again
            cudaFree(s.pairs);
            s.pairs = nullptr;
            s.pair_cap = size_t(found + found / 4 + 1024);
            CU(c, cudaMalloc(&s.pairs, s.pair_cap * sizeof(uint2)));
            continue;
        }
        if (found) {
            std::vector<uint2> host(found);
            CU(c, cudaMemcpyAsync(host.data(), s.pairs, found * sizeof(uint2), cudaMemcpyDeviceToHost, s.stream));
            CU(c, cudaStreamSynchronize(s.stream));
            for (const uint2& q : host) keys.push_back((unsigned long long)q.x << 32 | q.y);
        }
        return NBODY_OK;
    }
}

// Every shard's pairs: those whose lower index lies in the shard's own range, plus the lower-index
// partners of its big bodies.
int all_contacts(nbody_ctx* c, double margin, bool use_flags, std::vector<unsigned long long>& keys)
{
    const size_t ns = c->shards.size();
    std::vector<unsigned> n_cand(ns, 0), offsets(ns, 0);
    unsigned total = 0;
    bool compact = false;
    if (use_flags) {
        size_t flagged = 0;
        for (size_t k = 0; k < ns; ++k) {
            int rc = shard_flag(c, c->shards[k], margin, &n_cand[k]);
            if (rc != NBODY_OK) return rc;
            flagged += n_cand[k];
            c->cstat[0] += n_cand[k] + c->shards[k].n_big;
        }
        // F^2 tests among the candidates against F*N/2 of candidates x everybody (contacts.cuh)
        // One process per GPU: this rank knows only its OWN candidates, and a pair whose partner another rank flagged would be
        // found by nobody; candidates x everybody is complete for the own bodies, so the union over ranks stays the whole set.
        const bool all_candidates_here = !(c->rank_mode && c->world > 1);
        compact = all_candidates_here && flagged > 0 && 2 * flagged < c->n && env_int("NBODY_CONTACTS_COMPACT", 1) != 0;
        if (compact) {
            int rc = build_candidate_records(c, n_cand, offsets, &total);
            if (rc != NBODY_OK) return rc;
        }
    }
    if (compact) c->cstat[3] = 2;
    for (size_t k = 0; k < ns; ++k) {
        Shard& s = c->shards[k];
        if (s.n_local_real == 0) continue;
        const unsigned nc = use_flags ? n_cand[k] : unsigned(s.n_local_real);
        const unsigned n_big = use_flags ? s.n_big : 0;
        c->cstat[1] += nc + n_big;
        if (nc + n_big == 0) continue;
        int rc = shard_collect(c, s, [&]() -> int {
            int r;
            if (c->prec == NBODY_F32) {
                r = compact ? launch_pairs_compact(c, s, offsets[k], nc, total, margin)
                            : launch_pairs<float, false>(c, s, use_flags ? s.cand : nullptr, nc, margin);
                if (r == NBODY_OK) r = launch_pairs<float, true>(c, s, s.big, n_big, margin);
            } else {
                r = launch_pairs<double, false>(c, s, nullptr, nc, margin);  // fp64 contexts have no flags: every body is scanned
            }
            return r;
        }, keys);
        if (rc != NBODY_OK) return rc;
    }
    return NBODY_OK;
}

}  // namespace

extern "C" int nbody_contacts(nbody_ctx* c, double margin, unsigned* pairs, size_t cap, size_t* count)
{
    DeviceGuard device_guard_;
    int rc = check_ready(c, "contacts");
    if (rc != NBODY_OK) return rc;
    if (!count) return fail(c, NBODY_EINVAL, "contacts: count is required");
    *count = 0;
    if (!c->has_radius) return fail(c, NBODY_EINVAL, "contacts before set_radius");
    if (!(margin >= 0.0)) return fail(c, NBODY_EINVAL, "contacts: margin=%g", margin);
    if (c->n > 0xffffffffull) return fail(c, NBODY_EINVAL, "contacts: n exceeds 32-bit pair indices");
    const bool use_flags = c->near_valid;
    c->cstat[0] = c->cstat[1] = c->cstat[2] = 0;
    c->cstat[3] = use_flags;
    std::vector<unsigned long long> keys;
    try {
        rc = all_contacts(c, margin, use_flags, keys);
        if (rc != NBODY_OK) return rc;
        std::sort(keys.begin(), keys.end());  // the reference's visiting order: ascending i, then j (idx:301-305)
        keys.erase(std::unique(keys.begin(), keys.end()), keys.end());  // a big body's scan repeats pairs its partner also found
    } catch (const std::bad_alloc&) {
        return fail(c, NBODY_ENOMEM, "contacts: host allocation failed");
    }
    c->cstat[2] = (long long)keys.size();
    *count = keys.size();
    if (keys.size() > cap) return fail(c, NBODY_ENOMEM, "contacts: %zu pairs found, room for %zu", keys.size(), cap);
    if (!pairs && !keys.empty()) return fail(c, NBODY_EINVAL, "contacts: pairs is NULL");
    for (size_t k = 0; k < keys.size(); ++k) {
        pairs[2 * k] = unsigned(keys[k] >> 32);
        pairs[2 * k + 1] = unsigned(keys[k] & 0xffffffffull);
    }
    return NBODY_OK;
}

extern "C" int nbody_contact_stats(nbody_ctx* c, long long out[4])
{
    DeviceGuard device_guard_;
    if (!c || !out) return NBODY_EINVAL;
    for (int k = 0; k < 4; ++k) out[k] = c->cstat[k];
    return NBODY_OK;
}
