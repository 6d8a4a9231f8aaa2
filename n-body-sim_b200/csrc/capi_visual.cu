// libnbody.so -- renderer feed (history ring, ribbon trails), snapshots and the glue of the headless body
// renderer (include/nbody.h: nbody_history_*, nbody_trail_ribbons, nbody_snapshot_*, nbody_set_materials,
// nbody_render).  SURVEY.md 8f rows 3 and 4; kernels in trails.cuh and render.cu.
#include "capi_ctx.cuh"

#include "integrate.cuh"
#include "trails.cuh"

// ============================================================================
// renderer feed: position history ring + ribbon trails (trails.cuh)
// ============================================================================

extern "C" int nbody_history_enable(nbody_ctx* c, int depth)
{
    DeviceGuard device_guard_;
    if (!c) return NBODY_EINVAL;
    if (depth != 0 && (depth < 2 || depth > 256)) return fail(c, NBODY_EINVAL, "history_enable: depth=%d outside 2..256", depth);
    Shard& s = c->shards[0];
    CU(c, cudaSetDevice(s.device));
    CU(c, cudaStreamSynchronize(s.stream));
    cudaFree(s.hist);
    cudaFree(s.hist_alt);
    s.hist = s.hist_alt = nullptr;
    c->hist_depth = 0;
    c->hist_n = c->hist_prev_n = 0;
    c->hist_fresh_pending = false;
    c->hist_head = 0;
    if (depth == 0) return NBODY_OK;
    c->hist_stride = c->cap_pad;
    CU(c, cudaMalloc(&s.hist, size_t(depth) * c->hist_stride * 4 * c->esz));
    c->hist_depth = depth;
    const int smem = (depth + 1) * 512;  // trail_bodies() x sizeof(Real4) = 512 B per sample in either precision, one padding element per body
    if (c->prec == NBODY_F32) CU(c, cudaFuncSetAttribute(trail_ribbons_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    else CU(c, cudaFuncSetAttribute(trail_ribbons_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    return NBODY_OK;
}

extern "C" int nbody_history_push(nbody_ctx* c)
{
    DeviceGuard device_guard_;
    int rc = check_ready(c, "history_push");
    if (rc != NBODY_OK) return rc;
    if (c->hist_depth == 0) return fail(c, NBODY_EINVAL, "history_push before history_enable");
    Shard& s = c->shards[0];
    CU(c, cudaSetDevice(s.device));
    const int fill = c->hist_n != c->n;
    const int* fresh = (!fill && c->hist_fresh_pending) ? s.hist_src : nullptr;  // bodies new to a remapped ring
    c->hist_fresh_pending = false;
    c->hist_prev_n = 0;
    if (fill) c->hist_head = 0;
    else c->hist_head = (c->hist_head + c->hist_depth - 1) % c->hist_depth;  // unshift: the newest sample moves one slot back
    const unsigned blocks = unsigned((c->n + 255) / 256);
    if (c->prec == NBODY_F32)
        history_push_kernel<float><<<blocks, 256, 0, s.stream>>>((const float*)s.pos[c->cur], (float4*)s.hist, c->n, c->hist_stride, c->hist_head,
                                                                 c->hist_depth, fill, fresh);
    else
        history_push_kernel<double><<<blocks, 256, 0, s.stream>>>((const double*)s.pos[c->cur], (double4*)s.hist, c->n, c->hist_stride, c->hist_head,
                                                                  c->hist_depth, fill, fresh);
    c->launches++;
    CU(c, cudaGetLastError());
    c->hist_n = c->n;
    return NBODY_OK;
}

extern "C" int nbody_history_remap(nbody_ctx* c, const int* source)
{
    DeviceGuard device_guard_;
    int rc = check_ready(c, "history_remap");
    if (rc != NBODY_OK) return rc;
    if (!source) return fail(c, NBODY_EINVAL, "history_remap: source is NULL");
    const size_t old_n = c->hist_n ? c->hist_n : c->hist_prev_n;  // the list the ring was last pushed for
    if (c->hist_depth == 0 || old_n == 0) return fail(c, NBODY_EINVAL, "history_remap: no history to carry over (enable, then push)");
    for (size_t i = 0; i < c->n; ++i)
        if (source[i] >= 0 && size_t(source[i]) >= old_n)
            return fail(c, NBODY_EINVAL, "history_remap: source[%zu]=%d outside the previous list of %zu bodies", i, source[i], old_n);
    Shard& s = c->shards[0];
    CU(c, cudaSetDevice(s.device));
    const size_t ring_bytes = size_t(c->hist_depth) * c->hist_stride * 4 * c->esz;
    if (!s.hist_alt) CU(c, cudaMalloc(&s.hist_alt, ring_bytes));
    if (!s.hist_src) CU(c, cudaMalloc(&s.hist_src, c->cap_pad * sizeof(int)));
    CU(c, cudaMemcpyAsync(s.hist_src, source, c->n * sizeof(int), cudaMemcpyHostToDevice, s.stream));
    const unsigned blocks = unsigned((c->n + 255) / 256);
    if (c->prec == NBODY_F32)
        history_remap_kernel<float><<<blocks, 256, 0, s.stream>>>((const float4*)s.hist, (float4*)s.hist_alt, s.hist_src, c->n, c->hist_stride, c->hist_depth);
    else
        history_remap_kernel<double><<<blocks, 256, 0, s.stream>>>((const double4*)s.hist, (double4*)s.hist_alt, s.hist_src, c->n, c->hist_stride, c->hist_depth);
    c->launches++;
    CU(c, cudaGetLastError());
    CU(c, cudaStreamSynchronize(s.stream));  // `source` is borrowed for the call only
    std::swap(s.hist, s.hist_alt);
    c->hist_n = c->n;
    c->hist_prev_n = 0;
    c->hist_fresh_pending = true;
    return NBODY_OK;
}

namespace {
int history_ready(nbody_ctx* c, const char* what)
{
    int rc = check_ready(c, what);
    if (rc != NBODY_OK) return rc;
    if (c->hist_depth == 0 || c->hist_n != c->n || c->hist_fresh_pending)
        return fail(c, NBODY_EINVAL, "%s: no history for the current body list (enable, then push)", what);
    return NBODY_OK;
}
}  // namespace

extern "C" int nbody_history_download(nbody_ctx* c, void* out)
{
    DeviceGuard device_guard_;
    int rc = history_ready(c, "history_download");
    if (rc != NBODY_OK) return rc;
    if (!out) return fail(c, NBODY_EINVAL, "history_download: out is NULL");
    Shard& s = c->shards[0];
    CU(c, cudaSetDevice(s.device));
    const size_t elems = c->n * size_t(c->hist_depth) * 3;
    const size_t bytes = elems * c->esz;
    if (bytes > s.ribbons_bytes) {
        CU(c, cudaStreamSynchronize(s.stream));
        cudaFree(s.ribbons);
        s.ribbons = nullptr;
        s.ribbons_bytes = 0;
        CU(c, cudaMalloc(&s.ribbons, bytes));
        s.ribbons_bytes = bytes;
    }
    const unsigned blocks = unsigned((c->n * size_t(c->hist_depth) + 255) / 256);
    if (c->prec == NBODY_F32)
        history_gather_kernel<float><<<blocks, 256, 0, s.stream>>>((const float4*)s.hist, (float*)s.ribbons, c->n, c->hist_stride, c->hist_head, c->hist_depth);
    else
        history_gather_kernel<double><<<blocks, 256, 0, s.stream>>>((const double4*)s.hist, (double*)s.ribbons, c->n, c->hist_stride, c->hist_head, c->hist_depth);
    c->launches++;
    CU(c, cudaGetLastError());
    CU(c, cudaMemcpyAsync(out, s.ribbons, bytes, cudaMemcpyDeviceToHost, s.stream));
    CU(c, cudaStreamSynchronize(s.stream));
    return NBODY_OK;
}

namespace {
// Builds the ribbons into Shard::ribbons (device); shared by the download call and the timing hook.
int build_ribbons(nbody_ctx* c, const double camera[3])
{
    Shard& s = c->shards[0];
    CU(c, cudaSetDevice(s.device));
    const size_t bytes = c->n * size_t(c->hist_depth) * 6 * sizeof(float);
    if (bytes > s.ribbons_bytes) {
        CU(c, cudaStreamSynchronize(s.stream));
        cudaFree(s.ribbons);
        s.ribbons = nullptr;
        s.ribbons_bytes = 0;
        CU(c, cudaMalloc(&s.ribbons, bytes));
        s.ribbons_bytes = bytes;
    }
    const size_t bpb = c->prec == NBODY_F32 ? trail_bodies<float>() : trail_bodies<double>();
    const unsigned blocks = unsigned((c->n + bpb - 1) / bpb);
    const size_t smem = size_t(c->hist_depth + 1) * 512;
    if (c->prec == NBODY_F32)
        trail_ribbons_kernel<float><<<blocks, kTrailThreads, smem, s.stream>>>((const float4*)s.hist, (const float*)s.radius, s.ribbons, c->n, c->hist_stride,
                                                                               c->hist_head, c->hist_depth, c->hist_depth, float(camera[0]),
                                                                               float(camera[1]), float(camera[2]));
    else
        trail_ribbons_kernel<double><<<blocks, kTrailThreads, smem, s.stream>>>((const double4*)s.hist, (const double*)s.radius, s.ribbons, c->n,
                                                                                c->hist_stride, c->hist_head, c->hist_depth, c->hist_depth, camera[0],
                                                                                camera[1], camera[2]);
    c->launches++;
    CU(c, cudaGetLastError());
    return NBODY_OK;
}
}  // namespace

extern "C" int nbody_trail_ribbons(nbody_ctx* c, const double camera[3], float* out)
{
    DeviceGuard device_guard_;
    int rc = history_ready(c, "trail_ribbons");
    if (rc != NBODY_OK) return rc;
    if (!camera) return fail(c, NBODY_EINVAL, "trail_ribbons: camera is NULL");
    if (!c->has_radius) return fail(c, NBODY_EINVAL, "trail_ribbons before set_radius");
    rc = build_ribbons(c, camera);
    if (rc != NBODY_OK) return rc;
    Shard& s = c->shards[0];
    if (out) {  // out == NULL: build on the device only (timing)
        CU(c, cudaMemcpyAsync(out, s.ribbons, c->n * size_t(c->hist_depth) * 6 * sizeof(float), cudaMemcpyDeviceToHost, s.stream));
        CU(c, cudaStreamSynchronize(s.stream));
    }
    return NBODY_OK;
}

// ============================================================================
// snapshots
// ============================================================================

namespace {
struct SnapHeader {
    char magic[8];
    unsigned long long n, step;
    double dt, G, softening;
    unsigned prec, flags;
    unsigned long long reserved;
};
static_assert(sizeof(SnapHeader) == 64, "snapshot header is 64 bytes");
const char kSnapMagic[8] = {'N', 'B', 'O', 'D', 'Y', 'S', 'N', '1'};
}  // namespace

extern "C" int nbody_snapshot_save(nbody_ctx* c, const char* path, unsigned long long step, double dt, double G, double softening)
{
    DeviceGuard device_guard_;
    int rc = check_ready(c, "snapshot_save");
    if (rc != NBODY_OK) return rc;
    if (!path) return fail(c, NBODY_EINVAL, "snapshot_save: path is NULL");
    if (c->rank_mode && c->world > 1) return fail(c, NBODY_EINVAL, "snapshot_save: a rank holds only its own vel/acc; gather on the host instead");
    const size_t n = c->n, esz = c->esz;
    std::vector<char> pm, vel, acc, temp, rad;
    try {
        pm.resize(4 * n * esz);
        vel.resize(4 * n * esz);
        acc.resize(4 * n * esz);
        if (c->has_temp) temp.resize(n * esz);
        if (c->has_radius) rad.resize(n * esz);
    } catch (const std::bad_alloc&) {
        return fail(c, NBODY_ENOMEM, "snapshot_save: host allocation failed");
    }
    rc = nbody_download(c, pm.data(), vel.data(), acc.data(), c->has_temp ? temp.data() : nullptr);
    if (rc != NBODY_OK) return rc;
    if (c->has_radius) {
        Shard& s = c->shards[0];
        CU(c, cudaSetDevice(s.device));
        CU(c, cudaMemcpy(rad.data(), s.radius, n * esz, cudaMemcpyDeviceToHost));
    }
    SnapHeader h;
    memset(&h, 0, sizeof h);
    memcpy(h.magic, kSnapMagic, 8);
    h.n = n;
    h.step = step;
    h.dt = dt;
    h.G = G;
    h.softening = softening;
    h.prec = unsigned(c->prec);
    h.flags = (c->has_temp ? 1u : 0u) | (c->has_radius ? 2u : 0u);
    FILE* f = fopen(path, "wb");
    if (!f) return fail(c, NBODY_EIO, "snapshot_save: cannot open %s", path);
    bool ok = fwrite(&h, sizeof h, 1, f) == 1;
    for (const std::vector<char>* a : {&pm, &vel, &acc, &temp, &rad})
        if (ok && !a->empty()) ok = fwrite(a->data(), 1, a->size(), f) == a->size();
    ok = (fclose(f) == 0) && ok;
    return ok ? NBODY_OK : fail(c, NBODY_EIO, "snapshot_save: short write to %s", path);
}

extern "C" int nbody_snapshot_load(nbody_ctx* c, const char* path, unsigned long long* step, double* dt, double* G, double* softening)
{
    DeviceGuard device_guard_;
    if (!c) return NBODY_EINVAL;
    if (!path) return fail(c, NBODY_EINVAL, "snapshot_load: path is NULL");
    FILE* f = fopen(path, "rb");
    if (!f) return fail(c, NBODY_EIO, "snapshot_load: cannot open %s", path);
    SnapHeader h;
    if (fread(&h, sizeof h, 1, f) != 1 || memcmp(h.magic, kSnapMagic, 8) != 0) {
        fclose(f);
        return fail(c, NBODY_EIO, "snapshot_load: %s is not an NBODYSN1 snapshot", path);
    }
    if (h.prec != unsigned(c->prec) || h.n == 0 || h.n > c->capacity || (h.flags & ~3u)) {
        fclose(f);
        return fail(c, NBODY_EINVAL, "snapshot_load: n=%llu precision=%u flags=%u do not fit this context (capacity %zu, precision %d)", h.n, h.prec,
                    h.flags, c->capacity, int(c->prec));
    }
    const size_t n = size_t(h.n), esz = c->esz;
    std::vector<char> pm, vel, acc, temp, rad;
    try {
        pm.resize(4 * n * esz);
        vel.resize(4 * n * esz);
        acc.resize(4 * n * esz);
        if (h.flags & 1u) temp.resize(n * esz);
        if (h.flags & 2u) rad.resize(n * esz);
    } catch (const std::bad_alloc&) {
        fclose(f);
        return fail(c, NBODY_ENOMEM, "snapshot_load: host allocation failed");
    }
    bool ok = true;
    for (std::vector<char>* a : {&pm, &vel, &acc, &temp, &rad})
        if (ok && !a->empty()) ok = fread(a->data(), 1, a->size(), f) == a->size();
    char extra;
    ok = ok && fread(&extra, 1, 1, f) == 0;  // nothing may follow
    fclose(f);
    if (!ok) return fail(c, NBODY_EIO, "snapshot_load: %s is truncated or has trailing bytes", path);
    int rc = nbody_upload(c, n, pm.data(), vel.data(), acc.data(), (h.flags & 1u) ? temp.data() : nullptr);
    if (rc != NBODY_OK) return rc;
    if (h.flags & 2u) {
        rc = nbody_set_radius(c, rad.data());
        if (rc != NBODY_OK) return rc;
    }
    if (step) *step = h.step;
    if (dt) *dt = h.dt;
    if (G) *G = h.G;
    if (softening) *softening = h.softening;
    return NBODY_OK;
}

// ============================================================================
// headless body renderer (render.cu)
// ============================================================================

extern "C" int nbody_set_materials(nbody_ctx* c, const int* type, const float* seed)
{
    DeviceGuard device_guard_;
    int rc = check_ready(c, "set_materials");
    if (rc != NBODY_OK) return rc;
    Shard& s = c->shards[0];
    CU(c, cudaSetDevice(s.device));
    if (type)
        for (size_t i = 0; i < c->n; ++i)
            if (type[i] < 0 || type[i] > 2) return fail(c, NBODY_EINVAL, "set_materials: type[%zu] = %d (0 rock, 1 metal, 2 ice)", i, type[i]);
    if (!s.mat_type) {
        CU(c, cudaMalloc(&s.mat_type, c->cap_pad * sizeof(int)));
        CU(c, cudaMalloc(&s.mat_seed, c->cap_pad * sizeof(float)));
    }
    CU(c, cudaMemsetAsync(s.mat_type, 0, c->n * sizeof(int), s.stream));
    CU(c, cudaMemsetAsync(s.mat_seed, 0, c->n * sizeof(float), s.stream));
    if (type) CU(c, cudaMemcpyAsync(s.mat_type, type, c->n * sizeof(int), cudaMemcpyHostToDevice, s.stream));
    if (seed) CU(c, cudaMemcpyAsync(s.mat_seed, seed, c->n * sizeof(float), cudaMemcpyHostToDevice, s.stream));
    CU(c, cudaStreamSynchronize(s.stream));
    c->has_materials = true;
    return NBODY_OK;
}

extern "C" int nbody_render(nbody_ctx* c, int width, int height, const double cam_position[3], const double cam_target[3],
                            double fov_deg, float* rgba, int* ids)
{
    DeviceGuard device_guard_;
    int rc = check_ready(c, "render");
    if (rc != NBODY_OK) return rc;
    if (width < 1 || height < 1 || size_t(width) * size_t(height) > (size_t(1) << 28))
        return fail(c, NBODY_EINVAL, "render: %d x %d", width, height);
    if (!cam_position || !cam_target || !(fov_deg > 0.0 && fov_deg < 180.0)) return fail(c, NBODY_EINVAL, "render: camera");
    if (!c->has_radius) return fail(c, NBODY_EINVAL, "render before set_radius");
    if (c->rank_mode && c->world > 1) return fail(c, NBODY_EINVAL, "render: a rank holds only its own temperatures; render from a single-process context");
    if (c->n > 0xffffffffull) return fail(c, NBODY_EINVAL, "render: n exceeds the 32-bit body id of the depth buffer");
    // Matrix4.lookAt(position, target, up = +y) in double, stored as float (render.cuh)
    RenderCamera cam;
    double z[3] = {cam_position[0] - cam_target[0], cam_position[1] - cam_target[1], cam_position[2] - cam_target[2]};
    double l = sqrt(z[0] * z[0] + z[1] * z[1] + z[2] * z[2]);
    if (!(l > 0.0)) return fail(c, NBODY_EINVAL, "render: camera position equals its target");
    for (double& v : z) v /= l;
    double x[3] = {1.0 * z[2] - 0.0 * z[1], 0.0 * z[0] - 0.0 * z[2], 0.0 * z[1] - 1.0 * z[0]};
    l = sqrt(x[0] * x[0] + x[1] * x[1] + x[2] * x[2]);
    if (!(l > 0.0)) return fail(c, NBODY_EINVAL, "render: the camera looks along its up axis");
    for (double& v : x) v /= l;
    const double y[3] = {z[1] * x[2] - z[2] * x[1], z[2] * x[0] - z[0] * x[2], z[0] * x[1] - z[1] * x[0]};
    for (int k = 0; k < 3; ++k) {
        cam.pos[k] = float(cam_position[k]);
        cam.xa[k] = float(x[k]);
        cam.ya[k] = float(y[k]);
        cam.za[k] = float(z[k]);
    }
    cam.tan_half = float(tan(fov_deg * 3.14159265358979323846 / 360.0));
    cam.aspect = float(width) / float(height);
    cam.near = 0.1f;  // PerspectiveCamera(45, aspect, 0.1, 3000), idx:704

    Shard& s = c->shards[0];
    CU(c, cudaSetDevice(s.device));
    const size_t pixels = size_t(width) * size_t(height);
    if (pixels > s.image_pixels) {
        CU(c, cudaStreamSynchronize(s.stream));
        render_scratch_free(s.render_scratch);
        cudaFree(s.image);
        cudaFree(s.image_ids);
        s.image = nullptr;
        s.image_ids = nullptr;
        s.image_pixels = 0;
        CU(c, cudaMalloc(&s.image, pixels * sizeof(float4)));
        CU(c, cudaMalloc(&s.image_ids, pixels * sizeof(int)));
        s.image_pixels = pixels;
    }
    // temperatures: the renderer's `temperature` uniform (idx:553); sharded state is gathered on this device
    const void* temp = c->has_temp ? s.temp : nullptr;
    if (c->has_temp && c->shards.size() > 1) {
        if (!s.temp_full) CU(c, cudaMalloc(&s.temp_full, c->cap_pad * c->esz));
        for (Shard& o : c->shards) {
            if (o.n_local_real == 0) continue;
            CU(c, cudaSetDevice(o.device));
            CU(c, cudaStreamSynchronize(o.stream));
            CU(c, cudaMemcpy(static_cast<char*>(s.temp_full) + std::min(o.i_begin, c->n) * c->esz, o.temp, o.n_local_real * c->esz, cudaMemcpyDefault));
        }
        CU(c, cudaSetDevice(s.device));
        temp = s.temp_full;
    }
    const int* type = c->has_materials ? s.mat_type : nullptr;
    const float* seed = c->has_materials ? s.mat_seed : nullptr;
    cudaError_t e;
    if (c->prec == NBODY_F32)
        e = render_launch<float>((const float*)s.pos[c->cur], (const float*)s.radius, (const float*)temp, type, seed, c->n, cam, width, height, s.render_scratch,
                                 s.image, ids ? s.image_ids : nullptr, s.stream);
    else
        e = render_launch<double>((const double*)s.pos[c->cur], (const double*)s.radius, (const double*)temp, type, seed, c->n, cam, width, height,
                                  s.render_scratch, s.image, ids ? s.image_ids : nullptr, s.stream);
    c->launches += 4;  // bin count, scan, bin fill, tile render
    if (e != cudaSuccess) return fail(c, NBODY_ECUDA, "render: %s", cudaGetErrorString(e));
    if (rgba) CU(c, cudaMemcpyAsync(rgba, s.image, pixels * sizeof(float4), cudaMemcpyDeviceToHost, s.stream));
    if (ids) CU(c, cudaMemcpyAsync(ids, s.image_ids, pixels * sizeof(int), cudaMemcpyDeviceToHost, s.stream));
    if (rgba || ids) CU(c, cudaStreamSynchronize(s.stream));  // rgba == NULL: the frame stays on the device (timing)
    return NBODY_OK;
}
