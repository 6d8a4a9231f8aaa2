// nbody_napi.cc -- Node-API addon: a thin shim from JS typed arrays to the C ABI
// of libnbody.so (include/nbody.h).  This is the "N-API addon ... that the
// unchanged JS host calls with typed arrays" of BASELINE.json:north_star; the JS
// side is physics_shim.js, which replaces the body of PhysicsSystem.integrate
// (/root/reference/index.html:247-262).
//
// JS surface (all functions throw a JS Error carrying nbody_last_error on a
// non-zero status, because the reference's integrate() has no error channel):
//   create(precision: 'f32'|'f64', nGpus: int, capacity: int) -> ctx (external)
//   destroy(ctx)
//   upload(ctx, posMass, vel, acc|null, temp|null)        Float32Array/Float64Array
//   download(ctx, posMass|null, vel|null, acc|null, temp|null)
//   step(ctx, dt, G, softening, nsteps)                   nsteps x integrate(dt)
//   kick1Drift(ctx, dt) / force(ctx, G, softening) / kick2(ctx, dt, heatDecay)
//   energy(ctx, G, softening) -> {kinetic, potential, momentum: [px,py,pz]}
//   sync(ctx); info(ctx) -> {n, nPadded, nGpus, jChunks, iTile, kernelLaunches}
//   version() -> string; deviceCount() -> int
//   setRadius(ctx, radius|null)                            Body.radius, idx:221; enables contacts
//   contacts(ctx, margin, pairs: Uint32Array) -> count     overlap search of resolveCollisions, idx:301-313;
//                                                          count > pairs.length/2: nothing written, call again
//   historyEnable(ctx, depth) / historyPush(ctx) / historyDownload(ctx, out)      Body.history, idx:540, 564-565
//   historyRemap(ctx, source: Int32Array[n])               after a list change: body i continues row source[i] (-1: new body)
//   trailRibbons(ctx, camX, camY, camZ, out: Float32Array)  ribbon vertices, idx:567-592
//   snapshotSave(ctx, path, step, dt, G, softening) / snapshotLoad(ctx, path) -> {step, dt, G, softening}
//   setMaterials(ctx, type: Int32Array|null, seed: Float32Array|null)   Body.type and the material seed, idx:70-74, 492
//   render(ctx, width, height, [camX,camY,camZ, tgtX,tgtY,tgtZ, fovDeg] as 7 numbers, rgba: Float32Array w*h*4)
//                                                          planet shaders idx:83-202, headless
#include "napi_min.h"

#include <algorithm>
#include <cstdio>
#include <cstring>

#include "../../include/nbody.h"

namespace {

struct Handle {
    nbody_ctx* ctx;
    nbody_prec prec;
};

napi_value undefined(napi_env env)
{
    napi_value u = nullptr;
    napi_get_undefined(env, &u);
    return u;
}

napi_value fail(napi_env env, const char* what, int status, const nbody_ctx* ctx)
{
    char msg[640];
    const char* detail = ctx ? nbody_last_error(ctx) : "";
    snprintf(msg, sizeof msg, "nbody.%s: %s%s%s", what, nbody_strerror(status), detail[0] ? ": " : "", detail);
    napi_throw_error(env, "NBODY", msg);
    return undefined(env);
}

struct Args {
    size_t argc = 12;
    napi_value v[12] = {};
};

bool get_args(napi_env env, napi_callback_info info, Args& a, size_t need, const char* what)
{
    if (napi_get_cb_info(env, info, &a.argc, a.v, nullptr, nullptr) != napi_ok || a.argc < need) {
        char msg[96];
        snprintf(msg, sizeof msg, "nbody.%s: expected %zu arguments", what, need);
        napi_throw_error(env, "NBODY", msg);
        return false;
    }
    return true;
}

Handle* get_handle(napi_env env, napi_value v, const char* what)
{
    void* p = nullptr;
    if (napi_get_value_external(env, v, &p) != napi_ok || !p || !static_cast<Handle*>(p)->ctx) {
        char msg[96];
        snprintf(msg, sizeof msg, "nbody.%s: first argument is not a live context", what);
        napi_throw_error(env, "NBODY", msg);
        return nullptr;
    }
    return static_cast<Handle*>(p);
}

bool get_double(napi_env env, napi_value v, double* out) { return napi_get_value_double(env, v, out) == napi_ok; }

// null/undefined -> (nullptr, 0).  Otherwise the array must match the context precision.
bool get_array(napi_env env, napi_value v, nbody_prec prec, void** data, size_t* len, const char* what)
{
    *data = nullptr;
    *len = 0;
    napi_valuetype t = napi_undefined;
    napi_typeof(env, v, &t);
    if (t == napi_undefined || t == napi_null) return true;
    bool is_ta = false;
    napi_typedarray_type ty = napi_int8_array;
    if (napi_is_typedarray(env, v, &is_ta) != napi_ok || !is_ta ||
        napi_get_typedarray_info(env, v, &ty, len, data, nullptr, nullptr) != napi_ok ||
        ty != (prec == NBODY_F32 ? napi_float32_array : napi_float64_array)) {
        char msg[160];
        snprintf(msg, sizeof msg, "nbody.%s: arrays must be %s (the context precision)", what,
                 prec == NBODY_F32 ? "Float32Array" : "Float64Array");
        napi_throw_error(env, "NBODY", msg);
        return false;
    }
    return true;
}

void finalize_handle(napi_env, void* data, void*)
{
    Handle* h = static_cast<Handle*>(data);
    if (h->ctx) nbody_destroy(h->ctx);
    delete h;
}

// ---- lifetime ---------------------------------------------------------------

napi_value Create(napi_env env, napi_callback_info info)
{
    Args a;
    if (!get_args(env, info, a, 3, "create")) return undefined(env);
    char prec_s[8] = {};
    size_t n = 0;
    napi_get_value_string_utf8(env, a.v[0], prec_s, sizeof prec_s, &n);
    double n_gpus = 1, capacity = 0;
    if (!get_double(env, a.v[1], &n_gpus) || !get_double(env, a.v[2], &capacity) ||
        (strcmp(prec_s, "f32") != 0 && strcmp(prec_s, "f64") != 0) || capacity < 1) {
        napi_throw_error(env, "NBODY", "nbody.create(precision 'f32'|'f64', nGpus, capacity)");
        return undefined(env);
    }
    Handle* h = new Handle{nullptr, strcmp(prec_s, "f64") == 0 ? NBODY_F64 : NBODY_F32};
    const int rc = nbody_create(&h->ctx, h->prec, int(n_gpus), nullptr, size_t(capacity));
    if (rc != NBODY_OK) {
        delete h;
        return fail(env, "create", rc, nullptr);
    }
    napi_value ext = nullptr;
    napi_create_external(env, h, finalize_handle, nullptr, &ext);
    return ext;
}

napi_value Destroy(napi_env env, napi_callback_info info)
{
    Args a;
    if (!get_args(env, info, a, 1, "destroy")) return undefined(env);
    void* p = nullptr;
    if (napi_get_value_external(env, a.v[0], &p) == napi_ok && p) {
        Handle* h = static_cast<Handle*>(p);
        if (h->ctx) nbody_destroy(h->ctx);
        h->ctx = nullptr;  // the finalizer frees the Handle itself
    }
    return undefined(env);
}

// ---- state --------------------------------------------------------------------

napi_value Upload(napi_env env, napi_callback_info info)
{
    Args a;
    if (!get_args(env, info, a, 3, "upload")) return undefined(env);
    Handle* h = get_handle(env, a.v[0], "upload");
    if (!h) return undefined(env);
    void* p[4] = {};
    size_t len[4] = {};
    for (int k = 0; k < 4; ++k)
        if (size_t(k + 1) < a.argc && !get_array(env, a.v[k + 1], h->prec, &p[k], &len[k], "upload")) return undefined(env);
    const size_t n = len[0] / 4;
    if (!p[0] || !p[1] || len[0] % 4 || len[1] != len[0] || (p[2] && len[2] != len[0]) || (p[3] && len[3] != n)) {
        napi_throw_error(env, "NBODY", "nbody.upload: posMass[4n], vel[4n], acc[4n]|null, temp[n]|null");
        return undefined(env);
    }
    const int rc = nbody_upload(h->ctx, n, p[0], p[1], p[2], p[3]);
    return rc == NBODY_OK ? undefined(env) : fail(env, "upload", rc, h->ctx);
}

napi_value Download(napi_env env, napi_callback_info info)
{
    Args a;
    if (!get_args(env, info, a, 2, "download")) return undefined(env);
    Handle* h = get_handle(env, a.v[0], "download");
    if (!h) return undefined(env);
    nbody_info inf;
    nbody_get_info(h->ctx, &inf);
    void* p[4] = {};
    size_t len[4] = {};
    for (int k = 0; k < 4; ++k) {
        if (size_t(k + 1) >= a.argc) break;
        if (!get_array(env, a.v[k + 1], h->prec, &p[k], &len[k], "download")) return undefined(env);
        if (p[k] && len[k] < (k == 3 ? inf.n : 4 * inf.n)) {
            napi_throw_error(env, "NBODY", "nbody.download: array shorter than the uploaded body count");
            return undefined(env);
        }
    }
    const int rc = nbody_download(h->ctx, p[0], p[1], p[2], p[3]);
    return rc == NBODY_OK ? undefined(env) : fail(env, "download", rc, h->ctx);
}

// ---- the hot path ---------------------------------------------------------------

// step(ctx, dt, G, softening, nsteps): nsteps x PhysicsSystem.integrate(dt), idx:247-262
napi_value Step(napi_env env, napi_callback_info info)
{
    Args a;
    if (!get_args(env, info, a, 5, "step")) return undefined(env);
    Handle* h = get_handle(env, a.v[0], "step");
    if (!h) return undefined(env);
    double dt, G, eps;
    int32_t nsteps = 1;
    if (!get_double(env, a.v[1], &dt) || !get_double(env, a.v[2], &G) || !get_double(env, a.v[3], &eps) ||
        napi_get_value_int32(env, a.v[4], &nsteps) != napi_ok) {
        napi_throw_error(env, "NBODY", "nbody.step(ctx, dt, G, softening, nsteps): numbers expected");
        return undefined(env);
    }
    const int rc = nbody_step(h->ctx, dt, G, eps, nsteps);
    return rc == NBODY_OK ? undefined(env) : fail(env, "step", rc, h->ctx);
}

// stepHost(ctx, posMass, vel, acc, temp|null, dt, G, softening, nsteps): one frame of the shim in one call -- the
// typed arrays are uploaded as the body list, advanced by nsteps x integrate(dt) and overwritten in place (nbody_step_host)
napi_value StepHost(napi_env env, napi_callback_info info)
{
    Args a;
    if (!get_args(env, info, a, 9, "stepHost")) return undefined(env);
    Handle* h = get_handle(env, a.v[0], "stepHost");
    if (!h) return undefined(env);
    void* p[4] = {};
    size_t len[4] = {};
    for (int k = 0; k < 4; ++k)
        if (!get_array(env, a.v[k + 1], h->prec, &p[k], &len[k], "stepHost")) return undefined(env);
    const size_t n = len[0] / 4;
    if (!p[0] || !p[1] || !p[2] || len[0] % 4 || len[1] != len[0] || len[2] != len[0] || (p[3] && len[3] != n)) {
        napi_throw_error(env, "NBODY", "nbody.stepHost: posMass[4n], vel[4n], acc[4n], temp[n]|null");
        return undefined(env);
    }
    double dt, G, eps;
    int32_t nsteps = 1;
    if (!get_double(env, a.v[5], &dt) || !get_double(env, a.v[6], &G) || !get_double(env, a.v[7], &eps) ||
        napi_get_value_int32(env, a.v[8], &nsteps) != napi_ok) {
        napi_throw_error(env, "NBODY", "nbody.stepHost(ctx, posMass, vel, acc, temp, dt, G, softening, nsteps): numbers expected");
        return undefined(env);
    }
    const int rc = nbody_step_host(h->ctx, n, p[0], p[1], p[2], p[3], dt, G, eps, nsteps);
    return rc == NBODY_OK ? undefined(env) : fail(env, "stepHost", rc, h->ctx);
}

napi_value Kick1Drift(napi_env env, napi_callback_info info)  // idx:248-253
{
    Args a;
    if (!get_args(env, info, a, 2, "kick1Drift")) return undefined(env);
    Handle* h = get_handle(env, a.v[0], "kick1Drift");
    double dt;
    if (!h || !get_double(env, a.v[1], &dt)) return undefined(env);
    const int rc = nbody_kick1_drift(h->ctx, dt);
    return rc == NBODY_OK ? undefined(env) : fail(env, "kick1Drift", rc, h->ctx);
}

napi_value Force(napi_env env, napi_callback_info info)  // computeGravity, idx:264-291
{
    Args a;
    if (!get_args(env, info, a, 3, "force")) return undefined(env);
    Handle* h = get_handle(env, a.v[0], "force");
    double G, eps;
    if (!h || !get_double(env, a.v[1], &G) || !get_double(env, a.v[2], &eps)) return undefined(env);
    const int rc = nbody_force(h->ctx, G, eps);
    return rc == NBODY_OK ? undefined(env) : fail(env, "force", rc, h->ctx);
}

napi_value Kick2(napi_env env, napi_callback_info info)  // idx:258-261
{
    Args a;
    if (!get_args(env, info, a, 3, "kick2")) return undefined(env);
    Handle* h = get_handle(env, a.v[0], "kick2");
    double dt, decay;
    if (!h || !get_double(env, a.v[1], &dt) || !get_double(env, a.v[2], &decay)) return undefined(env);
    const int rc = nbody_kick2(h->ctx, dt, decay);
    return rc == NBODY_OK ? undefined(env) : fail(env, "kick2", rc, h->ctx);
}

napi_value SetHeatDecay(napi_env env, napi_callback_info info)  // CONFIG.heatDecay, idx:67
{
    Args a;
    if (!get_args(env, info, a, 2, "setHeatDecay")) return undefined(env);
    Handle* h = get_handle(env, a.v[0], "setHeatDecay");
    double decay;
    if (!h || !get_double(env, a.v[1], &decay)) return undefined(env);
    nbody_set_heat_decay(h->ctx, decay);
    return undefined(env);
}

// ---- diagnostics ------------------------------------------------------------------

napi_value Energy(napi_env env, napi_callback_info info)
{
    Args a;
    if (!get_args(env, info, a, 3, "energy")) return undefined(env);
    Handle* h = get_handle(env, a.v[0], "energy");
    double G, eps;
    if (!h || !get_double(env, a.v[1], &G) || !get_double(env, a.v[2], &eps)) return undefined(env);
    double ke = 0, pe = 0, mom[3] = {0, 0, 0};
    const int rc = nbody_energy(h->ctx, G, eps, &ke, &pe, mom);
    if (rc != NBODY_OK) return fail(env, "energy", rc, h->ctx);
    napi_value obj, v, arr;
    napi_create_object(env, &obj);
    napi_create_double(env, ke, &v);
    napi_set_named_property(env, obj, "kinetic", v);
    napi_create_double(env, pe, &v);
    napi_set_named_property(env, obj, "potential", v);
    napi_create_array_with_length(env, 3, &arr);
    for (uint32_t k = 0; k < 3; ++k) {
        napi_create_double(env, mom[k], &v);
        napi_set_element(env, arr, k, v);
    }
    napi_set_named_property(env, obj, "momentum", arr);
    return obj;
}

napi_value Sync(napi_env env, napi_callback_info info)
{
    Args a;
    if (!get_args(env, info, a, 1, "sync")) return undefined(env);
    Handle* h = get_handle(env, a.v[0], "sync");
    if (!h) return undefined(env);
    const int rc = nbody_sync(h->ctx);
    return rc == NBODY_OK ? undefined(env) : fail(env, "sync", rc, h->ctx);
}

napi_value Info(napi_env env, napi_callback_info info)
{
    Args a;
    if (!get_args(env, info, a, 1, "info")) return undefined(env);
    Handle* h = get_handle(env, a.v[0], "info");
    if (!h) return undefined(env);
    nbody_info inf;
    const int rc = nbody_get_info(h->ctx, &inf);
    if (rc != NBODY_OK) return fail(env, "info", rc, h->ctx);
    napi_value obj, v;
    napi_create_object(env, &obj);
    const struct { const char* k; double x; } rows[] = {
        {"n", double(inf.n)}, {"nPadded", double(inf.n_padded)}, {"nGpus", double(inf.n_gpus)},
        {"jChunks", double(inf.j_chunks)}, {"iTile", double(inf.i_tile)}, {"smCount", double(inf.sm_count)},
        {"kernelLaunches", double(inf.kernel_launches)}};
    for (const auto& r : rows) {
        napi_create_double(env, r.x, &v);
        napi_set_named_property(env, obj, r.k, v);
    }
    return obj;
}

napi_value Version(napi_env env, napi_callback_info)
{
    napi_value s = nullptr;
    napi_create_string_utf8(env, nbody_version(), NAPI_AUTO_LENGTH, &s);
    return s;
}

napi_value DeviceCount(napi_env env, napi_callback_info)
{
    napi_value v = nullptr;
    napi_create_int32(env, nbody_device_count(), &v);
    return v;
}

// ---- collision broad phase, renderer feed, snapshots --------

bool get_typed(napi_env env, napi_value v, napi_typedarray_type want, void** data, size_t* len)
{
    bool is_ta = false;
    napi_typedarray_type ty = napi_int8_array;
    return napi_is_typedarray(env, v, &is_ta) == napi_ok && is_ta &&
           napi_get_typedarray_info(env, v, &ty, len, data, nullptr, nullptr) == napi_ok && ty == want;
}

napi_value number(napi_env env, double x)
{
    napi_value v = nullptr;
    napi_create_double(env, x, &v);
    return v;
}

napi_value SetRadius(napi_env env, napi_callback_info info)  // Body.radius, idx:221
{
    Args a;
    if (!get_args(env, info, a, 2, "setRadius")) return undefined(env);
    Handle* h = get_handle(env, a.v[0], "setRadius");
    if (!h) return undefined(env);
    void* p = nullptr;
    size_t len = 0;
    if (!get_array(env, a.v[1], h->prec, &p, &len, "setRadius")) return undefined(env);
    nbody_info inf;
    nbody_get_info(h->ctx, &inf);
    if (p && len < inf.n) {
        napi_throw_error(env, "NBODY", "nbody.setRadius: array shorter than the uploaded body count");
        return undefined(env);
    }
    const int rc = nbody_set_radius(h->ctx, p);
    return rc == NBODY_OK ? undefined(env) : fail(env, "setRadius", rc, h->ctx);
}

napi_value Contacts(napi_env env, napi_callback_info info)  // the pair search of idx:301-313
{
    Args a;
    if (!get_args(env, info, a, 3, "contacts")) return undefined(env);
    Handle* h = get_handle(env, a.v[0], "contacts");
    double margin;
    void* p = nullptr;
    size_t len = 0;
    if (!h || !get_double(env, a.v[1], &margin)) return undefined(env);
    if (!get_typed(env, a.v[2], napi_uint32_array, &p, &len)) {
        napi_throw_error(env, "NBODY", "nbody.contacts(ctx, margin, pairs: Uint32Array)");
        return undefined(env);
    }
    size_t count = 0;
    const int rc = nbody_contacts(h->ctx, margin, static_cast<unsigned*>(p), len / 2, &count);
    if (rc != NBODY_OK && !(rc == NBODY_ENOMEM && count > len / 2)) return fail(env, "contacts", rc, h->ctx);
    return number(env, double(count));
}

napi_value HistoryEnable(napi_env env, napi_callback_info info)
{
    Args a;
    if (!get_args(env, info, a, 2, "historyEnable")) return undefined(env);
    Handle* h = get_handle(env, a.v[0], "historyEnable");
    int32_t depth = 0;
    if (!h || napi_get_value_int32(env, a.v[1], &depth) != napi_ok) return undefined(env);
    const int rc = nbody_history_enable(h->ctx, depth);
    return rc == NBODY_OK ? undefined(env) : fail(env, "historyEnable", rc, h->ctx);
}

napi_value HistoryPush(napi_env env, napi_callback_info info)  // history.unshift / pop, idx:564-565
{
    Args a;
    if (!get_args(env, info, a, 1, "historyPush")) return undefined(env);
    Handle* h = get_handle(env, a.v[0], "historyPush");
    if (!h) return undefined(env);
    const int rc = nbody_history_push(h->ctx);
    return rc == NBODY_OK ? undefined(env) : fail(env, "historyPush", rc, h->ctx);
}

napi_value HistoryRemap(napi_env env, napi_callback_info info)  // Body objects carry their history through list changes
{
    Args a;
    if (!get_args(env, info, a, 2, "historyRemap")) return undefined(env);
    Handle* h = get_handle(env, a.v[0], "historyRemap");
    if (!h) return undefined(env);
    nbody_info inf;
    nbody_get_info(h->ctx, &inf);
    void* p = nullptr;
    size_t len = 0;
    if (!get_typed(env, a.v[1], napi_int32_array, &p, &len) || len < inf.n) {
        napi_throw_error(env, "NBODY", "nbody.historyRemap(ctx, source: Int32Array[n])");
        return undefined(env);
    }
    const int rc = nbody_history_remap(h->ctx, static_cast<const int*>(p));
    return rc == NBODY_OK ? undefined(env) : fail(env, "historyRemap", rc, h->ctx);
}

napi_value HistoryDownload(napi_env env, napi_callback_info info)
{
    Args a;
    if (!get_args(env, info, a, 2, "historyDownload")) return undefined(env);
    Handle* h = get_handle(env, a.v[0], "historyDownload");
    if (!h) return undefined(env);
    void* p = nullptr;
    size_t len = 0;
    if (!get_array(env, a.v[1], h->prec, &p, &len, "historyDownload") || !p) return undefined(env);
    const int rc = nbody_history_download(h->ctx, p);  // the shim sizes `out` as n*depth*3
    return rc == NBODY_OK ? undefined(env) : fail(env, "historyDownload", rc, h->ctx);
}

napi_value TrailRibbons(napi_env env, napi_callback_info info)  // idx:567-592
{
    Args a;
    if (!get_args(env, info, a, 5, "trailRibbons")) return undefined(env);
    Handle* h = get_handle(env, a.v[0], "trailRibbons");
    double cam[3];
    void* p = nullptr;
    size_t len = 0;
    if (!h || !get_double(env, a.v[1], &cam[0]) || !get_double(env, a.v[2], &cam[1]) || !get_double(env, a.v[3], &cam[2])) return undefined(env);
    if (!get_typed(env, a.v[4], napi_float32_array, &p, &len)) {
        napi_throw_error(env, "NBODY", "nbody.trailRibbons(ctx, camX, camY, camZ, out: Float32Array)");
        return undefined(env);
    }
    const int rc = nbody_trail_ribbons(h->ctx, cam, static_cast<float*>(p));
    return rc == NBODY_OK ? undefined(env) : fail(env, "trailRibbons", rc, h->ctx);
}

napi_value SnapshotSave(napi_env env, napi_callback_info info)
{
    Args a;
    if (!get_args(env, info, a, 6, "snapshotSave")) return undefined(env);
    Handle* h = get_handle(env, a.v[0], "snapshotSave");
    char path[1024] = {};
    size_t n = 0;
    double step, dt, G, eps;
    if (!h || napi_get_value_string_utf8(env, a.v[1], path, sizeof path, &n) != napi_ok || !get_double(env, a.v[2], &step) ||
        !get_double(env, a.v[3], &dt) || !get_double(env, a.v[4], &G) || !get_double(env, a.v[5], &eps))
        return undefined(env);
    const int rc = nbody_snapshot_save(h->ctx, path, (unsigned long long)step, dt, G, eps);
    return rc == NBODY_OK ? undefined(env) : fail(env, "snapshotSave", rc, h->ctx);
}

napi_value SnapshotLoad(napi_env env, napi_callback_info info)
{
    Args a;
    if (!get_args(env, info, a, 2, "snapshotLoad")) return undefined(env);
    Handle* h = get_handle(env, a.v[0], "snapshotLoad");
    char path[1024] = {};
    size_t n = 0;
    if (!h || napi_get_value_string_utf8(env, a.v[1], path, sizeof path, &n) != napi_ok) return undefined(env);
    unsigned long long step = 0;
    double dt = 0, G = 0, eps = 0;
    const int rc = nbody_snapshot_load(h->ctx, path, &step, &dt, &G, &eps);
    if (rc != NBODY_OK) return fail(env, "snapshotLoad", rc, h->ctx);
    napi_value obj;
    napi_create_object(env, &obj);
    napi_set_named_property(env, obj, "step", number(env, double(step)));
    napi_set_named_property(env, obj, "dt", number(env, dt));
    napi_set_named_property(env, obj, "G", number(env, G));
    napi_set_named_property(env, obj, "softening", number(env, eps));
    return obj;
}

napi_value SetMaterials(napi_env env, napi_callback_info info)  // Body.type + material seed, idx:70-74, 492
{
    Args a;
    if (!get_args(env, info, a, 3, "setMaterials")) return undefined(env);
    Handle* h = get_handle(env, a.v[0], "setMaterials");
    if (!h) return undefined(env);
    nbody_info inf;
    nbody_get_info(h->ctx, &inf);
    void* p[2] = {nullptr, nullptr};
    const napi_typedarray_type want[2] = {napi_int32_array, napi_float32_array};
    for (int k = 0; k < 2; ++k) {
        napi_valuetype t = napi_undefined;
        napi_typeof(env, a.v[k + 1], &t);
        if (t == napi_undefined || t == napi_null) continue;
        size_t len = 0;
        if (!get_typed(env, a.v[k + 1], want[k], &p[k], &len) || len < inf.n) {
            napi_throw_error(env, "NBODY", "nbody.setMaterials(ctx, type: Int32Array[n]|null, seed: Float32Array[n]|null)");
            return undefined(env);
        }
    }
    const int rc = nbody_set_materials(h->ctx, static_cast<const int*>(p[0]), static_cast<const float*>(p[1]));
    return rc == NBODY_OK ? undefined(env) : fail(env, "setMaterials", rc, h->ctx);
}

// ---- the page's complete step on the device (include/nbody.h, nbody_world_*) ----------------

// worldEnable(ctx, seed, maxBodies [, fractureThreshold, minFragmentMass, mergeTime]): CONFIG idx:57, 62-64
napi_value WorldEnable(napi_env env, napi_callback_info info)
{
    Args a;
    if (!get_args(env, info, a, 3, "worldEnable")) return undefined(env);
    Handle* h = get_handle(env, a.v[0], "worldEnable");
    if (!h) return undefined(env);
    double seed = 0, max_bodies = 0, cfg[3] = {25.0, 0.2, 0.5};
    if (!get_double(env, a.v[1], &seed) || !get_double(env, a.v[2], &max_bodies) || seed < 0 || max_bodies < 1) {
        napi_throw_error(env, "NBODY", "nbody.worldEnable(ctx, seed, maxBodies[, fractureThreshold, minFragmentMass, mergeTime])");
        return undefined(env);
    }
    for (int k = 0; k < 3 && size_t(k + 3) < a.argc; ++k) get_double(env, a.v[k + 3], &cfg[k]);
    int rc = nbody_world_enable(h->ctx, static_cast<unsigned long long>(seed), size_t(max_bodies));
    if (rc == NBODY_OK) rc = nbody_world_config(h->ctx, cfg[0], cfg[1], cfg[2]);
    return rc == NBODY_OK ? undefined(env) : fail(env, "worldEnable", rc, h->ctx);
}

// worldSetBodies(ctx, id: Float64Array[n], type: Int32Array[n], radius: Float64Array[n]): Body.id / .type / .radius, idx:210-221
napi_value WorldSetBodies(napi_env env, napi_callback_info info)
{
    Args a;
    if (!get_args(env, info, a, 4, "worldSetBodies")) return undefined(env);
    Handle* h = get_handle(env, a.v[0], "worldSetBodies");
    if (!h) return undefined(env);
    nbody_info inf;
    nbody_get_info(h->ctx, &inf);
    void* p[3] = {};
    const napi_typedarray_type want[3] = {napi_float64_array, napi_int32_array, napi_float64_array};
    for (int k = 0; k < 3; ++k) {
        size_t len = 0;
        if (!get_typed(env, a.v[k + 1], want[k], &p[k], &len) || len < inf.n) {
            napi_throw_error(env, "NBODY", "nbody.worldSetBodies(ctx, id: Float64Array[n], type: Int32Array[n], radius: Float64Array[n])");
            return undefined(env);
        }
    }
    const int rc = nbody_world_set_bodies(h->ctx, static_cast<const double*>(p[0]), static_cast<const int*>(p[1]), static_cast<const double*>(p[2]));
    return rc == NBODY_OK ? undefined(env) : fail(env, "worldSetBodies", rc, h->ctx);
}

// worldStep(ctx, dt, G, softening, nsteps, collisions): nsteps x the COMPLETE integrate(dt), idx:247-262 with idx:256
napi_value WorldStep(napi_env env, napi_callback_info info)
{
    Args a;
    if (!get_args(env, info, a, 6, "worldStep")) return undefined(env);
    Handle* h = get_handle(env, a.v[0], "worldStep");
    if (!h) return undefined(env);
    double dt, G, eps;
    int32_t nsteps = 1, collisions = 1;
    if (!get_double(env, a.v[1], &dt) || !get_double(env, a.v[2], &G) || !get_double(env, a.v[3], &eps) ||
        napi_get_value_int32(env, a.v[4], &nsteps) != napi_ok || napi_get_value_int32(env, a.v[5], &collisions) != napi_ok) {
        napi_throw_error(env, "NBODY", "nbody.worldStep(ctx, dt, G, softening, nsteps, collisions: 0|1)");
        return undefined(env);
    }
    const int rc = nbody_world_step(h->ctx, dt, G, eps, nsteps, collisions ? 1 : 0);
    return rc == NBODY_OK ? undefined(env) : fail(env, "worldStep", rc, h->ctx);
}

// worldSync(ctx, id: Float64Array|null, type: Int32Array|null, radius: Float64Array|null) -> the body count the device holds now;
// the arrays (capacity entries) receive Body.id / .type / .radius of the current list.  nbody.download then reads the new list.
napi_value WorldSync(napi_env env, napi_callback_info info)
{
    Args a;
    if (!get_args(env, info, a, 1, "worldSync")) return undefined(env);
    Handle* h = get_handle(env, a.v[0], "worldSync");
    if (!h) return undefined(env);
    size_t n = 0;
    int rc = nbody_world_sync(h->ctx, &n, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr);
    if (rc != NBODY_OK) return fail(env, "worldSync", rc, h->ctx);
    void* p[3] = {};
    const napi_typedarray_type want[3] = {napi_float64_array, napi_int32_array, napi_float64_array};
    for (int k = 0; k < 3 && size_t(k + 1) < a.argc; ++k) {
        napi_valuetype t = napi_undefined;
        napi_typeof(env, a.v[k + 1], &t);
        if (t == napi_undefined || t == napi_null) continue;
        size_t len = 0;
        if (!get_typed(env, a.v[k + 1], want[k], &p[k], &len) || len < n) {
            napi_throw_error(env, "NBODY", "nbody.worldSync(ctx, id: Float64Array|null, type: Int32Array|null, radius: Float64Array|null): arrays too short");
            return undefined(env);
        }
    }
    rc = nbody_world_sync(h->ctx, nullptr, static_cast<double*>(p[0]), static_cast<int*>(p[1]), static_cast<double*>(p[2]), nullptr, nullptr, nullptr);
    if (rc != NBODY_OK) return fail(env, "worldSync", rc, h->ctx);
    napi_value out = nullptr;
    napi_create_double(env, double(n), &out);
    return out;
}

// worldFrame(ctx, dt, G, softening, nsteps, collisions, posMass, vel, acc, temp, id, type, radius [, stats: Float64Array[6]])
// -> the body count now.  The page's frame (idx:873-883) in one launch and one synchronise: the Float64Arrays (Int32Array for
// type) receive the state visuals.update reads; stats = merges, fractures, overlaps, flashes, explosions, contacts.
napi_value WorldFrame(napi_env env, napi_callback_info info)
{
    napi_value argv[14] = {};
    size_t argc = 14;
    const char* usage = "nbody.worldFrame(ctx, dt, G, softening, nsteps, collisions: 0|1, posMass, vel, acc: Float64Array[4*cap], temp, id: "
                        "Float64Array[cap], type: Int32Array[cap], radius: Float64Array[cap][, stats: Float64Array[6]])";
    if (napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr) != napi_ok || argc < 13) {
        napi_throw_error(env, "NBODY", usage);
        return undefined(env);
    }
    Handle* h = get_handle(env, argv[0], "worldFrame");
    if (!h) return undefined(env);
    double dt, G, eps;
    int32_t nsteps = 1, collisions = 1;
    if (!get_double(env, argv[1], &dt) || !get_double(env, argv[2], &G) || !get_double(env, argv[3], &eps) ||
        napi_get_value_int32(env, argv[4], &nsteps) != napi_ok || napi_get_value_int32(env, argv[5], &collisions) != napi_ok) {
        napi_throw_error(env, "NBODY", usage);
        return undefined(env);
    }
    void* p[8] = {};
    size_t len[8] = {};
    const napi_typedarray_type want[8] = {napi_float64_array, napi_float64_array, napi_float64_array, napi_float64_array,
                                          napi_float64_array, napi_int32_array,   napi_float64_array, napi_float64_array};
    for (int k = 0; k < 8 && size_t(k + 6) < argc; ++k) {
        if (k == 7) {
            napi_valuetype t = napi_undefined;
            napi_typeof(env, argv[13], &t);
            if (t == napi_undefined || t == napi_null) continue;
        }
        if (!get_typed(env, argv[k + 6], want[k], &p[k], &len[k]) || (k == 7 && len[k] < 6)) {
            napi_throw_error(env, "NBODY", usage);
            return undefined(env);
        }
    }
    size_t cap = len[3];  // rows every array has room for
    for (int k = 0; k < 7; ++k) cap = std::min(cap, k < 3 ? len[k] / 4 : len[k]);
    size_t n = 0, contacts = 0;
    long long counters[5] = {};
    const int rc = nbody_world_frame(h->ctx, dt, G, eps, nsteps, collisions ? 1 : 0, cap, &n, static_cast<double*>(p[0]), static_cast<double*>(p[1]),
                                     static_cast<double*>(p[2]), static_cast<double*>(p[3]), static_cast<double*>(p[4]), static_cast<int*>(p[5]),
                                     static_cast<double*>(p[6]), counters, &contacts, nullptr);
    if (rc != NBODY_OK) return fail(env, "worldFrame", rc, h->ctx);
    if (p[7]) {
        double* st = static_cast<double*>(p[7]);
        for (int k = 0; k < 5; ++k) st[k] = double(counters[k]);
        st[5] = double(contacts);
    }
    napi_value out = nullptr;
    napi_create_double(env, double(n), &out);
    return out;
}

napi_value Render(napi_env env, napi_callback_info info)  // planet shaders, idx:83-202
{
    Args a;
    a.argc = 11;
    napi_value argv[11] = {};
    size_t argc = 11;
    if (napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr) != napi_ok || argc < 11) {
        napi_throw_error(env, "NBODY", "nbody.render(ctx, width, height, camX, camY, camZ, tgtX, tgtY, tgtZ, fovDeg, rgba: Float32Array)");
        return undefined(env);
    }
    Handle* h = get_handle(env, argv[0], "render");
    if (!h) return undefined(env);
    double v[9];
    for (int k = 0; k < 9; ++k)
        if (!get_double(env, argv[k + 1], &v[k])) {
            napi_throw_error(env, "NBODY", "nbody.render: numbers expected");
            return undefined(env);
        }
    void* p = nullptr;
    size_t len = 0;
    if (!get_typed(env, argv[10], napi_float32_array, &p, &len) || v[0] < 1 || v[1] < 1 || len < size_t(v[0]) * size_t(v[1]) * 4) {
        napi_throw_error(env, "NBODY", "nbody.render: rgba must be a Float32Array of width*height*4");
        return undefined(env);
    }
    const int rc = nbody_render(h->ctx, int(v[0]), int(v[1]), &v[2], &v[5], v[8], static_cast<float*>(p), nullptr);
    return rc == NBODY_OK ? undefined(env) : fail(env, "render", rc, h->ctx);
}

}  // namespace

// Node loads the addon through this well-known symbol (what NAPI_MODULE_INIT expands to).
extern "C" __attribute__((visibility("default"))) napi_value napi_register_module_v1(napi_env env, napi_value exports)
{
    const struct { const char* name; napi_callback cb; } fns[] = {
        {"create", Create}, {"destroy", Destroy}, {"upload", Upload}, {"download", Download},
        {"step", Step}, {"stepHost", StepHost}, {"kick1Drift", Kick1Drift}, {"force", Force}, {"kick2", Kick2},
        {"setHeatDecay", SetHeatDecay}, {"energy", Energy}, {"sync", Sync}, {"info", Info},
        {"version", Version}, {"deviceCount", DeviceCount},
        {"setRadius", SetRadius}, {"contacts", Contacts}, {"historyEnable", HistoryEnable}, {"historyPush", HistoryPush},
        {"historyRemap", HistoryRemap},
        {"historyDownload", HistoryDownload}, {"trailRibbons", TrailRibbons}, {"snapshotSave", SnapshotSave},
        {"snapshotLoad", SnapshotLoad}, {"setMaterials", SetMaterials}, {"render", Render},
        {"worldEnable", WorldEnable}, {"worldSetBodies", WorldSetBodies}, {"worldStep", WorldStep}, {"worldSync", WorldSync}, {"worldFrame", WorldFrame}};
    for (const auto& f : fns) {
        napi_value fn = nullptr;
        if (napi_create_function(env, f.name, NAPI_AUTO_LENGTH, f.cb, nullptr, &fn) != napi_ok) return exports;
        napi_set_named_property(env, exports, f.name, fn);
    }
    return exports;
}
