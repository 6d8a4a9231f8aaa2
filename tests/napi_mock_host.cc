// napi_mock_host.cc -- TEST INFRASTRUCTURE.  A stand-in for the Node.js process:
// it implements the ~20 Node-API functions nbody_napi.node imports (a tiny
// tagged-value model, no JS engine), dlopens the addon, calls its
// napi_register_module_v1 and then drives the exported functions the way
// physics_shim.js does.  Node.js does not exist in this image; this is how the
// addon's glue code is executed on the GPU box (tests/test_gpu_addon.py).
//
//   napi_mock_host <nbody_napi.node> <in.bin> <out.bin> <nsteps> <step|split> <f64|f32>
//
// in.bin / out.bin: 4 x float64 header (n, dt, G, softening) then pos_mass[4n],
// vel[4n], acc[4n], temp[n] as float64 (converted here for f32 contexts).
// Build: g++ -std=c++17 -rdynamic napi_mock_host.cc -ldl   (-rdynamic exports napi_*).
#include <dlfcn.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <string>
#include <vector>

#include "../n-body-sim_b200/addon/napi_min.h"

struct napi_value__ {
    napi_valuetype type = napi_undefined;
    double num = 0;
    std::string str;
    void* ext = nullptr;
    napi_finalize fin = nullptr;
    bool is_typed = false;
    napi_typedarray_type ta_type = napi_float64_array;
    void* ta_data = nullptr;
    size_t ta_len = 0;
    napi_callback fn = nullptr;
    std::map<std::string, napi_value> props;
    std::vector<napi_value> elems;
};
struct napi_env__ {
    std::vector<std::unique_ptr<napi_value__>> heap;
    bool pending = false;
    std::string error;
    napi_value make(napi_valuetype t)
    {
        heap.emplace_back(new napi_value__());
        heap.back()->type = t;
        return heap.back().get();
    }
};
struct napi_callback_info__ {
    std::vector<napi_value> args;
};

#define EXPORT extern "C" __attribute__((visibility("default")))

EXPORT napi_status napi_create_function(napi_env env, const char*, size_t, napi_callback cb, void*, napi_value* result)
{
    *result = env->make(napi_function);
    (*result)->fn = cb;
    return napi_ok;
}
EXPORT napi_status napi_set_named_property(napi_env, napi_value object, const char* name, napi_value value)
{
    object->props[name] = value;
    return napi_ok;
}
EXPORT napi_status napi_get_cb_info(napi_env, napi_callback_info info, size_t* argc, napi_value* argv, napi_value*, void**)
{
    const size_t cap = *argc;
    for (size_t k = 0; k < cap && k < info->args.size(); ++k) argv[k] = info->args[k];
    *argc = info->args.size();
    return napi_ok;
}
EXPORT napi_status napi_typeof(napi_env, napi_value v, napi_valuetype* result)
{
    *result = v ? v->type : napi_undefined;
    return napi_ok;
}
EXPORT napi_status napi_is_typedarray(napi_env, napi_value v, bool* result)
{
    *result = v && v->is_typed;
    return napi_ok;
}
EXPORT napi_status napi_get_typedarray_info(napi_env, napi_value v, napi_typedarray_type* type, size_t* length, void** data,
                                            napi_value*, size_t*)
{
    if (!v || !v->is_typed) return napi_invalid_arg;
    if (type) *type = v->ta_type;
    if (length) *length = v->ta_len;
    if (data) *data = v->ta_data;
    return napi_ok;
}
EXPORT napi_status napi_get_value_double(napi_env, napi_value v, double* result)
{
    if (!v || v->type != napi_number) return napi_number_expected;
    *result = v->num;
    return napi_ok;
}
EXPORT napi_status napi_get_value_int32(napi_env, napi_value v, int32_t* result)
{
    if (!v || v->type != napi_number) return napi_number_expected;
    *result = int32_t(v->num);
    return napi_ok;
}
EXPORT napi_status napi_get_value_string_utf8(napi_env, napi_value v, char* buf, size_t bufsize, size_t* result)
{
    if (!v || v->type != napi_string) return napi_string_expected;
    const size_t n = v->str.size() < bufsize - 1 ? v->str.size() : bufsize - 1;
    memcpy(buf, v->str.data(), n);
    buf[n] = 0;
    if (result) *result = n;
    return napi_ok;
}
EXPORT napi_status napi_create_double(napi_env env, double value, napi_value* result)
{
    *result = env->make(napi_number);
    (*result)->num = value;
    return napi_ok;
}
EXPORT napi_status napi_create_int32(napi_env env, int32_t value, napi_value* result) { return napi_create_double(env, value, result); }
EXPORT napi_status napi_create_string_utf8(napi_env env, const char* str, size_t length, napi_value* result)
{
    *result = env->make(napi_string);
    (*result)->str = length == NAPI_AUTO_LENGTH ? std::string(str) : std::string(str, length);
    return napi_ok;
}
EXPORT napi_status napi_create_object(napi_env env, napi_value* result)
{
    *result = env->make(napi_object);
    return napi_ok;
}
EXPORT napi_status napi_create_array_with_length(napi_env env, size_t length, napi_value* result)
{
    *result = env->make(napi_object);
    (*result)->elems.resize(length);
    return napi_ok;
}
EXPORT napi_status napi_set_element(napi_env, napi_value object, uint32_t index, napi_value value)
{
    if (object->elems.size() <= index) object->elems.resize(index + 1);
    object->elems[index] = value;
    return napi_ok;
}
EXPORT napi_status napi_get_undefined(napi_env env, napi_value* result)
{
    *result = env->make(napi_undefined);
    return napi_ok;
}
EXPORT napi_status napi_create_external(napi_env env, void* data, napi_finalize fin, void*, napi_value* result)
{
    *result = env->make(napi_external);
    (*result)->ext = data;
    (*result)->fin = fin;
    return napi_ok;
}
EXPORT napi_status napi_get_value_external(napi_env, napi_value v, void** result)
{
    if (!v || v->type != napi_external) return napi_invalid_arg;
    *result = v->ext;
    return napi_ok;
}
EXPORT napi_status napi_throw_error(napi_env env, const char*, const char* msg)
{
    env->pending = true;
    env->error = msg;
    return napi_ok;
}

// ---- the "JS" side ------------------------------------------------------------

static napi_env g_env;
static napi_value g_exports;

static napi_value num(double x)
{
    napi_value v;
    napi_create_double(g_env, x, &v);
    return v;
}
static napi_value str(const char* s)
{
    napi_value v;
    napi_create_string_utf8(g_env, s, NAPI_AUTO_LENGTH, &v);
    return v;
}
static napi_value typed(bool f64, void* data, size_t len)
{
    napi_value v = g_env->make(napi_object);
    v->is_typed = true;
    v->ta_type = f64 ? napi_float64_array : napi_float32_array;
    v->ta_data = data;
    v->ta_len = len;
    return v;
}

static napi_value typed_as(napi_typedarray_type ty, void* data, size_t len)
{
    napi_value v = typed(true, data, len);
    v->ta_type = ty;
    return v;
}

// call exports[name](args...); a thrown "JS Error" ends the program unless expect_throw
static napi_value call(const char* name, std::vector<napi_value> args, bool expect_throw = false)
{
    auto it = g_exports->props.find(name);
    if (it == g_exports->props.end() || !it->second->fn) {
        fprintf(stderr, "mock host: addon does not export %s\n", name);
        exit(3);
    }
    napi_callback_info__ info{std::move(args)};
    g_env->pending = false;
    napi_value r = it->second->fn(g_env, &info);
    if (g_env->pending != expect_throw) {
        fprintf(stderr, "mock host: %s %s: %s\n", name, g_env->pending ? "threw" : "did not throw", g_env->error.c_str());
        exit(4);
    }
    return r;
}

static std::string g_out_path;  // collide mode writes <out>.pairs, <out>.ribbons and <out>.nbs next to the state
static bool g_collide = false;
static bool g_world = false;    // "world" mode: the page's complete step on the device (worldEnable / worldSetBodies / worldStep / worldSync)
static bool g_host = false;     // "host" mode: the frame as ONE addon call (stepHost), arrays in and out in place

template <typename T> static void run(bool f64, const char* prec, const std::vector<double>& in, size_t n, double dt, double G,
                                      double eps, int nsteps, bool split, std::vector<double>& out)
{
    std::vector<T> pm(4 * n), vel(4 * n), acc(4 * n), temp(n);
    for (size_t k = 0; k < 4 * n; ++k) {
        pm[k] = T(in[4 + k]);
        vel[k] = T(in[4 + 4 * n + k]);
        acc[k] = T(in[4 + 8 * n + k]);
    }
    for (size_t k = 0; k < n; ++k) temp[k] = T(in[4 + 12 * n + k]);

    napi_value ctx = call("create", {str(prec), num(1), num(double(n))});
    // error channel: a wrong array type must throw, not crash (the reference has no error path; the shim needs one)
    std::vector<float> wrong32(4 * n);
    std::vector<double> wrong64(4 * n);
    call("upload", {ctx, f64 ? typed(false, wrong32.data(), 4 * n) : typed(true, wrong64.data(), 4 * n),
                    typed(f64, vel.data(), 4 * n)}, true);
    call("step", {ctx, num(dt), num(G), num(eps), num(1)}, true);  // step before upload

    call("setHeatDecay", {ctx, num(0.998)});
    call("upload", {ctx, typed(f64, pm.data(), 4 * n), typed(f64, vel.data(), 4 * n), typed(f64, acc.data(), 4 * n),
                    typed(f64, temp.data(), n)});
    if (g_collide) {
        // physics_shim.js `collisions: 'gpu-broad-phase'`: the overlap search of resolveCollisions (idx:301-313) on the
        // device between force and kick-2, Body.history kept on the device, ribbons and a snapshot at the end
        const int depth = 8;
        std::vector<T> radius(n);
        for (size_t k = 0; k < n; ++k) radius[k] = T(1.5 * std::cbrt(3.0 * double(pm[4 * k + 3]) / (4.0 * 3.14159265358979323846)));
        std::vector<unsigned> pairs(2 * 4096);
        call("contacts", {ctx, num(0), typed_as(napi_uint32_array, pairs.data(), pairs.size())}, true);  // no radii yet: throws
        call("setRadius", {ctx, typed(f64, radius.data(), n)});
        call("historyEnable", {ctx, num(depth)});
        double count = 0;
        for (int s = 0; s < nsteps; ++s) {
            call("kick1Drift", {ctx, num(dt)});
            call("force", {ctx, num(G), num(eps)});
            std::vector<unsigned> tiny(2);
            count = call("contacts", {ctx, num(0), typed_as(napi_uint32_array, tiny.data(), tiny.size())})->num;  // too small: count only
            if (count > 1 && (tiny[0] || tiny[1])) exit(5);
            count = call("contacts", {ctx, num(0), typed_as(napi_uint32_array, pairs.data(), pairs.size())})->num;
            call("kick2", {ctx, num(dt), num(0.998)});
            call("historyPush", {ctx});
        }
        std::vector<float> ribbons(n * depth * 6);
        std::vector<T> hist(n * depth * 3);
        call("trailRibbons", {ctx, num(0), num(80), num(150), typed_as(napi_float32_array, ribbons.data(), ribbons.size())});
        call("historyDownload", {ctx, typed(f64, hist.data(), hist.size())});
        {   // historyRemap: the identity map keeps every ring (then one more push shifts it); a short array throws
            std::vector<int> source(n);
            for (size_t k = 0; k < n; ++k) source[k] = int(k);
            call("historyRemap", {ctx, typed_as(napi_int32_array, source.data(), n - 1)}, true);
            call("historyRemap", {ctx, typed_as(napi_int32_array, source.data(), n)});
            call("historyDownload", {ctx, typed(f64, hist.data(), hist.size())}, true);  // refused until the next push
            call("historyPush", {ctx});
            std::vector<T> hist2(n * depth * 3);
            call("historyDownload", {ctx, typed(f64, hist2.data(), hist2.size())});
            for (size_t k = 0; k < n; ++k)
                for (int i = 1; i < depth; ++i)
                    for (int c = 0; c < 3; ++c)
                        if (hist2[(k * depth + i) * 3 + c] != hist[(k * depth + i - 1) * 3 + c]) exit(7);
        }
        {   // a frame of the planet shaders through the addon (idx:83-202): type = k % 3, seed = k * 0.37
            std::vector<int> type(n);
            std::vector<float> seed(n), frame(96 * 54 * 4);
            for (size_t k = 0; k < n; ++k) {
                type[k] = int(k % 3);
                seed[k] = float(k) * 0.37f;
            }
            call("setMaterials", {ctx, typed_as(napi_int32_array, type.data(), n), typed_as(napi_float32_array, seed.data(), n)});
            call("render", {ctx, num(96), num(54), num(0), num(80), num(150), num(0), num(0), num(0), num(45),
                            typed_as(napi_float32_array, frame.data(), frame.size())});
            call("render", {ctx, num(96), num(54), num(0), num(80), num(150), num(0), num(0), num(0), num(45),
                            typed_as(napi_float32_array, frame.data(), 16)}, true);  // buffer too small: throws
            FILE* fr = fopen((g_out_path + ".frame").c_str(), "wb");
            fwrite(frame.data(), sizeof(float), frame.size(), fr);
            fclose(fr);
        }
        call("snapshotSave", {ctx, str((g_out_path + ".nbs").c_str()), num(nsteps), num(dt), num(G), num(eps)});
        napi_value meta = call("snapshotLoad", {ctx, str((g_out_path + ".nbs").c_str())});
        if (meta->props["step"]->num != nsteps || meta->props["dt"]->num != dt) exit(6);
        call("setRadius", {ctx, typed(f64, radius.data(), n)});
        FILE* fp = fopen((g_out_path + ".pairs").c_str(), "wb");
        fwrite(pairs.data(), sizeof(unsigned), size_t(2 * count), fp);
        fclose(fp);
        fp = fopen((g_out_path + ".ribbons").c_str(), "wb");
        fwrite(ribbons.data(), sizeof(float), ribbons.size(), fp);
        fclose(fp);
        fp = fopen((g_out_path + ".history").c_str(), "wb");
        for (T v : hist) {
            const double d = double(v);
            fwrite(&d, 8, 1, fp);
        }
        fclose(fp);
    } else if (g_world) {
        // fp64 only.  Radii so small that nothing touches: the complete step then equals the gravity-only contract
        std::vector<double> id(n), radius(n, 1e-6);
        std::vector<int> type(n);
        for (size_t k = 0; k < n; ++k) {
            id[k] = double(k + 1) / 1024.0;
            type[k] = int(k % 3);
        }
        call("worldStep", {ctx, num(dt), num(G), num(eps), num(1), num(1)}, true);  // before worldEnable: throws
        call("worldEnable", {ctx, num(12345), num(double(n)), num(25.0), num(0.2), num(0.5)});
        call("worldSetBodies", {ctx, typed_as(napi_float64_array, id.data(), n), typed_as(napi_int32_array, type.data(), n - 1),
                                typed_as(napi_float64_array, radius.data(), n)}, true);  // a short array throws
        call("worldSetBodies", {ctx, typed_as(napi_float64_array, id.data(), n), typed_as(napi_int32_array, type.data(), n),
                                typed_as(napi_float64_array, radius.data(), n)});
        call("worldStep", {ctx, num(dt), num(G), num(eps), num(nsteps), num(1)});
        std::vector<double> id2(n), radius2(n);
        std::vector<int> type2(n);
        const double n_now = call("worldSync", {ctx, typed_as(napi_float64_array, id2.data(), n), typed_as(napi_int32_array, type2.data(), n),
                                                 typed_as(napi_float64_array, radius2.data(), n)})->num;
        if (n_now != double(n) || id2 != id || type2 != type || radius2 != radius) exit(8);
        // worldFrame with nsteps = 0 returns what worldSync + download return, through one call
        std::vector<double> fpm(4 * n), fvel(4 * n), facc(4 * n), ftemp(n), fid(n), frad(n), stats(6, -1.0);
        std::vector<int> ftype(n);
        std::vector<double> dpm(4 * n), dvel(4 * n), dacc(4 * n), dtemp(n);
        call("download", {ctx, typed_as(napi_float64_array, dpm.data(), 4 * n), typed_as(napi_float64_array, dvel.data(), 4 * n),
                          typed_as(napi_float64_array, dacc.data(), 4 * n), typed_as(napi_float64_array, dtemp.data(), n)});
        call("worldFrame", {ctx, num(dt), num(G), num(eps), num(0), num(1), typed_as(napi_float64_array, fpm.data(), 4 * n),
                            typed_as(napi_float64_array, fvel.data(), 4 * n - 4), typed_as(napi_float64_array, facc.data(), 4 * n),
                            typed_as(napi_float64_array, ftemp.data(), n), typed_as(napi_float64_array, fid.data(), n),
                            typed_as(napi_int32_array, ftype.data(), n), typed_as(napi_float64_array, frad.data(), n)}, true);  // one row short: throws
        const double n_frame = call("worldFrame", {ctx, num(dt), num(G), num(eps), num(0), num(1), typed_as(napi_float64_array, fpm.data(), 4 * n),
                                                   typed_as(napi_float64_array, fvel.data(), 4 * n), typed_as(napi_float64_array, facc.data(), 4 * n),
                                                   typed_as(napi_float64_array, ftemp.data(), n), typed_as(napi_float64_array, fid.data(), n),
                                                   typed_as(napi_int32_array, ftype.data(), n), typed_as(napi_float64_array, frad.data(), n),
                                                   typed_as(napi_float64_array, stats.data(), 6)})->num;
        if (n_frame != double(n) || fpm != dpm || fvel != dvel || facc != dacc || ftemp != dtemp || fid != id || ftype != type || frad != radius) exit(9);
        for (double v : stats)
            if (v != 0.0) exit(10);  // nothing touched: no merges, fractures, overlaps, flashes, explosions, contacts
    } else if (g_host) {
        // the arrays uploaded above are handed over again and come back advanced; a bad array must throw, not crash
        call("stepHost", {ctx, typed(f64, pm.data(), 4 * n), typed(f64, vel.data(), 4 * n - 4), typed(f64, acc.data(), 4 * n),
                          typed(f64, temp.data(), n), num(dt), num(G), num(eps), num(nsteps)}, true);
        call("stepHost", {ctx, typed(f64, pm.data(), 4 * n), typed(f64, vel.data(), 4 * n), typed(f64, acc.data(), 4 * n),
                          typed(f64, temp.data(), n), num(dt), num(G), num(eps), num(nsteps)});
        std::vector<T> pm2 = pm, vel2 = vel, acc2 = acc, temp2 = temp;  // what stepHost wrote must equal what download reads
        call("download", {ctx, typed(f64, pm2.data(), 4 * n), typed(f64, vel2.data(), 4 * n), typed(f64, acc2.data(), 4 * n),
                          typed(f64, temp2.data(), n)});
        if (pm2 != pm || vel2 != vel || acc2 != acc || temp2 != temp) exit(7);
    } else if (!split) {
        call("step", {ctx, num(dt), num(G), num(eps), num(nsteps)});
    } else {
        for (int s = 0; s < nsteps; ++s) {  // physics_shim.js default mode with a no-op collision stage
            call("kick1Drift", {ctx, num(dt)});
            call("force", {ctx, num(G), num(eps)});
            call("kick2", {ctx, num(dt), num(0.998)});
        }
    }
    call("download", {ctx, typed(f64, pm.data(), 4 * n), typed(f64, vel.data(), 4 * n), typed(f64, acc.data(), 4 * n),
                      typed(f64, temp.data(), n)});
    napi_value e = call("energy", {ctx, num(G), num(eps)});
    napi_value inf = call("info", {ctx});
    napi_value ver = call("version", {});
    printf("{\"version\": \"%s\", \"n\": %.0f, \"kernel_launches\": %.0f, \"kinetic\": %.17g, \"potential\": %.17g}\n", ver->str.c_str(),
           inf->props["n"]->num, inf->props["kernelLaunches"]->num, e->props["kinetic"]->num, e->props["potential"]->num);
    call("destroy", {ctx});
    call("sync", {ctx}, true);  // a destroyed context is rejected
    if (ctx->fin) ctx->fin(g_env, ctx->ext, nullptr);  // what the GC would do

    out.assign(in.begin(), in.begin() + 4);
    out.resize(4 + 13 * n);
    for (size_t k = 0; k < 4 * n; ++k) {
        out[4 + k] = pm[k];
        out[4 + 4 * n + k] = vel[k];
        out[4 + 8 * n + k] = acc[k];
    }
    for (size_t k = 0; k < n; ++k) out[4 + 12 * n + k] = temp[k];
}

int main(int argc, char** argv)
{
    if (argc < 7) {
        fprintf(stderr, "usage: %s addon.node in.bin out.bin nsteps step|split f64|f32\n", argv[0]);
        return 2;
    }
    void* h = dlopen(argv[1], RTLD_NOW | RTLD_GLOBAL);
    if (!h) {
        fprintf(stderr, "dlopen %s: %s\n", argv[1], dlerror());
        return 2;
    }
    using init_fn = napi_value (*)(napi_env, napi_value);
    init_fn init = reinterpret_cast<init_fn>(dlsym(h, "napi_register_module_v1"));
    if (!init) {
        fprintf(stderr, "addon has no napi_register_module_v1\n");
        return 2;
    }
    napi_env__ env;
    g_env = &env;
    g_exports = init(g_env, env.make(napi_object));

    FILE* f = fopen(argv[2], "rb");
    if (!f) return 2;
    fseek(f, 0, SEEK_END);
    const long bytes = ftell(f);
    fseek(f, 0, SEEK_SET);
    std::vector<double> in(bytes / 8);
    if (fread(in.data(), 8, in.size(), f) != in.size()) return 2;
    fclose(f);
    const size_t n = size_t(in[0]);
    const double dt = in[1], G = in[2], eps = in[3];
    const int nsteps = atoi(argv[4]);
    const bool split = strcmp(argv[5], "split") == 0;
    g_collide = strcmp(argv[5], "collide") == 0;
    g_host = strcmp(argv[5], "host") == 0;
    g_world = strcmp(argv[5], "world") == 0;
    g_out_path = argv[3];
    const bool f64 = strcmp(argv[6], "f64") == 0;
    std::vector<double> out;
    if (f64)
        run<double>(true, "f64", in, n, dt, G, eps, nsteps, split, out);
    else
        run<float>(false, "f32", in, n, dt, G, eps, nsteps, split, out);
    f = fopen(argv[3], "wb");
    fwrite(out.data(), 8, out.size(), f);
    fclose(f);
    return 0;
}
