/*
 * napi_min.h -- the subset of Node-API (N-API version 8, a stable C ABI) that
 * nbody_napi.cc uses.  Neither Node.js nor node_api.h exists in this image, so
 * the addon ships these declarations itself; when node_api.h IS available,
 * compile with -DNBODY_HAVE_NODE_API_H and this file steps aside.  The napi_*
 * functions are left undefined in nbody_napi.node and resolved by the Node
 * process (or by tests/napi_mock_host.cc) at load time.
 */
#ifndef NBODY_NAPI_MIN_H_
#define NBODY_NAPI_MIN_H_
#ifdef NBODY_HAVE_NODE_API_H
#include <node_api.h>
#else
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct napi_env__* napi_env;
typedef struct napi_value__* napi_value;
typedef struct napi_callback_info__* napi_callback_info;

typedef enum {
    napi_ok = 0, napi_invalid_arg, napi_object_expected, napi_string_expected, napi_name_expected,
    napi_function_expected, napi_number_expected, napi_boolean_expected, napi_array_expected,
    napi_generic_failure, napi_pending_exception
} napi_status;

typedef enum {
    napi_undefined = 0, napi_null, napi_boolean, napi_number, napi_string, napi_symbol, napi_object,
    napi_function, napi_external, napi_bigint
} napi_valuetype;

typedef enum {
    napi_int8_array = 0, napi_uint8_array, napi_uint8_clamped_array, napi_int16_array,
    napi_uint16_array, napi_int32_array, napi_uint32_array, napi_float32_array,
    napi_float64_array, napi_bigint64_array, napi_biguint64_array
} napi_typedarray_type;

typedef napi_value (*napi_callback)(napi_env env, napi_callback_info info);
typedef void (*napi_finalize)(napi_env env, void* finalize_data, void* finalize_hint);

napi_status napi_create_function(napi_env env, const char* utf8name, size_t length, napi_callback cb,
                                 void* data, napi_value* result);
napi_status napi_set_named_property(napi_env env, napi_value object, const char* utf8name, napi_value value);
napi_status napi_get_cb_info(napi_env env, napi_callback_info cbinfo, size_t* argc, napi_value* argv,
                             napi_value* this_arg, void** data);
napi_status napi_typeof(napi_env env, napi_value value, napi_valuetype* result);
napi_status napi_is_typedarray(napi_env env, napi_value value, bool* result);
napi_status napi_get_typedarray_info(napi_env env, napi_value typedarray, napi_typedarray_type* type,
                                     size_t* length, void** data, napi_value* arraybuffer, size_t* byte_offset);
napi_status napi_get_value_double(napi_env env, napi_value value, double* result);
napi_status napi_get_value_int32(napi_env env, napi_value value, int32_t* result);
napi_status napi_get_value_string_utf8(napi_env env, napi_value value, char* buf, size_t bufsize, size_t* result);
napi_status napi_create_double(napi_env env, double value, napi_value* result);
napi_status napi_create_int32(napi_env env, int32_t value, napi_value* result);
napi_status napi_create_string_utf8(napi_env env, const char* str, size_t length, napi_value* result);
napi_status napi_create_object(napi_env env, napi_value* result);
napi_status napi_create_array_with_length(napi_env env, size_t length, napi_value* result);
napi_status napi_set_element(napi_env env, napi_value object, uint32_t index, napi_value value);
napi_status napi_get_undefined(napi_env env, napi_value* result);
napi_status napi_create_external(napi_env env, void* data, napi_finalize finalize_cb, void* finalize_hint,
                                 napi_value* result);
napi_status napi_get_value_external(napi_env env, napi_value value, void** result);
napi_status napi_throw_error(napi_env env, const char* code, const char* msg);

#define NAPI_AUTO_LENGTH ((size_t)-1)

#ifdef __cplusplus
}
#endif
#endif /* NBODY_HAVE_NODE_API_H */
#endif /* NBODY_NAPI_MIN_H_ */
