/*
 * node_api_decl.h — the subset of Node's stable C ABI (N-API, version 6: BigUint64Array) that
 * luxfhe_b200_napi.c uses, declared here ONLY so that the addon can be compiled and its exports checked on a box
 * without Node headers (this build image has no `node` and no `node_api.h`).  With Node installed the real
 * <node_api.h> is used instead (see the #if in luxfhe_b200_napi.c); the declarations below follow its documented
 * signatures, and the symbols are resolved by the node executable when it loads the addon.
 */
#ifndef LUX_NODE_API_DECL_H
#define LUX_NODE_API_DECL_H
#include <stdbool.h>
#include <stddef.h>
#include <stdint.h>

typedef struct napi_env__ *napi_env;
typedef struct napi_value__ *napi_value;
typedef struct napi_callback_info__ *napi_callback_info;
typedef struct napi_deferred__ *napi_deferred;
typedef struct napi_async_work__ *napi_async_work;
typedef struct napi_ref__ *napi_ref;

typedef enum { napi_ok = 0 } napi_status;
typedef enum {
    napi_int8_array,
    napi_uint8_array,
    napi_uint8_clamped_array,
    napi_int16_array,
    napi_uint16_array,
    napi_int32_array,
    napi_uint32_array,
    napi_float32_array,
    napi_float64_array,
    napi_bigint64_array,
    napi_biguint64_array
} napi_typedarray_type;
typedef enum {
    napi_undefined,
    napi_null,
    napi_boolean,
    napi_number,
    napi_string,
    napi_symbol,
    napi_object,
    napi_function,
    napi_external,
    napi_bigint
} napi_valuetype;
typedef enum { napi_default = 0 } napi_property_attributes;

typedef napi_value (*napi_callback)(napi_env env, napi_callback_info info);
typedef void (*napi_finalize)(napi_env env, void *finalize_data, void *finalize_hint);
typedef void (*napi_async_execute_callback)(napi_env env, void *data);
typedef void (*napi_async_complete_callback)(napi_env env, napi_status status, void *data);

typedef struct {
    const char *utf8name;
    napi_value name;
    napi_callback method;
    napi_callback getter;
    napi_callback setter;
    napi_value value;
    napi_property_attributes attributes;
    void *data;
} napi_property_descriptor;

#define NAPI_AUTO_LENGTH SIZE_MAX
#ifdef __cplusplus
#define LUX_NAPI_EXTERN extern "C"
#else
#define LUX_NAPI_EXTERN extern
#endif

LUX_NAPI_EXTERN napi_status napi_get_cb_info(napi_env env, napi_callback_info cbinfo, size_t *argc, napi_value *argv,
                                             napi_value *this_arg, void **data);
LUX_NAPI_EXTERN napi_status napi_throw_error(napi_env env, const char *code, const char *msg);
LUX_NAPI_EXTERN napi_status napi_typeof(napi_env env, napi_value value, napi_valuetype *result);
LUX_NAPI_EXTERN napi_status napi_get_undefined(napi_env env, napi_value *result);
LUX_NAPI_EXTERN napi_status napi_get_value_uint32(napi_env env, napi_value value, uint32_t *result);
LUX_NAPI_EXTERN napi_status napi_get_value_int32(napi_env env, napi_value value, int32_t *result);
LUX_NAPI_EXTERN napi_status napi_create_uint32(napi_env env, uint32_t value, napi_value *result);
LUX_NAPI_EXTERN napi_status napi_create_double(napi_env env, double value, napi_value *result);
LUX_NAPI_EXTERN napi_status napi_create_string_utf8(napi_env env, const char *str, size_t length, napi_value *result);
LUX_NAPI_EXTERN napi_status napi_get_named_property(napi_env env, napi_value object, const char *utf8name,
                                                    napi_value *result);
LUX_NAPI_EXTERN napi_status napi_has_named_property(napi_env env, napi_value object, const char *utf8name, bool *result);
LUX_NAPI_EXTERN napi_status napi_define_properties(napi_env env, napi_value object, size_t property_count,
                                                   const napi_property_descriptor *properties);
LUX_NAPI_EXTERN napi_status napi_create_external(napi_env env, void *data, napi_finalize finalize_cb,
                                                 void *finalize_hint, napi_value *result);
LUX_NAPI_EXTERN napi_status napi_get_value_external(napi_env env, napi_value value, void **result);
LUX_NAPI_EXTERN napi_status napi_get_typedarray_info(napi_env env, napi_value typedarray, napi_typedarray_type *type,
                                                     size_t *length, void **data, napi_value *arraybuffer,
                                                     size_t *byte_offset);
LUX_NAPI_EXTERN napi_status napi_create_arraybuffer(napi_env env, size_t byte_length, void **data, napi_value *result);
LUX_NAPI_EXTERN napi_status napi_create_typedarray(napi_env env, napi_typedarray_type type, size_t length,
                                                   napi_value arraybuffer, size_t byte_offset, napi_value *result);
LUX_NAPI_EXTERN napi_status napi_create_promise(napi_env env, napi_deferred *deferred, napi_value *promise);
LUX_NAPI_EXTERN napi_status napi_resolve_deferred(napi_env env, napi_deferred deferred, napi_value resolution);
LUX_NAPI_EXTERN napi_status napi_reject_deferred(napi_env env, napi_deferred deferred, napi_value rejection);
LUX_NAPI_EXTERN napi_status napi_create_error(napi_env env, napi_value code, napi_value msg, napi_value *result);
LUX_NAPI_EXTERN napi_status napi_create_async_work(napi_env env, napi_value async_resource,
                                                   napi_value async_resource_name, napi_async_execute_callback execute,
                                                   napi_async_complete_callback complete, void *data,
                                                   napi_async_work *result);
LUX_NAPI_EXTERN napi_status napi_create_reference(napi_env env, napi_value value, uint32_t initial_refcount, napi_ref *result);
LUX_NAPI_EXTERN napi_status napi_delete_reference(napi_env env, napi_ref ref);
LUX_NAPI_EXTERN napi_status napi_get_reference_value(napi_env env, napi_ref ref, napi_value *result);
LUX_NAPI_EXTERN napi_status napi_queue_async_work(napi_env env, napi_async_work work);
LUX_NAPI_EXTERN napi_status napi_delete_async_work(napi_env env, napi_async_work work);

/* module registration by well-known symbol (what NAPI_MODULE_INIT expands to in the real header) */
#define LUX_NAPI_INIT_SYMBOL napi_register_module_v1
#endif
