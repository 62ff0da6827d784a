/* Minimal declarations of the N-API surface napi/sail_napi.cc uses -- ONLY for a `g++ -fsyntax-only` check on
 * images without Node.js (tests/test_cpu_abi.py).  Not a Node header; signatures follow the public N-API docs. */
#ifndef SAIL_TEST_NODE_API_STUB_H
#define SAIL_TEST_NODE_API_STUB_H
#include <stddef.h>
#include <stdint.h>
typedef struct napi_env__* napi_env;
typedef struct napi_value__* napi_value;
typedef struct napi_callback_info__* napi_callback_info;
typedef enum { napi_ok = 0, napi_generic_failure = 9 } napi_status;
typedef enum { napi_default = 0 } napi_property_attributes;
typedef enum { napi_int8_array, napi_uint8_array, napi_uint8_clamped_array, napi_int16_array, napi_uint16_array, napi_int32_array,
               napi_uint32_array, napi_float32_array, napi_float64_array } napi_typedarray_type;
typedef enum { napi_undefined, napi_null, napi_boolean, napi_number, napi_string, napi_symbol, napi_object, napi_function, napi_external, napi_bigint } napi_valuetype;
#define NAPI_AUTO_LENGTH ((size_t)-1)
typedef napi_value (*napi_callback)(napi_env, napi_callback_info);
typedef void (*napi_finalize)(napi_env, void*, void*);
typedef struct { const char* utf8name; napi_value name; napi_callback method; napi_callback getter; napi_callback setter;
                 napi_value value; napi_property_attributes attributes; void* data; } napi_property_descriptor;
extern "C" {
napi_status napi_get_cb_info(napi_env, napi_callback_info, size_t*, napi_value*, napi_value*, void**);
napi_status napi_throw_error(napi_env, const char*, const char*);
napi_status napi_throw_type_error(napi_env, const char*, const char*);
napi_status napi_throw_range_error(napi_env, const char*, const char*);
napi_status napi_get_value_external(napi_env, napi_value, void**);
napi_status napi_get_typedarray_info(napi_env, napi_value, napi_typedarray_type*, size_t*, void**, napi_value*, size_t*);
napi_status napi_get_named_property(napi_env, napi_value, const char*, napi_value*);
napi_status napi_set_named_property(napi_env, napi_value, const char*, napi_value);
napi_status napi_get_value_double(napi_env, napi_value, double*);
napi_status napi_get_value_int32(napi_env, napi_value, int32_t*);
napi_status napi_get_element(napi_env, napi_value, uint32_t, napi_value*);
napi_status napi_create_object(napi_env, napi_value*);
napi_status napi_create_external(napi_env, void*, napi_finalize, void*, napi_value*);
napi_status napi_create_int32(napi_env, int32_t, napi_value*);
napi_status napi_create_arraybuffer(napi_env, size_t, void**, napi_value*);
napi_status napi_create_typedarray(napi_env, napi_typedarray_type, size_t, napi_value, size_t, napi_value*);
napi_status napi_typeof(napi_env, napi_value, napi_valuetype*);
napi_status napi_get_array_length(napi_env, napi_value, uint32_t*);
napi_status napi_get_boolean(napi_env, bool, napi_value*);
napi_status napi_create_string_utf8(napi_env, const char*, size_t, napi_value*);
napi_status napi_define_properties(napi_env, napi_value, size_t, const napi_property_descriptor*);
}
#define NAPI_MODULE(name, init) extern "C" napi_value sail_napi_register(napi_env env, napi_value exports) { return init(env, exports); }
#endif
