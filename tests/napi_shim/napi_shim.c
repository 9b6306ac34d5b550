/* napi_shim.c — TEST INFRASTRUCTURE: a stand-in for exactly the Node-API calls bindings/node/igpu_addon.c uses, so that the
 * addon's C code can be executed (marshalling, validation, results) on a machine without Node.js.  A toy value model
 * (numbers, strings, plain objects, ArrayBuffers, typed arrays, native functions) plus a small driver API that pytest calls
 * through ctypes (shim_*).  It says nothing about V8 itself: under a real Node.js the addon remains unverified.
 *
 * Built together with the addon into one shared object by tests/test_napi_addon.py:
 *   gcc -shared -fPIC -Iinclude -Ibindings/node tests/napi_shim/napi_shim.c bindings/node/igpu_addon.c -L<csrc> -ligpu_rollup
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "node_api_min.h"

typedef enum { V_UNDEFINED, V_NUMBER, V_STRING, V_OBJECT, V_ARRAYBUFFER, V_TYPEDARRAY, V_FUNCTION } vkind;

struct prop {
  char* name;
  struct napi_value__* value;
  struct prop* next;
};

struct napi_value__ {
  vkind kind;
  double num;
  char* str;
  struct prop* props;           /* V_OBJECT (insertion order, newest first) */
  void* data;                   /* V_ARRAYBUFFER */
  size_t bytes;
  int owned;
  napi_typedarray_type tt;      /* V_TYPEDARRAY */
  size_t length, byte_off;
  struct napi_value__* ab;
  napi_callback fn;             /* V_FUNCTION */
  struct napi_value__* gc_next; /* every value of an env, freed with it */
};

struct napi_env__ {
  struct napi_value__* all;
  char exception[512];
  char exception_code[64];
  int pending;
  struct napi_value__* exports;
};

struct napi_callback_info__ {
  size_t argc;
  napi_value* argv;
};

static const size_t ELEM[] = {1, 1, 1, 2, 2, 4, 4, 4, 8, 8, 8};

static napi_value mk(napi_env env, vkind k) {
  napi_value v = (napi_value)calloc(1, sizeof *v);
  v->kind = k;
  v->gc_next = env->all;
  env->all = v;
  return v;
}

/* ---- the Node-API subset -------------------------------------------------------------------------------------- */
napi_status napi_get_cb_info(napi_env env, napi_callback_info info, size_t* argc, napi_value* argv, napi_value* this_arg, void** data) {
  (void)this_arg, (void)data;
  const size_t want = *argc;
  for (size_t i = 0; i < want; ++i) argv[i] = i < info->argc ? info->argv[i] : mk(env, V_UNDEFINED);
  *argc = info->argc;
  return napi_ok;
}

napi_status napi_get_named_property(napi_env env, napi_value object, const char* name, napi_value* result) {
  if (!object || object->kind != V_OBJECT) return (napi_status)1; /* napi_invalid_arg / object_expected */
  for (struct prop* p = object->props; p; p = p->next)
    if (!strcmp(p->name, name)) {
      *result = p->value;
      return napi_ok;
    }
  *result = mk(env, V_UNDEFINED); /* like JS: a missing property reads as undefined */
  return napi_ok;
}

napi_status napi_set_named_property(napi_env env, napi_value object, const char* name, napi_value value) {
  (void)env;
  if (!object || object->kind != V_OBJECT) return (napi_status)1;
  for (struct prop* p = object->props; p; p = p->next)
    if (!strcmp(p->name, name)) {
      p->value = value;
      return napi_ok;
    }
  struct prop* p = (struct prop*)calloc(1, sizeof *p);
  p->name = strdup(name);
  p->value = value;
  p->next = object->props;
  object->props = p;
  return napi_ok;
}

napi_status napi_get_value_uint32(napi_env env, napi_value value, uint32_t* result) {
  (void)env;
  if (!value || value->kind != V_NUMBER) return (napi_status)6; /* napi_number_expected */
  *result = (uint32_t)(long long)value->num;
  return napi_ok;
}

napi_status napi_get_value_double(napi_env env, napi_value value, double* result) {
  (void)env;
  if (!value || value->kind != V_NUMBER) return (napi_status)6;
  *result = value->num;
  return napi_ok;
}

napi_status napi_get_typedarray_info(napi_env env, napi_value ta, napi_typedarray_type* type, size_t* length, void** data,
                                     napi_value* arraybuffer, size_t* byte_offset) {
  (void)env;
  if (!ta || ta->kind != V_TYPEDARRAY) return (napi_status)1;
  if (type) *type = ta->tt;
  if (length) *length = ta->length;
  if (data) *data = (char*)ta->ab->data + ta->byte_off;
  if (arraybuffer) *arraybuffer = ta->ab;
  if (byte_offset) *byte_offset = ta->byte_off;
  return napi_ok;
}

napi_status napi_get_arraybuffer_info(napi_env env, napi_value ab, void** data, size_t* byte_length) {
  (void)env;
  if (!ab || ab->kind != V_ARRAYBUFFER) return (napi_status)1;
  if (data) *data = ab->data;
  if (byte_length) *byte_length = ab->bytes;
  return napi_ok;
}

napi_status napi_create_object(napi_env env, napi_value* result) {
  *result = mk(env, V_OBJECT);
  return napi_ok;
}

napi_status napi_create_arraybuffer(napi_env env, size_t byte_length, void** data, napi_value* result) {
  napi_value v = mk(env, V_ARRAYBUFFER);
  v->data = calloc(byte_length ? byte_length : 1, 1); /* V8 hands out zero-filled buffers */
  v->bytes = byte_length;
  v->owned = 1;
  if (data) *data = v->data;
  *result = v;
  return napi_ok;
}

napi_status napi_create_typedarray(napi_env env, napi_typedarray_type type, size_t length, napi_value ab, size_t byte_offset,
                                   napi_value* result) {
  if (!ab || ab->kind != V_ARRAYBUFFER || (unsigned)type > napi_biguint64_array) return (napi_status)1;
  if (byte_offset % ELEM[type] || byte_offset + length * ELEM[type] > ab->bytes) {
    snprintf(env->exception, sizeof env->exception, "RangeError: typed array of %zu x %zu bytes at %zu does not fit its buffer of %zu",
             length, ELEM[type], byte_offset, ab->bytes);
    env->pending = 1;
    return (napi_status)10; /* napi_pending_exception */
  }
  napi_value v = mk(env, V_TYPEDARRAY);
  v->tt = type, v->length = length, v->byte_off = byte_offset, v->ab = ab;
  *result = v;
  return napi_ok;
}

napi_status napi_create_double(napi_env env, double value, napi_value* result) {
  napi_value v = mk(env, V_NUMBER);
  v->num = value;
  *result = v;
  return napi_ok;
}

napi_status napi_create_string_utf8(napi_env env, const char* str, size_t length, napi_value* result) {
  napi_value v = mk(env, V_STRING);
  const size_t n = length == (size_t)-1 ? strlen(str) : length; /* NAPI_AUTO_LENGTH */
  v->str = (char*)malloc(n + 1);
  memcpy(v->str, str, n);
  v->str[n] = 0;
  *result = v;
  return napi_ok;
}

napi_status napi_throw_error(napi_env env, const char* code, const char* msg) {
  if (!env->pending) { /* the first exception wins, like a JS throw that is never caught in between */
    snprintf(env->exception, sizeof env->exception, "%s", msg ? msg : "");
    snprintf(env->exception_code, sizeof env->exception_code, "%s", code ? code : "");
  }
  env->pending = 1;
  return napi_ok;
}

napi_status napi_define_properties(napi_env env, napi_value object, size_t n, const napi_property_descriptor* props) {
  for (size_t i = 0; i < n; ++i) {
    napi_value f = mk(env, V_FUNCTION);
    f->fn = props[i].method;
    napi_set_named_property(env, object, props[i].utf8name, f);
  }
  return napi_ok;
}

/* ---- driver API for the tests (ctypes) ---------------------------------------------------------------------------- */
napi_value napi_register_module_v1(napi_env env, napi_value exports); /* the addon's NAPI_MODULE_INIT */

napi_env shim_env_new(void) { return (napi_env)calloc(1, sizeof(struct napi_env__)); }

void shim_env_free(napi_env env) {
  for (napi_value v = env->all; v;) {
    napi_value nx = v->gc_next;
    for (struct prop* p = v->props; p;) {
      struct prop* pn = p->next;
      free(p->name);
      free(p);
      p = pn;
    }
    if (v->kind == V_ARRAYBUFFER && v->owned) free(v->data);
    free(v->str);
    free(v);
    v = nx;
  }
  free(env);
}

/* `require()` of the addon: NULL when module initialisation threw (text via shim_exception) */
napi_value shim_load(napi_env env) {
  napi_value exports = mk(env, V_OBJECT);
  napi_value r = napi_register_module_v1(env, exports);
  if (env->pending || !r) return NULL;
  env->exports = r;
  return r;
}

const char* shim_exception(napi_env env) { return env->pending ? env->exception : NULL; }
const char* shim_exception_code(napi_env env) { return env->exception_code; }
void shim_clear_exception(napi_env env) { env->pending = 0, env->exception[0] = 0, env->exception_code[0] = 0; }

napi_value shim_object(napi_env env) { return mk(env, V_OBJECT); }
napi_value shim_number(napi_env env, double x) {
  napi_value v;
  napi_create_double(env, x, &v);
  return v;
}
/* a typed array over memory the caller keeps alive (numpy): what `new Uint8Array(buffer)` is to the addon */
napi_value shim_typedarray(napi_env env, int type, void* data, size_t length) {
  napi_value ab = mk(env, V_ARRAYBUFFER);
  ab->data = data, ab->bytes = length * ELEM[type], ab->owned = 0;
  napi_value v = mk(env, V_TYPEDARRAY);
  v->tt = (napi_typedarray_type)type, v->length = length, v->ab = ab;
  return v;
}
int shim_set(napi_env env, napi_value obj, const char* name, napi_value v) { return napi_set_named_property(env, obj, name, v); }
napi_value shim_get(napi_env env, napi_value obj, const char* name) {
  napi_value v = NULL;
  if (napi_get_named_property(env, obj, name, &v) != napi_ok || v->kind == V_UNDEFINED) return NULL;
  return v;
}
int shim_kind(napi_value v) { return v ? (int)v->kind : -1; }
const char* shim_string(napi_value v) { return v && v->kind == V_STRING ? v->str : NULL; }
int shim_typedarray_info(napi_value v, int* type, size_t* length, void** data) {
  napi_typedarray_type t;
  if (napi_get_typedarray_info(NULL, v, &t, length, data, NULL, NULL) != napi_ok) return 1;
  *type = (int)t;
  return 0;
}
/* the i-th own property name of an object, oldest first; NULL past the end */
const char* shim_key(napi_value obj, size_t i) {
  size_t n = 0;
  for (struct prop* p = obj->props; p; p = p->next) ++n;
  if (i >= n) return NULL;
  struct prop* p = obj->props;
  for (size_t k = 0; k < n - 1 - i; ++k) p = p->next;
  return p->name;
}
/* exports[name](...argv): NULL when the call threw */
napi_value shim_call(napi_env env, const char* name, napi_value* argv, size_t argc) {
  napi_value f = shim_get(env, env->exports, name);
  if (!f || f->kind != V_FUNCTION) {
    napi_throw_error(env, "ERR_SHIM", "no such export");
    return NULL;
  }
  struct napi_callback_info__ info = {argc, argv};
  napi_value r = f->fn(env, &info);
  return env->pending ? NULL : r;
}
