/*
 * fake_napi.c — a minimal N-API HOST for bindings/node/luxfhe_b200_napi.c (TEST INFRASTRUCTURE).
 *
 * There is no `node` on the build or GPU boxes, so the addon cannot be loaded by its real host.  This file implements the
 * ~30 napi_* functions the addon imports (bindings/node/node_api_decl.h) over a tiny value model, and — the point of the
 * exercise — runs napi_queue_async_work the way libuv does: `execute` on a worker thread of its own, immediately, so that
 * several un-awaited jobs on ONE engine handle really run concurrently; `complete` later on the "loop" thread
 * (fakehost_drain).  tests/test_node_addon.py links it with the addon into one shared object and drives it through the
 * fakehost_* helpers with ctypes: createEngine / loadKey / two concurrent evalBatch jobs / pbsBatch on the GPU.
 */
#define _GNU_SOURCE
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../../bindings/node/node_api_decl.h"

typedef enum { V_UNDEFINED, V_NULL, V_NUMBER, V_STRING, V_OBJECT, V_EXTERNAL, V_ARRAYBUFFER, V_TYPEDARRAY, V_PROMISE, V_ERROR, V_FUNCTION } vkind;

typedef struct prop {
    char name[48];
    struct napi_value__ *value;
    struct prop *next;
} prop;

struct napi_value__ {
    vkind kind;
    double num;
    char *str;
    void *ptr;                 /* external data / buffer data */
    size_t len;                /* typed array: elements; array buffer: bytes */
    napi_typedarray_type ta_type;
    struct napi_value__ *buffer;
    prop *props;               /* object */
    napi_callback fn;          /* function */
    int state;                 /* promise: 0 pending, 1 resolved, 2 rejected */
    struct napi_value__ *settled;
    napi_finalize finalize;
};

struct napi_ref__ {
    napi_value v;
};
struct napi_deferred__ {
    napi_value promise;
};
struct napi_async_work__ {
    napi_async_execute_callback execute;
    napi_async_complete_callback complete;
    void *data;
    pthread_t thread;
    napi_env env;
    int started;
    struct napi_async_work__ *next;
};
struct napi_callback_info__ {
    size_t argc;
    napi_value *argv;
};
struct napi_env__ {
    char exception[700];
    int has_exception;
    struct napi_async_work__ *queued;
    pthread_mutex_t mtx;
    napi_value exports;
    int max_concurrent, running;
};

static napi_value mk(vkind k) {
    napi_value v = (napi_value)calloc(1, sizeof(struct napi_value__));
    v->kind = k;
    return v;
}

/* ---- N-API surface ------------------------------------------------------------------------------------------ */
napi_status napi_get_cb_info(napi_env env, napi_callback_info info, size_t *argc, napi_value *argv, napi_value *this_arg, void **data) {
    (void)env;
    const size_t cap = *argc;
    for (size_t i = 0; i < cap; i++) argv[i] = i < info->argc ? info->argv[i] : mk(V_UNDEFINED);
    *argc = info->argc;
    if (this_arg) *this_arg = NULL;
    if (data) *data = NULL;
    return napi_ok;
}
napi_status napi_throw_error(napi_env env, const char *code, const char *msg) {
    (void)code;
    snprintf(env->exception, sizeof(env->exception), "%s", msg);
    env->has_exception = 1;
    return napi_ok;
}
napi_status napi_typeof(napi_env env, napi_value v, napi_valuetype *out) {
    (void)env;
    switch (v ? v->kind : V_UNDEFINED) {
        case V_UNDEFINED: *out = napi_undefined; break;
        case V_NULL: *out = napi_null; break;
        case V_NUMBER: *out = napi_number; break;
        case V_STRING: *out = napi_string; break;
        case V_EXTERNAL: *out = napi_external; break;
        case V_FUNCTION: *out = napi_function; break;
        default: *out = napi_object; break;
    }
    return napi_ok;
}
napi_status napi_get_undefined(napi_env env, napi_value *out) {
    (void)env;
    *out = mk(V_UNDEFINED);
    return napi_ok;
}
napi_status napi_get_value_uint32(napi_env env, napi_value v, uint32_t *out) {
    (void)env;
    if (!v || v->kind != V_NUMBER) return (napi_status)1;
    *out = (uint32_t)(long long)v->num;
    return napi_ok;
}
napi_status napi_get_value_int32(napi_env env, napi_value v, int32_t *out) {
    (void)env;
    if (!v || v->kind != V_NUMBER) return (napi_status)1;
    *out = (int32_t)v->num;
    return napi_ok;
}
napi_status napi_create_uint32(napi_env env, uint32_t x, napi_value *out) {
    (void)env;
    *out = mk(V_NUMBER);
    (*out)->num = x;
    return napi_ok;
}
napi_status napi_create_double(napi_env env, double x, napi_value *out) {
    (void)env;
    *out = mk(V_NUMBER);
    (*out)->num = x;
    return napi_ok;
}
napi_status napi_create_string_utf8(napi_env env, const char *s, size_t n, napi_value *out) {
    (void)env;
    *out = mk(V_STRING);
    if (n == NAPI_AUTO_LENGTH) n = strlen(s);
    (*out)->str = (char *)malloc(n + 1);
    memcpy((*out)->str, s, n);
    (*out)->str[n] = 0;
    return napi_ok;
}
static prop *find_prop(napi_value o, const char *name) {
    for (prop *p = o ? o->props : NULL; p; p = p->next)
        if (!strcmp(p->name, name)) return p;
    return NULL;
}
static void set_prop(napi_value o, const char *name, napi_value v) {
    prop *p = find_prop(o, name);
    if (!p) {
        p = (prop *)calloc(1, sizeof(prop));
        snprintf(p->name, sizeof(p->name), "%s", name);
        p->next = o->props;
        o->props = p;
    }
    p->value = v;
}
napi_status napi_get_named_property(napi_env env, napi_value o, const char *name, napi_value *out) {
    (void)env;
    prop *p = find_prop(o, name);
    *out = p ? p->value : mk(V_UNDEFINED);
    return napi_ok;
}
napi_status napi_has_named_property(napi_env env, napi_value o, const char *name, bool *out) {
    (void)env;
    *out = find_prop(o, name) != NULL;
    return napi_ok;
}
napi_status napi_define_properties(napi_env env, napi_value o, size_t n, const napi_property_descriptor *d) {
    (void)env;
    for (size_t i = 0; i < n; i++) {
        napi_value f = mk(V_FUNCTION);
        f->fn = d[i].method;
        set_prop(o, d[i].utf8name, f);
    }
    return napi_ok;
}
napi_status napi_create_external(napi_env env, void *data, napi_finalize fin, void *hint, napi_value *out) {
    (void)env;
    (void)hint;
    *out = mk(V_EXTERNAL);
    (*out)->ptr = data;
    (*out)->finalize = fin;
    return napi_ok;
}
napi_status napi_get_value_external(napi_env env, napi_value v, void **out) {
    (void)env;
    if (!v || v->kind != V_EXTERNAL) return (napi_status)1;
    *out = v->ptr;
    return napi_ok;
}
napi_status napi_get_typedarray_info(napi_env env, napi_value v, napi_typedarray_type *type, size_t *length, void **data,
                                     napi_value *arraybuffer, size_t *byte_offset) {
    (void)env;
    if (!v || v->kind != V_TYPEDARRAY) return (napi_status)1;
    if (type) *type = v->ta_type;
    if (length) *length = v->len;
    if (data) *data = v->ptr;
    if (arraybuffer) *arraybuffer = v->buffer;
    if (byte_offset) *byte_offset = 0;
    return napi_ok;
}
napi_status napi_create_arraybuffer(napi_env env, size_t bytes, void **data, napi_value *out) {
    (void)env;
    *out = mk(V_ARRAYBUFFER);
    (*out)->ptr = calloc(1, bytes ? bytes : 1);
    (*out)->len = bytes;
    if (!(*out)->ptr) return (napi_status)1;
    *data = (*out)->ptr;
    return napi_ok;
}
napi_status napi_create_typedarray(napi_env env, napi_typedarray_type type, size_t length, napi_value ab, size_t off, napi_value *out) {
    (void)env;
    *out = mk(V_TYPEDARRAY);
    (*out)->ta_type = type;
    (*out)->len = length;
    (*out)->ptr = (char *)ab->ptr + off;
    (*out)->buffer = ab;
    return napi_ok;
}
napi_status napi_create_promise(napi_env env, napi_deferred *d, napi_value *promise) {
    (void)env;
    *promise = mk(V_PROMISE);
    *d = (napi_deferred)calloc(1, sizeof(struct napi_deferred__));
    (*d)->promise = *promise;
    return napi_ok;
}
napi_status napi_resolve_deferred(napi_env env, napi_deferred d, napi_value v) {
    (void)env;
    d->promise->state = 1;
    d->promise->settled = v;
    free(d);
    return napi_ok;
}
napi_status napi_reject_deferred(napi_env env, napi_deferred d, napi_value v) {
    (void)env;
    d->promise->state = 2;
    d->promise->settled = v;
    free(d);
    return napi_ok;
}
napi_status napi_create_error(napi_env env, napi_value code, napi_value msg, napi_value *out) {
    (void)env;
    (void)code;
    *out = mk(V_ERROR);
    (*out)->str = strdup(msg && msg->str ? msg->str : "");
    return napi_ok;
}
napi_status napi_create_reference(napi_env env, napi_value v, uint32_t rc, napi_ref *out) {
    (void)env;
    (void)rc;
    *out = (napi_ref)calloc(1, sizeof(struct napi_ref__));
    (*out)->v = v;
    return napi_ok;
}
napi_status napi_delete_reference(napi_env env, napi_ref r) {
    (void)env;
    free(r);
    return napi_ok;
}
napi_status napi_get_reference_value(napi_env env, napi_ref r, napi_value *out) {
    (void)env;
    *out = r->v;
    return napi_ok;
}
napi_status napi_create_async_work(napi_env env, napi_value res, napi_value name, napi_async_execute_callback ex,
                                   napi_async_complete_callback done, void *data, napi_async_work *out) {
    (void)res;
    (void)name;
    *out = (napi_async_work)calloc(1, sizeof(struct napi_async_work__));
    (*out)->execute = ex;
    (*out)->complete = done;
    (*out)->data = data;
    (*out)->env = env;
    return napi_ok;
}
static void *work_main(void *p) {
    napi_async_work w = (napi_async_work)p;
    pthread_mutex_lock(&w->env->mtx);
    if (++w->env->running > w->env->max_concurrent) w->env->max_concurrent = w->env->running;
    pthread_mutex_unlock(&w->env->mtx);
    w->execute(w->env, w->data);
    pthread_mutex_lock(&w->env->mtx);
    w->env->running--;
    pthread_mutex_unlock(&w->env->mtx);
    return NULL;
}
napi_status napi_queue_async_work(napi_env env, napi_async_work w) {
    pthread_mutex_lock(&env->mtx);
    w->next = env->queued;
    env->queued = w;
    pthread_mutex_unlock(&env->mtx);
    w->started = pthread_create(&w->thread, NULL, work_main, w) == 0;   /* libuv: a pool thread picks it up at once */
    return napi_ok;
}
napi_status napi_delete_async_work(napi_env env, napi_async_work w) {
    (void)env;
    free(w);
    return napi_ok;
}

/* ---- helpers for the driver (ctypes) ------------------------------------------------------------------------ */
napi_value napi_register_module_v1(napi_env env, napi_value exports);

napi_env fakehost_new(void) {
    napi_env env = (napi_env)calloc(1, sizeof(struct napi_env__));
    pthread_mutex_init(&env->mtx, NULL);
    env->exports = mk(V_OBJECT);
    napi_register_module_v1(env, env->exports);
    return env;
}
napi_value fakehost_number(double x) {
    napi_value v = mk(V_NUMBER);
    v->num = x;
    return v;
}
napi_value fakehost_null(void) { return mk(V_NULL); }
napi_value fakehost_object(void) { return mk(V_OBJECT); }
void fakehost_set(napi_value o, const char *name, napi_value v) { set_prop(o, name, v); }
/* a typed array VIEW of caller memory (zero copy, like a JS BigUint64Array over an ArrayBuffer) */
napi_value fakehost_typedarray(int type, void *data, size_t elements) {
    napi_value v = mk(V_TYPEDARRAY);
    v->ta_type = (napi_typedarray_type)type;
    v->ptr = data;
    v->len = elements;
    return v;
}
/* call exports.<name>(argv...); returns NULL when the addon threw (fakehost_exception has the message) */
napi_value fakehost_call(napi_env env, const char *name, size_t argc, napi_value *argv) {
    prop *p = find_prop(env->exports, name);
    if (!p || p->value->kind != V_FUNCTION) {
        napi_throw_error(env, NULL, "no such export");
        return NULL;
    }
    struct napi_callback_info__ info = {argc, argv};
    env->has_exception = 0;
    napi_value r = p->value->fn(env, &info);
    return env->has_exception ? NULL : r;
}
const char *fakehost_exception(napi_env env) { return env->has_exception ? env->exception : ""; }
/* the event loop: wait for every queued job's worker thread, then run its completion callback here */
int fakehost_drain(napi_env env) {
    int n = 0;
    for (;;) {
        pthread_mutex_lock(&env->mtx);
        napi_async_work w = env->queued;
        if (w) env->queued = w->next;
        pthread_mutex_unlock(&env->mtx);
        if (!w) break;
        if (w->started) pthread_join(w->thread, NULL);
        w->complete(env, w->started ? napi_ok : (napi_status)1, w->data);
        n++;
    }
    return n;
}
int fakehost_max_concurrent(napi_env env) { return env->max_concurrent; }
int fakehost_promise_state(napi_value p) { return p && p->kind == V_PROMISE ? p->state : -1; }
napi_value fakehost_promise_value(napi_value p) { return p ? p->settled : NULL; }
const char *fakehost_string(napi_value v) { return v && v->str ? v->str : ""; }
double fakehost_number_value(napi_value v) { return v ? v->num : 0; }
void *fakehost_data(napi_value v, size_t *elements) {
    if (!v || v->kind != V_TYPEDARRAY) return NULL;
    if (elements) *elements = v->len;
    return v->ptr;
}
/* garbage collection of an engine handle: runs the finalizer the addon registered */
void fakehost_finalize(napi_env env, napi_value external) {
    if (external && external->kind == V_EXTERNAL && external->finalize) {
        external->finalize(env, external->ptr, NULL);
        external->finalize = NULL;
        external->ptr = NULL;
    }
}
