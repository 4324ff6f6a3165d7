/*
 * luxfhe_b200_napi.c — Node N-API addon over the C ABI of libluxtfhe_b200.so (include/lux_tfhe.h, lux_radix.h).
 *
 * This is the host seam BASELINE.json's north_star names ("TypeScript host code drives batches through a Node
 * N-API C-ABI shim into CUDA").  It replaces, for the evaluation path, the object the reference's Go-wasm `main`
 * registers as globalThis.luxfhe (reference: packages/wasm/src/node.ts:48,57-96; interface
 * packages/wasm/src/index.ts:14-23) and the HTTP twin POST /evaluate (js/v2-sdk/src/luxd/client.ts:128-147).
 * What changes against that interface: keys are uploaded once and referred to by the engine handle
 * instead of being passed by value on every call; every op has a batch dimension; ciphertexts are BigUint64Array
 * views of the engine-native little-endian u64 layout (zero copy in, one ArrayBuffer out); calls that launch GPU work
 * return Promises and run on the libuv pool (napi_async_work), so Node's loop is never blocked; errors are JS
 * exceptions / rejections carrying lux_last_error().
 *
 * Not loadable in the build image (no node): tests/test_node_addon.py compiles it against node_api_decl.h, checks the
 * registration symbol and that every lux_* symbol it needs is exported by the library, and RUNS it on the GPU box inside
 * tests/napi_host/fake_napi.c, a minimal N-API host whose async work runs on real threads (two concurrent jobs per handle).
 */
#if defined(__has_include)
#if __has_include(<node_api.h>)
#include <node_api.h>
#define LUX_HAVE_NODE_API 1
#endif
#endif
#ifndef LUX_HAVE_NODE_API
#include "node_api_decl.h"
#endif

#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/lux_radix.h"
#include "../../include/lux_tfhe.h"

#define THROW(env, ...)                          \
    do {                                         \
        char _m[600];                            \
        snprintf(_m, sizeof(_m), __VA_ARGS__);   \
        napi_throw_error((env), NULL, _m);       \
        return NULL;                             \
    } while (0)

static napi_value undefined(napi_env env) {
    napi_value u;
    napi_get_undefined(env, &u);
    return u;
}

/* what a JS handle points at: one engine (one GPU) or a multi-GPU handle (lux_engine_create_multi); `e` is always usable for
 * parameter queries (GPU 0's engine of a multi handle).  The library serialises calls per engine, so several un-awaited
 * Promises on one handle are safe; every queued job also holds a reference to the handle, so the finalizer cannot run
 * while work is in flight. */
typedef struct {
    lux_engine *e;
    lux_multi *m;
} handle_t;

static void engine_finalize(napi_env env, void *data, void *hint) {
    (void)env;
    (void)hint;
    handle_t *h = (handle_t *)data;
    if (h->m)
        lux_multi_destroy(h->m);
    else
        lux_engine_destroy(h->e);
    free(h);
}

static handle_t *handle_arg(napi_env env, napi_value v) {
    void *p = NULL;
    if (napi_get_value_external(env, v, &p) != napi_ok || !p) {
        napi_throw_error(env, NULL, "LuxFHE not initialized: pass the handle returned by createEngine()");
        return NULL;
    }
    return (handle_t *)p;
}
static lux_engine *engine_arg(napi_env env, napi_value v) {
    handle_t *h = handle_arg(env, v);
    return h ? h->e : NULL;
}

/* BigUint64Array -> (pointer, words); zero copy */
static const uint64_t *u64_arg(napi_env env, napi_value v, size_t *words) {
    napi_typedarray_type t;
    void *data = NULL;
    size_t len = 0;
    if (napi_get_typedarray_info(env, v, &t, &len, &data, NULL, NULL) != napi_ok || t != napi_biguint64_array) return NULL;
    *words = len;
    return (const uint64_t *)data;
}

static int is_nullish(napi_env env, napi_value v) {
    napi_valuetype t;
    return napi_typeof(env, v, &t) != napi_ok || t == napi_undefined || t == napi_null;
}

static uint32_t u32_prop(napi_env env, napi_value obj, const char *name, uint32_t dflt) {
    bool has = false;
    napi_value v;
    uint32_t out = dflt;
    if (napi_has_named_property(env, obj, name, &has) == napi_ok && has && napi_get_named_property(env, obj, name, &v) == napi_ok)
        napi_get_value_uint32(env, v, &out);
    return out;
}

/* ---- synchronous entry points ------------------------------------------------------------------------------ */
static napi_value Version(napi_env env, napi_callback_info info) { /* LuxFHE.version(), index.ts:15 */
    (void)info;
    napi_value s;
    napi_create_string_utf8(env, lux_version(), NAPI_AUTO_LENGTH, &s);
    return s;
}

/* createEngine({n, N, k, ell, logB, ellKs, logBKs, order, arith}, device = 0) -> handle */
static napi_value CreateEngine(napi_env env, napi_callback_info info) {
    size_t argc = 2;
    napi_value argv[2];
    napi_get_cb_info(env, info, &argc, argv, NULL, NULL);
    if (argc < 1) THROW(env, "createEngine(params, device?)");
    lux_params p;
    p.n = u32_prop(env, argv[0], "n", 0);
    p.N = u32_prop(env, argv[0], "N", 0);
    p.k = u32_prop(env, argv[0], "k", 1);
    p.ell = u32_prop(env, argv[0], "ell", 0);
    p.logB = u32_prop(env, argv[0], "logB", 0);
    p.ell_ks = u32_prop(env, argv[0], "ellKs", 0);
    p.logB_ks = u32_prop(env, argv[0], "logBKs", 0);
    p.order = u32_prop(env, argv[0], "order", LUX_ORDER_LWEKS_PBS);
    p.arith = u32_prop(env, argv[0], "arith", LUX_ARITH_GOLDILOCKS_NTT);
    int32_t device = 0;
    if (argc > 1 && !is_nullish(env, argv[1])) napi_get_value_int32(env, argv[1], &device);
    const uint32_t gpus = u32_prop(env, argv[0], "gpus", 0);   /* gpus >= 1: one handle over that many GPUs (0xFFFFFFFF: all) */
    handle_t *hd = (handle_t *)calloc(1, sizeof(handle_t));
    if (!hd) THROW(env, "out of memory");
    if (gpus) {
        const int rc = lux_engine_create_multi(&p, gpus == 0xFFFFFFFFu ? 0 : (int)gpus, &hd->m);
        if (rc != LUX_OK) {
            free(hd);
            THROW(env, "createEngine failed (%d): %s", rc, lux_multi_last_error(NULL));
        }
        hd->e = lux_multi_engine(hd->m, 0);
    } else {
        const int rc = lux_engine_create(&p, device, &hd->e);
        if (rc != LUX_OK) {
            free(hd);
            THROW(env, "createEngine failed (%d): %s", rc, lux_last_error(NULL));
        }
    }
    napi_value h;
    if (napi_create_external(env, hd, engine_finalize, NULL, &h) != napi_ok) {
        engine_finalize(env, hd, NULL);
        THROW(env, "napi_create_external failed");
    }
    return h;
}

static napi_value CtWords(napi_env env, napi_callback_info info) {
    size_t argc = 1;
    napi_value argv[1], out;
    napi_get_cb_info(env, info, &argc, argv, NULL, NULL);
    lux_engine *e = argc ? engine_arg(env, argv[0]) : NULL;
    if (!e) return NULL;
    napi_create_uint32(env, (uint32_t)lux_ct_words(e), &out);
    return out;
}

static napi_value RadixBlocks(napi_env env, napi_callback_info info) {
    size_t argc = 1;
    napi_value argv[1], out;
    uint32_t bw = 0;
    napi_get_cb_info(env, info, &argc, argv, NULL, NULL);
    if (argc) napi_get_value_uint32(env, argv[0], &bw);
    napi_create_uint32(env, (uint32_t)lux_radix_blocks(bw), &out);
    return out;
}

/* loadKey(handle, which: 0 = bootstrap key, 1 = GLWE key switch key, 2 = LWE key switch key, words, domain = 0)
 * replaces the bootstrapKey argument of every LuxFHE.add/sub/eq/lt call (index.ts:19-22): uploaded once, converted
 * to the NTT domain on the device, resident in HBM afterwards */
static napi_value LoadKey(napi_env env, napi_callback_info info) {
    size_t argc = 4, words = 0;
    napi_value argv[4];
    napi_get_cb_info(env, info, &argc, argv, NULL, NULL);
    if (argc < 3) THROW(env, "loadKey(handle, which, words, domain?)");
    handle_t *hd = handle_arg(env, argv[0]);
    if (!hd) return NULL;
    lux_engine *e = hd->e;
    uint32_t which = 0, domain = LUX_DOMAIN_STANDARD;
    napi_get_value_uint32(env, argv[1], &which);
    const uint64_t *w = u64_arg(env, argv[2], &words);
    if (!w) THROW(env, "loadKey: key must be a BigUint64Array");
    if (argc > 3 && !is_nullish(env, argv[3])) napi_get_value_uint32(env, argv[3], &domain);
    int rc;
    if (which > 2) THROW(env, "loadKey: which must be 0, 1 or 2");
    if (hd->m) {   /* converted on GPU 0, then copied device to device to every other GPU of the handle */
        rc = which == 0 ? lux_multi_load_bsk(hd->m, w, words, (int)domain)
             : which == 1 ? lux_multi_load_glwe_ksk(hd->m, w, words, (int)domain)
                          : lux_multi_load_lwe_ksk(hd->m, w, words);
        if (rc == LUX_OK) rc = lux_multi_replicate_keys(hd->m);
        if (rc != LUX_OK) THROW(env, "loadKey failed (%d): %s", rc, lux_multi_last_error(hd->m));
        return undefined(env);
    }
    if (which == 0)
        rc = lux_load_bsk(e, w, words, (int)domain);
    else if (which == 1)
        rc = lux_load_glwe_ksk(e, w, words, (int)domain);
    else
        rc = lux_load_lwe_ksk(e, w, words);
    if (rc != LUX_OK) THROW(env, "loadKey failed (%d): %s", rc, lux_last_error(e));
    return undefined(env);
}

/* trivialEncrypt(handle, bitWidth, values: BigUint64Array) -> BigUint64Array [batch][blocks][ctWords] (no GPU work) */
static napi_value TrivialEncrypt(napi_env env, napi_callback_info info) {
    size_t argc = 3, batch = 0;
    napi_value argv[3];
    napi_get_cb_info(env, info, &argc, argv, NULL, NULL);
    if (argc < 3) THROW(env, "trivialEncrypt(handle, bitWidth, values)");
    lux_engine *e = engine_arg(env, argv[0]);
    if (!e) return NULL;
    uint32_t bw = 0;
    napi_get_value_uint32(env, argv[1], &bw);
    const uint64_t *vals = u64_arg(env, argv[2], &batch);
    const size_t blocks = lux_radix_blocks(bw);
    if (!vals || !blocks) THROW(env, "trivialEncrypt: bad bit width %u or values not a BigUint64Array", bw);
    const size_t words = batch * blocks * lux_ct_words(e);
    void *data = NULL;
    napi_value ab, out;
    if (napi_create_arraybuffer(env, words * 8, &data, &ab) != napi_ok) THROW(env, "out of memory");
    const int rc = lux_trivial_encrypt(e, bw, vals, batch, (uint64_t *)data);
    if (rc != LUX_OK) THROW(env, "trivialEncrypt failed (%d): %s", rc, lux_last_error(e));
    napi_create_typedarray(env, napi_biguint64_array, words, ab, 0, &out);
    return out;
}

/* ---- asynchronous GPU jobs (libuv pool) -------------------------------------------------------------------- */
typedef enum { JOB_PBS, JOB_EVAL, JOB_EVAL_SCALAR, JOB_SELECT, JOB_ERC20, JOB_GATE } job_kind;

typedef struct {
    job_kind kind;
    lux_engine *e;
    lux_multi *m;               /* non-NULL: the handle spans several GPUs, the library slices the batch */
    const uint64_t *in[3];      /* operand views (kept alive by refs) */
    const uint32_t *idx;        /* PBS: lut index per ciphertext, or NULL */
    uint64_t *out;              /* result ArrayBuffer data */
    size_t out_words, batch;
    uint32_t op, bit_width, n_luts;
    int rc;
    char err[512];
    napi_ref refs[6];
    int n_refs;
    napi_ref out_ref;
    napi_deferred deferred;
    napi_async_work work;
} job_t;

static void job_execute(napi_env env, void *data) { /* worker thread: no N-API calls here */
    (void)env;
    job_t *j = (job_t *)data;
    if (j->m && (j->kind == JOB_PBS || j->kind == JOB_EVAL || j->kind == JOB_SELECT || j->kind == JOB_ERC20)) {
        switch (j->kind) {
            case JOB_PBS: j->rc = lux_multi_pbs_batch(j->m, j->in[0], j->in[1], j->n_luts, j->idx, j->batch, j->out); break;
            case JOB_EVAL: j->rc = lux_multi_eval(j->m, (lux_op)j->op, j->bit_width, j->in[0], j->in[1], j->batch, j->out); break;
            case JOB_SELECT: j->rc = lux_multi_select(j->m, j->bit_width, j->in[0], j->in[1], j->in[2], j->batch, j->out); break;
            default:
                j->rc = lux_multi_erc20_transfer(j->m, j->in[0], j->in[1], j->in[2], j->batch, j->out, j->out + j->out_words / 2);
                break;
        }
        if (j->rc != LUX_OK) snprintf(j->err, sizeof(j->err), "lux error %d: %s", j->rc, lux_multi_last_error(j->m));
        return;
    }
    switch (j->kind) {
        case JOB_PBS: j->rc = lux_pbs_batch(j->e, j->in[0], j->in[1], j->n_luts, j->idx, j->batch, j->out); break;
        case JOB_EVAL: j->rc = lux_eval(j->e, (lux_op)j->op, j->bit_width, j->in[0], j->in[1], j->batch, j->out); break;
        case JOB_EVAL_SCALAR: j->rc = lux_eval_scalar(j->e, (lux_op)j->op, j->bit_width, j->in[0], j->in[1], j->batch, j->out); break;
        case JOB_SELECT: j->rc = lux_select(j->e, j->bit_width, j->in[0], j->in[1], j->in[2], j->batch, j->out); break;
        case JOB_ERC20:
            j->rc = lux_erc20_transfer(j->e, j->in[0], j->in[1], j->in[2], j->batch, j->out, j->out + j->out_words / 2);
            break;
        case JOB_GATE: j->rc = lux_gate_batch(j->e, (int)j->op, j->in[0], j->in[1], j->batch, j->out); break;
    }
    if (j->rc != LUX_OK) snprintf(j->err, sizeof(j->err), "lux error %d: %s", j->rc, lux_last_error(j->e));
}

static void job_complete(napi_env env, napi_status status, void *data) {
    job_t *j = (job_t *)data;
    napi_value result;
    if (status == napi_ok && j->rc == LUX_OK) {
        napi_get_reference_value(env, j->out_ref, &result);
        napi_resolve_deferred(env, j->deferred, result);
    } else {
        napi_value msg, err;
        napi_create_string_utf8(env, j->rc != LUX_OK ? j->err : "Evaluation failed: async work cancelled", NAPI_AUTO_LENGTH, &msg);
        napi_create_error(env, NULL, msg, &err);
        napi_reject_deferred(env, j->deferred, err);
    }
    for (int i = 0; i < j->n_refs; i++) napi_delete_reference(env, j->refs[i]);
    napi_delete_reference(env, j->out_ref);
    napi_delete_async_work(env, j->work);
    free(j);
}

static void job_hold(napi_env env, job_t *j, napi_value v) {
    if (j->n_refs < 6 && napi_create_reference(env, v, 1, &j->refs[j->n_refs]) == napi_ok) j->n_refs++;
}

/* allocate the result array, queue the job, return its Promise */
static napi_value job_submit(napi_env env, job_t *j, const char *name) {
    void *data = NULL;
    napi_value ab, out, promise, rname;
    if (napi_create_arraybuffer(env, j->out_words * 8, &data, &ab) != napi_ok ||
        napi_create_typedarray(env, napi_biguint64_array, j->out_words, ab, 0, &out) != napi_ok) {
        const size_t words = j->out_words;
        for (int i = 0; i < j->n_refs; i++) napi_delete_reference(env, j->refs[i]);
        free(j);
        THROW(env, "%s: cannot allocate %zu result words", name, words);
    }
    j->out = (uint64_t *)data;
    napi_create_reference(env, out, 1, &j->out_ref);
    napi_create_promise(env, &j->deferred, &promise);
    napi_create_string_utf8(env, name, NAPI_AUTO_LENGTH, &rname);
    napi_create_async_work(env, NULL, rname, job_execute, job_complete, j, &j->work);
    napi_queue_async_work(env, j->work);
    return promise;
}

/* the job keeps the JS handle alive (its finalizer destroys the engine) until the completion callback has run */
static job_t *job_new(napi_env env, job_kind kind, napi_value handle) {
    handle_t *hd = handle_arg(env, handle);
    job_t *j = hd ? (job_t *)calloc(1, sizeof(job_t)) : NULL;
    if (j) {
        j->kind = kind;
        j->e = hd->e;
        j->m = hd->m;
        job_hold(env, j, handle);
    }
    return j;
}
static void job_abandon(napi_env env, job_t *j) {
    for (int i = 0; i < j->n_refs; i++) napi_delete_reference(env, j->refs[i]);
    free(j);
}

/* pbsBatch(handle, lweIn, testVectors, nLuts, lutIdx: Uint32Array | null, batch) -> Promise<BigUint64Array> */
static napi_value PbsBatch(napi_env env, napi_callback_info info) {
    size_t argc = 6, w_in = 0, w_tv = 0;
    napi_value argv[6];
    napi_get_cb_info(env, info, &argc, argv, NULL, NULL);
    if (argc < 6) THROW(env, "pbsBatch(handle, lweIn, testVectors, nLuts, lutIdx, batch)");
    lux_engine *e = engine_arg(env, argv[0]);
    if (!e) return NULL;
    uint32_t n_luts = 0, batch = 0;
    const uint64_t *in = u64_arg(env, argv[1], &w_in), *tv = u64_arg(env, argv[2], &w_tv);
    napi_get_value_uint32(env, argv[3], &n_luts);
    napi_get_value_uint32(env, argv[5], &batch);
    lux_params p;
    lux_engine_params(e, &p);
    const size_t cw = lux_ct_words(e);
    if (!in || !tv || w_in != (size_t)batch * cw || w_tv != (size_t)n_luts * p.N)
        THROW(env, "pbsBatch: expected %zu ciphertext words and %zu test-vector words", (size_t)batch * cw, (size_t)n_luts * p.N);
    const uint32_t *idx = NULL;
    if (!is_nullish(env, argv[4])) {
        napi_typedarray_type t;
        size_t len = 0;
        void *d = NULL;
        if (napi_get_typedarray_info(env, argv[4], &t, &len, &d, NULL, NULL) != napi_ok || t != napi_uint32_array || len != batch)
            THROW(env, "pbsBatch: lutIdx must be a Uint32Array of length batch");
        idx = (const uint32_t *)d;
        for (size_t b = 0; b < len; b++)
            if (idx[b] >= n_luts) THROW(env, "pbsBatch: lutIdx[%zu] out of range", b);
    }
    job_t *j = job_new(env, JOB_PBS, argv[0]);
    if (!j) THROW(env, "out of memory");
    j->in[0] = in;
    j->in[1] = tv;
    j->idx = idx;
    j->n_luts = n_luts;
    j->batch = batch;
    j->out_words = (size_t)batch * cw;
    job_hold(env, j, argv[1]);
    job_hold(env, j, argv[2]);
    if (idx) job_hold(env, j, argv[4]);
    return job_submit(env, j, "luxfhe.pbsBatch");
}

/* evalBatch(handle, op, bitWidth, lhs, rhs | null, batch, scalar = false) -> Promise<BigUint64Array>
 * op = LUX_OP_* (lux_radix.h); replaces LuxFHE.add/sub/eq/lt (index.ts:19-22) and POST /evaluate (client.ts:128-147).
 * With scalar = true, rhs holds one plaintext u64 per integer (the (euintN, uintN) overloads; div/rem only exist so). */
static napi_value EvalBatch(napi_env env, napi_callback_info info) {
    size_t argc = 7, w_l = 0, w_r = 0;
    napi_value argv[7];
    napi_get_cb_info(env, info, &argc, argv, NULL, NULL);
    if (argc < 6) THROW(env, "evalBatch(handle, op, bitWidth, lhs, rhs, batch, scalar?)");
    lux_engine *e = engine_arg(env, argv[0]);
    if (!e) return NULL;
    uint32_t op = 0, bw = 0, batch = 0, scalar = 0;
    napi_get_value_uint32(env, argv[1], &op);
    napi_get_value_uint32(env, argv[2], &bw);
    napi_get_value_uint32(env, argv[5], &batch);
    if (argc > 6 && !is_nullish(env, argv[6])) napi_get_value_uint32(env, argv[6], &scalar);
    const size_t blocks = lux_radix_blocks(bw), cw = lux_ct_words(e);
    if (op >= LUX_OP_COUNT_ || !blocks) THROW(env, "evalBatch: unsupported op %u or bit width %u", op, bw);
    const uint64_t *lhs = u64_arg(env, argv[3], &w_l);
    const uint64_t *rhs = is_nullish(env, argv[4]) ? NULL : u64_arg(env, argv[4], &w_r);
    const size_t want = (size_t)batch * blocks * cw;
    if (!lhs || w_l != want) THROW(env, "evalBatch: lhs must hold %zu words", want);
    if (lux_op_is_unary((lux_op)op) ? rhs != NULL : (!rhs || w_r != (scalar ? (size_t)batch : want)))
        THROW(env, "evalBatch: rhs does not match op %u", op);
    if (lux_op_is_scalar_only((lux_op)op) && !scalar) THROW(env, "evalBatch: op %u needs a plaintext right operand (scalar = true)", op);
    job_t *j = job_new(env, scalar ? JOB_EVAL_SCALAR : JOB_EVAL, argv[0]);
    if (!j) THROW(env, "out of memory");
    j->in[0] = lhs;
    j->in[1] = rhs;
    j->op = op;
    j->bit_width = bw;
    j->batch = batch;
    j->out_words = lux_op_is_predicate((lux_op)op) ? (size_t)batch * cw : want;
    job_hold(env, j, argv[3]);
    if (rhs) job_hold(env, j, argv[4]);
    return job_submit(env, j, "luxfhe.evalBatch");
}

/* selectBatch(handle, bitWidth, cond, a, b, batch) -> Promise<BigUint64Array>     (FHE.select, WrappingERC20.sol:90-100) */
static napi_value SelectBatch(napi_env env, napi_callback_info info) {
    size_t argc = 6, w[3] = {0, 0, 0};
    napi_value argv[6];
    napi_get_cb_info(env, info, &argc, argv, NULL, NULL);
    if (argc < 6) THROW(env, "selectBatch(handle, bitWidth, cond, a, b, batch)");
    lux_engine *e = engine_arg(env, argv[0]);
    if (!e) return NULL;
    uint32_t bw = 0, batch = 0;
    napi_get_value_uint32(env, argv[1], &bw);
    napi_get_value_uint32(env, argv[5], &batch);
    const size_t blocks = lux_radix_blocks(bw), cw = lux_ct_words(e);
    job_t *j = job_new(env, JOB_SELECT, argv[0]);
    if (!j) THROW(env, "out of memory");
    for (int i = 0; i < 3; i++) j->in[i] = u64_arg(env, argv[2 + i], &w[i]);
    if (!blocks || !j->in[0] || !j->in[1] || !j->in[2] || w[0] != (size_t)batch * cw || w[1] != (size_t)batch * blocks * cw || w[2] != w[1]) {
        job_abandon(env, j);
        THROW(env, "selectBatch: operand sizes do not match bitWidth %u and batch %u", bw, batch);
    }
    j->bit_width = bw;
    j->batch = batch;
    j->out_words = w[1];
    for (int i = 0; i < 3; i++) job_hold(env, j, argv[2 + i]);
    return job_submit(env, j, "luxfhe.selectBatch");
}

/* erc20TransferBatch(handle, from, to, amount, batch) -> Promise<BigUint64Array>  [newFrom | newTo], euint64
 * (examples/erc20-tutorial/contracts/WrappingERC20.sol:86-102 as one fused circuit) */
static napi_value Erc20TransferBatch(napi_env env, napi_callback_info info) {
    size_t argc = 5, w[3] = {0, 0, 0};
    napi_value argv[5];
    napi_get_cb_info(env, info, &argc, argv, NULL, NULL);
    if (argc < 5) THROW(env, "erc20TransferBatch(handle, from, to, amount, batch)");
    lux_engine *e = engine_arg(env, argv[0]);
    if (!e) return NULL;
    uint32_t batch = 0;
    napi_get_value_uint32(env, argv[4], &batch);
    const size_t want = (size_t)batch * lux_radix_blocks(64) * lux_ct_words(e);
    job_t *j = job_new(env, JOB_ERC20, argv[0]);
    if (!j) THROW(env, "out of memory");
    for (int i = 0; i < 3; i++) j->in[i] = u64_arg(env, argv[1 + i], &w[i]);
    if (!j->in[0] || !j->in[1] || !j->in[2] || w[0] != want || w[1] != want || w[2] != want) {
        job_abandon(env, j);
        THROW(env, "erc20TransferBatch: every operand must hold %zu words", want);
    }
    j->batch = batch;
    j->out_words = 2 * want;
    for (int i = 0; i < 3; i++) job_hold(env, j, argv[1 + i]);
    return job_submit(env, j, "luxfhe.erc20TransferBatch");
}

/* gateBatch(handle, gate, a, b, batch) -> Promise<BigUint64Array>       (boolean gates, BASELINE config 1) */
static napi_value GateBatch(napi_env env, napi_callback_info info) {
    size_t argc = 5, w[2] = {0, 0};
    napi_value argv[5];
    napi_get_cb_info(env, info, &argc, argv, NULL, NULL);
    if (argc < 5) THROW(env, "gateBatch(handle, gate, a, b, batch)");
    lux_engine *e = engine_arg(env, argv[0]);
    if (!e) return NULL;
    uint32_t gate = 0, batch = 0;
    napi_get_value_uint32(env, argv[1], &gate);
    napi_get_value_uint32(env, argv[4], &batch);
    lux_params p;
    lux_engine_params(e, &p);
    const size_t want = (size_t)batch * (p.n + 1);
    job_t *j = job_new(env, JOB_GATE, argv[0]);
    if (!j) THROW(env, "out of memory");
    j->in[0] = u64_arg(env, argv[2], &w[0]);
    j->in[1] = u64_arg(env, argv[3], &w[1]);
    if (!j->in[0] || !j->in[1] || w[0] != want || w[1] != want) {
        job_abandon(env, j);
        THROW(env, "gateBatch: operands must hold %zu words", want);
    }
    j->op = gate;
    j->batch = batch;
    j->out_words = want;
    job_hold(env, j, argv[2]);
    job_hold(env, j, argv[3]);
    return job_submit(env, j, "luxfhe.gateBatch");
}

/* ---- registration ------------------------------------------------------------------------------------------ */
#define METHOD(name, fn) {name, NULL, fn, NULL, NULL, NULL, napi_default, NULL}
static napi_value lux_init(napi_env env, napi_value exports) {
    const napi_property_descriptor props[] = {
        METHOD("version", Version),           METHOD("createEngine", CreateEngine),
        METHOD("ctWords", CtWords),           METHOD("radixBlocks", RadixBlocks),
        METHOD("loadKey", LoadKey),           METHOD("trivialEncrypt", TrivialEncrypt),
        METHOD("pbsBatch", PbsBatch),         METHOD("evalBatch", EvalBatch),
        METHOD("selectBatch", SelectBatch),   METHOD("erc20TransferBatch", Erc20TransferBatch),
        METHOD("gateBatch", GateBatch),
    };
    napi_define_properties(env, exports, sizeof(props) / sizeof(props[0]), props);
    return exports;
}

#ifdef LUX_HAVE_NODE_API
NAPI_MODULE(NODE_GYP_MODULE_NAME, lux_init)
#else
/* what NAPI_MODULE_INIT provides: the well-known registration symbol node looks up with dlsym */
#ifdef __cplusplus
extern "C"
#endif
    __attribute__((visibility("default"))) napi_value
    napi_register_module_v1(napi_env env, napi_value exports) {
    return lux_init(env, exports);
}
#endif
