// multi.cu — several B200s of one box behind ONE handle of the C ABI (include/lux_tfhe.h, lux_engine_create_multi ...).
//
// SURVEY 8(b)'s boundary is `create(params, n_gpus)`: the reference's host is a Node process (packages/wasm/src/node.ts,
// js/v2-sdk/src/luxd/client.ts) with no torch and no NCCL of its own, so replication and sharding live inside the
// library.  One lux_engine per GPU; evaluation keys are converted once on GPU 0 and copied device to device
// (cudaMemcpyPeerAsync over NVLink) into the other engines; a batch call slices its batch into contiguous shares, one
// host thread per GPU runs the single-GPU entry point on its share (each engine has its own streams, pinned staging and
// lock), and the call returns when all shares are back.  No collective on the data path (north_star: independent
// ciphertexts of a batch are sharded across the GPUs of one box).
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <new>
#include <string>
#include <thread>
#include <vector>

#include "engine.h"
#include "../../include/lux_radix.h"

struct lux_multi {
    std::vector<lux_engine *> eng;
    std::mutex mtx;   // one multi-GPU call at a time per handle (the engines below are locked by their own calls)
    char err[512] = "";
};

static thread_local char g_multi_create_error[512] = "";

namespace {

// [lo, hi) of share g: whole kernel waves first (a share that is a multiple of the wave keeps every GPU's last wave full)
void share_of(size_t B, size_t n, size_t g, size_t unit, size_t &lo, size_t &hi) {
    if (unit == 0) unit = 1;
    const size_t units = (B + unit - 1) / unit;
    lo = std::min(B, units * g / n * unit);
    hi = std::min(B, units * (g + 1) / n * unit);
}

template <typename F>
int run_sharded(lux_multi *m, size_t B, size_t unit, F &&fn) {
    const size_t n = m->eng.size();
    std::vector<int> rc(n, LUX_OK);
    std::vector<std::thread> th;
    for (size_t g = 0; g < n; g++) {
        size_t lo, hi;
        share_of(B, n, g, unit, lo, hi);
        if (hi <= lo) continue;
        th.emplace_back([&, g, lo, hi] { rc[g] = fn(m->eng[g], lo, hi - lo); });
    }
    for (auto &t : th) t.join();
    for (size_t g = 0; g < n; g++)
        if (rc[g] != LUX_OK) {
            snprintf(m->err, sizeof(m->err), "GPU %d: %.480s", m->eng[g]->device, m->eng[g]->err);
            return rc[g];
        }
    return LUX_OK;
}

int replicate(lux_multi *m) {
    lux_engine *e0 = m->eng[0];
    const size_t bytes = lux_key_blob_bytes(e0);
    if (m->eng.size() == 1) return LUX_OK;
    void *src = nullptr;
    if (cudaSetDevice(e0->device) != cudaSuccess || cudaMalloc(&src, bytes) != cudaSuccess) {
        snprintf(m->err, sizeof(m->err), "cudaMalloc(%zu) for the key blob on GPU %d failed", bytes, e0->device);
        return LUX_ERR_OOM;
    }
    int rc = lux_export_keys_device(e0, src, bytes);
    std::vector<void *> dst(m->eng.size(), nullptr);
    for (size_t g = 1; g < m->eng.size() && rc == LUX_OK; g++) {
        lux_engine *e = m->eng[g];
        if (cudaSetDevice(e->device) != cudaSuccess || cudaMalloc(&dst[g], bytes) != cudaSuccess) {
            snprintf(m->err, sizeof(m->err), "cudaMalloc(%zu) for the key blob on GPU %d failed", bytes, e->device);
            rc = LUX_ERR_OOM;
            break;
        }
        int can = 0;
        cudaDeviceCanAccessPeer(&can, e->device, e0->device);
        if (can) {
            cudaError_t st = cudaDeviceEnablePeerAccess(e0->device, 0);
            if (st != cudaSuccess && st != cudaErrorPeerAccessAlreadyEnabled) can = 0;
            cudaGetLastError();
        }
        // cudaMemcpyPeer stages through the host when the GPUs are not peers; with NVLink it is a direct copy
        if (cudaMemcpyPeer(dst[g], e->device, src, e0->device, bytes) != cudaSuccess) {
            snprintf(m->err, sizeof(m->err), "key copy GPU %d -> GPU %d failed: %s", e0->device, e->device, cudaGetErrorString(cudaGetLastError()));
            rc = LUX_ERR_CUDA;
            break;
        }
        rc = lux_import_keys_device(e, dst[g], bytes);
        if (rc) snprintf(m->err, sizeof(m->err), "GPU %d: %.480s", e->device, e->err);
    }
    for (size_t g = 1; g < m->eng.size(); g++)
        if (dst[g]) {
            cudaSetDevice(m->eng[g]->device);
            cudaFree(dst[g]);
        }
    cudaSetDevice(e0->device);
    cudaFree(src);
    if (rc && !m->err[0]) snprintf(m->err, sizeof(m->err), "GPU %d: %.480s", e0->device, e0->err);
    return rc;
}

}  // namespace

extern "C" {

const char *lux_multi_last_error(const lux_multi *m) { return m ? m->err : g_multi_create_error; }

int lux_engine_create_multi(const lux_params *p, int n_gpus, lux_multi **out) {
    if (!p || !out) return LUX_ERR_INVALID;
    *out = nullptr;
    int have = 0;
    if (cudaGetDeviceCount(&have) != cudaSuccess || have == 0) {
        snprintf(g_multi_create_error, sizeof(g_multi_create_error), "no CUDA device");
        return LUX_ERR_CUDA;
    }
    if (n_gpus <= 0) n_gpus = have;
    if (n_gpus > have) {
        snprintf(g_multi_create_error, sizeof(g_multi_create_error), "%d GPUs requested, %d visible", n_gpus, have);
        return LUX_ERR_INVALID;
    }
    lux_multi *m = new (std::nothrow) lux_multi();
    if (!m) return LUX_ERR_OOM;
    for (int g = 0; g < n_gpus; g++) {
        lux_engine *e = nullptr;
        int rc = lux_engine_create(p, g, &e);
        if (rc) {
            snprintf(g_multi_create_error, sizeof(g_multi_create_error), "GPU %d: %.480s", g, lux_last_error(nullptr));
            for (auto *x : m->eng) lux_engine_destroy(x);
            delete m;
            return rc;
        }
        m->eng.push_back(e);
    }
    *out = m;
    return LUX_OK;
}

void lux_multi_destroy(lux_multi *m) {
    if (!m) return;
    m->mtx.lock();
    for (auto *e : m->eng) lux_engine_destroy(e);
    m->mtx.unlock();
    delete m;
}

int lux_multi_gpus(const lux_multi *m) { return m ? (int)m->eng.size() : 0; }
lux_engine *lux_multi_engine(lux_multi *m, int i) { return (m && i >= 0 && i < (int)m->eng.size()) ? m->eng[i] : nullptr; }

// keys: converted on GPU 0; lux_multi_replicate_keys copies the resident blob to the other GPUs (call it after the last load)
#define MULTI_LOAD(call)                                                                             \
    if (!m) return LUX_ERR_INVALID;                                                                  \
    std::lock_guard<std::mutex> lk(m->mtx);                                                          \
    int rc = (call);                                                                                 \
    if (rc) snprintf(m->err, sizeof(m->err), "GPU %d: %.480s", m->eng[0]->device, m->eng[0]->err);       \
    return rc;
int lux_multi_load_bsk(lux_multi *m, const uint64_t *bsk, size_t n_words, int domain) { MULTI_LOAD(lux_load_bsk(m->eng[0], bsk, n_words, domain)) }
int lux_multi_load_glwe_ksk(lux_multi *m, const uint64_t *ksk, size_t n_words, int domain) {
    MULTI_LOAD(lux_load_glwe_ksk(m->eng[0], ksk, n_words, domain))
}
int lux_multi_load_lwe_ksk(lux_multi *m, const uint64_t *ksk, size_t n_words) { MULTI_LOAD(lux_load_lwe_ksk(m->eng[0], ksk, n_words)) }
int lux_multi_replicate_keys(lux_multi *m) {
    if (!m) return LUX_ERR_INVALID;
    std::lock_guard<std::mutex> lk(m->mtx);
    m->err[0] = 0;
    return replicate(m);
}

// lux_pbs_batch over all GPUs: share g = ciphertexts [lo, hi), same test vectors everywhere
int lux_multi_pbs_batch(lux_multi *m, const uint64_t *lwe_in, const uint64_t *testvs, uint32_t n_luts, const uint32_t *lut_idx, size_t B,
                        uint64_t *lwe_out) {
    if (!m) return LUX_ERR_INVALID;
    std::lock_guard<std::mutex> lk(m->mtx);
    if (!lwe_in || !testvs || !lwe_out || n_luts == 0) {
        snprintf(m->err, sizeof(m->err), "lux_multi_pbs_batch: null/empty argument");
        return LUX_ERR_INVALID;
    }
    const size_t w = lux_ct_words(m->eng[0]);
    const size_t per = (B + m->eng.size() - 1) / m->eng.size();
    const size_t unit = lux_wave_ciphertexts(m->eng[0], per);
    return run_sharded(m, B, B >= 2 * unit * m->eng.size() ? unit : 1, [&](lux_engine *e, size_t lo, size_t nb) {
        return lux_pbs_batch(e, lwe_in + lo * w, testvs, n_luts, lut_idx ? lut_idx + lo : nullptr, nb, lwe_out + lo * w);
    });
}

// lux_eval / lux_select / lux_erc20_transfer over all GPUs: shares of whole integers
int lux_multi_eval(lux_multi *m, lux_op op, unsigned bit_width, const uint64_t *lhs, const uint64_t *rhs, size_t batch, uint64_t *out) {
    if (!m) return LUX_ERR_INVALID;
    std::lock_guard<std::mutex> lk(m->mtx);
    const size_t nb = lux_radix_blocks(bit_width), w = lux_ct_words(m->eng[0]);
    const size_t in_stride = nb * w, out_stride = (lux_op_is_predicate(op) ? 1 : nb) * w;
    return run_sharded(m, batch, 1, [&](lux_engine *e, size_t lo, size_t n) {
        return lux_eval(e, op, bit_width, lhs + lo * in_stride, rhs ? rhs + lo * in_stride : nullptr, n, out + lo * out_stride);
    });
}
int lux_multi_select(lux_multi *m, unsigned bit_width, const uint64_t *cond, const uint64_t *a, const uint64_t *b, size_t batch, uint64_t *out) {
    if (!m) return LUX_ERR_INVALID;
    std::lock_guard<std::mutex> lk(m->mtx);
    const size_t nb = lux_radix_blocks(bit_width), w = lux_ct_words(m->eng[0]);
    return run_sharded(m, batch, 1, [&](lux_engine *e, size_t lo, size_t n) {
        return lux_select(e, bit_width, cond + lo * w, a + lo * nb * w, b + lo * nb * w, n, out + lo * nb * w);
    });
}
int lux_multi_erc20_transfer(lux_multi *m, const uint64_t *from_bal, const uint64_t *to_bal, const uint64_t *amount, size_t batch,
                             uint64_t *new_from, uint64_t *new_to) {
    if (!m) return LUX_ERR_INVALID;
    std::lock_guard<std::mutex> lk(m->mtx);
    const size_t s = 32 * lux_ct_words(m->eng[0]);
    return run_sharded(m, batch, 1, [&](lux_engine *e, size_t lo, size_t n) {
        return lux_erc20_transfer(e, from_bal + lo * s, to_bal + lo * s, amount + lo * s, n, new_from + lo * s, new_to + lo * s);
    });
}

}  // extern "C"
