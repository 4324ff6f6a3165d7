// serial.cu — the engine-native wire container (include/lux_tfhe.h, "serialized keys and ciphertexts").
//
// The reference moves keys and ciphertexts as opaque bytes (packages/wasm/src/node.ts:65-96 Uint8Array in / out;
// js/v2-sdk/src/luxd/client.ts:128-147,232-244 number[] of bytes over JSON).  Those bytes are produced by the out-of-tree Go
// library and by tfhe-rs `safe_serialize` (relayer-sdk/src/test/keys/*.bin): neither format is in the tree, so this is NOT a
// converter for them (INTEGRATION.md states what one would need).  It is the part that can be done: a self-describing,
// versioned, checksummed container around the engine's word layouts, so that a key or ciphertext carries its parameter set,
// arithmetic, kind and width, and a caller that hands an S1 key to an S2 engine gets an error instead of garbage.
//
//   offset  0  magic "LUXFHE\r\n"          8 bytes
//           8  version (1)                 u32
//          12  kind (lux_blob_kind)        u32
//          16  lux_params                  9 x u32   n, N, k, ell, logB, ell_ks, logB_ks, order, arith
//          52  aux                         u32       ciphertexts: bit width (0 = single LWE); keys: LUX_DOMAIN_*
//          56  count                       u64       ciphertexts / integers in the payload (keys: 1)
//          64  payload_words               u64
//          72  payload_crc32               u32       IEEE 802.3 CRC of the payload bytes
//          76  header_crc32                u32       CRC of bytes 0..75
//          80  payload                     payload_words x u64, little-endian
// Host-only code (no CUDA calls); lux_load_key_blob is the one function that touches an engine.
#include <cstring>

#include "engine.h"
#include "../../include/lux_radix.h"

namespace {

constexpr unsigned char MAGIC[8] = {'L', 'U', 'X', 'F', 'H', 'E', '\r', '\n'};
constexpr size_t HEADER = 80;

uint32_t crc32(const unsigned char *p, size_t n, uint32_t crc = 0) {
    static uint32_t table[8][256];
    static bool ready = false;
    if (!ready) {   // slicing-by-8 tables (keys are tens of megabytes)
        for (uint32_t i = 0; i < 256; i++) {
            uint32_t c = i;
            for (int k = 0; k < 8; k++) c = (c >> 1) ^ (0xEDB88320u & (0u - (c & 1u)));
            table[0][i] = c;
        }
        for (uint32_t i = 0; i < 256; i++)
            for (int t = 1; t < 8; t++) table[t][i] = (table[t - 1][i] >> 8) ^ table[0][table[t - 1][i] & 0xFF];
        ready = true;
    }
    crc = ~crc;
    while (n >= 8) {
        uint32_t a, b;
        memcpy(&a, p, 4);
        memcpy(&b, p + 4, 4);
        a ^= crc;
        crc = table[7][a & 0xFF] ^ table[6][(a >> 8) & 0xFF] ^ table[5][(a >> 16) & 0xFF] ^ table[4][a >> 24] ^ table[3][b & 0xFF] ^
              table[2][(b >> 8) & 0xFF] ^ table[1][(b >> 16) & 0xFF] ^ table[0][b >> 24];
        p += 8;
        n -= 8;
    }
    while (n--) crc = (crc >> 8) ^ table[0][(crc ^ *p++) & 0xFF];
    return ~crc;
}

void put32(unsigned char *p, uint32_t v) { memcpy(p, &v, 4); }
void put64(unsigned char *p, uint64_t v) { memcpy(p, &v, 8); }
uint32_t get32(const unsigned char *p) {
    uint32_t v;
    memcpy(&v, p, 4);
    return v;
}
uint64_t get64(const unsigned char *p) {
    uint64_t v;
    memcpy(&v, p, 8);
    return v;
}

size_t ct_words(const lux_params &p) { return p.order == LUX_ORDER_LWEKS_PBS ? (size_t)p.k * p.N + 1 : (size_t)p.n + 1; }

// payload words of `count` objects of `kind`, or 0 when the combination is invalid
uint64_t payload_words(const lux_params &p, uint32_t kind, uint32_t aux, uint64_t count) {
    switch (kind) {
        case LUX_BLOB_LWE: return aux == 0 ? count * ct_words(p) : 0;
        case LUX_BLOB_RADIX: {
            const size_t nb = lux_radix_blocks(aux);
            return nb ? count * nb * ct_words(p) : 0;
        }
        case LUX_BLOB_BSK: return count == 1 ? (uint64_t)p.n * (p.k + 1) * p.ell * (p.k + 1) * p.N : 0;
        case LUX_BLOB_GLWE_KSK: return count == 1 ? (uint64_t)(p.k + 1) * p.ell * (p.k + 1) * p.N : 0;
        case LUX_BLOB_LWE_KSK: return count == 1 ? (uint64_t)p.k * p.N * p.ell_ks * (p.n + 1) : 0;
        default: return 0;
    }
}

bool same_params(const lux_params &a, const lux_params &b) {
    return a.n == b.n && a.N == b.N && a.k == b.k && a.ell == b.ell && a.logB == b.logB && a.ell_ks == b.ell_ks &&
           a.logB_ks == b.logB_ks && a.order == b.order && a.arith == b.arith;
}

}  // namespace

extern "C" {

size_t lux_serialized_bytes(const lux_params *p, uint32_t kind, uint32_t aux, uint64_t count) {
    if (!p) return 0;
    const uint64_t w = payload_words(*p, kind, aux, count);
    return w ? HEADER + 8 * (size_t)w : 0;
}

int lux_serialize(const lux_params *p, uint32_t kind, uint32_t aux, uint64_t count, const uint64_t *words, uint8_t *out, size_t cap) {
    if (!p || !words || !out) return LUX_ERR_INVALID;
    const uint64_t w = payload_words(*p, kind, aux, count);
    if (!w || cap < HEADER + 8 * (size_t)w) return LUX_ERR_INVALID;
    memcpy(out, MAGIC, 8);
    put32(out + 8, 1);
    put32(out + 12, kind);
    const uint32_t pv[9] = {p->n, p->N, p->k, p->ell, p->logB, p->ell_ks, p->logB_ks, p->order, p->arith};
    for (int i = 0; i < 9; i++) put32(out + 16 + 4 * i, pv[i]);
    put32(out + 52, aux);
    put64(out + 56, count);
    put64(out + 64, w);
    memcpy(out + HEADER, words, 8 * (size_t)w);   // little-endian host (x86-64 / aarch64): words are stored as they are
    put32(out + 72, crc32(out + HEADER, 8 * (size_t)w));
    put32(out + 76, crc32(out, 76));
    return LUX_OK;
}

int lux_blob_inspect(const uint8_t *blob, size_t bytes, lux_blob_info *info) {
    if (!blob || !info) return LUX_ERR_INVALID;
    memset(info, 0, sizeof(*info));
    if (bytes < HEADER || memcmp(blob, MAGIC, 8) != 0) return LUX_ERR_FORMAT;
    if (get32(blob + 76) != crc32(blob, 76)) return LUX_ERR_FORMAT;
    info->version = get32(blob + 8);
    if (info->version != 1) return LUX_ERR_FORMAT;
    info->kind = get32(blob + 12);
    uint32_t *pv[9] = {&info->params.n, &info->params.N, &info->params.k, &info->params.ell, &info->params.logB,
                       &info->params.ell_ks, &info->params.logB_ks, &info->params.order, &info->params.arith};
    for (int i = 0; i < 9; i++) *pv[i] = get32(blob + 16 + 4 * i);
    info->aux = get32(blob + 52);
    info->count = get64(blob + 56);
    info->payload_words = get64(blob + 64);
    if (info->payload_words != payload_words(info->params, info->kind, info->aux, info->count) || info->payload_words == 0) return LUX_ERR_FORMAT;
    if (bytes != HEADER + 8 * (size_t)info->payload_words) return LUX_ERR_FORMAT;
    return LUX_OK;
}

// expect: the parameter set the caller works with (NULL = accept any); expect_kind 0 = any.  The payload CRC is checked here.
int lux_deserialize(const uint8_t *blob, size_t bytes, const lux_params *expect, uint32_t expect_kind, uint64_t *words_out, size_t cap_words,
                    lux_blob_info *info) {
    lux_blob_info local;
    if (!info) info = &local;
    int rc = lux_blob_inspect(blob, bytes, info);
    if (rc) return rc;
    if (expect && !same_params(*expect, info->params)) return LUX_ERR_MISMATCH;
    if (expect_kind && expect_kind != info->kind) return LUX_ERR_MISMATCH;
    if (get32(blob + 72) != crc32(blob + HEADER, 8 * (size_t)info->payload_words)) return LUX_ERR_FORMAT;
    if (words_out) {
        if (cap_words < info->payload_words) return LUX_ERR_INVALID;
        memcpy(words_out, blob + HEADER, 8 * (size_t)info->payload_words);
    }
    return LUX_OK;
}

// one serialized evaluation key -> the engine (dispatch on the container's kind); a key of another parameter set or
// arithmetic is refused with LUX_ERR_MISMATCH before any byte reaches the GPU
int lux_load_key_blob(lux_engine *e, const uint8_t *blob, size_t bytes) {
    LUX_LOCK(e);
    if (!e || !blob) return LUX_ERR_INVALID;
    lux_blob_info info;
    int rc = lux_deserialize(blob, bytes, &e->p, 0, nullptr, 0, &info);
    if (rc == LUX_ERR_MISMATCH) LUX_FAIL(e, rc, "key container is for another parameter set / arithmetic (n=%u N=%u arith=%u) than this engine (n=%u N=%u arith=%u)",
                                         info.params.n, info.params.N, info.params.arith, e->p.n, e->p.N, e->p.arith);
    if (rc) LUX_FAIL(e, rc, "not a valid key container (bad magic, version, length or checksum)");
    const uint64_t *w = (const uint64_t *)(blob + HEADER);   // 80 is a multiple of 8: aligned whenever the blob is
    std::vector<uint64_t> copy;
    if (((uintptr_t)w & 7) != 0) {
        copy.resize(info.payload_words);
        memcpy(copy.data(), blob + HEADER, 8 * (size_t)info.payload_words);
        w = copy.data();
    }
    switch (info.kind) {
        case LUX_BLOB_BSK: return lux_load_bsk(e, w, info.payload_words, (int)info.aux);
        case LUX_BLOB_GLWE_KSK: return lux_load_glwe_ksk(e, w, info.payload_words, (int)info.aux);
        case LUX_BLOB_LWE_KSK: return lux_load_lwe_ksk(e, w, info.payload_words);
        default: LUX_FAIL(e, LUX_ERR_MISMATCH, "container kind %u is not an evaluation key", info.kind);
    }
}

}  // extern "C"
