/*
 * lux_radix.h — euintN operations over the batched programmable bootstrap (libluxtfhe_b200.so).
 *
 * Reference interfaces replaced (paths relative to /root/reference):
 *   packages/wasm/src/index.ts:19-22        LuxFHE.add / sub / eq / lt (ct1, ct2, bootstrapKey, secretKey)
 *   js/v2-sdk/src/luxd/client.ts:34-39,128-147   POST /evaluate {op: add|sub|eq|lt|gt|and|or|xor|not|shl|shr, left, right?, bitWidth}
 *   examples/erc20-tutorial/contracts/WrappingERC20.sol:86-102   FHE.le / sub / add / select on euint64
 *   mocks/foundry/src/MockFHE.sol:130-508   plaintext semantics of every op
 *   tests/fhevm-suite/e2e/test/fhevmOperations/fhevmOperations{1..13}.ts   2,394 known-answer tests (tests/golden/fhevm_kats.json)
 * The reference's own integer layer (Go luxfi/tfhe) is not in the tree; the circuits below are this
 * library's design, their *semantics* are pinned by the known-answer tests above.
 *
 * Encoding.  An euintN is bit_width/2 "blocks", least significant first; each block is one LWE
 * ciphertext of lux_ct_words() words encoding a value v in [0,16) as v * lux_delta(32) (2 message bits,
 * 2 carry bits, 1 padding bit — the PARAM_MESSAGE_2_CARRY_2 layout named at sdk/relayer/src/tfhe.ts:15).
 * Inputs must be clean (v < 4); outputs are clean.  An ebool is ONE block holding 0 or 1.
 * Buffers: integer-major, [batch][blocks][ct_words] u64; ebool results [batch][ct_words].
 *
 * Every function returns LUX_OK or a negative lux_status (lux_tfhe.h) and needs the evaluation keys of
 * the engine to be loaded.  Host pointers; the *_device variants take device pointers of the same layout.
 */
#ifndef LUX_RADIX_H
#define LUX_RADIX_H
#include "lux_tfhe.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    /* arithmetic, result euintN */
    LUX_OP_ADD = 0,
    LUX_OP_SUB = 1,
    LUX_OP_MUL = 2,
    LUX_OP_NEG = 3, /* unary */
    /* comparisons, result ebool */
    LUX_OP_EQ = 4,
    LUX_OP_NE = 5,
    LUX_OP_LT = 6,
    LUX_OP_LE = 7,
    LUX_OP_GT = 8,
    LUX_OP_GE = 9,
    LUX_OP_MIN = 10,
    LUX_OP_MAX = 11,
    /* bitwise, result euintN */
    LUX_OP_AND = 12,
    LUX_OP_OR = 13,
    LUX_OP_XOR = 14,
    LUX_OP_NOT = 15, /* unary */
    /* shifts / rotations by an encrypted amount (an euintN of the same width; taken mod bit_width) */
    LUX_OP_SHL = 16,
    LUX_OP_SHR = 17,
    LUX_OP_ROTL = 18,
    LUX_OP_ROTR = 19,
    /* division by a PLAINTEXT divisor (lux_eval_scalar only): the fhEVM `div`/`rem` overloads
     * (tests/fhevm-suite/e2e/test/fhevmOperations/fhevmOperations8.ts: `div (euint8, uint8)` ...) take a
     * plaintext right operand; the circuit is divisor-specific (multiply by a reciprocal, shift) */
    LUX_OP_DIV = 20,
    LUX_OP_REM = 21,
    LUX_OP_COUNT_ = 22
} lux_op;

/* blocks of an euint of this width (bit_width even, 2..256); 0 if unsupported */
size_t lux_radix_blocks(unsigned bit_width);
/* 1 when the op's result is an ebool (one block per integer) */
int lux_op_is_predicate(lux_op op);
int lux_op_is_unary(lux_op op);
/* 1 for ops that exist only with a plaintext right operand (div, rem) */
int lux_op_is_scalar_only(lux_op op);
/* programmable bootstraps one integer costs for this op and width (bench: ops/s <-> PBS/s), and the
 * number of sequential PBS levels (launch pairs) of the circuit */
uint64_t lux_eval_pbs_count(lux_op op, unsigned bit_width);
uint64_t lux_eval_depth(lux_op op, unsigned bit_width);
/* same for the plaintext-operand form; for div/rem the circuit (and its size) depends on the divisor */
uint64_t lux_eval_scalar_pbs_count(lux_op op, unsigned bit_width, uint64_t scalar);

/* out = op(lhs, rhs) for `batch` independent integers; rhs = NULL for unary ops.
 * replaces LuxFHE.add/sub/eq/lt (index.ts:19-22) and POST /evaluate (client.ts:128-147) */
int lux_eval(lux_engine *e, lux_op op, unsigned bit_width, const uint64_t *lhs, const uint64_t *rhs, size_t batch,
             uint64_t *out);
int lux_eval_device(lux_engine *e, lux_op op, unsigned bit_width, const uint64_t *d_lhs, const uint64_t *d_rhs,
                    size_t batch, uint64_t *d_out);
/* same with a plaintext right operand per integer (the `(euintN, uintN)` overloads of the fhEVM suite);
 * LUX_OP_DIV / LUX_OP_REM exist only in this form; a zero divisor is LUX_ERR_INVALID */
int lux_eval_scalar(lux_engine *e, lux_op op, unsigned bit_width, const uint64_t *lhs, const uint64_t *scalars,
                    size_t batch, uint64_t *out);
/* plaintext operands wider than 64 bits (the euint128 / euint256 div and rem overloads): scalar b is the
 * n_limbs little-endian words scalars[b * n_limbs ..], n_limbs in 1..4.  div/rem only. */
int lux_eval_scalar_wide(lux_engine *e, lux_op op, unsigned bit_width, const uint64_t *lhs, const uint64_t *scalars,
                         size_t n_limbs, size_t batch, uint64_t *out);
/* out = cond ? a : b     (FHE.select, WrappingERC20.sol:90-100); cond [batch][ct_words] ebool */
int lux_select(lux_engine *e, unsigned bit_width, const uint64_t *cond, const uint64_t *a, const uint64_t *b,
               size_t batch, uint64_t *out);
int lux_select_device(lux_engine *e, unsigned bit_width, const uint64_t *d_cond, const uint64_t *d_a,
                      const uint64_t *d_b, size_t batch, uint64_t *d_out);
/* zero-extend or truncate (FHE.asEuintN casts); pure data movement, no bootstrap */
int lux_cast(lux_engine *e, unsigned from_width, unsigned to_width, const uint64_t *in, size_t batch, uint64_t *out);
/* noise-free ("trivial") encryptions of plaintext values: mask 0, body = block * delta */
int lux_trivial_encrypt(lux_engine *e, unsigned bit_width, const uint64_t *values, size_t batch, uint64_t *out);

/* The confidential ERC20 transfer of examples/erc20-tutorial/contracts/WrappingERC20.sol:86-102 on
 * euint64, fused into one level-scheduled circuit:
 *   ok = le(amount, from);  new_from = select(ok, from - amount, from);  new_to = select(ok, to + amount, to) */
int lux_erc20_transfer(lux_engine *e, const uint64_t *from_bal, const uint64_t *to_bal, const uint64_t *amount,
                       size_t batch, uint64_t *new_from, uint64_t *new_to);
int lux_erc20_transfer_device(lux_engine *e, const uint64_t *d_from, const uint64_t *d_to, const uint64_t *d_amount,
                              size_t batch, uint64_t *d_new_from, uint64_t *d_new_to);
uint64_t lux_erc20_pbs_count(void);
uint64_t lux_erc20_depth(void);

/* the same three calls over every GPU of a lux_multi handle (lux_tfhe.h): shares of whole integers / transfers, one host
 * thread per GPU, no data-path collective.  BASELINE configs[2..4]: "batch-sharded over 2/4/8 GPUs". */
int lux_multi_eval(lux_multi *m, lux_op op, unsigned bit_width, const uint64_t *lhs, const uint64_t *rhs, size_t batch, uint64_t *out);
int lux_multi_select(lux_multi *m, unsigned bit_width, const uint64_t *cond, const uint64_t *a, const uint64_t *b, size_t batch,
                     uint64_t *out);
int lux_multi_erc20_transfer(lux_multi *m, const uint64_t *from_bal, const uint64_t *to_bal, const uint64_t *amount, size_t batch,
                             uint64_t *new_from, uint64_t *new_to);

/* device memory the radix layer may use for intermediate ciphertexts (default 16 GiB); batches that need
 * more are processed in chunks */
int lux_radix_set_pool_bytes(lux_engine *e, size_t bytes);

/* ---- introspection for the parity tests (no GPU needed) ----------------------------------------
 * The circuit of an op as a flat int64 stream, so that the CPU oracle can interpret exactly the
 * program the GPU executes (tests/ only):
 *   [n_inputs, blocks(in_0).., n_outputs, blocks(out_0).., n_levels,
 *    per level: n_pbs, then per pbs: lut_id, out_slot, konst, n_terms, (slot, coef)*,
 *    per output, per block: konst, n_terms, (slot, coef)*]
 * slots: input operand k block i = (sum of blocks of operands < k) + i; PBS outputs follow.
 * op = LUX_OP_* or LUX_PROG_SELECT / LUX_PROG_ERC20.  Returns the stream length (call with cap = 0 to size). */
#define LUX_PROG_SELECT 100
#define LUX_PROG_ERC20 101
/* OR-ed into `op`: the shallow variant of the circuit (carry chains resolved by a parallel-prefix scan: log2(blocks) + 2
 * levels instead of `blocks`, ~3.5x the bootstraps), which lux_eval* / lux_erc20_transfer* run on their own when
 * batch * blocks <= 296, i.e. while a level fits the latency-mode kernels and depth is what costs time */
#define LUX_PROG_SHALLOW 0x1000
/* op | LUX_PROG_NODUAL: the variant without dual tables (see lux_radix_lut_dual) */
#define LUX_PROG_NODUAL 0x2000
size_t lux_radix_program(int op, unsigned bit_width, int64_t *buf, size_t cap);
/* the divisor-specific circuit of LUX_OP_DIV / LUX_OP_REM (other ops: same as lux_radix_program) */
size_t lux_radix_program_scalar(int op, unsigned bit_width, const uint64_t *scalar_limbs, size_t n_limbs, int64_t *buf,
                                size_t cap);
/* the 16-entry table of lookup table `lut_id`; returns the number of tables when out == NULL */
int lux_radix_lut(int lut_id, uint64_t out[16]);
/* 1 when table `lut_id` is a signed table: its circuits feed it x in (-16, 16) and rely on the negacyclic extension
   T(x) = -T(x + 16) of the blind rotation for x < 0 (the padding bit is the sign bit); every other table is only ever fed
   x in [0, 16).  Test infrastructure (the CPU interpreters of oracle/radix.py) needs it to mirror that rule. */
int lux_radix_lut_signed(int lut_id);
/* 1 when table `lut_id` is a dual table (many-LUT bootstrap): its node has TWO output slots, out and out + 1, read from ONE
   blind rotation at coefficient 0 and at coefficient N/16 of the accumulator, i.e. T(x) and T(x + 1) for the (even) input x.
   Programs with such nodes run on engines of order LUX_ORDER_LWEKS_PBS; the others take the LUX_PROG_NODUAL variant. */
int lux_radix_lut_dual(int lut_id);

#ifdef __cplusplus
}
#endif
#endif
