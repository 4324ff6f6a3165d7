/**
 * @luxfhe/b200 — TypeScript host side of the B200 evaluation engine.
 *
 * Mirrors the evaluation half of the reference's `LuxFHE` object
 * (reference: packages/wasm/src/index.ts:14-23; Node loader packages/wasm/src/node.ts:29-96; HTTP twin
 * js/v2-sdk/src/luxd/client.ts:34-39,128-147) over the N-API addon in luxfhe_b200_napi.c.  Key generation,
 * encryption and decryption stay on the reference path (`generateKeys`, `encrypt`, `decrypt` are not here):
 * the caller hands over evaluation keys and ciphertexts as little-endian u64 arrays (SURVEY.md 8b "wire formats").
 *
 * Differences a caller of the reference sees:
 *   - `init()` returns an object bound to one GPU, or to several with `params.gpus` (the library replicates the keys and
 *     slices every batch itself); the bootstrap key is uploaded ONCE (`loadKeys`), which returns the handle that takes the
 *     place of the `bootstrapKey` string: `add(ct1, ct2, bootstrapKey, secretKey)` keeps its name and arity, with
 *     ciphertexts as base64 strings of the engine's self-describing container and `secretKey` ignored;
 *   - every op has a `*Batch` form that takes `batch` integers in one BigUint64Array
 *     ([batch][blocks][ctWords]); the single-ciphertext methods are the batch-of-one case;
 *   - evaluation methods return Promises (the addon runs them on the libuv pool; several un-awaited calls on one engine
 *     are safe: the library serialises them per engine);
 *   - errors keep the reference's messages where it has them ("LuxFHE not initialized. Call init() first.",
 *     "Evaluation failed: ...").
 */

/* eslint-disable @typescript-eslint/no-var-requires */
declare const require: (id: string) => any;
declare const Buffer: any;

export interface LuxParams {
  n: number; N: number; k: number; ell: number; logB: number; ellKs: number; logBKs: number;
  /** 0 = PBS then GLWE key switch (in-tree order), 1 = LWE key switch then PBS (KS_PBS sets) */
  order: 0 | 1 | 2;
  /** 0 = in-tree Goldilocks exact NTT, 1 = 2^64 torus with f64 FFT (the arithmetic of the named tfhe-rs set; ell/logB are level/base_log) */
  arith?: 0 | 1;
  /** >= 1: one handle over that many GPUs of the box (0xFFFFFFFF = all visible); absent / 0: the single GPU `device` */
  gpus?: number;
}

/** parameter sets of SURVEY.md Appendix B */
export const S1_BOOLEAN: LuxParams = { n: 728, N: 1024, k: 1, ell: 4, logB: 5, ellKs: 4, logBKs: 5, order: 0 };
export const S2_SHORTINT: LuxParams = { n: 918, N: 2048, k: 1, ell: 1, logB: 41, ellKs: 4, logBKs: 4, order: 1 };
/** PARAM_MESSAGE_2_CARRY_2_KS_PBS_TUNIFORM_2M128 on its own arithmetic (sdk/relayer/src/tfhe.ts:13-28) */
export const S2_TORUS: LuxParams = { n: 918, N: 2048, k: 1, ell: 1, logB: 23, ellKs: 4, logBKs: 4, order: 1, arith: 1 };

/** op codes of include/lux_radix.h (`lux_op`); names follow client.ts:35 and the fhEVM operator list */
export enum Op {
  add = 0, sub, mul, neg, eq, ne, lt, le, gt, ge, min, max, and, or, xor, not, shl, shr, rotl, rotr, div, rem,
}
export type OpName = keyof typeof Op;

/** gate codes of include/lux_tfhe.h (`lux_gate_batch`) */
export enum Gate { nand = 0, and, or, nor, xor, xnor }

export interface EvaluationKeys {
  /** GGSW array [n][k+1][ell][k+1][N], coefficient (0) or NTT (1) domain */
  bootstrapKey: BigUint64Array;
  bootstrapKeyDomain?: 0 | 1;
  /** order 1: LWE key switch key [kN][ellKs][n+1] */
  lweKeySwitchKey?: BigUint64Array;
  /** order 0: one GGSW */
  glweKeySwitchKey?: BigUint64Array;
  glweKeySwitchKeyDomain?: 0 | 1;
}

type Handle = unknown;
interface Native {
  version(): string;
  createEngine(p: LuxParams, device?: number): Handle;
  ctWords(h: Handle): number;
  radixBlocks(bitWidth: number): number;
  loadKey(h: Handle, which: 0 | 1 | 2, words: BigUint64Array, domain?: number): void;
  trivialEncrypt(h: Handle, bitWidth: number, values: BigUint64Array): BigUint64Array;
  pbsBatch(h: Handle, lwe: BigUint64Array, tvs: BigUint64Array, nLuts: number, idx: Uint32Array | null, batch: number): Promise<BigUint64Array>;
  evalBatch(h: Handle, op: number, bitWidth: number, lhs: BigUint64Array, rhs: BigUint64Array | null, batch: number, scalar?: number): Promise<BigUint64Array>;
  selectBatch(h: Handle, bitWidth: number, cond: BigUint64Array, a: BigUint64Array, b: BigUint64Array, batch: number): Promise<BigUint64Array>;
  erc20TransferBatch(h: Handle, from: BigUint64Array, to: BigUint64Array, amount: BigUint64Array, batch: number): Promise<BigUint64Array>;
  gateBatch(h: Handle, gate: number, a: BigUint64Array, b: BigUint64Array, batch: number): Promise<BigUint64Array>;
}

/** stands for the reference's `bootstrapKey: string` argument once the keys live on the GPU */
export interface KeyHandle { readonly id: number }
let keyIds = 0;

// ---- the engine's serialized container (tfhe_b200/csrc/serial.cu): 80-byte little-endian header + u64 words ----------------
const MAGIC = [0x4c, 0x55, 0x58, 0x46, 0x48, 0x45, 0x0d, 0x0a];   // "LUXFHE\r\n"
const BLOB_RADIX = 2;
let crcTable: Uint32Array | null = null;
function crc32(bytes: Uint8Array): number {
  if (!crcTable) {
    crcTable = new Uint32Array(256);
    for (let i = 0; i < 256; i++) {
      let c = i;
      for (let k = 0; k < 8; k++) c = (c & 1) ? (0xedb88320 ^ (c >>> 1)) : (c >>> 1);
      crcTable[i] = c >>> 0;
    }
  }
  let crc = 0xffffffff;
  for (let i = 0; i < bytes.length; i++) crc = crcTable[(crc ^ bytes[i]) & 0xff] ^ (crc >>> 8);
  return (crc ^ 0xffffffff) >>> 0;
}
function paramWords(p: LuxParams): number[] {
  return [p.n, p.N, p.k, p.ell, p.logB, p.ellKs, p.logBKs, p.order, p.arith ?? 0];
}
export function encodeContainer(p: LuxParams, bitWidth: number, count: number, words: BigUint64Array): Uint8Array {
  const out = new Uint8Array(80 + words.length * 8);
  const dv = new DataView(out.buffer);
  out.set(MAGIC, 0);
  dv.setUint32(8, 1, true);
  dv.setUint32(12, BLOB_RADIX, true);
  paramWords(p).forEach((v, i) => dv.setUint32(16 + 4 * i, v, true));
  dv.setUint32(52, bitWidth, true);
  dv.setBigUint64(56, BigInt(count), true);
  dv.setBigUint64(64, BigInt(words.length), true);
  out.set(new Uint8Array(words.buffer, words.byteOffset, words.length * 8), 80);
  dv.setUint32(72, crc32(out.subarray(80)), true);
  dv.setUint32(76, crc32(out.subarray(0, 76)), true);
  return out;
}
export function decodeContainer(bytes: Uint8Array, expect: LuxParams): { bitWidth: number; count: number; words: BigUint64Array } {
  const dv = new DataView(bytes.buffer, bytes.byteOffset, bytes.byteLength);
  if (bytes.length < 80 || MAGIC.some((m, i) => bytes[i] !== m) || dv.getUint32(8, true) !== 1 ||
      dv.getUint32(76, true) !== crc32(bytes.subarray(0, 76)))
    throw new Error('Evaluation failed: not a serialized LuxFHE ciphertext');
  if (dv.getUint32(12, true) !== BLOB_RADIX || paramWords(expect).some((v, i) => dv.getUint32(16 + 4 * i, true) !== v))
    throw new Error('Evaluation failed: ciphertext belongs to another parameter set or arithmetic');
  const nWords = Number(dv.getBigUint64(64, true));
  if (bytes.length !== 80 + nWords * 8 || dv.getUint32(72, true) !== crc32(bytes.subarray(80)))
    throw new Error('Evaluation failed: ciphertext is damaged (length / checksum)');
  const words = new BigUint64Array(nWords);       // copy: the payload of a base64-decoded Buffer is not 8-byte aligned
  new Uint8Array(words.buffer).set(bytes.subarray(80));
  return { bitWidth: dv.getUint32(52, true), count: Number(dv.getBigUint64(56, true)), words };
}

let native: Native | null = null;
let instance: LuxFHEB200 | null = null;

function loadNative(): Native {
  if (!native) native = require('./build/Release/luxfhe_b200.node') as Native;
  return native;
}

export class LuxFHEB200 {
  private readonly h: Handle;
  private readonly n: Native;
  readonly params: LuxParams;
  readonly ctWords: number;

  constructor(params: LuxParams, device = 0) {
    this.n = loadNative();
    this.h = this.n.createEngine(params, device);   // throws without an sm_100 device: there is no CPU fallback
    this.params = params;
    this.ctWords = this.n.ctWords(this.h);
  }

  version(): string { return this.n.version(); }

  private keyHandle: KeyHandle | null = null;

  /** upload once; converted on the device and kept resident in HBM.  Returns the handle that stands for the reference's
   *  per-call `bootstrapKey` string. */
  loadKeys(keys: EvaluationKeys): KeyHandle {
    this.keyHandle = { id: ++keyIds };
    this.n.loadKey(this.h, 0, keys.bootstrapKey, keys.bootstrapKeyDomain ?? 0);
    if (this.params.order === 1) {
      if (!keys.lweKeySwitchKey) throw new Error('loadKeys: lweKeySwitchKey is required for KS_PBS parameter sets');
      this.n.loadKey(this.h, 2, keys.lweKeySwitchKey);
    } else {
      if (!keys.glweKeySwitchKey) throw new Error('loadKeys: glweKeySwitchKey is required for PBS_KS parameter sets');
      this.n.loadKey(this.h, 1, keys.glweKeySwitchKey, keys.glweKeySwitchKeyDomain ?? 0);
    }
    return this.keyHandle;
  }

  blocks(bitWidth: number): number {
    const b = this.n.radixBlocks(bitWidth);
    if (!b) throw new Error(`Unsupported bit width: ${bitWidth}`);
    return b;
  }

  private batchOf(ct: BigUint64Array, bitWidth: number): number {
    const per = this.blocks(bitWidth) * this.ctWords;
    if (ct.length === 0 || ct.length % per !== 0) throw new Error(`ciphertext array is not a multiple of ${per} words`);
    return ct.length / per;
  }

  /** POST /evaluate {op, left, right?, bitWidth} (client.ts:128-147) for `batch` integers at once */
  evalBatch(op: OpName | Op, bitWidth: number, left: BigUint64Array, right: BigUint64Array | null = null): Promise<BigUint64Array> {
    const code = typeof op === 'number' ? op : Op[op];
    if (code === undefined) return Promise.reject(new Error(`Evaluation failed: unknown op ${String(op)}`));
    return this.n.evalBatch(this.h, code, bitWidth, left, right, this.batchOf(left, bitWidth), 0);
  }

  /** the (euintN, uintN) overloads: one plaintext right operand per integer; div / rem exist only in this form */
  evalScalarBatch(op: OpName | Op, bitWidth: number, left: BigUint64Array, scalars: BigUint64Array): Promise<BigUint64Array> {
    const code = typeof op === 'number' ? op : Op[op];
    return this.n.evalBatch(this.h, code, bitWidth, left, scalars, this.batchOf(left, bitWidth), 1);
  }

  // ---- the reference's four methods, same names AND arity (packages/wasm/src/index.ts:19-22):
  //        add(ct1, ct2, bootstrapKey, secretKey): string
  // ct1 / ct2 are serialized ciphertexts exactly as the reference passes them (strings), here the base64 of the
  // engine's self-describing container (lux_serialize, include/lux_tfhe.h): the bit width is read from its header and the
  // parameter set is checked against this engine.  `bootstrapKey` is the handle returned by loadKeys() (keys are resident
  // on the GPU, not re-sent per call) and is only checked for identity; `secretKey` is accepted and ignored (the
  // reference's wasm build needs it to re-encrypt, the server-key evaluation here never sees secrets).  The one change
  // that cannot be avoided: the result is a Promise<string>, because the GPU work runs off the event loop.
  add(ct1: string, ct2: string, bootstrapKey?: KeyHandle, _secretKey?: unknown) { return this.op2(Op.add, ct1, ct2, bootstrapKey); }
  sub(ct1: string, ct2: string, bootstrapKey?: KeyHandle, _secretKey?: unknown) { return this.op2(Op.sub, ct1, ct2, bootstrapKey); }
  eq(ct1: string, ct2: string, bootstrapKey?: KeyHandle, _secretKey?: unknown) { return this.op2(Op.eq, ct1, ct2, bootstrapKey); }
  lt(ct1: string, ct2: string, bootstrapKey?: KeyHandle, _secretKey?: unknown) { return this.op2(Op.lt, ct1, ct2, bootstrapKey); }

  private async op2(op: Op, ct1: string, ct2: string, key?: KeyHandle): Promise<string> {
    if (key !== undefined && key !== this.keyHandle) throw new Error('Evaluation failed: bootstrapKey is not the handle returned by loadKeys()');
    const a = decodeContainer(Buffer.from(ct1, 'base64'), this.params);
    const b = decodeContainer(Buffer.from(ct2, 'base64'), this.params);
    if (a.bitWidth !== b.bitWidth || a.count !== b.count) throw new Error('Evaluation failed: operands differ in width or count');
    const out = await this.evalBatch(op, a.bitWidth, a.words, b.words);
    const predicate = op >= Op.eq && op <= Op.ge;
    return Buffer.from(encodeContainer(this.params, predicate ? 2 : a.bitWidth, a.count, out)).toString('base64');
  }

  /** wrap engine-native words ([count][blocks][ctWords]) as the string form add/sub/eq/lt take */
  toCiphertext(bitWidth: number, words: BigUint64Array): string {
    return Buffer.from(encodeContainer(this.params, bitWidth, this.batchOf(words, bitWidth), words)).toString('base64');
  }
  fromCiphertext(ct: string): { bitWidth: number; count: number; words: BigUint64Array } {
    return decodeContainer(Buffer.from(ct, 'base64'), this.params);
  }

  /** cond ? a : b   (FHE.select) — cond is [batch][ctWords] */
  selectBatch(bitWidth: number, cond: BigUint64Array, a: BigUint64Array, b: BigUint64Array): Promise<BigUint64Array> {
    return this.n.selectBatch(this.h, bitWidth, cond, a, b, this.batchOf(a, bitWidth));
  }

  /** WrappingERC20.sol:86-102 on euint64: resolves to { from, to } = new balances */
  async erc20TransferBatch(from: BigUint64Array, to: BigUint64Array, amount: BigUint64Array): Promise<{ from: BigUint64Array; to: BigUint64Array }> {
    const out = await this.n.erc20TransferBatch(this.h, from, to, amount, this.batchOf(from, 64));
    return { from: out.subarray(0, out.length / 2), to: out.subarray(out.length / 2) };
  }

  /** programmable bootstrap of `batch` LWE ciphertexts with per-ciphertext lookup tables */
  pbsBatch(lwe: BigUint64Array, testVectors: BigUint64Array, lutIdx: Uint32Array | null = null): Promise<BigUint64Array> {
    const nLuts = testVectors.length / this.params.N;
    return this.n.pbsBatch(this.h, lwe, testVectors, nLuts, lutIdx, lwe.length / this.ctWords);
  }

  /** boolean gates on bits encoded m * q/4 (BASELINE config 1) */
  gateBatch(gate: Gate, a: BigUint64Array, b: BigUint64Array): Promise<BigUint64Array> {
    return this.n.gateBatch(this.h, gate, a, b, a.length / (this.params.n + 1));
  }

  trivialEncrypt(bitWidth: number, values: BigUint64Array): BigUint64Array {
    return this.n.trivialEncrypt(this.h, bitWidth, values);
  }
}

/** same contract as the reference's `init()` (index.ts:40-83): memoised, resolves to the engine object */
export async function init(params: LuxParams = S2_SHORTINT, device = 0): Promise<LuxFHEB200> {
  if (!instance) instance = new LuxFHEB200(params, device);
  return instance;
}

export function getLuxFHE(): LuxFHEB200 {
  if (!instance) throw new Error('LuxFHE not initialized. Call init() first.');
  return instance;
}

export function isInitialized(): boolean { return instance !== null; }

export default init;
