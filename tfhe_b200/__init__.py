"""tfhe_b200 — B200-native TFHE programmable-bootstrap engine (sm_100a CUDA behind a C ABI).

The product path is libluxtfhe_b200.so (include/lux_tfhe.h).  This package is the thin host-side
mirror of the reference's `LuxFHE` operator interface (packages/wasm/src/index.ts:14-23) over that
ABI via ctypes.  There is no CPU fallback: importing works without a GPU (so the ABI can be
inspected), but creating an engine raises unless a B200 is present and the library is built.
"""
from .engine import (Engine, MultiEngine, LuxError, Params, lib, lib_path, make_testv, delta, ORDER_PBS_GLWEKS, ORDER_LWEKS_PBS,
                     ORDER_PBS_LWEKS, ARITH_GOLDILOCKS_NTT, ARITH_TORUS64_FFT, DOMAIN_STANDARD, DOMAIN_NTT, GATES, Q,
                     S1_BOOLEAN, S2_SHORTINT, S2_TORUS, TEST_SMALL_O0, TEST_SMALL_O1, TEST_SMALL_TORUS,
                     TEST_SMALL_TORUS_O2)

__all__ = ["Engine", "MultiEngine", "LuxError", "Params", "lib", "lib_path", "make_testv", "delta", "ORDER_PBS_GLWEKS",
           "ORDER_LWEKS_PBS", "ORDER_PBS_LWEKS", "ARITH_GOLDILOCKS_NTT", "ARITH_TORUS64_FFT", "S2_TORUS",
           "TEST_SMALL_TORUS", "TEST_SMALL_TORUS_O2", "DOMAIN_STANDARD", "DOMAIN_NTT", "GATES", "Q", "S1_BOOLEAN", "S2_SHORTINT",
           "TEST_SMALL_O0", "TEST_SMALL_O1"]
