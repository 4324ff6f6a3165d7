"""The engine-native wire container (include/lux_tfhe.h "serialized keys and ciphertexts", tfhe_b200/csrc/serial.cu):
host-only, so everything but the key-loading test runs on the CPU box.  Replaces the bare little-endian word arrays of
round 1 on the by-value byte contract of packages/wasm/src/node.ts:65-96 / js/v2-sdk/src/luxd/client.ts:128-147."""
import struct
import zlib

import numpy as np
import pytest

import tfhe_b200 as T
from tfhe_b200 import engine as E

P = T.TEST_SMALL_TORUS


def rnd(n, seed=1):
    rs = np.random.RandomState(seed)
    return rs.randint(0, 2**63, n, dtype=np.int64).astype(np.uint64) * np.uint64(2) + np.uint64(1)


def test_round_trip_every_kind():
    cases = [(E.BLOB_LWE, 0, 5, 5 * P.ct_len), (E.BLOB_RADIX, 64, 2, 2 * 32 * P.ct_len), (E.BLOB_RADIX, 8, 1, 4 * P.ct_len),
             (E.BLOB_BSK, 0, 1, P.n * P.ggsw_len), (E.BLOB_LWE_KSK, 0, 1, P.lwe_ksk_len), (E.BLOB_GLWE_KSK, 1, 1, P.ggsw_len)]
    for kind, aux, count, words in cases:
        w = rnd(words, kind)
        blob = E.serialize(P, kind, w, aux, count)
        assert len(blob) == 80 + 8 * words
        out, info = E.deserialize(blob, P, kind)
        assert np.array_equal(out, w)
        assert (info["kind"], info["aux"], info["count"], info["version"]) == (kind, aux, count, 1)
        assert info["params"] == P


def test_header_layout_and_standard_crc():
    """the documented layout, and the checksums are the IEEE 802.3 CRC-32 (zlib's), so any host can produce containers"""
    w = rnd(3 * P.ct_len)
    blob = E.serialize(P, E.BLOB_LWE, w, 0, 3)
    assert blob[:8] == b"LUXFHE\r\n"
    version, kind = struct.unpack_from("<II", blob, 8)
    params = struct.unpack_from("<9I", blob, 16)
    aux, count, nwords, pcrc, hcrc = struct.unpack_from("<IQQII", blob, 52)
    assert (version, kind, aux, count, nwords) == (1, E.BLOB_LWE, 0, 3, 3 * P.ct_len)
    assert params == (P.n, P.N, P.k, P.ell, P.logB, P.ell_ks, P.logB_ks, P.order, P.arith)
    assert pcrc == zlib.crc32(blob[80:]) and hcrc == zlib.crc32(blob[:76])
    assert blob[80:] == w.astype("<u8").tobytes()


def test_damage_and_mismatch_are_told_apart():
    w = rnd(2 * P.ct_len)
    blob = E.serialize(P, E.BLOB_LWE, w, 0, 2)
    for pos in (0, 9, 20, 60, 75, 90, len(blob) - 1):          # magic, version, params, count, header crc, payload
        bad = bytearray(blob)
        bad[pos] ^= 0x40
        with pytest.raises(T.LuxError) as ei:
            E.deserialize(bytes(bad), P)
        assert ei.value.code == E.ERR_FORMAT, pos
    for cut in (0, 79, len(blob) - 8, len(blob) + 8):
        with pytest.raises(T.LuxError) as ei:
            E.deserialize((blob + b"\0" * 8)[:cut], P)
        assert ei.value.code == E.ERR_FORMAT
    with pytest.raises(T.LuxError) as ei:                       # an S2-shaped engine is handed a small-set ciphertext
        E.deserialize(blob, T.S2_TORUS)
    assert ei.value.code == E.ERR_MISMATCH
    other = T.Params(**{**P.__dict__, "arith": T.ARITH_GOLDILOCKS_NTT})
    with pytest.raises(T.LuxError) as ei:                       # same shape, other arithmetic
        E.deserialize(blob, other)
    assert ei.value.code == E.ERR_MISMATCH
    with pytest.raises(T.LuxError) as ei:
        E.deserialize(blob, P, E.BLOB_BSK)
    assert ei.value.code == E.ERR_MISMATCH
    with pytest.raises(T.LuxError):
        E.serialize(P, E.BLOB_RADIX, w, 7, 1)                   # odd width
    with pytest.raises(T.LuxError):
        E.serialize(P, E.BLOB_LWE, w[:-1], 0, 2)                # wrong word count


@pytest.mark.gpu
def test_engine_loads_key_containers_and_refuses_foreign_ones():
    from oracle import torus as TO

    ctx = TO.Ctx(TO.Params(n=P.n, N=P.N, k=P.k, ell=P.ell, logB=P.logB, ell_ks=P.ell_ks, logB_ks=P.logB_ks, order=P.order))
    ks = TO.keygen(ctx, 5, TO.S2_GLWE_NOISE, TO.S2_LWE_NOISE)
    TO.install_keys(ctx, ks)
    a, b = T.Engine(P), T.Engine(P)
    a.load_keyset(ks)
    b.load_key_blob(E.serialize(P, E.BLOB_BSK, ks.bsk, T.DOMAIN_STANDARD))
    b.load_key_blob(E.serialize(P, E.BLOB_LWE_KSK, ks.lwe_ksk))
    dl = T.delta(32, P.arith)
    tv = T.make_testv(P.N, 16, dl, list(range(16)), P.arith)
    cts = TO.lwe_encrypt_batch(3, ks.s_in, np.arange(6, dtype=np.uint64) * np.uint64(dl), TO.S2_GLWE_NOISE)
    wire = E.serialize(P, E.BLOB_LWE, cts, 0, 6)                 # ciphertexts travel as a container too
    got_cts, _ = E.deserialize(wire, P, E.BLOB_LWE)
    got = b.pbs_batch(got_cts.reshape(6, P.ct_len), tv)
    assert np.array_equal(got, a.pbs_batch(cts, tv))
    want, _ = ctx.pbs_batch(cts, tv)
    assert np.array_equal(got, want)
    foreign = T.Params(**{**P.__dict__, "n": P.n + 1})
    blob = E.serialize(foreign, E.BLOB_BSK, rnd(foreign.n * foreign.ggsw_len), T.DOMAIN_STANDARD)
    with pytest.raises(T.LuxError) as ei:
        b.load_key_blob(blob)
    assert ei.value.code == E.ERR_MISMATCH and "another parameter set" in str(ei.value)
    with pytest.raises(T.LuxError) as ei:
        b.load_key_blob(wire)                                   # a ciphertext is not a key
    assert ei.value.code == E.ERR_MISMATCH
