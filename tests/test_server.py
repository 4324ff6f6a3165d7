"""/evaluate-compatible server (tfhe_b200/server.py): protocol and request coalescing on the CPU box with a plaintext
stand-in for the engine (the wire format is bytes of u64 words, so a "ciphertext" here is just its blocks in the clear);
the GPU test drives the real engine through HTTP."""
import json
import threading
import urllib.error
import urllib.request
from http.server import ThreadingHTTPServer

import numpy as np
import pytest

from tfhe_b200 import server as S

W = 5   # words per stand-in ciphertext


def blocks(bw):
    return bw // 2 if bw in (8, 16, 32, 64) else 0


def enc(v, bw):
    out = np.zeros((blocks(bw), W), np.uint64)
    for i in range(blocks(bw)):
        out[i, 0] = (v >> (2 * i)) & 3
    return out.astype("<u8").tobytes()


def dec(b, bw):
    a = np.frombuffer(b, dtype="<u8").reshape(-1, W)
    return sum(int(a[i, 0]) << (2 * i) for i in range(a.shape[0]))


def plain_eval(op, bw, lhs, rhs):
    """batch evaluator with the semantics of MockFHE.sol on the stand-in encoding"""
    B, nb = lhs.shape[0], lhs.shape[1]
    val = lambda x: [sum(int(x[b, i, 0]) << (2 * i) for i in range(nb)) for b in range(B)]
    a, c = val(lhs), (val(rhs) if rhs is not None else [0] * B)
    f = {"add": lambda x, y: (x + y) % 2**bw, "sub": lambda x, y: (x - y) % 2**bw, "lt": lambda x, y: int(x < y),
         "eq": lambda x, y: int(x == y), "not": lambda x, y: x ^ (2**bw - 1)}
    if op not in f:
        raise S.BadRequest(f"unknown op {op}")
    res = [f[op](x, y) for x, y in zip(a, c)]
    out = np.zeros((B, 1 if op in S.PREDICATES else nb, W), np.uint64)
    for b, v in enumerate(res):
        for i in range(out.shape[1]):
            out[b, i, 0] = (v >> (2 * i)) & 3
    return out


@pytest.fixture()
def srv():
    co = S.EvalCoalescer(plain_eval, blocks, W, window_s=0.05, max_batch=64)
    httpd = ThreadingHTTPServer(("127.0.0.1", 0), S.make_handler(co))
    t = threading.Thread(target=httpd.serve_forever, daemon=True)
    t.start()
    yield f"http://127.0.0.1:{httpd.server_address[1]}", co
    httpd.shutdown()
    co.close()
    httpd.server_close()


def post(url, payload):
    req = urllib.request.Request(url + "/evaluate", data=json.dumps(payload).encode(), headers={"Content-Type": "application/json"})
    with urllib.request.urlopen(req, timeout=30) as r:
        return r.read()


def test_evaluate_protocol(srv):
    url, _ = srv
    # the request shape of client.ts:128-147: byte arrays as JSON number lists, raw bytes back
    out = post(url, {"op": "add", "left": list(enc(200, 8)), "right": list(enc(100, 8)), "bitWidth": 8})
    assert dec(out, 8) == (200 + 100) % 256
    assert dec(post(url, {"op": "lt", "left": list(enc(3, 16)), "right": list(enc(9, 16)), "bitWidth": 16}), 16) == 1
    assert dec(post(url, {"op": "not", "left": list(enc(0x0F, 8)), "bitWidth": 8}), 8) == 0xF0
    with urllib.request.urlopen(url + "/health", timeout=5) as r:
        assert json.loads(r.read())["status"] == "ok"


@pytest.mark.parametrize("payload,code", [
    ({"op": "add", "left": [1, 2, 3], "right": [1, 2, 3], "bitWidth": 8}, 400),     # wrong length
    ({"op": "add", "left": list(enc(1, 8)), "bitWidth": 8}, 400),                   # missing operand
    ({"op": "add", "left": list(enc(1, 8)), "right": list(enc(1, 8)), "bitWidth": 7}, 400),
    ({"op": "frobnicate", "left": list(enc(1, 8)), "right": list(enc(1, 8)), "bitWidth": 8}, 400),
])
def test_evaluate_errors_are_non_2xx_with_text(srv, payload, code):
    url, _ = srv
    with pytest.raises(urllib.error.HTTPError) as ei:
        post(url, payload)
    assert ei.value.code == code and len(ei.value.read()) > 0     # client: Error("Evaluation failed: " + text)


def test_key_and_encryption_routes_stay_on_the_reference_path(srv):
    url, _ = srv
    for path in ("/encrypt", "/decrypt"):
        with pytest.raises(urllib.error.HTTPError) as ei:
            urllib.request.urlopen(urllib.request.Request(url + path, data=b"{}"), timeout=5)
        assert ei.value.code == 501


def test_concurrent_requests_are_coalesced_into_batches(srv):
    url, co = srv
    n = 24
    outs = [None] * n

    def one(i):
        op = "add" if i % 2 == 0 else "sub"
        outs[i] = dec(post(url, {"op": op, "left": list(enc(50 + i, 8)), "right": list(enc(i, 8)), "bitWidth": 8}), 8)

    ts = [threading.Thread(target=one, args=(i,)) for i in range(n)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert outs == [(50 + 2 * i) % 256 if i % 2 == 0 else 50 for i in range(n)]
    assert sum(co.batches) == n and len(co.batches) < n      # fewer evaluator calls than requests
    assert max(co.batches) > 1


@pytest.mark.gpu
def test_server_drives_the_gpu_engine():
    import oracle as O
    import tfhe_b200 as T
    import tfhe_b200.radix as R
    from oracle.radix import join_blocks, split_blocks

    p = T.Params(n=32, N=512, k=1, ell=2, logB=16, ell_ks=8, logB_ks=4, order=T.ORDER_LWEKS_PBS)
    ctx = O.Ctx(O.Params(n=p.n, N=p.N, k=p.k, ell=p.ell, logB=p.logB, ell_ks=p.ell_ks, logB_ks=p.logB_ks, order=p.order))
    ks = O.keygen(ctx, 5, 1e-12, 1e-9)
    O.install_keys(ctx, ks)
    eng = T.Engine(p)
    eng.load_keyset(ks)
    rx = R.Radix(eng)
    co = S.EvalCoalescer(lambda op, bw, l, r: rx.eval(op, bw, l, r), R.blocks, p.ct_len, window_s=0.05, max_batch=64)
    httpd = ThreadingHTTPServer(("127.0.0.1", 0), S.make_handler(co))
    threading.Thread(target=httpd.serve_forever, daemon=True).start()
    url = f"http://127.0.0.1:{httpd.server_address[1]}"
    rng, delta = O.Rng(9), O.delta(32)

    def encrypt(v, bw):
        return np.stack([O.lwe_encrypt(rng, ks.s_in, (d * delta) % O.Q, 1e-9) for d in split_blocks(v, R.blocks(bw))]).astype("<u8").tobytes()

    def decrypt(b):
        a = np.frombuffer(b, dtype="<u8").reshape(-1, p.ct_len)
        return join_blocks([O.decode(O.lwe_decrypt(ks.s_in, c), delta, 32) for c in a])

    try:
        res = {}

        def one(i):
            res[i] = decrypt(post(url, {"op": "add", "left": list(encrypt(17 * i, 8)), "right": list(encrypt(90, 8)), "bitWidth": 8}))

        ts = [threading.Thread(target=one, args=(i,)) for i in range(6)]
        [t.start() for t in ts]
        [t.join() for t in ts]
        assert res == {i: (17 * i + 90) % 256 for i in range(6)}
        assert decrypt(post(url, {"op": "lt", "left": list(encrypt(5, 8)), "right": list(encrypt(9, 8)), "bitWidth": 8})) == 1
        assert len(co.batches) < 7
    finally:
        httpd.shutdown()
        co.close()
        httpd.server_close()


def test_hostile_requests_are_rejected_cheaply(srv):
    """a JSON integer as an operand must not become bytes(int) (that many zero bytes), a boolean is not a
    bit width, an unknown op is refused before it can open a coalescing window, and the body size is capped"""
    url, co = srv
    co2 = S.EvalCoalescer(plain_eval, blocks, W, window_s=0.01, max_batch=4)
    try:
        for bad in (20_000_000_000, "AAAA", {"0": 1}, [256] * (4 * W * 8), [-1] * (4 * W * 8), [0.5] * (4 * W * 8), [1, 2, 3]):
            with pytest.raises(S.BadRequest):
                co2.parse("add", 8, bad, list(enc(1, 8)))
        with pytest.raises(S.BadRequest):
            co2.parse("add", True, list(enc(1, 8)), list(enc(1, 8)))
        with pytest.raises(S.BadRequest):
            co2.parse("frobnicate", 8, list(enc(1, 8)), list(enc(1, 8)))
        lhs, rhs = co2.parse("add", 8, list(enc(3, 8)), list(enc(4, 8)))
        assert lhs.shape == (4, W) and int(lhs[0, 0]) == 3
    finally:
        co2.close()
    n_before = len(co.batches)
    req = urllib.request.Request(url + "/evaluate", data=json.dumps({"op": "add", "bitWidth": 8, "left": 20_000_000_000,
                                                                       "right": list(enc(1, 8))}).encode(), method="POST")
    with pytest.raises(urllib.error.HTTPError) as ei:
        urllib.request.urlopen(req, timeout=10)
    assert ei.value.code == 400
    big = urllib.request.Request(url + "/evaluate", data=b"x" * (2 * S.MAX_BLOCKS * W * 8 * 4 + 5000), method="POST")
    with pytest.raises((urllib.error.HTTPError, urllib.error.URLError, ConnectionError)) as ei:
        urllib.request.urlopen(big, timeout=10)
    if isinstance(ei.value, urllib.error.HTTPError):
        assert ei.value.code == 413
    assert len(co.batches) == n_before          # nothing reached the evaluator


def test_serialized_containers_in_and_out():
    """operands may be self-describing containers (lux_serialize): the server checks parameter set, kind and width and
    answers with a container; a container of another width or parameter set is a 400, not a garbage result"""
    import tfhe_b200 as T
    from tfhe_b200 import engine as E

    P = T.Params(n=3, N=512, k=1, ell=2, logB=16, ell_ks=8, logB_ks=4, order=T.ORDER_PBS_GLWEKS)   # ct words = n + 1 = 4
    w = P.ct_len

    def ev(op, bw, lhs, rhs):
        out = lhs.copy()
        out[:, :, 0] = (lhs[:, :, 0] + rhs[:, :, 0]) % 4
        return out[:, :1] if op in S.PREDICATES else out

    co = S.EvalCoalescer(ev, blocks, w, window_s=0.01, max_batch=8, params=P)
    try:
        a = np.zeros((4, w), np.uint64)
        a[:, 0] = [1, 2, 3, 0]
        b = np.ones((4, w), np.uint64)
        ca, cb = E.serialize(P, E.BLOB_RADIX, a, 8, 1), E.serialize(P, E.BLOB_RADIX, b, 8, 1)
        out = co.submit("add", 8, list(ca), list(cb)).result(timeout=10)
        words, info = E.deserialize(out, P, E.BLOB_RADIX)
        assert info["aux"] == 8 and words.reshape(4, w)[:, 0].tolist() == [2, 3, 0, 1]
        raw = co.submit("add", 8, list(a.tobytes()), list(b.tobytes())).result(timeout=10)       # bare words still work
        assert np.frombuffer(raw, "<u8").reshape(4, w)[:, 0].tolist() == [2, 3, 0, 1]
        pred = co.submit("lt", 8, list(ca), list(cb)).result(timeout=10)
        assert E.deserialize(pred, P, E.BLOB_RADIX)[1]["aux"] == 2
        with pytest.raises(S.BadRequest):
            co.submit("add", 16, list(ca), list(cb))                      # width in the request != width in the container
        other = T.Params(**{**P.__dict__, "logB": 15})
        with pytest.raises(S.BadRequest):
            co.submit("add", 8, list(E.serialize(other, E.BLOB_RADIX, a, 8, 1)), list(cb))
        bad = bytearray(ca)
        bad[200] ^= 1
        with pytest.raises(S.BadRequest):
            co.submit("add", 8, list(bad), list(cb))
    finally:
        co.close()
