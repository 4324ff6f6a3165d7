"""HTTP twin of the evaluation path: a `/evaluate`-compatible server over the GPU engine.

Reference protocol (js/v2-sdk/src/luxd/client.ts:128-147): `POST /evaluate` with JSON
`{op, left: number[], right?: number[], bitWidth}` -> raw ciphertext bytes, non-2xx + text on error (the client turns it
into `Error("Evaluation failed: " + text)`); `GET /health`; `GET /publickey` (client.ts:95-101).  The reference server
(Go luxd) is not in the tree; this one keeps its request/response shapes for the evaluation calls and nothing else:
key generation, /encrypt and /decrypt stay on the reference path (501 here), /publickey is a pass-through of a file the
operator provides.

Ciphertext bytes are the engine-native layout (little-endian u64 words, [blocks][ct_words]); the byte format of the Go
serializers is opaque in the tree, a converter is row N3.

What the GPU needs that the reference's one-request-one-op server does not: **request coalescing**.  A single euint8
add is 7 bootstraps on a device that runs 296 at a time for the same latency, so concurrent requests with the same
(op, bitWidth) are gathered for up to `window_s` (or until `max_batch`) and evaluated as ONE `lux_eval` batch.
"""
from __future__ import annotations

import json
import queue
import threading
import time
from concurrent.futures import Future
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer

import numpy as np

# op names of client.ts:35 plus the rest of lux_radix.h's lux_op (same spelling as tfhe_b200.radix.OPS)
UNARY = {"neg", "not"}
PREDICATES = {"eq", "ne", "lt", "le", "gt", "ge"}
OPS = {"add", "sub", "mul", "neg", "eq", "ne", "lt", "le", "gt", "ge", "min", "max", "and", "or", "xor", "not", "shl", "shr",
       "rotl", "rotr"}       # encrypted-operand ops of lux_radix.h (div / rem take a plaintext divisor: not an /evaluate op)
MAX_BLOCKS = 128             # euint256: the largest ciphertext an /evaluate request can carry


class BadRequest(ValueError):
    pass


class EvalCoalescer:
    """Gathers concurrent requests per (op, bitWidth) and evaluates each group as one batch.

    `evaluate(op, bit_width, lhs, rhs_or_None) -> out` is the batch evaluator ([batch][blocks][w] u64 arrays);
    the GPU server passes `Radix.eval`, the CPU tests pass a plaintext stand-in."""

    def __init__(self, evaluate, blocks, ct_words, window_s=0.002, max_batch=4096, params=None):
        self._evaluate, self._blocks, self.w = evaluate, blocks, ct_words
        self.params = params       # engine parameter set: operands may then be serialized containers (lux_serialize)
        self.window_s, self.max_batch = window_s, max_batch
        self._q: queue.Queue = queue.Queue()
        self.batches = []          # sizes of the batches evaluated so far (introspection for tests / metrics)
        self._worker = threading.Thread(target=self._run, daemon=True)
        self._worker.start()

    def close(self):
        self._q.put(None)
        self._worker.join(timeout=5)

    # ---- request side -----------------------------------------------------------------------------------
    def parse(self, op, bit_width, left, right):
        """validate one request, bytes -> u64 words; raises BadRequest with the text the client shows"""
        return self._parse(op, bit_width, left, right)[:2]

    def _parse(self, op, bit_width, left, right):
        # bool is an int subclass: True must not pass for bitWidth 1
        nb = self._blocks(bit_width) if isinstance(bit_width, int) and not isinstance(bit_width, bool) else 0
        if not nb:
            raise BadRequest(f"unsupported bitWidth {bit_width!r}")
        if not isinstance(op, str) or op not in OPS:
            raise BadRequest(f"unknown op {op!r}")     # rejected here, before the request can open a coalescing window
        want = nb * self.w * 8

        def words(name, v):
            # only a JSON array of exactly `want` byte values (or want + 80: a serialized container) is accepted:
            # bytes(int) would ALLOCATE that many bytes (a one-request memory DoS), a str or dict is not a ciphertext
            if isinstance(v, (bytes, bytearray)):
                b = bytes(v)
            elif isinstance(v, (list, tuple)):
                if len(v) != want and not (self.params is not None and len(v) == want + 80):
                    raise BadRequest(f"{name} has {len(v)} bytes, a euint{bit_width} ciphertext has {want}")
                try:
                    arr = np.asarray(v)
                    if arr.dtype.kind not in "iu" or arr.ndim != 1 or arr.min(initial=0) < 0 or arr.max(initial=0) > 255:
                        raise ValueError
                    b = arr.astype(np.uint8).tobytes()
                except (TypeError, ValueError, OverflowError):
                    raise BadRequest(f"{name} must be an array of bytes")
            else:
                raise BadRequest(f"{name} must be an array of bytes")
            if self.params is not None and len(b) == want + 80 and b[:8] == b"LUXFHE\r\n":
                # self-describing container: parameter set, arithmetic, kind and width are checked, not assumed
                from . import engine as E
                try:
                    wds, info = E.deserialize(b, self.params, E.BLOB_RADIX)
                except E.LuxError as ex:
                    raise BadRequest(f"{name}: {ex}")
                if info["aux"] != bit_width or info["count"] != 1:
                    raise BadRequest(f"{name} is a container of {info['count']} euint{info['aux']}, the request says one euint{bit_width}")
                containers.append(name)
                return wds.reshape(nb, self.w)
            if len(b) != want:
                raise BadRequest(f"{name} has {len(b)} bytes, a euint{bit_width} ciphertext has {want}")
            return np.frombuffer(b, dtype="<u8").reshape(nb, self.w)

        containers = []

        lhs = words("left", left)
        if op in UNARY:
            if right is not None:
                raise BadRequest(f"op {op} takes one operand")
            rhs = None
        else:
            if right is None:
                raise BadRequest(f"op {op} needs a right operand")
            rhs = words("right", right)
        return lhs, rhs, bool(containers)

    def submit(self, op, bit_width, left, right=None) -> Future:
        lhs, rhs, wrap = self._parse(op, bit_width, left, right)
        inner: Future = Future()
        self._q.put((op, int(bit_width), lhs, rhs, inner))
        if not wrap:
            return inner
        # operands came as containers: the answer goes back as one (a predicate is a euint2, i.e. one block)
        outer: Future = Future()

        def done(f):
            try:
                from . import engine as E
                raw = np.frombuffer(f.result(), dtype="<u8")
                outer.set_result(E.serialize(self.params, E.BLOB_RADIX, raw, 2 if op in PREDICATES else int(bit_width), 1))
            except Exception as ex:  # noqa: BLE001
                outer.set_exception(ex)

        inner.add_done_callback(done)
        return outer

    # ---- worker -----------------------------------------------------------------------------------------
    def _run(self):
        while True:
            item = self._q.get()                     # wait for the request that opens a window
            if item is None:
                return
            group = [item]
            deadline = time.monotonic() + self.window_s
            closing = False
            while len(group) < self.max_batch:
                rem = deadline - time.monotonic()
                if rem <= 0:
                    break
                try:
                    nxt = self._q.get(timeout=rem)
                except queue.Empty:
                    break
                if nxt is None:
                    closing = True
                    break
                group.append(nxt)
            by_key = {}                              # (op, bit_width) -> requests in arrival order
            for op, bw, lhs, rhs, fut in group:
                by_key.setdefault((op, bw), []).append((lhs, rhs, fut))
            for (op, bw), reqs in by_key.items():
                self._flush(op, bw, reqs)
            if closing:
                return

    def _flush(self, op, bw, reqs):
        for i in range(0, len(reqs), self.max_batch):
            chunk = reqs[i:i + self.max_batch]
            try:
                lhs = np.stack([r[0] for r in chunk])
                rhs = None if op in UNARY else np.stack([r[1] for r in chunk])
                out = np.asarray(self._evaluate(op, bw, lhs, rhs))
                self.batches.append(len(chunk))
                for k, (_, _, fut) in enumerate(chunk):
                    fut.set_result(np.ascontiguousarray(out[k]).astype("<u8").tobytes())
            except Exception as ex:   # engine errors reach every request of the batch as its own failure
                for _, _, fut in chunk:
                    if not fut.done():
                        fut.set_exception(ex)


def make_handler(coalescer: EvalCoalescer, public_key_path=None, version="tfhe_b200"):
    class Handler(BaseHTTPRequestHandler):
        protocol_version = "HTTP/1.1"

        def log_message(self, *a):   # quiet
            pass

        def _send(self, code, body: bytes, ctype):
            self.send_response(code)
            self.send_header("Content-Type", ctype)
            self.send_header("Content-Length", str(len(body)))
            self.end_headers()
            self.wfile.write(body)

        def do_GET(self):
            if self.path == "/health":
                self._send(200, json.dumps({"status": "ok", "engine": version, "batches": len(coalescer.batches)}).encode(),
                           "application/json")
            elif self.path == "/publickey":
                if public_key_path is None:
                    self._send(501, b"public key is served by the reference key service", "text/plain")
                else:
                    with open(public_key_path, "rb") as f:
                        self._send(200, f.read(), "application/octet-stream")
            else:
                self._send(404, b"not found", "text/plain")

        def do_POST(self):
            try:
                n = int(self.headers.get("Content-Length", "0"))
            except ValueError:
                n = -1
            # a request is two euint256 operands as JSON byte arrays (<= 4 characters per byte) plus a little framing
            limit = 2 * MAX_BLOCKS * coalescer.w * 8 * 4 + 4096
            if n < 0 or n > limit:
                self._send(413, f"request body of {n} bytes exceeds the {limit}-byte limit".encode(), "text/plain")
                self.close_connection = True
                return
            body = self.rfile.read(n)
            if self.path in ("/encrypt", "/decrypt", "/threshold/decrypt", "/verify"):
                self._send(501, b"key generation, encryption and decryption stay on the reference path", "text/plain")
                return
            if self.path != "/evaluate":
                self._send(404, b"not found", "text/plain")
                return
            try:
                req = json.loads(body)
                fut = coalescer.submit(req.get("op"), req.get("bitWidth"), req.get("left"), req.get("right"))
                self._send(200, fut.result(timeout=600), "application/octet-stream")
            except BadRequest as ex:
                self._send(400, str(ex).encode(), "text/plain")
            except json.JSONDecodeError:
                self._send(400, b"body is not JSON", "text/plain")
            except Exception as ex:   # LuxError etc.: the client prints "Evaluation failed: <text>"
                self._send(500, str(ex).encode(), "text/plain")

    return Handler


def serve(engine, host="127.0.0.1", port=8448, window_s=0.002, max_batch=4096, public_key_path=None):
    """blocking: serve /evaluate for one engine (one GPU, keys already loaded)"""
    from . import radix as R

    rx = R.Radix(engine)

    def evaluate(op, bw, lhs, rhs):
        if op not in R.OPS:
            raise BadRequest(f"unknown op {op}")
        return rx.eval(op, bw, lhs, rhs)

    co = EvalCoalescer(evaluate, R.blocks, engine.p.ct_len, window_s, max_batch, params=engine.p)
    srv = ThreadingHTTPServer((host, port), make_handler(co, public_key_path))
    try:
        srv.serve_forever()
    finally:
        co.close()
        srv.server_close()
