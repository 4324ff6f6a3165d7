"""N-API addon (bindings/node): there is no `node` in this image, so the addon cannot be loaded; what CAN be checked
on the CPU box is that it compiles against the N-API declarations, exports the registration symbol node looks up, and
that every lux_* function it calls is exported by the C-ABI library it links."""
import re
import shutil
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
SRC = ROOT / "bindings" / "node" / "luxfhe_b200_napi.c"
LIB = ROOT / "tfhe_b200" / "libluxtfhe_b200.so"


@pytest.fixture(scope="module")
def addon(tmp_path_factory):
    if not LIB.exists():
        pytest.skip("libluxtfhe_b200.so not built")
    out = tmp_path_factory.mktemp("node") / "luxfhe_b200.node"
    cmd = ["gcc", "-std=c11", "-Wall", "-Wextra", "-Werror", "-fPIC", "-shared", "-o", str(out), str(SRC),
           f"-L{LIB.parent}", f"-l:{LIB.name}"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return out


def _dyn_symbols(path):
    out = subprocess.run(["nm", "-D", str(path)], capture_output=True, text=True, check=True).stdout
    defined = {l.split()[-1] for l in out.splitlines() if " T " in l}
    undefined = {l.split()[-1] for l in out.splitlines() if " U " in l}
    return defined, undefined


def test_addon_exports_registration_symbol(addon):
    defined, _ = _dyn_symbols(addon)
    assert "napi_register_module_v1" in defined


def test_addon_links_only_exported_abi_functions(addon):
    _, undefined = _dyn_symbols(addon)
    lib_defined, _ = _dyn_symbols(LIB)
    lux = {s for s in undefined if s.startswith("lux_")}
    assert {"lux_engine_create", "lux_load_bsk", "lux_pbs_batch", "lux_eval", "lux_select", "lux_erc20_transfer"} <= lux
    assert lux <= lib_defined, lux - lib_defined
    # everything else it needs from the host process is N-API (resolved by node) or libc
    napi = {s for s in undefined if s.startswith("napi_")}
    decl = (ROOT / "bindings" / "node" / "node_api_decl.h").read_text()
    assert napi and all(re.search(rf"\b{s}\(", decl) for s in napi)


def test_ts_wrapper_names_every_op_of_the_abi():
    """index.ts's Op enum is positional: its order must be lux_radix.h's lux_op order"""
    hdr = (ROOT / "include" / "lux_radix.h").read_text()
    ops = [(m.group(1).lower(), int(m.group(2))) for m in re.finditer(r"LUX_OP_([A-Z]+) = (\d+)", hdr)]
    ts = (ROOT / "bindings" / "node" / "index.ts").read_text()
    body = re.search(r"export enum Op \{(.*?)\}", ts, re.S).group(1)
    names = [t.split("=")[0].strip() for t in body.replace("\n", " ").split(",") if t.strip()]
    assert names == [n for n, _ in sorted(ops, key=lambda x: x[1])]
    assert shutil.which("gcc")


# ---------------------------------------------------------------------------------------------------------------------------
# The addon EXECUTES: tests/napi_host/fake_napi.c is a minimal N-API host (value model + async work on real pthreads, the way
# libuv runs it).  Linked with the addon into one shared object and driven from here with ctypes, on the GPU box.
import ctypes as C

import numpy as np

HOST = ROOT / "tests" / "napi_host" / "fake_napi.c"
NAPI_BIGUINT64, NAPI_UINT32 = 10, 6       # napi_typedarray_type (node_api_decl.h)


@pytest.fixture(scope="module")
def host(tmp_path_factory):
    if not LIB.exists():
        pytest.skip("libluxtfhe_b200.so not built")
    out = tmp_path_factory.mktemp("napi_host") / "addon_in_fake_host.so"
    cmd = ["gcc", "-std=c11", "-Wall", "-Wextra", "-Werror", "-fPIC", "-shared", "-o", str(out), str(SRC), str(HOST),
           f"-L{LIB.parent}", f"-l:{LIB.name}", "-pthread", f"-Wl,-rpath,{LIB.parent}"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    L = C.CDLL(str(out))
    vp = C.c_void_p
    for name, res, args in [("fakehost_new", vp, []), ("fakehost_number", vp, [C.c_double]), ("fakehost_null", vp, []),
                            ("fakehost_object", vp, []), ("fakehost_set", None, [vp, C.c_char_p, vp]),
                            ("fakehost_typedarray", vp, [C.c_int, vp, C.c_size_t]),
                            ("fakehost_call", vp, [vp, C.c_char_p, C.c_size_t, C.POINTER(vp)]),
                            ("fakehost_exception", C.c_char_p, [vp]), ("fakehost_drain", C.c_int, [vp]),
                            ("fakehost_max_concurrent", C.c_int, [vp]), ("fakehost_promise_state", C.c_int, [vp]),
                            ("fakehost_promise_value", vp, [vp]), ("fakehost_string", C.c_char_p, [vp]),
                            ("fakehost_number_value", C.c_double, [vp]),
                            ("fakehost_data", vp, [vp, C.POINTER(C.c_size_t)]), ("fakehost_finalize", None, [vp, vp])]:
        f = getattr(L, name)
        f.restype, f.argtypes = res, args
    return L


class JS:
    """the few lines of `JavaScript` the test needs"""

    def __init__(self, L):
        self.L, self.env, self.keep = L, L.fakehost_new(), []

    def num(self, x):
        return self.L.fakehost_number(float(x))

    def u64(self, a):
        a = np.ascontiguousarray(a, np.uint64)
        self.keep.append(a)
        return self.L.fakehost_typedarray(NAPI_BIGUINT64, a.ctypes.data, a.size)

    def u32(self, a):
        a = np.ascontiguousarray(a, np.uint32)
        self.keep.append(a)
        return self.L.fakehost_typedarray(NAPI_UINT32, a.ctypes.data, a.size)

    def obj(self, **kw):
        o = self.L.fakehost_object()
        for k, v in kw.items():
            self.L.fakehost_set(o, k.encode(), self.num(v))
        return o

    def call(self, name, *args):
        argv = (C.c_void_p * len(args))(*args)
        r = self.L.fakehost_call(self.env, name.encode(), len(args), argv)
        exc = self.L.fakehost_exception(self.env).decode()
        if exc:
            raise RuntimeError(exc)
        return r

    def settle(self, promise):
        """-> numpy words of a resolved BigUint64Array promise, or raises with the rejection text"""
        st = self.L.fakehost_promise_state(promise)
        v = self.L.fakehost_promise_value(promise)
        if st == 2:
            raise RuntimeError(self.L.fakehost_string(v).decode())
        assert st == 1, "promise still pending"
        n = C.c_size_t()
        p = self.L.fakehost_data(v, C.byref(n))
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint64)), (n.value,)).copy()


@pytest.mark.gpu
def test_addon_runs_concurrent_jobs_on_one_handle(host):
    """two un-awaited evalBatch Promises and a pbsBatch Promise on ONE engine handle, executed by three worker threads at
    once (this used to race on the engine's scratch buffers); results equal the serial library calls,
    which tests/test_gpu_torus*.py pin to the oracle"""
    import tfhe_b200 as T
    import tfhe_b200.radix as R
    from oracle import torus as TO
    from oracle.radix import split_blocks

    p = T.TEST_SMALL_TORUS
    ctx = TO.Ctx(TO.Params(n=p.n, N=p.N, k=p.k, ell=p.ell, logB=p.logB, ell_ks=p.ell_ks, logB_ks=p.logB_ks, order=p.order))
    ks = TO.keygen(ctx, 61)
    TO.install_keys(ctx, ks)
    ref = T.Engine(p)
    ref.load_keyset(ks)
    rx = R.Radix(ref)
    dl = T.delta(32, p.arith)

    js = JS(host)
    assert b"tfhe_b200" in host.fakehost_string(js.call("version"))
    params = js.obj(n=p.n, N=p.N, k=p.k, ell=p.ell, logB=p.logB, ellKs=p.ell_ks, logBKs=p.logB_ks, order=p.order, arith=p.arith)
    h = js.call("createEngine", params)
    assert int(host.fakehost_number_value(js.call("ctWords", h))) == p.ct_len
    js.call("loadKey", h, js.num(0), js.u64(ks.bsk), js.num(0))
    js.call("loadKey", h, js.num(2), js.u64(ks.lwe_ksk))

    def enc(vals, width, seed):
        nb = R.blocks(width)
        dig = np.array([split_blocks(v, nb) for v in vals], dtype=np.uint64).reshape(-1)
        return TO.lwe_encrypt_batch(seed, ks.s_in, dig * np.uint64(dl)).reshape(len(vals), nb, p.ct_len)

    a, b = enc([71, 255, 0, 200, 13, 99], 8, 1), enc([66, 1, 0, 57, 250, 98], 8, 2)
    c, d = enc([40000, 7], 16, 3), enc([25536, 9], 16, 4)
    tv = T.make_testv(p.N, 16, dl, [(3 * x + 1) % 16 for x in range(16)], p.arith)
    cts = TO.lwe_encrypt_batch(5, ks.s_in, np.arange(9, dtype=np.uint64) % np.uint64(16) * np.uint64(dl))
    jobs = [js.call("evalBatch", h, js.num(R.OPS["add"]), js.num(8), js.u64(a), js.u64(b), js.num(6)),
            js.call("evalBatch", h, js.num(R.OPS["lt"]), js.num(16), js.u64(c), js.u64(d), js.num(2)),
            js.call("pbsBatch", h, js.u64(cts), js.u64(tv), js.num(1), host.fakehost_null(), js.num(9)),
            js.call("evalBatch", h, js.num(R.OPS["mul"]), js.num(8), js.u64(a), js.u64(b), js.num(6))]
    assert all(host.fakehost_promise_state(j) == 0 for j in jobs)            # nothing awaited yet
    assert host.fakehost_drain(js.env) == 4
    assert host.fakehost_max_concurrent(js.env) >= 2, "the fake host did not run the jobs concurrently"
    assert np.array_equal(js.settle(jobs[0]).reshape(6, 4, p.ct_len), rx.eval("add", 8, a, b))
    assert np.array_equal(js.settle(jobs[1]).reshape(2, 1, p.ct_len), rx.eval("lt", 16, c, d))
    assert np.array_equal(js.settle(jobs[2]).reshape(9, p.ct_len), ref.pbs_batch(cts, tv))
    assert np.array_equal(js.settle(jobs[3]).reshape(6, 4, p.ct_len), rx.eval("mul", 8, a, b))
    # errors: thrown synchronously for malformed arguments, rejected for failures inside the library
    with pytest.raises(RuntimeError, match="lhs must hold"):
        js.call("evalBatch", h, js.num(R.OPS["add"]), js.num(8), js.u64(a[:3]), js.u64(b), js.num(6))
    bare = js.call("createEngine", params)
    rej = js.call("evalBatch", bare, js.num(R.OPS["add"]), js.num(8), js.u64(a), js.u64(b), js.num(6))
    host.fakehost_drain(js.env)
    with pytest.raises(RuntimeError, match="lux error"):
        js.settle(rej)
    # one handle over every visible GPU (params.gpus): keys replicated and the batch sliced inside the library
    mh = js.call("createEngine", js.obj(n=p.n, N=p.N, k=p.k, ell=p.ell, logB=p.logB, ellKs=p.ell_ks, logBKs=p.logB_ks,
                                        order=p.order, arith=p.arith, gpus=0xFFFFFFFF))
    js.call("loadKey", mh, js.num(0), js.u64(ks.bsk), js.num(0))
    js.call("loadKey", mh, js.num(2), js.u64(ks.lwe_ksk))
    mj = js.call("evalBatch", mh, js.num(R.OPS["add"]), js.num(8), js.u64(a), js.u64(b), js.num(6))
    host.fakehost_drain(js.env)
    assert np.array_equal(js.settle(mj).reshape(6, 4, p.ct_len), rx.eval("add", 8, a, b))
    for handle in (h, bare, mh):
        host.fakehost_finalize(js.env, handle)
