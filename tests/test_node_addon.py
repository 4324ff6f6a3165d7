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
