"""CPU: the C-ABI library loads and exports every function include/fewtensorf.h declares, and the
ctypes mirrors have the sizes the C compiler gives the header's structs (no compute without a GPU)."""
import ctypes as C
import os
import re
import subprocess
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "fewtensorf.h")


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ftrf_[a-z0-9_]+)\s*\(", src)))


def test_every_declared_symbol_is_exported():
    from fewtensorf_b200 import _abi
    names = declared_functions()
    assert len(names) >= 15
    for n in names:
        assert hasattr(_abi.lib, n), n
        assert n in _abi.EXPORTS, f"{n} missing from _abi.EXPORTS"


def test_struct_sizes_match_the_c_compiler():
    from fewtensorf_b200 import _abi
    prog = '#include <stdio.h>\n#include "fewtensorf.h"\nint main(){printf("%zu %zu %zu %d\\n", sizeof(FtrfFactorSet), sizeof(FtrfField), sizeof(FtrfCall), FTRF_WS_COUNT);return 0;}\n'
    with tempfile.TemporaryDirectory() as d:
        open(os.path.join(d, "t.c"), "w").write(prog)
        subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), os.path.join(d, "t.c"), "-o", os.path.join(d, "t")], check=True)
        out = subprocess.run([os.path.join(d, "t")], capture_output=True, text=True, check=True).stdout.split()
    assert int(out[0]) == C.sizeof(_abi.FtrfFactorSet)
    assert int(out[1]) == C.sizeof(_abi.FtrfField)
    assert int(out[2]) == C.sizeof(_abi.FtrfCall)
    assert int(out[3]) == _abi.WS_COUNT


def test_error_convention_without_gpu():
    from fewtensorf_b200 import _abi
    assert _abi.lib.ftrf_abi_version() == 2
    rc = _abi.lib.ftrf_mse_loss(None, None, 0, None, None, None)
    assert rc == -1 and b"mse_loss" in _abi.lib.ftrf_last_error()
    assert _abi.lib.ftrf_workspace_layout(None, None, None) == 0


def test_test_build_is_a_superset_and_the_product_has_no_debug_switches():
    """libfewtensorf_b200_test.so (cross-check back ends, probes, unit-test entries) exports everything the product
    does; the product exports none of the switches."""
    from fewtensorf_b200 import _abi
    tl = _abi.load_test_lib()
    for n in _abi.EXPORTS:
        assert hasattr(tl, n), n
    for n in ("ftrf_set_shade_path", "ftrf_test_gemm3x", "ftrf_test_wgrad3x", "ftrf_test_wgrad_fused",
              "ftrf_debug_mma_probe", "ftrf_debug_set_app_bwd_tc", "ftrf_debug_set_runs"):
        assert hasattr(tl, n), n
        assert not hasattr(_abi.lib, n), n
