"""CPU: the C-ABI library loads and exports every symbol include/egoscaler_b200.h declares (no compute calls)."""
import ctypes
import os

from egoscaler_b200 import lib


def test_header_declares_functions():
    d = lib.declared_functions()
    names = [k for k in d if not k.endswith(":ret")]
    assert "esb_gemm_bf16" in names and "esb_traj_quantize" in names and len(names) >= 8


def test_library_exports_every_declared_symbol():
    assert os.path.exists(lib.LIB_PATH), "build the extension first (__graft_entry__.build())"
    dll = ctypes.CDLL(lib.LIB_PATH)
    for name in lib.declared_functions():
        if name.endswith(":ret"):
            continue
        assert hasattr(dll, name), f"{name} declared in the header but not exported"
    L = lib.lib()
    assert L.raw("esb_version")() >= 100


def test_no_alloc_entry_point_is_declared_and_its_test_is_collected_last():
    """esb_gemm_skinny_static_bf16 (no counterpart in the reference) sits in the header beside the other entry points, and its
    GPU test lives in the file pytest collects last."""
    d = lib.declared_functions()
    assert "esb_gemm_skinny_static_bf16" in d and "esb_gemm_skinny_static_scratch_bytes" in d
    here = os.path.dirname(os.path.abspath(__file__))
    files = sorted(f for f in os.listdir(here) if f.startswith("test_") and f.endswith(".py"))
    assert files[-1] == "test_zz_no_alloc.py"
