"""ctypes binding of ``libegoscaler_b200.so`` (the C ABI declared in ``include/egoscaler_b200.h``).

There is NO fallback: if the shared library is missing the import of anything that computes raises.
Build it with ``python -c "import __graft_entry__ as g; g.build()"`` or ``make -C egoscaler_b200/csrc``.
"""
from __future__ import annotations

import ctypes
import os
import re
from typing import Dict, List

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libegoscaler_b200.so")
HEADER_PATH = os.path.join(os.path.dirname(_HERE), "include", "egoscaler_b200.h")

_CTYPE = {
    "void*": ctypes.c_void_p, "const void*": ctypes.c_void_p, "float*": ctypes.c_void_p,
    "const float*": ctypes.c_void_p, "int64_t*": ctypes.c_void_p, "const int64_t*": ctypes.c_void_p,
    "int32_t*": ctypes.c_void_p, "const int32_t*": ctypes.c_void_p, "const uint8_t*": ctypes.c_void_p,
    "uint8_t*": ctypes.c_void_p, "double*": ctypes.c_void_p, "const double*": ctypes.c_void_p, "const char*": ctypes.c_char_p,
    "int": ctypes.c_int, "int64_t": ctypes.c_int64, "float": ctypes.c_float, "double": ctypes.c_double,
    "uint64_t": ctypes.c_uint64,
}


def declared_functions(header: str = HEADER_PATH) -> Dict[str, List[str]]:
    """Parse ``name -> [arg types]`` (+ return type under key ``name + ':ret'``) out of the public header."""
    src = open(header).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    out: Dict[str, List[str]] = {}
    for m in re.finditer(r"(?m)^\s*(int64_t|int|const char\*)\s+(esb_\w+)\s*\(([^;]*?)\)\s*;", src, flags=re.S):
        ret, name, args = m.group(1), m.group(2), m.group(3)
        types = []
        args = " ".join(args.split())
        if args and args != "void":
            for a in args.split(","):
                a = a.strip()
                t = a.rsplit(" ", 1)[0].strip() if " " in a else a
                if a.endswith("*"):
                    t = a
                t = t.replace(" *", "*")
                types.append(t)
        out[name] = types
        out[name + ":ret"] = [ret]
    return out


class EsbError(RuntimeError):
    pass


class _Lib:
    def __init__(self):
        if not os.path.exists(LIB_PATH):
            raise EsbError(
                f"{LIB_PATH} not found: the CUDA extension is required (no CPU fallback). "
                "Build it with `make -C egoscaler_b200/csrc` or `__graft_entry__.build()`.")
        self._dll = ctypes.CDLL(LIB_PATH)
        decl = declared_functions()
        self._fns = {}
        for name, types in decl.items():
            if name.endswith(":ret"):
                continue
            fn = getattr(self._dll, name)          # AttributeError if the .so does not export a declared symbol
            fn.argtypes = [_CTYPE[t] for t in types]
            fn.restype = {"const char*": ctypes.c_char_p, "int64_t": ctypes.c_int64}.get(decl[name + ":ret"][0], ctypes.c_int)
            self._fns[name] = fn

    def raw(self, name):
        return self._fns[name]

    def last_error(self) -> str:
        s = self._fns["esb_last_error"]()
        return s.decode() if s else ""

    def call(self, name, *args):
        rc = self._fns[name](*args)
        if rc != 0:
            raise EsbError(f"{name} failed with code {rc}: {self.last_error()}")
        return rc


_lib = None
# kernels launched per entry point (everything else launches exactly one); bench.py reports the total
KERNELS_PER_CALL = {"esb_cross_entropy": 3, "esb_attn_bwd": 3, "esb_attn_bwd_rope": 3, "esb_rgbd_to_points": 3}
launch_count = 0


def lib() -> _Lib:
    global _lib
    if _lib is None:
        _lib = _Lib()
    return _lib


def call(name, *args):
    global launch_count
    launch_count += KERNELS_PER_CALL.get(name, 1)
    return lib().call(name, *args)
