"""Loads the CPU oracle (oracle/libsfp_ref.so).  TEST INFRASTRUCTURE: import this only from
tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs."""
from __future__ import annotations

import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(_HERE, "libsfp_ref.so")
_api = None


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "sfp_ref.cpp")
    inc = os.path.join(_HERE, "..", "include")
    deps = [src] + [os.path.join(inc, h) for h in ("sfp.h", "sfp_math.h", "sfp_host.h")]
    stale = force or not os.path.exists(LIB) or any(
        os.path.getmtime(d) > os.path.getmtime(LIB) for d in deps)
    if stale:
        subprocess.run(["make", "-C", _HERE, "-B", "libsfp_ref.so"], check=True,
                       stdout=subprocess.DEVNULL)
    return LIB


def api():
    """The oracle bound through the same ctypes class as the product (prefix sfp_ref_)."""
    global _api
    if _api is None:
        from starframe_b200._ffi import Api
        build()
        _api = Api(LIB, "sfp_ref_")
    return _api
