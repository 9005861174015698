"""CPU-side checks of the drop-in boundary: both shared libraries load, export every symbol
include/sfp.h declares, and agree with the header on struct sizes.  No compute calls on the
product library here (no GPU in the CPU test tier)."""
import ctypes as C
import os
import re
import subprocess
import sys

import numpy as np
import pytest

import starframe_b200 as sf
from starframe_b200 import _ffi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "sfp.h")


def declared_symbols():
    src = open(HEADER).read()
    return sorted(set(re.findall(r"SFP_API\s+[\w\s\*]+?\b(sfp_\w+)\s*\(", src)))


def test_header_declares_what_python_binds():
    syms = declared_symbols()
    product = [s for s in syms if not s.startswith("sfp_ref_")]
    oracle = [s for s in syms if s.startswith("sfp_ref_")]
    assert sorted("sfp_" + n for n in _ffi.PRODUCT_SYMBOLS) == product
    assert sorted("sfp_ref_" + n for n in _ffi.ORACLE_SYMBOLS) == oracle


def test_product_library_exports_every_declared_symbol():
    if not os.path.exists(sf.LIB_PATH):
        import __graft_entry__ as g

        g.build()
    lib = C.CDLL(sf.LIB_PATH)
    for s in declared_symbols():
        if s.startswith("sfp_ref_"):
            continue
        assert hasattr(lib, s), s
    lib.sfp_backend.restype = C.c_char_p
    assert lib.sfp_backend() == b"cuda-sm_100a"


def test_product_library_does_not_link_or_name_the_oracle():
    out = subprocess.run(["nm", "-D", sf.LIB_PATH], capture_output=True, text=True).stdout
    assert "sfp_ref_" not in out
    ldd = subprocess.run(["ldd", sf.LIB_PATH], capture_output=True, text=True).stdout
    assert "libsfp_ref" not in ldd


def test_oracle_library_exports_every_declared_symbol(ref_api):
    for s in declared_symbols():
        if s.startswith("sfp_ref_"):
            assert hasattr(ref_api.lib, s), s
    assert ref_api.backend() == "cpu-oracle"


def test_struct_sizes_match_the_c_header(tmp_path):
    src = tmp_path / "sizes.c"
    src.write_text(
        '#include <stdio.h>\n#include "sfp.h"\nint main(void){printf("%zu %zu %zu %zu %zu %zu %zu\\n",'
        "sizeof(SfpParams),sizeof(SfpBodyDesc),sizeof(SfpJointDesc),sizeof(SfpBodyState),"
        "sizeof(SfpPair),sizeof(SfpContact),sizeof(SfpStats));return 0;}\n")
    exe = tmp_path / "sizes"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    got = [int(x) for x in subprocess.run([str(exe)], capture_output=True, text=True).stdout.split()]
    want = [d.itemsize for d in (_ffi.PARAMS_DTYPE, _ffi.BODY_DESC_DTYPE, _ffi.JOINT_DESC_DTYPE,
                                 _ffi.BODY_STATE_DTYPE, _ffi.PAIR_DTYPE, _ffi.CONTACT_DTYPE,
                                 _ffi.STATS_DTYPE)]
    assert got == want


def test_default_params_agree_between_libraries(ref_api):
    if not os.path.exists(sf.LIB_PATH):
        pytest.skip("product library not built")
    a = sf.product_api().default_params()  # pure host call, needs no GPU
    b = ref_api.default_params()
    assert a.tobytes() == b.tobytes()


def test_product_create_fails_loudly_without_a_gpu():
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(sf.SfpError) as e:
        sf.World(sf.product_api())
    assert e.value.code == _ffi.SFP_E_CUDA


def test_host_mirror_builds_a_scene():
    """the Physics / Body / Collider mirror only does bookkeeping until the first tick"""
    from starframe_b200.physics import Body, Collider, Material

    col = Collider.new_rounded_rect(1.0, 0.5, 0.1).with_material(Material(0.7, 0.5, 0.2))
    assert (col.hx, col.hy, col.radius) == pytest.approx((0.4, 0.15, 0.1))
    cap = Collider.new_capsule(2.0, 0.25)
    assert (cap.hx, cap.radius) == (1.0, 0.25)
    b = Body.new_dynamic(1.0, 2.0, 0.3).with_velocity(1.0, 0.0, 0.5)
    assert (b.vx, b.omega, b.is_static) == (1.0, 0.5, False)
    assert Body.new_static().is_static


def test_host_mirror_drives_the_oracle_end_to_end(ref_api):
    """the same mirror bound to the oracle library: tick, events, queries, per-body updates and the
    host-state round trip go through the calls the product library exports under the same names"""
    import starframe_b200 as sf

    ph = sf.Physics(api=ref_api)
    ph.add_body(sf.Body.new_static(0.0, -0.5), sf.Collider.new_rect(10.0, 1.0))
    ball = ph.add_body(sf.Body.new_dynamic(0.0, 0.5), sf.Collider.new_circle(0.5))
    for _ in range(30):
        ph.tick()
    ev = ph.contact_events()
    assert len(ev) == 1 and (ev[0]["a"], ev[0]["b"]) == (0, ball) and ev[0]["impulse"] > 0.0
    hit = ph.ray_cast([(0.0, 5.0)], [(0.0, -1.0)], 10.0)
    assert hit[0]["body"] == ball and hit[0]["t"] == pytest.approx(4.0, abs=1e-3)
    assert list(ph.query_aabb((-1.0, 0.0), (1.0, 1.0))) == [0, ball]
    st = ph.body_states()
    moved = st[[ball]].copy()
    moved["x"] = 3.0
    ph.set_body_states([ball], moved)
    assert ph.body_states()[ball]["x"] == np.float32(3.0)
    out = ph.tick_host(ph.body_states())
    assert out[ball]["x"] == pytest.approx(3.0, abs=1e-4) and out[ball]["y"] == pytest.approx(0.5, abs=1e-2)
