"""CPU-side: the C-ABI library loads, exports every symbol include/b2cuda.h declares, POD layouts match the ctypes
mirror, and the product refuses to run without a GPU (no CPU fallback)."""
import ctypes as C
import os
import re
import subprocess
import pytest
from box2d_b200 import capi, api

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    subprocess.check_call(["make", "-s", "-j8", "-C", os.path.join(ROOT, "box2d_b200", "csrc")])
    return C.CDLL(capi.product_library_path())


def test_exports_every_declared_symbol(lib):
    header = open(os.path.join(ROOT, "include", "b2cuda.h")).read()
    declared = set(re.findall(r"\b(b2cuda_\w+)\s*\(", header))
    assert declared, "no declarations parsed"
    assert declared == set(capi.exported_symbols()), declared ^ set(capi.exported_symbols())
    for sym in declared:
        assert hasattr(lib, sym), sym


def test_pod_sizes_match_header(tmp_path):
    names = ["B2CWorldDef", "B2CBodyDef", "B2CShapeDef", "B2CGeometry", "B2CRevoluteJointDef", "B2CBodyRead", "B2CManifoldPointRead",
             "B2CContactRead", "B2CContactTouchEvent", "B2CContactHitEvent", "B2CJointRead", "B2CJointDef", "B2CShapeRead", "B2CCounters", "B2CProfile"]
    src = '#include <stdio.h>\n#include "b2cuda.h"\nint main(){' + "".join(f'printf("{n} %zu\\n", sizeof({n}));' for n in names) + "return 0;}"
    c = tmp_path / "s.c"
    c.write_text(src)
    exe = tmp_path / "s"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(c), "-o", str(exe)])
    sizes = dict(line.split() for line in subprocess.check_output([str(exe)]).decode().splitlines())
    mirror = {"B2CWorldDef": capi.WorldDef, "B2CBodyDef": capi.BodyDef, "B2CShapeDef": capi.ShapeDef, "B2CGeometry": capi.Geometry,
              "B2CRevoluteJointDef": capi.RevoluteJointDef, "B2CJointDef": capi.JointDef, "B2CBodyRead": capi.BodyRead, "B2CManifoldPointRead": capi.ManifoldPointRead,
              "B2CContactRead": capi.ContactRead, "B2CContactTouchEvent": capi.ContactTouchEvent, "B2CContactHitEvent": capi.ContactHitEvent,
              "B2CJointRead": capi.JointRead, "B2CShapeRead": capi.ShapeRead, "B2CCounters": capi.Counters, "B2CProfile": capi.Profile}
    for n, ty in mirror.items():
        assert int(sizes[n]) == C.sizeof(ty), (n, sizes[n], C.sizeof(ty))


def test_defaults_match_reference_defs(lib):
    be = api.CudaBackend()
    wd, bd, sd, jd = capi.WorldDef(), capi.BodyDef(), capi.ShapeDef(), capi.RevoluteJointDef()
    assert be.fns["default_world_def"](C.byref(wd)) == 0 and be.fns["default_body_def"](C.byref(bd)) == 0
    assert be.fns["default_shape_def"](C.byref(sd)) == 0 and be.fns["default_revolute_joint_def"](C.byref(jd)) == 0
    ref = api.default_world_def()
    for f, _ in capi.WorldDef._fields_:
        a, b = getattr(wd, f), getattr(ref, f)
        assert (a.x, a.y) == (b.x, b.y) if f == "gravity" else a == b, f
    assert abs(sd.friction - 0.6) < 1e-7 and sd.density == 1.0 and sd.categoryBits == 2**64 - 1 and sd.invokeContactCreation == 1
    assert bd.gravityScale == 1.0 and bd.isAwake == 1 and bd.isEnabled == 1 and (bd.rotation.c, bd.rotation.s) == (1.0, 0.0)
    assert jd.constraintHertz == 60.0 and jd.constraintDampingRatio == 2.0 and jd.bodyIdA == -1
    for t in range(7):          # the generic joint definition: C defaults == the Python mirror's, field by field
        gd, ref = capi.JointDef(), api.default_joint_def(t)
        assert be.fns["default_joint_def"](t, C.byref(gd)) == 0
        assert bytes(gd) == bytes(ref), t
    assert be.fns["default_joint_def"](7, C.byref(capi.JointDef())) == -1


def test_no_gpu_means_loud_failure_not_fallback(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(api.B2Error) as e:
        api.World(api.CudaBackend(), __import__("box2d_b200").scenes.world_def())
    assert "no usable CUDA device" in str(e.value) or "rc=-3" in str(e.value)


def test_sleep_is_rejected_not_ignored(lib):
    be = api.CudaBackend()
    h = C.c_void_p()
    rc = be.fns["world_create"](C.byref(api.default_world_def()), 0, C.byref(h))   # default def has enableSleep=1
    assert rc in (-2, -3)
