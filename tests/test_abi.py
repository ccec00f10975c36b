"""CPU suite, part 3: the drop-in boundary.  The C-ABI library loads, exports every symbol
include/icx.h declares (and nothing with torch / C++ types in its signature), the Python binding
covers all of them, the AoS structs are byte-identical to the reference components, and -- there
being no GPU here -- every compute entry point refuses loudly instead of falling back to a CPU."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from intercession_b200 import abi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADERS = [os.path.join(ROOT, "include", "icx.h"), os.path.join(ROOT, "include", "icx_nccl.h")]


def declared_symbols():
    out = []
    for h in HEADERS:
        text = open(h).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        out += re.findall(r"ICX_API\s+[\w\s\*]+?\b(icx_\w+)\s*\(", text)
    return out


def test_header_symbols_exported_and_bound(built):
    syms = declared_symbols()
    assert len(syms) >= 25 and len(set(syms)) == len(syms)
    lib = abi.load()
    out = subprocess.run(["nm", "-D", "--defined-only", abi.LIB_PATH], capture_output=True, text=True, check=True).stdout
    exported = set(line.split()[-1] for line in out.splitlines() if " T " in line)
    for s in syms:
        assert s in exported, f"{s} declared in include/icx.h but not exported by libicx.so"
        assert s in abi.SIGNATURES, f"{s} has no ctypes signature in intercession_b200/abi.py"
        assert getattr(lib, s) is not None
    # nothing but the C ABI leaks out of the library (hidden visibility): no mangled C++ symbols
    leaked = [e for e in exported if e.startswith("_Z")]
    assert not leaked, leaked[:5]
    assert set(abi.SIGNATURES) <= set(syms)


def test_header_compiles_as_plain_c(tmp_path):
    src = tmp_path / "t.c"
    src.write_text('#include "icx.h"\n#include "icx_nccl.h"\n_Static_assert(sizeof(IcxTransform)==40,"T");_Static_assert(sizeof(IcxPhysics)==108,"P");\n'
                   '_Static_assert(sizeof(IcxCollider)==84,"C");_Static_assert(sizeof(IcxColliderComponent)==168,"CC");\n'
                   '_Static_assert(sizeof(IcxContact)==48,"K");int main(void){IcxConfig c; icx_config_default(&c); return (int)c.struct_size;}\n')
    subprocess.run(["gcc", "-std=c11", "-Wall", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"), "-c", str(src),
                    "-o", str(tmp_path / "t.o")], check=True)


def test_struct_layouts_match_reference_components():
    """SURVEY 8a sizes / offsets (probe of the reference structs under x86-64 g++)."""
    assert abi.transform_dt.itemsize == 40 and abi.physics_dt.itemsize == 108
    assert abi.collider_dt.itemsize == 84 and abi.collider_component_dt.itemsize == 168
    f = abi.physics_dt.fields
    assert [f[k][1] for k in ("velocity", "acceleration", "angularVelocity", "angularAcceleration", "linearDrag", "angularDrag",
                              "collisionLinearDrag", "collisionAngularDrag", "mass", "lockOrigin", "lockedOrigin",
                              "lockOrientation", "lockedOrientation", "isAsleep")] == [0, 12, 24, 36, 48, 52, 56, 60, 64, 68, 72, 84, 88, 104]
    f = abi.collider_dt.fields
    assert [f[k][1] for k in ("colliderType", "collisionType", "isActive", "useBehaviorsResponse", "localTransform",
                              "inheritOrientation", "influenceOrientation", "minParametricValue", "staticFriction",
                              "dynamicFriction", "restitution", "stiffness", "damping", "restLength")] == [0, 4, 8, 9, 12, 52, 53, 56, 60, 64, 68, 72, 76, 80]
    assert C.sizeof(abi.IcxConfig) == 60 and C.sizeof(abi.IcxStats) == 64


def test_entity_id_helpers():
    assert abi.widen_entity(0x1234) == 0x10000234
    assert abi.compose_entity(3, 0x12345, 7) == (3 << 28) | (0x12345 << 4) | 7


def test_no_cpu_fallback(built):
    """Without a CUDA device nothing computes: creation fails with ICX_E_CUDA, NULL worlds are rejected."""
    lib = abi.load()
    assert lib.icx_abi_version() == 2
    if lib.icx_device_count() > 0:
        pytest.skip("a CUDA device is present")
    cfg = abi.IcxConfig()
    lib.icx_config_default(C.byref(cfg))
    assert cfg.struct_size == C.sizeof(abi.IcxConfig) and tuple(cfg.gravity) == (0.0, -10.0, 0.0)
    h = C.c_void_p()
    assert lib.icx_world_create(C.byref(cfg), C.byref(h)) == abi.E_CUDA and not h.value
    assert lib.icx_step(None, 1.0 / 45) == abi.E_INVALID
    assert lib.icx_collision_engage(None) == abi.E_INVALID
    assert lib.icx_status_string(abi.E_CUDA) == b"CUDA error"
    cfg.struct_size = 4
    assert lib.icx_world_create(C.byref(cfg), C.byref(h)) == abi.E_INVALID
    from intercession_b200 import world
    with pytest.raises(world.IcxError):
        world.World(16)


def test_missing_library_fails_loudly(tmp_path):
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        abi.load(str(tmp_path / "libicx_missing.so"))
