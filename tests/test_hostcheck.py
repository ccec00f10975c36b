"""CPU suite, part 4: the PRODUCT's device arithmetic (intercession_b200/csrc/icx_narrow.cuh,
icx_response.cuh -- the very headers the kernels include) compiled for the host by
tests/hostcheck/hostcheck.cu and checked bit-for-bit against the golden KATs frozen from the
reference.  This pins operation order / zero signs of the CUDA code without a GPU; the `-m gpu`
tests then confirm the device build rounds identically (-fmad=false)."""
import ctypes as C
import os

import numpy as np
import pytest

from intercession_b200 import abi
import katgen
import refs
from util import fields_diff

HERE = os.path.dirname(os.path.abspath(__file__))
_vp, _u32 = C.c_void_p, C.c_uint32
SIGS = {"icxh_intersect": refs.REF_SIGS["icxref_intersect"], "icxh_response": refs.REF_SIGS["icxref_response"]}


@pytest.fixture(scope="module")
def hostlib(built):
    lib = refs._load(os.path.join(HERE, "hostcheck", "libicx_hostcheck.so"), SIGS)
    assert lib is not None, "tests/hostcheck/libicx_hostcheck.so not built (see __graft_entry__.build)"
    return lib


def test_product_math_matches_golden_kats(hostlib):
    g = np.load(os.path.join(HERE, "golden", "kat_procedures.npz"))
    cases, exp = katgen.all_cases(), g["out"]
    got = katgen.run_kat(hostlib, "icxh", cases)
    ok = exp["hit"] != katgen.REF_CRASHED
    # the product only implements the wired table entries: null entries report no hit
    d = fields_diff(got[ok], exp[ok], skip=("minParamA", "minParamB"))
    assert not d, d[:5]
    ray = (cases["cA"]["colliderType"] == abi.COLLIDER_RAY) | (cases["cB"]["colliderType"] == abi.COLLIDER_RAY)
    sel = ok & ray & (exp["hit"] == 1)
    assert sel.sum() > 5
    assert np.array_equal(got["minParamA"][sel].view(np.uint32), exp["minParamA"][sel].view(np.uint32))
    assert np.array_equal(got["minParamB"][sel].view(np.uint32), exp["minParamB"][sel].view(np.uint32))


def test_product_math_matches_oracle_fresh_seed(hostlib):
    cases = katgen.random_cases(90210, per_combo=64)
    exp = katgen.run_kat(refs.oracle(), "icxo", cases)
    got = katgen.run_kat(hostlib, "icxh", cases)
    finite = np.isfinite(exp["point"]).all(axis=1) & np.isfinite(exp["depth"])      # reference-UB cases (empty manifold)
    d = fields_diff(got[finite], exp[finite], skip=("minParamA", "minParamB"))
    assert not d, d[:5]
