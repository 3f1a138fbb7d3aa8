"""Parity against the REFERENCE's own outputs.

tests/golden/ref_cuda_vectors.npz holds what the reference's CUDA library
(/root/reference/cuda, compiled unmodified into oracle/_ref/libgpupoly_ref.so) returned on a
B200 for seeded inputs (generator: tests/golden/make_golden.py).  The CPU test pins the
oracle to those vectors; the GPU test pins the new CUDA path to them, and — when the
prebuilt reference library travelled to the box — runs both libraries side by side.
"""
import os

import numpy as np
import pytest

from oracle import oracle as orc
from tests.golden.make_golden import CASES

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden", "ref_cuda_vectors.npz")
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libgpupoly_ref.so")


@pytest.fixture(scope="module")
def golden():
    if not os.path.exists(GOLDEN):
        pytest.skip("golden vectors not generated yet")
    return np.load(GOLDEN)


def inputs(plan, params_digits, dpt):
    A, B, C2 = plan.uniform(2, 3, seed=101), plan.uniform(3, 2, seed=102), plan.uniform(2, 3, seed=103)
    s = plan.uniform(1, 1, seed=104)
    S = plan.uniform(2, 2 * params_digits, seed=105)
    Ss = plan.uniform(1, 2 * dpt, seed=106)
    return A, B, C2, s, S, Ss


@pytest.mark.parametrize("n,depth,bits,bb", CASES)
def test_oracle_matches_reference_cuda_outputs(golden, n, depth, bits, bb):
    key = f"n{n}_d{depth}_b{bits}_bb{bb}/"
    moduli = [int(x) for x in golden[key + "moduli"]]
    assert moduli == orc.prime_chain(n, bits, depth)
    plan = orc.Plan(n, moduli)
    A, B, C2, s, S, Ss = inputs(plan, plan.ell(bb), plan.dpt(bb))
    assert np.array_equal(plan.ntt(A), golden[key + "ntt_A"])
    assert np.array_equal(plan.intt(A), golden[key + "intt_A"])
    assert np.array_equal(plan.mul(A, B), golden[key + "mul_AB"])
    assert np.array_equal(plan.add(A, C2), golden[key + "add_AC"])
    assert np.array_equal(plan.sub(A, C2), golden[key + "sub_AC"])
    assert np.array_equal(plan.mul_scalar(A, s), golden[key + "mulscalar_As"])
    assert np.array_equal(plan.ntt(plan.decompose(plan.intt(A), bb)), golden[key + "decompose_A"])
    assert np.array_equal(plan.ntt(plan.decompose(plan.intt(A), bb, True)), golden[key + "small_decompose_A"])
    assert np.array_equal(plan.ntt(plan.gadget(2, bb)), golden[key + "gadget_2"])
    assert np.array_equal(plan.ntt(plan.gadget(2, bb, True)), golden[key + "small_gadget_2"])
    assert np.array_equal(plan.mul_decompose(S, A, bb), golden[key + "mul_decompose_SA"])
    assert np.array_equal(plan.mul_decompose(Ss, A, bb, True), golden[key + "mul_decompose_small_SsA"])


def _run_ours(n, depth, bits, bb, library=None):
    from tests.golden.make_golden import run_case

    if library is None:
        from mxx_b200 import ffi

        library = ffi.lib()
    return run_case(library, n, depth, bits, bb)


@pytest.mark.gpu
@pytest.mark.parametrize("n,depth,bits,bb", CASES)
def test_cuda_path_matches_reference_cuda_outputs(golden, n, depth, bits, bb):
    key = f"n{n}_d{depth}_b{bits}_bb{bb}/"
    ours = _run_ours(n, depth, bits, bb)
    for name, arr in ours.items():
        assert np.array_equal(arr, golden[key + name]), name


@pytest.mark.gpu
@pytest.mark.parametrize("n,depth,bits,bb", [(1024, 3, 24, 12), (16384, 2, 24, 12), (4096, 2, 51, 17), (32768, 1, 28, 14)])
def test_side_by_side_with_reference_library(n, depth, bits, bb):
    """larger rings than the committed vectors: both libraries on the same inputs, same process"""
    if not os.path.exists(REF_SO):
        pytest.skip("oracle/_ref/libgpupoly_ref.so not present (built only where /root/reference exists)")
    from mxx_b200 import ffi

    ref = ffi.bind(REF_SO, extensions=False)
    theirs = _run_ours(n, depth, bits, bb, library=ref)
    ours = _run_ours(n, depth, bits, bb)
    for name in ours:
        assert np.array_equal(ours[name], theirs[name]), name
