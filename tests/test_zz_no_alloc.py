"""gpu_matrix_decompose_base_noalloc (include/gpupoly.h; no counterpart in the reference):
the gadget decomposition with its scratch — the coefficient form of an EVAL input — in a static device array
instead of pool memory.  Compared bit for bit (integer work: tolerance zero, as everywhere in this suite) with
gpu_matrix_decompose_base through the C ABI, on two calls in a row that share the array, for both word widths, the block and
the cluster kernels; inputs beyond GPU_NOALLOC_SCRATCH_BYTES are rejected.  Named test_zz_* so that pytest collects it last."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

NOALLOC_SCRATCH_BYTES = 16 << 20  # GPU_NOALLOC_SCRATCH_BYTES, include/gpupoly.h

CASES = [
    # (n, depth, bits, bb, rows, cols)
    (1 << 12, 3, 24, 12, 2, 3),    # u32 words, one shared-memory block
    (1 << 14, 4, 28, 14, 1, 4),    # u32, C3-shaped ring
    (1 << 15, 2, 28, 14, 1, 2),    # u32, 2-CTA cluster kernels
    (1 << 13, 2, 51, 17, 2, 2),    # u64 words on the FP64 pipe
    (1 << 12, 2, 59, 20, 1, 3),    # u64 words, integer kernels
]


def _pair(mx, n, depth, bits, bb, rows, cols, seed):
    from oracle import oracle as orc

    moduli = mx.standard_moduli(n, depth, bits)
    params, plan = mx.GpuDCRTPolyParams(n, moduli, bb), orc.Plan(n, moduli)
    host = plan.uniform(rows, cols, seed=seed)
    return params, plan, host


@pytest.mark.parametrize("n,depth,bits,bb,rows,cols", CASES)
def test_noalloc_decompose_equals_decompose_base_twice(n, depth, bits, bb, rows, cols):
    import mxx_b200 as mx

    params, plan, _ = _pair(mx, n, depth, bits, bb, rows, cols, 1)
    ell = params.modulus_digits()
    out = mx.GpuDCRTPolyMatrix(params, rows * ell, cols)  # created once: the calls below allocate nothing
    for call, seed in enumerate((11, 12)):  # two calls in a row share the static array; different inputs
        host = plan.uniform(rows, cols, seed=seed)
        src = mx.GpuDCRTPolyMatrix.from_rns(params, host, True)  # EVAL input: the path that needs scratch
        want = src.decompose().to_rns()
        got = src.decompose_noalloc(out).to_rns()
        assert np.array_equal(got, want), f"call {call}"
        assert np.array_equal(src.to_rns(), host), "the input must stay untouched"
    # COEFF input: no scratch needed
    src = mx.GpuDCRTPolyMatrix.from_rns(params, plan.uniform(rows, cols, seed=13), False)
    assert np.array_equal(src.decompose_noalloc(out).to_rns(), src.decompose().to_rns())


def test_noalloc_decompose_matches_oracle_and_makes_no_pool_allocation():
    import mxx_b200 as mx

    n, depth, bits, bb, rows, cols = 1 << 12, 3, 24, 12, 1, 2
    params, plan, host = _pair(mx, n, depth, bits, bb, rows, cols, 21)
    src = mx.GpuDCRTPolyMatrix.from_rns(params, host, True)
    out = mx.GpuDCRTPolyMatrix(params, rows * params.modulus_digits(), cols)
    L = mx.ffi.lib()
    dev = params.gpu_ids[0]
    src.decompose_noalloc(out)  # first call: the driver loads the kernels' code (lazy module loading takes device memory)
    free0, total0 = C.c_size_t(), C.c_size_t()
    params.synchronize()
    assert L.gpu_device_mem_info(dev, C.byref(free0), C.byref(total0)) == 0
    src.decompose_noalloc(out)
    params.synchronize()
    free1 = C.c_size_t()
    assert L.gpu_device_mem_info(dev, C.byref(free1), C.byref(total0)) == 0
    assert free1.value >= free0.value, "the call must not take device memory"
    assert np.array_equal(plan.intt(out.to_rns()), plan.decompose(plan.intt(host), bb))


def test_noalloc_decompose_rejects_inputs_beyond_the_static_scratch():
    import mxx_b200 as mx

    n, depth, bits, bb = 1 << 14, 4, 28, 14
    moduli = mx.standard_moduli(n, depth, bits)
    params = mx.GpuDCRTPolyParams(n, moduli, bb)
    poly_bytes = depth * n * 4
    cols = NOALLOC_SCRATCH_BYTES // poly_bytes + 1  # one polynomial too many
    src = mx.GpuDCRTPolyMatrix.sample_distribution(params, 1, cols, 0, 0.0, [1, 2, 3, 4])
    out = mx.GpuDCRTPolyMatrix(params, params.modulus_digits(), cols)
    with pytest.raises(Exception, match="exceeds the static scratch"):
        src.decompose_noalloc(out)
    # exactly at the limit it runs, and equals the allocating entry point
    fit = mx.GpuDCRTPolyMatrix.sample_distribution(params, 1, cols - 1, 0, 0.0, [1, 2, 3, 4])
    assert fit.decompose_noalloc() == fit.decompose()
