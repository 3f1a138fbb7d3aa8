"""Compact wire format: oracle self-consistency on CPU; on the GPU the CUDA path against the
oracle, against the reference's own CUDA library (side by side), and the reference's round-trip tests
(src/matrix/gpu_dcrt_poly.rs:1964-2073)."""
import os

import numpy as np
import pytest

from oracle import oracle as orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CASES = [(8, 2, 17, 1), (16, 1, 52, 17), (32, 3, 24, 12), (64, 2, 51, 17), (16, 9, 28, 14), (8, 5, 59, 20)]


def small_signed(plan, moduli, rows, cols, bound, seed):
    """coefficient-form matrix whose centred coefficients are uniform in [-bound, bound]"""
    rng = np.random.default_rng(seed)
    vals = rng.integers(-bound, bound + 1, size=(rows, cols, plan.n))
    out = np.empty((rows, cols, len(moduli), plan.n), dtype=np.uint64)
    for l, q in enumerate(moduli):
        out[:, :, l, :] = np.mod(vals, int(q)).astype(np.uint64)
    return out


@pytest.mark.parametrize("n,depth,bits,bb", CASES)
def test_oracle_compact_roundtrip_and_width(n, depth, bits, bb):
    moduli = orc.prime_chain(n, bits, depth)
    plan = orc.Plan(n, moduli)
    full = plan.uniform(2, 2, seed=7)
    B, payload = orc.compact_encode(full, moduli)
    Q = 1
    for q in moduli:
        Q *= int(q)
    assert 2 <= B <= (Q >> 1).bit_length() + 1 and len(payload) == (4 * n * B + 7) // 8
    assert np.array_equal(orc.compact_decode(payload, B, 2, 2, moduli, n), full)
    small = small_signed(plan, moduli, 1, 3, 5, seed=8)
    B, payload = orc.compact_encode(small, moduli)
    assert B == 4  # |x| <= 5: three magnitude bits + sign
    assert np.array_equal(orc.compact_decode(payload, B, 1, 3, moduli, n), small)
    assert orc.compact_encode(np.zeros_like(small), moduli) == (0, b"")


def _gpu(n, depth, bits, bb, library=None):
    import mxx_b200 as mx

    moduli = orc.prime_chain(n, bits, depth)
    return mx, moduli, mx.GpuDCRTPolyParams(n, moduli, bb, library=library), orc.Plan(n, moduli)


def _envelope(data):
    from mxx_b200.matrix import _unvarint

    pos, fields = 2, []
    for _ in range(6):
        v, pos = _unvarint(data, pos)
        fields.append(v)
    return data[0], data[1], fields, data[pos:]


@pytest.mark.gpu
@pytest.mark.parametrize("n,depth,bits,bb", CASES + [(4096, 15, 24, 12), (1024, 16, 28, 14), (2048, 8, 51, 17)])
def test_cuda_compact_matches_oracle(n, depth, bits, bb):
    mx, moduli, params, plan = _gpu(n, depth, bits, bb)
    small_cases = n <= 64
    for host in (plan.uniform(2, 2, seed=11), small_signed(plan, moduli, 1, 3, 5, seed=12), small_signed(plan, moduli, 2, 1, (1 << 40), seed=13)):
        g = mx.GpuDCRTPolyMatrix.from_rns(params, host, False)
        data = g.to_compact_bytes()
        ver, fmt, (level, nrow, ncol, B, bpc, plen), payload = _envelope(data)
        assert (ver, fmt, level, nrow, ncol) == (1, 0, depth - 1, host.shape[0], host.shape[1]) and bpc == (B + 7) // 8 and plen == len(payload)
        if small_cases:
            wB, wpayload = orc.compact_encode(host, moduli)
            assert (B, payload) == (wB, wpayload)
        back = mx.GpuDCRTPolyMatrix.from_compact_bytes(params, data)
        assert not back.is_ntt and np.array_equal(back.to_rns(), host)
        # EVAL matrices travel as coefficients and come back in the evaluation domain
        e = mx.GpuDCRTPolyMatrix.from_rns(params, host, False)
        e.ntt_in_place()
        de = e.to_compact_bytes()
        assert de[1] == 1 and de[2:] == data[2:]
        assert mx.GpuDCRTPolyMatrix.from_compact_bytes(params, de) == e


@pytest.mark.gpu
def test_zero_compact_bytes_roundtrip():
    """gpu_dcrt_poly.rs:2050-2073"""
    mx, moduli, params, plan = _gpu(128, 2, 24, 12)
    for is_ntt in (False, True):
        data = mx.GpuDCRTPolyMatrix.zero_compact_bytes(params, 2, 3, 1, is_ntt, 9)
        dec = mx.GpuDCRTPolyMatrix.from_compact_bytes(params, data)
        assert dec == mx.GpuDCRTPolyMatrix.zero(params, 2, 3, 1, is_ntt)
    z = mx.GpuDCRTPolyMatrix.zero(params, 1, 2, 1, False)
    ver, fmt, (level, nrow, ncol, B, bpc, plen), payload = _envelope(z.to_compact_bytes())
    assert (B, bpc, plen) == (0, 0, 0)
    # native validation (MatrixSerde.cu:1671-1694): payload length must match the bit width
    import ctypes as C

    from mxx_b200 import ffi

    m = mx.GpuDCRTPolyMatrix(params, 1, 1, 1, False)
    buf = (C.c_uint8 * 16)()
    assert ffi.lib().gpu_matrix_load_compact_bytes(m.raw, buf, 16, 0) != 0
    assert ffi.lib().gpu_matrix_load_compact_bytes(m.raw, buf, 15, 9) != 0


@pytest.mark.gpu
@pytest.mark.parametrize("n,depth,bits,bb", [(64, 3, 24, 12), (128, 2, 51, 17), (256, 9, 28, 14)])
def test_compact_side_by_side_with_reference_library(n, depth, bits, bb):
    """the same matrix through the reference's CUDA library (oracle/_ref) and through ours: identical bytes"""
    ref_path = os.path.join(ROOT, "oracle", "_ref", "libgpupoly_ref.so")
    if not os.path.exists(ref_path):
        pytest.skip("reference CUDA library not built")
    from mxx_b200 import ffi

    ref = ffi.bind(ref_path, extensions=False)
    mx, moduli, params, plan = _gpu(n, depth, bits, bb)
    rparams = mx.GpuDCRTPolyParams(n, moduli, bb, library=ref)
    for host in (plan.uniform(2, 2, seed=21), small_signed(plan, moduli, 1, 2, 1000, seed=22)):
        ours = mx.GpuDCRTPolyMatrix.from_rns(params, host, False).to_compact_bytes()
        theirs = mx.GpuDCRTPolyMatrix.from_rns(rparams, host, False).to_compact_bytes()
        assert ours == theirs
        assert np.array_equal(mx.GpuDCRTPolyMatrix.from_compact_bytes(rparams, ours).to_rns(), host)


@pytest.mark.gpu
@pytest.mark.parametrize("n,depth,bits,bb", [(16, 2, 17, 1), (64, 3, 24, 12), (128, 2, 51, 17), (256, 9, 28, 14), (64, 5, 59, 20)])
def test_bigint_conversion_on_device(n, depth, bits, bb):
    """BigUint coefficients -> residues and back (CRT) on the device, against Python integers"""
    import random

    mx, moduli, params, plan = _gpu(n, depth, bits, bb)
    Q = 1
    for q in moduli:
        Q *= int(q)
    rnd = random.Random(5)
    for big in (Q, Q * Q + 12345, 1 << 40):
        vals = [rnd.randrange(big) for _ in range(2 * 3 * n)]
        want = np.array([[v % int(q) for v in vals] for q in moduli], dtype=np.uint64)  # [limb][poly*n]
        want = want.reshape(depth, 6, n).transpose(1, 0, 2).reshape(2, 3, depth, n)
        for is_ntt in (False, True):  # the tag only says which domain the caller's numbers live in
            g = mx.GpuDCRTPolyMatrix.from_biguints(params, vals, 2, 3, is_ntt=is_ntt)
            assert g.is_ntt == is_ntt and np.array_equal(g.to_rns(), want)
        g = mx.GpuDCRTPolyMatrix.from_biguints(params, vals, 2, 3, is_ntt=False)
        assert g.to_biguints() == [v % Q for v in vals]
    # evaluation-domain matrices reconstruct their COEFFICIENTS (coeffs() converts first, gpu.rs:959-961)
    host = plan.uniform(1, 2, seed=9)
    ge = mx.GpuDCRTPolyMatrix.from_rns(params, host, True)
    gc = mx.GpuDCRTPolyMatrix.from_rns(params, plan.intt(host), False)
    assert ge.to_biguints() == gc.to_biguints()
    # lower level: only the active limbs take part
    if depth > 1:
        vals = [rnd.randrange(1 << 70) for _ in range(n)]
        g = mx.GpuDCRTPolyMatrix.from_biguints(params, vals, 1, 1, is_ntt=False, level=depth - 2)
        q_low = 1
        for q in moduli[: depth - 1]:
            q_low *= int(q)
        assert g.to_biguints() == [v % q_low for v in vals]
