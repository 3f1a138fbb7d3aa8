"""GPU parity tests of the limb-sharded / 2-D sharded path on ONE device (every limb group is a context of its own on the
same GPU; the all-gather along the limb axis is replaced by device copies into the same layout), against the CPU oracle:

* gpu_matrix_mul_decompose_gathered: S * G^-1(other) computed per limb group from the gathered coefficient form equals the
  oracle's unsharded product on that group's limbs — u32, u64 and FP64-pipe contexts, rings of one block and of several
  (cluster kernels), both `small` variants, column sub-ranges, mixed tower widths;
* gpu_matrix_add_scaled_block against add(mul_scalar) of the oracle;
* the stacked BGG gate (mxx_b200/bgg.py StackedEncoding) equals the unstacked reference sequence.
The multi-process versions (NCCL, 2 GPUs) are in tests/nccl_assembly_check.py; the partition logic on CPU (gloo) in
tests/test_dist_gloo.py."""
import numpy as np
import pytest

from oracle import oracle as orc

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mx():
    import mxx_b200

    return mxx_b200


CASES = [
    # n, moduli spec (depth, bits) or explicit list, bb, groups, d, c, rows
    (1 << 12, (4, 24), 12, 2, 1, 3, 2),
    (1 << 12, (4, 28), 7, 4, 2, 2, 1),      # unaligned base, one tower per rank
    (1 << 13, (6, 24), 12, 3, 1, 2, 2),
    (1 << 14, (4, 28), 14, 2, 1, 2, 1),
    (1 << 15, (4, 28), 14, 2, 1, 2, 1),     # u32 cluster kernels
    (1 << 12, (4, 51), 17, 2, 1, 3, 2),     # FP64-pipe context
    (1 << 14, (4, 51), 17, 4, 1, 2, 1),     # FP64 cluster (2 blocks of 2^13)
    (1 << 12, (2, 59), 20, 2, 1, 2, 1),     # integer u64 context
]


def _gathered_device(mx, sub_params, coeff_full, groups, tpg):
    """(d, c, towers, n) host coefficient form -> the (groups*d, c) COEFF matrix an all-gather along the limb axis leaves.
    The words are written into the matrix's device storage as NCCL would write them (raw, in the context's word width):
    gpu_matrix_load_rns_batch would reduce the other groups' towers modulo THIS context's moduli."""
    import torch

    from mxx_b200.dist import device_tensor

    d, c = coeff_full.shape[0], coeff_full.shape[1]
    blocks = [coeff_full[:, :, g * tpg:(g + 1) * tpg] for g in range(groups)]
    host = np.ascontiguousarray(np.concatenate(blocks, axis=0))
    assert host.shape[0] == groups * d and host.shape[1] == c
    mat = mx.GpuDCRTPolyMatrix(sub_params, groups * d, c, tpg - 1, False)
    dst = device_tensor(mat)
    word = dst.numel() // host.size
    raw = host.astype(np.uint32 if word == 4 else np.uint64)
    dst.copy_(torch.from_numpy(raw.view(np.uint8).reshape(-1)))
    torch.cuda.synchronize()
    return mat, host


@pytest.mark.parametrize("n,spec,bb,groups,d,c,rows", CASES)
@pytest.mark.parametrize("small", [False, True])
def test_mul_decompose_gathered_matches_oracle(mx, n, spec, bb, groups, d, c, rows, small):
    moduli = mx.standard_moduli(n, spec[0], spec[1])
    towers = len(moduli)
    tpg = towers // groups
    plan = orc.Plan(n, moduli)
    ell = plan.ell(bb, small)
    S = plan.uniform(rows, d * ell, seed=101)
    other = plan.uniform(d, c, seed=102)
    want = plan.mul_decompose(S, other, bb, small)
    coeff = plan.intt(other)
    from mxx_b200.dist import towers_from_gathered_host

    for g in range(groups):
        sub = mx.GpuDCRTPolyParams(n, moduli[g * tpg:(g + 1) * tpg], bb)
        gathered, gathered_host = _gathered_device(mx, sub, coeff, groups, tpg)
        assert np.array_equal(towers_from_gathered_host(gathered_host, d), coeff)  # the layout helper agrees
        lhs = mx.GpuDCRTPolyMatrix.from_rns(sub, S[:, :, g * tpg:(g + 1) * tpg], True)
        got = lhs.mul_decompose_gathered(gathered, d, moduli, tpg, small=small).to_rns()
        assert np.array_equal(got, want[:, :, g * tpg:(g + 1) * tpg]), f"limb group {g}"
        if c >= 2:  # a column sub-range of the gathered operand (2-D partition: the block's own columns)
            part = lhs.mul_decompose_gathered(gathered, d, moduli, tpg, small=small, col0=1, ncol=c - 1).to_rns()
            assert np.array_equal(part, want[:, 1:, g * tpg:(g + 1) * tpg])


def test_mul_decompose_gathered_mixed_tower_widths(mx):
    """digit geometry follows the FULL basis: a 24-bit limb group next to a 28-bit one uses the 28-bit digit count"""
    n, bb = 1 << 12, 12
    moduli = mx.standard_moduli(n, 2, 28) + mx.standard_moduli(n, 2, 24)
    plan = orc.Plan(n, moduli)
    ell = plan.ell(bb)
    S, other = plan.uniform(2, ell, seed=7), plan.uniform(1, 3, seed=8)
    want, coeff = plan.mul_decompose(S, other, bb), plan.intt(other)
    for g in range(2):
        sub = mx.GpuDCRTPolyParams(n, moduli[2 * g:2 * g + 2], bb)
        gathered, _ = _gathered_device(mx, sub, coeff, 2, 2)
        lhs = mx.GpuDCRTPolyMatrix.from_rns(sub, S[:, :, 2 * g:2 * g + 2], True)
        got = lhs.mul_decompose_gathered(gathered, 1, moduli, 2).to_rns()
        assert np.array_equal(got, want[:, :, 2 * g:2 * g + 2])


def test_mul_decompose_gathered_rejects_bad_arguments(mx):
    n, bb = 1 << 12, 12
    moduli = mx.standard_moduli(n, 4, 24)
    sub = mx.GpuDCRTPolyParams(n, moduli[:2], bb)
    plan = orc.Plan(n, moduli)
    gathered = mx.GpuDCRTPolyMatrix.from_rns(sub, np.zeros((2, 2, 2, n), dtype=np.uint64), False)
    lhs = mx.GpuDCRTPolyMatrix.from_rns(sub, plan.uniform(1, plan.ell(bb), seed=1)[:, :, :2], True)
    with pytest.raises(mx.ffi.GpuError):  # wrong inner dimension
        bad = mx.GpuDCRTPolyMatrix.from_rns(sub, plan.uniform(1, plan.ell(bb) - 1, seed=1)[:, :, :2], True)
        bad.mul_decompose_gathered(gathered, 1, moduli, 2)
    with pytest.raises(mx.ffi.GpuError):  # a 40-bit tower cannot live in a u32 context
        lhs.mul_decompose_gathered(gathered, 1, moduli[:3] + [(1 << 40) - 87], 2)
    ok = lhs.mul_decompose_gathered(gathered, 1, moduli, 2)  # zero operand -> zero product
    assert not ok.to_rns().any()


@pytest.mark.parametrize("n,depth,bits", [(1 << 12, 3, 24), (1 << 12, 2, 51), (1 << 13, 2, 59)])
def test_add_scaled_block_matches_oracle(mx, n, depth, bits):
    moduli = mx.standard_moduli(n, depth, bits)
    params, plan = mx.GpuDCRTPolyParams(n, moduli, 12), orc.Plan(n, moduli)
    out, src, s = plan.uniform(3, 4, seed=1), plan.uniform(2, 5, seed=2), plan.uniform(1, 1, seed=3)
    g = mx.GpuDCRTPolyMatrix.from_rns(params, out, True)
    g.add_scaled_block_from(mx.GpuDCRTPolyMatrix.from_rns(params, src, True), mx.GpuDCRTPolyMatrix.from_rns(params, s, True), 1, 1, 0, 2, 2, 3)
    want = out.copy()
    want[1:3, 1:4] = plan.add(out[1:3, 1:4], plan.mul_scalar(src[0:2, 2:5], s))
    assert np.array_equal(g.to_rns(), want)


def test_stacked_gate_equals_reference_sequence(mx):
    from mxx_b200.bgg import BggEncoding, BggPublicKey, ShardedBggEvaluator, StackedEncoding
    from mxx_b200.dist import ColumnShardedOps

    n, depth, bits, bb, d = 1 << 12, 3, 24, 12, 1
    moduli = mx.standard_moduli(n, depth, bits)
    params, plan = mx.GpuDCRTPolyParams(n, moduli, bb), orc.Plan(n, moduli)
    m = d * plan.ell(bb)
    mk = lambda r, c, seed: mx.GpuDCRTPolyMatrix.from_rns(params, plan.uniform(r, c, seed=seed), True)  # noqa: E731
    left = BggEncoding(mk(1, m, 1), BggPublicKey(mk(d, m, 2)), mk(1, 1, 3))
    right = BggEncoding(mk(1, m, 4), BggPublicKey(mk(d, m, 5)), mk(1, 1, 6))
    want = left.mul(right, fused=False)  # the reference's operator sequence (encoding.rs:125-144)
    ev = ShardedBggEvaluator(ColumnShardedOps(params))  # one rank: the whole matrix is the local block
    got = ev.mul_gate_stacked(StackedEncoding.from_encoding(left), right.vector, right.pubkey.matrix, right.plaintext, m).to_encoding()
    assert got.vector == want.vector and got.pubkey.matrix == want.pubkey.matrix and got.plaintext == want.plaintext


@pytest.mark.parametrize("n,depth,bits", [(1 << 12, 3, 24), (1 << 15, 2, 28), (1 << 14, 2, 51)])
def test_intt_to_block_and_side_stream(mx, n, depth, bits):
    """gpu_matrix_intt_to: rows of a larger buffer <- coefficient form of src (src untouched), on the context stream and
    on a caller's side stream (gpu_thread_stream_override)"""
    import torch

    moduli = mx.standard_moduli(n, depth, bits)
    params, plan = mx.GpuDCRTPolyParams(n, moduli, 12), orc.Plan(n, moduli)
    ev = plan.uniform(2, 3, seed=5)
    src = mx.GpuDCRTPolyMatrix.from_rns(params, ev, True)
    buf = mx.GpuDCRTPolyMatrix.from_rns(params, np.zeros((6, 3, depth, n), dtype=np.uint64), False)
    buf.intt_rows_from(src, 2)
    want = np.zeros((6, 3, depth, n), dtype=np.uint64)
    want[2:4] = plan.intt(ev)
    assert np.array_equal(buf.to_rns(), want) and np.array_equal(src.to_rns(), ev)
    # the same on a side stream ordered by the caller's events
    side = torch.cuda.Stream()
    params.synchronize()
    L = params._L
    L.gpu_thread_stream_override(side.cuda_stream)
    try:
        buf.intt_rows_from(src, 4)
    finally:
        L.gpu_thread_stream_override(None)
    side.synchronize()
    want[4:6] = plan.intt(ev)
    assert np.array_equal(buf.to_rns(), want)
    coeff = mx.GpuDCRTPolyMatrix.from_rns(params, want[2:4], False)  # a COEFF source is copied
    buf.intt_rows_from(coeff, 0)
    want[0:2] = want[2:4]
    assert np.array_equal(buf.to_rns(), want)
    with pytest.raises(mx.ffi.GpuError):
        buf.intt_rows_from(src, 5)  # does not fit


def test_chunked_load_with_more_workers_than_ring_slots():
    """The chunked host->device load with its worker threads outnumbering the pinned ring slots, many loads in a row, in a
    fresh process (the worker count is fixed at first use).  Regression test for a worker that notified the caller's
    condition variable after the caller could leave (seen as memory corruption with 8 ranks x 31 workers on 24 slots)."""
    import os
    import subprocess
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = r'''
import sys, numpy as np
sys.path.insert(0, %r)
import mxx_b200 as mx
n, depth = 1 << 14, 8
moduli = mx.standard_moduli(n, depth, 24)
params = mx.GpuDCRTPolyParams(n, moduli, 12)
rng = np.random.default_rng(1)
host = rng.integers(0, min(moduli), size=(1, 160, depth, n), dtype=np.uint64)   # 21 M words: the chunked path
pin = mx.PinnedBuffer(host.shape)
pin.array[...] = host
g = mx.GpuDCRTPolyMatrix(params, 1, 160)
for i in range(40):
    g.load_rns(pin.array if i %% 2 else host, mx.GPU_POLY_FORMAT_EVAL, wait=True)
    (g + g).to_rns() if i %% 8 == 0 else None
assert np.array_equal(g.to_rns(), host)
print("loads ok")
''' % root
    env = dict(os.environ, MXX_LOAD_THREADS="24", MXX_LOAD_SLOTS="3", MXX_LOAD_CHUNK_LOG="18")
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0 and "loads ok" in r.stdout, r.stdout[-1500:] + r.stderr[-1500:]
