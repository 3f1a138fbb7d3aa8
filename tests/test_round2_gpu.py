"""Round-2 GPU parity tests (through the C ABI, against the CPU oracle):

* rings of 2 / 4 / 8 shared-memory blocks (the cluster / DSMEM kernels of ntt_cluster.cuh) at n = 2^15 .. 2^17 with
  L = 4 and 8 limbs on both word widths — NTT, INTT, decompose, mul_decompose compared with the ORACLE (not a round trip:
  a self-inverse stride bug would survive a round trip);
* the same rings with MXX_NTT_CLUSTER=0 (round-1 global passes) in a fresh process;
* two host threads inside one context and two contexts with different moduli on one device, repeated 20 times
  (the reference calls the library from rayon workers, src/sampler/trapdoor/gpu.rs:371-397, and per-gate threads,
  src/circuit/poly_circuit/eval.rs:562-575; its validation harness repeats GPU tests like
  scripts/lib/repo_validation.py:138-159);
* non-canonical host words are reduced on load (every kernel afterwards sees canonical residues);
* fill_gadget at a reduced level of a context with mixed modulus widths (cuda/src/matrix/MatrixDecompose.cu:346-355);
* one-row products wider than grid.y allows.
"""
import os
import subprocess
import sys
import threading

import numpy as np
import pytest

from oracle import oracle as orc

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def mx():
    import mxx_b200

    return mxx_b200


_cache = {}


def setup(mx, n, depth, bits, bb):
    key = (n, depth, bits, bb)
    if key not in _cache:
        moduli = mx.standard_moduli(n, depth, bits)
        _cache[key] = (mx.GpuDCRTPolyParams(n, moduli, bb), orc.Plan(n, moduli))
    return _cache[key]


def G(mx, params, host, is_ntt=True):
    return mx.GpuDCRTPolyMatrix.from_rns(params, host, is_ntt)


BIG = [
    # (n, depth, bits, bb)
    (1 << 15, 4, 28, 14),
    (1 << 15, 8, 24, 12),
    (1 << 16, 4, 28, 14),
    (1 << 16, 8, 28, 14),
    (1 << 16, 4, 24, 12),   # NOCSUB forward
    (1 << 17, 2, 28, 14),
    (1 << 14, 8, 51, 17),
    (1 << 15, 4, 51, 17),
    (1 << 16, 4, 51, 17),
    (1 << 16, 8, 51, 17),
    (1 << 16, 4, 59, 20),   # u64 integer kernels
    (1 << 15, 4, 55, 19),   # u64 integer NOCSUB
]


@pytest.mark.parametrize("n,depth,bits,bb", BIG)
def test_big_ring_ntt_intt_match_oracle(mx, n, depth, bits, bb):
    params, plan = setup(mx, n, depth, bits, bb)
    a = plan.uniform(1, 3, seed=n + bits + depth)
    mod = np.array(plan.moduli, dtype=np.uint64)
    a[0, 2, :, ::3] = (mod - 1)[:, None]  # a third of one polynomial at the largest residue
    g = G(mx, params, a, is_ntt=False)
    g.ntt_in_place()
    ev = g.to_rns()
    want = plan.ntt(a)
    assert np.array_equal(ev, want)
    h = G(mx, params, want, is_ntt=True)
    h.intt_all_in_place()
    assert np.array_equal(h.to_rns(), a)


@pytest.mark.parametrize("n,depth,bits,bb", [BIG[0], BIG[2], BIG[4], BIG[6], BIG[7], BIG[8], BIG[10]])
@pytest.mark.parametrize("small", [False, True])
def test_big_ring_decompose_and_mul_decompose_match_oracle(mx, n, depth, bits, bb, small):
    params, plan = setup(mx, n, depth, bits, bb)
    Mh = plan.uniform(1, 2, seed=8 + depth)
    gM = G(mx, params, Mh)
    dec = gM.small_decompose() if small else gM.decompose()
    assert np.array_equal(dec.to_rns(), plan.ntt(plan.decompose(plan.intt(Mh), bb, small)))
    ell = plan.ell(bb, small)
    S = plan.uniform(2, ell, seed=10)
    want = plan.mul_decompose(S, Mh, bb, small)
    gS = G(mx, params, S)
    assert np.array_equal(gS.mul_decompose(gM, fused=True, small=small).to_rns(), want)
    assert np.array_equal(gS.mul_decompose(gM, fused=False, small=small).to_rns(), want)


CLUSTER_OFF_SCRIPT = """
import sys
import numpy as np
sys.path.insert(0, {root!r})
import mxx_b200 as mx
from oracle import oracle as orc
for n, depth, bits, bb in [(1 << 15, 2, 28, 14), (1 << 16, 2, 28, 14), (1 << 14, 2, 51, 17), (1 << 15, 2, 51, 17), (1 << 16, 1, 59, 20)]:
    moduli = mx.standard_moduli(n, depth, bits)
    params, plan = mx.GpuDCRTPolyParams(n, moduli, bb), orc.Plan(n, moduli)
    a = plan.uniform(1, 2, seed=3)
    g = mx.GpuDCRTPolyMatrix.from_rns(params, a, False)
    g.ntt_in_place()
    assert np.array_equal(g.to_rns(), plan.ntt(a))
    g.intt_all_in_place()
    assert np.array_equal(g.to_rns(), a)
    ell = plan.ell(bb, False)
    S, other = plan.uniform(1, ell, seed=10), plan.uniform(1, 2, seed=11)
    got = mx.GpuDCRTPolyMatrix.from_rns(params, S, True).mul_decompose(mx.GpuDCRTPolyMatrix.from_rns(params, other, True), fused=True)
    assert np.array_equal(got.to_rns(), plan.mul_decompose(S, other, bb, False))
print("cluster knob ok")
"""


@pytest.mark.parametrize("knobs", [{"MXX_NTT_CLUSTER": "0"}, {"MXX_NTT_CLUSTER": "0", "MXX_NTT_F64": "0"}, {"MXX_NTT_F64": "0"}])
def test_cluster_knob_keeps_results(knobs):
    """MXX_NTT_CLUSTER=0 sends rings above one block back through global passes + block kernels: same results."""
    env = dict(os.environ, **knobs)
    out = subprocess.run([sys.executable, "-c", CLUSTER_OFF_SCRIPT.format(root=ROOT)], env=env, capture_output=True, text=True, timeout=900)
    assert out.returncode == 0 and "cluster knob ok" in out.stdout, out.stderr[-2000:]


# ---- host threads -------------------------------------------------------------------------------------------------
def _run_threads(fns):
    errs = []

    def wrap(f):
        def g():
            try:
                f()
            except BaseException as e:  # noqa: BLE001 - reported below
                errs.append(e)
        return g

    ts = [threading.Thread(target=wrap(f)) for f in fns]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    if errs:
        raise errs[0]


def test_two_host_threads_one_context(mx):
    """one thread runs the fused mul_decompose (auxiliary streams + fork/join events of the context) while another loads
    host data into other matrices of the SAME context and transforms them; 20 rounds, every result checked"""
    params, plan = setup(mx, 4096, 3, 24, 12)
    ell = plan.ell(12)
    S, other = plan.uniform(2, ell, seed=41), plan.uniform(1, 6, seed=42)
    want_md = plan.mul_decompose(S, other, 12)
    a = plan.uniform(2, 5, seed=43)
    want_ntt = plan.ntt(a)
    gS, gO = G(mx, params, S), G(mx, params, other)
    rounds = 20

    def worker_md():
        for _ in range(rounds):
            assert np.array_equal(gS.mul_decompose(gO, fused=True).to_rns(), want_md)

    def worker_load():
        for _ in range(rounds):
            m = G(mx, params, a, is_ntt=False)
            m.ntt_in_place()
            assert np.array_equal(m.to_rns(), want_ntt)
            assert np.array_equal((m + m).to_rns(), plan.add(want_ntt, want_ntt))

    _run_threads([worker_md, worker_load, worker_md])


def test_two_contexts_different_moduli_one_device(mx):
    """the reference keeps its NTT constants in device-global __constant__ arrays (cuda/src/matrix/MatrixNTT.cu:1-5), so two
    live contexts with different moduli on one device corrupt each other; here the tables are per context"""
    n = 2048
    p1, plan1 = setup(mx, n, 2, 28, 14)
    p2, plan2 = setup(mx, n, 2, 51, 17)
    p3, plan3 = setup(mx, n, 3, 24, 12)
    a1, a2, a3 = plan1.uniform(1, 4, seed=51), plan2.uniform(1, 4, seed=52), plan3.uniform(1, 4, seed=53)
    w1, w2, w3 = plan1.ntt(a1), plan2.ntt(a2), plan3.ntt(a3)
    c = np.ascontiguousarray
    m1 = plan1.mul(c(w1[:, :2]), c(w1.reshape(4, 1, 2, n)[:2]))
    m2 = plan2.mul(c(w2[:, :2]), c(w2.reshape(4, 1, 2, n)[:2]))
    m3 = None

    def worker(params, plan, a, w, prod):
        def run():
            for _ in range(20):
                g = G(mx, params, a, is_ntt=False)
                g.ntt_in_place()
                assert np.array_equal(g.to_rns(), w)
                if prod is not None:
                    lhs = G(mx, params, np.ascontiguousarray(w[:, :2]))
                    rhs = G(mx, params, np.ascontiguousarray(w.reshape(4, 1, w.shape[2], n)[:2]))
                    assert np.array_equal((lhs * rhs).to_rns(), prod)
                g.intt_all_in_place()
                assert np.array_equal(g.to_rns(), a)
        return run

    _run_threads([worker(p1, plan1, a1, w1, m1), worker(p2, plan2, a2, w2, m2), worker(p3, plan3, a3, w3, m3)])


# ---- input range -----------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("n,depth,bits,bb", [(1024, 2, 28, 14), (4096, 2, 51, 17), (1024, 2, 61, 20)])
def test_non_canonical_words_are_reduced_on_load(mx, n, depth, bits, bb):
    """host words >= q (but inside the storage word) behave as their residues: load reduces them, so the FP64-pipe and
    lazy integer kernels that assume canonical operands stay exact (the reference reduces with % inside every op)"""
    params, plan = setup(mx, n, depth, bits, bb)
    mod = np.array(plan.moduli, dtype=np.uint64)
    a = plan.uniform(1, 3, seed=61)
    word_max = np.uint64((1 << 32) - 1) if max(plan.moduli) < (1 << 30) else np.uint64((1 << 64) - 1)
    raw = a.copy()
    k = (word_max - raw) // mod[None, None, :, None]        # largest multiple of q that still fits the word
    raw = raw + (k // np.uint64(2)) * mod[None, None, :, None]
    raw[0, 1, :, 0] = word_max
    want = raw % mod[None, None, :, None]
    for is_ntt in (False, True):
        g = G(mx, params, raw, is_ntt=is_ntt)
        assert np.array_equal(g.to_rns(), want)
    g = G(mx, params, raw, is_ntt=False)
    g.ntt_in_place()
    assert np.array_equal(g.to_rns(), plan.ntt(want))
    e = G(mx, params, raw, is_ntt=True)
    assert np.array_equal((e * G(mx, params, np.ascontiguousarray(want.reshape(3, 1, depth, n)))).to_rns(),
                          plan.mul(want, np.ascontiguousarray(want.reshape(3, 1, depth, n))))


def test_fill_gadget_reduced_level_mixed_widths(mx):
    """a context whose LAST modulus is the widest: a level-0 gadget takes its digit count from the active limbs only"""
    n, bb = 256, 8
    moduli = orc.prime_chain(n, 20, 1) + orc.prime_chain(n, 28, 1)   # 20-bit first, 28-bit second
    params = mx.GpuDCRTPolyParams(n, moduli, bb)
    sub = orc.Plan(n, moduli[:1])
    dpt_active = -(-20 // bb)  # 3 digits for the 20-bit limb (the whole context would give ceil(28/8) = 4)
    from mxx_b200.ffi import check_status

    for small in (False, True):
        rows = 2
        out = mx.GpuDCRTPolyMatrix(params, rows, rows * dpt_active, level=0)
        fn = out._L.gpu_matrix_fill_small_gadget if small else out._L.gpu_matrix_fill_gadget
        check_status(fn(out.raw, bb), "gpu_matrix_fill_gadget", out._L)
        out.is_ntt = True
        assert np.array_equal(out.to_rns(), sub.ntt(sub.gadget(rows, bb, small=small)))
        bad = mx.GpuDCRTPolyMatrix(params, rows, rows * 4, level=0)
        assert fn(bad.raw, bb) != 0  # the context-wide digit count is NOT the accepted shape at this level


def test_one_row_product_wider_than_grid_y(mx):
    """1 x k times k x c with c > 65535 and c % 6 != 0 (one column per thread block row): launched in slabs"""
    n = 16
    moduli = orc.prime_chain(n, 24, 1)
    params, plan = mx.GpuDCRTPolyParams(n, moduli, 12), orc.Plan(n, moduli)
    c = 65535 + 12
    A, B = plan.uniform(1, 2, seed=71), plan.uniform(2, c, seed=72)
    assert np.array_equal((G(mx, params, A) * G(mx, params, B)).to_rns(), plan.mul(A, B))
