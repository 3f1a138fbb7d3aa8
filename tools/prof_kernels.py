"""Short driver for ncu captures: a few launches of each hot kernel at bench sizes.

    python tools/prof_kernels.py [ntt32|ntt64|md|mul|all]
"""
import os
import sys

import numpy as np

REPS = int(os.environ.get("PROF_REPS", "3"))
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import uniform_host  # noqa: E402
from mxx_b200 import GpuDCRTPolyMatrix, GpuDCRTPolyParams, ffi, standard_moduli  # noqa: E402


def ntt(n, depth, bits, polys, reps=REPS):
    moduli = standard_moduli(n, depth, bits)
    params = GpuDCRTPolyParams(n, moduli, 14)
    host = np.empty((1, polys, depth, n), dtype=np.uint64)
    uniform_host(host, moduli, seed=1)
    m = GpuDCRTPolyMatrix.from_rns(params, host, is_ntt=False)
    t = params.timer()
    for _ in range(reps):
        t.start()
        m.ntt_in_place()
        t.stop()
        f = t.elapsed_ms()
        t.start()
        m.intt_all_in_place()
        t.stop()
        i = t.elapsed_ms()
    coeffs = polys * depth * n
    print(f"ntt n={n} L={depth} bits={bits} polys={polys}: fwd {f:.3f} ms ({coeffs / f / 1e6:.1f} Gcoeff/s) inv {i:.3f} ms ({coeffs / i / 1e6:.1f} Gcoeff/s)")
    assert np.array_equal(m.to_rns(), host)


def md(reps=REPS):
    n, depth, bits, bb = 1 << 14, 15, 24, 12
    moduli = standard_moduli(n, depth, bits)
    params = GpuDCRTPolyParams(n, moduli, bb)
    S = np.empty((1, 30, depth, n), dtype=np.uint64)
    O = np.empty((1, 30, depth, n), dtype=np.uint64)
    uniform_host(S, moduli, 2)
    uniform_host(O, moduli, 3)
    gS, gO = GpuDCRTPolyMatrix.from_rns(params, S), GpuDCRTPolyMatrix.from_rns(params, O)
    t = params.timer()
    for _ in range(reps):
        t.start()
        out = gS.mul_decompose(gO, fused=True)
        t.stop()
        ms = t.elapsed_ms()
    print(f"mul_decompose C3-a fused: {ms:.3f} ms")
    return out


def c4(reps=REPS):
    """one C4 gate's fused mul_decompose: (2 x 24) * G^-1(1 x 24) at n = 2^15, L = 8 x 51-bit, base 2^17"""
    n, depth, bits, bb = 1 << 15, 8, 51, 17
    moduli = standard_moduli(n, depth, bits)
    params = GpuDCRTPolyParams(n, moduli, bb)
    ell = params.modulus_digits()
    S = np.empty((2, ell, depth, n), dtype=np.uint64)
    O = np.empty((1, ell, depth, n), dtype=np.uint64)
    uniform_host(S, moduli, 2)
    uniform_host(O, moduli, 3)
    gS, gO = GpuDCRTPolyMatrix.from_rns(params, S), GpuDCRTPolyMatrix.from_rns(params, O)
    t = params.timer()
    for _ in range(reps):
        t.start()
        out = gS.mul_decompose(gO, fused=True)
        t.stop()
        ms = t.elapsed_ms()
    print(f"mul_decompose C4 gate fused: {ms:.3f} ms")
    return out


def mul(reps=REPS):
    n, depth, bits = 1 << 14, 15, 24
    moduli = standard_moduli(n, depth, bits)
    params = GpuDCRTPolyParams(n, moduli, 12)
    A = np.empty((1, 30, depth, n), dtype=np.uint64)
    B = np.empty((30, 120, depth, n), dtype=np.uint64)
    uniform_host(A, moduli, 4)
    uniform_host(B, moduli, 5)
    gA, gB = GpuDCRTPolyMatrix.from_rns(params, A), GpuDCRTPolyMatrix.from_rns(params, B)
    t = params.timer()
    for _ in range(reps):
        t.start()
        out = gA * gB
        t.stop()
        ms = t.elapsed_ms()
    print(f"mul C1-R: {ms:.3f} ms")
    return out


if __name__ == "__main__":
    what = sys.argv[1] if len(sys.argv) > 1 else "all"
    if what in ("ntt32", "all"):
        ntt(1 << 14, 16, 28, 256)
    if what in ("ntt64", "all"):
        ntt(1 << 14, 8, 51, 256)
    if what in ("ntt15",):
        ntt(1 << 15, 8, 28, 256)
    if what in ("ntt15w",):
        ntt(1 << 15, 8, 51, 128)
    if what in ("ntt16w",):
        ntt(1 << 16, 4, 51, 128)
    if what in ("ntt16", "all"):
        ntt(1 << 16, 8, 28, 128)
    if what in ("md", "all"):
        md()
    if what in ("c4",):
        c4()
    if what in ("mul", "all"):
        mul()
    ffi.device_synchronize()
