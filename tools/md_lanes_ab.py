"""A/B of the column-chunk policy of the fused mul_decompose (one chunk vs two lanes on auxiliary streams) on one GPU, for
the shapes that matter: C3-a, one C4 gate (all 8 limbs) and the per-rank share of a limb-sharded C4 gate on 8 / 4 GPUs
(1 / 2 limbs, gathered source).  MXX_MUL_DECOMPOSE_LANES=1|2 selects the policy (fresh process per setting)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mxx_b200 import GpuDCRTPolyMatrix as M, GpuDCRTPolyParams, standard_moduli, ffi  # noqa: E402


def timed(params, fn, reps=30):
    for _ in range(5):
        fn()
    params.synchronize()
    t = params.timer()
    t.start()
    for _ in range(reps):
        fn()
    t.stop()
    return t.elapsed_ms() / reps


def main():
    L = ffi.lib()
    tag = os.environ.get("MXX_MUL_DECOMPOSE_LANES", "default")
    # C3-a
    n = 1 << 14
    p = GpuDCRTPolyParams(n, standard_moduli(n, 15, 24), 12)
    S, O = M.sample_distribution(p, 1, 30, 0, 0.0, [1, 1, 1, 1]), M.sample_distribution(p, 1, 30, 0, 0.0, [2, 2, 2, 2])
    out = M(p, 1, 30)
    ms = timed(p, lambda: ffi.check_status(L.gpu_matrix_mul_decompose(out.raw, S.raw, O.raw, 12, 0), "md"))
    print(f"[lanes {tag}] C3-a fused mul_decompose: {ms:.4f} ms")
    del S, O, out, p
    # one C4 gate, all limbs
    n = 1 << 15
    mod = standard_moduli(n, 8, 51)
    p = GpuDCRTPolyParams(n, mod, 17)
    S, O = M.sample_distribution(p, 2, 24, 0, 0.0, [3, 1, 1, 1]), M.sample_distribution(p, 1, 24, 0, 0.0, [4, 2, 2, 2])
    out = M(p, 2, 24)
    ms = timed(p, lambda: ffi.check_status(L.gpu_matrix_mul_decompose(out.raw, S.raw, O.raw, 17, 0), "md"))
    print(f"[lanes {tag}] C4 gate product, 8 limbs on one GPU: {ms:.4f} ms")
    del S, O, out, p
    for name, n, depth, bits, bb, r, d, c in (("C3-b", 1 << 14, 15, 24, 12, 4, 4, 120), ("C3-c", 1 << 14, 4, 51, 17, 2, 2, 24),
                                              ("C4 28-bit gate", 1 << 15, 16, 28, 14, 2, 1, 32), ("u32 2^16 L=4 d=1 c=16", 1 << 16, 4, 28, 14, 1, 1, 16),
                                              ("u32 2^13 L=8 c=64", 1 << 13, 8, 24, 12, 1, 1, 64)):
        p = GpuDCRTPolyParams(n, standard_moduli(n, depth, bits), bb)
        ell = p.modulus_digits()
        S, O = M.sample_distribution(p, r, d * ell, 0, 0.0, [7, 1, 1, 1]), M.sample_distribution(p, d, c, 0, 0.0, [8, 2, 2, 2])
        out = M(p, r, c)
        ms = timed(p, lambda: ffi.check_status(L.gpu_matrix_mul_decompose(out.raw, S.raw, O.raw, bb, 0), "md"), reps=10)
        print(f"[lanes {tag}] {name}: {ms:.4f} ms")
        del S, O, out, p
    n = 1 << 15
    for tpg in (1, 2):  # the per-rank share on 8 / 4 GPUs
        p = GpuDCRTPolyParams(n, mod[:tpg], 17)
        S = M.sample_distribution(p, 2, 24, 0, 0.0, [5, 1, 1, 1])
        gathered = M.sample_distribution(p, 8 // tpg, 24, 0, 0.0, [6, 2, 2, 2])
        gathered.intt_all_in_place()
        ms = timed(p, lambda: S.mul_decompose_gathered(gathered, 1, mod, tpg))
        print(f"[lanes {tag}] limb-sharded C4 gate product, {tpg} limb(s) per rank: {ms:.4f} ms")
        del S, gathered, p


if __name__ == "__main__":
    main()
