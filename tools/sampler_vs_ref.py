"""Side-by-side run of the samplers in libgpupoly.so and in the reference's CUDA library
(oracle/_ref/libgpupoly_ref.so) on the same seeds; reports how many words differ per call.
GPU box only:  gpurun -- 'python tools/sampler_vs_ref.py'"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from mxx_b200 import GpuDCRTPolyMatrix, GpuDCRTPolyParams, ffi  # noqa: E402
from mxx_b200.matrix import DIST_BIT, DIST_GAUSS, DIST_TERNARY, DIST_UNIFORM  # noqa: E402
from oracle import oracle as orc  # noqa: E402

CASES = [(128, 2, 16, 8), (64, 3, 24, 12), (256, 2, 51, 17), (64, 2, 28, 1), (1024, 4, 24, 6)]


def run(lib, n, depth, bits, bb):
    moduli = orc.prime_chain(n, bits, depth)
    params = GpuDCRTPolyParams(n, moduli, bb, library=lib)
    plan = orc.Plan(n, moduli)
    out = {}
    M = GpuDCRTPolyMatrix
    for name, dist, sigma in (("uniform", DIST_UNIFORM, 0.0), ("gauss", DIST_GAUSS, 4.578), ("gauss_big", DIST_GAUSS, 3.1e7),
                              ("bit", DIST_BIT, 0.0), ("ternary", DIST_TERNARY, 0.0)):
        out[name] = M.sample_distribution(params, 2, 3, dist, sigma, [7, 8, 9, 10]).to_rns()
        out[name + "_cols"] = M.sample_distribution_columns(params, 2, 5, 1, 3, dist, sigma, [7, 8, 9, 10]).to_rns()
    base = 1 << bb
    c = (base + 1.0) * 4.578
    src = M.from_rns(params, plan.uniform(2, 2, seed=5), False)
    out["gauss_samp"] = src.gauss_samp_gq_arb_base(c, 4.578, [1, 2, 3, 4]).to_rns()
    # perturbation: d = 2 (registers) and d = 5 (global scratch)
    for d in (2, 5):
        k = params.modulus_digits()
        r = M.sample_distribution(params, d, d * k, DIST_GAUSS, 4.578, [11, 0, 0, d])
        e = M.sample_distribution(params, d, d * k, DIST_GAUSS, 4.578, [12, 0, 0, d])
        a = (r * r.transpose()).into_coeff_domain()
        b = (r * e.transpose()).into_coeff_domain()
        dm = (e * e.transpose()).into_coeff_domain()
        s = 1.8 * (base + 1.0) * 4.578 * 4.578 * ((d * n * k) ** 0.5 + (2 * n) ** 0.5 + 4.7)
        cache = M.create_p1_covariance_cache(a, b, dm, c, s, 4.578)
        p2 = M.sample_distribution(params, d * k, d, DIST_GAUSS, (s * s - c * c) ** 0.5, [13, 0, 0, d])
        tp2 = r.concat_rows([e]) * p2
        out[f"p1_d{d}"] = M.sample_p1_full_cached(cache, tp2, [14, 0, 0, d]).to_rns()
    return out


def main():
    ref = ffi.bind(os.path.join(ROOT, "oracle", "_ref", "libgpupoly_ref.so"), extensions=False)
    bad = 0
    for case in CASES:
        mine, theirs = run(None, *case), run(ref, *case)
        for k in mine:
            diff = int((mine[k] != theirs[k]).sum())
            bad += diff
            print(case, k, mine[k].shape, "differing words:", diff, flush=True)
    print("TOTAL differing words:", bad)


if __name__ == "__main__":
    main()
