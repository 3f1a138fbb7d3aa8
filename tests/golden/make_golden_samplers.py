"""Generates tests/golden/ref_cuda_sampler_vectors.npz from the REFERENCE's own CUDA library
(oracle/_ref/libgpupoly_ref.so, built by oracle/Makefile from /root/reference/cuda where the
sources lie).  Must run on a machine with a GPU:

    gpurun -- 'python tests/golden/make_golden_samplers.py gpurun_out/ref_cuda_sampler_vectors.npz'

For fixed GpuRngSeeds it records what the reference library returns through its C ABI for
gpu_matrix_sample_distribution(_columns) (all four distributions), gpu_matrix_gauss_samp_gq_arb_base
and gpu_matrix_create_p1_covariance_cache + gpu_matrix_sample_p1_full_cached (d = 2: register
scratch, d = 5: global scratch).  `run_case(None, ...)` runs the same calls on libgpupoly.so, which
is how tests/test_samplers.py compares the two sample for sample.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

CASES = [  # (n, depth, bits, base_bits)
    (128, 2, 16, 8),   # the reference tests' parameters (src/sampler/trapdoor/gpu.rs:498-500)
    (64, 3, 24, 12),
    (64, 2, 51, 17),
    (32, 2, 28, 1),    # 28 digits per tower: the 64-digit code path
]
SIGMA = 4.578


def run_case(lib, n, depth, bits, bb, seed0=0x6D7878):
    from mxx_b200 import GpuDCRTPolyMatrix as M, GpuDCRTPolyParams
    from mxx_b200.matrix import DIST_BIT, DIST_GAUSS, DIST_TERNARY, DIST_UNIFORM
    from oracle import oracle as orc

    moduli = orc.prime_chain(n, bits, depth)
    params = GpuDCRTPolyParams(n, moduli, bb, library=lib)
    plan = orc.Plan(n, moduli)
    seed = [seed0, seed0 ^ 0x1111, 0x2222, 0x3333]
    out = {"seed": np.array(seed, dtype=np.uint64)}
    for name, dist, sigma in (("uniform", DIST_UNIFORM, 0.0), ("gauss", DIST_GAUSS, SIGMA), ("gauss_big", DIST_GAUSS, 3.1e7),
                              ("bit", DIST_BIT, 0.0), ("ternary", DIST_TERNARY, 0.0)):
        out[name] = M.sample_distribution(params, 2, 3, dist, sigma, seed).to_rns()
        out[name + "_cols"] = M.sample_distribution_columns(params, 2, 5, 1, 3, dist, sigma, seed).to_rns()
    base = 1 << bb
    c = (base + 1.0) * SIGMA
    src = plan.uniform(2, 2, seed=seed0 + 5)
    out["gauss_src"] = src
    out["gauss_samp"] = M.from_rns(params, src, False).gauss_samp_gq_arb_base(c, SIGMA, seed).to_rns()
    k = params.modulus_digits()
    for d in (2, 5):
        r = M.sample_distribution(params, d, d * k, DIST_GAUSS, SIGMA, [seed0, 1, 0, d])
        e = M.sample_distribution(params, d, d * k, DIST_GAUSS, SIGMA, [seed0, 2, 0, d])
        a = (r * r.transpose()).into_coeff_domain()
        b = (r * e.transpose()).into_coeff_domain()
        dm = (e * e.transpose()).into_coeff_domain()
        s = 1.8 * (base + 1.0) * SIGMA * SIGMA * ((d * n * k) ** 0.5 + (2 * n) ** 0.5 + 4.7)
        cache = M.create_p1_covariance_cache(a, b, dm, c, s, SIGMA)
        p2 = M.sample_distribution(params, d * k, d, DIST_GAUSS, (s * s - c * c) ** 0.5, [seed0, 3, 0, d])
        tp2 = r.concat_rows([e]) * p2
        out[f"p1_d{d}"] = M.sample_p1_full_cached(cache, tp2, [seed0, 4, 0, d]).to_rns()
    return out


def main():
    from mxx_b200 import ffi

    dst = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "tests", "golden", "ref_cuda_sampler_vectors.npz")
    lib = ffi.bind(os.path.join(ROOT, "oracle", "_ref", "libgpupoly_ref.so"), extensions=False)
    blob = {}
    for (n, depth, bits, bb) in CASES:
        for k, v in run_case(lib, n, depth, bits, bb).items():
            blob[f"n{n}_d{depth}_b{bits}_bb{bb}/{k}"] = v
    np.savez_compressed(dst, **blob)
    print("wrote", dst, len(blob), "arrays", os.path.getsize(dst), "bytes")


if __name__ == "__main__":
    main()
