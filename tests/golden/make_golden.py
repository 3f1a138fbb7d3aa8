"""Generates tests/golden/ref_cuda_vectors.npz from the REFERENCE's own CUDA library
(oracle/_ref/libgpupoly_ref.so, built by oracle/Makefile from /root/reference/cuda where the
sources lie).  Must run on a machine with a GPU:

    gpurun -- 'python tests/golden/make_golden.py gpurun_out/ref_cuda_vectors.npz'

Inputs are the oracle's seeded uniform matrices; outputs are what the reference library
returns through its C ABI for: NTT, INTT, mul, add, sub, mul_scalar, decompose (full and
small), gadget matrices and the mul_decompose call sequence of
src/matrix/gpu_dcrt_poly.rs:1414-1493.  The committed file pins the oracle (CPU test) and the
new CUDA path (GPU test) to the reference's results.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

CASES = [  # (n, depth, bits, base_bits)
    (4, 2, 17, 1),
    (4, 1, 52, 17),
    (16, 2, 17, 3),
    (64, 3, 24, 12),
    (256, 2, 28, 14),
    (128, 2, 51, 17),
    (64, 2, 59, 20),
]


def run_case(lib, n, depth, bits, bb):
    from mxx_b200 import GpuDCRTPolyMatrix, GpuDCRTPolyParams
    from oracle import oracle as orc

    moduli = orc.prime_chain(n, bits, depth)
    params = GpuDCRTPolyParams(n, moduli, bb, library=lib)
    plan = orc.Plan(n, moduli)
    G = lambda h, ntt=True: GpuDCRTPolyMatrix.from_rns(params, h, ntt)  # noqa: E731
    A, B, C2 = plan.uniform(2, 3, seed=101), plan.uniform(3, 2, seed=102), plan.uniform(2, 3, seed=103)
    s = plan.uniform(1, 1, seed=104)
    out = {"moduli": np.array(moduli, dtype=np.uint64)}
    g = G(A, False)
    g.ntt_in_place()
    out["ntt_A"] = g.to_rns()
    h = G(A, True)
    h.intt_all_in_place()
    out["intt_A"] = h.to_rns()
    out["mul_AB"] = (G(A) * G(B)).to_rns()
    out["add_AC"] = (G(A) + G(C2)).to_rns()
    out["sub_AC"] = (G(A) - G(C2)).to_rns()
    out["mulscalar_As"] = G(A).mul_scalar(G(s)).to_rns()
    out["decompose_A"] = G(A).decompose().to_rns()
    out["small_decompose_A"] = G(A).small_decompose().to_rns()
    out["gadget_2"] = GpuDCRTPolyMatrix.gadget_matrix(params, 2).to_rns()
    out["small_gadget_2"] = GpuDCRTPolyMatrix.small_gadget_matrix(params, 2).to_rns()
    ell = params.modulus_digits()
    S = plan.uniform(2, 2 * ell, seed=105)
    out["mul_decompose_SA"] = G(S).mul_decompose(G(A), fused=False).to_rns()
    Ss = plan.uniform(1, 2 * params.digits_per_tower, seed=106)
    out["mul_decompose_small_SsA"] = G(Ss).mul_decompose_small(G(A), fused=False).to_rns()
    return out


def main():
    from mxx_b200 import ffi

    dst = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "tests", "golden", "ref_cuda_vectors.npz")
    lib = ffi.bind(os.path.join(ROOT, "oracle", "_ref", "libgpupoly_ref.so"), extensions=False)
    blob = {}
    for (n, depth, bits, bb) in CASES:
        for k, v in run_case(lib, n, depth, bits, bb).items():
            blob[f"n{n}_d{depth}_b{bits}_bb{bb}/{k}"] = v
    np.savez_compressed(dst, **blob)
    print("wrote", dst, len(blob), "arrays")


if __name__ == "__main__":
    main()
