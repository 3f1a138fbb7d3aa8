"""One fused mul_decompose of a named shape, twice (warm-up + the call to profile):
   ncu --launch-skip <launches of one call> ... python tools/md_once.py c3b"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mxx_b200 import GpuDCRTPolyMatrix as M, GpuDCRTPolyParams, standard_moduli, ffi  # noqa: E402

SHAPES = {"c3a": (1 << 14, 15, 24, 12, 1, 1, 30), "c3b": (1 << 14, 15, 24, 12, 4, 4, 120), "c3c": (1 << 14, 4, 51, 17, 2, 2, 24),
          "c4": (1 << 15, 8, 51, 17, 2, 1, 24), "c4_28": (1 << 15, 16, 28, 14, 2, 1, 32)}
n, depth, bits, bb, r, d, c = SHAPES[sys.argv[1] if len(sys.argv) > 1 else "c3b"]
L = ffi.lib()
p = GpuDCRTPolyParams(n, standard_moduli(n, depth, bits), bb)
ell = p.modulus_digits()
S, O = M.sample_distribution(p, r, d * ell, 0, 0.0, [7, 1, 1, 1]), M.sample_distribution(p, d, c, 0, 0.0, [8, 2, 2, 2])
out = M(p, r, c)
p.synchronize()
before = L.gpu_kernel_launch_count()
for i in range(2):
    ffi.check_status(L.gpu_matrix_mul_decompose(out.raw, S.raw, O.raw, bb, 0), "md")
    p.synchronize()
    if i == 0:
        print("launches per call:", L.gpu_kernel_launch_count() - before, "launches before the first call:", before)
