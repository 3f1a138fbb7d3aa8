"""One Gaussian sample_distribution call (for ncu captures of the sampler kernels).  GPU box only."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mxx_b200 import GpuDCRTPolyMatrix as M, GpuDCRTPolyParams, ffi, standard_moduli  # noqa: E402

n, depth = 1 << 14, 10
params = GpuDCRTPolyParams(n, standard_moduli(n, depth, 24), 12)
polys = int(sys.argv[1]) if len(sys.argv) > 1 else 100
for rep in range(3):
    ffi.lib().gpu_device_synchronize()
    t0 = time.perf_counter()
    m = M.sample_distribution(params, 1, polys, 1, 1.2e8, [rep, 2, 3, 4])
    ffi.lib().gpu_device_synchronize()
    print("gauss %d polys: %.3f ms" % (polys, (time.perf_counter() - t0) * 1e3), flush=True)
