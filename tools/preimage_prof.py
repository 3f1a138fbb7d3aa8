"""Step-by-step wall-clock breakdown (device-synchronised) of the C5 trapdoor preimage, ours or the reference library.
GPU box only:  gpurun -- 'python tools/preimage_prof.py [ref]'"""
import math
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mxx_b200 import GpuDCRTPolyMatrix as M, GpuDCRTPolyParams, ffi, standard_moduli  # noqa: E402
from mxx_b200.trapdoor import GpuDCRTPolyTrapdoorSampler, derive_seed, preimage_smoothing_parameter  # noqa: E402

lib = ffi.bind(os.path.join(ROOT, "oracle", "_ref", "libgpupoly_ref.so"), extensions=False) if len(sys.argv) > 1 else None
L = lib or ffi.lib()
n, depth, bits, bb, d, cols, sigma = 1 << 14, 10, 24, 12, 1, 50, 4.578
params = GpuDCRTPolyParams(n, standard_moduli(n, depth, bits), bb, library=lib)
sampler = GpuDCRTPolyTrapdoorSampler(params, sigma)
td, a = sampler.trapdoor(params, d, seed=[1, 2, 3, 4])
u = M.sample_distribution(params, d, cols, 0, 0.0, [5, 6, 7, 8])
k = params.modulus_digits()
s = preimage_smoothing_parameter(1 << bb, sigma, d, n, k)
sl = math.sqrt(s * s - sampler.c ** 2)


def tick(label, fn, acc):
    L.gpu_device_synchronize()
    t0 = time.perf_counter()
    out = fn()
    L.gpu_device_synchronize()
    acc.append((label, (time.perf_counter() - t0) * 1e3))
    return out


for rep in range(3):
    acc = []
    seed = [20 + rep, 0, 0, 0]
    p2 = tick("p2 = Gauss(dk x cols, sigma_large)", lambda: M.sample_distribution(params, d * k, cols, 1, sl, derive_seed(seed, 20)), acc)
    tp2 = tick("tp2 = [R;E] * p2", lambda: td.r.concat_rows([td.e]) * p2, acc)
    cache = tick("p1 covariance cache", lambda: td.p1_covariance_cache(sampler.c, s, sigma), acc)
    p1 = tick("p1 = sample_p1_full_cached", lambda: M.sample_p1_full_cached(cache, tp2, derive_seed(seed, 21)), acc)
    img = tick("A_left*p1 + A_right*p2", lambda: a.slice(0, d, 0, 2 * d) * p1 + a.slice(0, d, 2 * d, 2 * d + d * k) * p2, acc)
    z = tick("gauss_samp_gq_arb_base(u - image)", lambda: (u - img).gauss_samp_gq_arb_base(sampler.c, sigma, derive_seed(seed, 22)), acc)
    rz = tick("R*z, E*z", lambda: (td.r * z, td.e * z), acc)
    tot = sum(t for _, t in acc)
    print(f"--- rep {rep}: total {tot:.2f} ms ({'reference' if lib else 'ours'})")
    for label, t in acc:
        print(f"  {t:9.3f} ms  {label}")
