"""Trapdoor generation and Gaussian preimage sampling over the gpu_* C ABI — host mirror of
src/sampler/trapdoor/gpu.rs (GpuDCRTTrapdoor :47-128, GpuDCRTPolyTrapdoorSampler :172-421,
sample_pert_square_mat_gpu_native_parts :423-474) and of the uniform sampler
src/sampler/gpu.rs:138-250.

Every random integer is drawn on the device (ChaCha streams keyed by a GpuRngSeed); this
module only sequences ABI calls, exactly like the Rust side.  Seeds are explicit arguments
(the Rust side draws them from the OS RNG, gpu.rs:138-142) so runs are reproducible.
"""
from __future__ import annotations

import math
import os

from .matrix import DIST_BIT, DIST_GAUSS, DIST_TERNARY, DIST_UNIFORM, GpuDCRTPolyMatrix

SPECTRAL_CONSTANT = 1.8  # src/sampler/trapdoor/gpu.rs:15


def random_gpu_rng_seed() -> bytes:  # src/sampler/gpu.rs:138-142
    return os.urandom(32)


def derive_seed(seed, tag: int) -> list[int]:
    """Independent sub-seeds from one caller seed: word 3 carries the call-site tag."""
    from .matrix import as_seed

    s = as_seed(seed)
    return [s.words[0], s.words[1], s.words[2], (s.words[3] + 0x9E3779B97F4A7C15 * (tag + 1)) & (2**64 - 1)]


def preimage_c(base: int, sigma: float) -> float:  # trapdoor/gpu.rs:17-19
    return (base + 1.0) * sigma


def preimage_smoothing_parameter(base: int, sigma: float, d: int, n: int, k: int) -> float:  # trapdoor/gpu.rs:21-27
    return SPECTRAL_CONSTANT * (base + 1.0) * sigma * sigma * (math.sqrt(d * n * k) + math.sqrt(2 * n) + 4.7)


class GpuDCRTPolyUniformSampler:
    """src/sampler/gpu.rs:17-55: FinRing / Gauss / Bit / Ternary matrices, returned in the evaluation domain."""

    FIN_RING, GAUSS, BIT, TERNARY = DIST_UNIFORM, DIST_GAUSS, DIST_BIT, DIST_TERNARY

    def sample_uniform(self, params, nrow, ncol, dist: int, sigma: float = 0.0, seed=None):
        return GpuDCRTPolyMatrix.sample_distribution(params, nrow, ncol, dist, sigma, seed if seed is not None else random_gpu_rng_seed())

    def sample_uniform_columns(self, params, nrow, total_ncol, col_start, col_len, dist: int, sigma: float = 0.0, seed=None):
        """src/sampler/gpu.rs:197-250: a column range of the matrix the seed defines (column-sharded generation)."""
        return GpuDCRTPolyMatrix.sample_distribution_columns(params, nrow, total_ncol, col_start, col_len, dist, sigma,
                                                             seed if seed is not None else random_gpu_rng_seed())


class GpuDCRTTrapdoor:
    """trapdoor/gpu.rs:47-80: R, E ~ D_sigma (d x d*k) and the coefficient-form Gram blocks A = R R^T, B = R E^T, D = E E^T."""

    def __init__(self, params, size: int, sigma: float, seed=None, r=None, e=None):
        seed = seed if seed is not None else random_gpu_rng_seed()
        k = params.modulus_digits()
        us = GpuDCRTPolyUniformSampler()
        self.r = r if r is not None else us.sample_uniform(params, size, size * k, DIST_GAUSS, sigma, derive_seed(seed, 0))
        self.e = e if e is not None else us.sample_uniform(params, size, size * k, DIST_GAUSS, sigma, derive_seed(seed, 1))
        rt, et = self.r.transpose(), self.e.transpose()
        self.a_mat_coeff = (self.r * rt).into_coeff_domain()
        self.b_mat_coeff = (self.r * et).into_coeff_domain()
        self.d_mat_coeff = (self.e * et).into_coeff_domain()
        self._cache = None  # (c, s, dgg_stddev, GpuP1CovarianceCache), trapdoor/gpu.rs:145-169

    def p1_covariance_cache(self, c: float, s: float, dgg_stddev: float):
        if self._cache is not None and self._cache[:3] == (c, s, dgg_stddev):
            return self._cache[3]
        cache = GpuDCRTPolyMatrix.create_p1_covariance_cache(self.a_mat_coeff, self.b_mat_coeff, self.d_mat_coeff, c, s, dgg_stddev)
        self._cache = (c, s, dgg_stddev, cache)
        return cache

    def to_compact_bytes(self) -> bytes:  # trapdoor/gpu.rs:82-97
        out = b""
        for m in (self.r, self.e):
            b = m.to_compact_bytes()
            out += len(b).to_bytes(8, "little") + b
        return out

    @classmethod
    def from_compact_bytes(cls, params, data: bytes, sigma: float = 0.0):  # trapdoor/gpu.rs:99-128
        mats, off = [], 0
        for _ in range(2):
            if off + 8 > len(data):
                return None
            ln = int.from_bytes(data[off:off + 8], "little")
            off += 8
            if off + ln > len(data):
                return None
            mats.append(GpuDCRTPolyMatrix.from_compact_bytes(params, data[off:off + ln]))
            off += ln
        if off != len(data):
            return None
        return cls(params, mats[0].nrow, sigma, r=mats[0], e=mats[1])


class GpuDCRTPolyTrapdoorSampler:
    """trapdoor/gpu.rs:172-421."""

    def __init__(self, params, sigma: float):
        self.sigma = float(sigma)
        self.base = 1 << params.base_bits
        self.c = preimage_c(self.base, self.sigma)

    def trapdoor(self, params, size: int, seed=None):  # :197-211: A = [A_bar | I | G - (A_bar R + E)]
        seed = seed if seed is not None else random_gpu_rng_seed()
        td = GpuDCRTTrapdoor(params, size, self.sigma, derive_seed(seed, 10))
        a_bar = GpuDCRTPolyUniformSampler().sample_uniform(params, size, size, DIST_UNIFORM, 0.0, derive_seed(seed, 11))
        g = GpuDCRTPolyMatrix.gadget_matrix(params, size)
        a0 = a_bar.concat_columns([GpuDCRTPolyMatrix.identity(params, size)])
        a1 = g - (a_bar * td.r + td.e)
        return td, a0.concat_columns([a1])

    def sample_perturbation(self, params, trapdoor, s, sigma_large, total_ncol, seed):  # :423-474
        d, dk = trapdoor.r.nrow, trapdoor.r.ncol
        padded = -(-total_ncol // d) * d
        p2 = GpuDCRTPolyUniformSampler().sample_uniform(params, dk, padded, DIST_GAUSS, sigma_large, derive_seed(seed, 20))
        tp2 = trapdoor.r.concat_rows([trapdoor.e]) * p2
        cache = trapdoor.p1_covariance_cache(self.c, s, self.sigma)
        p1 = GpuDCRTPolyMatrix.sample_p1_full_cached(cache, tp2, derive_seed(seed, 21))
        return p1, p2

    def preimage(self, params, trapdoor, public_matrix, target, seed=None):  # :224-369
        seed = seed if seed is not None else random_gpu_rng_seed()
        d, cols = public_matrix.nrow, target.ncol
        assert target.nrow == d, "Target matrix should have the same number of rows as the public matrix"
        n, k = params.ring_dimension, params.modulus_digits()
        s = preimage_smoothing_parameter(self.base, self.sigma, d, n, k)
        sigma_large = math.sqrt(s * s - self.c * self.c)
        p1, p2 = self.sample_perturbation(params, trapdoor, s, sigma_large, cols, seed)
        r1, r2 = p1.nrow, p2.nrow
        assert public_matrix.ncol == r1 + r2, "public matrix columns must match perturbation rows"
        image = public_matrix.slice(0, d, 0, r1) * p1 + public_matrix.slice(0, d, r1, r1 + r2) * p2
        if image.ncol != cols:
            image = image.slice_columns(0, cols)
        z_hat = (target - image).gauss_samp_gq_arb_base(self.c, self.sigma, derive_seed(seed, 22))
        out = GpuDCRTPolyMatrix(params, r1 + r2, cols, p1.level, p1.is_ntt)
        out.copy_block_from(p1, 0, 0, 0, 0, r1, cols)
        out.copy_block_from(p2, r1, 0, 0, 0, r2, cols)
        rz = trapdoor.r * z_hat
        out.add_block_from(rz, 0, 0, 0, 0, rz.nrow, cols)
        ez = trapdoor.e * z_hat
        out.add_block_from(ez, d, 0, 0, 0, ez.nrow, cols)
        out.add_block_from(z_hat, 2 * d, 0, 0, 0, z_hat.nrow, cols)
        return out

    def preimage_extend(self, params, trapdoor, public_matrix, ext_matrix, target, seed=None):  # :398-421
        seed = seed if seed is not None else random_gpu_rng_seed()
        d, n, k = public_matrix.nrow, params.ring_dimension, params.modulus_digits()
        s = preimage_smoothing_parameter(self.base, self.sigma, d, n, k)
        right = GpuDCRTPolyUniformSampler().sample_uniform(params, ext_matrix.ncol, target.ncol, DIST_GAUSS, s, derive_seed(seed, 30))
        left = self.preimage(params, trapdoor, public_matrix, target - ext_matrix * right, derive_seed(seed, 31))
        return left.concat_rows([right])
