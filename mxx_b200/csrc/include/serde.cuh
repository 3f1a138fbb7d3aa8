// serde.cuh — host u64 RNS words <-> device word-typed storage.
//
// Replaces serde_pack_u64_limb_to_packed_kernel / serde_unpack_packed_limb_to_u64_kernel
// (cuda/src/matrix/MatrixSerde.cu:110, 137) and their per-limb staging + 2D-copy loop
// (:566-741, :743-924).  The staging buffer holds the host layout [poly][limb][coeff]
// u64 in natural slot order; the device keeps EVAL data bit-reversed, so the EVAL tag
// applies the bit-reversal here (the only place natural order exists on the device).
#pragma once
#include "modarith.cuh"

namespace gp {

// S = source word type: u64 (the ABI's words) or u32 (words already narrowed on the host, MatrixHostLoad.inl).
// Every word is REDUCED into [0, q_limb) on the way in: the reference stores whatever it is given and reduces with `%`
// in every operation (cuda/src/matrix/MatrixUtils.cu:541-545), the kernels here assume canonical residues (the FP64-pipe
// ones need them), so the load is where non-canonical input is made canonical.  A u32 context keeps the low 32 bits of
// a wider word first (both load lanes agree on that; the reference truncates to the limb's byte width,
// cuda/src/matrix/MatrixSerde.cu:110-135).  `word0` = index of src[0] / dst[0] within the whole matrix (chunked loads).
template <typename T> GP_HD T load_canonical(T x, const LimbConst<T> &c) {
    if constexpr (sizeof(T) == 4) return csub<T>(x - mulhi(x, c.one_s) * c.q, c.q);
    else return red64<T>(x, c);
}
template <typename T, typename S, bool BITREV>
__global__ void pack_kernel(T *__restrict__ dst, const S *__restrict__ src, size_t total, uint32_t logn, size_t word0,
                            uint32_t limbs, const LimbConst<T> *__restrict__ lc) {
    const size_t step = (size_t)gridDim.x * GP_NT;
    const uint32_t mask = (1u << logn) - 1;
    for (size_t i = (size_t)blockIdx.x * GP_NT + GP_TID; i < total; i += step) {
        const uint32_t s = (uint32_t)i & mask;
        const size_t j = BITREV ? ((i & ~(size_t)mask) | (brev32(s) >> (32 - logn))) : i;
        const uint32_t limb = (uint32_t)(((word0 + i) >> logn) % limbs);
        dst[i] = load_canonical<T>((T)src[j], lc[limb]);
    }
}

template <typename T, bool BITREV>
__global__ void unpack_kernel(uint64_t *__restrict__ dst, const T *__restrict__ src, size_t total, uint32_t logn) {
    const size_t step = (size_t)gridDim.x * GP_NT;
    const uint32_t mask = (1u << logn) - 1;
    for (size_t i = (size_t)blockIdx.x * GP_NT + GP_TID; i < total; i += step) {
        const uint32_t s = (uint32_t)i & mask;
        const size_t j = BITREV ? ((i & ~(size_t)mask) | (brev32(s) >> (32 - logn))) : i;
        dst[i] = (uint64_t)src[j];
    }
}

// coefficient 0 of every (poly, limb): out[pl] = data[pl * n]
template <typename T>
__global__ void const_coeff_kernel(uint64_t *__restrict__ dst, const T *__restrict__ src, size_t count, uint32_t logn) {
    const size_t i = (size_t)blockIdx.x * GP_NT + GP_TID;
    if (i < count) dst[i] = (uint64_t)src[i << logn];
}

}  // namespace gp
