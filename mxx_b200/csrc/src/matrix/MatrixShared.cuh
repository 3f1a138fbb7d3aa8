// MatrixShared.cuh — host entry points shared between the translation units of the matrix layer.
// The layer is compiled as several translation units (one per kernel family, so that nvcc runs them in parallel);
// these are the only functions that cross a unit boundary.  T is the storage word of the context (u32 / u64).
#pragma once
#include <algorithm>
#include <atomic>
#include <cstring>
#include <exception>
#include <mutex>
#include <new>

#include "gpupoly_internal.cuh"
#include "arith.cuh"
#include "ntt.cuh"

namespace gpm {
// MatrixNTT.inl: dst = NTT / INTT(src) for `polys` polynomials of `limbs` limbs each (src may equal dst)
template <typename T> int run_ntt_fwd(GpuContext *ctx, const T *src, T *dst, size_t polys, uint32_t limbs, bool first_top_pass_done = false);
template <typename T> int run_ntt_inv(GpuContext *ctx, const T *src, T *dst, size_t polys, uint32_t limbs);
// in-place transform of a whole matrix with the format bookkeeping of gpu_matrix_ntt_all / intt_all
int matrix_transform(GpuMatrix *mat, bool inverse);
// MatrixNTTCluster.inl: the same for rings of 2^lc shared-memory blocks (pl = polys * limbs)
template <typename T> int run_cluster_fwd(GpuContext *ctx, const T *src, T *dst, size_t pl, uint32_t limbs, int lc);
template <typename T> int run_cluster_inv(GpuContext *ctx, const T *src, T *dst, size_t pl, uint32_t limbs, int lc);
// MatrixDecomposeCluster.inl: gadget digits + forward NTT for rings of 2^lc blocks (opl = output polys * limbs)
template <typename T> int run_cluster_decompose(GpuContext *ctx, const T *src, T *out, const gp::DecompGeom &g, size_t opl, int lc);
// MatrixArith.inl: out[:, out_col0 : out_col0+cols] = A(rows x inner) * B[:, b_col0 : b_col0+cols]
template <typename T> int run_matmul(GpuContext *ctx, T *out, const T *A, const T *B, gp::MatmulGeom g);
int binary_check(const char *name, GpuMatrix *out, const GpuMatrix *lhs, const GpuMatrix *rhs);
}  // namespace gpm
using namespace gpm;
