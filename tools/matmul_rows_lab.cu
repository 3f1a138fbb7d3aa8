// matmul_rows_lab.cu — development harness for the row-tiled product (r >= 2 output rows): times one build variant of
// matmul_thread (arith.cuh) on the shape of one column chunk of C3-b (S 4 x 120 times an 18-column digit block,
// 15 x 24-bit, n = 2^14) and checks it against the 4 x 4 half-vector tile.  Results: profiles/r02_matmul_rows_lab.txt.
//   -DLAB_V -DLAB_TR -DLAB_TC -DLAB_UNROLL -DLAB_SL (slot vectors per CTA) -DLAB_CG (column groups per CTA) -DLAB_MINB
// A CTA of SL * CG threads owns SL * V consecutive slots of one limb and CG adjacent column tiles (CG > 1: the groups
// read the same words of A, served from L1 after the first).  Build: tools/build_mmrows.sh "tag:flags" ...
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "arith.cuh"
using namespace gp;
#ifndef LAB_V
#define LAB_V 4
#endif
#ifndef LAB_TR
#define LAB_TR 4
#endif
#ifndef LAB_TC
#define LAB_TC 2
#endif
#ifndef LAB_UNROLL
#define LAB_UNROLL 2
#endif
#ifndef LAB_SL
#define LAB_SL 128
#endif
#ifndef LAB_CG
#define LAB_CG 1
#endif
#ifndef LAB_MINB
#define LAB_MINB 1
#endif
#ifndef LAB_R
#define LAB_R 4
#endif
#ifndef LAB_K
#define LAB_K 120
#endif
#ifndef LAB_C
#define LAB_C 18
#endif
#ifndef LAB_LIMBS
#define LAB_LIMBS 15
#endif
#ifndef LAB_LOGN
#define LAB_LOGN 14
#endif
#ifndef LAB_QBITS
#define LAB_QBITS 24
#endif
typedef uint32_t T;

// LAB_MINB = 0: __launch_bounds__(threads) alone, as the uncapped library kernels are declared (NOT the same register
// allocation as a minimum of 1: ptxas takes 162-166 registers for the wide tiles then, 254 with (threads, 1))
#if LAB_MINB == 0
#define LAB_BOUNDS __launch_bounds__(LAB_SL * LAB_CG)
#else
#define LAB_BOUNDS __launch_bounds__(LAB_SL * LAB_CG, LAB_MINB)
#endif
template <typename T, int V, int TR, int TC, int UNROLL, int SL, int CG>
__global__ void LAB_BOUNDS
matmul_cg_kernel(T *__restrict__ out, const T *__restrict__ A, const T *__restrict__ B, MatmulGeom g, const LimbConst<T> *__restrict__ lc) {
    const uint32_t sx = threadIdx.x % SL, cg = threadIdx.x / SL;
    const uint32_t by = blockIdx.x * CG + cg;
    if (by * TC >= g.cols) return;
    matmul_thread<T, V, TR, TC, UNROLL>(blockIdx.y, by, blockIdx.z, sx, SL, out, A, B, g, lc);
}
__global__ void fill_kernel(T *p, size_t nwords, T q, uint32_t salt) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nwords; i += (size_t)gridDim.x * blockDim.x) {
        uint64_t s = (i + salt) * 0x9E3779B97F4A7C15ull;
        s ^= s >> 29; s *= 0xBF58476D1CE4E5B9ull; s ^= s >> 32;
        p[i] = (T)(s % q);
    }
}

int main() {
    // default: a C3-b column chunk; -DLAB_R=2 -DLAB_K=32 -DLAB_C=32 -DLAB_LIMBS=16 -DLAB_LOGN=15 -DLAB_QBITS=28: the product of a C4 28-bit gate
    const uint32_t logn = LAB_LOGN, n = 1u << logn, limbs = LAB_LIMBS, r = LAB_R, k = LAB_K, c = LAB_C;
    const uint64_t q = LAB_QBITS == 24 ? 16515073ull : 268369921ull;
    std::vector<LimbConst<T>> lc(limbs);
    for (auto &x : lc) {
        x = LimbConst<T>{};
        x.q = (T)q; x.q2 = (T)(2 * q);
        x.mu64 = (uint64_t)((((unsigned __int128)1) << 64) / q);
        x.one_s = (T)((((unsigned __int128)1) << 32) / q);
        x.bits = LAB_QBITS;
    }
    const size_t pw = (size_t)limbs << logn;
    T *A, *B, *O, *O2; LimbConst<T> *L;
    cudaMalloc(&A, (size_t)r * k * pw * 4); cudaMalloc(&B, (size_t)k * c * pw * 4);
    cudaMalloc(&O, (size_t)r * c * pw * 4); cudaMalloc(&O2, (size_t)r * c * pw * 4);
    cudaMalloc(&L, limbs * sizeof(LimbConst<T>));
    fill_kernel<<<1184, 256>>>(B, (size_t)k * c * pw, (T)q, 1);
    fill_kernel<<<1184, 256>>>(A, (size_t)r * k * pw, (T)q, 2);
    cudaMemcpy(L, lc.data(), limbs * sizeof(LimbConst<T>), cudaMemcpyHostToDevice);
    MatmulGeom g{}; g.rows = r; g.inner = k; g.cols = c; g.b_cols = c; g.b_col0 = 0; g.out_cols = c; g.out_col0 = 0;
    g.limbs = limbs; g.logn = logn; g.fold = LAB_QBITS == 24 ? 65536 : 128;
    {   // comparison result: the 4 x 4 tile of half vectors, column tiles fastest
        const unsigned per_limb = (n / 2 + 127) / 128;
        matmul_kernel<T, 2, 4, 4, 2, 128, true><<<dim3((c + 3) / 4, limbs * per_limb, (r + 3) / 4), 128>>>(O2, A, B, g, L);
    }
    const unsigned threads = LAB_SL * LAB_CG, per_limb = (n / LAB_V + LAB_SL - 1) / LAB_SL;
    const unsigned col_tiles = (c + LAB_TC - 1) / LAB_TC;
    dim3 grid((col_tiles + LAB_CG - 1) / LAB_CG, limbs * per_limb, (r + LAB_TR - 1) / LAB_TR);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int i = 0; i < 8; i++) {
        cudaEventRecord(e0);
        matmul_cg_kernel<T, LAB_V, LAB_TR, LAB_TC, LAB_UNROLL, LAB_SL, LAB_CG><<<grid, threads>>>(O, A, B, g, L);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (i >= 2 && ms < best) best = ms;
    }
    std::vector<T> h1((size_t)r * c * pw), h2(h1.size());
    cudaMemcpy(h1.data(), O, h1.size() * 4, cudaMemcpyDeviceToHost); cudaMemcpy(h2.data(), O2, h2.size() * 4, cudaMemcpyDeviceToHost);
    size_t bad = 0;
    for (size_t i = 0; i < h1.size(); i++) bad += h1[i] != h2[i];
    cudaFuncAttributes fa; cudaFuncGetAttributes(&fa, matmul_cg_kernel<T, LAB_V, LAB_TR, LAB_TC, LAB_UNROLL, LAB_SL, LAB_CG>);
    const double bytes = 4.0 * pw * ((double)r * k + (double)k * c + (double)r * c), macs = (double)r * k * c * pw;
    printf("V=%d TR=%d TC=%d unroll=%d SL=%d CG=%d minb=%d regs=%d: %.4f ms  %.0f GB/s  %.2f TMAC/s  mismatches=%zu (%s)\n", LAB_V, LAB_TR,
           LAB_TC, LAB_UNROLL, LAB_SL, LAB_CG, LAB_MINB, fa.numRegs, best, bytes / best / 1e6, macs / best / 1e9, bad,
           cudaGetErrorString(cudaGetLastError()));
    return 0;
}
