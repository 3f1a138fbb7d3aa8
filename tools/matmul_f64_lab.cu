// matmul_f64_lab.cu — development harness for the FP64-pipe product (arith.cuh matmul_f64_thread) on the product of a C4 gate:
// S (2 x 24) times 24 x 24 digit polynomials, 8 x 51-bit limbs, n = 2^15 (1.2 GB of u64 words).  One build = one variant:
//   -DLAB_V -DLAB_TR -DLAB_TC -DLAB_UNROLL -DLAB_NT -DLAB_MINB (0 = __launch_bounds__(NT) alone)
// Checked against the shipped tile (V = 2, 2 x 4, unroll 1, 128 threads, 4 CTAs).  Results: profiles/r02_matmul_rows_lab.txt.
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "arith.cuh"
using namespace gp;
#ifndef LAB_V
#define LAB_V 2
#endif
#ifndef LAB_TR
#define LAB_TR 2
#endif
#ifndef LAB_TC
#define LAB_TC 4
#endif
#ifndef LAB_UNROLL
#define LAB_UNROLL 1
#endif
#ifndef LAB_NT
#define LAB_NT 128
#endif
#ifndef LAB_MINB
#define LAB_MINB 4
#endif
#if LAB_MINB == 0
#define LAB_BOUNDS __launch_bounds__(LAB_NT)
#else
#define LAB_BOUNDS __launch_bounds__(LAB_NT, LAB_MINB)
#endif
typedef uint64_t T;
__global__ void LAB_BOUNDS lab_kernel(T *__restrict__ out, const T *__restrict__ A, const T *__restrict__ B, MatmulGeom g,
                                      const LimbConst<F64> *__restrict__ lc) {
    matmul_f64_thread<LAB_V, LAB_TR, LAB_TC, LAB_UNROLL>(blockIdx.y, blockIdx.x, blockIdx.z, threadIdx.x, blockDim.x, out, A, B, g, lc);
}
__global__ void fill_kernel(T *p, size_t nwords, T q, uint32_t salt) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nwords; i += (size_t)gridDim.x * blockDim.x) {
        uint64_t s = (i + salt) * 0x9E3779B97F4A7C15ull;
        s ^= s >> 29; s *= 0xBF58476D1CE4E5B9ull; s ^= s >> 32;
        p[i] = s % q;
    }
}
int main() {
    const uint32_t logn = 15, n = 1u << logn, limbs = 8, r = 2, k = 24, c = 24;
    const uint64_t q = 2251799813554177ull;  // below 2^51
    std::vector<LimbConst<F64>> lc(limbs);
    for (auto &x : lc) {
        x = LimbConst<F64>{};
        x.q.v = (double)q;
        x.q2.v = 1.0 / (double)q;  // LimbConst<F64>::q2 carries 1/q
    }
    const size_t pw = (size_t)limbs << logn;
    T *A, *B, *O, *O2; LimbConst<F64> *L;
    cudaMalloc(&A, (size_t)r * k * pw * 8); cudaMalloc(&B, (size_t)k * c * pw * 8);
    cudaMalloc(&O, (size_t)r * c * pw * 8); cudaMalloc(&O2, (size_t)r * c * pw * 8);
    cudaMalloc(&L, limbs * sizeof(LimbConst<F64>));
    fill_kernel<<<1184, 256>>>(B, (size_t)k * c * pw, q, 1);
    fill_kernel<<<1184, 256>>>(A, (size_t)r * k * pw, q, 2);
    cudaMemcpy(L, lc.data(), limbs * sizeof(LimbConst<F64>), cudaMemcpyHostToDevice);
    MatmulGeom g{}; g.rows = r; g.inner = k; g.cols = c; g.b_cols = c; g.b_col0 = 0; g.out_cols = c; g.out_col0 = 0;
    g.limbs = limbs; g.logn = logn; g.fold = 3;
    {
        const unsigned per_limb = (n / 2 + 127) / 128;
        matmul_f64_kernel<2, 2, 4, 1, 128, true><<<dim3((c + 3) / 4, limbs * per_limb, (r + 1) / 2), 128>>>(O2, A, B, g, L);
    }
    const unsigned per_limb = (n / LAB_V + LAB_NT - 1) / LAB_NT;
    dim3 grid((c + LAB_TC - 1) / LAB_TC, limbs * per_limb, (r + LAB_TR - 1) / LAB_TR);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int i = 0; i < 8; i++) {
        cudaEventRecord(e0);
        lab_kernel<<<grid, LAB_NT>>>(O, A, B, g, L);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (i >= 2 && ms < best) best = ms;
    }
    std::vector<T> h1((size_t)r * c * pw), h2(h1.size());
    cudaMemcpy(h1.data(), O, h1.size() * 8, cudaMemcpyDeviceToHost); cudaMemcpy(h2.data(), O2, h2.size() * 8, cudaMemcpyDeviceToHost);
    size_t bad = 0;
    for (size_t i = 0; i < h1.size(); i++) bad += h1[i] != h2[i];
    cudaFuncAttributes fa; cudaFuncGetAttributes(&fa, lab_kernel);
    const double bytes = 8.0 * pw * ((double)r * k + (double)k * c + (double)r * c), macs = (double)r * k * c * pw;
    printf("f64 V=%d TR=%d TC=%d unroll=%d NT=%d minb=%d regs=%d: %.4f ms  %.0f GB/s  %.2f TMAC/s  mismatches=%zu (%s)\n", LAB_V, LAB_TR, LAB_TC,
           LAB_UNROLL, LAB_NT, LAB_MINB, fa.numRegs, best, bytes / best / 1e6, macs / best / 1e9, bad, cudaGetErrorString(cudaGetLastError()));
    return 0;
}
