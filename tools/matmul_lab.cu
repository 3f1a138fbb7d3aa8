// matmul_lab.cu — development harness: times one build variant of matmul_kernel (arith.cuh) on the C1-R shape.
//   -DLAB_V=4 -DLAB_TR=1 -DLAB_TC=4 -DGP_MM_UNROLL=2 -DGP_MM_THREADS=128
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "arith.cuh"
using namespace gp;
#ifndef LAB_V
#define LAB_V 4
#endif
#ifndef LAB_TR
#define LAB_TR 1
#endif
#ifndef LAB_TC
#define LAB_TC 4
#endif
#ifndef GP_MM_UNROLL
#define GP_MM_UNROLL 2
#endif
#ifndef GP_MM_THREADS
#define GP_MM_THREADS 128
#endif
typedef uint32_t T;
int main() {
    const uint32_t logn = 14, n = 1u << logn, limbs = 15, r = 1, k = 30, c = 120;
    const uint64_t q = 16515073ull;
    std::vector<LimbConst<T>> lc(limbs);
    for (auto &x : lc) {
        x = LimbConst<T>{};
        x.q = (T)q; x.q2 = (T)(2 * q);
        x.mu64 = (uint64_t)((((unsigned __int128)1) << 64) / q);
        x.one_s = (T)((((unsigned __int128)1) << 32) / q);
        x.bits = 24;
    }
    const size_t pw = (size_t)limbs << logn;
    T *A, *B, *O; LimbConst<T> *L;
    cudaMalloc(&A, r * k * pw * 4); cudaMalloc(&B, (size_t)k * c * pw * 4); cudaMalloc(&O, r * c * pw * 4); cudaMalloc(&L, limbs * sizeof(LimbConst<T>));
    cudaMemset(A, 1, r * k * pw * 4); cudaMemset(B, 2, (size_t)k * c * pw * 4);
    cudaMemcpy(L, lc.data(), limbs * sizeof(LimbConst<T>), cudaMemcpyHostToDevice);
    MatmulGeom g{}; g.rows = r; g.inner = k; g.cols = c; g.b_cols = c; g.b_col0 = 0; g.out_cols = c; g.out_col0 = 0; g.limbs = limbs; g.logn = logn; g.fold = 65536;
    const unsigned threads = GP_MM_THREADS, per_limb = (n / LAB_V + threads - 1) / threads;
    dim3 grid(limbs * per_limb, (c + LAB_TC - 1) / LAB_TC, (r + LAB_TR - 1) / LAB_TR);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int i = 0; i < 8; i++) {
        cudaEventRecord(e0);
        matmul_kernel<T, LAB_V, LAB_TR, LAB_TC, GP_MM_UNROLL, GP_MM_THREADS><<<grid, threads>>>(O, A, B, g, L);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (i >= 2 && ms < best) best = ms;
    }
    const double bytes = 4.0 * pw * (r * k + (double)k * c + r * c);
    printf("V=%d TR=%d TC=%d unroll=%d threads=%d: %.4f ms  %.1f GB/s  (%s)\n", LAB_V, LAB_TR, LAB_TC, GP_MM_UNROLL, GP_MM_THREADS, best, bytes / best / 1e6,
           cudaGetErrorString(cudaGetLastError()));
    return 0;
}
