// ntt_lab.cu — development harness: times ONE build variant of the compile-time-planned NTT kernels
// (mxx_b200/csrc/include/ntt_fast.cuh) on a C2-sized batch and prints a checksum, so that variants
// built with different -D flags can be compared in one gpurun call.  Not part of the product.
//
//   nvcc -std=c++17 -O3 -lineinfo -gencode arch=compute_100a,code=sm_100a -Imxx_b200/csrc/include \
//        [-DGP_...=...] -o /tmp/ntt_lab tools/ntt_lab.cu
//   ntt_lab [bits=28] [limbs=16] [polys=256]
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "host_tables.hpp"
#include "ntt_fast.cuh"

using namespace gp;
#ifndef LAB_LOGN
#define LAB_LOGN 14
#endif
// LAB_WIDE: u64 words, integer kernels; LAB_F64: u64 words, FP64-pipe kernels (moduli < 2^51)
#if defined(LAB_F64)
typedef uint64_t T;
typedef F64 K;
#elif defined(LAB_WIDE)
typedef uint64_t T;
typedef uint64_t K;
#else
typedef uint32_t T;
typedef uint32_t K;
#endif

static bool is_prime(uint64_t p) {
    if (p < 2) return false;
    for (uint64_t d = 2; d * d <= p; d++)
        if (p % d == 0) return false;
    return true;
}

#define CK(x)                                                                      \
    do {                                                                           \
        cudaError_t e_ = (x);                                                       \
        if (e_ != cudaSuccess) {                                                  \
            printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); \
            return 1;                                                              \
        }                                                                          \
    } while (0)

template <bool NOCSUB> static float run(const T *src_, T *dst_, unsigned grid, uint32_t limbs, const TwPair<K> *fwd, const TwPair<K> *inv,
                                        const LimbConst<K> *lc, bool inverse, int reps) {
    const K *src = (const K *)src_;
    K *dst = (K *)dst_;
    constexpr size_t smem = fast_smem_bytes<K, LAB_LOGN>();
    auto kf = fast_ntt_fwd_kernel<K, LAB_LOGN, NOCSUB>;
    auto ki = fast_ntt_inv_kernel<K, LAB_LOGN>;
    cudaFuncSetAttribute(kf, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(ki, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    float best = 1e30f;
    for (int r = 0; r < reps + 2; r++) {
        cudaEventRecord(e0);
        if (inverse)
            ki<<<grid, FastCfg<K, LAB_LOGN>::kThreads, smem>>>(src, dst, limbs, LAB_LOGN, inv, lc);
        else
            kf<<<grid, FastCfg<K, LAB_LOGN>::kThreads, smem>>>(src, dst, limbs, LAB_LOGN, fwd, lc);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (r >= 2 && ms < best) best = ms;
    }
    return best;
}

int main(int argc, char **argv) {
    const int bits = argc > 1 ? atoi(argv[1]) : 28, limbs = argc > 2 ? atoi(argv[2]) : 16, polys = argc > 3 ? atoi(argv[3]) : 256;
    const size_t n = (size_t)1 << LAB_LOGN;
    std::vector<uint64_t> moduli, psi;
    for (uint64_t c = (((1ull << bits) - 2) / (2 * n)) * (2 * n) + 1; (int)moduli.size() < limbs && c > 2 * n; c -= 2 * n)
        if (is_prime(c)) moduli.push_back(c);
    for (uint64_t q : moduli) psi.push_back(host::primitive_2n_root(q, n));
#if defined(LAB_F64)
    const host::Tables<K> t = host::f64_tables(host::build_tables<T>(LAB_LOGN, moduli, psi));
    const bool nocsub = false;
#else
    const host::Tables<K> t = host::build_tables<T>(LAB_LOGN, moduli, psi);
    const bool nocsub = host::nocsub_ok(moduli[0], sizeof(T) == 8);
#endif
    const size_t words = (size_t)polys * limbs * n;
    std::vector<T> h(words);
    uint64_t s = 0x9E3779B97F4A7C15ull;
    for (size_t i = 0; i < words; i++) {
        s ^= s << 13, s ^= s >> 7, s ^= s << 17;
        h[i] = (T)(s % moduli[(i >> LAB_LOGN) % limbs]);
    }
    T *d_a, *d_b, *d_c;
    TwPair<K> *d_f, *d_i;
    LimbConst<K> *d_l;
    CK(cudaMalloc(&d_a, words * sizeof(T)));
    CK(cudaMalloc(&d_b, words * sizeof(T)));
    CK(cudaMalloc(&d_c, words * sizeof(T)));
    CK(cudaMalloc(&d_f, t.fwd.size() * sizeof(TwPair<K>)));
    CK(cudaMalloc(&d_i, t.inv.size() * sizeof(TwPair<K>)));
    CK(cudaMalloc(&d_l, t.limb.size() * sizeof(LimbConst<K>)));
    CK(cudaMemcpy(d_a, h.data(), words * sizeof(T), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_f, t.fwd.data(), t.fwd.size() * sizeof(TwPair<K>), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_i, t.inv.data(), t.inv.size() * sizeof(TwPair<K>), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_l, t.limb.data(), t.limb.size() * sizeof(LimbConst<K>), cudaMemcpyHostToDevice));
    const unsigned grid = (unsigned)((size_t)polys * limbs);
    const float f = nocsub ? run<true>(d_a, d_b, grid, limbs, d_f, d_i, d_l, false, 5) : run<false>(d_a, d_b, grid, limbs, d_f, d_i, d_l, false, 5);
    const float iv = run<false>(d_b, d_c, grid, limbs, d_f, d_i, d_l, true, 5);
    CK(cudaDeviceSynchronize());
    std::vector<T> e(words), back(words);
    CK(cudaMemcpy(e.data(), d_b, words * sizeof(T), cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(back.data(), d_c, words * sizeof(T), cudaMemcpyDeviceToHost));
    uint64_t sum = 0;
    for (size_t i = 0; i < words; i++) sum = sum * 1000003ull + e[i];
    size_t bad = 0;
    for (size_t i = 0; i < words; i++) bad += back[i] != h[i];
    const double gc = (double)words / 1e6;
    printf("logn=%d bits=%d limbs=%d polys=%d nocsub=%d  fwd %.4f ms (%.1f Gcoeff/s)  inv %.4f ms (%.1f Gcoeff/s)  eval-checksum %016llx  roundtrip-mismatches %zu\n",
           LAB_LOGN, bits, limbs, polys, (int)nocsub, f, gc / f, iv, gc / iv, (unsigned long long)sum, bad);
    return bad ? 2 : 0;
}
