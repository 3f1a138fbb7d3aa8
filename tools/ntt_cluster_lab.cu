// ntt_cluster_lab.cu — development harness: the cluster / DSMEM NTT kernels (mxx_b200/csrc/include/ntt_cluster.cuh)
// against the round-1 schedule (global radix passes + block kernel, or the one-CTA-per-SM block kernel) on C2-sized
// batches: bit-exact comparison of the EVAL output, round trip, and CUDA-event timings of both.  Not part of the product.
//
//   nvcc -std=c++17 -O3 -lineinfo -gencode arch=compute_100a,code=sm_100a -Imxx_b200/csrc/include -o lab/ntt_cluster_lab tools/ntt_cluster_lab.cu
//   ntt_cluster_lab <u32|f64|u64> <logn> <limbs> <polys> [bits]
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "host_tables.hpp"
#include "ntt_cluster.cuh"

using namespace gp;

static bool is_prime(uint64_t p) {
    if (p < 2) return false;
    for (uint64_t d = 2; d * d <= p; d++)
        if (p % d == 0) return false;
    return true;
}
#define CK(x)                                                                                   \
    do {                                                                                        \
        cudaError_t e_ = (x);                                                                   \
        if (e_ != cudaSuccess) {                                                                \
            printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__);     \
            exit(1);                                                                            \
        }                                                                                       \
    } while (0)

template <typename F> static float best_ms(F f, int reps = 5) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    float best = 1e30f;
    for (int r = 0; r < reps + 2; r++) {
        cudaEventRecord(e0);
        f();
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (r >= 2 && ms < best) best = ms;
    }
    CK(cudaGetLastError());
    return best;
}

// W = storage word, K = kernel working type (W, or F64 over u64 words)
template <typename W, typename K> struct Lab {
    int logn, limbs, polys;
    size_t words;
    W *d_in, *d_a, *d_b, *d_c;
    TwPair<W> *d_fi, *d_ii;
    LimbConst<W> *d_li;
    TwPair<K> *d_f, *d_i;
    LimbConst<K> *d_l;
    bool nocsub;

    template <int LOGB, bool NOCSUB> void block_fwd(const W *src, W *dst, unsigned grid) {
        constexpr size_t smem = fast_smem_bytes<K, LOGB>();
        auto k = fast_ntt_fwd_kernel<K, LOGB, NOCSUB>;
        cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        k<<<grid, FastCfg<K, LOGB>::kThreads, smem>>>((const K *)src, (K *)dst, limbs, logn, d_f, d_l);
    }
    template <int LOGB> void block_inv(const W *src, W *dst, unsigned grid) {
        constexpr size_t smem = fast_smem_bytes<K, LOGB>();
        auto k = fast_ntt_inv_kernel<K, LOGB>;
        cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        k<<<grid, FastCfg<K, LOGB>::kThreads, smem>>>((const K *)src, (K *)dst, limbs, logn, d_i, d_l);
    }
    template <int R> void gfwd(W *d, uint32_t lspan) {
        const size_t total = ((size_t)polys * limbs) << (logn - R);
        ntt_fwd_global_pass<W, R><<<(unsigned)((total + 255) / 256), 256>>>(d, total, limbs, logn, lspan, d_fi, d_li);
    }
    template <int R> void ginv(W *d, uint32_t lt0) {
        const size_t total = ((size_t)polys * limbs) << (logn - R);
        ntt_inv_global_pass<W, R, true><<<(unsigned)((total + 255) / 256), 256>>>(d, total, limbs, logn, lt0, d_ii, d_li);
    }
    // round-1 schedule (MatrixNTT.inl run_ntt_fwd / run_ntt_inv)
    void base_fwd(const W *src, W *dst) {
        const int logb = fast_block_log<W>(logn), top = logn - logb;
        const unsigned grid = (unsigned)(((size_t)polys * limbs) << top);
        if (top > 0) {
            cudaMemcpyAsync(dst, src, words * sizeof(W), cudaMemcpyDeviceToDevice);
            if (top == 1) gfwd<1>(dst, logn);
            else if (top == 2) gfwd<2>(dst, logn);
            else gfwd<3>(dst, logn);
            src = dst;
        }
#define FWD(LB) (nocsub ? block_fwd<LB, true>(src, dst, grid) : block_fwd<LB, false>(src, dst, grid))
        if constexpr (sizeof(W) == 4) {
            if (logb == 14) FWD(14);
            else FWD(15);
        } else {
            if (logb == 13) FWD(13);
            else FWD(14);
        }
#undef FWD
    }
    void base_inv(const W *src, W *dst) {
        const int logb = fast_block_log<W>(logn), top = logn - logb;
        const unsigned grid = (unsigned)(((size_t)polys * limbs) << top);
        if constexpr (sizeof(W) == 4) {
            if (logb == 14) block_inv<14>(src, dst, grid);
            else block_inv<15>(src, dst, grid);
        } else {
            if (logb == 13) block_inv<13>(src, dst, grid);
            else block_inv<14>(src, dst, grid);
        }
        if (top == 1) ginv<1>(dst, logb);
        else if (top == 2) ginv<2>(dst, logb);
        else if (top == 3) ginv<3>(dst, logb);
    }
    template <int LOGB, int LC> void cl_fwd(const W *src, W *dst) {
        constexpr size_t smem = fast_smem_bytes<K, LOGB>();
        const unsigned grid = (unsigned)(((size_t)polys * limbs) << LC);
        if (nocsub) {
            auto k = fast_ntt_fwd_cluster_kernel<K, LOGB, LC, true>;
            cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            k<<<grid, FastCfg<K, LOGB>::kThreads, smem>>>((const K *)src, (K *)dst, limbs, d_f, d_l);
        } else {
            auto k = fast_ntt_fwd_cluster_kernel<K, LOGB, LC, false>;
            cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            k<<<grid, FastCfg<K, LOGB>::kThreads, smem>>>((const K *)src, (K *)dst, limbs, d_f, d_l);
        }
    }
    template <int LOGB, int LC> void cl_inv(const W *src, W *dst) {
        constexpr size_t smem = fast_smem_bytes<K, LOGB>();
        const unsigned grid = (unsigned)(((size_t)polys * limbs) << LC);
        auto k = fast_ntt_inv_cluster_kernel<K, LOGB, LC>;
        cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        k<<<grid, FastCfg<K, LOGB>::kThreads, smem>>>((const K *)src, (K *)dst, limbs, d_i, d_l);
    }
    void cluster_fwd(const W *src, W *dst) {
        constexpr int LB = sizeof(W) == 4 ? 14 : 13;
        switch (logn - LB) {
            case 1: cl_fwd<LB, 1>(src, dst); break;
            case 2: cl_fwd<LB, 2>(src, dst); break;
            default: cl_fwd<LB, 3>(src, dst); break;
        }
    }
    void cluster_inv(const W *src, W *dst) {
        constexpr int LB = sizeof(W) == 4 ? 14 : 13;
        switch (logn - LB) {
            case 1: cl_inv<LB, 1>(src, dst); break;
            case 2: cl_inv<LB, 2>(src, dst); break;
            default: cl_inv<LB, 3>(src, dst); break;
        }
    }

    int run(int bits, const char *name) {
        const size_t n = (size_t)1 << logn;
        std::vector<uint64_t> moduli, psi;
        for (uint64_t c = (((1ull << bits) - 2) / (2 * n)) * (2 * n) + 1; (int)moduli.size() < limbs && c > 2 * n; c -= 2 * n)
            if (is_prime(c)) moduli.push_back(c);
        for (uint64_t q : moduli) psi.push_back(host::primitive_2n_root(q, n));
        const host::Tables<W> ti = host::build_tables<W>(logn, moduli, psi);
        host::Tables<K> tk;
        if constexpr (is_f64<K>::value) {
            tk = host::f64_tables(ti);
            nocsub = false;
        } else {
            tk = ti;
            nocsub = host::nocsub_ok(moduli[0], sizeof(W) == 8);
        }
        words = (size_t)polys * limbs * n;
        std::vector<W> h(words);
        uint64_t s = 0x9E3779B97F4A7C15ull;
        for (size_t i = 0; i < words; i++) {
            s ^= s << 13, s ^= s >> 7, s ^= s << 17;
            h[i] = (W)(s % moduli[(i >> logn) % limbs]);
        }
        CK(cudaMalloc(&d_in, words * sizeof(W)));
        CK(cudaMalloc(&d_a, words * sizeof(W)));
        CK(cudaMalloc(&d_b, words * sizeof(W)));
        CK(cudaMalloc(&d_c, words * sizeof(W)));
        CK(cudaMalloc(&d_fi, ti.fwd.size() * sizeof(TwPair<W>)));
        CK(cudaMalloc(&d_ii, ti.inv.size() * sizeof(TwPair<W>)));
        CK(cudaMalloc(&d_li, ti.limb.size() * sizeof(LimbConst<W>)));
        CK(cudaMalloc(&d_f, tk.fwd.size() * sizeof(TwPair<K>)));
        CK(cudaMalloc(&d_i, tk.inv.size() * sizeof(TwPair<K>)));
        CK(cudaMalloc(&d_l, tk.limb.size() * sizeof(LimbConst<K>)));
        CK(cudaMemcpy(d_in, h.data(), words * sizeof(W), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(d_fi, ti.fwd.data(), ti.fwd.size() * sizeof(TwPair<W>), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(d_ii, ti.inv.data(), ti.inv.size() * sizeof(TwPair<W>), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(d_li, ti.limb.data(), ti.limb.size() * sizeof(LimbConst<W>), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(d_f, tk.fwd.data(), tk.fwd.size() * sizeof(TwPair<K>), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(d_i, tk.inv.data(), tk.inv.size() * sizeof(TwPair<K>), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(d_l, tk.limb.data(), tk.limb.size() * sizeof(LimbConst<K>), cudaMemcpyHostToDevice));

        const float bf = best_ms([&] { base_fwd(d_in, d_a); });
        CK(cudaDeviceSynchronize());
        const float cf = best_ms([&] { cluster_fwd(d_in, d_b); });
        CK(cudaDeviceSynchronize());
        std::vector<W> ea(words), eb(words), back(words);
        CK(cudaMemcpy(ea.data(), d_a, words * sizeof(W), cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(eb.data(), d_b, words * sizeof(W), cudaMemcpyDeviceToHost));
        size_t fwd_diff = 0;
        for (size_t i = 0; i < words; i++) fwd_diff += ea[i] != eb[i];
        // inverse: baseline works in place on a copy (its global passes are in place)
        const float bi = best_ms([&] { base_inv(d_a, d_c); });
        CK(cudaDeviceSynchronize());
        CK(cudaMemcpy(back.data(), d_c, words * sizeof(W), cudaMemcpyDeviceToHost));
        size_t base_rt = 0;
        for (size_t i = 0; i < words; i++) base_rt += back[i] != h[i];
        const float ci = best_ms([&] { cluster_inv(d_b, d_c); });
        CK(cudaDeviceSynchronize());
        CK(cudaMemcpy(back.data(), d_c, words * sizeof(W), cudaMemcpyDeviceToHost));
        size_t cl_rt = 0;
        for (size_t i = 0; i < words; i++) cl_rt += back[i] != h[i];
        const double gc = (double)words / 1e6;
        printf("%s logn=%d bits=%d limbs=%d polys=%d nocsub=%d | baseline fwd %.4f ms (%.1f Gc/s) inv %.4f ms (%.1f Gc/s) | cluster fwd %.4f ms (%.1f Gc/s) inv %.4f ms (%.1f Gc/s) | fwd-diff %zu baseline-roundtrip-bad %zu cluster-roundtrip-bad %zu\n",
               name, logn, bits, limbs, polys, (int)nocsub, bf, gc / bf, bi, gc / bi, cf, gc / cf, ci, gc / ci, fwd_diff, base_rt, cl_rt);
        fflush(stdout);
        return (fwd_diff || cl_rt) ? 2 : 0;
    }
};

int main(int argc, char **argv) {
    if (argc < 5) {
        printf("usage: ntt_cluster_lab <u32|f64|u64> <logn> <limbs> <polys> [bits]\n");
        return 1;
    }
    const char *ty = argv[1];
    const int logn = atoi(argv[2]), limbs = atoi(argv[3]), polys = atoi(argv[4]);
    if (!strcmp(ty, "u32")) {
        if (logn < 15 || logn > 17) return printf("u32: logn 15..17\n"), 1;
        Lab<uint32_t, uint32_t> lab{};
        lab.logn = logn, lab.limbs = limbs, lab.polys = polys;
        return lab.run(argc > 5 ? atoi(argv[5]) : 28, ty);
    }
    if (logn < 14 || logn > 16) return printf("u64/f64: logn 14..16\n"), 1;
    if (!strcmp(ty, "f64")) {
        Lab<uint64_t, F64> lab{};
        lab.logn = logn, lab.limbs = limbs, lab.polys = polys;
        return lab.run(argc > 5 ? atoi(argv[5]) : 51, ty);
    }
    Lab<uint64_t, uint64_t> lab{};
    lab.logn = logn, lab.limbs = limbs, lab.polys = polys;
    return lab.run(argc > 5 ? atoi(argv[5]) : 59, ty);
}
