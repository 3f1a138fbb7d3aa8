// modmul_peak.cu — register-resident microbenchmark of the modular primitives the DCRT path
// is built from, to obtain the INTEGER roofline denominators the kernels are reported against
// ("INT roof: measured, not assumed").
//
// Each thread runs ILP independent dependency chains of the exact butterfly / MAC code the
// kernels use (csrc/include/modarith.cuh, ntt_fast.cuh fwd_net / inv_net bodies), with no
// memory traffic in the loop.  Prints one JSON object:
//   {"sm_count":..,"clock_mhz":..,"bfly_u32_nocsub":<butterflies/s>, ...}
//
// build: make -C mxx_b200/csrc tools      run: mxx_b200/modmul_peak
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

#include "modarith.cuh"
#include "ntt_fast.cuh"

using namespace gp;

constexpr int kIlp = 8;      // independent butterflies per thread per round (16 values)
constexpr int kRounds = 512;

// forward Cooley-Tukey butterfly chains: (X, Y) -> (X + wY, X - wY + 2q)
template <typename T, bool NOCSUB>
__global__ void __launch_bounds__(256) bfly_fwd_kernel(T *out, T w, T ws, T q, int rounds) {
    T x[2 * kIlp];
#pragma unroll
    for (int i = 0; i < 2 * kIlp; i++) x[i] = (T)(threadIdx.x * 2654435761u + i * 40503u + blockIdx.x) % q;
    const T q2 = 2 * q;
    for (int r = 0; r < rounds; r++) {
#pragma unroll
        for (int i = 0; i < kIlp; i++) {
            const T X = NOCSUB ? x[2 * i] : csub_min<T>(x[2 * i], q2);
            const T t = mul_shoup_lazy<T>(x[2 * i + 1], w, ws, q);
            x[2 * i] = X + t;
            x[2 * i + 1] = X - t + q2;
        }
        if (NOCSUB) {  // keep the chain bounded the way the kernel does at its final store (amortised 1/14)
            if ((r & 15) == 15) {
#pragma unroll
                for (int i = 0; i < 2 * kIlp; i++) x[i] = reduce_lazy<T>(x[i], (T)(~(T)0 / q), q);
            }
        }
    }
    T acc = 0;
#pragma unroll
    for (int i = 0; i < 2 * kIlp; i++) acc ^= x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

// inverse Gentleman-Sande butterfly chains: (X, Y) -> (csub(X + Y), w (X - Y + 2q))
template <typename T>
__global__ void __launch_bounds__(256) bfly_inv_kernel(T *out, T w, T ws, T q, int rounds) {
    T x[2 * kIlp];
#pragma unroll
    for (int i = 0; i < 2 * kIlp; i++) x[i] = (T)(threadIdx.x * 2654435761u + i * 40503u + blockIdx.x) % q;
    const T q2 = 2 * q;
    for (int r = 0; r < rounds; r++) {
#pragma unroll
        for (int i = 0; i < kIlp; i++) {
            const T X = x[2 * i], Y = x[2 * i + 1];
            x[2 * i] = csub_min<T>(X + Y, q2);
            x[2 * i + 1] = mul_shoup_lazy<T>(X - Y + q2, w, ws, q);
        }
    }
    T acc = 0;
#pragma unroll
    for (int i = 0; i < 2 * kIlp; i++) acc ^= x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

// FP64-pipe butterflies (modarith.cuh f64_*; the kernels ntt_fast.cuh runs for u64 contexts with moduli < 2^51):
// forward = renormalise the multiplied leg, multiply, add/subtract; inverse = add (renormalised every other stage),
// subtract, renormalise, multiply
template <bool INVERSE>
__global__ void __launch_bounds__(256) bfly_f64_kernel(double *out, double w, double q, int rounds) {
    double x[2 * kIlp];
#pragma unroll
    for (int i = 0; i < 2 * kIlp; i++) x[i] = (double)((threadIdx.x * 2654435761u + i * 40503u + blockIdx.x) % 1000003u);
    const double qinv = 1.0 / q, wq = w / q;
    for (int r = 0; r < rounds; r++) {
#pragma unroll
        for (int i = 0; i < kIlp; i++) {
            const double X = x[2 * i], Y = x[2 * i + 1];
            if (INVERSE) {
                const double s = f64_add(X, Y);
                x[2 * i] = (r & 1) ? f64_renorm(s, q, qinv) : s;
                x[2 * i + 1] = f64_mulmod(f64_renorm(f64_add(X, -Y), q, qinv), w, wq, q);
            } else {
                const double t = f64_mulmod(f64_renorm(Y, q, qinv), w, wq, q);
                x[2 * i] = f64_add(X, t);
                x[2 * i + 1] = f64_add(X, -t);
            }
        }
        if (!INVERSE && (r & 3) == 3) {  // the nets renormalise every value at entry (at most 5 stages apart)
#pragma unroll
            for (int i = 0; i < kIlp; i++) x[2 * i] = f64_renorm(x[2 * i], q, qinv);
        }
    }
    double acc = 0;
#pragma unroll
    for (int i = 0; i < 2 * kIlp; i++) acc += x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

// lazy multiply-accumulate chains (matmul inner loop): acc += a * b, fold every `fold` terms
template <typename T>
__global__ void __launch_bounds__(256) mac_kernel(T *out, LimbConst<T> c, int rounds, int fold) {
    Acc<T> acc[kIlp];
    T a[kIlp], b[kIlp];
#pragma unroll
    for (int i = 0; i < kIlp; i++) {
        acc[i].clear();
        a[i] = (T)(threadIdx.x * 2654435761u + i * 40503u + blockIdx.x) % c.q;
        b[i] = (T)(threadIdx.x * 40503u + i * 2654435761u + 7u * blockIdx.x) % c.q;
    }
    int since = 0;
    for (int r = 0; r < rounds; r++) {
#pragma unroll
        for (int i = 0; i < kIlp; i++) {
            acc[i].mac(a[i], b[i]);
            a[i] += 1;  // keep the operands moving so nothing is hoisted (one ALU op, as cheap as an address update)
        }
        if (++since == fold) {
            since = 0;
#pragma unroll
            for (int i = 0; i < kIlp; i++) {
                acc[i].fold(c);
                a[i] = csub<T>(a[i], c.q);
            }
        }
    }
    T r = 0;
#pragma unroll
    for (int i = 0; i < kIlp; i++) r ^= acc[i].result(c);
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

template <typename F> static double time_ms(F launch, int reps = 5) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    launch();
    launch();
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int i = 0; i < reps; i++) {
        cudaEventRecord(e0);
        launch();
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    return best;
}

template <typename T> static LimbConst<T> limb_const(uint64_t q) {
    typedef unsigned __int128 u128;
    LimbConst<T> c{};
    c.q = (T)q;
    c.q2 = (T)(2 * q);
    const uint64_t r64 = (uint64_t)((((u128)1) << 64) % q);
    c.r64 = (T)r64;
    c.r64_s = (T)((((u128)r64) << (sizeof(T) * 8)) / q);
    c.one_s = (T)((((u128)1) << (sizeof(T) * 8)) / q);
    c.mu64 = (uint64_t)((((u128)1) << 64) / q);
    c.bits = 64 - __builtin_clzll(q);
    return c;
}

int main() {
    int dev = 0;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, dev) != cudaSuccess) {
        printf("{\"error\": \"no CUDA device\"}\n");
        return 1;
    }
    const int sms = prop.multiProcessorCount, threads = 256, blocks = sms * 8;
    void *out = nullptr;
    cudaMalloc(&out, (size_t)blocks * threads * 8);
    const uint64_t q32 = 16515073ull /* 24-bit, = 1 mod 2^15 */, q32b = 268369921ull /* 28-bit */, q64 = 2251799813554177ull /* 51-bit */;
    const uint32_t w32 = 1234567u, ws32 = (uint32_t)(((unsigned __int128)w32 << 32) / q32);
    const uint32_t w32b = 123456789u, ws32b = (uint32_t)(((unsigned __int128)w32b << 32) / q32b);
    const uint64_t w64 = 1234567890123ull, ws64 = (uint64_t)(((unsigned __int128)w64 << 64) / q64);
    const double bf = (double)blocks * threads * kIlp * kRounds;
    double t;
    printf("{\"sm_count\": %d, \"clock_mhz\": %d", sms, prop.clockRate / 1000);
    t = time_ms([&] { bfly_fwd_kernel<uint32_t, true><<<blocks, threads>>>((uint32_t *)out, w32, ws32, (uint32_t)q32, kRounds); });
    printf(", \"bfly_fwd_u32_nocsub\": %.4e", bf / (t * 1e-3));
    t = time_ms([&] { bfly_fwd_kernel<uint32_t, false><<<blocks, threads>>>((uint32_t *)out, w32b, ws32b, (uint32_t)q32b, kRounds); });
    printf(", \"bfly_fwd_u32\": %.4e", bf / (t * 1e-3));
    t = time_ms([&] { bfly_inv_kernel<uint32_t><<<blocks, threads>>>((uint32_t *)out, w32b, ws32b, (uint32_t)q32b, kRounds); });
    printf(", \"bfly_inv_u32\": %.4e", bf / (t * 1e-3));
    t = time_ms([&] { bfly_fwd_kernel<uint64_t, true><<<blocks, threads>>>((uint64_t *)out, w64, ws64, q64, kRounds); });
    printf(", \"bfly_fwd_u64_nocsub\": %.4e", bf / (t * 1e-3));
    t = time_ms([&] { bfly_fwd_kernel<uint64_t, false><<<blocks, threads>>>((uint64_t *)out, w64, ws64, q64, kRounds); });
    printf(", \"bfly_fwd_u64\": %.4e", bf / (t * 1e-3));
    t = time_ms([&] { bfly_inv_kernel<uint64_t><<<blocks, threads>>>((uint64_t *)out, w64, ws64, q64, kRounds); });
    printf(", \"bfly_inv_u64\": %.4e", bf / (t * 1e-3));
    t = time_ms([&] { bfly_f64_kernel<false><<<blocks, threads>>>((double *)out, (double)w64, (double)q64, kRounds); });
    printf(", \"bfly_fwd_f64\": %.4e", bf / (t * 1e-3));
    t = time_ms([&] { bfly_f64_kernel<true><<<blocks, threads>>>((double *)out, (double)w64, (double)q64, kRounds); });
    printf(", \"bfly_inv_f64\": %.4e", bf / (t * 1e-3));
    const LimbConst<uint32_t> c32 = limb_const<uint32_t>(q32);
    const LimbConst<uint64_t> c64 = limb_const<uint64_t>(q64);
    t = time_ms([&] { mac_kernel<uint32_t><<<blocks, threads>>>((uint32_t *)out, c32, kRounds, 16); });
    printf(", \"mac_u32\": %.4e", bf / (t * 1e-3));
    t = time_ms([&] { mac_kernel<uint64_t><<<blocks, threads>>>((uint64_t *)out, c64, kRounds, 16); });
    printf(", \"mac_u64\": %.4e", bf / (t * 1e-3));
    printf("}\n");
    cudaFree(out);
    return cudaDeviceSynchronize() == cudaSuccess ? 0 : 1;
}
