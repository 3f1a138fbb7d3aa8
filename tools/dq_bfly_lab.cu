// dq_bfly_lab.cu — development microbenchmark: u32 NTT butterflies whose Shoup quotient comes from ONE DFMA instead of
// IMAD.HI ("double-quotient" butterflies), against the integer butterflies the kernels use today.
//
// tools/pipe_lab.cu measured on B200: IMAD.HI.U32 issues at 0.89 warp-instr/clk/SM (IMAD.lo 1.93, DFMA 1.76, IADD3 3.6), so
// the integer butterfly (IMAD.HI + 2 IMAD + 3 ALU) is bound by the multiplier pipe at ~2.15 clk per warp-butterfly.
// Here the quotient estimate qh = rint(u * w/q) is the low word of fma(2^52 + u, w/q, 1.5*2^52 - 2^52*(w/q)) — exact when
// w/q is stored as a multiple of 2^-52 — and t = u*w - qh*q lands in [-q/2, q/2]: values are signed and lazy, the forward
// butterfly has no conditional subtraction (|x| grows by q/2 per stage), the inverse one reduces its sum leg every few stages.
// Register-resident dependency chains, no memory traffic.  Prints butterflies/s and butterflies/clk/SM.
//   nvcc -std=c++17 -O3 -lineinfo -gencode arch=compute_100a,code=sm_100a -Imxx_b200/csrc/include -o lab/dq_bfly_lab tools/dq_bfly_lab.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

#include "modarith.cuh"
#include "ntt_fast.cuh"

using namespace gp;

constexpr int kIlp = 8, kRounds = 512;

struct DqTw {  // a twiddle of the double-quotient butterfly
    uint32_t w;
    double winv;  // round(w/q * 2^52) / 2^52
    double c;     // 1.5 * 2^52 - 2^52 * winv  (exact)
};
__device__ __forceinline__ int32_t dq_mul(int32_t x, const DqTw &t, uint32_t q, uint32_t off) {
    const uint32_t u = (uint32_t)x + off;  // off = M*q >= |x|: u >= 0 and u = x (mod q)
    const double f = fma(__hiloint2double(0x43300000, (int)u), t.winv, t.c);
    const uint32_t qh = (uint32_t)__double2loint(f);
    return (int32_t)(u * t.w - qh * q);
}

template <int VARIANT>
__global__ void __launch_bounds__(256) dq_fwd_kernel(int32_t *out, DqTw tw, DqTw one, uint32_t q, uint32_t off, int rounds) {
    int32_t x[2 * kIlp];
#pragma unroll
    for (int i = 0; i < 2 * kIlp; i++) x[i] = (int32_t)((threadIdx.x * 2654435761u + i * 40503u + blockIdx.x) % q);
    for (int r = 0; r < rounds; r++) {
#pragma unroll
        for (int i = 0; i < kIlp; i++) {
            const int32_t t = dq_mul(x[2 * i + 1], tw, q, off);
            const int32_t X = x[2 * i];
            x[2 * i] = X + t;
            x[2 * i + 1] = X - t;
        }
        if ((r % 14) == 13) {  // the kernel's final canonicalisation, amortised over a 2^14 transform
#pragma unroll
            for (int i = 0; i < kIlp; i++) x[2 * i] = dq_mul(x[2 * i], one, q, off);
        }
    }
    int32_t acc = 0;
#pragma unroll
    for (int i = 0; i < 2 * kIlp; i++) acc ^= x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

// Gentleman-Sande: (X, Y) -> (X + Y, w (X - Y)); the sum leg is reduced every PERIOD stages
template <int PERIOD>
__global__ void __launch_bounds__(256) dq_inv_kernel(int32_t *out, DqTw tw, DqTw one, uint32_t q, uint32_t off, int rounds) {
    int32_t x[2 * kIlp];
#pragma unroll
    for (int i = 0; i < 2 * kIlp; i++) x[i] = (int32_t)((threadIdx.x * 2654435761u + i * 40503u + blockIdx.x) % q);
    for (int r = 0; r < rounds; r++) {
#pragma unroll
        for (int i = 0; i < kIlp; i++) {
            const int32_t X = x[2 * i], Y = x[2 * i + 1];
            int32_t s = X + Y;
            if ((r % PERIOD) == PERIOD - 1) s = dq_mul(s, one, q, off);
            x[2 * i] = s;
            x[2 * i + 1] = dq_mul(X - Y, tw, q, off);
        }
    }
    int32_t acc = 0;
#pragma unroll
    for (int i = 0; i < 2 * kIlp; i++) acc ^= x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

// today's integer butterflies (tools/modmul_peak.cu)
template <bool NOCSUB>
__global__ void __launch_bounds__(256) int_fwd_kernel(uint32_t *out, uint32_t w, uint32_t ws, uint32_t q, int rounds) {
    uint32_t x[2 * kIlp];
#pragma unroll
    for (int i = 0; i < 2 * kIlp; i++) x[i] = (threadIdx.x * 2654435761u + i * 40503u + blockIdx.x) % q;
    const uint32_t q2 = 2 * q;
    for (int r = 0; r < rounds; r++) {
#pragma unroll
        for (int i = 0; i < kIlp; i++) {
            const uint32_t X = NOCSUB ? x[2 * i] : csub_min<uint32_t>(x[2 * i], q2);
            const uint32_t t = mul_shoup_lazy<uint32_t>(x[2 * i + 1], w, ws, q);
            x[2 * i] = X + t;
            x[2 * i + 1] = X - t + q2;
        }
        if (NOCSUB && (r & 15) == 15) {
#pragma unroll
            for (int i = 0; i < 2 * kIlp; i++) x[i] = reduce_lazy<uint32_t>(x[i], (uint32_t)(~0u / q), q);
        }
    }
    uint32_t acc = 0;
#pragma unroll
    for (int i = 0; i < 2 * kIlp; i++) acc ^= x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}
__global__ void __launch_bounds__(256) int_inv_kernel(uint32_t *out, uint32_t w, uint32_t ws, uint32_t q, int rounds) {
    uint32_t x[2 * kIlp];
#pragma unroll
    for (int i = 0; i < 2 * kIlp; i++) x[i] = (threadIdx.x * 2654435761u + i * 40503u + blockIdx.x) % q;
    const uint32_t q2 = 2 * q;
    for (int r = 0; r < rounds; r++) {
#pragma unroll
        for (int i = 0; i < kIlp; i++) {
            const uint32_t X = x[2 * i], Y = x[2 * i + 1];
            x[2 * i] = csub_min<uint32_t>(X + Y, q2);
            x[2 * i + 1] = mul_shoup_lazy<uint32_t>(X - Y + q2, w, ws, q);
        }
    }
    uint32_t acc = 0;
#pragma unroll
    for (int i = 0; i < 2 * kIlp; i++) acc ^= x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
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
    return best;
}

static DqTw make_tw(uint32_t w, uint32_t q) {
    DqTw t;
    t.w = w;
    const unsigned __int128 num = ((unsigned __int128)w << 52);
    const uint64_t k = (uint64_t)((num + q / 2) / q);  // round(w/q * 2^52)
    t.winv = (double)k / 4503599627370496.0;
    t.c = 6755399441055744.0 - (double)k;  // 1.5 * 2^52 - k, exact
    return t;
}

// host check of the quotient trick on random operands: |t| <= q/2 + 1 and t = x*w (mod q)
static int self_check(uint32_t q, uint32_t off) {
    uint64_t s = 88172645463325252ull;
    int bad = 0;
    for (int it = 0; it < 2000000; it++) {
        s ^= s << 13, s ^= s >> 7, s ^= s << 17;
        const uint32_t w = (uint32_t)(s % q);
        s ^= s << 13, s ^= s >> 7, s ^= s << 17;
        const int64_t x = (int64_t)(s % (2ull * off)) - (int64_t)off;  // |x| <= off
        const DqTw t = make_tw(w, q);
        const uint32_t u = (uint32_t)((int32_t)x) + off;
        uint64_t bits = ((uint64_t)0x43300000u << 32) | u;
        double d;
        memcpy(&d, &bits, 8);
        const double f = __builtin_fma(d, t.winv, t.c);
        uint64_t fb;
        memcpy(&fb, &f, 8);
        const uint32_t qh = (uint32_t)fb;
        const int32_t r = (int32_t)(u * t.w - qh * q);
        const int64_t want = (int64_t)(((__int128)x * w % q + q) % q);
        const int64_t got = (((int64_t)r % q) + q) % q;
        if (got != want || r > (int64_t)q / 2 + 1 || r < -(int64_t)q / 2 - 1) bad++;
    }
    return bad;
}

int main() {
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, 0) != cudaSuccess) return 1;
    const int sms = prop.multiProcessorCount, threads = 256, blocks = sms * 8;
    void *out = nullptr;
    cudaMalloc(&out, (size_t)blocks * threads * 8);
    const double bf = (double)blocks * threads * kIlp * kRounds, clk = prop.clockRate * 1e3;
    auto report = [&](const char *name, double ms) {
        const double rate = bf / (ms * 1e-3);
        printf("%-44s %8.4f ms  %.4e butterflies/s  %6.2f /clk/SM\n", name, ms, rate, rate / clk / sms);
    };
    for (uint32_t q : {268369921u /* 28-bit */, 16515073u /* 24-bit */}) {
        const uint32_t off = (q >> 27) ? 7u * q : 64u * q;  // M*q <= 2^31 - q
        printf("q = %u (%d bits), offset %u, host self-check: %d bad of 2000000\n", q, 32 - __builtin_clz(q), off, self_check(q, off));
        const uint32_t w = 123456789u % q, ws = (uint32_t)(((unsigned __int128)w << 32) / q);
        const DqTw tw = make_tw(w, q), one = make_tw(1, q);
        report("integer fwd (csub)", time_ms([&] { int_fwd_kernel<false><<<blocks, threads>>>((uint32_t *)out, w, ws, q, kRounds); }));
        if (q < (1u << 26)) report("integer fwd (nocsub)", time_ms([&] { int_fwd_kernel<true><<<blocks, threads>>>((uint32_t *)out, w, ws, q, kRounds); }));
        report("integer inv", time_ms([&] { int_inv_kernel<<<blocks, threads>>>((uint32_t *)out, w, ws, q, kRounds); }));
        report("double-quotient fwd", time_ms([&] { dq_fwd_kernel<0><<<blocks, threads>>>((int32_t *)out, tw, one, q, off, kRounds); }));
        report("double-quotient inv, sum leg reduced /3", time_ms([&] { dq_inv_kernel<3><<<blocks, threads>>>((int32_t *)out, tw, one, q, off, kRounds); }));
        report("double-quotient inv, sum leg reduced /6", time_ms([&] { dq_inv_kernel<6><<<blocks, threads>>>((int32_t *)out, tw, one, q, off, kRounds); }));
    }
    cudaFree(out);
    return cudaDeviceSynchronize() == cudaSuccess ? 0 : 1;
}
