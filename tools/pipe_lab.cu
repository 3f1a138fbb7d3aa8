// pipe_lab.cu — development microbenchmark: issue rates of the integer / FP64 instructions the butterflies are made of,
// alone and interleaved, to decide which pipe bounds a butterfly formulation on B200 (sm_100a).  Register-resident
// dependency chains, 8 independent chains per thread.  Prints warp-instructions per clock per SM for each mix.
//   nvcc -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -o lab/pipe_lab tools/pipe_lab.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

constexpr int kChains = 8, kRounds = 2048;

enum Mix { IMAD_LO, IMAD_HI, IMAD_WIDE, IADD3, VIADDMNMX, LOP3, DFMA, DADD, MIX_IMAD_IADD, MIX_HI_LO_LO, MIX_BFLY, MIX_IMAD_DFMA, MIX_IADD_DFMA, MIX_IMAD_IADD_DFMA, MIX_HI_DFMA, NMIX };
static const char *kNames[NMIX] = {"imad.lo", "imad.hi", "imad.wide", "iadd3", "viaddmnmx(sub+min)", "lop3", "dfma", "dadd", "imad.lo+iadd3 (1:1)", "imad.hi+2 imad.lo", "butterfly hi+lo+lo+3alu", "imad.lo+dfma (1:1)", "iadd3+dfma (1:1)", "imad.lo+iadd3+dfma (1:1:1)", "imad.hi+dfma (1:1)"};
static const int kInstrPerRound[NMIX] = {1, 1, 1, 2, 1, 1, 1, 1, 2, 3, 6, 2, 4, 3, 2};

template <int M> __global__ void __launch_bounds__(256) k(uint32_t *out, uint32_t a, uint32_t b, double da, double db, int rounds) {
    uint32_t x[kChains], y[kChains];
    uint64_t w[kChains];
    double d[kChains];
#pragma unroll
    for (int i = 0; i < kChains; i++) {
        x[i] = threadIdx.x * 2654435761u + i * 40503u + blockIdx.x;
        y[i] = x[i] ^ 0x5bd1e995u;
        w[i] = x[i];
        d[i] = (double)(x[i] & 0xffff);
    }
    for (int r = 0; r < rounds; r++) {
#pragma unroll
        for (int i = 0; i < kChains; i++) {
            if (M == IMAD_LO) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x[i]) : "r"(a), "r"(b));
            if (M == IMAD_HI) asm volatile("mad.hi.u32 %0, %0, %1, %2;" : "+r"(x[i]) : "r"(a), "r"(b));
            if (M == IMAD_WIDE) {
                const uint32_t lo = (uint32_t)w[i];
                asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(w[i]) : "r"(lo), "r"(a));
            }
            if (M == IADD3) {
                asm volatile("add.u32 %0, %0, %1;" : "+r"(x[i]) : "r"(y[i]));
                asm volatile("add.u32 %0, %0, %1;" : "+r"(y[i]) : "r"(x[i]));
            }
            if (M == VIADDMNMX) {
                uint32_t t;
                asm volatile("sub.u32 %0, %1, %2;" : "=r"(t) : "r"(x[i]), "r"(a));
                asm volatile("min.u32 %0, %1, %2;" : "=r"(x[i]) : "r"(t), "r"(x[i]));
            }
            if (M == LOP3) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[i]) : "r"(a), "r"(b));
            if (M == DFMA) asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d[i]) : "d"(da), "d"(db));
            if (M == DADD) asm volatile("add.rn.f64 %0, %0, %1;" : "+d"(d[i]) : "d"(da));
            if (M == MIX_IMAD_IADD) {
                asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x[i]) : "r"(a), "r"(b));
                asm volatile("add.u32 %0, %0, %1;" : "+r"(y[i]) : "r"(x[i]));
            }
            if (M == MIX_HI_LO_LO) {
                uint32_t h, t;
                asm volatile("mul.hi.u32 %0, %1, %2;" : "=r"(h) : "r"(x[i]), "r"(b));
                asm volatile("mul.lo.u32 %0, %1, %2;" : "=r"(t) : "r"(h), "r"(a));
                asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x[i]) : "r"(b), "r"(t));
            }
            if (M == MIX_BFLY) {
                uint32_t h, t, X;
                asm volatile("mul.hi.u32 %0, %1, %2;" : "=r"(h) : "r"(y[i]), "r"(b));
                asm volatile("mul.lo.u32 %0, %1, %2;" : "=r"(t) : "r"(h), "r"(a));
                asm volatile("mad.lo.u32 %0, %1, %2, %0;" : "+r"(t) : "r"(y[i]), "r"(b));
                asm volatile("sub.u32 %0, %1, %2;" : "=r"(X) : "r"(x[i]), "r"(a));
                asm volatile("min.u32 %0, %1, %2;" : "=r"(X) : "r"(X), "r"(x[i]));
                asm volatile("add.u32 %0, %1, %2;" : "=r"(x[i]) : "r"(X), "r"(t));
                asm volatile("sub.u32 %0, %1, %2;" : "=r"(y[i]) : "r"(X), "r"(t));
            }
            if (M == MIX_IMAD_DFMA) {
                asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x[i]) : "r"(a), "r"(b));
                asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d[i]) : "d"(da), "d"(db));
            }
            if (M == MIX_IADD_DFMA) {
                asm volatile("add.u32 %0, %0, %1;" : "+r"(x[i]) : "r"(y[i]));
                asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d[i]) : "d"(da), "d"(db));
                asm volatile("add.u32 %0, %0, %1;" : "+r"(y[i]) : "r"(x[i]));
                asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d[i]) : "d"(da), "d"(db));
            }
            if (M == MIX_IMAD_IADD_DFMA) {
                asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x[i]) : "r"(a), "r"(b));
                asm volatile("add.u32 %0, %0, %1;" : "+r"(y[i]) : "r"(x[i]));
                asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d[i]) : "d"(da), "d"(db));
            }
            if (M == MIX_HI_DFMA) {
                asm volatile("mad.hi.u32 %0, %0, %1, %2;" : "+r"(x[i]) : "r"(a), "r"(b));
                asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d[i]) : "d"(da), "d"(db));
            }
        }
    }
    uint32_t acc = 0;
#pragma unroll
    for (int i = 0; i < kChains; i++) acc ^= x[i] ^ y[i] ^ (uint32_t)w[i] ^ (uint32_t)(w[i] >> 32) ^ (uint32_t)__double_as_longlong(d[i]);
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <int M> static void run(uint32_t *out, int sms, double clock_hz) {
    const int blocks = sms * 8, threads = 256;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    float best = 1e30f;
    for (int r = 0; r < 5; r++) {
        cudaEventRecord(e0);
        k<M><<<blocks, threads>>>(out, 2654435761u, 40503u, 1.0000001, 0.5, kRounds);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (r >= 1 && ms < best) best = ms;
    }
    const double warp_instr = (double)blocks * (threads / 32) * kChains * kRounds * kInstrPerRound[M];
    const double per_clk_sm = warp_instr / (best * 1e-3 * clock_hz) / sms;
    printf("%-32s %8.4f ms  %6.3f warp-instr/clk/SM  (%.2f clk per warp-instr per SMSP)\n", kNames[M], best, per_clk_sm, 4.0 / per_clk_sm);
    if constexpr (M + 1 < NMIX) run<M + 1>(out, sms, clock_hz);
}

int main() {
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, 0) != cudaSuccess) return printf("no device\n"), 1;
    uint32_t *out;
    cudaMalloc(&out, (size_t)prop.multiProcessorCount * 8 * 256 * 4);
    printf("device %s, %d SMs, clock %d MHz (rates assume the max clock)\n", prop.name, prop.multiProcessorCount, prop.clockRate / 1000);
    run<0>(out, prop.multiProcessorCount, prop.clockRate * 1e3);
    return cudaDeviceSynchronize() == cudaSuccess ? 0 : 1;
}
