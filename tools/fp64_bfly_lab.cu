// fp64_bfly_lab.cu — is the FP64 pipe a faster home for the 51-bit (u64-storage) butterfly than the
// integer pipe?  Register-resident chains, no memory traffic, same harness as tools/modmul_peak.cu.
//   modmul in doubles (exact for q < 2^50, |y| < 2^51):  h = y*w; l = fma(y,w,-h);
//   b = rint(y * (w/q)) via the 1.5*2^52 trick; c = fma(-b,q,h); t = c + l  in (-q, q) + rounding slack
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -Imxx_b200/csrc/include tools/fp64_bfly_lab.cu -o lab/fp64_bfly_lab
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

#include "modarith.cuh"
#include "ntt_fast.cuh"
using namespace gp;

constexpr int kIlp = 8;
constexpr int kRounds = 512;

__device__ __forceinline__ double mulmod_f64(double y, double w, double wq, double q) {
    const double magic = 6755399441055744.0;  // 1.5 * 2^52
    const double h = y * w;
    const double l = fma(y, w, -h);
    const double b = fma(y, wq, magic) - magic;
    const double c = fma(-b, q, h);
    return c + l;
}

// forward butterfly with one renormalisation of the X leg per stage (values stay in (-2q, 2q))
template <int RENORM>
__global__ void __launch_bounds__(256) bfly_f64_kernel(double *out, double w, double wq, double q, double qinv, int rounds) {
    double x[2 * kIlp];
#pragma unroll
    for (int i = 0; i < 2 * kIlp; i++) x[i] = (double)((threadIdx.x * 2654435761u + i * 40503u + blockIdx.x) % 1000003u);
    const double magic = 6755399441055744.0;
    for (int r = 0; r < rounds; r++) {
#pragma unroll
        for (int i = 0; i < kIlp; i++) {
            double X = x[2 * i];
            if (RENORM == 1) {  // X -= q * rint(X / q)
                const double k = fma(X, qinv, magic) - magic;
                X = fma(-k, q, X);
            } else if (RENORM == 2) {  // conditional: X in (-2q,2q) -> (-q,q)
                X = X >= q ? X - q : X;
                X = X <= -q ? X + q : X;
            }
            const double t = mulmod_f64(x[2 * i + 1], w, wq, q);
            x[2 * i] = X + t;
            x[2 * i + 1] = X - t;
        }
    }
    double acc = 0;
#pragma unroll
    for (int i = 0; i < 2 * kIlp; i++) acc += x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

// mixed: the same thread block runs u64 integer butterflies and FP64 butterflies side by side (do the pipes overlap?)
__global__ void __launch_bounds__(256) bfly_mixed_kernel(double *out, double w, double wq, double q, double qinv, uint64_t wi, uint64_t wsi,
                                                         uint64_t qi, int rounds) {
    double x[kIlp];
    uint64_t y[kIlp];
#pragma unroll
    for (int i = 0; i < kIlp; i++) {
        x[i] = (double)((threadIdx.x * 2654435761u + i * 40503u + blockIdx.x) % 1000003u);
        y[i] = (uint64_t)(threadIdx.x * 2654435761u + i * 40503u + blockIdx.x) % qi;
    }
    const double magic = 6755399441055744.0;
    const uint64_t q2 = 2 * qi;
    for (int r = 0; r < rounds; r++) {
#pragma unroll
        for (int i = 0; i < kIlp / 2; i++) {
            double X = x[2 * i];
            const double k = fma(X, qinv, magic) - magic;
            X = fma(-k, q, X);
            const double t = mulmod_f64(x[2 * i + 1], w, wq, q);
            x[2 * i] = X + t;
            x[2 * i + 1] = X - t;
            const uint64_t A = csub_min<uint64_t>(y[2 * i], q2);
            const uint64_t tt = mul_shoup_lazy<uint64_t>(y[2 * i + 1], wi, wsi, qi);
            y[2 * i] = A + tt;
            y[2 * i + 1] = A - tt + q2;
        }
    }
    double acc = 0;
#pragma unroll
    for (int i = 0; i < kIlp; i++) acc += x[i] + (double)y[i];
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

int main() {
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, 0) != cudaSuccess) return 1;
    const int sms = prop.multiProcessorCount, threads = 256, blocks = sms * 8;
    double *out;
    cudaMalloc(&out, (size_t)blocks * threads * 8);
    const uint64_t q64 = 562949953392641ull;  // 49-bit-ish odd modulus (value only matters for timing)
    const uint64_t w64 = 1234567890123ull, ws64 = (uint64_t)(((unsigned __int128)w64 << 64) / q64);
    const double q = (double)q64, w = (double)w64, wq = w / q, qinv = 1.0 / q;
    const double bf = (double)blocks * threads * kIlp * kRounds;
    double t;
    t = time_ms([&] { bfly_f64_kernel<0><<<blocks, threads>>>(out, w, wq, q, qinv, kRounds); });
    printf("fp64 butterfly, no renorm      : %.3e butterflies/s (%.2f /clk/SM)\n", bf / (t * 1e-3), bf / (t * 1e-3) / sms / (prop.clockRate * 1e3));
    t = time_ms([&] { bfly_f64_kernel<1><<<blocks, threads>>>(out, w, wq, q, qinv, kRounds); });
    printf("fp64 butterfly, rint renorm    : %.3e butterflies/s (%.2f /clk/SM)\n", bf / (t * 1e-3), bf / (t * 1e-3) / sms / (prop.clockRate * 1e3));
    t = time_ms([&] { bfly_f64_kernel<2><<<blocks, threads>>>(out, w, wq, q, qinv, kRounds); });
    printf("fp64 butterfly, cond renorm    : %.3e butterflies/s (%.2f /clk/SM)\n", bf / (t * 1e-3), bf / (t * 1e-3) / sms / (prop.clockRate * 1e3));
    t = time_ms([&] { bfly_mixed_kernel<<<blocks, threads>>>(out, w, wq, q, qinv, w64, ws64, q64, kRounds); });
    printf("mixed fp64 + u64 (half each)   : %.3e butterflies/s (%.2f /clk/SM)\n", bf / (t * 1e-3), bf / (t * 1e-3) / sms / (prop.clockRate * 1e3));
    cudaFree(out);
    return cudaDeviceSynchronize() == cudaSuccess ? 0 : 1;
}
