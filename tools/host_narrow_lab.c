// host_narrow_lab.c — how fast can the host cores narrow the ABI's u64 residue words to u32?
// (decides whether gpu_matrix_load_rns_batch should halve its PCIe bytes on the host side)
// build: gcc -O3 -pthread tools/host_narrow_lab.c -o lab/host_narrow_lab
#include <pthread.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

__attribute__((target_clones("avx512f", "avx2", "default"))) static void narrow(const uint64_t *src, uint32_t *dst, size_t n) {
    for (size_t i = 0; i < n; i++) dst[i] = (uint32_t)src[i];
}
typedef struct { const uint64_t *src; uint32_t *dst; size_t n; } Job;
static void *work(void *p) { Job *j = p; narrow(j->src, j->dst, j->n); return NULL; }
static double now(void) { struct timespec t; clock_gettime(CLOCK_MONOTONIC, &t); return t.tv_sec + 1e-9 * t.tv_nsec; }

int main(void) {
    const size_t words = (size_t)1 << 28;  // 2 GiB in, 1 GiB out
    uint64_t *src = aligned_alloc(4096, words * 8);
    uint32_t *dst = aligned_alloc(4096, words * 4);
    for (size_t i = 0; i < words; i++) src[i] = i * 2654435761u & 0xffffff;
    memset(dst, 0, words * 4);
    for (int T = 1; T <= 32; T *= 2) {
        pthread_t th[32]; Job jobs[32];
        double best = 1e9;
        for (int rep = 0; rep < 3; rep++) {
            double t0 = now();
            for (int t = 0; t < T; t++) {
                size_t a = words * t / T, b = words * (t + 1) / T;
                jobs[t] = (Job){src + a, dst + a, b - a};
                pthread_create(&th[t], NULL, work, &jobs[t]);
            }
            for (int t = 0; t < T; t++) pthread_join(th[t], NULL);
            double dt = now() - t0;
            if (dt < best) best = dt;
        }
        printf("threads %2d: %.1f GB/s of u64 input (%.1f ms per GiB in)\n", T, words * 8 / best / 1e9, best * 1e3 / 2);
    }
    printf("check %u\n", dst[12345]);
    return 0;
}
