// gpupoly_internal.cuh — host-side structures behind the opaque C-ABI handles.
//
// Replaces cuda/include/Runtime.cuh:77-106 and cuda/include/matrix/Matrix.cuh:34-69.
// Design differences (B200-first):
//   * ONE stream per context instead of one stream + event per limb per matrix;
//   * storage is word-typed (u32 when all moduli < 2^30, else u64), laid out
//     [poly][limb][coeff] contiguously — the host layout, narrowed — instead of
//     byte-packed limbs with a 2x over-allocation (MatrixData.cu:296-316);
//   * all NTT constants are per-context device tables, not device-global __constant__
//     arrays (the reference's two-contexts-one-device hazard, MatrixNTT.cu:1-5);
//   * EVAL data is kept in bit-reversed slot order on the device; the natural order
//     the boundary defines is produced/consumed only by load/store_rns_batch.
#pragma once
#include <stddef.h>
#include <stdint.h>

#include <atomic>
#include <mutex>
#include <string>
#include <vector>

#include <cuda_runtime.h>

#include "../../../include/gpupoly.h"
#include "modarith.cuh"

namespace gp {
int set_error(const char *msg);
int set_error(const std::string &msg);
int set_cuda_error(cudaError_t err, const char *where);
extern std::atomic<uint64_t> g_launches;
inline void count_launch(uint64_t n = 1) { g_launches.fetch_add(n, std::memory_order_relaxed); }
// Stream the launch helpers use: the context stream unless the calling thread has scoped an
// override (a context is shared by many host threads, so the override is thread-local).
extern thread_local cudaStream_t tl_stream;
struct StreamScope {
    cudaStream_t prev;
    explicit StreamScope(cudaStream_t s) : prev(tl_stream) { tl_stream = s; }
    ~StreamScope() { tl_stream = prev; }
};
}  // namespace gp

#define GP_CUDA(call, where)                                         \
    do {                                                             \
        cudaError_t _e = (call);                                     \
        if (_e != cudaSuccess) return gp::set_cuda_error(_e, where); \
    } while (0)

struct GpuContext {
    int N = 0;
    int logN = 0;
    int level = -1;  // limbs - 1
    std::vector<uint64_t> moduli;
    std::vector<uint64_t> psi;  // per limb
    std::vector<int> gpu_ids;
    uint32_t dnum = 1;
    int device = 0;
    bool wide = false;       // true: u64 storage, false: u32 storage
    uint32_t crt_bits = 0;   // widest modulus
    cudaStream_t stream = nullptr;
    // two auxiliary streams + events: independent column chunks of mul_decompose run on them so
    // that one chunk's tail wave overlaps the next chunk's kernels (MatrixDecompose.inl)
    cudaStream_t aux[2] = {nullptr, nullptr};
    cudaEvent_t ev_fork = nullptr, ev_join[2] = {nullptr, nullptr};
    int sm_count = 0;
    int max_smem_optin = 0;
    // device tables; element type follows `wide`
    void *d_fwd = nullptr;   // [L][N] TwPair: psi^bitrev(i)
    void *d_inv = nullptr;   // [L][N] TwPair: psi^-bitrev(i)
    void *d_limb = nullptr;  // [L] LimbConst
    // FP64-pipe NTT tables (TwPair<F64>, LimbConst<F64>), only for u64-word contexts with every modulus < 2^51
    void *d_fwd_f64 = nullptr, *d_inv_f64 = nullptr, *d_limb_f64 = nullptr;
    bool f64_ntt = false;  // MXX_NTT_F64=0 keeps the integer butterflies
    void *d_garner = nullptr;  // [L][L] u64: q_j^-1 mod q_i (compact wire format), built on first use
    std::mutex lazy_mutex;
    bool nocsub = false;             // forward NTT without conditional subtractions (36*qmax fits the word)
    bool fuse_decompose_top = true;  // MXX_DECOMP_FUSE_TOP=0: separate digit kernel + first global pass (A/B)
    bool force_generic_ntt = false;  // MXX_NTT_GENERIC=1: use only the runtime-planned kernels of ntt.cuh
    int ntt_block_override = 0;      // MXX_NTT_BLOCK_LOG
    bool ntt_cluster = true;         // MXX_NTT_CLUSTER=0: rings above one block go back to global passes + block kernel (A/B)
    bool smem_configured = false;  // >48 KB dynamic shared memory opted in for the NTT kernels
    bool f64_matmul = true;      // MXX_MATMUL_F64=0: keep the 128-bit integer MAC in FP64-pipe contexts (A/B)
    uint32_t f64_mac_fold = 1;   // reduced terms between renormalisations of the FP64 accumulator
    uint32_t mac_fold = 1;   // products that fit one lazy accumulator between folds
    // chunked host->device load pipeline (MatrixHostLoad.inl), built on first use
    void *load_pipe = nullptr;
    void (*load_pipe_free)(void *) = nullptr;
    size_t word_bytes() const { return wide ? 8 : 4; }
};

namespace gp {
inline cudaStream_t cur(const GpuContext *ctx) { return tl_stream ? tl_stream : ctx->stream; }
// Contexts that are alive.  A handle that outlives its context (host-language finalisers running in arbitrary order at
// process exit) must not touch the freed context or its stream: *_destroy consult this before any CUDA call.
bool context_alive(const GpuContext *ctx);
}  // namespace gp

struct GpuMatrix {
    GpuContext *ctx = nullptr;
    size_t rows = 0, cols = 0;
    int level = -1;
    int format = GPU_POLY_FORMAT_EVAL;
    void *data = nullptr;  // [rows*cols][level+1][N] words
    size_t limbs() const { return level < 0 ? 0 : (size_t)(level + 1); }
    size_t polys() const { return rows * cols; }
    size_t poly_words() const { return limbs() * (size_t)ctx->N; }
    size_t bytes() const { return polys() * poly_words() * ctx->word_bytes(); }
};

struct GpuEventSet {
    cudaEvent_t event = nullptr;
    int device = 0;
    void *staging = nullptr;       // device staging freed (stream-ordered) at creation; kept for debugging only
    void *host_bounce = nullptr;   // pinned bounce buffer to release after the wait
    // deferred host post-processing after the event completes (store paths)
    void (*post)(GpuEventSet *) = nullptr;
    void *post_a = nullptr, *post_b = nullptr;
    size_t post_n0 = 0, post_n1 = 0, post_n2 = 0, post_n3 = 0;
};
