// Runtime.cu — context, error channel, events, device queries, pinned memory.
//
// Replaces cuda/src/Runtime.cu of MachinaIO/mxx behind the same C ABI
// (include/gpupoly.h; reference prototypes cuda/include/Runtime.cuh:20-44).
// What gpu_context_create computes follows the reference's definitions
// (least primitive root, psi = g^((p-1)/2n): Runtime.cu:121-190; n^-1: :282-319) so
// the EVAL layout at the boundary is identical; how it is stored is new: per-context
// device tables of psi^bitrev(i) / psi^-bitrev(i) with Shoup companions, one stream.
#include <algorithm>
#include <cstring>
#include <exception>
#include <mutex>
#include <stdexcept>
#include <unordered_set>

#include "gpupoly_internal.cuh"
#include "host_tables.hpp"

namespace gp {
static thread_local std::string t_last_error;
std::atomic<uint64_t> g_launches{0};
thread_local cudaStream_t tl_stream = nullptr;

int set_error(const char *msg) {
    t_last_error = msg ? msg : "unknown error";
    return 1;
}
int set_error(const std::string &msg) {
    t_last_error = msg;
    return 1;
}
int set_cuda_error(cudaError_t err, const char *where) {
    t_last_error = std::string(where ? where : "cuda") + ": " + cudaGetErrorString(err);
    return 1;
}
}  // namespace gp

namespace {
using namespace gp::host;

template <typename T>
void upload_tables(GpuContext *ctx) {
    const Tables<T> t = build_tables<T>(ctx->logN, ctx->moduli, ctx->psi);
    void *d_fwd = nullptr, *d_inv = nullptr, *d_lc = nullptr;
    auto chk = [](cudaError_t e) { if (e != cudaSuccess) throw std::runtime_error(cudaGetErrorString(e)); };
    chk(cudaMalloc(&d_fwd, t.fwd.size() * sizeof(t.fwd[0])));
    ctx->d_fwd = d_fwd;
    chk(cudaMalloc(&d_inv, t.inv.size() * sizeof(t.inv[0])));
    ctx->d_inv = d_inv;
    chk(cudaMalloc(&d_lc, t.limb.size() * sizeof(t.limb[0])));
    ctx->d_limb = d_lc;
    chk(cudaMemcpy(d_fwd, t.fwd.data(), t.fwd.size() * sizeof(t.fwd[0]), cudaMemcpyHostToDevice));
    chk(cudaMemcpy(d_inv, t.inv.data(), t.inv.size() * sizeof(t.inv[0]), cudaMemcpyHostToDevice));
    chk(cudaMemcpy(d_lc, t.limb.data(), t.limb.size() * sizeof(t.limb[0]), cudaMemcpyHostToDevice));
    if constexpr (sizeof(T) == 8) {
        uint64_t qmax = 0;
        for (uint64_t q : ctx->moduli) qmax = std::max(qmax, q);
        const char *e = getenv("MXX_NTT_F64");
        if (gp::host::f64_ok(qmax, true) && !(e && atoi(e) == 0) && !ctx->force_generic_ntt) {
            const auto f = gp::host::f64_tables(t);
            chk(cudaMalloc(&ctx->d_fwd_f64, f.fwd.size() * sizeof(f.fwd[0])));
            chk(cudaMalloc(&ctx->d_inv_f64, f.inv.size() * sizeof(f.inv[0])));
            chk(cudaMalloc(&ctx->d_limb_f64, f.limb.size() * sizeof(f.limb[0])));
            chk(cudaMemcpy(ctx->d_fwd_f64, f.fwd.data(), f.fwd.size() * sizeof(f.fwd[0]), cudaMemcpyHostToDevice));
            chk(cudaMemcpy(ctx->d_inv_f64, f.inv.data(), f.inv.size() * sizeof(f.inv[0]), cudaMemcpyHostToDevice));
            chk(cudaMemcpy(ctx->d_limb_f64, f.limb.data(), f.limb.size() * sizeof(f.limb[0]), cudaMemcpyHostToDevice));
            ctx->f64_ntt = true;
        }
    }
}

std::mutex g_live_mutex;
std::unordered_set<const GpuContext *> g_live;

void free_context(GpuContext *ctx) {
    if (!ctx) return;
    {
        std::lock_guard<std::mutex> g(g_live_mutex);
        g_live.erase(ctx);
    }
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    if (ctx->load_pipe && ctx->load_pipe_free) ctx->load_pipe_free(ctx->load_pipe);
    if (ctx->d_fwd) cudaFree(ctx->d_fwd);
    if (ctx->d_inv) cudaFree(ctx->d_inv);
    if (ctx->d_limb) cudaFree(ctx->d_limb);
    if (ctx->d_garner) cudaFree(ctx->d_garner);
    if (ctx->d_fwd_f64) cudaFree(ctx->d_fwd_f64);
    if (ctx->d_inv_f64) cudaFree(ctx->d_inv_f64);
    if (ctx->d_limb_f64) cudaFree(ctx->d_limb_f64);
    for (int i = 0; i < 2; i++) {
        if (ctx->aux[i]) cudaStreamSynchronize(ctx->aux[i]);
        if (ctx->aux[i]) cudaStreamDestroy(ctx->aux[i]);
        if (ctx->ev_join[i]) cudaEventDestroy(ctx->ev_join[i]);
    }
    if (ctx->ev_fork) cudaEventDestroy(ctx->ev_fork);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

constexpr size_t kMaxLimbs = 64;  // Runtime.cuh:51
constexpr size_t kMaxGpus = 8;    // Runtime.cuh:52
}  // namespace

namespace gp {
bool context_alive(const GpuContext *ctx) {
    std::lock_guard<std::mutex> g(g_live_mutex);
    return g_live.count(ctx) != 0;
}
}  // namespace gp

extern "C" {

int gpu_context_create(uint32_t logN, uint32_t L, uint32_t dnum, const uint64_t *moduli, size_t moduli_len,
                       const int *gpu_ids, size_t gpu_ids_len, GpuContext **out_ctx) {
    GpuContext *ctx = nullptr;
    try {
        // validation rules of the reference: Runtime.cu:538-575, 192-227
        if (!out_ctx || !moduli || moduli_len == 0) return gp::set_error("invalid context arguments");
        *out_ctx = nullptr;
        if (moduli_len != (size_t)L + 1) return gp::set_error("moduli_len must equal L + 1");
        if (moduli_len > kMaxLimbs) return gp::set_error("moduli_len exceeds supported maximum");
        if (logN == 0 || logN >= 31) return gp::set_error("logN must be between 1 and 30");
        std::vector<int> gpus;
        if (gpu_ids_len == 0 || !gpu_ids) gpus.push_back(0);
        else gpus.assign(gpu_ids, gpu_ids + gpu_ids_len);
        if (gpus.size() > kMaxGpus) return gp::set_error("gpu count exceeds supported maximum");
        int device_count = 0;
        GP_CUDA(cudaGetDeviceCount(&device_count), "cudaGetDeviceCount");
        if (device_count <= 0) return gp::set_error("no CUDA device available");
        std::unordered_set<int> seen;
        for (int id : gpus) {
            if (id < 0 || id >= device_count) return gp::set_error("invalid gpu id in context creation");
            if (!seen.insert(id).second) return gp::set_error("duplicate gpu id in context creation");
        }
        const uint32_t resolved_dnum = dnum == 0 ? (uint32_t)gpus.size() : dnum;
        if (resolved_dnum == 0 || resolved_dnum > kMaxGpus) return gp::set_error("invalid dnum in context creation");

        ctx = new GpuContext();
        ctx->logN = (int)logN;
        ctx->N = 1 << logN;
        ctx->level = (int)L;
        ctx->moduli.assign(moduli, moduli + moduli_len);
        ctx->gpu_ids = gpus;
        ctx->dnum = resolved_dnum;
        // All limbs of a matrix live on ONE device, as in the reference's effective mode
        // ("single-device mode requires all limbs on one device", MatrixNTT.cu:495-498);
        // multi-GPU runs shard columns over processes (mxx_b200/dist.py).
        ctx->device = gpus[0];
        uint64_t qmax = 0;
        for (size_t i = 0; i < moduli_len; ++i) {
            const uint64_t q = moduli[i];
            if (q < 3) throw std::runtime_error("zero modulus in gpu context creation");
            if (q >> 62) throw std::runtime_error("modulus must be below 2^62 in gpu context creation");
            for (size_t j = 0; j < i; ++j)
                if (moduli[j] == q) throw std::runtime_error("CRT moduli must be pairwise coprime");
            qmax = std::max(qmax, q);
            ctx->psi.push_back(primitive_2n_root(q, (uint64_t)ctx->N));
        }
        ctx->crt_bits = 64 - (uint32_t)__builtin_clzll(qmax);
        ctx->wide = needs_wide(qmax);
        ctx->mac_fold = mac_fold_terms(qmax, ctx->wide);
        ctx->f64_mac_fold = f64_mac_fold_terms(qmax);
        if (const char *env = getenv("MXX_MATMUL_F64")) ctx->f64_matmul = env[0] != '0';
        ctx->nocsub = nocsub_ok(qmax, ctx->wide);
        if (const char *env = getenv("MXX_NTT_NOCSUB")) ctx->nocsub = ctx->nocsub && env[0] != '0';  // A/B: keep the conditional subtractions
        if (const char *env = getenv("MXX_NTT_GENERIC")) ctx->force_generic_ntt = env[0] == '1';
        if (const char *env = getenv("MXX_DECOMP_FUSE_TOP")) ctx->fuse_decompose_top = env[0] != '0';
        if (const char *env = getenv("MXX_NTT_BLOCK_LOG")) ctx->ntt_block_override = atoi(env);
        if (const char *env = getenv("MXX_NTT_CLUSTER")) ctx->ntt_cluster = env[0] != '0';
        GP_CUDA(cudaSetDevice(ctx->device), "cudaSetDevice");
        cudaDeviceProp prop;
        GP_CUDA(cudaGetDeviceProperties(&prop, ctx->device), "cudaGetDeviceProperties");
        ctx->sm_count = prop.multiProcessorCount;
        ctx->max_smem_optin = (int)prop.sharedMemPerBlockOptin;
        GP_CUDA(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking), "cudaStreamCreate");
        for (int i = 0; i < 2; i++) {
            GP_CUDA(cudaStreamCreateWithFlags(&ctx->aux[i], cudaStreamNonBlocking), "cudaStreamCreate(aux)");
            GP_CUDA(cudaEventCreateWithFlags(&ctx->ev_join[i], cudaEventDisableTiming), "cudaEventCreate(aux)");
        }
        GP_CUDA(cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming), "cudaEventCreate(fork)");
        if (ctx->wide) upload_tables<uint64_t>(ctx);
        else upload_tables<uint32_t>(ctx);
        {
            std::lock_guard<std::mutex> g(g_live_mutex);
            g_live.insert(ctx);
        }
        *out_ctx = ctx;
        return 0;
    } catch (const std::exception &e) {
        free_context(ctx);
        return gp::set_error(e.what());
    } catch (...) {
        free_context(ctx);
        return gp::set_error("unknown exception in gpu_context_create");
    }
}

void gpu_context_destroy(GpuContext *ctx) { free_context(ctx); }

int gpu_context_get_N(const GpuContext *ctx, int *out_N) {
    if (!ctx || !out_N) return gp::set_error("invalid gpu_context_get_N arguments");
    *out_N = ctx->N;
    return 0;
}

void *gpu_context_stream(const GpuContext *ctx) { return ctx ? (void *)ctx->stream : nullptr; }

// extension: the CUDA stream the CALLING THREAD's kernel launches of this library go to (nullptr: back to each context's own
// stream).  Lets a caller run an operation on a side stream of its own — the operand exchange of a limb-sharded gate one
// gate ahead of the product (mxx_b200/dist.py) — with the ordering against the context stream done by the caller's events.
// Matrices are still allocated on the context stream: only operations on buffers that already exist belong on a side stream.
int gpu_thread_stream_override(void *stream) {
    gp::tl_stream = (cudaStream_t)stream;
    return 0;
}

int gpu_event_set_wait(GpuEventSet *events) {
    if (!events) return gp::set_error("invalid gpu_event_set_wait arguments");
    if (events->event) {
        GP_CUDA(cudaSetDevice(events->device), "cudaSetDevice");
        GP_CUDA(cudaEventSynchronize(events->event), "cudaEventSynchronize");
    }
    return 0;
}

void gpu_event_set_destroy(GpuEventSet *events) {
    if (!events) return;
    if (events->event) {
        cudaSetDevice(events->device);
        cudaEventDestroy(events->event);
    }
    delete events;
}

int gpu_device_count(int *out_count) {
    if (!out_count) return gp::set_error("invalid gpu_device_count arguments");
    int count = 0;
    cudaError_t err = cudaGetDeviceCount(&count);
    if (err == cudaErrorNoDevice) {
        *out_count = 0;
        return 0;
    }
    if (err != cudaSuccess) return gp::set_cuda_error(err, "cudaGetDeviceCount");
    *out_count = count;
    return 0;
}

int gpu_device_mem_info(int device, size_t *out_free, size_t *out_total) {
    if (!out_free || !out_total) return gp::set_error("invalid gpu_device_mem_info arguments");
    int current = 0;
    GP_CUDA(cudaGetDevice(&current), "cudaGetDevice");
    GP_CUDA(cudaSetDevice(device), "cudaSetDevice");
    size_t f = 0, t = 0;
    cudaError_t err = cudaMemGetInfo(&f, &t);
    cudaError_t restore = cudaSetDevice(current);
    if (err != cudaSuccess) return gp::set_cuda_error(err, "cudaMemGetInfo");
    if (restore != cudaSuccess) return gp::set_cuda_error(restore, "cudaSetDevice");
    *out_free = f;
    *out_total = t;
    return 0;
}

int gpu_device_synchronize(void) {
    GP_CUDA(cudaDeviceSynchronize(), "cudaDeviceSynchronize");
    return 0;
}

int gpu_device_reset(void) {
    GP_CUDA(cudaDeviceReset(), "cudaDeviceReset");
    return 0;
}

const char *gpu_last_error(void) { return gp::t_last_error.c_str(); }

void *gpu_pinned_alloc(size_t bytes) {
    if (bytes == 0) return nullptr;
    void *ptr = nullptr;
    cudaError_t err = cudaMallocHost(&ptr, bytes);
    if (err != cudaSuccess) {
        gp::set_cuda_error(err, "cudaMallocHost");
        return nullptr;
    }
    return ptr;
}

void gpu_pinned_free(void *ptr) {
    if (!ptr) return;
    cudaError_t err = cudaFreeHost(ptr);
    if (err != cudaSuccess) gp::set_cuda_error(err, "cudaFreeHost");
}

uint64_t gpu_kernel_launch_count(void) { return gp::g_launches.load(); }

// ---- extension: CUDA-event stopwatch on the context stream (bench.py times kernels with it) ----
struct GpuTimer {
    cudaEvent_t start = nullptr, stop = nullptr;
    int device = 0;
};
int gpu_timer_create(const GpuContext *ctx, GpuTimer **out) {
    if (!ctx || !out) return gp::set_error("invalid gpu_timer_create arguments");
    GP_CUDA(cudaSetDevice(ctx->device), "cudaSetDevice");
    GpuTimer *t = new GpuTimer();
    t->device = ctx->device;
    cudaError_t e = cudaEventCreate(&t->start);
    if (e == cudaSuccess) e = cudaEventCreate(&t->stop);
    if (e != cudaSuccess) {
        if (t->start) cudaEventDestroy(t->start);
        delete t;
        return gp::set_cuda_error(e, "cudaEventCreate");
    }
    *out = t;
    return 0;
}
int gpu_timer_record(const GpuContext *ctx, GpuTimer *t, int stop) {
    if (!ctx || !t) return gp::set_error("invalid gpu_timer_record arguments");
    GP_CUDA(cudaSetDevice(ctx->device), "cudaSetDevice");
    GP_CUDA(cudaEventRecord(stop ? t->stop : t->start, ctx->stream), "cudaEventRecord");
    return 0;
}
int gpu_timer_elapsed_ms(GpuTimer *t, float *out_ms) {
    if (!t || !out_ms) return gp::set_error("invalid gpu_timer_elapsed_ms arguments");
    GP_CUDA(cudaSetDevice(t->device), "cudaSetDevice");
    GP_CUDA(cudaEventSynchronize(t->stop), "cudaEventSynchronize");
    GP_CUDA(cudaEventElapsedTime(out_ms, t->start, t->stop), "cudaEventElapsedTime");
    return 0;
}
void gpu_timer_destroy(GpuTimer *t) {
    if (!t) return;
    cudaSetDevice(t->device);
    cudaEventDestroy(t->start);
    cudaEventDestroy(t->stop);
    delete t;
}
int gpu_context_synchronize(const GpuContext *ctx) {
    if (!ctx) return gp::set_error("invalid gpu_context_synchronize arguments");
    GP_CUDA(cudaSetDevice(ctx->device), "cudaSetDevice");
    GP_CUDA(cudaStreamSynchronize(ctx->stream), "cudaStreamSynchronize");
    return 0;
}

}  // extern "C"
