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

namespace gp {
static thread_local std::string t_last_error;
std::atomic<uint64_t> g_launches{0};

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
typedef unsigned __int128 u128;

uint64_t mulmod(uint64_t a, uint64_t b, uint64_t q) { return (uint64_t)(((u128)a * b) % q); }
uint64_t powmod(uint64_t b, uint64_t e, uint64_t q) {
    uint64_t r = 1 % q, c = b % q;
    for (; e; e >>= 1) {
        if (e & 1) r = mulmod(r, c, q);
        c = mulmod(c, c, q);
    }
    return r;
}
bool invmod(uint64_t a, uint64_t q, uint64_t &out) {
    __int128 t = 0, nt = 1, r = q, nr = a % q;
    while (nr != 0) {
        __int128 k = r / nr, tmp = t - k * nt;
        t = nt; nt = tmp;
        tmp = r - k * nr; r = nr; nr = tmp;
    }
    if (r != 1) return false;
    if (t < 0) t += q;
    out = (uint64_t)t;
    return true;
}
// distinct prime factors of m (trial division; 2 stripped first, odd candidates after)
std::vector<uint64_t> prime_factors(uint64_t m) {
    std::vector<uint64_t> f;
    if ((m & 1) == 0) {
        f.push_back(2);
        while ((m & 1) == 0) m >>= 1;
    }
    for (uint64_t d = 3; d * d <= m; d += 2) {
        if (m % d) continue;
        f.push_back(d);
        while (m % d == 0) m /= d;
    }
    if (m > 1) f.push_back(m);
    return f;
}
// least primitive root — same value the reference finds (Runtime.cu:121-165)
uint64_t least_primitive_root(uint64_t p) {
    const uint64_t phi = p - 1;
    const std::vector<uint64_t> fac = prime_factors(phi);
    for (uint64_t g = 2; g < p; ++g) {
        bool ok = true;
        for (uint64_t f : fac)
            if (powmod(g, phi / f, p) == 1) { ok = false; break; }
        if (ok) return g;
    }
    throw std::runtime_error("failed to find primitive root");
}
uint64_t primitive_2n_root(uint64_t p, uint64_t n) {  // Runtime.cu:167-190
    const uint64_t order = 2 * n;
    if (p % order != 1) throw std::runtime_error("modulus is not congruent to 1 mod 2N");
    const uint64_t root = powmod(least_primitive_root(p), (p - 1) / order, p);
    if (powmod(root, n, p) != p - 1) throw std::runtime_error("computed root is not a primitive 2N-th root");
    return root;
}
uint32_t bitrev(uint32_t x, int bits) { return bits == 0 ? 0 : (__builtin_bitreverse32(x) >> (32 - bits)); }

template <typename T> T shoup(uint64_t w, uint64_t q) { return (T)((((u128)w) << (sizeof(T) * 8)) / q); }

template <typename T>
void build_tables(GpuContext *ctx) {
    const size_t n = (size_t)ctx->N, L = ctx->moduli.size();
    std::vector<gp::TwPair<T>> fwd(L * n), inv(L * n);
    std::vector<gp::LimbConst<T>> lc(L);
    for (size_t l = 0; l < L; ++l) {
        const uint64_t q = ctx->moduli[l], psi = ctx->psi[l];
        uint64_t ipsi = 0, ninv = 0;
        if (!invmod(psi, q, ipsi) || !invmod(n % q, q, ninv)) throw std::runtime_error("failed to invert NTT constants");
        uint64_t w = 1, iw = 1;
        for (size_t i = 0; i < n; ++i) {
            const size_t r = bitrev((uint32_t)i, ctx->logN);
            fwd[l * n + r] = {(T)w, shoup<T>(w, q)};
            inv[l * n + r] = {(T)iw, shoup<T>(iw, q)};
            w = mulmod(w, psi, q);
            iw = mulmod(iw, ipsi, q);
        }
        gp::LimbConst<T> &c = lc[l];
        c.q = (T)q;
        c.q2 = (T)(2 * q);
        c.ninv = (T)ninv;
        c.ninv_s = shoup<T>(ninv, q);
        const uint64_t nw = mulmod(ninv, n > 1 ? (uint64_t)inv[l * n + 1].w : 1, q);
        c.nw = (T)nw;
        c.nw_s = shoup<T>(nw, q);
        const uint64_t r64 = (uint64_t)((((u128)1) << 64) % q);
        c.r64 = (T)r64;
        c.r64_s = shoup<T>(r64, q);
        c.mu64 = (uint64_t)((((u128)1) << 64) / q);
        c.bits = 64 - (uint32_t)__builtin_clzll(q);
        c.pad = 0;
    }
    void *d_fwd = nullptr, *d_inv = nullptr, *d_lc = nullptr;
    auto chk = [](cudaError_t e) { if (e != cudaSuccess) throw std::runtime_error(cudaGetErrorString(e)); };
    chk(cudaMalloc(&d_fwd, fwd.size() * sizeof(fwd[0])));
    ctx->d_fwd = d_fwd;
    chk(cudaMalloc(&d_inv, inv.size() * sizeof(inv[0])));
    ctx->d_inv = d_inv;
    chk(cudaMalloc(&d_lc, lc.size() * sizeof(lc[0])));
    ctx->d_limb = d_lc;
    chk(cudaMemcpy(d_fwd, fwd.data(), fwd.size() * sizeof(fwd[0]), cudaMemcpyHostToDevice));
    chk(cudaMemcpy(d_inv, inv.data(), inv.size() * sizeof(inv[0]), cudaMemcpyHostToDevice));
    chk(cudaMemcpy(d_lc, lc.data(), lc.size() * sizeof(lc[0]), cudaMemcpyHostToDevice));
}

void free_context(GpuContext *ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    if (ctx->d_fwd) cudaFree(ctx->d_fwd);
    if (ctx->d_inv) cudaFree(ctx->d_inv);
    if (ctx->d_limb) cudaFree(ctx->d_limb);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

constexpr size_t kMaxLimbs = 64;  // Runtime.cuh:51
constexpr size_t kMaxGpus = 8;    // Runtime.cuh:52
}  // namespace

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
        ctx->wide = qmax >= (1ull << 30);
        {  // products that fit a lazy accumulator (2^64 resp. 2^128) on top of one folded residue
            const u128 q2 = (u128)(qmax - 1) * (qmax - 1);
            const u128 room = ctx->wide ? (~(u128)0 - qmax) : ((u128)~0ull - qmax);
            u128 terms = room / q2;
            ctx->mac_fold = (uint32_t)std::min<u128>(std::max<u128>(terms, 1), 1u << 16);
        }
        GP_CUDA(cudaSetDevice(ctx->device), "cudaSetDevice");
        cudaDeviceProp prop;
        GP_CUDA(cudaGetDeviceProperties(&prop, ctx->device), "cudaGetDeviceProperties");
        ctx->sm_count = prop.multiProcessorCount;
        ctx->max_smem_optin = (int)prop.sharedMemPerBlockOptin;
        GP_CUDA(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking), "cudaStreamCreate");
        if (ctx->wide) build_tables<uint64_t>(ctx);
        else build_tables<uint32_t>(ctx);
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

}  // extern "C"
