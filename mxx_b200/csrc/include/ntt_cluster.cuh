// ntt_cluster.cuh — negacyclic NTT / INTT of rings LARGER than one CTA's shared-memory block in ONE pass over HBM:
// a thread-block cluster of C = 2^LC CTAs owns one (poly, limb); every CTA keeps a 2^LOGB-word block in its own
// shared memory (64 KB: three CTAs per SM, so one CTA's global phase still overlaps its neighbours' butterflies)
// and the top LC stages are exchanged through DISTRIBUTED SHARED MEMORY instead of a round trip through HBM.
//
// Replaces the in-place radix-2^k global passes of ntt.cuh (ntt_fwd_global_pass / ntt_inv_global_pass: one extra
// read + write of the whole batch per pass) that the round-1 schedule ran before / after the block kernel, i.e. the
// reference's one-global-pass-per-stage loop (cuda/src/matrix/MatrixNTT.cu:685-704).
//
//   forward   rank r loads, for its quarter (1/C) of the in-block offsets e, the C legs e + j*B straight from global
//             memory (coalesced 128-bit), runs the radix-C Cooley-Tukey net in registers and PUSHES leg j into rank
//             j's shared memory (st.shared::cluster, 128-bit);  barrier.cluster;  every rank then owns sub-transform
//             `rank` of size B and finishes it with the block schedule of ntt_fast.cuh (pass 1 reads shared memory).
//   inverse   the mirror image: block schedule with the last local pass written back to shared memory;
//             barrier.cluster;  rank r PULLS leg j of its offsets from rank j (ld.shared::cluster, 128-bit), runs the
//             radix-C Gentleman-Sande net (n^-1 merged) and stores canonical residues to global memory;
//             barrier.cluster before exit (a CTA's shared memory must outlive its peers' reads).
//
// The same templates compile as host code (one-thread CTAs run phase by phase, emu.cu) for the CPU test-suite.
#pragma once
#include "ntt_fast.cuh"

namespace gp {

// ---- 16-byte vectors of the storage word -----------------------------------------------------------------
template <typename T> struct ClVec { static constexpr int kVec = sizeof(T) == 4 ? 4 : 2; };
GP_HD void gld_vec(const uint32_t *p, uint32_t (&x)[4]) { gld4(p, x); }
GP_HD void gld_vec(const uint64_t *p, uint64_t (&x)[2]) {
    const ulonglong2 a = *reinterpret_cast<const ulonglong2 *>(p);
    x[0] = a.x, x[1] = a.y;
}
GP_HD void gst_vec(uint32_t *p, const uint32_t (&x)[4]) { gst4(p, x); }
GP_HD void gst_vec(uint64_t *p, const uint64_t (&x)[2]) { *reinterpret_cast<ulonglong2 *>(p) = make_ulonglong2(x[0], x[1]); }

// vector forms of the source functors (ntt.cuh PlainSrc / DigitSrc, ntt_fast.cuh F64DigitSrc)
template <typename T, int V> GP_HD void src_ldv(const PlainSrc<T> &s, uint32_t i, T (&x)[V]) { gld_vec(s.p + i, x); }
template <int V> GP_HD void src_ldv(const PlainSrc<F64> &s, uint32_t i, F64 (&x)[V]) {
    uint64_t u[V];
    gld_vec(reinterpret_cast<const uint64_t *>(s.p) + i, u);
#pragma unroll
    for (int k = 0; k < V; k++) {
        uint64_t r = u[k];
        if (r >= s.q4) r -= s.q4;
        if (r >= (s.q4 >> 1)) r -= s.q4 >> 1;
        x[k].v = f64_from_word(r);
    }
}
template <typename T, int V> GP_HD void src_ldv(const DigitSrc<T> &s, uint32_t i, T (&x)[V]) {
    T raw[V];
    gld_vec(s.p + i, raw);
#pragma unroll
    for (int k = 0; k < V; k++) x[k] = s.xform(raw[k]);
}
template <int V> GP_HD void src_ldv(const F64DigitSrc &s, uint32_t i, F64 (&x)[V]) {
    uint64_t u[V];
    gld_vec(s.p + i, u);
#pragma unroll
    for (int k = 0; k < V; k++) x[k].v = f64_from_word((u[k] >> s.shift) & s.mask);
}

// ---- cluster primitives -------------------------------------------------------------------------------------
// (host bodies are empty: the host pass of nvcc parses the kernels, and emu.cu runs the phases with ClusterPeersHost)
GP_HD uint32_t cl_rank() {
    uint32_t r = 0;
#ifdef __CUDA_ARCH__
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
#endif
    return r;
}
GP_HD uint32_t cl_map(const void *smem_ptr, uint32_t rank) {
    uint32_t r = 0;
#ifdef __CUDA_ARCH__
    const uint32_t a = (uint32_t)__cvta_generic_to_shared(smem_ptr);
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(rank));
#endif
    return r;
}
GP_HD void cl_arrive() {
#ifdef __CUDA_ARCH__
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
#endif
}
GP_HD void cl_wait() {
#ifdef __CUDA_ARCH__
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
#endif
}
GP_HD void cl_st4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
#ifdef __CUDA_ARCH__
    asm volatile("st.shared::cluster.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
#endif
}
GP_HD void cl_st2(uint32_t addr, uint32_t a, uint32_t b) {
#ifdef __CUDA_ARCH__
    asm volatile("st.shared::cluster.v2.u32 [%0], {%1, %2};" ::"r"(addr), "r"(a), "r"(b) : "memory");
#endif
}
GP_HD void cl_ld4(uint32_t addr, uint32_t &a, uint32_t &b, uint32_t &c, uint32_t &d) {
#ifdef __CUDA_ARCH__
    asm volatile("ld.shared::cluster.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "r"(addr) : "memory");
#else
    a = b = c = d = 0;
#endif
}
GP_HD void cl_ld2(uint32_t addr, uint32_t &a, uint32_t &b) {
#ifdef __CUDA_ARCH__
    asm volatile("ld.shared::cluster.v2.u32 {%0, %1}, [%2];" : "=r"(a), "=r"(b) : "r"(addr) : "memory");
#else
    a = b = 0;
#endif
}

// The block buffers of all ranks of this cluster (same padded layout as SmemView<T, LOGB>): vector access to rank j's
// buffer at padded word index i (a multiple of the vector length; a vector never straddles a pad).
template <typename T, int LOGB, int LC> struct ClusterPeers {
    static constexpr int C = 1 << LC, V = ClVec<T>::kVec;
    static constexpr uint32_t kPlaneBytes = FastGeom<LOGB>::kPlane * 4u;
    uint32_t addr[C];  // shared::cluster byte address of every rank's buffer
    GP_HD explicit ClusterPeers(void *smem) {
#pragma unroll
        for (int j = 0; j < C; j++) addr[j] = cl_map(smem, (uint32_t)j);
    }
    GP_HD void stv(int j, uint32_t i, const uint32_t (&x)[4]) const { cl_st4(addr[j] + 4u * i, x[0], x[1], x[2], x[3]); }
    GP_HD void stv(int j, uint32_t i, const uint64_t (&x)[2]) const {
        cl_st2(addr[j] + 4u * i, (uint32_t)x[0], (uint32_t)x[1]);
        cl_st2(addr[j] + kPlaneBytes + 4u * i, (uint32_t)(x[0] >> 32), (uint32_t)(x[1] >> 32));
    }
    GP_HD void stv(int j, uint32_t i, const F64 (&x)[2]) const {
        const uint64_t u[2] = {f64_bits(x[0].v), f64_bits(x[1].v)};
        stv(j, i, u);
    }
    GP_HD void ldv(int j, uint32_t i, uint32_t (&x)[4]) const { cl_ld4(addr[j] + 4u * i, x[0], x[1], x[2], x[3]); }
    GP_HD void ldv(int j, uint32_t i, uint64_t (&x)[2]) const {
        uint32_t l0, l1, h0, h1;
        cl_ld2(addr[j] + 4u * i, l0, l1);
        cl_ld2(addr[j] + kPlaneBytes + 4u * i, h0, h1);
        x[0] = ((uint64_t)h0 << 32) | l0, x[1] = ((uint64_t)h1 << 32) | l1;
    }
    GP_HD void ldv(int j, uint32_t i, F64 (&x)[2]) const {
        uint64_t u[2];
        ldv(j, i, u);
        x[0].v = f64_from_bits(u[0]), x[1].v = f64_from_bits(u[1]);
    }
};
// host emulation (emu.cu): plain pointers to the ranks' buffers
template <typename T, int LOGB, int LC> struct ClusterPeersHost {
    void *base[1 << LC];
    template <int VV> GP_HD void stv(int j, uint32_t i, const T (&x)[VV]) const {
        const SmemView<T, LOGB> sm(base[j]);
        for (int k = 0; k < VV; k++) sm.st(i + k, x[k]);
    }
    template <int VV> GP_HD void ldv(int j, uint32_t i, T (&x)[VV]) const {
        const SmemView<T, LOGB> sm(base[j]);
        for (int k = 0; k < VV; k++) x[k] = sm.ld(i + k);
    }
};

// ---- shared-memory passes of arbitrary radix (pass 1 of a forward block / last pass of an inverse block when the
// block's data already sits in shared memory) ---------------------------------------------------------------------
template <typename T, int LOGB, int LSPAN, int LOGR, bool NOCSUB>
GP_HD void f_mid_r(const SmemView<T, LOGB> &sm, const TwPair<T> *__restrict__ tw, uint32_t nblk, uint32_t sb, T q, T q2) {
    constexpr int LST = LSPAN - LOGR, R = 1 << LOGR;
    constexpr uint32_t stride = 1u << LST, groups = 1u << (LOGB - LOGR);
    constexpr uint32_t per = LOGB - LSPAN;
    static_assert(LST >= 6 || LSPAN <= 6, "radix legs must not straddle a pad");
#pragma unroll 2
    for (uint32_t g = GP_TID; g < groups; g += GP_NT) {
        const uint32_t b = g >> LST, e = g & (stride - 1);
        const uint32_t base = padw((b << LSPAN) + e);
        T x[R];
#pragma unroll
        for (int j = 0; j < R; j++) x[j] = sm.ld(base + joff<LST>(j));
        fwd_net<T, LOGR, NOCSUB>(x, tw, ((nblk + sb) << per) + b, q, q2);
#pragma unroll
        for (int j = 0; j < R; j++) sm.st(base + joff<LST>(j), x[j]);
    }
}
template <typename T, int LOGB, int LT0, int LOGR>
GP_HD void i_mid_r(const SmemView<T, LOGB> &sm, const TwPair<T> *__restrict__ tw, uint32_t nblk, uint32_t sb, const LimbConst<T> &c) {
    constexpr int LSPAN = LT0 + LOGR, R = 1 << LOGR;
    constexpr uint32_t t0 = 1u << LT0, groups = 1u << (LOGB - LOGR);
    constexpr uint32_t per = LOGB - LSPAN;
    static_assert(LT0 >= 6 || LSPAN <= 6, "radix legs must not straddle a pad");
#pragma unroll 2
    for (uint32_t g = GP_TID; g < groups; g += GP_NT) {
        const uint32_t b = g >> LT0, e = g & (t0 - 1);
        const uint32_t base = padw((b << LSPAN) + e);
        T x[R];
#pragma unroll
        for (int j = 0; j < R; j++) x[j] = sm.ld(base + joff<LT0>(j));
        inv_net<T, LOGR, false>(x, tw, ((nblk + sb) << per) + b, c);
#pragma unroll
        for (int j = 0; j < R; j++) sm.st(base + joff<LT0>(j), x[j]);
    }
}

// ---- forward ------------------------------------------------------------------------------------------------------
// Phase A (push): `src` addresses the whole 2^(LOGB+LC)-word transform.  WAIT_FIRST: the cluster barrier that says
// "every peer CTA is running" is waited for after the first batch of global loads has been issued.
template <typename T, int LOGB, int LC, bool NOCSUB, class Src, class Peers>
GP_HD void cl_fwd_push(const Src &src, const Peers &peers, uint32_t rank, const TwPair<T> *__restrict__ tw,
                       T q, T q2) {
    constexpr int C = 1 << LC, V = ClVec<T>::kVec;
    constexpr uint32_t B = 1u << LOGB, per = B >> LC, vecs = per / V;
    // every thread runs the same number of iterations (the aligned cluster barrier below sits inside the loop)
    constexpr uint32_t NT = FastCfg<T, LOGB>::kThreads;
    static_assert(vecs % NT == 0, "push plan: whole iterations for every thread");
    // U iterations are batched: all their global loads (U*C 128-bit loads per thread) are issued before the first
    // butterfly.  ncu put 20 % of the u32 2^15 kernel's stall samples on the first butterfly after a 2-load iteration
    // (profiles/r02_ncu_cluster_summary.txt), but batching 2 / 4 / 8 / 16 loads measured the same within 1 % (16: slower,
    // registers) — the other two resident CTAs already fill those slots (profiles/r02_push_sweep.txt)
#ifndef GP_PUSH_LOADS
#define GP_PUSH_LOADS 4
#endif
    constexpr uint32_t iters = vecs / NT;
    constexpr uint32_t U = (GP_PUSH_LOADS / C) < 1 ? 1u : ((GP_PUSH_LOADS / C) > iters ? iters : (uint32_t)(GP_PUSH_LOADS / C));
    static_assert(iters % U == 0, "push plan: whole batches");
    bool first = true;
#pragma unroll 1
    for (uint32_t v0 = GP_TID; v0 < vecs; v0 += U * GP_NT) {
        T x[U][V][C];
#pragma unroll
        for (uint32_t u = 0; u < U; u++) {
            const uint32_t e = rank * per + V * (v0 + u * GP_NT);
#pragma unroll
            for (int j = 0; j < C; j++) {
                T t[V];
                src_ldv(src, e + (uint32_t)j * B, t);
#pragma unroll
                for (int k = 0; k < V; k++) x[u][k][j] = t[k];
            }
        }
#pragma unroll
        for (uint32_t u = 0; u < U; u++) {
#pragma unroll
            for (int k = 0; k < V; k++) fwd_net<T, LC, NOCSUB>(x[u][k], tw, 1u, q, q2);
        }
        if (first) {
            cl_wait();
            first = false;
        }
#pragma unroll
        for (uint32_t u = 0; u < U; u++) {
            const uint32_t at = padw(rank * per + V * (v0 + u * GP_NT));
#pragma unroll
            for (int j = 0; j < C; j++) {
                T t[V];
#pragma unroll
                for (int k = 0; k < V; k++) t[k] = x[u][k][j];
                peers.stv(j, at, t);
            }
        }
    }
}
// Phase B (local): sub-transform `rank` of size 2^LOGB, data already in shared memory in the lazy range of the nets
template <typename T, int LOGB, int LC, bool NOCSUB>
GP_HD void cl_fwd_local(void *smem, T *__restrict__ dst, uint32_t rank, const TwPair<T> *__restrict__ tw, const LimbConst<T> &c) {
    const SmemView<T, LOGB> sm(smem);
    constexpr uint32_t nblk = 1u << LC;
    f_mid_r<T, LOGB, LOGB, FastGeom<LOGB>::kR1, NOCSUB>(sm, tw, nblk, rank, c.q, c.q2);
    GP_SYNC();
    fast_fwd_tail<T, LOGB, NOCSUB>(sm, dst, nblk, rank, tw, c);
}

// ---- inverse ------------------------------------------------------------------------------------------------------
// Phase A (local): the low LOGB stages of sub-block `rank`, result left in shared memory in [0, 2q)
template <typename T, int LOGB, int LC>
GP_HD void cl_inv_local(const T *__restrict__ src, void *smem, uint32_t rank, const TwPair<T> *__restrict__ tw, const LimbConst<T> &c) {
    const SmemView<T, LOGB> sm(smem);
    constexpr uint32_t nblk = 1u << LC;
    constexpr int R1 = FastGeom<LOGB>::kR1;
    fast_inv_head<T, LOGB>(src, sm, nblk, rank, tw, c);
    i_mid_r<T, LOGB, LOGB - R1, R1>(sm, tw, nblk, rank, c);
}
// Phase B (pull): the top LC stages with n^-1 merged; canonical residues to global memory
template <typename T, int LOGB, int LC, class Peers>
GP_HD void cl_inv_pull(const Peers &peers, T *__restrict__ dst, uint32_t rank, const TwPair<T> *__restrict__ tw,
                       const LimbConst<T> &c) {
    constexpr int C = 1 << LC, V = ClVec<T>::kVec;
    constexpr uint32_t B = 1u << LOGB, per = B >> LC, vecs = per / V;
#pragma unroll 2
    for (uint32_t v = GP_TID; v < vecs; v += GP_NT) {
        const uint32_t e = rank * per + V * v, at = padw(e);
        T x[V][C];
#pragma unroll
        for (int j = 0; j < C; j++) {
            T t[V];
            peers.ldv(j, at, t);
#pragma unroll
            for (int k = 0; k < V; k++) x[k][j] = t[k];
        }
#pragma unroll
        for (int k = 0; k < V; k++) inv_net<T, LC, true>(x[k], tw, 1u, c);
#pragma unroll
        for (int j = 0; j < C; j++) {
            if constexpr (is_f64<T>::value) {
                uint64_t u[V];
#pragma unroll
                for (int k = 0; k < V; k++) u[k] = f64_canonical(x[k][j].v, c.q.v, c.q2.v);
                gst_vec(reinterpret_cast<uint64_t *>(dst) + e + (size_t)j * B, u);
            } else {
                T t[V];
#pragma unroll
                for (int k = 0; k < V; k++) t[k] = csub_min<T>(x[k][j], c.q);
                gst_vec(dst + e + (size_t)j * B, t);
            }
        }
    }
}

// ---- digit source of one output (poly, limb) of a decomposition: geometry as fast_decompose_cta (ntt_fast.cuh) -------
template <typename T> struct DigitSrcOf { typedef DigitSrc<T> type; };
template <> struct DigitSrcOf<F64> { typedef F64DigitSrc type; };
template <typename T>
GP_HD typename DigitSrcOf<T>::type make_digit_src(const T *__restrict__ src, const DecompGeom &g, size_t pl, uint32_t logn,
                                                  const LimbConst<T> *__restrict__ lc) {
    const DigitSpec ds = digit_spec(g, pl, logn, lc);
    if constexpr (is_f64<T>::value) {
        return make_f64_digit_src(src, ds);
    } else {
        return make_int_digit_src<T>(src, ds, lc[ds.limb].q, lc[ds.limb].one_s);
    }
}

// launch shape: cluster_block_log / cluster_log live in ntt_fast.cuh (host planning code uses them without this header)

#if defined(__CUDACC__) && !defined(GP_EMU_BUILD)  // the kernels proper (emu.cu drives the phases above on the host)
template <typename T, int LOGB, int LC, bool NOCSUB, class Src>
__device__ __forceinline__ void cl_fwd_cta(void *smem, const Src &src, T *dst_poly, const TwPair<T> *__restrict__ tw, const LimbConst<T> &c) {
    const uint32_t rank = cl_rank();
    const ClusterPeers<T, LOGB, LC> peers(smem);
    cl_arrive();  // "this CTA is running": peers wait for it before their first push
    cl_fwd_push<T, LOGB, LC, NOCSUB>(src, peers, rank, tw, c.q, c.q2);
    cl_arrive();  // release: my pushes are visible to a peer that has waited
    cl_wait();
    cl_fwd_local<T, LOGB, LC, NOCSUB>(smem, dst_poly + ((size_t)rank << LOGB), rank, tw, c);
}

template <typename T, int LOGB, int LC, bool NOCSUB>
__global__ void __cluster_dims__(1 << LC, 1, 1) __launch_bounds__(FastCfg<T, LOGB>::kThreads, FastCfg<T, LOGB>::kMinBlocks)
fast_ntt_fwd_cluster_kernel(const T *src, T *dst, uint32_t limbs, const TwPair<T> *__restrict__ tw_all,
                            const LimbConst<T> *__restrict__ lc) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int logn = LOGB + LC;
    const size_t pl = blockIdx.x >> LC;
    const uint32_t limb = (uint32_t)(pl % limbs);
    const LimbConst<T> c = lc[limb];
    const TwPair<T> *tw = tw_all + ((size_t)limb << logn);
    if constexpr (is_f64<T>::value) {
        PlainSrc<T> s{src + (pl << logn), 4 * (uint64_t)c.q.v};
        cl_fwd_cta<T, LOGB, LC, NOCSUB>(smem_raw, s, dst + (pl << logn), tw, c);
    } else {
        PlainSrc<T> s{src + (pl << logn)};
        cl_fwd_cta<T, LOGB, LC, NOCSUB>(smem_raw, s, dst + (pl << logn), tw, c);
    }
}

template <typename T, int LOGB, int LC>
__global__ void __cluster_dims__(1 << LC, 1, 1) __launch_bounds__(FastCfg<T, LOGB>::kThreads, FastCfg<T, LOGB>::kMinBlocks)
fast_ntt_inv_cluster_kernel(const T *src, T *dst, uint32_t limbs, const TwPair<T> *__restrict__ tw_all,
                            const LimbConst<T> *__restrict__ lc) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int logn = LOGB + LC;
    const uint32_t rank = cl_rank();
    const size_t pl = blockIdx.x >> LC;
    const uint32_t limb = (uint32_t)(pl % limbs);
    const LimbConst<T> c = lc[limb];
    const TwPair<T> *tw = tw_all + ((size_t)limb << logn);
    const ClusterPeers<T, LOGB, LC> peers(smem_raw);
    cl_inv_local<T, LOGB, LC>(src + (pl << logn) + ((size_t)rank << LOGB), smem_raw, rank, tw, c);
    cl_arrive();
    cl_wait();
    cl_inv_pull<T, LOGB, LC>(peers, dst + (pl << logn), rank, tw, c);
    cl_arrive();  // my reads of the peers' buffers are done
    cl_wait();    // ... and theirs of mine
}

// fused gadget decomposition + forward NTT (digit extraction is the load functor of the push phase)
template <typename T, int LOGB, int LC, bool NOCSUB>
__global__ void __cluster_dims__(1 << LC, 1, 1) __launch_bounds__(FastCfg<T, LOGB>::kThreads, FastCfg<T, LOGB>::kMinBlocks)
fast_decompose_cluster_kernel(const T *__restrict__ src, T *__restrict__ out, DecompGeom g, const TwPair<T> *__restrict__ tw_all,
                              const LimbConst<T> *__restrict__ lc) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int logn = LOGB + LC;
    const size_t pl = blockIdx.x >> LC;  // out (poly, limb)
    const uint32_t limb = (uint32_t)(pl % g.limbs);
    const auto dsrc = make_digit_src<T>(src, g, pl, logn, lc);
    cl_fwd_cta<T, LOGB, LC, NOCSUB>(smem_raw, dsrc, out + (pl << logn), tw_all + ((size_t)limb << logn), lc[limb]);
}
#endif  // __CUDACC__ && !GP_EMU_BUILD

}  // namespace gp
