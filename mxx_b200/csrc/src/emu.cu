// emu.cu — HOST emulation harness for the block-level kernel templates (tests only).
//
// The kernels in include/{ntt,arith}.cuh are written against GP_TID / GP_NT / GP_SYNC;
// compiled for the host they run as one-thread blocks.  This file drives them over all
// block indices with the same pass schedule the launcher uses (MatrixNTT.inl,
// MatrixDecompose.inl, MatrixArith.inl), so the CPU test-suite can check index math,
// twiddle addressing and lazy-reduction bounds against the oracle without a GPU.
//
// It is NOT part of libgpupoly.so and is not a CPU fallback: the product library has no
// host compute path and fails loudly without a CUDA device.
//
// Data crosses this interface as u64 words in DEVICE order ([poly][limb][coeff], EVAL
// bit-reversed); `wide` picks the storage width exactly like gpu_context_create does.
#define GP_EMU_BUILD 1
#include <stdint.h>
#include <string.h>

#include <vector>

#include "arith.cuh"
#include "host_tables.hpp"
#include "ntt.cuh"
#include "ntt_fast.cuh"
#include "ntt_cluster.cuh"
#include "sampling.cuh"

namespace {
using namespace gp;

size_t g_f64_blocks = 0;  // blocks run through F64 kernels (tests assert the path was really taken)
int g_fuse_top = 1;  // emu_set_fuse_top: digits extracted inside the first global pass (the launcher's default)
int g_cluster = 1;  // emu_set_cluster: rings of 2..8 blocks through the cluster / DSMEM phases (the launcher's default)
int g_f64 = 0;  // emu_set_f64: drive the FP64-pipe kernels (type F64) where the product library would

template <typename T> struct Emu {
    int logn;
    uint32_t limbs;
    host::Tables<T> tab;
    host::Tables<F64> ftab;
    bool f64 = false;
    Emu(int logn_, uint32_t limbs_, const uint64_t *moduli) : logn(logn_), limbs(limbs_) {
        std::vector<uint64_t> m(moduli, moduli + limbs_), psi;
        for (uint64_t q : m) psi.push_back(host::primitive_2n_root(q, (uint64_t)1 << logn_));
        tab = host::build_tables<T>(logn_, m, psi);
        uint64_t qmax = 0;
        for (uint64_t q : m) qmax = q > qmax ? q : qmax;
        nocsub = host::nocsub_ok(qmax, sizeof(T) == 8);
        if constexpr (sizeof(T) == 8) {
            if (g_f64 && host::f64_ok(qmax, true)) {
                ftab = host::f64_tables(tab);
                f64 = true;
            }
        }
    }
    // ---- compile-time-planned kernels (ntt_fast.cuh), same selection rule as MatrixNTT.inl ----
    bool fast_ok(int logb_cap) const { return logb_cap == 0 && fast_supported<T>(fast_block_log<T>(logn)); }
    template <int LOGB> void fast_fwd_blocks(std::vector<T> &d, size_t nbid, bool nocsub) {
        std::vector<uint32_t> sm(fast_smem_bytes<T, LOGB>() / 4 + 4);
        if constexpr (sizeof(T) == 8) {
            if (f64) {
                g_f64_blocks += nbid;
                for (size_t bid = 0; bid < nbid; bid++)
                    fast_fwd_cta<F64, LOGB, false>((uint32_t)bid, sm.data(), (const F64 *)d.data(), (F64 *)d.data(), limbs, logn, ftab.fwd.data(), ftab.limb.data());
                return;
            }
        }
        for (size_t bid = 0; bid < nbid; bid++) {
            if (nocsub) fast_fwd_cta<T, LOGB, true>((uint32_t)bid, sm.data(), d.data(), d.data(), limbs, logn, tab.fwd.data(), tab.limb.data());
            else fast_fwd_cta<T, LOGB, false>((uint32_t)bid, sm.data(), d.data(), d.data(), limbs, logn, tab.fwd.data(), tab.limb.data());
        }
    }
    template <int LOGB> void fast_inv_blocks(std::vector<T> &d, size_t nbid) {
        std::vector<uint32_t> sm(fast_smem_bytes<T, LOGB>() / 4 + 4);
        if constexpr (sizeof(T) == 8) {
            if (f64) {
                g_f64_blocks += nbid;
                for (size_t bid = 0; bid < nbid; bid++)
                    fast_inv_cta<F64, LOGB>((uint32_t)bid, sm.data(), (const F64 *)d.data(), (F64 *)d.data(), limbs, logn, ftab.inv.data(), ftab.limb.data());
                return;
            }
        }
        for (size_t bid = 0; bid < nbid; bid++)
            fast_inv_cta<T, LOGB>((uint32_t)bid, sm.data(), d.data(), d.data(), limbs, logn, tab.inv.data(), tab.limb.data());
    }
    void fast_fwd_dispatch(std::vector<T> &d, size_t nbid, int logb, bool nocsub) {
        switch (logb) {
            case 12: fast_fwd_blocks<12>(d, nbid, nocsub); break;
            case 13: fast_fwd_blocks<13>(d, nbid, nocsub); break;
            case 14: fast_fwd_blocks<14>(d, nbid, nocsub); break;
            default:
                if constexpr (sizeof(T) == 4) fast_fwd_blocks<15>(d, nbid, nocsub);
        }
    }
    void fast_inv_dispatch(std::vector<T> &d, size_t nbid, int logb) {
        switch (logb) {
            case 12: fast_inv_blocks<12>(d, nbid); break;
            case 13: fast_inv_blocks<13>(d, nbid); break;
            case 14: fast_inv_blocks<14>(d, nbid); break;
            default:
                if constexpr (sizeof(T) == 4) fast_inv_blocks<15>(d, nbid);
        }
    }
    // ---- cluster / DSMEM kernels (ntt_cluster.cuh), phase by phase: every rank's push (pull) runs before any rank's
    // local (store) phase, which is what the cluster barrier guarantees on the device ----
    int cluster_lc(int logb_cap) const { return (g_cluster && logb_cap == 0) ? cluster_log<T>(logn) : 0; }
    template <typename K, int LB, int LC, bool NOCSUB, class MakeSrc>
    void cluster_fwd_run(K *dst, size_t pl_count, const TwPair<K> *tw_all, const LimbConst<K> *lc, MakeSrc make_src) {
        constexpr int C = 1 << LC;
        std::vector<std::vector<uint32_t>> bufs(C, std::vector<uint32_t>(fast_smem_bytes<K, LB>() / 4 + 4));
        ClusterPeersHost<K, LB, LC> peers;
        for (int j = 0; j < C; j++) peers.base[j] = bufs[j].data();
        for (size_t pl = 0; pl < pl_count; pl++) {
            const uint32_t limb = (uint32_t)(pl % limbs);
            const LimbConst<K> c = lc[limb];
            const TwPair<K> *tw = tw_all + ((size_t)limb << logn);
            const auto src = make_src(pl, c);
            for (uint32_t r = 0; r < (uint32_t)C; r++) cl_fwd_push<K, LB, LC, NOCSUB>(src, peers, r, tw, c.q, c.q2);
            for (uint32_t r = 0; r < (uint32_t)C; r++)
                cl_fwd_local<K, LB, LC, NOCSUB>(bufs[r].data(), dst + (pl << logn) + ((size_t)r << LB), r, tw, c);
        }
    }
    template <typename K, int LB, int LC>
    void cluster_inv_run(const K *src, K *dst, size_t pl_count, const TwPair<K> *tw_all, const LimbConst<K> *lc) {
        constexpr int C = 1 << LC;
        std::vector<std::vector<uint32_t>> bufs(C, std::vector<uint32_t>(fast_smem_bytes<K, LB>() / 4 + 4));
        ClusterPeersHost<K, LB, LC> peers;
        for (int j = 0; j < C; j++) peers.base[j] = bufs[j].data();
        for (size_t pl = 0; pl < pl_count; pl++) {
            const uint32_t limb = (uint32_t)(pl % limbs);
            const LimbConst<K> c = lc[limb];
            const TwPair<K> *tw = tw_all + ((size_t)limb << logn);
            for (uint32_t r = 0; r < (uint32_t)C; r++)
                cl_inv_local<K, LB, LC>(src + (pl << logn) + ((size_t)r << LB), bufs[r].data(), r, tw, c);
            for (uint32_t r = 0; r < (uint32_t)C; r++) cl_inv_pull<K, LB, LC>(peers, dst + (pl << logn), r, tw, c);
        }
    }
    // MakeSrc for the working type K (plain words of `d`, or digits of `s` under geometry g)
    template <typename K, int LC, class MakeSrc> void cluster_fwd_any(K *dst, size_t pl_count, MakeSrc make_src) {
        constexpr int LB = sizeof(T) == 4 ? 14 : 13;
        if constexpr (is_f64<K>::value) {
            g_f64_blocks += pl_count << LC;
            cluster_fwd_run<K, LB, LC, false>(dst, pl_count, ftab.fwd.data(), ftab.limb.data(), make_src);
        } else {
            if (nocsub) cluster_fwd_run<K, LB, LC, true>(dst, pl_count, tab.fwd.data(), tab.limb.data(), make_src);
            else cluster_fwd_run<K, LB, LC, false>(dst, pl_count, tab.fwd.data(), tab.limb.data(), make_src);
        }
    }
    template <typename K, class MakeSrc> void cluster_fwd_lc(int lc, K *dst, size_t pl_count, MakeSrc make_src) {
        switch (lc) {
            case 1: cluster_fwd_any<K, 1>(dst, pl_count, make_src); break;
            case 2: cluster_fwd_any<K, 2>(dst, pl_count, make_src); break;
            default: cluster_fwd_any<K, 3>(dst, pl_count, make_src); break;
        }
    }
    void cluster_fwd(std::vector<T> &d, size_t polys, int lc) {
        const size_t pl = polys * limbs;
        const int ln = logn;
        if constexpr (sizeof(T) == 8) {
            if (f64) {
                const F64 *base = (const F64 *)d.data();
                cluster_fwd_lc<F64>(lc, (F64 *)d.data(), pl, [base, ln](size_t p, const LimbConst<F64> &c) { return PlainSrc<F64>{base + (p << ln), 4 * (uint64_t)c.q.v}; });
                return;
            }
        }
        const T *base = d.data();
        cluster_fwd_lc<T>(lc, d.data(), pl, [base, ln](size_t p, const LimbConst<T> &) { return PlainSrc<T>{base + (p << ln)}; });
    }
    void cluster_decompose(const std::vector<T> &s, std::vector<T> &o, const DecompGeom &g, size_t opl, int lc) {
        const int ln = logn;
        if constexpr (sizeof(T) == 8) {
            if (f64) {
                const F64 *sp = (const F64 *)s.data();
                const LimbConst<F64> *lcf = ftab.limb.data();
                cluster_fwd_lc<F64>(lc, (F64 *)o.data(), opl, [sp, g, ln, lcf](size_t p, const LimbConst<F64> &) { return make_digit_src<F64>(sp, g, p, (uint32_t)ln, lcf); });
                return;
            }
        }
        const T *sp = s.data();
        const LimbConst<T> *lct = tab.limb.data();
        cluster_fwd_lc<T>(lc, o.data(), opl, [sp, g, ln, lct](size_t p, const LimbConst<T> &) { return make_digit_src<T>(sp, g, p, (uint32_t)ln, lct); });
    }
    void cluster_inv(std::vector<T> &d, size_t polys, int lc) {
        const size_t pl = polys * limbs;
        constexpr int LB = sizeof(T) == 4 ? 14 : 13;
#define EMU_CINV(K, LC, TW, LCT) cluster_inv_run<K, LB, LC>((const K *)d.data(), (K *)d.data(), pl, TW, LCT)
        if constexpr (sizeof(T) == 8) {
            if (f64) {
                g_f64_blocks += pl << lc;
                switch (lc) {
                    case 1: EMU_CINV(F64, 1, ftab.inv.data(), ftab.limb.data()); break;
                    case 2: EMU_CINV(F64, 2, ftab.inv.data(), ftab.limb.data()); break;
                    default: EMU_CINV(F64, 3, ftab.inv.data(), ftab.limb.data()); break;
                }
                return;
            }
        }
        switch (lc) {
            case 1: EMU_CINV(T, 1, tab.inv.data(), tab.limb.data()); break;
            case 2: EMU_CINV(T, 2, tab.inv.data(), tab.limb.data()); break;
            default: EMU_CINV(T, 3, tab.inv.data(), tab.limb.data()); break;
        }
#undef EMU_CINV
    }
    bool nocsub = false;
    int block_log(int logb_cap) const {
        if (fast_ok(logb_cap)) return fast_block_log<T>(logn);
        int lb = ntt_block_log<T>(logn);
        if (logb_cap > 0 && logb_cap < lb) lb = logb_cap;
        return lb;
    }
    void fwd(std::vector<T> &d, size_t polys, int logb_cap, bool first_top_pass_done = false) {
        if (const int lc = cluster_lc(logb_cap); lc > 0 && !first_top_pass_done) return cluster_fwd(d, polys, lc);
        const int logb = block_log(logb_cap), top = logn - logb;
        const size_t pl = polys * limbs;
        std::vector<T> sm(ntt_smem_bytes<T>(logb) / sizeof(T) + 1);
        if (top > 0) {
            int passes[8], np = 0;
            split_top(top, passes, &np);
            uint32_t lspan = (uint32_t)logn;
            for (int i = 0; i < np; i++) {
                const size_t total = (i == 0 && first_top_pass_done) ? 0 : pl << (logn - passes[i]);
                for (size_t gid = 0; gid < total; gid++) switch (passes[i]) {
                        case 1: fwd_global_thread<T, 1>(gid, d.data(), limbs, logn, lspan, tab.fwd.data(), tab.limb.data()); break;
                        case 2: fwd_global_thread<T, 2>(gid, d.data(), limbs, logn, lspan, tab.fwd.data(), tab.limb.data()); break;
                        case 3: fwd_global_thread<T, 3>(gid, d.data(), limbs, logn, lspan, tab.fwd.data(), tab.limb.data()); break;
                        default: fwd_global_thread<T, 4>(gid, d.data(), limbs, logn, lspan, tab.fwd.data(), tab.limb.data()); break;
                    }
                lspan -= passes[i];
            }
        }
        if (fast_ok(logb_cap)) {
            fast_fwd_dispatch(d, pl << top, logb, nocsub);
            return;
        }
        for (size_t bid = 0; bid < (pl << top); bid++)
            ntt_fwd_cta<T>((uint32_t)bid, sm.data(), d.data(), d.data(), limbs, logn, logb, tab.fwd.data(), tab.limb.data());
    }
    void inv(std::vector<T> &d, size_t polys, int logb_cap) {
        if (const int lc = cluster_lc(logb_cap); lc > 0) return cluster_inv(d, polys, lc);
        const int logb = block_log(logb_cap), top = logn - logb;
        const size_t pl = polys * limbs;
        std::vector<T> sm(ntt_smem_bytes<T>(logb) / sizeof(T) + 1);
        if (fast_ok(logb_cap)) {
            fast_inv_dispatch(d, pl << top, logb);
        } else {
            for (size_t bid = 0; bid < (pl << top); bid++)
                ntt_inv_cta<T>((uint32_t)bid, sm.data(), d.data(), d.data(), limbs, logn, logb, tab.inv.data(), tab.limb.data());
        }
        if (top > 0) {
            int passes[8], np = 0;
            split_top(top, passes, &np);
            uint32_t lt0 = (uint32_t)logb;
            for (int i = np - 1; i >= 0; i--) {
                const size_t total = pl << (logn - passes[i]);
                const bool fin = i == 0;
                for (size_t gid = 0; gid < total; gid++) {
#define INV_CASE(R)                                                                                              \
    case R:                                                                                                      \
        if (fin) inv_global_thread<T, R, true>(gid, d.data(), limbs, logn, lt0, tab.inv.data(), tab.limb.data()); \
        else inv_global_thread<T, R, false>(gid, d.data(), limbs, logn, lt0, tab.inv.data(), tab.limb.data());    \
        break;
                    switch (passes[i]) {
                        INV_CASE(1)
                        INV_CASE(2)
                        INV_CASE(3)
                        default:
                            if (fin) inv_global_thread<T, 4, true>(gid, d.data(), limbs, logn, lt0, tab.inv.data(), tab.limb.data());
                            else inv_global_thread<T, 4, false>(gid, d.data(), limbs, logn, lt0, tab.inv.data(), tab.limb.data());
                    }
#undef INV_CASE
                }
                lt0 += passes[i];
            }
        }
    }
};

template <typename T> std::vector<T> narrow(const uint64_t *p, size_t n) {
    std::vector<T> v(n);
    for (size_t i = 0; i < n; i++) v[i] = (T)p[i];
    return v;
}
template <typename T> void widen(const std::vector<T> &v, uint64_t *p) {
    for (size_t i = 0; i < v.size(); i++) p[i] = (uint64_t)v[i];
}

template <typename T>
int do_ntt(int logn, uint32_t limbs, const uint64_t *moduli, uint64_t *data, size_t polys, int inverse, int logb_cap) {
    Emu<T> e(logn, limbs, moduli);
    auto d = narrow<T>(data, (polys * limbs) << logn);
    if (inverse) e.inv(d, polys, logb_cap);
    else e.fwd(d, polys, logb_cap);
    widen(d, data);
    return 0;
}

template <typename T>
int do_decompose(int logn, uint32_t limbs, const uint64_t *moduli, const uint64_t *src, size_t rows, size_t src_cols,
                 size_t src_col0, size_t cols, uint32_t base_bits, int small, int to_eval, uint64_t *out, int logb_cap) {
    Emu<T> e(logn, limbs, moduli);
    uint64_t qmax = 0;
    for (uint32_t i = 0; i < limbs; i++) qmax = moduli[i] > qmax ? moduli[i] : qmax;
    const uint32_t crt_bits = 64 - (uint32_t)__builtin_clzll(qmax);
    DecompGeom g{};
    g.limbs = limbs;
    g.src_cols = (uint32_t)src_cols;
    g.src_col0 = (uint32_t)src_col0;
    g.cols = (uint32_t)cols;
    g.dpt = (crt_bits + base_bits - 1) / base_bits;
    g.ell = small ? g.dpt : g.dpt * limbs;
    g.base_bits = base_bits;
    const size_t opolys = rows * g.ell * cols, opl = opolys * limbs;
    auto s = narrow<T>(src, (rows * src_cols * limbs) << logn);
    std::vector<T> o(opl << logn);
    const int logb = e.block_log(logb_cap), top = logn - logb;
    std::vector<T> sm(ntt_smem_bytes<T>(logb) / sizeof(T) + 1);
    if (const int lc = e.cluster_lc(logb_cap); to_eval && lc > 0) {
        e.cluster_decompose(s, o, g, opl, lc);
    } else if (to_eval && top == 0 && e.fast_ok(logb_cap)) {
        std::vector<uint32_t> fsm((sizeof(T) / 4) * ((1u << logb) + (1u << (logb - 4))) + 8);
#define DEC_CASE(LB)                                                                                                          \
    case LB:                                                                                                                  \
        for (size_t bid = 0; bid < opl; bid++) {                                                                              \
            if constexpr (sizeof(T) == 8) {                                                                                   \
                if (e.f64) {                                                                                                  \
                    g_f64_blocks++;                                                                                           \
                    fast_decompose_cta<F64, LB, false>((uint32_t)bid, fsm.data(), (const F64 *)s.data(), (F64 *)o.data(), g, e.ftab.fwd.data(), e.ftab.limb.data()); \
                    continue;                                                                                                 \
                }                                                                                                             \
            }                                                                                                                 \
            if (e.nocsub) fast_decompose_cta<T, LB, true>((uint32_t)bid, fsm.data(), s.data(), o.data(), g, e.tab.fwd.data(), e.tab.limb.data()); \
            else fast_decompose_cta<T, LB, false>((uint32_t)bid, fsm.data(), s.data(), o.data(), g, e.tab.fwd.data(), e.tab.limb.data());         \
        }                                                                                                                     \
        break;
        switch (logb) {
            DEC_CASE(12)
            DEC_CASE(13)
            DEC_CASE(14)
            default:
                if constexpr (sizeof(T) == 4) {
                    for (size_t bid = 0; bid < opl; bid++) {
                        if (e.nocsub) fast_decompose_cta<T, 15, true>((uint32_t)bid, fsm.data(), s.data(), o.data(), g, e.tab.fwd.data(), e.tab.limb.data());
                        else fast_decompose_cta<T, 15, false>((uint32_t)bid, fsm.data(), s.data(), o.data(), g, e.tab.fwd.data(), e.tab.limb.data());
                    }
                }
        }
#undef DEC_CASE
    } else if (to_eval && top == 0) {
        for (size_t bid = 0; bid < opl; bid++)
            decompose_cta<T, true>((uint32_t)bid, sm.data(), s.data(), o.data(), g, logn, logb, e.tab.fwd.data(), e.tab.limb.data());
    } else if (to_eval && g_fuse_top) {
        // the launcher's fused digit extraction + first global pass (MatrixDecompose.inl run_decompose)
        int passes[8], np = 0;
        split_top(top, passes, &np);
#define EMU_DFG(R)                                                                                                    \
    case R: {                                                                                                         \
        const size_t nb = opl << ((logn - R) - decompose_top_chunk_log<R>((uint32_t)logn));                            \
        for (size_t bid = 0; bid < nb; bid++)                                                                         \
            decompose_fwd_global_cta<T, R>(bid, s.data(), o.data(), g, (uint32_t)logn, e.tab.fwd.data(), e.tab.limb.data()); \
    } break;
        switch (passes[0]) {
            EMU_DFG(1)
            EMU_DFG(2)
            EMU_DFG(3)
            default:
                EMU_DFG(4)
        }
#undef EMU_DFG
        e.fwd(o, opolys, logb_cap, true);
    } else {
        for (size_t bid = 0; bid < (opl << top); bid++)
            decompose_cta<T, false>((uint32_t)bid, sm.data(), s.data(), o.data(), g, logn, logb, e.tab.fwd.data(), e.tab.limb.data());
        if (to_eval) e.fwd(o, opolys, logb_cap);
    }
    widen(o, out);
    return 0;
}

template <typename T, int V, int TR, int TC>
void matmul_grid(std::vector<T> &out, const std::vector<T> &A, const std::vector<T> &B, const MatmulGeom &g,
                 const LimbConst<T> *lc) {
    const uint32_t threads = 4;  // a small block so that several CTAs per limb exist even for tiny rings
    const uint32_t per_limb = ((1u << g.logn) / V + threads - 1) / threads;
    for (uint32_t bz = 0; bz < (g.rows + TR - 1) / TR; bz++)
        for (uint32_t by = 0; by < (g.cols + TC - 1) / TC; by++)
            for (uint32_t bx = 0; bx < g.limbs * per_limb; bx++)
                for (uint32_t t = 0; t < threads; t++)
                    matmul_thread<T, V, TR, TC>(bx, by, bz, t, threads, out.data(), A.data(), B.data(), g, lc);
}

template <int V, int TR, int TC>
void matmul_f64_grid(std::vector<uint64_t> &out, const std::vector<uint64_t> &A, const std::vector<uint64_t> &B, const MatmulGeom &g,
                     const LimbConst<F64> *lc) {
    const uint32_t threads = 4;
    const uint32_t per_limb = ((1u << g.logn) / V + threads - 1) / threads;
    for (uint32_t bz = 0; bz < (g.rows + TR - 1) / TR; bz++)
        for (uint32_t by = 0; by < (g.cols + TC - 1) / TC; by++)
            for (uint32_t bx = 0; bx < g.limbs * per_limb; bx++)
                for (uint32_t t = 0; t < threads; t++)
                    matmul_f64_thread<V, TR, TC>(bx, by, bz, t, threads, out.data(), A.data(), B.data(), g, lc);
}

template <typename T>
int do_matmul(int logn, uint32_t limbs, const uint64_t *moduli, uint64_t *out, const uint64_t *A, const uint64_t *B,
              uint32_t rows, uint32_t inner, uint32_t cols, uint32_t b_cols, uint32_t b_col0, uint32_t out_cols,
              uint32_t out_col0, uint32_t fold, int variant) {
    Emu<T> e(logn, limbs, moduli);
    uint64_t qmax = 0;
    for (uint32_t i = 0; i < limbs; i++) qmax = moduli[i] > qmax ? moduli[i] : qmax;
    MatmulGeom g{};
    g.rows = rows; g.inner = inner; g.cols = cols; g.b_cols = b_cols; g.b_col0 = b_col0;
    g.out_cols = out_cols; g.out_col0 = out_col0; g.limbs = limbs; g.logn = (uint32_t)logn;
    g.fold = fold ? fold : host::mac_fold_terms(qmax, sizeof(T) == 8);
    const size_t pw = (size_t)limbs << logn;
    auto a = narrow<T>(A, (size_t)rows * inner * pw), b = narrow<T>(B, (size_t)inner * b_cols * pw);
    auto o = narrow<T>(out, (size_t)rows * out_cols * pw);
    constexpr int VMAX = 16 / (int)sizeof(T);
    if constexpr (sizeof(T) == 8) {
        if (e.f64) {  // the FP64-pipe product the launcher picks for these contexts (MatrixArith.inl run_matmul)
            g.fold = fold ? fold : host::f64_mac_fold_terms(qmax);
            g_f64_blocks++;
            switch (variant) {
                case 0: matmul_f64_grid<VMAX, 1, 6>(o, a, b, g, e.ftab.limb.data()); break;
                case 1: matmul_f64_grid<VMAX, 2, 4>(o, a, b, g, e.ftab.limb.data()); break;
                case 2: matmul_f64_grid<VMAX / 2, 4, 4>(o, a, b, g, e.ftab.limb.data()); break;
                default: matmul_f64_grid<VMAX, 1, 1>(o, a, b, g, e.ftab.limb.data()); break;
            }
            widen(o, out);
            return 0;
        }
    }
    switch (variant) {
        case 0: matmul_grid<T, VMAX, 1, 4>(o, a, b, g, e.tab.limb.data()); break;
        case 1: matmul_grid<T, VMAX, 2, 4>(o, a, b, g, e.tab.limb.data()); break;
        case 2:  // the four-row tiles of run_matmul: 4 x 2 full vectors for 32-bit words, 4 x 4 half vectors for 64-bit words
            if constexpr (sizeof(T) == 4) matmul_grid<T, VMAX, 4, 2>(o, a, b, g, e.tab.limb.data());
            else matmul_grid<T, VMAX / 2, 4, 4>(o, a, b, g, e.tab.limb.data());
            break;
        default: matmul_grid<T, 1, 1, 1>(o, a, b, g, e.tab.limb.data()); break;
    }
    widen(o, out);
    return 0;
}

bool wide_for(const uint64_t *moduli, uint32_t limbs) {
    uint64_t qmax = 0;
    for (uint32_t i = 0; i < limbs; i++) qmax = moduli[i] > qmax ? moduli[i] : qmax;
    return host::needs_wide(qmax);
}

// ---- samplers (sampling.cuh): the same per-element functions the kernels call, looped over
// the kernels' index space.  Coefficient-domain u64 words in host layout.
template <typename T> int do_sample_distribution(int logn, uint32_t limbs, const uint64_t *moduli, int dist, double sigma,
                                                 const uint64_t *seed, size_t rows, size_t local_cols, size_t full_cols,
                                                 size_t col_offset, uint64_t *out) {
    Emu<T> e(logn, limbs, moduli);
    const size_t n = (size_t)1 << logn;
    Seed256 sd;
    for (int i = 0; i < 4; i++) sd.w[i] = seed[i];
    for (size_t poly = 0; poly < rows * local_cols; poly++) {
        const size_t row = poly / local_cols, col = poly - row * local_cols;
        const uint64_t gp = row * full_cols + col_offset + col;
        for (size_t c = 0; c < n; c++) {
            if (dist == kDistUniform) {
                for (uint32_t l = 0; l < limbs; l++)
                    out[(poly * limbs + l) * n + c] = draw_uniform_residue(sd, gp, c, l, (uint64_t)e.tab.limb[l].q);
            } else {
                const int64_t z = draw_small_integer(sd, dist, sigma, gp, c);
                for (uint32_t l = 0; l < limbs; l++) out[(poly * limbs + l) * n + c] = small_to_residue<T>(z, e.tab.limb[l]);
            }
        }
    }
    return 0;
}
template <typename T> int do_gauss_samp(int logn, uint32_t limbs, const uint64_t *moduli, const uint64_t *src, size_t rows,
                                        size_t cols, uint32_t base_bits, uint32_t dpt, double c, const uint64_t *seed,
                                        uint64_t *out) {
    Emu<T> e(logn, limbs, moduli);
    const size_t n = (size_t)1 << logn, ell = (size_t)dpt * limbs;
    Seed256 sd;
    for (int i = 0; i < 4; i++) sd.w[i] = seed[i];
    for (size_t poly = 0; poly < rows * cols; poly++) {
        const size_t row = poly / cols, col = poly - row * cols;
        for (uint32_t tower = 0; tower < limbs; tower++)
            for (size_t ci = 0; ci < n; ci++) {
                int64_t digits[kMaxGaussDigits];
                const uint64_t v = src[(poly * limbs + tower) * n + ci];
                if (dpt <= 8) gq_sample_digits<8>(sd, v, (uint64_t)e.tab.limb[tower].q, base_bits, dpt, c, tower, poly, ci, digits);
                else gq_sample_digits<kMaxGaussDigits>(sd, v, (uint64_t)e.tab.limb[tower].q, base_bits, dpt, c, tower, poly, ci, digits);
                for (uint32_t dg = 0; dg < dpt; dg++) {
                    const size_t op = (row * ell + (size_t)tower * dpt + dg) * cols + col;
                    for (uint32_t l = 0; l < limbs; l++) out[(op * limbs + l) * n + ci] = small_to_residue<T>(digits[dg], e.tab.limb[l]);
                }
            }
    }
    return 0;
}
template <typename T> int do_sample_p1(int logn, uint32_t limbs, const uint64_t *moduli, const uint64_t *a, const uint64_t *b,
                                       const uint64_t *dm, const uint64_t *tp2, uint32_t d, size_t cols, double sigma, double s,
                                       double dgg, const uint64_t *seed, uint64_t *out) {
    Emu<T> e(logn, limbs, moduli);
    const size_t n = (size_t)1 << logn;
    const uint32_t m = 2 * d;
    const uint64_t q0 = moduli[0];
    Seed256 sd;
    for (int i = 0; i < 4; i++) sd.w[i] = seed[i];
    std::vector<double> cov((size_t)m * m * n), upd((size_t)m * m * n), sdv((size_t)m * n), mean(m);
    std::vector<int64_t> zs(m);
    const double c_scale = -(sigma * sigma) / (s * s - sigma * sigma);
    for (size_t ci = 0; ci < n; ci++) {
        auto load = [&](int which, uint32_t i, uint32_t j) -> int64_t {
            const uint64_t *base = which == 0 ? a : (which == 1 ? b : dm);
            return centred(base[((size_t)i * d + j) * limbs * n + ci], q0);
        };
        p1_factor_coeff(load, d, n, ci, sigma, s, dgg, cov.data(), sdv.data(), upd.data());
    }
    for (size_t col = 0; col < cols; col++)
        for (size_t ci = 0; ci < n; ci++) {
            auto load = [&](uint32_t r) -> int64_t { return centred(tp2[((size_t)r * cols + col) * limbs * n + ci], q0); };
            p1_sample_column(load, sd, m, n, col, ci, c_scale, sdv.data(), upd.data(), mean.data(), zs.data(), 1);
            for (uint32_t r = 0; r < m; r++)
                for (uint32_t l = 0; l < limbs; l++)
                    out[(((size_t)r * cols + col) * limbs + l) * n + ci] = small_to_residue<T>(zs[r], e.tab.limb[l]);
        }
    return 0;
}
// the lane-balanced kernels' thread bodies, walked over their thread index space (K tasks per thread, tests only)
template <typename T, int K>
int do_sample_gauss_batched(int logn, uint32_t limbs, const uint64_t *moduli, double sigma, const uint64_t *seed, size_t rows,
                            size_t local_cols, size_t full_cols, size_t col_offset, uint64_t *out) {
    Emu<T> e(logn, limbs, moduli);
    Seed256 sd;
    for (int i = 0; i < 4; i++) sd.w[i] = seed[i];
    SampleGeom g;
    g.polys = rows * local_cols, g.local_cols = local_cols, g.full_cols = full_cols, g.col_offset = col_offset;
    g.n = 1u << logn, g.limbs = limbs;
    if (g.n % (32 * K)) return 1;
    std::vector<T> dst(g.polys * limbs * g.n);
    const size_t threads = g.polys * g.n / K;
    for (size_t tid = 0; tid < threads; tid++) sample_gauss_batch_thread<T, K>(tid, dst.data(), g, sigma, sd, e.tab.limb.data());
    for (size_t i = 0; i < dst.size(); i++) out[i] = (uint64_t)dst[i];
    return 0;
}
}  // namespace

extern "C" {
void emu_stream_words(const uint64_t *seed, uint64_t s0, uint64_t s1, uint64_t s2, uint64_t domain, size_t count, uint64_t *out) {
    Seed256 sd;
    for (int i = 0; i < 4; i++) sd.w[i] = seed[i];
    ChaChaStream g;
    g.open(sd, s0, s1, s2, domain);
    for (size_t i = 0; i < count; i++) out[i] = g.next();
}
int emu_sample_distribution(int force_wide, int logn, uint32_t limbs, const uint64_t *moduli, int dist, double sigma,
                            const uint64_t *seed, size_t rows, size_t local_cols, size_t full_cols, size_t col_offset,
                            uint64_t *out) {
    try {
        if (force_wide || wide_for(moduli, limbs))
            return do_sample_distribution<uint64_t>(logn, limbs, moduli, dist, sigma, seed, rows, local_cols, full_cols, col_offset, out);
        return do_sample_distribution<uint32_t>(logn, limbs, moduli, dist, sigma, seed, rows, local_cols, full_cols, col_offset, out);
    } catch (...) { return 1; }
}
int emu_gauss_samp(int force_wide, int logn, uint32_t limbs, const uint64_t *moduli, const uint64_t *src, size_t rows, size_t cols,
                   uint32_t base_bits, uint32_t dpt, double c, const uint64_t *seed, uint64_t *out) {
    try {
        if (force_wide || wide_for(moduli, limbs))
            return do_gauss_samp<uint64_t>(logn, limbs, moduli, src, rows, cols, base_bits, dpt, c, seed, out);
        return do_gauss_samp<uint32_t>(logn, limbs, moduli, src, rows, cols, base_bits, dpt, c, seed, out);
    } catch (...) { return 1; }
}
#define EMU_BATCH_SWITCH(CALL)          \
    switch (batch) {                    \
        case 2: return CALL(2);         \
        case 4: return CALL(4);         \
        case 8: return CALL(8);         \
        default: return 1;              \
    }
int emu_sample_gauss_batched(int batch, int force_wide, int logn, uint32_t limbs, const uint64_t *moduli, double sigma,
                             const uint64_t *seed, size_t rows, size_t local_cols, size_t full_cols, size_t col_offset, uint64_t *out) {
    if (force_wide || wide_for(moduli, limbs)) {
#define EMU_CALL(K) do_sample_gauss_batched<uint64_t, K>(logn, limbs, moduli, sigma, seed, rows, local_cols, full_cols, col_offset, out)
        EMU_BATCH_SWITCH(EMU_CALL)
#undef EMU_CALL
    }
#define EMU_CALL(K) do_sample_gauss_batched<uint32_t, K>(logn, limbs, moduli, sigma, seed, rows, local_cols, full_cols, col_offset, out)
    EMU_BATCH_SWITCH(EMU_CALL)
#undef EMU_CALL
}
// stream words each (poly, tower, coeff) task of the G-lattice sampler consumes, in the kernel's thread order: the input of
// the warp model in tests/test_samplers.py (a warp runs max-over-lanes trips; a lane is busy for its own count)
int emu_gauss_samp_words(int logn, uint32_t limbs, const uint64_t *moduli, const uint64_t *src, size_t polys, uint32_t base_bits,
                         uint32_t dpt, double c, const uint64_t *seed, uint32_t *words) {
    if (dpt > 8) return 1;
    const size_t n = (size_t)1 << logn;
    Seed256 sd;
    for (int i = 0; i < 4; i++) sd.w[i] = seed[i];
    for (size_t poly = 0; poly < polys; poly++)
        for (uint32_t tower = 0; tower < limbs; tower++)
            for (size_t ci = 0; ci < n; ci++) {
                int64_t digits[8];
                const size_t idx = (poly * limbs + tower) * n + ci;
                gq_sample_digits<8>(sd, src[idx], moduli[tower], base_bits, dpt, c, tower, poly, ci, digits, &words[idx]);
            }
    return 0;
}
int emu_sample_p1(int force_wide, int logn, uint32_t limbs, const uint64_t *moduli, const uint64_t *a, const uint64_t *b,
                  const uint64_t *dm, const uint64_t *tp2, uint32_t d, size_t cols, double sigma, double s, double dgg,
                  const uint64_t *seed, uint64_t *out) {
    try {
        if (force_wide || wide_for(moduli, limbs))
            return do_sample_p1<uint64_t>(logn, limbs, moduli, a, b, dm, tp2, d, cols, sigma, s, dgg, seed, out);
        return do_sample_p1<uint32_t>(logn, limbs, moduli, a, b, dm, tp2, d, cols, sigma, s, dgg, seed, out);
    } catch (...) { return 1; }
}
void emu_set_f64(int on) { g_f64 = on; }
void emu_set_cluster(int on) { g_cluster = on; }
void emu_set_fuse_top(int on) { g_fuse_top = on; }
size_t emu_f64_blocks(void) { return g_f64_blocks; }
// force_wide: 0 = the context rule (u32 words when all moduli < 2^30), 1 = u64 words regardless
int emu_ntt(int force_wide, int logn, uint32_t limbs, const uint64_t *moduli, uint64_t *data, size_t polys, int inverse,
            int logb_cap) {
    try {
        if (force_wide || wide_for(moduli, limbs)) return do_ntt<uint64_t>(logn, limbs, moduli, data, polys, inverse, logb_cap);
        return do_ntt<uint32_t>(logn, limbs, moduli, data, polys, inverse, logb_cap);
    } catch (...) { return 1; }
}
int emu_decompose(int force_wide, int logn, uint32_t limbs, const uint64_t *moduli, const uint64_t *src, size_t rows,
                  size_t src_cols, size_t src_col0, size_t cols, uint32_t base_bits, int small, int to_eval, uint64_t *out,
                  int logb_cap) {
    try {
        if (force_wide || wide_for(moduli, limbs))
            return do_decompose<uint64_t>(logn, limbs, moduli, src, rows, src_cols, src_col0, cols, base_bits, small, to_eval, out, logb_cap);
        return do_decompose<uint32_t>(logn, limbs, moduli, src, rows, src_cols, src_col0, cols, base_bits, small, to_eval, out, logb_cap);
    } catch (...) { return 1; }
}
int emu_matmul(int force_wide, int logn, uint32_t limbs, const uint64_t *moduli, uint64_t *out, const uint64_t *A,
               const uint64_t *B, uint32_t rows, uint32_t inner, uint32_t cols, uint32_t b_cols, uint32_t b_col0,
               uint32_t out_cols, uint32_t out_col0, uint32_t fold, int variant) {
    try {
        if (force_wide || wide_for(moduli, limbs))
            return do_matmul<uint64_t>(logn, limbs, moduli, out, A, B, rows, inner, cols, b_cols, b_col0, out_cols, out_col0, fold, variant);
        return do_matmul<uint32_t>(logn, limbs, moduli, out, A, B, rows, inner, cols, b_cols, b_col0, out_cols, out_col0, fold, variant);
    } catch (...) { return 1; }
}
}
