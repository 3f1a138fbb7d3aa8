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
#include <stdint.h>
#include <string.h>

#include <vector>

#include "arith.cuh"
#include "host_tables.hpp"
#include "ntt.cuh"
#include "ntt_fast.cuh"

namespace {
using namespace gp;

template <typename T> struct Emu {
    int logn;
    uint32_t limbs;
    host::Tables<T> tab;
    Emu(int logn_, uint32_t limbs_, const uint64_t *moduli) : logn(logn_), limbs(limbs_) {
        std::vector<uint64_t> m(moduli, moduli + limbs_), psi;
        for (uint64_t q : m) psi.push_back(host::primitive_2n_root(q, (uint64_t)1 << logn_));
        tab = host::build_tables<T>(logn_, m, psi);
        uint64_t qmax = 0;
        for (uint64_t q : m) qmax = q > qmax ? q : qmax;
        nocsub = host::nocsub_ok(qmax, sizeof(T) == 8);
    }
    // ---- compile-time-planned kernels (ntt_fast.cuh), same selection rule as MatrixNTT.inl ----
    bool fast_ok(int logb_cap) const { return logb_cap == 0 && fast_supported<T>(fast_block_log<T>(logn)); }
    template <int LOGB> void fast_fwd_blocks(std::vector<T> &d, size_t nbid, bool nocsub) {
        std::vector<uint32_t> sm(fast_smem_bytes<T, LOGB>() / 4 + 4);
        for (size_t bid = 0; bid < nbid; bid++) {
            if (nocsub) fast_fwd_cta<T, LOGB, true>((uint32_t)bid, sm.data(), d.data(), d.data(), limbs, logn, tab.fwd.data(), tab.limb.data());
            else fast_fwd_cta<T, LOGB, false>((uint32_t)bid, sm.data(), d.data(), d.data(), limbs, logn, tab.fwd.data(), tab.limb.data());
        }
    }
    template <int LOGB> void fast_inv_blocks(std::vector<T> &d, size_t nbid) {
        std::vector<uint32_t> sm(fast_smem_bytes<T, LOGB>() / 4 + 4);
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
    bool nocsub = false;
    int block_log(int logb_cap) const {
        if (fast_ok(logb_cap)) return fast_block_log<T>(logn);
        int lb = ntt_block_log<T>(logn);
        if (logb_cap > 0 && logb_cap < lb) lb = logb_cap;
        return lb;
    }
    void fwd(std::vector<T> &d, size_t polys, int logb_cap) {
        const int logb = block_log(logb_cap), top = logn - logb;
        const size_t pl = polys * limbs;
        std::vector<T> sm(ntt_smem_bytes<T>(logb) / sizeof(T) + 1);
        if (top > 0) {
            int passes[8], np = 0;
            split_top(top, passes, &np);
            uint32_t lspan = (uint32_t)logn;
            for (int i = 0; i < np; i++) {
                const size_t total = pl << (logn - passes[i]);
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
    DecompGeom g;
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
    if (to_eval && top == 0 && e.fast_ok(logb_cap)) {
        std::vector<uint32_t> fsm((sizeof(T) / 4) * ((1u << logb) + (1u << (logb - 4))) + 8);
#define DEC_CASE(LB)                                                                                                          \
    case LB:                                                                                                                  \
        for (size_t bid = 0; bid < opl; bid++) {                                                                              \
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
    switch (variant) {
        case 0: matmul_grid<T, VMAX, 1, 4>(o, a, b, g, e.tab.limb.data()); break;
        case 1: matmul_grid<T, VMAX, 2, 4>(o, a, b, g, e.tab.limb.data()); break;
        case 2: matmul_grid<T, VMAX / 2, 4, 4>(o, a, b, g, e.tab.limb.data()); break;
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
}  // namespace

extern "C" {
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
