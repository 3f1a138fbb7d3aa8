// MatrixDecomposeCluster.inl — host side of the fused gadget decomposition + cluster NTT (ntt_cluster.cuh
// fast_decompose_cluster_kernel): rings of 2..8 blocks, digits extracted by the cluster's push phase.
// Own translation unit (MatrixDecomposeCluster.cu).

namespace {
template <typename K, int LOGB, int LC, bool NOCSUB>
int launch_cluster_decompose(GpuContext *ctx, const K *src, K *out, const gp::DecompGeom &g, size_t opl) {
    static std::atomic<uint32_t> done{0};
    constexpr size_t smem = gp::fast_smem_bytes<K, LOGB>();
    auto kernel = gp::fast_decompose_cluster_kernel<K, LOGB, LC, NOCSUB>;
    if (optin_smem(kernel, smem, ctx->device, done)) return 1;
    kernel<<<(unsigned)(opl << LC), gp::FastCfg<K, LOGB>::kThreads, smem, gp::cur(ctx)>>>(src, out, g, fwd_tab<K>(ctx), limb_tab<K>(ctx));
    gp::count_launch();
    GP_CHECK_LAUNCH("fast_decompose_cluster_kernel");
    return 0;
}
template <typename K> int dispatch_cluster_decompose(GpuContext *ctx, const K *src, K *out, const gp::DecompGeom &g, size_t opl, int lc) {
    constexpr int LB = sizeof(K) == 4 ? 14 : 13;
    if constexpr (gp::is_f64<K>::value) {
#define GP_CALL(LC) launch_cluster_decompose<K, LB, LC, false>(ctx, src, out, g, opl)
        GP_CLUSTER_SWITCH(lc, GP_CALL)
#undef GP_CALL
    } else {
#define GP_CALL(LC) (ctx->nocsub ? launch_cluster_decompose<K, LB, LC, true>(ctx, src, out, g, opl) : launch_cluster_decompose<K, LB, LC, false>(ctx, src, out, g, opl))
        GP_CLUSTER_SWITCH(lc, GP_CALL)
#undef GP_CALL
    }
}

}  // namespace

template <typename T> int gpm::run_cluster_decompose(GpuContext *ctx, const T *src, T *out, const gp::DecompGeom &g, size_t opl, int lc) {
    if constexpr (sizeof(T) == 8) {
        if (ctx->f64_ntt) return dispatch_cluster_decompose<gp::F64>(ctx, (const gp::F64 *)src, (gp::F64 *)out, g, opl, lc);
    }
    return dispatch_cluster_decompose<T>(ctx, src, out, g, opl, lc);
}
template int gpm::run_cluster_decompose<uint32_t>(GpuContext *, const uint32_t *, uint32_t *, const gp::DecompGeom &, size_t, int);
template int gpm::run_cluster_decompose<uint64_t>(GpuContext *, const uint64_t *, uint64_t *, const gp::DecompGeom &, size_t, int);
