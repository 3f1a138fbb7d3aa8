// MatrixDeferred.inl — symbols of the reference ABI that are outside the hot path
// (SURVEY §8f "next" rows: samplers, trapdoor perturbation).  They
// are exported so the Rust crate links against this library unchanged; each fails loudly
// (non-zero + message) rather than falling back to any CPU path.
// Reference implementations: cuda/src/matrix/MatrixSampling.cu, MatrixTrapdoor.cu,
// MatrixSerde.cu:1066-1970.

#define GP_NOT_BUILT(name) return gp::set_error(name ": not built in this library yet (outside the DCRT matrix-mul hot path)")

extern "C" int gpu_matrix_gauss_samp_gq_arb_base(GpuMatrix *, uint32_t, double, double, GpuRngSeed, GpuMatrix *) {
    GP_NOT_BUILT("gpu_matrix_gauss_samp_gq_arb_base");
}
extern "C" int gpu_matrix_create_p1_covariance_cache(const GpuMatrix *, const GpuMatrix *, const GpuMatrix *, double, double,
                                                     double, GpuP1CovarianceCache **out_cache) {
    if (out_cache) *out_cache = nullptr;
    GP_NOT_BUILT("gpu_matrix_create_p1_covariance_cache");
}
extern "C" void gpu_matrix_destroy_p1_covariance_cache(GpuP1CovarianceCache *) {}
extern "C" int gpu_matrix_sample_p1_full_cached(const GpuP1CovarianceCache *, const GpuMatrix *, GpuRngSeed, GpuMatrix *) {
    GP_NOT_BUILT("gpu_matrix_sample_p1_full_cached");
}
extern "C" int gpu_matrix_sample_distribution(GpuMatrix *, int, double, GpuRngSeed) {
    GP_NOT_BUILT("gpu_matrix_sample_distribution");
}
extern "C" int gpu_matrix_sample_distribution_columns(GpuMatrix *, int, double, GpuRngSeed, size_t, size_t) {
    GP_NOT_BUILT("gpu_matrix_sample_distribution_columns");
}
