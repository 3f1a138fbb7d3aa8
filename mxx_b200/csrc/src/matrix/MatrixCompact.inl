// MatrixCompact.inl — gpu_matrix_store_compact_bytes / gpu_matrix_load_compact_bytes.
// Replaces gpu_poly_store_compact_bytes / gpu_poly_load_compact_bytes of cuda/src/matrix/MatrixSerde.cu
// (:1066-1630, :1632-1944; the matrix entry points :1946-1970 forward to them).  Kernels: csrc/include/compact.cuh.

namespace {

typedef unsigned __int128 u128_t;

// Q = product of the first `limbs` moduli as little-endian 64-bit words (MatrixSerde.cu:472-520)
std::vector<uint64_t> modulus_words(const GpuContext *ctx, size_t limbs, size_t words) {
    std::vector<uint64_t> q(words, 0);
    q[0] = 1;
    for (size_t i = 0; i < limbs; i++) {
        uint64_t carry = 0;
        for (size_t w = 0; w < words; w++) {
            const u128_t t = (u128_t)q[w] * ctx->moduli[i] + carry;
            q[w] = (uint64_t)t;
            carry = (uint64_t)(t >> 64);
        }
    }
    return q;
}

size_t compact_words(const GpuContext *ctx, size_t limbs) {
    size_t bits = 0;
    for (size_t i = 0; i < limbs; i++) bits += 64 - (size_t)__builtin_clzll(ctx->moduli[i]);
    return std::max<size_t>(1, (bits + 63) / 64);  // MatrixSerde.cu:1122-1129
}

// q_j^-1 mod q_i for j < i, [j * L + i]; built once per context
int ensure_garner(GpuContext *ctx) {
    std::lock_guard<std::mutex> lock(ctx->lazy_mutex);
    if (ctx->d_garner) return 0;
    const size_t L = ctx->moduli.size();
    std::vector<uint64_t> tab(L * L, 0);
    for (size_t j = 0; j < L; j++)
        for (size_t i = j + 1; i < L; i++) {
            const uint64_t q = ctx->moduli[i];
            // q_j^-1 mod q_i by Fermat (the moduli are distinct primes)
            uint64_t base = ctx->moduli[j] % q, e = q - 2, r = 1;
            for (; e; e >>= 1) {
                if (e & 1) r = (uint64_t)((u128_t)r * base % q);
                base = (uint64_t)((u128_t)base * base % q);
            }
            if ((u128_t)r * (ctx->moduli[j] % q) % q != 1) return gp::set_error("CRT moduli are not invertible modulo each other");
            tab[j * L + i] = r;
        }
    void *d = nullptr;
    GP_CUDA(cudaMalloc(&d, tab.size() * sizeof(uint64_t)), "cudaMalloc(garner)");
    cudaError_t e = cudaMemcpy(d, tab.data(), tab.size() * sizeof(uint64_t), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
        cudaFree(d);
        return gp::set_cuda_error(e, "cudaMemcpy(garner)");
    }
    ctx->d_garner = d;
    return 0;
}

template <typename T>
int run_compact_store(GpuContext *ctx, GpuMatrix *mat, uint8_t *payload_out, size_t capacity, uint16_t *out_bits, uint16_t *out_bpc,
                      size_t *out_len) {
    const size_t limbs = mat->limbs(), count = mat->polys() << ctx->logN, words = compact_words(ctx, limbs);
    if (words > (size_t)gp::kCompactMaxWords) return gp::set_error("words_per_coeff exceeds kernel limit in gpu_matrix_store_compact_bytes");
    std::vector<uint64_t> q = modulus_words(ctx, limbs, words), h(words);
    for (size_t w = 0; w < words; w++) h[w] = (q[w] >> 1) | (w + 1 < words ? q[w + 1] << 63 : 0);  // floor(Q / 2)
    cudaStream_t st = ctx->stream;
    Scratch consts, mag, sign, maxb, payload;
    if (consts.alloc(2 * words * sizeof(uint64_t), st) || mag.alloc(count * words * sizeof(uint64_t), st) || sign.alloc(count, st) ||
        maxb.alloc(sizeof(unsigned int), st))
        return 1;
    std::vector<uint64_t> both(q);
    both.insert(both.end(), h.begin(), h.end());
    GP_CUDA(cudaMemcpyAsync(consts.p, both.data(), both.size() * sizeof(uint64_t), cudaMemcpyHostToDevice, st), "cudaMemcpyAsync(Q)");
    GP_CUDA(cudaMemsetAsync(maxb.p, 0, sizeof(unsigned int), st), "cudaMemsetAsync(max bits)");
    gp::CompactGeom g{(uint32_t)limbs, (uint32_t)ctx->logN, (uint32_t)words, count};
    const unsigned threads = 128, grid = (unsigned)((count + threads - 1) / threads);
    const auto *lc = (const gp::LimbConst<T> *)ctx->d_limb;
    const uint64_t *qw = (const uint64_t *)consts.p;
    if (limbs <= 8 && words <= 8)
        gp::compact_reconstruct_kernel<T, 8><<<grid, threads, 0, st>>>((const T *)mat->data, g, lc, (const uint64_t *)ctx->d_garner,
                                                                      (uint32_t)ctx->moduli.size(), qw, qw + words, (uint64_t *)mag.p,
                                                                      (uint8_t *)sign.p, (unsigned int *)maxb.p);
    else
        gp::compact_reconstruct_kernel<T, gp::kCompactMaxWords><<<grid, threads, 0, st>>>(
            (const T *)mat->data, g, lc, (const uint64_t *)ctx->d_garner, (uint32_t)ctx->moduli.size(), qw, qw + words, (uint64_t *)mag.p,
            (uint8_t *)sign.p, (unsigned int *)maxb.p);
    gp::count_launch();
    GP_CHECK_LAUNCH("compact_reconstruct_kernel");
    unsigned int max_bits = 0;
    GP_CUDA(cudaMemcpyAsync(&max_bits, maxb.p, sizeof(unsigned int), cudaMemcpyDeviceToHost, st), "cudaMemcpyAsync(max bits)");
    GP_CUDA(cudaStreamSynchronize(st), "cudaStreamSynchronize(compact)");  // the bit width sizes the payload: a host value
    const unsigned int B = max_bits == 0 ? 0 : max_bits + 1;  // MatrixSerde.cu:1552-1553
    if (B > 0xffffu) return gp::set_error("centered max coeff bits exceed u16 range in gpu_matrix_store_compact_bytes");
    const size_t len = (count * (size_t)B + 7) / 8;
    if (len > capacity) return gp::set_error("payload capacity too small in gpu_matrix_store_compact_bytes");
    if (len) {
        const size_t padded = ((len + 7) / 8 + 2) * 8;
        if (payload.alloc(padded, st)) return 1;
        GP_CUDA(cudaMemsetAsync(payload.p, 0, padded, st), "cudaMemsetAsync(payload)");
        gp::compact_pack_kernel<<<grid, threads, 0, st>>>((const uint64_t *)mag.p, (const uint8_t *)sign.p, count, (uint32_t)words, B,
                                                          (unsigned long long *)payload.p);
        gp::count_launch();
        GP_CHECK_LAUNCH("compact_pack_kernel");
        GP_CUDA(cudaMemcpyAsync(payload_out, payload.p, len, cudaMemcpyDeviceToHost, st), "cudaMemcpyAsync(payload D2H)");
        GP_CUDA(cudaStreamSynchronize(st), "cudaStreamSynchronize(compact payload)");
    }
    *out_bits = (uint16_t)B;
    *out_bpc = (uint16_t)((B + 7) / 8);
    *out_len = len;
    return 0;
}

template <typename T>
int run_compact_load(GpuContext *ctx, GpuMatrix *mat, const uint8_t *payload, size_t len, uint32_t B) {
    const size_t limbs = mat->limbs(), count = mat->polys() << ctx->logN;
    cudaStream_t st = ctx->stream;
    Scratch dev;
    if (len) {
        if (dev.alloc(len, st)) return 1;
        GP_CUDA(cudaMemcpyAsync(dev.p, payload, len, cudaMemcpyHostToDevice, st), "cudaMemcpyAsync(payload H2D)");
    }
    gp::CompactGeom g{(uint32_t)limbs, (uint32_t)ctx->logN, 0, count};
    const unsigned threads = 128, grid = (unsigned)((count + threads - 1) / threads);
    gp::compact_unpack_kernel<T><<<grid, threads, 0, st>>>((T *)mat->data, g, (const gp::LimbConst<T> *)ctx->d_limb, (const uint8_t *)dev.p, len, B);
    gp::count_launch();
    GP_CHECK_LAUNCH("compact_unpack_kernel");
    // the caller may release `payload` as soon as the call returns (gpu_dcrt_poly.rs:1033-1036)
    if (len) GP_CUDA(cudaStreamSynchronize(st), "cudaStreamSynchronize(compact load)");
    return 0;
}
}  // namespace

extern "C" int gpu_matrix_store_compact_bytes(GpuMatrix *mat, uint8_t *payload_out, size_t payload_capacity, uint16_t *out_max_coeff_bits,
                                              uint16_t *out_bytes_per_coeff, size_t *out_payload_len) {
    if (!mat || !payload_out || !out_max_coeff_bits || !out_bytes_per_coeff || !out_payload_len)
        return gp::set_error("invalid gpu_matrix_store_compact_bytes arguments");  // MatrixSerde.cu:1074-1077
    *out_max_coeff_bits = 0;
    *out_bytes_per_coeff = 0;
    *out_payload_len = 0;
    GpuContext *ctx = mat->ctx;
    if (!ctx) return gp::set_error("invalid matrix in gpu_matrix_store_compact_bytes");
    if (mat->polys() == 0 || mat->level < 0) return 0;
    if (bind_device(ctx)) return 1;
    // the reference converts an EVAL matrix to coefficient form IN PLACE (MatrixSerde.cu:1108-1115)
    if (mat->format == GPU_POLY_FORMAT_EVAL && gpu_matrix_intt_all(mat)) return 1;
    if (ensure_garner(ctx)) return 1;
    return GP_DISPATCH_WIDTH(ctx, return run_compact_store<T>(ctx, mat, payload_out, payload_capacity, out_max_coeff_bits, out_bytes_per_coeff,
                                                              out_payload_len););
}

extern "C" int gpu_matrix_load_compact_bytes(GpuMatrix *mat, const uint8_t *payload, size_t payload_len, uint16_t max_coeff_bits) {
    if (!mat || !mat->ctx) return gp::set_error("invalid gpu_matrix_load_compact_bytes arguments");
    GpuContext *ctx = mat->ctx;
    const size_t count = mat->polys() << ctx->logN;
    mat->format = GPU_POLY_FORMAT_COEFF;
    if (count == 0 || mat->level < 0) return 0;
    if (max_coeff_bits == 0 && payload_len != 0)
        return gp::set_error("payload_len must be zero when max_coeff_bits is zero");  // MatrixSerde.cu:1671-1676
    const size_t expected = (count * (size_t)max_coeff_bits + 7) / 8;
    if (payload_len != expected) return gp::set_error("payload_len mismatch in gpu_matrix_load_compact_bytes");  // :1683-1694
    if (payload_len && !payload) return gp::set_error("null payload in gpu_matrix_load_compact_bytes");
    if (bind_device(ctx)) return 1;
    return GP_DISPATCH_WIDTH(ctx, return run_compact_load<T>(ctx, mat, payload, payload_len, max_coeff_bits););
}

// ---- extension: big-integer coefficients <-> residues on the device ------------------------------
// Replaces the per-coefficient BigUint `%` of GpuDCRTPolyMatrix::from_cpu_matrix (gpu_dcrt_poly.rs:793-803) /
// GpuDCRTPoly::from_coeffs (gpu.rs:903-920) and the host CRT reconstruction of GpuDCRTPoly::coeffs (gpu.rs:959-994).
extern "C" int gpu_matrix_load_bigint_batch(GpuMatrix *mat, const uint64_t *words, size_t words_per_coeff, int format,
                                            GpuEventSet **out_events) {
    if (!mat || !mat->ctx || !out_events) return gp::set_error("invalid gpu_matrix_load_bigint_batch arguments");
    *out_events = nullptr;
    if (format != GPU_POLY_FORMAT_COEFF && format != GPU_POLY_FORMAT_EVAL) return gp::set_error("invalid format in gpu_matrix_load_bigint_batch");
    GpuContext *ctx = mat->ctx;
    const size_t count = mat->polys() << ctx->logN;
    mat->format = format;
    if (count == 0 || mat->level < 0) return 0;
    if (!words || words_per_coeff == 0) return gp::set_error("null words in gpu_matrix_load_bigint_batch");
    if (bind_device(ctx)) return 1;
    cudaStream_t st = ctx->stream;
    Scratch dev;
    if (dev.alloc(count * words_per_coeff * sizeof(uint64_t), st)) return 1;
    GP_CUDA(cudaMemcpyAsync(dev.p, words, count * words_per_coeff * sizeof(uint64_t), cudaMemcpyHostToDevice, st), "cudaMemcpyAsync(bigint H2D)");
    gp::CompactGeom g{(uint32_t)mat->limbs(), (uint32_t)ctx->logN, 0, count};
    const unsigned threads = 128, grid = (unsigned)((count + threads - 1) / threads);
    int stt = GP_DISPATCH_WIDTH(
        ctx, const auto *lc = (const gp::LimbConst<T> *)ctx->d_limb;
        if (format == GPU_POLY_FORMAT_EVAL)
            gp::bigint_load_kernel<T, true><<<grid, threads, 0, st>>>((T *)mat->data, g, lc, (const uint64_t *)dev.p, (uint32_t)words_per_coeff);
        else
            gp::bigint_load_kernel<T, false><<<grid, threads, 0, st>>>((T *)mat->data, g, lc, (const uint64_t *)dev.p, (uint32_t)words_per_coeff);
        gp::count_launch(); GP_CHECK_LAUNCH("bigint_load_kernel"); return 0;);
    if (stt) return stt;
    return make_event_set(ctx, out_events);
}

// words_out[(poly * n + j) * words_per_coeff + w]: the coefficient in [0, Q) (Q = product of the active moduli);
// requires coefficient form (as GpuDCRTPoly::coeffs converts first); words_per_coeff >= ceil(bits(Q) / 64).
extern "C" int gpu_matrix_store_bigint_batch(const GpuMatrix *mat, uint64_t *words_out, size_t words_per_coeff, GpuEventSet **out_events) {
    if (!mat || !mat->ctx || !out_events) return gp::set_error("invalid gpu_matrix_store_bigint_batch arguments");
    *out_events = nullptr;
    GpuContext *ctx = mat->ctx;
    const size_t count = mat->polys() << ctx->logN;
    if (count == 0 || mat->level < 0) return 0;
    if (mat->format != GPU_POLY_FORMAT_COEFF) return gp::set_error("gpu_matrix_store_bigint_batch requires coeff format");
    if (!words_out) return gp::set_error("null words_out in gpu_matrix_store_bigint_batch");
    const size_t limbs = mat->limbs(), words = compact_words(ctx, limbs);
    if (words_per_coeff < words) return gp::set_error("words_per_coeff too small in gpu_matrix_store_bigint_batch");
    if (words > (size_t)gp::kCompactMaxWords) return gp::set_error("words_per_coeff exceeds kernel limit in gpu_matrix_store_bigint_batch");
    if (bind_device(ctx)) return 1;
    if (ensure_garner(ctx)) return 1;
    cudaStream_t st = ctx->stream;
    std::vector<uint64_t> q = modulus_words(ctx, limbs, words);
    Scratch consts, mag;
    if (consts.alloc(2 * words * sizeof(uint64_t), st) || mag.alloc(count * words * sizeof(uint64_t), st)) return 1;
    std::vector<uint64_t> both(q);
    both.insert(both.end(), q.begin(), q.end());
    GP_CUDA(cudaMemcpyAsync(consts.p, both.data(), both.size() * sizeof(uint64_t), cudaMemcpyHostToDevice, st), "cudaMemcpyAsync(Q)");
    GP_CUDA(cudaStreamSynchronize(st), "cudaStreamSynchronize(Q)");  // `both` is a stack object
    gp::CompactGeom g{(uint32_t)limbs, (uint32_t)ctx->logN, (uint32_t)words, count};
    const unsigned threads = 128, grid = (unsigned)((count + threads - 1) / threads);
    const uint64_t *qw = (const uint64_t *)consts.p;
    int stt = GP_DISPATCH_WIDTH(
        ctx, const auto *lc = (const gp::LimbConst<T> *)ctx->d_limb;
        if (limbs <= 8 && words <= 8)
            gp::compact_reconstruct_kernel<T, 8, false><<<grid, threads, 0, st>>>((const T *)mat->data, g, lc, (const uint64_t *)ctx->d_garner,
                (uint32_t)ctx->moduli.size(), qw, qw + words, (uint64_t *)mag.p, nullptr, nullptr);
        else
            gp::compact_reconstruct_kernel<T, gp::kCompactMaxWords, false><<<grid, threads, 0, st>>>((const T *)mat->data, g, lc,
                (const uint64_t *)ctx->d_garner, (uint32_t)ctx->moduli.size(), qw, qw + words, (uint64_t *)mag.p, nullptr, nullptr);
        gp::count_launch(); GP_CHECK_LAUNCH("compact_reconstruct_kernel"); return 0;);
    if (stt) return stt;
    if (words_per_coeff > words) std::fill_n(words_out, count * words_per_coeff, (uint64_t)0);
    GP_CUDA(cudaMemcpy2DAsync(words_out, words_per_coeff * 8, mag.p, words * 8, words * 8, count, cudaMemcpyDeviceToHost, st),
            "cudaMemcpy2DAsync(bigint D2H)");
    return make_event_set(ctx, out_events);
}
