// compact.cuh — the compact wire format of a matrix, on the device.
//
// Replaces serde_reconstruct_rns_to_words_kernel (cuda/src/matrix/MatrixSerde.cu:162), serde_center_coeff_words_kernel
// (:280), serde_pack_centered_coeff_words_bits_kernel (:340) and serde_unpack_packed_coeffs_mod_kernel (:388).
// Format (defined by those kernels and by gpu_poly_store/load_compact_bytes, :1066-1944):
//   * every coefficient (all polys of the matrix, row-major, coefficient form) is CRT-reconstructed to
//     X in [0, Q), Q = product of the active moduli, and centred: negative iff X > floor(Q/2), magnitude Q - X;
//   * B = 0 if every magnitude is 0, else (largest magnitude bit width) + 1;
//   * coefficient k occupies bits [k*B, (k+1)*B) of the payload, least-significant bit first inside a byte:
//     bits 0..B-2 the magnitude, bit B-1 the sign; payload length = ceil(count * B / 8).
// Here: one thread per coefficient does Garner's mixed-radix conversion with Shoup-free Barrett products, keeps
// the multiword magnitude in a scratch row, and a second kernel scatters B bits per coefficient with at most
// three 64-bit atomic ORs — no per-bit loops, no per-byte gather.
#pragma once
#include "modarith.cuh"

namespace gp {

constexpr int kCompactMaxWords = 64;  // 64 limbs x 62 bits

struct CompactGeom {
    uint32_t limbs, logn, words;  // active limbs, log2 n, 64-bit words per reconstructed coefficient
    size_t count;                 // coefficients in the matrix (polys * n)
};

// (a * b) mod q for a, b < q through the context's Barrett constants
template <typename T> GP_HD uint64_t cmul(uint64_t a, uint64_t b, const LimbConst<T> &c) { return (uint64_t)mul_mod<T>((T)a, (T)b, c); }

// Pass 1: residues -> centred magnitude words (scratch[k][words]) + sign byte; max bit width by atomicMax.
// garner[j * limbs_total + i] = q_j^-1 mod q_i (j < i); qwords / hwords = Q and floor(Q/2), little endian.
// CENTRE = false keeps X itself (big-integer store): no sign, no width.
template <typename T, int WMAX, bool CENTRE = true>
__global__ void compact_reconstruct_kernel(const T *__restrict__ data, CompactGeom g, const LimbConst<T> *__restrict__ lc,
                                           const uint64_t *__restrict__ garner, uint32_t garner_stride,
                                           const uint64_t *__restrict__ qwords, const uint64_t *__restrict__ hwords,
                                           uint64_t *__restrict__ mag, uint8_t *__restrict__ sign, unsigned int *__restrict__ max_bits) {
    const size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned int width = 0;
    if (k < g.count) {
        const size_t n = (size_t)1 << g.logn, poly = k >> g.logn, j = k & (n - 1);
        const T *p = data + ((poly * g.limbs) << g.logn) + j;
        // Garner: v_i = (((x_i - v_0) q_0^-1 - v_1) q_1^-1 - ...) mod q_i
        uint64_t v[WMAX];
        for (uint32_t i = 0; i < g.limbs; i++) {
            const LimbConst<T> c = lc[i];
            const uint64_t q = (uint64_t)c.q;
            uint64_t t = (uint64_t)p[(size_t)i << g.logn];
            for (uint32_t a = 0; a < i; a++) {
                const uint64_t va = v[a] >= q ? v[a] % q : v[a];
                t = t >= va ? t - va : t + q - va;
                t = cmul<T>(t, garner[(size_t)a * garner_stride + i], c);
            }
            v[i] = t;
        }
        // X = v_0 + q_0 (v_1 + q_1 (v_2 + ...)): Horner from the top, multiword times word
        uint64_t X[WMAX];
        for (uint32_t w = 0; w < g.words; w++) X[w] = 0;
        for (int i = (int)g.limbs - 1; i >= 0; i--) {
            const uint64_t q = (uint64_t)lc[i].q;
            uint64_t carry = v[i];
            for (uint32_t w = 0; w < g.words; w++) {
                const uint64_t lo = X[w] * q, hi = mulhi(X[w], q);
                const uint64_t s = lo + carry;
                carry = hi + (s < lo);
                X[w] = s;
            }
        }
        // centre: negative iff X > floor(Q/2)
        int cmp = 0;
        for (int w = (int)g.words - 1; w >= 0 && cmp == 0; w--) cmp = X[w] > hwords[w] ? 1 : (X[w] < hwords[w] ? -1 : 0);
        const bool neg = CENTRE && cmp > 0;
        if (neg) {
            uint64_t borrow = 0;
            for (uint32_t w = 0; w < g.words; w++) {
                const uint64_t a = qwords[w], b = X[w];
                const uint64_t d = a - b - borrow;
                borrow = (a < b) || (a == b && borrow);
                X[w] = d;
            }
        }
        for (int w = (int)g.words - 1; w >= 0; w--)
            if (X[w]) {
                width = (unsigned)w * 64u + (64u - (unsigned)__clzll((long long)X[w]));
                break;
            }
        for (uint32_t w = 0; w < g.words; w++) mag[k * g.words + w] = X[w];
        if (CENTRE) sign[k] = neg ? 1 : 0;
    }
    // one atomic per warp
    for (int o = 16; o; o >>= 1) width = max(width, __shfl_xor_sync(0xffffffffu, width, o));
    if (CENTRE && (threadIdx.x & 31) == 0 && width) atomicMax(max_bits, width);
}

// Pass 2: scatter B = bits per coefficient into the (zeroed, 8-byte padded) payload.
__global__ void compact_pack_kernel(const uint64_t *__restrict__ mag, const uint8_t *__restrict__ sign, size_t count, uint32_t words,
                                    uint32_t B, unsigned long long *__restrict__ payload) {
    const size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= count) return;
    const uint32_t mbits = B - 1;  // magnitude bits
    const size_t base = k * (size_t)B;
    const uint32_t mwords = (mbits + 63) / 64;
    for (uint32_t w = 0; w <= mwords; w++) {
        // chunk w covers coefficient bits [64 w, 64 w + 64) intersected with [0, B)
        const uint32_t lo_bit = w * 64;
        if (lo_bit >= B) break;
        uint64_t chunk = (w < words && w < mwords) ? mag[k * words + w] : 0;
        if (lo_bit + 64 > mbits && mbits >= lo_bit) {
            const uint32_t keep = mbits - lo_bit;  // magnitude bits in this chunk (< 64 here)
            chunk = keep ? (chunk & ((~0ull) >> (64 - keep))) : 0;
            if (keep < 64) chunk |= (uint64_t)(sign[k] & 1) << keep;  // the sign bit sits right above the magnitude
        }
        const size_t bit = base + lo_bit;
        const size_t word = bit >> 6;
        const uint32_t sh = (uint32_t)(bit & 63);
        if (chunk) {
            atomicOr(payload + word, (unsigned long long)(chunk << sh));
            if (sh) {
                const uint64_t spill = chunk >> (64 - sh);
                if (spill) atomicOr(payload + word + 1, (unsigned long long)spill);
            }
        }
    }
}

// Load: B bits per coefficient -> residues of +-magnitude in every active limb (coefficient form).
template <typename T>
__global__ void compact_unpack_kernel(T *__restrict__ data, CompactGeom g, const LimbConst<T> *__restrict__ lc,
                                      const uint8_t *__restrict__ payload, size_t payload_len, uint32_t B) {
    const size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= g.count) return;
    const size_t n = (size_t)1 << g.logn, poly = k >> g.logn, j = k & (n - 1);
    T *p = data + ((poly * g.limbs) << g.logn) + j;
    if (B == 0) {
        for (uint32_t i = 0; i < g.limbs; i++) p[(size_t)i << g.logn] = 0;
        return;
    }
    const uint32_t mbits = B - 1;
    const size_t base = k * (size_t)B;
    auto bits_at = [&](size_t bit, uint32_t cnt) -> uint64_t {  // cnt <= 64 bits starting at `bit`, zero past the end
        uint64_t out = 0;
        const size_t byte0 = bit >> 3;
        const uint32_t sh = (uint32_t)(bit & 7);
        // up to 9 bytes cover 64 bits at an arbitrary bit offset
        unsigned __int128 acc = 0;
        for (uint32_t b = 0; b < 9; b++) {
            const size_t at = byte0 + b;
            if (at < payload_len) acc |= (unsigned __int128)payload[at] << (8 * b);
        }
        out = (uint64_t)(acc >> sh);
        return cnt >= 64 ? out : (out & ((1ull << cnt) - 1));
    };
    const bool neg = bits_at(base + mbits, 1) != 0;
    const uint32_t mwords = (mbits + 63) / 64;
    for (uint32_t i = 0; i < g.limbs; i++) {
        const LimbConst<T> c = lc[i];
        const uint64_t q = (uint64_t)c.q;
        uint64_t r = 0;
        for (int w = (int)mwords - 1; w >= 0; w--) {
            const uint32_t cnt = (uint32_t)w * 64 + 64 <= mbits ? 64u : mbits - (uint32_t)w * 64;
            const uint64_t word = bits_at(base + (size_t)w * 64, cnt);
            // r = (r * 2^64 + word) mod q
            r = cmul<T>(r, (uint64_t)c.r64, c);
            const uint64_t lo = (uint64_t)red64<T>(word, c);
            r += lo;
            r = r >= q ? r - q : r;
        }
        if (neg && r) r = q - r;
        p[(size_t)i << g.logn] = (T)r;
    }
}

// Big-integer coefficients -> residues (the host loop `value % modulus` of
// GpuDCRTPolyMatrix::from_cpu_matrix, src/matrix/gpu_dcrt_poly.rs:793-803, and of GpuDCRTPoly::from_coeffs,
// src/poly/dcrt/gpu.rs:903-920, moved onto the device).  words[(poly * n + j) * wpc + w], little endian;
// BITREV: the input is EVAL data in natural slot order, the device keeps it bit-reversed.
template <typename T, bool BITREV>
__global__ void bigint_load_kernel(T *__restrict__ data, CompactGeom g, const LimbConst<T> *__restrict__ lc,
                                   const uint64_t *__restrict__ words, uint32_t wpc) {
    const size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= g.count) return;
    const size_t n = (size_t)1 << g.logn, poly = k >> g.logn;
    const uint32_t j = (uint32_t)(k & (n - 1));
    const uint32_t slot = BITREV ? (g.logn ? brev32(j) >> (32 - g.logn) : 0u) : j;
    T *p = data + ((poly * g.limbs) << g.logn) + slot;
    const uint64_t *src = words + k * wpc;
    for (uint32_t i = 0; i < g.limbs; i++) {
        const LimbConst<T> c = lc[i];
        const uint64_t q = (uint64_t)c.q;
        uint64_t r = 0;
        for (int w = (int)wpc - 1; w >= 0; w--) {  // r = (r * 2^64 + word) mod q
            r = cmul<T>(r, (uint64_t)c.r64, c);
            r += (uint64_t)red64<T>(src[w], c);
            r = r >= q ? r - q : r;
        }
        p[(size_t)i << g.logn] = (T)r;
    }
}

}  // namespace gp
