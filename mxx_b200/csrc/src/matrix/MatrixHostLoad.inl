// MatrixHostLoad.inl — chunked, two-lane host->device load for large matrices on u32 contexts.
//
// The reference ABI hands residues over as little-endian u64 words (src/matrix/gpu_dcrt_poly.rs:
// 769-817, 1916-1920) even when every modulus fits 30 bits, so gpu_matrix_load_rns_batch of the
// C1-R operand is PCIe-bound on words that are half zero bytes.  This pipeline cuts the bytes that
// cross PCIe: the matrix is split into row chunks that are consumed from both ends at once —
//   lane N: host worker threads narrow a chunk to u32 into a pinned ring slot (vectorised; reads
//           the caller's buffer whether pinned or pageable), then it is DMA'd at half size;
//   lane R: chunks are DMA'd as they are (u64) while the copy engine has room (pinned callers only);
// each chunk is finished by the pack kernel (narrow / bit-reverse into device order).  Whichever lane
// is free takes the next chunk, so the split adapts to the host's core count and to PCIe.
// No arithmetic happens on the host: the workers only drop the zero upper halves of the words.
// The reference stages every limb separately (MatrixSerde.cu:566-741).

#include <condition_variable>
#include <deque>
#include <functional>
#include <thread>

namespace {

__attribute__((target_clones("avx512f", "avx2", "default"))) void narrow_words_plain(const uint64_t *src, uint32_t *dst, size_t n) {
    for (size_t i = 0; i < n; i++) dst[i] = (uint32_t)src[i];
}
#if defined(__x86_64__)
}  // namespace
#include <immintrin.h>
namespace {
// AVX-512: 16 words per iteration, streaming (non-temporal) stores so the ring slot is written without a
// read-for-ownership — the host's memory bandwidth, shared with the DMA engine, is what bounds the load
__attribute__((target("avx512f"))) void narrow_words_avx512_nt(const uint64_t *src, uint32_t *dst, size_t n) {
    size_t i = 0;
    if (((uintptr_t)dst & 63) == 0) {
        for (; i + 16 <= n; i += 16) {
            const __m256i lo = _mm512_cvtepi64_epi32(_mm512_loadu_si512((const void *)(src + i)));
            const __m256i hi = _mm512_cvtepi64_epi32(_mm512_loadu_si512((const void *)(src + i + 8)));
            _mm512_stream_si512((__m512i *)(dst + i), _mm512_inserti64x4(_mm512_castsi256_si512(lo), hi, 1));
        }
        _mm_sfence();
    }
    for (; i < n; i++) dst[i] = (uint32_t)src[i];
}
#endif
void narrow_words(const uint64_t *src, uint32_t *dst, size_t n) {
#if defined(__x86_64__)
    static const bool nt = __builtin_cpu_supports("avx512f") && !(getenv("MXX_LOAD_NT") && atoi(getenv("MXX_LOAD_NT")) == 0);
    if (nt) return narrow_words_avx512_nt(src, dst, n);
#endif
    narrow_words_plain(src, dst, n);
}

// process-wide worker threads (started on first use, parked on a condition variable)
class HostPool {
  public:
    static HostPool &get() {
        static HostPool *pool = new HostPool();  // intentionally leaked: workers may outlive static destructors
        return *pool;
    }
    int size() const { return (int)workers_.size(); }
    void submit(std::function<void()> fn) {
        {
            std::lock_guard<std::mutex> g(m_);
            q_.push_back(std::move(fn));
        }
        cv_.notify_one();
    }

  private:
    HostPool() {
        int want = (int)std::thread::hardware_concurrency() - 1;
        if (const char *e = getenv("MXX_LOAD_THREADS")) want = atoi(e);
        want = std::max(0, std::min(want, 32));
        for (int i = 0; i < want; i++) workers_.emplace_back([this] { run(); });
        for (auto &t : workers_) t.detach();
    }
    void run() {
        for (;;) {
            std::function<void()> fn;
            {
                std::unique_lock<std::mutex> g(m_);
                cv_.wait(g, [this] { return !q_.empty(); });
                fn = std::move(q_.front());
                q_.pop_front();
            }
            fn();
        }
    }
    std::vector<std::thread> workers_;
    std::mutex m_;
    std::condition_variable cv_;
    std::deque<std::function<void()>> q_;
};

constexpr int kMaxSlots = 64;
constexpr size_t kHybridMinWords = (size_t)16 << 20;
inline size_t env_size(const char *name, size_t dflt, size_t lo, size_t hi) {
    const char *e = getenv(name);
    if (!e) return dflt;
    const long v = atol(e);
    return (size_t)std::min<long>((long)hi, std::max<long>((long)lo, v));
}

struct LoadPipe {
    int device = 0;
    std::mutex busy;  // one hybrid load at a time per context; others take the simple path
    cudaStream_t lane[4] = {nullptr, nullptr, nullptr, nullptr};  // R0 R1 N0 N1
    cudaEvent_t lane_done[4] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t start = nullptr;
    cudaEvent_t slot_ev[kMaxSlots] = {};
    size_t kChunkWords = 0;  // host words per chunk (MXX_LOAD_CHUNK_LOG, default 2^21 = 16 MB in, 8 MB narrowed)
    int kSlots = 0;          // pinned ring slots (MXX_LOAD_SLOTS, default 24); measured sweep: gpurun_out/load_lab*.log -> profiles/r01_load_pipeline.txt
    uint32_t *pinned = nullptr;    // kSlots * kChunkWords u32
    uint64_t *dev_raw[2] = {nullptr, nullptr};
    uint32_t *dev_narrow[2] = {nullptr, nullptr};
    ~LoadPipe() {
        cudaSetDevice(device);
        for (auto s : lane)
            if (s) { cudaStreamSynchronize(s); cudaStreamDestroy(s); }
        for (auto e : lane_done) if (e) cudaEventDestroy(e);
        for (auto e : slot_ev) if (e) cudaEventDestroy(e);
        if (start) cudaEventDestroy(start);
        if (pinned) cudaFreeHost(pinned);
        for (auto p : dev_raw) if (p) cudaFree(p);
        for (auto p : dev_narrow) if (p) cudaFree(p);
    }
};

LoadPipe *get_load_pipe(GpuContext *ctx) {
    std::lock_guard<std::mutex> g(ctx->lazy_mutex);
    if (ctx->load_pipe) return (LoadPipe *)ctx->load_pipe;
    auto *p = new (std::nothrow) LoadPipe();
    if (!p) return nullptr;
    p->device = ctx->device;
    p->kChunkWords = (size_t)1 << env_size("MXX_LOAD_CHUNK_LOG", 21, 16, 24);
    p->kSlots = (int)env_size("MXX_LOAD_SLOTS", 24, 2, kMaxSlots);
    const size_t kChunkWords = p->kChunkWords;
    const int kSlots = p->kSlots;
    bool ok = true;
    for (int i = 0; i < 4 && ok; i++) {
        ok = cudaStreamCreateWithFlags(&p->lane[i], cudaStreamNonBlocking) == cudaSuccess &&
             cudaEventCreateWithFlags(&p->lane_done[i], cudaEventDisableTiming) == cudaSuccess;
    }
    ok = ok && cudaEventCreateWithFlags(&p->start, cudaEventDisableTiming) == cudaSuccess;
    for (int i = 0; i < kSlots && ok; i++) ok = cudaEventCreateWithFlags(&p->slot_ev[i], cudaEventDisableTiming) == cudaSuccess;
    ok = ok && cudaHostAlloc((void **)&p->pinned, (size_t)kSlots * kChunkWords * 4, cudaHostAllocDefault) == cudaSuccess;
    for (int i = 0; i < 2 && ok; i++) {
        ok = cudaMalloc((void **)&p->dev_raw[i], kChunkWords * 8) == cudaSuccess &&
             cudaMalloc((void **)&p->dev_narrow[i], kChunkWords * 4) == cudaSuccess;
    }
    if (!ok) {
        cudaGetLastError();
        delete p;
        return nullptr;
    }
    ctx->load_pipe = p;
    ctx->load_pipe_free = [](void *q) { delete (LoadPipe *)q; };
    return p;
}

template <typename S>
int launch_pack_from(GpuContext *ctx, cudaStream_t stream, uint32_t *dst, const S *src, size_t words, bool bitrev, size_t word0,
                     uint32_t limbs) {
    const unsigned threads = 256, grid = stride_grid(ctx, words, threads);
    const auto *lc = (const gp::LimbConst<uint32_t> *)ctx->d_limb;
    if (bitrev) gp::pack_kernel<uint32_t, S, true><<<grid, threads, 0, stream>>>(dst, src, words, (uint32_t)ctx->logN, word0, limbs, lc);
    else gp::pack_kernel<uint32_t, S, false><<<grid, threads, 0, stream>>>(dst, src, words, (uint32_t)ctx->logN, word0, limbs, lc);
    gp::count_launch();
    GP_CHECK_LAUNCH("pack_kernel(chunk)");
    return 0;
}

// returns 0 done, 1 error, -1 not applicable (caller uses the single-DMA path)
int load_hybrid(GpuContext *ctx, GpuMatrix *mat, const uint64_t *host, size_t words, bool bitrev) {
    if (ctx->wide || words < kHybridMinWords || gp::tl_stream) return -1;
    HostPool &pool = HostPool::get();
    const int T = pool.size();
    if (T <= 0) return -1;
    LoadPipe *pipe = get_load_pipe(ctx);
    if (!pipe) return -1;
    std::unique_lock<std::mutex> only(pipe->busy, std::try_to_lock);
    if (!only.owns_lock()) return -1;

    const size_t kChunkWords = pipe->kChunkWords;
    const int kSlots = pipe->kSlots;
    const size_t n = (size_t)ctx->N;
    const size_t rows_per_chunk = std::max<size_t>(1, kChunkWords / n);
    if (rows_per_chunk * n > kChunkWords) return -1;  // a single row larger than a chunk: n > 2^21
    const size_t chunk_words = rows_per_chunk * n;
    const size_t chunks = (words + chunk_words - 1) / chunk_words;
    cudaPointerAttributes attr{};
    const bool pinned_src = cudaPointerGetAttributes(&attr, host) == cudaSuccess && attr.type == cudaMemoryTypeHost;
    cudaGetLastError();

    // shared work interval [lo, hi): lane R takes from the front, the workers from the back
    struct Shared {
        std::mutex m;
        std::condition_variable cv;
        size_t lo = 0, hi = 0;
        std::deque<int> free_slots;
        std::deque<std::pair<int, size_t>> ready;  // (slot, chunk)
        int workers_live = 0;
    } sh;
    sh.hi = chunks;
    for (int i = 0; i < kSlots; i++) sh.free_slots.push_back(i);
    sh.workers_live = T;
    uint32_t *ring = pipe->pinned;
    auto chunk_len = [&](size_t c) { return std::min(chunk_words, words - c * chunk_words); };
    for (int w = 0; w < T; w++) {
        pool.submit([&sh, ring, host, chunk_words, kChunkWords, &chunk_len] {
            for (;;) {
                int slot;
                size_t c;
                {
                    std::unique_lock<std::mutex> g(sh.m);
                    sh.cv.wait(g, [&] { return sh.lo >= sh.hi || !sh.free_slots.empty(); });
                    if (sh.lo >= sh.hi) break;
                    slot = sh.free_slots.front();
                    sh.free_slots.pop_front();
                    c = --sh.hi;
                }
                narrow_words(host + c * chunk_words, ring + (size_t)slot * kChunkWords, chunk_len(c));
                {
                    std::lock_guard<std::mutex> g(sh.m);
                    sh.ready.emplace_back(slot, c);
                }
                sh.cv.notify_all();
            }
            {
                // notify UNDER the lock: the waiter (drain) owns `sh` on its stack and destroys it as soon as it sees
                // workers_live == 0, so nothing of `sh` may be touched after the mutex is released
                std::lock_guard<std::mutex> g(sh.m);
                sh.workers_live--;
                sh.cv.notify_all();
            }
        });
    }
    // from here on every exit path must wait for the workers (they reference this frame)
    auto drain = [&] {
        std::unique_lock<std::mutex> g(sh.m);
        sh.lo = sh.hi;  // nothing more to claim
        sh.cv.notify_all();
        sh.cv.wait(g, [&] { return sh.workers_live == 0; });
    };

    int status = 0;
    cudaError_t err = cudaEventRecord(pipe->start, ctx->stream);
    for (int i = 0; i < 4 && err == cudaSuccess; i++) err = cudaStreamWaitEvent(pipe->lane[i], pipe->start, 0);
    if (err != cudaSuccess) {
        drain();
        return gp::set_cuda_error(err, "load pipeline start");
    }
    uint32_t *dst = (uint32_t *)mat->data;
    std::deque<std::pair<cudaEvent_t, int>> slot_inflight;  // narrow copies whose ring slot is still being read
    cudaEvent_t raw_ev[2] = {pipe->lane_done[0], pipe->lane_done[1]};
    bool raw_busy[2] = {false, false};
    size_t enq_raw = 0, enq_narrow = 0, finished = 0;
    while (finished < chunks && status == 0) {
        bool progressed = false;
        // 1. ring slots whose DMA has completed go back to the workers
        while (!slot_inflight.empty()) {
            const cudaError_t q = cudaEventQuery(slot_inflight.front().first);
            if (q == cudaErrorNotReady) break;
            if (q != cudaSuccess) {  // a sticky fault: nothing will ever complete, do not spin on it
                status = gp::set_cuda_error(q, "load pipeline (ring slot)");
                break;
            }
            {
                std::lock_guard<std::mutex> g(sh.m);
                sh.free_slots.push_back(slot_inflight.front().second);
            }
            sh.cv.notify_all();
            slot_inflight.pop_front();
            progressed = true;
        }
        // 2. narrowed chunks -> lane N
        std::pair<int, size_t> item;
        bool have = false;
        {
            std::lock_guard<std::mutex> g(sh.m);
            if (!sh.ready.empty()) {
                item = sh.ready.front();
                sh.ready.pop_front();
                have = true;
            }
        }
        if (have) {
            const int l = (int)(enq_narrow & 1);
            cudaStream_t s = pipe->lane[2 + l];
            const size_t len = chunk_len(item.second);
            err = cudaMemcpyAsync(pipe->dev_narrow[l], ring + (size_t)item.first * kChunkWords, len * 4, cudaMemcpyHostToDevice, s);
            if (err == cudaSuccess) err = cudaEventRecord(pipe->slot_ev[item.first], s);
            if (err != cudaSuccess) { status = gp::set_cuda_error(err, "narrow-lane DMA"); break; }
            slot_inflight.emplace_back(pipe->slot_ev[item.first], item.first);
            if (launch_pack_from<uint32_t>(ctx, s, dst + item.second * chunk_words, pipe->dev_narrow[l], len, bitrev, item.second * chunk_words, (uint32_t)mat->limbs())) { status = 1; break; }
            enq_narrow++;
            finished++;
            progressed = true;
        }
        // 3. lane R takes a front chunk when one of its two buffers is idle
        if (pinned_src) {
            for (int l = 0; l < 2; l++) {
                if (!raw_busy[l]) continue;
                const cudaError_t q = cudaEventQuery(raw_ev[l]);
                if (q == cudaSuccess) raw_busy[l] = false;
                else if (q != cudaErrorNotReady && status == 0) status = gp::set_cuda_error(q, "load pipeline (raw lane)");
            }
            const int l = (int)(enq_raw & 1);
            if (!raw_busy[l]) {
                size_t c = 0;
                bool got = false;
                {
                    std::lock_guard<std::mutex> g(sh.m);
                    if (sh.lo < sh.hi) {
                        c = sh.lo++;
                        got = true;
                    }
                }
                if (got) {
                    cudaStream_t s = pipe->lane[l];
                    const size_t len = chunk_len(c);
                    err = cudaMemcpyAsync(pipe->dev_raw[l], host + c * chunk_words, len * 8, cudaMemcpyHostToDevice, s);
                    if (err != cudaSuccess) { status = gp::set_cuda_error(err, "raw-lane DMA"); break; }
                    if (launch_pack_from<uint64_t>(ctx, s, dst + c * chunk_words, pipe->dev_raw[l], len, bitrev, c * chunk_words, (uint32_t)mat->limbs())) { status = 1; break; }
                    err = cudaEventRecord(raw_ev[l], s);
                    if (err != cudaSuccess) { status = gp::set_cuda_error(err, "raw-lane event"); break; }
                    raw_busy[l] = true;
                    enq_raw++;
                    finished++;
                    progressed = true;
                }
            }
        }
        if (!progressed) {
            std::unique_lock<std::mutex> g(sh.m);
            sh.cv.wait_for(g, std::chrono::microseconds(50), [&] { return !sh.ready.empty(); });
        }
    }
    drain();
    // the ring slots must be idle before the next load reuses them; the lanes join the context stream
    for (int i = 0; i < 4; i++) {
        cudaError_t e = cudaEventRecord(pipe->lane_done[i], pipe->lane[i]);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(ctx->stream, pipe->lane_done[i], 0);
        if (e != cudaSuccess && status == 0) status = gp::set_cuda_error(e, "load pipeline join");
    }
    for (auto &p : slot_inflight) cudaEventSynchronize(p.first);
    return status;
}

}  // namespace
