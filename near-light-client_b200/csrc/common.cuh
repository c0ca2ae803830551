// common.cuh -- context, error handling and launch bookkeeping shared by the nlc_commit translation units.
#pragma once
#include <cuda_runtime.h>
#include <nvtx3/nvToolsExt.h> // header-only; ranges are no-ops unless a profiler (nsys) is attached

#include <atomic>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <exception>
#include <map>
#include <mutex>
#include <stdexcept>
#include <string>
#include <tuple>
#include <vector>

#include "../../include/nlc_commit.h"
#include "staging.h"

typedef uint64_t u64;
typedef uint32_t u32;

struct NlcError : std::runtime_error {
    using std::runtime_error::runtime_error;
};

#define NLC_CUDA(expr)                                                                                       \
    do {                                                                                                     \
        cudaError_t _e = (expr);                                                                             \
        if (_e != cudaSuccess)                                                                               \
            throw NlcError(std::string(#expr) + " failed: " + cudaGetErrorString(_e) + " at " + __FILE__ +   \
                           ":" + std::to_string(__LINE__));                                                  \
    } while (0)

#define NLC_CHECK(cond, msg)                                                              \
    do {                                                                                  \
        if (!(cond)) throw NlcError(std::string("nlc_commit: ") + (msg) + " [" #cond "]"); \
    } while (0)

// Table cache key: (kind, log size, digit bits, inverse?, shift/scale value)
typedef std::tuple<int, int, int, int, u64, u64> TableKey;

struct nlc_ctx {
    int device = 0;
    cudaStream_t own_stream = nullptr;
    cudaStream_t stream = nullptr;
    // host-fed commits upload column chunks on copy_stream while the compute stream transforms the previous chunk
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t ev_ready = nullptr;
    std::vector<cudaEvent_t> chunk_events;
    cudaEvent_t chunk_event(size_t i) {
        while (chunk_events.size() <= i) {
            cudaEvent_t e;
            NLC_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
            chunk_events.push_back(e);
        }
        return chunk_events[i];
    }
    std::recursive_mutex mu;
    std::atomic<u64> launches{0};
    std::map<TableKey, u64*> tables; // device-resident twiddle / scale tables, built once
    size_t table_bytes = 0;
    std::vector<TableKey> pw_lru; // caller-keyed power tables, least recently used first (ntt.cu:pw_table)
    size_t pw_bytes = 0;
    // small pinned staging area + device scratch for transcript-sized transfers
    HostStager stager; // pageable host columns -> HBM through pinned slots (staging.h)
    // a commit that also materialises struct fields on the host (nlc_commit_from_values_materialised): device -> host
    // traffic runs on its own stream, through its own stager, driven by a worker thread while the caller's thread
    // keeps feeding the compute stream
    HostStager dl_stager;
    Downloader downloader;
    cudaStream_t dl_stream = nullptr;
    std::vector<cudaEvent_t> dl_events;
    cudaEvent_t dl_event(size_t i) {
        while (dl_events.size() <= i) {
            cudaEvent_t e;
            NLC_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
            dl_events.push_back(e);
        }
        return dl_events[i];
    }
    u64* pinned = nullptr;
    size_t pinned_words = 0;
    int sm_count = 148;
    bool ntt_attrs = false; // cudaFuncSetAttribute is per device: remembered per context
    // TimingTree-style phase timing (plonky2's "IFFT" / "FFT + blinding" / "build Merkle tree" scopes),
    // measured with CUDA events on the launching stream when enabled
    bool timing = false;
    std::vector<std::pair<int, std::pair<cudaEvent_t, cudaEvent_t>>> timing_events;
    // plonky2's TimingTree scopes ("IFFT", "FFT + blinding", "build Merkle tree") as NVTX ranges (SURVEY.md section 5:
    // tracing), and -- when nlc_timing_enable is on -- as CUDA-event pairs on the launching stream
    bool nvtx_open = false;
    void phase_begin(int phase) {
        static const char* const names[] = {"nlc: IFFT", "nlc: FFT + blinding (coset LDE)", "nlc: leaf hashing",
                                            "nlc: Merkle digest levels", "nlc: leaf hashing (column chunk)"};
        nvtxRangePushA(phase >= 0 && phase < 5 ? names[phase] : "nlc: phase");
        nvtx_open = true;
        if (!timing) return;
        cudaEvent_t a, b;
        cudaEventCreate(&a);
        cudaEventCreate(&b);
        cudaEventRecord(a, stream);
        timing_events.push_back({phase, {a, b}});
    }
    void phase_end() {
        if (nvtx_open) {
            nvtxRangePop();
            nvtx_open = false;
        }
        if (!timing || timing_events.empty()) return;
        cudaEventRecord(timing_events.back().second.second, stream);
    }

    // Device buffers come from the stream-ordered pool (cudaMallocAsync) THROUGH a size-keyed cache of this context:
    // a proof commits a handful of shapes over and over (wires 135, Z 20, quotient 16 columns, FRI rounds), and when
    // their buffers are released in one order and requested in another the driver pool has to remap physical pages
    // into new contiguous ranges -- measured: a 2^16 x 124 trace commit 4.5 ms when its blocks come back as they were,
    // 35-140 ms when the pool has to grow or remap.  A freed block therefore waits here, keyed by its (rounded) size,
    // and the next request of exactly that size takes it back without a driver call.  Safe because every reuse is ordered on
    // this context's stream: the other streams of a call (copy, download) are joined before the call returns, and
    // nlc_set_stream synchronises the old stream before switching.  Budget NLC_CACHE_MB (default 16384): blocks beyond
    // it go back to the driver; everything goes back when an allocation fails, and at nlc_destroy.
    std::multimap<size_t, void*> cache_free; // rounded bytes -> block
    std::map<void*, size_t> cache_live;      // block -> rounded bytes
    size_t cache_bytes = 0, cache_budget = (size_t)16384 << 20;
    u64 cache_hits = 0, cache_misses = 0, cache_miss_us = 0; // nlc_cache_stats
    static size_t cache_round(size_t bytes) {
        const size_t g = bytes < ((size_t)1 << 20) ? 512 : ((size_t)2 << 20);
        return (bytes + g - 1) / g * g;
    }
    void cache_release_all() {
        for (auto& kv : cache_free) cudaFreeAsync(kv.second, stream);
        cache_free.clear();
        cache_bytes = 0;
    }
    u64* dmalloc(size_t words) {
        if (words == 0) words = 1;
        size_t bytes = cache_round(words * sizeof(u64));
        void* p = nullptr;
        auto it = cache_free.find(bytes); // exact (rounded) size only: a proof's multiset of shapes is then served from the
        if (it != cache_free.end()) {     // second proof on, whatever the order of requests and releases
            p = it->second;
            bytes = it->first;
            cache_bytes -= bytes;
            cache_free.erase(it);
            cache_hits++;
        } else {
            cache_misses++;
            const auto t0 = std::chrono::steady_clock::now();
            struct Timer {
                u64& acc;
                std::chrono::steady_clock::time_point t0;
                ~Timer() { acc += (u64)std::chrono::duration_cast<std::chrono::microseconds>(std::chrono::steady_clock::now() - t0).count(); }
            } timer{cache_miss_us, t0};
            cudaError_t e = cudaMallocAsync(&p, bytes, stream);
            if (e == cudaErrorMemoryAllocation) { // give the cached blocks back and try once more
                cudaGetLastError();
                cache_release_all();
                cudaStreamSynchronize(stream);
                e = cudaMallocAsync(&p, bytes, stream);
            }
            NLC_CUDA(e);
        }
        cache_live[p] = bytes;
        return (u64*)p;
    }
    void dfree(void* p) {
        if (!p) return;
        auto it = cache_live.find(p);
        if (it == cache_live.end()) {
            cudaFreeAsync(p, stream);
            return;
        }
        const size_t bytes = it->second;
        cache_live.erase(it);
        if (bytes > cache_budget) {
            cudaFreeAsync(p, stream);
            return;
        }
        cache_free.emplace(bytes, p);
        cache_bytes += bytes;
        while (cache_bytes > cache_budget && !cache_free.empty()) { // over budget: the largest blocks go first
            auto last = std::prev(cache_free.end());
            cudaFreeAsync(last->second, stream);
            cache_bytes -= last->first;
            cache_free.erase(last);
        }
    }
    u64* pin(size_t words) {
        if (words > pinned_words) {
            if (pinned) cudaFreeHost(pinned);
            pinned = nullptr;
            size_t w = words < (1u << 16) ? (1u << 16) : words;
            NLC_CUDA(cudaMallocHost((void**)&pinned, w * sizeof(u64)));
            pinned_words = w;
        }
        return pinned;
    }
};

// RAII guard: selects the context's device and serialises calls on the context.
struct CtxGuard {
    nlc_ctx* c;
    int prev = -1;
    int unwinding;
    explicit CtxGuard(nlc_ctx* ctx) : c(ctx), unwinding(std::uncaught_exceptions()) {
        c->mu.lock();
        cudaGetDevice(&prev);
        if (prev != c->device) cudaSetDevice(c->device);
    }
    ~CtxGuard() {
        // a call that fails drops its half-filled staging slot while it still holds the context: once the lock is
        // gone another thread's upload may own the stager
        if (std::uncaught_exceptions() > unwinding) c->stager.abandon();
        if (prev >= 0 && prev != c->device) cudaSetDevice(prev);
        c->mu.unlock();
    }
};

#define NLC_LAUNCH(ctx, kernel, grid, block, smem, ...)                         \
    do {                                                                        \
        kernel<<<(grid), (block), (smem), (ctx)->stream>>>(__VA_ARGS__);        \
        (ctx)->launches.fetch_add(1, std::memory_order_relaxed);                \
        NLC_CUDA(cudaGetLastError());                                           \
    } while (0)

inline std::string& nlc_thread_error() {
    static thread_local std::string e;
    return e;
}
// Every extern "C" entry point wraps its body with this: exceptions become error codes.
template <class F>
static inline int nlc_guarded(nlc_ctx* ctx, F&& f) {
    try {
        f();
        return 0;
    } catch (const std::exception& e) {
        // per calling thread: rayon workers share a context, each reads back its own failure (nlc_last_error)
        (void)ctx;
        nlc_thread_error() = e.what();
        cudaGetLastError(); // clear a non-sticky error so later calls can proceed
        return 1;
    }
}

static inline int ilog2(size_t n) {
    int l = 0;
    while (((size_t)1 << l) < n) l++;
    return l;
}

// ---- ntt.cu ----
// forward/inverse NTT of `n_cols` columns of 2^log_n elements (column j at src + j*src_stride).
// ROWS: out is a row-major (rows x row_width) matrix; row = row_base + bitrev order position, column
//       offset col0.  COLS: the same matrix stored column-major (column stride dst_stride).
//       NATURAL: out is column-contiguous, natural order.
// COLS output re-blocked by leaf-row shard: leaf k of column c goes to
//   base(k >> log_per) + c * dst_stride + (k & (2^log_per - 1)),  base(g) = peer[g] if peers are given (the receive
// buffers of the other GPUs, written over NVLink) else dst + g * shard_stride (the send buffer of an all-to-all).
struct NttScatter {
    int log_per = -1; // < 0: off
    size_t shard_stride = 0;
    u64* peer[16] = {};
    // NATURAL output only: store every result into n_bcast buffers (peer[0..n_bcast), same offsets) instead of dst --
    // the all-gather of a column-sharded iFFT done by the transform's own stores over NVLink
    int n_bcast = 0;
};
// Several forward transforms of the SAME source columns in one launch per pass (the 2^r cosets of an LDE; fast path,
// COLS output): transform z multiplies its input by pre + z * z_pre (one power table per coset, back to back), and lands
// at leaf rows row_base + z * z_dst.  The passes before the last one run in place on the destination columns, so no
// scratch is needed and a launch carries 2^log_z times the CTAs (narrow batches fill the GPU; launch ramps amortise).
struct NttCosets {
    int log_z = 0;
    size_t z_src = 0, z_dst = 0, z_pre = 0;
};
struct NttJob {
    const u64* src = nullptr;
    size_t src_stride = 0;
    size_t n_cols = 0;
    int log_n = 0;
    bool inverse = false;      // use inverse roots and scale by 1/n
    u64 pre_shift = 1;         // multiply input j by pre_shift^j (coset_fft)
    u64 post_shift = 1;        // NATURAL only: multiply output k by post_shift^k (coset_ifft)
    bool rows = false;
    bool cols = false;         // COLS: column-major out, index-bit-reversed order, column stride dst_stride
    u64* dst = nullptr;
    size_t dst_stride = 0;     // NATURAL / COLS: column stride
    size_t row_width = 0;      // ROWS: elements per row
    size_t row_base = 0;       // ROWS / COLS: first row (bit-reversed position)
    size_t col0 = 0;           // ROWS / COLS: first column
    u64* scratch = nullptr;    // n_cols * 2^log_n words, needed when log_n > one digit
    NttScatter scatter;        // COLS only
    // COLS, forward, no scatter: 2^log_cosets transforms at once; pre_tables = their power tables back to back (device,
    // 2^log_n words each, replaces pre_shift), coset z at leaf rows row_base + z * 2^log_n
    int log_cosets = 0;
    const u64* pre_tables = nullptr;
};
void ntt_run(nlc_ctx* ctx, const NttJob& job);
const u64* ntt_coset_tables(nlc_ctx* ctx, int log_n, int rate_bits, size_t kr0, size_t count);
bool ntt_can_batch_cosets(int log_n);
size_t ntt_scratch_words(size_t n_cols, int log_n);

// ---- merkle.cu ----
// Build MerkleTree::new over device-resident leaves; digests in plonky2's interleaved layout.
// elem_stride == 0: row-major leaves (leaf k at leaves + k*leaf_len); elem_stride > 0: column-major
// (element j of leaf k at leaves[j*elem_stride + k]), the layout of a PolynomialBatch's LDE.
void merkle_build(nlc_ctx* ctx, const u64* leaves, size_t n_leaves, size_t leaf_len, size_t elem_stride, int cap_height,
                  u64* digests, u64* cap);
void merkle_levels(nlc_ctx* ctx, size_t n_leaves, int cap_height, u64* digests, u64* cap);
void merkle_leaf_absorb(nlc_ctx* ctx, const u64* lde, size_t n_leaves, size_t elem_stride, size_t col0, size_t n_cols,
                        bool first, bool last, u64* state, int cap_height, u64* digests, u64* cap);
// device copy of leaf + siblings for many leaf indices (indices in device memory)
void merkle_open_many(nlc_ctx* ctx, const u64* leaves, const u64* digests, size_t n_leaves, size_t leaf_len,
                      size_t elem_stride, int cap_height, const u64* idx_dev, int shift_right, size_t n_idx, u64* out,
                      size_t out_stride, size_t out_offset);
void poseidon_permute_many(nlc_ctx* ctx, u64* states_dev, size_t count);
void hash_many(nlc_ctx* ctx, const u64* in_dev, size_t count, size_t len, bool or_noop, u64* out_dev);
void two_to_one_many(nlc_ctx* ctx, const u64* l_dev, const u64* r_dev, size_t count, u64* out_dev);
