// staging.h -- pageable host memory <-> HBM at PCIe speed (host side only, no kernels).
//
// The plonky2 caller hands nlc_commit_from_values a Vec<PolynomialValues<F>>: W separate, PAGEABLE heap vectors
// (vendor/starkyx/starkyx/src/plonky2/stark/prover.rs:91-93).  cudaMemcpyAsync from pageable memory goes through
// the driver's single-threaded bounce buffer (measured here: 8.4 GB/s, 15.9 ms for the 71 MB of config 2 against
// 11.6 ms from pinned memory).  The stager copies the columns into a ring of 4 MiB pinned slots with a small pool of
// host threads and issues one asynchronous H2D per slot, so the CPU fills slot k+1 while the DMA engine drains
// slot k.  Pinned / registered sources skip it.
#pragma once
#include <cuda_runtime.h>
#if defined(__x86_64__) && defined(__SSE2__)
#include <emmintrin.h>
#endif
#include <cstdint>

#include <condition_variable>
#include <deque>
#include <exception>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <mutex>
#include <stdexcept>
#include <thread>
#include <vector>

struct CopyPool {
    std::vector<std::thread> workers;
    std::mutex mu;
    std::condition_variable cv_work, cv_done;
    const std::function<void(size_t)>* fn = nullptr;
    size_t n_tasks = 0, next = 0, done = 0;
    unsigned long long generation = 0;
    bool stop = false;

    explicit CopyPool(int threads) {
        for (int t = 0; t < threads; t++) workers.emplace_back([this] { run(); });
    }
    ~CopyPool() {
        {
            std::lock_guard<std::mutex> l(mu);
            stop = true;
        }
        cv_work.notify_all();
        for (auto& w : workers) w.join();
    }
    void run() {
        std::unique_lock<std::mutex> l(mu);
        for (;;) {
            cv_work.wait(l, [&] { return stop || (fn && next < n_tasks); });
            if (stop) return;
            while (fn && next < n_tasks) {
                const size_t i = next++;
                const auto* f = fn;
                l.unlock();
                (*f)(i);
                l.lock();
                if (++done == n_tasks) cv_done.notify_all();
            }
        }
    }
    // runs f(0..n-1) on the workers and the calling thread; returns when all are done
    void parallel_for(size_t n, const std::function<void(size_t)>& f) {
        if (n == 0) return;
        std::unique_lock<std::mutex> l(mu);
        fn = &f;
        n_tasks = n;
        next = 0;
        done = 0;
        cv_work.notify_all();
        while (next < n_tasks) {
            const size_t i = next++;
            l.unlock();
            f(i);
            l.lock();
            ++done;
        }
        cv_done.wait(l, [&] { return done == n_tasks; });
        fn = nullptr;
    }
};

struct HostStager {
    static constexpr size_t SLOT_BYTES = (size_t)4 << 20;
    static constexpr size_t GRANULE = (size_t)512 << 10;
    static constexpr int N_SLOTS = 6;
    void* slot[N_SLOTS] = {};
    cudaEvent_t ev[N_SLOTS] = {};
    bool busy[N_SLOTS] = {};
    int cur = 0;
    size_t fill = 0; // bytes already placed in slot[cur]
    char* dst_base = nullptr;
    struct Granule {
        char* dst;
        const char* src;
        size_t len;
        size_t rows = 1, src_pitch = 0; // rows > 1: `rows` pieces of `len` bytes, src_pitch apart (0: packed) ...
        size_t dst_pitch = 0;           // ... written dst_pitch apart (0: packed)
    };
    static void run_granule(const Granule& g) {
        const size_t sp = g.src_pitch ? g.src_pitch : g.len, dp = g.dst_pitch ? g.dst_pitch : g.len;
        for (size_t r = 0; r < g.rows; r++) memcpy(g.dst + r * dp, g.src + r * sp, g.len);
    }
    // The same for device -> pageable reads, whose destination is a large buffer of the caller that nobody reads back
    // soon: non-temporal stores write whole lines without first reading them (no read-for-ownership: two thirds of a
    // cached copy's memory traffic) and leave the caches to the ring.  Granules are 512 KiB, below the size at which
    // the C library's memcpy switches to such stores by itself.  NLC_NT_STORES=0 turns it off.
    static void stream_copy(char* d, const char* s, size_t n) {
#if defined(__x86_64__) && defined(__SSE2__)
        static const bool on = [] { const char* e = getenv("NLC_NT_STORES"); return !(e && atoi(e) == 0); }();
        if (on && n >= 4096) {
            const size_t head = (16 - ((uintptr_t)d & 15)) & 15;
            memcpy(d, s, head);
            d += head, s += head, n -= head;
            const size_t lines = n / 64;
            for (size_t i = 0; i < lines; i++, d += 64, s += 64) {
                const __m128i a = _mm_loadu_si128((const __m128i*)s), b = _mm_loadu_si128((const __m128i*)(s + 16)),
                              c = _mm_loadu_si128((const __m128i*)(s + 32)), e = _mm_loadu_si128((const __m128i*)(s + 48));
                _mm_stream_si128((__m128i*)d, a);
                _mm_stream_si128((__m128i*)(d + 16), b);
                _mm_stream_si128((__m128i*)(d + 32), c);
                _mm_stream_si128((__m128i*)(d + 48), e);
            }
            n -= lines * 64;
            _mm_sfence();
        }
#endif
        memcpy(d, s, n);
    }
    static void run_granule_out(const Granule& g) {
        const size_t sp = g.src_pitch ? g.src_pitch : g.len, dp = g.dst_pitch ? g.dst_pitch : g.len;
        for (size_t r = 0; r < g.rows; r++) stream_copy(g.dst + r * dp, g.src + r * sp, g.len);
    }
    std::vector<Granule> pending;
    CopyPool* pool = nullptr;

    static bool pageable(const void* p) {
        cudaPointerAttributes a;
        if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
            cudaGetLastError();
            return true;
        }
        return a.type == cudaMemoryTypeUnregistered;
    }
    void init() {
        if (pool) return;
        int threads = 8;
        if (const char* e = getenv("NLC_COPY_THREADS")) threads = atoi(e);
        const int hw = (int)std::thread::hardware_concurrency();
        if (hw > 0 && threads > hw) threads = hw;
        if (threads < 1) threads = 1;
        pool = new CopyPool(threads - 1);
        for (int i = 0; i < N_SLOTS; i++) {
            if (cudaHostAlloc(&slot[i], SLOT_BYTES, cudaHostAllocDefault) != cudaSuccess)
                throw std::runtime_error("HostStager: cudaHostAlloc failed");
            cudaEventCreateWithFlags(&ev[i], cudaEventDisableTiming);
        }
    }
    void destroy() {
        if (!pool) return;
        delete pool;
        pool = nullptr;
        for (int i = 0; i < N_SLOTS; i++) {
            cudaFreeHost(slot[i]);
            cudaEventDestroy(ev[i]);
        }
    }
    // drop anything not yet flushed (error paths: the destination may be gone)
    void abandon() {
        pending.clear();
        fill = 0;
        dst_base = nullptr;
    }
    // flush slot[cur]: memcpy the pending granules in parallel, one async H2D, move to the next slot
    void flush(cudaStream_t stream) {
        if (fill == 0) return;
        pool->parallel_for(pending.size(), [&](size_t i) { run_granule(pending[i]); });
        pending.clear();
        if (cudaMemcpyAsync(dst_base, slot[cur], fill, cudaMemcpyHostToDevice, stream) != cudaSuccess)
            throw std::runtime_error("HostStager: cudaMemcpyAsync failed");
        cudaEventRecord(ev[cur], stream);
        busy[cur] = true;
        cur = (cur + 1) % N_SLOTS;
        fill = 0;
        dst_base = nullptr;
    }
    // append `bytes` from pageable `src`, destined for device address `dst`; consecutive calls with adjacent
    // destinations share a slot (and a DMA), anything else starts a new one
    void upload(cudaStream_t stream, void* dst, const void* src, size_t bytes) {
        init();
        char* d = (char*)dst;
        const char* s = (const char*)src;
        while (bytes) {
            if (fill && (dst_base + fill != d || fill == SLOT_BYTES)) flush(stream);
            if (fill == 0) {
                if (busy[cur]) {
                    cudaEventSynchronize(ev[cur]);
                    busy[cur] = false;
                }
                dst_base = d;
            }
            size_t take = SLOT_BYTES - fill < bytes ? SLOT_BYTES - fill : bytes;
            for (size_t o = 0; o < take; o += GRANULE) {
                const size_t len = take - o < GRANULE ? take - o : GRANULE;
                pending.push_back({(char*)slot[cur] + fill + o, s + o, len});
            }
            fill += take;
            d += take;
            s += take;
            bytes -= take;
        }
    }

    // `n_rows` strips of `row_bytes` bytes, `src_pitch` apart in pageable memory (a column range of a row-major
    // trace matrix), packed back to back at device address `dst`
    void upload_rows(cudaStream_t stream, void* dst, const void* src, size_t src_pitch, size_t row_bytes, size_t n_rows) {
        init();
        char* d = (char*)dst;
        const char* s = (const char*)src;
        if (row_bytes == 0 || row_bytes > SLOT_BYTES) { // degenerate shapes: row by row
            for (size_t r = 0; r < n_rows; r++) upload(stream, d + r * row_bytes, s + r * src_pitch, row_bytes);
            return;
        }
        const size_t rows_per_granule = GRANULE / row_bytes ? GRANULE / row_bytes : 1;
        while (n_rows) {
            if (fill && (dst_base + fill != d || SLOT_BYTES - fill < row_bytes)) flush(stream);
            if (fill == 0) {
                if (busy[cur]) {
                    cudaEventSynchronize(ev[cur]);
                    busy[cur] = false;
                }
                dst_base = d;
            }
            size_t take = (SLOT_BYTES - fill) / row_bytes;
            if (take > n_rows) take = n_rows;
            for (size_t o = 0; o < take; o += rows_per_granule) {
                const size_t k = take - o < rows_per_granule ? take - o : rows_per_granule;
                pending.push_back({(char*)slot[cur] + fill + o * row_bytes, s + o * src_pitch, row_bytes, k, src_pitch});
            }
            fill += take * row_bytes;
            d += take * row_bytes;
            s += take * src_pitch;
            n_rows -= take;
        }
    }

    // device -> pageable host, strided destination: n_rows packed rows of row_bytes at device address `src` go to
    // dst + r * dst_pitch (a column range of a row-major host matrix).  Synchronous, pipelined like download().
    void download_rows(cudaStream_t stream, const void* src, void* dst, size_t dst_pitch, size_t row_bytes, size_t n_rows) {
        init();
        flush(stream);
        if (!row_bytes || !n_rows) return;
        const size_t rows_per_slot = SLOT_BYTES / row_bytes;
        if (rows_per_slot == 0) throw std::runtime_error("HostStager: row wider than a staging slot");
        const size_t rows_per_granule = GRANULE / row_bytes ? GRANULE / row_bytes : 1;
        const char* s = (const char*)src;
        size_t issued = 0, drained = 0;
        int head = cur, tail = cur;
        size_t rows_of[N_SLOTS] = {}, first_of[N_SLOTS] = {};
        while (drained < n_rows) {
            while (issued < n_rows && !(issued > drained && head == tail)) {
                if (busy[head]) {
                    cudaEventSynchronize(ev[head]);
                    busy[head] = false;
                }
                const size_t take = n_rows - issued < rows_per_slot ? n_rows - issued : rows_per_slot;
                if (cudaMemcpyAsync(slot[head], s + issued * row_bytes, take * row_bytes, cudaMemcpyDeviceToHost, stream) != cudaSuccess)
                    throw std::runtime_error("HostStager: cudaMemcpyAsync (D2H) failed");
                cudaEventRecord(ev[head], stream);
                rows_of[head] = take;
                first_of[head] = issued;
                issued += take;
                head = (head + 1) % N_SLOTS;
            }
            cudaEventSynchronize(ev[tail]);
            pending.clear();
            for (size_t o = 0; o < rows_of[tail]; o += rows_per_granule) {
                const size_t k = rows_of[tail] - o < rows_per_granule ? rows_of[tail] - o : rows_per_granule;
                pending.push_back({(char*)dst + (first_of[tail] + o) * dst_pitch, (const char*)slot[tail] + o * row_bytes,
                                   row_bytes, k, 0, dst_pitch});
            }
            pool->parallel_for(pending.size(), [&](size_t i) { run_granule_out(pending[i]); });
            pending.clear();
            drained += rows_of[tail];
            tail = (tail + 1) % N_SLOTS;
        }
        cur = head;
    }

    // device -> pageable host: `segs` lists the host destinations of consecutive device bytes starting at `src`
    // (one segment = one Vec of the caller).  DMA into slot k+1 overlaps the host threads scattering slot k.
    // Synchronous: the data is in place on return.
    struct Seg {
        void* dst;
        size_t len;
    };
    void download(cudaStream_t stream, const void* src, const std::vector<Seg>& segs) {
        init();
        flush(stream);
        size_t total = 0;
        for (const Seg& g : segs) total += g.len;
        const char* s = (const char*)src;
        size_t issued = 0, drained = 0, seg_i = 0, seg_off = 0;
        int head = cur, tail = cur;
        size_t len_of[N_SLOTS] = {};
        while (drained < total) {
            while (issued < total && issued - drained < (size_t)N_SLOTS * SLOT_BYTES && !(issued > drained && head == tail)) {
                if (busy[head]) {
                    cudaEventSynchronize(ev[head]);
                    busy[head] = false;
                }
                const size_t take = total - issued < SLOT_BYTES ? total - issued : SLOT_BYTES;
                if (cudaMemcpyAsync(slot[head], s + issued, take, cudaMemcpyDeviceToHost, stream) != cudaSuccess)
                    throw std::runtime_error("HostStager: cudaMemcpyAsync (D2H) failed");
                cudaEventRecord(ev[head], stream);
                len_of[head] = take;
                issued += take;
                head = (head + 1) % N_SLOTS;
            }
            cudaEventSynchronize(ev[tail]);
            pending.clear();
            size_t left = len_of[tail], so = 0;
            while (left) {
                const Seg& g = segs[seg_i];
                size_t take = g.len - seg_off < left ? g.len - seg_off : left;
                for (size_t o = 0; o < take; o += GRANULE) {
                    const size_t len = take - o < GRANULE ? take - o : GRANULE;
                    pending.push_back({(char*)g.dst + seg_off + o, (const char*)slot[tail] + so + o, len});
                }
                so += take;
                left -= take;
                seg_off += take;
                if (seg_off == g.len) seg_i++, seg_off = 0;
            }
            pool->parallel_for(pending.size(), [&](size_t i) { run_granule_out(pending[i]); });
            pending.clear();
            drained += len_of[tail];
            tail = (tail + 1) % N_SLOTS;
        }
        cur = head;
    }
};

// A worker thread that runs device->host materialisation jobs in FIFO order while the calling thread keeps feeding the
// compute stream (a commit that also hands back `polynomials` / `leaves` / `digests`): each job typically waits for a
// CUDA event on the worker's own stream, then drains a buffer through the worker's own HostStager.
struct Downloader {
    std::thread worker;
    std::mutex mu;
    std::condition_variable cv_job, cv_idle;
    std::deque<std::function<void()>> jobs;
    bool stop = false, busy = false;
    std::exception_ptr error;
    int device = 0;

    void start(int dev) {
        if (worker.joinable()) return;
        device = dev;
        worker = std::thread([this] { run(); });
    }
    void run() {
        cudaSetDevice(device);
        std::unique_lock<std::mutex> l(mu);
        for (;;) {
            cv_job.wait(l, [&] { return stop || !jobs.empty(); });
            if (stop && jobs.empty()) return;
            std::function<void()> f = std::move(jobs.front());
            jobs.pop_front();
            busy = true;
            l.unlock();
            try {
                if (!error) f(); // after a failure the remaining jobs of the call are dropped
            } catch (...) {
                std::lock_guard<std::mutex> g(mu);
                if (!error) error = std::current_exception();
            }
            l.lock();
            busy = false;
            if (jobs.empty()) cv_idle.notify_all();
        }
    }
    void submit(std::function<void()> f) {
        {
            std::lock_guard<std::mutex> l(mu);
            jobs.push_back(std::move(f));
        }
        cv_job.notify_one();
    }
    // waits until every submitted job has run; rethrows the first failure
    void drain() {
        std::unique_lock<std::mutex> l(mu);
        cv_idle.wait(l, [&] { return jobs.empty() && !busy; });
        if (error) {
            std::exception_ptr e = error;
            error = nullptr;
            std::rethrow_exception(e);
        }
    }
    void shutdown() {
        if (!worker.joinable()) return;
        {
            std::lock_guard<std::mutex> l(mu);
            stop = true;
        }
        cv_job.notify_all();
        worker.join();
    }
};
