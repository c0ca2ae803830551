// api.cu -- the extern "C" boundary of libnlc_commit.so (include/nlc_commit.h).
//
// Every entry point: lock the context, select its device, run on its stream, turn C++ exceptions into
// a non-zero return code + nlc_last_error.  There is no CPU implementation behind any of them: if no
// CUDA device is usable nlc_init fails and no other call can be made.
#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "batch.cuh"

namespace {

thread_local std::string g_init_error;

// Large pageable sources go through the pinned-slot stager (staging.h); everything else is a plain async copy.
void h2d_on(nlc_ctx* ctx, cudaStream_t stream, void* dst, const void* src, size_t bytes, bool flush = true) {
    if (!bytes) return;
    if (bytes >= ((size_t)256 << 10) && HostStager::pageable(src)) {
        ctx->stager.upload(stream, dst, src, bytes);
        if (flush) ctx->stager.flush(stream);
    } else {
        ctx->stager.flush(stream);
        NLC_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, stream));
    }
}
void h2d(nlc_ctx* ctx, void* dst, const void* src, size_t bytes) { h2d_on(ctx, ctx->stream, dst, src, bytes); }
void d2h(nlc_ctx* ctx, void* dst, const void* src, size_t bytes) {
    if (!bytes) return;
    if (bytes >= ((size_t)1 << 20) && HostStager::pageable(dst))
        ctx->stager.download(ctx->stream, src, {{dst, bytes}}); // synchronous; callers sync anyway
    else
        NLC_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, ctx->stream));
}
void sync(nlc_ctx* ctx) { NLC_CUDA(cudaStreamSynchronize(ctx->stream)); }

// RAII device buffer on the context's stream-ordered pool
struct DevBuf {
    nlc_ctx* ctx;
    u64* p;
    DevBuf(nlc_ctx* c, size_t words) : ctx(c), p(c->dmalloc(words)) {}
    ~DevBuf() { ctx->dfree(p); }
    DevBuf(const DevBuf&) = delete;
    DevBuf& operator=(const DevBuf&) = delete;
};

void upload_cols(nlc_ctx* ctx, const u64* const* cols, size_t n_cols, size_t n, u64* dev) {
    NLC_CHECK(cols != nullptr, "null column table");
    for (size_t j = 0; j < n_cols; j++) {
        NLC_CHECK(cols[j] != nullptr, "null column pointer");
        h2d_on(ctx, ctx->stream, dev + j * n, cols[j], n * sizeof(u64), false); // adjacent columns share a slot
    }
    ctx->stager.flush(ctx->stream);
}

enum FftKind { FFT, IFFT, COSET_FFT, COSET_IFFT };

int fft_host(nlc_ctx* ctx, u64* data, size_t n_cols, int log_n, FftKind kind, u64 shift) {
    if (!ctx) return 1;
    return nlc_guarded(ctx, [&] {
        CtxGuard g(ctx);
        NLC_CHECK(data != nullptr || n_cols == 0, "null data");
        NLC_CHECK(log_n >= 0 && log_n <= 32, "log_n out of range (two-adicity is 32)");
        if (n_cols == 0) return;
        const size_t n = (size_t)1 << log_n;
        DevBuf in(ctx, n_cols * n), out(ctx, n_cols * n), scratch(ctx, ntt_scratch_words(n_cols, log_n));
        h2d(ctx, in.p, data, n_cols * n * sizeof(u64));
        NttJob job;
        job.src = in.p;
        job.src_stride = n;
        job.n_cols = n_cols;
        job.log_n = log_n;
        job.inverse = (kind == IFFT || kind == COSET_IFFT);
        if (kind == COSET_FFT) job.pre_shift = gl::canon(shift);
        if (kind == COSET_IFFT) {
            NLC_CHECK(gl::canon(shift) != 0, "coset shift must be non-zero");
            job.post_shift = gl::h_inv(gl::canon(shift));
        }
        job.rows = false;
        job.dst = out.p;
        job.dst_stride = n;
        job.scratch = scratch.p;
        ntt_run(ctx, job);
        d2h(ctx, data, out.p, n_cols * n * sizeof(u64));
        sync(ctx);
    });
}

nlc_batch* commit_dev(nlc_ctx* ctx, const u64* src_dev, bool is_values, size_t W, int log_n, int rate_bits,
                      int cap_height, int shard_log = 0, size_t shard_idx = 0) {
    nlc_batch* b = batch_alloc(ctx, W, log_n, rate_bits, cap_height, shard_log, shard_idx);
    try {
        const size_t n = (size_t)1 << log_n;
        if (is_values)
            batch_ifft_values(b, src_dev);
        else
            NLC_CUDA(cudaMemcpyAsync(b->coeffs, src_dev, W * n * sizeof(u64), cudaMemcpyDeviceToDevice, ctx->stream));
        batch_lde_and_tree(b);
    } catch (...) {
        batch_release(b);
        throw;
    }
    return b;
}

// Host images of the struct fields, filled WHILE the commit runs (all optional)
struct Materialise {
    u64* const* coeffs_cols = nullptr; // polynomials: W host vectors of n
    u64* leaves = nullptr;             // merkle_tree.leaves flattened: N x W row-major
    u64* digests = nullptr;            // merkle_tree.digests
    u64* cap = nullptr;                // merkle_tree.cap
    bool any() const { return coeffs_cols || leaves || digests || cap; }
};

// Runs on the downloader thread: D2H of `bytes` contiguous device bytes into host segments / a strided host matrix,
// after `after` has completed; pinned destinations take plain asynchronous copies, pageable ones the stager.
void dl_wait(nlc_ctx* ctx, cudaEvent_t after) { NLC_CUDA(cudaStreamWaitEvent(ctx->dl_stream, after, 0)); }
void dl_segments(nlc_ctx* ctx, const u64* src, const std::vector<HostStager::Seg>& segs) {
    if (segs.empty()) return;
    if (HostStager::pageable(segs[0].dst)) {
        ctx->dl_stager.download(ctx->dl_stream, src, segs);
    } else {
        const char* s = (const char*)src;
        for (const auto& g : segs) {
            NLC_CUDA(cudaMemcpyAsync(g.dst, s, g.len, cudaMemcpyDeviceToHost, ctx->dl_stream));
            s += g.len;
        }
        NLC_CUDA(cudaStreamSynchronize(ctx->dl_stream));
    }
}
// A row-major host matrix as the source of a commit (starkyx's AirTrace, trace/mod.rs:19-24): columns
// [col0, col0 + W) of `rows` (n x width).  Chunk c of the pipeline below is then the [c0, c0 + nc) strip of EVERY row:
// it crosses PCIe as n pieces of nc words (gathered into the pinned ring by the host copy threads when the matrix is
// pageable, a 2-D DMA otherwise) and one transpose pass on the compute stream turns it into nc columns.
struct RowSource {
    const u64* rows;
    size_t width, col0;
};
int commit_host(nlc_ctx* ctx, const u64* const* cols, size_t W, int log_n, int rate_bits, int cap_height,
                bool is_values, nlc_batch** out, int shard_log = 0, size_t shard_idx = 0, const Materialise* mat = nullptr,
                const RowSource* rs = nullptr) {
    if (!ctx) return 1;
    return nlc_guarded(ctx, [&] {
        CtxGuard g(ctx);
        NLC_CHECK(out != nullptr, "null output handle");
        *out = nullptr;
        const bool want = mat && mat->any();
        size_t n_dl_ev = 0;
        u64* slab = nullptr; // row-major image of the LDE (leaves materialisation)
        if (want) ctx->downloader.start(ctx->device);
        NLC_CHECK(W > 0 && log_n >= 0 && log_n <= 32, "bad shape");
        const size_t n = (size_t)1 << log_n;
        nlc_batch* b = batch_alloc(ctx, W, log_n, rate_bits, cap_height, shard_log, shard_idx);
        try {
            // Columns are independent until the leaf hash, and the leaf sponge absorbs them in order: upload them in
            // chunks on the copy stream and run IFFT + coset NTTs + the sponge blocks of chunk c on the compute
            // stream while chunk c+1 crosses PCIe (inner chunks are whole 8-column sponge blocks; the 12-word
            // sponge states wait in HBM between chunks).  Only the first, short chunk's upload is exposed.
            // Chunk sizes: 16 columns first, then growing by half (each upload hides behind the previous chunk's compute
            // as long as a column computes 1.5x as long as it uploads: measured 1.45-1.95x at rate_bits 1), tripling
            // at rate_bits >= 2 where compute is >= 8x the upload, also from pageable memory (fewer, fuller NTT grids).
            // When the host wants merkle_tree.leaves (N x W row-major, the largest field by far) the sponge is NOT
            // streamed: rows are whole only once every column is extended, so all transforms run first; one transpose
            // pass then writes the row-major image into a second HBM buffer (0.2 ms for config 2), and the copy engine
            // moves it to the host under the one leaf-hash launch and the digest levels.  (Transposing slab by slab on
            // the download stream would wait for the hash to END: its CTAs hold every SM's registers until no unit
            // is left.)
            const bool want_leaves = want && mat->leaves;
            const bool stream_hash = W >= 32 && !want_leaves;
            const size_t chunk = W <= 8 ? W : (W + 4) / 5; // narrow batches: plain fifths, one leaf-hash launch
            // Growth stops at an eighth of the batch (whole 8-column sponge blocks), but never below what fills the
            // NTT grids (2^23 elements): once chunk k+1 uploads faster than chunk k computes the compute stream
            // never waits, and what is exposed at the end is one capped chunk's compute, not half the batch's
            // (2^16 x 2502 at rate_bits 1: 56.6 -> 48 ms against 47.0 ms resident).
            const size_t chunk_cap = std::max<size_t>((W / 8 + 7) / 8 * 8, std::max<size_t>(16, ((size_t)1 << 23) >> log_n));
            const size_t N = b->n_local();
            DevBuf vals(ctx, is_values ? W * n : 1);
            DevBuf sponge(ctx, stream_hash ? 12 * N : 1);
            // the chunk sequence (also used below): 16 columns first, then growing
            auto next_chunk = [&](size_t c0, size_t prev) {
                size_t want = chunk;
                if (stream_hash) want = c0 == 0 ? 16 : std::min(chunk_cap, rate_bits >= 2 ? 3 * prev : (prev * 3 / 2 + 7) / 8 * 8);
                size_t nc = W - c0 < want ? W - c0 : want;
                if (stream_hash && W - c0 - nc < 16) nc = W - c0; // no tiny last chunk
                return nc;
            };
            size_t max_nc = 0;
            for (size_t c0 = 0, nc = 0; rs && c0 < W; c0 += nc) max_nc = std::max(max_nc, nc = next_chunk(c0, nc));
            DevBuf strip0(ctx, rs ? n * max_nc : 1), strip1(ctx, rs ? n * max_nc : 1); // row-major strips, alternating
            const bool rows_pageable = rs && HostStager::pageable(rs->rows);
            std::vector<cudaEvent_t> transposed;
            if (want_leaves) slab = ctx->dmalloc(N * W); // row-major image of the LDE, freed once it is on the host
            // jobs for the downloader thread (it owns dl_stream): after `ev`, drain one piece
            auto after_compute = [&]() {
                cudaEvent_t e = ctx->dl_event(n_dl_ev++);
                NLC_CUDA(cudaEventRecord(e, ctx->stream));
                return e;
            };
            auto dl_coeffs = [&](size_t c0, size_t nc) {
                if (!want || !mat->coeffs_cols) return;
                cudaEvent_t e = after_compute();
                std::vector<HostStager::Seg> segs;
                for (size_t j = c0; j < c0 + nc; j++) segs.push_back({mat->coeffs_cols[j], n * sizeof(u64)});
                const u64* src = b->coeffs + c0 * n;
                ctx->downloader.submit([ctx, e, src, segs] {
                    dl_wait(ctx, e);
                    dl_segments(ctx, src, segs);
                });
            };
            auto dl_leaves = [&]() { // after the last column's LDE
                if (!want_leaves) return;
                batch_leaf_rows(b, 0, N, slab);
                cudaEvent_t e = after_compute();
                u64* dst = mat->leaves;
                const u64* src = slab;
                const size_t bytes = N * W * sizeof(u64);
                ctx->downloader.submit([ctx, e, src, dst, bytes] {
                    dl_wait(ctx, e);
                    dl_segments(ctx, src, {{dst, bytes}});
                });
            };
            auto dl_tree = [&]() {
                if (!want || (!mat->digests && !mat->cap)) return;
                cudaEvent_t e = after_compute();
                const u64 *dg = b->digests, *cp = b->cap;
                const size_t dw = b->n_digest_words, cw = (size_t)4 << b->cap_local();
                u64 *dgo = mat->digests, *cpo = mat->cap;
                ctx->downloader.submit([ctx, e, dg, cp, dw, cw, dgo, cpo] {
                    dl_wait(ctx, e);
                    if (dgo && dw) dl_segments(ctx, dg, {{dgo, dw * sizeof(u64)}});
                    if (cpo) {
                        NLC_CUDA(cudaMemcpyAsync(cpo, cp, cw * sizeof(u64), cudaMemcpyDeviceToHost, ctx->dl_stream));
                        NLC_CUDA(cudaStreamSynchronize(ctx->dl_stream));
                    }
                });
            };
            u64* stage = is_values ? vals.p : b->coeffs;
            NLC_CUDA(cudaEventRecord(ctx->ev_ready, ctx->stream)); // buffers allocated on the compute stream
            NLC_CUDA(cudaStreamWaitEvent(ctx->copy_stream, ctx->ev_ready, 0));
            NLC_CHECK(rs || cols != nullptr, "null column table");
            for (size_t j = 0; !rs && j < W; j++) NLC_CHECK(cols[j] != nullptr, "null column pointer"); // before any upload is issued
            size_t n_ev = 0;
            for (size_t c0 = 0, nc = 0, k = 0; c0 < W; c0 += nc, k++) {
                nc = next_chunk(c0, nc);
                u64* strip = k & 1 ? strip1.p : strip0.p;
                if (rs) {
                    // the transpose that last read this strip buffer must be done before it is overwritten
                    if (k >= 2) NLC_CUDA(cudaStreamWaitEvent(ctx->copy_stream, transposed[k - 2], 0));
                    const u64* src = rs->rows + rs->col0 + c0;
                    if (rows_pageable)
                        ctx->stager.upload_rows(ctx->copy_stream, strip, src, rs->width * sizeof(u64), nc * sizeof(u64), n);
                    else
                        NLC_CUDA(cudaMemcpy2DAsync(strip, nc * sizeof(u64), src, rs->width * sizeof(u64), nc * sizeof(u64), n,
                                                   cudaMemcpyHostToDevice, ctx->copy_stream));
                } else
                    for (size_t j = c0; j < c0 + nc; j++) {
                        NLC_CHECK(cols[j] != nullptr, "null column pointer");
                        h2d_on(ctx, ctx->copy_stream, stage + j * n, cols[j], n * sizeof(u64), false);
                    }
                ctx->stager.flush(ctx->copy_stream);
                cudaEvent_t e = ctx->chunk_event(n_ev++);
                NLC_CUDA(cudaEventRecord(e, ctx->copy_stream));
                NLC_CUDA(cudaStreamWaitEvent(ctx->stream, e, 0));
                if (rs) {
                    transpose_trace_cols(ctx, strip, nc, 0, nc, n, stage + c0 * n, n);
                    transposed.push_back(ctx->chunk_event(n_ev++));
                    NLC_CUDA(cudaEventRecord(transposed.back(), ctx->stream));
                }
                if (is_values) batch_ifft_cols(b, vals.p + c0 * n, c0, nc);
                dl_coeffs(c0, nc); // the coefficients of this chunk cross PCIe while later chunks compute
                batch_lde_cols(b, c0, nc);
                if (stream_hash)
                    merkle_leaf_absorb(ctx, b->lde, N, N, c0, nc, c0 == 0, c0 + nc == W, sponge.p, b->cap_local(),
                                       b->digests, b->cap);
            }
            dl_leaves(); // every column is extended: the leaf rows cross PCIe under the hashing below
            if (stream_hash)
                merkle_levels(ctx, N, b->cap_local(), b->digests, b->cap);
            else
                batch_tree(b);
            dl_tree();
            sync(ctx);
            if (want) ctx->downloader.drain();
            ctx->dfree(slab);
            slab = nullptr;
        } catch (...) {
            if (want) {
                try {
                    ctx->downloader.drain();
                } catch (...) {
                }
                cudaStreamSynchronize(ctx->dl_stream);
            }
            ctx->dfree(slab);
            // uploads of earlier chunks may still be in flight on the copy stream into `vals` / b->coeffs: the
            // stream-ordered frees below are ordered on the compute stream only, so drain the copies first or the
            // pool could hand this memory to the next commit while a stale DMA lands in it
            ctx->stager.abandon();
            cudaStreamSynchronize(ctx->copy_stream);
            cudaStreamSynchronize(ctx->stream);
            batch_release(b);
            throw;
        }
        *out = b;
    });
}

} // namespace

extern "C" {

int nlc_init(int device, nlc_ctx** out) {
    if (!out) return 1;
    *out = nullptr;
    nlc_ctx* ctx = nullptr;
    try {
        int count = 0;
        cudaError_t e = cudaGetDeviceCount(&count);
        if (e != cudaSuccess || count == 0)
            throw NlcError(std::string("nlc_init: no CUDA device (") + cudaGetErrorString(e) +
                           "); this library has no CPU fallback");
        if (device < 0) { // NLC_DEVICE: pick the GPU without touching the caller's code, else the current one
            const char* e = getenv("NLC_DEVICE");
            if (e && *e) {
                char* end = nullptr;
                const long d = strtol(e, &end, 10);
                NLC_CHECK(end && *end == 0 && d >= 0, "NLC_DEVICE must be a non-negative device index");
                device = (int)d;
            } else {
                NLC_CUDA(cudaGetDevice(&device));
            }
        }
        NLC_CHECK(device < count, "device index out of range");
        int prev = 0;
        cudaGetDevice(&prev);
        NLC_CUDA(cudaSetDevice(device));
        ctx = new nlc_ctx();
        ctx->device = device;
        cudaDeviceProp prop;
        NLC_CUDA(cudaGetDeviceProperties(&prop, device));
        NLC_CHECK(prop.major >= 10, "nlc_commit kernels are built for sm_100a (Blackwell) only");
        ctx->sm_count = prop.multiProcessorCount;
        NLC_CUDA(cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking));
        ctx->stream = ctx->own_stream;
        NLC_CUDA(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
        NLC_CUDA(cudaStreamCreateWithFlags(&ctx->dl_stream, cudaStreamNonBlocking));
        NLC_CUDA(cudaEventCreateWithFlags(&ctx->ev_ready, cudaEventDisableTiming));
        cudaMemPool_t pool;
        NLC_CUDA(cudaDeviceGetDefaultMemPool(&pool, device));
        unsigned long long thresh = ~0ull; // keep freed blocks in the pool: commits reuse the same sizes
        NLC_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thresh));
        if (const char* e = getenv("NLC_CACHE_MB")) ctx->cache_budget = (size_t)strtoull(e, nullptr, 10) << 20;
        cudaSetDevice(prev);
        *out = ctx;
        return 0;
    } catch (const std::exception& e) {
        g_init_error = e.what();
        cudaGetLastError();
        delete ctx;
        return 1;
    }
}

void nlc_destroy(nlc_ctx* ctx) {
    if (!ctx) return;
    {
        CtxGuard g(ctx);
        cudaStreamSynchronize(ctx->stream);
        ctx->cache_release_all();
        for (auto& kv : ctx->tables) cudaFree(kv.second);
        if (ctx->pinned) cudaFreeHost(ctx->pinned);
        ctx->downloader.shutdown();
        ctx->stager.destroy();
        ctx->dl_stager.destroy();
        if (ctx->dl_stream) cudaStreamDestroy(ctx->dl_stream);
        for (cudaEvent_t e : ctx->dl_events) cudaEventDestroy(e);
        if (ctx->own_stream) cudaStreamDestroy(ctx->own_stream);
        if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
        if (ctx->ev_ready) cudaEventDestroy(ctx->ev_ready);
        for (cudaEvent_t e : ctx->chunk_events) cudaEventDestroy(e);
    }
    delete ctx;
}

const char* nlc_last_error(nlc_ctx* ctx) { return ctx ? nlc_thread_error().c_str() : g_init_error.c_str(); }
int nlc_device(nlc_ctx* ctx) { return ctx ? ctx->device : -1; }

int nlc_set_stream(nlc_ctx* ctx, void* cuda_stream) {
    if (!ctx) return 1;
    return nlc_guarded(ctx, [&] {
        CtxGuard g(ctx);
        NLC_CUDA(cudaStreamSynchronize(ctx->stream));
        ctx->stream = cuda_stream ? (cudaStream_t)cuda_stream : ctx->own_stream;
    });
}
int nlc_cache_stats(nlc_ctx* ctx, uint64_t out[4]) {
    if (!ctx || !out) return 1;
    CtxGuard g(ctx);
    out[0] = ctx->cache_hits;
    out[1] = ctx->cache_misses;
    out[2] = ctx->cache_miss_us;
    out[3] = ctx->cache_bytes;
    return 0;
}
int nlc_wait_stream(nlc_ctx* ctx, void* producer_stream) {
    if (!ctx) return 1;
    return nlc_guarded(ctx, [&] {
        CtxGuard g(ctx);
        if ((cudaStream_t)producer_stream == ctx->stream) return;
        // cudaEventRecord on stream 0 records on the legacy NULL stream (torch's default), which the context's
        // non-blocking stream does not order against by itself
        NLC_CUDA(cudaEventRecord(ctx->ev_ready, (cudaStream_t)producer_stream));
        NLC_CUDA(cudaStreamWaitEvent(ctx->stream, ctx->ev_ready, 0));
    });
}
int nlc_synchronize(nlc_ctx* ctx) {
    if (!ctx) return 1;
    return nlc_guarded(ctx, [&] {
        CtxGuard g(ctx);
        sync(ctx);
    });
}
uint64_t nlc_launch_count(nlc_ctx* ctx) { return ctx ? ctx->launches.load() : 0; }

int nlc_timing_enable(nlc_ctx* ctx, int on) {
    if (!ctx) return 1;
    CtxGuard g(ctx);
    ctx->timing = on != 0;
    return 0;
}
int nlc_timing_read(nlc_ctx* ctx, nlc_timing* out) {
    if (!ctx || !out) return 1;
    return nlc_guarded(ctx, [&] {
        CtxGuard g(ctx);
        sync(ctx);
        memset(out, 0, sizeof *out);
        for (auto& ev : ctx->timing_events) {
            float ms = 0;
            NLC_CUDA(cudaEventElapsedTime(&ms, ev.second.first, ev.second.second));
            switch (ev.first) {
                case 0: out->ifft_ms += ms; break;
                case 1: out->lde_ms += ms; break;
                case 2: out->leaf_hash_ms += ms; out->commits++; break; // one leaf-hash launch per commit
                case 4: out->leaf_hash_ms += ms; break;                  // chunked leaf sponge (host-fed commits)
                default: out->merkle_levels_ms += ms; break;
            }
            cudaEventDestroy(ev.second.first);
            cudaEventDestroy(ev.second.second);
        }
        ctx->timing_events.clear();
    });
}

// ---- a1 / a2 ----
int nlc_commit_from_values(nlc_ctx* ctx, const uint64_t* const* cols, size_t n_cols, int log_n, int rate_bits,
                           int cap_height, nlc_batch** out) {
    return commit_host(ctx, cols, n_cols, log_n, rate_bits, cap_height, true, out);
}
int nlc_commit_from_coeffs(nlc_ctx* ctx, const uint64_t* const* cols, size_t n_cols, int log_n, int rate_bits,
                           int cap_height, nlc_batch** out) {
    return commit_host(ctx, cols, n_cols, log_n, rate_bits, cap_height, false, out);
}
int nlc_commit_from_values_dev(nlc_ctx* ctx, const uint64_t* values_dev, size_t n_cols, int log_n, int rate_bits,
                               int cap_height, nlc_batch** out) {
    if (!ctx) return 1;
    return nlc_guarded(ctx, [&] {
        CtxGuard g(ctx);
        NLC_CHECK(out && values_dev, "null argument");
        *out = commit_dev(ctx, values_dev, true, n_cols, log_n, rate_bits, cap_height);
    });
}
int nlc_commit_from_coeffs_dev(nlc_ctx* ctx, const uint64_t* coeffs_dev, size_t n_cols, int log_n, int rate_bits,
                               int cap_height, nlc_batch** out) {
    if (!ctx) return 1;
    return nlc_guarded(ctx, [&] {
        CtxGuard g(ctx);
        NLC_CHECK(out && coeffs_dev, "null argument");
        *out = commit_dev(ctx, coeffs_dev, false, n_cols, log_n, rate_bits, cap_height);
    });
}
int nlc_commit_from_values_materialised(nlc_ctx* ctx, const uint64_t* const* cols, size_t n_cols, int log_n, int rate_bits,
                                        int cap_height, uint64_t* const* coeffs_out_cols, uint64_t* leaves_out,
                                        uint64_t* digests_out, uint64_t* cap_out, nlc_batch** out) {
    Materialise m;
    m.coeffs_cols = coeffs_out_cols;
    m.leaves = leaves_out;
    m.digests = digests_out;
    m.cap = cap_out;
    return commit_host(ctx, cols, n_cols, log_n, rate_bits, cap_height, true, out, 0, 0, &m);
}
// ---- 8e.2: one coset shard of a commit ----
static int shard_log_of(size_t n_shards) {
    int l = 0;
    while (((size_t)1 << l) < n_shards) l++;
    return ((size_t)1 << l) == n_shards ? l : -1;
}
int nlc_commit_from_values_shard(nlc_ctx* ctx, const uint64_t* const* cols, size_t n_cols, int log_n, int rate_bits,
                                 int cap_height, size_t shard_idx, size_t n_shards, nlc_batch** out) {
    if (!ctx) return 1;
    const int sl = shard_log_of(n_shards);
    if (sl < 0) return nlc_guarded(ctx, [&] { NLC_CHECK(false, "n_shards must be a power of two"); });
    return commit_host(ctx, cols, n_cols, log_n, rate_bits, cap_height, true, out, sl, shard_idx);
}
int nlc_commit_from_values_shard_dev(nlc_ctx* ctx, const uint64_t* values_dev, size_t n_cols, int log_n,
                                     int rate_bits, int cap_height, size_t shard_idx, size_t n_shards,
                                     nlc_batch** out) {
    if (!ctx) return 1;
    return nlc_guarded(ctx, [&] {
        CtxGuard g(ctx);
        NLC_CHECK(out && values_dev, "null argument");
        const int sl = shard_log_of(n_shards);
        NLC_CHECK(sl >= 0, "n_shards must be a power of two");
        *out = commit_dev(ctx, values_dev, true, n_cols, log_n, rate_bits, cap_height, sl, shard_idx);
    });
}
int nlc_ifft_columns_bcast_dev(nlc_ctx* ctx, const uint64_t* values_dev, size_t n_cols, int log_n, size_t col0,
                               size_t n_peers, uint64_t* const* peer_coeffs) {
    if (!ctx) return 1;
    return nlc_guarded(ctx, [&] {
        CtxGuard g(ctx);
        NLC_CHECK(values_dev && peer_coeffs, "null argument");
        NLC_CHECK(n_cols > 0, "PolynomialBatch: at least one polynomial required");
        NLC_CHECK(log_n >= 0 && log_n <= 32, "log_n out of range (two-adicity is 32)");
        NLC_CHECK(n_peers >= 1 && n_peers <= 16, "1..16 coefficient buffers");
        for (size_t p = 0; p < n_peers; p++) NLC_CHECK(peer_coeffs[p] != nullptr, "null peer buffer");
        ifft_columns_bcast_dev(ctx, values_dev, n_cols, log_n, col0, n_peers, peer_coeffs);
    });
}
int nlc_commit_from_coeffs_shard_dev(nlc_ctx* ctx, uint64_t* coeffs_dev, size_t n_cols, int log_n, int rate_bits,
                                     int cap_height, size_t shard_idx, size_t n_shards, nlc_batch** out) {
    if (!ctx) return 1;
    return nlc_guarded(ctx, [&] {
        CtxGuard g(ctx);
        NLC_CHECK(out && coeffs_dev, "null argument");
        *out = nullptr;
        const int sl = shard_log_of(n_shards);
        NLC_CHECK(sl >= 0, "n_shards must be a power of two");
        *out = commit_from_coeffs_shard_dev(ctx, coeffs_dev, n_cols, log_n, rate_bits, cap_height, sl, shard_idx);
    });
}
// ---- 8e.3: column shards for the transforms, exchange (caller), leaf-row shards for the tree ----
int nlc_lde_columns_dev(nlc_ctx* ctx, const uint64_t* src_dev, size_t n_cols, int log_n, int rate_bits, int is_values,
                        uint64_t* coeffs_out_dev, uint64_t* lde_out_dev) {
    if (!ctx) return 1;
    return nlc_guarded(ctx, [&] {
        CtxGuard g(ctx);
        NLC_CHECK(src_dev && lde_out_dev, "null argument");
        NLC_CHECK(n_cols > 0, "PolynomialBatch: at least one polynomial required");
        NLC_CHECK(log_n >= 0 && rate_bits >= 0 && log_n + rate_bits <= 32, "degree_log + rate_bits exceeds the two-adicity");
        lde_columns_dev(ctx, src_dev, n_cols, log_n, rate_bits, is_values != 0, coeffs_out_dev, lde_out_dev);
    });
}
int nlc_lde_columns_scatter_dev(nlc_ctx* ctx, const uint64_t* src_dev, size_t n_cols, int log_n, int rate_bits,
                                int is_values, uint64_t* coeffs_out_dev, uint64_t* send_dev, size_t n_shards,
                                uint64_t* const* peer_rows, size_t col0) {
    if (!ctx) return 1;
    return nlc_guarded(ctx, [&] {
        CtxGuard g(ctx);
        NLC_CHECK(src_dev && (send_dev || peer_rows), "null argument");
        NLC_CHECK(n_cols > 0, "PolynomialBatch: at least one polynomial required");
        NLC_CHECK(log_n >= 0 && rate_bits >= 0 && log_n + rate_bits <= 32, "degree_log + rate_bits exceeds the two-adicity");
        const int sl = shard_log_of(n_shards);
        NLC_CHECK(sl >= 0 && n_shards <= 16, "n_shards must be a power of two, at most 16");
        NLC_CHECK(sl <= log_n + rate_bits, "more shards than leaves");
        NttScatter sc;
        sc.log_per = log_n + rate_bits - sl;
        sc.shard_stride = n_cols << sc.log_per;
        if (peer_rows)
            for (size_t p = 0; p < n_shards; p++) {
                NLC_CHECK(peer_rows[p] != nullptr, "null peer buffer");
                sc.peer[p] = peer_rows[p];
            }
        lde_columns_dev(ctx, src_dev, n_cols, log_n, rate_bits, is_values != 0, coeffs_out_dev, send_dev, &sc,
                        peer_rows ? col0 : 0);
    });
}
// receive buffers other processes can write: plain cudaMalloc memory (stream-ordered pools cannot be exported)
int nlc_peer_alloc(nlc_ctx* ctx, size_t words, uint64_t** ptr_out, unsigned char handle_out[64]) {
    if (!ctx) return 1;
    return nlc_guarded(ctx, [&] {
        CtxGuard g(ctx);
        NLC_CHECK(ptr_out && handle_out && words, "null argument");
        static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size");
        void* p = nullptr;
        NLC_CUDA(cudaMalloc(&p, words * sizeof(u64)));
        cudaIpcMemHandle_t h;
        cudaError_t e = cudaIpcGetMemHandle(&h, p);
        if (e != cudaSuccess) {
            cudaFree(p);
            NLC_CUDA(e);
        }
        memcpy(handle_out, &h, 64);
        *ptr_out = (u64*)p;
    });
}
int nlc_peer_open(nlc_ctx* ctx, const unsigned char handle[64], uint64_t** ptr_out) {
    if (!ctx) return 1;
    return nlc_guarded(ctx, [&] {
        CtxGuard g(ctx);
        NLC_CHECK(ptr_out && handle, "null argument");
        cudaIpcMemHandle_t h;
        memcpy(&h, handle, 64);
        void* p = nullptr;
        NLC_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
        *ptr_out = (u64*)p;
    });
}
int nlc_peer_close(nlc_ctx* ctx, uint64_t* ptr) {
    if (!ctx) return 1;
    return nlc_guarded(ctx, [&] {
        CtxGuard g(ctx);
        if (ptr) NLC_CUDA(cudaIpcCloseMemHandle(ptr));
    });
}
int nlc_peer_free(nlc_ctx* ctx, uint64_t* ptr) {
    if (!ctx) return 1;
    return nlc_guarded(ctx, [&] {
        CtxGuard g(ctx);
        NLC_CUDA(cudaStreamSynchronize(ctx->stream));
        if (ptr) NLC_CUDA(cudaFree(ptr));
    });
}
int nlc_commit_from_lde_rows_dev(nlc_ctx* ctx, uint64_t* lde_rows_dev, size_t n_cols, int log_n, int rate_bits,
                                 int cap_height, size_t shard_idx, size_t n_shards, nlc_batch** out) {
    if (!ctx) return 1;
    return nlc_guarded(ctx, [&] {
        CtxGuard g(ctx);
        NLC_CHECK(out && lde_rows_dev, "null argument");
        *out = nullptr;
        const int sl = shard_log_of(n_shards);
        NLC_CHECK(sl >= 0, "n_shards must be a power of two");
        *out = commit_from_lde_rows_dev(ctx, lde_rows_dev, n_cols, log_n, rate_bits, cap_height, sl, shard_idx);
    });
}
int nlc_batch_shard(const nlc_batch* b, size_t* shard_idx, size_t* n_shards) {
    if (!b) return 1;
    if (shard_idx) *shard_idx = b->shard_idx;
    if (n_shards) *n_shards = (size_t)1 << b->shard_log;
    return 0;
}

int nlc_commit_from_trace_rows(nlc_ctx* ctx, const uint64_t* rows, size_t width, size_t col0, size_t n_cols, int log_n,
                               int rate_bits, int cap_height, nlc_batch** out) {
    if (!ctx) return 1;
    const int rc = nlc_guarded(ctx, [&] {
        NLC_CHECK(out && rows, "null argument");
        *out = nullptr;
        NLC_CHECK(n_cols > 0 && col0 + n_cols <= width, "column range outside the trace");
        NLC_CHECK(log_n >= 0 && log_n <= 32, "bad shape");
    });
    if (rc) return rc;
    // Only the [col0, col0 + n_cols) strip of each row crosses PCIe (ed25519: 1 651 of 4 153 columns), and it does so in
    // COLUMN chunks through the pipeline of commit_host: while chunk c+1 uploads, chunk c is transposed, inverse-
    // transformed, extended and absorbed by the leaf sponge -- the commit of a wide trace ends one chunk after its upload
    // instead of a whole commit after it (2^16 x 800 from pageable rows: ~32 -> ~20 ms).
    const RowSource rs{rows, width, col0};
    return commit_host(ctx, nullptr, n_cols, log_n, rate_bits, cap_height, true, out, 0, 0, nullptr, &rs);
}

// SURVEY.md section 5 "checkpoint / resume": a PolynomialBatch that was built earlier and serialised (circuit data:
// constants_sigmas_commitment).  The LDE and the tree are recomputed from the coefficients (bit-identical by
// construction, a few ms) and checked against what the serialised struct carried.
int nlc_batch_import(nlc_ctx* ctx, const uint64_t* const* coeff_cols, size_t n_cols, int log_n, int rate_bits,
                     int cap_height, const uint64_t* expect_cap, const uint64_t* expect_digests, nlc_batch** out) {
    if (!ctx) return 1;
    nlc_batch* b = nullptr;
    int rc = commit_host(ctx, coeff_cols, n_cols, log_n, rate_bits, cap_height, false, &b);
    if (rc) return rc;
    rc = nlc_guarded(ctx, [&] {
        CtxGuard g(ctx);
        NLC_CHECK(out != nullptr, "null output handle");
        *out = nullptr;
        if (expect_cap) {
            const size_t words = (size_t)4 << cap_height;
            DevBuf e(ctx, words);
            h2d(ctx, e.p, expect_cap, words * sizeof(u64));
            NLC_CHECK(!device_words_differ(ctx, e.p, b->cap, words),
                      "PolynomialBatch import: the MerkleCap recomputed from the coefficients differs from the serialised one");
        }
        if (expect_digests && b->n_digest_words) {
            DevBuf e(ctx, b->n_digest_words);
            h2d(ctx, e.p, expect_digests, b->n_digest_words * sizeof(u64));
            NLC_CHECK(!device_words_differ(ctx, e.p, b->digests, b->n_digest_words),
                      "PolynomialBatch import: the digests recomputed from the coefficients differ from the serialised ones");
        }
        *out = b;
    });
    if (rc) nlc_batch_free(b);
    return rc;
}
void nlc_batch_free(nlc_batch* b) {
    if (!b) return;
    CtxGuard g(b->ctx);
    batch_release(b);
}

int nlc_batch_info(const nlc_batch* b, size_t* n_cols, int* log_n, int* rate_bits, int* cap_height) {
    if (!b) return 1;
    if (n_cols) *n_cols = b->W;
    if (log_n) *log_n = b->log_n;
    if (rate_bits) *rate_bits = b->rate_bits;
    if (cap_height) *cap_height = b->cap_height;
    return 0;
}
int nlc_batch_cap(nlc_batch* b, uint64_t* out) {
    if (!b) return 1;
    return nlc_guarded(b->ctx, [&] {
        CtxGuard g(b->ctx);
        d2h(b->ctx, out, b->cap, ((size_t)4 << b->cap_local()) * sizeof(u64));
        sync(b->ctx);
    });
}
int nlc_batch_coeffs(nlc_batch* b, uint64_t* const* out_cols) {
    if (!b) return 1;
    return nlc_guarded(b->ctx, [&] {
        CtxGuard g(b->ctx);
        const size_t n = (size_t)1 << b->log_n;
        NLC_CHECK(out_cols != nullptr, "null column table");
        NLC_CHECK(b->coeffs != nullptr, "a leaf-row shard holds no coefficients: they stay with the column owners");
        if (b->W * n * sizeof(u64) >= ((size_t)1 << 20) && HostStager::pageable(out_cols[0])) {
            std::vector<HostStager::Seg> segs; // W separate Vecs on the host, one contiguous matrix on the device
            for (size_t j = 0; j < b->W; j++) segs.push_back({out_cols[j], n * sizeof(u64)});
            b->ctx->stager.download(b->ctx->stream, b->coeffs, segs);
        } else {
            for (size_t j = 0; j < b->W; j++) d2h(b->ctx, out_cols[j], b->coeffs + j * n, n * sizeof(u64));
        }
        sync(b->ctx);
    });
}
int nlc_batch_leaves(nlc_batch* b, uint64_t* out) {
    if (!b) return 1;
    return nlc_guarded(b->ctx, [&] {
        CtxGuard g(b->ctx);
        // the device keeps the LDE column-major; plonky2's row-major leaves are made on demand, a slab at a time
        nlc_ctx* ctx = b->ctx;
        const size_t N = b->n_local();
        size_t slab = ((size_t)8 << 20) / b->W; // ~64 MiB of rows
        if (slab < 32) slab = 32;
        if (slab > N) slab = N;
        DevBuf rows(ctx, slab * b->W);
        for (size_t r0 = 0; r0 < N; r0 += slab) {
            const size_t nr = N - r0 < slab ? N - r0 : slab;
            batch_leaf_rows(b, r0, nr, rows.p);
            d2h(ctx, out + r0 * b->W, rows.p, nr * b->W * sizeof(u64));
            sync(ctx); // the slab buffer is reused
        }
    });
}
int nlc_batch_digests(nlc_batch* b, uint64_t* out) {
    if (!b) return 1;
    return nlc_guarded(b->ctx, [&] {
        CtxGuard g(b->ctx);
        d2h(b->ctx, out, b->digests, b->n_digest_words * sizeof(u64));
        sync(b->ctx);
    });
}
int nlc_batch_lde_rows(nlc_batch* b, const uint64_t* idx, size_t m, size_t step, uint64_t* out) {
    if (!b) return 1;
    return nlc_guarded(b->ctx, [&] {
        nlc_ctx* ctx = b->ctx;
        CtxGuard g(ctx);
        NLC_CHECK(b->shard_log == 0, "get_lde_values on a coset shard: rows of other shards live on other GPUs");
        if (!m) return;
        DevBuf idx_dev(ctx, m), rows(ctx, m * b->W);
        h2d(ctx, idx_dev.p, idx, m * sizeof(u64));
        batch_gather_lde_rows(b, idx_dev.p, 0, m, step, rows.p);
        d2h(ctx, out, rows.p, m * b->W * sizeof(u64));
        sync(ctx);
    });
}
int nlc_batch_lde_rows_range(nlc_batch* b, size_t i_start, size_t count, size_t step, uint64_t* out) {
    if (!b) return 1;
    return nlc_guarded(b->ctx, [&] {
        nlc_ctx* ctx = b->ctx;
        CtxGuard g(ctx);
        NLC_CHECK(b->shard_log == 0, "get_lde_values on a coset shard: rows of other shards live on other GPUs");
        NLC_CHECK(out || !count, "null argument");
        if (!count) return;
        DevBuf rows(ctx, count * b->W);
        batch_gather_lde_rows(b, nullptr, i_start, count, step, rows.p);
        d2h(ctx, out, rows.p, count * b->W * sizeof(u64));
        sync(ctx);
    });
}
int nlc_batch_open(nlc_batch* b, size_t leaf_index, uint64_t* row_out, uint64_t* siblings_out) {
    if (!b) return 1;
    return nlc_guarded(b->ctx, [&] {
        nlc_ctx* ctx = b->ctx;
        CtxGuard g(ctx);
        NLC_CHECK(b->shard_log == 0, "MerkleTree::prove on a coset shard is not supported");
        const int log_N = b->log_n + b->rate_bits;
        NLC_CHECK(leaf_index < ((size_t)1 << log_N), "leaf index out of range");
        const size_t sib = (size_t)4 * (log_N - b->cap_height);
        DevBuf idx_dev(ctx, 1), buf(ctx, b->W + sib);
        u64 idx = leaf_index;
        h2d(ctx, idx_dev.p, &idx, sizeof(u64));
        merkle_open_many(ctx, b->lde, b->digests, (size_t)1 << log_N, b->W, (size_t)1 << log_N, b->cap_height, idx_dev.p, 0, 1, buf.p,
                         b->W + sib, 0);
        if (row_out) d2h(ctx, row_out, buf.p, b->W * sizeof(u64));
        if (siblings_out) d2h(ctx, siblings_out, buf.p + b->W, sib * sizeof(u64));
        sync(ctx);
    });
}
const uint64_t* nlc_batch_lde_dev(const nlc_batch* b) { return b ? b->lde : nullptr; }
const uint64_t* nlc_batch_coeffs_dev(const nlc_batch* b) { return b ? b->coeffs : nullptr; }
const uint64_t* nlc_batch_cap_dev(const nlc_batch* b) { return b ? b->cap : nullptr; }
int nlc_batch_eval(nlc_batch* b, const uint64_t z[2], uint64_t* out) {
    if (!b) return 1;
    return nlc_guarded(b->ctx, [&] {
        nlc_ctx* ctx = b->ctx;
        CtxGuard g(ctx);
        NLC_CHECK(b->coeffs != nullptr, "a leaf-row shard holds no coefficients: they stay with the column owners");
        DevBuf res(ctx, 2 * b->W);
        batch_eval_ext(b, z, res.p);
        d2h(ctx, out, res.p, 2 * b->W * sizeof(u64));
        sync(ctx);
    });
}

// ---- a3 ----
int nlc_fft(nlc_ctx* ctx, uint64_t* data, size_t n_cols, int log_n) { return fft_host(ctx, data, n_cols, log_n, FFT, 1); }
int nlc_ifft(nlc_ctx* ctx, uint64_t* data, size_t n_cols, int log_n) {
    return fft_host(ctx, data, n_cols, log_n, IFFT, 1);
}
int nlc_coset_fft(nlc_ctx* ctx, uint64_t* data, size_t n_cols, int log_n, uint64_t shift) {
    return fft_host(ctx, data, n_cols, log_n, COSET_FFT, shift);
}
int nlc_coset_ifft(nlc_ctx* ctx, uint64_t* data, size_t n_cols, int log_n, uint64_t shift) {
    return fft_host(ctx, data, n_cols, log_n, COSET_IFFT, shift);
}
int nlc_lde_onto_coset(nlc_ctx* ctx, const uint64_t* in, uint64_t* out, size_t n_cols, int log_n, int rate_bits) {
    if (!ctx) return 1;
    return nlc_guarded(ctx, [&] {
        CtxGuard g(ctx);
        NLC_CHECK(in && out, "null argument");
        NLC_CHECK(log_n >= 0 && rate_bits >= 0 && log_n + rate_bits <= 32, "log_n + rate_bits out of range");
        if (!n_cols) return;
        const size_t n = (size_t)1 << log_n, N = n << rate_bits;
        const int log_N = log_n + rate_bits;
        DevBuf vals(ctx, n_cols * n), padded(ctx, n_cols * N), res(ctx, n_cols * N);
        DevBuf scratch(ctx, std::max(ntt_scratch_words(n_cols, log_n), ntt_scratch_words(n_cols, log_N)));
        h2d(ctx, vals.p, in, n_cols * n * sizeof(u64));
        NLC_CUDA(cudaMemsetAsync(padded.p, 0, n_cols * N * sizeof(u64), ctx->stream));
        NttJob inv;
        inv.src = vals.p;
        inv.src_stride = n;
        inv.n_cols = n_cols;
        inv.log_n = log_n;
        inv.inverse = true;
        inv.dst = padded.p;
        inv.dst_stride = N; // coefficients land in the first n slots of each zero-padded column
        inv.scratch = scratch.p;
        ntt_run(ctx, inv);
        NttJob fwd;
        fwd.src = padded.p;
        fwd.src_stride = N;
        fwd.n_cols = n_cols;
        fwd.log_n = log_N;
        fwd.pre_shift = gl::GENERATOR;
        fwd.dst = res.p;
        fwd.dst_stride = N;
        fwd.scratch = scratch.p;
        ntt_run(ctx, fwd);
        d2h(ctx, out, res.p, n_cols * N * sizeof(u64));
        sync(ctx);
    });
}

// ---- quotient_polys without the AIR evaluation (prover.rs:253-358, 158-178) ----
static void quotient_check(size_t n_ch, int log_n, int qdb, size_t qdf, int rate_bits) {
    NLC_CHECK(n_ch > 0, "no challenges");
    NLC_CHECK(log_n >= 0 && qdb >= 0 && log_n + qdb <= 32, "log_n + quotient_degree_bits out of range");
    NLC_CHECK(qdb <= rate_bits, "Having constraints of degree higher than the rate is not supported yet.");
    NLC_CHECK(qdf >= 1 && qdf <= ((size_t)1 << qdb), "quotient_degree_factor exceeds 2^quotient_degree_bits");
}
int nlc_quotient_domain(nlc_ctx* ctx, int log_n, int quotient_degree_bits, uint64_t* coset_x, uint64_t* lagrange_first,
                        uint64_t* lagrange_last, uint64_t* zh_inv) {
    if (!ctx) return 1;
    return nlc_guarded(ctx, [&] {
        CtxGuard g(ctx);
        const int qdb = quotient_degree_bits;
        NLC_CHECK(log_n >= 0 && qdb >= 0 && log_n + qdb <= 32, "log_n + quotient_degree_bits out of range");
        const size_t n = (size_t)1 << log_n, size = n << qdb;
        const int log_size = log_n + qdb;
        if (zh_inv) {
            const std::vector<u64> inv = zero_poly_on_coset_inverses(log_n, qdb);
            memcpy(zh_inv, inv.data(), inv.size() * sizeof(u64));
        }
        if (coset_x) {
            DevBuf x(ctx, size);
            coset_points(ctx, log_size, x.p);
            d2h(ctx, coset_x, x.p, size * sizeof(u64));
            sync(ctx);
        }
        if (lagrange_first || lagrange_last) {
            // PolynomialValues::selector(degree, i).lde_onto_coset(quotient_degree_bits), i = 0 and degree - 1
            DevBuf vals(ctx, 2 * n), padded(ctx, 2 * size), res(ctx, 2 * size);
            DevBuf scratch(ctx, std::max(ntt_scratch_words(2, log_n), ntt_scratch_words(2, log_size)));
            NLC_CUDA(cudaMemsetAsync(vals.p, 0, 2 * n * sizeof(u64), ctx->stream));
            NLC_CUDA(cudaMemsetAsync(padded.p, 0, 2 * size * sizeof(u64), ctx->stream));
            const u64 one = 1;
            NLC_CUDA(cudaMemcpyAsync(vals.p, &one, sizeof(u64), cudaMemcpyHostToDevice, ctx->stream));
            NLC_CUDA(cudaMemcpyAsync(vals.p + 2 * n - 1, &one, sizeof(u64), cudaMemcpyHostToDevice, ctx->stream));
            NttJob inv;
            inv.src = vals.p;
            inv.src_stride = n;
            inv.n_cols = 2;
            inv.log_n = log_n;
            inv.inverse = true;
            inv.dst = padded.p;
            inv.dst_stride = size;
            inv.scratch = scratch.p;
            ntt_run(ctx, inv);
            NttJob fwd;
            fwd.src = padded.p;
            fwd.src_stride = size;
            fwd.n_cols = 2;
            fwd.log_n = log_size;
            fwd.pre_shift = gl::GENERATOR;
            fwd.dst = res.p;
            fwd.dst_stride = size;
            fwd.scratch = scratch.p;
            ntt_run(ctx, fwd);
            if (lagrange_first) d2h(ctx, lagrange_first, res.p, size * sizeof(u64));
            if (lagrange_last) d2h(ctx, lagrange_last, res.p + size, size * sizeof(u64));
            sync(ctx);
        }
    });
}

int nlc_quotient_commit(nlc_ctx* ctx, const uint64_t* evals, size_t n_challenges, int log_n, int quotient_degree_bits,
                        size_t quotient_degree_factor, int divide_by_zh, int rate_bits, int cap_height,
                        nlc_batch** out) {
    if (!ctx) return 1;
    return nlc_guarded(ctx, [&] {
        CtxGuard g(ctx);
        NLC_CHECK(out && evals, "null argument");
        *out = nullptr;
        quotient_check(n_challenges, log_n, quotient_degree_bits, quotient_degree_factor, rate_bits);
        const size_t words = n_challenges << (log_n + quotient_degree_bits);
        DevBuf ev(ctx, words);
        h2d(ctx, ev.p, evals, words * sizeof(u64));
        *out = quotient_commit_dev(ctx, ev.p, n_challenges, log_n, quotient_degree_bits, quotient_degree_factor,
                                   divide_by_zh != 0, rate_bits, cap_height);
        sync(ctx);
    });
}
int nlc_quotient_commit_dev(nlc_ctx* ctx, const uint64_t* evals_dev, size_t n_challenges, int log_n,
                            int quotient_degree_bits, size_t quotient_degree_factor, int divide_by_zh, int rate_bits,
                            int cap_height, nlc_batch** out) {
    if (!ctx) return 1;
    return nlc_guarded(ctx, [&] {
        CtxGuard g(ctx);
        NLC_CHECK(out && evals_dev, "null argument");
        *out = nullptr;
        quotient_check(n_challenges, log_n, quotient_degree_bits, quotient_degree_factor, rate_bits);
        *out = quotient_commit_dev(ctx, evals_dev, n_challenges, log_n, quotient_degree_bits, quotient_degree_factor,
                                   divide_by_zh != 0, rate_bits, cap_height);
    });
}

// ---- AIR evaluation on the device (prover.rs:306-352) ----
int nlc_air_eval_dev(nlc_ctx* ctx, const nlc_air_program* air, nlc_batch* const* trace_data, int n_trace_data,
                     const uint64_t* public_inputs, const uint64_t* global_values, const uint64_t* challenges,
                     const uint64_t* alphas, size_t n_alphas, int quotient_degree_bits, uint64_t* evals_out_dev) {
    if (!ctx) return 1;
    return nlc_guarded(ctx, [&] {
        CtxGuard g(ctx);
        NLC_CHECK(evals_out_dev != nullptr, "null argument");
        air_eval_dev(ctx, air, trace_data, n_trace_data, public_inputs, global_values, challenges, alphas, n_alphas,
                     quotient_degree_bits, evals_out_dev);
    });
}
int nlc_air_eval(nlc_ctx* ctx, const nlc_air_program* air, nlc_batch* const* trace_data, int n_trace_data,
                 const uint64_t* public_inputs, const uint64_t* global_values, const uint64_t* challenges,
                 const uint64_t* alphas, size_t n_alphas, int quotient_degree_bits, uint64_t* evals_out) {
    if (!ctx) return 1;
    return nlc_guarded(ctx, [&] {
        CtxGuard g(ctx);
        NLC_CHECK(evals_out && trace_data && n_trace_data > 0 && trace_data[0], "null argument");
        NLC_CHECK(quotient_degree_bits >= 0 && quotient_degree_bits <= 32, "quotient_degree_bits out of range");
        const size_t words = n_alphas << (trace_data[0]->log_n + quotient_degree_bits);
        DevBuf ev(ctx, words);
        air_eval_dev(ctx, air, trace_data, n_trace_data, public_inputs, global_values, challenges, alphas, n_alphas,
                     quotient_degree_bits, ev.p);
        d2h(ctx, evals_out, ev.p, words * sizeof(u64));
        sync(ctx);
    });
}
int nlc_stark_quotient_commit(nlc_ctx* ctx, const nlc_air_program* air, nlc_batch* const* trace_data, int n_trace_data,
                              const uint64_t* public_inputs, const uint64_t* global_values, const uint64_t* challenges,
                              const uint64_t* alphas, size_t n_alphas, size_t quotient_degree_factor, int cap_height,
                              nlc_batch** out) {
    if (!ctx) return 1;
    return nlc_guarded(ctx, [&] {
        CtxGuard g(ctx);
        NLC_CHECK(out && trace_data && n_trace_data > 0 && trace_data[0], "null argument");
        *out = nullptr;
        int qdb = 0;
        while (((size_t)1 << qdb) < quotient_degree_factor) qdb++; // log2_ceil (prover.rs:266)
        const int log_n = trace_data[0]->log_n, rate_bits = trace_data[0]->rate_bits;
        quotient_check(n_alphas, log_n, qdb, quotient_degree_factor, rate_bits);
        DevBuf ev(ctx, n_alphas << (log_n + qdb));
        air_eval_dev(ctx, air, trace_data, n_trace_data, public_inputs, global_values, challenges, alphas, n_alphas, qdb, ev.p);
        *out = quotient_commit_dev(ctx, ev.p, n_alphas, log_n, qdb, quotient_degree_factor, true, rate_bits, cap_height);
        sync(ctx);
    });
}

// ---- plonk-side quotient path (plonk.cu) ----
static void same_ctx(nlc_ctx* ctx, std::initializer_list<nlc_batch*> bs) {
    for (nlc_batch* b : bs) NLC_CHECK(b && b->ctx == ctx, "commitment missing or from another context");
}
int nlc_plonk_zs_partial_products_commit(nlc_ctx* ctx, const nlc_plonk_circuit* circuit, nlc_batch* constants_sigmas,
                                         nlc_batch* wires, const uint64_t* betas, const uint64_t* gammas, int cap_height,
                                         nlc_batch** out) {
    if (!ctx) return 1;
    return nlc_guarded(ctx, [&] {
        CtxGuard g(ctx);
        NLC_CHECK(out != nullptr, "null output handle");
        *out = nullptr;
        same_ctx(ctx, {constants_sigmas, wires});
        *out = plonk_zs_partial_products_commit_dev(ctx, circuit, constants_sigmas, wires, betas, gammas, cap_height);
        sync(ctx);
    });
}
int nlc_plonk_vanishing_eval(nlc_ctx* ctx, const nlc_plonk_circuit* circuit, const nlc_air_program* gate_constraints,
                             nlc_batch* constants_sigmas, nlc_batch* wires, nlc_batch* zs_partial_products,
                             const uint64_t public_inputs_hash[4], const uint64_t* betas, const uint64_t* gammas,
                             const uint64_t* alphas, uint64_t* evals_out) {
    if (!ctx) return 1;
    return nlc_guarded(ctx, [&] {
        CtxGuard g(ctx);
        NLC_CHECK(circuit && evals_out, "null argument");
        same_ctx(ctx, {constants_sigmas, wires, zs_partial_products});
        int qdb = 0;
        while ((1 << qdb) < circuit->quotient_degree_factor) qdb++;
        const size_t words = circuit->num_challenges << (wires->log_n + qdb);
        DevBuf ev(ctx, words);
        plonk_vanishing_dev(ctx, circuit, gate_constraints, constants_sigmas, wires, zs_partial_products, public_inputs_hash,
                            betas, gammas, alphas, ev.p);
        d2h(ctx, evals_out, ev.p, words * sizeof(u64));
        sync(ctx);
    });
}
int nlc_plonk_quotient_commit(nlc_ctx* ctx, const nlc_plonk_circuit* circuit, const nlc_air_program* gate_constraints,
                              nlc_batch* constants_sigmas, nlc_batch* wires, nlc_batch* zs_partial_products,
                              const uint64_t public_inputs_hash[4], const uint64_t* betas, const uint64_t* gammas,
                              const uint64_t* alphas, int cap_height, nlc_batch** out) {
    if (!ctx) return 1;
    return nlc_guarded(ctx, [&] {
        CtxGuard g(ctx);
        NLC_CHECK(circuit && out, "null argument");
        *out = nullptr;
        same_ctx(ctx, {constants_sigmas, wires, zs_partial_products});
        int qdb = 0;
        while ((1 << qdb) < circuit->quotient_degree_factor) qdb++;
        const int log_n = wires->log_n, rate_bits = wires->rate_bits;
        quotient_check(circuit->num_challenges, log_n, qdb, (size_t)circuit->quotient_degree_factor, rate_bits);
        DevBuf ev(ctx, circuit->num_challenges << (log_n + qdb));
        plonk_vanishing_dev(ctx, circuit, gate_constraints, constants_sigmas, wires, zs_partial_products, public_inputs_hash,
                            betas, gammas, alphas, ev.p);
        *out = quotient_commit_dev(ctx, ev.p, circuit->num_challenges, log_n, qdb, (size_t)circuit->quotient_degree_factor, true,
                                   rate_bits, cap_height);
        sync(ctx);
    });
}

// ---- a4 ----
int nlc_poseidon_permute(nlc_ctx* ctx, uint64_t* states, size_t count) {
    if (!ctx) return 1;
    return nlc_guarded(ctx, [&] {
        CtxGuard g(ctx);
        if (!count) return;
        DevBuf st(ctx, count * 12);
        h2d(ctx, st.p, states, count * 12 * sizeof(u64));
        poseidon_permute_many(ctx, st.p, count);
        d2h(ctx, states, st.p, count * 12 * sizeof(u64));
        sync(ctx);
    });
}
static int hash_host(nlc_ctx* ctx, const uint64_t* in, size_t count, size_t len, bool or_noop, uint64_t* out) {
    if (!ctx) return 1;
    return nlc_guarded(ctx, [&] {
        CtxGuard g(ctx);
        if (!count) return;
        DevBuf din(ctx, count * len), dout(ctx, count * 4);
        h2d(ctx, din.p, in, count * len * sizeof(u64));
        hash_many(ctx, din.p, count, len, or_noop, dout.p);
        d2h(ctx, out, dout.p, count * 4 * sizeof(u64));
        sync(ctx);
    });
}
int nlc_hash_no_pad(nlc_ctx* ctx, const uint64_t* in, size_t count, size_t len, uint64_t* out) {
    return hash_host(ctx, in, count, len, false, out);
}
int nlc_hash_or_noop(nlc_ctx* ctx, const uint64_t* in, size_t count, size_t len, uint64_t* out) {
    return hash_host(ctx, in, count, len, true, out);
}
int nlc_two_to_one(nlc_ctx* ctx, const uint64_t* left, const uint64_t* right, size_t count, uint64_t* out) {
    if (!ctx) return 1;
    return nlc_guarded(ctx, [&] {
        CtxGuard g(ctx);
        if (!count) return;
        DevBuf l(ctx, count * 4), r(ctx, count * 4), o(ctx, count * 4);
        h2d(ctx, l.p, left, count * 4 * sizeof(u64));
        h2d(ctx, r.p, right, count * 4 * sizeof(u64));
        two_to_one_many(ctx, l.p, r.p, count, o.p);
        d2h(ctx, out, o.p, count * 4 * sizeof(u64));
        sync(ctx);
    });
}

// ---- a5 ----
int nlc_merkle_new(nlc_ctx* ctx, const uint64_t* leaves, size_t n_leaves, size_t leaf_len, int cap_height,
                   nlc_tree** out) {
    if (!ctx) return 1;
    return nlc_guarded(ctx, [&] {
        CtxGuard g(ctx);
        NLC_CHECK(out && leaves, "null argument");
        *out = nullptr;
        const int h = ilog2(n_leaves);
        NLC_CHECK(n_leaves > 0 && ((size_t)1 << h) == n_leaves, "MerkleTree::new: number of leaves must be a power of two");
        NLC_CHECK(cap_height >= 0 && cap_height <= h, "cap_height should be at most log2(leaves.len())");
        nlc_tree* t = new nlc_tree();
        t->ctx = ctx;
        t->n_leaves = n_leaves;
        t->leaf_len = leaf_len;
        t->cap_height = cap_height;
        try {
            t->leaves = ctx->dmalloc(n_leaves * leaf_len);
            t->digests = ctx->dmalloc(8 * (n_leaves - ((size_t)1 << cap_height)));
            t->cap = ctx->dmalloc((size_t)4 << cap_height);
            h2d(ctx, t->leaves, leaves, n_leaves * leaf_len * sizeof(u64));
            merkle_build(ctx, t->leaves, n_leaves, leaf_len, 0, cap_height, t->digests, t->cap);
            sync(ctx);
        } catch (...) {
            ctx->dfree(t->leaves);
            ctx->dfree(t->digests);
            ctx->dfree(t->cap);
            delete t;
            throw;
        }
        *out = t;
    });
}
int nlc_tree_cap(nlc_tree* t, uint64_t* out) {
    if (!t) return 1;
    return nlc_guarded(t->ctx, [&] {
        CtxGuard g(t->ctx);
        d2h(t->ctx, out, t->cap, ((size_t)4 << t->cap_height) * sizeof(u64));
        sync(t->ctx);
    });
}
int nlc_tree_digests(nlc_tree* t, uint64_t* out) {
    if (!t) return 1;
    return nlc_guarded(t->ctx, [&] {
        CtxGuard g(t->ctx);
        d2h(t->ctx, out, t->digests, 8 * (t->n_leaves - ((size_t)1 << t->cap_height)) * sizeof(u64));
        sync(t->ctx);
    });
}
int nlc_tree_open(nlc_tree* t, size_t leaf_index, uint64_t* leaf_out, uint64_t* siblings_out) {
    if (!t) return 1;
    return nlc_guarded(t->ctx, [&] {
        nlc_ctx* ctx = t->ctx;
        CtxGuard g(ctx);
        NLC_CHECK(leaf_index < t->n_leaves, "leaf index out of range");
        const size_t sib = (size_t)4 * (ilog2(t->n_leaves) - t->cap_height);
        DevBuf idx_dev(ctx, 1), buf(ctx, t->leaf_len + sib);
        u64 idx = leaf_index;
        h2d(ctx, idx_dev.p, &idx, sizeof(u64));
        merkle_open_many(ctx, t->leaves, t->digests, t->n_leaves, t->leaf_len, 0, t->cap_height, idx_dev.p, 0, 1, buf.p,
                         t->leaf_len + sib, 0);
        if (leaf_out) d2h(ctx, leaf_out, buf.p, t->leaf_len * sizeof(u64));
        if (siblings_out) d2h(ctx, siblings_out, buf.p + t->leaf_len, sib * sizeof(u64));
        sync(ctx);
    });
}
void nlc_tree_free(nlc_tree* t) {
    if (!t) return;
    CtxGuard g(t->ctx);
    if (t->owns_leaves) t->ctx->dfree(t->leaves);
    t->ctx->dfree(t->digests);
    t->ctx->dfree(t->cap);
    delete t;
}

// ---- a9 ----
void nlc_challenger_init(nlc_challenger* c) {
    if (c) memset(c, 0, sizeof *c);
}
int nlc_challenger_observe(nlc_ctx* ctx, nlc_challenger* c, const uint64_t* elems, size_t n) {
    if (!ctx) return 1;
    return nlc_guarded(ctx, [&] {
        CtxGuard g(ctx);
        NLC_CHECK(c && (elems || !n), "null argument");
        NLC_CHECK(c->input_len < 8 && c->output_len <= 8, "corrupt challenger buffers");
        if (!n) return;
        const size_t cw = (sizeof(nlc_challenger) + 7) / 8;
        DevBuf cd(ctx, cw), ed(ctx, n);
        h2d(ctx, cd.p, c, sizeof *c);
        h2d(ctx, ed.p, elems, n * sizeof(u64));
        challenger_observe_dev(ctx, (nlc_challenger*)cd.p, ed.p, n);
        d2h(ctx, c, cd.p, sizeof *c);
        sync(ctx);
    });
}
int nlc_challenger_get(nlc_ctx* ctx, nlc_challenger* c, size_t n, uint64_t* out) {
    if (!ctx) return 1;
    return nlc_guarded(ctx, [&] {
        CtxGuard g(ctx);
        NLC_CHECK(c && (out || !n), "null argument");
        NLC_CHECK(c->input_len < 8 && c->output_len <= 8, "corrupt challenger buffers");
        if (!n) return;
        const size_t cw = (sizeof(nlc_challenger) + 7) / 8;
        DevBuf cd(ctx, cw), od(ctx, n);
        h2d(ctx, cd.p, c, sizeof *c);
        challenger_get_dev(ctx, (nlc_challenger*)cd.p, n, od.p);
        d2h(ctx, c, cd.p, sizeof *c);
        d2h(ctx, out, od.p, n * sizeof(u64));
        sync(ctx);
    });
}

// ---- a7 / a8 / a10 ----
int nlc_fri_params_constant_arity(nlc_fri_params* p, int degree_bits, int rate_bits, int cap_height,
                                  int proof_of_work_bits, int num_query_rounds, int arity_bits, int final_poly_bits) {
    if (!p || arity_bits < 1 || arity_bits > 8 || degree_bits < 0 || rate_bits < 0) return 1;
    memset(p, 0, sizeof *p);
    p->degree_bits = degree_bits;
    p->rate_bits = rate_bits;
    p->cap_height = cap_height;
    p->proof_of_work_bits = proof_of_work_bits;
    p->num_query_rounds = num_query_rounds;
    // FriReductionStrategy::ConstantArityBits: keep reducing while the result stays above the final
    // polynomial size and the round's tree still has at least 2^cap_height leaves
    int d = degree_bits;
    while (d > final_poly_bits && d + rate_bits - arity_bits >= cap_height) {
        if (p->n_reduction_rounds == NLC_MAX_FRI_ROUNDS) return 1; // never hand back fewer rounds than plonky2 computes
        p->reduction_arity_bits[p->n_reduction_rounds++] = arity_bits;
        d -= arity_bits;
    }
    return 0;
}

size_t nlc_fri_proof_words(const nlc_fri_params* p, nlc_batch* const* oracles, int n_oracles) {
    if (!p || (!oracles && n_oracles)) return 0;
    return fri_proof_words_impl(p, oracles, n_oracles, nullptr, nullptr, nullptr, nullptr);
}

int nlc_prove_openings(nlc_ctx* ctx, const nlc_fri_instance* inst, nlc_batch* const* oracles, int n_oracles,
                       nlc_challenger* challenger, const nlc_fri_params* p, uint64_t* proof_out) {
    if (!ctx) return 1;
    return nlc_guarded(ctx, [&] {
        CtxGuard g(ctx);
        NLC_CHECK(inst && oracles && challenger && p && proof_out, "null argument");
        NLC_CHECK(challenger->input_len < 8 && challenger->output_len <= 8, "corrupt challenger buffers");
        nlc_fri_state* s = fri_state_new(ctx, inst, oracles, n_oracles, p);
        try {
            const size_t cw = (sizeof(nlc_challenger) + 7) / 8;
            DevBuf cd(ctx, cw);
            h2d(ctx, cd.p, challenger, sizeof *challenger);
            fri_prove_openings(s, inst, (nlc_challenger*)cd.p);
            FriView v = fri_view(s);
            u64 resp = 0;
            d2h(ctx, proof_out, v.proof, v.proof_words * sizeof(u64));
            d2h(ctx, challenger, cd.p, sizeof *challenger);
            d2h(ctx, &resp, v.resp, sizeof(u64));
            sync(ctx);
            const int pb = p->proof_of_work_bits;
            NLC_CHECK(pb == 0 || (resp >> (64 - pb)) == 0, "proof-of-work response check failed");
        } catch (...) {
            fri_state_release(s);
            throw;
        }
        fri_state_release(s);
    });
}

int nlc_fri_begin(nlc_ctx* ctx, const nlc_fri_instance* inst, nlc_batch* const* oracles, int n_oracles,
                  const uint64_t alpha[2], const nlc_fri_params* p, nlc_fri_state** out) {
    if (!ctx) return 1;
    return nlc_guarded(ctx, [&] {
        CtxGuard g(ctx);
        NLC_CHECK(inst && oracles && alpha && p && out, "null argument");
        *out = nullptr;
        nlc_fri_state* s = fri_state_new(ctx, inst, oracles, n_oracles, p);
        try {
            u64 a[2] = {gl::canon(alpha[0]), gl::canon(alpha[1])};
            h2d(ctx, fri_view(s).alpha, a, sizeof a);
            fri_begin(s, inst);
            sync(ctx);
        } catch (...) {
            fri_state_release(s);
            throw;
        }
        *out = s;
    });
}
int nlc_fri_commit_round(nlc_fri_state* s, uint64_t* cap_out) {
    if (!s) return 1;
    nlc_ctx* ctx = fri_view(s).ctx;
    return nlc_guarded(ctx, [&] {
        CtxGuard g(ctx);
        fri_commit_round(s);
        FriView v = fri_view(s);
        const size_t cap_words = (size_t)4 << v.cap_height;
        if (cap_out) d2h(ctx, cap_out, v.proof + (size_t)(v.round - 1) * cap_words, cap_words * sizeof(u64));
        sync(ctx);
    });
}
int nlc_fri_fold_round(nlc_fri_state* s, const uint64_t beta[2]) {
    if (!s) return 1;
    nlc_ctx* ctx = fri_view(s).ctx;
    return nlc_guarded(ctx, [&] {
        CtxGuard g(ctx);
        NLC_CHECK(beta != nullptr, "null argument");
        u64 b[2] = {gl::canon(beta[0]), gl::canon(beta[1])};
        h2d(ctx, fri_view(s).beta, b, sizeof b);
        fri_fold_round(s);
        sync(ctx);
    });
}
int nlc_fri_final_poly(nlc_fri_state* s, uint64_t* out) {
    if (!s) return 1;
    nlc_ctx* ctx = fri_view(s).ctx;
    return nlc_guarded(ctx, [&] {
        CtxGuard g(ctx);
        fri_final_poly_to_proof(s);
        FriView v = fri_view(s);
        d2h(ctx, out, v.proof + v.final_off, 2 * fri_final_poly_len(s) * sizeof(u64));
        sync(ctx);
    });
}
int nlc_fri_pow(nlc_ctx* ctx, const nlc_challenger* c, int proof_of_work_bits, uint64_t* witness_out) {
    if (!ctx) return 1;
    return nlc_guarded(ctx, [&] {
        CtxGuard g(ctx);
        NLC_CHECK(c && witness_out, "null argument");
        NLC_CHECK(c->input_len < 8, "corrupt challenger buffers");
        NLC_CHECK(proof_of_work_bits >= 0 && proof_of_work_bits < 64, "proof_of_work_bits out of range");
        const size_t cw = (sizeof(nlc_challenger) + 7) / 8;
        DevBuf cd(ctx, cw), w(ctx, 1);
        h2d(ctx, cd.p, c, sizeof *c);
        fri_pow_dev(ctx, (const nlc_challenger*)cd.p, proof_of_work_bits, w.p);
        d2h(ctx, witness_out, w.p, sizeof(u64));
        sync(ctx);
    });
}
int nlc_fri_queries(nlc_fri_state* s, const uint64_t* x_indices, size_t n_queries, uint64_t* out) {
    if (!s) return 1;
    nlc_ctx* ctx = fri_view(s).ctx;
    return nlc_guarded(ctx, [&] {
        CtxGuard g(ctx);
        FriView v = fri_view(s);
        NLC_CHECK(x_indices && out, "null argument");
        NLC_CHECK(n_queries <= (size_t)v.num_queries, "more queries than num_query_rounds");
        if (!n_queries) return;
        std::vector<u64> idx(n_queries);
        for (size_t i = 0; i < n_queries; i++) idx[i] = x_indices[i] & (v.N - 1);
        h2d(ctx, v.xidx, idx.data(), n_queries * sizeof(u64));
        fri_queries(s, n_queries);
        d2h(ctx, out, v.proof + v.query_off, n_queries * v.query_words * sizeof(u64));
        sync(ctx);
    });
}
void nlc_fri_state_free(nlc_fri_state* s) {
    if (!s) return;
    CtxGuard g(fri_view(s).ctx);
    fri_state_release(s);
}

} // extern "C"
