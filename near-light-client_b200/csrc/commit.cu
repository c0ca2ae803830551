// commit.cu -- PolynomialBatch::from_values / from_coeffs on the device, and the accessors that
// materialise the fields of plonky2's PolynomialBatch / MerkleTree structs on demand.
//
// Replaces plonky2::fri::oracle::PolynomialBatch::{from_values, from_coeffs, lde_values,
// get_lde_values} (plonky2 0.1.4 @ d2598bd), called in tree at
// /root/reference/vendor/starkyx/starkyx/src/plonky2/stark/prover.rs:91-93,171-178,289-294 and
// plonky2/stark/config.rs:92-94.
//
// Data layout in HBM (one nlc_batch):
//   coeffs  : W x n   column-major (polynomial j contiguous), natural coefficient order
//   lde     : W x N   column-major, lde[j*N + k] = column j of leaf k, leaf k = LDE row bitrev_N(k)  (N = n << rate_bits)
//             -- the order a decimation-in-frequency NTT produces by itself, so the LDE is never transposed or
//             bit-reversed in HBM; the leaf hash reads it with coalesced column loads, and plonky2's row-major
//             `leaves` are produced only when the host asks for them (batch_leaf_rows)
//   digests : 2*(N - 2^cap) x 4 in plonky2's interleaved layout;  cap : 2^cap x 4
// The LDE is computed coset by coset: LDE point i = m*2^r + k lies on the coset 7*w_N^k*<w_n>, and
// after the index bit reversal coset k is the contiguous leaf range [bitrev_r(k)*n, +n).  So each
// coset is one size-n NTT of coeff_j*(7 w_N^k)^j written straight into that range of every column: no zero
// padding, no size-N transform, no transpose or bit-reversal pass.
#include <algorithm>

#include "batch.cuh"

namespace {

__global__ void k_transpose_cols(const u64* __restrict__ rows, size_t width, size_t col0, size_t n_cols, size_t n,
                                 u64* __restrict__ out, size_t out_stride) {
    __shared__ u64 tile[32][33];
    const size_t r0 = (size_t)blockIdx.x * 32, c0 = (size_t)blockIdx.y * 32;
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        size_t r = r0 + i, c = c0 + threadIdx.x;
        if (r < n && c < n_cols) tile[i][threadIdx.x] = rows[r * width + col0 + c];
    }
    __syncthreads();
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        size_t c = c0 + i, r = r0 + threadIdx.x;
        if (r < n && c < n_cols) out[c * out_stride + r] = tile[threadIdx.x][i];
    }
}

// quotient_values of prover.rs:312-352 turned into columns: evals is the size x n_ch row-major matrix the AIR
// evaluation produced (one row per LDE point, natural order); column j, natural order, optionally multiplied by
// 1/Z_H(x_i) = zh_inv[i mod 2^qdb] (ZeroPolyOnCoset::eval_inverse, prover.rs:336-340)
__global__ void k_quotient_cols(const u64* __restrict__ evals, size_t n_ch, size_t size, const u64* __restrict__ zh_inv,
                                u64 zh_mask, u64* __restrict__ out) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < size; i += stride) {
        const u64 d = zh_inv ? zh_inv[i & zh_mask] : 1;
        for (size_t j = 0; j < n_ch; j++) {
            const u64 v = gl::dcanon(evals[i * n_ch + j]);
            out[j * size + i] = zh_inv ? gl::mul(v, d) : v;
        }
    }
}

// trim_to_len (prover.rs:161-165): every coefficient of column j at index >= keep must be zero
__global__ void k_trim_check(const u64* __restrict__ coeffs, size_t n_cols, size_t size, size_t keep, int* __restrict__ bad) {
    const size_t tail = size - keep, total = n_cols * tail, stride = (size_t)gridDim.x * blockDim.x;
    int any = 0;
    for (size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += stride)
        any |= gl::dcanon(coeffs[(t / tail) * size + keep + t % tail]) != 0;
    if (__syncthreads_or(any) && threadIdx.x == 0) atomicOr(bad, 1);
}

// x_i = shift * w^i, i < size (cyclic_subgroup_coset_known_order, prover.rs:298-302)
__global__ void k_coset_points(u64 shift, u64 w, size_t size, u64* __restrict__ out) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < size; i += stride)
        out[i] = gl::mul(shift, gl::pow(w, i));
}

// out[i*W + j] = lde[j*N + bitrev_N(idx[i]*step)]
// idx == nullptr: the arithmetic stripe idx[i] = i_start + i (get_lde_values_packed over a whole range, one launch)
__global__ void k_gather_lde_rows(const u64* __restrict__ lde, size_t W, int log_N, const u64* __restrict__ idx,
                                  u64 i_start, size_t step, u64* __restrict__ out) {
    const size_t i = blockIdx.x;
    const u64 pt = ((idx ? idx[i] : i_start + i) * step) & (((u64)1 << log_N) - 1);
    const u64 row = log_N ? (__brevll(pt) >> (64 - log_N)) : 0;
    for (size_t j = threadIdx.x; j < W; j += blockDim.x) out[i * W + j] = lde[(j << log_N) + row];
}

// out[(r - row0)*W + j] = lde[j*N + r], 32x32 tiles through shared memory (both sides coalesced)
__global__ void k_leaf_rows(const u64* __restrict__ lde, size_t W, size_t N, size_t row0, size_t n_rows,
                            u64* __restrict__ out) {
    __shared__ u64 tile[32][33];
    const size_t r0 = (size_t)blockIdx.x * 32, c0 = (size_t)blockIdx.y * 32;
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        size_t c = c0 + i, r = r0 + threadIdx.x;
        if (r < n_rows && c < W) tile[i][threadIdx.x] = lde[c * N + row0 + r];
    }
    __syncthreads();
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        size_t r = r0 + i, c = c0 + threadIdx.x;
        if (r < n_rows && c < W) out[r * W + c] = tile[threadIdx.x][i];
    }
}

__global__ void k_words_differ(const u64* __restrict__ a, const u64* __restrict__ b, size_t words, int* __restrict__ flag) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    int any = 0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < words; i += stride) any |= a[i] != b[i];
    if (__syncthreads_or(any) && threadIdx.x == 0) atomicOr(flag, 1);
}

// One block per polynomial: sum_j c_j z^j over F_{p^2}.
__global__ void __launch_bounds__(256) k_eval_ext(const u64* __restrict__ coeffs, size_t n, gl::Ext2 z, u64* out) {
    __shared__ gl::Ext2 part[256];
    const u64* c = coeffs + blockIdx.x * n;
    const size_t per = (n + 255) / 256;
    const size_t lo = threadIdx.x * per, hi = lo + per < n ? lo + per : n;
    gl::Ext2 acc = {0, 0};
    for (size_t j = hi; j-- > lo;) {
        acc = gl::ext_mul(acc, z);
        acc.a0 = gl::add(acc.a0, gl::canon(c[j]));
    }
    if (lo < n) acc = gl::ext_mul(acc, gl::ext_pow(z, lo));
    part[threadIdx.x] = acc;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if ((int)threadIdx.x < s) part[threadIdx.x] = gl::ext_add(part[threadIdx.x], part[threadIdx.x + s]);
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        out[2 * blockIdx.x] = part[0].a0;
        out[2 * blockIdx.x + 1] = part[0].a1;
    }
}

} // namespace

nlc_batch* batch_alloc(nlc_ctx* ctx, size_t W, int log_n, int rate_bits, int cap_height, int shard_log,
                       size_t shard_idx) {
    NLC_CHECK(W > 0, "PolynomialBatch: at least one polynomial required");
    NLC_CHECK(log_n >= 0 && rate_bits >= 0 && log_n + rate_bits <= 32, "degree_log + rate_bits exceeds the two-adicity");
    NLC_CHECK(cap_height >= 0 && cap_height <= log_n + rate_bits, "cap_height should be at most log2(leaves.len())");
    NLC_CHECK(shard_log >= 0 && shard_log <= rate_bits && shard_log <= cap_height,
              "coset sharding needs n_shards <= 2^rate_bits and n_shards <= 2^cap_height");
    NLC_CHECK(shard_idx < ((size_t)1 << shard_log), "shard index out of range");
    nlc_batch* b = new nlc_batch();
    b->ctx = ctx;
    b->W = W;
    b->log_n = log_n;
    b->rate_bits = rate_bits;
    b->cap_height = cap_height;
    b->shard_log = shard_log;
    b->shard_idx = shard_idx;
    const size_t n = (size_t)1 << log_n, N = b->n_local();
    try {
        b->coeffs = ctx->dmalloc(W * n);
        b->lde = ctx->dmalloc(W * N);
        b->n_digest_words = 8 * (N - ((size_t)1 << b->cap_local()));
        b->digests = ctx->dmalloc(b->n_digest_words);
        b->cap = ctx->dmalloc((size_t)4 << b->cap_local());
    } catch (...) {
        batch_release(b);
        throw;
    }
    return b;
}

void batch_release(nlc_batch* b) {
    if (!b) return;
    nlc_ctx* ctx = b->ctx;
    if (b->owns_coeffs) ctx->dfree(b->coeffs);
    if (b->owns_lde) ctx->dfree(b->lde);
    ctx->dfree(b->digests);
    ctx->dfree(b->cap);
    delete b;
}

// "FFT + blinding" (+ "transpose LDEs" + reverse_index_bits, which never happen) of from_coeffs for the columns
// [c0, c0 + nc): columns are independent until the leaf hash, so a host-fed commit can run this on one chunk while
// the next chunk is still crossing PCIe.
void batch_lde_cols(nlc_batch* b, size_t c0, size_t nc, const NttScatter* sc) {
    nlc_ctx* ctx = b->ctx;
    const int log_n = b->log_n, r = b->rate_bits;
    const size_t n = (size_t)1 << log_n, N = b->n_local();
    const u64 w_N = gl::h_root_of_unity(log_n + r);
    const size_t cosets_local = ((size_t)1 << r) >> b->shard_log, kr0 = b->shard_idx * cosets_local;
    if (!sc && ntt_can_batch_cosets(log_n)) {
        // all of this shard's cosets in one launch per pass, in place on the LDE columns (no scratch): 2 launches
        // instead of 2 * 2^r, each with 2^r times the CTAs
        NttJob job;
        job.src = b->coeffs + c0 * n;
        job.src_stride = n;
        job.n_cols = nc;
        job.log_n = log_n;
        job.cols = true;
        job.dst = b->lde;
        job.dst_stride = N;
        job.row_base = 0;
        job.col0 = c0;
        job.log_cosets = r - b->shard_log;
        job.pre_tables = ntt_coset_tables(ctx, log_n, r, kr0, cosets_local);
        ctx->phase_begin(1);
        ntt_run(ctx, job);
        ctx->phase_end();
        return;
    }
    u64* scratch = nullptr;
    const size_t sw = ntt_scratch_words(nc, log_n);
    if (sw) scratch = ctx->dmalloc(sw);
    ctx->phase_begin(1);
    for (size_t k = 0; k < ((size_t)1 << r); k++) {
        size_t kr = 0; // bitrev_r(k): coset k is the leaf range [kr*n, (kr+1)*n)
        for (int t = 0; t < r; t++) kr |= ((k >> t) & 1) << (r - 1 - t);
        if (kr < kr0 || kr >= kr0 + cosets_local) continue; // another shard's coset
        NttJob job;
        job.src = b->coeffs + c0 * n;
        job.src_stride = n;
        job.n_cols = nc;
        job.log_n = log_n;
        job.pre_shift = gl::h_mul(gl::GENERATOR, gl::h_pow(w_N, k));
        job.cols = true;
        job.dst = b->lde;
        job.dst_stride = N;
        job.row_base = (kr - kr0) * n;
        job.col0 = c0;
        job.scratch = scratch;
        if (sc) { // re-blocked by leaf-row shard: columns are 2^log_per apart inside a shard's block
            job.scatter = *sc;
            job.dst_stride = (size_t)1 << sc->log_per;
        }
        ntt_run(ctx, job);
    }
    ctx->phase_end();
    ctx->dfree(scratch);
}

// "build Merkle tree"
void batch_tree(nlc_batch* b) {
    const size_t N = b->n_local();
    merkle_build(b->ctx, b->lde, N, b->W, N, b->cap_local(), b->digests, b->cap);
}

void batch_lde_and_tree(nlc_batch* b) {
    batch_lde_cols(b, 0, b->W);
    batch_tree(b);
}

// "IFFT" of from_values for the columns [c0, c0 + nc): values (column-major, device; column c0 first) -> b->coeffs.
void batch_ifft_cols(nlc_batch* b, const u64* values_dev, size_t c0, size_t nc) {
    nlc_ctx* ctx = b->ctx;
    const size_t n = (size_t)1 << b->log_n;
    u64* scratch = nullptr;
    const size_t sw = ntt_scratch_words(nc, b->log_n);
    if (sw) scratch = ctx->dmalloc(sw);
    NttJob job;
    job.src = values_dev;
    job.src_stride = n;
    job.n_cols = nc;
    job.log_n = b->log_n;
    job.inverse = true;
    job.rows = false;
    job.dst = b->coeffs + c0 * n;
    job.dst_stride = n;
    job.scratch = scratch;
    ctx->phase_begin(0);
    ntt_run(ctx, job);
    ctx->phase_end();
    ctx->dfree(scratch);
}
void batch_ifft_values(nlc_batch* b, const u64* values_dev) { batch_ifft_cols(b, values_dev, 0, b->W); }

// rows_dev: n_rows x width row-major; columns [col0, col0 + n_cols) -> out_dev[c * out_stride + r]
void transpose_trace_cols(nlc_ctx* ctx, const u64* rows_dev, size_t width, size_t col0, size_t n_cols, size_t n_rows,
                          u64* out_dev, size_t out_stride) {
    if (!n_rows || !n_cols) return;
    dim3 grid((unsigned)((n_rows + 31) / 32), (unsigned)((n_cols + 31) / 32));
    NLC_LAUNCH(ctx, k_transpose_cols, grid, dim3(32, 8), 0, rows_dev, width, col0, n_cols, n_rows, out_dev, out_stride);
}

void batch_gather_lde_rows(nlc_batch* b, const u64* idx_dev, u64 i_start, size_t m, size_t step, u64* out_dev) {
    if (!m) return;
    NLC_LAUNCH(b->ctx, k_gather_lde_rows, (unsigned)m, 128, 0, b->lde, b->W, b->log_n + b->rate_bits, idx_dev, i_start,
               step, out_dev);
}

bool device_words_differ(nlc_ctx* ctx, const u64* a_dev, const u64* b_dev, size_t words) {
    if (!words) return false;
    int* flag = (int*)ctx->dmalloc(1);
    int bad = 0;
    try {
        NLC_CUDA(cudaMemsetAsync(flag, 0, sizeof(int), ctx->stream));
        const unsigned grid = (unsigned)std::min<size_t>((words + 255) / 256, (size_t)ctx->sm_count * 8);
        NLC_LAUNCH(ctx, k_words_differ, grid, 256, 0, a_dev, b_dev, words, flag);
        NLC_CUDA(cudaMemcpyAsync(&bad, flag, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
        NLC_CUDA(cudaStreamSynchronize(ctx->stream));
    } catch (...) {
        ctx->dfree(flag);
        throw;
    }
    ctx->dfree(flag);
    return bad != 0;
}

void batch_eval_ext(nlc_batch* b, const u64 z[2], u64* out_dev) {
    gl::Ext2 ze = {gl::canon(z[0]), gl::canon(z[1])};
    NLC_LAUNCH(b->ctx, k_eval_ext, (unsigned)b->W, 256, 0, b->coeffs, (size_t)1 << b->log_n, ze, out_dev);
}

void batch_leaf_rows(nlc_batch* b, size_t row0, size_t n_rows, u64* out_dev, cudaStream_t on) {
    if (!n_rows) return;
    const size_t N = b->n_local();
    dim3 grid((unsigned)((n_rows + 31) / 32), (unsigned)((b->W + 31) / 32));
    k_leaf_rows<<<grid, dim3(32, 8), 0, on ? on : b->ctx->stream>>>(b->lde, b->W, N, row0, n_rows, out_dev);
    b->ctx->launches.fetch_add(1, std::memory_order_relaxed);
    NLC_CUDA(cudaGetLastError());
}

// ---- the constraint-free parts of quotient_polys (prover.rs:253-358) ----
std::vector<u64> zero_poly_on_coset_inverses(int log_n, int qdb) {
    // ZeroPolyOnCoset::new(n_log, rate_bits): Z_H(g w_N^i) = g^n w_{2^qdb}^(i mod 2^qdb) - 1
    u64 g_pow_n = gl::GENERATOR;
    for (int i = 0; i < log_n; i++) g_pow_n = gl::h_mul(g_pow_n, g_pow_n);
    const u64 w = gl::h_root_of_unity(qdb);
    std::vector<u64> inv((size_t)1 << qdb);
    u64 x = 1;
    for (auto& v : inv) {
        v = gl::h_inv(gl::sub(gl::h_mul(g_pow_n, x), 1));
        x = gl::h_mul(x, w);
    }
    return inv;
}

void coset_points(nlc_ctx* ctx, int log_size, u64* out_dev) {
    const size_t size = (size_t)1 << log_size;
    const unsigned grid = (unsigned)std::min<size_t>((size + 255) / 256, (size_t)ctx->sm_count * 8);
    NLC_LAUNCH(ctx, k_coset_points, grid, 256, 0, gl::GENERATOR, gl::h_root_of_unity(log_size), size, out_dev);
}

nlc_batch* quotient_commit_dev(nlc_ctx* ctx, const u64* evals_dev, size_t n_ch, int log_n, int qdb, size_t qdf,
                               bool divide_by_zh, int rate_bits, int cap_height) {
    const size_t n = (size_t)1 << log_n, size = n << qdb, keep = n * qdf;
    const int log_size = log_n + qdb;
    u64* vals = ctx->dmalloc(n_ch * size);
    u64* coef = ctx->dmalloc(n_ch * size);
    u64* scratch = ctx->dmalloc(ntt_scratch_words(n_ch, log_size));
    u64* small = ctx->dmalloc(((size_t)1 << qdb) + 1); // zh_inv, then the trim flag
    nlc_batch* b = nullptr;
    try {
        int* flag = (int*)(small + ((size_t)1 << qdb));
        const u64* zh = nullptr;
        if (divide_by_zh) {
            const std::vector<u64> inv = zero_poly_on_coset_inverses(log_n, qdb);
            // pageable source: the runtime stages it before returning, `inv` may go out of scope
            NLC_CUDA(cudaMemcpyAsync(small, inv.data(), inv.size() * sizeof(u64), cudaMemcpyHostToDevice, ctx->stream));
            zh = small;
        }
        const unsigned grid = (unsigned)std::min<size_t>((size + 255) / 256, (size_t)ctx->sm_count * 8);
        NLC_LAUNCH(ctx, k_quotient_cols, grid, 256, 0, evals_dev, n_ch, size, zh, (u64)(((size_t)1 << qdb) - 1), vals);
        NttJob job; // values.coset_ifft(F::coset_shift()), prover.rs:353-357
        job.src = vals;
        job.src_stride = size;
        job.n_cols = n_ch;
        job.log_n = log_size;
        job.inverse = true;
        job.post_shift = gl::h_inv(gl::GENERATOR);
        job.dst = coef;
        job.dst_stride = size;
        job.scratch = scratch;
        ntt_run(ctx, job);
        if (keep < size) {
            NLC_CUDA(cudaMemsetAsync(flag, 0, sizeof(int), ctx->stream));
            const size_t total = n_ch * (size - keep);
            const unsigned g2 = (unsigned)std::min<size_t>((total + 255) / 256, (size_t)ctx->sm_count * 8);
            NLC_LAUNCH(ctx, k_trim_check, g2, 256, 0, coef, n_ch, size, keep, flag);
            int bad = 0;
            NLC_CUDA(cudaMemcpyAsync(&bad, flag, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
            NLC_CUDA(cudaStreamSynchronize(ctx->stream));
            NLC_CHECK(bad == 0, "Quotient has failed, the vanishing polynomial is not divisible by Z_H");
        }
        // quotient_poly.chunks(degree) (prover.rs:166): polynomial j, chunk c -> column j*qdf + c
        b = batch_alloc(ctx, n_ch * qdf, log_n, rate_bits, cap_height);
        NLC_CUDA(cudaMemcpy2DAsync(b->coeffs, keep * sizeof(u64), coef, size * sizeof(u64), keep * sizeof(u64), n_ch,
                                   cudaMemcpyDeviceToDevice, ctx->stream));
        batch_lde_and_tree(b);
    } catch (...) {
        if (b) batch_release(b);
        ctx->dfree(small);
        ctx->dfree(scratch);
        ctx->dfree(coef);
        ctx->dfree(vals);
        throw;
    }
    ctx->dfree(small);
    ctx->dfree(scratch);
    ctx->dfree(coef);
    ctx->dfree(vals);
    return b;
}

// ---- column shards for the transforms, leaf-row shards for the tree ----
// "IFFT" + "FFT + blinding" of from_values / from_coeffs for the columns one rank owns: lde_out = n_cols x N
// column-major in leaf order (what batch_lde_cols writes), no tree.
void lde_columns_dev(nlc_ctx* ctx, const u64* src_dev, size_t n_cols, int log_n, int rate_bits, bool is_values,
                     u64* coeffs_out_dev, u64* lde_out_dev, const NttScatter* sc, size_t col0) {
    const size_t n = (size_t)1 << log_n;
    nlc_batch view; // a window over the caller's buffers: only the fields the two column passes read
    view.ctx = ctx;
    view.W = n_cols;
    view.log_n = log_n;
    view.rate_bits = rate_bits;
    u64* own = coeffs_out_dev ? nullptr : ctx->dmalloc(n_cols * n);
    view.coeffs = coeffs_out_dev ? coeffs_out_dev : own;
    view.lde = lde_out_dev;
    try {
        if (is_values)
            batch_ifft_cols(&view, src_dev, 0, n_cols);
        else if (view.coeffs != src_dev)
            NLC_CUDA(cudaMemcpyAsync(view.coeffs, src_dev, n_cols * n * sizeof(u64), cudaMemcpyDeviceToDevice, ctx->stream));
        if (sc && col0) { // peers address columns globally: shift the base so that local column j lands at col0 + j
            NttScatter shifted = *sc;
            for (auto& p : shifted.peer)
                if (p) p += col0 << sc->log_per;
            batch_lde_cols(&view, 0, n_cols, &shifted);
        } else {
            batch_lde_cols(&view, 0, n_cols, sc);
        }
    } catch (...) {
        ctx->dfree(own);
        throw;
    }
    ctx->dfree(own);
}

void ifft_columns_bcast_dev(nlc_ctx* ctx, const u64* values_dev, size_t n_cols, int log_n, size_t col0, size_t n_peers,
                            u64* const* peer_coeffs) {
    const size_t n = (size_t)1 << log_n;
    u64* scratch = nullptr;
    const size_t sw = ntt_scratch_words(n_cols, log_n);
    if (sw) scratch = ctx->dmalloc(sw);
    NttJob job;
    job.src = values_dev;
    job.src_stride = n;
    job.n_cols = n_cols;
    job.log_n = log_n;
    job.inverse = true;
    job.dst = peer_coeffs[0] + col0 * n; // unused by the stores, kept for the out-of-place check
    job.dst_stride = n;
    job.scratch = scratch;
    job.scatter.n_bcast = (int)n_peers;
    for (size_t p = 0; p < n_peers; p++) job.scatter.peer[p] = peer_coeffs[p] + col0 * n;
    ctx->phase_begin(0);
    try {
        ntt_run(ctx, job);
    } catch (...) {
        ctx->dfree(scratch);
        throw;
    }
    ctx->phase_end();
    ctx->dfree(scratch);
}

nlc_batch* commit_from_coeffs_shard_dev(nlc_ctx* ctx, u64* coeffs_dev, size_t W, int log_n, int rate_bits, int cap_height,
                                        int shard_log, size_t shard_idx) {
    NLC_CHECK(W > 0, "PolynomialBatch: at least one polynomial required");
    NLC_CHECK(log_n >= 0 && rate_bits >= 0 && log_n + rate_bits <= 32, "degree_log + rate_bits exceeds the two-adicity");
    NLC_CHECK(cap_height >= 0 && cap_height <= log_n + rate_bits, "cap_height should be at most log2(leaves.len())");
    NLC_CHECK(shard_log >= 0 && shard_log <= rate_bits && shard_log <= cap_height,
              "coset sharding needs n_shards <= 2^rate_bits and n_shards <= 2^cap_height");
    NLC_CHECK(shard_idx < ((size_t)1 << shard_log), "shard index out of range");
    nlc_batch* b = new nlc_batch();
    b->ctx = ctx;
    b->W = W;
    b->log_n = log_n;
    b->rate_bits = rate_bits;
    b->cap_height = cap_height;
    b->shard_log = shard_log;
    b->shard_idx = shard_idx;
    b->coeffs = coeffs_dev;
    b->owns_coeffs = false;
    const size_t N = b->n_local();
    try {
        b->lde = ctx->dmalloc(W * N);
        b->n_digest_words = 8 * (N - ((size_t)1 << b->cap_local()));
        b->digests = ctx->dmalloc(b->n_digest_words);
        b->cap = ctx->dmalloc((size_t)4 << b->cap_local());
        batch_lde_and_tree(b);
    } catch (...) {
        batch_release(b);
        throw;
    }
    return b;
}

// The tree over the leaves [shard_idx, shard_idx + 1) * N >> shard_log, whose rows (all W columns) sit column-major in
// lde_rows_dev: digests of the shard's cap subtrees and its 2^(cap_height - shard_log) cap entries.
nlc_batch* commit_from_lde_rows_dev(nlc_ctx* ctx, u64* lde_rows_dev, size_t W, int log_n, int rate_bits, int cap_height,
                                    int shard_log, size_t shard_idx) {
    NLC_CHECK(W > 0, "PolynomialBatch: at least one polynomial required");
    NLC_CHECK(log_n >= 0 && rate_bits >= 0 && log_n + rate_bits <= 32, "degree_log + rate_bits exceeds the two-adicity");
    NLC_CHECK(cap_height >= 0 && cap_height <= log_n + rate_bits, "cap_height should be at most log2(leaves.len())");
    NLC_CHECK(shard_log >= 0 && shard_log <= cap_height, "leaf-row sharding needs n_shards <= 2^cap_height");
    NLC_CHECK(shard_idx < ((size_t)1 << shard_log), "shard index out of range");
    nlc_batch* b = new nlc_batch();
    b->ctx = ctx;
    b->W = W;
    b->log_n = log_n;
    b->rate_bits = rate_bits;
    b->cap_height = cap_height;
    b->shard_log = shard_log;
    b->shard_idx = shard_idx;
    b->lde = lde_rows_dev;
    b->owns_lde = false;
    const size_t N = b->n_local();
    try {
        b->n_digest_words = 8 * (N - ((size_t)1 << b->cap_local()));
        b->digests = ctx->dmalloc(b->n_digest_words);
        b->cap = ctx->dmalloc((size_t)4 << b->cap_local());
        batch_tree(b);
    } catch (...) {
        batch_release(b);
        throw;
    }
    return b;
}
