// ntt.cu -- Goldilocks NTT for sm_100a: bit-reversal-free multi-digit (four-step style) layout.
//
// Replaces plonky2_field's fft_classic / ifft_with_options (field/src/fft.rs @ d2598bd) and the
// coset / LDE wrappers of field/src/polynomial/mod.rs, reached from
// /root/reference/vendor/starkyx/starkyx/src/plonky2/stark/prover.rs:91,171,280-284,356.
//
// A size-2^log_n transform is split into digits S_0 | S_1 | ... | S_last (each <= 10 bits).  A "hi"
// pass takes one digit of a contiguous 2^m block: it gathers 2^S strided elements x T neighbours into
// shared memory (coalesced T*8-byte segments), runs a pure 2^S-point decimation-in-frequency NTT,
// multiplies by the inter-digit twiddle w_{2^m}^{k_hi * i_low} and stores IN PLACE.  The "lo" pass takes
// the last digit on contiguous data and writes either
//   ROWS    : row-major leaves (row = index-bit-reversed evaluation point, exactly plonky2's
//             `leaves` after transpose + reverse_index_bits_in_place), several columns per CTA so a
//             row segment is one coalesced store; or
//   NATURAL : column-contiguous natural order (T bit-reversal-related blocks per CTA, so the
//             scattered positions come out as contiguous T-element segments).
// No separate bit-reversal or transpose pass ever touches HBM.
#include "common.cuh"
#include "gl.cuh"

namespace {

enum { KIND_ROOTS = 1, KIND_TW = 2, KIND_PW = 3 };
constexpr int NTT_THREADS = 256;

__device__ __forceinline__ u32 brev(u32 x, int bits) { return bits ? (__brev(x) >> (32 - bits)) : 0; }

// out[i] = scale * base^i
__global__ void k_pow_table(u64* out, size_t n, u64 base, u64 scale) {
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i < n) out[i] = gl::mul(scale, gl::pow(base, i));
}
// tw[pos << L | i_low] = root^(bitrev_S(pos) * i_low), root = w_{2^m}^{+-1}
__global__ void k_tw_table(u64* out, int m, int S, u64 root) {
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= ((size_t)1 << m)) return;
    const int L = m - S;
    u64 pos = i >> L, low = i & (((u64)1 << L) - 1);
    out[i] = gl::pow(root, (u64)brev((u32)pos, S) * low);
}

// One digit of a 2^m block, strided ("hi") pass.  smem: [2^S][T].
__global__ void __launch_bounds__(NTT_THREADS)
k_ntt_hi(const u64* __restrict__ src, size_t src_stride, u64* __restrict__ dst, size_t dst_stride, int m, int S,
         int log_T, const u64* __restrict__ pre, const u64* __restrict__ tw, const u64* __restrict__ roots) {
    extern __shared__ u64 sm[];
    const int L = m - S;
    const int T = 1 << log_T;
    const u32 tiles_per_block = 1u << (L - log_T);
    const size_t tile = blockIdx.x;
    const size_t base = (tile / tiles_per_block) << m;
    const u32 low0 = (u32)(tile % tiles_per_block) << log_T;
    const u64* s = src + blockIdx.y * src_stride + base;
    u64* d = dst + blockIdx.y * dst_stride + base;
    const u32 total = (u32)T << S;

    for (u32 idx = threadIdx.x; idx < total; idx += NTT_THREADS) {
        u32 t = idx & (T - 1), ih = idx >> log_T;
        size_t off = ((size_t)ih << L) + low0 + t;
        u64 v = gl::canon(s[off]);
        if (pre) v = gl::mul(v, pre[base + off]);
        sm[idx] = v;
    }
    __syncthreads();
    for (int st = 0; st < S; st++) {
        const u32 half = 1u << (S - 1 - st);
        for (u32 b = threadIdx.x; b < (total >> 1); b += NTT_THREADS) {
            u32 t = b & (T - 1), j = b >> log_T;
            u32 jj = j & (half - 1);
            u32 i0 = ((j - jj) << 1) + jj, i1 = i0 + half;
            u64 a = sm[(i0 << log_T) + t], c = sm[(i1 << log_T) + t];
            sm[(i0 << log_T) + t] = gl::add(a, c);
            sm[(i1 << log_T) + t] = gl::mul(gl::sub(a, c), roots[jj << st]);
        }
        __syncthreads();
    }
    for (u32 idx = threadIdx.x; idx < total; idx += NTT_THREADS) {
        u32 t = idx & (T - 1), pos = idx >> log_T;
        size_t off = ((size_t)pos << L) + low0 + t;
        d[off] = gl::mul(sm[idx], tw[off]);
    }
}

// Last digit on contiguous data.  smem: [G][2^S + 1].
enum { OUT_NATURAL = 0, OUT_ROWS = 1, OUT_COLS = 2 };
template <int MODE>
__global__ void __launch_bounds__(NTT_THREADS)
k_ntt_lo(const u64* __restrict__ src, size_t src_stride, int log_n, int S, int log_G, size_t n_cols,
         const u64* __restrict__ pre, const u64* __restrict__ post, u64 scale, const u64* __restrict__ roots,
         u64* __restrict__ dst, size_t dst_stride, size_t row_width, size_t row_base, size_t col0) {
    extern __shared__ u64 sm[];
    const int G = 1 << log_G;
    const u32 len = 1u << S, pitch = len + 1;
    const u32 total = (u32)G << S;
    const int log_nb = log_n - S;

    for (u32 idx = threadIdx.x; idx < total; idx += NTT_THREADS) {
        u32 g = idx >> S, e = idx & (len - 1);
        u64 v = 0;
        if (MODE != OUT_NATURAL) {
            size_t col = (size_t)blockIdx.y * G + g;
            size_t i = ((size_t)blockIdx.x << S) + e;
            if (col < n_cols) {
                v = gl::canon(src[col * src_stride + i]);
                if (pre) v = gl::mul(v, pre[i]);
            }
        } else {
            size_t blk = brev((u32)(blockIdx.x * G + g), log_nb);
            size_t i = (blk << S) + e;
            v = gl::canon(src[blockIdx.y * src_stride + i]);
            if (pre) v = gl::mul(v, pre[i]);
        }
        sm[g * pitch + e] = v;
    }
    __syncthreads();
    for (int st = 0; st < S; st++) {
        const u32 half = 1u << (S - 1 - st);
        for (u32 b = threadIdx.x; b < (total >> 1); b += NTT_THREADS) {
            u32 g = b >> (S - 1), j = b & ((len >> 1) - 1);
            u32 jj = j & (half - 1);
            u32 i0 = ((j - jj) << 1) + jj, i1 = i0 + half;
            u64* row = sm + g * pitch;
            u64 a = row[i0], c = row[i1];
            row[i0] = gl::add(a, c);
            row[i1] = gl::mul(gl::sub(a, c), roots[jj << st]);
        }
        __syncthreads();
    }
    if (MODE == OUT_COLS) { // column-major, index-bit-reversed: position q of this block is contiguous in its column
        for (u32 idx = threadIdx.x; idx < total; idx += NTT_THREADS) {
            u32 g = idx >> S, q = idx & (len - 1);
            size_t col = (size_t)blockIdx.y * G + g;
            if (col >= n_cols) continue;
            u64 v = sm[g * pitch + q];
            if (scale != 1) v = gl::mul(v, scale);
            dst[(col0 + col) * dst_stride + row_base + ((size_t)blockIdx.x << S) + q] = v;
        }
        return;
    }
    for (u32 idx = threadIdx.x; idx < total; idx += NTT_THREADS) {
        u32 g = idx & (G - 1), q = idx >> log_G;
        u64 v = sm[g * pitch + q];
        if (scale != 1) v = gl::mul(v, scale);
        if (MODE == OUT_ROWS) {
            size_t col = (size_t)blockIdx.y * G + g;
            if (col < n_cols) {
                size_t row = row_base + ((size_t)blockIdx.x << S) + q;
                dst[row * row_width + col0 + col] = v;
            }
        } else {
            size_t k = ((size_t)brev(q, S) << log_nb) + (size_t)blockIdx.x * G + g;
            if (post) v = gl::mul(v, post[k]);
            dst[blockIdx.y * dst_stride + k] = v;
        }
    }
}

u64* get_table(nlc_ctx* ctx, const TableKey& key, size_t words) {
    auto it = ctx->tables.find(key);
    if (it != ctx->tables.end()) return it->second;
    u64* p = nullptr;
    NLC_CUDA(cudaMalloc((void**)&p, (words ? words : 1) * sizeof(u64)));
    ctx->tables[key] = p;
    ctx->table_bytes += words * sizeof(u64);
    return nullptr; // caller fills, then looks it up again
}

u64* roots_table(nlc_ctx* ctx, int S, bool inv) {
    TableKey key{KIND_ROOTS, S, 0, inv ? 1 : 0, 0, 0};
    size_t words = S >= 1 ? ((size_t)1 << (S - 1)) : 1;
    if (u64* p = get_table(ctx, key, words)) return p;
    u64* p = ctx->tables[key];
    u64 w = gl::h_root_of_unity(S);
    if (inv) w = gl::h_inv(w);
    NLC_LAUNCH(ctx, k_pow_table, (unsigned)((words + 255) / 256), 256, 0, p, words, w, (u64)1);
    return p;
}
u64* tw_table(nlc_ctx* ctx, int m, int S, bool inv) {
    TableKey key{KIND_TW, m, S, inv ? 1 : 0, 0, 0};
    size_t words = (size_t)1 << m;
    if (u64* p = get_table(ctx, key, words)) return p;
    u64* p = ctx->tables[key];
    u64 w = gl::h_root_of_unity(m);
    if (inv) w = gl::h_inv(w);
    NLC_LAUNCH(ctx, k_tw_table, (unsigned)((words + 255) / 256), 256, 0, p, m, S, w);
    return p;
}
u64* pw_table(nlc_ctx* ctx, int log_n, u64 base, u64 scale) {
    TableKey key{KIND_PW, log_n, 0, 0, base, scale};
    size_t words = (size_t)1 << log_n;
    if (u64* p = get_table(ctx, key, words)) return p;
    u64* p = ctx->tables[key];
    NLC_LAUNCH(ctx, k_pow_table, (unsigned)((words + 255) / 256), 256, 0, p, words, base, scale);
    return p;
}

std::vector<int> plan_digits(int log_n) {
    if (log_n <= 10) return {log_n};
    int d = (log_n + 9) / 10;
    std::vector<int> v(d, log_n / d);
    for (int i = 0; i < log_n % d; i++) v[i]++;
    return v;
}

void ensure_attrs(nlc_ctx* ctx) {
    if (ctx->ntt_attrs) return;
    const int big = 160 * 1024;
    NLC_CUDA(cudaFuncSetAttribute(k_ntt_hi, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
    NLC_CUDA(cudaFuncSetAttribute(k_ntt_lo<OUT_NATURAL>, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
    NLC_CUDA(cudaFuncSetAttribute(k_ntt_lo<OUT_ROWS>, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
    NLC_CUDA(cudaFuncSetAttribute(k_ntt_lo<OUT_COLS>, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
    ctx->ntt_attrs = true;
}

} // namespace

size_t ntt_scratch_words(size_t n_cols, int log_n) {
    return plan_digits(log_n).size() > 1 ? n_cols << log_n : 0;
}

void ntt_run(nlc_ctx* ctx, const NttJob& job) {
    ensure_attrs(ctx);
    const int log_n = job.log_n;
    const size_t n = (size_t)1 << log_n;
    NLC_CHECK(log_n >= 0 && log_n <= 32, "log_n out of range (two-adicity is 32)");
    NLC_CHECK(job.n_cols > 0 && job.n_cols <= 65535u, "column count out of range");
    const std::vector<int> digits = plan_digits(log_n);
    const int d = (int)digits.size();
    const bool inv = job.inverse;
    u64 scale = 1;
    if (inv) scale = gl::h_inv(gl::canon((u64)(n % gl::P)));

    const u64* pre = nullptr;
    if (gl::canon(job.pre_shift) != 1) pre = pw_table(ctx, log_n, gl::canon(job.pre_shift), 1);
    const u64* post = nullptr;
    if (!job.rows && !job.cols && gl::canon(job.post_shift) != 1) { // fold 1/n into the post table
        post = pw_table(ctx, log_n, gl::canon(job.post_shift), scale);
        scale = 1;
    }

    const u64* cur = job.src;
    size_t cur_stride = job.src_stride;
    int m = log_n;
    for (int lvl = 0; lvl + 1 < d; lvl++) {
        NLC_CHECK(job.scratch != nullptr, "ntt_run: scratch required for multi-digit transform");
        const int S = digits[lvl], L = m - S;
        int log_T = S >= 9 ? 3 : 4;
        if (log_T > L) log_T = L;
        const u64* tw = tw_table(ctx, m, S, inv);
        const u64* roots = roots_table(ctx, S, inv);
        dim3 grid((unsigned)((n >> m) << (L - log_T)), (unsigned)job.n_cols);
        size_t smem = ((size_t)1 << (S + log_T)) * sizeof(u64);
        NLC_LAUNCH(ctx, k_ntt_hi, grid, NTT_THREADS, smem, cur, cur_stride, job.scratch, n, m, S, log_T,
                   lvl == 0 ? pre : nullptr, tw, roots);
        cur = job.scratch;
        cur_stride = n;
        m = L;
    }
    const int S = digits[d - 1];
    const u64* roots = roots_table(ctx, S, inv);
    const u64* lo_pre = d == 1 ? pre : nullptr;
    if (job.rows || job.cols) {
        int log_G = S >= 9 ? 3 : 4;
        if (job.cols) log_G = 0; // a column is contiguous: no need to gang columns for coalescing
        while (log_G > 0 && ((size_t)1 << (log_G - 1)) >= job.n_cols) log_G--;
        const size_t G = (size_t)1 << log_G;
        dim3 grid((unsigned)(n >> S), (unsigned)((job.n_cols + G - 1) / G));
        size_t smem = G * (((size_t)1 << S) + 1) * sizeof(u64);
        if (job.cols)
            NLC_LAUNCH(ctx, k_ntt_lo<OUT_COLS>, grid, NTT_THREADS, smem, cur, cur_stride, log_n, S, log_G, job.n_cols,
                       lo_pre, (const u64*)nullptr, scale, roots, job.dst, job.dst_stride, (size_t)0, job.row_base,
                       job.col0);
        else
            NLC_LAUNCH(ctx, k_ntt_lo<OUT_ROWS>, grid, NTT_THREADS, smem, cur, cur_stride, log_n, S, log_G, job.n_cols,
                       lo_pre, (const u64*)nullptr, scale, roots, job.dst, (size_t)0, job.row_width, job.row_base,
                       job.col0);
    } else {
        NLC_CHECK((const void*)cur != (const void*)job.dst, "ntt_run: NATURAL output must be out of place");
        int log_G = S >= 10 ? 3 : 4;
        if (log_G > log_n - S) log_G = log_n - S;
        dim3 grid((unsigned)((n >> S) >> log_G), (unsigned)job.n_cols);
        size_t smem = ((size_t)1 << log_G) * (((size_t)1 << S) + 1) * sizeof(u64);
        NLC_LAUNCH(ctx, k_ntt_lo<OUT_NATURAL>, grid, NTT_THREADS, smem, cur, cur_stride, log_n, S, log_G, job.n_cols, lo_pre,
                   post, scale, roots, job.dst, job.dst_stride, (size_t)0, (size_t)0, (size_t)0);
    }
}
