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
#ifndef NTT_MINB
#define NTT_MINB 3
#endif

__device__ __forceinline__ u32 brev(u32 x, int bits) { return bits ? (__brev(x) >> (32 - bits)) : 0; }

// out[i] = scale * base^i
__global__ void k_pow_table(u64* out, size_t n, u64 base, u64 scale) {
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i < n) out[i] = gl::mul(scale, gl::pow(base, i));
}
// tw[pos << L | i_low] = root^(bitrev_S(pos) * i_low), root = w_{2^m}^{+-1}
__global__ void k_tw_table(u64* out, int m, int S, u64 root, u64 scale) {
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= ((size_t)1 << m)) return;
    const int L = m - S;
    u64 pos = i >> L, low = i & (((u64)1 << L) - 1);
    out[i] = gl::mul(scale, gl::pow(root, (u64)brev((u32)pos, S) * low));
}

// ---- the 2^S-point transform of a tile held in shared memory ---------------------------------------------
// A decimation-in-frequency NTT in radix-2^R register steps (R <= 4).  With L = 2^(S-st) the length of the
// current sub-problems and stride = L / 2^R, one task loads x[k*stride + b], k < 2^R, runs the 2^R-point DIF
// in registers (twiddles w_{2^R}^e, e < 2^(R-1)), multiplies register k by the inter-step twiddle
// w_L^(b * bitrev_R(k)) and stores it back to slot k: the sub-problem of frequency class m = bitrev_R(k) then
// occupies the contiguous block k, which makes the final order the same index-bit-reversed order the radix-2
// cascade gives -- at (S/R - 1) full multiplications per element between steps instead of S/2.
// `roots` holds all 2^S powers of w_{2^S} (or its inverse).
// The twiddles INSIDE a register step are powers of w_{2^R} = w_16^(2^(4-R)), and plonky2's w_16 =
// primitive_root_of_unity(4) is 2^156 (2 has order 192 in F_p^*): multiplication by w_{2^R}^e is a shift by
// s = 156 e 2^(4-R) mod 192 bits (gl::mul_pow2), negated when s >= 96 (2^96 == -1) by swapping the subtraction.
template <int R, bool INV>
__device__ __forceinline__ void dft_regs(u64 (&v)[1 << R]) {
#pragma unroll
    for (int t = 0; t < R; t++) {
        const int half = 1 << (R - 1 - t);
#pragma unroll
        for (int k = 0; k < (1 << R); k++) {
            if (k & half) continue;
            const u64 a = v[k], c = v[k + half];
            v[k] = gl::dadd(a, c);
            const int e = ((k & (half - 1)) << t) << (4 - R); // exponent of w_16
            const int s = ((INV ? 192 - 156 : 156) * e) % 192;
            v[k + half] = s >= 96 ? gl::mul_pow2(gl::dsub(c, a), s - 96) : gl::mul_pow2(gl::dsub(a, c), s);
        }
    }
}

__device__ __forceinline__ u32 pad16(u32 i) { return i + (i >> 4); } // LO layout: one pad word per 16

// HI layout: element i of transform t at sm[i*T + t] (t fastest: what the strided gather delivers);
// LO layout: element i of transform g at sm[g*pitch + pad16(i)].
template <int R, bool HI, bool INV>
__device__ __forceinline__ void fft_step(u64* sm, int S, int st, int log_tr, u32 pitch, const u64* __restrict__ roots) {
    const int L_log = S - st, s_log = L_log - R;
    const u32 tasks = ((u32)1 << (S + log_tr)) >> R;
    for (u32 tau = threadIdx.x; tau < tasks; tau += NTT_THREADS) {
        u32 t, b, p;
        if (HI) {
            t = tau & (((u32)1 << log_tr) - 1);
            const u32 rest = tau >> log_tr;
            b = rest & (((u32)1 << s_log) - 1);
            p = rest >> s_log;
        } else {
            b = tau & (((u32)1 << s_log) - 1);
            const u32 rest = tau >> s_log;
            p = rest & (((u32)1 << st) - 1);
            t = rest >> st;
        }
        const u32 i0 = (p << L_log) + b;
        u64 v[1 << R];
#pragma unroll
        for (int k = 0; k < (1 << R); k++) {
            const u32 i = i0 + ((u32)k << s_log);
            v[k] = sm[HI ? (i << log_tr) + t : t * pitch + pad16(i)];
        }
        dft_regs<R, INV>(v);
        if (s_log > 0) {
#pragma unroll
            for (int k = 1; k < (1 << R); k++) {
                const u32 m = __brev((u32)k) >> (32 - R);
                v[k] = gl::mul(v[k], __ldg(roots + ((size_t)(b * m) << st)));
            }
        }
#pragma unroll
        for (int k = 0; k < (1 << R); k++) {
            const u32 i = i0 + ((u32)k << s_log);
            sm[HI ? (i << log_tr) + t : t * pitch + pad16(i)] = v[k];
        }
    }
}

// radix plan of a 2^S-point tile transform, packed 4 bits per step (host side: plan_steps)
template <bool HI, bool INV>
__device__ __forceinline__ void fft_tile(u64* sm, int S, u32 plan, int log_tr, u32 pitch, const u64* __restrict__ roots) {
    int st = 0;
    while (plan) {
        const int R = plan & 15;
        plan >>= 4;
        switch (R) {
        case 4: fft_step<4, HI, INV>(sm, S, st, log_tr, pitch, roots); break;
        case 3: fft_step<3, HI, INV>(sm, S, st, log_tr, pitch, roots); break;
        case 2: fft_step<2, HI, INV>(sm, S, st, log_tr, pitch, roots); break;
        default: fft_step<1, HI, INV>(sm, S, st, log_tr, pitch, roots); break;
        }
        st += R;
        __syncthreads();
    }
}

// One digit of a 2^m block, strided ("hi") pass.  smem: [2^S][T].
template <bool INV>
__global__ void __launch_bounds__(NTT_THREADS, NTT_MINB)
k_ntt_hi(const u64* __restrict__ src, size_t src_stride, u64* __restrict__ dst, size_t dst_stride, int m, int S,
         int log_T, u32 plan, const u64* __restrict__ pre, const u64* __restrict__ tw, const u64* __restrict__ roots) {
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

    _Pragma("unroll 4") for (u32 idx = threadIdx.x; idx < total; idx += NTT_THREADS) {
        u32 t = idx & (T - 1), ih = idx >> log_T;
        size_t off = ((size_t)ih << L) + low0 + t;
        u64 v = gl::canon(s[off]);
        if (pre) v = gl::mul(v, pre[base + off]);
        sm[idx] = v;
    }
    __syncthreads();
    fft_tile<true, INV>(sm, S, plan, log_T, 0, roots);
    _Pragma("unroll 4") for (u32 idx = threadIdx.x; idx < total; idx += NTT_THREADS) {
        u32 t = idx & (T - 1), pos = idx >> log_T;
        size_t off = ((size_t)pos << L) + low0 + t;
        d[off] = gl::mul(sm[idx], tw[off]);
    }
}

// Last digit on contiguous data.  A CTA takes Gb = 2^log_Gb blocks of 2^S elements in each of Gc = 2^log_Gc
// columns (transform g = gb*Gc + gc).  smem: [Gb*Gc][pitch].
//   COLS / ROWS : the Gb blocks are consecutive; NATURAL : Gc = 1 and the blocks are the bit-reversal-related
//   ones whose outputs interleave into contiguous Gb-element segments.
enum { OUT_NATURAL = 0, OUT_ROWS = 1, OUT_COLS = 2, OUT_SCATTER = 3 }; // SCATTER: COLS re-blocked by leaf-row shard
template <int MODE, bool INV>
__global__ void __launch_bounds__(NTT_THREADS, NTT_MINB)
k_ntt_lo(const u64* __restrict__ src, size_t src_stride, int log_n, int S, int log_Gb, int log_Gc, u32 plan,
         size_t n_cols, const u64* __restrict__ pre, const u64* __restrict__ post, u64 scale,
         const u64* __restrict__ roots, u64* __restrict__ dst, size_t dst_stride, size_t row_width, size_t row_base,
         size_t col0, const NttScatter sc) {
    extern __shared__ u64 sm[];
    const int log_G = log_Gb + log_Gc;
    const u32 len = 1u << S, pitch = pad16(len) + 1;
    const u32 total = (u32)1 << (S + log_G);
    const int log_nb = log_n - S;
    const u32 Gc_mask = (1u << log_Gc) - 1;

    _Pragma("unroll 4") for (u32 idx = threadIdx.x; idx < total; idx += NTT_THREADS) {
        const u32 g = idx >> S, e = idx & (len - 1);
        const size_t col = ((size_t)blockIdx.y << log_Gc) + (g & Gc_mask);
        size_t blk = ((size_t)blockIdx.x << log_Gb) + (g >> log_Gc);
        if (MODE == OUT_NATURAL) blk = brev((u32)blk, log_nb);
        const size_t i = (blk << S) + e;
        u64 v = 0;
        if (col < n_cols) {
            v = gl::canon(src[col * src_stride + i]);
            if (pre) v = gl::mul(v, pre[i]);
        }
        sm[g * pitch + pad16(e)] = v;
    }
    __syncthreads();
    fft_tile<false, INV>(sm, S, plan, log_G, pitch, roots);
    if (MODE == OUT_COLS || MODE == OUT_SCATTER) { // column-major, index-bit-reversed: the tile is contiguous in its column
        _Pragma("unroll 4") for (u32 idx = threadIdx.x; idx < total; idx += NTT_THREADS) {
            const u32 g = idx >> S, q = idx & (len - 1);
            const size_t col = ((size_t)blockIdx.y << log_Gc) + (g & Gc_mask);
            if (col >= n_cols) continue;
            const size_t blk = ((size_t)blockIdx.x << log_Gb) + (g >> log_Gc);
            u64 v = sm[g * pitch + pad16(q)];
            if (scale != 1) v = gl::mul(v, scale);
            const size_t leaf = row_base + (blk << S) + q;
            if (MODE == OUT_COLS) {
                dst[(col0 + col) * dst_stride + leaf] = v;
            } else { // re-blocked by leaf-row shard, possibly straight into a peer GPU's buffer
                const size_t gd = leaf >> sc.log_per;
                u64* base = sc.peer[0] ? sc.peer[gd] : dst + gd * sc.shard_stride;
                base[(col0 + col) * dst_stride + (leaf & (((size_t)1 << sc.log_per) - 1))] = v;
            }
        }
    } else if (MODE == OUT_ROWS) { // row-major: the Gc columns of a row are adjacent words
        _Pragma("unroll 4") for (u32 idx = threadIdx.x; idx < total; idx += NTT_THREADS) {
            const u32 gc = idx & Gc_mask, rest = idx >> log_Gc;
            const u32 q = rest & (len - 1), gb = rest >> S;
            const size_t col = ((size_t)blockIdx.y << log_Gc) + gc;
            if (col >= n_cols) continue;
            u64 v = sm[((gb << log_Gc) + gc) * pitch + pad16(q)];
            if (scale != 1) v = gl::mul(v, scale);
            const size_t row = row_base + ((((size_t)blockIdx.x << log_Gb) + gb) << S) + q;
            dst[row * row_width + col0 + col] = v;
        }
    } else { // natural order: output k = bitrev_S(q) * 2^log_nb + block index, contiguous over the Gb blocks
        const u32 G = 1u << log_G;
        _Pragma("unroll 4") for (u32 idx = threadIdx.x; idx < total; idx += NTT_THREADS) {
            const u32 g = idx & (G - 1), q = idx >> log_G;
            u64 v = sm[g * pitch + pad16(q)];
            if (scale != 1) v = gl::mul(v, scale);
            const size_t k = ((size_t)brev(q, S) << log_nb) + ((size_t)blockIdx.x << log_G) + g;
            if (post) v = gl::mul(v, post[k]);
            if (sc.n_bcast == 0) {
                dst[blockIdx.y * dst_stride + k] = v;
            } else {
                for (int p = 0; p < sc.n_bcast; p++) sc.peer[p][blockIdx.y * dst_stride + k] = v;
            }
        }
    }
}


// ---------------------------------------------------------------------------------------------------------------
// Fast path (digits of 4..8 bits = two register steps R1 + R2, 4096-element tiles, 256 threads).  Same pass
// structure and the same HBM layouts as the generic kernels above, but every index is a compile-time constant plus
// one per-thread base, and the data makes ONE shared-memory round trip per pass instead of three:
//   hi: global -> registers (the first register step works on exactly the elements a thread gathers: tile index
//       tau + k * TASKS) -> twiddle -> shared -> registers (second step) -> inter-digit twiddle -> global;
//   lo: global -> registers -> twiddle -> shared -> registers -> shared (in place: a thread owns its 2^R2 slots)
//       -> coalesced global stores.
// ---------------------------------------------------------------------------------------------------------------
constexpr u32 F_TILE_LOG = 12, F_TILE = 1u << F_TILE_LOG;

template <int R>
__device__ __forceinline__ constexpr u32 brev_c(u32 k) {
    u32 r = 0;
    for (int i = 0; i < R; i++) r |= ((k >> i) & 1u) << (R - 1 - i);
    return r;
}

template <int R1, int R2, bool INV>
__global__ void __launch_bounds__(NTT_THREADS, NTT_MINB)
k_ntt_hi_f(const u64* __restrict__ src, size_t src_stride, u64* __restrict__ dst, size_t dst_stride, int m,
           const u64* __restrict__ pre, const u64* __restrict__ tw, const u64* __restrict__ roots, int canon_in,
           const NttCosets cz) {
    constexpr int S = R1 + R2, LOG_T = F_TILE_LOG - S;
    constexpr u32 T = 1u << LOG_T, TASKS1 = F_TILE >> R1, TASKS2 = F_TILE >> R2;
    extern __shared__ u64 sm[]; // [2^S][T], t fastest
    const int L = m - S;
    // several transforms of the same columns in one launch (the 2^r cosets of an LDE): the coset index is the FASTEST
    // part of blockIdx.x, so the CTAs that read the same coefficient tile run together and share it through L2
    const size_t z = blockIdx.x & (((size_t)1 << cz.log_z) - 1);
    const size_t tile = blockIdx.x >> cz.log_z;
    const size_t base = (tile >> (L - LOG_T)) << m;
    const u32 low0 = ((u32)tile & ((1u << (L - LOG_T)) - 1)) << LOG_T;
    const u64* s = src + blockIdx.y * src_stride + z * cz.z_src + base + low0;
    u64* d = dst + blockIdx.y * dst_stride + z * cz.z_dst + base + low0;
    const u64* pr = pre ? pre + z * cz.z_pre + base + low0 : nullptr;
    const u64* twp = tw + low0;
    const size_t kstride = (size_t)1 << (L + R2);

#pragma unroll
    for (u32 j = 0; j < TASKS1 / NTT_THREADS; j++) {
        const u32 tau = threadIdx.x + j * NTT_THREADS;
        const u32 t = tau & (T - 1), b = tau >> LOG_T; // b < 2^R2
        const size_t off0 = ((size_t)b << L) + t;
        u64 v[1 << R1];
#pragma unroll
        for (int k = 0; k < (1 << R1); k++) v[k] = s[off0 + k * kstride];
        if (pr) {
#pragma unroll
            for (int k = 0; k < (1 << R1); k++) v[k] = gl::mul(v[k], __ldg(pr + off0 + k * kstride));
        } else if (canon_in) { // caller data may be any representative; later passes read their own canonical output
#pragma unroll
            for (int k = 0; k < (1 << R1); k++) v[k] = gl::dcanon(v[k]);
        }
        dft_regs<R1, INV>(v);
#pragma unroll
        for (int k = 1; k < (1 << R1); k++) v[k] = gl::mul(v[k], __ldg(roots + b * brev_c<R1>(k)));
#pragma unroll
        for (int k = 0; k < (1 << R1); k++) sm[tau + k * TASKS1] = v[k];
    }
    __syncthreads();
#pragma unroll
    for (u32 j = 0; j < TASKS2 / NTT_THREADS; j++) {
        const u32 tau = threadIdx.x + j * NTT_THREADS;
        const u32 t = tau & (T - 1), p = tau >> LOG_T; // p < 2^R1: which sub-problem
        const u32 i0 = (p << (R2 + LOG_T)) + t;
        u64 v[1 << R2];
#pragma unroll
        for (int k = 0; k < (1 << R2); k++) v[k] = sm[i0 + (k << LOG_T)];
        dft_regs<R2, INV>(v);
        const size_t off0 = ((size_t)(p << R2) << L) + t;
        const size_t pstride = (size_t)1 << L;
#pragma unroll
        for (int k = 0; k < (1 << R2); k++) d[off0 + k * pstride] = gl::mul(v[k], __ldg(twp + off0 + k * pstride));
    }
}

// lo-pass shared layout: transform g at sm[g * pitch + pidx(e)], two pad words per 16 (keeps 16-byte alignment of
// every 16-element run and makes both the strided and the contiguous register steps conflict-free)
template <int S>
struct LoGeom {
    static constexpr u32 pitch = (1u << S) + (S >= 4 ? (1u << (S - 3)) : 0) + 2;
    __device__ __forceinline__ static u32 pidx(u32 e) { return e + 2 * (e >> 4); }
};

template <int R1, int R2, int MODE, bool INV>
__global__ void __launch_bounds__(NTT_THREADS, NTT_MINB)
k_ntt_lo_f(const u64* __restrict__ src, size_t src_stride, int log_n, const u64* __restrict__ post, u64 scale,
           const u64* __restrict__ roots, u64* __restrict__ dst, size_t dst_stride, size_t row_base, size_t col0,
           const NttScatter sc, const NttCosets cz) {
    constexpr int S = R1 + R2, LOG_G = F_TILE_LOG - S;
    constexpr u32 G = 1u << LOG_G, LEN = 1u << S, TASKS1 = F_TILE >> R1, TASKS2 = F_TILE >> R2;
    constexpr u32 pitch = LoGeom<S>::pitch;
    extern __shared__ u64 sm[];
    const int log_nb = log_n - S;
    const size_t z = blockIdx.x & (((size_t)1 << cz.log_z) - 1); // coset of a batched LDE (COLS only; 0 otherwise)
    const size_t bx = blockIdx.x >> cz.log_z;
    const u64* s = src + blockIdx.y * src_stride + z * cz.z_src;
    row_base += z * cz.z_dst;

#pragma unroll
    for (u32 j = 0; j < TASKS1 / NTT_THREADS; j++) {
        const u32 tau = threadIdx.x + j * NTT_THREADS;
        const u32 b = tau & ((1u << R2) - 1), g = tau >> R2;
        size_t blk = (bx << LOG_G) + g;
        if (MODE == OUT_NATURAL) blk = brev((u32)blk, log_nb);
        const u64* in = s + (blk << S) + b;
        u64 v[1 << R1];
#pragma unroll
        for (int k = 0; k < (1 << R1); k++) v[k] = in[k << R2]; // an earlier pass of this transform wrote it: canonical
        dft_regs<R1, INV>(v);
#pragma unroll
        for (int k = 1; k < (1 << R1); k++) v[k] = gl::mul(v[k], __ldg(roots + b * brev_c<R1>(k)));
#pragma unroll
        for (int k = 0; k < (1 << R1); k++) sm[g * pitch + LoGeom<S>::pidx(b + (k << R2))] = v[k];
    }
    __syncthreads();
#pragma unroll
    for (u32 j = 0; j < TASKS2 / NTT_THREADS; j++) {
        const u32 tau = threadIdx.x + j * NTT_THREADS;
        const u32 p = tau & ((1u << R1) - 1), g = tau >> R1;
        u64* row = sm + g * pitch;
        u64 v[1 << R2];
#pragma unroll
        for (int k = 0; k < (1 << R2); k++) v[k] = row[LoGeom<S>::pidx((p << R2) + k)];
        dft_regs<R2, INV>(v);
#pragma unroll
        for (int k = 0; k < (1 << R2); k++) row[LoGeom<S>::pidx((p << R2) + k)] = v[k]; // in place: the thread owns these slots
    }
    __syncthreads();
    if (MODE == OUT_COLS || MODE == OUT_SCATTER) {
        u64* out = dst + (col0 + blockIdx.y) * dst_stride;
#pragma unroll
        for (u32 j = 0; j < F_TILE / NTT_THREADS; j++) {
            const u32 idx = threadIdx.x + j * NTT_THREADS;
            const u32 g = idx >> S, q = idx & (LEN - 1);
            u64 v = sm[g * pitch + LoGeom<S>::pidx(q)];
            if (scale != 1) v = gl::mul(v, scale);
            const size_t leaf = row_base + (((bx << LOG_G) + g) << S) + q;
            if (MODE == OUT_COLS) {
                out[leaf] = v;
            } else { // re-blocked by leaf-row shard, possibly straight into a peer GPU's buffer
                const size_t gd = leaf >> sc.log_per;
                u64* pbase = sc.peer[0] ? sc.peer[gd] : dst + gd * sc.shard_stride;
                pbase[(col0 + blockIdx.y) * dst_stride + (leaf & (((size_t)1 << sc.log_per) - 1))] = v;
            }
        }
    } else { // natural order: output k = bitrev_S(q) * 2^log_nb + block index, contiguous over the G blocks
        u64* out = dst + blockIdx.y * dst_stride + (bx << LOG_G);
#pragma unroll
        for (u32 j = 0; j < F_TILE / NTT_THREADS; j++) {
            const u32 idx = threadIdx.x + j * NTT_THREADS;
            const u32 g = idx & (G - 1), q = idx >> LOG_G;
            u64 v = sm[g * pitch + LoGeom<S>::pidx(q)];
            if (scale != 1) v = gl::mul(v, scale);
            const size_t k = ((size_t)brev(q, S) << log_nb) + g;
            if (post) v = gl::mul(v, __ldg(post + (bx << LOG_G) + k));
            if (sc.n_bcast == 0) {
                out[k] = v;
            } else { // every rank's coefficient matrix receives this rank's columns (local store + NVLink stores)
                const size_t off = blockIdx.y * dst_stride + (bx << LOG_G) + k;
                for (int p = 0; p < sc.n_bcast; p++) sc.peer[p][off] = v;
            }
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
    size_t words = (size_t)1 << S; // all powers: the inter-step twiddles of fft_step index the whole circle
    if (u64* p = get_table(ctx, key, words)) return p;
    u64* p = ctx->tables[key];
    u64 w = gl::h_root_of_unity(S);
    if (inv) w = gl::h_inv(w);
    NLC_LAUNCH(ctx, k_pow_table, (unsigned)((words + 255) / 256), 256, 0, p, words, w, (u64)1);
    return p;
}
// scale != 1: the 1/n of an inverse transform rides on the first inter-digit twiddle (every element meets it once)
u64* tw_table(nlc_ctx* ctx, int m, int S, bool inv, u64 scale = 1) {
    TableKey key{KIND_TW, m, S, inv ? 1 : 0, 0, scale};
    size_t words = (size_t)1 << m;
    if (u64* p = get_table(ctx, key, words)) return p;
    u64* p = ctx->tables[key];
    u64 w = gl::h_root_of_unity(m);
    if (inv) w = gl::h_inv(w);
    NLC_LAUNCH(ctx, k_tw_table, (unsigned)((words + 255) / 256), 256, 0, p, m, S, w, scale);
    return p;
}
// Power tables are keyed by a shift the CALLER may choose (nlc_coset_fft / nlc_coset_ifft take any shift): they are kept
// least-recently-used within a byte budget, so a caller that walks through many shifts cannot grow the context without
// bound.  The library's own shifts (7, 7 w_N^k, 7^-1, the FRI round shifts) are a fixed small set that fits the budget and
// therefore stays resident; an evicted table is simply rebuilt on its next use.  cudaFree synchronises the device, so no
// kernel in flight can still be reading an evicted table.
constexpr size_t PW_BUDGET_BYTES = (size_t)1 << 30;
u64* pw_table(nlc_ctx* ctx, int log_n, u64 base, u64 scale) {
    TableKey key{KIND_PW, log_n, 0, 0, base, scale};
    size_t words = (size_t)1 << log_n;
    auto& lru = ctx->pw_lru;
    for (size_t i = 0; i < lru.size(); i++)
        if (lru[i] == key) { // hit: most recently used goes to the back
            lru.erase(lru.begin() + i);
            lru.push_back(key);
            return ctx->tables[key];
        }
    while (!lru.empty() && ctx->pw_bytes + words * sizeof(u64) > PW_BUDGET_BYTES) {
        const TableKey old = lru.front();
        lru.erase(lru.begin());
        auto it = ctx->tables.find(old);
        const size_t old_bytes = sizeof(u64) << std::get<1>(old);
        cudaFree(it->second);
        ctx->tables.erase(it);
        ctx->pw_bytes -= old_bytes;
        ctx->table_bytes -= old_bytes;
    }
    get_table(ctx, key, words);
    u64* p = ctx->tables[key];
    lru.push_back(key);
    ctx->pw_bytes += words * sizeof(u64);
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

// register-radix steps of a 2^S-point tile transform, 4 bits each, first step in the low nibble
u32 plan_steps(int S) {
    static const u32 plans[11] = {0, 0x1, 0x2, 0x3, 0x4, 0x23, 0x33, 0x34, 0x44, 0x333, 0x334};
    return plans[S];
}

void ensure_attrs(nlc_ctx* ctx) {
    if (ctx->ntt_attrs) return;
    const int big = 160 * 1024;
    NLC_CUDA(cudaFuncSetAttribute(k_ntt_hi<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
    NLC_CUDA(cudaFuncSetAttribute(k_ntt_hi<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
    NLC_CUDA(cudaFuncSetAttribute((k_ntt_lo<OUT_NATURAL, false>), cudaFuncAttributeMaxDynamicSharedMemorySize, big));
    NLC_CUDA(cudaFuncSetAttribute((k_ntt_lo<OUT_NATURAL, true>), cudaFuncAttributeMaxDynamicSharedMemorySize, big));
    NLC_CUDA(cudaFuncSetAttribute((k_ntt_lo<OUT_ROWS, false>), cudaFuncAttributeMaxDynamicSharedMemorySize, big));
    NLC_CUDA(cudaFuncSetAttribute((k_ntt_lo<OUT_ROWS, true>), cudaFuncAttributeMaxDynamicSharedMemorySize, big));
    NLC_CUDA(cudaFuncSetAttribute((k_ntt_lo<OUT_COLS, false>), cudaFuncAttributeMaxDynamicSharedMemorySize, big));
    NLC_CUDA(cudaFuncSetAttribute((k_ntt_lo<OUT_COLS, true>), cudaFuncAttributeMaxDynamicSharedMemorySize, big));
    NLC_CUDA(cudaFuncSetAttribute((k_ntt_lo<OUT_SCATTER, false>), cudaFuncAttributeMaxDynamicSharedMemorySize, big));
    ctx->ntt_attrs = true;
}

} // namespace

size_t ntt_scratch_words(size_t n_cols, int log_n) {
    return plan_digits(log_n).size() > 1 ? n_cols << log_n : 0;
}

// ---- fast path dispatch ----
namespace {
std::vector<int> plan_digits_fast(int log_n) { // digits of 4..8 bits, smaller ones first (the last digit is >= 6)
    const int d = (log_n + 7) / 8;
    std::vector<int> v(d, log_n / d);
    for (int i = 0; i < log_n % d; i++) v[d - 1 - i]++;
    return v;
}

template <bool INV>
void launch_hi_f(nlc_ctx* ctx, int S, dim3 grid, const u64* src, size_t src_stride, u64* dst, size_t dst_stride, int m,
                 const u64* pre, const u64* tw, const u64* roots, int canon_in, const NttCosets& cz) {
    const size_t smem = (size_t)F_TILE * sizeof(u64);
    switch (S) {
    case 4: NLC_LAUNCH(ctx, (k_ntt_hi_f<2, 2, INV>), grid, NTT_THREADS, smem, src, src_stride, dst, dst_stride, m, pre, tw, roots, canon_in, cz); break;
    case 5: NLC_LAUNCH(ctx, (k_ntt_hi_f<3, 2, INV>), grid, NTT_THREADS, smem, src, src_stride, dst, dst_stride, m, pre, tw, roots, canon_in, cz); break;
    case 6: NLC_LAUNCH(ctx, (k_ntt_hi_f<3, 3, INV>), grid, NTT_THREADS, smem, src, src_stride, dst, dst_stride, m, pre, tw, roots, canon_in, cz); break;
    case 7: NLC_LAUNCH(ctx, (k_ntt_hi_f<4, 3, INV>), grid, NTT_THREADS, smem, src, src_stride, dst, dst_stride, m, pre, tw, roots, canon_in, cz); break;
    default: NLC_LAUNCH(ctx, (k_ntt_hi_f<4, 4, INV>), grid, NTT_THREADS, smem, src, src_stride, dst, dst_stride, m, pre, tw, roots, canon_in, cz); break;
    }
}
template <int MODE, bool INV>
void launch_lo_f(nlc_ctx* ctx, int S, dim3 grid, const u64* src, size_t src_stride, int log_n, const u64* post, u64 scale,
                 const u64* roots, u64* dst, size_t dst_stride, size_t row_base, size_t col0, const NttScatter& sc,
                 const NttCosets& cz = NttCosets()) {
    switch (S) {
    case 6: NLC_LAUNCH(ctx, (k_ntt_lo_f<3, 3, MODE, INV>), grid, NTT_THREADS, (size_t)(F_TILE >> 6) * LoGeom<6>::pitch * 8, src, src_stride, log_n, post, scale, roots, dst, dst_stride, row_base, col0, sc, cz); break;
    case 7: NLC_LAUNCH(ctx, (k_ntt_lo_f<4, 3, MODE, INV>), grid, NTT_THREADS, (size_t)(F_TILE >> 7) * LoGeom<7>::pitch * 8, src, src_stride, log_n, post, scale, roots, dst, dst_stride, row_base, col0, sc, cz); break;
    default: NLC_LAUNCH(ctx, (k_ntt_lo_f<4, 4, MODE, INV>), grid, NTT_THREADS, (size_t)(F_TILE >> 8) * LoGeom<8>::pitch * 8, src, src_stride, log_n, post, scale, roots, dst, dst_stride, row_base, col0, sc, cz); break;
    }
}

void ntt_run_fast(nlc_ctx* ctx, const NttJob& job) {
    const int log_n = job.log_n;
    const size_t n = (size_t)1 << log_n;
    const std::vector<int> digits = plan_digits_fast(log_n);
    const int d = (int)digits.size();
    const bool inv = job.inverse;
    u64 scale = inv ? gl::h_inv(gl::canon((u64)(n % gl::P))) : 1;
    const bool batched = job.log_cosets > 0 || job.pre_tables;
    if (batched)
        NLC_CHECK(job.cols && !inv && job.scatter.log_per < 0 && job.pre_tables, "coset batches: forward COLS transforms only");
    const u64* pre = nullptr;
    if (batched) pre = job.pre_tables;
    else if (gl::canon(job.pre_shift) != 1) pre = pw_table(ctx, log_n, gl::canon(job.pre_shift), 1);
    const u64* post = nullptr;
    if (!job.cols && gl::canon(job.post_shift) != 1) post = pw_table(ctx, log_n, gl::canon(job.post_shift), 1);
    NLC_CHECK(batched || job.scratch != nullptr, "ntt_run: scratch required for multi-digit transform");

    // a coset batch works in place on its destination columns (the lo pass reads a tile whole before it writes it)
    u64* work = batched ? job.dst + job.col0 * job.dst_stride + job.row_base : job.scratch;
    const size_t work_stride = batched ? job.dst_stride : n;
    NttCosets cz;
    cz.log_z = job.log_cosets;
    cz.z_dst = n;
    cz.z_pre = n;
    const u64* cur = job.src;
    size_t cur_stride = job.src_stride;
    int m = log_n;
    for (int lvl = 0; lvl + 1 < d; lvl++) {
        const int S = digits[lvl], L = m - S, log_T = (int)F_TILE_LOG - S;
        NLC_CHECK(L >= log_T, "ntt fast path: tile wider than the stride");
        const u64* tw = tw_table(ctx, m, S, inv, lvl == 0 ? scale : 1); // 1/n folded into the first inter-digit twiddle
        const u64* roots = roots_table(ctx, S, inv);
        dim3 grid((unsigned)(((n >> m) << (L - log_T)) << cz.log_z), (unsigned)job.n_cols);
        cz.z_src = lvl == 0 ? 0 : n; // every coset reads the same coefficients, then its own block of the destination
        if (inv)
            launch_hi_f<true>(ctx, S, grid, cur, cur_stride, work, work_stride, m, lvl == 0 ? pre : nullptr, tw, roots, lvl == 0, cz);
        else
            launch_hi_f<false>(ctx, S, grid, cur, cur_stride, work, work_stride, m, lvl == 0 ? pre : nullptr, tw, roots, lvl == 0, cz);
        cur = work;
        cur_stride = work_stride;
        m = L;
    }
    scale = 1;
    const int S = digits[d - 1];
    const u64* roots = roots_table(ctx, S, inv);
    dim3 grid((unsigned)((n >> F_TILE_LOG) << cz.log_z), (unsigned)job.n_cols);
    cz.z_src = n;
    if (job.cols && job.scatter.log_per >= 0) {
        NLC_CHECK(!inv, "scatter output is for forward transforms");
        launch_lo_f<OUT_SCATTER, false>(ctx, S, grid, cur, cur_stride, log_n, nullptr, scale, roots, job.dst, job.dst_stride,
                                        job.row_base, job.col0, job.scatter);
    } else if (job.cols) {
        if (inv)
            launch_lo_f<OUT_COLS, true>(ctx, S, grid, cur, cur_stride, log_n, nullptr, scale, roots, job.dst, job.dst_stride,
                                        job.row_base, job.col0, job.scatter);
        else
            launch_lo_f<OUT_COLS, false>(ctx, S, grid, cur, cur_stride, log_n, nullptr, scale, roots, job.dst,
                                         job.dst_stride, job.row_base, job.col0, job.scatter, batched ? cz : NttCosets());
    } else {
        NLC_CHECK((const void*)cur != (const void*)job.dst || job.scatter.n_bcast, "ntt_run: NATURAL output must be out of place");
        if (inv)
            launch_lo_f<OUT_NATURAL, true>(ctx, S, grid, cur, cur_stride, log_n, post, scale, roots, job.dst, job.dst_stride,
                                           0, 0, job.scatter);
        else
            launch_lo_f<OUT_NATURAL, false>(ctx, S, grid, cur, cur_stride, log_n, post, scale, roots, job.dst,
                                            job.dst_stride, 0, 0, job.scatter);
    }
}

// power tables of the cosets [k0, k0 + count) of an LDE in LEAF order: table z = powers of 7 * w_N^bitrev_r(k0 + z)
// (coset k of the domain 7<w_N> is the leaf range [bitrev_r(k) n, +n): SURVEY.md 8e.2), back to back; cached
enum { KIND_PWC = 4 };
} // namespace

const u64* ntt_coset_tables(nlc_ctx* ctx, int log_n, int rate_bits, size_t kr0, size_t count) {
    const size_t n = (size_t)1 << log_n;
    TableKey key{KIND_PWC, log_n, rate_bits, (int)kr0, (u64)count, 0};
    if (u64* p = get_table(ctx, key, count * n)) return p;
    u64* p = ctx->tables[key];
    const u64 w_N = gl::h_root_of_unity(log_n + rate_bits);
    for (size_t z = 0; z < count; z++) {
        size_t k = 0;
        for (int t = 0; t < rate_bits; t++) k |= (((kr0 + z) >> t) & 1) << (rate_bits - 1 - t);
        NLC_LAUNCH(ctx, k_pow_table, (unsigned)((n + 255) / 256), 256, 0, p + z * n, n,
                   gl::h_mul(gl::GENERATOR, gl::h_pow(w_N, k)), (u64)1);
    }
    return p;
}
bool ntt_can_batch_cosets(int log_n) { return log_n >= 12; }

namespace {
} // namespace

void ntt_run(nlc_ctx* ctx, const NttJob& job) {
    NLC_CHECK(job.log_n >= 0 && job.log_n <= 32, "log_n out of range (two-adicity is 32)");
    NLC_CHECK(job.n_cols > 0 && job.n_cols <= 65535u, "column count out of range");
    if (job.log_n >= 12 && !job.rows) return ntt_run_fast(ctx, job);
    NLC_CHECK(job.log_cosets == 0 && !job.pre_tables, "coset batches need the fast path (log_n >= 12)");
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
        int log_T = 12 - S; // 4096-element tiles: one radix-16 task per thread
        if (log_T < 3) log_T = 3;
        if (log_T > L) log_T = L;
        const u64* tw = tw_table(ctx, m, S, inv);
        const u64* roots = roots_table(ctx, S, inv);
        dim3 grid((unsigned)((n >> m) << (L - log_T)), (unsigned)job.n_cols);
        size_t smem = ((size_t)1 << (S + log_T)) * sizeof(u64);
        auto k_hi = inv ? k_ntt_hi<true> : k_ntt_hi<false>;
        NLC_LAUNCH(ctx, k_hi, grid, NTT_THREADS, smem, cur, cur_stride, job.scratch, n, m, S, log_T, plan_steps(S),
                   lvl == 0 ? pre : nullptr, tw, roots);
        cur = job.scratch;
        cur_stride = n;
        m = L;
    }
    const int S = digits[d - 1];
    const u64* roots = roots_table(ctx, S, inv);
    const u64* lo_pre = d == 1 ? pre : nullptr;
    const int log_nb = log_n - S;
    int log_G = 12 - S; // transforms per CTA
    if (log_G < 2) log_G = 2;
    const size_t pitch = ((size_t)1 << S) + ((size_t)1 << S >> 4) + 1;
    if (job.rows || job.cols) {
        int log_Gc = 0;
        if (job.rows) // gang columns so a row segment is one store (FRI: the two planes of an extension element)
            while (log_Gc < log_G && ((size_t)1 << log_Gc) < job.n_cols) log_Gc++;
        int log_Gb = log_G - log_Gc;
        if (log_Gb > log_nb) log_Gb = log_nb;
        dim3 grid((unsigned)((n >> S) >> log_Gb), (unsigned)((job.n_cols + ((size_t)1 << log_Gc) - 1) >> log_Gc));
        size_t smem = ((size_t)1 << (log_Gb + log_Gc)) * pitch * sizeof(u64);
        if (job.cols && job.scatter.log_per >= 0) {
            NLC_CHECK(!inv, "scatter output is for forward transforms"); // the shard is picked per element: tiles may span shards
            NLC_LAUNCH(ctx, (k_ntt_lo<OUT_SCATTER, false>), grid, NTT_THREADS, smem, cur, cur_stride, log_n, S, log_Gb,
                       log_Gc, plan_steps(S), job.n_cols, lo_pre, (const u64*)nullptr, scale, roots, job.dst,
                       job.dst_stride, (size_t)0, job.row_base, job.col0, job.scatter);
        } else if (job.cols) {
            auto k_lo = inv ? k_ntt_lo<OUT_COLS, true> : k_ntt_lo<OUT_COLS, false>;
            NLC_LAUNCH(ctx, k_lo, grid, NTT_THREADS, smem, cur, cur_stride, log_n, S, log_Gb, log_Gc,
                       plan_steps(S), job.n_cols, lo_pre, (const u64*)nullptr, scale, roots, job.dst, job.dst_stride,
                       (size_t)0, job.row_base, job.col0, job.scatter);
        } else {
            auto k_lo = inv ? k_ntt_lo<OUT_ROWS, true> : k_ntt_lo<OUT_ROWS, false>;
            NLC_LAUNCH(ctx, k_lo, grid, NTT_THREADS, smem, cur, cur_stride, log_n, S, log_Gb, log_Gc,
                       plan_steps(S), job.n_cols, lo_pre, (const u64*)nullptr, scale, roots, job.dst, (size_t)0,
                       job.row_width, job.row_base, job.col0, NttScatter());
        }
    } else {
        NLC_CHECK((const void*)cur != (const void*)job.dst || job.scatter.n_bcast, "ntt_run: NATURAL output must be out of place");
        if (log_G > log_nb) log_G = log_nb;
        dim3 grid((unsigned)((n >> S) >> log_G), (unsigned)job.n_cols);
        size_t smem = ((size_t)1 << log_G) * pitch * sizeof(u64);
        auto k_lo = inv ? k_ntt_lo<OUT_NATURAL, true> : k_ntt_lo<OUT_NATURAL, false>;
        NLC_LAUNCH(ctx, k_lo, grid, NTT_THREADS, smem, cur, cur_stride, log_n, S, log_G, 0,
                   plan_steps(S), job.n_cols, lo_pre, post, scale, roots, job.dst, job.dst_stride, (size_t)0,
                   (size_t)0, (size_t)0, job.scatter);
    }
}
