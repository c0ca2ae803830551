// merkle.cu -- Poseidon leaf hashing and level-by-level Merkle digests for sm_100a.
//
// Replaces plonky2::hash::merkle_tree::{MerkleTree::new, fill_digests_buf, fill_subtree,
// MerkleTree::prove} and PoseidonHash::{hash_no_pad, hash_or_noop, two_to_one}
// (plonky2 0.1.4 @ d2598bd), whose result is read in tree at
// /root/reference/vendor/starkyx/starkyx/src/plonky2/stark/prover.rs:95,180,210 (merkle_tree.cap).
//
// Digests are written straight into plonky2's interleaved `digests` layout, so the host struct can
// be filled by one D2H copy and MerkleTree::prove's index arithmetic works unchanged:
//   inside the chunk of cap subtree s, the sibling pair q of layer i (0 = leaf digests) sits at
//   2*((q << (i+1)) + (1<<i) - 1) + {0,1};  the subtree root goes to cap[s].
#include "common.cuh"
#include "poseidon.cuh"
#include "transcript.cuh"

namespace {

constexpr int HASH_THREADS = 128;
// the limb-resident partial rounds want ~96 registers: 5 CTAs of 128 threads per SM (measured best, tools/ubench/ubench_perm.cu;
// throughput is flat from 4 to 8 CTAs/SM: the integer pipes, not latency, are the limit)
constexpr int HASH_MINB = 5;

struct TreeGeom {
    int sub_height;       // log2(leaves per cap subtree)
    size_t sub_digests;   // digests per subtree = 2*(2^sub_height - 1)
};

// slot of node `j` (global index within its layer) of layer `i`; layer sub_height = cap.
__device__ __forceinline__ u64* node_slot(u64* digests, u64* cap, TreeGeom g, int layer, size_t j) {
    const int rem = g.sub_height - layer; // log2(nodes of this layer per subtree)
    const size_t s = j >> rem;
    if (rem == 0) return cap + s * 4;
    const size_t jl = j & (((size_t)1 << rem) - 1);
    const size_t q = jl >> 1;
    const size_t idx = 2 * ((q << (layer + 1)) + ((size_t)1 << layer) - 1) + (jl & 1);
    return digests + (s * g.sub_digests + idx) * 4;
}

// hash_or_noop of each leaf (one thread per leaf), written to its layer-0 slot.  Element j of leaf k is
// leaves[k * leaf_stride + j * elem_stride]:
//   COLS  (leaf_stride 1, elem_stride N): the column-major LDE of a PolynomialBatch, exactly as the NTT left it --
//         the 8 loads of a sponge block are 8 fully coalesced 256-byte warp requests, nothing is transposed;
//   !COLS (leaf_stride leaf_len, elem_stride 1): row-major leaves (FRI round trees, nlc_merkle_new).
// Leaves are dealt to warps in units of 32 consecutive leaves.  A warp starts on unit = its global warp index and then
// claims the next unit from a device counter (one atomicAdd per warp per unit, i.e. per >= 32 x ceil(W/8) permutations):
// with a static grid-stride deal the last round ran on the CTAs with the lowest indices only, 3 of them on some SMs and
// 2 on others, and the SMs with 2 idled for the difference (ncu: SMs active 95.8 % of the kernel); claimed units even
// that out and let warps drift apart, so that they do not all sit in the S-box (multiplier) phase at the same time.
__device__ __forceinline__ size_t next_unit(u32* __restrict__ ctr, u32 n_warps) {
    u32 nx = 0;
    if ((threadIdx.x & 31) == 0) nx = atomicAdd(ctr, 1u);
    return (size_t)n_warps + __shfl_sync(0xffffffffu, nx, 0);
}

template <bool COLS>
__global__ void __launch_bounds__(HASH_THREADS, HASH_MINB)
k_leaf_hash(const u64* __restrict__ leaves, size_t n_leaves, size_t leaf_len, size_t leaf_stride, size_t elem_stride,
            TreeGeom g, u64* digests, u64* cap, u32* __restrict__ ctr) {
    const u32 n_warps = gridDim.x * (HASH_THREADS / 32);
    const size_t n_units = (n_leaves + 31) >> 5;
    for (size_t unit = blockIdx.x * (size_t)(HASH_THREADS / 32) + (threadIdx.x >> 5); unit < n_units;
         unit = next_unit(ctr, n_warps)) {
        const size_t row = (unit << 5) + (threadIdx.x & 31);
        if (row < n_leaves) { // whole warps reach next_unit together
            const u64* in = leaves + row * (COLS ? 1 : leaf_stride);
            const size_t es = COLS ? elem_stride : 1;
            u64 s[poseidon::WIDTH];
#pragma unroll
            for (int i = 0; i < poseidon::WIDTH; i++) s[i] = 0;
            if (leaf_len <= 4) { // hash_or_noop: short rows are copied, zero padded
#pragma unroll
                for (int i = 0; i < 4; i++)
                    if ((size_t)i < leaf_len) s[i] = gl::canon(in[i * es]);
            } else {
                // one call site for the permutation (its two round bodies are ~29 KB of code: a second inlined copy
                // for the ragged last block would not fit the instruction cache beside the first)
#pragma unroll 1
                for (size_t off = 0; off < leaf_len; off += poseidon::RATE) {
                    const size_t rem = leaf_len - off;
#pragma unroll
                    for (int i = 0; i < poseidon::RATE; i++)
                        if ((size_t)i < rem) s[i] = in[(off + i) * es]; // overwrite mode
                    poseidon::permute_nc(s);
                }
#pragma unroll
                for (int i = 0; i < 4; i++) s[i] = gl::canon(s[i]);
            }
            u64* out = node_slot(digests, cap, g, 0, row);
            reinterpret_cast<ulonglong2*>(out)[0] = make_ulonglong2(s[0], s[1]);
            reinterpret_cast<ulonglong2*>(out)[1] = make_ulonglong2(s[2], s[3]);
        }
    }
}

// The same sponge, fed a column range at a time (host-fed commits: columns [col0, col0 + n_cols) have just been
// extended while later columns are still crossing PCIe).  Every chunk but the last covers whole 8-element blocks;
// between chunks the 12-word sponge state of each leaf waits in `state` (state[i * n_leaves + row]: coalesced).
// Same permutations, same order, same result as k_leaf_hash<true> over all columns (leaf_len > 4 only).
__global__ void __launch_bounds__(HASH_THREADS, HASH_MINB)
k_leaf_absorb(const u64* __restrict__ lde, size_t n_leaves, size_t elem_stride, size_t col0, size_t n_cols, bool first,
              bool last, u64* __restrict__ state, TreeGeom g, u64* digests, u64* cap, u32* __restrict__ ctr) {
    const u32 n_warps = gridDim.x * (HASH_THREADS / 32);
    const size_t n_units = (n_leaves + 31) >> 5;
    for (size_t unit = blockIdx.x * (size_t)(HASH_THREADS / 32) + (threadIdx.x >> 5); unit < n_units;
         unit = next_unit(ctr, n_warps)) {
        const size_t row = (unit << 5) + (threadIdx.x & 31);
        if (row < n_leaves) {
            const u64* in = lde + col0 * elem_stride + row;
            u64 s[poseidon::WIDTH];
#pragma unroll
            for (int i = 0; i < poseidon::WIDTH; i++) s[i] = first ? 0 : state[i * n_leaves + row];
#pragma unroll 1
            for (size_t off = 0; off < n_cols; off += poseidon::RATE) { // a ragged block only ends the last chunk
                const size_t rem = n_cols - off;
#pragma unroll
                for (int i = 0; i < poseidon::RATE; i++)
                    if ((size_t)i < rem) s[i] = in[(off + i) * elem_stride];
                poseidon::permute_nc(s);
            }
            if (last) {
                u64* out = node_slot(digests, cap, g, 0, row);
                reinterpret_cast<ulonglong2*>(out)[0] = make_ulonglong2(gl::canon(s[0]), gl::canon(s[1]));
                reinterpret_cast<ulonglong2*>(out)[1] = make_ulonglong2(gl::canon(s[2]), gl::canon(s[3]));
            } else {
#pragma unroll
                for (int i = 0; i < poseidon::WIDTH; i++) state[i * n_leaves + row] = s[i];
            }
        }
    }
}

// layer (i) -> layer (i+1): one thread per parent.
__global__ void __launch_bounds__(HASH_THREADS, HASH_MINB)
k_merkle_level(u64* digests, u64* cap, TreeGeom g, int layer, size_t n_parents) {
    const size_t j = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (j >= n_parents) return;
    const u64* kids = node_slot(digests, cap, g, layer, 2 * j); // the sibling pair is contiguous
    u64 l[4], r[4], o[4];
    const ulonglong2* k2 = reinterpret_cast<const ulonglong2*>(kids);
    ulonglong2 a = k2[0], b = k2[1], c = k2[2], d = k2[3];
    l[0] = a.x; l[1] = a.y; l[2] = b.x; l[3] = b.y;
    r[0] = c.x; r[1] = c.y; r[2] = d.x; r[3] = d.y;
    poseidon::two_to_one(l, r, o);
    u64* out = node_slot(digests, cap, g, layer + 1, j);
    reinterpret_cast<ulonglong2*>(out)[0] = make_ulonglong2(o[0], o[1]);
    reinterpret_cast<ulonglong2*>(out)[1] = make_ulonglong2(o[2], o[3]);
}

// Narrow layers (fewer parents than the GPU has threads): one permutation per half-warp
// (transcript::warp_permute: 12 lanes, S-boxes across lanes, MDS by shuffles), two per warp, spread over all SMs.
// There the latency of ONE permutation is all that matters, and the 12-lane form is several times shorter than
// the one-thread form (measured: 32 us per layer with one thread per parent, 36 us with all of a subtree's
// half-warps crowded onto one SM).
constexpr int COOP_THREADS = 128;
// measured per layer (ncu, config 2): one thread per parent 23.8 us at 16384 parents; half-warp form 42 / 27 / 17 / 13 us
// at 8192 / 4096 / 2048 / <= 1024 parents -- a lone warp per sub-partition finishes its permutations in ~19 us, so the
// half-warp form only pays below 4096 parents
#ifndef NLC_COOP_MAX_PARENTS
#define NLC_COOP_MAX_PARENTS 2048
#endif
constexpr size_t COOP_MAX_PARENTS = NLC_COOP_MAX_PARENTS;
__global__ void __launch_bounds__(COOP_THREADS) k_merkle_level_coop(u64* digests, u64* cap, TreeGeom g, int layer,
                                                                     size_t n_parents) {
    const int lane = threadIdx.x & 31, l16 = lane & 15;
    const size_t j = (blockIdx.x * (size_t)COOP_THREADS + threadIdx.x) >> 4; // one parent per half-warp
    const bool active = j < n_parents;
    const u64* kids = node_slot(digests, cap, g, layer, 2 * (active ? j : 0));
    u64 x = 0;
    if (active && l16 < 8) x = kids[l16];
    x = transcript::warp_permute(x, lane);
    if (active && l16 < 4) node_slot(digests, cap, g, layer + 1, j)[l16] = x;
}

// MerkleTree::get + MerkleTree::prove for many indices: out[i*stride + offset ..] = leaf row, siblings.
__global__ void k_open_many(const u64* __restrict__ leaves, const u64* __restrict__ digests, size_t leaf_len,
                            size_t leaf_stride, size_t elem_stride, TreeGeom g, const u64* __restrict__ idx,
                            int shift_right, u64 idx_mask, size_t n_idx, u64* out, size_t out_stride,
                            size_t out_offset) {
    const size_t qi = blockIdx.x;
    if (qi >= n_idx) return;
    const size_t leaf = (size_t)((idx[qi] & idx_mask) >> shift_right);
    u64* o = out + qi * out_stride + out_offset;
    const u64* row = leaves + leaf * leaf_stride;
    for (size_t i = threadIdx.x; i < leaf_len; i += blockDim.x) o[i] = gl::canon(row[i * elem_stride]);
    o += leaf_len;
    const size_t s = leaf >> g.sub_height;
    size_t pair_index = leaf & (((size_t)1 << g.sub_height) - 1);
    const u64* tree = digests + s * g.sub_digests * 4;
    for (int i = threadIdx.x >> 2; i < g.sub_height; i += blockDim.x >> 2) {
        size_t p = pair_index >> i; // index within layer i
        size_t parity = p & 1;
        size_t q = p >> 1;
        size_t sib = 2 * ((q << (i + 1)) + ((size_t)1 << i) - 1) + (1 - parity);
        o[(size_t)i * 4 + (threadIdx.x & 3)] = tree[sib * 4 + (threadIdx.x & 3)];
    }
}

__global__ void k_permute_many(u64* states, size_t count) {
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= count) return;
    u64 s[poseidon::WIDTH];
#pragma unroll
    for (int k = 0; k < poseidon::WIDTH; k++) s[k] = states[i * poseidon::WIDTH + k];
    poseidon::permute(s);
#pragma unroll
    for (int k = 0; k < poseidon::WIDTH; k++) states[i * poseidon::WIDTH + k] = s[k];
}

__global__ void k_two_to_one_many(const u64* l, const u64* r, size_t count, u64* out) {
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= count) return;
    u64 a[4], b[4], o[4];
    for (int k = 0; k < 4; k++) { a[k] = l[i * 4 + k]; b[k] = r[i * 4 + k]; }
    poseidon::two_to_one(a, b, o);
    for (int k = 0; k < 4; k++) out[i * 4 + k] = o[k];
}

} // namespace

// unit counter of one leaf-hash launch: stream-ordered, so concurrent launches on different streams never share one
static u32* unit_counter(nlc_ctx* ctx) {
    u32* c = (u32*)ctx->dmalloc(1);
    NLC_CUDA(cudaMemsetAsync(c, 0, sizeof(u32), ctx->stream));
    return c;
}

// Grid: HASH_MINB CTAs per SM (all resident at once), each warp claiming 32-leaf units until none are left.  The kernel
// is bound by each SM's integer pipes, not by thread count -- measured with the static deal: 5 CTAs/SM 7.41 ms,
// 4 CTAs/SM 7.51 ms, 3 CTAs/SM 7.75 ms for config 2.
static unsigned leaf_grid(nlc_ctx* ctx, size_t n_leaves) {
    const size_t want = (n_leaves + HASH_THREADS - 1) / HASH_THREADS;
    const size_t full = (size_t)ctx->sm_count * HASH_MINB;
    return (unsigned)(want < full ? want : full);
}

static TreeGeom tree_geom(size_t n_leaves, int cap_height) {
    const int h = ilog2(n_leaves);
    NLC_CHECK(((size_t)1 << h) == n_leaves, "MerkleTree::new: number of leaves must be a power of two");
    NLC_CHECK(cap_height >= 0 && cap_height <= h,
              "cap_height should be at most log2(leaves.len())"); // merkle_tree.rs assert in MerkleTree::new
    TreeGeom g;
    g.sub_height = h - cap_height;
    g.sub_digests = 2 * (((size_t)1 << g.sub_height) - 1);
    return g;
}

// digest layers above the leaf digests, down to the cap
void merkle_levels(nlc_ctx* ctx, size_t n_leaves, int cap_height, u64* digests, u64* cap) {
    const TreeGeom g = tree_geom(n_leaves, cap_height);
    ctx->phase_begin(3);
    for (int layer = 0; layer < g.sub_height; layer++) {
        const size_t parents = n_leaves >> (layer + 1);
        if (parents > COOP_MAX_PARENTS) {
            // fewer parents than the GPU has lanes: one-warp CTAs, so every warp gets a sub-partition of its own
            const int threads = parents >= (size_t)ctx->sm_count * 4 * 32 * 4 ? HASH_THREADS : 32;
            NLC_LAUNCH(ctx, k_merkle_level, (unsigned)((parents + threads - 1) / threads), threads, 0, digests, cap, g,
                       layer, parents);
        } else
            NLC_LAUNCH(ctx, k_merkle_level_coop, (unsigned)((parents * 16 + COOP_THREADS - 1) / COOP_THREADS),
                       COOP_THREADS, 0, digests, cap, g, layer, parents);
    }
    ctx->phase_end();
}

void merkle_build(nlc_ctx* ctx, const u64* leaves, size_t n_leaves, size_t leaf_len, size_t elem_stride, int cap_height,
                  u64* digests, u64* cap) {
    const TreeGeom g = tree_geom(n_leaves, cap_height);
    u32* ctr = unit_counter(ctx);
    ctx->phase_begin(2);
    if (elem_stride)
        NLC_LAUNCH(ctx, k_leaf_hash<true>, leaf_grid(ctx, n_leaves), HASH_THREADS, 0, leaves, n_leaves, leaf_len,
                   (size_t)1, elem_stride, g, digests, cap, ctr);
    else
        NLC_LAUNCH(ctx, k_leaf_hash<false>, leaf_grid(ctx, n_leaves), HASH_THREADS, 0, leaves, n_leaves, leaf_len,
                   leaf_len, (size_t)1, g, digests, cap, ctr);
    ctx->phase_end();
    ctx->dfree(ctr);
    merkle_levels(ctx, n_leaves, cap_height, digests, cap);
}

// One chunk of the leaf sponge over a column-major LDE (k_leaf_absorb); `state` = 12 * n_leaves words.
void merkle_leaf_absorb(nlc_ctx* ctx, const u64* lde, size_t n_leaves, size_t elem_stride, size_t col0, size_t n_cols,
                        bool first, bool last, u64* state, int cap_height, u64* digests, u64* cap) {
    NLC_CHECK(last || n_cols % poseidon::RATE == 0, "leaf absorb: inner chunks must be whole sponge blocks");
    const TreeGeom g = tree_geom(n_leaves, cap_height);
    u32* ctr = unit_counter(ctx);
    ctx->phase_begin(4);
    NLC_LAUNCH(ctx, k_leaf_absorb, leaf_grid(ctx, n_leaves), HASH_THREADS, 0, lde, n_leaves, elem_stride, col0, n_cols,
               first, last, state, g, digests, cap, ctr);
    ctx->phase_end();
    ctx->dfree(ctr);
}

void merkle_open_many(nlc_ctx* ctx, const u64* leaves, const u64* digests, size_t n_leaves, size_t leaf_len,
                      size_t elem_stride, int cap_height, const u64* idx_dev, int shift_right, size_t n_idx, u64* out,
                      size_t out_stride, size_t out_offset) {
    if (n_idx == 0) return;
    TreeGeom g;
    g.sub_height = ilog2(n_leaves) - cap_height;
    g.sub_digests = 2 * (((size_t)1 << g.sub_height) - 1);
    // idx values are x_index of the ORIGINAL domain already reduced mod N by the caller; the leaf of
    // this tree is x_index >> shift_right (fri_prover_query_round).
    NLC_LAUNCH(ctx, k_open_many, (unsigned)n_idx, 128, 0, leaves, digests, leaf_len, elem_stride ? (size_t)1 : leaf_len,
               elem_stride ? elem_stride : (size_t)1, g, idx_dev, shift_right, ~(u64)0, n_idx, out, out_stride,
               out_offset);
}

void poseidon_permute_many(nlc_ctx* ctx, u64* states_dev, size_t count) {
    if (!count) return;
    NLC_LAUNCH(ctx, k_permute_many, (unsigned)((count + 127) / 128), 128, 0, states_dev, count);
}

void hash_many(nlc_ctx* ctx, const u64* in_dev, size_t count, size_t len, bool or_noop, u64* out_dev) {
    if (!count) return;
    // hash_no_pad == hash_or_noop for len > 4; for len <= 4 hash_no_pad still permutes.  Reuse the leaf
    // kernel with a degenerate geometry (every "leaf" is its own cap entry: sub_height 0).
    TreeGeom g;
    g.sub_height = 0;
    g.sub_digests = 0;
    if (or_noop || len > 4) {
        u32* ctr = unit_counter(ctx);
        NLC_LAUNCH(ctx, k_leaf_hash<false>, leaf_grid(ctx, count), HASH_THREADS, 0, in_dev, count, len, len, (size_t)1, g,
                   (u64*)nullptr, out_dev, ctr);
        ctx->dfree(ctr);
    } else {
        // short hash_no_pad: pad into full states, permute, take 4 lanes
        u64* st = ctx->dmalloc(count * 12);
        NLC_CUDA(cudaMemsetAsync(st, 0, count * 12 * sizeof(u64), ctx->stream));
        if (len) NLC_CUDA(cudaMemcpy2DAsync(st, 12 * sizeof(u64), in_dev, len * sizeof(u64), len * sizeof(u64), count,
                                            cudaMemcpyDeviceToDevice, ctx->stream));
        if (len) poseidon_permute_many(ctx, st, count);
        NLC_CUDA(cudaMemcpy2DAsync(out_dev, 4 * sizeof(u64), st, 12 * sizeof(u64), 4 * sizeof(u64), count,
                                   cudaMemcpyDeviceToDevice, ctx->stream));
        ctx->dfree(st);
    }
}

void two_to_one_many(nlc_ctx* ctx, const u64* l_dev, const u64* r_dev, size_t count, u64* out_dev) {
    if (!count) return;
    NLC_LAUNCH(ctx, k_two_to_one_many, (unsigned)((count + 127) / 128), 128, 0, l_dev, r_dev, count, out_dev);
}
