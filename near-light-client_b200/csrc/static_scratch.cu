// static_scratch.cu -- nlc_merkle_cap_static_dev: a Merkle cap over row-major device leaves that allocates NOTHING.
//
// No counterpart in the reference: added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  It is a variant of merkle.cu's
// k_leaf_hash<false> + k_merkle_level (MerkleTree::new) whose scratch -- the leaf-unit counter
// and the interleaved `digests` buffer that merkle_build takes from the context's allocator -- lives in ONE file-scope
// __device__ array of fixed size (below).  Every GPU that loads the library holds that array until the process exits.
//
// Rules this file follows:
//   * in HOST code `g_scratch` is only a stand-in for the symbol, never a device address: the address comes from
//     cudaGetSymbolAddress, once per device with that device current (scratch_ptr), is checked, and is what the
//     kernels and cudaMemsetAsync receive;
//   * the entry point calls no cudaMalloc / cudaFree / allocating routine (no ctx->dmalloc, no staging, no events);
//   * every call shares the array and finds the previous call's contents in it: the unit counter is cleared in stream
//     order before each launch, and every digest slot is written before it is read (layer by layer), so nothing else
//     has to start at zero.  Calls that OVERLAP on one GPU (two streams, two contexts on the same device) race on it.
#include "common.cuh"
#include "poseidon.cuh"

#include <mutex>

namespace {

constexpr size_t STATIC_MAX_LEAVES = NLC_STATIC_MAX_LEAVES;       // 2^16 leaves
constexpr size_t HEADER_WORDS = 32;                               // 256 bytes: word 0 holds the u32 unit counter
constexpr size_t SCRATCH_WORDS = HEADER_WORDS + 8 * STATIC_MAX_LEAVES; // + digests: < 2 digests (8 words) per leaf
static __device__ __align__(256) u64 g_scratch[SCRATCH_WORDS];     // 4 MiB + 256 B per GPU

constexpr int THREADS = 128;
constexpr int MINB = 5;

struct Geom {
    int sub_height;     // log2(leaves per cap subtree)
    size_t sub_digests; // digests per subtree = 2 * (2^sub_height - 1)
};

// plonky2's interleaved `digests` layout (merkle_tree.rs fill_subtree; the same arithmetic as merkle.cu:node_slot):
// node j of layer i; layer sub_height is the cap
__device__ __forceinline__ u64* slot(u64* digests, u64* cap, Geom g, int layer, size_t j) {
    const int rem = g.sub_height - layer;
    const size_t s = j >> rem;
    if (rem == 0) return cap + s * 4;
    const size_t jl = j & (((size_t)1 << rem) - 1);
    const size_t idx = 2 * (((jl >> 1) << (layer + 1)) + ((size_t)1 << layer) - 1) + (jl & 1);
    return digests + (s * g.sub_digests + idx) * 4;
}

// hash_or_noop of every row-major leaf; warps claim 32-leaf units from the counter in scratch[0]
__global__ void __launch_bounds__(THREADS, MINB)
k_static_leaf_hash(const u64* __restrict__ leaves, size_t n_leaves, size_t leaf_len, Geom g, u64* scratch, u64* cap) {
    u32* ctr = reinterpret_cast<u32*>(scratch);
    u64* digests = scratch + HEADER_WORDS;
    const u32 n_warps = gridDim.x * (THREADS / 32);
    const size_t n_units = (n_leaves + 31) >> 5;
    const int lane = threadIdx.x & 31;
    for (size_t unit = blockIdx.x * (size_t)(THREADS / 32) + (threadIdx.x >> 5); unit < n_units;) {
        const size_t row = (unit << 5) + lane;
        if (row < n_leaves) {
            const u64* in = leaves + row * leaf_len;
            u64 s[poseidon::WIDTH];
#pragma unroll
            for (int i = 0; i < poseidon::WIDTH; i++) s[i] = 0;
            if (leaf_len <= 4) {
#pragma unroll
                for (int i = 0; i < 4; i++)
                    if ((size_t)i < leaf_len) s[i] = gl::canon(in[i]);
            } else {
#pragma unroll 1
                for (size_t off = 0; off < leaf_len; off += poseidon::RATE) {
                    const size_t rem = leaf_len - off;
#pragma unroll
                    for (int i = 0; i < poseidon::RATE; i++)
                        if ((size_t)i < rem) s[i] = in[off + i]; // overwrite mode
                    poseidon::permute_nc(s);
                }
#pragma unroll
                for (int i = 0; i < 4; i++) s[i] = gl::canon(s[i]);
            }
            u64* out = slot(digests, cap, g, 0, row);
            reinterpret_cast<ulonglong2*>(out)[0] = make_ulonglong2(s[0], s[1]);
            reinterpret_cast<ulonglong2*>(out)[1] = make_ulonglong2(s[2], s[3]);
        }
        u32 nx = 0;
        if (lane == 0) nx = atomicAdd(ctr, 1u);
        unit = (size_t)n_warps + __shfl_sync(0xffffffffu, nx, 0);
    }
}

// layer i -> layer i + 1, one thread per parent
__global__ void __launch_bounds__(THREADS, MINB)
k_static_level(u64* scratch, u64* cap, Geom g, int layer, size_t n_parents) {
    u64* digests = scratch + HEADER_WORDS;
    const size_t j = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (j >= n_parents) return;
    const ulonglong2* k2 = reinterpret_cast<const ulonglong2*>(slot(digests, cap, g, layer, 2 * j));
    const ulonglong2 a = k2[0], b = k2[1], c = k2[2], d = k2[3];
    u64 l[4] = {a.x, a.y, b.x, b.y}, r[4] = {c.x, c.y, d.x, d.y}, o[4];
    poseidon::two_to_one(l, r, o);
    u64* out = slot(digests, cap, g, layer + 1, j);
    reinterpret_cast<ulonglong2*>(out)[0] = make_ulonglong2(o[0], o[1]);
    reinterpret_cast<ulonglong2*>(out)[1] = make_ulonglong2(o[2], o[3]);
}

// device address of g_scratch on `device` (which must be current): looked up once per device
constexpr int MAX_DEVICES = 64;
std::mutex g_mu;
u64* g_ptr[MAX_DEVICES] = {};

u64* scratch_ptr(int device) {
    NLC_CHECK(device >= 0 && device < MAX_DEVICES, "static scratch: device ordinal out of range");
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_ptr[device]) {
        int cur = -1;
        NLC_CUDA(cudaGetDevice(&cur));
        NLC_CHECK(cur == device, "static scratch: the context's device is not current");
        u64* p = nullptr;
        NLC_CUDA(cudaGetSymbolAddress((void**)&p, g_scratch));
        NLC_CHECK(p != nullptr, "static scratch: cudaGetSymbolAddress returned a null address");
        g_ptr[device] = p;
    }
    return g_ptr[device];
}

} // namespace

extern "C" int nlc_merkle_cap_static_dev(nlc_ctx* ctx, const uint64_t* leaves_dev, size_t n_leaves, size_t leaf_len,
                                         int cap_height, uint64_t* cap_out_dev) {
    if (!ctx) return 1;
    return nlc_guarded(ctx, [&] {
        CtxGuard guard(ctx); // takes the context's lock and makes its device current
        NLC_CHECK(leaves_dev && cap_out_dev, "null argument");
        NLC_CHECK(n_leaves > 0 && n_leaves <= STATIC_MAX_LEAVES,
                  "nlc_merkle_cap_static_dev: n_leaves exceeds NLC_STATIC_MAX_LEAVES (65536), the size of the static "
                  "scratch array; use nlc_merkle_new for larger trees");
        const int h = ilog2(n_leaves);
        NLC_CHECK(((size_t)1 << h) == n_leaves, "MerkleTree::new: number of leaves must be a power of two");
        NLC_CHECK(cap_height >= 0 && cap_height <= h, "cap_height should be at most log2(leaves.len())");
        u64* scratch = scratch_ptr(ctx->device);
        Geom g;
        g.sub_height = h - cap_height;
        g.sub_digests = 2 * (((size_t)1 << g.sub_height) - 1);
        NLC_CUDA(cudaMemsetAsync(scratch, 0, HEADER_WORDS * sizeof(u64), ctx->stream)); // the unit counter starts at zero
        const size_t want = (n_leaves + THREADS - 1) / THREADS, full = (size_t)ctx->sm_count * MINB;
        NLC_LAUNCH(ctx, k_static_leaf_hash, (unsigned)(want < full ? want : full), THREADS, 0, leaves_dev, n_leaves,
                   leaf_len, g, scratch, cap_out_dev);
        for (int layer = 0; layer < g.sub_height; layer++) {
            const size_t parents = n_leaves >> (layer + 1);
            const int threads = parents >= (size_t)ctx->sm_count * 512 ? THREADS : 32;
            NLC_LAUNCH(ctx, k_static_level, (unsigned)((parents + threads - 1) / threads), threads, 0, scratch,
                       cap_out_dev, g, layer, parents);
        }
    });
}
