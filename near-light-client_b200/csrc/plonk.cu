// plonk.cu -- the plonk-side quotient path on the device: the permutation argument's partial
// products and Z polynomials, and the vanishing polynomial over the quotient domain.
//
// Replaces, in plonky2 0.1.4 @ d2598bd (pinned by /root/reference/Cargo.lock:4864-4886; source NOT in tree; reached from
// /root/reference/nearx/src/test_utils.rs:62 `circuit.prove` and every plonky2x circuit build):
//   plonk/prover.rs   wires_permutation_partial_products_and_zs / all_wires_permutation_partial_products
//                     (per row: quotient_chunk_products of (w + beta k_j x + gamma) / (w + beta sigma_j + gamma) over
//                     chunks of quotient_degree_factor routed wires; Z(g x) = Z(x) * prod of the row's chunks)
//   plonk/prover.rs   compute_quotient_polys (the loop over the points of the quotient domain: get_lde_values of the
//                     three commitments, eval_vanishing_poly_base_batch, the 1/Z_H scaling)
//   plonk/vanishing_poly.rs  eval_vanishing_poly_base_batch: L_0(x) (Z(x) - 1), check_partial_products,
//                     evaluate_gate_constraints_base_batch, reduce_with_powers_multi over the alphas
// The gate constraints arrive as the same bytecode the STARK path uses (air.cuh): the host records every gate's
// filtered constraints against a recording parser, sums them per constraint index, and emits them HIGHEST INDEX FIRST,
// so that the interpreter's Horner update acc = acc * alpha + c continues seamlessly into the permutation terms and
// ends with sum_k term_k alpha^k -- reduce_with_powers_multi's value for the order
// [z_1 terms | partial-product terms | gate constraint terms].
//
// Nothing leaves HBM: wire / sigma values on the subgroup H are recovered from the resident coefficient columns by
// one forward NTT, the LDE rows are the resident column-major `lde` of the three batches read in leaf order.
#include "air.cuh"

namespace {

constexpr int PP_THREADS = 128;
constexpr int MAX_CHUNKS = 64;

struct PermArgs {
    const u64* wires;    // num_routed x n values on H (natural order), column-major
    const u64* sigmas;   // num_routed x n
    const u64* subgroup; // w_n^i
    const u64* k_is;
    const u64* betas;
    const u64* gammas;
    u32 num_routed, chunk, n_chunks, num_ch;
    size_t n;
    u64* chunk_q;   // (num_ch * n_chunks) x n: quotient chunk products
    u64* row_total; // num_ch x n: product of a row's chunks = Z(g x) / Z(x)
};

// wires_permutation_partial_products_and_zs, the per-row part: one thread per (row, challenge).
__global__ void __launch_bounds__(PP_THREADS) k_pp_rows(const PermArgs A) {
    const size_t i = blockIdx.x * (size_t)PP_THREADS + threadIdx.x;
    const u32 ch = blockIdx.y;
    if (i >= A.n) return;
    const u64 beta = A.betas[ch], gamma = A.gammas[ch];
    const u64 bx = gl::mul(beta, A.subgroup[i]);
    u64 num[MAX_CHUNKS], den[MAX_CHUNKS];
    for (u32 c = 0; c < A.n_chunks; c++) {
        u64 pn = 1, pd = 1;
        const u32 j1 = min((c + 1) * A.chunk, A.num_routed);
        for (u32 j = c * A.chunk; j < j1; j++) {
            const u64 w = gl::dadd(gl::dcanon(A.wires[j * A.n + i]), gamma);
            pn = gl::mul(pn, gl::dadd(w, gl::mul(A.k_is[j], bx)));
            pd = gl::mul(pd, gl::dadd(w, gl::mul(beta, gl::dcanon(A.sigmas[j * A.n + i]))));
        }
        num[c] = pn;
        den[c] = pd;
    }
    // batch inversion of the chunk denominators (one field inversion per row): in a field prod(num_j / den_j) over a
    // chunk equals prod(num_j) / prod(den_j) exactly, so this is plonky2's per-wire batch_multiplicative_inverse value
    u64 pre[MAX_CHUNKS];
    u64 run = 1;
    for (u32 c = 0; c < A.n_chunks; c++) {
        pre[c] = run;
        run = gl::mul(run, den[c]);
    }
    u64 inv = gl::inv(run);
    u64 total = 1;
    for (u32 c = A.n_chunks; c-- > 0;) {
        const u64 dinv = gl::mul(inv, pre[c]);
        inv = gl::mul(inv, den[c]);
        num[c] = gl::mul(num[c], dinv);
    }
    for (u32 c = 0; c < A.n_chunks; c++) {
        A.chunk_q[((size_t)ch * A.n_chunks + c) * A.n + i] = num[c];
        total = gl::mul(total, num[c]);
    }
    A.row_total[(size_t)ch * A.n + i] = total;
}

// exclusive prefix PRODUCTS of `len` elements per row of a (rows x len) matrix, blocks of 1024: data[i] <- prod_{i' < i}
// within the block, block totals to `totals` (rows x n_blocks)
__global__ void __launch_bounds__(1024) k_scan_block(u64* data, size_t len, u64* totals, size_t n_blocks) {
    __shared__ u64 sh[1024];
    const size_t i = blockIdx.x * (size_t)1024 + threadIdx.x;
    u64* row = data + blockIdx.y * len;
    const u64 v = i < len ? row[i] : 1;
    sh[threadIdx.x] = v;
    __syncthreads();
    u64 acc = v;
    for (int off = 1; off < 1024; off <<= 1) { // Hillis-Steele inclusive scan
        u64 o = (int)threadIdx.x >= off ? sh[threadIdx.x - off] : 1;
        __syncthreads();
        acc = gl::mul(acc, o);
        sh[threadIdx.x] = acc;
        __syncthreads();
    }
    if (i < len) row[i] = threadIdx.x ? sh[threadIdx.x - 1] : 1;
    if (threadIdx.x == 1023 && totals) totals[blockIdx.y * n_blocks + blockIdx.x] = acc;
}
__global__ void k_scan_apply(u64* data, size_t len, const u64* block_prefix, size_t n_blocks) {
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= len) return;
    u64* row = data + blockIdx.y * len;
    row[i] = gl::mul(row[i], block_prefix[blockIdx.y * n_blocks + i / 1024]);
}

void exclusive_prefix_products(nlc_ctx* ctx, u64* data, size_t rows, size_t len) {
    const size_t n_blocks = (len + 1023) / 1024;
    if (n_blocks == 1) {
        NLC_LAUNCH(ctx, k_scan_block, dim3(1, (unsigned)rows), 1024, 0, data, len, (u64*)nullptr, (size_t)1);
        return;
    }
    u64* totals = ctx->dmalloc(rows * n_blocks);
    try {
        NLC_LAUNCH(ctx, k_scan_block, dim3((unsigned)n_blocks, (unsigned)rows), 1024, 0, data, len, totals, n_blocks);
        exclusive_prefix_products(ctx, totals, rows, n_blocks);
        NLC_LAUNCH(ctx, k_scan_apply, dim3((unsigned)((len + 255) / 256), (unsigned)rows), 256, 0, data, len, totals, n_blocks);
    } catch (...) {
        ctx->dfree(totals);
        throw;
    }
    ctx->dfree(totals);
}

// partial_products_and_z_gx + the swap that puts Z(x) in the row (prover.rs): values in COMMITMENT order
//   [Z_0 .. Z_{c-1} | pp(0)_0 .. pp(0)_{np-1} | pp(1)_0 .. ]  (zs_range, then partial_products_range)
__global__ void __launch_bounds__(PP_THREADS) k_pp_finish(const PermArgs A, const u64* __restrict__ z, u64* __restrict__ out) {
    const size_t i = blockIdx.x * (size_t)PP_THREADS + threadIdx.x;
    const u32 ch = blockIdx.y;
    if (i >= A.n) return;
    const u32 num_prods = A.n_chunks - 1;
    u64 acc = z[(size_t)ch * A.n + i];
    out[(size_t)ch * A.n + i] = acc;
    for (u32 c = 0; c < num_prods; c++) {
        acc = gl::mul(acc, A.chunk_q[((size_t)ch * A.n_chunks + c) * A.n + i]);
        out[((size_t)A.num_ch + (size_t)ch * num_prods + c) * A.n + i] = acc;
    }
}

struct VanishArgs {
    const u64* const* wires;  // LDE columns (leaf order) of the routed wires
    const u64* const* sigmas; // LDE columns of the sigma polynomials
    const u64* const* zs_pp;  // LDE columns of zs_partial_products: num_ch Zs, then num_ch * num_prods partial products
    const u64* k_is;
    const u64* betas;
    const u64* gammas;
    u32 num_routed, chunk, n_chunks, num_ch;
};

// eval_vanishing_poly_base_batch for every point of the quotient domain; one thread per leaf k of the domain prefix.
// out[i * num_ch + j] = sum_k term_k alpha_j^k, BEFORE the multiplication by 1/Z_H (quotient_commit_dev applies it).
template <int NA>
__global__ void __launch_bounds__(AIR_THREADS, air_min_blocks(NA)) k_plonk_vanishing(const AirArgs A, const VanishArgs V) {
    const size_t size = (size_t)1 << A.log_size;
    const size_t stride = (size_t)gridDim.x * AIR_THREADS;
    u64* regs = A.regs + (size_t)blockIdx.x * AIR_THREADS + threadIdx.x;
    const u32 num_prods = V.n_chunks - 1;
    for (size_t k = (size_t)blockIdx.x * AIR_THREADS + threadIdx.x; k < size; k += stride) {
        const size_t i = A.log_size ? (__brevll(k) >> (64 - A.log_size)) : 0;
        const size_t i_next = (i + ((size_t)1 << A.qdb)) & (size - 1);
        const size_t k_next = A.log_size ? (__brevll(i_next) >> (64 - A.log_size)) : 0;
        u64 acc[NA];
#pragma unroll
        for (int j = 0; j < NA; j++) acc[j] = 0;
        auto fold = [&](const u64 term) {
#pragma unroll
            for (int j = 0; j < NA; j++)
                if (NA < MAX_ALPHAS || (u32)j < A.n_alphas) acc[j] = gl::dadd(gl::mul(acc[j], A.alphas[j]), term);
        };
        // gate constraint terms, highest index first (the host emits them in that order)
        air_run<NA>(A, regs, k, k_next, 0, 0, 0, acc);
        // check_partial_products, last challenge and last chunk first
        const u64 x = A.coset_x[i]; // shifted_x = g * w^i
        for (u32 ch = V.num_ch; ch-- > 0;) {
            const u64 beta = V.betas[ch], gamma = V.gammas[ch];
            const u64 bx = gl::mul(beta, x);
            const u64 z_x = gl::dcanon(V.zs_pp[ch][k]), z_gx = gl::dcanon(V.zs_pp[ch][k_next]);
            for (u32 c = V.n_chunks; c-- > 0;) {
                u64 pn = 1, pd = 1;
                const u32 j1 = min((c + 1) * V.chunk, V.num_routed);
                for (u32 j = c * V.chunk; j < j1; j++) {
                    const u64 w = gl::dadd(gl::dcanon(V.wires[j][k]), gamma);
                    pn = gl::mul(pn, gl::dadd(w, gl::mul(V.k_is[j], bx)));
                    pd = gl::mul(pd, gl::dadd(w, gl::mul(beta, gl::dcanon(V.sigmas[j][k]))));
                }
                const u64 prev = c == 0 ? z_x : gl::dcanon(V.zs_pp[V.num_ch + ch * num_prods + c - 1][k]);
                const u64 next = c == num_prods ? z_gx : gl::dcanon(V.zs_pp[V.num_ch + ch * num_prods + c][k]);
                fold(gl::dsub(gl::mul(prev, pn), gl::mul(next, pd)));
            }
        }
        // vanishing_z_1_terms: L_0(x) (Z(x) - 1); L_0 on the coset is the LDE of the first Lagrange selector
        const u64 l0 = A.lagrange_first[i];
        for (u32 ch = V.num_ch; ch-- > 0;) fold(gl::mul(l0, gl::dsub(gl::dcanon(V.zs_pp[ch][k]), 1)));
#pragma unroll
        for (int j = 0; j < NA; j++)
            if (NA < MAX_ALPHAS || (u32)j < A.n_alphas) A.out[i * A.n_alphas + j] = acc[j];
    }
}

__global__ void k_subgroup(u64* out, size_t n, u64 w) {
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i < n) out[i] = gl::pow(w, i);
}

void check_circuit(const nlc_plonk_circuit* c, const nlc_batch* cs, const nlc_batch* wires) {
    NLC_CHECK(c && cs && wires, "null argument");
    NLC_CHECK(c->num_routed_wires >= 1 && c->num_routed_wires <= wires->W, "num_routed_wires exceeds the wires commitment");
    NLC_CHECK(c->num_constants + c->num_routed_wires <= cs->W, "constants_sigmas commitment narrower than constants + sigmas");
    NLC_CHECK(c->num_challenges >= 1 && c->num_challenges <= (size_t)MAX_ALPHAS, "num_challenges out of range (1..8)");
    NLC_CHECK(c->quotient_degree_factor >= 2, "max_degree > 1"); // check_partial_products' debug_assert
    const size_t n_chunks = (c->num_routed_wires + c->quotient_degree_factor - 1) / c->quotient_degree_factor;
    NLC_CHECK(c->num_partial_products + 1 == n_chunks, "num_partial_products must be ceil(num_routed_wires / quotient_degree_factor) - 1");
    NLC_CHECK(n_chunks <= (size_t)MAX_CHUNKS, "too many partial-product chunks");
    NLC_CHECK(c->k_is != nullptr, "null k_is");
    NLC_CHECK(cs->log_n == wires->log_n && cs->rate_bits == wires->rate_bits, "commitments of different shapes");
    NLC_CHECK(cs->shard_log == 0 && wires->shard_log == 0 && cs->owns_lde && wires->owns_lde && cs->coeffs && wires->coeffs,
              "the plonk quotient path needs whole (unsharded) commitments");
}

} // namespace

// all_wires_permutation_partial_products + the column order of the prover's `zs_partial_products` + from_values
nlc_batch* plonk_zs_partial_products_commit_dev(nlc_ctx* ctx, const nlc_plonk_circuit* c, nlc_batch* cs, nlc_batch* wires,
                                                const u64* betas, const u64* gammas, int cap_height) {
    check_circuit(c, cs, wires);
    NLC_CHECK(betas && gammas, "null challenges");
    const int log_n = wires->log_n;
    const size_t n = (size_t)1 << log_n, R = c->num_routed_wires, nch = c->num_challenges;
    const size_t n_chunks = c->num_partial_products + 1, W_out = nch * n_chunks;
    u64 *wv = nullptr, *sv = nullptr, *scratch = nullptr, *small = nullptr, *chunk_q = nullptr, *row_total = nullptr, *vals = nullptr;
    nlc_batch* out = nullptr;
    try {
        wv = ctx->dmalloc(R * n);
        sv = ctx->dmalloc(R * n);
        scratch = ctx->dmalloc(std::max<size_t>(ntt_scratch_words(R, log_n), 1));
        // values on H of the routed wires and of the sigmas: one forward transform of the resident coefficients
        for (int which = 0; which < 2; which++) {
            NttJob job;
            job.src = which ? cs->coeffs + c->num_constants * n : wires->coeffs;
            job.src_stride = n;
            job.n_cols = R;
            job.log_n = log_n;
            job.dst = which ? sv : wv;
            job.dst_stride = n;
            job.scratch = scratch;
            ntt_run(ctx, job);
        }
        std::vector<u64> host(R + 2 * nch);
        for (size_t j = 0; j < R; j++) host[j] = gl::canon(c->k_is[j]);
        for (size_t j = 0; j < nch; j++) host[R + j] = gl::canon(betas[j]), host[R + nch + j] = gl::canon(gammas[j]);
        small = ctx->dmalloc(host.size());
        NLC_CUDA(cudaMemcpyAsync(small, host.data(), host.size() * sizeof(u64), cudaMemcpyHostToDevice, ctx->stream));
        NLC_CUDA(cudaStreamSynchronize(ctx->stream));
        chunk_q = ctx->dmalloc(W_out * n);
        row_total = ctx->dmalloc(nch * n);
        vals = ctx->dmalloc(W_out * n);
        PermArgs A;
        A.wires = wv;
        A.sigmas = sv;
        u64* sub = ctx->dmalloc(n); // the subgroup H in natural order: w_n^i
        try {
            NLC_LAUNCH(ctx, k_subgroup, (unsigned)((n + 255) / 256), 256, 0, sub, n, gl::h_root_of_unity(log_n));
            A.subgroup = sub;
            A.k_is = small;
            A.betas = small + R;
            A.gammas = small + R + nch;
            A.num_routed = (u32)R;
            A.chunk = (u32)c->quotient_degree_factor;
            A.n_chunks = (u32)n_chunks;
            A.num_ch = (u32)nch;
            A.n = n;
            A.chunk_q = chunk_q;
            A.row_total = row_total;
            dim3 grid((unsigned)((n + PP_THREADS - 1) / PP_THREADS), (unsigned)nch);
            NLC_LAUNCH(ctx, k_pp_rows, grid, PP_THREADS, 0, A);
            exclusive_prefix_products(ctx, row_total, nch, n); // Z(x_i) = prod_{i' < i} row_total(i'), Z(x_0) = 1
            NLC_LAUNCH(ctx, k_pp_finish, grid, PP_THREADS, 0, A, row_total, vals);
        } catch (...) {
            ctx->dfree(sub);
            throw;
        }
        ctx->dfree(sub);
        out = batch_alloc(ctx, W_out, log_n, wires->rate_bits, cap_height);
        batch_ifft_values(out, vals);
        batch_lde_and_tree(out);
    } catch (...) {
        if (out) batch_release(out);
        for (u64* p : {wv, sv, scratch, small, chunk_q, row_total, vals}) ctx->dfree(p);
        throw;
    }
    for (u64* p : {wv, sv, scratch, small, chunk_q, row_total, vals}) ctx->dfree(p);
    return out;
}

// compute_quotient_polys up to (not including) the 1/Z_H scaling: out = size x num_challenges, natural order of i
void plonk_vanishing_dev(nlc_ctx* ctx, const nlc_plonk_circuit* c, const nlc_air_program* gates, nlc_batch* cs, nlc_batch* wires,
                         nlc_batch* zs_pp, const u64* public_inputs_hash, const u64* betas, const u64* gammas,
                         const u64* alphas, u64* out_dev) {
    check_circuit(c, cs, wires);
    NLC_CHECK(zs_pp && gates && betas && gammas && alphas, "null argument");
    const size_t R = c->num_routed_wires, nch = c->num_challenges, n_chunks = c->num_partial_products + 1;
    NLC_CHECK(zs_pp->W == nch * n_chunks, "zs_partial_products commitment must hold num_challenges * (num_partial_products + 1) polynomials");
    NLC_CHECK(zs_pp->log_n == wires->log_n && zs_pp->rate_bits == wires->rate_bits && zs_pp->shard_log == 0 && zs_pp->owns_lde,
              "commitments of different shapes");
    const int log_n = wires->log_n, rate_bits = wires->rate_bits;
    int qdb = 0;
    while ((1 << qdb) < c->quotient_degree_factor) qdb++; // log2_ceil(quotient_degree_factor)
    NLC_CHECK(qdb <= rate_bits, "Having constraints of degree higher than the rate is not supported yet.");
    // LOCAL column j of the gate program: constants_sigmas column j for j < W_cs, then the wires
    std::vector<const u64*> cols;
    const size_t N = cs->n_local();
    for (size_t j = 0; j < cs->W; j++) cols.push_back(cs->lde + j * N);
    for (size_t j = 0; j < wires->W; j++) cols.push_back(wires->lde + j * N);
    const size_t n_gate_cols = cols.size();
    validate_program(gates, n_gate_cols, true);
    // ... followed by the pointer tables of the permutation argument
    const size_t o_w = cols.size();
    for (size_t j = 0; j < R; j++) cols.push_back(wires->lde + j * N);
    const size_t o_s = cols.size();
    for (size_t j = 0; j < R; j++) cols.push_back(cs->lde + (c->num_constants + j) * N);
    const size_t o_z = cols.size();
    for (size_t j = 0; j < zs_pp->W; j++) cols.push_back(zs_pp->lde + j * N);
    std::vector<u64> extra(R + 2 * nch);
    for (size_t j = 0; j < R; j++) extra[j] = c->k_is[j];
    for (size_t j = 0; j < nch; j++) extra[R + j] = betas[j], extra[R + nch + j] = gammas[j];
    const size_t size = (size_t)1 << (log_n + qdb);
    const u64 *x, *lf, *ll;
    quotient_tables_dev(ctx, log_n, qdb, &x, &lf, &ll);
    const u64* extra_dev = nullptr;
    NLC_CHECK(gates->n_public <= 4, "the gate program's public inputs are the 4 elements of public_inputs_hash");
    AirDevice dev(ctx, gates, cols, public_inputs_hash, nullptr, nullptr, alphas, nch, size, extra.data(), extra.size(), &extra_dev,
                  air_min_blocks(air_na(nch)));
    AirArgs& A = dev.A;
    A.log_size = log_n + qdb;
    A.qdb = qdb;
    A.coset_x = x;
    A.lagrange_first = lf;
    A.lagrange_last = ll;
    A.last = 0;
    A.out = out_dev;
    VanishArgs V;
    V.wires = A.cols + o_w;
    V.sigmas = A.cols + o_s;
    V.zs_pp = A.cols + o_z;
    V.k_is = extra_dev;
    V.betas = extra_dev + R;
    V.gammas = extra_dev + R + nch;
    V.num_routed = (u32)R;
    V.chunk = (u32)c->quotient_degree_factor;
    V.n_chunks = (u32)n_chunks;
    V.num_ch = (u32)nch;
    switch (air_na(nch)) {
    case 1: NLC_LAUNCH(ctx, k_plonk_vanishing<1>, (unsigned)dev.blocks, AIR_THREADS, 0, A, V); break;
    case 2: NLC_LAUNCH(ctx, k_plonk_vanishing<2>, (unsigned)dev.blocks, AIR_THREADS, 0, A, V); break;
    case 3: NLC_LAUNCH(ctx, k_plonk_vanishing<3>, (unsigned)dev.blocks, AIR_THREADS, 0, A, V); break;
    case 4: NLC_LAUNCH(ctx, k_plonk_vanishing<4>, (unsigned)dev.blocks, AIR_THREADS, 0, A, V); break;
    default: NLC_LAUNCH(ctx, k_plonk_vanishing<MAX_ALPHAS>, (unsigned)dev.blocks, AIR_THREADS, 0, A, V); break;
    }
}
