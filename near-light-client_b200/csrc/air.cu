// air.cu -- AIR constraint evaluation over the quotient domain, on the device.
//
// Replaces the hot loop of starkyx's quotient_polys
// (/root/reference/vendor/starkyx/starkyx/src/plonky2/stark/prover.rs:306-352): for every point i of the quotient
// domain  local_vars = get_trace_values_packed(i), next_vars = get_trace_values_packed(i + next_step)
// (prover.rs:289-294,326-327), `stark.air().eval(&mut parser)` (prover.rs:334) feeding a ConstraintConsumer
// (plonky2/parser/consumer.rs:54-78), and the accumulators of every alpha.  The AIR arrives as a BYTECODE: the
// caller runs its `air.eval(..)` ONCE against a recording AirParser (the way RecursiveStarkParser records a circuit,
// plonky2/parser/mod.rs:30-38) and hands the resulting straight-line program over; any AIR that is written against
// the AirParser trait (air/parser.rs) -- Chip<L> constraint lists (chip/mod.rs:58-66), ArithmeticConstraint
// (chip/arithmetic/mod.rs:13-50), the field instructions (chip/field/*.rs) -- compiles unchanged.
//
// Data never leaves HBM: the trace LDEs are the resident column-major, index-bit-reversed `lde` of each nlc_batch,
// and the quotient domain g<w_size> is exactly the leaf prefix [0, size) of that order (point i = leaf
// bitrev_{log size}(i)), so one thread per LEAF reads its local row with fully coalesced column loads; the next row
// (i + next_step) differs from it in a few middle bits, the same for a whole warp, so those loads coalesce too.
#include "air.cuh"

namespace {

enum { KIND_QX = 10, KIND_QLF = 11, KIND_QLL = 12 };

// One thread per leaf k of the quotient-domain prefix.
template <int NA>
__global__ void __launch_bounds__(AIR_THREADS, air_min_blocks(NA)) k_air_eval(const AirArgs A) {
    const size_t size = (size_t)1 << A.log_size;
    const size_t stride = (size_t)gridDim.x * AIR_THREADS;
    u64* regs = A.regs + (size_t)blockIdx.x * AIR_THREADS + threadIdx.x;
    for (size_t k = (size_t)blockIdx.x * AIR_THREADS + threadIdx.x; k < size; k += stride) {
        const size_t i = A.log_size ? (__brevll(k) >> (64 - A.log_size)) : 0;
        const size_t i_next = (i + ((size_t)1 << A.qdb)) & (size - 1);
        const size_t k_next = A.log_size ? (__brevll(i_next) >> (64 - A.log_size)) : 0;
        const u64 z_last = gl::sub(A.coset_x[i], A.last);
        const u64 l_first = A.lagrange_first[i], l_last = A.lagrange_last[i];
        u64 acc[NA];
#pragma unroll
        for (int j = 0; j < NA; j++) acc[j] = 0;
        air_run<NA>(A, regs, k, k_next, z_last, l_first, l_last, acc);
#pragma unroll
        for (int j = 0; j < NA; j++)
            if (NA < MAX_ALPHAS || (u32)j < A.n_alphas) A.out[i * A.n_alphas + j] = acc[j];
    }
}

__global__ void k_selector_pair(u64* vals, size_t n) { // PolynomialValues::selector(n, 0) and selector(n, n - 1)
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i < 2 * n) vals[i] = (i == 0 || i == 2 * n - 1) ? 1 : 0;
}

// device-resident, cached per (log_n, qdb): coset points and the two Lagrange selectors on the quotient domain
const u64* cached(nlc_ctx* ctx, int kind, int log_n, int qdb) {
    auto it = ctx->tables.find(TableKey{kind, log_n, qdb, 0, 0, 0});
    return it == ctx->tables.end() ? nullptr : it->second;
}
u64* reserve(nlc_ctx* ctx, int kind, int log_n, int qdb, size_t words) {
    u64* p = nullptr;
    NLC_CUDA(cudaMalloc((void**)&p, words * sizeof(u64)));
    ctx->tables[TableKey{kind, log_n, qdb, 0, 0, 0}] = p;
    ctx->table_bytes += words * sizeof(u64);
    return p;
}

} // namespace

// The tables prover.rs:278-302 builds before its loop, left in HBM (nlc_quotient_domain copies them to the host).
void quotient_tables_dev(nlc_ctx* ctx, int log_n, int qdb, const u64** x, const u64** lf, const u64** ll) {
    const size_t n = (size_t)1 << log_n, size = n << qdb;
    const int log_size = log_n + qdb;
    if (!cached(ctx, KIND_QX, log_n, qdb)) {
        u64* xd = reserve(ctx, KIND_QX, log_n, qdb, size);
        coset_points(ctx, log_size, xd);
        u64* lfd = reserve(ctx, KIND_QLF, log_n, qdb, 2 * size); // first and last back to back
        ctx->tables[TableKey{KIND_QLL, log_n, qdb, 0, 0, 0}] = lfd + size;
        // PolynomialValues::selector(degree, i).lde_onto_coset(quotient_degree_bits), i = 0 and degree - 1
        u64* vals = ctx->dmalloc(2 * n);
        u64* padded = ctx->dmalloc(2 * size);
        u64* scratch = ctx->dmalloc(std::max<size_t>(std::max(ntt_scratch_words(2, log_n), ntt_scratch_words(2, log_size)), 1));
        try {
            NLC_LAUNCH(ctx, k_selector_pair, (unsigned)((2 * n + 255) / 256), 256, 0, vals, n);
            NLC_CUDA(cudaMemsetAsync(padded, 0, 2 * size * sizeof(u64), ctx->stream));
            NttJob inv;
            inv.src = vals;
            inv.src_stride = n;
            inv.n_cols = 2;
            inv.log_n = log_n;
            inv.inverse = true;
            inv.dst = padded;
            inv.dst_stride = size; // coefficients land in the first n slots of each zero-padded column
            inv.scratch = scratch;
            ntt_run(ctx, inv);
            NttJob fwd;
            fwd.src = padded;
            fwd.src_stride = size;
            fwd.n_cols = 2;
            fwd.log_n = log_size;
            fwd.pre_shift = gl::GENERATOR;
            fwd.dst = lfd;
            fwd.dst_stride = size;
            fwd.scratch = scratch;
            ntt_run(ctx, fwd);
        } catch (...) {
            ctx->dfree(scratch);
            ctx->dfree(padded);
            ctx->dfree(vals);
            throw;
        }
        ctx->dfree(scratch);
        ctx->dfree(padded);
        ctx->dfree(vals);
    }
    *x = cached(ctx, KIND_QX, log_n, qdb);
    *lf = cached(ctx, KIND_QLF, log_n, qdb);
    *ll = cached(ctx, KIND_QLL, log_n, qdb);
}

// prover.rs:306-352 without the 1/Z_H scaling (nlc_quotient_commit_dev applies it): out = size x n_alphas raw accumulators.
void air_eval_dev(nlc_ctx* ctx, const nlc_air_program* prog, nlc_batch* const* batches, int n_batches, const u64* publics,
                  const u64* globals, const u64* challenges, const u64* alphas, size_t n_alphas, int qdb, u64* out_dev) {
    NLC_CHECK(prog && prog->code && prog->n_ops > 0, "empty AIR program");
    NLC_CHECK(n_batches > 0 && batches, "no trace commitments");
    NLC_CHECK(n_alphas >= 1 && n_alphas <= MAX_ALPHAS, "num_challenges out of range (1..8)");
    const int log_n = batches[0]->log_n, rate_bits = batches[0]->rate_bits;
    NLC_CHECK(qdb >= 0 && qdb <= rate_bits, "Having constraints of degree higher than the rate is not supported yet.");
    std::vector<const u64*> cols;
    for (int b = 0; b < n_batches; b++) {
        NLC_CHECK(batches[b] && batches[b]->ctx == ctx, "trace commitment belongs to another context");
        NLC_CHECK(batches[b]->log_n == log_n && batches[b]->rate_bits == rate_bits, "trace commitments of different shapes");
        NLC_CHECK(batches[b]->shard_log == 0 && batches[b]->owns_lde, "AIR evaluation needs whole (unsharded) commitments");
        const size_t N = batches[b]->n_local();
        for (size_t j = 0; j < batches[b]->W; j++) cols.push_back(batches[b]->lde + j * N);
    }
    validate_program(prog, cols.size(), false);
    const size_t size = (size_t)1 << (log_n + qdb);
    const u64 *x, *lf, *ll;
    quotient_tables_dev(ctx, log_n, qdb, &x, &lf, &ll);
    AirDevice dev(ctx, prog, cols, publics, globals, challenges, alphas, n_alphas, size, nullptr, 0, nullptr,
                  air_min_blocks(air_na(n_alphas)));
    AirArgs& A = dev.A;
    A.log_size = log_n + qdb;
    A.qdb = qdb;
    A.coset_x = x;
    A.lagrange_first = lf;
    A.lagrange_last = ll;
    A.last = gl::h_inv(gl::h_root_of_unity(log_n));
    A.out = out_dev;
    switch (air_na(n_alphas)) {
    case 1: NLC_LAUNCH(ctx, k_air_eval<1>, (unsigned)dev.blocks, AIR_THREADS, 0, A); break;
    case 2: NLC_LAUNCH(ctx, k_air_eval<2>, (unsigned)dev.blocks, AIR_THREADS, 0, A); break;
    case 3: NLC_LAUNCH(ctx, k_air_eval<3>, (unsigned)dev.blocks, AIR_THREADS, 0, A); break;
    case 4: NLC_LAUNCH(ctx, k_air_eval<4>, (unsigned)dev.blocks, AIR_THREADS, 0, A); break;
    default: NLC_LAUNCH(ctx, k_air_eval<MAX_ALPHAS>, (unsigned)dev.blocks, AIR_THREADS, 0, A); break;
    }
}
