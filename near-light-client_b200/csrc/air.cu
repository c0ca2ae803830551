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
#include "batch.cuh"

namespace {

enum { KIND_QX = 10, KIND_QLF = 11, KIND_QLL = 12 };

struct AirArgs {
    const u32* code;          // n_ops x 4 words {op, dst, a, b}
    u32 n_ops;
    const u64* pool;          // 64-bit constants
    const u64* const* cols;   // device pointers to the columns of the concatenated trace (lde + j*N of every batch)
    const u64* publics;
    const u64* globals;
    const u64* challenges;
    const u64* alphas;
    u32 n_alphas;
    u32 n_regs;
    int log_size;             // log2(quotient domain) = log_n + quotient_degree_bits
    int qdb;
    const u64* coset_x;       // g * w_size^i, natural order
    const u64* lagrange_first;
    const u64* lagrange_last;
    u64 last;                 // w_n^-1: the last point of the subgroup (z_last = x - last)
    u64* regs;                // n_regs x n_threads scratch (register r of thread t at regs[r * n_threads + t])
    u64* out;                 // size x n_alphas row-major, natural order of i (the `quotient_values` before transpose)
};

constexpr int AIR_THREADS = 128;
constexpr int MAX_ALPHAS = 8;

// One thread per leaf k of the quotient-domain prefix; the virtual registers of the program live in a
// thread-interleaved scratch matrix (coalesced), the program counter is uniform across the grid.
__global__ void __launch_bounds__(AIR_THREADS) k_air_eval(const AirArgs A) {
    const size_t size = (size_t)1 << A.log_size;
    const size_t stride = (size_t)gridDim.x * AIR_THREADS;
    u64* regs = A.regs + (size_t)blockIdx.x * AIR_THREADS + threadIdx.x;
    for (size_t k = (size_t)blockIdx.x * AIR_THREADS + threadIdx.x; k < size; k += stride) {
        const size_t i = A.log_size ? (__brevll(k) >> (64 - A.log_size)) : 0;
        const size_t i_next = (i + ((size_t)1 << A.qdb)) & (size - 1);
        const size_t k_next = A.log_size ? (__brevll(i_next) >> (64 - A.log_size)) : 0;
        const u64 z_last = gl::sub(A.coset_x[i], A.last);
        const u64 l_first = A.lagrange_first[i], l_last = A.lagrange_last[i];
        u64 acc[MAX_ALPHAS];
#pragma unroll
        for (int j = 0; j < MAX_ALPHAS; j++) acc[j] = 0;
        for (u32 pc = 0; pc < A.n_ops; pc++) {
            const uint4 ins = __ldg(reinterpret_cast<const uint4*>(A.code) + pc);
            const u32 op = ins.x;
            u64 r;
            switch (op) {
            case NLC_AIR_LOCAL: r = gl::dcanon(A.cols[ins.z][k]); break;
            case NLC_AIR_NEXT: r = gl::dcanon(A.cols[ins.z][k_next]); break;
            case NLC_AIR_PUBLIC: r = A.publics[ins.z]; break;
            case NLC_AIR_GLOBAL: r = A.globals[ins.z]; break;
            case NLC_AIR_CHALLENGE: r = A.challenges[ins.z]; break;
            case NLC_AIR_CONST: r = A.pool[ins.z]; break;
            case NLC_AIR_ADD: r = gl::dadd(regs[ins.z * stride], regs[ins.w * stride]); break;
            case NLC_AIR_SUB: r = gl::dsub(regs[ins.z * stride], regs[ins.w * stride]); break;
            case NLC_AIR_MUL: r = gl::mul(regs[ins.z * stride], regs[ins.w * stride]); break;
            case NLC_AIR_NEG: r = gl::neg(regs[ins.z * stride]); break;
            case NLC_AIR_MUL_CONST: r = gl::mul(regs[ins.z * stride], A.pool[ins.w]); break;
            case NLC_AIR_ADD_CONST: r = gl::dadd(regs[ins.z * stride], A.pool[ins.w]); break;
            case NLC_AIR_SUB_CONST: r = gl::dsub(regs[ins.z * stride], A.pool[ins.w]); break;
            default: { // ConstraintConsumer (consumer.rs:54-78): filter, then acc_j = acc_j * alpha_j + c
                u64 c = regs[ins.z * stride];
                if (op == NLC_AIR_CONSTRAINT_TRANSITION) c = gl::mul(c, z_last);
                else if (op == NLC_AIR_CONSTRAINT_FIRST_ROW) c = gl::mul(c, l_first);
                else if (op == NLC_AIR_CONSTRAINT_LAST_ROW) c = gl::mul(c, l_last);
#pragma unroll
                for (int j = 0; j < MAX_ALPHAS; j++)
                    if ((u32)j < A.n_alphas) acc[j] = gl::dadd(gl::mul(acc[j], A.alphas[j]), c);
                continue;
            }
            }
            regs[ins.y * stride] = r;
        }
#pragma unroll
        for (int j = 0; j < MAX_ALPHAS; j++)
            if ((u32)j < A.n_alphas) A.out[i * A.n_alphas + j] = acc[j];
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
    // validate the program on the host: every operand in range, every source register written before it is read
    {
        std::vector<char> written(prog->n_regs, 0);
        for (u32 pc = 0; pc < prog->n_ops; pc++) {
            const u32 op = prog->code[4 * pc], dst = prog->code[4 * pc + 1], a = prog->code[4 * pc + 2], b = prog->code[4 * pc + 3];
            auto reg = [&](u32 r) { NLC_CHECK(r < prog->n_regs && written[r], "AIR program reads a register before writing it"); };
            switch (op) {
            case NLC_AIR_LOCAL: case NLC_AIR_NEXT: NLC_CHECK(a < cols.size(), "AIR program: trace column out of range"); break;
            case NLC_AIR_PUBLIC: NLC_CHECK(a < prog->n_public, "AIR program: public input out of range"); break;
            case NLC_AIR_GLOBAL: NLC_CHECK(a < prog->n_global, "AIR program: global value out of range"); break;
            case NLC_AIR_CHALLENGE: NLC_CHECK(a < prog->n_challenge, "AIR program: challenge out of range"); break;
            case NLC_AIR_CONST: NLC_CHECK(a < prog->n_pool, "AIR program: constant out of range"); break;
            case NLC_AIR_ADD: case NLC_AIR_SUB: case NLC_AIR_MUL: reg(a); reg(b); break;
            case NLC_AIR_NEG: reg(a); break;
            case NLC_AIR_MUL_CONST: case NLC_AIR_ADD_CONST: case NLC_AIR_SUB_CONST:
                reg(a); NLC_CHECK(b < prog->n_pool, "AIR program: constant out of range"); break;
            case NLC_AIR_CONSTRAINT: case NLC_AIR_CONSTRAINT_TRANSITION: case NLC_AIR_CONSTRAINT_FIRST_ROW:
            case NLC_AIR_CONSTRAINT_LAST_ROW: reg(a); continue;
            default: NLC_CHECK(false, "AIR program: unknown opcode");
            }
            NLC_CHECK(dst < prog->n_regs, "AIR program: destination register out of range");
            written[dst] = 1;
        }
    }
    const size_t size = (size_t)1 << (log_n + qdb);
    const u64 *x, *lf, *ll;
    quotient_tables_dev(ctx, log_n, qdb, &x, &lf, &ll);
    size_t blocks = (size + AIR_THREADS - 1) / AIR_THREADS;
    const size_t full = (size_t)ctx->sm_count * 8;
    if (blocks > full) blocks = full;
    const size_t n_threads = blocks * AIR_THREADS;
    const size_t n_small = prog->n_pool + prog->n_public + prog->n_global + prog->n_challenge + n_alphas;
    u64* code_dev = ctx->dmalloc(((size_t)prog->n_ops * 4 + 1) / 2 + 2);
    u64* small = ctx->dmalloc(n_small + 1);
    u64* cols_dev = ctx->dmalloc(cols.size());
    u64* regs = ctx->dmalloc(std::max<size_t>((size_t)prog->n_regs * n_threads, 1));
    try {
        // pageable sources: the runtime stages them before returning
        NLC_CUDA(cudaMemcpyAsync(code_dev, prog->code, (size_t)prog->n_ops * 16, cudaMemcpyHostToDevice, ctx->stream));
        std::vector<u64> host(n_small + 1, 0);
        size_t o = 0;
        auto put = [&](const u64* p, size_t cnt) {
            NLC_CHECK(p || !cnt, "null input array");
            for (size_t q = 0; q < cnt; q++) host[o + q] = gl::canon(p[q]);
            o += cnt;
            return o - cnt;
        };
        const size_t o_pool = put(prog->pool, prog->n_pool), o_pub = put(publics, prog->n_public),
                     o_glob = put(globals, prog->n_global), o_chal = put(challenges, prog->n_challenge),
                     o_alpha = put(alphas, n_alphas);
        NLC_CUDA(cudaMemcpyAsync(small, host.data(), host.size() * sizeof(u64), cudaMemcpyHostToDevice, ctx->stream));
        NLC_CUDA(cudaMemcpyAsync(cols_dev, cols.data(), cols.size() * sizeof(u64), cudaMemcpyHostToDevice, ctx->stream));
        NLC_CUDA(cudaStreamSynchronize(ctx->stream)); // `host` / `cols` are stack-lifetime staging sources
        AirArgs A;
        A.code = reinterpret_cast<const u32*>(code_dev);
        A.n_ops = prog->n_ops;
        A.pool = small + o_pool;
        A.cols = reinterpret_cast<const u64* const*>(cols_dev);
        A.publics = small + o_pub;
        A.globals = small + o_glob;
        A.challenges = small + o_chal;
        A.alphas = small + o_alpha;
        A.n_alphas = (u32)n_alphas;
        A.n_regs = prog->n_regs;
        A.log_size = log_n + qdb;
        A.qdb = qdb;
        A.coset_x = x;
        A.lagrange_first = lf;
        A.lagrange_last = ll;
        A.last = gl::h_inv(gl::h_root_of_unity(log_n));
        A.regs = regs;
        A.out = out_dev;
        NLC_LAUNCH(ctx, k_air_eval, (unsigned)blocks, AIR_THREADS, 0, A);
    } catch (...) {
        ctx->dfree(regs);
        ctx->dfree(cols_dev);
        ctx->dfree(small);
        ctx->dfree(code_dev);
        throw;
    }
    ctx->dfree(regs);
    ctx->dfree(cols_dev);
    ctx->dfree(small);
    ctx->dfree(code_dev);
}
