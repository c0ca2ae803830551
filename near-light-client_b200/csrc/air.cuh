// air.cuh -- the constraint-program interpreter shared by the STARK (air.cu) and plonk (plonk.cu) quotient kernels:
// program operands on the device (AirArgs / AirDevice), host-side validation, and the per-point interpreter air_run.
// The bytecode is the recording of an AIR's / a gate set's `eval` against a recording parser (include/nlc_commit.h,
// nlc_air_program; starkyx air/parser.rs:6-75).
#pragma once
#include "batch.cuh"

namespace {

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
    u64* regs;                // (n_regs + 1) x n_threads scratch (register r of thread t at regs[r * n_threads + t])
    u64* out;                 // size x n_alphas row-major, natural order of i (the `quotient_values` before transpose)
};

constexpr int AIR_THREADS = 128;
constexpr int MAX_ALPHAS = 8;
// The interpreter kernels are compiled per accumulator count NA: NA = 1..4 hold exactly n_alphas accumulators (a starky
// or plonk proof has 2 or 3: 12 fewer live registers than the generic form),
// NA = MAX_ALPHAS serves 5..8 with a bound check per accumulator.
constexpr int air_na(size_t n_alphas) { return n_alphas <= 4 ? (int)n_alphas : MAX_ALPHAS; }
constexpr int air_min_blocks(int) { return 8; } // 64 registers per thread (measured: 10 CTAs of 48 registers are 8-15 % slower)

// The interpreter: runs the straight-line program for ONE point (leaf k, next row at leaf k_next); the virtual
// registers of the program live in a thread-interleaved scratch matrix (coalesced), the program counter is uniform
// across the grid.  CONSTRAINT ops fold into acc_j = acc_j * alpha_j + c (ConstraintConsumer, consumer.rs:54-78).
//
// The code it runs is the LOWERED form AirDevice makes of a validated program (lower_program): every register operand
// is already the element offset of that register's row in the scratch matrix (register r -> r * n_threads), so an
// operand fetch is one IMAD.WIDE + one LDG instead of a 64-bit multiply; a DOT_CONST whose terms name the accumulator
// first parks it in a slot of its own (device flag AIR_DEV_PARK_LAST), so that its term loop is plain loads, and odd
// term counts are padded with a zero coefficient.  Measured on the demo gate set (4698 words, 4320 of them dot-product
// terms): 379 k -> 230 k executed instructions per warp.
constexpr u32 AIR_DEV_PARK_LAST = 8; // device-only instruction flag (bit 3 of the flag byte), set by lower_program
constexpr u32 AIR_DEV_END = 0xff;     // device-only opcode: end of the lowered program
// d = a * b + c with a, b 32-bit: ONE IMAD.WIDE.U32 (the compiler's own rendering of `c += (u64)a * b` drags zero
// carry words along)
__device__ __forceinline__ u64 mad_wide(u32 a, u32 b, u64 c) {
    u64 d;
    asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(d) : "r"(a), "r"(b), "l"(c));
    return d;
}
// scratch-matrix accesses as explicit global-space instructions: `g` below is deliberately opaque to the compiler, which
// would otherwise fall back to generic loads
__device__ __forceinline__ u64 reg_ld(const u64* g, u32 off) {
    u64 v;
    asm volatile("ld.global.u64 %0, [%1];" : "=l"(v) : "l"(g + off) : "memory");
    return v;
}
__device__ __forceinline__ void reg_st(u64* g, u32 off, u64 v) {
    asm volatile("st.global.u64 [%0], %1;" ::"l"(g + off), "l"(v) : "memory");
}
template <int NA>
__device__ __forceinline__ void air_run(const AirArgs& A, u64* g, const size_t k, const size_t k_next,
                                        const u64 z_last, const u64 l_first, const u64 l_last, u64 (&acc)[NA]) {
    u64 last = 0; // the accumulator: the previous instruction's result (NLC_AIR_FLAG_*)
    // the thread's column of the scratch matrix as ONE opaque 64-bit value: an operand address is then a single
    // IMAD.WIDE (offset * 8 + g) instead of the re-associated (base + (offset + thread) * 8) in four instructions
    asm volatile("" : "+l"(g));
    // the lowered code carries, in the upper half of its first word, the distance to the next instruction (a DOT_CONST
    // skips its term words) and ends with an AIR_DEV_END word: no program counter, no bound check, no clamp -- the next
    // instruction word is on its way while this one executes
    // one instruction; `cur` points at its word (a DOT_CONST reads its terms from cur + 1)
    auto exec = [&](const uint4 ins, const uint4* const cur) {
        const u32 op = ins.x & 0xffu, flags = ins.x >> 8;
        u64 r;
        if (op >= NLC_AIR_ADD) { // register operands
            u64 ra = last, rb = last;
            if (!(flags & NLC_AIR_FLAG_A_LAST) && op != NLC_AIR_DOT_CONST) ra = reg_ld(g, ins.z);
            if (op <= NLC_AIR_MUL) {
                if (!(flags & NLC_AIR_FLAG_B_LAST)) rb = reg_ld(g, ins.w);
                r = op == NLC_AIR_MUL ? gl::mul(ra, rb) : op == NLC_AIR_ADD ? gl::dadd(ra, rb) : gl::dsub(ra, rb);
            } else if (op == NLC_AIR_DOT_CONST) { // ins.z terms in the following words; 96-bit products summed in two 64-bit lanes
                if (flags & AIR_DEV_PARK_LAST) reg_st(g, ins.w, last);
                u64 s0 = 0, s1 = 0;
                const uint4* __restrict__ tw = cur + 1;
                const u32 words = (ins.z + 1) >> 1;
#pragma unroll 2
                for (u32 t = 0; t < words; t++) {
                    const uint4 w = __ldg(tw + t);
                    const u64 v0 = reg_ld(g, w.x), v1 = reg_ld(g, w.z);
                    s0 = mad_wide((u32)v0, w.y, s0);
                    s1 = mad_wide((u32)(v0 >> 32), w.y, s1);
                    s0 = mad_wide((u32)v1, w.w, s0);
                    s1 = mad_wide((u32)(v1 >> 32), w.w, s1);
                }
                // s0 + 2^32 s1 with s0, s1 < 2^60 (<= 2^12 terms, constants < 2^16):  2^32 s1 = (s1_lo << 32) + s1_hi 2^64
                const unsigned __int128 T = (unsigned __int128)s0 + ((u64)(u32)s1 << 32) + (s1 >> 32) * gl::EPS;
                r = gl::dcanon(gl::reduce128_dev_nc((u64)T, (u64)(T >> 64)));
            } else if (op >= NLC_AIR_CONSTRAINT && op <= NLC_AIR_CONSTRAINT_LAST_ROW) {
                // ConstraintConsumer (consumer.rs:54-78): filter, then acc_j = acc_j * alpha_j + c
                u64 c = ra;
                if (op == NLC_AIR_CONSTRAINT_TRANSITION) c = gl::mul(c, z_last);
                else if (op == NLC_AIR_CONSTRAINT_FIRST_ROW) c = gl::mul(c, l_first);
                else if (op == NLC_AIR_CONSTRAINT_LAST_ROW) c = gl::mul(c, l_last);
#pragma unroll
                for (int j = 0; j < NA; j++)
                    if (NA < MAX_ALPHAS || (u32)j < A.n_alphas) acc[j] = gl::dadd(gl::mul(acc[j], A.alphas[j]), c);
                return;
            } else if (op == NLC_AIR_NEG) {
                r = gl::neg(ra);
            } else { // one register and one pooled constant (the same for the whole grid: the branches are uniform)
                const u64 c = A.pool[ins.w];
                switch (op) {
                case NLC_AIR_MUL_CONST: r = (c >> 32) ? gl::mul(ra, c) : gl::mul_u32(ra, (u32)c); break;
                case NLC_AIR_ADD_CONST: r = gl::dadd(ra, c); break;
                case NLC_AIR_SUB_CONST: r = gl::dsub(ra, c); break;
                default: r = gl::dadd((c >> 32) ? gl::mul(ra, c) : gl::mul_u32(ra, (u32)c), last); break; // MULADD_CONST
                }
            }
        } else {
            switch (op) {
            case NLC_AIR_LOCAL: r = gl::dcanon(A.cols[ins.z][k]); break;
            case NLC_AIR_NEXT: r = gl::dcanon(A.cols[ins.z][k_next]); break;
            case NLC_AIR_PUBLIC: r = A.publics[ins.z]; break;
            case NLC_AIR_GLOBAL: r = A.globals[ins.z]; break;
            case NLC_AIR_CHALLENGE: r = A.challenges[ins.z]; break;
            default: r = A.pool[ins.z]; break; // NLC_AIR_CONST
            }
        }
        if (!(flags & NLC_AIR_FLAG_NO_STORE)) reg_st(g, ins.y, r);
        last = r;
    };
    // the dispatch loop, two instructions per trip so that the instruction words alternate between two register sets
    // instead of being copied (nxt -> ins) every time
    const uint4* cp = reinterpret_cast<const uint4*>(A.code);
    uint4 wa = __ldg(cp), wb;
    for (;;) {
        if ((wa.x & 0xffu) == AIR_DEV_END) break;
        const uint4* cur = cp;
        cp += wa.x >> 16;
        wb = __ldg(cp);
        exec(wa, cur);
        if ((wb.x & 0xffu) == AIR_DEV_END) break;
        cur = cp;
        cp += wb.x >> 16;
        wa = __ldg(cp);
        exec(wb, cur);
    }
}

// validate a program on the host: every operand in range, every source register written before it is read
void validate_program(const nlc_air_program* prog, size_t n_cols, bool plain_constraints_only) {
    NLC_CHECK(prog && (prog->code || prog->n_ops == 0), "null AIR program");
    std::vector<char> written(prog->n_regs, 0);
    bool have_last = false;
    for (u32 pc = 0; pc < prog->n_ops; pc++) {
        const u32 word = prog->code[4 * pc], op = word & 0xffu, flags = word >> 8;
        const u32 dst = prog->code[4 * pc + 1], a = prog->code[4 * pc + 2], b = prog->code[4 * pc + 3];
        NLC_CHECK((flags & ~7u) == 0, "AIR program: unknown instruction flag");
        auto reg = [&](u32 r, u32 last_flag) {
            if (flags & last_flag) {
                NLC_CHECK(have_last, "AIR program reads the accumulator before any result exists");
                return;
            }
            NLC_CHECK(r < prog->n_regs && written[r], "AIR program reads a register before writing it");
        };
        switch (op) {
        case NLC_AIR_NEXT: NLC_CHECK(!plain_constraints_only, "gate constraints read the local row only");
        // fall through
        case NLC_AIR_LOCAL: NLC_CHECK(a < n_cols, "AIR program: trace column out of range"); break;
        case NLC_AIR_PUBLIC: NLC_CHECK(a < prog->n_public, "AIR program: public input out of range"); break;
        case NLC_AIR_GLOBAL: NLC_CHECK(a < prog->n_global, "AIR program: global value out of range"); break;
        case NLC_AIR_CHALLENGE: NLC_CHECK(a < prog->n_challenge, "AIR program: challenge out of range"); break;
        case NLC_AIR_CONST: NLC_CHECK(a < prog->n_pool, "AIR program: constant out of range"); break;
        case NLC_AIR_ADD: case NLC_AIR_SUB: case NLC_AIR_MUL: reg(a, NLC_AIR_FLAG_A_LAST); reg(b, NLC_AIR_FLAG_B_LAST); break;
        case NLC_AIR_NEG: reg(a, NLC_AIR_FLAG_A_LAST); break;
        case NLC_AIR_DOT_CONST: {
            const u32 words = (a + 1) / 2;
            NLC_CHECK(a >= 1 && a <= (1u << 12) && pc + words < prog->n_ops + 0u && pc + words >= pc, "AIR program: dot product runs past the end");
            for (u32 t = 0; t < a; t++) {
                const u32 r = prog->code[4 * (pc + 1 + t / 2) + 2 * (t & 1)], cst = prog->code[4 * (pc + 1 + t / 2) + 2 * (t & 1) + 1];
                NLC_CHECK(cst < (1u << 16), "AIR program: dot-product constant too large");
                if (r & NLC_AIR_DOT_TERM_LAST) NLC_CHECK(have_last, "AIR program reads the accumulator before any result exists");
                else NLC_CHECK(r < prog->n_regs && written[r], "AIR program reads a register before writing it");
            }
            pc += words;
            break;
        }
        case NLC_AIR_MULADD_CONST: NLC_CHECK(have_last, "AIR program reads the accumulator before any result exists");
        // fall through
        case NLC_AIR_MUL_CONST: case NLC_AIR_ADD_CONST: case NLC_AIR_SUB_CONST:
            reg(a, NLC_AIR_FLAG_A_LAST); NLC_CHECK(b < prog->n_pool, "AIR program: constant out of range"); break;
        case NLC_AIR_CONSTRAINT_TRANSITION: case NLC_AIR_CONSTRAINT_FIRST_ROW: case NLC_AIR_CONSTRAINT_LAST_ROW:
            NLC_CHECK(!plain_constraints_only, "gate constraints carry no row filters");
        // fall through
        case NLC_AIR_CONSTRAINT: reg(a, NLC_AIR_FLAG_A_LAST); continue;
        default: NLC_CHECK(false, "AIR program: unknown opcode");
        }
        if (!(flags & NLC_AIR_FLAG_NO_STORE)) {
            NLC_CHECK(dst < prog->n_regs, "AIR program: destination register out of range");
            written[dst] = 1;
        }
        have_last = true;
    }
}

// The device form of a VALIDATED program for a grid of n_threads (see air_run): register operands become row offsets
// into the scratch matrix, accumulator terms of a DOT_CONST go through the parking slot (row n_regs), odd term counts
// are padded.  A renaming of operands: results are unchanged.
inline std::vector<u32> lower_program(const nlc_air_program* prog, size_t n_threads) {
    NLC_CHECK(((size_t)prog->n_regs + 1) * n_threads < ((size_t)1 << 32), "AIR program: register file too large for 32-bit offsets");
    std::vector<u32> code(prog->code, prog->code + (size_t)prog->n_ops * 4);
    code.insert(code.end(), {AIR_DEV_END | (1u << 16), 0u, 0u, 0u});
    const u32 nt = (u32)n_threads, park = prog->n_regs * nt;
    for (u32 pc = 0; pc < prog->n_ops; pc++) {
        u32* w = code.data() + 4 * (size_t)pc;
        const u32 op = w[0] & 0xffu, flags = w[0] >> 8;
        w[0] |= (op == NLC_AIR_DOT_CONST ? 1u + (w[2] + 1) / 2 : 1u) << 16; // distance to the next instruction (<= 2049 words)
        if (op < NLC_AIR_ADD) { // loads: a = column / table index
            w[1] *= nt;
            continue;
        }
        if (op == NLC_AIR_DOT_CONST) {
            const u32 terms = w[2], words = (terms + 1) / 2;
            bool parks = false;
            u32 first = 0;
            for (u32 t = 0; t < 2 * words; t++) {
                u32* term = code.data() + 4 * (size_t)(pc + 1 + t / 2) + 2 * (t & 1);
                if (t >= terms) { // padding: any readable slot, coefficient 0
                    term[0] = first;
                    term[1] = 0;
                } else if (term[0] & NLC_AIR_DOT_TERM_LAST) {
                    term[0] = park;
                    parks = true;
                } else
                    term[0] *= nt;
                if (t == 0) first = term[0];
            }
            if (parks) w[0] |= AIR_DEV_PARK_LAST << 8;
            w[1] *= nt;
            w[3] = park;
            pc += words;
            continue;
        }
        if (!(flags & NLC_AIR_FLAG_A_LAST)) w[2] *= nt;
        if (op <= NLC_AIR_MUL && !(flags & NLC_AIR_FLAG_B_LAST)) w[3] *= nt;
        if (!(op >= NLC_AIR_CONSTRAINT && op <= NLC_AIR_CONSTRAINT_LAST_ROW)) w[1] *= nt;
    }
    return code;
}

// A program and its small operands resident on the device for one launch; frees in stream order on destruction.
struct AirDevice {
    nlc_ctx* ctx;
    u64 *code_dev = nullptr, *small = nullptr, *cols_dev = nullptr, *regs = nullptr;
    AirArgs A{};
    size_t blocks = 0;
    AirDevice(nlc_ctx* c, const nlc_air_program* prog, const std::vector<const u64*>& cols, const u64* publics, const u64* globals,
              const u64* challenges, const u64* alphas, size_t n_alphas, size_t n_points, const u64* extra = nullptr, size_t n_extra = 0,
              const u64** extra_dev = nullptr, int ctas_per_sm = 8)
        : ctx(c) {
        blocks = (n_points + AIR_THREADS - 1) / AIR_THREADS;
        const size_t full = (size_t)ctx->sm_count * ctas_per_sm;
        if (blocks > full) blocks = full;
        // the scratch matrix is (n_regs + 1) x threads words: a program with very many live values runs on fewer CTAs
        // (the kernels stride over the points) rather than on more memory -- at most NLC_AIR_SCRATCH_MB (default 8192),
        // which also keeps every row offset of the lowered program inside 32 bits
        const char* e = getenv("NLC_AIR_SCRATCH_MB");
        const size_t budget_words = ((size_t)std::min<long>(std::max<long>(e ? atol(e) : 8192, 1), 16384) << 20) / sizeof(u64);
        const size_t fit = budget_words / ((size_t)prog->n_regs + 1) / AIR_THREADS;
        NLC_CHECK(fit >= 1, "AIR program: too many registers for the scratch budget (NLC_AIR_SCRATCH_MB)");
        if (blocks > fit) blocks = fit;
        const size_t n_threads = blocks * AIR_THREADS;
        const size_t n_small = prog->n_pool + prog->n_public + prog->n_global + prog->n_challenge + n_alphas + n_extra;
        try {
            code_dev = ctx->dmalloc(((size_t)prog->n_ops * 4 + 1) / 2 + 2);
            small = ctx->dmalloc(n_small + 1);
            cols_dev = ctx->dmalloc(cols.size() + 1);
            regs = ctx->dmalloc(((size_t)prog->n_regs + 1) * n_threads); // + the parking slot of lower_program
            // pageable sources: the runtime stages them before returning
            const std::vector<u32> code = lower_program(prog, n_threads);
            NLC_CUDA(cudaMemcpyAsync(code_dev, code.data(), code.size() * sizeof(u32), cudaMemcpyHostToDevice, ctx->stream));
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
                         o_alpha = put(alphas, n_alphas), o_extra = put(extra, n_extra);
            NLC_CUDA(cudaMemcpyAsync(small, host.data(), host.size() * sizeof(u64), cudaMemcpyHostToDevice, ctx->stream));
            if (!cols.empty())
                NLC_CUDA(cudaMemcpyAsync(cols_dev, cols.data(), cols.size() * sizeof(u64), cudaMemcpyHostToDevice, ctx->stream));
            NLC_CUDA(cudaStreamSynchronize(ctx->stream)); // `host` / `cols` are stack-lifetime staging sources
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
            A.regs = regs;
            if (extra_dev) *extra_dev = small + o_extra;
        } catch (...) {
            release();
            throw;
        }
    }
    void release() {
        ctx->dfree(regs);
        ctx->dfree(cols_dev);
        ctx->dfree(small);
        ctx->dfree(code_dev);
        regs = cols_dev = small = code_dev = nullptr;
    }
    ~AirDevice() { release(); }
    AirDevice(const AirDevice&) = delete;
    AirDevice& operator=(const AirDevice&) = delete;
};

} // namespace
