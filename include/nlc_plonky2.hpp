// nlc_plonky2.hpp -- the host side above the C ABI, in C++: plonky2's interface for the commitment hot path.
//
// The reference's host language is Rust (no cargo / rustc in this image), so the compiled-language mirror of
// the plonky2 API that the unchanged starkyx prover calls is this header-only C++ layer over include/nlc_commit.h:
// same type and method names, same argument meaning, and plonky2's error contract at this boundary
// (infallible-or-panic: a non-zero return code becomes a thrown plonky2::Panic carrying nlc_last_error).
// Call sites mirrored: /root/reference/vendor/starkyx/starkyx/src/plonky2/stark/prover.rs:50-225
// (from_values :91-93, observe_cap :94-100, from_coeffs :171-178, prove_openings :200-206,
// get_lde_values_packed :289-294, lde_onto_coset / coset_ifft :280-284,356) and
// plonky2/stark/config.rs:50-94 (FriConfig knobs, fri_params).  The Rust binding of the same symbols is in
// INTEGRATION.md; tests/cpp/test_plonky2_mirror.cpp drives this header the way prover.rs drives plonky2.
#pragma once
#include <algorithm>
#include <cstdint>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "nlc_commit.h"

namespace plonky2 {

using F = uint64_t;                 // GoldilocksField, #[repr(transparent)] over u64
using FE = std::pair<F, F>;         // QuadraticExtension<GoldilocksField> (a0, a1), W = 7
constexpr F ORDER = 0xFFFFFFFF00000001ULL;
constexpr F COSET_SHIFT = 7;        // F::coset_shift() == MULTIPLICATIVE_GROUP_GENERATOR

struct Panic : std::runtime_error {
    using std::runtime_error::runtime_error;
};

inline size_t log2_strict(size_t n) {
    size_t l = 0;
    while (((size_t)1 << l) < n) l++;
    if (((size_t)1 << l) != n) throw Panic("Not a power of two: " + std::to_string(n)); // plonky2_util::log2_strict
    return l;
}

// One nlc_ctx per (thread, device); rayon workers / tokio blocking threads each get their own.
class Context {
  public:
    explicit Context(int device = -1) {
        if (nlc_init(device, &ctx_) != 0) throw Panic(std::string("nlc_init: ") + nlc_last_error(nullptr));
    }
    ~Context() { nlc_destroy(ctx_); }
    Context(const Context&) = delete;
    Context& operator=(const Context&) = delete;
    nlc_ctx* raw() const { return ctx_; }
    void check(int rc) const {
        if (rc != 0) throw Panic(std::string("nlc_commit: ") + nlc_last_error(ctx_));
    }
    static Context& current() {
        thread_local Context c;
        return c;
    }

  private:
    nlc_ctx* ctx_ = nullptr;
};

// ---- field/src/polynomial/mod.rs -------------------------------------------------------------------------
struct PolynomialCoeffs;
struct PolynomialValues {
    std::vector<F> values;
    PolynomialValues() = default;
    explicit PolynomialValues(std::vector<F> v) : values(std::move(v)) {}
    size_t len() const { return values.size(); }
    static PolynomialValues selector(size_t len, size_t index) {
        std::vector<F> v(len, 0);
        v.at(index) = 1;
        return PolynomialValues(std::move(v));
    }
    inline PolynomialCoeffs ifft() const;
    inline PolynomialCoeffs coset_ifft(F shift) const;
    // evaluations on the coset shift * <w_N>, N = len << rate_bits (prover.rs:281-284)
    PolynomialValues lde_onto_coset(size_t rate_bits) const {
        Context& c = Context::current();
        std::vector<F> out(values.size() << rate_bits);
        c.check(nlc_lde_onto_coset(c.raw(), values.data(), out.data(), 1, (int)log2_strict(values.size()), (int)rate_bits));
        return PolynomialValues(std::move(out));
    }
};

struct PolynomialCoeffs {
    std::vector<F> coeffs;
    PolynomialCoeffs() = default;
    explicit PolynomialCoeffs(std::vector<F> c) : coeffs(std::move(c)) {}
    size_t len() const { return coeffs.size(); }
    PolynomialValues fft() const {
        Context& c = Context::current();
        std::vector<F> d = coeffs;
        c.check(nlc_fft(c.raw(), d.data(), 1, (int)log2_strict(d.size())));
        return PolynomialValues(std::move(d));
    }
    PolynomialValues coset_fft(F shift) const {
        Context& c = Context::current();
        std::vector<F> d = coeffs;
        c.check(nlc_coset_fft(c.raw(), d.data(), 1, (int)log2_strict(d.size()), shift));
        return PolynomialValues(std::move(d));
    }
    PolynomialCoeffs lde(size_t rate_bits) const { // zero-pad to len << rate_bits
        std::vector<F> d = coeffs;
        d.resize(coeffs.size() << rate_bits, 0);
        return PolynomialCoeffs(std::move(d));
    }
    // Split into degree-n chunks (prover.rs:166: quotient_poly.chunks(degree))
    std::vector<PolynomialCoeffs> chunks(size_t chunk_size) const {
        std::vector<PolynomialCoeffs> out;
        for (size_t o = 0; o < coeffs.size(); o += chunk_size)
            out.emplace_back(std::vector<F>(coeffs.begin() + o, coeffs.begin() + std::min(o + chunk_size, coeffs.size())));
        return out;
    }
};

inline PolynomialCoeffs PolynomialValues::ifft() const {
    Context& c = Context::current();
    std::vector<F> d = values;
    c.check(nlc_ifft(c.raw(), d.data(), 1, (int)log2_strict(d.size())));
    return PolynomialCoeffs(std::move(d));
}
inline PolynomialCoeffs PolynomialValues::coset_ifft(F shift) const {
    Context& c = Context::current();
    std::vector<F> d = values;
    c.check(nlc_coset_ifft(c.raw(), d.data(), 1, (int)log2_strict(d.size()), shift));
    return PolynomialCoeffs(std::move(d));
}

// ---- hash/hash_types.rs, hash/poseidon.rs, hash/hashing.rs -------------------------------------------------
struct HashOut {
    F elements[4];
    bool operator==(const HashOut& o) const {
        return elements[0] == o.elements[0] && elements[1] == o.elements[1] && elements[2] == o.elements[2] &&
               elements[3] == o.elements[3];
    }
};
struct MerkleCap {
    std::vector<HashOut> hashes; // MerkleCap(pub Vec<H::Hash>)
    size_t len() const { return hashes.size(); }
    size_t height() const { return log2_strict(hashes.size()); }
    std::vector<F> flatten() const {
        std::vector<F> v;
        for (const HashOut& h : hashes) v.insert(v.end(), h.elements, h.elements + 4);
        return v;
    }
};
struct MerkleProof {
    std::vector<HashOut> siblings; // bottom-up
};

struct PoseidonPermutation {
    static void permute(F (&state)[12]) {
        Context& c = Context::current();
        c.check(nlc_poseidon_permute(c.raw(), state, 1));
    }
};
struct PoseidonHash {
    static HashOut hash_no_pad(const std::vector<F>& input) {
        Context& c = Context::current();
        HashOut h;
        c.check(nlc_hash_no_pad(c.raw(), input.data(), 1, input.size(), h.elements));
        return h;
    }
    static HashOut hash_or_noop(const std::vector<F>& input) {
        Context& c = Context::current();
        HashOut h;
        c.check(nlc_hash_or_noop(c.raw(), input.data(), 1, input.size(), h.elements));
        return h;
    }
    static HashOut two_to_one(const HashOut& left, const HashOut& right) {
        Context& c = Context::current();
        HashOut h;
        c.check(nlc_two_to_one(c.raw(), left.elements, right.elements, 1, h.elements));
        return h;
    }
};

// ---- hash/merkle_tree.rs -----------------------------------------------------------------------------------
class MerkleTree {
  public:
    MerkleCap cap;
    // MerkleTree::new(leaves, cap_height): leaves of equal length, a power-of-two number of them
    MerkleTree(const std::vector<std::vector<F>>& leaves, size_t cap_height) {
        if (leaves.empty()) throw Panic("MerkleTree::new: no leaves");
        leaf_len_ = leaves[0].size();
        n_leaves_ = leaves.size();
        std::vector<F> flat;
        flat.reserve(n_leaves_ * leaf_len_);
        for (const auto& l : leaves) {
            if (l.size() != leaf_len_) throw Panic("MerkleTree::new: ragged leaves");
            flat.insert(flat.end(), l.begin(), l.end());
        }
        Context& c = Context::current();
        nlc_tree* t = nullptr;
        c.check(nlc_merkle_new(c.raw(), flat.data(), n_leaves_, leaf_len_, (int)cap_height, &t));
        tree_.reset(t, nlc_tree_free);
        cap.hashes.resize((size_t)1 << cap_height);
        c.check(nlc_tree_cap(t, cap.hashes[0].elements));
        cap_height_ = cap_height;
    }
    std::vector<F> get(size_t i) const {
        std::vector<F> leaf(leaf_len_);
        Context::current().check(nlc_tree_open(tree_.get(), i, leaf.data(), nullptr));
        return leaf;
    }
    MerkleProof prove(size_t leaf_index) const {
        MerkleProof p;
        p.siblings.resize(log2_strict(n_leaves_) - cap_height_);
        Context::current().check(nlc_tree_open(tree_.get(), leaf_index, nullptr,
                                               p.siblings.empty() ? nullptr : p.siblings[0].elements));
        return p;
    }
    std::vector<HashOut> digests() const {
        std::vector<HashOut> d(2 * (n_leaves_ - ((size_t)1 << cap_height_)));
        if (!d.empty()) Context::current().check(nlc_tree_digests(tree_.get(), d[0].elements));
        return d;
    }

  private:
    std::shared_ptr<nlc_tree> tree_;
    size_t n_leaves_ = 0, leaf_len_ = 0, cap_height_ = 0;
};

// ---- iop/challenger.rs ---------------------------------------------------------------------------------------
class Challenger {
  public:
    Challenger() { nlc_challenger_init(&c_); }
    void observe_element(F e) { observe_elements({e}); }
    void observe_elements(const std::vector<F>& elems) {
        Context& c = Context::current();
        if (!elems.empty()) c.check(nlc_challenger_observe(c.raw(), &c_, elems.data(), elems.size()));
    }
    void observe_extension_element(const FE& e) { observe_elements({e.first, e.second}); }
    void observe_extension_elements(const std::vector<FE>& es) {
        std::vector<F> flat;
        for (const FE& e : es) flat.push_back(e.first), flat.push_back(e.second);
        observe_elements(flat);
    }
    void observe_hash(const HashOut& h) { observe_elements(std::vector<F>(h.elements, h.elements + 4)); }
    void observe_cap(const MerkleCap& cap) { observe_elements(cap.flatten()); }
    F get_challenge() { return get_n_challenges(1)[0]; }
    std::vector<F> get_n_challenges(size_t n) {
        Context& c = Context::current();
        std::vector<F> out(n);
        if (n) c.check(nlc_challenger_get(c.raw(), &c_, n, out.data()));
        return out;
    }
    FE get_extension_challenge() {
        auto v = get_n_challenges(2);
        return {v[0], v[1]};
    }
    nlc_challenger* raw() { return &c_; }

  private:
    nlc_challenger c_;
};

// ---- fri/mod.rs, fri/reduction_strategies.rs, fri/structure.rs, fri/proof.rs ---------------------------------
struct FriReductionStrategy {
    int arity_bits = 4, final_poly_bits = 5;
    static FriReductionStrategy ConstantArityBits(int arity_bits, int final_poly_bits) {
        return FriReductionStrategy{arity_bits, final_poly_bits};
    }
};
struct FriConfig;
struct FriParams {
    nlc_fri_params c{}; // config { rate_bits, cap_height, proof_of_work_bits, num_query_rounds }, degree_bits, arities
    bool hiding = false;
    size_t degree_bits() const { return (size_t)c.degree_bits; }
    size_t lde_bits() const { return (size_t)(c.degree_bits + c.rate_bits); }
    size_t lde_size() const { return (size_t)1 << lde_bits(); }
    std::vector<size_t> reduction_arity_bits() const {
        return std::vector<size_t>(c.reduction_arity_bits, c.reduction_arity_bits + c.n_reduction_rounds);
    }
    size_t final_poly_len() const {
        size_t s = 0;
        for (int i = 0; i < c.n_reduction_rounds; i++) s += (size_t)c.reduction_arity_bits[i];
        return (size_t)1 << ((size_t)c.degree_bits - s);
    }
};
struct FriConfig {
    size_t rate_bits, cap_height, proof_of_work_bits;
    FriReductionStrategy reduction_strategy;
    size_t num_query_rounds;
    // starkyx StarkyConfig::standard_fast_config (plonky2/stark/config.rs:58-73)
    static FriConfig standard_fast_config() { return {1, 4, 16, FriReductionStrategy::ConstantArityBits(4, 5), 84}; }
    // plonky2 CircuitConfig::standard_recursion_config().fri_config
    static FriConfig standard_recursion_config() { return {3, 4, 16, FriReductionStrategy::ConstantArityBits(4, 5), 28}; }
    FriParams fri_params(size_t degree_bits, bool hiding) const { // config.rs:75-77
        if (hiding) throw Panic("hiding / blinding is not used on this path (all in-tree call sites pass false)");
        FriParams p;
        if (nlc_fri_params_constant_arity(&p.c, (int)degree_bits, (int)rate_bits, (int)cap_height,
                                          (int)proof_of_work_bits, (int)num_query_rounds,
                                          reduction_strategy.arity_bits, reduction_strategy.final_poly_bits) != 0)
            throw Panic("bad FriConfig");
        return p;
    }
};
struct FriPolynomialInfo {
    size_t oracle_index, polynomial_index;
    static std::vector<FriPolynomialInfo> from_range(size_t oracle_index, size_t begin, size_t end) {
        std::vector<FriPolynomialInfo> v;
        for (size_t i = begin; i < end; i++) v.push_back({oracle_index, i});
        return v;
    }
};
struct FriBatchInfo {
    FE point;
    std::vector<FriPolynomialInfo> polynomials;
};
struct FriOracleInfo {
    size_t num_polys;
    bool blinding;
};
struct FriInstanceInfo {
    std::vector<FriOracleInfo> oracles;
    std::vector<FriBatchInfo> batches;
};
struct FriInitialTreeProof {
    std::vector<std::pair<std::vector<F>, MerkleProof>> evals_proofs; // one per oracle
};
struct FriQueryStep {
    std::vector<FE> evals;
    MerkleProof merkle_proof;
};
struct FriQueryRound {
    FriInitialTreeProof initial_trees_proof;
    std::vector<FriQueryStep> steps;
};
struct FriProof {
    std::vector<MerkleCap> commit_phase_merkle_caps;
    std::vector<FriQueryRound> query_round_proofs;
    std::vector<FE> final_poly; // PolynomialCoeffs<F::Extension>
    F pow_witness = 0;
    std::vector<F> flat;        // the flat layout of include/nlc_commit.h (what crosses the ABI)
};

// ---- fri/oracle.rs: PolynomialBatch --------------------------------------------------------------------------
// The tables quotient_polys builds before its evaluation loop (prover.rs:278-302), computed on the device:
// coset[i] = g w^i over the quotient domain, lagrange_first / lagrange_last = selector(n, 0 | n-1)
// .lde_onto_coset(quotient_degree_bits), and ZeroPolyOnCoset::eval_inverse(i) = z_h_inverses[i % 2^bits].
struct QuotientDomain {
    std::vector<F> coset, z_h_inverses;
    PolynomialValues lagrange_first, lagrange_last;
    QuotientDomain(size_t degree_bits, size_t quotient_degree_bits)
        : coset((size_t)1 << (degree_bits + quotient_degree_bits)),
          z_h_inverses((size_t)1 << quotient_degree_bits),
          lagrange_first(std::vector<F>((size_t)1 << (degree_bits + quotient_degree_bits))),
          lagrange_last(std::vector<F>((size_t)1 << (degree_bits + quotient_degree_bits))) {
        Context& c = Context::current();
        c.check(nlc_quotient_domain(c.raw(), (int)degree_bits, (int)quotient_degree_bits, coset.data(),
                                    lagrange_first.values.data(), lagrange_last.values.data(), z_h_inverses.data()));
    }
    F eval_inverse(size_t i) const { return z_h_inverses[i % z_h_inverses.size()]; }
};

class PolynomialBatch {
  public:
    struct Tree { // the fields of merkle_tree the prover reads (prover.rs:95,180,210); leaves / digests on demand
        MerkleCap cap;
    } merkle_tree;
    size_t degree_log = 0, rate_bits = 0;
    bool blinding = false;

    // PolynomialBatch::from_values(values, rate_bits, blinding, cap_height, timing, fft_root_table)  prover.rs:91-93
    static PolynomialBatch from_values(const std::vector<PolynomialValues>& values, size_t rate_bits, bool blinding,
                                       size_t cap_height) {
        std::vector<const uint64_t*> cols;
        for (const auto& v : values) cols.push_back(v.values.data());
        return commit(cols, values.empty() ? 0 : values[0].len(), lens_equal(values), rate_bits, blinding, cap_height, true);
    }
    // PolynomialBatch::from_coeffs(polynomials, rate_bits, blinding, cap_height, timing, fft_root_table)  :171-178
    static PolynomialBatch from_coeffs(const std::vector<PolynomialCoeffs>& polynomials, size_t rate_bits, bool blinding,
                                       size_t cap_height) {
        std::vector<const uint64_t*> cols;
        for (const auto& p : polynomials) cols.push_back(p.coeffs.data());
        return commit(cols, polynomials.empty() ? 0 : polynomials[0].len(), lens_equal(polynomials), rate_bits, blinding,
                      cap_height, false);
    }
    // A PolynomialBatch read back from serialised circuit data (plonky2's Read::read_polynomial_batch; circuits are
    // serialised by /root/reference/scripts/build-circuit.sh:10-18): `polynomials` + the serialised MerkleCap.  The LDE
    // and the tree are recomputed on the GPU from the coefficients and the cap must match (panics otherwise).
    static PolynomialBatch from_serialized(const std::vector<PolynomialCoeffs>& polynomials, size_t rate_bits,
                                           size_t cap_height, const MerkleCap* cap = nullptr) {
        if (polynomials.empty()) throw Panic("PolynomialBatch: at least one polynomial required");
        if (!lens_equal(polynomials)) throw Panic("All polynomials in a batch must have the same length");
        if (cap && cap->hashes.size() != ((size_t)1 << cap_height)) throw Panic("serialised cap has the wrong size");
        std::vector<const uint64_t*> cols;
        for (const auto& p : polynomials) cols.push_back(p.coeffs.data());
        Context& c = Context::current();
        nlc_batch* h = nullptr;
        const int log_n = (int)log2_strict(polynomials[0].len());
        c.check(nlc_batch_import(c.raw(), cols.data(), cols.size(), log_n, (int)rate_bits, (int)cap_height,
                                 cap ? cap->hashes[0].elements : nullptr, nullptr, &h));
        PolynomialBatch b;
        b.h_.reset(h, nlc_batch_free);
        b.n_polys_ = cols.size();
        b.degree_log = (size_t)log_n;
        b.rate_bits = rate_bits;
        b.cap_height_ = cap_height;
        b.merkle_tree.cap.hashes.resize((size_t)1 << cap_height);
        c.check(nlc_batch_cap(h, b.merkle_tree.cap.hashes[0].elements));
        return b;
    }
    // SURVEY.md 8f.1 -- what prover.rs:336-357 and :158-178 do with the constraint evaluations: `quotient_values` is
    // the row-per-point Vec<Vec<F>> (num_challenges values per row, size = n << log2_ceil(quotient_degree_factor)
    // rows) flattened; per challenge coset_ifft(F::coset_shift()), trim_to_len(n * factor).expect("Quotient has
    // failed ..."), chunks(n), from_coeffs(chunks, rate_bits, false, cap_height).  divide_by_zh: rows are the raw
    // accumulators, multiplied by ZeroPolyOnCoset::eval_inverse first (prover.rs:336-340).
    static PolynomialBatch from_quotient_values(const std::vector<F>& quotient_values, size_t num_challenges,
                                                size_t degree_log, size_t quotient_degree_factor, size_t rate_bits,
                                                size_t cap_height, bool divide_by_zh = false) {
        size_t qdb = 0;
        while (((size_t)1 << qdb) < quotient_degree_factor) qdb++; // log2_ceil
        if (num_challenges == 0 || quotient_values.size() != (num_challenges << (degree_log + qdb)))
            throw Panic("quotient_values must hold (n << log2_ceil(quotient_degree_factor)) rows of num_challenges");
        Context& c = Context::current();
        nlc_batch* h = nullptr;
        c.check(nlc_quotient_commit(c.raw(), quotient_values.data(), num_challenges, (int)degree_log, (int)qdb,
                                    quotient_degree_factor, divide_by_zh ? 1 : 0, (int)rate_bits, (int)cap_height, &h));
        PolynomialBatch b;
        b.h_.reset(h, nlc_batch_free);
        b.n_polys_ = num_challenges * quotient_degree_factor;
        b.degree_log = degree_log;
        b.rate_bits = rate_bits;
        b.cap_height_ = cap_height;
        b.merkle_tree.cap.hashes.resize((size_t)1 << cap_height);
        c.check(nlc_batch_cap(h, b.merkle_tree.cap.hashes[0].elements));
        return b;
    }
    // from_values that also fills the PUBLIC struct fields a patched plonky2 keeps on the host -- `polynomials`,
    // merkle_tree.leaves (flat N x W), merkle_tree.digests -- while the commit runs.
    struct HostFields {
        std::vector<F> leaves;       // N x W row-major, row k = leaf k (index-bit-reversed LDE point)
        std::vector<HashOut> digests;
    };
    static PolynomialBatch from_values_materialised(const std::vector<PolynomialValues>& values, size_t rate_bits, bool blinding,
                                                    size_t cap_height, HostFields* fields = nullptr) {
        if (blinding) throw Panic("blinding = true is not used on this path (prover.rs:91-93,171-178 pass false)");
        if (values.empty()) throw Panic("PolynomialBatch: at least one polynomial required");
        if (!lens_equal(values)) throw Panic("All polynomials in a batch must have the same length");
        Context& c = Context::current();
        const size_t n = values[0].len(), W = values.size();
        const int log_n = (int)log2_strict(n);
        const size_t N = n << rate_bits;
        if (cap_height > (size_t)log_n + rate_bits) throw Panic("cap_height should be at most log2(leaves.len())");
        PolynomialBatch b;
        b.polys_.assign(W, PolynomialCoeffs(std::vector<F>(n)));
        std::vector<const uint64_t*> cols;
        std::vector<uint64_t*> cout;
        for (const auto& v : values) cols.push_back(v.values.data());
        for (auto& p : b.polys_) cout.push_back(p.coeffs.data());
        b.merkle_tree.cap.hashes.resize((size_t)1 << cap_height);
        if (fields) {
            fields->leaves.assign(N * W, 0);
            fields->digests.assign(2 * (N - ((size_t)1 << cap_height)), HashOut());
        }
        nlc_batch* h = nullptr;
        c.check(nlc_commit_from_values_materialised(c.raw(), cols.data(), W, log_n, (int)rate_bits, (int)cap_height, cout.data(),
                                                    fields ? fields->leaves.data() : nullptr,
                                                    fields && !fields->digests.empty() ? fields->digests[0].elements : nullptr,
                                                    b.merkle_tree.cap.hashes[0].elements, &h));
        b.h_.reset(h, nlc_batch_free);
        b.n_polys_ = W;
        b.degree_log = (size_t)log_n;
        b.rate_bits = rate_bits;
        b.cap_height_ = cap_height;
        return b;
    }
    // wrap a handle produced by one of the quotient-side entry points (stark:: / plonk:: below)
    static PolynomialBatch adopt(nlc_batch* h) {
        Context& c = Context::current();
        PolynomialBatch b;
        b.h_.reset(h, nlc_batch_free);
        size_t W = 0;
        int log_n = 0, r = 0, cap = 0;
        c.check(nlc_batch_info(h, &W, &log_n, &r, &cap));
        b.n_polys_ = W;
        b.degree_log = (size_t)log_n;
        b.rate_bits = (size_t)r;
        b.cap_height_ = (size_t)cap;
        b.merkle_tree.cap.hashes.resize((size_t)1 << cap);
        c.check(nlc_batch_cap(h, b.merkle_tree.cap.hashes[0].elements));
        return b;
    }
    size_t num_polys() const { return n_polys_; }
    // `polynomials` field (proof.rs:307-310 evaluates them at zeta): fetched from HBM on first use
    const std::vector<PolynomialCoeffs>& polynomials() {
        if (polys_.empty()) {
            polys_.assign(n_polys_, PolynomialCoeffs(std::vector<F>((size_t)1 << degree_log)));
            std::vector<uint64_t*> out;
            for (auto& p : polys_) out.push_back(p.coeffs.data());
            Context::current().check(nlc_batch_coeffs(h_.get(), out.data()));
        }
        return polys_;
    }
    // leaves[reverse_bits(index * step, lde_bits)]
    std::vector<F> get_lde_values(size_t index, size_t step) const {
        std::vector<F> row(n_polys_);
        const uint64_t idx = index;
        Context::current().check(nlc_batch_lde_rows(h_.get(), &idx, 1, step, row.data()));
        return row;
    }
    // P::WIDTH consecutive rows starting at index_start, one packed value (WIDTH lanes) per polynomial (:289-294)
    std::vector<std::vector<F>> get_lde_values_packed(size_t index_start, size_t step, size_t width) const {
        std::vector<uint64_t> idx(width);
        for (size_t i = 0; i < width; i++) idx[i] = index_start + i;
        std::vector<F> rows(width * n_polys_);
        Context::current().check(nlc_batch_lde_rows(h_.get(), idx.data(), width, step, rows.data()));
        std::vector<std::vector<F>> packed(n_polys_, std::vector<F>(width));
        for (size_t i = 0; i < width; i++)
            for (size_t j = 0; j < n_polys_; j++) packed[j][i] = rows[i * n_polys_ + j];
        return packed;
    }
    // every polynomial at the extension point z (StarkOpeningSet::new, proof.rs:300-328), on the device
    std::vector<FE> eval_all(const FE& z) const {
        const uint64_t zz[2] = {z.first, z.second};
        std::vector<F> out(2 * n_polys_);
        Context::current().check(nlc_batch_eval(h_.get(), zz, out.data()));
        std::vector<FE> r(n_polys_);
        for (size_t j = 0; j < n_polys_; j++) r[j] = {out[2 * j], out[2 * j + 1]};
        return r;
    }
    std::pair<std::vector<F>, MerkleProof> open(size_t leaf_index) const { // merkle_tree.get + merkle_tree.prove
        std::vector<F> row(n_polys_);
        MerkleProof p;
        p.siblings.resize(degree_log + rate_bits - cap_height_);
        Context::current().check(nlc_batch_open(h_.get(), leaf_index, row.data(),
                                                p.siblings.empty() ? nullptr : p.siblings[0].elements));
        return {row, p};
    }
    nlc_batch* raw() const { return h_.get(); }

    // PolynomialBatch::prove_openings(instance, oracles, challenger, fri_params, timing)  prover.rs:200-206
    static FriProof prove_openings(const FriInstanceInfo& instance, const std::vector<const PolynomialBatch*>& oracles,
                                   Challenger& challenger, const FriParams& fri_params) {
        if (oracles.empty()) throw Panic("prove_openings: no oracles");
        Context& c = Context::current();
        std::vector<uint64_t> pts;
        std::vector<uint32_t> offs{0}, po, pi;
        for (const FriBatchInfo& b : instance.batches) {
            pts.push_back(b.point.first), pts.push_back(b.point.second);
            for (const FriPolynomialInfo& i : b.polynomials) po.push_back((uint32_t)i.oracle_index), pi.push_back((uint32_t)i.polynomial_index);
            offs.push_back((uint32_t)po.size());
        }
        nlc_fri_instance ci{(int)instance.batches.size(), pts.data(), offs.data(), po.data(), pi.data()};
        std::vector<nlc_batch*> hs;
        for (const PolynomialBatch* o : oracles) hs.push_back(o->raw());
        FriProof proof;
        proof.flat.resize(nlc_fri_proof_words(&fri_params.c, hs.data(), (int)hs.size()));
        c.check(nlc_prove_openings(c.raw(), &ci, hs.data(), (int)hs.size(), challenger.raw(), &fri_params.c, proof.flat.data()));
        parse(proof, fri_params, oracles);
        return proof;
    }

  private:
    std::shared_ptr<nlc_batch> h_;
    size_t n_polys_ = 0, cap_height_ = 0;
    std::vector<PolynomialCoeffs> polys_;

    template <class V> static bool lens_equal(const std::vector<V>& v) {
        for (const auto& x : v)
            if (x.len() != v[0].len()) return false;
        return true;
    }
    static PolynomialBatch commit(const std::vector<const uint64_t*>& cols, size_t n, bool equal, size_t rate_bits,
                                  bool blinding, size_t cap_height, bool is_values) {
        if (blinding) throw Panic("blinding = true is not used on this path (prover.rs:91-93,171-178 pass false)");
        if (cols.empty()) throw Panic("PolynomialBatch: at least one polynomial required");
        if (!equal) throw Panic("All polynomials in a batch must have the same length");
        Context& c = Context::current();
        nlc_batch* h = nullptr;
        const int log_n = (int)log2_strict(n);
        c.check((is_values ? nlc_commit_from_values : nlc_commit_from_coeffs)(c.raw(), cols.data(), cols.size(), log_n,
                                                                             (int)rate_bits, (int)cap_height, &h));
        PolynomialBatch b;
        b.h_.reset(h, nlc_batch_free);
        b.n_polys_ = cols.size();
        b.degree_log = (size_t)log_n;
        b.rate_bits = rate_bits;
        b.cap_height_ = cap_height;
        b.merkle_tree.cap.hashes.resize((size_t)1 << cap_height);
        c.check(nlc_batch_cap(h, b.merkle_tree.cap.hashes[0].elements));
        return b;
    }
    static void take_hashes(const std::vector<F>& flat, size_t& w, size_t count, std::vector<HashOut>& out) {
        out.resize(count);
        for (size_t i = 0; i < count; i++, w += 4)
            for (int k = 0; k < 4; k++) out[i].elements[k] = flat[w + k];
    }
    // slice the flat proof (layout: include/nlc_commit.h at nlc_fri_proof_words) into plonky2's FriProof
    static void parse(FriProof& p, const FriParams& fp, const std::vector<const PolynomialBatch*>& oracles) {
        const std::vector<F>& f = p.flat;
        size_t w = 0;
        const size_t cap = (size_t)fp.c.cap_height, log_N = fp.lde_bits();
        for (int r = 0; r < fp.c.n_reduction_rounds; r++) {
            MerkleCap mc;
            take_hashes(f, w, (size_t)1 << cap, mc.hashes);
            p.commit_phase_merkle_caps.push_back(mc);
        }
        for (size_t i = 0; i < fp.final_poly_len(); i++, w += 2) p.final_poly.push_back({f[w], f[w + 1]});
        p.pow_witness = f[w++];
        for (int q = 0; q < fp.c.num_query_rounds; q++) {
            FriQueryRound qr;
            for (const PolynomialBatch* o : oracles) {
                std::vector<F> leaf(f.begin() + w, f.begin() + w + o->num_polys());
                w += o->num_polys();
                MerkleProof mp;
                take_hashes(f, w, log_N - cap, mp.siblings);
                qr.initial_trees_proof.evals_proofs.push_back({leaf, mp});
            }
            size_t lg = log_N;
            for (int r = 0; r < fp.c.n_reduction_rounds; r++) {
                const size_t ab = (size_t)fp.c.reduction_arity_bits[r];
                FriQueryStep st;
                for (size_t i = 0; i < ((size_t)1 << ab); i++, w += 2) st.evals.push_back({f[w], f[w + 1]});
                take_hashes(f, w, lg - ab - cap, st.merkle_proof.siblings);
                qr.steps.push_back(st);
                lg -= ab;
            }
            p.query_round_proofs.push_back(qr);
        }
        if (w != f.size()) throw Panic("FriProof: flat layout mismatch");
    }
};

// ---- starkyx's quotient with the AIR evaluated on the device --------------------------------------
// An AIR recorded once against a recording AirParser (vendor/starkyx/starkyx/src/air/parser.rs:6-75): the straight-line
// program of include/nlc_commit.h (nlc_air_program).  `emit` appends one instruction {op, dst, a, b}.
struct AirProgram {
    std::vector<uint32_t> code;
    std::vector<F> pool;
    uint32_t n_regs = 0, n_public = 0, n_global = 0, n_challenge = 0;
    void emit(uint32_t op, uint32_t dst, uint32_t a, uint32_t b = 0) {
        code.insert(code.end(), {op, dst, a, b});
        if ((op < NLC_AIR_CONSTRAINT || op == NLC_AIR_MULADD_CONST) && dst + 1 > n_regs) n_regs = dst + 1;
    }
    nlc_air_program c() const {
        nlc_air_program p{};
        p.code = code.data();
        p.n_ops = (uint32_t)(code.size() / 4);
        p.n_regs = n_regs;
        p.pool = pool.data();
        p.n_pool = (uint32_t)pool.size();
        p.n_public = n_public;
        p.n_global = n_global;
        p.n_challenge = n_challenge;
        return p;
    }
};
namespace stark {
inline std::vector<nlc_batch*> handles(const std::vector<const PolynomialBatch*>& bs) {
    std::vector<nlc_batch*> hs;
    for (const PolynomialBatch* b : bs) hs.push_back(b->raw());
    return hs;
}
// the constraint accumulators of prover.rs:306-335 for every point of the quotient domain, before 1/Z_H
inline std::vector<F> air_eval(const AirProgram& air, const std::vector<const PolynomialBatch*>& trace_data,
                               const std::vector<F>& public_inputs, const std::vector<F>& global_values,
                               const std::vector<F>& challenges, const std::vector<F>& alphas, size_t quotient_degree_bits) {
    Context& c = Context::current();
    const nlc_air_program p = air.c();
    auto hs = handles(trace_data);
    std::vector<F> out(alphas.size() << (trace_data.at(0)->degree_log + quotient_degree_bits));
    c.check(nlc_air_eval(c.raw(), &p, hs.data(), (int)hs.size(), public_inputs.data(), global_values.data(), challenges.data(),
                         alphas.data(), alphas.size(), (int)quotient_degree_bits, out.data()));
    return out;
}
// quotient_polys + trim / chunks / from_coeffs (prover.rs:144-178), all on the device
inline PolynomialBatch quotient_commitment(const AirProgram& air, const std::vector<const PolynomialBatch*>& trace_data,
                                           const std::vector<F>& public_inputs, const std::vector<F>& global_values,
                                           const std::vector<F>& challenges, const std::vector<F>& alphas,
                                           size_t quotient_degree_factor, size_t cap_height) {
    Context& c = Context::current();
    const nlc_air_program p = air.c();
    auto hs = handles(trace_data);
    nlc_batch* h = nullptr;
    c.check(nlc_stark_quotient_commit(c.raw(), &p, hs.data(), (int)hs.size(), public_inputs.data(), global_values.data(),
                                      challenges.data(), alphas.data(), alphas.size(), quotient_degree_factor, (int)cap_height, &h));
    return PolynomialBatch::adopt(h);
}
} // namespace stark

// ---- plonk::prover::prove between the wires commitment and the openings ---------------------------
namespace plonk {
struct CircuitShape { // the fields of CommonCircuitData the quotient path reads
    size_t num_constants = 0, num_routed_wires = 0, num_challenges = 0, num_partial_products = 0, quotient_degree_factor = 0;
    std::vector<F> k_is;
    nlc_plonk_circuit c() const {
        nlc_plonk_circuit p{};
        p.num_constants = num_constants;
        p.num_routed_wires = num_routed_wires;
        p.num_challenges = num_challenges;
        p.num_partial_products = num_partial_products;
        p.quotient_degree_factor = (int)quotient_degree_factor;
        p.k_is = k_is.data();
        return p;
    }
};
// all_wires_permutation_partial_products + the zs_partial_products commitment (plonk/prover.rs)
inline PolynomialBatch zs_partial_products_commitment(const CircuitShape& circuit, const PolynomialBatch& constants_sigmas,
                                                      const PolynomialBatch& wires, const std::vector<F>& betas,
                                                      const std::vector<F>& gammas, size_t cap_height) {
    Context& c = Context::current();
    const nlc_plonk_circuit p = circuit.c();
    nlc_batch* h = nullptr;
    c.check(nlc_plonk_zs_partial_products_commit(c.raw(), &p, constants_sigmas.raw(), wires.raw(), betas.data(), gammas.data(),
                                                 (int)cap_height, &h));
    return PolynomialBatch::adopt(h);
}
// compute_quotient_polys + the quotient commitment
inline PolynomialBatch quotient_commitment(const CircuitShape& circuit, const AirProgram& gate_constraints,
                                           const PolynomialBatch& constants_sigmas, const PolynomialBatch& wires,
                                           const PolynomialBatch& zs_partial_products, const HashOut& public_inputs_hash,
                                           const std::vector<F>& betas, const std::vector<F>& gammas,
                                           const std::vector<F>& alphas, size_t cap_height) {
    Context& c = Context::current();
    const nlc_plonk_circuit p = circuit.c();
    const nlc_air_program g = gate_constraints.c();
    nlc_batch* h = nullptr;
    c.check(nlc_plonk_quotient_commit(c.raw(), &p, &g, constants_sigmas.raw(), wires.raw(), zs_partial_products.raw(),
                                      public_inputs_hash.elements, betas.data(), gammas.data(), alphas.data(), (int)cap_height, &h));
    return PolynomialBatch::adopt(h);
}
} // namespace plonk

} // namespace plonky2
