// tests/cpp/test_plonky2_mirror.cpp -- drives include/nlc_plonky2.hpp the way starkyx's prover drives plonky2
// (/root/reference/vendor/starkyx/starkyx/src/plonky2/stark/prover.rs:50-225): trace commit -> observe cap ->
// alphas -> quotient values -> coset_ifft -> chunks -> from_coeffs -> observe cap -> zeta -> openings ->
// prove_openings, and checks every artefact bit for bit against the CPU oracle (oracle/oracle.h), whose restated
// plonky2 verifier must also accept the device's proof.  Built and run by tests/test_gpu_cpp_mirror.py.
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "nlc_plonky2.hpp"
#include "oracle.h"

using namespace plonky2;

#define CHECK(cond)                                                      \
    do {                                                                 \
        if (!(cond)) {                                                   \
            std::fprintf(stderr, "FAILED %s:%d: %s\n", __FILE__, __LINE__, #cond); \
            std::exit(1);                                                \
        }                                                                \
    } while (0)

static uint64_t rng_state = 0x9E3779B97F4A7C15ULL;
static F rand_field() { // splitmix64, rejected to [0, p)
    for (;;) {
        uint64_t z = (rng_state += 0x9E3779B97F4A7C15ULL);
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
        z ^= z >> 31;
        if (z < ORDER) return z;
    }
}
static F mulmod(F a, F b) { return (F)((unsigned __int128)a * b % ORDER); }
static F powmod(F b, uint64_t e) {
    F r = 1;
    for (; e; e >>= 1, b = mulmod(b, b))
        if (e & 1) r = mulmod(r, b);
    return r;
}

struct OracleBatch { // host arrays of the oracle's PolynomialBatch
    size_t W;
    int log_n, r, cap;
    std::vector<uint64_t> coeffs, leaves, digests, capv;
    orc_batch view() const { return {W, coeffs.data(), leaves.data(), digests.data(), capv.data()}; }
};
static OracleBatch oracle_commit(const std::vector<std::vector<F>>& cols, int r, int cap, bool is_values) {
    OracleBatch b;
    b.W = cols.size();
    const size_t n = cols[0].size();
    b.log_n = (int)log2_strict(n), b.r = r, b.cap = cap;
    const size_t N = n << r;
    std::vector<uint64_t> in;
    for (const auto& c : cols) in.insert(in.end(), c.begin(), c.end());
    b.coeffs.resize(b.W * n), b.leaves.resize(N * b.W), b.digests.resize(8 * (N - ((size_t)1 << cap))), b.capv.resize((size_t)4 << cap);
    if (is_values)
        orc_commit_from_values(in.data(), b.W, b.log_n, r, cap, b.coeffs.data(), b.leaves.data(), b.digests.data(), b.capv.data());
    else {
        b.coeffs = in;
        orc_commit_from_coeffs(in.data(), b.W, b.log_n, r, cap, b.leaves.data(), b.digests.data(), b.capv.data());
    }
    return b;
}

int main() {
    orc_poseidon_init();
    CHECK(orc_poseidon_selftest() == 0);
    const FriConfig fri_config = FriConfig::standard_fast_config(); // starkyx config.rs:58-73
    const size_t degree_bits = 10, degree = (size_t)1 << degree_bits, W = 11, quotient_degree_factor = 2, num_challenges = 2;
    const size_t rate_bits = fri_config.rate_bits, cap_height = fri_config.cap_height;

    // ---- trace commitment (prover.rs:86-100) ----
    std::vector<PolynomialValues> trace_cols;
    std::vector<std::vector<F>> raw_cols;
    for (size_t j = 0; j < W; j++) {
        std::vector<F> v(degree);
        for (F& x : v) x = rand_field();
        raw_cols.push_back(v);
        trace_cols.emplace_back(v);
    }
    PolynomialBatch trace_commitment = PolynomialBatch::from_values(trace_cols, rate_bits, false, cap_height);
    OracleBatch otrace = oracle_commit(raw_cols, (int)rate_bits, (int)cap_height, true);
    CHECK(trace_commitment.merkle_tree.cap.flatten() == otrace.capv);
    CHECK(trace_commitment.degree_log == degree_bits);
    for (size_t j = 0; j < W; j++)
        CHECK(std::memcmp(trace_commitment.polynomials()[j].coeffs.data(), otrace.coeffs.data() + j * degree, degree * 8) == 0);

    Challenger challenger;
    orc_challenger och;
    orc_challenger_init(&och);
    challenger.observe_cap(trace_commitment.merkle_tree.cap);
    orc_challenger_observe(&och, otrace.capv.data(), otrace.capv.size());
    std::vector<F> alphas = challenger.get_n_challenges(num_challenges);
    std::vector<F> oalphas(num_challenges);
    orc_challenger_get_n(&och, num_challenges, oalphas.data());
    CHECK(alphas == oalphas);

    // ---- get_lde_values_packed in the quotient loop (prover.rs:289-294) ----
    const size_t log_N = degree_bits + rate_bits;
    auto bitrev = [&](size_t x) { size_t r = 0; for (size_t b = 0; b < log_N; b++) r |= ((x >> b) & 1) << (log_N - 1 - b); return r; };
    auto packed = trace_commitment.get_lde_values_packed(8, 1, 4);
    for (size_t i = 0; i < 4; i++)
        for (size_t j = 0; j < W; j++) CHECK(packed[j][i] == otrace.leaves[bitrev(8 + i) * W + j]);
    CHECK(trace_commitment.get_lde_values(5, 2) == std::vector<F>(otrace.leaves.begin() + bitrev(10) * W, otrace.leaves.begin() + (bitrev(10) + 1) * W));

    // ---- quotient polys: values on the coset of size degree * 2^q -> coset_ifft -> chunks(degree) (prover.rs:147-178,352-358) ----
    std::vector<PolynomialCoeffs> all_quotient_chunks;
    std::vector<std::vector<F>> oquot_cols;
    for (size_t a = 0; a < num_challenges; a++) {
        std::vector<F> qv(degree * quotient_degree_factor);
        for (F& x : qv) x = mulmod(rand_field(), alphas[a]);
        PolynomialCoeffs quotient_poly = PolynomialValues(qv).coset_ifft(COSET_SHIFT);
        orc_coset_ifft(qv.data(), (int)log2_strict(qv.size()), COSET_SHIFT);
        CHECK(quotient_poly.coeffs == qv);
        for (PolynomialCoeffs& chunk : quotient_poly.chunks(degree)) {
            oquot_cols.push_back(chunk.coeffs);
            all_quotient_chunks.push_back(std::move(chunk));
        }
    }
    PolynomialBatch quotient_commitment = PolynomialBatch::from_coeffs(all_quotient_chunks, rate_bits, false, cap_height);
    OracleBatch oquot = oracle_commit(oquot_cols, (int)rate_bits, (int)cap_height, false);
    CHECK(quotient_commitment.merkle_tree.cap.flatten() == oquot.capv);
    {   // the same commit from the row-per-point quotient_values in one call (prover.rs:336-357):
        // rows = values of the quotient polynomials on the coset, then again as raw accumulators q(x) * Z_H(x)
        QuotientDomain dom(degree_bits, 1);
        size_t w = 0;
        std::vector<F> qcoef[2];
        for (size_t a = 0; a < num_challenges; a++)
            for (size_t c = 0; c < quotient_degree_factor; c++, w++)
                qcoef[a].insert(qcoef[a].end(), all_quotient_chunks[w].coeffs.begin(), all_quotient_chunks[w].coeffs.end());
        const size_t size = degree * quotient_degree_factor;
        std::vector<F> rows(size * num_challenges), raw(size * num_challenges);
        for (size_t a = 0; a < num_challenges; a++) {
            PolynomialValues v = PolynomialCoeffs(qcoef[a]).coset_fft(COSET_SHIFT);
            for (size_t i = 0; i < size; i++) {
                rows[i * num_challenges + a] = v.values[i];
                const F xn = powmod(dom.coset[i], degree), z_h = xn ? xn - 1 : ORDER - 1;
                CHECK(mulmod(z_h, dom.eval_inverse(i)) == 1);
                raw[i * num_challenges + a] = mulmod(v.values[i], z_h);
            }
        }
        PolynomialBatch q1 = PolynomialBatch::from_quotient_values(rows, num_challenges, degree_bits, quotient_degree_factor, rate_bits, cap_height);
        PolynomialBatch q2 = PolynomialBatch::from_quotient_values(raw, num_challenges, degree_bits, quotient_degree_factor, rate_bits, cap_height, true);
        CHECK(q1.merkle_tree.cap.flatten() == oquot.capv);
        CHECK(q2.merkle_tree.cap.flatten() == oquot.capv);
        std::vector<uint64_t> ox(size), of(size), ol(size), oz(2);
        orc_quotient_domain((int)degree_bits, 1, ox.data(), of.data(), ol.data(), oz.data());
        CHECK(dom.coset == ox && dom.lagrange_first.values == of && dom.lagrange_last.values == ol && dom.z_h_inverses == oz);
    }
    challenger.observe_cap(quotient_commitment.merkle_tree.cap);
    orc_challenger_observe(&och, oquot.capv.data(), oquot.capv.size());

    // ---- an AIR evaluated on the device (prover.rs:306-352), driven from C++ ----
    {   // trace columns (a, b, c = a * b, d): constraints  a * b - c  (all rows)  and  (d_next - d - 1) (transition);
        // quotient_degree_factor 1 (degree-2 constraints), 2 alphas: acc = (a b - c) alpha + (d' - d - 1) z_last
        const size_t n = degree;
        std::vector<std::vector<F>> cols(4, std::vector<F>(n));
        for (size_t i = 0; i < n; i++) {
            cols[0][i] = rand_field();
            cols[1][i] = rand_field();
            cols[2][i] = mulmod(cols[0][i], cols[1][i]);
            cols[3][i] = (F)(1000 + i);
        }
        std::vector<PolynomialValues> tv(cols.begin(), cols.end());
        PolynomialBatch::HostFields fields;
        PolynomialBatch t = PolynomialBatch::from_values_materialised(tv, rate_bits, false, cap_height, &fields);
        OracleBatch ot = oracle_commit(cols, (int)rate_bits, (int)cap_height, true);
        CHECK(t.merkle_tree.cap.flatten() == ot.capv);       // fields written back while the commit ran
        CHECK(fields.leaves == ot.leaves);
        CHECK(std::memcmp(fields.digests[0].elements, ot.digests.data(), ot.digests.size() * 8) == 0);
        for (size_t j = 0; j < 4; j++) CHECK(std::memcmp(t.polynomials()[j].coeffs.data(), ot.coeffs.data() + j * n, n * 8) == 0);
        AirProgram air;  // what a recording AirParser produces for this AIR
        air.pool = {1};
        air.emit(NLC_AIR_LOCAL, 0, 0);
        air.emit(NLC_AIR_LOCAL, 1, 1);
        air.emit(NLC_AIR_MUL, 0, 0, 1);
        air.emit(NLC_AIR_LOCAL, 1, 2);
        air.emit(NLC_AIR_SUB, 0, 0, 1);
        air.emit(NLC_AIR_CONSTRAINT, 0, 0);
        air.emit(NLC_AIR_NEXT, 0, 3);
        air.emit(NLC_AIR_LOCAL, 1, 3);
        air.emit(NLC_AIR_SUB, 0, 0, 1);
        air.emit(NLC_AIR_SUB_CONST, 0, 0, 0);
        air.emit(NLC_AIR_CONSTRAINT_TRANSITION, 0, 0);
        const std::vector<F> al = {rand_field(), rand_field()};
        std::vector<F> acc = stark::air_eval(air, {&t}, {}, {}, {}, al, 0);
        // the same accumulators from the oracle's LDE rows: point i of the quotient domain = LDE point i * 2^rate_bits
        const size_t lN = degree_bits + rate_bits;
        auto brev = [&](size_t x) { size_t r = 0; for (size_t b = 0; b < lN; b++) r |= ((x >> b) & 1) << (lN - 1 - b); return r; };
        const F w = powmod(1753635133440165772ULL, (uint64_t)1 << (32 - degree_bits));
        const F last = powmod(w, ORDER - 2);
        F x = COSET_SHIFT;
        auto sub = [](F a, F b) { return a >= b ? a - b : a + (ORDER - b); };
        std::vector<F> want(acc.size());
        for (size_t i = 0; i < n; i++, x = mulmod(x, w)) {
            const F* row = &ot.leaves[brev(i << rate_bits) * 4];
            const F* nxt = &ot.leaves[brev(((i + 1) % n) << rate_bits) * 4];
            const F c0 = sub(mulmod(row[0], row[1]), row[2]);
            const F c1 = mulmod(sub(sub(nxt[3], row[3]), 1), sub(x, last));
            for (size_t a = 0; a < 2; a++) want[i * 2 + a] = (F)(((unsigned __int128)mulmod(c0, al[a]) + c1) % ORDER);
        }
        CHECK(acc == want);
        // ... and the quotient commitment in one call: evaluate, 1/Z_H, coset_ifft, chunks, from_coeffs
        PolynomialBatch q = stark::quotient_commitment(air, {&t}, {}, {}, {}, al, 1, cap_height);
        OracleBatch oq;
        oq.W = 2;
        oq.log_n = (int)degree_bits;
        oq.coeffs.resize(2 * n);
        oq.leaves.resize((n << rate_bits) * 2);
        oq.digests.resize(8 * ((n << rate_bits) - ((size_t)1 << cap_height)));
        oq.capv.resize((size_t)4 << cap_height);
        CHECK(orc_quotient_commit(want.data(), 2, (int)degree_bits, 0, 1, 1, (int)rate_bits, (int)cap_height, oq.coeffs.data(),
                                  oq.leaves.data(), oq.digests.data(), oq.capv.data()) == 0);
        CHECK(q.num_polys() == 2 && q.merkle_tree.cap.flatten() == oq.capv);
    }

    // ---- zeta, openings (prover.rs:183-193; proof.rs:300-328) ----
    FE zeta = challenger.get_extension_challenge();
    uint64_t ozeta[2];
    orc_challenger_get_ext(&och, ozeta);
    CHECK(zeta.first == ozeta[0] && zeta.second == ozeta[1]);
    const F g = powmod(1753635133440165772ULL, (uint64_t)1 << (32 - degree_bits)); // F::primitive_root_of_unity(degree_bits)
    FE zeta_next = {mulmod(zeta.first, g), mulmod(zeta.second, g)};
    std::vector<FE> local_values = trace_commitment.eval_all(zeta), next_values = trace_commitment.eval_all(zeta_next),
                    quotient_polys = quotient_commitment.eval_all(zeta);
    std::vector<uint64_t> openings; // to_fri_openings order: [local ++ quotient] at zeta, [next] at g * zeta
    auto push = [&](const std::vector<FE>& v) { for (const FE& e : v) openings.push_back(e.first), openings.push_back(e.second); };
    push(local_values), push(quotient_polys), push(next_values);
    for (size_t j = 0; j < W; j++) {
        uint64_t z2[2] = {zeta.first, zeta.second}, o2[2];
        orc_eval_poly_ext(otrace.coeffs.data() + j * degree, degree, z2, o2);
        CHECK(local_values[j].first == o2[0] && local_values[j].second == o2[1]);
    }
    challenger.observe_extension_elements(local_values);
    challenger.observe_extension_elements(quotient_polys);
    challenger.observe_extension_elements(next_values);
    orc_challenger_observe(&och, openings.data(), openings.size());
    orc_challenger verifier_ch = och; // the verifier's transcript right after observe_openings

    // ---- prove_openings (prover.rs:195-206; instance built as in plonky2/stark/mod.rs:56-96) ----
    FriInstanceInfo instance;
    instance.oracles = {{W, false}, {all_quotient_chunks.size(), false}};
    FriBatchInfo zeta_batch{zeta, FriPolynomialInfo::from_range(0, 0, W)};
    auto qinfo = FriPolynomialInfo::from_range(1, 0, all_quotient_chunks.size());
    zeta_batch.polynomials.insert(zeta_batch.polynomials.end(), qinfo.begin(), qinfo.end());
    instance.batches = {zeta_batch, FriBatchInfo{zeta_next, FriPolynomialInfo::from_range(0, 0, W)}};
    FriParams fri_params = fri_config.fri_params(degree_bits, false);
    FriProof opening_proof = PolynomialBatch::prove_openings(instance, {&trace_commitment, &quotient_commitment}, challenger, fri_params);

    std::vector<uint64_t> pts = {zeta.first, zeta.second, zeta_next.first, zeta_next.second};
    std::vector<uint32_t> offs{0}, po, pi;
    for (const FriBatchInfo& b : instance.batches) {
        for (const auto& i : b.polynomials) po.push_back((uint32_t)i.oracle_index), pi.push_back((uint32_t)i.polynomial_index);
        offs.push_back((uint32_t)po.size());
    }
    orc_fri_instance oinst{2, pts.data(), offs.data(), po.data(), pi.data()};
    orc_fri_params op;
    orc_fri_params_constant_arity(&op, (int)degree_bits, (int)rate_bits, (int)cap_height, (int)fri_config.proof_of_work_bits,
                                  (int)fri_config.num_query_rounds, 4, 5);
    CHECK(op.n_rounds == fri_params.c.n_reduction_rounds);
    orc_batch obs[2] = {otrace.view(), oquot.view()};
    std::vector<uint64_t> oproof(orc_fri_proof_words(&op, obs, 2));
    CHECK(oproof.size() == opening_proof.flat.size());
    CHECK(orc_prove_openings(&oinst, obs, 2, &och, &op, oproof.data()) == 0);
    CHECK(oproof == opening_proof.flat); // bit-exact FriProof
    CHECK(opening_proof.query_round_proofs.size() == fri_config.num_query_rounds);
    CHECK(opening_proof.final_poly.size() == fri_params.final_poly_len());
    CHECK(std::memcmp(challenger.raw()->sponge_state, och.state, sizeof och.state) == 0); // transcripts stay in step

    // the restated plonky2 verifier accepts the device's proof
    const uint64_t* caps[2] = {otrace.capv.data(), oquot.capv.data()};
    const size_t Ws[2] = {W, all_quotient_chunks.size()};
    CHECK(orc_fri_verify(&oinst, openings.data(), caps, Ws, 2, &verifier_ch, &op, opening_proof.flat.data()) == 0);

    // ---- Merkle openings of the committed batch; MerkleTree::new; hashing ----
    auto opened = trace_commitment.open(77);
    CHECK(orc_merkle_verify(opened.first.data(), W, 77, otrace.capv.data(), (int)cap_height,
                            opened.second.siblings[0].elements, (int)opened.second.siblings.size()) == 0);
    std::vector<std::vector<F>> leaves(64, std::vector<F>(9));
    std::vector<uint64_t> flat_leaves;
    for (auto& l : leaves)
        for (F& x : l) x = rand_field(), flat_leaves.push_back(x);
    MerkleTree tree(leaves, 2);
    std::vector<uint64_t> od(8 * (64 - 4)), oc(16);
    orc_merkle_new(flat_leaves.data(), 64, 9, 2, od.data(), oc.data());
    CHECK(tree.cap.flatten() == oc);
    CHECK(tree.get(13) == leaves[13]);
    MerkleProof mp = tree.prove(13);
    CHECK(orc_merkle_verify(leaves[13].data(), 9, 13, oc.data(), 2, mp.siblings[0].elements, (int)mp.siblings.size()) == 0);
    uint64_t oh[4];
    orc_hash_no_pad(leaves[0].data(), 9, oh);
    CHECK(std::memcmp(PoseidonHash::hash_no_pad(leaves[0]).elements, oh, 32) == 0);
    HashOut l = PoseidonHash::hash_or_noop({1, 2, 3}), rr = PoseidonHash::hash_no_pad({4, 5, 6, 7, 8});
    orc_two_to_one(l.elements, rr.elements, oh);
    CHECK(std::memcmp(PoseidonHash::two_to_one(l, rr).elements, oh, 32) == 0);

    // ---- lde_onto_coset / selector (prover.rs:280-284) ----
    PolynomialValues lagrange_first = PolynomialValues::selector(degree, 0).lde_onto_coset(1);
    std::vector<uint64_t> sel(degree, 0), olde(degree * 2);
    sel[0] = 1;
    orc_lde_onto_coset(sel.data(), (int)degree_bits, 1, olde.data());
    CHECK(lagrange_first.values == olde);

    // ---- plonky2's contract at this boundary: inconsistent input panics ----
    bool panicked = false;
    try {
        std::vector<PolynomialValues> ragged = {PolynomialValues(std::vector<F>(8, 1)), PolynomialValues(std::vector<F>(16, 1))};
        PolynomialBatch::from_values(ragged, 1, false, 0);
    } catch (const Panic&) { panicked = true; }
    CHECK(panicked);
    panicked = false;
    try { PolynomialBatch::from_values({PolynomialValues(std::vector<F>(8, 1))}, 1, false, 9); } // cap_height > log2(leaves)
    catch (const Panic&) { panicked = true; }
    CHECK(panicked);

    std::puts("plonky2 C++ mirror: all checks passed");
    return 0;
}
