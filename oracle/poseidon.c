/*
 * oracle/poseidon.c -- width-12 Poseidon over Goldilocks and plonky2's sponge modes, CPU restatement.
 *
 * TEST INFRASTRUCTURE ONLY (see oracle/gl.h).
 *
 * Restates plonky2 0.1.4 @ d2598bd (Cargo.lock:4864-4866; source absent from /root/reference):
 *   plonky2/src/hash/poseidon.rs            Poseidon::poseidon (4 full + 22 partial + 4 full, x^7,
 *                                           circulant+diagonal MDS) -- the *naive* round form, which
 *                                           upstream's optimised partial rounds equal algebraically
 *   plonky2/src/hash/poseidon_goldilocks.rs MDS_MATRIX_CIRC / MDS_MATRIX_DIAG / ALL_ROUND_CONSTANTS
 *   plonky2/src/hash/hashing.rs             hash_n_to_m_no_pad (overwrite-mode sponge), compress
 *   plonky2/src/plonk/config.rs             Hasher::hash_or_noop
 * In-tree binding of this hasher: vendor/starkyx/starkyx/src/plonky2/stark/config.rs:24,104
 * (CurtaPoseidonGoldilocksConfig::Hasher = PoseidonHash).
 *
 * Round constants are regenerated here at start-up exactly as upstream generated them
 * (ChaCha8Rng::seed_from_u64(0), 360 x gen_range(0..p)) and then
 * checked against upstream's own known-answer vectors (poseidon_goldilocks.rs test_vectors) by
 * orc_poseidon_selftest().  The product ships an independently generated table
 * (tools/gen_poseidon_constants.py); tests compare the two.
 */
#include "oracle.h"
#include "poseidon_fast.h"
#include <string.h>

#define W 12
#define N_ROUNDS 30
#define HALF_FULL 4
#define N_PARTIAL 22

static const uint64_t MDS_CIRC[W] = {17, 15, 41, 16, 2, 28, 13, 13, 39, 18, 34, 20};
static const uint64_t MDS_DIAG[W] = {8, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};

static gl_t RC[N_ROUNDS * W];
static int rc_ready = 0;

/* ---- ChaCha8Rng::seed_from_u64(0) stream (rand_core 0.6 seed_from_u64 + rand_chacha 0.3) ---- */
static inline uint32_t rotl32(uint32_t x, int k) { return (x << k) | (x >> (32 - k)); }
static inline uint32_t rotr32(uint32_t x, int k) { k &= 31; return k ? (x >> k) | (x << (32 - k)) : x; }
#define QR(a, b, c, d)                 \
    a += b; d ^= a; d = rotl32(d, 16); \
    c += d; b ^= c; b = rotl32(b, 12); \
    a += b; d ^= a; d = rotl32(d, 8);  \
    c += d; b ^= c; b = rotl32(b, 7);

typedef struct { uint32_t key[8]; uint64_t counter; uint32_t buf[16]; int pos; } chacha8_t;

static void chacha8_block(chacha8_t* g) {
    uint32_t s[16] = {0x61707865, 0x3320646e, 0x79622d32, 0x6b206574};
    for (int i = 0; i < 8; i++) s[4 + i] = g->key[i];
    s[12] = (uint32_t)g->counter; s[13] = (uint32_t)(g->counter >> 32); s[14] = 0; s[15] = 0;
    uint32_t x[16];
    memcpy(x, s, sizeof x);
    for (int r = 0; r < 4; r++) { /* 4 double rounds = 8 rounds */
        QR(x[0], x[4], x[8], x[12]) QR(x[1], x[5], x[9], x[13]) QR(x[2], x[6], x[10], x[14]) QR(x[3], x[7], x[11], x[15])
        QR(x[0], x[5], x[10], x[15]) QR(x[1], x[6], x[11], x[12]) QR(x[2], x[7], x[8], x[13]) QR(x[3], x[4], x[9], x[14])
    }
    for (int i = 0; i < 16; i++) g->buf[i] = x[i] + s[i];
    g->counter++;
    g->pos = 0;
}
static void chacha8_seed_from_u64(chacha8_t* g, uint64_t state) {
    for (int i = 0; i < 8; i++) { /* PCG32 expands the u64 into the 32-byte key */
        state = state * 6364136223846793005ULL + 11634580027462260723ULL;
        uint32_t xorshifted = (uint32_t)(((state >> 18) ^ state) >> 27);
        g->key[i] = rotr32(xorshifted, (int)(state >> 59));
    }
    g->counter = 0;
    g->pos = 16;
}
static uint32_t chacha8_next_u32(chacha8_t* g) {
    if (g->pos >= 16) chacha8_block(g);
    return g->buf[g->pos++];
}
static uint64_t chacha8_next_u64(chacha8_t* g) {
    uint64_t lo = chacha8_next_u32(g), hi = chacha8_next_u32(g);
    return lo | (hi << 32);
}
/* rand 0.8 UniformInt<u64>::sample_single_inclusive(0, p-1): widening-multiply rejection sampler */
static uint64_t gen_range_p(chacha8_t* g) {
    const uint64_t zone = GL_P - 1; /* (p << lzcnt(p)) - 1 with lzcnt(p) = 0 */
    for (;;) {
        uint64_t v = chacha8_next_u64(g);
        __uint128_t m = (__uint128_t)v * GL_P;
        if ((uint64_t)m <= zone) return (uint64_t)(m >> 64);
    }
}

void orc_poseidon_init(void) {
    if (rc_ready) return;
    chacha8_t g;
    chacha8_seed_from_u64(&g, 0);
    for (int i = 0; i < N_ROUNDS * W; i++) RC[i] = gen_range_p(&g);
    rc_ready = 1;
}
const uint64_t* orc_poseidon_round_constants(void) { orc_poseidon_init(); return RC; }

/* ---- permutation (poseidon.rs: constant_layer, sbox_layer / sbox_monomial, mds_layer) ---- */
static inline gl_t sbox7(gl_t x) {
    gl_t x2 = gl_sqr(x), x4 = gl_sqr(x2), x3 = gl_mul(x, x2);
    return gl_mul(x3, x4);
}
static inline void mds_layer(gl_t* s) {
    /* the circulant entries are tiny, so (like upstream's mds_row_shf on 32-bit halves) the two 12-term dot
     * products fit in plain u64: 12 * 2^32 * 41 < 2^42.  Doubled arrays avoid the index modulo; the loops
     * vectorise. */
    uint64_t lo[2 * W], hi[2 * W];
    for (int i = 0; i < W; i++) {
        lo[i] = lo[i + W] = s[i] & 0xFFFFFFFFULL;
        hi[i] = hi[i + W] = s[i] >> 32;
    }
    for (int r = 0; r < W; r++) {
        uint64_t al = 0, ah = 0;
        for (int i = 0; i < W; i++) {
            al += lo[i + r] * MDS_CIRC[i];
            ah += hi[i + r] * MDS_CIRC[i];
        }
        al += lo[r] * MDS_DIAG[r];
        ah += hi[r] * MDS_DIAG[r];
        s[r] = gl_reduce128((__uint128_t)al + ((__uint128_t)ah << 32));
    }
}
/* naive form: every round = constant layer, S-box layer, dense MDS (the definition; used as a cross-check) */
void orc_poseidon_permute_naive(uint64_t* s) {
    if (!rc_ready) orc_poseidon_init();
    for (int i = 0; i < W; i++) s[i] = gl_canon(s[i]);
    int round = 0;
    for (int r = 0; r < HALF_FULL; r++, round++) {
        for (int i = 0; i < W; i++) s[i] = sbox7(gl_add(s[i], RC[round * W + i]));
        mds_layer(s);
    }
    for (int r = 0; r < N_PARTIAL; r++, round++) {
        for (int i = 0; i < W; i++) s[i] = gl_add(s[i], RC[round * W + i]);
        s[0] = sbox7(s[0]);
        mds_layer(s);
    }
    for (int r = 0; r < HALF_FULL; r++, round++) {
        for (int i = 0; i < W; i++) s[i] = sbox7(gl_add(s[i], RC[round * W + i]));
        mds_layer(s);
    }
}

/* sum of up to 12 products of canonical elements: < 12 * 2^128, kept as (lo128, carry count) */
static inline gl_t dot_reduce(const gl_t* a, const uint64_t* b, int n, __uint128_t init) {
    __uint128_t acc = init;
    uint64_t over = 0;
    for (int i = 0; i < n; i++) {
        __uint128_t t = (__uint128_t)a[i] * b[i];
        acc += t;
        over += acc < t;
    }
    /* 2^128 == -2^32 (mod p) since 2^96 == -1 */
    gl_t r = gl_reduce128(acc);
    return gl_sub(r, gl_reduce128((__uint128_t)over << 32));
}

/* poseidon.rs Poseidon::poseidon with the fast partial rounds (partial_first_constant_layer,
 * mds_partial_layer_init, 22 x {sbox_monomial(state[0]) + constant, mds_partial_layer_fast}) -- what the
 * reference CPU prover actually runs; constants in poseidon_fast.h (tools/gen_poseidon_fast.py). */
void orc_poseidon_permute(uint64_t* s) {
    if (!rc_ready) orc_poseidon_init();
    for (int i = 0; i < W; i++) s[i] = gl_canon(s[i]);
    int round = 0;
    for (int r = 0; r < HALF_FULL; r++, round++) {
        for (int i = 0; i < W; i++) s[i] = sbox7(gl_add(s[i], RC[round * W + i]));
        mds_layer(s);
    }
    for (int i = 0; i < W; i++) s[i] = gl_add(s[i], PF_FIRST[i]);
    {
        gl_t t[W];
        t[0] = s[0];
        for (int r = 1; r < W; r++) t[r] = dot_reduce(s + 1, PF_INIT[r - 1], W - 1, 0);
        memcpy(s, t, sizeof t);
    }
    for (int r = 0; r < N_PARTIAL; r++) {
        gl_t s0 = gl_add(sbox7(s[0]), PF_POST[r]);
        gl_t d = dot_reduce(s + 1, PF_W_HAT[r], W - 1, (__uint128_t)s0 * (MDS_CIRC[0] + MDS_DIAG[0]));
        for (int i = 1; i < W; i++) s[i] = gl_add(s[i], gl_mul(s0, PF_VS[r][i - 1]));
        s[0] = d;
    }
    round += N_PARTIAL;
    for (int r = 0; r < HALF_FULL; r++, round++) {
        for (int i = 0; i < W; i++) s[i] = sbox7(gl_add(s[i], RC[round * W + i]));
        mds_layer(s);
    }
}

/* ---- sponge modes (hashing.rs) ---- */
#define RATE 8
void orc_hash_no_pad(const uint64_t* in, size_t len, uint64_t* out4) {
    gl_t st[W] = {0};
    for (size_t off = 0; off < len; off += RATE) {
        size_t c = len - off < RATE ? len - off : RATE;
        for (size_t i = 0; i < c; i++) st[i] = gl_canon(in[off + i]); /* overwrite mode; other lanes keep */
        orc_poseidon_permute(st);
    }
    memcpy(out4, st, 4 * sizeof(gl_t)); /* len == 0: no permutation, output zeros (perm.squeeze()) */
}
void orc_two_to_one(const uint64_t* l, const uint64_t* r, uint64_t* out4) {
    gl_t st[W] = {0};
    for (int i = 0; i < 4; i++) { st[i] = gl_canon(l[i]); st[4 + i] = gl_canon(r[i]); }
    orc_poseidon_permute(st);
    memcpy(out4, st, 4 * sizeof(gl_t));
}
void orc_hash_or_noop(const uint64_t* in, size_t len, uint64_t* out4) {
    if (len <= 4) {
        for (size_t i = 0; i < 4; i++) out4[i] = i < len ? gl_canon(in[i]) : 0;
    } else {
        orc_hash_no_pad(in, len, out4);
    }
}

/* Upstream KATs: plonky2/src/hash/poseidon_goldilocks.rs `test_vectors` (first three entries), plus
 * spot checks of ALL_ROUND_CONSTANTS (first 16, last 4).  Returns 0 when everything matches. */
int orc_poseidon_selftest(void) {
    orc_poseidon_init();
    static const uint64_t rc_head[16] = {
        0xb585f766f2144405ULL, 0x7746a55f43921ad7ULL, 0xb2fb0d31cee799b4ULL, 0x0f6760a4803427d7ULL,
        0xe10d666650f4e012ULL, 0x8cae14cb07d09bf1ULL, 0xd438539c95f63e9fULL, 0xef781c7ce35b4c3dULL,
        0xcdc4a239b0c44426ULL, 0x277fa208bf337bffULL, 0xe17653a29da578a1ULL, 0xc54302f225db2c76ULL,
        0x86287821f722c881ULL, 0x59cd1a8a41c18e55ULL, 0xc3b919ad495dc574ULL, 0xa484c4c5ef6a0781ULL};
    static const uint64_t rc_tail[4] = {0x4543d9df5476d3cbULL, 0xf172d73e004fc90dULL, 0xdfd1c4febcc81238ULL,
                                        0xbc8dfb627fe558fcULL};
    for (int i = 0; i < 16; i++) if (RC[i] != rc_head[i]) return 1;
    for (int i = 0; i < 4; i++) if (RC[356 + i] != rc_tail[i]) return 2;
    static const uint64_t kat_zero[12] = {
        0x3c18a9786cb0b359ULL, 0xc4055e3364a246c3ULL, 0x7953db0ab48808f4ULL, 0xc71603f33a1144caULL,
        0xd7709673896996dcULL, 0x46a84e87642f44edULL, 0xd032648251ee0b3cULL, 0x1c687363b207df62ULL,
        0xdf8565563e8045feULL, 0x40f5b37ff4254daeULL, 0xd070f637b431067cULL, 0x1792b1c4342109d7ULL};
    static const uint64_t kat_iota[12] = {
        0xd64e1e3efc5b8e9eULL, 0x53666633020aaa47ULL, 0xd40285597c6a8825ULL, 0x613a4f81e81231d2ULL,
        0x414754bfebd051f0ULL, 0xcb1f8980294a023fULL, 0x6eb2a9e4d54a9d0fULL, 0x1902bc3af467e056ULL,
        0xf045d5eafdc6021fULL, 0xe4150f77caaa3be5ULL, 0xc9bfd01d39b50cceULL, 0x5c0a27fcb0e1459bULL};
    static const uint64_t kat_neg1[12] = {
        0xbe0085cfc57a8357ULL, 0xd95af71847d05c09ULL, 0xcf55a13d33c1c953ULL, 0x95803a74f4530e82ULL,
        0xfcd99eb30a135df1ULL, 0xe095905e913a3029ULL, 0xde0392461b42919bULL, 0x7d3260e24e81d031ULL,
        0x10d3d0465d9deaa0ULL, 0xa87571083dfc2a47ULL, 0xe18263681e9958f8ULL, 0xe28e96f1ae5e60d3ULL};
    uint64_t s[12];
    memset(s, 0, sizeof s);
    orc_poseidon_permute(s);
    if (memcmp(s, kat_zero, sizeof s)) return 3;
    for (int i = 0; i < 12; i++) s[i] = (uint64_t)i;
    orc_poseidon_permute(s);
    if (memcmp(s, kat_iota, sizeof s)) return 4;
    for (int i = 0; i < 12; i++) s[i] = GL_P - 1;
    orc_poseidon_permute(s);
    if (memcmp(s, kat_neg1, sizeof s)) return 5;
    /* the definition (naive rounds) and the fast partial rounds must agree, on the KATs and on a pseudo-random walk */
    uint64_t a[12], b[12];
    for (int i = 0; i < 12; i++) a[i] = b[i] = GL_P - 1 - (uint64_t)i * 0x9E3779B97F4A7C15ULL % GL_P;
    for (int it = 0; it < 64; it++) {
        orc_poseidon_permute(a);
        orc_poseidon_permute_naive(b);
        if (memcmp(a, b, sizeof a)) return 6;
    }
    memset(s, 0, sizeof s);
    orc_poseidon_permute_naive(s);
    if (memcmp(s, kat_zero, sizeof s)) return 7;
    return 0;
}
