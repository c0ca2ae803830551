#!/usr/bin/env python3
"""tools/proof_phases.py -- where the wall time of a whole proof goes (host timers around every stage, device synchronised
after each): plonk.PlonkProver.prove on the demo circuit and starky.StarkyProver.prove on the FpMul chip."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as graft  # noqa: E402

P = 0xFFFFFFFF00000001


def main():
    import torch
    pkg = graft.load_package()
    P2, K, S = pkg.plonky2, pkg.plonk, pkg.starky
    ctx = P2.default_context()
    degree_bits = int(sys.argv[1]) if len(sys.argv) > 1 else 16
    cfg = K.CircuitConfig.standard_recursion_config()
    fc = cfg.fri_config
    r, cap = fc.rate_bits, fc.cap_height
    common, consts, sig, wires, pih = K.build_demo_circuit(cfg, degree_bits, 7)
    pd = K.ProverOnlyCircuitData(P2.PolynomialBatch.from_values(list(np.concatenate([consts, sig])), r, False, cap), sig)
    cols = list(wires)
    common.compile_gate_constraints()
    wc = zs = q = None
    for rep in range(4):
        t = {}
        c0 = ctx.cache_stats()
        def lap(name, t0):
            ctx.synchronize()
            t[name] = t.get(name, 0.0) + (time.perf_counter() - t0) * 1e3
        wc = zs = q = None  # a prover drops the previous proof's batches before it starts the next
        ctx.synchronize()
        T0 = time.perf_counter()
        t0 = time.perf_counter(); ch = P2.Challenger(); wc = P2.PolynomialBatch.from_values(cols, r, False, cap); lap("wires commit (135 pageable columns in)", t0)
        t0 = time.perf_counter(); ch.observe_hash(pd.circuit_digest); ch.observe_hash(pih); ch.observe_cap(wc.merkle_tree.cap)
        betas = ch.get_n_challenges(2); gammas = ch.get_n_challenges(2); lap("transcript", t0)
        t0 = time.perf_counter(); zs = K.PlonkProver.zs_partial_products_commitment(common, pd, wc, betas, gammas); lap("partial products + Z + commit", t0)
        t0 = time.perf_counter(); ch.observe_cap(zs.merkle_tree.cap); alphas = ch.get_n_challenges(2); lap("transcript", t0)
        t0 = time.perf_counter(); q = K.PlonkProver.quotient_commitment(common, pd, pih, wc, zs, betas, gammas, alphas); lap("vanishing polynomial + quotient commit", t0)
        t0 = time.perf_counter(); ch.observe_cap(q.merkle_tree.cap); zeta = ch.get_extension_challenge(); lap("transcript", t0)
        g = pow(1753635133440165772, 1 << (32 - degree_bits), P)
        t0 = time.perf_counter(); op = K.OpeningSet.new(zeta, g, pd.constants_sigmas_commitment, wc, zs, q, common); lap("OpeningSet::new (5 batch evaluations)", t0)
        t0 = time.perf_counter()
        for b in op.to_fri_openings():
            ch.observe_extension_elements(b)
        lap("transcript", t0)
        t0 = time.perf_counter()
        inst = common.get_fri_instance(zeta, g)
        lap("get_fri_instance (host)", t0)
        t0 = time.perf_counter()
        P2.PolynomialBatch.prove_openings(inst, [pd.constants_sigmas_commitment, wc, zs, q], ch, fc.fri_params(degree_bits))
        lap("prove_openings", t0)
        total = (time.perf_counter() - T0) * 1e3
        if rep >= 1:
            c1 = ctx.cache_stats()
            print(f"plonk proof, degree 2^{degree_bits}: {total:.2f} ms   buffer cache: {c1['hits'] - c0['hits']} hits, "
                  f"{c1['misses'] - c0['misses']} driver allocations ({(c1['miss_us'] - c0['miss_us']) / 1e3:.2f} ms), {c1['cached_bytes'] >> 20} MiB cached")
            for k, v in t.items():
                print(f"  {k:50s} {v:7.2f} ms")
    n = 1 << degree_bits
    chip = S.FpMulChip()
    base = chip.generate_trace(np.random.default_rng(1), 1 << 10)
    trace = np.ascontiguousarray(np.tile(base, (n // base.shape[0], 1)))
    stark = S.Starky(chip)
    scfg = S.StarkyConfig.standard_fast_config(n)
    sfc = scfg.fri_config
    stark.program(0)
    tc = q = None
    for rep in range(4):
        t = {}
        def lap(name, t0):
            ctx.synchronize()
            t[name] = t.get(name, 0.0) + (time.perf_counter() - t0) * 1e3
        tc = q = None
        ctx.synchronize()
        T0 = time.perf_counter()
        t0 = time.perf_counter(); ch = P2.Challenger(); ch.observe_elements([])
        tc = P2.PolynomialBatch.from_trace_rows(trace, 0, chip.width(), sfc.rate_bits, sfc.cap_height); lap("trace commit (row-major AirTrace in)", t0)
        t0 = time.perf_counter(); ch.observe_cap(tc.merkle_tree.cap); lap("transcript", t0)
        t0 = time.perf_counter(); q = S.StarkyProver.quotient_commitment(stark, scfg, [tc], [], [], [], ch); lap("alphas + AIR evaluation + quotient commit", t0)
        t0 = time.perf_counter(); ch.observe_cap(q.merkle_tree.cap); zeta = ch.get_extension_challenge(); lap("transcript", t0)
        g = pow(1753635133440165772, 1 << (32 - degree_bits), P)
        t0 = time.perf_counter(); op = S.StarkOpeningSet.new(zeta, g, [tc], q); lap("StarkOpeningSet::new", t0)
        t0 = time.perf_counter()
        for b in op.to_fri_openings():
            ch.observe_extension_elements(b)
        lap("transcript", t0)
        t0 = time.perf_counter()
        P2.PolynomialBatch.prove_openings(stark.fri_instance(zeta, g, scfg), [tc, q], ch, scfg.fri_params()); lap("prove_openings", t0)
        total = (time.perf_counter() - T0) * 1e3
        if rep >= 1:
            print(f"starky proof, 2^{degree_bits} x {chip.width()}: {total:.2f} ms")
            for k, v in t.items():
                print(f"  {k:50s} {v:7.2f} ms")


if __name__ == "__main__":
    main()
