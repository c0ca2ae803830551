#!/usr/bin/env python3
"""tools/air_time.py -- time of the two constraint-interpreter calls alone (JSON lines to stdout).  Run on the GPU box:
    python tools/air_time.py [--degree-bits 16] [--only plonk|stark]
  plonk : nlc_plonk_vanishing_eval of the demo circuit (7 gates incl. Poseidon; 2^(degree_bits+3) points)
  stark : nlc_air_eval of the FpMul chip (2^16 rows x 124 columns, 2^17 points)
Times include the D2H of the accumulators (8 / 2 MB), as bench.py's `air_eval` / `quotient_commit` legs do; `digest` is a
checksum of the accumulators to compare builds by."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as graft  # noqa: E402

SETTINGS = [0]


def timed(fn, reps=5):
    import torch
    fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        out = fn()
        torch.cuda.synchronize()
        ts.append(time.perf_counter() - t0)
    return float(np.median(ts)) * 1e3, out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--degree-bits", type=int, default=16)
    ap.add_argument("--only", default="", choices=["", "plonk", "stark"])
    args = ap.parse_args()
    settings = SETTINGS
    pkg = graft.load_package()
    P2, K, S = pkg.plonky2, pkg.plonk, pkg.starky
    # plonk
    cfg = K.CircuitConfig.standard_recursion_config()
    r, cap = cfg.fri_config.rate_bits, cfg.fri_config.cap_height
    common, consts, sig, wires, pih = K.build_demo_circuit(cfg, args.degree_bits, 0xC0FFEE + 30)
    pd = K.ProverOnlyCircuitData(P2.PolynomialBatch.from_values(list(np.concatenate([consts, sig])), r, False, cap), sig)
    gw = P2.PolynomialBatch.from_values(list(wires), r, False, cap)
    betas, gammas, alphas = [3, 5], [7, 11], [13, 17]
    gz = K.PlonkProver.zs_partial_products_commitment(common, pd, gw, betas, gammas)
    prog = common.compile_gate_constraints()
    ref = None
    for s in settings if args.only != "stark" else []:
        ms, out = timed(lambda: K.PlonkProver.vanishing_values(common, pd, pih, gw, gz, betas, gammas, alphas))
        if ref is None:
            ref = out
        print(json.dumps({"kernel": "k_plonk_vanishing", "points": out.shape[0], "ops": prog.n_ops, "regs": prog.n_regs,
                          "ms": round(ms, 3), "digest": int(np.bitwise_xor.reduce(out.ravel()))}), flush=True)
    # stark
    n = 1 << 16
    rng = np.random.default_rng(0xC0FFEE + 20)
    chip = S.FpMulChip()
    base = chip.generate_trace(rng, 1 << 10)
    trace = np.ascontiguousarray(np.tile(base, (n // base.shape[0], 1)))
    stark = S.Starky(chip)
    fc = S.StarkyConfig.standard_fast_config(n).fri_config
    tb = P2.PolynomialBatch.from_trace_rows(trace, 0, chip.width(), fc.rate_bits, fc.cap_height)
    sprog = stark.program(0)
    ref = None
    for s in settings if args.only != "plonk" else []:
        ms, out = timed(lambda: S.StarkyProver.air_eval(stark, [tb], [], [], [], [3, 5]))
        if ref is None:
            ref = out
        print(json.dumps({"kernel": "k_air_eval", "points": out.shape[0], "ops": sprog.n_ops, "regs": sprog.n_regs,
                          "ms": round(ms, 3), "digest": int(np.bitwise_xor.reduce(out.ravel()))}), flush=True)


if __name__ == "__main__":
    main()
