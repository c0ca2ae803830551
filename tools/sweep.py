#!/usr/bin/env python3
"""tools/sweep.py -- the non-headline BASELINE.json configs on one B200 (JSON lines to stdout).

  sweep   : configs[4], commit sweep 2^14..2^22 rows x 64..512 cols, rate_bits 1 and 3, cap_height 4
            (SURVEY.md 8d table): resident-input commit time per cell, cells that exceed HBM skipped.
  replay  : configs[0]/[2] cannot run here (no Rust toolchain, no plonky2x sources, live NEAR RPC); as SURVEY.md 8d
            prescribes they are replaced by replaying the SEQUENCE OF COMMIT SHAPES one such proof issues:
              plonk : standard_recursion_config proof at degree 2^16 -- wires 135, Z/partial products 20,
                      quotient 16 (from_coeffs), openings over the 4 oracles (constants/sigmas 84) at zeta, g*zeta
              stark : starky standard_fast_config proof, 2^16 x 418 trace (SHA-256 AIR width, sha256/air.rs:221),
                      quotient 2^16 x 4, openings at zeta and g*zeta (prover.rs:50-225)
            each timed end to end through the host mirror (host columns in, FriProof out) and, with --cpu, on the
            CPU oracle for the same sequence.
bench.py stays the contract line; the output of this tool is committed under profiles/.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as graft  # noqa: E402
from bench import synth_values, algorithmic_bytes  # noqa: E402

P = 0xFFFFFFFF00000001
HBM_BUDGET = 150e9  # bytes a cell may occupy (coeffs + lde + digests + input + NTT scratch)


def cell_bytes(log_n, W, r):
    n, N = 1 << log_n, 1 << (log_n + r)
    return 8 * W * n * 3 + 8 * W * N + 64 * N


def run_sweep(args):
    import ctypes as C
    import torch
    pkg = graft.load_package()
    P2, ffi = pkg.plonky2, pkg._ffi
    L = ffi.lib()
    ctx = P2.Context(0)
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    ctx.set_stream(stream.cuda_stream)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for r in (3, 1):
        for log_n in range(args.min_log, args.max_log + 1):
            for W in (64, 128, 256, 512):
                if cell_bytes(log_n, W, r) > HBM_BUDGET:
                    print(json.dumps({"cell": f"2^{log_n}x{W}_r{r}", "skipped": "exceeds HBM budget"}), flush=True)
                    continue
                n = 1 << log_n
                g = torch.Generator(device="cuda").manual_seed(0xC0FFEE + 4)
                dev = torch.randint(0, 1 << 62, (W, n), dtype=torch.int64, device="cuda", generator=g)
                times = []
                for it in range(args.warmup + args.steps):
                    flush.zero_()
                    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    h = ffi.vp()
                    a.record(stream)
                    ctx._check(L.nlc_commit_from_values_dev(ctx._h, ffi.vp(dev.data_ptr()), W, log_n, r, 4, C.byref(h)))
                    b.record(stream)
                    torch.cuda.synchronize()
                    L.nlc_batch_free(h)
                    if it >= args.warmup:
                        times.append(a.elapsed_time(b))
                ms = float(np.median(times))
                ab = algorithmic_bytes(log_n, W, r, 4)
                print(json.dumps({"cell": f"2^{log_n}x{W}_r{r}", "ms": ms, "elements_per_s": W * n / ms * 1e3,
                                  "lde_elements_per_s": W * (n << r) / ms * 1e3,
                                  "commit_algorithmic_gbs": ab["commit"] / ms / 1e6}), flush=True)
                del dev
                torch.cuda.empty_cache()


def proof_shapes(kind):
    """(log_n, rate_bits, cap, pow_bits, queries, oracles[(name, W, is_values, timed)], next_polys[(oracle, count)])."""
    if kind == "plonk":
        return 16, 3, 4, 16, 28, [("constants_sigmas", 84, True, False), ("wires", 135, True, True),
                                  ("zs_partial_products", 20, True, True), ("quotient", 16, False, True)], [(2, 2)]
    return 16, 1, 4, 16, 84, [("trace", 418, True, True), ("quotient", 4, False, True)], [(0, 418)]


def replay(kind, mod, commit_v, commit_c, make_inst, make_params, challenger, prove, seed):
    log_n, r, cap, pow_bits, queries, oracles, nxt = proof_shapes(kind)
    n = 1 << log_n
    mats = [synth_values(W, n, seed + i) for i, (_, W, _, _) in enumerate(oracles)]
    batches, t_commit = [], 0.0
    ch = challenger()
    for (name, W, is_values, timed), m in zip(oracles, mats):
        t0 = time.perf_counter()
        b = (commit_v if is_values else commit_c)(m, r, cap)
        c = np.array(b.merkle_tree.cap if hasattr(b, "merkle_tree") else b.cap)
        if timed:
            t_commit += time.perf_counter() - t0
        ch.observe_cap(c)
        batches.append(b)
    zeta = (0x1234567, 0x89ABCDE)
    g = pow(1753635133440165772, 1 << (32 - log_n), P)
    zeta_next = (zeta[0] * g % P, zeta[1] * g % P)
    all_polys = [(o, i) for o, (_, W, _, _) in enumerate(oracles) for i in range(W)]
    next_polys = [(o, i) for o, cnt in nxt for i in range(cnt)]
    inst = make_inst([(zeta, all_polys), (zeta_next, next_polys)])
    params = make_params(log_n, r, cap, pow_bits, queries)
    t0 = time.perf_counter()
    proof = prove(inst, batches, ch, params)
    t_open = time.perf_counter() - t0
    return t_commit, t_open, proof


def run_replay(args):
    pkg = graft.load_package()
    P2 = pkg.plonky2
    out = {}
    for kind in ("plonk", "stark"):
        def gpu_once():
            return replay(
                kind, P2,
                lambda m, r, cap: P2.PolynomialBatch.from_values(list(m), r, False, cap),
                lambda m, r, cap: P2.PolynomialBatch.from_coeffs(list(m), r, False, cap),
                lambda bs: P2.FriInstanceInfo([], [P2.FriBatchInfo(pt, [P2.FriPolynomialInfo(o, i) for o, i in polys])
                                                   for pt, polys in bs]),
                lambda log_n, r, cap, pw, q: P2.FriConfig(r, cap, pw, P2.FriReductionStrategy.ConstantArityBits(4, 5),
                                                          q).fri_params(log_n),
                P2.Challenger, P2.PolynomialBatch.prove_openings, 0xC0FFEE + 10)
        gpu_once()
        runs = [gpu_once() for _ in range(args.steps)]
        tc, to = float(np.median([x[0] for x in runs])), float(np.median([x[1] for x in runs]))
        line = {"replay": kind, "shape": proof_shapes(kind)[:5], "oracles": [o[:2] for o in proof_shapes(kind)[5]],
                "gpu_commit_s": tc, "gpu_prove_openings_s": to, "gpu_total_s": tc + to,
                "note": "host columns in, caps + FriProof out (H2D/D2H inside)"}
        if args.cpu:
            O = graft.load_oracle()
            c = replay(kind, O, O.commit_from_values, O.commit_from_coeffs, O.FriInstance,
                       lambda log_n, r, cap, pw, q: O.fri_params(log_n, r, cap, pw, q, 4, 5),
                       O.Challenger, O.prove_openings, 0xC0FFEE + 10)
            gp, cp = runs[-1][2], c[2]
            same = bool(np.array_equal(np.asarray(gp.flat), np.asarray(cp)))
            line.update({"cpu_commit_s": c[0], "cpu_prove_openings_s": c[1], "cpu_total_s": c[0] + c[1],
                         "cpu_cores": O.get_threads(), "cpu_kind": "port (oracle, OpenMP)",
                         "proof_bit_exact": same, "speedup": (c[0] + c[1]) / (tc + to)})
        print(json.dumps(line), flush=True)
        out[kind] = line
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("mode", choices=["sweep", "replay"])
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--min-log", type=int, default=14)
    ap.add_argument("--max-log", type=int, default=22)
    ap.add_argument("--cpu", action="store_true", help="replay: also time the CPU oracle on the same sequence")
    args = ap.parse_args()
    (run_sweep if args.mode == "sweep" else run_replay)(args)


if __name__ == "__main__":
    main()
