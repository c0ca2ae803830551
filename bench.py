#!/usr/bin/env python3
"""bench.py -- LDE + Merkle commit throughput (BASELINE.json metric) on N B200s.

A "step" is one PolynomialBatch::from_values commit (iFFT -> coset LDE -> Poseidon leaf hashing -> Merkle
digests down to the cap) of one synthetic trace matrix.  Default workload = BASELINE.json configs[1]:
2^16 rows x 135 columns, rate_bits 3, cap_height 4.

  value    : input trace elements / s, whole job, inputs resident in HBM (nlc_commit_from_values_dev)
  e2e      : same metric through the host-buffer C-ABI call a plonky2 caller makes
             (nlc_commit_from_values with W host column pointers + nlc_batch_cap): H2D and D2H inside
  roofline : the dominant kernel (Poseidon leaf hashing), algorithmic bytes / CUDA-event time vs measured HBM
  cpu_baseline / --impl reference : the CPU oracle (plonky2's algorithm restated in C + OpenMP, all host cores)

Multi-GPU: independent commit instances per rank (Verify map/reduce proofs, independent STARK commits):
no data-path collective, weak scaling; time = max over ranks.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
import __graft_entry__ as graft  # noqa: E402

P = 0xFFFFFFFF00000001
WORKLOADS = {
    # name: (log_n, W, rate_bits, cap_height)
    "commit_2^16x135_r3_cap4": (16, 135, 3, 4),   # BASELINE.json configs[1] (plonk wires commit)
    "commit_2^16x20_r3_cap4": (16, 20, 3, 4),
    "commit_2^20x400_r1_cap4": (20, 400, 1, 4),   # configs[3] (starkyx trace commitment)
    "commit_2^14x64_r3_cap4": (14, 64, 3, 4),
    "commit_2^16x2502_r1_cap4": (16, 2502, 1, 4),  # ed25519 AIR trace in the tree (machine/ec/builder.rs:301-303)
    "commit_2^17x912_r1_cap4": (17, 912, 1, 4),    # SHA-256 AIR trace (sha256/air.rs:221-222)
}
DEFAULT_WORKLOAD = "commit_2^16x135_r3_cap4"


def synth_values(W: int, n: int, seed: int) -> np.ndarray:
    """i.i.d. uniform in [0, p): PCG64(seed), reject >= p."""
    rng = np.random.Generator(np.random.PCG64(seed))
    a = rng.integers(0, 1 << 64, size=(W, n), dtype=np.uint64)
    bad = a >= np.uint64(P)
    while bad.any():
        a[bad] = rng.integers(0, 1 << 64, size=int(bad.sum()), dtype=np.uint64)
        bad = a >= np.uint64(P)
    return a


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.index)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def ncu_traffic(workload: str, kernel: str):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel`, from the committed `ncu --set full`
    capture of this workload (profiles/traffic.json names the capture); None when no capture exists."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return json.load(f)[workload][kernel]["dram_bytes_per_launch"]
    except Exception:
        return None


def ncu_pipes(workload: str, kernel: str):
    """Pipe utilisation of `kernel` from the same committed capture (explains an HBM fraction of ~1 %)."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return json.load(f)[workload][kernel]["pipes_pct_of_peak"]
    except Exception:
        return None


def kernel_source_hash(files):
    """sha256 over the kernel sources a committed ncu capture describes: a capture whose hash differs from the tree's
    is reported as stale instead of being passed off as a measurement of the current kernel."""
    import hashlib
    h = hashlib.sha256()
    for f in files:
        with open(os.path.join(ROOT, "near-light-client_b200", "csrc", f), "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()[:16]


def ncu_capture_meta(workload: str, kernel: str):
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            d = json.load(f)[workload][kernel]
        return d.get("source_files"), d.get("source_sha256_16"), d.get("capture")
    except Exception:
        return None, None, None


def algorithmic_bytes(log_n, W, r, cap):
    n, N = 1 << log_n, 1 << (log_n + r)
    return {
        "commit": 8 * W * n + 8 * W * n + 8 * W * N + 64 * (N - (1 << cap)),  # SURVEY.md 8(d) B_commit
        "ntt": 8 * W * n + 8 * W * N,                                          # B_ntt (LDE only)
        "ifft": 16 * W * n,
        "leaf_hash": 8 * W * N + 32 * N,                                       # read every LDE row once, write N digests
        "levels": 96 * (N - (1 << cap)),
    }


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


def _gpu_replay_once(kind, P2):
    return replay(
        kind, P2,
        lambda m, r, cap: P2.PolynomialBatch.from_values(list(m), r, False, cap),
        lambda m, r, cap: P2.PolynomialBatch.from_coeffs(list(m), r, False, cap),
        lambda bs: P2.FriInstanceInfo([], [P2.FriBatchInfo(pt, [P2.FriPolynomialInfo(o, i) for o, i in polys])
                                           for pt, polys in bs]),
        lambda log_n, r, cap, pw, q: P2.FriConfig(r, cap, pw, P2.FriReductionStrategy.ConstantArityBits(4, 5),
                                                  q).fri_params(log_n),
        P2.Challenger, P2.PolynomialBatch.prove_openings, 0xC0FFEE + 10)


def replay_gpu(kind, steps, keep_proof=False):
    """One proof's commit path (configs[0]/[2] replaced by their sequence of commit shapes) through
    the host mirror: pageable host columns in, caps + FriProof out."""
    P2 = graft.load_package().plonky2
    _gpu_replay_once(kind, P2)
    runs = [_gpu_replay_once(kind, P2) for _ in range(steps)]
    tc, to = float(np.median([x[0] for x in runs])), float(np.median([x[1] for x in runs]))
    line = {"replay": kind, "shape": proof_shapes(kind)[:5], "oracles": [o[:2] for o in proof_shapes(kind)[5]],
            "gpu_commit_s": tc, "gpu_prove_openings_s": to, "gpu_total_s": tc + to,
            "note": "host columns in, caps + FriProof out (H2D/D2H inside)"}
    return (line, runs[-1][2]) if keep_proof else line


def run_replay_workload(args):
    """--workload replay_plonk / replay_stark: time per proof of the commit path, GPU vs the CPU oracle."""
    kind = args.workload.split("_")[1]
    line, gproof = replay_gpu(kind, args.steps, keep_proof=True)
    O = graft.load_oracle()
    O.set_threads(len(os.sched_getaffinity(0)))
    c = replay(kind, O, O.commit_from_values, O.commit_from_coeffs, O.FriInstance,
               lambda log_n, r, cap, pw, q: O.fri_params(log_n, r, cap, pw, q, 4, 5),
               O.Challenger, O.prove_openings, 0xC0FFEE + 10)
    same = bool(np.array_equal(np.asarray(gproof.flat), np.asarray(c[2])))
    out = {"metric": "commit-path time per proof (commits + prove_openings), host to host", "value": line["gpu_total_s"],
           "unit": "s", "n_gpus": 1, "steps": args.steps, "warmup": 1, "ms_per_step": line["gpu_total_s"] * 1e3,
           "higher_is_better": False, "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
           "config": {"workload": args.workload, "shape": line["shape"], "oracles": line["oracles"]},
           "gpu_commit_s": line["gpu_commit_s"], "gpu_prove_openings_s": line["gpu_prove_openings_s"],
           "cpu_baseline": {"value": c[0] + c[1], "unit": "s", "cores": O.get_threads(), "kind": "port",
                            "sample": f"the same proof once: commits {c[0]:.2f} s, prove_openings {c[1]:.2f} s",
                            "proof_bit_exact": same}}
    print(json.dumps(out), flush=True)


def run_stark_proof(args):
    """--workload stark_proof_fpmul | stark_proof_ed_add | stark_proof_ed_scalar_mul: ONE WHOLE starky proof (prover.rs:50-225) of a field-arithmetic
    chip (chip/field/mul.rs FpMulInstruction<Fp25519>: 124 columns, 31 degree-2 constraints) or of the ed25519
    point-addition gadget (chip/ec/edwards/add.rs: 800 columns, 248 constraints) or of the ed25519 scalar multiplication
    (machine/ec/builder.rs scalar_mul_batch: 1651 columns, 720 constraints incl. transitions) at 2^16 rows under
    standard_fast_config -- trace commit from the row-major AirTrace, AIR evaluation over the quotient domain ON THE
    DEVICE (csrc/air.cu), quotient commit, openings, prove_openings -- against the same sequence driven through the CPU
    oracle (direct AIR evaluation, oracle/air.py).  The `air_eval` leg is timed alone on both sides."""
    import torch
    from oracle import air as OA
    pkg = graft.load_package()
    P2, S = pkg.plonky2, pkg.starky
    O = graft.load_oracle()
    O.set_threads(len(os.sched_getaffinity(0)))
    log_n = 16
    n = 1 << log_n
    rng = np.random.default_rng(0xC0FFEE + 20)
    if args.workload == "stark_proof_ed_scalar_mul":  # machine/ec/builder.rs scalar_mul_batch: 2^16 x 1651, 720 constraints
        chip, what = S.EdScalarMulChip(), "valid ed25519 scalar-multiplication trace: 256 multiplications of 256 rows, two distinct ones repeated"
        trace = chip.generate_trace(rng, n, distinct=2)
    else:
        if args.workload == "stark_proof_ed_add":  # chip/ec/edwards/add.rs ed_add: 800 columns, 248 constraints
            chip, base_rows, what = S.EdAddChip(), 1 << 6, "valid ed25519 point-addition trace: 2^6 additions of multiples of the base point, tiled"
        else:
            chip, base_rows, what = S.FpMulChip(), 1 << 10, "valid FpMul trace: 2^10 random multiplications mod 2^255-19, tiled"
        base = chip.generate_trace(rng, base_rows)
        trace = np.ascontiguousarray(np.tile(base, (n // base.shape[0], 1)))  # rows are independent: repetition stays valid
    stark = S.Starky(chip)
    cfg = S.StarkyConfig.standard_fast_config(n)
    fc = cfg.fri_config

    def gpu_once():
        t0 = time.perf_counter()
        proof = S.StarkyProver.prove(cfg, stark, trace, [])
        return time.perf_counter() - t0, proof
    gpu_once()
    runs = [gpu_once() for _ in range(max(args.steps, 1))]
    t_gpu = float(np.median([r[0] for r in runs]))
    proof = runs[-1][1]
    # the AIR evaluation alone (accumulators of every alpha for all 2^17 points, result left on the host)
    tb = P2.PolynomialBatch.from_trace_rows(trace, 0, chip.width(), fc.rate_bits, fc.cap_height)
    alphas = [3, 5]
    S.StarkyProver.air_eval(stark, [tb], [], [], [], alphas)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(3):
        acc_gpu = S.StarkyProver.air_eval(stark, [tb], [], [], [], alphas)
    t_air_gpu = (time.perf_counter() - t0) / 3
    # -- the oracle-driven proof, in the order of prover.rs:50-225
    t0 = time.perf_counter()
    ch = O.Challenger()
    ch.observe_elements([])
    tref = O.commit_from_values(np.ascontiguousarray(trace.T), fc.rate_bits, fc.cap_height)
    t_commit_cpu = time.perf_counter() - t0
    ch.observe_cap(tref.cap)
    al = ch.get_n_challenges(cfg.num_challenges)
    t1 = time.perf_counter()
    qdf = chip.quotient_degree_factor()
    qdb = max(qdf - 1, 0).bit_length()
    qv = OA.quotient_values(chip, [tref], [], [], [], al, log_n, fc.rate_bits)
    t_air_cpu = time.perf_counter() - t1
    qref = O.quotient_commit(qv, log_n, qdb, qdf, False, fc.rate_bits, fc.cap_height)
    ch.observe_cap(qref.cap)
    zeta = ch.get_extension_challenge()
    g = pow(1753635133440165772, 1 << (32 - log_n), P)
    zeta_next = (zeta[0] * g % P, zeta[1] * g % P)
    ev = lambda b, z: [O.eval_poly_ext(c, np.array(z, np.uint64)) for c in b.coeffs]
    local, nxt, quot = ev(tref, zeta), ev(tref, zeta_next), ev(qref, zeta)
    for batch in (local + quot, nxt):
        ch.observe_elements([c for e in batch for c in e])
    params = O.fri_params(log_n, fc.rate_bits, fc.cap_height, fc.proof_of_work_bits, fc.num_query_rounds, 4, 5)
    W, Q = chip.width(), qref.coeffs.shape[0]
    inst = O.FriInstance([(zeta, [(0, i) for i in range(W)] + [(1, i) for i in range(Q)]), (zeta_next, [(0, i) for i in range(W)])])
    oproof = O.prove_openings(inst, [tref, qref], ch, params)
    t_cpu = time.perf_counter() - t0
    acc_cpu = OA.quotient_values(chip, [tref], [], [], [], alphas, log_n, fc.rate_bits, divide_by_zh=False)
    same = {"trace_cap": bool(np.array_equal(proof.trace_caps[0], tref.cap)),
            "quotient_cap": bool(np.array_equal(proof.quotient_polys_cap, qref.cap)),
            "opening_proof": bool(np.array_equal(np.asarray(proof.opening_proof.flat), np.asarray(oproof))),
            "air_accumulators": bool(np.array_equal(acc_gpu, acc_cpu))}
    prog = stark.program(0)
    out = {"metric": "STARK proof time (trace rows in, StarkProof out)", "value": t_gpu, "unit": "s", "n_gpus": 1,
           "steps": args.steps, "warmup": 1, "ms_per_step": t_gpu * 1e3, "higher_is_better": False, "scaling": "weak",
           "vs_baseline": None, "dtype": "u64", "data": f"synthetic ({what})",
           "config": {"workload": args.workload, "rows": n, "cols": W, "rate_bits": fc.rate_bits, "cap_height": fc.cap_height,
                      "num_challenges": cfg.num_challenges, "queries": fc.num_query_rounds, "constraints": prog.n_constraints,
                      "bytecode_ops": prog.n_ops, "bytecode_registers": prog.n_regs},
           "air_eval": {"gpu_s": t_air_gpu, "cpu_oracle_s": t_air_cpu, "points": n << qdb,
                        "note": "GPU: nlc_air_eval incl. the D2H of the accumulators; oracle: numpy/C vector ops over the whole domain"},
           "cpu_baseline": {"value": t_cpu, "unit": "s", "cores": O.get_threads(), "kind": "port",
                            "sample": f"the same proof once through the oracle: trace commit {t_commit_cpu:.2f} s, AIR evaluation {t_air_cpu:.2f} s, total {t_cpu:.2f} s",
                            "bit_exact": same}}
    print(json.dumps(out), flush=True)
    if not all(same.values()):
        raise SystemExit(f"bench.py: stark proof differs from the oracle's: {same}")


def run_plonk_proof(args):
    """--workload plonk_proof_demo: ONE WHOLE plonk proof from the witness on (plonk/prover.rs `prove`) under
    standard_recursion_config at degree 2^16 -- wires commit (135), partial products + Z (parallel prefix product on the
    device) and their commit (20), the vanishing polynomial over the 2^19-point quotient domain with the gate constraints
    run from bytecode, quotient commit (16), openings, prove_openings over the four oracles -- against the same sequence
    through the CPU oracle (oracle/plonk.py).  Circuit: plonk.build_demo_circuit (arithmetic, extension-arithmetic,
    base-sum, constant, public-input gates; random copy constraints; a satisfying witness)."""
    import torch
    from oracle import plonk as OP
    pkg = graft.load_package()
    P2, K = pkg.plonky2, pkg.plonk
    O = graft.load_oracle()
    O.set_threads(len(os.sched_getaffinity(0)))
    degree_bits = args.plonk_degree_bits
    cfg = K.CircuitConfig.standard_recursion_config()
    fc = cfg.fri_config
    r, cap = fc.rate_bits, fc.cap_height
    t0 = time.perf_counter()
    common, consts, sig, wires, pih = K.build_demo_circuit(cfg, degree_bits, 0xC0FFEE + 30)
    t_build = time.perf_counter() - t0
    cs_vals = np.concatenate([consts, sig])
    pd = K.ProverOnlyCircuitData(P2.PolynomialBatch.from_values(list(cs_vals), r, False, cap), sig)
    wire_cols = list(wires)

    def gpu_once():
        t0 = time.perf_counter()
        proof = K.PlonkProver.prove(common, pd, wire_cols, pih)
        return time.perf_counter() - t0, proof
    gpu_once()
    runs = [gpu_once() for _ in range(max(args.steps, 1))]
    t_gpu = float(np.median([x[0] for x in runs]))
    proof = runs[-1][1]
    # the two new legs alone
    gw = P2.PolynomialBatch.from_values(wire_cols, r, False, cap)
    betas, gammas, alphas = [3, 5], [7, 11], [13, 17]

    def timed(fn, reps=3):
        fn()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(reps):
            out = fn()
        torch.cuda.synchronize()
        return (time.perf_counter() - t0) / reps, out
    t_pp_gpu, gz = timed(lambda: K.PlonkProver.zs_partial_products_commitment(common, pd, gw, betas, gammas))
    t_q_gpu, _ = timed(lambda: K.PlonkProver.quotient_commitment(common, pd, pih, gw, gz, betas, gammas, alphas))
    # -- the oracle-driven proof
    t0 = time.perf_counter()
    rcs = O.commit_from_values(cs_vals, r, cap)  # preprocessing, not part of `prove`
    t_pre = time.perf_counter() - t0
    t0 = time.perf_counter()
    ch = O.Challenger()
    ch.observe_elements(pd.circuit_digest)
    ch.observe_elements(pih)
    rw = O.commit_from_values(wires, r, cap)
    ch.observe_cap(rw.cap)
    b2, g2 = ch.get_n_challenges(2), ch.get_n_challenges(2)
    t1 = time.perf_counter()
    vals = OP.partial_products_and_zs(common, wires, sig, b2, g2)
    t_pp_cpu = time.perf_counter() - t1
    rz = O.commit_from_values(vals, r, cap)
    ch.observe_cap(rz.cap)
    a2 = ch.get_n_challenges(2)
    t1 = time.perf_counter()
    qv = OP.vanishing_values(common, rcs, rw, rz, pih, b2, g2, a2, r)
    t_van_cpu = time.perf_counter() - t1
    rq = O.quotient_commit(qv, degree_bits, 3, 8, False, r, cap)
    ch.observe_cap(rq.cap)
    zeta = ch.get_extension_challenge()
    g = pow(1753635133440165772, 1 << (32 - degree_bits), P)
    zeta_next = (zeta[0] * g % P, zeta[1] * g % P)
    ev = lambda b, z: [O.eval_poly_ext(c, np.array(z, np.uint64)) for c in b.coeffs]
    cs_z, w_z, zp_z, q_z, zp_n = ev(rcs, zeta), ev(rw, zeta), ev(rz, zeta), ev(rq, zeta), ev(rz, zeta_next)
    for batch in (cs_z + w_z + zp_z + q_z, zp_n[:2]):
        ch.observe_elements([c for e in batch for c in e])
    params = O.fri_params(degree_bits, r, cap, fc.proof_of_work_bits, fc.num_query_rounds, 4, 5)
    widths = [rcs.coeffs.shape[0], rw.coeffs.shape[0], rz.coeffs.shape[0], rq.coeffs.shape[0]]
    inst = O.FriInstance([(zeta, [(oi, i) for oi, wd in enumerate(widths) for i in range(wd)]), (zeta_next, [(2, 0), (2, 1)])])
    oproof = O.prove_openings(inst, [rcs, rw, rz, rq], ch, params)
    t_cpu = time.perf_counter() - t0
    same = {"wires_cap": bool(np.array_equal(proof.wires_cap, rw.cap)),
            "zs_partial_products_cap": bool(np.array_equal(proof.plonk_zs_partial_products_cap, rz.cap)),
            "quotient_cap": bool(np.array_equal(proof.quotient_polys_cap, rq.cap)),
            "opening_proof": bool(np.array_equal(np.asarray(proof.opening_proof.flat), np.asarray(oproof)))}
    prog = common.compile_gate_constraints()
    out = {"metric": "plonk proof time (witness columns in, proof out)", "value": t_gpu, "unit": "s", "n_gpus": 1,
           "steps": args.steps, "warmup": 1, "ms_per_step": t_gpu * 1e3, "higher_is_better": False, "scaling": "weak",
           "vs_baseline": None, "dtype": "u64", "data": "synthetic (demo circuit with a satisfying witness)",
           "config": {"workload": args.workload, "degree_bits": degree_bits, "num_wires": cfg.num_wires,
                      "num_routed_wires": cfg.num_routed_wires, "rate_bits": r, "cap_height": cap, "queries": fc.num_query_rounds,
                      "gates": [gt.id() for gt in common.gates], "gate_constraints": common.num_gate_constraints,
                      "bytecode_ops": prog.n_ops, "bytecode_registers": prog.n_regs},
           "zs_partial_products_commit": {"gpu_s": t_pp_gpu, "cpu_oracle_partial_products_s": t_pp_cpu},
           "quotient_commit": {"gpu_s": t_q_gpu, "cpu_oracle_vanishing_s": t_van_cpu, "points": 1 << (degree_bits + 3)},
           "cpu_baseline": {"value": t_cpu, "unit": "s", "cores": O.get_threads(), "kind": "port",
                            "sample": f"the same proof once through the oracle ({t_cpu:.2f} s; numpy-vector restatement of the "
                                      f"quotient side, C/OpenMP commits); circuit build {t_build:.1f} s and the constants/sigmas "
                                      f"commit ({t_pre:.2f} s) are preprocessing on both sides", "bit_exact": same}}
    print(json.dumps(out), flush=True)
    if not all(same.values()):
        raise SystemExit(f"bench.py: plonk proof differs from the oracle's: {same}")


def oracle_quality(O):
    """Single-thread rates of the CPU port, so that the quality of the baseline is on record beside every GPU/CPU
    factor: Poseidon permutations/s/core (fast partial rounds, as upstream runs them) and FFT butterflies/s/core."""
    rng = np.random.default_rng(3)
    threads = O.get_threads()
    O.set_threads(1)
    leaves = O.random_field(rng, (1 << 10, 135))
    O.merkle_new(leaves[:16], 0)
    t0 = time.perf_counter()
    O.merkle_new(leaves, 4)
    dt = time.perf_counter() - t0
    perms = (1 << 10) * 17 + (1 << 10) - 16
    col = O.random_field(rng, (1 << 16,))
    O.fft(col)
    t0 = time.perf_counter()
    for _ in range(4):
        O.fft(col)
    df = (time.perf_counter() - t0) / 4
    O.set_threads(threads)
    return {"poseidon_perms_per_s_per_core": perms / dt, "fft_butterflies_per_s_per_core": 16 * (1 << 15) / df,
            "note": "C restatement (scalar 64-bit mulq, fast partial rounds, OpenMP); NOT plonky2's AVX2/asm build"}


def run_reference(args, rank, world):
    """The reference's CPU algorithm for the path (oracle port; plonky2 itself cannot be built here), on the
    WORKLOAD'S OWN SIZE by default; --ref-sample-log-rows N is an explicit opt-in to a bounded row sample."""
    if rank != 0:
        return
    O = graft.load_oracle()
    log_n, W, r, cap = WORKLOADS[args.workload]
    full_log_n = log_n
    sampled = args.ref_sample_log_rows is not None and args.ref_sample_log_rows < log_n
    if sampled:
        log_n = args.ref_sample_log_rows  # bounded sample: same columns / rate / cap, fewer rows
    n = 1 << log_n
    values = synth_values(W, n, 0xC0FFEE + 1)
    # torchrun exports OMP_NUM_THREADS=1 to every rank: the reference arm runs on rank 0 alone and takes every core
    O.set_threads(len(os.sched_getaffinity(0)))
    cores = O.get_threads()
    for _ in range(max(1, min(args.warmup, 1))):
        O.commit_from_values(values, r, cap)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        O.commit_from_values(values, r, cap)
    dt = (time.perf_counter() - t0) / args.steps
    val = W * n / dt
    config = {"workload": args.workload, "rows": 1 << full_log_n, "cols": W, "rate_bits": r, "cap_height": cap}
    sample = f"{args.steps} full commits of {args.workload} (OpenMP over columns/rows/subtrees)"
    if sampled:
        config["sample_rows"] = n
        sample = (f"{args.steps} commits of 2^{log_n} of the workload's 2^{full_log_n} rows x {W} cols, rate_bits {r} "
                  f"(same per-row work; OpenMP over columns/rows/subtrees)")
    line = {"impl": "reference", "metric": "LDE+Merkle commit throughput (input trace elements/s)", "value": val,
            "unit": "elements/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64",
            "data": "synthetic", "config": config,
            "cpu_baseline": {"value": val, "unit": "elements/s", "cores": cores, "kind": "port", "sample": sample,
                             "port_quality": oracle_quality(O)},
            "e2e": {"value": val, "unit": "elements/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


class Strong:
    """ONE commit split over all ranks (strong scaling), measured inside the same process group as the weak-scaling
    line so that the driver's SCALE record carries it: CUDA-event timed, L2 flushed, own clock sample.
      coset  (world <= 2^rate_bits): rank p uploads + inverse-transforms its columns only, the iFFT's
             stores put the coefficients into every rank's matrix over NVLink, barrier, every rank extends + hashes its
             cosets, cap entries all-gathered;
      rows   (world > 2^rate_bits): column shards for the transforms, last NTT pass stores into the
             peers' leaf-row buffers over NVLink, barrier, leaf-row shards for the tree, cap all-gather."""

    def __init__(self, pkg, ctx, stream, rank, world, workload, flush, device_values=False):
        """workload: a WORKLOADS name or a (log_n, W, rate_bits, cap_height) tuple; device_values: this rank's columns
        are drawn on the device (tools/sweep.py: cells too large to generate on every rank's host; no 1-GPU check)."""
        import ctypes as C
        import torch
        self.C, self.torch = C, torch
        self.pkg, self.ctx, self.stream, self.rank, self.world, self.flush = pkg, ctx, stream, rank, world, flush
        self.ffi, self.L = pkg._ffi, pkg._ffi.lib()
        self.workload = workload
        self.log_n, self.W, self.r, self.cap = WORKLOADS[workload] if isinstance(workload, str) else workload
        self.n, self.N = 1 << self.log_n, 1 << (self.log_n + self.r)
        self.by_rows = world > (1 << self.r)
        S = pkg.sharding
        if self.by_rows:
            S.check_row_plan(self.cap, world)
            self.split = S.column_split(self.W, world)
            self.per_rows = self.N // world
            self.peers = S.PeerRows(ctx, self.W * self.per_rows)
        else:
            S.check_plan(self.r, self.cap, world)
            self.per_cols, self.split = S.uniform_column_split(self.W, world)
            self.peers = S.PeerRows(ctx, world * self.per_cols * self.n)
        self.c0, self.nc = self.split[rank]
        self.peer_table = (self.ffi.vp * world)(*[self.ffi.vp(p) for p in self.peers.ptrs])
        self.tiny = torch.zeros(1, dtype=torch.int32, device="cuda")
        self.cap_local = torch.empty(((1 << self.cap) // world, 4), dtype=torch.int64, device="cuda")
        self.cap_full = torch.empty((1 << self.cap, 4), dtype=torch.int64, device="cuda")
        self.cudart = C.CDLL("libcudart.so.12")
        if device_values:
            g = torch.Generator(device="cuda").manual_seed(0xC0FFEE + 4 + rank)
            self.dev_mine = torch.randint(0, 1 << 62, (max(self.nc, 1), self.n), dtype=torch.int64, device="cuda", generator=g)
            self.host_mine, self.full_values = None, None
        else:
            # every rank holds the same trace (seed without the rank); only this rank's columns are ever touched
            full = synth_values(self.W, self.n, 0xC0FFEE + 1)
            self.host_mine = torch.empty((max(self.nc, 1), self.n), dtype=torch.int64).pin_memory()
            if self.nc:
                self.host_mine.numpy().view(np.uint64)[:self.nc] = full[self.c0:self.c0 + self.nc]
            self.dev_mine = self.host_mine.cuda(non_blocking=True)
            self.full_values = full
        self.my_coeffs = torch.empty((max(self.nc, 1), self.n), dtype=torch.int64, device="cuda") if self.by_rows else None
        self.batch = None

    def free_prev(self):
        if self.batch is not None:
            self.L.nlc_batch_free(self.batch)
            self.batch = None

    def commit(self, src_cols):
        import torch.distributed as dist
        C, ffi, L, ctx = self.C, self.ffi, self.L, self.ctx
        self.free_prev()
        h = ffi.vp()
        if self.by_rows:
            ctx._check(L.nlc_lde_columns_scatter_dev(ctx._h, ffi.vp(src_cols.data_ptr()), self.nc, self.log_n, self.r, 1,
                                                     ffi.vp(self.my_coeffs.data_ptr()), None, self.world, self.peer_table,
                                                     self.c0))
            dist.all_reduce(self.tiny)  # barrier in stream order: every rank's stores have landed
            ctx._check(L.nlc_commit_from_lde_rows_dev(ctx._h, ffi.vp(self.peers.own), self.W, self.log_n, self.r, self.cap,
                                                      self.rank, self.world, C.byref(h)))
        else:
            if self.nc:
                ctx._check(L.nlc_ifft_columns_bcast_dev(ctx._h, ffi.vp(src_cols.data_ptr()), self.nc, self.log_n, self.c0,
                                                        self.world, self.peer_table))
            dist.all_reduce(self.tiny)
            ctx._check(L.nlc_commit_from_coeffs_shard_dev(ctx._h, ffi.vp(self.peers.own), self.W, self.log_n, self.r,
                                                          self.cap, self.rank, self.world, C.byref(h)))
        self.batch = h
        self.cudart.cudaMemcpyAsync(C.c_void_p(self.cap_local.data_ptr()), C.c_void_p(L.nlc_batch_cap_dev(h)),
                                    C.c_size_t(self.cap_local.numel() * 8), 3, C.c_void_p(self.stream.cuda_stream))
        dist.all_gather_into_tensor(self.cap_full, self.cap_local)  # also the "buffers free" barrier for the next commit

    def step_resident(self):
        self.commit(self.dev_mine)

    def step_e2e(self):
        """this rank's columns from pinned host memory over PCIe, the full cap back on the host"""
        self.commit(self.host_mine.cuda(non_blocking=True))
        return self.cap_full.cpu().numpy().view(np.uint64)

    def measure(self, steps):
        import torch.distributed as dist
        torch = self.torch
        for _ in range(3):
            self.flush.zero_()
            self.step_resident()
        torch.cuda.synchronize()
        sampler = ClockSampler(torch.cuda.current_device())
        sampler.start()
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        dist.barrier()
        torch.cuda.synchronize()
        for a, b in evs:
            self.flush.zero_()
            a.record(self.stream)
            self.step_resident()
            b.record(self.stream)
        torch.cuda.synchronize()
        dist.barrier()
        t = torch.tensor([sum(a.elapsed_time(b) for a, b in evs) / steps], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        # a timed region shorter than nvidia-smi's start-up + sampling period would carry no clock sample: keep the same
        # commits running (untimed; the count follows from the all-reduced time, so every rank runs as many) while it samples
        extra = 0 if steps * ms >= 400.0 else min(2000, int(400.0 / max(ms, 0.05)))
        for _ in range(extra):
            self.flush.zero_()
            self.step_resident()
        torch.cuda.synchronize()
        dist.barrier()
        clocks = sampler.stop()
        if extra:
            clocks["window"] = f"the {steps} timed commits + {extra} more of the same, untimed"
        # phases (this rank's CUDA events; max over ranks per phase)
        self.ctx.timing_enable(True)
        self.ctx.timing_read()
        for _ in range(3):
            self.flush.zero_()
            self.step_resident()
        ph = self.ctx.timing_read()
        self.ctx.timing_enable(False)
        k = max(ph["commits"], 1)
        names = ["IFFT", "FFT + blinding", "leaf hashing", "digest levels"]
        pt = torch.tensor([ph[nm] / k for nm in names], dtype=torch.float64, device="cuda")
        dist.all_reduce(pt, op=dist.ReduceOp.MAX)
        phases = {nm: float(v) for nm, v in zip(names, pt.tolist())}
        if self.host_mine is None:  # sweep cells drawn on the device: resident figures only
            self.free_prev()
            return ms, phases, None, clocks, None
        # e2e: host columns in, cap out
        for _ in range(2):
            self.step_e2e()
        dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(steps):
            cap_host = self.step_e2e()
        torch.cuda.synchronize()
        te = torch.tensor([(time.perf_counter() - t0) / steps * 1e3], dtype=torch.float64, device="cuda")
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
        self.free_prev()
        return ms, phases, float(te.item()), clocks, cap_host


def n1_resident_ms(pkg, ctx, stream, workload, flush, values, steps=3):
    """The same commit on ONE GPU, resident input, in this process: the denominator of the strong-scaling factor."""
    import ctypes as C
    import torch
    ffi, L = pkg._ffi, pkg._ffi.lib()
    log_n, W, r, cap = WORKLOADS[workload]
    dev = torch.from_numpy(values.view(np.int64)).cuda()
    torch.cuda.synchronize()
    h = [None]

    def step():
        if h[0] is not None:
            L.nlc_batch_free(h[0])
        hh = ffi.vp()
        ctx._check(L.nlc_commit_from_values_dev(ctx._h, ffi.vp(dev.data_ptr()), W, log_n, r, cap, C.byref(hh)))
        h[0] = hh
    for _ in range(2):
        step()
    torch.cuda.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for a, b in evs:
        flush.zero_()
        a.record(stream)
        step()
        b.record(stream)
    torch.cuda.synchronize()
    cap_host = np.zeros((1 << cap, 4), np.uint64)
    ctx._check(L.nlc_batch_cap(h[0], cap_host.ctypes.data_as(ffi.u64p)))
    L.nlc_batch_free(h[0])
    del dev
    torch.cuda.empty_cache()
    return sum(a.elapsed_time(b) for a, b in evs) / steps, cap_host


def strong_scaling_block(pkg, ctx, stream, rank, world, flush, steps):
    """`strong` of the bench line at N > 1: configs[1] and configs[3] as ONE commit over all ranks."""
    import torch
    import torch.distributed as dist
    out = {}
    for name in ("commit_2^16x135_r3_cap4", "commit_2^20x400_r1_cap4"):
        log_n, W, r, cap = WORKLOADS[name]
        if not (world > (1 << r) or world <= (1 << min(r, cap))):
            continue
        st = Strong(pkg, ctx, stream, rank, world, name, flush)
        ms, phases, e2e_ms, clocks, cap_sharded = st.measure(max(3, min(steps, 10)))
        n1_ms, cap_n1 = n1_resident_ms(pkg, ctx, stream, name, flush, st.full_values)
        same = torch.tensor([int(np.array_equal(cap_sharded, cap_n1))], dtype=torch.int32, device="cuda")
        dist.all_reduce(same, op=dist.ReduceOp.MIN)
        resid = max(phases, key=phases.get)
        accounted = sum(phases.values())
        out[name] = {"split": ("column shards -> NTT stores into peers over NVLink -> leaf-row shards (8e.3)" if st.by_rows else
                               "column-sharded upload + iFFT, coefficient all-gather by the iFFT's NVLink stores, coset shards (8e.2)"),
                     "ms_per_commit": ms, "phases_ms_max_over_ranks": phases,
                     "unaccounted_ms (barriers, cap all-gather, launch gaps)": ms - accounted,
                     "e2e_ms": e2e_ms, "e2e_h2d_bytes_per_rank": 8 * st.nc * st.n, "n1_resident_ms_same_process": n1_ms,
                     "speedup_vs_1gpu": n1_ms / ms, "efficiency": n1_ms / ms / world, "largest_phase": resid,
                     "cap_equals_1gpu_commit_on_every_rank": bool(same.item()), "clocks": clocks}
        st.peers.free()
        del st
        torch.cuda.empty_cache()
        dist.barrier()
    return out


def sharded_parity_block(pkg, ctx, rank, world):
    """Multi-GPU parity where the driver has the GPUs (tests/test_gpu_multi.py is skipped on its 1-GPU test box): one
    coset-sharded (both forms) and one row-sharded (p2p and nccl) commit of small traces against the ORACLE (checker
    leg, after every timed region): cap on every rank, this rank's leaves and digests."""
    import torch
    import torch.distributed as dist
    O = graft.load_oracle()
    P2, S = pkg.plonky2, pkg.sharding
    res = {}
    rng = np.random.default_rng(5)
    cases = {"coset": (135, 12, 3, 4), "coset_bcast": (135, 12, 3, 4), "rows_p2p": (418, 12, 1, 4), "rows_nccl": (418, 12, 1, 4)}
    refs = {}
    for key, (W, log_n, r, cap) in cases.items():
        shape = (W, log_n, r, cap)
        if shape not in refs:
            values = O.random_field(rng, (W, 1 << log_n))
            refs[shape] = (values, O.commit_from_values(values, r, cap))
        values, ref = refs[shape]
        lo, hi = S.leaf_range(log_n, r, rank, world)
        per_d = ref.digests.shape[0] // world
        ok = True
        try:
            if key.startswith("coset"):
                if world > (1 << r):
                    res[key] = None
                    continue
                f = P2.PolynomialBatch.from_values_sharded if key == "coset" else P2.PolynomialBatch.from_values_coset_sharded
                b = f(list(values), r, False, cap, ctx=ctx)
            else:
                b = P2.PolynomialBatch.from_values_row_sharded(list(values), r, False, cap, ctx=ctx, exchange=key.split("_")[1])
            ok = (np.array_equal(b.merkle_tree.cap, ref.cap) and np.array_equal(b.merkle_tree.leaves, ref.leaves[lo:hi])
                  and np.array_equal(b.merkle_tree.digests, ref.digests[per_d * rank:per_d * (rank + 1)]))
            b.free()
            del b
        except Exception as e:  # a failing rank must not hang the others in a collective: report, then fail below
            print(f"[rank {rank}] sharded parity {key}: {e!r}", file=sys.stderr, flush=True)
            ok = False
        t = torch.tensor([int(ok)], dtype=torch.int32, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        res[key] = bool(t.item())
        dist.barrier()
    return res


def run_ours(args, rank, local_rank, world):
    import ctypes as C

    import torch
    import torch.distributed as dist

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    pkg = graft.load_package()
    P2, ffi = pkg.plonky2, pkg._ffi
    L = ffi.lib()
    ctx = P2.Context(local_rank)
    # torch's default stream is the NULL stream, which the C ABI reads as "use your own stream": make a real
    # stream current so that the library's kernels and torch.cuda.Event live on the same one
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    assert stream.cuda_stream != 0
    ctx.set_stream(stream.cuda_stream)

    log_n, W, r, cap = WORKLOADS[args.workload]
    n, N = 1 << log_n, 1 << (log_n + r)
    elems = W * n
    host = torch.empty((W, n), dtype=torch.int64).pin_memory()
    host.numpy().view(np.uint64)[:] = synth_values(W, n, 0xC0FFEE + 1 + (0 if args.shard_commit else rank))
    dev = host.cuda(non_blocking=True)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2
    cap_host = np.zeros((1 << cap, 4), np.uint64)
    col_table = (ffi.u64p * W)(*[C.cast(host[j].data_ptr(), ffi.u64p) for j in range(W)])

    def barrier():
        if world > 1:
            dist.barrier()

    state = {"batch": None}

    def free_prev():
        if state["batch"] is not None:
            L.nlc_batch_free(state["batch"])
            state["batch"] = None

    sharded = bool(args.shard_commit) and world > 1
    by_rows = sharded and world > (1 << r)  # more ranks than cosets: SURVEY.md 8(e).3
    if sharded:  # ONE commit split by coset over the ranks, cap entries all-gathered over NCCL
        if by_rows:  # column shards for the transforms, one all-to-all, leaf-row shards for the tree
            pkg.sharding.check_row_plan(cap, world)
            split = pkg.sharding.column_split(W, world)
            c0, nc = split[rank]
            per = N // world
            my_coeffs = torch.empty((nc, n), dtype=torch.int64, device="cuda")
            if args.exchange == "p2p":  # receive buffers every rank can write, registered once (like NCCL's own)
                peers = pkg.sharding.PeerRows(ctx, W * per)
                peer_table = (ffi.vp * world)(*[ffi.vp(p) for p in peers.ptrs])
                tiny = torch.zeros(1, dtype=torch.int32, device="cuda")
            else:
                if args.exchange == "nccl-pack":
                    my_lde = torch.empty((nc, N), dtype=torch.int64, device="cuda")
                send = torch.empty((world, nc, per), dtype=torch.int64, device="cuda")
                rows = torch.empty((W, per), dtype=torch.int64, device="cuda")
            out_split, in_split = [c * per for _, c in split], [nc * per] * world
        else:
            pkg.sharding.check_plan(r, cap, world)
        cap_local = torch.empty(((1 << cap) // world, 4), dtype=torch.int64, device="cuda")
        cap_full = torch.empty((1 << cap, 4), dtype=torch.int64, device="cuda")
        cudart = C.CDLL("libcudart.so.12")

    def rows_commit(src_cols):
        """8(e).3 from this rank's columns (device): transforms (+ exchange), tree, cap all-gather."""
        h = ffi.vp()
        if args.exchange == "p2p":  # the last NTT pass stores into the peers' receive buffers over NVLink
            ctx._check(L.nlc_lde_columns_scatter_dev(ctx._h, ffi.vp(src_cols.data_ptr()), nc, log_n, r, 1,
                                                     ffi.vp(my_coeffs.data_ptr()), None, world, peer_table, c0))
            dist.all_reduce(tiny)  # barrier in stream order: every rank's stores have landed
            rows_ptr = peers.own
        elif args.exchange == "nccl":  # ... into the send layout, then one all-to-all
            ctx._check(L.nlc_lde_columns_scatter_dev(ctx._h, ffi.vp(src_cols.data_ptr()), nc, log_n, r, 1,
                                                     ffi.vp(my_coeffs.data_ptr()), ffi.vp(send.data_ptr()), world, None, 0))
            dist.all_to_all_single(rows.view(-1), send.view(-1), output_split_sizes=out_split,
                                   input_split_sizes=in_split)
            rows_ptr = rows.data_ptr()
        else:  # "nccl-pack": plain LDE, a re-blocking pass, all-to-all (the unfused baseline)
            ctx._check(L.nlc_lde_columns_dev(ctx._h, ffi.vp(src_cols.data_ptr()), nc, log_n, r, 1,
                                             ffi.vp(my_coeffs.data_ptr()), ffi.vp(my_lde.data_ptr())))
            send.copy_(my_lde.view(nc, world, per).permute(1, 0, 2))
            dist.all_to_all_single(rows.view(-1), send.view(-1), output_split_sizes=out_split,
                                   input_split_sizes=in_split)
            rows_ptr = rows.data_ptr()
        ctx._check(L.nlc_commit_from_lde_rows_dev(ctx._h, ffi.vp(rows_ptr), W, log_n, r, cap, rank, world, C.byref(h)))
        cudart.cudaMemcpyAsync(C.c_void_p(cap_local.data_ptr()), C.c_void_p(L.nlc_batch_cap_dev(h)),
                               C.c_size_t(cap_local.numel() * 8), 3, C.c_void_p(stream.cuda_stream))
        dist.all_gather_into_tensor(cap_full, cap_local)
        return h

    def step_resident():
        free_prev()
        h = ffi.vp()
        if by_rows:
            h = rows_commit(dev[c0:c0 + nc])
        elif sharded:
            ctx._check(L.nlc_commit_from_values_shard_dev(ctx._h, ffi.vp(dev.data_ptr()), W, log_n, r, cap, rank, world,
                                                          C.byref(h)))
            # cudaMemcpyAsync(dst, src, bytes, D2D = 3, stream): the shard's cap entries into the all-gather input
            cudart.cudaMemcpyAsync(C.c_void_p(cap_local.data_ptr()), C.c_void_p(L.nlc_batch_cap_dev(h)),
                                   C.c_size_t(cap_local.numel() * 8), 3, C.c_void_p(stream.cuda_stream))
            dist.all_gather_into_tensor(cap_full, cap_local)
        else:
            ctx._check(L.nlc_commit_from_values_dev(ctx._h, ffi.vp(dev.data_ptr()), W, log_n, r, cap, C.byref(h)))
        state["batch"] = h

    def step_e2e():
        free_prev()
        h = ffi.vp()
        if by_rows:  # this rank's columns from pinned host memory, the full cap back on the host
            h = rows_commit(host[c0:c0 + nc].cuda(non_blocking=True))
            state["batch"] = h
            cap_host[:] = cap_full.cpu().numpy().view(np.uint64)
        elif sharded:
            ctx._check(L.nlc_commit_from_values_shard(ctx._h, col_table, W, log_n, r, cap, rank, world, C.byref(h)))
            state["batch"] = h
            local = np.zeros(((1 << cap) // world, 4), np.uint64)
            ctx._check(L.nlc_batch_cap(h, local.ctypes.data_as(ffi.u64p)))
            cap_host[:] = pkg.sharding.gather_cap(local, cap, device=local_rank)
        else:
            ctx._check(L.nlc_commit_from_values(ctx._h, col_table, W, log_n, r, cap, C.byref(h)))
            state["batch"] = h
            ctx._check(L.nlc_batch_cap(h, cap_host.ctypes.data_as(ffi.u64p)))

    for _ in range(max(args.warmup, 3)):
        flush.zero_()
        step_resident()
    torch.cuda.synchronize()

    # ---- timed region: exactly K steps, device time from CUDA events on the launching stream ----
    sampler = ClockSampler(local_rank)
    sampler.start()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    launches0 = ctx.launch_count()
    barrier()
    torch.cuda.synchronize()
    for a, b in evs:
        flush.zero_()  # L2 flush between timed iterations (outside the event pair)
        a.record(stream)
        step_resident()
        b.record(stream)
    torch.cuda.synchronize()
    barrier()
    launches = ctx.launch_count() - launches0
    clocks = sampler.stop()
    ms_total = sum(a.elapsed_time(b) for a, b in evs)
    t = torch.tensor([ms_total], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_per_step = float(t.item()) / args.steps
    value = (1 if sharded else world) * elems / (ms_per_step * 1e-3)

    # ---- per-phase device time (plonky2 TimingTree scopes) for the roofline of the dominant kernel ----
    ctx.timing_enable(True)
    ctx.timing_read()
    for _ in range(3):
        flush.zero_()
        step_resident()
    phases = ctx.timing_read()
    ctx.timing_enable(False)
    k = max(phases["commits"], 1)
    ab = algorithmic_bytes(log_n, W, r, cap)
    peak, peak_src = measured_peaks()
    leaf_ms = phases["leaf hashing"] / k
    ntt_ms = (phases["IFFT"] + phases["FFT + blinding"]) / k
    achieved = ab["leaf_hash"] / (leaf_ms * 1e-3) / 1e9 if leaf_ms > 0 else 0.0
    perms = N * ((W + 7) // 8 if W > 4 else 0)
    roofline = {"bound": "hbm", "kernel": "k_leaf_hash (Poseidon sponge over LDE rows; INT-pipe limited)",
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": ncu_traffic(args.workload, "k_leaf_hash"), "algorithmic_bytes_per_launch": ab["leaf_hash"],
                "ncu_pipes_pct_of_peak": ncu_pipes(args.workload, "k_leaf_hash"),
                "peak_source": peak_src, "ms_per_launch": leaf_ms,
                "poseidon_perms_per_s": perms / (leaf_ms * 1e-3) if leaf_ms > 0 else None}
    # The pipe that actually bounds the hash: the busiest integer pipe of the committed ncu capture.  Its utilisation IS
    # the fraction of the pipe-bound rate; peak = the permutation rate the same instruction stream would reach at 100 %.
    pipes = roofline["ncu_pipes_pct_of_peak"] or {}
    busiest = max(((v, k) for k, v in pipes.items() if k in ("alu", "fmaheavy (IMAD)")), default=None)
    if busiest and roofline["poseidon_perms_per_s"]:
        roofline["int_pipe"] = {"bound": busiest[1], "achieved": roofline["poseidon_perms_per_s"],
                                "peak": roofline["poseidon_perms_per_s"] * 100.0 / busiest[0], "unit": "permutations/s",
                                "frac": busiest[0] / 100.0,
                                "source": "ncu pipe-active cycles of profiles/traffic.json (capture of this kernel, this workload)"}
        files, sha, cap_name = ncu_capture_meta(args.workload, "k_leaf_hash")
        roofline["int_pipe"]["capture"] = cap_name
        # the fraction is a property of the CAPTURED kernel: flag it when the kernel sources have changed since
        roofline["int_pipe"]["stale_capture"] = (sha is None) or (kernel_source_hash(files) != sha)
    roofline_ntt = {"bound": "hbm", "kernel": "k_ntt_* (iFFT + 2^r coset NTTs into leaf rows)",
                    "achieved": ab["ntt"] / (ntt_ms * 1e-3) / 1e9 if ntt_ms > 0 else 0.0, "peak": peak, "unit": "GB/s",
                    "ms": ntt_ms}
    roofline_ntt["frac"] = roofline_ntt["achieved"] / peak
    # the roof that actually binds the transforms: the ALU pipe (64-bit modular add/sub and shift twiddles are carry
    # chains); its utilisation in the committed capture of the two pass kernels, weighted by their durations
    ntt_caps = [(ncu_pipes(args.workload, kk), ncu_capture_meta(args.workload, kk)) for kk in ("k_ntt_hi_f", "k_ntt_lo_f")]
    if all(c[0] for c in ntt_caps) and ntt_ms > 0:
        try:
            with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
                tj = json.load(f)[args.workload]
            wts = [tj[kk]["duration_us"] for kk in ("k_ntt_hi_f", "k_ntt_lo_f")]
            alu = sum(c[0]["alu"] * w for c, w in zip(ntt_caps, wts)) / sum(wts)
            fmah = sum(c[0]["fmaheavy (IMAD)"] * w for c, w in zip(ntt_caps, wts)) / sum(wts)
            bfly = W * (n * log_n // 2) * (1 + (1 << r))
            files, sha, cap_name = ntt_caps[0][1]
            roofline_ntt["ncu_pipes_pct_of_peak"] = {kk: c[0] for kk, c in zip(("k_ntt_hi_f", "k_ntt_lo_f"), ntt_caps)}
            roofline_ntt["traffic"] = {kk: tj[kk]["dram_bytes_per_launch"] for kk in ("k_ntt_hi_f", "k_ntt_lo_f")}
            roofline_ntt["int_pipe"] = {"bound": "alu", "achieved": bfly / (ntt_ms * 1e-3), "peak": bfly / (ntt_ms * 1e-3) * 100.0 / alu,
                                        "unit": "butterflies/s", "frac": alu / 100.0, "fmaheavy_frac": fmah / 100.0,
                                        "capture": cap_name, "stale_capture": (sha is None) or (kernel_source_hash(files) != sha)}
        except Exception:
            pass
    commit_gbs = ab["commit"] / (ms_per_step * 1e-3) / 1e9

    # ---- e2e through the host-buffer C-ABI call (pinned host columns in, cap out) ----
    for _ in range(2):
        step_e2e()
    barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_e2e()
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    te = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_val = (1 if sharded else world) * elems / (float(te.item()) / args.steps)
    free_prev()
    # the same call fed from PAGEABLE host vectors (what a Rust Vec<PolynomialValues> is): csrc/staging.h
    e2e_pageable = None
    if not sharded:
        pageable = host.numpy().view(np.uint64).copy()
        ptab = (ffi.u64p * W)(*[pageable[j].ctypes.data_as(ffi.u64p) for j in range(W)])

        def step_pageable():
            free_prev()
            h = ffi.vp()
            ctx._check(L.nlc_commit_from_values(ctx._h, ptab, W, log_n, r, cap, C.byref(h)))
            state["batch"] = h
            ctx._check(L.nlc_batch_cap(h, cap_host.ctypes.data_as(ffi.u64p)))
        step_pageable()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            step_pageable()
        torch.cuda.synchronize()
        e2e_pageable = (time.perf_counter() - t0) / args.steps * 1e3
        free_prev()

    # what a patched plonky2 that keeps its public struct fields on the host pays: pageable columns in,
    # (i) cap + `polynomials` out (proof.rs:307-310 reads them), (ii) additionally merkle_tree.leaves + digests out --
    # written back WHILE the commit runs (nlc_commit_from_values_materialised) against the same fields fetched after it
    e2e_mat = None
    if not sharded and world == 1:
        coeffs_h = np.zeros((W, n), np.uint64)
        leaves_h = np.zeros((N, W), np.uint64)
        digests_h = np.zeros((2 * (N - (1 << cap)), 4), np.uint64)
        ctab = (ffi.u64p * W)(*[coeffs_h[j].ctypes.data_as(ffi.u64p) for j in range(W)])
        u = lambda a: a.ctypes.data_as(ffi.u64p)

        def timed(fn):
            fn()  # touches every destination page once
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(args.steps):
                fn()
            torch.cuda.synchronize()
            return (time.perf_counter() - t0) / args.steps * 1e3

        def overlapped(full):
            def fn():
                free_prev()
                h = ffi.vp()
                ctx._check(L.nlc_commit_from_values_materialised(ctx._h, ptab, W, log_n, r, cap, ctab,
                                                                 u(leaves_h) if full else None,
                                                                 u(digests_h) if full else None, u(cap_host), C.byref(h)))
                state["batch"] = h
            return fn

        def serial(full):
            def fn():
                free_prev()
                h = ffi.vp()
                ctx._check(L.nlc_commit_from_values(ctx._h, ptab, W, log_n, r, cap, C.byref(h)))
                state["batch"] = h
                ctx._check(L.nlc_batch_cap(h, u(cap_host)))
                ctx._check(L.nlc_batch_coeffs(h, ctab))
                if full:
                    ctx._check(L.nlc_batch_leaves(h, u(leaves_h)))
                    ctx._check(L.nlc_batch_digests(h, u(digests_h)))
            return fn

        b_i = 32 * (1 << cap) + 8 * W * n
        b_ii = b_i + 8 * N * W + digests_h.nbytes
        e2e_mat = {"host_memory": "pageable (in and out)",
                   "polynomials_and_cap": {"ms_per_step": timed(overlapped(False)), "after_the_commit_ms": timed(serial(False)),
                                           "d2h_bytes_per_step": b_i},
                   "polynomials_leaves_digests_cap": {"ms_per_step": timed(overlapped(True)),
                                                      "after_the_commit_ms": timed(serial(True)), "d2h_bytes_per_step": b_ii},
                   "h2d_bytes_per_step": 8 * elems}
        free_prev()
        del coeffs_h, leaves_h, digests_h
        # the same with the struct fields living in page-locked memory (a patched plonky2 that allocates `leaves` /
        # `digests` / `polynomials` with cudaHostAlloc): plain DMA, no host-side copy threads
        pin = lambda shape: torch.zeros(shape, dtype=torch.int64).pin_memory().numpy().view(np.uint64)
        coeffs_h, leaves_h, digests_h = pin((W, n)), pin((N, W)), pin((2 * (N - (1 << cap)), 4))
        ctab = (ffi.u64p * W)(*[coeffs_h[j].ctypes.data_as(ffi.u64p) for j in range(W)])
        e2e_mat["pinned_outputs"] = {"polynomials_and_cap_ms": timed(overlapped(False)),
                                     "polynomials_leaves_digests_cap_ms": timed(overlapped(True)),
                                     "after_the_commit_ms": timed(serial(True))}
        free_prev()
        del coeffs_h, leaves_h, digests_h

    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        O = graft.load_oracle()
        O.set_threads(len(os.sched_getaffinity(0)))
        vals = host.numpy().view(np.uint64)
        O.commit_from_values(vals[:8], r, cap)  # warm tables
        t0 = time.perf_counter()
        ref = O.commit_from_values(vals, r, cap)
        dt = time.perf_counter() - t0
        ok = bool(np.array_equal(ref.cap, cap_host))
        cpu_baseline = {"value": elems / dt, "unit": "elements/s", "cores": O.get_threads(), "kind": "port",
                        "sample": f"1 full commit of {args.workload} ({dt:.2f} s); cap equals the GPU's: {ok}",
                        "port_quality": oracle_quality(O)}

    line = None
    if rank == 0:
        line = {"metric": "LDE+Merkle commit throughput (input trace elements/s)", "value": value,
                "unit": "elements/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
                "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong" if sharded else "weak", "vs_baseline": None,
                "dtype": "u64", "data": "synthetic",
                "config": {"workload": args.workload, "rows": n, "cols": W, "rate_bits": r, "cap_height": cap,
                           "parallelism": (f"one commit, column shards x{world} -> exchange over NVLink -> leaf-row shards + "
                                           f"all-gather of the cap (SURVEY 8e.3; exchange={args.exchange})" if by_rows else
                                           f"one commit, coset shards x{world} + NCCL all-gather of the cap" if sharded
                                           else f"instances x{world}"), "l2": "flushed between timed iterations (256 MiB write)"},
                "clocks": clocks, "gpu_launches": int(launches),
                "e2e": {"value": e2e_val, "unit": "elements/s", "h2d_bytes_per_step": 8 * elems,
                        "d2h_bytes_per_step": 32 << cap, "ms_per_step": float(te.item()) / args.steps * 1e3,
                        "host_memory": "pinned", "pageable_host_ms_per_step": e2e_pageable},
                "e2e_materialised": e2e_mat,
                "roofline": roofline, "roofline_ntt": roofline_ntt,
                "commit_algorithmic_gbs": commit_gbs, "commit_frac_of_hbm": commit_gbs / peak,
                "phases_ms": {kk: (v / k if kk != "commits" else v) for kk, v in phases.items()},
                "cpu_baseline": cpu_baseline}

    # The single-commit blocks come after every number of the contract line is final.  They use CUDA IPC, peer stores and
    # extra collectives; if one of them fails or stalls on a box (no peer access, a rank out of memory), the contract
    # line is still printed -- with the failure in it -- instead of being lost, and the process ends non-zero.
    strong = parity = None
    extra_error = None
    if world > 1 and not sharded and not args.no_strong:
        # a rank that fails says so through a marker file the other ranks' watchers poll: they may be inside a collective
        # the failed rank never joins, and torchrun would otherwise end them before rank 0 has printed anything
        marker = os.path.join(tempfile.gettempdir(), f"nlc_bench_failed_{os.getppid()}_{os.environ.get('MASTER_PORT', '0')}")
        if rank == 0 and os.path.exists(marker):
            os.remove(marker)
        dist.barrier()
        done = threading.Event()

        def watch():
            t0 = time.time()
            while not done.wait(0.5):
                why = None
                if os.path.exists(marker):
                    why = "another rank failed: " + open(marker).read()[:400]
                elif time.time() - t0 > args.strong_timeout:
                    why = f"single-commit blocks did not finish within {args.strong_timeout} s"
                if why is not None:
                    if rank == 0:
                        line["strong"] = {"error": why}
                        line["sharded_parity"] = None
                        print(json.dumps(line), flush=True)
                    os._exit(3)
        threading.Thread(target=watch, daemon=True).start()
        try:
            if args.inject_strong_failure == rank:
                raise RuntimeError("injected failure (--inject-strong-failure)")
            strong = strong_scaling_block(pkg, ctx, stream, rank, world, flush, args.steps)
            ctx.set_stream(None)  # the mirror's sharded commits bring their own stream
            parity = sharded_parity_block(pkg, ctx, rank, world)
            ctx.set_stream(stream.cuda_stream)
        except Exception as e:
            extra_error = repr(e)
            print(f"[rank {rank}] single-commit blocks: {extra_error}", file=sys.stderr, flush=True)
            with open(marker, "w") as f:
                f.write(f"rank {rank}: {extra_error}")
            if rank != 0:
                time.sleep(3.0)  # rank 0's watcher prints the line first
                os._exit(3)
        done.set()

    if rank == 0:
        if extra_error is not None:
            line["strong"] = {"error": extra_error}
            line["sharded_parity"] = parity
        elif strong is not None:
            line["strong"] = strong
            line["sharded_parity"] = parity
        print(json.dumps(line), flush=True)
    if extra_error is not None:
        os._exit(3)  # the other ranks may be inside a collective this rank will never join
    if world > 1:
        dist.destroy_process_group()
    if parity is not None and any(v is False for v in parity.values()):
        raise SystemExit(f"bench.py: sharded parity FAILED: {parity}")
    if strong is not None and not all(v["cap_equals_1gpu_commit_on_every_rank"] for v in strong.values()):
        raise SystemExit("bench.py: a sharded commit's cap differs from the single-GPU commit")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS) + ["replay_plonk", "replay_stark", "stark_proof_fpmul", "stark_proof_ed_add", "stark_proof_ed_scalar_mul", "plonk_proof_demo"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--shard-commit", action="store_true",
                    help="N > 1: split ONE commit by LDE coset over the ranks (strong scaling, N <= 2^rate_bits) instead "
                         "of one independent commit per rank")
    ap.add_argument("--no-strong", action="store_true", help="N > 1: skip the single-commit strong-scaling block")
    ap.add_argument("--inject-strong-failure", type=int, default=-1,
                    help="debug: make this rank fail at the start of the single-commit blocks (checks that the contract line survives)")
    ap.add_argument("--strong-timeout", type=float, default=600.0,
                    help="N > 1: seconds the single-commit blocks may take before the contract line is printed without them")
    ap.add_argument("--exchange", default="p2p", choices=["p2p", "nccl", "nccl-pack"],
                    help="--shard-commit with more ranks than cosets (SURVEY 8e.3): how leaf rows reach their shard")
    ap.add_argument("--plonk-degree-bits", type=int, default=16, help="--workload plonk_proof_demo: log2 rows of the circuit")
    ap.add_argument("--ref-sample-log-rows", type=int, default=None,
                    help="reference arm: opt in to a bounded row sample (log2 rows); default = the workload's own size")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", 0))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    if args.workload.startswith("replay_"):
        if rank == 0:
            run_replay_workload(args)
    elif args.workload in ("stark_proof_fpmul", "stark_proof_ed_add", "stark_proof_ed_scalar_mul"):
        if rank == 0:
            run_stark_proof(args)
    elif args.workload == "plonk_proof_demo":
        if rank == 0:
            run_plonk_proof(args)
    elif args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
