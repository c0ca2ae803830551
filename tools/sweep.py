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
            each timed end to end through the host mirror (host columns in, FriProof out).  The CPU-oracle leg and
            the bit-exact proof comparison live in bench.py (--workload replay_plonk / replay_stark), the only
            place besides tests/ that may run oracle/.
  air     : the trace shapes of the AIRs that are actually in the tree (SURVEY.md 8d, config 4 note): ed25519
            2^16 x 1651 / 2502 (machine/ec/builder.rs:301-303), SHA-256 2^17 x 418 / 912 (sha256/air.rs:221-222),
            SHA-512 2^16 x 815 / 1782 (sha512/air.rs:223-224), starky standard_fast_config (rate_bits 1, cap 4):
            resident commit, host-fed commit from pinned columns, and the quotient commit (2 challenges,
            quotient_degree_factor 2: scale by 1/Z_H, coset_ifft, chunks, from_coeffs -- nlc_quotient_commit_dev).
  multi   : configs[4] as written -- the same sweep on 1/2/4/8 GPUs, launched like bench.py
            (python -m torch.distributed.run --nproc-per-node N tools/sweep.py multi): per cell
              instances  -- one independent commit per rank (no data-path collective), whole-job elements/s;
              one_commit -- ONE commit split over the ranks (bench.Strong: by coset while N <= 2^rate_bits, column ->
                            row shards with the exchange fused into the NTT stores beyond), including the cells that do
                            not fit one GPU (2^22 x 512 at rate_bits 3 is 174 GB: SURVEY.md H7);
            device time (CUDA events, L2 flushed), max over ranks; one clock sample over the whole run in the last line.
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


AIR_SHAPES = [("ed25519_2^16x1651", 16, 1651), ("ed25519_2^16x2502", 16, 2502), ("sha256_2^17x418", 17, 418),
              ("sha256_2^17x912", 17, 912), ("sha512_2^16x815", 16, 815), ("sha512_2^16x1782", 16, 1782)]


def run_air(args):
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
    r, cap = 1, 4

    def timed(fn):
        times = []
        for it in range(args.warmup + args.steps):
            flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            h = ffi.vp()
            a.record(stream)
            fn(h)
            b.record(stream)
            torch.cuda.synchronize()
            L.nlc_batch_free(h)
            if it >= args.warmup:
                times.append(a.elapsed_time(b))
        return float(np.median(times))

    for name, log_n, W in AIR_SHAPES:
        n = 1 << log_n
        g = torch.Generator(device="cuda").manual_seed(0xC0FFEE + 4)
        dev = torch.randint(0, 1 << 62, (W, n), dtype=torch.int64, device="cuda", generator=g)
        ms = timed(lambda h: ctx._check(L.nlc_commit_from_values_dev(ctx._h, ffi.vp(dev.data_ptr()), W, log_n, r, cap,
                                                                     C.byref(h))))
        host = dev.cpu().pin_memory()
        table = (ffi.u64p * W)(*[C.cast(host[j].data_ptr(), ffi.u64p) for j in range(W)])
        ms_host = timed(lambda h: ctx._check(L.nlc_commit_from_values(ctx._h, table, W, log_n, r, cap, C.byref(h))))
        # quotient: 2 challenges, factor 2 -> size 2n rows; any rows divide when factor == 2^bits
        ev = torch.randint(0, 1 << 62, (2 * n, 2), dtype=torch.int64, device="cuda", generator=g)
        ms_q = timed(lambda h: ctx._check(L.nlc_quotient_commit_dev(ctx._h, ffi.vp(ev.data_ptr()), 2, log_n, 1, 2, 1, r,
                                                                    cap, C.byref(h))))
        print(json.dumps({"air": name, "rate_bits": r, "cap_height": cap, "commit_resident_ms": ms,
                          "commit_host_pinned_ms": ms_host, "elements_per_s": W * n / ms * 1e3,
                          "e2e_elements_per_s": W * n / ms_host * 1e3, "h2d_bytes": 8 * W * n,
                          "quotient_commit_ms": ms_q}), flush=True)
        del dev, host, ev
        torch.cuda.empty_cache()


def run_multi(args):
    import ctypes as C
    import torch
    import torch.distributed as dist
    import bench
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pkg = graft.load_package()
    P2, ffi, S = pkg.plonky2, pkg._ffi, pkg.sharding
    L = ffi.lib()
    ctx = P2.Context(local)
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    ctx.set_stream(stream.cuda_stream)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    sampler = bench.ClockSampler(local)
    if rank == 0:
        sampler.start()
    cap = 4

    def max_over_ranks(x):
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    widths = [int(w) for w in args.widths.split(",")]
    for r in (3, 1):
        for log_n in range(args.min_log, args.max_log + 1):
            for W in widths:
                n, N = 1 << log_n, 1 << (log_n + r)
                line = {"cell": f"2^{log_n}x{W}_r{r}", "n_gpus": world}
                if cell_bytes(log_n, W, r) <= HBM_BUDGET:
                    g = torch.Generator(device="cuda").manual_seed(0xC0FFEE + 4 + rank)
                    dev = torch.randint(0, 1 << 62, (W, n), dtype=torch.int64, device="cuda", generator=g)
                    times = []
                    for it in range(args.warmup + args.steps):
                        flush.zero_()
                        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                        h = ffi.vp()
                        a.record(stream)
                        ctx._check(L.nlc_commit_from_values_dev(ctx._h, ffi.vp(dev.data_ptr()), W, log_n, r, cap, C.byref(h)))
                        b.record(stream)
                        torch.cuda.synchronize()
                        L.nlc_batch_free(h)
                        if it >= args.warmup:
                            times.append(a.elapsed_time(b))
                    ms1 = max_over_ranks(float(np.median(times)))
                    line["instances"] = {"ms_per_commit_max_over_ranks": ms1, "elements_per_s": world * W * n / ms1 * 1e3}
                    del dev
                    torch.cuda.empty_cache()
                else:
                    ms1 = None
                    line["instances"] = {"skipped": "one commit exceeds one GPU's HBM budget"}
                # one commit over all ranks
                by_rows = world > (1 << r)
                per_rank = (8 * W * n * 2 + 8 * W * N / world * (2 if by_rows else 1) + 64 * N / world + 8 * W * n / world * 3)
                if world == 1:
                    pass
                elif per_rank > HBM_BUDGET:
                    line["one_commit"] = {"skipped": "exceeds the HBM budget per rank"}
                elif by_rows and (W < world or (N // world) < (1 << 12)):
                    line["one_commit"] = {"skipped": "fewer columns than ranks / shard smaller than an NTT tile"}
                else:
                    st = bench.Strong(pkg, ctx, stream, rank, world, (log_n, W, r, cap), flush, device_values=True)
                    ms, phases, _, _, _ = st.measure(args.steps)
                    line["one_commit"] = {"split": "column -> leaf-row shards (8e.3)" if st.by_rows else "coset shards (8e.2)",
                                          "ms_per_commit": ms, "elements_per_s": W * n / ms * 1e3, "phases_ms_max_over_ranks": phases}
                    if ms1:
                        line["one_commit"]["speedup_vs_1gpu"] = ms1 / ms
                    st.peers.free()
                    del st
                    torch.cuda.empty_cache()
                dist.barrier()
                if rank == 0:
                    print(json.dumps(line), flush=True)
    if rank == 0:
        print(json.dumps({"clocks_over_the_whole_run": sampler.stop(), "n_gpus": world}), flush=True)
    dist.destroy_process_group()


def run_replay(args):
    """GPU side of the proof-shape replay (bench.py --workload replay_* adds the CPU oracle leg and the bit-exact check)."""
    import bench
    for kind in ("plonk", "stark"):
        print(json.dumps(bench.replay_gpu(kind, args.steps)), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("mode", choices=["sweep", "replay", "air", "multi"])
    ap.add_argument("--widths", default="64,128,256,512")
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--min-log", type=int, default=14)
    ap.add_argument("--max-log", type=int, default=22)
    args = ap.parse_args()
    {"sweep": run_sweep, "replay": run_replay, "air": run_air, "multi": run_multi}[args.mode](args)


if __name__ == "__main__":
    main()
