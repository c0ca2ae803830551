#!/usr/bin/env python3
"""tools/update_traffic.py CAPTURE.ncu-rep [...] -- rebuild profiles/traffic.json from `ncu --set full` captures.

Per kernel (first launch of each name in the captures): DRAM bytes per launch, pipe-active percentages (ALU, FMA-heavy,
issue slots, DRAM), duration, registers, executed warp instructions, and a hash of the kernel's source files so that
bench.py can flag a capture that no longer describes the tree (`stale_capture`)."""
import csv
import hashlib
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WORKLOAD = "commit_2^16x135_r3_cap4"
SOURCES = {"k_leaf_hash": ["poseidon.cuh", "merkle.cu", "gl.cuh", "poseidon_constants.h"],
           "k_ntt_hi_f": ["ntt.cu", "gl.cuh"], "k_ntt_lo_f": ["ntt.cu", "gl.cuh"]}


def sha(files):
    h = hashlib.sha256()
    for f in files:
        with open(os.path.join(ROOT, "near-light-client_b200", "csrc", f), "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()[:16]


def num(x):
    return float(x.replace(",", ""))


def main():
    out = {"_source": "ncu --set full --clock-control none --import-source on, python bench.py --steps 2 --warmup 1 "
                      "--no-cpu-baseline; dram__bytes_read.sum + dram__bytes_write.sum per launch; pipes: pipe-active "
                      "cycles of the same launch; rebuilt by tools/update_traffic.py", WORKLOAD: {}}
    for rep in sys.argv[1:]:
        raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(raw.splitlines()))
        hdr, units = rows[0], rows[1]
        col = {h: i for i, h in enumerate(hdr)}
        for r in rows[2:]:
            name = r[col["Kernel Name"]]
            key = next((k for k in SOURCES if k in name), None)
            if key is None:
                continue
            g = lambda m: num(r[col[m]])
            dur_now = g("gpu__time_duration.sum") * {"ms": 1e3, "us": 1.0, "ns": 1e-3, "s": 1e6}[units[col["gpu__time_duration.sum"]]]
            if key in out[WORKLOAD] and out[WORKLOAD][key]["duration_us"] >= dur_now:
                continue  # keep the longest launch of each kernel family (the batched LDE passes, not the iFFT's)
            scale = lambda m: {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1}[units[col[m]]]
            rd, wr = g("dram__bytes_read.sum") * scale("dram__bytes_read.sum"), g("dram__bytes_write.sum") * scale("dram__bytes_write.sum")
            dur = g("gpu__time_duration.sum") * {"ms": 1e3, "us": 1.0, "ns": 1e-3, "s": 1e6}[units[col["gpu__time_duration.sum"]]]
            out[WORKLOAD][key] = {
                "kernel": name[:60], "dram_bytes_per_launch": int(rd + wr), "read": int(rd), "write": int(wr),
                "pipes_pct_of_peak": {"alu": g("sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active"),
                                      "fmaheavy (IMAD)": g("sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_active")
                                      if "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_active" in col
                                      else g("sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed"),
                                      "issue_slots": g("smsp__issue_active.avg.pct_of_peak_sustained_active"),
                                      "dram": g("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed")},
                "sm_active_pct_of_elapsed": 100.0 * g("sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_elapsed")
                / max(g("sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active"), 1e-9),
                "duration_us": dur, "registers": int(g("launch__registers_per_thread")),
                "instructions": int(g("smsp__inst_executed.sum")), "source_files": SOURCES[key],
                "source_sha256_16": sha(SOURCES[key]), "capture": "profiles/" + os.path.basename(rep).replace(".ncu-rep", ".txt")}
    with open(os.path.join(ROOT, "profiles", "traffic.json"), "w") as f:
        json.dump(out, f, indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
