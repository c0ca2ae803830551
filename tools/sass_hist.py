#!/usr/bin/env python3
"""tools/sass_hist.py -- SASS opcode histograms of the hot kernels (cuobjdump -sass of libnlc_commit.so), written as
markdown: per loop body of k_leaf_hash<true> (the Poseidon full-round and partial-round bodies are the two rolled
loops inside permute_nc), and whole-kernel for the straight-line NTT pass kernels.  With --ncu REP.ncu-rep the stall
samples of the source page are aggregated per opcode class beside them (the hot-spot table)."""
import collections
import csv
import io
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "near-light-client_b200", "libnlc_commit.so")


def opclass(ins):
    ins = re.sub(r"^@!?U?P\d\s+", "", ins.strip())
    op = ins.split()[0].rstrip(";")
    parts = op.split(".")
    key = parts[0]
    if key == "IMAD":
        sub = [p for p in parts[1:] if p in ("WIDE", "IADD", "MOV", "SHL", "X", "HI")]
        key = "IMAD." + ".".join(sub) if sub else "IMAD"
    if key == "IADD3" and "X" in parts:
        key = "IADD3.X"
    return key


PIPE = {"IMAD.WIDE": ("fmaheavy", 4), "IMAD.WIDE.X": ("fmaheavy", 4), "IMAD.HI": ("fmaheavy", 4)}


def pipe_of(k):
    if k in PIPE:
        return PIPE[k]
    if k.startswith("IMAD") or k in ("IMUL",):
        return ("fmaheavy", 2)
    if k in ("IADD3", "IADD3.X", "LOP3", "SHF", "LEA", "SEL", "ISETP", "VIADD", "PRMT", "MOV", "IABS", "PLOP3", "VIMNMX"):
        return ("alu", 2)
    return ("other", 0)


def functions():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    cur, body = None, {}
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            body[cur] = []
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,5})\*/\s+(.*?);", line)
        if m and cur:
            body[cur].append((int(m.group(1), 16), m.group(2)))
    return body


def hist(rows):
    c = collections.Counter(opclass(i) for _, i in rows)
    return c


def table(c, title):
    tot = sum(c.values())
    clk = collections.Counter()
    for k, v in c.items():
        p, w = pipe_of(k)
        clk[p] += v * w
    lines = [f"### {title}", "", f"{tot} instructions; pipe occupancy if issued alone: ALU {clk['alu']} clk, FMA-heavy {clk['fmaheavy']} clk "
             f"(16-lane pipes: 2 clk per warp instruction, IMAD.WIDE 4), issue {tot} clk", "", "| opcode | count | share |", "|---|---|---|"]
    for k, v in c.most_common():
        lines.append(f"| `{k}` | {v} | {100.0 * v / tot:.1f}% |")
    return "\n".join(lines) + "\n"


def loops(rows):
    """backward branches -> (target, branch address) loop bodies, innermost first"""
    out = []
    for a, ins in rows:
        m = re.search(r"BRA\S*\s+(?:!?U?P\d,?\s*)?(0x[0-9a-f]+)", ins)
        if m and int(m.group(1), 16) < a:
            out.append((int(m.group(1), 16), a))
    return sorted(out, key=lambda t: t[1] - t[0])


def main():
    body = functions()
    md = ["# SASS opcode histograms (cuobjdump -sass near-light-client_b200/libnlc_commit.so; tools/sass_hist.py)\n"]
    leaf = next(k for k in body if "k_leaf_hashILb1" in k)
    rows = body[leaf]
    lp = [l for l in loops(rows) if l[1] - l[0] > 0x800]  # the two round bodies (the sponge / unit loops enclose them)
    lp = sorted(lp, key=lambda t: t[1] - t[0])[:2]
    names = ["Poseidon partial round (22 per permutation)", "Poseidon full round (8 per permutation)"]
    per = []
    for (lo, hi), nm in zip(lp, names):
        c = hist([r for r in rows if lo <= r[0] <= hi])
        per.append(sum(c.values()))
        md.append(table(c, f"k_leaf_hash<true>: {nm}, loop body 0x{lo:x}..0x{hi:x}"))
    if len(per) == 2:
        md.append(f"Per permutation: 22 x {per[0]} + 8 x {per[1]} = {22 * per[0] + 8 * per[1]} instructions in the round bodies "
                  f"(ncu: smsp__inst_executed / permutations = 20.1 k including the sponge loads and the loop overhead).\n")
    for pat, nm in (("k_ntt_hi_fILi4ELi4ELb0", "k_ntt_hi_f<4,4,false>: one 8-bit strided pass over a 4096-element tile (256 threads, 16 elements each)"),
                    ("k_ntt_lo_fILi4ELi4ELi2ELb0", "k_ntt_lo_f<4,4,COLS,false>: the last 8-bit pass over a 4096-element tile")):
        k = next(k for k in body if pat in k)
        c = hist(body[k])
        md.append(table(c, nm + f" -- {sum(c.values())} instructions per thread = {sum(c.values()) / 16:.0f} per element"))
    if "--ncu" in sys.argv:
        rep = sys.argv[sys.argv.index("--ncu") + 1]
        raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
        rr = list(csv.reader(io.StringIO(raw)))
        hdr = rr[1]
        ix = {h: i for i, h in enumerate(hdr)}
        agg = collections.defaultdict(collections.Counter)
        keys = ["# Samples", "Instructions Executed", "stall_math", "stall_dispatch", "stall_wait", "stall_not_selected", "stall_selected",
                "stall_short_sb", "stall_long_sb", "stall_no_inst"]
        for r in rr[2:]:
            if len(r) <= ix["Source"] or not r[ix["Source"]].strip():
                continue
            k = opclass(r[ix["Source"]])
            for kk in keys:
                try:
                    agg[k][kk] += float(r[ix[kk]])
                except (ValueError, KeyError):
                    pass
        tot = sum(v["# Samples"] for v in agg.values())
        md.append(f"### ncu source page of k_leaf_hash<true> ({os.path.basename(rep)}): warp-stall samples per opcode class ({int(tot)} samples)\n")
        md.append("| opcode | executed (M warp-instr) | samples | share | math-pipe throttle | dispatch | wait | not selected | selected (issued) | short scoreboard |")
        md.append("|---|---|---|---|---|---|---|---|---|---|")
        for k, v in sorted(agg.items(), key=lambda kv: -kv[1]["# Samples"])[:14]:
            md.append(f"| `{k}` | {v['Instructions Executed'] / 1e6:.1f} | {int(v['# Samples'])} | {100 * v['# Samples'] / tot:.1f}% | {int(v['stall_math'])} | "
                      f"{int(v['stall_dispatch'])} | {int(v['stall_wait'])} | {int(v['stall_not_selected'])} | {int(v['stall_selected'])} | {int(v['stall_short_sb'])} |")
        md.append("")
    print("\n".join(md))


if __name__ == "__main__":
    main()
