"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel launches, total/avg time, share."""
import csv, sys, collections, re
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 14 and r[0].isdigit()]
tot = collections.defaultdict(float); cnt = collections.Counter()
for r in rows:
    name = re.sub(r"\(.*", "", r[4]); name = name.replace("void ", "").replace("<unnamed>::", "")
    tot[name] += float(r[14].replace(",", "")) / 1e3; cnt[name] += 1
s = sum(tot.values())
print("| kernel | launches | total us | avg us | share |\n|---|---|---|---|---|")
for k, v in sorted(tot.items(), key=lambda kv: -kv[1]):
    print(f"| `{k[:60]}` | {cnt[k]} | {v:.1f} | {v / cnt[k]:.1f} | {100 * v / s:.1f}% |")
