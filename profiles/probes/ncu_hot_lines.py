"""Warp-stall samples per CUDA source line from an ncu report captured with --import-source on.
Usage: python profiles/probes/ncu_hot_lines.py report.ncu-rep [top]"""
import csv, collections, subprocess, sys
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hi = next(i for i, r in enumerate(rows) if "Warp Stall Sampling (All Samples)" in r)
h = rows[hi]; sa = h.index("Warp Stall Sampling (All Samples)"); ie = h.index("Instructions Executed")
def num(x):
    try: return int(x)
    except Exception: return 0
per = collections.OrderedDict(); cur = None; sass = []
for r in rows[hi + 1:]:
    if len(r) <= sa: continue
    if r[0] != "":
        cur = (r[0], r[1].strip()); per.setdefault(cur, [0, 0]); per[cur][0] += num(r[sa]); per[cur][1] += num(r[ie])
    else:
        sass.append((num(r[sa]), cur[0] if cur else "?", r[3].strip()))
tot = sum(v[0] for v in per.values())
print("total samples %d" % tot)
for k, v in sorted(per.items(), key=lambda kv: -kv[1][0])[:top]:
    print("%6d %5.1f%%  inst %9d  L%-4s %s" % (v[0], 100 * v[0] / max(tot, 1), v[1], k[0], k[1][:120]))
print("--- hottest SASS")
for s in sorted(sass, reverse=True)[:16]:
    print("%6d L%-4s %s" % s)
