"""Static SASS instruction counts per kernel of mjrl_b200/libmjrl_b200.so -> profiles/r02_sass_summary.txt.
Usage: python profiles/probes/sass_summary.py  (needs cuobjdump; runs without a GPU)."""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
LIB = os.path.join(ROOT, "mjrl_b200", "libmjrl_b200.so")
COLS = [("tcgen05.mma", r"\bUTC\w*MMA"), ("tcgen05.ld/st", r"\b(LDTM|STTM)"), ("tcgen05 alloc/commit", r"\b(UTCATOMSWS|UTCBAR)"),
        ("TMA bulk", r"\b(UBLKCP|UTMALDG|UTMASTG|UBLKPF|UBLKRED)"), ("mbarrier", r"\bSYNCS"), ("HMMA", r"\bHMMA"), ("DMMA", r"\bDMMA"),
        ("FFMA2/FMUL2/FADD2", r"\b(FFMA2|FMUL2|FADD2)"), ("FFMA", r"\bFFMA\b"), ("MUFU", r"\bMUFU")]
sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
names = subprocess.run(["c++filt"], input="\n".join(re.findall(r"Function : (\S+)", sass)), capture_output=True, text=True).stdout.split("\n")
counts, order, cur = collections.OrderedDict(), [], None
it = iter(names)
for line in sass.split("\n"):
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = next(it)
        counts[cur] = collections.Counter()
        continue
    if cur is None:
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if not m:
        continue
    op = m.group(1)
    counts[cur]["total"] += 1
    for name, pat in COLS:
        if re.search(pat, op):
            counts[cur][name] += 1
out = ["SASS instruction summary of mjrl_b200/libmjrl_b200.so (cuobjdump -sass, round 2; profiles/probes/sass_summary.py).",
       "Static instructions per kernel; columns: " + ", ".join(n for n, _ in COLS) + ", total.", ""]
tot = collections.Counter()
for k, c in counts.items():
    out.append("%-110s" % k[:110] + "".join("%7d" % c[n] for n, _ in COLS) + "  total %d" % c["total"])
    tot.update(c)
out += ["", "library totals: " + ", ".join("%s %d" % (n, tot[n]) for n, _ in COLS) + ", instructions %d" % tot["total"]]
tck = [k for k, c in counts.items() if c["tcgen05.mma"]]
out += ["kernels issuing tcgen05.mma (%d): " % len(tck)] + ["  " + k[:140] for k in tck]
open(os.path.join(ROOT, "profiles", "r02_sass_summary.txt"), "w").write("\n".join(out) + "\n")
print("\n".join(out[-(len(tck) + 3):]))
