"""Per-kernel SASS evidence: counts of the Blackwell-specific mnemonics in libtfdl_kernels.so
(UTC*MMA = tcgen05.mma, LDTM = tcgen05.ld, UTMALDG/UTMASTG = TMA tensor load/store, UBLKCP = bulk copy,
LDGSTS = cp.async, multimem, RED/ATOM). Run: python tools/sass_summary.py > profiles/sass_summary.md"""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "tensorflow-distributed-learning_b200", "_native", "libtfdl_kernels.so")
out = subprocess.run(["cuobjdump", "-sass", lib], stdout=subprocess.PIPE, text=True).stdout
pats = collections.OrderedDict([
    ("UTC*MMA", r"\bUTC[A-Z]*MMA"), ("LDTM", r"\bLDTM"), ("UTMALDG", r"\bUTMALDG"), ("UTMASTG", r"\bUTMASTG"),
    ("UBLKCP", r"\bUBLKCP"), ("LDGSTS", r"\bLDGSTS"), ("SYNCS", r"\bSYNCS"), ("UTCBAR", r"\bUTCBAR"),
    ("MULTIMEM", r"MULTIMEM|\.MMLT|LDGMC|REDG?\.MC|STG\.MC"), ("RED/ATOM", r"\b(RED|REDG|ATOMG|ATOM)\b"), ("HMMA", r"\bHMMA"),
])
cur, counts = None, collections.OrderedDict()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1)
        counts[cur] = collections.Counter()
        continue
    if cur:
        for k, p in pats.items():
            if re.search(p, line):
                counts[cur][k] += 1
demangle = subprocess.run(["c++filt"], input="\n".join(counts), stdout=subprocess.PIPE, text=True).stdout.splitlines()
print("| kernel | " + " | ".join(pats) + " |")
print("|---|" + "---|" * len(pats))
for (k, c), name in zip(counts.items(), demangle):
    short = re.sub(r"\(.*", "", name.replace("(anonymous namespace)::", "")).replace("tfdl::", "")
    print(f"| {short} | " + " | ".join(str(c.get(p, 0)) for p in pats) + " |")
