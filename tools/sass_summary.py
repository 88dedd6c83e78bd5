"""profiles/sass_summary.txt: per-kernel counts of the SASS mnemonics that prove the Blackwell paths (B200_PROFILING.md):
UTCHMMA (tcgen05.mma), LDTM / STTM (tcgen05.ld / st), UBLKCP (cp.async.bulk), UTCBAR (tcgen05.commit), SYNCS (mbarrier),
USETMAXREG (setmaxnreg), MUFU.   usage: python tools/sass_summary.py [libbfk.so] > profiles/sass_summary.txt"""
import collections
import re
import subprocess
import sys

lib = sys.argv[1] if len(sys.argv) > 1 else "bayesflow_b200/libbfk.so"
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
keys = ["UTCHMMA", "LDTM", "STTM", "UBLKCP", "UTCBAR", "SYNCS", "USETMAXREG", "MUFU", "UTMALDG", "HMMA", "STL", "LDL"]
per = collections.OrderedDict()
cur = None
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        cur = re.sub(r"\(.*", "", cur)
        per[cur] = collections.Counter()
        continue
    if cur is None:
        continue
    m = re.search(r"/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m:
        op = m.group(1)
        per[cur]["_total"] += 1
        for k in keys:
            if op.startswith(k) or (k == "HMMA" and op.startswith("HMMA")):
                per[cur][k] += 1
print(f"# {lib}: SASS mnemonic counts per kernel (cuobjdump -sass, sm_100a); kernels without any tensor / TMEM / bulk-copy "
      "instruction are listed at the end in one line")
print("%-96s %7s " % ("kernel", "instrs") + " ".join("%8s" % k for k in keys))
plain = []
tot = collections.Counter()
for name, c in per.items():
    tot.update(c)
    if not any(c[k] for k in ("UTCHMMA", "LDTM", "STTM", "UBLKCP", "UTCBAR")):
        plain.append(name)
        continue
    print("%-96s %7d " % (name[:96], c["_total"]) + " ".join("%8d" % c[k] for k in keys))
print("%-96s %7d " % ("TOTAL (all kernels)", tot["_total"]) + " ".join("%8d" % tot[k] for k in keys))
print("# kernels without tensor-path instructions (%d): %s" % (len(plain), ", ".join(sorted(set(n.split("<")[0] for n in plain)))))
