"""Per-kernel counts of the SASS mnemonics that show which hardware paths a kernel uses (tcgen05 MMA =
UTCHMMA / UTCQMMA..., TMEM load = LDTM, TMA = UTMALDG / UTMAPF / UBLKCP, legacy tensor core = HMMA).
    python profiles/sass_summary.py > profiles/r2_sass_summary.txt       (needs only cuobjdump, no GPU)"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "mil-nce_howto100m_b200", "libb200head.so")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
PAT = ["UTCHMMA.2CTA", "UTCHMMA", "UTCMMA", "LDTM", "STTM", "UTMALDG", "UTMAPF", "UBLKCP", "UTMASTG", "HMMA", "SYNCS", "ATOMG", "REDG", "RED."]
cur = None
counts = collections.OrderedDict()
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        cur = re.sub(r"\(.*", "", cur)
        counts[cur] = collections.Counter()
        continue
    if cur is None:
        continue
    ins = re.search(r"\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]+)", line)
    if not ins:
        continue
    op = ins.group(1)
    counts[cur]["_total"] += 1
    for p in PAT:
        if op.startswith(p):
            counts[cur][p] += 1
            break
print(f"# SASS summary of {os.path.basename(lib)} (cuobjdump -sass; instruction counts per kernel)")
print("# UTCHMMA = tcgen05.mma (kind::f16 / tf32), .2CTA = cta_group::2; LDTM = tcgen05.ld; UTMALDG = TMA tensor load;")
print("# UBLKCP = cp.async.bulk; UTMAPF = TMA L2 prefetch; HMMA = legacy mma.sync tensor-core instruction")
for k, c in counts.items():
    items = ", ".join(f"{p} {c[p]}" for p in PAT if c[p])
    print(f"{k}\n    instructions {c['_total']}; {items if items else 'no tensor-core / TMA instructions'}")
