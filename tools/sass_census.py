"""SASS opcode census of libpwc_b200.so per kernel (what proves tcgen05 / TMEM / TMA: UTCHMMA, LDTM, UTMALDG, UTMASTG; legacy
tensor path: HMMA; cp.async: LDGSTS).  python tools/sass_census.py > profiles/rN_sass_census.txt   (no GPU needed)"""
import os, re, subprocess, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from opticalflow_b200 import _lib
out = subprocess.run(["/usr/local/cuda/bin/cuobjdump", "-sass", _lib.LIB_PATH], capture_output=True, text=True).stdout
dem = lambda n: subprocess.run(["/usr/local/cuda/bin/cu++filt", n], capture_output=True, text=True).stdout.strip() or n
KEYS = ["UTCHMMA", "LDTM", "UTMALDG", "UTMASTG", "UBLKCP", "HMMA", "LDGSTS", "LDSM", "FFMA", "ATOM", "RED"]
cur, rows = None, {}
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1); rows[cur] = dict.fromkeys(KEYS, 0); rows[cur]["total"] = 0
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and cur:
        op = m.group(1)
        rows[cur]["total"] += 1
        for k in KEYS:
            if op.startswith(k):
                rows[cur][k] += 1
archs = set(re.findall(r"arch = (sm_\w+)", out))
print(f"# {os.path.basename(_lib.LIB_PATH)}: architectures {sorted(archs)}; {len(rows)} kernels")
print(f"{'kernel':88s} {'total':>7s} " + " ".join(f"{k:>7s}" for k in KEYS))
tot = dict.fromkeys(KEYS, 0)
for fn, r in sorted(rows.items(), key=lambda kv: -kv[1]["total"]):
    name = dem(fn).replace("void ", "").replace("pwc::", "")[:88]
    print(f"{name:88s} {r['total']:7d} " + " ".join(f"{r[k]:7d}" for k in KEYS))
    for k in KEYS:
        tot[k] += r[k]
print(f"{'ALL':88s} {sum(r['total'] for r in rows.values()):7d} " + " ".join(f"{tot[k]:7d}" for k in KEYS))
