"""Instruction shares of a kernel between its CTA barriers, from an ncu report: python tools/sass_regions.py X.ncu-rep"""
import csv, subprocess, sys, io
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
h = rows[1]
ci, si = h.index('Instructions Executed'), h.index('# Samples')
body = rows[2:]
tot = sum(float(r[ci] or 0) for r in body)
print(rows[0][1][:80], 'total warp instr', tot)
reg = acc = smp = start = 0
hist = {}
for i, r in enumerate(body):
    ins = r[1]
    n, s = float(r[ci] or 0), float(r[si] or 0)
    acc += n; smp += s
    op = (ins.split()[1] if ins.startswith('@') else ins.split()[0]).split('.')[0]
    hist.setdefault(reg, {}); hist[reg][op] = hist[reg].get(op, 0) + n
    if 'BAR.SYNC' in ins or i == len(body) - 1:
        print(f"region {reg}: sass {start}-{i} instr {acc/tot*100:5.1f}% ({acc/1e6:.2f}M) samples {smp:.0f}")
        print("   ", ", ".join(f"{k}:{v/tot*100:.1f}" for k, v in sorted(hist[reg].items(), key=lambda kv: -kv[1])[:14]))
        reg += 1; acc = 0; smp = 0; start = i + 1
