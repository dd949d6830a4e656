"""Label and condense ncu output of tools/prof_forward.py (run ON the GPU box, where the .ncu-rep lives):
  python tools/ncu_summary.py launches <ncu --csv log> <names file> <out.txt>     # gpu__time_duration.sum per launch
  python tools/ncu_summary.py full <raw page csv> <names file> <out.txt> <out.json>  # --set full capture of every launch
The names file is gpurun_out/forward_launch_names.txt (one line per plan op; ops that launch two kernels -- the deterministic
two-pass split-K -- are expanded with a '+reduce' row by matching kernel names)."""
import csv, json, sys


def read_names(path):
    return [l.split()[0] for l in open(path) if l.strip()]


def label(kernels, names):
    """kernels: demangled names in launch order; names: plan ops in launch order; the split-K reduce kernel follows its conv"""
    out, i = [], 0
    for k in kernels:
        if "wm_splitk_reduce_kernel" in k and out:
            out.append(out[-1].split("+")[0] + "+reduce")
            continue
        out.append(names[i] if i < len(names) else "?")
        i += 1
    return out


def short(k):
    k = k.replace("void pwc::", "").replace("pwc::", "")
    return k.split("(")[0][:44]


def launches(log, names_path, out):
    rows = [r for r in csv.reader(l for l in open(log) if not l.startswith("==")) if r]
    h = rows[0]
    ik, iv, im = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Name")
    data = [(r[ik], float(r[iv].replace(",", ""))) for r in rows[1:] if len(r) > iv and r[im] == "gpu__time_duration.sum"]
    unit = rows[1][h.index("Metric Unit")] if len(rows) > 1 else "ns"
    scale = {"ns": 1e-3, "us": 1.0, "usecond": 1.0, "nsecond": 1e-3, "ms": 1e3, "msecond": 1e3}.get(unit, 1e-3)
    labs = label([k for k, _ in data], read_names(names_path))
    tot = sum(v for _, v in data) * scale
    with open(out, "w") as f:
        f.write("# ncu --metrics gpu__time_duration.sum --clock-control none, ONE forward of the mixed plan launch by launch (batch 8, 1024x448);\n")
        f.write("# cold-cache, serialised: compare SHARES with profiles/r2_layer_times_mixed_v*.txt, not absolutes\n")
        f.write(f"# {len(data)} launches, sum {tot:.1f} us\n")
        for (k, v), n in zip(data, labs):
            f.write(f"{n:28s} {short(k):46s} {v * scale:8.1f} us {100 * v * scale / tot:5.1f}%\n")
    print(open(out).read()[:600])


def full(raw, names_path, out_txt, out_json):
    rows = list(csv.reader(open(raw)))
    h = rows[0]
    col = {k: i for i, k in enumerate(h)}
    data = rows[2:]
    labs = label([r[col["Kernel Name"]] for r in data], read_names(names_path))
    want = [("us", "gpu__time_duration.sum", 1e-3), ("tensor%", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", 1),
            ("dram%", "dram__throughput.avg.pct_of_peak_sustained_elapsed", 1), ("l2%", "lts__throughput.avg.pct_of_peak_sustained_elapsed", 1),
            ("rd MB", "dram__bytes_read.sum", None), ("wr MB", "dram__bytes_write.sum", None),
            ("regs", "launch__registers_per_thread", 1), ("grid", "launch__grid_size", 1)]
    units = rows[1]

    def val(r, key, scale):
        if key not in col:
            return float("nan")
        try:
            v = float(r[col[key]].replace(",", ""))
        except ValueError:
            return float("nan")
        if scale is None:                      # bytes -> MB by the unit row
            u = units[col[key]].lower()
            return v * {"byte": 1e-6, "kbyte": 1e-3, "mbyte": 1.0, "gbyte": 1e3}.get(u, 1e-6)
        if key == "gpu__time_duration.sum":
            u = units[col[key]].lower()
            return v * {"ns": 1e-3, "nsecond": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3}.get(u, 1e-3)
        return v * scale
    js = {}
    with open(out_txt, "w") as f:
        f.write("# ncu --set full --clock-control none, ONE mixed-plan forward (batch 8, 1024x448), launch by launch; DRAM MB read / written per launch\n")
        f.write(f"{'layer':24s} {'kernel':40s} " + " ".join(f"{n:>8s}" for n, _, _ in want) + "\n")
        for r, n in zip(data, labs):
            vals = [val(r, k, s) for _, k, s in want]
            f.write(f"{n:24s} {short(r[col['Kernel Name']])[:40]:40s} " + " ".join(f"{v:8.1f}" for v in vals) + "\n")
            js[n] = {"kernel": short(r[col["Kernel Name"]]), "us": vals[0], "read": vals[4] * 1e6, "write": vals[5] * 1e6}
    json.dump(js, open(out_json, "w"), indent=1)
    print(open(out_txt).read()[:800])


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(*sys.argv[2:5])
    else:
        full(*sys.argv[2:6])
