#!/usr/bin/env python
"""Benchmark of the B200 PWC-Net inference path (contract: see DESIGN.md "Measurement").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--mode bf16|fp32]

Metric (BASELINE.json): PWC-Net frame-pairs/s at 1024x448, batch 8 per GPU (configs[2]); N>1 ranks each run their own
batch (pairs are independent: weak scaling, no collective on the data path -- NCCL is only used for the timing barrier).
One JSON line is printed by rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time
import warnings

warnings.filterwarnings("ignore")
ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

H_DEF, W_DEF, B_DEF = 448, 1024, 8
GFLOP_PER_PAIR_1024x448 = 168.24          # SURVEY.md Appendix A (2 * 84.12 GMAC)
LEVELS = {6: (196, 7, 16), 5: (128, 14, 32), 4: (96, 28, 64), 3: (64, 56, 128), 2: (32, 112, 256)}


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "bf16_tflops": d["bf16_tflops"],
                "bf16_tflops_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]), "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        super().__init__(daemon=True)
        self.index, self.rows, self.proc = index, [], None

    def run(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                self.rows.append([c.strip() for c in line.split(",")])
        except Exception:  # noqa: BLE001
            pass

    def stop(self):
        if self.proc:
            self.proc.terminate()
        sm = sorted(int(float(r[0])) for r in self.rows if r and r[0].replace(".", "").isdigit())
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            for nm, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        mx = max((int(float(r[1])) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()), default=None)
        pw = [float(r[2]) for r in self.rows if len(r) > 2 and r[2].replace(".", "").isdigit()]
        return {"sm_mhz": (sm[len(sm) // 2] if sm else None), "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm), "power_w_max": (max(pw) if pw else None)}


def _cpu_forward_fn(H, W):
    """The reference's CPU forward on 1 pair: the UNMODIFIED models/PWCNet.py from oracle/_ref/refpkg (kind 'reference',
    SURVEY App. D harness) when oracle/build_ref.sh has run, else the oracle port (kind 'port')."""
    import torch
    from oracle import pwcnet_oracle as O
    x = O.synthetic_input(1, H, W, 1)
    try:
        from oracle import ref_runtime as R
        if R.available():
            m = R.reference_model(0)

            def fwd():
                with R.cpu_execution(), torch.no_grad():
                    return m(x)
            return fwd, "reference"
    except Exception as e:  # noqa: BLE001
        sys.stderr.write(f"bench: reference package unusable ({e}); timing the oracle port\n")
    sd = O.reference_like_state_dict(0)
    return (lambda: O.forward(sd, x)), "port"


def cpu_reference_forward_timer(H, W, reps, threads=None):
    """Times the reference's CPU forward on a bounded sample: `reps` single 1-pair forwards at HxW after 2 warm-ups, with all
    host threads, plus one 1-thread figure.  Returns (pairs/s, cores, sample, kind, pairs/s on 1 thread)."""
    import torch
    cores = threads or os.cpu_count() or 1
    fwd, kind = _cpu_forward_fn(H, W)
    torch.set_num_threads(cores)
    fwd(); fwd()
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter(); fwd(); ts.append(time.perf_counter() - t0)
    ts.sort()
    med = ts[len(ts) // 2]
    torch.set_num_threads(1)
    t0 = time.perf_counter(); fwd(); t1 = time.perf_counter() - t0
    torch.set_num_threads(cores)
    return (1.0 / med, cores, f"{reps} forwards of 1 pair {W}x{H} fp32 (median, after 2 warm-ups), torch CPU {torch.__version__}, "
            f"{cores} threads", kind, 1.0 / t1)


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path on the host cores -- the unmodified
    models/PWCNet.py (oracle/_ref/refpkg, kind 'reference'; its correlation runs as the reference's torch shift-and-dot / C
    because correlation_cuda is CUDA-only), else the oracle port.  Each step = 1 pair, a bounded sample of the batch-8 workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    fwd, kind = _cpu_forward_fn(args.height, args.width)
    t0 = time.perf_counter(); fwd(); t_first = time.perf_counter() - t0
    warm = max(args.warmup, 1)
    budget = 150.0                                           # seconds for the whole run
    steps = max(1, min(max(args.steps, 1), int((budget - warm * t_first) / max(t_first, 1e-3))))
    for _ in range(warm - 1):
        fwd()
    t0 = time.perf_counter()
    for _ in range(steps):
        fwd()
    dt = time.perf_counter() - t0
    v = steps / dt
    line = {"impl": "reference", "metric": "pwcnet_frame_pairs_per_s", "value": v, "unit": "pairs/s", "n_gpus": args.gpus,
            "steps": steps, "warmup": warm, "ms_per_step": 1e3 * dt / steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"PWCDCNet forward {args.width}x{args.height}, random-init weights, 1 pair per step "
                                   f"(bounded sample of the batch-{args.batch} workload)"},
            "cpu_baseline": {"value": v, "unit": "pairs/s", "cores": cores, "kind": kind,
                             "sample": f"{steps} x 1 pair {args.width}x{args.height} fp32, torch CPU {torch.__version__}, {cores} threads"},
            "e2e": {"value": v, "unit": "pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def gpu_reference_baseline(dev, B, H, W, steps=5):
    """The reference itself on the same B200 (SURVEY 8d 'GPU baselines'): unmodified models/PWCNet.py in eager fp32 with its
    own correlation_cuda (oracle/_ref, compiled for sm_100a), and its warp + correlation + LeakyReLU per pyramid level.
    CUDA-event timed.  Returns a dict or {'unavailable': why}."""
    import torch
    import torch.nn.functional as F
    try:
        from oracle import ref_runtime as R
        from oracle import pwcnet_oracle as O
        if not R.cuda_ext_available():
            return {"unavailable": "oracle/_ref not built (oracle/build_ref.sh runs in the build container)"}
        out = {}
        m = R.reference_model(0, dev)
        x = O.synthetic_input(B, H, W, 1).to(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for name, tf32, bench in (("eager_fp32", False, False), ("eager_tf32_cudnn_benchmark", True, True)):
            torch.backends.cudnn.allow_tf32 = tf32
            torch.backends.cudnn.benchmark = bench
            with torch.no_grad():
                for _ in range(3):
                    m(x)
                torch.cuda.synchronize()
                e0.record()
                for _ in range(steps):
                    m(x)
                e1.record(); e1.synchronize()
            out[name] = {"value": B * steps / (e0.elapsed_time(e1) * 1e-3), "unit": "pairs/s", "ms_per_step": e0.elapsed_time(e1) / steps}
        torch.backends.cudnn.allow_tf32 = True
        torch.backends.cudnn.benchmark = False
        _, corr_mod, _ = R.load()
        corr = corr_mod.Correlation(pad_size=4, kernel_size=1, max_displacement=4, stride1=1, stride2=1, corr_multiply=1)
        lv, tot = {}, 0.0
        for lvl, (C, h, w) in LEVELS.items():
            f1, f2, flow = O.corr_microbench_inputs(lvl, B, h, w, C, with_flow=(lvl < 6))
            f1, f2 = f1.to(dev), f2.to(dev)
            fl = None if flow is None else flow.to(dev) * O.FLOW_SCALE[lvl]

            def step():
                with torch.no_grad():
                    w2 = f2 if fl is None else m.warp(f2, fl)
                    return F.leaky_relu(corr(f1, w2), 0.1)
            for _ in range(3):
                step()
            torch.cuda.synchronize()
            e0.record()
            for _ in range(10):
                step()
            e1.record(); e1.synchronize()
            t = e0.elapsed_time(e1) * 1e-3 / 10
            nbytes = 4.0 * B * h * w * (2 * C + 81 + (2 if lvl < 6 else 0))
            lv[f"L{lvl}"] = {"us": round(t * 1e6, 1), "GBps": round(nbytes / t / 1e9, 1)}
            tot += t
        out["warp_corr_lrelu"] = {"levels": lv, "total_us": round(tot * 1e6, 1),
                                  "what": "reference PWCDCNet.warp + correlation_cuda.forward + leaky_relu, eager, fp32 NCHW"}
        out["what"] = "unmodified /root/reference models/PWCNet.py + its correlation_cuda (7-token F8 patch) on this GPU, eager"
        del m
        torch.cuda.empty_cache()
        return out
    except Exception as e:  # noqa: BLE001
        return {"unavailable": f"{type(e).__name__}: {e}"}


def ncu_traffic(layer):
    """DRAM bytes per launch of `layer` from the committed ncu --set full capture (profiles/r2_ncu_traffic.json), or None --
    also None when the kernel's source files changed since the capture (the entry carries their sha256)."""
    import hashlib
    p = os.path.join(ROOT, "profiles", "r2_ncu_traffic.json")
    try:
        e = json.load(open(p)).get(layer, {})
        m = hashlib.sha256()
        for f in e["source_files"]:
            m.update(open(os.path.join(ROOT, f), "rb").read())
        return e.get("bytes") if m.hexdigest()[:16] == e["source_sha"] else None
    except Exception:  # noqa: BLE001
        return None


def corr_microbench(dev, B, iters=20):
    """Fused warp+corr kernel over the 5 pyramid-level shapes of 1024x448 (BASELINE configs[1]); CUDA-event time of
    the kernel alone, L2 flushed between launches.  Algorithmic bytes: 4*B*H*W*(2C+81+(2 if L<6))."""
    import torch
    from opticalflow_b200 import ops
    from oracle import pwcnet_oracle as O
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    res, tot_b, tot_t = {}, 0.0, 0.0
    for lvl, (C, H, W) in LEVELS.items():
        f1, f2, flow = O.corr_microbench_inputs(lvl, B, H, W, C, with_flow=(lvl < 6))
        nhwc = lambda t: t.permute(0, 2, 3, 1).contiguous().to(dev)  # noqa: E731
        a, b = nhwc(f1), nhwc(f2)
        fl = None if flow is None else nhwc(flow)
        out = torch.empty(B, H, W, 81, device=dev)
        scale = O.FLOW_SCALE.get(lvl, 1.0)
        for _ in range(3):
            ops.warp_corr_lrelu(a, b, fl, scale, layout="nhwc", out=out)
        ts = []
        for _ in range(iters):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); ops.warp_corr_lrelu(a, b, fl, scale, layout="nhwc", out=out); e1.record()
            e1.synchronize(); ts.append(e0.elapsed_time(e1) * 1e-3)
        t = sum(ts) / len(ts)
        nbytes = 4.0 * B * H * W * (2 * C + 81 + (2 if lvl < 6 else 0))
        res[f"L{lvl}"] = {"us": round(t * 1e6, 2), "GBps": round(nbytes / t / 1e9, 1)}
        tot_b += nbytes; tot_t += t
    return res, tot_b, tot_t


class InFlight:
    """K device-resident steps (one batch each) over `n` plans of the same model on `n` streams, steps dealt round-robin:
    with n = 2 the next forward starts while the previous one is in its small level-4..6 layers (driver.FlowPipeline does the
    same for host batches).  n = 1 is one CUDA-graph replay after another on one stream."""

    def __init__(self, model, B, H, W, dev, x, n):
        import torch
        self.torch, self.dev = torch, dev
        self.plans = [model._plan(B, H, W, dev, instance=k) for k in range(n)]
        for p in self.plans:
            p.pack_weights(model)
            if hasattr(p, "x_static"):
                p.x_static.copy_(x)
        self.x = x
        self.streams = [torch.cuda.Stream(dev) for _ in range(n)] if n > 1 else [None]
        self.e0, self.e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def _one(self, p):
        return p.run() if hasattr(p, "x_static") else p.run(self.x)

    def warm(self, k):
        for p in self.plans:
            for _ in range(k):
                self._one(p)
        self.torch.cuda.synchronize()

    def timed(self, K):
        torch = self.torch
        cur = torch.cuda.current_stream(self.dev)
        self.e0.record()
        if self.streams[0] is None:
            for _ in range(K):
                self._one(self.plans[0])
        else:
            for st in self.streams:
                st.wait_event(self.e0)
            for k in range(K):
                i = k % len(self.plans)
                with torch.cuda.stream(self.streams[i]):
                    self._one(self.plans[i])
            for st in self.streams:
                cur.wait_stream(st)
        self.e1.record(); self.e1.synchronize()
        return self.e0.elapsed_time(self.e1) * 1e-3


def run_ours(args):
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback (use --impl reference)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    from opticalflow_b200 import _lib
    from opticalflow_b200.models.PWCNet import PWCDCNet
    from oracle import pwcnet_oracle as O   # synthetic-input generator + cpu_baseline only
    _lib.load()
    peaks = load_peaks()
    B, H, W = args.batch, args.height, args.width
    torch.manual_seed(0)
    model = PWCDCNet(conv_mode=args.mode).eval().to(dev)
    x_host = O.synthetic_input(B, H, W, 1 + rank).pin_memory()
    x = x_host.to(dev)
    plan = model._plan(B, H, W, dev)
    plan.pack_weights(model)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # device-resident leg: the batch sits in the plan's own input buffer (conv1a reads it in place), no per-step copy
    n_fl = max(1, args.inflight) if hasattr(plan, "run_from") else 1
    fl = InFlight(model, B, H, W, dev, x, n_fl)
    fl1 = fl if n_fl == 1 else InFlight(model, B, H, W, dev, x, 1)      # one forward at a time, reported beside it
    fl.warm(max(args.warmup, 3))
    launches_per_step = plan.launches
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start(); time.sleep(0.3)
    # ---- device-resident throughput ----
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_dev = fl.timed(args.steps)
    barrier()
    t_dev1 = fl1.timed(args.steps) if fl1 is not fl else t_dev
    barrier()
    # ---- end to end through the public API: pinned host batches -> FlowPipeline (H2D | forward | D2H on three
    #      streams, the package's streaming driver) -> flow on the host; every step uploads its own input ----
    from opticalflow_b200 import driver
    pipe = driver.FlowPipeline(model, B, H, W, dev, inflight=n_fl)
    x_hosts = [x_host, O.synthetic_input(B, H, W, 101 + rank).pin_memory()]
    got = []
    pipe.run((x_hosts[i % 2] for i in range(3)))
    barrier()
    e0.record()
    pipe.run((x_hosts[i % 2] for i in range(args.steps)), sink=lambda i, f: got.append(float(f[0, 0, 0, 0])))
    e1.record(); e1.synchronize()
    t_e2e = e0.elapsed_time(e1) * 1e-3
    assert len(got) == args.steps
    barrier()
    # the same pairs handed over as uint8 BGR frames (what the reference drivers hold before frame_to_tensor): /255 + pad +
    # cat run on the GPU, 6 bytes per pixel pair go over PCIe instead of 24 -- reported beside the fp32-input figure
    t_u8 = None
    if args.mode != "fp32":
        gen = torch.Generator().manual_seed(7 + rank)
        u8_hosts = [torch.randint(0, 256, (B, 2, H, W, 3), dtype=torch.uint8, generator=gen).pin_memory() for _ in range(2)]
        pipe8 = driver.FlowPipeline(model, B, H, W, dev, input="u8", inflight=n_fl)
        got8 = []
        pipe8.run((u8_hosts[i % 2] for i in range(3)))
        barrier()
        e0.record()
        pipe8.run((u8_hosts[i % 2] for i in range(args.steps)), sink=lambda i, f: got8.append(float(f[0, 0, 0, 0])))
        e1.record(); e1.synchronize()
        t_u8 = e0.elapsed_time(e1) * 1e-3
        assert len(got8) == args.steps
        del pipe8
        barrier()
    # unpipelined variant (one blocking model(x) call per step), reported beside it
    out_host = torch.empty(B, 2, H // 4, W // 4).pin_memory()
    for _ in range(2):
        out_host.copy_(model(x_host.to(dev, non_blocking=True)), non_blocking=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        out_host.copy_(model(x_host.to(dev, non_blocking=True)), non_blocking=True)
    e1.record(); e1.synchronize()
    t_sync = e0.elapsed_time(e1) * 1e-3
    barrier()
    clocks = sampler.stop() if sampler else None
    # ---- sustained leg (BASELINE configs[4] "sustained pairs/s"): the same graph back to back for >= args.sustain seconds,
    #      then the uint8-input pipeline for as long, clocks and power sampled throughout ----
    sus = None
    if args.sustain > 0:
        n_s = max(args.steps, int(args.sustain / (t_dev / args.steps)) + 1)
        sampler2 = ClockSampler(local) if rank == 0 else None
        if sampler2:
            sampler2.start(); time.sleep(0.2)
        barrier()
        t_sus = fl.timed(n_s)
        barrier()
        n_p = max(args.steps, int(args.sustain / (t_e2e / args.steps)) + 1)
        pipe_s = pipe
        hosts_s = x_hosts
        if args.mode != "fp32":
            pipe_s = driver.FlowPipeline(model, B, H, W, dev, input="u8", inflight=n_fl)
            hosts_s = u8_hosts
            pipe_s.run((hosts_s[i % 2] for i in range(3)))
        barrier()
        e0.record()
        pipe_s.run((hosts_s[i % 2] for i in range(n_p)))
        e1.record(); e1.synchronize()
        t_sus_e2e = e0.elapsed_time(e1) * 1e-3
        barrier()
        ck2 = sampler2.stop() if sampler2 else None
        if world > 1:
            t_sus, t_sus_e2e = driver.max_over_ranks([t_sus, t_sus_e2e], device=dev)
        sus = {"value": B * world * n_s / t_sus, "e2e": B * world * n_p / t_sus_e2e, "unit": "pairs/s", "seconds": [t_sus, t_sus_e2e],
               "steps": [n_s, n_p], "e2e_input": "uint8 frames (FlowPipeline input='u8')" if args.mode != "fp32" else "fp32 [B,6,H,W]",
               "sm_mhz_median": ck2 and ck2["sm_mhz"], "power_w_max": ck2 and ck2["power_w_max"],
               "reasons": ck2 and ck2["reasons"], "vs_burst": (B * world * n_s / t_sus) / (B * world * args.steps / t_dev)}
    if world > 1:
        t_dev, t_e2e, t_sync, t_u8x, t_dev1 = driver.max_over_ranks([t_dev, t_e2e, t_sync, t_u8 if t_u8 is not None else 0.0, t_dev1],
                                                                    device=dev)
        t_u8 = t_u8x if t_u8 is not None else None
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    pairs = B * world * args.steps
    value = pairs / t_dev
    gflop_pair = GFLOP_PER_PAIR_1024x448 * (H * W) / (448.0 * 1024.0)
    step_tflops = value * gflop_pair / 1e3 / world
    # ---- per-kernel roofline, measured live: CUDA events around every launch of one forward (same stream) ----
    roof, roof_corr, kern_table = None, None, None
    if hasattr(plan, "time_ops"):
        ops_t = plan.time_ops(reps=3)
        is_tc = lambda n: not (n.startswith("warp_corr") or n.startswith("deconv") or n.startswith("conv1a.") or n == "unpack_flow" or n.startswith("c1_copy"))  # noqa: E731
        tc = [(n, t, f) for (n, t, f) in ops_t if is_tc(n)]
        tc_t, tc_f = sum(t for _, t, _ in tc), sum(f for _, _, f in tc)
        all_t = sum(t for _, t, _ in ops_t)
        top = max(tc, key=lambda r: r[1])
        peak = peaks["bf16_tflops"]
        roof = {"bound": "tensor", "achieved": tc_f / tc_t / 1e12, "peak": peak, "unit": "TFLOP/s",
                "frac": tc_f / tc_t / 1e12 / peak, "traffic": ncu_traffic(top[0]),
                "traffic_note": f"DRAM bytes of ONE launch of the top layer ({top[0]}), ncu --set full (profiles/r2_ncu_traffic.json)",
                "kernel": "conv_tc_wm_kernel / conv_tc_halo_kernel / conv_tc_kernel (tcgen05 3x3 convolutions) + "
                          "conv_mma_small_kernel (level-6 layers)",
                "launches_per_step": len(tc), "share_of_step": tc_t / all_t,
                # frac is against the pool's measured cuBLAS bf16 figure (MEASURED_PEAKS.json, taken on another box at its
                # power-capped clocks): a launch timed alone on a box holding higher clocks can exceed it, so the nominal dense
                # bf16 peak (2250 TFLOP/s, 148 SMs x 8192 FLOP/cycle x 1.86 GHz) is quoted beside it
                "top_launch": {"layer": top[0], "us": round(top[1] * 1e6, 1), "TFLOPs": round(top[2] / top[1] / 1e12, 1),
                               "frac": top[2] / top[1] / 1e12 / peak, "frac_of_nominal_2250": top[2] / top[1] / 1e12 / 2250.0},
                "note": f"algorithmic conv FLOPs of the {len(tc)} tensor-core launches of one step / their summed CUDA-event "
                        f"durations (each launch captured 8x into a CUDA graph and timed alone); peak = bf16 burst ({peaks['source']}); whole-step figure "
                        f"{step_tflops:.1f} TFLOP/s incl. all other kernels"}
        kern_table = {n: round(t * 1e6, 1) for n, t, _ in ops_t}
        # FLOP-weighted fractions per layer group (north_star: "the decoder convs at 50% or more of tensor-pipe peak")
        import re as _re

        def grp(pat):
            sel = [(t, f) for n, t, f in ops_t if _re.match(pat, n)]
            tt, ff = sum(t for t, _ in sel), sum(f for _, f in sel)
            return {"us": round(tt * 1e6, 1), "TFLOPs": round(ff / tt / 1e12, 1), "frac": ff / tt / 1e12 / peak} if tt > 0 else None
        roof["groups"] = {"decoder_conv2-6_0-4": grp(r"conv[2-6]_[0-4]\."), "decoder_level2": grp(r"conv2_[0-4]\."),
                          "decoder_level3": grp(r"conv3_[0-4]\."), "decoder_levels4-6": grp(r"conv[4-6]_[0-4]\."),
                          "context_dc_conv1-6": grp(r"dc_conv[1-6]\."), "pyramid": grp(r"conv[1-6](a|aa|b)\."),
                          "note": "algorithmic FLOPs / summed launch times; split-precision layers issue 2-3x the tensor work per algorithmic FLOP"}
    corr_res, corr_bytes, corr_t = corr_microbench(dev, B)
    roof_corr_f32 = {"bound": "hbm", "achieved": corr_bytes / corr_t / 1e9, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                     "frac": corr_bytes / corr_t / 1e9 / peaks["hbm_gbs"], "traffic": None, "levels": corr_res,
                     "kernel": "warp_corr_nhwc_kernel (fp32 features, CUDA cores)",
                     "note": f"fused warp+corr kernel alone, 5 pyramid levels, batch {B}, fp32 NHWC, L2 flushed between "
                             f"launches; algorithmic bytes 4*B*H*W*(2C+81+2); peak {peaks['source']}"}
    roof_corr = roof_corr_f32
    if hasattr(plan, "time_ops"):
        # the cost-volume launches of the timed forward itself: bf16 features inside the concat buffers
        lv, cb, ct = {}, 0.0, 0.0
        for n_, t_, _ in ops_t:
            if n_.startswith("warp_corr"):
                l = int(n_[-1]); C_, H_, W_ = LEVELS[l][0], H >> l, W >> l
                by = float(B) * H_ * W_ * (2 * C_ * 2 + 81 * 2 + (8 if l < 6 else 0))
                lv[f"L{l}"] = {"us": round(t_ * 1e6, 2), "GBps": round(by / t_ / 1e9, 1)}
                cb += by; ct += t_
        roof_corr = {"bound": "hbm", "achieved": cb / ct / 1e9, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                     "frac": cb / ct / 1e9 / peaks["hbm_gbs"], "traffic": ncu_traffic("warp_corr2"),
                     "traffic_note": "DRAM bytes of the level-2 launch alone (ncu --set full, profiles/r2_ncu_traffic.json)", "levels": lv,
                     "kernel": "warp_corr_tc_kernel (levels 2,3: tcgen05) / warp_corr_mma_kernel (levels 4-6: mma.sync)",
                     "note": "the 5 cost-volume launches of one forward (bf16 features in the concat buffers); each launch captured 8x into "
                             "a CUDA graph and timed with CUDA events (no host launch cost); algorithmic bytes B*H*W*(2*C*2 + 81*2 + 8)"}
    if roof is None:
        roof = {"bound": "tensor", "achieved": step_tflops, "peak": peaks["bf16_tflops_sustained"], "unit": "TFLOP/s",
                "frac": step_tflops / peaks["bf16_tflops_sustained"], "traffic": None,
                "note": "whole forward (fp32 CUDA-core mode): conv FLOPs / step time"}
    # host-core baseline: rank 0 at N = 1 only (under torchrun the other ranks' processes share the cores)
    cpu_v, cpu_cores, cpu_sample, cpu_kind, cpu_v1 = cpu_reference_forward_timer(H, W, reps=5) if world == 1 else \
        (None, None, "not measured at N > 1", "reference", None)
    gpu_base = gpu_reference_baseline(dev, B, H, W) if world == 1 else None
    parity = None
    if world == 1:
        # parity of exactly what was benched: this mode, this batch, this shape, against the CPU oracle on the first pairs
        npar = min(B, 2)
        got_f = model(x).float().cpu()[:npar]
        ref_f = O.forward({k: v.detach().cpu() for k, v in model.state_dict().items()}, x_host[:npar])
        d_ = (got_f - ref_f)
        epe = d_.pow(2).sum(1).sqrt().mean().item()
        parity = {"vs": "oracle/pwcnet_oracle.forward (CPU fp32, pinned to the unmodified reference)", "pairs_checked": npar,
                  "epe_delta_px_over_20": epe, "epe_delta_px": 20.0 * epe, "max_abs_px": 20.0 * d_.abs().max().item(),
                  "flow_max_px": 20.0 * ref_f.abs().max().item(),
                  "bar": "fp32: max-abs <= 1e-3 px; 16-bit conv modes: mean EPE delta <= 0.02 px on the x20 flow (script_pwc.py:72)"}
    other_modes = None
    if world == 1 and not args.no_other_modes:
        # the other conv modes on the same input, device-resident leg only (same protocol: CUDA-graph replay, CUDA events) + parity
        other_modes = {}
        ref_f = O.forward({k: v.detach().cpu() for k, v in model.state_dict().items()}, x_host[:min(B, 2)])
        for om in ("f16", "bf16", "fp32"):
            if om == args.mode:
                continue
            try:
                mo = PWCDCNet(conv_mode=om).eval().to(dev)
                mo.load_state_dict(model.state_dict())
                po = mo._plan(B, H, W, dev)
                po.pack_weights(mo)
                flo = InFlight(mo, B, H, W, dev, x, n_fl if hasattr(po, "run_from") else 1)
                nst = args.steps if om != "fp32" else max(3, args.steps // 5)
                flo.warm(3)
                to = flo.timed(nst)
                go = flo._one(po).float().cpu()[:min(B, 2)]
                dd = go - ref_f
                other_modes[om] = {"value": B * nst / to, "unit": "pairs/s", "ms_per_step": 1e3 * to / nst, "steps": nst,
                                   "plans_in_flight": len(flo.plans),
                                   "epe_delta_px": 20.0 * dd.pow(2).sum(1).sqrt().mean().item(),
                                   "max_abs_px": 20.0 * dd.abs().max().item()}
                del mo, po, flo
                torch.cuda.empty_cache()
            except Exception as ex:  # noqa: BLE001
                other_modes[om] = {"error": f"{type(ex).__name__}: {ex}"}
    if sus is not None and roof is not None:
        sus["step_tflops"] = sus["value"] * gflop_pair / 1e3 / world
        sus["frac_of_sustained_peak"] = sus["step_tflops"] / peaks["bf16_tflops_sustained"]
    line = {
        "metric": "pwcnet_frame_pairs_per_s", "value": value, "unit": "pairs/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": 1e3 * t_dev / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": {"bf16": "bf16", "f16": "f16", "mixed": "f16", "fp32": "f32"}.get(args.mode, args.mode), "data": "synthetic",
        "config": {"workload": f"PWCDCNet forward, batch {B}/GPU, {W}x{H}, random-init weights, conv_mode={args.mode}" +
                               (" (IEEE-half storage; feature pyramid, level 5/4 decoders and up_flow/up_feat in split precision: "
                                "two half planes per tensor, three tensor-core products per K step)" if args.mode == "mixed" else ""),
                   "l2": "activation working set per step >> 126 MB L2 (no flush needed)", "parallelism": f"pairs sharded x{world}, no collective",
                   "plans_in_flight": n_fl,
                   "plans_in_flight_note": "steps alternate between two plans (own buffers + CUDA graph) on two streams, in the device-"
                                           "resident leg and in FlowPipeline alike; value_one_forward_at_a_time is the single-plan figure"},
        "value_one_forward_at_a_time": pairs / t_dev1,
        "e2e": {"value": pairs / t_e2e, "unit": "pairs/s", "h2d_bytes_per_step": x_host.numel() * 4 * world,
                "d2h_bytes_per_step": out_host.numel() * 4 * world,
                "api": "opticalflow_b200.driver.FlowPipeline.run (copy/compute overlapped)",
                "blocking_model_call_value": pairs / t_sync,
                "uint8_frames_input": None if t_u8 is None else {
                    "value": pairs / t_u8, "unit": "pairs/s", "h2d_bytes_per_step": 6 * B * H * W * world,
                    "api": "FlowPipeline(input='u8'): uint8 BGR frame pairs in, pre-processing kernel + forward on the GPU"}},
        "gpu_launches": launches_per_step * args.steps * world,
        "roofline": roof, "roofline_corr": roof_corr, "roofline_corr_fp32": roof_corr_f32,
        "step_tflops": step_tflops,
        "cpu_baseline": {"value": cpu_v, "unit": "pairs/s", "cores": cpu_cores, "kind": cpu_kind, "sample": cpu_sample,
                         "one_thread_value": cpu_v1},
        "gpu_baseline": gpu_base,
        "clocks": clocks, "sustained": sus, "parity": parity, "other_modes": other_modes,
    }
    if kern_table and os.environ.get("PWC_BENCH_KERNELS"):
        line["kernel_us"] = kern_table
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--mode", default=os.environ.get("PWC_BENCH_MODE", "mixed"), choices=["fp32", "bf16", "f16", "mixed"])
    ap.add_argument("--sustain", type=float, default=float(os.environ.get("PWC_BENCH_SUSTAIN", "3")),
                    help="seconds of the back-to-back sustained leg (0 = off)")
    ap.add_argument("--no-other-modes", action="store_true", help="skip the short device-resident legs of the other conv modes")
    ap.add_argument("--inflight", type=int, default=int(os.environ.get("PWC_BENCH_INFLIGHT", "2")),
                    help="plans (forwards) kept in flight on their own streams, steps dealt round-robin; 1 = one forward at a time")
    ap.add_argument("--batch", type=int, default=B_DEF)
    ap.add_argument("--height", type=int, default=H_DEF)
    ap.add_argument("--width", type=int, default=W_DEF)
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
