"""Batching driver for the PWC-Net forward: shards independent frame pairs over the GPUs of one box (one process per
GPU, no data-path collective) and streams batches through a copy/compute pipeline.

Replaces the per-pair loops of the reference drivers (inference.py:88-91, inference_kitti.py:160-175,
pwc_extract_flow_video.py:75-90: read two frames -> `net(im_all)` -> write flow), which run one pair at a time on
one GPU.  Pairs do not interact, so the only cross-rank step is the timing reduce in bench.py.
"""
import os

import torch


# ------------------------------------------------------------------------------------------------------------------
# Sharding (host logic; covered by world_size-2 gloo tests on CPU)
# ------------------------------------------------------------------------------------------------------------------
def shard_range(n_items, rank, world):
    """Contiguous balanced slice [lo, hi) of n_items owned by `rank`: the first n_items % world ranks get one more."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError(f"bad rank/world {rank}/{world}")
    if n_items < 0:
        raise ValueError("n_items < 0")
    q, r = divmod(n_items, world)
    lo = rank * q + min(rank, r)
    return lo, lo + q + (1 if rank < r else 0)


def batches(lo, hi, batch):
    """[(start, count)] covering [lo, hi) in batches of `batch`; the last one may be ragged."""
    if batch < 1:
        raise ValueError("batch < 1")
    return [(s, min(batch, hi - s)) for s in range(lo, hi, batch)]


def video_shard(n_frames, rank, world):
    """Frames [first, last] (inclusive) of the contiguous chunk of a video stream owned by `rank`: the n_frames - 1
    consecutive pairs are split as shard_range does and each chunk takes ONE extra frame at its end (the frame it shares
    with the next rank), so every pair (t, t+1) is computed by exactly one rank and no frame is exchanged
    (pwc_extract_flow_video.py:process_video walks the pairs in order).  Returns None for an empty chunk."""
    if n_frames < 0:
        raise ValueError("n_frames < 0")
    lo, hi = shard_range(max(n_frames - 1, 0), rank, world)      # pairs [lo, hi): pair t = frames (t, t+1)
    return (lo, hi) if hi > lo else None


def dist_env():
    """(rank, local_rank, world) from the torchrun environment (1-process defaults)."""
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")),
            int(os.environ.get("WORLD_SIZE", "1")))


def max_over_ranks(values, device=None):
    """Element-wise MAX of a list of floats over all ranks (identity without a process group)."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return list(values)
    t = torch.tensor(list(values), dtype=torch.float64, device=device if device is not None else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.tolist()


def sum_over_ranks(values, device=None):
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return list(values)
    t = torch.tensor(list(values), dtype=torch.float64, device=device if device is not None else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return t.tolist()


# ------------------------------------------------------------------------------------------------------------------
# Copy/compute pipeline on one GPU
# ------------------------------------------------------------------------------------------------------------------
def bind_host_to_gpu(device_index):
    """Pin this process (and so its first-touch pinned allocations) to the CPU cores NVML reports as nearest to GPU
    `device_index` (a torch device index: CUDA_VISIBLE_DEVICES is honoured through the device UUID).  With 8 ranks each
    uploading 11 MB per pair, host buffers on the wrong socket halve the aggregate H2D rate.  Returns the core list, or
    None when NVML / the affinity call is unavailable or the intersection with the allowed cores is empty (no change)."""
    try:
        import pynvml
        pynvml.nvmlInit()
        props = torch.cuda.get_device_properties(device_index)
        try:
            h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + str(props.uuid)).encode())
        except Exception:  # noqa: BLE001
            h = pynvml.nvmlDeviceGetHandleByIndex(device_index)
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        near = {64 * i + b for i, w in enumerate(mask) for b in range(64) if (int(w) >> b) & 1}
        allowed = os.sched_getaffinity(0) & near
        if not allowed:
            return None
        os.sched_setaffinity(0, allowed)
        return sorted(allowed)
    except Exception:  # noqa: BLE001
        return None


class FlowPipeline:
    """Streams host batches [B,6,H,W] fp32 (pinned) through `model` with H2D, compute and D2H on three streams.

    Slot s = i % depth owns a device input buffer and a device+pinned output buffer.  Per batch i:
      copy-in stream : wait(in_free[s])  -> H2D x_i             -> record(in_ready[s])
      compute stream : wait(in_ready[s]) -> conv1a straight from the slot (bf16 plan; else a d2d copy into the plan's
                       static input), record(in_free[s]), graph replay of the other launches -> copy flow2 into out_dev[s] (after wait(out_free[s])) -> record(out_ready[s])
      copy-out stream: wait(out_ready[s]) -> D2H into out_host[s] -> record(out_done[s]), record(out_free[s])
    so the 22 MB/pair input upload of batch i+1 overlaps the forward of batch i.

    inflight = 2 (default for the graph plans): batches alternate between TWO plans (own activation buffers and CUDA graph) on two
    compute streams, so the forward of batch i+1 starts while batch i is still in its small pyramid-level 4-6 / decoder-level 4-6
    layers, whose grids leave most SMs idle (a fifth of the step at ~10% of the tensor peak): +5-12% pairs/s on B200
    (tools/two_plans.py).  A batch's latency grows accordingly; inflight = 1 restores one forward at a time.
    """

    def __init__(self, model, B, H, W, device, depth=None, input="f32", swap_rb=True, inflight=None):
        """input='f32': host batches are the model's own input, [B,6,H,W] fp32 (H, W multiples of 64).
        input='u8' : host batches are the frames as the reference drivers hold them before frame_to_tensor /
        pad_to_multiple_of_64 (pwc_extract_flow_video.py:27-42, inference_kitti.py:53-63): uint8 [B,2,H,W,3] (frame 1,
        frame 2; BGR if swap_rb); /255, channel swap, replicate pad and cat run as ONE kernel on the GPU
        (prepost.frames_to_input), so a pair costs 6*H*W bytes of upload instead of 24*Hp*Wp.  Flows come back at a
        quarter of the PADDED size, like the reference model's output before unpad."""
        if device.type != "cuda":
            raise RuntimeError("FlowPipeline: CUDA device required -- this build has no CPU fallback")
        if input not in ("f32", "u8"):
            raise ValueError("FlowPipeline: input must be 'f32' or 'u8'")
        self.model, self.B, self.device = model, B, device
        self.input, self.swap_rb, self.in_h, self.in_w = input, bool(swap_rb), H, W
        if input == "u8":
            from . import prepost
            ph, pw = prepost.pad_amounts(H, W)
            H, W = H + ph, W + pw
        self.H, self.W = H, W
        with torch.cuda.device(device):
            self.plan = model._plan(B, H, W, device)
            self.plan.pack_weights(model)
            if inflight is None:
                inflight = 2 if hasattr(self.plan, "run_from") else 1
            self.inflight = max(1, int(inflight))
            # slots: the host hands batch i over once batch i - depth has left the device, so two forwards only overlap if more
            # than two batches may be queued
            self.depth = depth = max(2, int(depth)) if depth is not None else 2 * self.inflight
            self.inflight = min(self.inflight, depth)
            self.plans = [self.plan] + [model._plan(B, H, W, device, instance=k) for k in range(1, self.inflight)]
            self.s_cmp = [torch.cuda.Stream(device) for _ in range(self.inflight)] if self.inflight > 1 else None
            self.s_in, self.s_out = torch.cuda.Stream(device), torch.cuda.Stream(device)
            self.x_dev = [torch.empty(B, 6, H, W, device=device) for _ in range(depth)]
            self.u8_dev = [torch.empty(B, 2, self.in_h, self.in_w, 3, device=device, dtype=torch.uint8) for _ in range(depth)] \
                if input == "u8" else None
            self.o_dev = [torch.empty(B, 2, H // 4, W // 4, device=device) for _ in range(depth)]
            self.o_host = [torch.empty(B, 2, H // 4, W // 4).pin_memory() for _ in range(depth)]
            ev = lambda: [torch.cuda.Event() for _ in range(depth)]  # noqa: E731
            self.in_ready, self.in_free, self.out_ready, self.out_free, self.out_done = ev(), ev(), ev(), ev(), ev()
            for pl in self.plans:                          # capture the graph(s) outside the pipeline
                pl.pack_weights(model)
                pl.run(self.x_dev[0].zero_())
                if hasattr(pl, "run_from"):
                    pl.run_from(self.x_dev[0])
            torch.cuda.synchronize(device)
        self.launches_per_batch = self.plan.launches

    @torch.no_grad()
    def run(self, host_batches, sink=None, uploaded=None):
        """host_batches: iterable of pinned [B,6,H,W] fp32 tensors.  sink(i, flow_host) is called once batch i's flow2
        ([B,2,H/4,W/4] fp32, pinned) is on the host.  The tensor handed to sink is a pipeline-owned staging buffer that is
        VALID ONLY DURING THE CALL: the next D2H copy into it is enqueued as soon as sink returns -- clone it to keep it.
        uploaded(i, event): optional callback, called right after batch i's H2D copy has been enqueued; `event` (a
        torch.cuda.Event) completes when the copy has finished reading the host batch -- a loader that recycles its
        pinned buffers must wait for it (event.synchronize()) before overwriting batch i.  Returns the batch count."""
        cur = torch.cuda.current_stream(self.device)
        pending = []                                     # (i, slot) whose D2H is in flight
        n = 0
        with torch.cuda.device(self.device):
            for pl in self.plans:
                pl.pack_weights(self.model)              # no-op unless the parameters changed since the last run
            if self.s_cmp is not None:
                for st in self.s_cmp:
                    st.wait_stream(cur)                  # what the caller enqueued so far (and the packing) comes first
            for i, xh in enumerate(host_batches):
                s = i % self.depth
                plan = self.plans[i % self.inflight]
                cs = self.s_cmp[i % self.inflight] if self.s_cmp is not None else cur
                slot = self.u8_dev[s] if self.input == "u8" else self.x_dev[s]
                if xh.shape != slot.shape or xh.dtype != slot.dtype:
                    raise RuntimeError(f"FlowPipeline: batch {i} has shape {tuple(xh.shape)} / {xh.dtype}")
                if len(pending) == self.depth:           # slot s is about to be reused: hand its flow to the sink
                    j, sj = pending.pop(0)
                    self.out_done[sj].synchronize()
                    if sink is not None:
                        sink(j, self.o_host[sj])
                with torch.cuda.stream(self.s_in):
                    self.s_in.wait_event(self.in_free[s])
                    slot.copy_(xh, non_blocking=True)
                    self.in_ready[s].record(self.s_in)
                if uploaded is not None:
                    ev = torch.cuda.Event()
                    ev.record(self.s_in)
                    uploaded(i, ev)
                with torch.cuda.stream(cs):
                    cs.wait_event(self.in_ready[s])
                    if self.input == "u8":               # /255 + swap + pad + cat into the fp32 slot; the u8 slot is free again
                        from . import prepost
                        prepost.frames_to_input(self.u8_dev[s][:, 0], self.u8_dev[s][:, 1], swap_rb=self.swap_rb, out=self.x_dev[s])
                        self.in_free[s].record(cs)
                        flow2 = plan.run_from(self.x_dev[s]) if hasattr(plan, "run_from") else plan.run(self.x_dev[s])
                    elif hasattr(plan, "run_from"):      # bf16 plan: conv1a reads the slot in place, the rest is one graph
                        flow2 = plan.run_from(self.x_dev[s], release=self.in_free[s])
                    elif hasattr(plan, "x_static"):      # graph plan: release the slot right after the d2d copy
                        plan.x_static.copy_(self.x_dev[s], non_blocking=True)
                        self.in_free[s].record(cs)
                        flow2 = plan.run()
                    else:
                        flow2 = plan.run(self.x_dev[s])
                        self.in_free[s].record(cs)
                    cs.wait_event(self.out_free[s])
                    self.o_dev[s].copy_(flow2, non_blocking=True)
                    self.out_ready[s].record(cs)
                with torch.cuda.stream(self.s_out):
                    self.s_out.wait_event(self.out_ready[s])
                    self.o_host[s].copy_(self.o_dev[s], non_blocking=True)
                    self.out_done[s].record(self.s_out)
                    self.out_free[s].record(self.s_out)
                pending.append((i, s))
                n += 1
            for j, sj in pending:
                self.out_done[sj].synchronize()
                if sink is not None:
                    sink(j, self.o_host[sj])
            if self.s_cmp is not None:
                for st in self.s_cmp:
                    cur.wait_stream(st)                  # later work on the caller's stream follows the pipeline's
        return n


def run_sharded(model, n_pairs, load_batch, sink, B, H, W, device, rank=None, world=None):
    """Each rank runs its contiguous slice of `n_pairs` through a FlowPipeline.  load_batch(start, count) returns a
    pinned [B,6,H,W] fp32 tensor (rows >= count are padding); sink(start, count, flow_host) receives the flows of
    pairs [start, start+count).  Returns the number of pairs this rank processed."""
    r, _, w = dist_env()
    rank = r if rank is None else rank
    world = w if world is None else world
    lo, hi = shard_range(n_pairs, rank, world)
    todo = batches(lo, hi, B)
    if not todo:
        return 0
    if world > 1:
        bind_host_to_gpu(device.index if device.index is not None else torch.cuda.current_device())
    pipe = FlowPipeline(model, B, H, W, device)
    pipe.run((load_batch(s, c) for s, c in todo),
             sink=lambda i, f: sink(todo[i][0], todo[i][1], f[:todo[i][1]]))
    return hi - lo


# ------------------------------------------------------------------------------------------------------------------
# Video stream: uint8 frames in, flows out (pwc_extract_flow_video.py:process_video / process_frame_pair :190-215)
# ------------------------------------------------------------------------------------------------------------------
class VideoFlowPipeline:
    """Consecutive frame pairs (t, t+1) of one uint8 video, batched B pairs at a time, everything after the H2D copy of
    the raw frames on the GPU: frame_to_tensor + pad_to_multiple_of_64 + cat (ONE kernel, prepost.frames_to_input, reading
    frames t and t+1 of the same upload through two base pointers), the forward, then unpad (+ optional resize to the
    frame size) and the x`mult` flow scale (ONE kernel, prepost.postprocess_flow).  Per pair 3 bytes/pixel go up instead
    of the 24 bytes/pixel of the fp32 [1,6,H,W] tensor the reference uploads.  Three streams as in FlowPipeline."""

    def __init__(self, model, B, H, W, device, swap_rb=True, mult=20.0, full_res=False, depth=None, share_pyramid=None, inflight=None):
        """inflight (default 2 for the graph plans; depth defaults to 2 * inflight slots): batches alternate between that many plans on
        their own compute streams, as in FlowPipeline.
        share_pyramid (default: on for the 16-bit conv modes): frame t is image 2 of pair t-1 and image 1 of pair t, so the plan
        runs the feature pyramid once per frame (B+1 images per batch instead of 2B) and the cost volumes read frame t+1's
        features in place (engine_bf16.PlanBF16(video=True))."""
        from . import prepost
        if device.type != "cuda":
            raise RuntimeError("VideoFlowPipeline: CUDA device required -- this build has no CPU fallback")
        self.pp = prepost
        self.model, self.B, self.H, self.W, self.device = model, B, H, W, device
        self.swap_rb, self.mult, self.full_res = swap_rb, float(mult), full_res
        self.ph, self.pw = prepost.pad_amounts(H, W)
        Hp, Wp = H + self.ph, W + self.pw
        self.Hp, self.Wp = Hp, Wp
        # the reference cuts the image-pixel pads from the quarter-resolution flow (unpad quirk, see prepost.unpad)
        self.oh, self.ow = (H, W) if full_res else (Hp // 4 - self.ph, Wp // 4 - self.pw)
        with torch.cuda.device(device):
            self.share = (getattr(model, "conv_mode", "") != "fp32") if share_pyramid is None else bool(share_pyramid)
            mk = lambda k: model._plan(B, Hp, Wp, device, video=True, instance=k) if self.share \
                else model._plan(B, Hp, Wp, device, instance=k)      # noqa: E731
            self.plan = mk(0)
            self.plan.pack_weights(model)
            if inflight is None:
                inflight = 2 if hasattr(self.plan, "x_static") and hasattr(self.plan, "run_from") else 1
            self.inflight = max(1, int(inflight))
            self.depth = depth = max(2, int(depth)) if depth is not None else 2 * self.inflight
            self.inflight = min(self.inflight, depth)
            self.plans = [self.plan] + [mk(k) for k in range(1, self.inflight)]
            self.s_cmp = [torch.cuda.Stream(device) for _ in range(self.inflight)] if self.inflight > 1 else None
            self.s_in, self.s_out = torch.cuda.Stream(device), torch.cuda.Stream(device)
            self.f_dev = [torch.empty(B + 1, H, W, 3, device=device, dtype=torch.uint8) for _ in range(depth)]
            self.x_devs = [pl.x_static if hasattr(pl, "x_static") else torch.empty(B, 6, Hp, Wp, device=device) for pl in self.plans]
            self.x_dev = self.x_devs[0]
            self.o_dev = [torch.empty(B, 2, self.oh, self.ow, device=device) for _ in range(depth)]
            self.o_host = [torch.empty(B, 2, self.oh, self.ow).pin_memory() for _ in range(depth)]
            ev = lambda: [torch.cuda.Event() for _ in range(depth)]  # noqa: E731
            self.in_ready, self.in_free, self.out_ready, self.out_free, self.out_done = ev(), ev(), ev(), ev(), ev()
            for pl, xd in zip(self.plans, self.x_devs):
                pl.pack_weights(model)
                pl.run(xd.zero_())
            torch.cuda.synchronize(device)
        self.launches_per_batch = self.plan.launches + 2

    def _compute(self, s, k=0):
        f = self.f_dev[s]
        plan, x_dev = self.plans[k], self.x_devs[k]
        if self.share:
            self.pp.frames_to_stack(f, swap_rb=self.swap_rb, out=x_dev)
        else:
            self.pp.frames_to_input(f[:-1], f[1:], swap_rb=self.swap_rb, out=x_dev)
        flow2 = plan.run() if hasattr(plan, "x_static") else plan.run(x_dev)
        lib, _lib = self.pp._lib.load(), self.pp._lib
        h, w = self.Hp // 4, self.Wp // 4
        e = lib.pwc_flow_postprocess_fwd(_lib.ptr(flow2), self.B, h, w, h - self.ph, w - self.pw, _lib.ptr(self.o_dev[s]),
                                         self.oh, self.ow, self.mult, 0, _lib.stream_handle())
        _lib.check(e, "pwc_flow_postprocess_fwd")

    @torch.no_grad()
    def run(self, frames, sink=None):
        """frames: pinned uint8 tensor [N,H,W,3] (N >= 2).  sink(t0, count, flow_host) receives the flows of pairs
        (t0+k, t0+k+1), k < count, as [count,2,oh,ow] fp32 (a pinned staging buffer, valid only during the sink call: the
        next D2H copy into it is enqueued as soon as sink returns).  `frames` must stay untouched until run() returns.
        Returns N-1."""
        if frames.dtype != torch.uint8 or frames.dim() != 4 or tuple(frames.shape[1:]) != (self.H, self.W, 3):
            raise RuntimeError(f"VideoFlowPipeline: uint8 [N,{self.H},{self.W},3] frames expected, got {tuple(frames.shape)}")
        n_pairs = frames.shape[0] - 1
        todo = batches(0, n_pairs, self.B)
        cur = torch.cuda.current_stream(self.device)
        pending = []

        def deliver(j, sj):
            self.out_done[sj].synchronize()
            if sink is not None:
                sink(todo[j][0], todo[j][1], self.o_host[sj][:todo[j][1]])

        with torch.cuda.device(self.device):
            for pl in self.plans:
                pl.pack_weights(self.model)              # no-op unless the parameters changed since the last run
            if self.s_cmp is not None:
                for st in self.s_cmp:
                    st.wait_stream(cur)
            for i, (t0, cnt) in enumerate(todo):
                s = i % self.depth
                k = i % self.inflight
                cs = self.s_cmp[k] if self.s_cmp is not None else cur
                if len(pending) == self.depth:
                    deliver(*pending.pop(0))
                with torch.cuda.stream(self.s_in):
                    self.s_in.wait_event(self.in_free[s])
                    self.f_dev[s][:cnt + 1].copy_(frames[t0:t0 + cnt + 1], non_blocking=True)
                    if cnt < self.B:                      # ragged tail: repeat the last frame (its flows are dropped)
                        self.f_dev[s][cnt + 1:] = self.f_dev[s][cnt:cnt + 1]
                    self.in_ready[s].record(self.s_in)
                with torch.cuda.stream(cs):
                    cs.wait_event(self.in_ready[s])
                    cs.wait_event(self.out_free[s])
                    self._compute(s, k)
                    self.in_free[s].record(cs)
                    self.out_ready[s].record(cs)
                with torch.cuda.stream(self.s_out):
                    self.s_out.wait_event(self.out_ready[s])
                    self.o_host[s].copy_(self.o_dev[s], non_blocking=True)
                    self.out_done[s].record(self.s_out)
                    self.out_free[s].record(self.s_out)
                pending.append((i, s))
            for j, sj in pending:
                deliver(j, sj)
            if self.s_cmp is not None:
                for st in self.s_cmp:
                    cur.wait_stream(st)
        return n_pairs


def run_video_sharded(model, n_frames, load_frames, sink, B, H, W, device, rank=None, world=None, **kw):
    """Each rank runs its chunk of one video through a VideoFlowPipeline.  load_frames(first, last) returns the pinned uint8
    tensor [last - first + 1, H, W, 3] of frames first..last; sink(t0, count, flow_host) receives the flows of the pairs
    (t0+k, t0+k+1) in stream numbering.  Returns the number of pairs this rank processed."""
    r, _, w = dist_env()
    rank = r if rank is None else rank
    world = w if world is None else world
    chunk = video_shard(n_frames, rank, world)
    if chunk is None:
        return 0
    first, last = chunk
    pipe = VideoFlowPipeline(model, B, H, W, device, **kw)
    return pipe.run(load_frames(first, last), sink=lambda t0, c, f: sink(first + t0, c, f))
