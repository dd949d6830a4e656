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
class FlowPipeline:
    """Streams host batches [B,6,H,W] fp32 (pinned) through `model` with H2D, compute and D2H on three streams.

    Slot s = i % depth owns a device input buffer and a device+pinned output buffer.  Per batch i:
      copy-in stream : wait(in_free[s])  -> H2D x_i             -> record(in_ready[s])
      compute stream : wait(in_ready[s]) -> d2d into the plan's static input, record(in_free[s])
                       graph replay -> copy flow2 into out_dev[s] (after wait(out_free[s])) -> record(out_ready[s])
      copy-out stream: wait(out_ready[s]) -> D2H into out_host[s] -> record(out_done[s]), record(out_free[s])
    so the 22 MB/pair input upload of batch i+1 overlaps the forward of batch i.
    """

    def __init__(self, model, B, H, W, device, depth=2):
        if device.type != "cuda":
            raise RuntimeError("FlowPipeline: CUDA device required -- this build has no CPU fallback")
        self.model, self.B, self.H, self.W, self.device, self.depth = model, B, H, W, device, depth
        with torch.cuda.device(device):
            self.plan = model._plan(B, H, W, device)
            self.plan.pack_weights(model)
            self.s_in, self.s_out = torch.cuda.Stream(device), torch.cuda.Stream(device)
            self.x_dev = [torch.empty(B, 6, H, W, device=device) for _ in range(depth)]
            self.o_dev = [torch.empty(B, 2, H // 4, W // 4, device=device) for _ in range(depth)]
            self.o_host = [torch.empty(B, 2, H // 4, W // 4).pin_memory() for _ in range(depth)]
            ev = lambda: [torch.cuda.Event() for _ in range(depth)]  # noqa: E731
            self.in_ready, self.in_free, self.out_ready, self.out_free, self.out_done = ev(), ev(), ev(), ev(), ev()
            self.plan.run(self.x_dev[0].zero_())          # capture the graph outside the pipeline
            torch.cuda.synchronize(device)
        self.launches_per_batch = self.plan.launches

    @torch.no_grad()
    def run(self, host_batches, sink=None):
        """host_batches: iterable of pinned [B,6,H,W] fp32 tensors.  sink(i, flow_host) is called once batch i's flow2
        ([B,2,H/4,W/4] fp32, pinned, reused after `depth` more batches) is on the host.  Returns the batch count."""
        cs = torch.cuda.current_stream(self.device)
        pending = []                                     # (i, slot) whose D2H is in flight
        n = 0
        with torch.cuda.device(self.device):
            for i, xh in enumerate(host_batches):
                s = i % self.depth
                if xh.shape != self.x_dev[s].shape or xh.dtype != torch.float32:
                    raise RuntimeError(f"FlowPipeline: batch {i} has shape {tuple(xh.shape)} / {xh.dtype}")
                if len(pending) == self.depth:           # slot s is about to be reused: hand its flow to the sink
                    j, sj = pending.pop(0)
                    self.out_done[sj].synchronize()
                    if sink is not None:
                        sink(j, self.o_host[sj])
                with torch.cuda.stream(self.s_in):
                    self.s_in.wait_event(self.in_free[s])
                    self.x_dev[s].copy_(xh, non_blocking=True)
                    self.in_ready[s].record(self.s_in)
                cs.wait_event(self.in_ready[s])
                if hasattr(self.plan, "x_static"):       # graph plan: release the slot right after the d2d copy
                    self.plan.x_static.copy_(self.x_dev[s], non_blocking=True)
                    self.in_free[s].record(cs)
                    flow2 = self.plan.run()
                else:
                    flow2 = self.plan.run(self.x_dev[s])
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
    pipe = FlowPipeline(model, B, H, W, device)
    pipe.run((load_batch(s, c) for s, c in todo),
             sink=lambda i, f: sink(todo[i][0], todo[i][1], f[:todo[i][1]]))
    return hi - lo
