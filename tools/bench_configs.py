"""BASELINE.json configs[3] and configs[4] on N GPUs (one rank per GPU under torchrun; pairs / stream chunks sharded, no collective
on the data path -- NCCL is only the timing barrier / max-reduce):
  kitti32  : KITTI-shape 1242x375 pairs (padded to 1280x384), global batch 32 sharded 32/N per rank (inference_kitti.py path),
             through driver.FlowPipeline with the model's own fp32 input AND with uint8 frames in (pre-processing on the GPU)
  video1080: one 1920x1080 uint8 stream cut into N contiguous chunks (+1 shared frame), driver.VideoFlowPipeline (shared pyramid,
             quarter-resolution flows out), run for >= SECONDS seconds: sustained pairs/s (pwc_extract_flow_video.py path)
    torchrun --nproc-per-node N tools/bench_configs.py [mode] [seconds]
Device-side timing (CUDA events), max over ranks; rank 0 prints one JSON line per configuration."""
import json, os, sys, time, warnings
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
warnings.filterwarnings("ignore")
import torch
import torch.distributed as dist
from opticalflow_b200 import driver
from opticalflow_b200.models.PWCNet import PWCDCNet

mode = sys.argv[1] if len(sys.argv) > 1 else "mixed"
seconds = float(sys.argv[2]) if len(sys.argv) > 2 else 3.0
rank, local, world = driver.dist_env()
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
    driver.bind_host_to_gpu(local)


def barrier():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def timed(fn):
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); n = fn(); e1.record(); e1.synchronize()
    t = e0.elapsed_time(e1) * 1e-3
    if world > 1:
        t = driver.max_over_ranks([t], device=dev)[0]
    barrier()
    return n, t


torch.manual_seed(0)
m = PWCDCNet(conv_mode=mode).eval().to(dev)
out = []
# ---------------- configs[3]: KITTI shape, global batch 32 ----------------
GB, H, W = 32, 375, 1242
Bk = GB // world
g = torch.Generator().manual_seed(100 + rank)
u8 = [torch.randint(0, 256, (Bk, 2, H, W, 3), dtype=torch.uint8, generator=g).pin_memory() for _ in range(2)]
pipe8 = driver.FlowPipeline(m, Bk, H, W, dev, input="u8")
f32 = [torch.rand(Bk, 6, pipe8.H, pipe8.W, generator=g).pin_memory() for _ in range(2)]
pipe32 = driver.FlowPipeline(m, Bk, pipe8.H, pipe8.W, dev)
for name, pipe, hosts in (("uint8_frames", pipe8, u8), ("fp32_input", pipe32, f32)):
    pipe.run((hosts[i % 2] for i in range(3)))
    _, t1 = timed(lambda: pipe.run((hosts[i % 2] for i in range(4))))
    iters = max(4, int(seconds / (t1 / 4)) + 1)
    n, t = timed(lambda: pipe.run((hosts[i % 2] for i in range(iters))))
    out.append({"config": "kitti32", "workload": f"1242x375 pairs padded to {pipe8.W}x{pipe8.H}, global batch {GB} = {Bk}/GPU x {world} GPUs, "
                f"conv_mode={mode}, input {name}, FlowPipeline (H2D | forward | D2H overlapped)", "n_gpus": world,
                "pairs_per_s": GB * iters / t, "ms_per_global_batch": 1e3 * t / iters, "seconds": t, "input": name})
del pipe8, pipe32
m.release_plans()
torch.cuda.empty_cache()
# ---------------- configs[4]: 1080p video stream, sustained ----------------
H, W, B, NF = 1080, 1920, 4, 97                      # each rank: a chunk of 97 frames (96 pairs), looped
frames = torch.randint(0, 256, (NF, H, W, 3), dtype=torch.uint8, generator=g).pin_memory()
vp = driver.VideoFlowPipeline(m, B, H, W, dev)
vp.run(frames[:2 * B + 1])
_, t1 = timed(lambda: vp.run(frames))
loops = max(1, int(seconds / t1) + 1)


def stream():
    n = 0
    for _ in range(loops):
        n += vp.run(frames)
    return n
n, t = timed(stream)
out.append({"config": "video1080", "workload": f"uint8 1920x1080 stream (padded to {vp.Wp}x{vp.Hp}), {world} chunks of {NF} frames looped {loops}x, "
            f"batch {B} pairs, shared pyramid, quarter-resolution flows out, conv_mode={mode}", "n_gpus": world,
            "pairs_per_s": n * world / t, "seconds": t, "pairs": n * world})
if rank == 0:
    for o in out:
        print(json.dumps(o), flush=True)
if world > 1:
    dist.destroy_process_group()
