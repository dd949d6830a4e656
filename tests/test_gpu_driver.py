"""The streaming driver returns, for every batch, exactly what a blocking model(x) call returns."""
import pytest
import torch

from oracle import pwcnet_oracle as O

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("mode,inflight", [("bf16", None), ("bf16", 1), ("mixed", None), ("mixed", 3), ("fp32", None)])
def test_pipeline_matches_blocking_forward(mode, inflight):
    """inflight None = the default (two plans on two compute streams for the graph plans, one for fp32); every plan instance
    must return exactly what the blocking call's plan returns."""
    from opticalflow_b200 import driver
    from opticalflow_b200.models.PWCNet import PWCDCNet
    dev = torch.device("cuda", 0)
    torch.manual_seed(1)
    m = PWCDCNet(conv_mode=mode).eval().to(dev)
    B, H, W = 2, 128, 192
    xs = [O.synthetic_input(B, H, W, 10 + i).pin_memory() for i in range(5)]
    want = [m(x.to(dev)).cpu() for x in xs]
    got = {}
    pipe = driver.FlowPipeline(m, B, H, W, dev, inflight=inflight)
    assert pipe.inflight == (inflight if inflight else (1 if mode == "fp32" else 2)) and pipe.depth >= pipe.inflight
    n = pipe.run(iter(xs), sink=lambda i, f: got.__setitem__(i, f.clone()))
    assert n == 5 and sorted(got) == list(range(5))
    for i in range(5):
        assert torch.equal(got[i], want[i]), f"batch {i} differs"
    # a second pass over the same pipeline object (slots and events reused)
    got.clear()
    pipe.run(iter(xs[::-1]), sink=lambda i, f: got.__setitem__(i, f.clone()))
    for i in range(5):
        assert torch.equal(got[i], want[4 - i])


def test_run_sharded_ragged_tail():
    from opticalflow_b200 import driver
    from opticalflow_b200.models.PWCNet import PWCDCNet
    dev = torch.device("cuda", 0)
    torch.manual_seed(2)
    m = PWCDCNet(conv_mode="bf16").eval().to(dev)
    B, H, W, n_pairs = 4, 64, 128, 10
    allx = O.synthetic_input(n_pairs, H, W, 5)
    want = torch.cat([m(allx[s:s + 1].to(dev).expand(1, -1, -1, -1)).cpu() for s in range(n_pairs)])
    bufs = [torch.zeros(B, 6, H, W).pin_memory() for _ in range(3)]

    def load(start, count, _k=[0]):
        b = bufs[_k[0] % 3]; _k[0] += 1
        b.zero_(); b[:count] = allx[start:start + count]
        return b

    out = torch.zeros(n_pairs, 2, H // 4, W // 4)
    seen = []

    def sink(start, count, flow):
        out[start:start + count] = flow
        seen.append((start, count))

    for rank in range(2):                                   # both shards of a world of 2, run one after the other
        driver.run_sharded(m, n_pairs, load, sink, B, H, W, dev, rank=rank, world=2)
    assert sorted(seen) == [(0, 4), (4, 1), (5, 4), (9, 1)]
    # per-pair results do not depend on which batch/rank computed them (tiling may differ with the batch size, so the
    # bar is the bf16-mode EPE bar rather than bit equality)
    epe = (out - want).pow(2).sum(1).sqrt().mean().item()
    print("batch-1 vs batch-4 plan: max abs", (out - want).abs().max().item(), "mean EPE", epe)
    assert epe <= 0.02


@pytest.mark.parametrize("share", [True, False])
@pytest.mark.parametrize("full_res", [False, True])
def test_video_pipeline_matches_reference_steps(full_res, share):
    """uint8 frames -> flows through VideoFlowPipeline == the reference's per-pair steps (frame_to_tensor, pad, cat,
    model, unpad[, flow_resize], x20) with each step done by its own oracle / blocking call."""
    from opticalflow_b200 import driver, prepost
    from opticalflow_b200.models.PWCNet import PWCDCNet
    from oracle import prepost_oracle as PO
    dev = torch.device("cuda", 0)
    torch.manual_seed(3)
    m = PWCDCNet(conv_mode="bf16").eval().to(dev)
    B, H, W, N = 3, 100, 150, 8                              # pads to 128 x 192; 7 pairs = 2 full batches + ragged tail
    g = torch.Generator().manual_seed(9)
    frames = torch.randint(0, 256, (N, H, W, 3), dtype=torch.uint8, generator=g).pin_memory()
    pipe = driver.VideoFlowPipeline(m, B, H, W, dev, swap_rb=True, mult=20.0, full_res=full_res, share_pyramid=share)
    assert pipe.plan.video == share and pipe.plan.NP == (B + 1 if share else 2 * B)
    got = {}
    n = pipe.run(frames, sink=lambda t0, c, f: got.update({t0 + k: f[k].clone() for k in range(c)}))
    assert n == N - 1 and sorted(got) == list(range(N - 1))
    ph, pw = prepost.pad_amounts(H, W)
    m1 = PWCDCNet(conv_mode="bf16").eval().to(dev)
    m1.load_state_dict(m.state_dict())
    for t in range(N - 1):
        x, rph, rpw = PO.frames_to_input(frames[t].numpy(), frames[t + 1].numpy())
        assert (rph, rpw) == (ph, pw)
        # the pipeline runs a batch-B plan: compare with the same pair inside a batch of B (tiling depends on the batch)
        flow = m(x.expand(B, -1, -1, -1).contiguous().to(dev))[:1].cpu()
        ref = PO.unpad(flow, ph, pw) * 20.0
        if full_res:
            ref = PO.flow_resize(ref.clone(), H, W)
        assert got[t].shape == ref.shape[1:]
        # the video plan's stride-2 pyramid layers read one tensor of B + 1 images, the pair plan's two tensors of B: the small
        # levels take different tcgen05 kernels (chunk-major against tap-major fp32 accumulation, tools/dbg_s2_plan.py), so a
        # bf16 rounding of a level-5/6 feature can flip (1 ulp) and move the x20 flow by tenths of a pixel on a 30-pixel flow
        # (pairs 4 and 5 of this clip: mean 0.2 / 0.13 px, the others bit-identical; tools/dbg_s2_video.py).  The bar is 1.5% of
        # the flow range in the mean (full_res scales flows and differences alike); a plumbing error is tens of percent
        d = got[t] - ref[0]
        epe = d.pow(2).sum(0).sqrt().mean().item()
        print("pair", t, "max abs", d.abs().max().item(), "mean EPE", epe, "flow max", ref.abs().max().item())
        assert epe <= 0.015 * max(1.0, ref.abs().max().item()), f"pair {t}"


def test_run_from_reads_the_callers_tensor_in_place():
    """plan.run_from(x): conv1a on the caller's buffer + the rest as a graph == plan.run(x); the release event fires."""
    from opticalflow_b200.models.PWCNet import PWCDCNet
    dev = torch.device("cuda", 0)
    torch.manual_seed(2)
    m = PWCDCNet(conv_mode="bf16").eval().to(dev)
    B, H, W = 2, 64, 128
    plan = m._plan(B, H, W, dev)
    plan.pack_weights(m)
    for seed in (3, 4):
        x = O.synthetic_input(B, H, W, seed).to(dev)
        want = plan.run(x).clone()
        plan.x_static.zero_()                               # prove the static input is not what run_from reads
        ev = torch.cuda.Event()
        got = plan.run_from(x, release=ev).clone()
        ev.synchronize()
        assert torch.equal(got, want)
    with pytest.raises(RuntimeError):
        plan.run_from(x[:, :, :, ::2])


@pytest.mark.parametrize("hw", [(128, 192), (100, 150)])
def test_pipeline_uint8_input_matches_reference_preprocessing(hw):
    """FlowPipeline(input='u8'): uint8 BGR frame pairs in, the flow of model(cat(frame_to_tensor + pad)) out -- the
    pre-processing of pwc_extract_flow_video.py:27-42,197-205 restated by oracle/prepost_oracle.py."""
    from opticalflow_b200 import driver
    from opticalflow_b200.models.PWCNet import PWCDCNet
    from oracle import prepost_oracle as PO
    dev = torch.device("cuda", 0)
    torch.manual_seed(3)
    m = PWCDCNet(conv_mode="bf16").eval().to(dev)
    B, (H, W) = 2, hw
    g = torch.Generator().manual_seed(8)
    batches = [torch.randint(0, 256, (B, 2, H, W, 3), dtype=torch.uint8, generator=g).pin_memory() for _ in range(4)]
    want = []
    for xb in batches:
        xs = [PO.frames_to_input(xb[k, 0].numpy(), xb[k, 1].numpy())[0] for k in range(B)]
        want.append(m(torch.cat(xs, 0).to(dev)).cpu())
    got = {}
    pipe = driver.FlowPipeline(m, B, H, W, dev, input="u8")
    assert pipe.run(iter(batches), sink=lambda i, f: got.__setitem__(i, f.clone())) == 4
    for i in range(4):
        assert got[i].shape == want[i].shape and torch.equal(got[i], want[i]), f"batch {i} differs"
    with pytest.raises(RuntimeError):
        pipe.run(iter([batches[0].float()]))
