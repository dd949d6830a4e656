"""Host logic of the sharded driver: pairs are split over ranks with no data-path collective.
CPU only: world_size-2 gloo process groups stand in for the one-process-per-GPU launch."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from opticalflow_b200 import driver


def test_shard_range_partitions_exactly():
    for n in (0, 1, 2, 7, 8, 9, 200, 1001):
        for world in (1, 2, 3, 4, 8):
            got = [driver.shard_range(n, r, world) for r in range(world)]
            assert got[0][0] == 0 and got[-1][1] == n
            assert all(got[i][1] == got[i + 1][0] for i in range(world - 1))          # contiguous, disjoint
            sizes = [hi - lo for lo, hi in got]
            assert max(sizes) - min(sizes) <= 1 and sizes == sorted(sizes, reverse=True)


def test_shard_range_rejects_bad_arguments():
    with pytest.raises(ValueError):
        driver.shard_range(10, 2, 2)
    with pytest.raises(ValueError):
        driver.shard_range(-1, 0, 2)
    with pytest.raises(ValueError):
        driver.batches(0, 10, 0)


def test_batches_cover_ragged_tail_and_empty():
    assert driver.batches(3, 3, 8) == []
    assert driver.batches(0, 8, 8) == [(0, 8)]
    assert driver.batches(5, 22, 8) == [(5, 8), (13, 8), (21, 1)]


def test_video_shard_covers_every_pair_once():
    for n_frames in (0, 1, 2, 3, 9, 100, 1001):
        for world in (1, 2, 3, 8):
            chunks = [driver.video_shard(n_frames, r, world) for r in range(world)]
            pairs = []
            for ch in chunks:
                if ch is not None:
                    first, last = ch
                    assert 0 <= first < last <= n_frames - 1
                    pairs.extend(range(first, last))                   # pair t = frames (t, t+1), both inside the chunk
            assert pairs == list(range(max(n_frames - 1, 0)))
            live = [c for c in chunks if c is not None]
            assert all(live[i][1] == live[i + 1][0] for i in range(len(live) - 1))   # neighbours share exactly one frame
    with pytest.raises(ValueError):
        driver.video_shard(-1, 0, 1)
    from opticalflow_b200.models.PWCNet import PWCDCNet
    with pytest.raises(RuntimeError):
        driver.VideoFlowPipeline(PWCDCNet(), 1, 64, 64, torch.device("cpu"))


def test_pipeline_and_model_refuse_cpu():
    from opticalflow_b200.models.PWCNet import PWCDCNet
    m = PWCDCNet()
    with pytest.raises(RuntimeError):
        driver.FlowPipeline(m, 1, 64, 64, torch.device("cpu"))
    with pytest.raises(RuntimeError):
        m(torch.zeros(1, 6, 64, 64))


def test_reduce_helpers_without_process_group():
    assert driver.max_over_ranks([1.5, 2.0]) == [1.5, 2.0]
    assert driver.sum_over_ranks([3.0]) == [3.0]


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_pairs, batch, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        assert driver.dist_env() == (rank, rank, world)
        lo, hi = driver.shard_range(n_pairs, rank, world)
        done = []
        for s, c in driver.batches(lo, hi, batch):          # stand-in for load -> forward -> sink
            done.extend(range(s, s + c))
        chunk = driver.video_shard(n_pairs + 1, rank, world)         # the same pairs seen as a stream of n_pairs + 1 frames
        assert (chunk is None and lo == hi) or chunk == (lo, hi)
        t_max = driver.max_over_ranks([float(rank + 1), 10.0 - rank])
        n_sum = driver.sum_over_ranks([float(len(done))])
        gathered = [None] * world
        dist.all_gather_object(gathered, done)               # test-only: the data path itself has no collective
        if rank == 0:
            q.put((gathered, t_max, n_sum))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_pairs,batch", [(200, 8), (17, 8), (1, 8)])
def test_two_ranks_gloo_cover_all_pairs_once(n_pairs, batch):
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_pairs, batch, q)) for r in range(world)]
    for p in procs:
        p.start()
    gathered, t_max, n_sum = q.get(timeout=120)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    flat = [i for part in gathered for i in part]
    assert sorted(flat) == list(range(n_pairs)) and len(set(flat)) == n_pairs
    assert t_max == [2.0, 10.0]                              # max over ranks, as bench.py times a step
    assert n_sum == [float(n_pairs)]


def test_bind_host_to_gpu_is_harmless_without_a_gpu():
    """No NVML / no CUDA device: the NUMA binding helper reports None and leaves the affinity mask alone."""
    import os
    from opticalflow_b200 import driver
    before = os.sched_getaffinity(0)
    try:
        cpus = driver.bind_host_to_gpu(0)
        assert cpus is None or (isinstance(cpus, list) and set(cpus) <= before)
    finally:
        os.sched_setaffinity(0, before)
