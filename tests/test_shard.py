"""Host logic of the multi-GPU path (mfcc_b200/shard.py) under gloo, world_size 2, on
CPU.  The compute on each rank is the C oracle standing in for the CUDA operator: what
is tested is the partitioning, the reductions and the gather, i.e. everything the N > 1
bench and a sharded job do around the (collective-free) kernel."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from mfcc_b200 import shard


def test_shard_clips_partitions_exactly():
    for n in (0, 1, 7, 8, 10000, 16667):
        for w in (1, 2, 3, 4, 8):
            for layout in ("block", "cyclic"):
                got = sorted(i for r in range(w) for i in shard.shard_clips(n, r, w, layout))
                assert got == list(range(n))
                sizes = [len(shard.shard_clips(n, r, w, layout)) for r in range(w)]
                assert max(sizes) - min(sizes) <= 1
    assert list(shard.shard_clips(10, 1, 4, "cyclic")) == [1, 5, 9]
    assert shard.shard_clips(10, 1, 4, "block") == range(3, 6)
    with pytest.raises(ValueError):
        shard.shard_clips(4, 4, 4)
    assert [shard.stream_owner(s, 8) for s in (0, 7, 8, 4095)] == [0, 7, 0, 7]


def test_split_long_clip_reassembles(wav):
    from oracle import ref_c as RC
    RC.build()
    cfg = RC.RefConfig(nfilters=32, nceptrums=32)
    x = wav[:60000]
    whole = RC.run(cfg, x)
    for parts in (1, 2, 3, 8):
        got = []
        for f0, f1, s0, s1, skip in shard.split_long_clip(len(x), parts):
            piece = RC.run(cfg, x[s0:s1])
            assert piece.shape[0] == skip + f1 - f0
            got.append(piece[skip:])
        assert np.array_equal(np.concatenate(got), whole)
    assert shard.split_long_clip(100, 4) == []


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world_size, port, n_clips, layout, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world_size)
    try:
        from oracle import ref_c as RC
        cfg = RC.RefConfig(nfilters=32, nceptrums=16)
        rng = np.random.default_rng(1234)                     # same data on every rank
        x = rng.integers(-3000, 3000, (n_clips, 2000)).astype(np.int16)
        mine = shard.shard_clips(n_clips, rank, world_size, layout)
        assert shard.world() == (rank, world_size)
        local = RC.run(cfg, x[mine.start:mine.stop:mine.step]) if len(mine) else \
            np.zeros((0, cfg.num_frames(2000), 16), np.int16)
        frames = shard.reduce_sum(local.shape[0] * local.shape[1])
        slowest = shard.reduce_max(float(rank + 1))
        shard.barrier()
        full = shard.gather_cepstra(torch.from_numpy(local), n_clips, layout)
        if rank == 0:
            want = RC.run(cfg, x)
            q.put((frames == want.shape[0] * want.shape[1], slowest == float(world_size),
                   bool(np.array_equal(full.numpy(), want))))
        else:
            assert full is None
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("layout,n_clips", [("block", 7), ("cyclic", 7), ("block", 1)])
def test_two_ranks_gloo(layout, n_clips):
    from oracle import ref_c as RC
    RC.build()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_clips, layout, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert q.get(timeout=5) == (True, True, True)
