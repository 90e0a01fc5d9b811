"""N > 1 host logic on CPU (gloo, world_size 2): VFO sharding, FFT frame time-sharding, stream broadcast."""
import os
import socket

import numpy as np
import pytest

from sdrplusplusbrown_b200 import configs, sharding


def test_vfo_sharding_partitions_and_balances():
    for cfg, n in [("cfg3", 256), ("cfg4", 1024), ("cfg2", 1)]:
        vfos = configs.vfo_list(cfg, n)
        sr = configs.CONFIGS[cfg]["sample_rate"]
        for world in (1, 2, 4, 8):
            shards = sharding.shard_vfos(vfos, sr, world)
            flat = sorted(i for s in shards for i in s)
            assert flat == list(range(len(vfos)))                   # every VFO exactly once
            loads = [sum(sharding.vfo_cost(sr, vfos[i][1], vfos[i][2]) for i in s) for s in shards]
            if len(vfos) >= 8 * world:
                assert max(loads) <= 1.05 * (sum(loads) / world)     # balanced by cost (WFM > SSB > AM)


def test_frame_time_sharding_covers_every_frame_once():
    interval, nz = 333333, 65536
    pos, seen = 0, {r: [] for r in range(4)}
    all_frames = []
    for count in [1000000, 12000, 333333, 65535, 1, 2000000, 500000]:
        fr = sharding.frames_in_block(pos, count, interval, nz)
        all_frames += fr
        for r in range(4):
            seen[r] += sharding.frames_for_rank(pos, count, interval, nz, r, 4)
        pos += count
    total = (pos - nz) // interval + 1
    assert all_frames == list(range(total))
    merged = sorted(f for r in seen for f in seen[r])
    assert merged == list(range(total))
    for r in range(4):
        assert all(f % 4 == r for f in seen[r])


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from sdrplusplusbrown_b200 import synth
        cfg = configs.CONFIGS["cfg3"]
        sr = cfg["sample_rate"]
        my_vfos = sharding.plan("cfg3", world, 16)[rank]
        interval, nz = int(round(sr / cfg["fft_rate"])), cfg["fft_size"]
        pos, my_frames, sums = 0, [], []
        for step in range(3):
            n = 300000
            if rank == 0:
                blk = torch.from_numpy(synth.make_iq(pos, n, sr).view(np.float32).reshape(-1, 2).copy())
            else:
                blk = torch.zeros((n, 2), dtype=torch.float32)
            sharding.broadcast_block(blk, src=0)                    # the path's one exchange step
            sums.append(float(blk.double().sum()))
            my_frames += sharding.frames_for_rank(pos, n, interval, nz, rank, world)
            pos += n
        everyone = sharding.gather_frame_indices(my_frames)
        q.put((rank, len(my_vfos), sums, everyone))
    finally:
        dist.destroy_process_group()


def test_broadcast_and_sharding_world_size_2_gloo():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=180) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (r0, nv0, s0, e0), (r1, nv1, s1, e1) = res
    assert nv0 + nv1 == 16 and abs(nv0 - nv1) <= 1               # VFO bank split across ranks
    assert s0 == s1                                              # both ranks hold rank 0's blocks after the broadcast
    assert e0 == e1
    frames = sorted(f for part in e0 for f in part)
    assert frames == list(range(len(frames))) and len(frames) >= 1
    assert all(f % 2 == 0 for f in e0[0]) and all(f % 2 == 1 for f in e0[1])
