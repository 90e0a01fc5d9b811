"""Host-side multi-GPU logic: which rank owns which VFO, which FFT frames a rank computes,
and the per-block stream broadcast.  Pure Python + torch.distributed, so the N > 1 path is testable on CPU with
the gloo backend; on GPUs the same code runs over NCCL/NVLink.

* VFOs are independent given the same input (the reference fans the block out with a Splitter memcpy,
  routing/splitter.h:74-83), so they are partitioned across ranks, balanced by their per-sample cost.
* FFT frames are independent windows [f*I, f*I + nz) (iq_frontend.h:88-92): frame f belongs to rank f % world.
* The one exchange step is the broadcast of each input block from the ingest rank.
"""
import math

from . import configs


def vfo_cost(in_sr, out_sr, bw):
    """FP32 instructions per input sample for one VFO: rotation + stage-1 taps dominate;
    the tail scales with the output rate."""
    ratio = max(1.0, in_sr / out_sr)
    power = 2 ** int(math.floor(math.log2(ratio))) if ratio >= 2 else 1
    stage1 = 12.0 if power >= 16 else 4.0 + 2.0 * 35.0   # unfused small ratios pay the long /2 filter at full rate
    chan_taps = 3.8 * out_sr / (bw / 20.0) if bw != out_sr else 0.0
    tail = (2.0 * chan_taps + 2.0 * 130.0) * out_sr / in_sr
    return stage1 + tail


def shard_vfos(vfos, in_sr, world):
    """Greedy longest-processing-time partition.  vfos: [(mode, out_sr, bw, offset), ...].
    Returns a list of `world` lists of indices into `vfos`; every index appears exactly once."""
    order = sorted(range(len(vfos)), key=lambda i: -vfo_cost(in_sr, vfos[i][1], vfos[i][2]))
    loads = [0.0] * world
    shards = [[] for _ in range(world)]
    for i in order:
        r = min(range(world), key=lambda k: (loads[k], k))
        shards[r].append(i)
        loads[r] += vfo_cost(in_sr, vfos[i][1], vfos[i][2])
    for s in shards:
        s.sort()
    return shards


def frame_owner(frame_index, world):
    return frame_index % world


def frames_in_block(stream_pos, count, interval, nz):
    """Global indices of the FFT frames that complete inside the block [stream_pos, stream_pos + count)
    (the Reshaper's keep/skip framing, buffer/reshaper.h:101-128)."""
    first = max(0, -(-(stream_pos - nz + 1) // interval)) if stream_pos + 1 > nz else 0
    out = []
    f = first
    while f * interval + nz <= stream_pos + count:
        if f * interval + nz > stream_pos:
            out.append(f)
        f += 1
    return out


def frames_for_rank(stream_pos, count, interval, nz, rank, world):
    return [f for f in frames_in_block(stream_pos, count, interval, nz) if frame_owner(f, world) == rank]


def broadcast_block(block, src=0, group=None):
    """Every rank ends up with the ingest rank's block (tensor is overwritten in place on the other ranks)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.broadcast(block, src=src, group=group)
    return block


def gather_frame_indices(local_indices, group=None):
    """All ranks' frame indices at every rank (small control-plane exchange used by tests and by a display owner)."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return [list(local_indices)]
    out = [None] * dist.get_world_size(group)
    dist.all_gather_object(out, list(local_indices), group=group)
    return out


def plan(cfg_name, world, count=None):
    """Per-rank VFO lists for a BASELINE configuration."""
    cfg = configs.CONFIGS[cfg_name]
    vfos = configs.vfo_list(cfg_name, count)
    shards = shard_vfos(vfos, cfg["sample_rate"], world)
    return [[vfos[i] for i in s] for s in shards]
