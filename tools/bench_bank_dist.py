#!/usr/bin/env python3
"""Sharded VFO bank over N GPUs of one node: the whole bank of a configuration (cfg3: 256 NFM VFOs at
61.44 MS/s; cfg4: 1024 mixed AM/SSB/WFM VFOs at 122.88 MS/s + 2^20-point FFT frames sharded by time) is dealt out
by cost (sharding.shard_vfos: every shard the same mix of modes); rank 0 holds the stream and broadcasts every block (NCCL over NVLink, double buffered
on a communication stream beside the kernels of the previous block); every rank runs its shard through the front end.
STRONG scaling: the work per block is fixed, the time per block is the slowest rank's.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29533 \
        tools/bench_bank_dist.py --cfg cfg3 [--steps 10] [--batch 16777216] [--demod]
    python tools/bench_bank_dist.py --cfg cfg3           (N = 1)

One JSON line on rank 0.  Not the contract benchmark (bench.py is); this is the evidence for DESIGN.md's statement that
the broadcast hides behind the kernels once a shard holds tens of VFOs.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cfg", default="cfg3")
    ap.add_argument("--vfos", default="0", help="BASELINE config[4] (cfg5): this many NFM VFOs at cfg3's rate instead of the configuration's own bank (1 ... 4096); "
                    "a comma-separated list runs one measurement per count in the same process group (one JSON line each)")
    ap.add_argument("--batch", type=int, default=1 << 24)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--demod", action="store_true")
    ap.add_argument("--no-fft", action="store_true")
    ap.add_argument("--wfm-stereo", type=int, default=0, help="WFM demodulators decode stereo (pilot PLL); the radio's default is mono")
    ap.add_argument("--defer-tail", type=int, default=1)
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    # keep stdout for the JSON line only (NCCL prints its banner there)
    out_fd = os.dup(1)
    os.dup2(2, 1)
    import torch
    import torch.distributed as dist
    from sdrplusplusbrown_b200 import binding, configs, synth

    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    cfg = configs.CONFIGS[args.cfg]
    sr = cfg["sample_rate"]
    B = args.batch
    def measure(nv):
        fft_rate = 0.01 if args.no_fft else cfg["fft_rate"]
        fe = binding.FrontEnd(sr, fft_size=cfg["fft_size"], fft_rate=fft_rate, fft_window=cfg["window"], max_block=B, device=local)
        if world > 1:
            fe.set_frame_shard(rank, world)
        vl = configs.vfo_list(args.cfg)
        if nv > 0:
            # cfg5: the NFM bank of cfg3 swept in size; beyond the 256 grid points the offsets are spread over the band
            import numpy as np
            if nv <= len(vl):
                vl = vl[:nv]
            else:
                rng = np.random.default_rng(4096)
                vl = [("NFM48", 48000.0, 12500.0, float(f)) for f in np.sort(rng.uniform(-0.45 * sr, 0.45 * sr, nv))]
        # deal the bank out by cost (WFM > SSB > AM; sharding.shard_vfos, the same function the gloo tests cover).
        # For cfg4 that is 64 AM + 48 SSB + 16 WFM on every one of 8 ranks -- the mix the recorded 46.4 x run had, dealt there
        # mode by mode (the j-th VFO of a mode to rank j % N); plain i % N puts cfg4's 128 WFM channels on one rank.
        from sdrplusplusbrown_b200 import sharding
        mine = sorted(sharding.shard_vfos(vl, sr, world)[rank])
        demod_of = {"NFM48": (binding.DEMOD_NFM, 1, 0, 0, 0), "NFM": (binding.DEMOD_NFM, 1, 0, 0, 0),
                    "AM": (binding.DEMOD_AM, 1, None, None, None), "USB": (binding.DEMOD_USB, 0, None, None, 0),
                    "WFM": (binding.DEMOD_WFM, args.wfm_stereo, 1, 0, 0)}
        for i in mine:
            mode, out_sr, bw, off = vl[i]
            v = fe.vfos.add(out_sr, bw, off)
            if args.demod:
                kind, p0, p1, p2, p3 = demod_of[mode]
                if mode == "AM":
                    p1, p2, p3 = 50.0 / out_sr, 5.0 / out_sr, 100.0 / out_sr
                if mode == "USB":
                    p1, p2 = 50.0 / out_sr, 5.0 / out_sr
                fe.vfos.set_demod(v, kind, p0, p1, p2, p3)
        fe.vfos.set_option(binding.OPT_DEFER_TAIL, args.defer_tail)

        if rank == 0:
            src = synth.make_iq_torch(B, sr, dev, tones=[(vl[0][3] + 700.0, 0.05), (3.3e6, 0.2)])
            bufs = [src, src]
        else:
            bufs = [torch.empty((B, 2), dtype=torch.float32, device=dev) for _ in range(2)]
        compute = torch.cuda.Stream(device=dev, priority=-1)
        comm = torch.cuda.Stream(device=dev, priority=-3) if world > 1 else None      # above the compute stream: see bench.py
        ev_ready = [torch.cuda.Event() for _ in bufs]
        ev_free = [torch.cuda.Event() for _ in bufs]
        ev_b0 = [torch.cuda.Event(enable_timing=True) for _ in range(64)]
        ev_b1 = [torch.cuda.Event(enable_timing=True) for _ in range(64)]

        def bcast(i, timed=False):
            b = bufs[i % 2]
            with torch.cuda.stream(comm):
                comm.wait_event(ev_free[i % 2])
                if timed and i < 64:
                    ev_b0[i].record(comm)
                dist.broadcast(b, src=0)
                if timed and i < 64:
                    ev_b1[i].record(comm)
                ev_ready[i % 2].record(comm)

        def run(n, timed=False):
            if world > 1:
                for e in ev_free:
                    e.record(compute)
                bcast(0, timed)
            for i in range(n):
                if world > 1 and i + 1 < n:
                    bcast(i + 1, timed)
                if world > 1:
                    compute.wait_event(ev_ready[i % 2])
                with torch.cuda.stream(compute):
                    fe.process(bufs[i % 2], B, stream=compute)
                    if world > 1:
                        ev_free[i % 2].record(compute)

        def barrier():
            torch.cuda.synchronize(dev)
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize(dev)

        run(max(args.warmup, 3))
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(compute)
        run(args.steps, timed=True)
        if mine:
            fe.vfos.out_ptr(0)       # order the last deferred stages into the compute stream
        e1.record(compute)
        barrier()
        ms = e0.elapsed_time(e1)
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        bc = None
        if world > 1:
            n = min(args.steps, 64)
            bc = sum(ev_b0[i].elapsed_time(ev_b1[i]) for i in range(n)) / n
        if rank == 0:
            line = {"cfg": args.cfg if nv == 0 else "cfg5 (%d NFM VFOs at %s's rate)" % (nv, args.cfg), "n_gpus": world, "vfos_total": len(vl), "vfos_per_gpu": len(mine), "demod": bool(args.demod), "wfm_stereo": bool(args.wfm_stereo) if args.demod else None,
                    "batch": B, "steps": args.steps, "ms_per_block": ms / args.steps,
                    "input_MSps": B * args.steps / (ms * 1e-3) / 1e6, "x_real_time": B * args.steps / (ms * 1e-3) / sr,
                    "broadcast_ms_per_block_rank0": bc, "scaling": "strong", "defer_tail": args.defer_tail,
                    "fft": None if args.no_fft else {"size": cfg["fft_size"], "rate": cfg["fft_rate"], "frames": "time-sharded f % N == rank"}}
            os.write(out_fd, (json.dumps(line) + "\n").encode())
        del fe
        torch.cuda.synchronize(dev)

    for nv in [int(v) for v in str(args.vfos).split(",")]:
        measure(nv)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
