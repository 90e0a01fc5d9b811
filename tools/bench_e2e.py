#!/usr/bin/env python
"""The host-fed path (b200_frontend_process_host: pinned host block in, dB rows and every VFO's IF output back out) on a
BASELINE configuration, inline and in streaming mode, for a few chunk sizes.  One JSON line per variant.

    python tools/bench_e2e.py [cfg4] [--steps 4]
"""
import argparse, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
from sdrplusplusbrown_b200 import binding, build, configs, synth

ap = argparse.ArgumentParser()
ap.add_argument("workload", nargs="?", default="cfg4")
ap.add_argument("--steps", type=int, default=4)
ap.add_argument("--batch", type=int, default=1 << 24)
ap.add_argument("--chunks", default="0,32,16")
args = ap.parse_args()
build.build()
dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
compute = torch.cuda.Stream(dev)
cfg = configs.CONFIGS[args.workload]
src = synth.make_iq_torch(args.batch, cfg["sample_rate"], dev, tones=[(0.11 * cfg["sample_rate"], 0.2)])
for defer in (False, True):
    for mib in [int(c) for c in args.chunks.split(",")]:
        r = bench.host_path(binding, torch, args.workload, args.batch, 0, 0, 1, src, compute, args.steps, None, dev, defer_tail=defer, chunk_mib=mib)
        r.update({"workload": args.workload, "chunk_mib": mib if mib else "automatic"})
        r.pop("path", None)
        print(json.dumps(r), flush=True)
