#!/usr/bin/env python
"""Copy/kernel timeline of one b200_frontend_process_host call on a BASELINE bank configuration (diagnosis only).

    python tools/timeline_e2e_bank.py [cfg4] [--streaming] [--min-us 150]
"""
import argparse, ctypes, json, os, sys, tempfile, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
from torch.profiler import ProfilerActivity, profile
import bench
from sdrplusplusbrown_b200 import binding, configs, synth

ap = argparse.ArgumentParser()
ap.add_argument("workload", nargs="?", default="cfg4")
ap.add_argument("--streaming", action="store_true")
ap.add_argument("--batch", type=int, default=1 << 24)
ap.add_argument("--min-us", type=float, default=150.0)
ap.add_argument("--all-until-us", type=float, default=0.0, help="list every record that starts before this time")
args = ap.parse_args()
cfg = configs.CONFIGS[args.workload]
sr, Be = cfg["sample_rate"], args.batch
dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
j = bench.Job(binding, torch, args.workload, Be, 0, 0, 1, defer_tail=args.streaming)
fe = j.fe
h_in = torch.empty((Be, 2), dtype=torch.float32).pin_memory()
h_in.copy_(synth.make_iq_torch(Be, sr, dev, tones=[(3.3e6, 0.2)]).cpu())
mf = Be // int(round(sr / cfg["fft_rate"])) + 2
h_frames = torch.empty((mf, cfg["fft_size"]), dtype=torch.float32).pin_memory()
sizes = [fe.vfos.max_out(v, Be) for v in j.ids]
h_iq = torch.empty((sum(sizes), 2), dtype=torch.float32).pin_memory()
optr = (ctypes.c_void_p * (max(j.ids) + 1))()
pos = 0
for v, n in zip(j.ids, sizes):
    optr[v] = h_iq.data_ptr() + pos * 8
    pos += n
counts = np.zeros(max(j.ids) + 1, np.int32)
s = torch.cuda.Stream()

def step():
    rc = binding.lib.b200_frontend_process_host(fe.h, h_in.data_ptr(), Be, h_frames.data_ptr(), mf, optr, None, counts.ctypes.data, s.cuda_stream)
    assert rc >= 0, binding.lib.b200_last_error().decode()

for _ in range(3):
    step()
torch.cuda.synchronize()
t0 = time.perf_counter()
step()
wall = (time.perf_counter() - t0) * 1e3
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    step()
    torch.cuda.synchronize()
path = os.path.join(tempfile.gettempdir(), "b200_e2e_bank.json")
prof.export_chrome_trace(path)
ev = [e for e in json.load(open(path))["traceEvents"] if e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset") and "dur" in e]
ev.sort(key=lambda e: e["ts"])
t0 = ev[0]["ts"]
print("one call of %d samples, streaming %d: %.2f ms of wall time unprofiled" % (Be, args.streaming, wall))
print("%-52s %7s %10s %9s" % ("activity (> %g us)" % args.min_us, "stream", "start us", "dur us"))
for e in ev:
    if e["dur"] > args.min_us or e["ts"] - t0 < args.all_until_us:
        print("%-52s %7s %10.1f %9.1f" % (e["name"][:52], e.get("args", {}).get("stream", "?"), e["ts"] - t0, e["dur"]))
print("records %d, span %.1f us" % (len(ev), ev[-1]["ts"] + ev[-1]["dur"] - t0))
# idle time of the GPU: gaps between the union of all activities
gaps, end = 0.0, ev[0]["ts"]
for e in ev:
    if e["ts"] > end:
        gaps += e["ts"] - end
    end = max(end, e["ts"] + e["dur"])
print("idle inside the span: %.1f us" % gaps)
fe.vfos.set_option(binding.OPT_DEFER_TAIL, 0)
