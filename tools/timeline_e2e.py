#!/usr/bin/env python
"""Copy/kernel timeline of one b200_frontend_process_host call (diagnosis only; never a reported number)."""
import ctypes, json, os, sys, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from torch.profiler import ProfilerActivity, profile
from sdrplusplusbrown_b200 import binding, configs, synth

cfg = configs.CONFIGS["cfg2"]
sr = cfg["sample_rate"]
Be = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 24
dev = torch.device("cuda", 0)
fe = binding.FrontEnd(sr, fft_size=cfg["fft_size"], fft_rate=cfg["fft_rate"], fft_window=cfg["window"], max_block=Be, device=0)
mode, out_sr, bw, off = configs.vfo_list("cfg2")[0]
v = fe.vfos.add(out_sr, bw, off)
fe.vfos.set_demod(v, binding.DEMOD_USB, 0, 50.0 / out_sr, 5.0 / out_sr, 0)
h_in = torch.empty((Be, 2), dtype=torch.float32).pin_memory()
h_in.copy_(synth.make_iq_torch(Be, sr, dev, tones=[(3.3e6, 0.2)]).cpu())
mf = Be // int(round(sr / cfg["fft_rate"])) + 2
h_frames = torch.empty((mf, cfg["fft_size"]), dtype=torch.float32).pin_memory()
mo = fe.vfos.max_out(v, Be)
h_iq = torch.empty((mo, 2), dtype=torch.float32).pin_memory()
h_au = torch.empty((mo, 2), dtype=torch.float32).pin_memory()
optr = (ctypes.c_void_p * (v + 1))(); optr[v] = h_iq.data_ptr()
aptr = (ctypes.c_void_p * (v + 1))(); aptr[v] = h_au.data_ptr()
counts = np.zeros(v + 1, np.int32)
s = torch.cuda.Stream()

def step():
    rc = binding.lib.b200_frontend_process_host(fe.h, h_in.data_ptr(), Be, h_frames.data_ptr(), mf, optr, aptr, counts.ctypes.data, s.cuda_stream)
    assert rc >= 0, binding.lib.b200_last_error().decode()

for _ in range(2):
    step()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    step()
    torch.cuda.synchronize()
path = os.path.join(tempfile.gettempdir(), "b200_e2e.json")
prof.export_chrome_trace(path)
ev = [e for e in json.load(open(path))["traceEvents"] if e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset") and "dur" in e]
ev.sort(key=lambda e: e["ts"])
t0 = ev[0]["ts"]
big = [e for e in ev if e["dur"] > 20.0]
print("%-52s %7s %10s %9s" % ("activity (> 20 us)", "stream", "start us", "dur us"))
for e in big:
    print("%-52s %7s %10.1f %9.1f" % (e["name"][:52], e.get("args", {}).get("stream", "?"), e["ts"] - t0, e["dur"]))
print("records %d, span %.1f us" % (len(ev), ev[-1]["ts"] + ev[-1]["dur"] - t0))
h2d = [e for e in ev if "HtoD" in e["name"] and e["dur"] > 50]
print("H2D: %d copies, %.1f us total, %.1f GB/s" % (len(h2d), sum(e["dur"] for e in h2d), Be * 8 / 1e3 / max(sum(e["dur"] for e in h2d), 1)))
