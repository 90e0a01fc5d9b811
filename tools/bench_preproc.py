#!/usr/bin/env python3
"""Pre-processing chain at cfg4's rate (122.88 MS/s): power-of-two decimator + DC blocker (single-pass chained scan) on a
16 Mi-sample resident block.  One JSON line per configuration: step time and the DC blocker's own time / algorithmic GB/s."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from sdrplusplusbrown_b200 import binding, synth  # noqa: E402

sr = 122.88e6
n = 1 << 24
x = synth.make_iq_torch(n, sr, torch.device("cuda"), dc=(0.05, -0.02))
hbm = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(
    os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")) else 6650.0
for dec, dc in ((1, True), (4, True), (4, False), (16, True)):
    fe = binding.FrontEnd(sr, decim_ratio=dec, dc_blocking=dc, fft_size=1024, fft_rate=1.0, max_block=n)
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for _ in range(3):
            fe.process(x, n, stream=s)
        torch.cuda.synchronize()
        binding.profile_enable(True)
        binding.profile_read_kind(binding.PROF_DC)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        for _ in range(10):
            fe.process(x, n, stream=s)
        e1.record(s)
    torch.cuda.synchronize()
    ms, k, by, fl = binding.profile_read_kind(binding.PROF_DC)
    binding.profile_enable(False)
    step = e0.elapsed_time(e1) / 10
    gbs = by / 1e9 / (ms * 1e-3) if ms else None
    print(json.dumps({"sample_rate": sr, "block": n, "decim": dec, "dc_block": dc, "step_ms": step, "x_real_time": n / (step * 1e-3) / sr,
                      "dc_kernel_ms": ms / k if k else None, "dc_algorithmic_GBps": gbs, "dc_hbm_frac": gbs / hbm if gbs else None}), flush=True)

# the three-kernel form of the same scan (what b200_dc_block_noalloc runs: reduce, scan, apply; 24 bytes per sample)
d_out = torch.empty_like(x)
d_off = torch.zeros(2, dtype=torch.float64, device="cuda")
for _ in range(3):
    binding.dc_block_noalloc(x, d_out, n, 50.0 / sr, d_off)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    binding.dc_block_noalloc(x, d_out, n, 50.0 / sr, d_off)
e1.record()
torch.cuda.synchronize()
ms3 = e0.elapsed_time(e1) / 10
print(json.dumps({"three_kernel_dc_ms": ms3, "algorithmic_GBps": 16.0 * n / 1e9 / (ms3 * 1e-3)}))
