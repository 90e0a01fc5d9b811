#!/usr/bin/env python
"""Host-side cost of one b200_frontend_process call (descriptor building + launches), measured with blocks so small
that the GPU is never the bottleneck.  Diagnosis only."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from sdrplusplusbrown_b200 import binding, configs, synth

cfg = configs.CONFIGS["cfg2"]
sr = cfg["sample_rate"]
for nv, with_fft in ((0, True), (1, False), (1, True), (8, True)):
    B = 400000
    fe = binding.FrontEnd(sr, fft_size=cfg["fft_size"], fft_rate=cfg["fft_rate"] if with_fft else 0.05, fft_window=cfg["window"], max_block=B, device=0)
    mode, out_sr, bw, off = configs.vfo_list("cfg2")[0]
    for i in range(nv):
        fe.vfos.add(out_sr, bw, off + 1e3 * i)
    x = synth.make_iq_torch(B, sr, torch.device("cuda", 0), tones=[(3.3e6, 0.2)])
    s = torch.cuda.Stream()
    for _ in range(20):
        fe.process(x, B, stream=s)
    torch.cuda.synchronize()
    n = 300
    t0 = time.perf_counter()
    for _ in range(n):
        fe.process(x, B, stream=s)
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    print("vfos %d fft %d: host %.1f us/call (enqueue only), %.1f us/call incl. drain" % (nv, with_fft, (t1 - t0) / n * 1e6, (t2 - t0) / n * 1e6))
