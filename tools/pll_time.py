#!/usr/bin/env python
"""Time of the broadcast-FM stereo chain for a few VFOs (the pilot PLL walk dominates): kernel records of one block."""
import json, os, re, sys, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from torch.profiler import ProfilerActivity, profile
from sdrplusplusbrown_b200 import binding, synth

sr, B, V = 4e6, 1 << 22, int(sys.argv[1]) if len(sys.argv) > 1 else 16
dev = torch.device("cuda", 0)
b = binding.VfoBank(sr, max_block=B)
if "exact" in sys.argv[2:]:
    b.set_rotator_mode(binding.ROTATOR_EXACT)       # the loop in the reference's operation order (pllKernel)
for i in range(V):
    v = b.add(500e3, 150e3, -1.5e6 + 200e3 * i)
    b.set_demod(v, binding.DEMOD_WFM, 1, 1, 0, 0)
if "nopilot" in sys.argv[2:]:
    x = synth.make_iq_torch(B, sr, dev, tones=[(1e5, 0.2)])          # nothing at 19 kHz: the loop sits at a frequency limit
else:
    # a broadcast-FM carrier with a 19 kHz pilot under every VFO: the loop locks
    t = torch.arange(B, device=dev, dtype=torch.float64) / sr
    mpx = 0.1 * torch.sin(2 * torch.pi * 19000.0 * t) + 0.4 * torch.sin(2 * torch.pi * 1000.0 * t)
    ph = 2 * torch.pi * 75e3 * torch.cumsum(mpx, 0) / sr
    z = torch.zeros(B, dtype=torch.complex128, device=dev)
    for i in range(V):
        z += (0.5 / V) * torch.exp(1j * (ph + 2 * torch.pi * (-1.5e6 + 200e3 * i) * t))
    x = torch.view_as_real(z.to(torch.complex64)).contiguous()
    del z, t, mpx, ph
for _ in range(3):
    b.process(x, B)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    b.process(x, B)
    torch.cuda.synchronize()
path = os.path.join(tempfile.gettempdir(), "pll.json")
prof.export_chrome_trace(path)
ev = [e for e in json.load(open(path))["traceEvents"] if e.get("cat") == "kernel"]
tot = {}
for e in ev:
    mm = re.search(r"(\w+)(<[^(]*>)?\(", e["name"].replace("(anonymous namespace)", "anon"))
    n = (mm.group(1) if mm else e["name"])[:40]
    tot[n] = tot.get(n, 0.0) + e["dur"]
n_out = B * 500e3 / sr
for k, v in sorted(tot.items(), key=lambda kv: -kv[1])[:8]:
    print("%-42s %9.1f us" % (k, v))
pll = tot.get("pllKernel", 0) + tot.get("pllErrKernel", 0)     # exact banks / default banks
print("pll walk: %.1f cycles/sample at 1.965 GHz (%d samples per VFO)" % (pll * 1965.0 / n_out, n_out))
