#!/usr/bin/env python
"""Step time of the bench.py workload (cfg2: FFT branch with attached waterfall + 1 USB VFO, 64 Mi samples resident)
for each setting of the two scheduling switches: B200_OPT_BRANCH_ORDER (0 = FFT branch enqueued first, 1 = right behind
the channelizer's full-rate kernel) and B200_OPT_FFT_CTAS_PER_SM (0 = as many as fit, 1), and B200_OPT_DEFER_TAIL (later stages of block k beside the full-rate stage of block k+1).  CUDA events, no profiler.

    python tools/order_sweep.py [--batch N] [--steps K] [--timeline ORDER,CTAS]
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=1 << 26)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--vfos", type=int, default=1)
    ap.add_argument("--cfg", default="cfg2")
    args = ap.parse_args()
    import torch
    from sdrplusplusbrown_b200 import binding, configs, synth

    cfg = configs.CONFIGS[args.cfg]
    sr = cfg["sample_rate"]
    B = args.batch
    dev = torch.device("cuda", 0)
    fe = binding.FrontEnd(sr, fft_size=cfg["fft_size"], fft_rate=cfg["fft_rate"], fft_window=cfg["window"], max_block=B, device=0)
    vl = configs.vfo_list(args.cfg)
    for i in range(args.vfos):
        mode, out_sr, bw, off = vl[i % len(vl)]
        fe.vfos.add(out_sr, bw, off + 1e3 * (i // len(vl)))
    width = 1333
    wf = binding.Waterfall(cfg["fft_size"], width, device=0)
    wf.set_view(0, cfg["fft_size"])
    wf.set_range(-120.0, 0.0)
    max_frames = B // int(round(sr / cfg["fft_rate"])) + 2
    d_zoom = torch.empty((max_frames, width), dtype=torch.float32, device=dev)
    d_rgba = torch.empty((max_frames, width), dtype=torch.int32, device=dev)
    fe.attach_waterfall(wf, d_zoom.data_ptr(), d_rgba.data_ptr(), max_frames)
    x = synth.make_iq_torch(B, sr, dev, tones=[(3.3e6, 0.2)])
    s = torch.cuda.Stream(device=dev)

    def run(k):
        with torch.cuda.stream(s):
            for _ in range(k):
                fe.process(x, B, stream=s)

    for defer, order, ctas in [(0, 0, 0), (0, 1, 0), (0, 0, 1), (0, 1, 1), (1, 0, 0), (1, 1, 0), (1, 0, 1), (1, 1, 1)]:
        if True:
            fe.vfos.set_option(binding.OPT_DEFER_TAIL, defer)
            fe.vfos.set_option(binding.OPT_BRANCH_ORDER, order)
            fe.vfos.set_option(binding.OPT_FFT_CTAS_PER_SM, ctas)
            run(5)
            torch.cuda.synchronize()
            best = 1e9
            for rep in range(3):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                with torch.cuda.stream(s):
                    e0.record(s)
                run(args.steps)
                with torch.cuda.stream(s):
                    e1.record(s)
                torch.cuda.synchronize()
                best = min(best, e0.elapsed_time(e1) / args.steps)
            print("defer_tail %d  order %d  fft_ctas_per_sm %d  %.1f us/step  %.1f GS/s" % (defer, order, ctas, best * 1e3, B / best / 1e6), flush=True)


if __name__ == "__main__":
    main()
