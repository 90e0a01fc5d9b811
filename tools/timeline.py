#!/usr/bin/env python
"""Kernel timeline of one bench step (diagnosis only; never a reported number).

Runs the bench.py workload (cfg2 front end: FFT branch + 1 USB VFO + waterfall row reduction) a few times under
torch.profiler (CUPTI activity records see every kernel of the process, the library's included) and prints, for the
last step, each kernel / copy with its stream, start offset and duration, so that overlap between the FFT branch and
the channelizer branch and the gaps between small launches can be read directly.

    python tools/timeline.py [--batch N] [--steps K] [--vfos V] [--no-fft] [--no-vfo]
"""
import argparse
import json
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=1 << 26)
    ap.add_argument("--steps", type=int, default=4)
    ap.add_argument("--vfos", type=int, default=1)
    ap.add_argument("--no-fft", action="store_true")
    ap.add_argument("--no-vfo", action="store_true")
    ap.add_argument("--demod", action="store_true")
    ap.add_argument("--cfg", default="cfg2")
    ap.add_argument("--attach", action="store_true", help="waterfall attached to the FFT branch, as bench.py runs it")
    ap.add_argument("--order", type=int, default=0, help="B200_OPT_BRANCH_ORDER")
    ap.add_argument("--fft-ctas", type=int, default=0, help="B200_OPT_FFT_CTAS_PER_SM")
    ap.add_argument("--defer", type=int, default=0, help="B200_OPT_DEFER_TAIL")
    ap.add_argument("--room", type=int, default=0, help="B200_OPT_LEAVE_ROOM")
    args = ap.parse_args()
    import torch
    from torch.profiler import ProfilerActivity, profile
    from sdrplusplusbrown_b200 import binding, configs, synth

    cfg = configs.CONFIGS[args.cfg]
    sr = cfg["sample_rate"]
    B = args.batch
    dev = torch.device("cuda", 0)
    fft_rate = 0.05 if args.no_fft else cfg["fft_rate"]
    fe = binding.FrontEnd(sr, fft_size=cfg["fft_size"], fft_rate=fft_rate, fft_window=cfg["window"], max_block=B, device=0)
    vl = configs.vfo_list(args.cfg)
    if not args.no_vfo:
        for i in range(args.vfos):
            mode, out_sr, bw, off = vl[i % len(vl)]
            v = fe.vfos.add(out_sr, bw, off + 1e3 * (i // len(vl)))
            if args.demod:
                fe.vfos.set_demod(v, binding.DEMOD_USB, 0, 50.0 / out_sr, 5.0 / out_sr, 0)
    width = 1333
    wf = binding.Waterfall(cfg["fft_size"], width, device=0)
    wf.set_view(0, cfg["fft_size"])
    wf.set_range(-120.0, 0.0)
    max_frames = B // int(round(sr / cfg["fft_rate"])) + 2
    d_zoom = torch.empty((max_frames, width), dtype=torch.float32, device=dev)
    d_rgba = torch.empty((max_frames, width), dtype=torch.int32, device=dev)
    x = synth.make_iq_torch(B, sr, dev, tones=[(3.3e6, 0.2)])
    s = torch.cuda.Stream(device=dev, priority=-1)

    fe.vfos.set_option(binding.OPT_BRANCH_ORDER, args.order)
    fe.vfos.set_option(binding.OPT_FFT_CTAS_PER_SM, args.fft_ctas)
    fe.vfos.set_option(binding.OPT_DEFER_TAIL, args.defer)
    fe.vfos.set_option(binding.OPT_LEAVE_ROOM, args.room)
    if args.attach:
        fe.attach_waterfall(wf, d_zoom.data_ptr(), d_rgba.data_ptr(), max_frames)

    def step():
        with torch.cuda.stream(s):
            fe.process(x, B, stream=s)
            nf = fe.frame_count()
            if nf and not args.attach:
                wf.push(fe.frame_ptr(0), nf, d_zoom.data_ptr(), d_rgba.data_ptr(), stream=s)

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    marks = []
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for i in range(args.steps):
            step()
        torch.cuda.synchronize()
    path = os.path.join(tempfile.gettempdir(), "b200_timeline.json")
    prof.export_chrome_trace(path)
    tr = json.load(open(path))
    ev = [e for e in tr["traceEvents"] if e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset") and "dur" in e]
    ev.sort(key=lambda e: e["ts"])
    if not ev:
        print("no GPU activity records (CUPTI unavailable?)")
        return
    # split into steps: a step begins at every rotator-checkpoint kernel (or FFT kernel when there is no VFO)
    per = len(ev) // args.steps
    last = ev[-2 * per:] if args.defer else ev[-per:]      # deferred tail: two steps, to see what overlaps what
    t0 = last[0]["ts"]
    print("%-58s %7s %10s %9s" % ("kernel / copy", "stream", "start us", "dur us"))
    for e in last:
        name = e["name"]
        if len(name) > 58:
            name = name[:55] + "..."
        print("%-58s %7s %10.1f %9.1f" % (name, e.get("args", {}).get("stream", "?"), e["ts"] - t0, e["dur"]))
    span = last[-1]["ts"] + last[-1]["dur"] - t0
    busy = sum(e["dur"] for e in last)
    print("step span %.1f us, sum of durations %.1f us, %d records/step" % (span, busy, per))
    whole = ev[-1]["ts"] + ev[-1]["dur"] - ev[0]["ts"]
    print("all %d steps: %.1f us/step" % (args.steps, whole / args.steps))


if __name__ == "__main__":
    main()
