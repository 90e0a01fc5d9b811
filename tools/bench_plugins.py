#!/usr/bin/env python3
"""Rates of the two front-end plug-ins on one GPU, device-resident input, CUDA events, with the oracle
(CPU, one thread, the reference's own code when oracle/_ref is there) timed beside them on a bounded sample.

    python tools/bench_plugins.py > profiles/r02_plugins.jsonl

detector: dsp::detector::SignalDetector (one dB row of sampleRate / 10 points per 0.1 s of signal + line candidates +
history average); logmmse: dsp::IFNRLogMMSE (20 ms frames, hop 10 ms).  x_real_time = seconds of signal per second."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    import torch
    from sdrplusplusbrown_b200 import binding, synth
    import oracle_lib as ol
    orc = ol.Oracle("ref") if os.path.exists(ol.REF_SO) else ol.Oracle("port")
    dev = torch.device("cuda", 0)
    s = torch.cuda.Stream(device=dev)

    def timed(fn, reps):
        fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(s):
            e0.record(s)
            for _ in range(reps):
                fn()
            e1.record(s)
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    for sr in (2.4e6, 20e6, 61.44e6, 122.88e6):
        n = int(sr * (1 / 10.0))
        rows = 10
        x = synth.make_iq_torch(n * rows, sr, dev, tones=[(0.11 * sr, 0.05)])
        d = binding.SignalDetector(sr)
        for _ in range(3):                       # past the 20 rows the detection waits for: the timed blocks aggregate too
            d.process(x, stream=s)
        ms = timed(lambda: d.process(x, stream=s), 3)
        line = {"plugin": "signal detector", "sample_rate": sr, "row_points": n, "rows_per_block": rows, "ms_per_block": ms,
                "ms_per_row": ms / rows, "x_real_time": (rows * 0.1) / (ms * 1e-3)}
        if sr <= 20e6:
            xh = x[:n].cpu().numpy().view(np.complex64).reshape(-1)
            od = ol.SigDet(orc, sr)
            t0 = time.perf_counter()
            od.process(xh)
            line["cpu_oracle"] = {"kind": orc.kind, "cores": 1, "ms_per_row": (time.perf_counter() - t0) * 1e3, "sample": "one row"}
        print(json.dumps(line), flush=True)
        del d, x

    for sr in (2400000, 20000000, 61440000):
        blk = sr // 10
        x = synth.make_iq_torch(blk * 4, float(sr), dev, tones=[(0.11 * sr, 0.05)])
        d = binding.LogMmse(sr, max_block=blk)
        blocks = [x[i * blk:(i + 1) * blk] for i in range(4)]
        p, cnt = binding.C.c_void_p(), binding.C.c_int()

        def feed():
            for b in blocks:
                binding._chk(binding.lib.b200_logmmse_process(d.h, b.data_ptr(), blk, binding.C.byref(p), binding.C.byref(cnt), s.cuda_stream))
        ms = timed(feed, 3)
        line = {"plugin": "logmmse", "sample_rate": sr, "fft_size": d.fft_size(), "ms_per_100ms_block": ms / 4, "x_real_time": 0.4 / (ms * 1e-3)}
        if sr <= 2400000:
            xh = x.cpu().numpy().view(np.complex64).reshape(-1)
            od = ol.LogMmse(orc, sr)
            od.process(xh)                        # sampling + first frames
            t0 = time.perf_counter()
            od.process(xh[:blk * 2])
            line["cpu_oracle"] = {"kind": orc.kind, "cores": 1, "ms_per_100ms_block": (time.perf_counter() - t0) * 1e3 / 2, "sample": "two 100 ms blocks"}
        print(json.dumps(line), flush=True)
        del d, x

    for sr in (240000, 2400000, 10000000):
        d = binding.FftCompressor(sr, 10.0, max_block=1 << 20)
        n = d.n
        x = synth.make_iq_torch(n * 4, float(sr), dev, tones=[(0.11 * sr, 0.05)])
        blocks = [x[i * n:(i + 1) * n] for i in range(4)]
        p, cnt = binding.C.c_void_p(), binding.C.c_int()

        def feed_c():
            for b in blocks:
                binding._chk(binding.lib.b200_fftcomp_process(d.h, b.data_ptr(), n, binding.C.byref(p), binding.C.byref(cnt), s.cuda_stream))
        for _ in range(3):
            feed_c()                              # past the ten frames of delay
        ms = timed(feed_c, 3)
        line = {"plugin": "server spectral compressor", "sample_rate": sr, "fft_size": n, "ms_per_frame": ms / 4, "x_real_time": (4 * n / sr) / (ms * 1e-3)}
        if sr <= 2400000:
            xh = x.cpu().numpy().view(np.complex64).reshape(-1)
            oc = ol.FftCompressor(orc, sr, 10.0)
            for _ in range(3):
                for i in range(4):
                    oc.process(xh[i * n:(i + 1) * n])
            t0 = time.perf_counter()
            for i in range(4):
                oc.process(xh[i * n:(i + 1) * n])
            line["cpu_oracle"] = {"kind": orc.kind, "cores": 1, "ms_per_frame": (time.perf_counter() - t0) * 1e3 / 4, "sample": "four frames"}
        print(json.dumps(line), flush=True)
        del d, x


if __name__ == "__main__":
    main()
