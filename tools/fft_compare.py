#!/usr/bin/env python
"""Timing comparison only: this library's windowed FFT -> dB kernels against cuFFT (torch.fft) on the same batch.
cuFFT is NOT part of the product path; torch.fft here does the bare c2c transform (no window, no |X|^2, no log), so
it has strictly less work to do.  Prints one JSON line."""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from sdrplusplusbrown_b200 import binding, synth


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=65536)
    ap.add_argument("--frames", type=int, default=201)
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--prefetch", type=int, default=0, help="B200_OPT_FFT_DIRECT_PREFETCH of the one-pass kernel")
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    N, F = a.n, a.frames
    if binding.lib.b200_set_global_option(binding.OPT_FFT_DIRECT_PREFETCH, a.prefetch) != 0:
        raise RuntimeError(binding.lib.b200_last_error().decode())
    sr = float(N * 60)                      # interval == N: every sample belongs to exactly one frame
    fe = binding.FrontEnd(sr, fft_size=N, fft_rate=60.0, fft_window=1, max_block=N * F, device=0)
    x = synth.make_iq_torch(N * F, sr, dev, tones=[(0.1 * sr, 0.2)])
    s = torch.cuda.Stream()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(s):
        for _ in range(3):
            fe.process(x, N * F, stream=s)
        e0.record(s)
        for _ in range(a.reps):
            fe.process(x, N * F, stream=s)
        e1.record(s)
    torch.cuda.synchronize()
    ours = e0.elapsed_time(e1) / a.reps
    assert fe.frame_count() == F
    xc = torch.view_as_complex(x).reshape(F, N)
    with torch.cuda.stream(s):
        for _ in range(3):
            y = torch.fft.fft(xc, dim=1)
        e0.record(s)
        for _ in range(a.reps):
            y = torch.fft.fft(xc, dim=1)
        e1.record(s)
    torch.cuda.synchronize()
    cufft = e0.elapsed_time(e1) / a.reps
    byts = F * (8 * N + 4 * N)
    print(json.dumps({"fft_size": N, "frames": F, "prefetch": a.prefetch, "ours_ms": ours, "ours_what": "window + FFT + |X|^2 + 10log10 -> float dB rows",
                      "ours_algorithmic_GBps": byts / 1e6 / ours, "cufft_ms": cufft, "cufft_what": "bare c2c FFT (complex out)",
                      "ratio_ours_over_cufft": ours / cufft}))


if __name__ == "__main__":
    main()
