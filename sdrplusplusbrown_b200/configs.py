"""The BASELINE.json configurations as data, shared by tests and bench.py."""

# radio-module IF rates / default bandwidths (decoder_modules/radio/src/demodulators/*.h:
# nfm.h:65-67, am.h:76-78, usb.h:70-72, wfm.h:261)
MODES = {
    "NFM": dict(out_sr=50000.0, bw=12500.0),
    "NFM48": dict(out_sr=48000.0, bw=12500.0),   # BASELINE config 3 quotes 61.44M -> 48k
    "AM": dict(out_sr=15000.0, bw=10000.0),
    "USB": dict(out_sr=24000.0, bw=2800.0),
    "LSB": dict(out_sr=24000.0, bw=2800.0),
    "WFM": dict(out_sr=500000.0, bw=150000.0),
}

WINDOW_RECT, WINDOW_BLACKMAN, WINDOW_NUTTALL = 0, 1, 2

CONFIGS = {
    # configs[0]: baseband file stand-in, 2.4 MS/s, 8192-pt FFT (Nuttall, 20 fps), one NFM 12.5 kHz VFO
    "cfg1": dict(sample_rate=2.4e6, fft_size=8192, fft_rate=20.0, window=WINDOW_NUTTALL,
                 vfos=[("NFM", 100e3)]),
    # configs[1]: 20 MS/s, 65536-pt Blackman at 60 fps, one USB VFO  (bench workload at N=1)
    "cfg2": dict(sample_rate=20e6, fft_size=65536, fft_rate=60.0, window=WINDOW_BLACKMAN,
                 vfos=[("USB", 1.25e6)]),
    # configs[2]: 61.44 MS/s, 256 NFM VFOs (61.44M -> 48k), sharded over GPUs
    "cfg3": dict(sample_rate=61.44e6, fft_size=65536, fft_rate=60.0, window=WINDOW_BLACKMAN,
                 vfos="nfm_grid_256"),
    # configs[3]: 122.88 MS/s, 1024 mixed AM/SSB/WFM VFOs + 2^20-pt FFT
    "cfg4": dict(sample_rate=122.88e6, fft_size=1 << 20, fft_rate=60.0, window=WINDOW_BLACKMAN,
                 vfos="mixed_1024"),
}


def vfo_list(cfg_name, count=None, seed=1234):
    """Expand a config's VFO description into [(mode, out_sr, bw, offset_hz), ...]."""
    import numpy as np
    cfg = CONFIGS[cfg_name]
    spec = cfg["vfos"]
    sr = cfg["sample_rate"]
    rng = np.random.default_rng(seed)
    out = []
    if isinstance(spec, list):
        for mode, off in spec:
            m = MODES[mode]
            out.append((mode, m["out_sr"], m["bw"], float(off)))
    elif spec == "nfm_grid_256":
        n = count or 256
        # 200 kHz grid inside +-28 MHz, jittered by U(-5k, 5k)
        grid = (np.arange(n) - n / 2.0 + 0.5) * 200e3
        jit = rng.uniform(-5e3, 5e3, n)
        for i in range(n):
            out.append(("NFM48", 48000.0, 12500.0, float(grid[i] + jit[i])))
    elif spec == "mixed_1024":
        n = count or 1024
        span = 0.9 * sr
        offs = (np.arange(n) - n / 2.0 + 0.5) * (span / n) + rng.uniform(-2e3, 2e3, n)
        for i in range(n):
            # 512 AM + 384 SSB + 128 WFM per 1024, interleaved over the band.  (Sharding by i % 8 would put one mode on each
            # rank -- all WFM channels on rank 7; tools/bench_bank_dist.py deals the bank out mode by mode instead.)
            k = i % 8
            mode = "WFM" if k == 7 else ("USB" if k in (1, 3, 5) else "AM")
            m = MODES[mode]
            out.append((mode, m["out_sr"], m["bw"], float(offs[i])))
    if count is not None:
        out = out[:count]
    return out
