"""Synthetic cf32 IQ for tests, golden vectors and the benchmark.

White Gaussian noise (sigma per component) + complex tones + a DC offset, values
kept inside (-1, 1).  The numpy generator is counter based (splitmix64 hash of the
absolute sample index -> Box-Muller) so any slice of the stream can be produced
independently and identically on every machine; tests and tests/golden rely on that.
"""
import numpy as np

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def _splitmix64(x):
    x = (x + np.uint64(0x9E3779B97F4A7C15)) & _M64
    z = x
    z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _M64
    z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _M64
    return z ^ (z >> np.uint64(31))


def gaussian_iq(start, count, seed, sigma=0.1):
    """complex64[count]: independent N(0, sigma^2) per component for absolute indices start..start+count."""
    with np.errstate(over="ignore"):
        n = np.arange(start, start + count, dtype=np.uint64)
        base = _splitmix64(n * np.uint64(2) + np.uint64(seed) * np.uint64(0xD1342543DE82EF95))
        h1 = _splitmix64(base)
        h2 = _splitmix64(base ^ np.uint64(0xA5A5A5A5A5A5A5A5))
    u1 = ((h1 >> np.uint64(11)).astype(np.float64) + 1.0) * (1.0 / 9007199254740993.0)   # (0,1)
    u2 = (h2 >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)           # [0,1)
    r = sigma * np.sqrt(-2.0 * np.log(u1))
    out = np.empty(count, np.complex64)
    out.real = (r * np.cos(2.0 * np.pi * u2)).astype(np.float32)
    out.imag = (r * np.sin(2.0 * np.pi * u2)).astype(np.float32)
    return out


def make_iq(start, count, sample_rate, seed=0x5D2B200, sigma=0.1, tones=(), dc=(0.01, -0.02)):
    """Noise + tones [(freq_hz, amplitude), ...] + DC, complex64[count], for absolute sample indices."""
    x = gaussian_iq(start, count, seed, sigma).astype(np.complex128)
    if tones:
        n = np.arange(start, start + count, dtype=np.float64)
        for f, a in tones:
            # reduce the phase modulo 1 turn in float64 before the trig call
            ph = np.mod(n * (f / sample_rate), 1.0)
            x += a * np.exp(2j * np.pi * ph)
    x += complex(dc[0], dc[1])
    return np.clip(x.real, -0.999, 0.999).astype(np.float32) + 1j * np.clip(x.imag, -0.999, 0.999).astype(np.float32)


def make_iq_torch(count, sample_rate, device, seed=0x5D2B200, sigma=0.1, tones=(), dc=(0.01, -0.02), chunk=1 << 24):
    """Same kind of signal generated on `device` with torch (bench input; not bit-identical to make_iq)."""
    import torch
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    out = torch.empty((count, 2), dtype=torch.float32, device=device)
    for s in range(0, count, chunk):
        n = min(chunk, count - s)
        blk = torch.randn((n, 2), generator=g, device=device, dtype=torch.float32) * sigma
        if tones:
            idx = torch.arange(s, s + n, device=device, dtype=torch.float64)
            for f, a in tones:
                ph = torch.remainder(idx * (f / sample_rate), 1.0) * (2.0 * np.pi)
                blk[:, 0] += (a * torch.cos(ph)).float()
                blk[:, 1] += (a * torch.sin(ph)).float()
        blk[:, 0] += dc[0]
        blk[:, 1] += dc[1]
        out[s:s + n] = blk.clamp_(-0.999, 0.999)
    return out
