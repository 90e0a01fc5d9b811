"""Golden-vector cases for the IQ-front-end path.

Each case is a deterministic scenario (seeded synthetic IQ from sdrplusplusbrown_b200.synth)
run through an oracle (tests/oracle_lib.py).  tests/golden/make_golden.py runs every case
through oracle/_ref (the reference's own headers compiled in place) and stores the results;
tests compare the port oracle, _ref and the CUDA path against the stored arrays.
"""
import zlib

import numpy as np

import oracle_lib as ol
from sdrplusplusbrown_b200 import synth


def _crc(x):
    return np.array([zlib.crc32(np.ascontiguousarray(x).tobytes())], np.uint32)


def _blocks(x, size):
    return [x[i:i + size] for i in range(0, len(x), size)]


# name -> (in_sr, out_sr, bw, offset, total samples, block size, tone offset from centre)
VFO_GOLDEN = {
    "vfo_cfg1_nfm": (2.4e6, 50e3, 12.5e3, 100e3, 120000, 12000, 1000.0),
    "vfo_cfg2_usb": (20e6, 24e3, 2.8e3, 1.25e6, 1000000, 500000, 1400.0),
    "vfo_cfg3_nfm48": (61.44e6, 48e3, 12.5e3, -3.3e6, 1000000, 1000000, -2000.0),
    "vfo_cfg4_am": (122.88e6, 15e3, 10e3, 7.7e6, 1000000, 250000, 500.0),
    "vfo_cfg4_ssb": (122.88e6, 24e3, 2.8e3, -31.0e6, 1000000, 333333, 700.0),
    "vfo_cfg4_wfm": (122.88e6, 500e3, 150e3, -20e6, 300000, 100000, 30e3),
}


def vfo_input(name):
    in_sr, out_sr, bw, off, n, blk, tone = VFO_GOLDEN[name]
    return synth.make_iq(0, n, in_sr, tones=[(off + tone, 0.05), (off + 40 * bw, 0.3)])


def _run_vfo(orc, name):
    in_sr, out_sr, bw, off, n, blk, tone = VFO_GOLDEN[name]
    x = vfo_input(name)
    v = ol.Vfo(orc, in_sr, out_sr, bw, off)
    outs, counts, phases = [], [], []
    for b in _blocks(x, blk):
        phases.append(v.phase())
        y = v.process(b)
        outs.append(y)
        counts.append(len(y))
    phases.append(v.phase())
    return dict(out=np.concatenate(outs).astype(np.complex64), counts=np.array(counts, np.int32),
                phases=np.array(phases, np.complex64), input_crc=_crc(x))


def preproc_input():
    return synth.make_iq(0, 24000, 8e6, tones=[(1e6, 0.1)], dc=(0.05, -0.08))


def _run_preproc(orc, name):
    x = preproc_input()
    res = dict(input_crc=_crc(x))
    for tag, (dec, dc, inv) in dict(dc=(1, 1, 0), d4_dc_inv=(4, 1, 1), d16=(16, 0, 0)).items():
        p = ol.Preproc(orc, 8e6, dec, dc, inv)
        res["out_" + tag] = np.concatenate([p.process(b) for b in _blocks(x, 7000)]).astype(np.complex64)
    return res


FFT_GOLDEN = {
    # name -> (sample_rate, N, fft_rate, window, zoom (offset, width, outSize), wfMin, wfMax)
    "fft_cfg1_8192_nuttall": (2.4e6, 8192, 20.0, 2, (1024, 6144, 1000), -110.0, -20.0),
    "fft_cfg2_65536_blackman": (20e6, 65536, 60.0, 1, (0, 65536, 1333), -120.0, -10.0),
    "fft_nz_lt_n_rect": (48e3, 4096, 20.0, 0, (100, 3000, 640), -100.0, 0.0),
    "fft_cfg4_1048576_blackman": (122.88e6, 1 << 20, 60.0, 1, (0, 1 << 20, 1333), -120.0, -10.0),
}


def fft_input(name):
    sr, N, rate, wt, zoom, lo, hi = FFT_GOLDEN[name]
    import oracle_lib
    interval = int(round(sr / rate))
    nz = min(interval, N)
    return synth.make_iq(0, nz, sr, tones=[(0.11 * sr, 0.2), (-0.31 * sr, 0.01)]), nz


def _run_fft(orc, name):
    sr, N, rate, wt, zoom, lo, hi = FFT_GOLDEN[name]
    x, nz = fft_input(name)
    skip, nz2 = orc.reshape_params(sr, N, rate)
    assert nz2 == nz
    w = orc.fft_window(wt, nz)
    db = orc.fft_frame(x, w, N)
    zo = orc.do_zoom(zoom[0], zoom[1], db, zoom[2])
    sid, uf = orc.zoom_indices(zoom[0], zoom[1], N, zoom[2])
    pal = orc.palette_indices(zo, lo, hi)
    stride = max(1, N // 4096)
    return dict(db_sub=db[::stride].copy(), db_peak=np.array([db.max(), float(np.argmax(db))], np.float32),
                zoom=zo, zoom_sid=sid, zoom_uf=uf, palette=pal, window_sub=w[::max(1, nz // 1024)].copy(),
                reshape=np.array([skip, nz], np.int32), input_crc=_crc(x))


def demod_input(sr):
    x = synth.make_iq(0, 6000, sr, sigma=0.02, tones=[(900.0, 0.3), (-1500.0, 0.05)])
    x[3000:3004] = 0
    return x


def _run_demod(orc, name):
    if name == "demod_nfm":
        sr = 50e3
        d = ol.Fm(orc, sr, 12.5e3, True, False)
    elif name == "demod_usb":
        sr = 24e3
        d = ol.Ssb(orc, 0, 2800.0, sr, 50.0 / sr, 5.0 / sr)
    elif name == "demod_lsb":
        sr = 24e3
        d = ol.Ssb(orc, 1, 2800.0, sr, 50.0 / sr, 5.0 / sr)
    elif name == "demod_am":
        sr = 15e3
        d = ol.Am(orc, 1, 10e3, 50.0 / sr, 5.0 / sr, 100.0 / sr, sr)
    elif name == "demod_am_carrier":
        sr = 15e3
        d = ol.Am(orc, 0, 10e3, 50.0 / sr, 5.0 / sr, 100.0 / sr, sr)
    x = demod_input(sr)
    out = np.concatenate([d.process(b) for b in _blocks(x, 2500)])
    assert np.array_equal(out[:, 0], out[:, 1])
    return dict(mono=out[:, 0].copy(), input_crc=_crc(x))


def _run_taps(orc, name):
    res = {}
    for tag, (cut, tr, sr) in dict(nfm=(6250.0, 625.0, 50e3), usb=(1400.0, 140.0, 24e3), am=(5000.0, 500.0, 15e3),
                                   wfm=(75e3, 7.5e3, 500e3), resamp_4_5=(24e3, 2.4e3, 240e3)).items():
        res["lp_" + tag] = orc.lowpass(cut, tr, sr)
    for k in (1, 5, 10, 13):
        for s, (d, t) in enumerate(orc.decim_plan(1 << k)):
            res["plan%d_s%d_d%d" % (1 << k, s, d)] = t
    return res


def sigdet_input(sr, rows, seed=11):
    """Noise plus three voice-like groups of spectral lines (harmonics 110 / 170 / 260 Hz apart, i.e. 11 / 17 / 26 of the
    detector's 10 Hz bins, strongest line last) whose levels wander from row to row, and one plain carrier."""
    n = int(sr * (1 / 10.0)) * rows
    rng = np.random.default_rng(seed)
    t = np.arange(n, dtype=np.float64) / sr
    x = synth.gaussian_iq(0, n, 0x5D2B200 + seed, 0.05).astype(np.complex128)
    for (f0, df, k) in ((-0.31 * sr, 110.0, 9), (0.07 * sr, 170.0, 7), (0.33 * sr, 260.0, 6)):
        for h in range(k):
            a = 0.004 * (1.0 + h) * (1.0 + 0.5 * np.sin(2 * np.pi * (0.7 + 0.1 * h) * t + rng.uniform(0, 6.28)))
            x += a * np.exp(2j * np.pi * ((f0 + h * df) * t + rng.uniform(0, 1)))
    x += 0.05 * np.exp(2j * np.pi * (-0.12 * sr) * t)
    return x.astype(np.complex64)


def _run_sigdet(orc, name):
    sr, rows = 51200.0, 25
    x = sigdet_input(sr, rows)
    d = ol.SigDet(orc, sr)
    assert d.n == 5120
    done = 0
    for b in _blocks(x, 7001):                      # ragged against the 5120-sample rows
        done = d.process(b)
    hist = d.history()
    row0 = ol.sigdet_row(orc, x[:d.n])
    cand0 = ol.sigdet_candidates(orc, row0)
    marked = np.flatnonzero(hist != -20.0)
    return dict(rows=np.array([done, hist.shape[0]], np.int32), row0_sub=row0[::4].copy(), cand0=cand0,
                smoothed=d.result(), hist_marked=marked.astype(np.int32), hist_values=hist.ravel()[marked].copy(),
                input_crc=_crc(x))


def logmmse_input(sr, seconds, seed=1):
    """white noise with an amplitude-modulated carrier and a weaker keyed one: what the noise reduction is meant to clean up"""
    n = int(sr * seconds)
    t = np.arange(n, dtype=np.float64) / sr
    x = synth.gaussian_iq(0, n, 0x5D2B200 + 77 + seed, 0.05).astype(np.complex128)
    x += 0.3 * np.exp(2j * np.pi * (0.0625 * sr) * t) * (1 + 0.5 * np.sin(2 * np.pi * 3 * t))
    x += 0.02 * np.exp(2j * np.pi * (-0.21 * sr) * t) * (np.sin(2 * np.pi * 1.5 * t) > 0)
    return x.astype(np.complex64)


def _run_logmmse(orc, name):
    sr = 24000
    x = logmmse_input(sr, 2.5)
    d = ol.LogMmse(orc, sr)
    outs = [d.process(b) for b in _blocks(x, 1000)]
    y = np.concatenate(outs)
    return dict(counts=np.array([len(o) for o in outs], np.int32), out_sub=y[::16].copy(), out_crc=_crc(y), noise=d.noise()[::2].copy(),
                input_crc=_crc(x))


def _run_fftcomp(orc, name):
    sr = 48000
    x = logmmse_input(sr, 2.0, seed=6)
    c = ol.FftCompressor(orc, sr, 10.0)
    assert c.n == 2048
    outs = [c.process(b) for b in _blocks(x, c.n)]
    y = np.concatenate(outs)
    return dict(counts=np.array([len(o) for o in outs], np.int32), out_first=outs[12].copy(), out_crc=_crc(y),
                zeroed=np.array([int(np.count_nonzero(o == 0)) for o in outs if len(o)], np.int32), noise_figure=c.noise_figure(),
                input_crc=_crc(x))


CASES = {}
for _n in VFO_GOLDEN:
    CASES[_n] = _run_vfo
CASES["preproc"] = _run_preproc
for _n in FFT_GOLDEN:
    CASES[_n] = _run_fft
for _n in ("demod_nfm", "demod_usb", "demod_lsb", "demod_am", "demod_am_carrier"):
    CASES[_n] = _run_demod
CASES["taps"] = _run_taps
CASES["sigdet_51200"] = _run_sigdet
CASES["logmmse_24000"] = _run_logmmse
CASES["fftcomp_48000"] = _run_fftcomp


def run_case(orc, name):
    return CASES[name](orc, name)
