"""CPU tests of the oracle itself: the self-contained port (oracle/oracle_port.cpp) is pinned
bit-for-bit against the reference's own headers compiled in place (oracle/_ref) whenever that
library is available, and both are pinned against the committed vectors in tests/golden/."""
import os

import numpy as np
import pytest

import oracle_lib as ol
from sdrplusplusbrown_b200 import synth

PORT = ol.Oracle("port")
REF = ol.Oracle("ref") if ol.have_ref() else None
needs_ref = pytest.mark.skipif(REF is None, reason="oracle/_ref not built and /root/reference absent")

VFO_CASES = [
    # in_sr, out_sr, bw, offset, n
    (2.4e6, 50e3, 12.5e3, 100e3, 60000),        # cfg1 NFM: /32 then 2/3, 304-tap FIR
    (20e6, 24e3, 2.8e3, 1.25e6, 300000),        # cfg2 USB: /512 then 8000/13021
    (61.44e6, 48e3, 12.5e3, -3.3e6, 300000),    # cfg3 NFM: /1024 then 4/5
    (122.88e6, 15e3, 10e3, 7.7e6, 400000),      # cfg4 AM: /8192 DECIM_ONLY
    (122.88e6, 500e3, 150e3, -20e6, 100000),    # cfg4 WFM: /128 then 25/48
    (48e3, 48e3, 48e3, 1e3, 5000),              # NONE mode, filter bypassed (bw == outSR)
    (24e3, 48e3, 12e3, 500.0, 5000),            # RESAMP_ONLY (interpolating)
]


def _split(x, sizes):
    out, i, k = [], 0, 0
    while i < len(x):
        n = sizes[k % len(sizes)]
        out.append(x[i:i + n])
        i += n
        k += 1
    return out


@needs_ref
def test_decim_plans_port_equals_ref():
    for k in range(1, 14):
        p, r = PORT.decim_plan(1 << k), REF.decim_plan(1 << k)
        assert len(p) == len(r) > 0
        for (dp, tp), (dr, tr) in zip(p, r):
            assert dp == dr
            assert np.array_equal(tp.view(np.uint32), tr.view(np.uint32))


@needs_ref
def test_taps_and_windows_port_equals_ref():
    for args in [(6250.0, 625.0, 50e3), (1400.0, 140.0, 24e3), (24000.0, 2400.0, 240000.0), (0.25, 0.1, 1.0)]:
        a, b = PORT.lowpass(*args), REF.lowpass(*args)
        assert len(a) == len(b) and np.array_equal(a.view(np.uint32), b.view(np.uint32))
    for wt in (0, 1, 2):
        for nz in (8192, 12000, 65536):
            a, b = PORT.fft_window(wt, nz), REF.fft_window(wt, nz)
            assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
    for sr, size, rate in [(2.4e6, 8192, 20.0), (20e6, 65536, 60.0), (122.88e6, 1 << 20, 60.0), (48e3, 65536, 20.0)]:
        assert PORT.reshape_params(sr, size, rate) == REF.reshape_params(sr, size, rate)


@needs_ref
@pytest.mark.parametrize("case", VFO_CASES)
def test_vfo_port_equals_ref_bitwise(case):
    in_sr, out_sr, bw, off, n = case
    x = synth.make_iq(0, n, in_sr, tones=[(off + 300.0, 0.05)])
    vp, vr = ol.Vfo(PORT, in_sr, out_sr, bw, off), ol.Vfo(REF, in_sr, out_sr, bw, off)
    assert vp.describe() == vr.describe()
    assert np.array_equal(vp.resamp_taps().view(np.uint32), vr.resamp_taps().view(np.uint32))
    assert np.array_equal(vp.chan_taps().view(np.uint32), vr.chan_taps().view(np.uint32))
    for blk in _split(x, [12000, 7777, 333, 65536]):
        a, b = vp.process(blk), vr.process(blk)
        assert len(a) == len(b)
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
        assert vp.phase() == vr.phase()


def test_vfo_counts_are_chunk_invariant():
    # 3 M samples 61.44M -> 48k give the same total count for any block partition
    in_sr, out_sr, bw, off = 61.44e6, 48e3, 12.5e3, 2.0e6
    n = 600000
    x = synth.make_iq(0, n, in_sr)
    totals = []
    for sizes in ([n], [65536], [12000], [7777], [333, 100001]):
        v = ol.Vfo(PORT, in_sr, out_sr, bw, off)
        totals.append(sum(len(v.process(b)) for b in _split(x, sizes)))
    assert len(set(totals)) == 1


@needs_ref
def test_vfo_reconfigure_port_equals_ref():
    in_sr = 2.4e6
    x = synth.make_iq(0, 96000, in_sr, tones=[(100e3, 0.1)])
    vp, vr = ol.Vfo(PORT, in_sr, 50e3, 12.5e3, 100e3), ol.Vfo(REF, in_sr, 50e3, 12.5e3, 100e3)
    blocks = _split(x, [12000])
    for i, blk in enumerate(blocks):
        if i == 2:
            vp.set_offset(-50e3); vr.set_offset(-50e3)
        if i == 3:
            vp.set_bandwidth(8000.0); vr.set_bandwidth(8000.0)
        if i == 4:
            vp.set_out_samplerate(24e3, 2.8e3); vr.set_out_samplerate(24e3, 2.8e3)
        if i == 5:
            vp.set_bandwidth(24e3); vr.set_bandwidth(24e3)   # filter bypass
        if i == 6:
            vp.reset(); vr.reset()
        if i == 7:
            vp.set_in_samplerate(1.2e6); vr.set_in_samplerate(1.2e6)
        a, b = vp.process(blk), vr.process(blk)
        assert len(a) == len(b), i
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32)), i


@needs_ref
def test_preproc_dcblock_decim_port_equals_ref():
    x = synth.make_iq(0, 50000, 8e6)
    for dec, dc, inv in [(1, 1, 0), (4, 1, 1), (16, 0, 0), (2, 1, 0), (1, 0, 1)]:
        p, r = ol.Preproc(PORT, 8e6, dec, dc, inv), ol.Preproc(REF, 8e6, dec, dc, inv)
        for blk in _split(x, [12000, 5001]):
            a, b = p.process(blk), r.process(blk)
            assert len(a) == len(b) and np.array_equal(a.view(np.uint32), b.view(np.uint32))
    for ratio in (2, 8, 64, 2048):
        p, r = ol.Decim(PORT, ratio), ol.Decim(REF, ratio)
        for blk in _split(x, [20000, 4999]):
            a, b = p.process(blk), r.process(blk)
            assert len(a) == len(b) and np.array_equal(a.view(np.uint32), b.view(np.uint32))


@needs_ref
def test_demods_port_equals_ref():
    x = synth.make_iq(0, 20000, 50e3, sigma=0.02, tones=[(1000.0, 0.3)])
    x[5000:5010] = 0   # exercise the AGC's inAmp == 0 branch
    pairs = [
        (ol.Fm(PORT, 50e3, 12.5e3), ol.Fm(REF, 50e3, 12.5e3)),
        (ol.Fm(PORT, 50e3, 12.5e3, False, False), ol.Fm(REF, 50e3, 12.5e3, False, False)),
        (ol.Fm(PORT, 50e3, 12.5e3, True, True), ol.Fm(REF, 50e3, 12.5e3, True, True)),
        (ol.Fm(PORT, 50e3, 12.5e3, False, True), ol.Fm(REF, 50e3, 12.5e3, False, True)),
        (ol.Ssb(PORT, 0, 2800.0, 24e3, 50.0 / 24e3, 5.0 / 24e3), ol.Ssb(REF, 0, 2800.0, 24e3, 50.0 / 24e3, 5.0 / 24e3)),
        (ol.Ssb(PORT, 1, 2800.0, 24e3, 50.0 / 24e3, 5.0 / 24e3), ol.Ssb(REF, 1, 2800.0, 24e3, 50.0 / 24e3, 5.0 / 24e3)),
        (ol.Am(PORT, 1, 10e3, 50.0 / 15e3, 5.0 / 15e3, 100.0 / 15e3, 15e3), ol.Am(REF, 1, 10e3, 50.0 / 15e3, 5.0 / 15e3, 100.0 / 15e3, 15e3)),
        (ol.Am(PORT, 0, 10e3, 50.0 / 15e3, 5.0 / 15e3, 100.0 / 15e3, 15e3), ol.Am(REF, 0, 10e3, 50.0 / 15e3, 5.0 / 15e3, 100.0 / 15e3, 15e3)),
    ]
    for p, r in pairs:
        for blk in _split(x, [4800, 333, 7000]):
            a, b = p.process(blk), r.process(blk)
            assert a.shape == b.shape and np.array_equal(a.view(np.uint32), b.view(np.uint32)), type(p)


def test_fft_frame_against_numpy_float64():
    # second opinion: numpy float64 FFT of the fp32-windowed frame
    for N, nz, wt in [(8192, 8192, 2), (65536, 65536, 1), (1024, 600, 0)]:
        x = synth.make_iq(0, nz, 2.4e6, tones=[(300e3, 0.2)])
        w = PORT.fft_window(wt, nz)
        db, spec = PORT.fft_frame(x, w, N, want_spectrum=True)
        xw = np.zeros(N, np.complex128)
        xw[:nz] = (x.real * w).astype(np.float32) + 1j * (x.imag * w).astype(np.float32)
        ref = np.fft.fft(xw)
        err = np.sqrt(np.mean(np.abs(spec - ref) ** 2) / np.mean(np.abs(ref) ** 2))
        assert err < 1e-6      # fp32 narrowing only
        ref_db = 10 * np.log10(np.abs(ref / N) ** 2 + 1e-300)
        assert np.max(np.abs(db - ref_db)) < 1e-3
    # an exactly-zero bin reads the clamped floor, not -inf (VOLK generic log2 clamp)
    db = PORT.fft_frame(np.zeros(64, np.complex64), np.ones(64, np.float32), 64)
    assert np.all(np.isfinite(db)) and np.allclose(db, -382.3, atol=0.1)


def test_zoom_and_palette_properties():
    rng = np.random.default_rng(3)
    row = rng.uniform(-120, -20, 65536).astype(np.float32)
    out = PORT.do_zoom(1000, 40000, row, 1333)
    s, u = PORT.zoom_indices(1000, 40000, 65536, 1333)
    assert s[0] == 1000 and np.all(np.diff(s) >= 0) and np.all(u >= 1)
    ref = np.array([row[a:a + b].max() for a, b in zip(s, u)], np.float32)
    assert np.array_equal(out, ref)
    idx = PORT.palette_indices(out, -100.0, -30.0)
    assert idx.min() >= 0 and idx.max() <= 999
    lut = PORT.make_palette(np.array([[0, 0, 32], [0, 0, 255], [255, 255, 0], [255, 0, 0]], np.float32))
    assert lut.shape == (1000,) and np.all((lut >> 24) == 255)


GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _golden_files():
    if not os.path.isdir(GOLDEN_DIR):
        return []
    return sorted(f for f in os.listdir(GOLDEN_DIR) if f.endswith(".npz"))


@pytest.mark.parametrize("fname", _golden_files())
def test_oracles_match_golden(fname):
    import golden_cases
    g = np.load(os.path.join(GOLDEN_DIR, fname))
    for orc in [PORT] + ([REF] if REF is not None else []):
        got = golden_cases.run_case(orc, fname[:-4])
        assert set(got.keys()) == set(g.files)
        for k in g.files:
            a, b = np.asarray(got[k]), g[k]
            assert a.shape == b.shape and a.dtype == b.dtype, (fname, k, orc.kind)
            assert a.tobytes() == b.tobytes(), (fname, k, orc.kind)


def test_vfo_signal_info_against_numpy():
    rng = np.random.default_rng(5)
    row = rng.uniform(-110, -70, 65536).astype(np.float32)
    row[40000:40050] += 50
    wb, center, bw = 20e6, 2.2e6, 30e3
    st, snr = PORT.vfo_signal_info(row, wb, center, bw)
    idx = lambda f: int(np.clip(int(f / (wb / 2) * 32768 + 32768), 0, 65536))
    a, b, c, d = idx(center - bw), idx(center - bw / 2), idx(center + bw / 2), idx(center + bw)
    side = np.concatenate([row[a:b], row[c + 1:d]]).astype(np.float64)
    avg = side.mean()
    srt = np.sort(side.astype(np.float32))
    lower = len(srt) // 4
    qavg = srt[srt <= srt[lower]].astype(np.float64).sum() / lower
    mx = row[b:c + 1].max()
    assert abs(st - (mx - (avg - qavg))) < 1e-3 and abs(snr - (mx - avg - (avg - qavg))) < 1e-3
    if REF is not None:
        assert REF.vfo_signal_info(row, wb, center, bw) == (st, snr)


@needs_ref
def test_radio_chain_blocks_port_equals_ref_bitwise():
    """IF chain (noise blanker, squelch), AF chain (stereo resampler, de-emphasis) and BroadcastFM (mono / stereo):
    the self-contained restatement against the reference's own headers, bit for bit, over ragged blocks."""
    rng = np.random.default_rng(21)
    x = ((rng.standard_normal(9000) + 1j * rng.standard_normal(9000)) * 0.05).astype(np.complex64)
    x[100:110] *= 40
    x[3000:3005] = 0
    for cfg in [(1, 500 / 50e3, 3.0, 0, 0.0), (0, 0.01, 1.0, 1, -30.0), (1, 0.02, 2.0, 1, -10.0), (1, 0.02, 2.0, 1, -40.0)]:
        p, r = ol.IfChain(PORT, *cfg), ol.IfChain(REF, *cfg)
        for i in range(0, len(x), 1700):
            assert p.process(x[i:i + 1700]).tobytes() == r.process(x[i:i + 1700]).tobytes(), cfg
    au = rng.standard_normal((9000, 2)).astype(np.float32)
    for a, b, tau in [(50e3, 48e3, 0.0), (24e3, 48e3, 50e-6), (15e3, 48e3, 75e-6), (250e3, 48e3, 50e-6), (48e3, 48e3, 50e-6), (50e3, 11025.0, 50e-6), (50e3, 12500.0, 0.0)]:
        p, r = ol.AfPost(PORT, a, b, tau), ol.AfPost(REF, a, b, tau)
        for i in range(0, 9000, 2500):
            yp, yr = p.process(au[i:i + 2500]), r.process(au[i:i + 2500])
            assert yp.shape == yr.shape and yp.tobytes() == yr.tobytes(), (a, b, tau)
    sr, n = 500e3, 40000
    t = np.arange(n) / sr
    L, R = 0.5 * np.sin(2 * np.pi * 1000 * t), 0.4 * np.sin(2 * np.pi * 2500 * t)
    mpx = 0.45 * (L + R) + 0.1 * np.sin(2 * np.pi * 19000 * t) + 0.45 * (L - R) * np.sin(2 * np.pi * 38000 * t)
    iq = np.exp(1j * 2 * np.pi * 75e3 * np.cumsum(mpx) / sr).astype(np.complex64)
    for stereo, lp in [(0, 1), (1, 1), (1, 0)]:
        p, r = ol.Wfm(PORT, 75e3, sr, stereo, lp), ol.Wfm(REF, 75e3, sr, stereo, lp)
        for i in range(0, n, 17000):
            assert p.process(iq[i:i + 17000]).tobytes() == r.process(iq[i:i + 17000]).tobytes(), (stereo, lp)


@pytest.mark.parametrize("bins", [9, 15, 31, 32])
def test_fmif_port_against_numpy_model_and_ref(bins):
    """FM IF noise reduction at the radio module's four presets (radio/src/main.cpp:11-16): the restatement against a
    float64 model (wherever the strongest bin is not a near-tie) and, bit for bit, against the reference's own header
    compiled on oracle/shim/fftw3.h."""
    rng = np.random.default_rng(100 + bins)
    n = 6000
    t = np.arange(n)
    x = (0.4 * np.exp(2j * np.pi * (0.07 * t + 0.02 * np.sin(2 * np.pi * t / 700.0) * 50))
         + 0.1 * (rng.standard_normal(n) + 1j * rng.standard_normal(n))).astype(np.complex64)
    x[2000:2100] = 0                                     # all bins equal: index 0 wins, output 0
    p = ol.FmIf(PORT, bins)
    r = ol.FmIf(REF, bins) if REF is not None else None
    hist = None
    for i in range(0, n, 1700):
        blk = x[i:i + 1700]
        y = p.process(blk)
        model, margin, hist = ol.fmif_model(blk, bins, hist)
        ok = margin > 1e-5
        assert ok.sum() > 0.98 * len(blk) - 100
        e = y[ok] - model[ok]
        assert np.sqrt(np.mean(np.abs(e) ** 2)) <= 1e-6 * max(1e-9, np.sqrt(np.mean(np.abs(model[ok]) ** 2))) + 1e-12
        if r is not None:
            assert r.process(blk).tobytes() == y.tobytes()
    assert not np.any(ol.FmIf(PORT, bins).process(np.zeros(10, np.complex64)))


def test_wfm_rds_branch_and_volume_port_equals_ref():
    """BroadcastFM with rdsOut (mono and stereo): audio and the 5 kS/s RDS baseband, port against the reference headers
    bit for bit; a 57 kHz subcarrier in the MPX must come out of the branch at 0 Hz.  audio::Volume likewise."""
    sr, n = 500e3, 60000
    t = np.arange(n) / sr
    mpx = 0.3 * np.sin(2 * np.pi * 1000 * t) + 0.1 * np.sin(2 * np.pi * 19000 * t) + 0.05 * np.cos(2 * np.pi * 57000 * t)
    iq = np.exp(1j * 2 * np.pi * 75e3 * np.cumsum(mpx) / sr).astype(np.complex64)
    for stereo, lp in [(0, 1), (1, 1)]:
        p = ol.WfmRds(PORT, 75e3, sr, stereo, lp)
        r = ol.WfmRds(REF, 75e3, sr, stereo, lp) if REF is not None else None
        plain = ol.Wfm(PORT, 75e3, sr, stereo, lp)
        rds_all = []
        for i in range(0, n, 17001):
            a, rds = p.process(iq[i:i + 17001])
            assert a.tobytes() == plain.process(iq[i:i + 17001]).tobytes()      # the branch does not touch the audio
            if r is not None:
                a2, rds2 = r.process(iq[i:i + 17001])
                assert a.tobytes() == a2.tobytes() and rds.tobytes() == rds2.tobytes(), (stereo, lp)
            rds_all.append(rds)
        rds = np.concatenate(rds_all)
        assert abs(len(rds) - n * 5000 / sr) <= 2
        tail = rds[len(rds) // 2:]
        # 0.05 cos(57 kHz) -> 0.025 at DC after the -57 kHz shift (the other half lands at -114 kHz and is filtered out)
        assert abs(abs(np.mean(tail)) - 0.025) < 0.004
        assert np.std(tail - np.mean(tail)) < 0.01
    au = np.random.default_rng(3).standard_normal((1000, 2)).astype(np.float32)
    for vol, muted in [(0.5, 0), (1.0, 0), (0.37, 0), (0.8, 1)]:
        y = ol.volume(PORT, vol, muted, au)
        assert np.array_equal(y, au * (np.float32(0) if muted else np.float32(vol) ** 2)) or \
            np.allclose(y, au * (0.0 if muted else vol ** 2), rtol=2e-7, atol=0)
        if REF is not None:
            assert ol.volume(REF, vol, muted, au).tobytes() == y.tobytes()


@needs_ref
@pytest.mark.parametrize("stype", [0, 1, 2, 3])
def test_wav_writer_port_equals_ref_bytewise(stype, tmp_path):
    """The recorder's file writer (wav::Writer over riff::Writer): the restatement writes the same bytes as the reference's
    own wav.cpp / riff.cpp for every sample type, mono and stereo, including values that saturate."""
    rng = np.random.default_rng(40 + stype)
    for ch in (1, 2):
        x = (rng.standard_normal((10001, ch)) * 0.4).astype(np.float32)
        x[5] = 1.0; x[6] = -1.0; x[7] = 0.999999; x[9] = 0.5 / 32767.0; x[10] = 1.5 / 32767.0; x[11] = 2.5 / 32767.0
        if stype in (1, 2):
            x[8] = 1.7; x[12] = -3.0           # beyond full scale: clamped (the uint8 path has no clamp and is left in range)
        a, b = tmp_path / ("p%d.wav" % ch), tmp_path / ("r%d.wav" % ch)
        ol.wav_write(PORT, a, ch, 48000, stype, x, 4096)
        ol.wav_write(REF, b, ch, 48000, stype, x, 4096)
        pa, pb = open(a, "rb").read(), open(b, "rb").read()
        assert len(pa) == 44 + len(x) * ch * (1 if stype == 0 else 2 if stype == 1 else 4)
        assert pa == pb, (stype, ch)


@needs_ref
def test_server_sample_packets_port_equals_ref():
    """The SDR++ server's sample packets (SampleStreamCompressor / SampleStreamDecompressor): restatement against the
    reference's headers byte for byte, for int8, int16 and float32 payloads, an empty block and a block of negatives."""
    rng = np.random.default_rng(55)
    blocks = [((rng.standard_normal(5000) + 1j * rng.standard_normal(5000)) * 0.3).astype(np.complex64),
              np.zeros(0, np.complex64),
              (-(np.abs(rng.standard_normal(300)) + 0.1) * (1 + 1j)).astype(np.complex64),
              np.concatenate([np.full(64, 0.25 + 0.25j), np.full(8, 0.1 - 0.05j)]).astype(np.complex64)]
    for x in blocks:
        for t in (0, 1, 2):
            p, r = ol.stream_compress(PORT, x, t), ol.stream_compress(REF, x, t)
            assert p.tobytes() == r.tobytes(), (len(x), t)
            assert len(p) == 8 + len(x) * (2, 4, 8)[t]
            if len(x):
                dp, dr = ol.stream_decompress(PORT, p), ol.stream_decompress(REF, p)
                assert dp.tobytes() == dr.tobytes() and len(dp) == len(x)
                if t == 2:
                    assert np.array_equal(dp, x)
                elif x.real.max() > 0:
                    # the scaler is the largest SIGNED value: what lies within +-scaler comes back to one step, the rest clips
                    xf, df = x.view(np.float32), dp.view(np.float32)
                    mx = xf.max()
                    step = mx / (128.0 if t == 0 else 32768.0)
                    inside = np.abs(xf) <= mx * (1 - 2.0 / 128)
                    assert inside.any() and np.max(np.abs(df[inside] - xf[inside])) <= 1.01 * step
