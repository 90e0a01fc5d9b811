"""GPU parity: IQ front end (pre-processing, FFT frames, waterfall zoom/palette) against the oracle.

Tolerances (SURVEY §8c / BASELINE.md §4): pre-processed stream -80 dB per block, counts exact;
complex spectrum -80 dB vs the float64 oracle FFT; dB rows within 0.01 dB for bins within 80 dB of the
frame peak; zoomed rows select identical bins; palette index +-1."""
import os

import numpy as np
import pytest

import golden_cases
import oracle_lib as ol
from gpu_util import err_db, split, to_dev
from sdrplusplusbrown_b200 import synth

pytestmark = pytest.mark.gpu

PORT = ol.Oracle("port")
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _fe(*a, **k):
    from sdrplusplusbrown_b200 import binding
    return binding.FrontEnd(*a, **k)


@pytest.mark.parametrize("cfg", [(1, 1, 0), (4, 1, 1), (16, 0, 0), (2, 1, 0), (1, 0, 1), (1, 0, 0), (64, 1, 0)])
def test_preproc_chain_matches_oracle(cfg):
    dec, dc, inv = cfg
    sr = 8e6
    x = synth.make_iq(0, 200000, sr, tones=[(1e6, 0.1)], dc=(0.05, -0.08))
    o = ol.Preproc(PORT, sr, dec, dc, inv)
    fe = _fe(sr, decim_ratio=dec, dc_blocking=bool(dc), invert_iq=bool(inv), fft_size=1024, fft_rate=20.0, max_block=70000)
    for blk in split(x, [65536, 7001, 33]):
        ref = o.process(blk)
        fe.process(to_dev(blk))
        got = fe.preproc_out()
        assert len(got) == len(ref)
        if len(ref):
            assert err_db(got, ref) <= -80.0, cfg


def test_dc_blocker_long_stream_state_carry():
    sr = 20e6
    o = ol.Preproc(PORT, sr, 1, 1, 0)
    fe = _fe(sr, dc_blocking=True, fft_size=1024, fft_rate=60.0)
    for i in range(3):
        blk = synth.make_iq(i * 1000000, 1000000, sr, dc=(0.2, -0.1))
        ref = o.process(blk)
        fe.process(to_dev(blk))
        assert err_db(fe.preproc_out(), ref) <= -80.0


@pytest.mark.parametrize("N", [32, 256, 512, 1024, 2048, 4096, 8192, 16384, 32768, 65536, 1 << 17, 1 << 18, 1 << 20])
def test_fft_complex_spectrum(N):
    from sdrplusplusbrown_b200 import binding
    import torch
    nz = N if N < 4096 else N - 37
    x = synth.make_iq(0, nz, 2.4e6, tones=[(300e3, 0.2), (-777e3, 0.05)])
    d_out = torch.empty((N, 2), dtype=torch.float32, device="cuda")
    binding.fft_forward(to_dev(x), nz, N, d_out)
    got = d_out.cpu().numpy().view(np.complex64).reshape(-1)
    xw = np.zeros(N, np.complex128)
    xw[:nz] = x
    ref = np.fft.fft(xw)
    assert err_db(got, ref) <= -80.0


FFT_CASES = [
    (2.4e6, 8192, 20.0, 2, [12000]),            # cfg1: nz = N, 20 fps
    (20e6, 65536, 60.0, 1, [1000000]),          # cfg2: two-pass, I = 333333
    (48e3, 4096, 20.0, 0, [1000, 777]),         # nz (2400) < N, frames straddle blocks
    (2.4e6, 1024, 2000.0, 1, [5000]),           # nz == I (no skip): back-to-back frames
    # the sizes above 2^16 (two passes of up to 1024 points each), window + dB epilogue and framing included:
    (122.88e6, 1 << 20, 60.0, 1, [1000000]),    # cfg4: I = 2 048 000, nz = 2^20 -- every frame spans two or three blocks
    (122.88e6, 1 << 17, 60.0, 1, [1 << 20]),
    (61.44e6, 1 << 19, 60.0, 2, [777777]),      # the largest size the reference's display menu offers (display.cpp:54-66)
    (30.72e6, 1 << 18, 60.0, 0, [1 << 19, 333]),
]


@pytest.mark.parametrize("case", FFT_CASES)
def test_fft_frames_match_oracle(case):
    sr, N, rate, wt, sizes = case
    skip, nz = PORT.reshape_params(sr, N, rate)
    interval = nz + skip
    n = interval * 3 + nz + 17
    x = synth.make_iq(0, n, sr, tones=[(0.11 * sr, 0.2), (-0.31 * sr, 0.01)])
    w = PORT.fft_window(wt, nz)
    fe = _fe(sr, fft_size=N, fft_rate=rate, fft_window=wt, max_block=max(sizes))
    rows, idx = [], []
    for blk in split(x, sizes):
        fe.process(to_dev(blk))
        for i, r in enumerate(fe.frames()):
            rows.append(r)
            idx.append(fe.frame_index(i))
    assert idx == list(range(4))
    for f, row in zip(idx, rows):
        ref = PORT.fft_frame(x[f * interval:f * interval + nz], w, N)
        strong = ref > ref.max() - 80.0
        assert np.max(np.abs(row[strong] - ref[strong])) <= 0.01
        assert np.all(np.isfinite(row))


def test_fft_zero_input_reads_clamped_floor():
    fe = _fe(48e3, fft_size=1024, fft_rate=40.0, fft_window=0, max_block=4096)
    fe.process(to_dev(np.zeros(2400, np.complex64)))
    rows = fe.frames()
    assert len(rows) == 2
    assert np.allclose(rows[0], -382.3, atol=0.1)   # VOLK generic: log2(0) clamped to -127


def test_fft_frame_sharding_partitions_frames():
    sr, N, rate = 2.4e6, 1024, 2000.0
    x = synth.make_iq(0, 1200 * 9, sr, tones=[(1e5, 0.1)])
    full = _fe(sr, fft_size=N, fft_rate=rate, max_block=20000)
    full.process(to_dev(x))
    all_rows = full.frames()
    assert len(all_rows) == 9
    seen = {}
    for r in range(2):
        fe = _fe(sr, fft_size=N, fft_rate=rate, max_block=20000)
        fe.set_frame_shard(r, 2)
        for blk in split(x, [5000, 3333]):
            fe.process(to_dev(blk))
            for i, row in enumerate(fe.frames()):
                seen[fe.frame_index(i)] = row
    assert sorted(seen) == list(range(9))
    for f in range(9):
        assert np.max(np.abs(seen[f] - all_rows[f])) <= 1e-4


@pytest.mark.parametrize("name", sorted(golden_cases.FFT_GOLDEN))
def test_fft_zoom_palette_against_golden(name):
    import torch
    from sdrplusplusbrown_b200 import binding
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    sr, N, rate, wt, zoom, lo, hi = golden_cases.FFT_GOLDEN[name]
    x, nz = golden_cases.fft_input(name)
    fe = _fe(sr, fft_size=N, fft_rate=rate, fft_window=wt, max_block=max(nz, 1024))
    fe.process(to_dev(x))
    assert fe.frame_count() == 1
    row = fe.frames()[0]
    stride = max(1, N // 4096)
    ref_sub = g["db_sub"]
    strong = ref_sub > g["db_peak"][0] - 80.0
    assert np.max(np.abs(row[::stride][strong] - ref_sub[strong])) <= 0.01
    assert int(np.argmax(row)) == int(g["db_peak"][1])
    wf = binding.Waterfall(N, zoom[2])
    wf.set_view(zoom[0], zoom[1])
    wf.set_range(lo, hi)
    sid, uf = wf.zoom_table()
    assert np.array_equal(sid, g["zoom_sid"]) and np.array_equal(uf, g["zoom_uf"])   # identical bin selection
    d_zoom = torch.empty(zoom[2], dtype=torch.float32, device="cuda")
    d_rgba = torch.empty(zoom[2], dtype=torch.int32, device="cuda")
    colors = np.array([[0, 0, 0], [255, 255, 255]], np.float32)
    lut = binding.Waterfall.make_palette(colors)
    assert np.array_equal(lut, PORT.make_palette(colors))
    wf.set_palette(np.arange(1000, dtype=np.uint32))    # identity LUT: rgba output = palette index
    wf.push(fe.frame_ptr(0), 1, d_zoom.data_ptr(), d_rgba.data_ptr())
    z = d_zoom.cpu().numpy()
    zs = g["zoom"] > g["db_peak"][0] - 80.0
    assert np.max(np.abs(z[zs] - g["zoom"][zs])) <= 0.01
    pal = d_rgba.cpu().numpy()
    assert np.max(np.abs(pal[zs] - g["palette"][zs])) <= 1
    # and the device zoom is exactly the max over the reference's bins of the device row
    ref_z = np.array([row[a:a + b].max() for a, b in zip(sid, uf)], np.float32)
    assert np.array_equal(z, ref_z)


def test_preproc_against_golden():
    g = np.load(os.path.join(GOLDEN, "preproc.npz"))
    x = golden_cases.preproc_input()
    for tag, (dec, dc, inv) in dict(dc=(1, 1, 0), d4_dc_inv=(4, 1, 1), d16=(16, 0, 0)).items():
        fe = _fe(8e6, decim_ratio=dec, dc_blocking=bool(dc), invert_iq=bool(inv), fft_size=1024, fft_rate=20.0, max_block=8192)
        outs = []
        for blk in split(x, [7000]):
            fe.process(to_dev(blk))
            outs.append(fe.preproc_out())
        got = np.concatenate(outs)
        assert len(got) == len(g["out_" + tag])
        assert err_db(got, g["out_" + tag]) <= -80.0


def test_frontend_feeds_vfos_and_fft_together():
    """cfg1 end to end: 2.4 MS/s, 8192-pt Nuttall FFT at 20 fps + one NFM VFO, file-source block size 12000."""
    sr = 2.4e6
    x = synth.make_iq(0, 240000, sr, tones=[(100e3 + 1e3, 0.05)])
    fe = _fe(sr, fft_size=8192, fft_rate=20.0, fft_window=2, max_block=12000)
    v = fe.vfos.add(50e3, 12.5e3, 100e3)
    o = ol.Vfo(PORT, sr, 50e3, 12.5e3, 100e3)
    w = PORT.fft_window(2, 8192)
    nframes = 0
    for i, blk in enumerate(split(x, [12000])):
        fe.vfos.set_phase(v, o.phase())
        ref = o.process(blk)
        fe.process(to_dev(blk))
        assert fe.vfos.out_count(v) == len(ref) == 250
        assert err_db(fe.vfos.output(v), ref) <= -80.0
        for j, row in enumerate(fe.frames()):
            f = fe.frame_index(j)
            refrow = PORT.fft_frame(x[f * 120000:f * 120000 + 8192], w, 8192)
            strong = refrow > refrow.max() - 80.0
            assert np.max(np.abs(row[strong] - refrow[strong])) <= 0.01
            nframes += 1
    assert nframes == 2


def test_vfo_signal_info_smoothing_and_hold_match_oracle():
    """calculateVFOSignalInfo (strength / SNR), FFT smoothing and peak hold of WaterFall::pushFFT."""
    import torch
    from sdrplusplusbrown_b200 import binding
    sr, N, W = 2.4e6, 8192, 1000
    x = synth.make_iq(0, 8192 * 6, sr, tones=[(100e3 + 1e3, 0.05), (-400e3, 0.2)])
    fe = _fe(sr, fft_size=N, fft_rate=sr / 8192.0, fft_window=2, max_block=len(x))
    fe.process(to_dev(x))
    rows = fe.frames()
    assert len(rows) == 6
    wf = binding.Waterfall(N, W)
    wf.set_view(0, N)
    d_sig = torch.empty((6, 2), dtype=torch.float32, device="cuda")
    for center, bw in [(100e3, 12.5e3), (-400e3, 50e3), (900e3, 200e3)]:
        assert wf.vfo_signal(fe.frame_ptr(0), 6, sr, center, bw, d_sig.data_ptr())
        got = d_sig.cpu().numpy()
        for f in range(6):
            ref = PORT.vfo_signal_info(rows[f], sr, center, bw)
            assert ref is not None
            assert abs(got[f, 0] - ref[0]) <= 1e-3 and abs(got[f, 1] - ref[1]) <= 1e-3
    assert not wf.vfo_signal(fe.frame_ptr(0), 6, sr, 100e3, 100.0, d_sig.data_ptr())     # band narrower than a bin: the reference returns false
    assert PORT.vfo_signal_info(rows[0], sr, 100e3, 100.0) is None
    # zoom -> smoothing + hold across the six rows, state carried inside the handle across two calls
    d_zoom = torch.empty((6, W), dtype=torch.float32, device="cuda")
    d_hold = torch.empty((6, W), dtype=torch.float32, device="cuda")
    wf.push(fe.frame_ptr(0), 6, d_zoom.data_ptr(), None)
    zoomed = d_zoom.cpu().numpy().copy()
    wf.smooth_hold(d_zoom.data_ptr(), 4, True, 0.3, True, 0.5, d_hold.data_ptr())
    wf.smooth_hold(d_zoom[4:].data_ptr(), 2, True, 0.3, True, 0.5, d_hold[4:].data_ptr())
    sm, hd = d_zoom.cpu().numpy(), d_hold.cpu().numpy()
    sbuf = np.full(W, -1000.0, np.float32)
    hbuf = np.full(W, -1000.0, np.float32)
    for f in range(6):
        latest = zoomed[f].copy()
        PORT.smooth_hold(latest, sbuf, hbuf, True, 0.3, True, 0.5)
        assert np.max(np.abs(sm[f] - latest)) <= 1e-3
        assert np.max(np.abs(hd[f][1:] - hbuf[1:])) <= 1e-3
        assert hd[f][0] == -1000.0        # pixel 0 is never updated by the reference's hold loop


def test_process_host_pipelined_long_block_matches_one_shot_device_path():
    """b200_frontend_process_host cuts a long host block into chunks (copy of chunk k+1 behind the kernels of chunk k).
    Frames, VFO output and counts must equal those of the same block processed in one device-resident call."""
    from sdrplusplusbrown_b200 import binding
    sr, N, rate = 20e6, 65536, 60.0
    n = 5 * (1 << 21) + 12345                       # five and a bit chunks
    x = synth.make_iq(7, n, sr, sigma=0.05, tones=[(2.5e6 + 900.0, 0.2), (-3.3e6, 0.1)])
    a = _fe(sr, fft_size=N, fft_rate=rate, fft_window=1, max_block=n)
    b = _fe(sr, fft_size=N, fft_rate=rate, fft_window=1, max_block=n)
    va = a.vfos.add(24e3, 2.8e3, 2.5e6)
    vb = b.vfos.add(24e3, 2.8e3, 2.5e6)
    for fe, v in ((a, va), (b, vb)):
        fe.vfos.set_rotator_mode(binding.ROTATOR_EXACT)
    a.process(to_dev(x))
    ref_frames = np.stack(a.frames())
    ref_out = a.vfos.output(va)
    got = b.process_host(x, max_frames=len(ref_frames) + 2, vfos=[vb])
    assert got["n_frames"] == len(ref_frames) == b.frame_count()
    assert len(got["vfo"][vb]) == len(ref_out)
    assert err_db(got["vfo"][vb], ref_out) <= -80.0
    peak = ref_frames.max(axis=1, keepdims=True)
    m = ref_frames > peak - 80.0
    assert np.max(np.abs(got["frames"][m] - ref_frames[m])) <= 0.01
    # the rows of every chunk are still resident, and the pre-processed block is the whole staged block
    assert np.array_equal(np.stack(b.frames()), got["frames"])
    assert len(b.preproc_out()) == n
    # a second long call continues the stream
    y = synth.make_iq(8, 2 * (1 << 21) + 5, sr, sigma=0.05, tones=[(2.5e6 + 900.0, 0.2)])
    a.process(to_dev(y))
    got2 = b.process_host(y, max_frames=64, vfos=[vb])
    assert got2["n_frames"] == a.frame_count()
    assert err_db(got2["vfo"][vb], a.vfos.output(va)) <= -80.0


def test_waterfall_history_and_full_regeneration_match_oracle():
    """WaterFall::rawFFTs + updateWaterfallFb: rows pushed in ragged batches are kept (newest `height`), and a view /
    range change regenerates every stored line -- newest first, opaque black past the stored lines -- with the oracle's
    doZoom bins and palette indices."""
    import torch
    from sdrplusplusbrown_b200 import binding
    N, W, H = 4096, 333, 12
    rng = np.random.default_rng(11)
    rows = (rng.standard_normal((20, N)) * 12.0 - 60.0).astype(np.float32)
    d_rows = torch.from_numpy(rows).cuda()
    wf = binding.Waterfall(N, W)
    wf.set_history(H)
    wf.set_palette(np.arange(1000, dtype=np.uint32))          # identity LUT: rgba = palette index
    wf.set_range(-100.0, -20.0)
    d_zoom = torch.empty((H, W), dtype=torch.float32, device="cuda")
    d_rgba = torch.empty((H, W), dtype=torch.int32, device="cuda")

    def check(n_pushed, view, lo, hi):
        wf.set_view(*view)
        wf.set_range(lo, hi)
        d_rgba.fill_(7)
        stored = wf.regenerate(d_zoom.data_ptr(), d_rgba.data_ptr())
        torch.cuda.synchronize()
        assert stored == min(n_pushed, H)
        z, p = d_zoom.cpu().numpy(), d_rgba.cpu().numpy()
        for i in range(stored):
            src = rows[n_pushed - 1 - i]                      # row 0 = newest
            ref_z = PORT.do_zoom(view[0], view[1], src, W)
            assert np.array_equal(z[i], ref_z), i
            ref_p = PORT.palette_indices(ref_z, lo, hi)
            assert np.max(np.abs(p[i] - ref_p)) <= 1
        if stored < H:
            assert np.all(p[stored:].view(np.uint32) == 0xFF000000)

    wf.push(d_rows[0:5].data_ptr(), 5, None, None)            # history only, no per-row output
    check(5, (0, N), -100.0, -20.0)
    wf.push(d_rows[5:6].data_ptr(), 1, None, None)
    wf.push(d_rows[6:17].data_ptr(), 11, None, None)          # wraps the ring
    check(17, (1000, 1500), -90.0, -30.0)                     # zoomed-in view, new range
    wf.push(d_rows[17:20].data_ptr(), 3, None, None)
    check(20, (0, N), -100.0, -20.0)
    wf.set_history(4)                                         # a batch longer than the history keeps its newest rows
    d_zoom4 = torch.empty((4, W), dtype=torch.float32, device="cuda")
    wf.push(d_rows[0:9].data_ptr(), 9, None, None)
    wf.set_view(0, N)
    assert wf.regenerate(d_zoom4.data_ptr(), None) == 4
    z = d_zoom4.cpu().numpy()
    for i in range(4):
        assert np.array_equal(z[i], PORT.do_zoom(0, N, rows[8 - i], W))


@pytest.mark.parametrize("n", [300000, 2 * (1 << 22) + 777])
def test_process_host_int16_ingest_is_bit_identical_to_the_converted_block(n):
    """int16 I/Q crosses PCIe as 4 bytes per sample and is converted on the device with the file source's arithmetic
    (volk_16i_s32f_convert_32f, scalar 32768): outputs must equal the cf32 call on the host-converted block -- bit for
    bit for a one-shot block, to rounding for one long enough to take the pipelined path."""
    from sdrplusplusbrown_b200 import binding
    import torch
    sr, N, rate = 2.4e6, 8192, 20.0
    rng = np.random.default_rng(5)
    xi = rng.integers(-32768, 32767, size=(n, 2), dtype=np.int16)
    t = np.arange(n)
    xi[:, 0] = np.clip(xi[:, 0] // 16 + (9000 * np.cos(2 * np.pi * 100.7e3 / sr * t)).astype(np.int32), -32768, 32767).astype(np.int16)
    xi[:, 1] = np.clip(xi[:, 1] // 16 + (9000 * np.sin(2 * np.pi * 100.7e3 / sr * t)).astype(np.int32), -32768, 32767).astype(np.int16)
    xf = (xi.astype(np.float32) / np.float32(32768.0)).view(np.complex64).reshape(-1)
    outs = []
    for kind in ("f32", "i16"):
        fe = _fe(sr, fft_size=N, fft_rate=rate, fft_window=2, max_block=n)
        v = fe.vfos.add(50e3, 12.5e3, 100e3)
        fe.vfos.set_demod(v, binding.DEMOD_NFM, 1, 0, 0, 0)
        if kind == "f32":
            outs.append(fe.process_host(xf, max_frames=64, vfos=[v], audio=[v]))
        else:
            outs.append(fe.process_host(xi, max_frames=64, vfos=[v], audio=[v], i16_scalar=32768.0))
    a, b = outs
    assert a["n_frames"] == b["n_frames"] > 0
    assert len(a["vfo"][v]) == len(b["vfo"][v]) > 0
    na = len(a["vfo"][v])
    if n < (1 << 22):
        # same cut of the block on both paths: the same kernels see the same numbers
        assert np.array_equal(a["frames"], b["frames"])
        assert np.array_equal(a["vfo"][v], b["vfo"][v])
        assert np.array_equal(a["audio"][v][:na], b["audio"][v][:na])
    else:
        # the two paths chunk by bytes (16 MiB), i.e. at different sample counts: equal up to summation order at the cuts
        assert np.max(np.abs(a["frames"] - b["frames"])) <= 1e-3
        assert err_db(b["vfo"][v], a["vfo"][v]) <= -100.0
        assert err_db(b["audio"][v][:na, 0], a["audio"][v][:na, 0]) <= -80.0
    # the device-resident conversion entry point
    d_in = torch.from_numpy(xi[:1000].copy()).cuda()
    d_out = torch.empty((1000, 2), dtype=torch.float32, device="cuda")
    assert binding.lib.b200_convert_i16(d_in.data_ptr(), 1000, 32768.0, d_out.data_ptr(), None) == 0
    torch.cuda.synchronize()
    assert np.array_equal(d_out.cpu().numpy(), xi[:1000].astype(np.float32) / np.float32(32768.0))


def test_process_host_int8_ingest_matches_the_converted_block():
    """int8 I/Q (the server stream's PCM_TYPE_I8, scalar 128 / scaler): 2 bytes per sample over PCIe, converted on the
    device; a one-shot block must equal the cf32 call on the host-converted block bit for bit."""
    from sdrplusplusbrown_b200 import binding
    sr, N, rate, n = 2.4e6, 8192, 20.0, 250000
    rng = np.random.default_rng(9)
    t = np.arange(n)
    xi = np.empty((n, 2), np.int8)
    xi[:, 0] = np.clip(rng.integers(-8, 8, n) + (90 * np.cos(2 * np.pi * 100.7e3 / sr * t)).astype(np.int32), -128, 127)
    xi[:, 1] = np.clip(rng.integers(-8, 8, n) + (90 * np.sin(2 * np.pi * 100.7e3 / sr * t)).astype(np.int32), -128, 127)
    scalar = np.float32(128.0 / 0.75)
    xf = (xi.astype(np.float32) / scalar).view(np.complex64).reshape(-1)
    outs = []
    for kind in ("f32", "i8"):
        fe = _fe(sr, fft_size=N, fft_rate=rate, fft_window=2, max_block=n)
        v = fe.vfos.add(50e3, 12.5e3, 100e3)
        if kind == "f32":
            outs.append(fe.process_host(xf, max_frames=8, vfos=[v]))
        else:
            outs.append(fe.process_host(xi, max_frames=8, vfos=[v], i8_scalar=float(scalar)))
    a, b = outs
    assert a["n_frames"] == b["n_frames"] > 0
    assert np.array_equal(a["frames"], b["frames"])
    assert len(a["vfo"][v]) == len(b["vfo"][v]) > 0
    assert np.array_equal(a["vfo"][v], b["vfo"][v])


def test_attached_waterfall_rows_equal_a_separate_push():
    """b200_frontend_attach_waterfall runs the zoom + palette reduction inside the FFT branch; its rows must be the
    rows a separate b200_waterfall_push of the same frames gives, call after call (with a VFO present, so that the
    branch really runs on its side stream)."""
    import torch
    from sdrplusplusbrown_b200 import binding
    sr, N, rate, W = 20e6, 65536, 60.0, 1333
    n = 1500000
    x = synth.make_iq(2, 2 * n, sr, sigma=0.05, tones=[(2.5e6, 0.2), (-6.1e6, 0.02)])
    fe = _fe(sr, fft_size=N, fft_rate=rate, fft_window=1, max_block=n)
    fe.vfos.add(24e3, 2.8e3, 2.5e6)
    wf = binding.Waterfall(N, W)
    wf.set_view(8192, 40000)
    wf.set_range(-110.0, -10.0)
    wf.set_palette(np.arange(1000, dtype=np.uint32))
    max_rows = n // int(round(sr / rate)) + 2
    d_zoom = torch.full((max_rows, W), 123.0, dtype=torch.float32, device="cuda")
    d_rgba = torch.zeros((max_rows, W), dtype=torch.int32, device="cuda")
    fe.attach_waterfall(wf, d_zoom.data_ptr(), d_rgba.data_ptr(), max_rows)
    ref_zoom = torch.empty((max_rows, W), dtype=torch.float32, device="cuda")
    ref_rgba = torch.empty((max_rows, W), dtype=torch.int32, device="cuda")
    for blk in (x[:n], x[n:]):
        fe.process(to_dev(blk))
        nf = fe.frame_count()
        assert nf >= 4
        torch.cuda.synchronize()
        got_z, got_p = d_zoom[:nf].cpu().numpy(), d_rgba[:nf].cpu().numpy()
        wf.push(fe.frame_ptr(0), nf, ref_zoom.data_ptr(), ref_rgba.data_ptr())
        torch.cuda.synchronize()
        assert np.array_equal(got_z, ref_zoom[:nf].cpu().numpy())
        assert np.array_equal(got_p, ref_rgba[:nf].cpu().numpy())
    fe.attach_waterfall(None, None, None, 0)
    d_zoom.fill_(7.0)
    fe.process(to_dev(x[:n]))
    torch.cuda.synchronize()
    assert float(d_zoom.min()) == 7.0          # detached: nothing written


@pytest.mark.parametrize("stype,np_type", [(0, np.uint8), (1, np.int16), (2, np.int32)])
def test_recorder_pcm_conversion_matches_the_reference_writer(stype, np_type, tmp_path):
    """b200_convert_pcm against wav::Writer::write (utils/wav.cpp:148-178): the converted samples are the bytes of the
    data chunk the reference's writer produces, bit for bit, including rounding ties and values beyond full scale."""
    import torch
    from sdrplusplusbrown_b200 import binding
    rng = np.random.default_rng(70 + stype)
    x = (rng.standard_normal(100003) * 0.4).astype(np.float32)
    x[:8] = [1.0, -1.0, 0.999999, 0.5 / 32767.0, 1.5 / 32767.0, 2.5 / 32767.0, -0.5 / 32767.0, 0.0]
    if stype:
        x[8:12] = [1.7, -3.0, 1.00001, -1.00001]         # clamped; the uint8 path has no clamp and stays in range
    else:
        x = np.clip(x, -1.0, 0.99)
    path = tmp_path / "ref.wav"
    ol.wav_write(PORT, path, 1, 48000, stype, x, 8192)
    want = np.frombuffer(open(path, "rb").read()[44:], np_type)
    d_in = torch.from_numpy(x).cuda()
    d_out = torch.zeros(len(x), dtype={0: torch.uint8, 1: torch.int16, 2: torch.int32}[stype], device="cuda")
    assert binding.lib.b200_convert_pcm(d_in.data_ptr(), d_out.data_ptr(), len(x), stype, None) == 0
    torch.cuda.synchronize()
    got = d_out.cpu().numpy()
    assert got.dtype == want.dtype and np.array_equal(got, want)
    assert binding.lib.b200_convert_pcm(d_in.data_ptr(), d_out.data_ptr(), len(x), 3, None) != 0     # float files need no conversion
    assert binding.lib.b200_convert_pcm(d_in.data_ptr(), d_out.data_ptr(), 0, stype, None) == 0


@pytest.mark.parametrize("pcm", [0, 1, 2])
def test_server_sample_packet_matches_the_reference_compressor(pcm):
    """b200_stream_compress against SampleStreamCompressor::process: same packet bytes (header, scaler, payload) for a noisy
    block, an all-negative block, a one-sample block and an empty one; and the packet fed back through the int16 / int8
    ingest of the front end gives what SampleStreamDecompressor gives."""
    import torch
    from sdrplusplusbrown_b200 import binding
    rng = np.random.default_rng(90 + pcm)
    blocks = [((rng.standard_normal(300001) + 1j * rng.standard_normal(300001)) * 0.3).astype(np.complex64),
              (-(np.abs(rng.standard_normal(4097)) + 0.1) * (1 + 0.5j)).astype(np.complex64),
              np.array([0.3 - 0.7j], np.complex64), np.zeros(0, np.complex64)]
    for x in blocks:
        want = ol.stream_compress(PORT, x, pcm)
        d_in = torch.from_numpy(x.view(np.float32).copy()).cuda() if len(x) else torch.zeros(2, device="cuda")
        d_out = torch.zeros(8 + 8 * max(len(x), 1), dtype=torch.uint8, device="cuda")
        n = binding.lib.b200_stream_compress(d_in.data_ptr(), len(x), pcm, d_out.data_ptr(), None)
        torch.cuda.synchronize()
        assert n == len(want)
        got = d_out.cpu().numpy()[:n]
        assert np.array_equal(got, want), (pcm, len(x))
    assert binding.lib.b200_stream_compress(0, 5, pcm, 0, None) < 0
    if pcm < 2:
        # receiving side: header read on the host, payload through the integer ingest
        x = blocks[0][:65536]
        pkt = ol.stream_compress(PORT, x, pcm)
        scaler = pkt[4:8].view(np.float32)[0]
        ref = ol.stream_decompress(PORT, pkt)
        fe = binding.FrontEnd(2.4e6, fft_size=8192, fft_rate=20.0, fft_window=2, max_block=65536, device=0)
        v = fe.vfos.add(2.4e6, 2.4e6, 0.0)             # no translation, no resampling: the VFO hands the ingested block through
        payload = pkt[8:].view(np.int8 if pcm == 0 else np.int16).reshape(-1, 2)
        kw = dict(i8_scalar=float(np.float32(128.0) / scaler)) if pcm == 0 else dict(i16_scalar=float(np.float32(32768.0) / scaler))
        res = fe.process_host(payload, vfos=(v,), **kw)
        got = res["vfo"][v] if isinstance(res, dict) and "vfo" in res else res[1][v]
        assert err_db(got, ref) <= -120.0


@pytest.mark.parametrize("pinned", [True, False])
def test_process_host_returns_a_large_bank_in_one_scatter(pinned):
    """With 16 or more VFOs b200_frontend_process_host stores all IF outputs of a chunk into the callers' buffers with
    one kernel when those buffers are page-locked; pageable buffers take one copy per VFO.  Either way the bytes equal the
    device-resident outputs of the same blocks, call after call (the second call continues every buffer's stream)."""
    import ctypes
    import torch
    from sdrplusplusbrown_b200 import binding
    sr, N, rate = 2.4e6, 8192, 20.0
    nv = 24
    n = 300000
    offs = [(-1.0e6 + 80e3 * k + 7.0) for k in range(nv)]
    x = synth.make_iq(3, 2 * n, sr, sigma=0.05, tones=[(o + 900.0, 0.02) for o in offs])
    a = _fe(sr, fft_size=N, fft_rate=rate, fft_window=2, max_block=n)
    b = _fe(sr, fft_size=N, fft_rate=rate, fft_window=2, max_block=n)
    for fe in (a, b):
        fe.vfos.set_rotator_mode(binding.ROTATOR_EXACT)
    va = [a.vfos.add(50e3 if k % 3 else 24e3, 12.5e3 if k % 3 else 2.8e3, o) for k, o in enumerate(offs)]
    vb = [b.vfos.add(50e3 if k % 3 else 24e3, 12.5e3 if k % 3 else 2.8e3, o) for k, o in enumerate(offs)]
    sizes = [b.vfos.max_out(v, n) for v in vb]
    total = sum(sizes)
    if pinned:
        store = torch.zeros((total, 2), dtype=torch.float32).pin_memory()
        base, keep = store.data_ptr(), store
        view = store.numpy().view(np.complex64).reshape(-1)
    else:
        view = np.zeros(total, np.complex64)
        base, keep = view.ctypes.data, view
    optr = (ctypes.c_void_p * (max(vb) + 1))()
    starts, pos = [], 0
    for v, m in zip(vb, sizes):
        optr[v] = base + 8 * pos
        starts.append(pos)
        pos += m
    counts = np.zeros(max(vb) + 1, np.int32)
    for blk in (x[:n], x[n:]):
        a.process(to_dev(blk))
        xin = np.ascontiguousarray(blk)
        rc = binding.lib.b200_frontend_process_host(b.h, xin.ctypes.data, len(xin), None, 0, optr, None, counts.ctypes.data, None)
        assert rc == 0, binding.lib.b200_last_error()
        for v_a, v_b, st in zip(va, vb, starts):
            want = a.vfos.output(v_a)
            assert counts[v_b] == len(want) > 0
            assert np.array_equal(view[st:st + len(want)], want)
    assert keep is not None


@pytest.mark.parametrize("streaming", [False, True])
def test_process_host_chunked_large_bank_inline_and_streaming(streaming):
    """A long host block through a bank of 24 VFOs, cut into chunks (B200_OPT_HOST_CHUNK_MIB), inline and in streaming mode
    (B200_OPT_DEFER_TAIL: every chunk's later stages AND its copies back to the callers' page-locked buffers run on the
    bank's own stream beside the next chunk's full-rate kernels).  Counts, frames and every VFO's output must equal the same
    blocks processed device-resident in one call each; the second call continues every stream."""
    import ctypes
    import torch
    from sdrplusplusbrown_b200 import binding
    sr, N, rate = 2.4e6, 8192, 20.0
    nv, n = 24, 700000                               # 1 MiB chunks = 131072 samples: five chunks and a bit
    offs = [(-1.0e6 + 80e3 * k + 7.0) for k in range(nv)]
    x = synth.make_iq(5, 2 * n, sr, sigma=0.05, tones=[(o + 900.0, 0.02) for o in offs])
    a = _fe(sr, fft_size=N, fft_rate=rate, fft_window=2, max_block=n)
    b = _fe(sr, fft_size=N, fft_rate=rate, fft_window=2, max_block=n)
    for fe in (a, b):
        fe.vfos.set_rotator_mode(binding.ROTATOR_EXACT)
    va = [a.vfos.add(50e3 if k % 3 else 24e3, 12.5e3 if k % 3 else 2.8e3, o) for k, o in enumerate(offs)]
    vb = [b.vfos.add(50e3 if k % 3 else 24e3, 12.5e3 if k % 3 else 2.8e3, o) for k, o in enumerate(offs)]
    sizes = [b.vfos.max_out(v, n) for v in vb]
    store = torch.zeros((sum(sizes), 2), dtype=torch.float32).pin_memory()
    view = store.numpy().view(np.complex64).reshape(-1)
    optr = (ctypes.c_void_p * (max(vb) + 1))()
    starts, pos = [], 0
    for v, m in zip(vb, sizes):
        optr[v] = store.data_ptr() + 8 * pos
        starts.append(pos)
        pos += m
    counts = np.zeros(max(vb) + 1, np.int32)
    mf = 64
    h_frames = torch.zeros((mf, N), dtype=torch.float32).pin_memory()
    try:
        b.vfos.set_option(binding.OPT_HOST_CHUNK_MIB, 1)
        b.vfos.set_option(binding.OPT_DEFER_TAIL, int(streaming))
        for blk in (x[:n], x[n:]):
            b.vfos.set_option(binding.OPT_DEFER_TAIL, 0)
            a.process(to_dev(blk))
            ref_frames = np.stack(a.frames())
            wants = [a.vfos.output(v) for v in va]
            b.vfos.set_option(binding.OPT_DEFER_TAIL, int(streaming))
            xin = torch.from_numpy(np.ascontiguousarray(blk).view(np.float32).reshape(-1, 2)).pin_memory()
            view[:] = 0
            rc = binding.lib.b200_frontend_process_host(b.h, xin.data_ptr(), len(blk), h_frames.data_ptr(), mf, optr, None, counts.ctypes.data, None)
            assert rc == 0, binding.lib.b200_last_error()
            assert b.frame_count() == len(ref_frames)
            got_frames = h_frames.numpy()[:len(ref_frames)]
            m = ref_frames > ref_frames.max(axis=1, keepdims=True) - 80.0
            assert np.max(np.abs(got_frames[m] - ref_frames[m])) <= 0.01
            for v_b, st, want in zip(vb, starts, wants):
                assert counts[v_b] == len(want) > 0
                assert err_db(view[st:st + len(want)], want) <= -100.0, (v_b, err_db(view[st:st + len(want)], want))
    finally:
        b.vfos.set_option(binding.OPT_DEFER_TAIL, 0)
        b.vfos.set_option(binding.OPT_HOST_CHUNK_MIB, 0)


@pytest.mark.parametrize("tuned_to_signal", [False, True])
def test_reference_snr_threshold_scenario_on_a_synthetic_cw_capture(tuned_to_signal):
    """The reference's only numeric test, re-created (tests/baseband_file_test.cpp:36-88): a baseband capture with CW
    stations is played through the front end, the radio VFO is USB (2.8 kHz wide, lower edge on the tuned frequency,
    waterfall.cpp:1595-1598), and the largest selectedVFOSNR seen over the run must stay below 20 dB when tuned to a
    clear frequency and exceed 20 dB when tuned to a station.  Its WAV fixtures are absent from the mount, so the
    capture is synthetic: int16 I/Q, noise plus keyed carriers, played through the int16 ingest in the file source's
    block size; SNR = calculateVFOSignalInfo over every dB row (waterfall.cpp:688-759), on the device and -- to pin the
    numbers -- from the oracle on the same rows."""
    import torch
    from sdrplusplusbrown_b200 import binding
    sr, N, rate = 192000.0, 8192, 20.0
    blk = int(np.float32(sr) / np.float32(200.0))
    n = blk * 400                                            # two seconds
    t = np.arange(n) / sr
    rng = np.random.default_rng(1397)
    x = (rng.standard_normal(n) + 1j * rng.standard_normal(n)) * 0.02
    for f0, amp, wpm in ((14.9e3, 0.25, 0.11), (-31.2e3, 0.1, 0.07)):      # two keyed carriers ("CW contest")
        keyed = (np.floor(t / wpm) % 3 != 2).astype(np.float64)
        x += amp * keyed * np.exp(2j * np.pi * f0 * t)
    xi = np.round(np.stack([x.real, x.imag], 1) * 20000.0).clip(-32768, 32767).astype(np.int16)
    tuned = 14.0e3 if tuned_to_signal else 60.0e3           # USB: the channel spans [tuned, tuned + 2.8 kHz]
    bw = 2800.0
    center = tuned + bw / 2.0
    fe = _fe(sr, fft_size=N, fft_rate=rate, fft_window=2, max_block=blk)
    v = fe.vfos.add(24e3, bw, center)
    fe.vfos.set_demod(v, binding.DEMOD_USB, 0, 50.0 / 24e3, 5.0 / 24e3, 0)
    wf = binding.Waterfall(N, 1000)
    wf.set_view(0, N)
    d_sig = torch.empty((4, 2), dtype=torch.float32, device="cuda")
    max_snr, max_ref, rows_seen = 0.0, 0.0, 0
    for i in range(0, n, blk):
        res = fe.process_host(xi[i:i + blk], max_frames=4, vfos=[v], audio=[v], i16_scalar=32768.0)
        k = res["n_frames"]
        if k:
            assert wf.vfo_signal(fe.frame_ptr(0), k, sr, center, bw, d_sig.data_ptr())
            got = d_sig[:k].cpu().numpy()
            for f in range(k):
                ref = PORT.vfo_signal_info(res["frames"][f], sr, center, bw)
                assert ref is not None and abs(got[f, 1] - ref[1]) <= 1e-3
                max_ref = max(max_ref, ref[1])
            max_snr = max(max_snr, float(got[:, 1].max()))
            rows_seen += k
    assert rows_seen >= 38 and max_snr != 0.0                # "MaxSNR is not captured" in the reference's test
    assert abs(max_snr - max_ref) <= 1e-3
    if tuned_to_signal:
        assert max_snr > 20.0, max_snr
    else:
        assert max_snr < 20.0, max_snr
