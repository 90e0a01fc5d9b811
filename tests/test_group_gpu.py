"""GPU parity: the front-end group (b200_frontend_group_*), i.e. one IQFrontEnd over several GPUs in one process.

With one member the group must behave exactly like a plain front end (this part runs on a one-GPU box).  With two
members (skipped below two GPUs) every block is broadcast to the second GPU, VFOs are dealt out over both, FFT frames
alternate between them, and the dB rows come back to the ingest GPU in frame order: VFO outputs and rows must equal a
single-GPU front end's, and the frame-order recurrences (smoothing, peak hold) run on the gathered rows must equal
the single-GPU ones.  Both transports are run: ncclBroadcast and the peer-copy fallback.

Reference: Splitter fan-out (core/src/dsp/routing/splitter.h:74-83) inside IQFrontEnd
(core/src/signal_path/iq_frontend.cpp:46,79,159-179); pushFFT's recurrences (gui/widgets/waterfall.cpp:1247-1273)."""
import os
import subprocess
import sys

import numpy as np
import pytest

from gpu_util import err_db, to_dev
from sdrplusplusbrown_b200 import synth

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SPECS = [(48e3, 12.5e3, 2.2e6 + 11.0), (15e3, 10e3, -7.1e6 + 3.0), (24e3, 2.8e3, 9.9e6), (500e3, 150e3, -20e6), (48e3, 12.5e3, -0.4e6), (50e3, 12.5e3, 27e6 + 5.0)]


def _gpus():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


def _run_pair(devices, dc=False, decim=1, i16=False):
    """group over `devices` against a plain front end on device 0, three ragged blocks; returns nothing, asserts"""
    import torch
    from sdrplusplusbrown_b200 import binding
    sr, N, rate, W = 61.44e6, 65536, 240.0, 777           # I = 256000: frames alternate between the members
    sizes = [1000000, 333333, 1 << 20]
    ref = binding.FrontEnd(sr, decim_ratio=decim, dc_blocking=dc, fft_size=N, fft_rate=rate, fft_window=1, max_block=max(sizes), device=devices[0])
    grp = binding.FrontEndGroup(devices, sr, decim_ratio=decim, dc_blocking=dc, fft_size=N, fft_rate=rate, fft_window=1, max_block=max(sizes))
    ref.vfos.set_rotator_mode(binding.ROTATOR_EXACT)
    for b in grp.banks:
        b.set_rotator_mode(binding.ROTATOR_EXACT)
    scale = 1.0 / decim
    rid = [ref.vfos.add(o, bw, off * scale) for o, bw, off in SPECS]
    gid = [grp.add_vfo(o, bw, off * scale) for o, bw, off in SPECS]
    members = {grp.location(v)[0] for v in gid}
    assert members == set(range(len(devices)))                      # the dealing-out uses every member
    wf_r, wf_g = binding.Waterfall(N, W, device=devices[0]), binding.Waterfall(N, W, device=devices[0])
    for wf in (wf_r, wf_g):
        wf.set_view(1000, 50000)
        wf.set_range(-110.0, -10.0)
    max_rows = max(sizes) // 256000 + 3
    with torch.cuda.device(devices[0]):
        zr = torch.empty((max_rows, W), dtype=torch.float32, device="cuda")
        zg = torch.empty((max_rows, W), dtype=torch.float32, device="cuda")
        hr = torch.empty((max_rows, W), dtype=torch.float32, device="cuda")
        hg = torch.empty((max_rows, W), dtype=torch.float32, device="cuda")
    ref.attach_waterfall(wf_r, zr.data_ptr(), None, max_rows)
    grp.attach_waterfall(wf_g, zg.data_ptr(), None, max_rows)
    pos, seen = 0, []
    for n in sizes:
        x = synth.make_iq(pos, n, sr, sigma=0.05, tones=[(off + 500.0, 0.05) for _, _, off in SPECS], dc=(0.03, -0.02))
        pos += n
        if i16:
            xi = np.round(x.view(np.float32).reshape(-1, 2) * 32767.0).astype(np.int16)
            xf = (xi.astype(np.float32) / np.float32(32768.0)).view(np.complex64).reshape(-1)
            with torch.cuda.device(devices[0]):
                ref.process(to_dev(xf))
            grp.process_host(xi, i16_scalar=32768.0)
        else:
            with torch.cuda.device(devices[0]):
                d = to_dev(x)
                ref.process(d)
                grp.process(d)
        grp.sync()
        torch.cuda.synchronize(devices[0])
        for a, b in zip(rid, gid):
            want = ref.vfos.output(a)
            assert grp.out_count(b) == len(want)
            got = grp.output(b)
            if len(want):
                assert err_db(got, want) <= -100.0          # same kernels; later-stage lane splits may differ with the launch shape
        rows_r, rows_g = ref.frames(), grp.frames()
        assert len(rows_r) == len(rows_g)
        assert [grp.frame_index(i) for i in range(len(rows_g))] == [ref.frame_index(i) for i in range(len(rows_r))]
        for a, b in zip(rows_r, rows_g):
            assert np.array_equal(a, b)                     # same FFT kernels on the same samples: same bits
        seen += [grp.frame_index(i) for i in range(len(rows_g))]
        k = len(rows_g)
        if k:
            # zoomed rows, then the frame-order recurrences on the owner: same state evolution as on one GPU
            torch.cuda.synchronize(devices[0])
            assert np.array_equal(zr[:k].cpu().numpy(), zg[:k].cpu().numpy())
            wf_r.smooth_hold(zr.data_ptr(), k, True, 0.3, True, 0.5, hr.data_ptr())
            wf_g.smooth_hold(zg.data_ptr(), k, True, 0.3, True, 0.5, hg.data_ptr())
            torch.cuda.synchronize(devices[0])
            assert np.array_equal(zr[:k].cpu().numpy(), zg[:k].cpu().numpy())
            assert np.array_equal(hr[:k].cpu().numpy(), hg[:k].cpu().numpy())
    assert seen == list(range(len(seen))) and len(seen) >= 8
    if len(devices) > 1:
        assert grp.broadcast_ms() > 0.0
    return grp.uses_nccl()


@pytest.mark.parametrize("cfg", [dict(), dict(dc=True), dict(decim=2, dc=True), dict(i16=True)])
def test_group_of_one_is_a_plain_front_end(cfg):
    _run_pair([0], **cfg)


_CHILD = """
import sys
sys.path.insert(0, %r); sys.path.insert(0, %r)
import test_group_gpu as t
used = t._run_pair([0, 1], **%r)
print("TRANSPORT", "nccl" if used else "peer")
"""


@pytest.mark.skipif(_gpus() < 2, reason="needs two GPUs")
@pytest.mark.parametrize("transport", ["nccl", "peer"])
@pytest.mark.parametrize("cfg", [dict(), dict(decim=2, dc=True), dict(i16=True)])
def test_group_of_two_equals_one_gpu(cfg, transport):
    """each case in its own process: the transport is chosen when libb200dsp first looks for libnccl"""
    env = dict(os.environ)
    if transport == "peer":
        env["B200_GROUP_NO_NCCL"] = "1"
    r = subprocess.run([sys.executable, "-c", _CHILD % (ROOT, os.path.join(ROOT, "tests"), cfg)], capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert ("TRANSPORT " + transport) in r.stdout, r.stdout[-500:]
