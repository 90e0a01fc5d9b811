"""The allocation-free DC blocker entry point (b200_dc_block_noalloc) against the regular front end.

Who needs it: callers that may not call cudaMalloc while they run (hard real-time DSP threads, sandboxed
plug-ins with a fixed memory budget, processes whose device memory is owned by another allocator).  The entry
point keeps its scan scratch in a fixed array inside the library, so the checks here are: same output as
IQFrontEnd's DC blocker path on two consecutive blocks of one stream (the second call reuses the array the
first one dirtied and continues from the carried offset), a fresh start when no offset is passed, and a clear
refusal of blocks beyond the fixed limit.  Runs last (file name) so a failure here cannot hide other tests."""
import numpy as np
import pytest

from gpu_util import err_db, to_dev
from sdrplusplusbrown_b200 import synth

pytestmark = pytest.mark.gpu

TOL_DB = -80.0      # same tolerance as the other pre-processing parity tests


def test_noalloc_dc_block_matches_frontend_on_two_calls_in_a_row():
    import torch
    from sdrplusplusbrown_b200 import binding
    sr = 8e6
    n = 300000
    blocks = [synth.make_iq(i * n, n, sr, tones=[(1e6, 0.1)], dc=(0.05, -0.08)) for i in range(2)]
    # existing entry point: the front end with DC blocking on, state carried inside the handle
    fe = binding.FrontEnd(sr, dc_blocking=True, fft_size=1024, fft_rate=20.0, max_block=n)
    want = []
    for blk in blocks:
        fe.process(to_dev(blk))
        want.append(fe.preproc_out())
    # new entry point: caller-owned input, output and carried offset; nothing else
    d_off = torch.zeros(2, dtype=torch.float64, device="cuda")
    for blk, ref in zip(blocks, want):
        d_in = to_dev(blk)
        d_out = torch.empty_like(d_in)
        binding.dc_block_noalloc(d_in, d_out, n, 50.0 / sr, d_off)
        torch.cuda.synchronize()
        got = d_out.cpu().numpy().view(np.complex64).reshape(-1)
        assert err_db(got, ref) <= TOL_DB
    assert abs(float(d_off[0]) - 0.05) < 5e-3 and abs(float(d_off[1]) + 0.08) < 5e-3   # settled on the DC level


def test_noalloc_dc_block_without_offset_starts_from_zero_each_call():
    import torch
    from sdrplusplusbrown_b200 import binding
    sr = 2.4e6
    x = synth.make_iq(0, 50001, sr, dc=(0.2, 0.1))
    fe = binding.FrontEnd(sr, dc_blocking=True, fft_size=1024, fft_rate=20.0, max_block=len(x))
    fe.process(to_dev(x))
    ref = fe.preproc_out()
    d_in = to_dev(x)
    for _ in range(2):      # the second call must not see the first call's state
        d_out = torch.empty_like(d_in)
        binding.dc_block_noalloc(d_in, d_out, len(x), 50.0 / sr, None)
        torch.cuda.synchronize()
        assert err_db(d_out.cpu().numpy().view(np.complex64).reshape(-1), ref) <= TOL_DB


def test_noalloc_dc_block_refuses_oversized_blocks():
    import torch
    from sdrplusplusbrown_b200 import binding
    d = torch.zeros((16, 2), dtype=torch.float32, device="cuda")
    with pytest.raises(binding.B200Error) as e:
        binding.dc_block_noalloc(d, d, binding.DC_NOALLOC_MAX_SAMPLES + 1, 1e-5, None)
    assert "exceeds the fixed scratch limit" in str(e.value)
    binding.dc_block_noalloc(d, d, 0, 1e-5, None)    # an empty block is fine


def test_noalloc_dc_block_at_the_size_limit_in_place_against_the_oracle():
    """The largest block the fixed array serves (its last chunk slot), run in place (d_out == d_in), and checked against
    the CPU oracle's DCBlocker itself rather than against the front end; then a second, ragged block continues the
    stream from the carried offset."""
    import torch
    import oracle_lib as ol
    from sdrplusplusbrown_b200 import binding
    sr = 20e6
    rate = 50.0 / sr
    n = binding.DC_NOALLOC_MAX_SAMPLES
    port = ol.Oracle("port")
    o = ol.DcBlock(port, rate)
    d_off = torch.zeros(2, dtype=torch.float64, device="cuda")
    for k, m in enumerate((n, 2049)):
        x = synth.make_iq(k * n, m, sr, dc=(0.2, -0.1))
        ref = o.process(x)
        d = to_dev(x)
        binding.dc_block_noalloc(d, d, m, rate, d_off)          # in place
        torch.cuda.synchronize()
        got = d.cpu().numpy().view(np.complex64).reshape(-1)
        assert err_db(got, ref) <= TOL_DB
        # the last samples of the block went through the array's last chunk
        assert err_db(got[-2048:], ref[-2048:]) <= TOL_DB


def test_noalloc_dc_block_refuses_rates_the_scan_cannot_represent():
    import torch
    from sdrplusplusbrown_b200 import binding
    d = torch.zeros((16, 2), dtype=torch.float32, device="cuda")
    for bad in (-1e-3, 0.5, float("nan")):
        with pytest.raises(binding.B200Error):
            binding.dc_block_noalloc(d, d, 16, bad, None)
    binding.dc_block_noalloc(d, d, 16, 0.0, None)            # rate 0: out = in - offset, offset never moves
