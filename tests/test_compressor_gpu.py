"""GPU parity of the server stream's spectral compressor against the oracle, through the C ABI.

Reference: core/src/dsp/compression/experimental_fft_compressor.h:66-88,116-189,198-427 (core/src/server.cpp:101-110).

Frame counts per call follow integer bookkeeping: bit-exact.  The noise-floor chain runs in the reference's summation
order on the device, so what separates the two sides is the rounding of the two transforms per frame (fp32 passes here,
float64 DFT in the oracle) and of log2f: a bin whose level sits within that rounding of its threshold may be kept on one
side and zeroed on the other, and the mask's 37-bin widening (:165) turns one such bin into up to 37.  Asserted per frame:
the kept / zeroed pattern agrees on >= 99 % of the bins (measured: 99.8-100 %), the kept bins'
values agree to -80 dB of the frame's RMS over the bins kept on both sides, the noise figure to 0.01 dB."""
import numpy as np
import pytest

import golden_cases
import oracle_lib as ol
from gpu_util import err_db, to_dev

pytestmark = pytest.mark.gpu

PORT = ol.Oracle("port")


@pytest.mark.parametrize("sr,loss", [(48000, 10.0), (240000, 10.0), (240000, 30.0), (2400000, 30.0), (192000, 1.0), (192000, 0.0)])
def test_frames_against_the_oracle(sr, loss):
    from sdrplusplusbrown_b200 import binding
    x = golden_cases.logmmse_input(sr, 1.6 if sr > 1e6 else 3.0, seed=3)
    dev = binding.FftCompressor(sr, loss, max_block=1 << 17)
    orc = ol.FftCompressor(PORT, sr, loss)
    n = dev.n
    assert n == orc.n
    if sr == 240000:
        dev.set_masked([-30000, -20000, 50000, 52000]); orc.set_masked([-30000, -20000, 50000, 52000])
    frames, worst_pat, worst_err, zeroed = 0, 1.0, -np.inf, []
    for k, i in enumerate(range(0, len(x) - n + 1, n)):
        ref = orc.process(x[i:i + n])
        got = dev.process(to_dev(x[i:i + n])) if k % 2 else dev.process_host(x[i:i + n])
        assert len(got) == len(ref) == (n if k >= 9 else 0), k
        if not len(ref):
            continue
        frames += 1
        za, zb = ref == 0, got == 0
        worst_pat = min(worst_pat, float(np.mean(za == zb)))
        both = ~za & ~zb
        if both.any():
            worst_err = max(worst_err, err_db(got[both], ref[both]))
        zeroed.append(float(zb.mean()))
    assert frames >= 3
    print("sr %d loss %g: %d frames, pattern agreement %.5f, kept-bin error %.1f dB, zeroed %.3f ... %.3f" %
          (sr, loss, frames, worst_pat, worst_err, zeroed[0], zeroed[-1]))
    assert worst_pat >= 0.99
    assert worst_err <= -80.0
    nf_ref, nf_got = orc.noise_figure(), dev.noise_figure()
    assert len(nf_ref) == len(nf_got) == (30 if loss > 0 else 0)
    if loss > 0:
        assert np.max(np.abs(nf_ref - nf_got)) <= 0.01


def test_blocks_smaller_and_larger_than_a_frame():
    """one frame leaves per call however much arrives (run() :431-452); the rest waits in the input buffer"""
    from sdrplusplusbrown_b200 import binding
    sr = 48000
    x = golden_cases.logmmse_input(sr, 2.5, seed=9)
    dev, orc = binding.FftCompressor(sr, 10.0, max_block=5000), ol.FftCompressor(PORT, sr, 10.0)
    pos, k, out_d, out_o = 0, 0, [], []
    sizes = [1000, 2048, 3000, 500, 2048, 2048]
    while pos < len(x) and k < 60:
        m = min(sizes[k % len(sizes)], len(x) - pos)
        a, b = dev.process_host(x[pos:pos + m]), orc.process(x[pos:pos + m])
        assert len(a) == len(b)
        out_d.append(a); out_o.append(b)
        pos += m
        k += 1
    got, ref = np.concatenate(out_d), np.concatenate(out_o)
    assert len(ref) >= 10 * 2048
    za, zb = ref == 0, got == 0
    assert np.mean(za == zb) >= 0.99
    assert err_db(got[~za & ~zb], ref[~za & ~zb]) <= -80.0
    with pytest.raises(binding.B200Error):
        dev.process_host(np.zeros(5001, np.complex64))
    with pytest.raises(binding.B200Error):
        binding.FftCompressor(8000)            # 256-bin frames: shorter than the 400-bin averages of the noise tracker
