"""CPU tests of the ExperimentalFFTCompressor oracle (the server stream's spectral compressor): the
self-contained restatement (oracle/port_compressor.cpp) against the reference's own header + utils/arrays.cpp compiled
in place into oracle/_ref (oracle/ref_compressor.cpp) -- bit for bit.  The committed vectors
(tests/golden/fftcomp_48000.npz) are checked by test_oracle_cpu.py::test_oracles_match_golden.

Reference: core/src/dsp/compression/experimental_fft_compressor.h:66-88,116-189,198-427; core/src/server.cpp:101-110."""
import numpy as np
import pytest

import golden_cases
import oracle_lib as ol

PORT = ol.Oracle("port")
REF = ol.Oracle("ref") if ol.have_ref() else None
needs_ref = pytest.mark.skipif(REF is None, reason="oracle/_ref not built and /root/reference absent")


@needs_ref
@pytest.mark.parametrize("sr,loss", [(48000, 1.0), (240000, 10.0), (2400000, 30.0), (192000, 0.0)])
def test_port_equals_reference_bit_for_bit(sr, loss):
    x = golden_cases.logmmse_input(sr, 1.6 if sr > 1e6 else 3.0, seed=3)
    a, b = ol.FftCompressor(PORT, sr, loss), ol.FftCompressor(REF, sr, loss)
    assert a.n == b.n == 2 ** int(np.floor(np.log2(sr * 50 // 1000)))
    if sr == 240000:
        a.set_masked([-30000, -20000, 50000, 52000]); b.set_masked([-30000, -20000, 50000, 52000])
    n, frames = a.n, 0
    for k, i in enumerate(range(0, len(x) - a.n + 1, a.n)):
        blk = x[i:i + n] if k % 3 else x[i:i + n]          # one frame per call, as the server's resampler delivers them
        ya, yb = a.process(blk), b.process(blk)
        assert len(ya) == len(yb) == (n if k >= 9 else 0), k
        assert np.array_equal(ya, yb), k
        frames += len(ya) // n
    assert frames >= 3
    assert np.array_equal(a.noise_figure(), b.noise_figure())
    assert len(a.noise_figure()) == (30 if loss > 0 else 0)


def test_carrier_bins_survive_noise_bins_are_zeroed_and_amplitudes_are_companded():
    sr = 240000
    x = golden_cases.logmmse_input(sr, 3.0, seed=4)
    c = ol.FftCompressor(PORT, sr, 30.0)          # (at loss rate 1 this signal's noise bins all pass: the mask is widened by 37 bins)
    n = c.n
    frames = [c.process(x[i:i + n]) for i in range(0, len(x) - n + 1, n)]
    last = frames[-1]
    assert len(last) == n
    zero = last == 0
    assert 0.5 < zero.mean() < 0.999                      # most of the band is noise
    k = n // 2 + int(round(0.0625 * sr / (sr / n)))       # the carrier, DC in the middle after swapfft
    assert not zero[k]
    # |out| = |X|^(1/4): the frame that went out is the one from ten calls earlier
    j = len(frames) - 1 - 9
    X = np.fft.fftshift(np.fft.fft(x[j * n:(j + 1) * n].astype(np.complex128)))
    assert abs(abs(last[k]) - abs(X[k]) ** 0.25) < 1e-3 * abs(X[k]) ** 0.25
