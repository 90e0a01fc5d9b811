"""CPU tests of the SignalDetector oracle: the self-contained restatement
(oracle/port_detector.cpp) against the reference's own signal_detector.cpp + utils/arrays.cpp compiled in place into
oracle/_ref (oracle/ref_detector.cpp) -- bit for bit -- and the float64 DFT both use (oracle/fft64.h) against numpy.
The committed vectors (tests/golden/sigdet_51200.npz) are checked by test_oracle_cpu.py::test_oracles_match_golden.

Reference: core/src/dsp/detector/signal_detector.cpp (run :359-414, getLineCandidates :151-267,
addSingleFFTRow :454-481, aggregateAndDetect :416-452)."""
import numpy as np
import pytest

import golden_cases
import oracle_lib as ol

PORT = ol.Oracle("port")
REF = ol.Oracle("ref") if ol.have_ref() else None
needs_ref = pytest.mark.skipif(REF is None, reason="oracle/_ref not built and /root/reference absent")


def _numpy_row(x):
    n = len(x)
    k = np.arange(n)
    w = (0.42 - 0.5 * np.cos(2 * np.pi * k / n) + 0.08 * np.cos(4 * np.pi * k / n)).astype(np.float32)
    xw = (x.real * w).astype(np.float64) + 1j * (x.imag * w).astype(np.float64)
    X = np.fft.fft(xw).astype(np.complex64)
    half = n // 2
    Y = X.copy()
    Y[:half], Y[half:2 * half] = X[half:2 * half], X[:half]          # swapfft: an odd last bin stays in place
    return 20.0 * np.log10(np.abs(Y).astype(np.float64) + 1e-10)


@pytest.mark.parametrize("n", [5120, 10001, 24000, 4099])     # 2^10*5, 73*137 (odd), 2^6*3*5^3, prime
def test_row_against_numpy(n):
    x = golden_cases.sigdet_input(n * 10.0, 1, seed=n)
    assert len(x) == n
    row = ol.sigdet_row(PORT, x)
    ref = _numpy_row(x)
    strong = ref > ref.max() - 100.0
    assert np.abs(row - ref)[strong].max() < 2e-3


@needs_ref
@pytest.mark.parametrize("sr", [51200.0, 100010.0, 240000.0])
def test_port_equals_reference_bit_for_bit(sr):
    rows = 24
    x = golden_cases.sigdet_input(sr, rows, seed=int(sr) % 97)
    n = int(sr * (1 / 10.0))
    assert np.array_equal(ol.sigdet_row(PORT, x[:n]), ol.sigdet_row(REF, x[:n]))
    row = ol.sigdet_row(REF, x[n:2 * n])
    assert np.array_equal(ol.sigdet_candidates(PORT, row), ol.sigdet_candidates(REF, row))
    a, b = ol.SigDet(PORT, sr), ol.SigDet(REF, sr)
    assert a.n == b.n == n
    pos, k = 0, 0
    sizes = [n // 3 + 1, 2 * n + 7, 1000, n]
    while pos < len(x):
        m = min(sizes[k % len(sizes)], len(x) - pos)
        ra, rb = a.process(x[pos:pos + m]), b.process(x[pos:pos + m])
        assert ra == rb
        pos += m
        k += 1
        if k == 9:
            a.set_center_frequency(7.1e6)         # clears the sample buffer; the row history stays
            b.set_center_frequency(7.1e6)
    assert ra >= rows - 2
    assert np.array_equal(a.history(), b.history())
    ra, rb = a.result(), b.result()
    assert len(ra) == n and np.array_equal(ra, rb)


def test_detection_needs_twenty_one_rows_and_runs_every_tenth():
    sr = 51200.0
    x = golden_cases.sigdet_input(sr, 45)
    n = 5120
    d = ol.SigDet(PORT, sr)
    seen = []
    for r in range(45):
        d.process(x[r * n:(r + 1) * n])
        seen.append(len(d.result()) > 0)
    # MIN_DETECT_FFT_ROWS = 20 (strictly more), DETECT_INTERVAL_FRAMES = 10: first detection with row 21 (index 20)
    assert seen.index(True) == 20
    # detections ran with rows 21, 31 and 41; rows 46..50 leave the result alone, row 51 renews it
    before = d.result().copy()
    d.process(x[:n * 5])
    assert np.array_equal(before, d.result())
    d.process(x[n * 5:n * 6])
    assert not np.array_equal(before, d.result())
    # marked line positions carry -normalised level (positive above the local mean is negative here), the rest -20
    h = d.history()
    assert h.shape == (51, n)
    assert np.count_nonzero(h != -20.0) > 0
