"""CPU tests of the logmmse oracle: the self-contained restatement (oracle/port_logmmse.cpp) against
the reference's own if_nr.h / logmmse.h / bgnoise.h / utils/arrays.cpp compiled in place into oracle/_ref
(oracle/ref_logmmse.cpp) -- bit for bit.  The committed vectors (tests/golden/logmmse_24000.npz) are checked by
test_oracle_cpu.py::test_oracles_match_golden.

Reference: misc_modules/noise_reduction_logmmse/src/if_nr.h:63-113, logmmse.h:117-143,150-242,244-296,298-405,
bgnoise.h:27-59, core/src/utils/arrays.cpp:60-431 (E1 table), :433-470, :1068-1092."""
import numpy as np
import pytest

import golden_cases
import oracle_lib as ol

PORT = ol.Oracle("port")
REF = ol.Oracle("ref") if ol.have_ref() else None
needs_ref = pytest.mark.skipif(REF is None, reason="oracle/_ref not built and /root/reference absent")


def test_exponential_integral_table_against_scipy():
    """the reference tabulates scipy.special.expn(1, x) at x = (i / 1000)^4 and interpolates linearly in x^(1/4)"""
    from scipy.special import exp1
    i = np.arange(1, 2515)
    x = ((i / 1000.0) ** 4).astype(np.float32)
    got = np.array([PORT.lib.orc_expn(float(v)) for v in x])
    # (the index x^(1/4) * 1000 is formed in fp32: up to 2.4e-4 off at the table's end, where E1 falls 6.5 % per entry)
    assert np.max(np.abs(got - exp1(x.astype(np.float64))) / exp1(x.astype(np.float64))) < 5e-5
    i = i[i >= 100]                                      # (below, E1 ~ -ln x bends too much for a chord in x^(1/4))
    mid = (((i + 0.5) / 1000.0) ** 4).astype(np.float32)
    got = np.array([PORT.lib.orc_expn(float(v)) for v in mid])
    assert np.max(np.abs(got - exp1(mid.astype(np.float64))) / exp1(mid.astype(np.float64))) < 2e-3     # the chord's error


@needs_ref
def test_exponential_integral_equals_reference_table():
    xs = np.concatenate([np.linspace(1e-9, 40, 100001).astype(np.float32), ((np.arange(1, 2515) / 1000.0) ** 4).astype(np.float32)])
    a = np.array([PORT.lib.orc_expn(float(v)) for v in xs], np.float32)
    b = np.array([REF.lib.orc_expn(float(v)) for v in xs], np.float32)
    assert np.array_equal(a, b)


@needs_ref
@pytest.mark.parametrize("sr,seconds,blk", [(48000, 3.0, 4800), (240000, 2.5, 12000), (31250, 3.0, 7001)])
def test_port_equals_reference_bit_for_bit(sr, seconds, blk):
    """long enough for the noise tracker to run (more than 100 frames kept, logmmse.h:166), a retune reset and a
    transmit hold in the middle"""
    x = golden_cases.logmmse_input(sr, seconds)
    a, b = ol.LogMmse(PORT, sr), ol.LogMmse(REF, sr)
    total = 0
    for k, i in enumerate(range(0, len(x), blk)):
        if k == 12:
            a.set_hold(True); b.set_hold(True)
        if k == 15:
            a.set_hold(False); b.set_hold(False)
        if i >= int(0.7 * len(x)) and i - blk < int(0.7 * len(x)):
            a.reset(); b.reset()
        ya, yb = a.process(x[i:i + blk]), b.process(x[i:i + blk])
        assert len(ya) == len(yb)
        assert np.array_equal(ya, yb), k
        total += len(ya)
    assert total > 0.5 * len(x)
    assert np.array_equal(a.noise(), b.noise()) and len(a.noise()) == 2 * (int(0.02 * sr) + int(0.02 * sr) % 2)


def test_noise_goes_down_and_the_carrier_stays():
    sr = 48000
    x = golden_cases.logmmse_input(sr, 4.0)
    d = ol.LogMmse(PORT, sr)
    y = np.concatenate([d.process(x[i:i + 4800]) for i in range(0, len(x), 4800)])
    X, Y = np.abs(np.fft.fft(x[-sr:])), np.abs(np.fft.fft(y[-sr:]))
    k = int(0.0625 * sr)
    assert 3.5 < Y[k] / X[k] < 4.5                       # the module's fixed gain of 4 (if_nr.h:106)
    assert np.median(Y) / np.median(X) < 4.0 * 0.25      # the floor: at least 12 dB below the carrier's gain
