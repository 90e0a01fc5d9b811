"""GPU parity of the logmmse front-end noise reduction against the oracle, through the C ABI.

Reference: misc_modules/noise_reduction_logmmse/src/if_nr.h:63-113 over logmmse.h:117-143,150-242,244-296,298-405 and
bgnoise.h:27-59 (the module's front-end pre-processor, main.cpp:165).

Output counts per call follow integer bookkeeping: bit-exact.  Samples: every per-bin recurrence runs the reference's
sequence of fp32 operations, the transforms (length 2 * Slen, not a power of two) differ from the oracle's float64 DFT
by fp32 rounding: each call's output at least 80 dB below its RMS (north star tolerance) against the oracle -- while the
noise tracker (a histogram mode, logmmse.h:219) takes the same decisions on both sides, which these signals are chosen
for; the noise estimate itself is compared too."""
import numpy as np
import pytest

import golden_cases
import oracle_lib as ol
from gpu_util import err_db, to_dev

pytestmark = pytest.mark.gpu

PORT = ol.Oracle("port")
TOL_DB = -80.0


@pytest.mark.parametrize("sr,seconds,blk", [(48000, 3.0, 4800), (240000, 2.5, 12000), (31250, 2.5, 7001), (2400000, 1.5, 120000)])
def test_stream_against_the_oracle(sr, seconds, blk):
    from sdrplusplusbrown_b200 import binding
    x = golden_cases.logmmse_input(sr, seconds)
    dev = binding.LogMmse(sr, max_block=blk)
    orc = ol.LogMmse(PORT, sr)
    assert dev.fft_size() == 2 * (int(0.02 * sr) + int(0.02 * sr) % 2)
    worst, total = -np.inf, 0
    for k, i in enumerate(range(0, len(x), blk)):
        ref = orc.process(x[i:i + blk])
        got = dev.process(to_dev(x[i:i + blk]))
        assert len(got) == len(ref), k
        if len(ref):
            e = err_db(got, ref)
            worst = max(worst, e)
            assert e <= TOL_DB, (k, e)
        total += len(ref)
    assert total > 0.4 * len(x)
    n_ref, n_got = orc.noise(), dev.noise()
    assert len(n_ref) == len(n_got) == dev.fft_size()
    ne = err_db(n_got, n_ref)
    print("sr %d: worst block error %.1f dB, noise estimate %.1f dB" % (sr, worst, ne))
    assert ne <= -60.0


def test_reset_hold_and_host_blocks():
    from sdrplusplusbrown_b200 import binding
    sr, blk = 48000, 4800
    x = golden_cases.logmmse_input(sr, 4.0, seed=5)
    dev = binding.LogMmse(sr, max_block=blk)
    orc = ol.LogMmse(PORT, sr)
    for k, i in enumerate(range(0, len(x), blk)):
        if k == 14:
            dev.set_hold(True); orc.set_hold(True)            # transmitting: the noise histories freeze (if_nr.h:46-48)
        if k == 18:
            dev.set_hold(False); orc.set_hold(False)
        if k == 26:
            dev.reset(); orc.reset()                          # retune: start over (if_nr.h:58-61, main.cpp:171)
        ref = orc.process(x[i:i + blk])
        got = dev.process_host(x[i:i + blk])
        assert len(got) == len(ref), k
        if len(ref):
            assert err_db(got, ref) <= TOL_DB, k
        if k in (26, 27):
            assert len(got) == 0                              # collecting again
    with pytest.raises(binding.B200Error):
        dev.process_host(np.zeros(blk + 1, np.complex64))     # beyond max_block
    with pytest.raises(binding.B200Error):
        binding.LogMmse(500000000)                            # transform of 20 M points: beyond the 2^24 served


def test_noise_goes_down_and_the_carrier_stays():
    from sdrplusplusbrown_b200 import binding
    sr = 240000
    x = golden_cases.logmmse_input(sr, 3.0)
    dev = binding.LogMmse(sr, max_block=24000)
    y = np.concatenate([dev.process(to_dev(x[i:i + 24000])) for i in range(0, len(x), 24000)])
    X, Y = np.abs(np.fft.fft(x[-sr:])), np.abs(np.fft.fft(y[-sr:]))
    k = int(0.0625 * sr)
    assert 3.5 < Y[k] / X[k] < 4.5
    assert np.median(Y) / np.median(X) < 1.0
