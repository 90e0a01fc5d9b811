"""GPU parity: FM / AM / SSB demodulators + AGC fed by the VFO bank, against the oracle.
Tolerance: audio error at least 80 dB below the block's audio RMS; counts exact."""
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


def _demod_pair(name):
    from sdrplusplusbrown_b200 import binding as B
    if name == "demod_nfm":
        sr = 50e3
        return sr, ol.Fm(PORT, sr, 12.5e3, True, False), (B.DEMOD_NFM, 1, 0, 0, 0), 12.5e3
    if name == "demod_usb":
        sr = 24e3
        return sr, ol.Ssb(PORT, 0, 2800.0, sr, 50.0 / sr, 5.0 / sr), (B.DEMOD_USB, 0, 50.0 / sr, 5.0 / sr, 0), 2800.0
    if name == "demod_lsb":
        sr = 24e3
        return sr, ol.Ssb(PORT, 1, 2800.0, sr, 50.0 / sr, 5.0 / sr), (B.DEMOD_LSB, 0, 50.0 / sr, 5.0 / sr, 0), 2800.0
    if name == "demod_am":
        sr = 15e3
        return sr, ol.Am(PORT, 1, 10e3, 50.0 / sr, 5.0 / sr, 100.0 / sr, sr), (B.DEMOD_AM, 1, 50.0 / sr, 5.0 / sr, 100.0 / sr), 10e3
    if name == "demod_am_carrier":
        sr = 15e3
        return sr, ol.Am(PORT, 0, 10e3, 50.0 / sr, 5.0 / sr, 100.0 / sr, sr), (B.DEMOD_AM, 0, 50.0 / sr, 5.0 / sr, 100.0 / sr), 10e3
    raise KeyError(name)


DEMODS = ["demod_nfm", "demod_usb", "demod_lsb", "demod_am", "demod_am_carrier"]


@pytest.mark.parametrize("mode", ["exact", "fast"])
@pytest.mark.parametrize("name", DEMODS)
def test_vfo_plus_demod_chain_matches_oracle(name, mode):
    """VFO (channel filter on) -> demodulator, both against the oracle chain, blocks of ragged size.
    exact: free-running.  fast: the VFO's rotator phase is taken from the oracle per block; the SSB demodulator's
    own rotator runs in closed form (parallel prep / AGC walk / finish kernels)."""
    from sdrplusplusbrown_b200 import binding
    out_sr, o_dem, (kind, p0, p1, p2, p3), bw = _demod_pair(name)
    in_sr = out_sr * 32
    off = 37e3 + 11.0
    n = 32 * 9000
    x = synth.make_iq(0, n, in_sr, sigma=0.02, tones=[(off + 900.0, 0.3), (off - 1500.0, 0.05)])
    x[100000:100200] = 0
    o_vfo = ol.Vfo(PORT, in_sr, out_sr, bw, off)
    b = binding.VfoBank(in_sr, max_block=200000)
    if mode == "exact":
        b.set_rotator_mode(binding.ROTATOR_EXACT)
    v = b.add(out_sr, bw, off)
    b.set_demod(v, kind, p0, p1, p2, p3)
    for blk in split(x, [96000, 4800 * 32 + 17, 33]):
        if mode == "fast":
            b.set_phase(v, o_vfo.phase())
        iq = o_vfo.process(blk)
        ref = o_dem.process(iq) if len(iq) else np.zeros((0, 2), np.float32)
        b.process(to_dev(blk))
        assert b.out_count(v) == len(ref)
        if len(ref):
            got = b.audio(v)
            assert np.array_equal(got[:, 0], got[:, 1])
            e = err_db(got[:, 0], ref[:, 0])
            assert e <= -80.0, (name, e)


@pytest.mark.parametrize("name", DEMODS)
def test_demod_against_golden(name):
    """Demodulator alone (bypass VFO: inSR == outSR == bw would change the demod bandwidth, so the VFO here
    has the demod's bandwidth and the golden input is passed through the same oracle VFO first)."""
    from sdrplusplusbrown_b200 import binding
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    sr, o_dem, (kind, p0, p1, p2, p3), bw = _demod_pair(name)
    x = golden_cases.demod_input(sr)
    # run the committed golden through the *demod only*: compare via the oracle demod fed with the GPU VFO output
    b = binding.VfoBank(sr, max_block=10000)
    b.set_rotator_mode(binding.ROTATOR_EXACT)
    v = b.add(sr, bw, 0.0)            # NONE mode + channel filter (bw != sr)
    b.set_demod(v, kind, p0, p1, p2, p3)
    o_vfo = ol.Vfo(PORT, sr, sr, bw, 0.0)
    for blk in split(x, [2500]):
        iq = o_vfo.process(blk)
        ref = o_dem.process(iq)
        b.process(to_dev(blk))
        got = b.audio(v)
        assert len(got) == len(ref)
        assert err_db(got[:, 0], ref[:, 0]) <= -80.0
    # and the oracle demod itself is pinned to the golden file in tests/test_oracle_cpu.py
    assert g["mono"].shape == (6000,)


def test_agc_start_ramp_and_zero_input():
    """AGC quirks (agc.h:96-134): zero samples leave the average untouched and pass with gain 1; the 4800-sample
    start ramp keeps its last value for the rest of the call in which it ends."""
    from sdrplusplusbrown_b200 import binding
    sr = 24e3
    x = synth.make_iq(0, 12000, sr, sigma=0.01, tones=[(700.0, 0.2)])
    x[2000:2100] = 0
    o_vfo = ol.Vfo(PORT, sr, sr, 2800.0, 0.0)
    o_dem = ol.Ssb(PORT, 2, 2800.0, sr, 50.0 / sr, 5.0 / sr)     # DSB: no rotation, isolates the AGC
    b = binding.VfoBank(sr, max_block=10000)
    v = b.add(sr, 2800.0, 0.0)
    b.set_demod(v, binding.DEMOD_DSB, 0, 50.0 / sr, 5.0 / sr, 0)
    for blk in split(x, [4000, 1000, 7000]):     # the ramp ends inside the 2nd block
        ref = o_dem.process(o_vfo.process(blk))
        b.process(to_dev(blk))
        got = b.audio(v)
        assert err_db(got[:, 0], ref[:, 0]) <= -80.0


def test_many_demods_in_one_bank():
    from sdrplusplusbrown_b200 import binding
    in_sr = 2.4e6
    specs = [("demod_nfm", 100e3 + 7.0), ("demod_usb", -200e3 + 3.0), ("demod_am", 300e3 + 1.0), ("demod_lsb", -50e3 + 9.0), ("demod_am_carrier", 410e3)]
    x = synth.make_iq(0, 480000, in_sr, sigma=0.02, tones=[(o + 800.0, 0.1) for _, o in specs])
    b = binding.VfoBank(in_sr, max_block=240000)
    b.set_rotator_mode(binding.ROTATOR_EXACT)
    pairs = []
    for name, off in specs:
        out_sr, o_dem, (kind, p0, p1, p2, p3), bw = _demod_pair(name)
        v = b.add(out_sr, bw, off)
        b.set_demod(v, kind, p0, p1, p2, p3)
        pairs.append((v, ol.Vfo(PORT, in_sr, out_sr, bw, off), o_dem, name))
    for blk in split(x, [240000]):
        b.process(to_dev(blk))
        for v, o_vfo, o_dem, name in pairs:
            ref = o_dem.process(o_vfo.process(blk))
            got = b.audio(v)
            assert len(got) == len(ref)
            assert err_db(got[:, 0], ref[:, 0]) <= -80.0, name
