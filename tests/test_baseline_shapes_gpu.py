"""GPU parity at the launch shapes of BASELINE configs 3, 4 and 5 (the banks every many-VFO number is quoted on).

The small-bank tests in test_vfo_gpu.py run the grouped first-stage kernel with two VFOs per CTA on 150 k-sample
blocks.  Here the whole bank of a configuration is built from `configs.vfo_list` and fed 1 M-sample blocks, so the
first stage runs in the shape the bench runs it in (eight VFOs per CTA, two per register pass where the plan allows
it, the 8192-sample tile of the /8192 plan, hundreds of groups per launch) and the later stages run as launches of
256 ... 4096 jobs.  A spread of VFOs (first, last, every k-th, at least one per mode) is compared with its own
single-VFO oracle: counts bit-exact, output at least 80 dB below the block's RMS (north star tolerance) -- FAST with
the phase taken from the oracle at each block start, EXACT free-running with the carried phase equal to
the oracle's bit for bit.

Reference: dsp::channel::RxVFO::process (core/src/dsp/channel/rx_vfo.h:89-100) per VFO behind one Splitter
(routing/splitter.h:59-92)."""
import numpy as np
import pytest

import oracle_lib as ol
from gpu_util import err_db, to_dev
from sdrplusplusbrown_b200 import configs, synth

pytestmark = pytest.mark.gpu

TOL_DB = -80.0
PORT = ol.Oracle("port")

BANKS = [
    # config, VFO count, which VFOs get an oracle
    ("cfg3", 256, lambda n: sorted(set([0, n - 1] + list(range(5, n, 16))))),
    ("cfg4", 1024, lambda n: sorted(set([0, 1, 7, n - 1, n - 2] + list(range(3, n, 40)) + list(range(4, n, 72)) + list(range(15, n, 128))))),
    ("cfg3", 4096, lambda n: sorted(set([0, n - 1] + list(range(100, n, 300))))),     # cfg5's largest bank
]


@pytest.mark.parametrize("mode", ["fast", "exact"])
@pytest.mark.parametrize("bank", BANKS, ids=["cfg3_256", "cfg4_1024", "cfg5_4096"])
def test_whole_bank_at_bench_block_sizes(bank, mode):
    from sdrplusplusbrown_b200 import binding
    cfg_name, count, pick = bank
    sr = configs.CONFIGS[cfg_name]["sample_rate"]
    if count > 256 and cfg_name == "cfg3":
        # cfg5 sweeps the NFM bank past the 256 grid points of cfg3: same modes, offsets spread over the band
        rng = np.random.default_rng(4096)
        vfos = [("NFM48", 48000.0, 12500.0, float(f)) for f in np.sort(rng.uniform(-0.45 * sr, 0.45 * sr, count))]
    else:
        vfos = configs.vfo_list(cfg_name, count)
    assert len(vfos) == count
    checked = pick(count)
    assert {vfos[i][0] for i in checked} == {v[0] for v in vfos}        # every mode of the bank has an oracle
    sizes = [1000000, 1 << 20]        # the reference's stream block, then a power of two (another carried offset parity)
    exact = mode == "exact"
    b = binding.VfoBank(sr, max_block=max(sizes))
    if exact:
        b.set_rotator_mode(binding.ROTATOR_EXACT)
    ids = [b.add(o, bw, off) for (_, o, bw, off) in vfos]
    oracles = {i: ol.Vfo(PORT, sr, vfos[i][1], vfos[i][2], vfos[i][3]) for i in checked}
    tones = [(vfos[i][3] + 0.1 * vfos[i][2], 0.02) for i in checked]
    pos, worst = 0, -np.inf
    for n in sizes:
        x = synth.make_iq(pos, n, sr, tones=tones)
        pos += n
        if not exact:
            for i in checked:
                b.set_phase(ids[i], oracles[i].phase())
        b.process(to_dev(x))
        for i in checked:
            ref = oracles[i].process(x)
            assert b.out_count(ids[i]) == len(ref), (i, vfos[i])
            e = err_db(b.output(ids[i]), ref)
            worst = max(worst, e)
            assert e <= TOL_DB, (i, vfos[i], e)
            if exact:
                p, q = b.get_phase(ids[i]), oracles[i].phase()
                assert np.float32(p.real) == np.float32(q.real) and np.float32(p.imag) == np.float32(q.imag)
    # the VFOs without an oracle: counts must follow the same integer state machine as a checked VFO of their mode
    by_mode = {}
    for i in checked:
        by_mode.setdefault(vfos[i][0], b.out_count(ids[i]))
    for i in range(count):
        assert b.out_count(ids[i]) == by_mode[vfos[i][0]], i
    print("%s x%d %s: %d VFOs checked, worst block error %.1f dB" % (cfg_name, count, mode, len(checked), worst))


def test_cfg4_bank_with_demodulators_at_bench_block_size():
    """cfg4's mixed bank with the radio demodulators attached (AM, USB, mono WFM) on a 1 M-sample block: audio of a
    spread of VFOs against the oracle's demodulators fed with the oracle's VFO output.  EXACT rotator, free-running."""
    from sdrplusplusbrown_b200 import binding
    sr = configs.CONFIGS["cfg4"]["sample_rate"]
    vfos = configs.vfo_list("cfg4", 256)
    b = binding.VfoBank(sr, max_block=1000000)
    b.set_rotator_mode(binding.ROTATOR_EXACT)
    ids = []
    for (m, o, bw, off) in vfos:
        v = b.add(o, bw, off)
        if m == "AM":
            b.set_demod(v, binding.DEMOD_AM, 1, 50.0 / o, 5.0 / o, 100.0 / o)
        elif m == "USB":
            b.set_demod(v, binding.DEMOD_USB, 0, 50.0 / o, 5.0 / o, 0)
        else:
            b.set_demod(v, binding.DEMOD_WFM, 0, 1, 0, 0)
        ids.append(v)
    checked = [0, 1, 7, 100, 129, 255, 254]
    tones = [(vfos[i][3] + 0.1 * vfos[i][2], 0.05) for i in checked]
    x = synth.make_iq(0, 1000000, sr, tones=tones)
    b.process(to_dev(x))
    for i in checked:
        m, o, bw, off = vfos[i]
        ov = ol.Vfo(PORT, sr, o, bw, off)
        y = ov.process(x)
        if m == "AM":
            d = ol.Am(PORT, 1, bw, 50.0 / o, 5.0 / o, 100.0 / o, o)
        elif m == "USB":
            d = ol.Ssb(PORT, 0, bw, o, 50.0 / o, 5.0 / o)
        else:
            d = ol.Wfm(PORT, bw / 2.0, o, False, True)       # deviation = bandwidth / 2 (radio/src/demodulators/wfm.h:76)
        assert b.out_count(ids[i]) == len(y)
        assert err_db(b.output(ids[i]), y) <= TOL_DB
        ref = d.process(y)
        got = b.audio(ids[i])
        assert got.shape[0] == ref.shape[0]
        assert err_db(got[:, 0], ref[:, 0]) <= TOL_DB, (i, m)


@pytest.mark.parametrize("mode", ["fast", "exact"])
def test_wfm_channels_on_long_blocks_take_the_shared_sample_resampler(mode):
    """cfg4's WFM channels (122.88 MS/s -> /128 -> 25/48 polyphase, 146 taps per phase -> 500 kS/s) on blocks long enough
    (>= 4800 outputs) for the polyphase path in which a thread computes three outputs of a residue class from one walk
    over their joint window (fir_jobs.cu: sharedPoly) -- the path the 16 Mi-sample bench blocks run.  Two ragged blocks,
    so the carried resampler phase / offset enter the second one; against the oracle per VFO."""
    from sdrplusplusbrown_b200 import binding
    sr = configs.CONFIGS["cfg4"]["sample_rate"]
    specs = [(500e3, 150e3, -20e6 + 11.0), (500e3, 150e3, 7.3e6), (500e3, 150e3, 41e6 - 5.0), (15e3, 10e3, 1.1e6), (24e3, 2.8e3, -3.3e6)]
    sizes = [(1 << 21) + 12345, 1500000]
    exact = mode == "exact"
    b = binding.VfoBank(sr, max_block=max(sizes))
    if exact:
        b.set_rotator_mode(binding.ROTATOR_EXACT)
    ids = [b.add(*sp) for sp in specs]
    oracles = [ol.Vfo(PORT, sr, *sp) for sp in specs]
    d = b.describe(ids[0])
    assert (d["interp"], d["decim"], d["taps_per_phase"]) == (25, 48, 146)
    pos = 0
    for n in sizes:
        x = synth.make_iq(pos, n, sr, tones=[(sp[2] + 0.1 * sp[1], 0.05) for sp in specs])
        pos += n
        if not exact:
            for i, o in zip(ids, oracles):
                b.set_phase(i, o.phase())
        b.process(to_dev(x))
        for i, o in zip(ids, oracles):
            ref = o.process(x)
            assert b.out_count(i) == len(ref)
            assert err_db(b.output(i), ref) <= TOL_DB
    assert b.out_count(ids[0]) >= 4800
