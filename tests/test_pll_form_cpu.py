"""The pilot PLL's error form (csrc/demod.cu, pllErrKernel) against the reference's recurrence (loop/pll.h:64-70,
phase_control_loop.h:58-66), both restated in float32 numpy (tools/pll_sim.py): the algebra the kernel rests on, checked
without a GPU.  The kernel itself is compared with the oracle in tests/test_demod_gpu.py."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
import pll_sim  # noqa: E402


@pytest.mark.parametrize("name,amp,noise,fp", pll_sim.SCENARIOS)
def test_error_form_tracks_the_reference_recurrence(name, amp, noise, fp):
    """Phases within 2e-6 rad of the reference's at every sample, in lock and out of lock, and the carried state too."""
    d, nfull, nchunks, dphase, dfreq = pll_sim.scenario(name, amp, noise, fp, n=20000)
    assert d.max() < 2e-6, (name, d.max(), int(d.argmax()))
    assert min(dphase, 2 * np.pi - dphase) < 2e-6 and dfreq < 1e-6
    if name == "clean pilot":
        assert nfull <= 3                      # a clean pilot is walked with the one-instruction chain
    if name in ("no pilot", "pilot off limits"):
        assert nfull == nchunks                # the clamp acts: every chunk needs it in the chain
