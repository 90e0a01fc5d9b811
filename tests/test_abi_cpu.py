"""CPU-side checks of the drop-in boundary: the C-ABI library builds, loads and exports every
symbol include/b200dsp.h declares; without a GPU every entry point that needs one fails loudly
(there is no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    txt = open(os.path.join(ROOT, "include", "b200dsp.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    names = re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", txt)
    return sorted(set(names))


def test_library_exports_every_declared_symbol():
    from sdrplusplusbrown_b200 import binding
    names = _declared()
    assert len(names) >= 40
    for n in names:
        assert hasattr(binding.lib, n), "libb200dsp.so does not export %s" % n
    # and the binding knows every one of them
    assert set(names) == set(binding.EXPORTS), set(names) ^ set(binding.EXPORTS)


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from sdrplusplusbrown_b200 import binding
    with pytest.raises(binding.B200Error) as e:
        binding.VfoBank(2.4e6)
    assert "no CUDA device" in str(e.value) or "CUDA" in str(e.value)
    with pytest.raises(binding.B200Error):
        binding.FrontEnd(2.4e6, fft_size=8192)


def test_product_does_not_reference_oracle():
    # the shipped sources must never include, link or load anything under oracle/
    import glob
    pat = re.compile(r"oracle/|libport_oracle|libref_oracle|oracle_lib|orc_[a-z]+")
    for path in glob.glob(os.path.join(ROOT, "sdrplusplusbrown_b200", "**", "*"), recursive=True) + glob.glob(os.path.join(ROOT, "include", "**", "*"), recursive=True):
        if os.path.isdir(path) or path.endswith((".so", ".o", ".pyc", ".log")):
            continue
        txt = open(path, errors="ignore").read()
        for line in txt.splitlines():
            s = line.strip()
            if s.startswith(("//", "#", "*", "/*", '"""')) or "tests/" in s:
                continue
            assert not pat.search(s), (path, s)


def test_host_design_matches_oracle_taps():
    """The product's host-side designers (csrc/host_design.h) against the oracle, no GPU needed:
    compiled into a tiny standalone program that dumps taps/windows/state-machine walks."""
    import subprocess
    import tempfile

    import numpy as np

    import oracle_lib as ol
    src = r'''
#include <cstdio>
#include "host_design.h"
using namespace b200::design;
static void dump(const char* name, const std::vector<float>& v) { printf("%s %zu", name, v.size()); for (float f : v) { uint32_t u; memcpy(&u, &f, 4); printf(" %08x", u); } printf("\n"); }
int main() {
    dump("lp_nfm", lowPass(6250.0, 625.0, 50e3));
    dump("lp_usb", lowPass(1400.0, 140.0, 24e3));
    dump("win_blackman", fftWindow(1, 1000));
    dump("win_nuttall", fftWindow(2, 777));
    dump("win_rect", fftWindow(0, 16));
    ResampPlan p = planResampler(61.44e6, 48e3);
    printf("plan %d %d %d %d %zu\n", p.mode, p.ratio, p.interp, p.decim, p.taps.size());
    dump("resamp", p.taps);
    p = planResampler(20e6, 24e3);
    printf("plan %d %d %d %d %zu\n", p.mode, p.ratio, p.interp, p.decim, p.taps.size());
    // polyphase walk: L=8000 M=13021 over ragged blocks
    int phase = 0, offset = 0, total = 0;
    int blocks[] = { 39063, 1, 0, 977, 39062, 5 };
    for (int n : blocks) { total += polyOutCount(phase, offset, 8000, 13021, n); polyNextState(phase, offset, 8000, 13021, n); }
    printf("poly %d %d %d\n", total, phase, offset);
    int off = 0, tot = 0;
    int blocks2[] = { 12000, 333, 1, 0, 65536 };
    for (int n : blocks2) { tot += decimOutCount(off, 64, n); off = decimNextOffset(off, 64, n); }
    printf("decim %d %d\n", tot, off);
    const PlanDef* pl = findPlan(1024);
    printf("plan1024 %d %d %d %d\n", pl->nstages, pl->stages[0].decim, pl->stages[0].ntaps, pl->stages[2].ntaps);
    return 0;
}
'''
    with tempfile.TemporaryDirectory() as td:
        open(os.path.join(td, "t.cpp"), "w").write(src)
        exe = os.path.join(td, "t")
        subprocess.check_call(["g++", "-O1", "-ffp-contract=off", "-std=c++17", "-I", os.path.join(ROOT, "sdrplusplusbrown_b200", "csrc"),
                               os.path.join(td, "t.cpp"), "-o", exe])
        out = subprocess.check_output([exe], text=True).splitlines()
    port = ol.Oracle("port")
    vals = {}
    plans = []
    for line in out:
        parts = line.split()
        if parts[0] == "plan":
            plans.append([int(x) for x in parts[1:]])
        elif parts[0] in ("poly", "decim", "plan1024"):
            vals[parts[0]] = [int(x) for x in parts[1:]]
        else:
            vals[parts[0]] = np.array([int(x, 16) for x in parts[2:]], np.uint32)
    assert np.array_equal(vals["lp_nfm"], port.lowpass(6250.0, 625.0, 50e3).view(np.uint32))
    assert np.array_equal(vals["lp_usb"], port.lowpass(1400.0, 140.0, 24e3).view(np.uint32))
    assert np.array_equal(vals["win_blackman"], port.fft_window(1, 1000).view(np.uint32))
    assert np.array_equal(vals["win_nuttall"], port.fft_window(2, 777).view(np.uint32))
    assert np.array_equal(vals["win_rect"], port.fft_window(0, 16).view(np.uint32))
    v = ol.Vfo(port, 61.44e6, 48e3, 12.5e3, 0.0)
    d = v.describe()
    assert plans[0] == [d["mode"], d["ratio"], d["interp"], d["decim"], d["resamp_taps"]]
    assert np.array_equal(vals["resamp"], v.resamp_taps().view(np.uint32))
    v2 = ol.Vfo(port, 20e6, 24e3, 2.8e3, 0.0)
    d2 = v2.describe()
    assert plans[1] == [d2["mode"], d2["ratio"], d2["interp"], d2["decim"], d2["resamp_taps"]]
    assert vals["plan1024"] == [3, 64, 257, 69]
    # integer walks against the oracle's own loops (decimating FIR /64 and polyphase 8000/13021 via real blocks)
    dec = ol.Decim(port, 2)   # any decimator exercises the same loop; /64 single stage is checked through counts below
    n_out = 0
    off = 0
    for n in [12000, 333, 1, 0, 65536]:
        k = 0
        while off < n:
            off += 64
            k += 1
        off -= n
        n_out += k
    assert vals["decim"] == [n_out, off]
    ph, of, tot = 0, 0, 0
    for n in [39063, 1, 0, 977, 39062, 5]:
        while of < n:
            tot += 1
            ph += 13021
            of += ph // 8000
            ph %= 8000
        of -= n
    assert vals["poly"] == [tot, ph, of]
