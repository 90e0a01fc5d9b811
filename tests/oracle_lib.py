"""ctypes binding of oracle/oracle_api.h (TEST INFRASTRUCTURE — checker only).

Two builds export the same API:
  * ``port``: oracle/libport_oracle.so — self-contained restatement, builds anywhere.
  * ``ref`` : oracle/_ref/libref_oracle.so — the reference's own headers compiled in
    place (needs /root/reference at build time; the built .so travels to the GPU box).
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
PORT_SO = os.path.join(ORACLE_DIR, "libport_oracle.so")
REF_SO = os.path.join(ORACLE_DIR, "_ref", "libref_oracle.so")
REF_ROOT = os.environ.get("REF_ROOT", "/root/reference")

_vp, _i, _d, _f = C.c_void_p, C.c_int, C.c_double, C.c_float

_SIGS = {
    "orc_kind": (C.c_char_p, []),
    "orc_vfo_create": (_vp, [_d, _d, _d, _d]),
    "orc_vfo_destroy": (None, [_vp]),
    "orc_vfo_process": (_i, [_vp, _i, _vp, _vp]),
    "orc_vfo_set_offset": (None, [_vp, _d]),
    "orc_vfo_set_bandwidth": (None, [_vp, _d]),
    "orc_vfo_set_in_samplerate": (None, [_vp, _d]),
    "orc_vfo_set_out_samplerate": (None, [_vp, _d, _d]),
    "orc_vfo_reset": (None, [_vp]),
    "orc_vfo_get_phase": (None, [_vp, _vp]),
    "orc_vfo_set_phase": (None, [_vp, _f, _f]),
    "orc_vfo_get_phase_delta": (None, [_vp, _vp]),
    "orc_vfo_describe": (None, [_vp, _vp]),
    "orc_vfo_get_resamp_taps": (_i, [_vp, _vp, _i]),
    "orc_vfo_get_chan_taps": (_i, [_vp, _vp, _i]),
    "orc_decim_create": (_vp, [_i]),
    "orc_decim_destroy": (None, [_vp]),
    "orc_decim_process": (_i, [_vp, _i, _vp, _vp]),
    "orc_decim_plan": (_i, [_i, _vp, _vp, _i]),
    "orc_decim_stage_taps": (_i, [_i, _i, _vp, _i]),
    "orc_dcblock_create": (_vp, [_d]),
    "orc_dcblock_destroy": (None, [_vp]),
    "orc_dcblock_process": (_i, [_vp, _i, _vp, _vp]),
    "orc_preproc_create": (_vp, [_d, _i, _i, _i]),
    "orc_preproc_destroy": (None, [_vp]),
    "orc_preproc_process": (_i, [_vp, _i, _vp, _vp]),
    "orc_lowpass_taps": (_i, [_d, _d, _d, _vp, _i]),
    "orc_fft_window": (None, [_i, _i, _vp]),
    "orc_reshape_params": (None, [_d, _i, _d, _vp, _vp]),
    "orc_fft_frame": (None, [_vp, _vp, _i, _i, _vp, _vp]),
    "orc_do_zoom": (None, [_i, _i, _i, _i, _vp, _vp]),
    "orc_zoom_indices": (None, [_i, _i, _i, _i, _vp, _vp]),
    "orc_palette_indices": (None, [_vp, _i, _f, _f, _vp]),
    "orc_make_palette": (None, [_vp, _i, _vp]),
    "orc_vfo_signal_info": (_i, [_vp, _i, _d, _d, _d, _vp, _vp]),
    "orc_smooth_hold": (None, [_vp, _vp, _vp, _i, _i, _f, _i, _f]),
    "orc_fm_create": (_vp, [_d, _d, _i, _i]),
    "orc_fm_destroy": (None, [_vp]),
    "orc_fm_process": (_i, [_vp, _i, _vp, _vp]),
    "orc_ssb_create": (_vp, [_i, _d, _d, _d, _d]),
    "orc_ssb_destroy": (None, [_vp]),
    "orc_ssb_process": (_i, [_vp, _i, _vp, _vp]),
    "orc_am_create": (_vp, [_i, _d, _d, _d, _d, _d]),
    "orc_am_destroy": (None, [_vp]),
    "orc_am_process": (_i, [_vp, _i, _vp, _vp]),
    "orc_wfm_create": (_vp, [_d, _d, _i, _i]),
    "orc_wfm_destroy": (None, [_vp]),
    "orc_wfm_process": (_i, [_vp, _i, _vp, _vp]),
    "orc_wfm_create_rds": (_vp, [_d, _d, _i, _i]),
    "orc_wfm_process_rds": (_i, [_vp, _i, _vp, _vp, _vp]),
    "orc_fmif_create": (_vp, [_i]),
    "orc_fmif_destroy": (None, [_vp]),
    "orc_fmif_process": (_i, [_vp, _i, _vp, _vp]),
    "orc_volume_process": (_i, [_d, _i, _i, _vp, _vp]),
    "orc_stream_compress": (_i, [_i, _i, _vp, _vp]),
    "orc_stream_decompress": (_i, [_i, _vp, _vp]),
    "orc_wav_write": (_i, [C.c_char_p, _i, _d, _i, _vp, _i, _i]),
    "orc_ifchain_create": (_vp, [_i, _d, _d, _i, _d]),
    "orc_ifchain_destroy": (None, [_vp]),
    "orc_ifchain_process": (_i, [_vp, _i, _vp, _vp]),
    "orc_afpost_create": (_vp, [_d, _d, _d]),
    "orc_afpost_destroy": (None, [_vp]),
    "orc_afpost_process": (_i, [_vp, _i, _vp, _vp]),
    "orc_agc_create": (_vp, [_d, _d, _d, _d, _d, _d]),
    "orc_agc_destroy": (None, [_vp]),
    "orc_agc_process": (_i, [_vp, _i, _vp, _vp]),
    "orc_sigdet_create": (_vp, [_d]),
    "orc_sigdet_destroy": (None, [_vp]),
    "orc_sigdet_fft_size": (_i, [_vp]),
    "orc_sigdet_set_center_frequency": (None, [_vp, _d]),
    "orc_sigdet_process": (_i, [_vp, _i, _vp]),
    "orc_sigdet_result": (_i, [_vp, _vp, _i]),
    "orc_sigdet_history": (_i, [_vp, _i, _vp, _i]),
    "orc_sigdet_row": (None, [_vp, _i, _vp]),
    "orc_sigdet_candidates": (None, [_vp, _i, _vp]),
    "orc_sigdet_add_row": (None, [_vp, _vp, _i]),
    "orc_logmmse_create": (_vp, [_i]),
    "orc_logmmse_destroy": (None, [_vp]),
    "orc_logmmse_reset": (None, [_vp]),
    "orc_logmmse_set_hold": (None, [_vp, _i]),
    "orc_logmmse_process": (_i, [_vp, _i, _vp, _vp]),
    "orc_logmmse_noise": (_i, [_vp, _vp, _i]),
    "orc_expn": (_f, [_f]),
    "orc_fftcomp_create": (_vp, [_i, _d]),
    "orc_fftcomp_destroy": (None, [_vp]),
    "orc_fftcomp_fft_size": (_i, [_vp]),
    "orc_fftcomp_set_masked": (None, [_vp, _vp, _i]),
    "orc_fftcomp_process": (_i, [_vp, _i, _vp, _vp]),
    "orc_fftcomp_noise_figure": (_i, [_vp, _vp, _i]),
    "orc_graph_run": (_d, [_d, _i, _vp, _vp, _vp, _i, _i, _vp]),
}


def build(which="port"):
    """Compile the oracle library if its .so is missing (or stale for the port).
    "port_fast" / "ref_fast" are the -O3 -march=native timing builds (bench.py only, never parity)."""
    if which == "port_fast":
        subprocess.check_call(["make", "-s", "-C", ORACLE_DIR, "fast"])
        return os.path.join(ORACLE_DIR, "libport_oracle_fast.so")
    if which == "ref_fast":
        p = os.path.join(ORACLE_DIR, "_ref", "libref_oracle_fast.so")
        if os.path.isdir(os.path.join(REF_ROOT, "core", "src", "dsp")):
            subprocess.check_call(["make", "-s", "-C", ORACLE_DIR, "ref_fast", "REF_ROOT=" + REF_ROOT])
        return p if os.path.exists(p) else None
    if which == "port":
        subprocess.check_call(["make", "-s", "-C", ORACLE_DIR, "port"])
        return PORT_SO
    if os.path.isdir(os.path.join(REF_ROOT, "core", "src", "dsp")):
        subprocess.check_call(["make", "-s", "-C", ORACLE_DIR, "ref", "REF_ROOT=" + REF_ROOT])
    return REF_SO if os.path.exists(REF_SO) else None


class Oracle:
    def __init__(self, which="port"):
        path = build(which)
        if path is None or not os.path.exists(path):
            raise FileNotFoundError("oracle library '%s' is not built and cannot be built here" % which)
        self.lib = C.CDLL(path)
        for name, (res, args) in _SIGS.items():
            fn = getattr(self.lib, name)
            fn.restype = res
            fn.argtypes = args
        self.kind = self.lib.orc_kind().decode()

    # ------------------------------------------------------------ helpers
    @staticmethod
    def _c(a):
        a = np.ascontiguousarray(a)
        return a, a.ctypes.data

    def lowpass(self, cutoff, trans, sr):
        n = self.lib.orc_lowpass_taps(cutoff, trans, sr, None, 0)
        out = np.zeros(n, np.float32)
        self.lib.orc_lowpass_taps(cutoff, trans, sr, out.ctypes.data, n)
        return out

    def fft_window(self, wtype, nz):
        out = np.zeros(nz, np.float32)
        self.lib.orc_fft_window(wtype, nz, out.ctypes.data)
        return out

    def reshape_params(self, sr, size, rate):
        skip, nz = C.c_int(), C.c_int()
        self.lib.orc_reshape_params(sr, size, rate, C.addressof(skip), C.addressof(nz))
        return skip.value, nz.value

    def fft_frame(self, iq, window, N, want_spectrum=False):
        iq = np.ascontiguousarray(iq, np.complex64)
        window = np.ascontiguousarray(window, np.float32)
        nz = len(window)
        db = np.zeros(N, np.float32)
        spec = np.zeros(N, np.complex64) if want_spectrum else None
        self.lib.orc_fft_frame(iq.ctypes.data, window.ctypes.data, nz, N, db.ctypes.data, spec.ctypes.data if want_spectrum else None)
        return (db, spec) if want_spectrum else db

    def do_zoom(self, offset, width, row, out_size):
        row = np.ascontiguousarray(row, np.float32)
        out = np.zeros(out_size, np.float32)
        self.lib.orc_do_zoom(offset, width, len(row), out_size, row.ctypes.data, out.ctypes.data)
        return out

    def zoom_indices(self, offset, width, in_size, out_size):
        s = np.zeros(out_size, np.int32)
        u = np.zeros(out_size, np.int32)
        self.lib.orc_zoom_indices(offset, width, in_size, out_size, s.ctypes.data, u.ctypes.data)
        return s, u

    def palette_indices(self, row, wf_min, wf_max):
        row = np.ascontiguousarray(row, np.float32)
        idx = np.zeros(len(row), np.int32)
        self.lib.orc_palette_indices(row.ctypes.data, len(row), wf_min, wf_max, idx.ctypes.data)
        return idx

    def make_palette(self, colors):
        colors = np.ascontiguousarray(colors, np.float32)
        lut = np.zeros(1000, np.uint32)
        self.lib.orc_make_palette(colors.ctypes.data, colors.shape[0], lut.ctypes.data)
        return lut

    def vfo_signal_info(self, row, whole_bw, center, bw):
        row = np.ascontiguousarray(row, np.float32)
        st, sn = C.c_float(), C.c_float()
        ok = self.lib.orc_vfo_signal_info(row.ctypes.data, len(row), whole_bw, center, bw, C.addressof(st), C.addressof(sn))
        return (st.value, sn.value) if ok else None

    def smooth_hold(self, latest, smoothing_buf, hold, smoothing, alpha, do_hold, hold_speed):
        """in place on the three float32 arrays"""
        self.lib.orc_smooth_hold(latest.ctypes.data, smoothing_buf.ctypes.data, hold.ctypes.data, len(latest), int(smoothing), alpha, int(do_hold), hold_speed)

    def decim_plan(self, ratio):
        d = np.zeros(8, np.int32)
        t = np.zeros(8, np.int32)
        n = self.lib.orc_decim_plan(ratio, d.ctypes.data, t.ctypes.data, 8)
        taps = []
        for s in range(n):
            buf = np.zeros(t[s], np.float32)
            self.lib.orc_decim_stage_taps(ratio, s, buf.ctypes.data, int(t[s]))
            taps.append(buf)
        return [(int(d[s]), taps[s]) for s in range(n)]


class _Block:
    """Streaming block wrapper: process(block) -> output array, state kept in the oracle."""
    out_dtype = np.complex64
    out_per_in = 1.0

    def __init__(self, orc, handle, prefix):
        self.orc, self.h, self.prefix = orc, handle, prefix

    def process(self, x):
        x = np.ascontiguousarray(x, np.complex64)
        out = np.zeros(int(len(x) * self.out_per_in) + 1024, self.out_dtype)
        n = getattr(self.orc.lib, self.prefix + "_process")(self.h, len(x), x.ctypes.data, out.ctypes.data)
        return out[:n].copy()

    def close(self):
        if self.h:
            getattr(self.orc.lib, self.prefix + "_destroy")(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Vfo(_Block):
    def __init__(self, orc, in_sr, out_sr, bw, offset):
        self.out_per_in = max(1.0, 1.05 * out_sr / in_sr)
        super().__init__(orc, orc.lib.orc_vfo_create(in_sr, out_sr, bw, offset), "orc_vfo")

    def describe(self):
        info = np.zeros(7, np.int32)
        self.orc.lib.orc_vfo_describe(self.h, info.ctypes.data)
        return dict(mode=int(info[0]), ratio=int(info[1]), interp=int(info[2]), decim=int(info[3]),
                    resamp_taps=int(info[4]), taps_per_phase=int(info[5]), chan_taps=int(info[6]))

    def phase(self):
        p = np.zeros(2, np.float32)
        self.orc.lib.orc_vfo_get_phase(self.h, p.ctypes.data)
        return complex(p[0], p[1])

    def set_phase(self, z):
        self.orc.lib.orc_vfo_set_phase(self.h, float(np.float32(z.real)), float(np.float32(z.imag)))

    def phase_delta(self):
        p = np.zeros(2, np.float32)
        self.orc.lib.orc_vfo_get_phase_delta(self.h, p.ctypes.data)
        return complex(p[0], p[1])

    def resamp_taps(self):
        n = self.orc.lib.orc_vfo_get_resamp_taps(self.h, None, 0)
        out = np.zeros(n, np.float32)
        self.orc.lib.orc_vfo_get_resamp_taps(self.h, out.ctypes.data, n)
        return out

    def chan_taps(self):
        n = self.orc.lib.orc_vfo_get_chan_taps(self.h, None, 0)
        out = np.zeros(n, np.float32)
        self.orc.lib.orc_vfo_get_chan_taps(self.h, out.ctypes.data, n)
        return out

    def set_offset(self, v): self.orc.lib.orc_vfo_set_offset(self.h, v)
    def set_bandwidth(self, v): self.orc.lib.orc_vfo_set_bandwidth(self.h, v)
    def set_in_samplerate(self, v): self.orc.lib.orc_vfo_set_in_samplerate(self.h, v)
    def set_out_samplerate(self, sr, bw): self.orc.lib.orc_vfo_set_out_samplerate(self.h, sr, bw)
    def reset(self): self.orc.lib.orc_vfo_reset(self.h)


class Decim(_Block):
    def __init__(self, orc, ratio):
        super().__init__(orc, orc.lib.orc_decim_create(ratio), "orc_decim")


class DcBlock(_Block):
    def __init__(self, orc, rate):
        super().__init__(orc, orc.lib.orc_dcblock_create(rate), "orc_dcblock")


class Preproc(_Block):
    def __init__(self, orc, sr, decim, dc, invert):
        super().__init__(orc, orc.lib.orc_preproc_create(sr, decim, int(dc), int(invert)), "orc_preproc")


class _Demod(_Block):
    def process(self, x):
        x = np.ascontiguousarray(x, np.complex64)
        out = np.zeros((len(x) + 16, 2), np.float32)
        n = getattr(self.orc.lib, self.prefix + "_process")(self.h, len(x), x.ctypes.data, out.ctypes.data)
        return out[:n].copy()


class Fm(_Demod):
    def __init__(self, orc, sr, bw, low_pass=True, high_pass=False):
        super().__init__(orc, orc.lib.orc_fm_create(sr, bw, int(low_pass), int(high_pass)), "orc_fm")


class Ssb(_Demod):
    def __init__(self, orc, mode, bw, sr, attack, decay):
        super().__init__(orc, orc.lib.orc_ssb_create(mode, bw, sr, attack, decay), "orc_ssb")


class Am(_Demod):
    def __init__(self, orc, agc_mode, bw, attack, decay, dc_rate, sr):
        super().__init__(orc, orc.lib.orc_am_create(agc_mode, bw, attack, decay, dc_rate, sr), "orc_am")


class Wfm(_Block):
    """demod::BroadcastFM (mono or stereo, no RDS): complex IF in, stereo audio out at the same rate"""
    def __init__(self, orc, deviation, samplerate, stereo, low_pass=True):
        super().__init__(orc, orc.lib.orc_wfm_create(deviation, samplerate, int(stereo), int(low_pass)), "orc_wfm")

    def process(self, x):
        x = np.ascontiguousarray(x, np.complex64)
        out = np.zeros((len(x), 2), np.float32)
        n = self.orc.lib.orc_wfm_process(self.h, len(x), x.ctypes.data, out.ctypes.data)
        return out[:n]


class WfmRds(_Block):
    """demod::BroadcastFM with the RDS branch on: process() returns (stereo audio, complex RDS baseband at 5 kS/s)"""
    def __init__(self, orc, deviation, samplerate, stereo, low_pass=True):
        super().__init__(orc, orc.lib.orc_wfm_create_rds(deviation, samplerate, int(stereo), int(low_pass)), "orc_wfm")

    def process(self, x):
        x = np.ascontiguousarray(x, np.complex64)
        out = np.zeros((len(x), 2), np.float32)
        rds = np.zeros(len(x) + 1024, np.complex64)       # the resampler's stages also use it as scratch
        n = self.orc.lib.orc_wfm_process_rds(self.h, len(x), x.ctypes.data, out.ctypes.data, rds.ctypes.data)
        return out, rds[:n].copy()


class FmIf(_Block):
    """noise_reduction::FMIF: complex in, complex out, same count"""
    def __init__(self, orc, bins):
        super().__init__(orc, orc.lib.orc_fmif_create(int(bins)), "orc_fmif")

    def process(self, x):
        x = np.ascontiguousarray(x, np.complex64)
        out = np.zeros(len(x), np.complex64)
        self.orc.lib.orc_fmif_process(self.h, len(x), x.ctypes.data, out.ctypes.data)
        return out


def stream_compress(orc, x, pcm_type):
    """SampleStreamCompressor::process: complex64 block -> the server's sample packet bytes"""
    x = np.ascontiguousarray(x, np.complex64)
    out = np.zeros(8 + 8 * len(x), np.uint8)
    n = orc.lib.orc_stream_compress(len(x), int(pcm_type), x.ctypes.data, out.ctypes.data)
    return out[:n].copy()


def stream_decompress(orc, pkt):
    pkt = np.ascontiguousarray(pkt, np.uint8)
    out = np.zeros(len(pkt), np.complex64)
    n = orc.lib.orc_stream_decompress(len(pkt), pkt.ctypes.data, out.ctypes.data)
    return out[:n].copy()


def wav_write(orc, path, channels, samplerate, sample_type, samples, block_len=4096):
    """wav::Writer: `samples` float32 [count, channels] written as a WAV file of the given sample type (0 u8, 1 i16, 2 i32, 3 f32)"""
    x = np.ascontiguousarray(samples, np.float32).reshape(-1, channels)
    rc = orc.lib.orc_wav_write(str(path).encode(), int(channels), float(samplerate), int(sample_type), x.ctypes.data, len(x), int(block_len))
    if rc != 0:
        raise RuntimeError("orc_wav_write failed")


def volume(orc, vol, muted, x):
    """audio::Volume on a stereo block"""
    x = np.ascontiguousarray(x, np.float32).reshape(-1, 2)
    out = np.zeros_like(x)
    orc.lib.orc_volume_process(float(vol), int(muted), len(x), x.ctypes.data, out.ctypes.data)
    return out


class IfChain(_Block):
    """radio IF chain: NoiseBlanker -> Squelch on the VFO output (complex in, complex out, same count)"""
    def __init__(self, orc, nb_on, nb_rate, nb_level, sq_on, sq_level):
        super().__init__(orc, orc.lib.orc_ifchain_create(int(nb_on), nb_rate, nb_level, int(sq_on), sq_level), "orc_ifchain")

    def process(self, x):
        x = np.ascontiguousarray(x, np.complex64)
        out = np.zeros(len(x), np.complex64)
        self.orc.lib.orc_ifchain_process(self.h, len(x), x.ctypes.data, out.ctypes.data)
        return out


class AfPost(_Block):
    """radio AF chain: stereo in at the demodulator's rate -> stereo out at the audio rate (+ de-emphasis)"""
    def __init__(self, orc, in_sr, audio_sr, tau=0.0):
        self.ratio = audio_sr / in_sr
        super().__init__(orc, orc.lib.orc_afpost_create(in_sr, audio_sr, tau), "orc_afpost")

    def process(self, x):
        x = np.ascontiguousarray(x, np.float32).reshape(-1, 2)
        out = np.zeros((int(len(x) * max(self.ratio, 1.0) * 1.05) + 1024, 2), np.float32)   # stages also use `out` as scratch
        n = self.orc.lib.orc_afpost_process(self.h, len(x), x.ctypes.data, out.ctypes.data)
        return out[:n].copy()


class Agc(_Block):
    def __init__(self, orc, set_point, attack, decay, max_gain, max_out, init_gain):
        super().__init__(orc, orc.lib.orc_agc_create(set_point, attack, decay, max_gain, max_out, init_gain), "orc_agc")

    def process(self, x):
        x = np.ascontiguousarray(x, np.float32)
        out = np.zeros(len(x), np.float32)
        self.orc.lib.orc_agc_process(self.h, len(x), x.ctypes.data, out.ctypes.data)
        return out


class SigDet:
    """dsp::detector::SignalDetector (dsp/detector/signal_detector.cpp): feed blocks, read sigs_smoothed."""

    def __init__(self, orc, sample_rate):
        self.orc = orc
        self.h = orc.lib.orc_sigdet_create(float(sample_rate))
        self.n = orc.lib.orc_sigdet_fft_size(self.h)

    def process(self, x):
        x = np.ascontiguousarray(x, np.complex64)
        return self.orc.lib.orc_sigdet_process(self.h, len(x), x.ctypes.data)

    def set_center_frequency(self, f):
        self.orc.lib.orc_sigdet_set_center_frequency(self.h, float(f))

    def add_row(self, row):
        row = np.ascontiguousarray(row, np.float32)
        self.orc.lib.orc_sigdet_add_row(self.h, row.ctypes.data, len(row))

    def result(self):
        n = self.orc.lib.orc_sigdet_result(self.h, None, 0)
        out = np.zeros(n, np.float32)
        if n:
            self.orc.lib.orc_sigdet_result(self.h, out.ctypes.data, n)
        return out

    def history(self):
        k = self.orc.lib.orc_sigdet_history(self.h, -1, None, 0)
        rows = np.zeros((k, self.n), np.float32)
        for r in range(k):
            self.orc.lib.orc_sigdet_history(self.h, r, rows[r].ctypes.data, self.n)
        return rows

    def close(self):
        if self.h:
            self.orc.lib.orc_sigdet_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class LogMmse(_Block):
    """dsp::IFNRLogMMSE (noise_reduction_logmmse/src/if_nr.h): blocks in, de-noised blocks out (fewer or more samples per call)."""

    def __init__(self, orc, sample_rate):
        super().__init__(orc, orc.lib.orc_logmmse_create(int(sample_rate)), "orc_logmmse")
        self.rate = int(sample_rate)

    def process(self, x):
        x = np.ascontiguousarray(x, np.complex64)
        out = np.zeros(len(x) + 40 * (self.rate // 100) + 1024, np.complex64)
        n = self.orc.lib.orc_logmmse_process(self.h, len(x), x.ctypes.data, out.ctypes.data)
        return out[:n].copy()

    def reset(self):
        self.orc.lib.orc_logmmse_reset(self.h)

    def set_hold(self, hold):
        self.orc.lib.orc_logmmse_set_hold(self.h, int(hold))

    def noise(self):
        n = self.orc.lib.orc_logmmse_noise(self.h, None, 0)
        out = np.zeros(n, np.float32)
        if n:
            self.orc.lib.orc_logmmse_noise(self.h, out.ctypes.data, n)
        return out


class FftCompressor(_Block):
    """dsp::compression::ExperimentalFFTCompressor: blocks in, one frequency-domain frame (or nothing) out per call"""

    def __init__(self, orc, sample_rate, loss_rate=1.0):
        super().__init__(orc, orc.lib.orc_fftcomp_create(int(sample_rate), float(loss_rate)), "orc_fftcomp")
        self.n = orc.lib.orc_fftcomp_fft_size(self.h)

    def process(self, x):
        x = np.ascontiguousarray(x, np.complex64)
        out = np.zeros(self.n, np.complex64)
        k = self.orc.lib.orc_fftcomp_process(self.h, len(x), x.ctypes.data, out.ctypes.data)
        return out[:k].copy()

    def set_masked(self, from_to_hz):
        a = np.ascontiguousarray(from_to_hz, np.int32)
        self.orc.lib.orc_fftcomp_set_masked(self.h, a.ctypes.data, len(a))

    def noise_figure(self):
        out = np.zeros(64, np.float32)
        k = self.orc.lib.orc_fftcomp_noise_figure(self.h, out.ctypes.data, 64)
        return out[:k].copy()


def graph_run(orc, in_sr, vfos, block, n_blocks):
    """the reference's Splitter + one-thread-per-RxVFO graph on `block` fed n_blocks times: seconds, or -1 (port)"""
    block = np.ascontiguousarray(block, np.complex64)
    o = np.array([v[0] for v in vfos], np.float64)
    b = np.array([v[1] for v in vfos], np.float64)
    f = np.array([v[2] for v in vfos], np.float64)
    return orc.lib.orc_graph_run(float(in_sr), len(vfos), o.ctypes.data, b.ctypes.data, f.ctypes.data, len(block), int(n_blocks), block.ctypes.data)


def sigdet_row(orc, x):
    x = np.ascontiguousarray(x, np.complex64)
    out = np.zeros(len(x), np.float32)
    orc.lib.orc_sigdet_row(x.ctypes.data, len(x), out.ctypes.data)
    return out


def sigdet_candidates(orc, row):
    row = np.ascontiguousarray(row, np.float32)
    out = np.zeros(len(row), np.float32)
    orc.lib.orc_sigdet_candidates(row.ctypes.data, len(row), out.ctypes.data)
    return out


def fmif_model(x, bins, hist=None):
    """float64 numpy model of noise_reduction::FMIF (fm_if.h:45-80): returns (out, relative margin between the two
    strongest bins of every window).  The window is Nuttall over bins-1 (fm_if.h:118); the backward DFT of a single
    bin evaluated at element bins/2 is X[idx] * exp(+2 pi i idx (bins//2) / bins)."""
    n = np.arange(bins)
    c = [0.355768, 0.487396, 0.144232, 0.012604]
    win = sum(((-1) ** k) * c[k] * np.cos(2 * np.pi * k * n / (bins - 1)) for k in range(4))
    hist = np.zeros(bins - 1, np.complex128) if hist is None else hist
    buf = np.concatenate([hist, x.astype(np.complex128)])
    seg = np.lib.stride_tricks.sliding_window_view(buf, bins)[:len(x)]
    # fp32 window product as the reference forms it, then the transform in float64
    segw = (seg.real.astype(np.float32) * win.astype(np.float32)).astype(np.float64) + \
        1j * (seg.imag.astype(np.float32) * win.astype(np.float32)).astype(np.float64)
    F = np.exp(-2j * np.pi * np.outer(n, n) / bins)
    X = segw @ F
    mag = np.abs(X)
    idx = np.argmax(mag, axis=1)
    top2 = np.sort(mag, axis=1)[:, -2:]
    margin = (top2[:, 1] - top2[:, 0]) / np.maximum(top2[:, 1], 1e-30)
    out = X[np.arange(len(x)), idx] * np.exp(2j * np.pi * idx * (bins // 2) / bins)
    return out, margin, buf[len(x):]


def have_ref():
    return os.path.exists(REF_SO) or os.path.isdir(os.path.join(REF_ROOT, "core", "src", "dsp"))
