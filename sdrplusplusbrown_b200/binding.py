"""ctypes binding of include/b200dsp.h (host side used by tests and bench.py).

Importing this module loads ``libb200dsp.so``; if the library is missing it is built with nvcc,
and if that fails the import fails — there is no CPU fallback for any process call.
Device buffers are torch CUDA tensors (torch is the allocator / stream provider only).
"""
import ctypes as C
import os

import numpy as np

from . import build as _build

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb200dsp.so")

def _ensure_built():
    """Build the library when it is missing or older than its sources -- under a file lock, so that several ranks
    importing this module at once (torchrun) do not run nvcc into the same output at the same time."""
    if os.path.exists(LIB_PATH) and not _build.stale():
        return
    import fcntl
    with open(os.path.join(_HERE, ".build.lock"), "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            if not os.path.exists(LIB_PATH) or _build.stale():      # another rank may have built it while we waited
                _build.build()
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)


_ensure_built()
lib = C.CDLL(LIB_PATH)

_vp, _i, _d, _f, _pp = C.c_void_p, C.c_int, C.c_double, C.c_float, C.POINTER(C.c_void_p)

_SIGS = {
    "b200_last_error": (C.c_char_p, []),
    "b200_version": (C.c_char_p, []),
    "b200_launch_count": (C.c_uint64, []),
    "b200_host_alloc": (_vp, [C.c_size_t]),
    "b200_host_free": (None, [_vp]),
    "b200_device_to_host": (_i, [_vp, _vp, C.c_size_t, _vp]),
    "b200_profile_enable": (_i, [_i]),
    "b200_profile_read": (_i, [C.POINTER(_d), C.POINTER(C.c_longlong), C.POINTER(_d)]),
    "b200_profile_read_kind": (_i, [_i, C.POINTER(_d), C.POINTER(C.c_longlong), C.POINTER(_d), C.POINTER(_d)]),
    "b200_measure_fp32_peak": (_i, [_i, C.POINTER(_d)]),
    "b200_vfo_bank_create": (_i, [_pp, _d, _i, _i]),
    "b200_vfo_bank_destroy": (_i, [_vp]),
    "b200_vfo_bank_add": (_i, [_vp, _d, _d, _d]),
    "b200_vfo_bank_remove": (_i, [_vp, _i]),
    "b200_vfo_bank_count": (_i, [_vp]),
    "b200_vfo_set_offset": (_i, [_vp, _i, _d]),
    "b200_vfo_set_bandwidth": (_i, [_vp, _i, _d]),
    "b200_vfo_set_out_samplerate": (_i, [_vp, _i, _d, _d]),
    "b200_vfo_bank_set_in_samplerate": (_i, [_vp, _d]),
    "b200_vfo_reset": (_i, [_vp, _i]),
    "b200_vfo_get_phase": (_i, [_vp, _i, C.POINTER(_d), C.POINTER(_d)]),
    "b200_vfo_set_phase": (_i, [_vp, _i, _d, _d]),
    "b200_vfo_describe": (_i, [_vp, _i, _vp]),
    "b200_vfo_bank_set_rotator_mode": (_i, [_vp, _i]),
    "b200_vfo_bank_set_option": (_i, [_vp, _i, _i]),
    "b200_set_global_option": (_i, [_i, _i]),
    "b200_vfo_get_taps": (_i, [_vp, _i, _i, _vp, _i]),
    "b200_vfo_bank_process": (_i, [_vp, _vp, _i, _vp]),
    "b200_vfo_out_count": (_i, [_vp, _i]),
    "b200_vfo_out_ptr": (_vp, [_vp, _i]),
    "b200_vfo_bank_process_host": (_i, [_vp, _vp, _i, _vp, _vp, _vp]),
    "b200_vfo_max_out": (_i, [_vp, _i, _i]),
    "b200_vfo_set_demod": (_i, [_vp, _i, _i, _d, _d, _d, _d]),
    "b200_vfo_set_audio_post": (_i, [_vp, _i, _d, _d]),
    "b200_vfo_set_if_chain": (_i, [_vp, _i, _i, _d, _d, _i, _d]),
    "b200_vfo_set_if_nr": (_i, [_vp, _i, _i, _i]),
    "b200_vfo_set_volume": (_i, [_vp, _i, _d, _i]),
    "b200_vfo_rds_ptr": (_vp, [_vp, _i]),
    "b200_vfo_if_ptr": (_vp, [_vp, _i]),
    "b200_vfo_rds_count": (_i, [_vp, _i]),
    "b200_vfo_audio_ptr": (_vp, [_vp, _i]),
    "b200_vfo_audio_count": (_i, [_vp, _i]),
    "b200_vfo_audio_max_out": (_i, [_vp, _i, _i]),
    "b200_stream_compress": (_i, [_vp, _i, _i, _vp, _vp]),
    "b200_convert_pcm": (_i, [_vp, _vp, _i, _i, _vp]),
    "b200_frontend_create": (_i, [_pp, _d, _i, _i, _i, _i, _d, _i, _i, _i]),
    "b200_frontend_destroy": (_i, [_vp]),
    "b200_frontend_set_sample_rate": (_i, [_vp, _d]),
    "b200_frontend_set_decimation": (_i, [_vp, _i]),
    "b200_frontend_set_dc_blocking": (_i, [_vp, _i]),
    "b200_frontend_set_invert_iq": (_i, [_vp, _i]),
    "b200_frontend_set_fft_size": (_i, [_vp, _i]),
    "b200_frontend_set_fft_rate": (_i, [_vp, _d]),
    "b200_frontend_set_fft_window": (_i, [_vp, _i]),
    "b200_frontend_effective_samplerate": (_d, [_vp]),
    "b200_frontend_set_frame_shard": (_i, [_vp, _i, _i]),
    "b200_frontend_vfo_bank": (_vp, [_vp]),
    "b200_frontend_process": (_i, [_vp, _vp, _i, _vp]),
    "b200_frontend_preproc_out": (_i, [_vp, _pp, C.POINTER(_i)]),
    "b200_frontend_frame_count": (_i, [_vp]),
    "b200_frontend_frame_ptr": (_vp, [_vp, _i]),
    "b200_frontend_frame_index": (C.c_int64, [_vp, _i]),
    "b200_frontend_attach_waterfall": (_i, [_vp, _vp, _i, _vp, _vp, _i]),
    "b200_frontend_process_host": (_i, [_vp, _vp, _i, _vp, _i, _vp, _vp, _vp, _vp]),
    "b200_frontend_process_host_i16": (_i, [_vp, _vp, _i, _f, _vp, _i, _vp, _vp, _vp, _vp]),
    "b200_frontend_process_host_i8": (_i, [_vp, _vp, _i, _f, _vp, _i, _vp, _vp, _vp, _vp]),
    "b200_convert_i16": (_i, [_vp, _i, _f, _vp, _vp]),
    "b200_frontend_group_create": (_i, [_pp, _vp, _i, _d, _i, _i, _i, _i, _d, _i, _i]),
    "b200_frontend_group_destroy": (_i, [_vp]),
    "b200_frontend_group_size": (_i, [_vp]),
    "b200_frontend_group_uses_nccl": (_i, [_vp]),
    "b200_frontend_group_member": (_vp, [_vp, _i]),
    "b200_frontend_group_add_vfo": (_i, [_vp, _d, _d, _d]),
    "b200_frontend_group_remove_vfo": (_i, [_vp, _i]),
    "b200_frontend_group_vfo_location": (_i, [_vp, _i, C.POINTER(_i), C.POINTER(_i)]),
    "b200_frontend_group_attach_waterfall": (_i, [_vp, _vp, _i, _vp, _vp, _i]),
    "b200_frontend_group_process": (_i, [_vp, _vp, _i, _vp]),
    "b200_frontend_group_process_host": (_i, [_vp, _vp, _i, _vp]),
    "b200_frontend_group_process_host_i16": (_i, [_vp, _vp, _i, _f, _vp]),
    "b200_frontend_group_frame_count": (_i, [_vp]),
    "b200_frontend_group_frame_ptr": (_vp, [_vp, _i]),
    "b200_frontend_group_frame_index": (C.c_int64, [_vp, _i]),
    "b200_frontend_group_sync": (_i, [_vp]),
    "b200_frontend_group_broadcast_ms": (_i, [_vp, C.POINTER(_d)]),
    "b200_sigdet_create": (_i, [_pp, _d, _i]),
    "b200_sigdet_destroy": (None, [_vp]),
    "b200_sigdet_set_sample_rate": (_i, [_vp, _d]),
    "b200_sigdet_set_center_frequency": (_i, [_vp, _d]),
    "b200_sigdet_set_enabled": (_i, [_vp, _i]),
    "b200_sigdet_is_enabled": (_i, [_vp]),
    "b200_sigdet_fft_size": (_i, [_vp]),
    "b200_sigdet_process": (_i, [_vp, _vp, _i, _vp]),
    "b200_sigdet_process_host": (_i, [_vp, _vp, _i, _vp]),
    "b200_sigdet_add_row": (_i, [_vp, _vp, _vp]),
    "b200_sigdet_sync": (_i, [_vp]),
    "b200_sigdet_rows": (C.c_int64, [_vp]),
    "b200_sigdet_result": (_i, [_vp, _vp, _i]),
    "b200_sigdet_result_ptr": (_vp, [_vp]),
    "b200_sigdet_last_row_ptr": (_vp, [_vp]),
    "b200_sigdet_history": (_i, [_vp, _i, _vp, _i]),
    "b200_frontend_detector": (_vp, [_vp]),
    "b200_frontend_set_detector": (_i, [_vp, _i]),
    "b200_logmmse_create": (_i, [_pp, _i, _i, _i]),
    "b200_logmmse_destroy": (None, [_vp]),
    "b200_logmmse_reset": (_i, [_vp]),
    "b200_logmmse_set_hold": (_i, [_vp, _i]),
    "b200_logmmse_fft_size": (_i, [_vp]),
    "b200_logmmse_max_out": (_i, [_vp]),
    "b200_logmmse_process": (_i, [_vp, _vp, _i, _pp, C.POINTER(_i), _vp]),
    "b200_logmmse_process_host": (_i, [_vp, _vp, _i, _vp, _i, C.POINTER(_i), _vp]),
    "b200_logmmse_sync": (_i, [_vp]),
    "b200_logmmse_noise": (_i, [_vp, _vp, _i]),
    "b200_fftcomp_create": (_i, [_pp, _i, _d, _i, _i]),
    "b200_fftcomp_destroy": (None, [_vp]),
    "b200_fftcomp_fft_size": (_i, [_vp]),
    "b200_fftcomp_set_loss_rate": (_i, [_vp, _d]),
    "b200_fftcomp_set_masked": (_i, [_vp, _vp, _i]),
    "b200_fftcomp_process": (_i, [_vp, _vp, _i, _pp, C.POINTER(_i), _vp]),
    "b200_fftcomp_process_host": (_i, [_vp, _vp, _i, _vp, C.POINTER(_i), _vp]),
    "b200_fftcomp_sync": (_i, [_vp]),
    "b200_fftcomp_noise_figure": (_i, [_vp, _vp, _i]),
    "b200_fft_forward": (_i, [_vp, _i, _i, _vp, _vp]),
    "b200_dc_block_noalloc": (_i, [_vp, _vp, _i, _d, _vp, _vp]),
    "b200_waterfall_create": (_i, [_pp, _i, _i, _i]),
    "b200_waterfall_destroy": (_i, [_vp]),
    "b200_waterfall_set_view": (_i, [_vp, _i, _i]),
    "b200_waterfall_set_range": (_i, [_vp, _f, _f]),
    "b200_waterfall_set_palette": (_i, [_vp, _vp]),
    "b200_waterfall_make_palette": (_i, [_vp, _i, _vp]),
    "b200_waterfall_push": (_i, [_vp, _vp, _i, _vp, _vp, _vp]),
    "b200_waterfall_set_history": (_i, [_vp, _i]),
    "b200_waterfall_regenerate": (_i, [_vp, _vp, _vp, _vp]),
    "b200_waterfall_zoom_table": (_i, [_vp, _vp, _vp]),
    "b200_waterfall_vfo_signal": (_i, [_vp, _vp, _i, _d, _d, _d, _vp, _vp]),
    "b200_waterfall_smooth_hold": (_i, [_vp, _vp, _i, _i, _f, _i, _f, _vp, _vp]),
}

EXPORTS = sorted(_SIGS.keys())
for _name, (_res, _args) in _SIGS.items():
    _fn = getattr(lib, _name)
    _fn.restype = _res
    _fn.argtypes = _args


ROTATOR_FAST, ROTATOR_EXACT = 0, 1
OPT_FUSED_STAGE1, OPT_CALIBRATION_STEPS, OPT_TILE_SAMPLES, OPT_FFT_CTAS_PER_SM, OPT_BRANCH_ORDER, OPT_DEFER_TAIL, OPT_VFOS_PER_PASS, OPT_VFOS_PER_CTA, OPT_FFT_DIRECT_MAX_LOG, OPT_SPLIT_RUNS, OPT_LEAVE_ROOM, OPT_PARAM_TABLES, OPT_HOST_CHUNK_MIB, OPT_FFT_DIRECT_PREFETCH = 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14
DEMOD_NONE, DEMOD_NFM, DEMOD_AM, DEMOD_USB, DEMOD_LSB, DEMOD_DSB, DEMOD_WFM = 0, 1, 2, 3, 4, 5, 6


class B200Error(RuntimeError):
    pass


def _chk(rc):
    if rc < 0:
        raise B200Error("libb200dsp error %d: %s" % (rc, lib.b200_last_error().decode()))
    return rc


def _stream_ptr(stream):
    if stream is None:
        import torch
        return torch.cuda.current_stream().cuda_stream
    return getattr(stream, "cuda_stream", stream)


def _dev_to_numpy(ptr, count, dtype, device):
    """Copy `count` elements of `dtype` from a raw device pointer to a numpy array."""
    import torch
    out = np.empty(count, dtype)
    if count == 0:
        return out
    torch.cuda.synchronize(device)
    rt = C.CDLL("libcudart.so.12") if not hasattr(_dev_to_numpy, "_rt") else _dev_to_numpy._rt
    _dev_to_numpy._rt = rt
    rt.cudaMemcpy.argtypes = [_vp, _vp, C.c_size_t, _i]
    rc = rt.cudaMemcpy(out.ctypes.data, ptr, out.nbytes, 2)
    if rc != 0:
        raise B200Error("cudaMemcpy D2H failed: %d" % rc)
    return out


class VfoBank:
    """N RxVFOs fed by one input stream (dsp::channel::RxVFO, channel/rx_vfo.h)."""

    def __init__(self, in_samplerate, max_block=1000000, device=0, _handle=None):
        self.device = device
        self.owned = _handle is None
        if _handle is None:
            h = _vp()
            _chk(lib.b200_vfo_bank_create(C.byref(h), float(in_samplerate), int(max_block), int(device)))
            self.h = h
        else:
            self.h = _vp(_handle)

    def close(self):
        if self.h and self.owned:
            lib.b200_vfo_bank_destroy(self.h)
        self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def add(self, out_sr, bw, offset):
        return _chk(lib.b200_vfo_bank_add(self.h, float(out_sr), float(bw), float(offset)))

    def remove(self, v): _chk(lib.b200_vfo_bank_remove(self.h, v))
    def count(self): return _chk(lib.b200_vfo_bank_count(self.h))
    def set_offset(self, v, off): _chk(lib.b200_vfo_set_offset(self.h, v, float(off)))
    def set_bandwidth(self, v, bw): _chk(lib.b200_vfo_set_bandwidth(self.h, v, float(bw)))
    def set_out_samplerate(self, v, sr, bw): _chk(lib.b200_vfo_set_out_samplerate(self.h, v, float(sr), float(bw)))
    def set_in_samplerate(self, sr): _chk(lib.b200_vfo_bank_set_in_samplerate(self.h, float(sr)))
    def reset(self, v): _chk(lib.b200_vfo_reset(self.h, v))

    def get_phase(self, v):
        re, im = _d(), _d()
        _chk(lib.b200_vfo_get_phase(self.h, v, C.byref(re), C.byref(im)))
        return complex(re.value, im.value)

    def set_phase(self, v, z): _chk(lib.b200_vfo_set_phase(self.h, v, float(z.real), float(z.imag)))
    def set_rotator_mode(self, mode): _chk(lib.b200_vfo_bank_set_rotator_mode(self.h, int(mode)))
    def set_option(self, opt, value): _chk(lib.b200_vfo_bank_set_option(self.h, int(opt), int(value)))

    def describe(self, v):
        info = np.zeros(8, np.int32)
        _chk(lib.b200_vfo_describe(self.h, v, info.ctypes.data))
        return dict(mode=int(info[0]), ratio=int(info[1]), interp=int(info[2]), decim=int(info[3]),
                    resamp_taps=int(info[4]), taps_per_phase=int(info[5]), chan_taps=int(info[6]), fused=int(info[7]))

    def taps(self, v, which):
        n = _chk(lib.b200_vfo_get_taps(self.h, v, which, None, 0))
        out = np.zeros(n, np.float32)
        _chk(lib.b200_vfo_get_taps(self.h, v, which, out.ctypes.data, n))
        return out

    def set_demod(self, v, kind, p0=0.0, p1=0.0, p2=0.0, p3=0.0):
        _chk(lib.b200_vfo_set_demod(self.h, v, int(kind), float(p0), float(p1), float(p2), float(p3)))

    def set_if_chain(self, v, nb=False, nb_rate=0.0, nb_level=0.0, squelch=False, squelch_level=0.0):
        """radio IF chain in front of the demodulator: noise blanker -> squelch"""
        _chk(lib.b200_vfo_set_if_chain(self.h, v, int(nb), float(nb_rate), float(nb_level), int(squelch), float(squelch_level)))

    def set_if_nr(self, v, enabled=True, bins=32):
        """FM IF noise reduction (noise_reduction::FMIF), the last block of the IF chain"""
        _chk(lib.b200_vfo_set_if_nr(self.h, v, int(enabled), int(bins)))

    def set_volume(self, v, volume=1.0, muted=False):
        """sink-side audio::Volume behind the AF chain"""
        _chk(lib.b200_vfo_set_volume(self.h, v, float(volume), int(muted)))

    def if_output(self, v):
        """the demodulator's input of the last call: the IF chain's output"""
        n = self.out_count(v)
        p = lib.b200_vfo_if_ptr(self.h, v)
        if n == 0 or not p:
            return np.zeros(0, np.complex64)
        return _dev_to_numpy(p, n, np.complex64, self.device)

    def rds_count(self, v): return _chk(lib.b200_vfo_rds_count(self.h, v))

    def rds(self, v):
        """5 kS/s complex RDS baseband of a WFM demodulator created with p2 != 0"""
        n = self.rds_count(v)
        p = lib.b200_vfo_rds_ptr(self.h, v)
        if n == 0 or not p:
            return np.zeros(0, np.complex64)
        return _dev_to_numpy(p, n, np.complex64, self.device)

    def set_audio_post(self, v, audio_samplerate=0.0, deemph_tau=0.0):
        """radio AF chain behind the demodulator: resampler to the audio rate + de-emphasis (0 = leave out)"""
        _chk(lib.b200_vfo_set_audio_post(self.h, v, float(audio_samplerate), float(deemph_tau)))

    def process(self, d_in, count=None, stream=None):
        """d_in: torch CUDA float32 tensor [count, 2] (or any tensor whose data_ptr holds cf32)."""
        if count is None:
            count = d_in.shape[0]
        _chk(lib.b200_vfo_bank_process(self.h, d_in.data_ptr() if count else None, int(count), _stream_ptr(stream)))

    def out_count(self, v): return _chk(lib.b200_vfo_out_count(self.h, v))
    def out_ptr(self, v): return lib.b200_vfo_out_ptr(self.h, v)
    def audio_ptr(self, v): return lib.b200_vfo_audio_ptr(self.h, v)

    def output(self, v):
        n = self.out_count(v)
        return _dev_to_numpy(self.out_ptr(v), n, np.complex64, self.device)

    def audio_count(self, v): return _chk(lib.b200_vfo_audio_count(self.h, v))
    def audio_max_out(self, v, count): return _chk(lib.b200_vfo_audio_max_out(self.h, v, int(count)))

    def audio(self, v):
        n = self.audio_count(v)
        p = self.audio_ptr(v)
        if not p:
            raise B200Error("VFO %d has no demodulator" % v)
        return _dev_to_numpy(p, 2 * n, np.float32, self.device).reshape(n, 2)

    def max_out(self, v, count): return _chk(lib.b200_vfo_max_out(self.h, v, int(count)))

    def process_host(self, x, vfos, stream=None):
        """Host-buffer variant: x complex64 numpy; returns {vfo: complex64 array}."""
        x = np.ascontiguousarray(x, np.complex64)
        nslots = max(vfos) + 1
        bufs = {v: np.empty(self.max_out(v, len(x)), np.complex64) for v in vfos}
        ptrs = (_vp * nslots)()
        for v in vfos:
            ptrs[v] = bufs[v].ctypes.data
        counts = np.zeros(nslots, np.int32)
        _chk(lib.b200_vfo_bank_process_host(self.h, x.ctypes.data, len(x), ptrs, counts.ctypes.data, _stream_ptr(stream)))
        return {v: bufs[v][:counts[v]].copy() for v in vfos}


class FftCompressor:
    """dsp::compression::ExperimentalFFTCompressor (dsp/compression/experimental_fft_compressor.h) on one GPU."""

    def __init__(self, sample_rate, loss_rate=1.0, max_block=1000000, device=0):
        h = _vp()
        _chk(lib.b200_fftcomp_create(C.byref(h), int(sample_rate), float(loss_rate), int(max_block), int(device)))
        self.h = h
        self.device = device
        self.n = lib.b200_fftcomp_fft_size(h)

    def close(self):
        if self.h:
            lib.b200_fftcomp_destroy(self.h)
        self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_loss_rate(self, r): _chk(lib.b200_fftcomp_set_loss_rate(self.h, float(r)))

    def set_masked(self, from_to_hz):
        a = np.ascontiguousarray(from_to_hz, np.int32)
        _chk(lib.b200_fftcomp_set_masked(self.h, a.ctypes.data, len(a)))

    def process(self, d_in, count=None, stream=None):
        if count is None:
            count = d_in.shape[0]
        p, n = _vp(), _i()
        _chk(lib.b200_fftcomp_process(self.h, d_in.data_ptr() if count else None, int(count), C.byref(p), C.byref(n), _stream_ptr(stream)))
        _chk(lib.b200_fftcomp_sync(self.h))
        return _dev_to_numpy(p.value, n.value, np.complex64, self.device) if n.value else np.zeros(0, np.complex64)

    def process_host(self, x, stream=None):
        x = np.ascontiguousarray(x, np.complex64)
        out = np.zeros(self.n, np.complex64)
        n = _i()
        _chk(lib.b200_fftcomp_process_host(self.h, x.ctypes.data, len(x), out.ctypes.data, C.byref(n), _stream_ptr(stream)))
        return out[:n.value].copy()

    def noise_figure(self):
        out = np.zeros(30, np.float32)
        k = _chk(lib.b200_fftcomp_noise_figure(self.h, out.ctypes.data, 30))
        return out[:k].copy()


class LogMmse:
    """dsp::IFNRLogMMSE (misc_modules/noise_reduction_logmmse/src/if_nr.h) on one GPU."""

    def __init__(self, sample_rate, max_block=1000000, device=0):
        h = _vp()
        _chk(lib.b200_logmmse_create(C.byref(h), int(sample_rate), int(max_block), int(device)))
        self.h = h
        self.device = device

    def close(self):
        if self.h:
            lib.b200_logmmse_destroy(self.h)
        self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def reset(self): _chk(lib.b200_logmmse_reset(self.h))
    def set_hold(self, hold): _chk(lib.b200_logmmse_set_hold(self.h, int(hold)))
    def fft_size(self): return lib.b200_logmmse_fft_size(self.h)

    def process(self, d_in, count=None, stream=None):
        """device block in; the de-noised samples of this call as a numpy array"""
        if count is None:
            count = d_in.shape[0]
        p, n = _vp(), _i()
        _chk(lib.b200_logmmse_process(self.h, d_in.data_ptr() if count else None, int(count), C.byref(p), C.byref(n), _stream_ptr(stream)))
        _chk(lib.b200_logmmse_sync(self.h))
        return _dev_to_numpy(p.value, n.value, np.complex64, self.device) if n.value else np.zeros(0, np.complex64)

    def process_host(self, x, stream=None):
        x = np.ascontiguousarray(x, np.complex64)
        cap = lib.b200_logmmse_max_out(self.h)
        out = np.zeros(cap, np.complex64)
        n = _i()
        _chk(lib.b200_logmmse_process_host(self.h, x.ctypes.data, len(x), out.ctypes.data, cap, C.byref(n), _stream_ptr(stream)))
        return out[:n.value].copy()

    def noise(self):
        n = _chk(lib.b200_logmmse_noise(self.h, None, 0))
        out = np.zeros(n, np.float32)
        if n:
            _chk(lib.b200_logmmse_noise(self.h, out.ctypes.data, n))
        return out


class SignalDetector:
    """dsp::detector::SignalDetector (dsp/detector/signal_detector.{h,cpp}) on one GPU."""

    def __init__(self, sample_rate=0.0, device=0, _handle=None):
        self.device = device
        self._owned = _handle is None
        if _handle is None:
            h = _vp()
            _chk(lib.b200_sigdet_create(C.byref(h), float(sample_rate), int(device)))
            _handle = h
        self.h = _handle

    def close(self):
        if self.h and self._owned:
            lib.b200_sigdet_destroy(self.h)
        self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_sample_rate(self, sr): _chk(lib.b200_sigdet_set_sample_rate(self.h, float(sr)))
    def set_center_frequency(self, f): _chk(lib.b200_sigdet_set_center_frequency(self.h, float(f)))
    def set_enabled(self, e): _chk(lib.b200_sigdet_set_enabled(self.h, int(e)))
    def is_enabled(self): return bool(lib.b200_sigdet_is_enabled(self.h))
    def fft_size(self): return lib.b200_sigdet_fft_size(self.h)
    def rows(self): return int(lib.b200_sigdet_rows(self.h))

    def process(self, d_in, count=None, stream=None):
        if count is None:
            count = d_in.shape[0]
        _chk(lib.b200_sigdet_process(self.h, d_in.data_ptr() if count else None, int(count), _stream_ptr(stream)))
        return self.rows()

    def process_host(self, x, stream=None):
        x = np.ascontiguousarray(x, np.complex64)
        _chk(lib.b200_sigdet_process_host(self.h, x.ctypes.data, len(x), _stream_ptr(stream)))
        return self.rows()

    def add_row(self, d_row, stream=None):
        _chk(lib.b200_sigdet_add_row(self.h, d_row.data_ptr(), _stream_ptr(stream)))

    def result(self):
        n = _chk(lib.b200_sigdet_result(self.h, None, 0))
        out = np.zeros(n, np.float32)
        if n:
            _chk(lib.b200_sigdet_result(self.h, out.ctypes.data, n))
        return out

    def last_row(self):
        _chk(lib.b200_sigdet_sync(self.h))
        return _dev_to_numpy(lib.b200_sigdet_last_row_ptr(self.h), self.fft_size(), np.float32, self.device)

    def history(self):
        n = self.fft_size()
        k = _chk(lib.b200_sigdet_history(self.h, -1, None, 0))
        rows = np.zeros((k, n), np.float32)
        for r in range(k):
            _chk(lib.b200_sigdet_history(self.h, r, rows[r].ctypes.data, n))
        return rows


class FrontEnd:
    """IQFrontEnd (signal_path/iq_frontend.{h,cpp}) on one GPU."""

    def __init__(self, sample_rate, decim_ratio=1, dc_blocking=False, invert_iq=False, fft_size=65536, fft_rate=20.0,
                 fft_window=2, max_block=1000000, device=0):
        h = _vp()
        _chk(lib.b200_frontend_create(C.byref(h), float(sample_rate), int(decim_ratio), int(dc_blocking), int(invert_iq),
                                      int(fft_size), float(fft_rate), int(fft_window), int(max_block), int(device)))
        self.h = h
        self.device = device
        self.fft_size = fft_size
        self.vfos = VfoBank(0, device=device, _handle=lib.b200_frontend_vfo_bank(self.h))

    def close(self):
        if self.h:
            lib.b200_frontend_destroy(self.h)
        self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def detector(self):
        """the front end's own detector pre-processor (IQFrontEnd::detectorPreprocessor), disabled until set_detector(True)"""
        h = lib.b200_frontend_detector(self.h)
        if not h:
            raise RuntimeError(lib.b200_last_error().decode())
        return SignalDetector(device=self.device, _handle=_vp(h))

    def set_detector(self, enabled): _chk(lib.b200_frontend_set_detector(self.h, int(enabled)))
    def set_sample_rate(self, sr): _chk(lib.b200_frontend_set_sample_rate(self.h, float(sr)))
    def set_decimation(self, r): _chk(lib.b200_frontend_set_decimation(self.h, int(r)))
    def set_dc_blocking(self, e): _chk(lib.b200_frontend_set_dc_blocking(self.h, int(e)))
    def set_invert_iq(self, e): _chk(lib.b200_frontend_set_invert_iq(self.h, int(e)))

    def set_fft_size(self, n):
        _chk(lib.b200_frontend_set_fft_size(self.h, int(n)))
        self.fft_size = n

    def set_fft_rate(self, r): _chk(lib.b200_frontend_set_fft_rate(self.h, float(r)))
    def set_fft_window(self, w): _chk(lib.b200_frontend_set_fft_window(self.h, int(w)))
    def set_frame_shard(self, rank, world): _chk(lib.b200_frontend_set_frame_shard(self.h, int(rank), int(world)))
    def effective_samplerate(self): return lib.b200_frontend_effective_samplerate(self.h)

    def process(self, d_in, count=None, stream=None):
        if count is None:
            count = d_in.shape[0]
        _chk(lib.b200_frontend_process(self.h, d_in.data_ptr() if count else None, int(count), _stream_ptr(stream)))

    def frame_count(self): return _chk(lib.b200_frontend_frame_count(self.h))
    def frame_ptr(self, i): return lib.b200_frontend_frame_ptr(self.h, i)
    def frame_index(self, i): return lib.b200_frontend_frame_index(self.h, i)

    def frames(self):
        n = self.frame_count()
        return [_dev_to_numpy(self.frame_ptr(i), self.fft_size, np.float32, self.device) for i in range(n)]

    def attach_waterfall(self, wf, d_zoom_ptr, d_rgba_ptr, max_rows):
        """run wf's row reduction inside the FFT branch: frame i of a call -> row i of the two device buffers"""
        _chk(lib.b200_frontend_attach_waterfall(self.h, wf.h if wf is not None else None, int(wf.w) if wf is not None else 0,
                                                d_zoom_ptr, d_rgba_ptr, int(max_rows)))
        self._wf = wf      # keep it alive

    def preproc_out(self):
        p, n = _vp(), _i()
        _chk(lib.b200_frontend_preproc_out(self.h, C.byref(p), C.byref(n)))
        return _dev_to_numpy(p.value, n.value, np.complex64, self.device)

    def process_host(self, x, max_frames=0, vfos=(), audio=(), stream=None, i16_scalar=None, i8_scalar=None):
        """x: complex64 block, or (with i16_scalar / i8_scalar) an int16 / int8 array [n, 2] of I/Q converted on the device"""
        if i8_scalar is not None:
            x = np.ascontiguousarray(x, np.int8).reshape(-1, 2)
        elif i16_scalar is not None:
            x = np.ascontiguousarray(x, np.int16).reshape(-1, 2)
        else:
            x = np.ascontiguousarray(x, np.complex64)
        frames = np.empty((max_frames, self.fft_size), np.float32) if max_frames else None
        nslots = (max(list(vfos) + list(audio)) + 1) if (vfos or audio) else 0
        obufs = {v: np.empty(self.vfos.max_out(v, len(x)), np.complex64) for v in vfos}
        abufs = {v: np.empty((self.vfos.audio_max_out(v, len(x)), 2), np.float32) for v in audio}
        optrs = (_vp * max(nslots, 1))()
        aptrs = (_vp * max(nslots, 1))()
        for v in vfos:
            optrs[v] = obufs[v].ctypes.data
        for v in audio:
            aptrs[v] = abufs[v].ctypes.data
        counts = np.zeros(max(nslots, 1), np.int32)
        if i8_scalar is not None:
            _chk(lib.b200_frontend_process_host_i8(self.h, x.ctypes.data, len(x), float(i8_scalar), frames.ctypes.data if max_frames else None, max_frames,
                                                   optrs if vfos else None, aptrs if audio else None, counts.ctypes.data, _stream_ptr(stream)))
        elif i16_scalar is not None:
            _chk(lib.b200_frontend_process_host_i16(self.h, x.ctypes.data, len(x), float(i16_scalar), frames.ctypes.data if max_frames else None, max_frames,
                                                    optrs if vfos else None, aptrs if audio else None, counts.ctypes.data, _stream_ptr(stream)))
        else:
            _chk(lib.b200_frontend_process_host(self.h, x.ctypes.data, len(x), frames.ctypes.data if max_frames else None, max_frames,
                                                optrs if vfos else None, aptrs if audio else None, counts.ctypes.data, _stream_ptr(stream)))
        nf = self.frame_count()
        return dict(frames=frames[:min(nf, max_frames)] if max_frames else None, n_frames=nf,
                    vfo={v: obufs[v][:counts[v]] for v in vfos}, audio={v: abufs[v] for v in audio})


class FrontEndGroup:
    """IQFrontEnd over several GPUs in one process (b200_frontend_group_*): member 0 ingests and owns the display."""

    def __init__(self, devices, sample_rate, decim_ratio=1, dc_blocking=False, invert_iq=False, fft_size=65536, fft_rate=20.0,
                 fft_window=2, max_block=1000000):
        h = _vp()
        devs = (C.c_int * len(devices))(*devices)
        _chk(lib.b200_frontend_group_create(C.byref(h), devs, len(devices), float(sample_rate), int(decim_ratio), int(dc_blocking), int(invert_iq),
                                            int(fft_size), float(fft_rate), int(fft_window), int(max_block)))
        self.h, self.devices, self.fft_size = h, list(devices), fft_size
        self.banks = [VfoBank(0, device=d, _handle=lib.b200_frontend_vfo_bank(lib.b200_frontend_group_member(h, m))) for m, d in enumerate(devices)]

    def close(self):
        if self.h:
            lib.b200_frontend_group_destroy(self.h)
        self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def uses_nccl(self): return bool(_chk(lib.b200_frontend_group_uses_nccl(self.h)))
    def add_vfo(self, out_sr, bw, offset): return _chk(lib.b200_frontend_group_add_vfo(self.h, float(out_sr), float(bw), float(offset)))
    def remove_vfo(self, v): _chk(lib.b200_frontend_group_remove_vfo(self.h, int(v)))

    def location(self, v):
        m, loc = _i(), _i()
        _chk(lib.b200_frontend_group_vfo_location(self.h, int(v), C.byref(m), C.byref(loc)))
        return m.value, loc.value

    def attach_waterfall(self, wf, d_zoom_ptr, d_rgba_ptr, max_rows):
        _chk(lib.b200_frontend_group_attach_waterfall(self.h, wf.h if wf is not None else None, int(wf.w) if wf is not None else 0, d_zoom_ptr, d_rgba_ptr, int(max_rows)))
        self._wf = wf

    def process(self, d_in, count=None, stream=None):
        if count is None:
            count = d_in.shape[0]
        _chk(lib.b200_frontend_group_process(self.h, d_in.data_ptr() if count else None, int(count), _stream_ptr(stream)))

    def process_host(self, x, stream=None, i16_scalar=None):
        if i16_scalar is not None:
            x = np.ascontiguousarray(x, np.int16).reshape(-1, 2)
            _chk(lib.b200_frontend_group_process_host_i16(self.h, x.ctypes.data, len(x), float(i16_scalar), _stream_ptr(stream)))
        else:
            x = np.ascontiguousarray(x, np.complex64)
            _chk(lib.b200_frontend_group_process_host(self.h, x.ctypes.data, len(x), _stream_ptr(stream)))
        self.sync()           # the host block may be reused by the caller

    def sync(self): _chk(lib.b200_frontend_group_sync(self.h))
    def frame_count(self): return _chk(lib.b200_frontend_group_frame_count(self.h))
    def frame_ptr(self, i): return lib.b200_frontend_group_frame_ptr(self.h, i)
    def frame_index(self, i): return lib.b200_frontend_group_frame_index(self.h, i)

    def frames(self):
        self.sync()
        return [_dev_to_numpy(self.frame_ptr(i), self.fft_size, np.float32, self.devices[0]) for i in range(self.frame_count())]

    def output(self, v):
        """IF output of group VFO v (fetched from the member that owns it)"""
        self.sync()
        m, loc = self.location(v)
        return self.banks[m].output(loc)

    def out_count(self, v):
        m, loc = self.location(v)
        return self.banks[m].out_count(loc)

    def broadcast_ms(self):
        r = _d()
        _chk(lib.b200_frontend_group_broadcast_ms(self.h, C.byref(r)))
        return r.value


class Waterfall:
    """doZoom + palette mapping of WaterFall::pushFFT (gui/widgets/waterfall.cpp:98-154,1229-1235), headless."""

    def __init__(self, raw_fft_size, data_width, device=0):
        h = _vp()
        _chk(lib.b200_waterfall_create(C.byref(h), int(raw_fft_size), int(data_width), int(device)))
        self.h, self.n, self.w, self.device = h, raw_fft_size, data_width, device

    def close(self):
        if self.h:
            lib.b200_waterfall_destroy(self.h)
        self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_view(self, offset, width): _chk(lib.b200_waterfall_set_view(self.h, int(offset), int(width)))
    def set_range(self, lo, hi): _chk(lib.b200_waterfall_set_range(self.h, float(lo), float(hi)))

    def set_palette(self, lut):
        lut = np.ascontiguousarray(lut, np.uint32)
        assert lut.shape == (1000,)
        _chk(lib.b200_waterfall_set_palette(self.h, lut.ctypes.data))

    @staticmethod
    def make_palette(colors):
        colors = np.ascontiguousarray(colors, np.float32)
        lut = np.zeros(1000, np.uint32)
        _chk(lib.b200_waterfall_make_palette(colors.ctypes.data, colors.shape[0], lut.ctypes.data))
        return lut

    def push(self, d_rows_ptr, nrows, d_zoom_ptr, d_rgba_ptr, stream=None):
        _chk(lib.b200_waterfall_push(self.h, d_rows_ptr, int(nrows), d_zoom_ptr, d_rgba_ptr, _stream_ptr(stream)))

    def set_history(self, height):
        _chk(lib.b200_waterfall_set_history(self.h, int(height)))

    def regenerate(self, d_zoom_ptr, d_rgba_ptr, stream=None):
        """all stored raw rows through the current view/range/palette; returns the number of stored rows"""
        return _chk(lib.b200_waterfall_regenerate(self.h, d_zoom_ptr, d_rgba_ptr, _stream_ptr(stream)))

    def vfo_signal(self, d_rows_ptr, nrows, whole_bw, center, bw, d_out_ptr, stream=None):
        """returns False where the reference's calculateVFOSignalInfo returns false"""
        return _chk(lib.b200_waterfall_vfo_signal(self.h, d_rows_ptr, int(nrows), float(whole_bw), float(center), float(bw), d_out_ptr, _stream_ptr(stream))) == 0

    def smooth_hold(self, d_zoom_ptr, nrows, smoothing, alpha, hold, hold_speed, d_hold_ptr=None, stream=None):
        _chk(lib.b200_waterfall_smooth_hold(self.h, d_zoom_ptr, int(nrows), int(smoothing), float(alpha), int(hold), float(hold_speed), d_hold_ptr, _stream_ptr(stream)))

    def zoom_table(self):
        s = np.zeros(self.w, np.int32)
        u = np.zeros(self.w, np.int32)
        _chk(lib.b200_waterfall_zoom_table(self.h, s.ctypes.data, u.ctypes.data))
        return s, u


def profile_enable(on):
    _chk(lib.b200_profile_enable(int(on)))


def profile_read():
    """(kernel ms, launches, algorithmic bytes) of the fused xlate+decimate kernel since the last read."""
    ms, n, by = _d(), C.c_longlong(), _d()
    _chk(lib.b200_profile_read(C.byref(ms), C.byref(n), C.byref(by)))
    return ms.value, n.value, by.value


PROF_STAGE1, PROF_FFT, PROF_DC = 0, 1, 2


def profile_read_kind(kind):
    """(ms, launches, algorithmic bytes, algorithmic flops) of one kernel family since the last read."""
    ms, n, by, fl = _d(), C.c_longlong(), _d(), _d()
    _chk(lib.b200_profile_read_kind(int(kind), C.byref(ms), C.byref(n), C.byref(by), C.byref(fl)))
    return ms.value, n.value, by.value, fl.value


def measure_fp32_peak(device=0):
    """FFMA per second of `device`, measured now (three-register FFMA chains on every SM)."""
    r = _d()
    _chk(lib.b200_measure_fp32_peak(int(device), C.byref(r)))
    return r.value


DC_NOALLOC_MAX_SAMPLES = 1 << 24


def dc_block_noalloc(d_in, d_out, count, rate, d_offset=None, stream=None):
    """DC blocker through the allocation-free entry point (fixed scratch inside the library, <= 2**24 samples
    per call).  d_in / d_out: CUDA float32 tensors [count, 2]; d_offset: CUDA float64 tensor [2] carried
    between calls, or None to start from zero."""
    _chk(lib.b200_dc_block_noalloc(d_in.data_ptr(), d_out.data_ptr(), int(count), float(rate),
                                   d_offset.data_ptr() if d_offset is not None else None, _stream_ptr(stream)))


def fft_forward(d_in, nz, fft_size, d_out, stream=None):
    _chk(lib.b200_fft_forward(d_in.data_ptr(), int(nz), int(fft_size), d_out.data_ptr(), _stream_ptr(stream)))
