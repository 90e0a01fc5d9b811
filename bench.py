#!/usr/bin/env python3
"""Benchmark of the IQ-front-end hot path (BASELINE.json: complex MS/s through FFT + N-VFO channelizer).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path (C ABI of libb200dsp.so)
    python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU blocks on the host cores

Workload: BASELINE config[3] (the largest configuration that fits one GPU) -- 122.88 MS/s IQ, 1024 mixed RxVFOs
(512 AM 15 kS/s, 384 USB 24 kS/s, 128 WFM 500 kS/s: xlate, power-of-two decimator, polyphase resampler, channel filter)
and a 2^20-point Blackman FFT waterfall at 60 fps (window + FFT + log power + max-bin zoom + palette).  The metric is
"through FFT + N-VFO channelizer", so the timed step ends at every VFO's IF output and the waterfall rows.
A step is one pass of that path over one 16 Mi-sample batch of synthetic cf32 IQ that is resident in HBM when the timed
region starts.  STRONG scaling: with N > 1 ranks the same bank is dealt out over the ranks by cost
(sharding.shard_vfos) and the FFT frames by time (f % N == rank); rank 0 owns the stream and every step's batch is
NCCL-broadcast to the other ranks (double buffered, behind the previous step's kernels).  `value` is the stream rate
the whole job sustains: batch samples x steps / time (max over ranks).

Beside the headline: `roofline` holds the two dominant kernels (the fused first-stage bank kernel against the FP32
roof measured in the same run, the FFT passes against the measured HBM copy peak); `extra.cfg2` / `extra.cfg3` are the
same measurement on BASELINE config[1] (one USB VFO + 65536-pt FFT at 20 MS/s, HBM-bound) and config[2] (256 NFM VFOs at
61.44 MS/s); `cpu_baseline` is the reference's blocks on ALL host cores (RxVFO::process over a thread pool).

One JSON line on stdout (rank 0).  See the task contract for the keys.
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import tempfile
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from sdrplusplusbrown_b200 import configs, sharding  # noqa: E402

METRIC = "complex MS/s through FFT + N-VFO channelizer"
UNIT = "MS/s"
WORKLOAD = "cfg4"
DATA_WIDTH = 1333                    # waterfall pixels per row (a typical dataWidth)
WF_MIN, WF_MAX = -120.0, -10.0
E2E_DEFER_TAIL = True                # the host-fed path in streaming mode (B200_OPT_DEFER_TAIL)?  tools/bench_e2e.py measures both
# first stage of each power-of-two plan: (decimation, taps) -- FMA lane operations per input sample and VFO in the
# fused kernel = 4 (rotation) + 2 * taps / decimation
STAGE1 = {16: (8, 54), 32: (8, 44), 64: (8, 36), 128: (16, 72), 256: (32, 143), 512: (32, 129), 1024: (64, 257),
          2048: (64, 329), 4096: (64, 400), 8192: (128, 726)}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json copy bandwidth)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def workload_text(name):
    cfg = configs.CONFIGS[name]
    vl = configs.vfo_list(name)
    modes = {}
    for m, _, _, _ in vl:
        modes[m] = modes.get(m, 0) + 1
    return "BASELINE %s: %.2f MS/s synthetic IQ, %d RxVFOs (%s), %d-pt FFT waterfall at %g fps (+zoom to %d px, palette)" % (
        name, cfg["sample_rate"] / 1e6, len(vl), ", ".join("%d %s" % (n, m) for m, n in sorted(modes.items())), cfg["fft_size"], cfg["fft_rate"], DATA_WIDTH)


def make_config(name):
    """The same dict in both arms (the driver compares them); run-specific sizes live in `run`."""
    cfg = configs.CONFIGS[name]
    return {"workload": workload_text(name), "sample_rate": cfg["sample_rate"], "fft_size": cfg["fft_size"], "fft_rate": cfg["fft_rate"],
            "window": {0: "rectangular", 1: "blackman", 2: "nuttall"}[cfg["window"]], "vfos": len(configs.vfo_list(name)),
            "demod": False, "rotator": "fast (closed form, calibrated step)",
            "parallelism": "VFO bank dealt out over the GPUs by cost, every batch broadcast from rank 0, FFT frames time-sharded"}


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    QUERY = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown," \
            "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, device):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(device), "--query-gpu=" + self.QUERY, "--format=csv,noheader,nounits", "-lms", "50"],
                                      stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        for line in self.f.read().splitlines():
            parts = [x.strip() for x in line.split(",")]
            if len(parts) < 8:
                continue
            try:
                sm.append(float(parts[0])); mx.append(float(parts[1]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), parts[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        try:
            os.unlink(self.f.name)
        except OSError:
            pass
        if sm:
            out["sm_mhz"] = float(np.median(sm))
            out["sm_max_mhz"] = float(max(mx))
            out["samples"] = len(sm)
        out["reasons"] = sorted(reasons)
        return out


# ------------------------------------------------------------------------------------------------ CPU (reference) arm
def _oracle():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as ol
    # timing builds (-O3 -march=native); the reference's own headers when they were compiled here
    # (oracle/_ref travels to the GPU box as a prebuilt .so), else the self-contained port.
    # NOTE: -march=native code built in this container may not run on the GPU box's CPU -> fall back.
    for which in ("ref_fast", "ref", "port_fast", "port"):
        try:
            if which.endswith("_fast"):
                # an illegal instruction cannot be caught in-process: probe the -march=native build in a child
                probe = "import sys; sys.path.insert(0, %r); import oracle_lib as ol, numpy as np; o = ol.Oracle(%r); " \
                        "v = ol.Vfo(o, 2.4e6, 50e3, 12.5e3, 1e5); v.process(np.ones(50000, np.complex64))" % (os.path.join(ROOT, "tests"), which)
                if subprocess.run([sys.executable, "-c", probe], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, timeout=120).returncode != 0:
                    continue
            o = ol.Oracle(which)
            o.lowpass(1000.0, 100.0, 48000.0)
            return ol, o
        except Exception:
            continue
    raise RuntimeError("no oracle library could be loaded")


def host_cores():
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except Exception:
        return max(1, os.cpu_count() or 1)


def cpu_model():
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except Exception:
        pass
    return "unknown"


class CpuPath:
    """The reference's blocks for the workload in the reference's best case: RxVFO::process
    (channel/rx_vfo.h:89-100) called directly, no stream threads and no Splitter copies, the VFOs dealt round robin over a
    pool of one thread per host core (the reference itself runs one thread per block, dsp/block.h:82-84), 1 M-sample
    blocks (STREAM_BUFFER_SIZE, dsp/stream.h:14); the FFT branch (window, FFT, log power, zoom, palette: one thread in
    the reference) is one more task of the same pool.  The oracle calls release the GIL (ctypes)."""

    def __init__(self, cfg_name=WORKLOAD, threads=None):
        self.ol, self.orc = _oracle()
        cfg = configs.CONFIGS[cfg_name]
        self.sr = cfg["sample_rate"]
        self.N = cfg["fft_size"]
        self.skip, self.nz = self.orc.reshape_params(self.sr, self.N, cfg["fft_rate"])
        self.interval = self.nz + self.skip
        self.window = self.orc.fft_window(cfg["window"], self.nz)
        self.vfos = [self.ol.Vfo(self.orc, self.sr, out_sr, bw, off) for _, out_sr, bw, off in configs.vfo_list(cfg_name)]
        self.threads = threads or host_cores()
        self.pool = ThreadPoolExecutor(max_workers=self.threads)
        self.pos = 0
        self.buf = np.zeros(0, np.complex64)
        self.buf_start = 0
        self.next_frame = 0
        self.kind = "reference" if self.orc.kind == "reference" else "port"

    def _vfo_task(self, k, x):
        for v in self.vfos[k::self.threads]:
            for i in range(0, len(x), 1000000):
                v.process(x[i:i + 1000000])

    def _fft_task(self, x):
        # Reshaper framing: frame f = samples [f*I, f*I + nz) of the stream (iq_frontend.h:88-92); samples are kept until
        # the frame they belong to is complete (the reference's ring buffer does the same)
        self.buf = np.concatenate([self.buf, x]) if len(self.buf) else x
        end = self.buf_start + len(self.buf)
        while self.next_frame * self.interval + self.nz <= end:
            s = self.next_frame * self.interval - self.buf_start
            db = self.orc.fft_frame(self.buf[s:s + self.nz], self.window, self.N)
            z = self.orc.do_zoom(0, self.N, db, DATA_WIDTH)
            self.orc.palette_indices(z, WF_MIN, WF_MAX)
            self.next_frame += 1
        drop = min(len(self.buf), max(0, self.next_frame * self.interval - self.buf_start))
        self.buf = self.buf[drop:].copy()
        self.buf_start += drop

    def step(self, x):
        """x: complex64 numpy block.  Returns when every task has finished (the slowest thread bounds the rate)."""
        futs = [self.pool.submit(self._fft_task, x)] + [self.pool.submit(self._vfo_task, k, x) for k in range(self.threads)]
        for fu in futs:
            fu.result()
        self.pos += len(x)


def cpu_sample_text(n, cfg_name, threads):
    return ("%d-sample steps (%.1f ms of signal) of %s; reference blocks compiled from its own headers against generic-C VOLK and a float64 "
            "radix-2 FFT stand-in (real VOLK SIMD / FFTW are not installable here, so this under-states the reference); RxVFO::process "
            "over a pool of %d threads = all host cores (%s)" % (n, 1e3 * n / configs.CONFIGS[cfg_name]["sample_rate"], cfg_name, threads, cpu_model()))


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from sdrplusplusbrown_b200 import synth
    cfg = configs.CONFIGS[WORKLOAD]
    sample = args.cpu_samples
    x = synth.make_iq(0, sample, cfg["sample_rate"], tones=[(1.25e6 + 700.0, 0.05), (3.3e6, 0.2)])
    cpu = CpuPath()
    for _ in range(args.warmup):
        cpu.step(x)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu.step(x)
    dt = time.perf_counter() - t0
    value = sample * args.steps / dt / 1e6
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": make_config(WORKLOAD),
        "run": {"batch_samples_per_step": sample, "l2_policy": "n/a (CPU arm)"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cpu.threads, "kind": cpu.kind, "sample": cpu_sample_text(sample, WORKLOAD, cpu.threads)},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    # the graph as the reference itself runs it -- a Splitter copying every block to one stream per VFO,
    # one thread per RxVFO -- for as many VFOs of the workload as the host has cores (reference blocks only; a few seconds)
    if cpu.kind == "reference":
        try:
            vl = configs.vfo_list(WORKLOAD)
            n_v = min(cpu.threads, len(vl))
            pick = [vl[(i * len(vl)) // n_v] for i in range(n_v)]
            blk = x[:min(len(x), 1000000)]
            secs = cpu.ol.graph_run(cpu.orc, cfg["sample_rate"], [(v[1], v[2], v[3]) for v in pick], blk, 2)
            if secs > 0:
                line["cpu_baseline"]["threaded_graph"] = {
                    "value": 2 * len(blk) / secs / 1e6, "unit": UNIT, "vfos": n_v, "threads": n_v + 1,
                    "what": "dsp::routing::Splitter + one dsp::channel::RxVFO thread per VFO + Null sinks (iq_frontend.cpp:159-180), %d of the workload's "
                            "%d VFOs (one per host core), 1 M-sample blocks, no FFT branch; scale by %d / %d for the whole bank on this host"
                            % (n_v, len(vl), n_v, len(vl))}
        except Exception as e:       # the measurement is an extra: never fail the arm over it
            line["cpu_baseline"]["threaded_graph"] = {"unavailable": str(e)}
    # the FFT branch on ONE thread, as the reference runs it (IQFrontEnd::handler on the FFT sink's thread,
    # signal_path/iq_frontend.cpp:267-287: window, forward FFT, log power; then the waterfall's zoom and palette), with the
    # oracle's FFT stand-in -- and numpy's pocketfft in float32 on the same windowed frame as a sanity comparison of the
    # transform alone (FFTW itself is not installable here)
    try:
        nz, N, I = cpu.nz, cpu.N, cpu.interval
        frame = np.resize(x, nz).astype(np.complex64)
        reps = 2 if N >= (1 << 18) else 8
        cpu.orc.fft_frame(frame, cpu.window, N)
        t0 = time.perf_counter()
        for _ in range(reps):
            db = cpu.orc.fft_frame(frame, cpu.window, N)
            cpu.orc.palette_indices(cpu.orc.do_zoom(0, N, db, DATA_WIDTH), WF_MIN, WF_MAX)
        t_frame = (time.perf_counter() - t0) / reps
        xw = np.zeros(N, np.complex64)
        xw[:nz] = frame * np.asarray(cpu.window, np.float32)[:nz]
        np.fft.fft(xw)
        t0 = time.perf_counter()
        for _ in range(reps):
            y = np.fft.fft(xw)
        t_np = (time.perf_counter() - t0) / reps
        line["cpu_baseline"]["fft_branch"] = {
            "ms_per_frame": 1e3 * t_frame, "frames_per_s": 1.0 / t_frame, "stream_rate_it_sustains": I / t_frame / 1e6, "unit": UNIT, "threads": 1,
            "numpy_pocketfft_ms_per_transform": 1e3 * t_np, "numpy_dtype": str(y.dtype),
            "what": "window + %d-point forward FFT (oracle's float64 stand-in for FFTW) + log power + zoom to %d px + palette on one thread; a frame every "
                    "%d samples, so the branch alone keeps up with this stream rate; numpy: the bare transform of the same windowed frame" % (N, DATA_WIDTH, I)}
    except Exception as e:
        line["cpu_baseline"]["fft_branch"] = {"unavailable": str(e)}
    print(json.dumps(line))


def cpu_baseline(args):
    from sdrplusplusbrown_b200 import synth
    cfg = configs.CONFIGS[WORKLOAD]
    n = args.cpu_samples
    x = synth.make_iq(0, n, cfg["sample_rate"], tones=[(1.25e6 + 700.0, 0.05), (3.3e6, 0.2)])
    cpu = CpuPath()
    cpu.step(x[:200000])
    t0 = time.perf_counter()
    reps = 0
    while True:
        cpu.step(x)
        reps += 1
        if time.perf_counter() - t0 > args.cpu_seconds:
            break
    dt = time.perf_counter() - t0
    return {"value": n * reps / dt / 1e6, "unit": UNIT, "cores": cpu.threads, "kind": cpu.kind, "nproc": os.cpu_count(), "cpu_model": cpu_model(),
            "sample": "%d x %s, %.1f s of wall time" % (reps, cpu_sample_text(n, WORKLOAD, cpu.threads), dt)}


# ------------------------------------------------------------------------------------------------ GPU arm
class Job:
    """One front end with this rank's share of a configuration's VFO bank and its waterfall reduction attached."""

    def __init__(self, binding, torch, cfg_name, B, local, rank, world, defer_tail=True, vfo_indices=None, demod=None):
        cfg = configs.CONFIGS[cfg_name]
        self.cfg, self.name, self.B = cfg, cfg_name, B
        self.sr = cfg["sample_rate"]
        dev = torch.device("cuda", local)
        self.fe = binding.FrontEnd(self.sr, fft_size=cfg["fft_size"], fft_rate=cfg["fft_rate"], fft_window=cfg["window"], max_block=B, device=local)
        if world > 1:
            self.fe.set_frame_shard(rank, world)
        vl = configs.vfo_list(cfg_name)
        self.mine = sorted(sharding.shard_vfos(vl, self.sr, world)[rank]) if vfo_indices is None else list(vfo_indices)
        self.ids = []
        self.lane_ops_per_sample = 0.0          # FMA lane operations per input sample of this rank's fused first stages
        for i in self.mine:
            _, out_sr, bw, off = vl[i]
            v = self.fe.vfos.add(out_sr, bw, off)
            self.ids.append(v)
            d = self.fe.vfos.describe(v)
            if d["fused"]:
                D, T = STAGE1[d["ratio"]]
                self.lane_ops_per_sample += 4.0 + 2.0 * T / D
            if demod is not None:
                # the radio module's demodulator for the VFO's mode, with the module's defaults (demodulators/*.h)
                mode = vl[i][0]
                if mode == "AM":
                    self.fe.vfos.set_demod(v, binding.DEMOD_AM, 1, 50.0 / out_sr, 5.0 / out_sr, 100.0 / out_sr)
                elif mode in ("USB", "LSB"):
                    self.fe.vfos.set_demod(v, binding.DEMOD_USB if mode == "USB" else binding.DEMOD_LSB, 0, 50.0 / out_sr, 5.0 / out_sr, 0)
                elif mode == "WFM":
                    self.fe.vfos.set_demod(v, binding.DEMOD_WFM, 1 if demod == "stereo" else 0, 1, 0, 0)
                else:
                    self.fe.vfos.set_demod(v, binding.DEMOD_NFM, 1, 0, 0, 0)
        self.fe.vfos.set_option(binding.OPT_DEFER_TAIL, 1 if defer_tail else 0)
        self.wf = binding.Waterfall(cfg["fft_size"], DATA_WIDTH, device=local)
        self.wf.set_view(0, cfg["fft_size"])
        self.wf.set_range(WF_MIN, WF_MAX)
        self.max_frames = B // int(round(self.sr / cfg["fft_rate"])) + 2
        self.d_zoom = torch.empty((self.max_frames, DATA_WIDTH), dtype=torch.float32, device=dev)
        self.d_rgba = torch.empty((self.max_frames, DATA_WIDTH), dtype=torch.int32, device=dev)
        self.fe.attach_waterfall(self.wf, self.d_zoom.data_ptr(), self.d_rgba.data_ptr(), self.max_frames)

    def join(self):
        """orders the last steps' deferred stages into the caller's stream"""
        if self.ids:
            self.fe.vfos.out_ptr(self.ids[0])


def kernel_rooflines(binding, steps, ms_total, lane_ops, ffma_peak, hbm_peak, hbm_src, fused_name, fft_name):
    """roofline entries of the two dominant kernel families from the library's CUDA-event hooks (timed on the launch stream)"""
    kms, kl, kb, kf = binding.profile_read_kind(binding.PROF_STAGE1)
    fms, fl, fb, ff = binding.profile_read_kind(binding.PROF_FFT)
    out = {}
    if kms > 0:
        gbs = kb / 1e9 / (kms * 1e-3)
        tfl = kf / 1e12 / (kms * 1e-3)
        ops = lane_ops / (kms * 1e-3)
        fp32_frac = tfl / (2.0 * ffma_peak / 1e12) if ffma_peak else None
        hbm_frac = gbs / hbm_peak
        fp32_bound = fp32_frac is not None and fp32_frac >= hbm_frac
        out["fused"] = {
            "kernel": fused_name, "bound": "fp32" if fp32_bound else "hbm",
            "achieved": tfl if fp32_bound else gbs, "peak": (2.0 * ffma_peak / 1e12) if fp32_bound else hbm_peak, "unit": "TFLOP/s" if fp32_bound else "GB/s",
            "frac": fp32_frac if fp32_bound else hbm_frac,
            "fp32": {"achieved_tflops": tfl, "peak_tflops": 2.0 * ffma_peak / 1e12 if ffma_peak else None, "frac": fp32_frac,
                     "flops": "SURVEY 8(d): 6 + 4 T/D per input sample and VFO", "fma_lane_ops_per_s": ops,
                     "frac_by_instruction_count": ops / ffma_peak if ffma_peak else None,
                     "peak_source": "FMA rate measured in this run (b200_measure_fp32_peak: packed FFMA2 chains on every SM, the form that reaches 128 FMA/clk/SM), x2 flop"},
            "hbm": {"achieved_gbs": gbs, "peak_gbs": hbm_peak, "frac": hbm_frac, "peak_source": hbm_src},
            "launches": int(kl), "avg_launch_ms": kms / kl if kl else None, "algorithmic_bytes_per_launch": kb / kl if kl else None,
            "ms_per_step": kms / steps, "share_of_step": kms / ms_total if ms_total else None,
        }
    if fms > 0:
        gbs = fb / 1e9 / (fms * 1e-3)
        out["fft"] = {"kernel": fft_name, "bound": "hbm", "achieved": gbs, "peak": hbm_peak, "unit": "GB/s", "frac": gbs / hbm_peak, "peak_source": hbm_src,
                      "bytes": "8 per windowed sample + 4 per dB value written, per frame", "calls": int(fl), "avg_call_ms": fms / fl if fl else None,
                      "algorithmic_bytes_per_call": fb / fl if fl else None, "ms_per_step": fms / steps, "share_of_step": fms / ms_total if ms_total else None,
                      "gflops": ff / 1e9 / (fms * 1e-3)}
    return out


def run_gpu(args):
    import torch
    import torch.distributed as dist
    from sdrplusplusbrown_b200 import binding, synth

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (there is no CPU fallback)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    cfg = configs.CONFIGS[WORKLOAD]
    sr = cfg["sample_rate"]
    B = args.batch
    hbm_peak, hbm_src = peaks()
    vl = configs.vfo_list(WORKLOAD)

    job = Job(binding, torch, WORKLOAD, B, local, rank, world, defer_tail=args.defer_tail)
    fe = job.fe

    # ---- synthetic stream, resident in HBM (rank 0 is the ingest GPU)
    tones = [(vl[i][3] + 0.1 * vl[i][2], 0.02) for i in range(0, len(vl), 64)] + [(3.3e6, 0.2)]
    if rank == 0:
        src = synth.make_iq_torch(B, sr, dev, tones=tones)
        bufs = [src, src]          # the ingest GPU channelizes the stream in place and broadcasts from it
    else:
        src = None
        bufs = [torch.empty((B, 2), dtype=torch.float32, device=dev) for _ in range(2)]
    compute = torch.cuda.Stream(device=dev, priority=-1)      # above the library's deferred-tail stream (lowest priority)
    # the broadcast's NCCL kernel needs a few CTAs beside a bank kernel that fills every SM: above the compute stream, or it
    # only runs in the gaps between the compute kernels (measured at N = 8: 2.0 ms per broadcast and 3.65 ms per step with
    # the communication stream at the default priority)
    comm = torch.cuda.Stream(device=dev, priority=-3) if world > 1 else None
    ev_ready = [torch.cuda.Event() for _ in bufs]
    ev_free = [torch.cuda.Event() for _ in bufs]
    nb = 64
    ev_b0 = [torch.cuda.Event(enable_timing=True) for _ in range(nb)]
    ev_b1 = [torch.cuda.Event(enable_timing=True) for _ in range(nb)]

    def bcast(i, timed=False):
        """rank 0 -> all: batch for step i into buffer i % 2 (NCCL over NVLink), on the comm stream."""
        b = bufs[i % 2]
        with torch.cuda.stream(comm):
            comm.wait_event(ev_free[i % 2])
            if timed and i < nb:
                ev_b0[i].record(comm)
            dist.broadcast(b, src=0)
            if timed and i < nb:
                ev_b1[i].record(comm)
            ev_ready[i % 2].record(comm)

    def run(j, nsteps, timed=False):
        if world > 1:
            for e in ev_free:
                e.record(compute)
            bcast(0, timed)
        for i in range(nsteps):
            if world > 1 and i + 1 < nsteps:
                bcast(i + 1, timed)
            if world > 1:
                compute.wait_event(ev_ready[i % 2])
            with torch.cuda.stream(compute):
                j.fe.process(bufs[i % 2], j.B, stream=compute)      # FFT rows -> zoom + palette rows inside the FFT branch (attached waterfall)
                if world > 1:
                    ev_free[i % 2].record(compute)

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def timed_run(j, nsteps, warm):
        run(j, warm)
        j.join()
        barrier()
        binding.profile_enable(True)
        for k in (binding.PROF_STAGE1, binding.PROF_FFT, binding.PROF_DC):
            binding.profile_read_kind(k)
        l0 = binding.lib.b200_launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record(compute)
        run(j, nsteps, timed=True)
        j.join()            # e1 fires behind ALL work of the K steps, deferred stages included
        e1.record(compute)
        barrier()
        ms = e0.elapsed_time(e1)
        launches = binding.lib.b200_launch_count() - l0
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()), ms, int(launches)

    clocks = ClockSampler(local) if rank == 0 else None      # spans warm-up, the timed region and a short tail of identical steps
    warm = max(args.warmup, 3)
    ms, ms_local, launches = timed_run(job, args.steps, warm)
    value = B * args.steps / (ms * 1e-3) / 1e6
    ffma_peak = binding.measure_fp32_peak(local)
    roof = kernel_rooflines(binding, args.steps, ms_local, job.lane_ops_per_sample * B * args.steps, ffma_peak, hbm_peak, hbm_src,
                            "fusedBankKernel<ratio> (xlate + first decimating FIR stage of every VFO of this GPU's shard: /8192 AM, /4096 USB, /128 WFM plans; "
                            "packed FP32, pairs of VFOs per FFMA2)",
                            "fftPass1Kernel + fftPass2Kernel (window, 2^20-pt FFT as 1024 x 1024, log power)")
    binding.profile_enable(False)
    # the FFT branch by itself (a front end without VFOs on the same resident batch): inside the step its passes share the
    # SMs with the bank kernel, so the events around them mostly measure waiting
    if roof.get("fft") is not None:
        jf = Job(binding, torch, WORKLOAD, B, local, rank, world, defer_tail=False, vfo_indices=[])
        _, ms_f, _ = timed_run(jf, max(5, args.steps // 2), 3)
        rf = kernel_rooflines(binding, max(5, args.steps // 2), ms_f, 0.0, ffma_peak, hbm_peak, hbm_src, "", roof["fft"]["kernel"])
        binding.profile_enable(False)
        if rf.get("fft"):
            roof["fft"]["alone"] = {k: rf["fft"][k] for k in ("achieved", "frac", "avg_call_ms", "calls", "gflops")}
            roof["fft"]["alone"]["note"] = "same frames, same resident batch, FFT branch without the VFO bank beside it (zoom + palette still attached)"
        del jf
    bc_ms = None
    if world > 1:
        n = min(args.steps, nb)
        bc_ms = sum(ev_b0[i].elapsed_time(ev_b1[i]) for i in range(n)) / n
    # the timed region can be shorter than nvidia-smi's sampling period: keep the identical load running for ~0.6 s
    # (same step count on every rank) so the clock record holds samples taken under this workload
    run(job, min(2000, int(600.0 / max(ms / args.steps, 1e-3)) + 1))
    job.join()
    barrier()
    clk = clocks.stop() if clocks else None

    # ---- N > 1: the sharded job's results against a single-GPU run of the same block, bit for bit
    self_check = None
    if world > 1:
        self_check = sharded_self_check(binding, torch, dist, bufs[0] if rank == 0 else bufs[(args.steps - 1) % 2], B, local, rank, world, compute)

    # ---- end to end through the host-buffer C-ABI call (H2D of the block + D2H of frames and every VFO's IF output inside)
    Be = args.e2e_batch
    e2e = host_path(binding, torch, WORKLOAD, Be, local, rank, world, src if src is not None else bufs[0], compute, args.e2e_steps, dist, dev)

    # ---- the lighter BASELINE configurations as extra keys (single GPU only: they are parity / regression lines)
    extra = {}
    if world == 1 and args.extras:
        del job
        torch.cuda.empty_cache()
        # the headline bank with the radio module's demodulators behind every VFO (AM, USB + AGC, WFM mono / stereo): outside
        # the metric ("through FFT + N-VFO channelizer"), reported beside it -- the AGC and the stereo pilot PLL are
        # sequential walks per VFO
        for key, kind in (("cfg4_demod_wfm_mono", "mono"), ("cfg4_demod_wfm_stereo", "stereo")):
            jd = Job(binding, torch, WORKLOAD, B, local, 0, 1, defer_tail=args.defer_tail, demod=kind)
            kd = max(3, args.steps // 4)
            msd, _, ld = timed_run(jd, kd, 2)
            binding.profile_enable(False)
            extra[key] = {"workload": workload_text(WORKLOAD) + ", every VFO with its demodulator (WFM %s)" % kind, "value": B * kd / (msd * 1e-3) / 1e6, "unit": UNIT,
                          "ms_per_step": msd / kd, "steps": kd, "x_real_time": B * kd / (msd * 1e-3) / sr, "gpu_launches": ld}
            del jd
            torch.cuda.empty_cache()
        # the headline bank with the rotators in EXACT mode (what the drop-in IQFrontEnd header runs: the reference's fp32 phase
        # recursion replayed bit for bit, one thread per VFO walking the block) next to the FAST mode the headline is quoted on
        je = Job(binding, torch, WORKLOAD, B, local, 0, 1, defer_tail=args.defer_tail)
        je.fe.vfos.set_rotator_mode(binding.ROTATOR_EXACT)
        ke = 3
        mse, _, le = timed_run(je, ke, 1)
        binding.profile_enable(False)
        extra["cfg4_exact_rotator"] = {"workload": workload_text(WORKLOAD) + ", rotators in B200_ROTATOR_EXACT mode", "value": B * ke / (mse * 1e-3) / 1e6, "unit": UNIT,
                                       "ms_per_step": mse / ke, "steps": ke, "x_real_time": B * ke / (mse * 1e-3) / sr, "gpu_launches": le}
        del je
        torch.cuda.empty_cache()
        for name, Bx, fused_name, fft_name in (("cfg2", 1 << 26, "fusedStreamKernel<512> (single-VFO variant, 16-byte staging)", "fftColsKernel + fftRowsKernel (65536 = 256 x 256)"),
                                               ("cfg3", 1 << 24, "fusedBankKernel<1024> (256 NFM VFOs)", "fftColsKernel + fftRowsKernel (65536 = 256 x 256)")):
            cx = configs.CONFIGS[name]
            jx = Job(binding, torch, name, Bx, local, 0, 1, defer_tail=args.defer_tail)
            sx = synth.make_iq_torch(Bx, cx["sample_rate"], dev, tones=[(configs.vfo_list(name)[0][3] + 700.0, 0.05), (3.3e6, 0.2)])
            bufs[0] = bufs[1] = sx
            kx = max(5, args.steps // 2)
            msx, msx_local, lx = timed_run(jx, kx, 3)
            rx = kernel_rooflines(binding, kx, msx_local, jx.lane_ops_per_sample * Bx * kx, ffma_peak, hbm_peak, hbm_src, fused_name, fft_name)
            binding.profile_enable(False)
            extra[name] = {"workload": workload_text(name), "value": Bx * kx / (msx * 1e-3) / 1e6, "unit": UNIT, "ms_per_step": msx / kx, "steps": kx,
                           "batch_samples_per_step": Bx, "x_real_time": Bx * kx / (msx * 1e-3) / cx["sample_rate"], "gpu_launches": lx, "roofline": rx}
            del jx, sx
            torch.cuda.empty_cache()

    if rank == 0:
        fused = roof.get("fused", {})
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": warm,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": make_config(WORKLOAD),
            "run": {"batch_samples_per_step": B, "l2_policy": "batch (%.0f MB) larger than L2 (126 MB); %d VFO output streams and the FFT intermediates on top" % (B * 8 / 1e6, len(job_mine_count(vl, sr, world, 0))),
                    "vfos_on_rank0": len(job_mine_count(vl, sr, world, 0)), "x_real_time": value * 1e6 / sr,
                    "stage_overlap": ("B200_OPT_DEFER_TAIL: the later stages of step k run on their own stream beside the full-rate stage of step k+1; "
                                      "all K steps' outputs are complete when the timed region ends") if args.defer_tail else "none",
                    "broadcast_ms_per_step_rank0": bc_ms},
            "clocks": clk,
            "e2e": e2e,
            "gpu_launches": int(launches),
            "roofline": dict(fused, **{"traffic": _traffic(), "fft": roof.get("fft"), "fp32_peak_ffma_per_s": ffma_peak}),
        }
        if self_check is not None:
            line["self_check"] = self_check
        if extra:
            line["extra"] = extra
        if args.cpu_baseline and world == 1:
            line["cpu_baseline"] = cpu_baseline(args)
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def job_mine_count(vl, sr, world, rank):
    return sharding.shard_vfos(vl, sr, world)[rank]


def host_path(binding, torch, name, Be, local, rank, world, d_src, compute, steps, dist, dev, defer_tail=E2E_DEFER_TAIL, chunk_mib=0):
    """b200_frontend_process_host on this rank's shard: pinned host input, H2D, kernels, D2H of the dB rows and of every
    VFO's IF output.  With N ranks every rank feeds its own copy of the block over its own PCIe link (the host-fed path
    has no broadcast; SURVEY 8(e): "replicas per source")."""
    cfg = configs.CONFIGS[name]
    sr = cfg["sample_rate"]
    j = Job(binding, torch, name, Be, local, rank, world, defer_tail=defer_tail)
    fe2 = j.fe
    fe2.vfos.set_option(binding.OPT_HOST_CHUNK_MIB, chunk_mib)
    h_in = torch.empty((Be, 2), dtype=torch.float32).pin_memory()
    n_src = int(d_src.shape[0])
    for pos in range(0, Be, n_src):                      # a host block longer than the resident batch repeats it
        n = min(n_src, Be - pos)
        h_in[pos:pos + n].copy_(d_src[:n].cpu())
    mf = Be // int(round(sr / cfg["fft_rate"])) + 2
    h_frames = torch.empty((mf, cfg["fft_size"]), dtype=torch.float32).pin_memory()
    nslots = max(j.ids) + 1 if j.ids else 1
    sizes = [fe2.vfos.max_out(v, Be) for v in j.ids]
    h_iq = torch.empty((max(1, sum(sizes)), 2), dtype=torch.float32).pin_memory()
    optr = (ctypes.c_void_p * nslots)()
    pos = 0
    for v, n in zip(j.ids, sizes):
        optr[v] = h_iq.data_ptr() + pos * 8
        pos += n
    counts = np.zeros(nslots, np.int32)
    # the caller's stream at the default priority, i.e. level with the library's stream for the later stages and the copies
    # back: a chunk's results then cross PCIe beside the next chunk's full-rate kernels.  (On the timed loop's stream, which
    # sits above it, they waited for the whole call's full-rate kernels: 687 against 963 MS/s.)
    hs = torch.cuda.Stream(device=dev) if defer_tail else compute

    def step():
        rc = binding.lib.b200_frontend_process_host(fe2.h, h_in.data_ptr(), Be, h_frames.data_ptr(), mf, optr, None, counts.ctypes.data, hs.cuda_stream)
        if rc < 0:
            raise RuntimeError(binding.lib.b200_last_error().decode())

    for _ in range(2):
        step()
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    torch.cuda.synchronize(dev)
    te = time.perf_counter() - t0
    tt = torch.tensor([te], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    te = float(tt.item())
    d2h = fe2.frame_count() * cfg["fft_size"] * 4 + int(counts.sum()) * 8
    fe2.vfos.set_option(binding.OPT_HOST_CHUNK_MIB, 0)
    fe2.vfos.set_option(binding.OPT_DEFER_TAIL, 0)
    return {"value": Be * steps / te / 1e6, "unit": UNIT, "h2d_bytes_per_step": Be * 8, "d2h_bytes_per_step": int(d2h), "batch_samples": Be, "streaming_mode": bool(defer_tail),
            "path": "b200_frontend_process_host (pinned host buffers; H2D of the block, kernels, D2H of the dB rows and of every VFO's IF output, per call)"}


def sharded_self_check(binding, torch, dist, d_block, B, local, rank, world, compute, per_rank=3):
    """Every rank runs a fresh front end over its shard on the block it last received; rank 0 runs the same block through
    ONE front end that holds the picked VFOs of every rank and all FFT frames.  The sharded outputs must equal the
    single-GPU ones (bit for bit where the launch shape does not enter the order of summation, to -100 dB elsewhere); a
    block that was broadcast into a buffer still being read, or a frame computed by the wrong rank, fails this at once."""
    cfg = configs.CONFIGS[WORKLOAD]
    sr = cfg["sample_rate"]
    vl = configs.vfo_list(WORKLOAD)
    shards = sharding.shard_vfos(vl, sr, world)
    # the block of the last step is in different buffers on different ranks; make it the same bytes everywhere
    blk = d_block if rank == 0 else torch.empty_like(d_block)
    if rank == 0:
        blk = d_block.clone()
    dist.broadcast(blk, src=0)
    picks = [sorted(s)[:: max(1, len(s) // per_rank)][:per_rank] for s in shards]
    j = Job(binding, torch, WORKLOAD, B, local, rank, world, defer_tail=False)
    with torch.cuda.stream(compute):
        j.fe.process(blk, B, stream=compute)
    torch.cuda.synchronize()
    mine = {}
    for i in picks[rank]:
        v = j.ids[j.mine.index(i)]
        mine[i] = j.fe.vfos.output(v)
    rows = {int(j.fe.frame_index(k)): r for k, r in enumerate(j.fe.frames())}
    gathered = [None] * world
    dist.all_gather_object(gathered, {"vfo": mine, "rows": {f: r[:: max(1, len(r) // 4096)].copy() for f, r in rows.items()}})
    res = None
    if rank == 0:
        allp = [i for p in picks for i in p]
        ref = Job(binding, torch, WORKLOAD, B, local, 0, 1, defer_tail=False, vfo_indices=allp)
        with torch.cuda.stream(compute):
            ref.fe.process(blk, B, stream=compute)
        torch.cuda.synchronize()
        nv, nf, nv_exact, nf_exact = 0, 0, 0, 0
        worst_db, worst_row = -400.0, 0.0
        structure_ok = True
        for g in gathered:
            for i, y in g["vfo"].items():
                want = ref.fe.vfos.output(ref.ids[allp.index(i)])
                nv += 1
                if want.shape != y.shape or len(want) == 0:
                    structure_ok = False
                    continue
                nv_exact += int(np.array_equal(want.view(np.uint32), y.view(np.uint32)))
                num = float(np.sqrt(np.mean(np.abs(want.astype(np.complex128) - y) ** 2)))
                den = float(np.sqrt(np.mean(np.abs(want.astype(np.complex128)) ** 2)))
                worst_db = max(worst_db, 20.0 * np.log10(max(num, 1e-30) / max(den, 1e-30)))
        ref_rows = {int(ref.fe.frame_index(k)): r for k, r in enumerate(ref.fe.frames())}
        seen = set()
        for r, g in enumerate(gathered):
            for f, sub in g["rows"].items():
                nf += 1
                if (f % world != r) or f not in ref_rows:
                    structure_ok = False
                    continue
                want = ref_rows[f][:: max(1, len(ref_rows[f]) // 4096)]
                nf_exact += int(np.array_equal(want.view(np.uint32), sub.view(np.uint32)))
                worst_row = max(worst_row, float(np.max(np.abs(want - sub))))
                seen.add(f)
        structure_ok = structure_ok and seen == set(ref_rows)
        # The first stage and the FFT give the same bits in any launch; the later stages split long filters over lanes by
        # a balance taken over the whole launch, so a VFO's last bits can depend on how many VFOs share the launch.
        ok = structure_ok and worst_db <= -100.0 and worst_row <= 1e-3
        res = {"vfos_compared": nv, "vfos_bit_exact": nv_exact, "vfo_worst_error_db": worst_db, "frames_compared": nf, "frames_bit_exact": nf_exact,
               "rows_worst_abs_diff_db": worst_row, "every_frame_on_its_rank_once": bool(structure_ok), "ok": bool(ok),
               "tolerance": "VFO output error <= -100 dB of its RMS, dB rows within 0.001 dB, every frame f computed by rank f % N exactly once"}
        if not ok:
            raise RuntimeError("sharded run differs from the single-GPU run: %r" % res)
    return res


def _traffic():
    """dram bytes per launch of the dominant kernel from the committed ncu --set full capture, if any."""
    p = os.path.join(ROOT, "profiles", "fused_kernel_traffic.json")
    if os.path.exists(p):
        try:
            return json.load(open(p)).get("dram_bytes_per_launch")
        except Exception:
            return None
    return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=1 << 24, help="samples per step, device resident (default 16 Mi = 134 MB > L2)")
    ap.add_argument("--e2e-batch", type=int, default=1 << 25)
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--cpu-samples", type=int, default=1000000, help="samples per CPU-arm step (bounded sample of the workload: one reference stream block)")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--no-cpu-baseline", dest="cpu_baseline", action="store_false")
    ap.add_argument("--no-extras", dest="extras", action="store_false", help="skip the cfg2 / cfg3 extra keys")
    ap.add_argument("--no-defer-tail", dest="defer_tail", action="store_false",
                    help="run the VFOs' later stages inline in the caller's stream (library default) instead of the streaming mode")
    args = ap.parse_args()
    # The contract is ONE JSON line on stdout.  Libraries below us write there too (NCCL prints its version banner on
    # stdout when NCCL_DEBUG is set on the box): send file descriptor 1 to stderr for the duration of the run and give
    # Python's print() the real stdout back.
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(real_stdout, "w", buffering=1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)
    sys.stdout.flush()


if __name__ == "__main__":
    main()
