#!/usr/bin/env python3
"""Benchmark of the IQ-front-end hot path (BASELINE.json: complex MS/s through FFT + N-VFO channelizer).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path (C ABI of libb200dsp.so)
    python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU blocks on the host cores

Workload at every N (per GPU; weak scaling): BASELINE config[1] -- 20 MS/s IQ, 65536-point Blackman FFT
waterfall at 60 fps (window + FFT + log power + max-bin zoom + palette) and one USB RxVFO channelizer (20 M -> 24 k:
xlate, /512 decimator, 8000/13021 polyphase resampler, 651-tap channel filter).  The metric is "through FFT +
N-VFO channelizer", so the timed step ends at the VFO's IF output; the radio module's SSB demodulator + AGC that
the VFO feeds is timed in the same run as the extra key `with_demod` (its AGC is a strictly sequential
recurrence per VFO, a latency chain that says nothing about the channelizer).
A step is one pass of that path over one batch of synthetic cf32 IQ that is resident in HBM when the timed
region starts (batch > L2).  With N > 1 ranks rank 0 owns the stream: every step's batch is NCCL-broadcast
to the other ranks (double buffered, overlapped with the previous step's kernels), each rank channelizes it
for its own VFO shard and computes the FFT frames f with f % N == rank.  `value` counts the samples every
rank pushed through its channelizer shard.

One JSON line on stdout (rank 0).  See the task contract for the keys.
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from sdrplusplusbrown_b200 import configs  # noqa: E402

METRIC = "complex MS/s through FFT + N-VFO channelizer"
UNIT = "MS/s"
WORKLOAD = "cfg2"
DATA_WIDTH = 1333                    # waterfall pixels per row (a typical dataWidth)
WF_MIN, WF_MAX = -120.0, -10.0


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json copy bandwidth)"
    return 6650.0, "fallback (B200_PROFILING.md)"


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


class CpuPath:
    """The reference's blocks for the workload, run the way the reference runs them: one thread per block
    (dsp/block.h:82-84) -- the VFO(+demod) chain on one thread, the FFT branch (window, FFT, log power, zoom,
    palette) on another; the Splitter's memcpy fan-out is replaced by sharing the block."""

    def __init__(self, cfg_name=WORKLOAD, demod=False):
        self.ol, self.orc = _oracle()
        cfg = configs.CONFIGS[cfg_name]
        self.sr = cfg["sample_rate"]
        self.N = cfg["fft_size"]
        self.skip, self.nz = self.orc.reshape_params(self.sr, self.N, cfg["fft_rate"])
        self.interval = self.nz + self.skip
        self.window = self.orc.fft_window(cfg["window"], self.nz)
        self.vfos = []
        for mode, out_sr, bw, off in configs.vfo_list(cfg_name):
            v = self.ol.Vfo(self.orc, self.sr, out_sr, bw, off)
            kind = {"USB": 0, "LSB": 1}.get(mode)
            d = self.ol.Ssb(self.orc, kind, bw, out_sr, 50.0 / out_sr, 5.0 / out_sr) if (kind is not None and demod) else None
            self.vfos.append((v, d))
        self.pos = 0
        self.kind = "reference" if self.orc.kind == "reference" else "port"

    def step(self, x):
        """x: complex64 numpy block.  Returns after both branches finished (the slower thread bounds the rate)."""
        def vfo_branch():
            for v, d in self.vfos:
                for i in range(0, len(x), 1000000):          # the reference's streams carry <= 1 M samples (dsp/stream.h:14)
                    y = v.process(x[i:i + 1000000])
                    if d is not None and len(y):
                        d.process(y)

        def fft_branch():
            f = -(-self.pos // self.interval)
            while f * self.interval + self.nz <= self.pos + len(x):
                s = f * self.interval - self.pos
                db = self.orc.fft_frame(x[s:s + self.nz], self.window, self.N)
                z = self.orc.do_zoom(0, self.N, db, DATA_WIDTH)
                self.orc.palette_indices(z, WF_MIN, WF_MAX)
                f += 1

        t1 = threading.Thread(target=vfo_branch)
        t2 = threading.Thread(target=fft_branch)
        t1.start(); t2.start(); t1.join(); t2.join()
        self.pos += len(x)

    @property
    def threads(self):
        return 2


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from sdrplusplusbrown_b200 import synth
    cfg = configs.CONFIGS[WORKLOAD]
    sample = args.cpu_samples
    x = synth.make_iq(0, sample, cfg["sample_rate"], tones=[(1.25e6 + 700.0, 0.05), (3.3e6, 0.2)])
    cpu = CpuPath(demod=args.demod)
    for _ in range(args.warmup):
        cpu.step(x)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu.step(x)
    dt = time.perf_counter() - t0
    value = sample * args.steps / dt / 1e6
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": _config(args, sample, 1),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cpu.threads, "kind": cpu.kind,
                         "sample": "%d samples per step (%.2f s of signal), reference blocks on generic-C VOLK / float64 FFT stand-ins" % (sample, sample / cfg["sample_rate"])},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def _config(args, batch, n_gpus):
    cfg = configs.CONFIGS[WORKLOAD]
    return {
        "workload": "BASELINE config[1]: 20 MS/s synthetic IQ, 65536-pt Blackman FFT waterfall at 60 fps (+zoom to %d px, palette), "
                    "1 USB SSB RxVFO channelizer (20M->24k), per GPU" % DATA_WIDTH,
        "sample_rate": cfg["sample_rate"], "fft_size": cfg["fft_size"], "fft_rate": cfg["fft_rate"], "window": "blackman",
        "vfos_per_gpu": 1, "vfo": "USB out 24 kS/s bw 2.8 kHz", "demod": bool(getattr(args, "demod", False)),
        "batch_samples_per_step": batch, "l2_policy": ("batch (%.0f MB) larger than L2 (126 MB)" % (batch * 8 / 1e6)) if args.impl == "b200" else "n/a (CPU arm)",
        "rotator": "fast (closed form, calibrated step)",
        "stage_overlap": ("B200_OPT_DEFER_TAIL: the VFO's later stages of step k run on their own stream beside the full-rate stage of step k+1; "
                          "all K steps' outputs are complete when the timed region ends (device-wide synchronize)") if (args.impl == "b200" and args.defer_tail) else "none",
        "parallelism": "vfo-shard x%d + stream broadcast, fft frames time-sharded" % n_gpus if n_gpus > 1 else "single GPU",
    }


# ------------------------------------------------------------------------------------------------ GPU arm
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

    # ---- handles: front end (FFT branch + VFO bank + demod) and waterfall reduction
    fe = binding.FrontEnd(sr, fft_size=cfg["fft_size"], fft_rate=cfg["fft_rate"], fft_window=cfg["window"], max_block=B, device=local)
    if world > 1:
        fe.set_frame_shard(rank, world)
    mode, out_sr, bw, off = configs.vfo_list(WORKLOAD)[0]
    # every rank's shard holds one USB VFO; shards sit 50 kHz apart so no two ranks do identical work
    v = fe.vfos.add(out_sr, bw, off + 50e3 * rank)
    if args.demod:
        fe.vfos.set_demod(v, binding.DEMOD_USB, 0, 50.0 / out_sr, 5.0 / out_sr, 0)
    fe.vfos.set_option(binding.OPT_DEFER_TAIL, 1 if args.defer_tail else 0)
    wf = binding.Waterfall(cfg["fft_size"], DATA_WIDTH, device=local)
    wf.set_view(0, cfg["fft_size"])
    wf.set_range(WF_MIN, WF_MAX)
    max_frames = B // int(round(sr / cfg["fft_rate"])) + 2
    d_zoom = torch.empty((max_frames, DATA_WIDTH), dtype=torch.float32, device=dev)
    d_rgba = torch.empty((max_frames, DATA_WIDTH), dtype=torch.int32, device=dev)
    fe.attach_waterfall(wf, d_zoom.data_ptr(), d_rgba.data_ptr(), max_frames)

    # ---- synthetic stream, resident in HBM (rank 0 is the ingest GPU)
    tones = [(off + 700.0 + 50e3 * r, 0.05) for r in range(world)] + [(3.3e6, 0.2)]
    if rank == 0:
        src = synth.make_iq_torch(B, sr, dev, tones=tones)
        bufs = [src, src]          # the ingest GPU channelizes the stream in place and broadcasts from it
    else:
        src = None
        bufs = [torch.empty((B, 2), dtype=torch.float32, device=dev) for _ in range(2)]
    compute = torch.cuda.Stream(device=dev)
    comm = torch.cuda.Stream(device=dev) if world > 1 else None
    ev_ready = [torch.cuda.Event() for _ in bufs]
    ev_free = [torch.cuda.Event() for _ in bufs]

    def bcast(i):
        """rank 0 -> all: batch for step i into buffer i % 2 (NCCL over NVLink), on the comm stream."""
        b = bufs[i % 2]
        with torch.cuda.stream(comm):
            comm.wait_event(ev_free[i % 2])
            dist.broadcast(b, src=0)
            ev_ready[i % 2].record(comm)

    def step(i):
        b = bufs[i % len(bufs)]
        if world > 1:
            compute.wait_event(ev_ready[i % 2])
        with torch.cuda.stream(compute):
            fe.process(b, B, stream=compute)      # FFT rows -> zoom + palette rows inside the FFT branch (attached waterfall)
            if world > 1:
                ev_free[i % 2].record(compute)

    def run(nsteps):
        if world > 1:
            for e in ev_free:
                e.record(compute)
            bcast(0)
        for i in range(nsteps):
            if world > 1 and i + 1 < nsteps:
                bcast(i + 1)
            step(i)

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    clocks = ClockSampler(local) if rank == 0 else None      # spans warm-up, the timed region and a short tail of identical steps
    run(max(args.warmup, 3))
    barrier()
    binding.profile_enable(True)
    binding.profile_read()
    launches0 = binding.lib.b200_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record(compute)
    run(args.steps)
    fe.vfos.out_ptr(v)            # orders the last steps' deferred stages into the compute stream: e1 fires behind ALL work of the K steps
    e1.record(compute)
    barrier()
    ms = e0.elapsed_time(e1)
    launches = binding.lib.b200_launch_count() - launches0
    kms, klaunches, kbytes = binding.profile_read()
    binding.profile_enable(False)
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    value = world * B * args.steps / (ms * 1e-3) / 1e6
    # the timed region can be shorter than nvidia-smi's sampling period: keep the identical load running for ~0.6 s
    # (same step count on every rank) so the clock record holds samples taken under this workload
    run(min(5000, int(600.0 / max(ms / args.steps, 1e-3)) + 1))
    barrier()
    clk = clocks.stop() if clocks else None

    # ---- same step with the VFO's SSB demodulator + AGC attached (reported beside the headline)
    with_demod = None
    if not args.demod:
        fe.vfos.set_demod(v, binding.DEMOD_USB, 0, 50.0 / out_sr, 5.0 / out_sr, 0)
        run(3)
        barrier()
        d0, d1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        d0.record(compute)
        run(max(3, args.steps // 2))
        fe.vfos.out_ptr(v)
        d1.record(compute)
        barrier()
        td = torch.tensor([d0.elapsed_time(d1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(td, op=dist.ReduceOp.MAX)
        nd = max(3, args.steps // 2)
        with_demod = {"value": world * B * nd / (float(td.item()) * 1e-3) / 1e6, "unit": UNIT, "ms_per_step": float(td.item()) / nd,
                      "note": "step + USB SSB demodulator (xlate, real part, AGC walk) on the VFO output"}
        fe.vfos.set_demod(v, binding.DEMOD_NONE)

    # ---- the dominant kernel by itself: the same VFO bank on the same resident block, run without the FFT branch and with
    # its later stages inline, so that nothing shares the SMs with the full-rate kernel (reported beside the in-step figure)
    barrier()
    fe.vfos.set_option(binding.OPT_DEFER_TAIL, 0)
    b0 = bufs[0]
    with torch.cuda.stream(compute):
        for _ in range(3):
            fe.vfos.process(b0, B, stream=compute)
    barrier()
    binding.profile_enable(True)
    binding.profile_read()
    with torch.cuda.stream(compute):
        for _ in range(20):
            fe.vfos.process(b0, B, stream=compute)
    barrier()
    kms_alone, kl_alone, kb_alone = binding.profile_read()
    binding.profile_enable(False)
    fe.vfos.set_option(binding.OPT_DEFER_TAIL, 1 if args.defer_tail else 0)

    # ---- end to end through the host-buffer C-ABI call (H2D of the block + D2H of frames, VFO IQ and audio inside)
    Be = args.e2e_batch
    fe2 = binding.FrontEnd(sr, fft_size=cfg["fft_size"], fft_rate=cfg["fft_rate"], fft_window=cfg["window"], max_block=Be, device=local)
    if world > 1:
        fe2.set_frame_shard(rank, world)
    v2 = fe2.vfos.add(out_sr, bw, off + 50e3 * rank)
    fe2.vfos.set_demod(v2, binding.DEMOD_USB, 0, 50.0 / out_sr, 5.0 / out_sr, 0)   # the drop-in path returns audio too
    h_in = torch.empty((Be, 2), dtype=torch.float32).pin_memory()
    h_in.copy_((src if src is not None else bufs[0])[:Be].cpu())
    mf = Be // int(round(sr / cfg["fft_rate"])) + 2
    h_frames = torch.empty((mf, cfg["fft_size"]), dtype=torch.float32).pin_memory()
    mo = fe2.vfos.max_out(v2, Be)
    h_iq = torch.empty((mo, 2), dtype=torch.float32).pin_memory()
    h_au = torch.empty((mo, 2), dtype=torch.float32).pin_memory()
    optr = (ctypes.c_void_p * (v2 + 1))(); optr[v2] = h_iq.data_ptr()
    aptr = (ctypes.c_void_p * (v2 + 1))(); aptr[v2] = h_au.data_ptr()
    counts = np.zeros(v2 + 1, np.int32)

    def e2e_step():
        rc = binding.lib.b200_frontend_process_host(fe2.h, h_in.data_ptr(), Be, h_frames.data_ptr(), mf, optr, aptr, counts.ctypes.data, compute.cuda_stream)
        if rc < 0:
            raise RuntimeError(binding.lib.b200_last_error().decode())

    for _ in range(2):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.e2e_steps):
        e2e_step()
    torch.cuda.synchronize(dev)
    te = time.perf_counter() - t0
    tt = torch.tensor([te], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    te = float(tt.item())
    e2e_value = world * Be * args.e2e_steps / te / 1e6
    nf2 = fe2.frame_count()
    d2h = nf2 * cfg["fft_size"] * 4 + int(counts[v2]) * 8 * 2

    # ---- the same call fed with int16 I/Q (what the reference's file source and most SDR hardware deliver; converted on
    # the device): reported beside e2e, not instead of it
    h_i16 = (h_in * 32767.0).to(torch.int16).pin_memory()

    def e2e_i16_step():
        rc = binding.lib.b200_frontend_process_host_i16(fe2.h, h_i16.data_ptr(), Be, ctypes.c_float(32768.0), h_frames.data_ptr(), mf, optr, aptr,
                                                        counts.ctypes.data, compute.cuda_stream)
        if rc < 0:
            raise RuntimeError(binding.lib.b200_last_error().decode())

    for _ in range(2):
        e2e_i16_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.e2e_steps):
        e2e_i16_step()
    torch.cuda.synchronize(dev)
    ti = time.perf_counter() - t0
    tti = torch.tensor([ti], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tti, op=dist.ReduceOp.MAX)
    e2e_i16_value = world * Be * args.e2e_steps / float(tti.item()) / 1e6

    if rank == 0:
        peak, which = peaks()
        achieved = (kbytes / 1e9) / (kms * 1e-3) if kms > 0 else 0.0
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": _config(args, B, world),
            "clocks": clk,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": Be * 8, "d2h_bytes_per_step": d2h,
                    "batch_samples": Be, "path": "b200_frontend_process_host (pinned host buffers, H2D + kernels + D2H per call)",
                    "int16_input": {"value": e2e_i16_value, "unit": UNIT, "h2d_bytes_per_step": Be * 4,
                                    "path": "b200_frontend_process_host_i16 (int16 I/Q over PCIe, converted on the device)"}},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "kernel": "fusedStreamKernel<512> (xlate + /32 decimating FIR, stage 1 of the VFO; single-VFO variant of the fused kernel with 16-byte staging)",
                         "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak if peak else None,
                         "peak_source": which, "traffic": _traffic(), "launches": int(klaunches), "avg_launch_ms": kms / klaunches if klaunches else None,
                         "algorithmic_bytes_per_launch": kbytes / klaunches if klaunches else None,
                         "kernel_alone": {"achieved": (kb_alone / 1e9) / (kms_alone * 1e-3) if kms_alone > 0 else None,
                                          "frac": ((kb_alone / 1e9) / (kms_alone * 1e-3)) / peak if (kms_alone > 0 and peak) else None,
                                          "avg_launch_ms": kms_alone / kl_alone if kl_alone else None, "launches": int(kl_alone),
                                          "note": "same kernel, same resident block, measured in this run with CUDA events: the VFO bank alone (no FFT branch, later stages inline), so nothing shares the SMs with it; `achieved` / `frac` above are its figures inside the timed step, where the FFT passes and the previous block's tail run beside it"},
                         "share_of_step": (kms / ms) if ms else None},
        }
        if with_demod is not None:
            line["with_demod"] = with_demod
        if args.cpu_baseline and world == 1:
            line["cpu_baseline"] = cpu_baseline(args)
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def _traffic():
    """dram bytes per launch of the dominant kernel from the committed ncu --set full capture, if any."""
    p = os.path.join(ROOT, "profiles", "fused_kernel_traffic.json")
    if os.path.exists(p):
        try:
            return json.load(open(p)).get("dram_bytes_per_launch")
        except Exception:
            return None
    return None


def cpu_baseline(args):
    from sdrplusplusbrown_b200 import synth
    cfg = configs.CONFIGS[WORKLOAD]
    n = args.cpu_samples
    x = synth.make_iq(0, n, cfg["sample_rate"], tones=[(1.25e6 + 700.0, 0.05), (3.3e6, 0.2)])
    cpu = CpuPath(demod=args.demod)
    cpu.step(x[:1000000])
    t0 = time.perf_counter()
    reps = 0
    while True:
        cpu.step(x)
        reps += 1
        if time.perf_counter() - t0 > args.cpu_seconds:
            break
    dt = time.perf_counter() - t0
    return {"value": n * reps / dt / 1e6, "unit": UNIT, "cores": cpu.threads, "kind": cpu.kind,
            "sample": "%d x %d samples (%.1f s of CPU time) of the same workload; reference blocks compiled from its own headers against "
                      "generic-C VOLK and a float64 radix-2 FFT stand-in (real VOLK SIMD / FFTW are not installable here, so this under-states the reference)" % (reps, n, dt)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=1 << 26, help="samples per step, device resident (default 64 Mi = 512 MiB > L2)")
    ap.add_argument("--e2e-batch", type=int, default=1 << 24)
    ap.add_argument("--e2e-steps", type=int, default=5)
    ap.add_argument("--cpu-samples", type=int, default=1 << 24, help="samples per CPU-arm step (bounded sample of the workload)")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--no-cpu-baseline", dest="cpu_baseline", action="store_false")
    ap.add_argument("--no-defer-tail", dest="defer_tail", action="store_false",
                    help="run the VFO's later stages inline in the caller's stream (library default) instead of the streaming mode")
    ap.add_argument("--demod", action="store_true", help="include the SSB demodulator + AGC in the timed step")
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
