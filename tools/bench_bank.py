#!/usr/bin/env python3
"""VFO-bank throughput sweep (BASELINE configs[2]/[4]: N NFM RxVFOs at 61.44 MS/s, and the mixed cfg4 bank).

    python tools/bench_bank.py [--cfg cfg3] [--counts 1,2,4,...] [--batch 16777216] [--steps 10]

Per VFO count: whole-bank input rate (complex MS/s), the fused xlate+decimate kernel's time, its algorithmic
HBM rate and its FP32 instruction rate against the B200 CUDA-core peak.  One JSON line per count on stdout.
Not the contract benchmark (bench.py is); this is the evidence for DESIGN.md's many-VFO numbers.
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from sdrplusplusbrown_b200 import configs  # noqa: E402

# FP32 instructions per input sample per VFO in the fused kernel: rotation (4) + 2 per tap/decimation
STAGE1 = {16: (8, 54), 32: (8, 44), 64: (8, 36), 128: (16, 72), 256: (32, 143), 512: (32, 129), 1024: (64, 257),
          2048: (64, 329), 4096: (64, 400), 8192: (128, 726)}


def main():
    import torch
    from sdrplusplusbrown_b200 import binding, synth
    ap = argparse.ArgumentParser()
    ap.add_argument("--cfg", default="cfg3")
    ap.add_argument("--counts", default="1,4,16,32,64,256")
    ap.add_argument("--batch", type=int, default=1 << 24)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--demod", action="store_true")
    ap.add_argument("--tile", type=int, default=0, help="fused-kernel tile samples (0 = library default)")
    ap.add_argument("--vb", type=int, default=0, help="VFOs per register pass of the grouped first-stage kernel (0 = library default, 2, 4)")
    ap.add_argument("--gv", type=int, default=32, help="VFOs per CTA of the grouped first-stage kernel")
    ap.add_argument("--ptab", type=int, default=1, help="rotation tables in the kernel parameter block (1) or in shared memory (0)")
    ap.add_argument("--split", type=int, default=1, help="split plans whose loop body exceeds the instruction cache into two launches (library default 1)")
    args = ap.parse_args()
    cfg = configs.CONFIGS[args.cfg]
    sr = cfg["sample_rate"]
    dev = torch.device("cuda", 0)
    x = synth.make_iq_torch(args.batch, sr, dev, tones=[(1e6, 0.1)])
    peak_fp32 = 148 * 128 * 1.965e9          # FP32 lanes x max SM clock, instructions/s (FFMA counts as one)
    hbm = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
    for n in [int(c) for c in args.counts.split(",")]:
        vfos = configs.vfo_list(args.cfg, n)
        bank = binding.VfoBank(sr, max_block=args.batch)
        bank.set_option(binding.OPT_TILE_SAMPLES, args.tile)
        bank.set_option(binding.OPT_VFOS_PER_PASS, args.vb)
        bank.set_option(binding.OPT_VFOS_PER_CTA, args.gv)
        bank.set_option(binding.OPT_SPLIT_RUNS, args.split)
        bank.set_option(binding.OPT_PARAM_TABLES, args.ptab)
        instr = 0.0
        for mode, out_sr, bw, off in vfos:
            v = bank.add(out_sr, bw, off)
            d = bank.describe(v)
            if d["fused"]:
                D, T = STAGE1[d["ratio"]]
                instr += 4.0 + 2.0 * T / D
            if args.demod:
                kind = {"NFM48": binding.DEMOD_NFM, "NFM": binding.DEMOD_NFM, "AM": binding.DEMOD_AM, "USB": binding.DEMOD_USB, "LSB": binding.DEMOD_LSB,
                        "WFM": binding.DEMOD_WFM}[mode]
                if kind == binding.DEMOD_WFM:
                    bank.set_demod(v, kind, 0, 1, 0, 0)          # the radio's defaults: mono, low-pass on, no RDS
                elif kind == binding.DEMOD_NFM:
                    bank.set_demod(v, kind, 1, 0)
                elif kind == binding.DEMOD_AM:
                    bank.set_demod(v, kind, 1, 50.0 / out_sr, 5.0 / out_sr, 100.0 / out_sr)
                else:
                    bank.set_demod(v, kind, 0, 50.0 / out_sr, 5.0 / out_sr)
        s = torch.cuda.Stream()
        with torch.cuda.stream(s):
            for _ in range(3):
                bank.process(x, args.batch, stream=s)
        torch.cuda.synchronize()
        binding.profile_enable(True)
        binding.profile_read()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(s):
            e0.record(s)
            for _ in range(args.steps):
                bank.process(x, args.batch, stream=s)
            e1.record(s)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / args.steps
        kms, kl, kb = binding.profile_read()
        binding.profile_enable(False)
        kms_per = kms / max(kl, 1) * (kl / args.steps)     # fused-kernel ms per step (all plan groups)
        rate = args.batch / (ms * 1e-3) / 1e6
        line = {
            "cfg": args.cfg, "vfos": n, "batch": args.batch, "ms_per_step": ms, "input_MSps": rate, "vfo_MSps": rate * n,
            "fused_ms_per_step": kms_per, "fused_share": kms_per / ms,
            "fused_hbm_GBps": (kb / args.steps) / 1e9 / (kms_per * 1e-3) if kms_per > 0 else None,
            "fused_hbm_frac": ((kb / args.steps) / 1e9 / (kms_per * 1e-3)) / hbm if kms_per > 0 else None,
            "fused_fp32_instr_per_s": instr * args.batch / (kms_per * 1e-3) if kms_per > 0 else None,
            "fused_fp32_frac": (instr * args.batch / (kms_per * 1e-3)) / peak_fp32 if kms_per > 0 else None,
            "demod": args.demod, "vfos_per_pass": args.vb, "vfos_per_cta": args.gv, "split_runs": args.split,
        }
        print(json.dumps(line), flush=True)
        bank.close()
        del bank
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
