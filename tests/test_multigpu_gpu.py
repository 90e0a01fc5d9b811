"""Two ranks on two GPUs (skipped on a one-GPU box): the sharded job of bench.py -- VFO bank dealt out over the ranks,
every block broadcast from rank 0 over NCCL, FFT frames sharded by time -- must give, VFO by VFO and dB row by dB row,
what a single GPU gives for the same block (bench.py: sharded_self_check; bit for bit where the launch shape does not
enter the order of summation, -100 dB elsewhere).  Reference fan-out this replaces:
Splitter::run (core/src/dsp/routing/splitter.h:74-83)."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _gpus():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


@pytest.mark.skipif(_gpus() < 2, reason="needs two GPUs")
def test_two_rank_sharded_run_equals_single_gpu_bit_for_bit():
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1", "--master-port", "29617",
           os.path.join(ROOT, "bench.py"), "--gpus", "2", "--steps", "3", "--warmup", "3", "--no-extras", "--no-cpu-baseline", "--batch", str(1 << 23),
           "--e2e-batch", str(1 << 22), "--e2e-steps", "1"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    line = json.loads([x for x in r.stdout.splitlines() if x.startswith("{")][-1])
    assert line["n_gpus"] == 2 and line["scaling"] == "strong"
    sc = line["self_check"]
    assert sc["ok"] is True and sc["vfos_compared"] >= 4 and sc["frames_compared"] >= 2
    assert sc["vfo_worst_error_db"] <= -100.0 and sc["rows_worst_abs_diff_db"] <= 1e-3 and sc["every_frame_on_its_rank_once"]
    assert line["run"]["broadcast_ms_per_step_rank0"] > 0
