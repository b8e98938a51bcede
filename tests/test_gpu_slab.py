"""Multi-GPU z-slab parity: N ranks, one slab each, NCCL halo planes + IPC advection
gathers, against the single-domain CPU oracle, bit-exact.  Needs >= 2 GPUs (gpurun --gpus 2); skipped on
the 1-GPU round-end box.  The worker is tests/slab_worker.py."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpu():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


@pytest.mark.parametrize("n", [2, 4, 8])
def test_slab_parity(n):
    if _ngpu() < n:
        pytest.skip(f"needs {n} GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
           "--master-addr", "127.0.0.1", "--master-port", str(29610 + n), os.path.join(ROOT, "tests", "slab_worker.py")]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT)
    out_dir = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(out_dir):                              # the whole SLAB-PARITY log travels back from the GPU box
        tag = os.environ.get("FLUID_SLAB_LOG_TAG", "")
        with open(os.path.join(out_dir, f"slab_parity_x{n}{tag}.log"), "w") as f:
            f.write(res.stdout + "\n--- stderr ---\n" + res.stderr[-6000:])
    fails = [l for l in res.stdout.splitlines() if "SLAB-PARITY FAIL" in l]
    tail = "\n".join(fails[:40]) + "\n...\n" + (res.stdout + res.stderr)[-3000:]
    assert res.returncode == 0, tail
    assert "SLAB-PARITY OK" in res.stdout and not fails, tail
