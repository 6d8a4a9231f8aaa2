"""The one-process-per-GPU model (nbody_create_rank under torchrun, CUDA-IPC peer mappings) on 2 and 4 GPUs: bit-identical to
one GPU with ranks deliberately out of step around nbody_upload, sharded uploads, own-range downloads, complete contact lists.
The driver's one-GPU lease skips it; scripts/gpu_suite.sh runs it on a multi-GPU lease (profiles/)."""
import os
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world", [2, 4])
def test_rank_mode_equals_one_gpu(nb, world):
    if nb.device_count() < world:
        pytest.skip(f"needs >= {world} GPUs")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
                        "--master-port", str(_free_port()), os.path.join(ROOT, "tests", "rank_mode_worker.py")],
                       capture_output=True, text=True, cwd=ROOT, timeout=900)
    tail = r.stdout[-3000:] + r.stderr[-1500:]
    assert r.returncode == 0 and "RANK-MODE OK" in r.stdout, tail
