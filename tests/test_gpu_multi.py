"""Multi-GPU parity, collected by pytest: launches tests/mgpu_parity.py under torch.distributed.run on 2 and 4 GPUs of
the box (skipped when fewer are visible).  Every rank owns one slab of the last axis; halos, the decomposed pressure solve
and the all-reduced reductions are compared with the single-domain oracle on identical initial arrays -- small grids
(2-D, TMA path, generic path, closure) and 256 x 256 x 512 Float32 + Smagorinsky against the tiled 64^3 oracle run.
Variants: the transpose-free last-axis solve with flag barriers (default; also with forced sub-chunks and with
ncclAllReduce barriers), the distributed FFT in its three transpose forms (packed / pitched peer copies, scatter kernels),
NCCL send/recv transposes, the bit-exact kernels."""
import os
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("world", [2, 4])
@pytest.mark.parametrize("variant", ["default", "line_chunks", "fft_packed", "fft_pitched_copies", "fft_scatter_kernels", "nccl_barrier", "nccl_sendrecv", "exact"])
def test_slab_path_matches_oracle(world, variant):
    import torch

    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs, {torch.cuda.device_count()} visible")
    env = dict(os.environ, **{"default": {}, "line_chunks": {"MGPU_ZLCHUNKS": "2"}, "fft_packed": {"MGPU_ZLINE": "0", "MGPU_TMODE": "2"},
                             "fft_pitched_copies": {"MGPU_ZLINE": "0", "MGPU_TMODE": "1"}, "fft_scatter_kernels": {"MGPU_ZLINE": "0", "MGPU_TMODE": "0"},
                             "nccl_barrier": {"MGPU_FASTBAR": "0"}, "nccl_sendrecv": {"MGPU_P2P": "0"}, "exact": {"MGPU_EXACT": "1"}}[variant])
    if variant != "default":
        env["MGPU_BIG"] = "0" if variant == "nccl_sendrecv" else "1"
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(_free_port()), os.path.join(ROOT, "tests", "mgpu_parity.py")]
    res = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=900, cwd=ROOT)
    tail = "\n".join((res.stdout + res.stderr).splitlines()[-40:])
    assert res.returncode == 0 and "FAIL" not in res.stdout, tail
    assert "[mgpu x%d]" % world in res.stdout, tail
