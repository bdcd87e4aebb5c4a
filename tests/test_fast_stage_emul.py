"""CPU check of the fused stage kernel's CODE: tools/fast_stage_emul.cu compiles csrc/fast_stage.cuh for the host, steps
the per-thread objects of every block between the block barriers (TMA boxes copied by hand with the kernel's own box
coordinates) and compares k_i, u' and the Poisson right-hand side with a plain periodic triple loop in double precision
-- all six kernel variants (first / middle / general stage, with and without closure), several grids and z-chunkings
(chunk boundaries, a last chunk that is shorter, the minimum chunk of 2 planes).  No GPU involved: nvcc only compiles."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def emul(tmp_path_factory):
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    if not os.path.exists(nvcc):
        nvcc = shutil.which("nvcc")
    if not nvcc:
        pytest.skip("nvcc not available")
    exe = str(tmp_path_factory.mktemp("emul") / "fast_stage_emul")
    subprocess.run([nvcc, "-std=c++17", "-O1", "-gencode", "arch=compute_100a,code=sm_100a", "-o", exe,
                    os.path.join(ROOT, "tools", "fast_stage_emul.cu")], check=True)
    return exe


@pytest.mark.parametrize("args", [(64, 32, 12, 5), (32, 16, 6, 2), (64, 48, 9, 4), (32, 16, 8, 8)], ids=str)
def test_fast_stage_code_on_host(emul, args):
    res = subprocess.run([emul] + [str(a) for a in args], capture_output=True, text=True)
    assert res.returncode == 0 and "all variants agree" in res.stdout, res.stdout + res.stderr
