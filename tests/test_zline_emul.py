"""CPU check of the transpose-free last-axis pressure solve's CODE (csrc/zline.cuh): tools/zline_emul.cu compiles the
data routines of the kernels for the host, runs the passes for P emulated ranks with C sub-chunks each (exchange areas,
chain over the chunks, the constant mode's column) and compares with the reference's definition -- transform along the
last axis, divide by -(ax+ay+az), zero the constant mode, transform back (src/pressure.jl:323-344) -- evaluated as a
plain DFT in double precision.  Float64 data must agree to 1e-11, Float32 data to 5e-6 (storage rounding; measured 3e-8).
No GPU involved: nvcc only compiles."""
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
    exe = str(tmp_path_factory.mktemp("emul") / "zline_emul")
    subprocess.run([nvcc, "-std=c++17", "-O1", "-gencode", "arch=compute_100a,code=sm_100a", "-o", exe,
                    os.path.join(ROOT, "tools", "zline_emul.cu")], check=True)
    return exe


# nx ny n P C: one rank, ranks only, ranks x sub-chunks, sizes that are not powers of two, the 2-D case (ny = 1), a line
# long enough that the lowest modes couple every chunk with every other one, chunks so long (and modes so stiff) that
# rho^j underflows inside a chunk, chunks of two planes and of one
@pytest.mark.parametrize("args", [(16, 8, 32, 1, 1), (16, 8, 32, 4, 1), (16, 8, 64, 4, 2), (12, 6, 48, 2, 3), (20, 1, 60, 3, 2),
                                  (32, 1, 64, 2, 4), (64, 4, 256, 8, 2), (16, 4, 1024, 2, 1), (256, 1, 64, 2, 1), (16, 8, 16, 4, 2), (8, 4, 8, 8, 1)], ids=str)
def test_line_solve_code_on_host(emul, args):
    res = subprocess.run([emul] + [str(a) for a in args], capture_output=True, text=True)
    assert res.returncode == 0 and "line solve agrees" in res.stdout, res.stdout + res.stderr
