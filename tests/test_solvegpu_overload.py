"""include/mgcg.hpp: the C++ overload SolveGpu(thrust::device_vector<double>, ...) of the reference's SimpleConjugateGradient program
(SimpleConjugateGradient.cu:12-22, called at :231), through tests/harness/simple_cg_main.cu — a caller shaped like the reference's main.
CPU: the header and the caller compile with nvcc for sm_100a.  GPU: built, linked against libMgcgGpu.so and run — N = 10 against
the known answer of SURVEY.md 8c case (1) (tests/golden/tridiag10.json), and bit-identical to SolveGpu_c on the same device arrays."""
import json
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "harness", "simple_cg_main.cu")
NVCC = ["nvcc", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-I" + os.path.join(ROOT, "include")]


def test_header_and_caller_compile(tmp_path):
    subprocess.run(NVCC + ["-c", SRC, "-o", str(tmp_path / "simple_cg_main.o")], check=True, timeout=300)


@pytest.mark.gpu
def test_overload_runs_like_the_reference_program(tmp_path):
    from conjugategradient_b200 import _lib
    libdir = os.path.dirname(_lib.LIB_PATH)
    exe = str(tmp_path / "simple_cg_main")
    subprocess.run(NVCC + [SRC, "-o", exe, "-L" + libdir, "-lMgcgGpu", "-Xlinker", "-rpath=" + libdir], check=True, timeout=300)
    out = subprocess.run([exe, "10"], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stdout + out.stderr
    head, xs = out.stdout.strip().splitlines()
    assert "iteration=10 " in head and "same_as_SolveGpu_c=1" in head          # 10 iterations done == iteration index 9
    golden = json.load(open(os.path.join(ROOT, "tests", "golden", "tridiag10.json")))
    assert golden["iteration_index"] == 9
    x = [float(t) for t in xs[2:].split()]
    assert len(x) == 10 and all(abs(a - b) <= 5e-7 * max(1.0, abs(b)) for a, b in zip(x, golden["x"]))
    out = subprocess.run([exe, "4096"], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0 and "same_as_SolveGpu_c=1" in out.stdout, out.stdout + out.stderr
