"""The C++ caller-side harness (tests/harness/mgcg_main.cpp): the reference's MgcgMain flow — CPU, single-GPU and
host-threaded multi-device solvers through the unchanged MgcgGpu.dll export names — compiled with g++ and run here."""
import os
import subprocess

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def harness(tmp_path_factory):
    from conjugategradient_b200 import _lib
    from oracle import oracle
    oracle.lib()
    exe = str(tmp_path_factory.mktemp("harness") / "mgcg_main")
    subprocess.run(["g++", "-O2", "-std=c++17", "-Wall", "-o", exe, os.path.join(ROOT, "tests", "harness", "mgcg_main.cpp"), "-ldl", "-lpthread"],
                   check=True)
    return exe, _lib.LIB_PATH, os.path.join(ROOT, "oracle", "liboracle.so")


@pytest.mark.parametrize("count,maxnz,min_it,devices", [(2000, 160, 0, 0), (2000, 160, 200, 3), (5001, 40, 0, 4), (20000, 160, 0, 2)])
def test_mgcg_main_flow(harness, count, maxnz, min_it, devices):
    exe, lib, orc = harness
    out = subprocess.run([exe, lib, orc, str(count), str(maxnz), str(min_it), str(devices)], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and "HARNESS_OK" in out.stdout, out.stdout[-3000:] + out.stderr[-2000:]
    if (count, maxnz, min_it) == (2000, 160, 0):
        assert "Iteration=26" in out.stdout          # SURVEY.md §8c known answer (2)
