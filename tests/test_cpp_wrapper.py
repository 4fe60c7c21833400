"""include/shelfi_ckks.hpp — the C++ mirror of the reference's class (include/ckks.h:27-54): builds against the
library on the CPU box; on a GPU box the reference's main.cpp scenario runs and asserts."""
import subprocess

import pytest

from conftest import FIXTURE_DIRS, ROOT, has_gpu

SRC = ROOT / "tests" / "cpp" / "main_example.cpp"
EXE = ROOT / "tests" / "cpp" / "_build" / "main_example"


def build():
    EXE.parent.mkdir(exist_ok=True)
    lib = ROOT / "fhe-fed_b200" / "lib"
    subprocess.run(["g++", "-std=c++17", "-O2", "-Wall", "-o", str(EXE), str(SRC), f"-L{lib}", "-lshelfi_b200",
                    f"-Wl,-rpath,{lib}", "-L/usr/local/cuda/lib64", "-Wl,-rpath,/usr/local/cuda/lib64"], check=True)


def test_cpp_wrapper_builds_and_links():
    build()
    assert EXE.exists()


@pytest.mark.skipif(has_gpu(), reason="no-GPU failure mode")
def test_cpp_example_fails_loudly_without_gpu():
    build()
    r = subprocess.run([str(EXE), str(FIXTURE_DIRS["pybind"]) + "/"], capture_output=True, text=True)
    assert r.returncode != 0 and "no CPU path" in (r.stderr + r.stdout)


@pytest.mark.gpu
def test_main_cpp_scenario_on_gpu():
    build()
    r = subprocess.run([str(EXE), str(FIXTURE_DIRS["pybind"]) + "/"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
