"""The C++ mirror of the reference's Store interface (include/docagents_b200.hpp) and its contract
test: compiled on every run (CPU), executed on a GPU."""
import os
import subprocess

import pytest

import oracle

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _build(tmp_path, built_lib):
    oracle.build()
    exe = tmp_path / "contract_test"
    cmd = ["/usr/bin/g++", "-std=c++17", "-O1", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"),
           os.path.join(ROOT, "tests", "cpp", "contract_test.cpp"), "-o", str(exe),
           built_lib, os.path.join(ROOT, "oracle", "liboracle.so"),
           f"-Wl,-rpath,{os.path.dirname(built_lib)}", f"-Wl,-rpath,{os.path.join(ROOT, 'oracle')}"]
    subprocess.run(cmd, check=True)
    return exe


def test_cpp_mirror_compiles_and_links(tmp_path, built_lib):
    assert os.path.exists(_build(tmp_path, built_lib))


@pytest.mark.gpu
def test_cpp_contract_replays_reference_tests(tmp_path, built_lib):
    exe = _build(tmp_path, built_lib)
    out = subprocess.run([str(exe)], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "contract_test ok" in out.stdout
