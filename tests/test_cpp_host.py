"""The C++ host mirror (include/igs.hpp): compiles everywhere (CPU check), runs on the GPU box."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def build(tmp_path, built_lib, orc):
    exe = str(tmp_path / "test_scoring_engine")
    pkg_dir = os.path.dirname(built_lib)
    cmd = ["g++", "-std=c++17", "-O1", "-Wall", "-pthread", "-I", os.path.join(ROOT, "include"),
           os.path.join(ROOT, "tests", "cpp", "test_scoring_engine.cc"), "-o", exe,
           "-L", pkg_dir, "-ligs", "-L", os.path.join(ROOT, "oracle"), "-loracle_igs",
           f"-Wl,-rpath,{pkg_dir}", f"-Wl,-rpath,{os.path.join(ROOT, 'oracle')}"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    return exe


def test_cpp_mirror_compiles_and_links(tmp_path, built_lib, orc):
    exe = build(tmp_path, built_lib, orc)
    r = subprocess.run([exe, "--compile-only-check"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


@pytest.mark.gpu
def test_cpp_mirror_scores_like_the_reference(tmp_path, built_lib, orc):
    exe = build(tmp_path, built_lib, orc)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "ok" in r.stdout
