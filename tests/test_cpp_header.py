"""include/canopy-cpp.h compiles as plain C++17 against libcanopy.so and behaves: lowering works on a CPU box,
running without a GPU fails loudly with CANOPY_E_CUDA (no CPU fallback), validity errors carry their status."""
import os
import subprocess

import torch

from helpers import ROOT


def test_cpp_header_smoke(tmp_path):
    exe = str(tmp_path / "header_smoke")
    lib_dir = os.path.join(ROOT, "canopy_b200")
    subprocess.check_call(["g++", "-std=c++17", "-Wall", "-I", os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "cpp", "header_smoke.cpp"), "-o", exe,
                           os.path.join(lib_dir, "libcanopy.so"), f"-Wl,-rpath,{lib_dir}"])
    out = subprocess.check_output([exe], text=True)
    assert "layers=1 gates=1 tallies=2" in out
    assert "validity: status 2" in out
    assert "session: fingerprint 1 done=0 tallies=2" in out
    assert "update_p: status 1" in out      # an event cannot turn constant without a new lowering
    if not torch.cuda.is_available():
        assert "run: status 5" in out and "no CPU fallback" in out


def test_c_header_is_plain_c(tmp_path):
    src = tmp_path / "t.c"
    src.write_text('#include "canopy.h"\nint main(void){ canopy_mc_opts o; canopy_mc_opts_default(&o); return (int)canopy_spec_version() - 2; }\n')
    exe = str(tmp_path / "t")
    lib_dir = os.path.join(ROOT, "canopy_b200")
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-pedantic", "-I", os.path.join(ROOT, "include"), str(src), "-o", exe,
                           os.path.join(lib_dir, "libcanopy.so"), f"-Wl,-rpath,{lib_dir}"])
    assert subprocess.call([exe]) == 0
