"""include/canopy-cpp.h compiles as plain C++17 against libcanopy.so and behaves: lowering works on a CPU box,
running without a GPU fails loudly with CANOPY_E_CUDA (no CPU fallback), validity errors carry their status."""
import os
import re
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
    assert "psource: n=3 deviates=2 mean_ok=1 repeat=1 fresh=1" in out
    if not torch.cuda.is_available():
        assert "run: status 5" in out and "no CPU fallback" in out
        assert "uncertainty: status 5" in out


def test_c_header_is_plain_c(tmp_path):
    src = tmp_path / "t.c"
    src.write_text('#include "canopy.h"\nint main(void){ canopy_mc_opts o; canopy_mc_opts_default(&o); return (int)canopy_spec_version() - CANOPY_SPEC_VERSION; }\n')
    exe = str(tmp_path / "t")
    lib_dir = os.path.join(ROOT, "canopy_b200")
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-pedantic", "-I", os.path.join(ROOT, "include"), str(src), "-o", exe,
                           os.path.join(lib_dir, "libcanopy.so"), f"-Wl,-rpath,{lib_dir}"])
    assert subprocess.call([exe]) == 0


def test_cmake_package_roundtrip(tmp_path):
    """SURVEY.md 8b packaging: `cmake --install` gives lib `canopy` + exactly the two public headers + a relocatable
    find_package(canopy) that sets CANOPY_INCLUDE_DIR / CANOPY_LIB_DIR / CANOPY_BIN_DIR (reference
    cmake/canopy-config.cmake.in:8-10); a C++23 consumer links the imported target and runs."""
    import shutil
    import pytest
    if not (shutil.which("cmake") and shutil.which("ninja") and shutil.which("nvcc")):
        pytest.skip("cmake / ninja / nvcc not on PATH")
    build, prefix, moved = tmp_path / "b", tmp_path / "prefix", tmp_path / "moved"
    subprocess.check_call(["cmake", "-S", ROOT, "-B", str(build), "-G", "Ninja", f"-DCMAKE_INSTALL_PREFIX={prefix}",
                           "-DBUILD_TESTS=OFF"], stdout=subprocess.DEVNULL)
    subprocess.check_call(["cmake", "--build", str(build), "-j", "8"], stdout=subprocess.DEVNULL)
    subprocess.check_call(["cmake", "--install", str(build)], stdout=subprocess.DEVNULL)
    assert sorted(os.listdir(prefix / "include")) == ["canopy-cpp.h", "canopy.h"]
    flags = (build / "build.ninja").read_text()
    assert "arch=compute_100a,code=[sm_100a]" in flags and "-lineinfo" in flags
    shutil.move(str(prefix), str(moved))     # relocatable: the package must not remember its install prefix
    cbuild = tmp_path / "c"
    cfg = subprocess.check_output(["cmake", "-S", os.path.join(ROOT, "tests", "cpp", "consumer"), "-B", str(cbuild), "-G", "Ninja",
                                   f"-DCMAKE_PREFIX_PATH={moved}"], text=True)
    assert f"CANOPY_INCLUDE_DIR={moved}/include" in cfg
    lib_dir = re.search(r"CANOPY_LIB_DIR=(\S+)", cfg).group(1)      # lib or lib64 (GNUInstallDirs)
    assert lib_dir.startswith(f"{moved}/lib")
    assert f"CANOPY_BIN_DIR={moved}/bin" in cfg
    subprocess.check_call(["cmake", "--build", str(cbuild)], stdout=subprocess.DEVNULL)
    out = subprocess.check_output([str(cbuild / "consumer")], text=True)
    assert "layers=1 gates=1 tallies=2" in out and "validity: status 2" in out
    # the installed library exports the same C ABI as the Makefile build
    nm = subprocess.check_output(["nm", "-D", "--defined-only", os.path.join(lib_dir, "libcanopy.so")], text=True)
    assert " T canopy_plan_create" in nm and " T canopy_mc_run" in nm
