"""The top-level CMakeLists.txt builds the C-ABI library (CUDA, sm_100a), the C++ drop-in class and its caller-pattern
link test -- what the reference's CMakeLists.txt:51-79 gains is in integration/ORB_SLAM2.CMakeLists.patch.
nvcc cross-compiles without a GPU, so this runs on CPU (about a minute and a half: one nvcc pass over borb_api.cu)."""
import ctypes
import re
import shutil
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]


@pytest.mark.skipif(shutil.which("cmake") is None or not Path("/usr/local/cuda/bin/nvcc").exists(), reason="cmake or nvcc missing")
def test_cmake_configures_and_builds(tmp_path):
    build = tmp_path / "build"
    cfg = subprocess.run(["cmake", "-S", str(ROOT), "-B", str(build), "-DCMAKE_CUDA_COMPILER=/usr/local/cuda/bin/nvcc"],
                         capture_output=True, text=True)
    assert cfg.returncode == 0, cfg.stdout[-2000:] + cfg.stderr[-2000:]
    bld = subprocess.run(["cmake", "--build", str(build), "-j", "4"], capture_output=True, text=True)
    assert bld.returncode == 0, bld.stdout[-3000:] + bld.stderr[-3000:]
    lib = build / "liborb_b200.so"
    assert lib.exists() and (build / "libORBextractor_b200.so").exists() and (build / "frame_like").exists()
    # every symbol include/borb.h declares is exported, and the cubin is sm_100a only
    h = ctypes.CDLL(str(lib))
    text = (ROOT / "include" / "borb.h").read_text()
    for sym in sorted(set(re.findall(r"\b(borb_[a-z0-9_]+)\s*\(", text))):
        assert hasattr(h, sym), sym
    elf = subprocess.run(["/usr/local/cuda/bin/cuobjdump", "-lelf", str(lib)], capture_output=True, text=True).stdout
    assert "sm_100a" in elf and not re.search(r"sm_(?!100a)\d+", elf), elf


def test_integration_patch_names_the_reference_lines():
    patch = (ROOT / "integration" / "ORB_SLAM2.CMakeLists.patch").read_text()
    assert "-src/ORBextractor.cc" in patch and "+ORBextractor_b200" in patch and "add_subdirectory" in patch
    ref = Path("/root/reference/CMakeLists.txt")
    if ref.exists():  # the context lines of the patch are the reference's own lines
        lines = set(l.rstrip() for l in ref.read_text().splitlines())
        for l in patch.splitlines():
            if l.startswith(" ") and l.strip() and not l.startswith(" #"):
                assert l[1:].rstrip() in lines, f"context line not in the reference: {l!r}"
