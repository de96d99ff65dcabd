"""CPU: the C-ABI library loads, exports every symbol include/vepb200.h declares, and refuses to run
without a GPU (no silent CPU fallback).  No compute calls here."""
import ctypes
import os
import re
import subprocess

import pytest

from video_edge_ai_proxy_b200 import _abi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "vepb200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(vep_[a-z_0-9]+)\s*\(", src)))


def test_header_symbols_are_exported_and_bound():
    names = _declared()
    assert len(names) >= 20
    L = _abi.lib()
    for n in names:
        assert hasattr(L, n), f"{n} declared in vepb200.h but not exported"
    assert sorted(_abi.SIGNATURES) == names, "ctypes binding and header disagree"
    exported = subprocess.check_output(["nm", "-D", "--defined-only", _abi.LIB_PATH], text=True)
    extra = set(re.findall(r" T (vep_[a-z_0-9]+)", exported)) - set(names)
    assert not extra, f"exported but undeclared: {extra}"


def test_header_is_plain_c():
    """The boundary is a C ABI: the header must compile as C99 on its own, and a C caller must link."""
    import shutil
    import tempfile
    gcc = shutil.which("gcc")
    if not gcc:
        pytest.skip("no gcc")
    with tempfile.TemporaryDirectory() as d:
        src = os.path.join(d, "t.c")
        with open(src, "w") as f:
            f.write('#include "vepb200.h"\n#include <stdio.h>\n'
                    'int main(void){ vep_frame_desc d; vep_frame_meta m; (void)d; (void)m;\n'
                    ' printf("%d %s %d\\n", vep_abi_version(), vep_status_name(VEP_ERR_BUSY), (int)sizeof(vep_frame_desc)); return 0; }\n')
        exe = os.path.join(d, "t")
        lib_dir = os.path.dirname(_abi.LIB_PATH)
        subprocess.check_call([gcc, "-std=c99", "-Wall", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"), src, "-o", exe,
                               "-L", lib_dir, "-lvepb200", f"-Wl,-rpath,{lib_dir}"])
        out = subprocess.check_output([exe], text=True).split()
        assert out == ["1", "VEP_ERR_BUSY", "72"]


def test_abi_version_and_status_names():
    L = _abi.lib()
    assert L.vep_abi_version() == 1
    assert L.vep_status_name(0) == b"VEP_OK"
    assert L.vep_status_name(-2) == b"VEP_ERR_GEOMETRY"


def test_struct_layouts_match_header():
    assert ctypes.sizeof(_abi.FrameDesc) == 72      # static_assert'ed in vep_abi.cu
    assert ctypes.sizeof(_abi.FrameMeta) == 64
    assert _abi.FrameDesc.dst.offset == 40 and _abi.FrameDesc.width.offset == 52
    assert _abi.FrameMeta.time_base.offset == 40 and _abi.FrameMeta.frame_type.offset == 56
    assert ctypes.sizeof(_abi.SubmitItem) == 112 and _abi.SubmitItem.plane.offset == 8 and _abi.SubmitItem.meta.offset == 48


def test_library_carries_sm100a_code_only():
    out = subprocess.run(["cuobjdump", "-lelf", _abi.LIB_PATH], capture_output=True, text=True)
    if out.returncode != 0:
        pytest.skip("cuobjdump unavailable")
    assert "sm_100a" in out.stdout
    assert not re.search(r"sm_(?!100a)\d+", out.stdout)


def test_no_gpu_means_loud_failure():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from video_edge_ai_proxy_b200 import FramePath, VepError
    with pytest.raises(VepError) as e:
        FramePath(0)
    assert e.value.status == _abi.VEP_ERR_CUDA and "no CPU path" in str(e.value)


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "video_edge_ai_proxy_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", text, flags=re.M), f
                assert "swscale_live" not in text and "libyuv2bgr_port" not in text, f


def test_missing_library_fails_loudly(monkeypatch):
    """No silent fallback: without libvepb200.so the binding raises, it does not convert on the CPU."""
    monkeypatch.setattr(_abi, "_lib", None)
    monkeypatch.setattr(_abi, "LIB_PATH", os.path.join(ROOT, "video_edge_ai_proxy_b200", "no_such_lib.so"))
    with pytest.raises(ImportError) as e:
        _abi.lib()
    assert "no CPU or PyTorch fallback" in str(e.value)
