"""GPU: a plain C99 program drives the whole path through the C ABI (no Python, no torch in the loop)
and checks every byte against the plain-C oracle."""
import os
import shutil
import subprocess

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _build_and_run(tmp_path, name, marker):
    import torch
    if not torch.cuda.is_available():
        pytest.fail("marked gpu but no CUDA device")
    gcc = shutil.which("gcc")
    if not gcc:
        pytest.skip("no gcc on this machine")
    from oracle import port
    port.build()
    lib_dir = os.path.join(ROOT, "video_edge_ai_proxy_b200")
    ora_dir = os.path.join(ROOT, "oracle", "_build")
    exe = str(tmp_path / name)
    subprocess.check_call([gcc, "-std=c99", "-O1", "-Wall", "-I", os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "c", name + ".c"), "-o", exe,
                           "-L", lib_dir, "-lvepb200", "-L", ora_dir, "-lyuv2bgr_port",
                           f"-Wl,-rpath,{lib_dir}", f"-Wl,-rpath,{ora_dir}"])
    out = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0 and marker in out.stdout, out.stdout + out.stderr


def test_c_program_roundtrip(tmp_path):
    _build_and_run(tmp_path, "abi_roundtrip", "C_ABI_ROUNDTRIP_OK")


def test_c_program_batch_tickets_pinned_pool_and_slots(tmp_path):
    """The round-2 entry points from plain C: vep_stream_open_many, vep_submit_batch + helper threads, ticketed ticks,
    vep_host_register + VEP_SUBMIT_PINNED, vep_acquire_input / vep_commit_input."""
    _build_and_run(tmp_path, "abi_batch", "C_ABI_BATCH_OK")
