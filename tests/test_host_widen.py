"""Host half of rti_ptm_compute_host's narrow transfer, without a GPU: the int16 -> int32
widening loop and the Wideners thread protocol (csrc/host_widen.h, host_widen.cpp) are plain C++, so the
check driver tools/bench_widen.cpp is built with g++ here and run: every pointer alignment and ragged
length, tiles in order through a reused three-tile staging ring, skipped tiles left untouched, and
destruction before all announced tiles were published (must join, not hang)."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(shutil.which("g++") is None, reason="needs g++")
def test_widening_loop_and_thread_protocol(tmp_path):
    exe = str(tmp_path / "widen_check")
    subprocess.run(["g++", "-O2", "-std=c++17", "-pthread", os.path.join(ROOT, "tools", "bench_widen.cpp"),
                    os.path.join(ROOT, "smartphone-based-rti_b200", "csrc", "host_widen.cpp"), "-o", exe],
                   check=True, timeout=300)
    out = subprocess.run([exe, "--check"], check=True, timeout=120, capture_output=True, text=True).stdout
    assert out.strip().endswith("check ok"), out
