"""Host check of csrc/jpeg_unstuff.h (the per-lane word stepper of k_unstuff) against the byte loop that follows
stbi__grow_buffer_unsafe (jpeg.c:167-187) on random streams dense in FF / 00 / fill runs / markers."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(shutil.which("g++") is None, reason="g++ not available")
def test_word_stepper_matches_byte_loop(tmp_path):
    exe = tmp_path / "unstuff_check"
    subprocess.run(["g++", "-O1", "-I", os.path.join(ROOT, "image_codecs_b200", "csrc"),
                    os.path.join(ROOT, "tests", "native", "unstuff_check.cpp"), "-o", str(exe)], check=True)
    out = subprocess.run([str(exe), "60000"], capture_output=True, text=True)
    assert out.returncode == 0, out.stdout[-2000:]
    assert "ok 60000 cases" in out.stdout
