"""csrc/sincosf_glibc.h (the device-side restatement of glibc's sinf/cosf used by the descriptor
rotation) against the live libm on a strided sweep of every float in [0, 2*pi]; the exhaustive
sweep (stride 1, 1.09e9 values, 0 mismatches) is recorded in DESIGN.md."""
import json
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_sincosf_matches_libm(tmp_path):
    exe = str(tmp_path / "sincosf_check")
    subprocess.run(["g++", "-O2", "-ffp-contract=off", "-o", exe, os.path.join(ROOT, "tools", "sincosf_check.cpp"), "-lm"], check=True)
    out = subprocess.run([exe, "101"], capture_output=True, text=True)
    res = json.loads(out.stdout.strip().splitlines()[-1])
    assert out.returncode == 0 and res["sin_mismatch"] == 0 and res["cos_mismatch"] == 0 and res["n"] > 10_000_000
