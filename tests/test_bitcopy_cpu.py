"""CPU check of the host batch coalescer's bit-range append (csrc/bitcopy.h), compiled with g++ -fsanitize=address."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_bitcopy_against_bit_model(tmp_path):
    exe = str(tmp_path / "bitcopy_test")
    subprocess.run(["g++", "-O1", "-g", "-fsanitize=address", "-std=c++17", "-I", os.path.join(ROOT, "datafusion_upstream_b200", "csrc"),
                    os.path.join(ROOT, "tests", "cpp", "bitcopy_test.cpp"), "-o", exe], check=True)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and r.stdout.strip() == "OK", r.stdout + r.stderr
