"""The plain-C caller (examples/ffi_demo.c: no Python, no CUDA headers) compiled, linked and RUN against libfcb200.so."""
import os
import subprocess

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_c_host_program_runs_a_filtered_search(built, tmp_path):
    from financial_chatbot_llm_b200 import _lib
    lib_dir = os.path.dirname(_lib.LIB_PATH)
    out = tmp_path / "ffi_demo"
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"),
                    os.path.join(ROOT, "examples", "ffi_demo.c"), "-L", lib_dir, "-lfcb200", "-L", "/usr/local/cuda/lib64", "-lcudart",
                    "-Wl,-rpath," + lib_dir, "-o", str(out)], check=True)
    run = subprocess.run([str(out)], capture_output=True, text=True, timeout=120)
    assert run.returncode == 0, run.stdout + run.stderr          # 5 results, the nearest is the row the query was built from
    lines = run.stdout.strip().splitlines()
    scores = [float(l.rsplit(" ", 1)[1]) for l in lines]
    assert len(lines) == 5 and lines[0].startswith("0: id 2000 score ") and scores[0] > 0.8 > scores[1]
    assert scores == sorted(scores, reverse=True)
