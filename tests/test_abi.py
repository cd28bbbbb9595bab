"""The C-ABI library builds for sm_100a without a GPU, loads, and exports every symbol
include/fcb200.h declares (no compute calls here)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "fcb200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(fcb_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(built):
    from financial_chatbot_llm_b200 import _lib
    lib = ctypes.CDLL(_lib.LIB_PATH)
    names = declared_symbols()
    assert "fcb_search" in names and "fcb_open" in names and len(names) >= 10
    for n in names:
        assert hasattr(lib, n), f"{n} declared in fcb200.h but not exported"
    assert sorted(_lib.SIGNATURES) == names, "ctypes binding and header disagree"


def test_struct_layouts_match_header(built):
    from financial_chatbot_llm_b200 import _lib
    assert ctypes.sizeof(_lib.FcbFilter) == 40
    assert ctypes.sizeof(_lib.FcbConfig) == 16
    assert ctypes.sizeof(_lib.FcbShard) == 7 * 8 + 3 * 8 + 2 * 4
    assert _lib.load().fcb_abi_version() == _lib.ABI_VERSION


def test_no_cpu_fallback(built):
    """Without a GPU the product must fail loudly, never route to a CPU path."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from financial_chatbot_llm_b200._lib import FcbError
    from financial_chatbot_llm_b200.engine import Engine
    with pytest.raises(FcbError, match="no CUDA device"):
        Engine()


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "financial-chatbot-llm_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f
                assert "libcanon" not in src, f


def test_plain_c_caller_compiles_and_links(built, tmp_path):
    """The boundary is a C ABI: the header compiles as C99 and as C++, and examples/ffi_demo.c (a host with no Python
    and no CUDA headers) compiles and links against the library. Running it needs a GPU (tests/test_gpu_search.py)."""
    import subprocess
    from financial_chatbot_llm_b200 import _lib
    root = ROOT
    inc = os.path.join(root, "include")
    for compiler, std in (("gcc", "-std=c99"), ("g++", "-std=c++11")):
        src = tmp_path / ("hdr.c" if compiler == "gcc" else "hdr.cpp")
        src.write_text('#include "fcb200.h"\nint main(void) { return (int)sizeof(fcb_filter) - 40; }\n')
        subprocess.run([compiler, std, "-Wall", "-Werror", "-I", inc, "-fsyntax-only", str(src)], check=True)
    cuda_lib = "/usr/local/cuda/lib64"
    if not os.path.exists(os.path.join(cuda_lib, "libcudart.so")):
        pytest.skip("no libcudart to link the demo against")
    lib_dir = os.path.dirname(_lib.LIB_PATH)
    out = tmp_path / "ffi_demo"
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", inc, os.path.join(root, "examples", "ffi_demo.c"), "-L", lib_dir,
                    "-lfcb200", "-L", cuda_lib, "-lcudart", "-Wl,-rpath," + lib_dir, "-o", str(out)], check=True)
    assert out.exists()
