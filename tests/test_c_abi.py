"""The C ABI (include/tfd_b200.h): valid C99 and C++17, and in one-to-one correspondence with the symbols the built
library exports and with the ctypes binding -- an integrator linking without Python sees exactly this surface."""
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "tfd_b200.h")
LIB = os.path.join(ROOT, "tensorflow_distributed_b200", "_native", "libtfd_b200.so")


def _declared():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)  # comments mention entry points by name
    return set(re.findall(r"\b(tfd_[a-z0-9_]+)\s*\(", text))


@pytest.mark.skipif(shutil.which("gcc") is None, reason="needs gcc")
@pytest.mark.parametrize("lang,std", [("c", "c99"), ("c++", "c++17")])
def test_header_compiles_standalone(lang, std):
    r = subprocess.run(["gcc", "-fsyntax-only", "-x", lang, "-std=" + std, "-Wall", "-Werror", HEADER],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


@pytest.mark.skipif(not os.path.exists(LIB) or shutil.which("nm") is None, reason="needs the built library and nm")
def test_every_exported_symbol_is_declared_and_vice_versa():
    out = subprocess.run(["nm", "-D", "--defined-only", LIB], capture_output=True, text=True, check=True).stdout
    exported = {l.split()[-1] for l in out.splitlines() if " T tfd_" in l}
    assert exported == _declared()


def test_ctypes_binding_only_calls_declared_entry_points():
    src = open(os.path.join(ROOT, "tensorflow_distributed_b200", "ops", "native.py")).read()
    used = set(re.findall(r"\.(tfd_[a-z0-9_]+)\b", src))
    assert used and used <= _declared()


@pytest.mark.skipif(not os.path.exists(LIB) or shutil.which("gcc") is None or not os.path.isdir("/usr/local/cuda/include"),
                    reason="needs gcc, the CUDA headers and the built library")
def test_c_example_compiles_and_links_against_the_library(tmp_path):
    """examples/c_api_dense_step.c is plain C99 that uses only the public header; running it needs a B200."""
    exe = str(tmp_path / "dense_step")
    r = subprocess.run(["gcc", "-Wall", "-Werror", "-std=c99", os.path.join(ROOT, "examples", "c_api_dense_step.c"),
                        "-I" + os.path.join(ROOT, "include"), "-I/usr/local/cuda/include", "-L" + os.path.dirname(LIB),
                        "-ltfd_b200", "-L/usr/local/cuda/lib64", "-lcudart", "-lm", "-o", exe],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
