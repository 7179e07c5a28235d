"""The reference-named C++ host layer (adapters/cloudclean_gpu.hpp) over the C ABI, checked on cuda:0 by the
C++ test program adapters/test_adapters.cpp (brute-force CPU restatements inside the test)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "adapters", "test_adapters")


def test_adapter_header_compiles_standalone():
    """CPU: the header needs nothing but the C ABI header and the standard library."""
    src = '#include "adapters/cloudclean_gpu.hpp"\nint main() { return sizeof(ccgpu::Normal) == 32 ? 0 : 1; }\n'
    subprocess.run(["g++", "-std=c++14", "-fsyntax-only", "-I", ROOT, "-x", "c++", "-"], input=src.encode(), check=True, cwd=ROOT)


@pytest.mark.gpu
def test_adapters_on_gpu():
    if not os.path.exists(EXE):
        subprocess.check_call(["make", "-C", ROOT, "adapters"])
    r = subprocess.run([EXE], capture_output=True, text=True, timeout=600)
    print(r.stdout)
    print(r.stderr)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-1000:]
    assert "all checks passed" in r.stdout
