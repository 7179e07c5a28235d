"""The C-ABI library loads on a CPU-only box and exports every symbol include/ccgpu.h declares (no compute calls)."""
import ctypes
import os
import re

import cloudclean_b200 as cc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, "include", "ccgpu.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ccg_[a-z0-9_]+)\s*\(", src)))


def test_header_matches_binding_list():
    assert header_symbols() == sorted(cc.SYMBOLS)


def test_library_exports_every_symbol():
    lib = cc.load_library()
    for name in header_symbols():
        assert getattr(lib, name) is not None
    assert lib.ccg_version() == 100


def test_no_cpu_fallback_without_a_device():
    """without a usable GPU the product refuses to create a context (it never computes on the CPU)"""
    import torch
    if torch.cuda.is_available():
        return
    lib = cc.load_library()
    h = ctypes.c_void_p()
    rc = lib.ccg_create(ctypes.byref(h), 0)
    assert rc == -2 and not h.value            # CCG_ERR_CUDA
    try:
        cc.Context(0)
    except cc.CcgError as e:
        assert e.code == -2
    else:
        raise AssertionError("Context() must fail without a GPU")


def test_null_ctx_is_rejected():
    lib = cc.load_library()
    assert lib.ccg_set_cloud(None, None, 0, 32) == -1
    assert lib.ccg_knn(None, 5, None, None) == -1
    assert lib.ccg_last_error(None) == b"invalid context"
    assert lib.ccg_cloud_size(None) == -1


def test_product_never_imports_the_oracle():
    """the product path must not reach into oracle/ (or any CPU fallback)"""
    for dirpath, _, files in os.walk(os.path.join(ROOT, "cloudclean_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in text and "ccoracle" not in text, f
    for dirpath, _, files in os.walk(os.path.join(ROOT, "adapters")):
        for f in files:
            if f.endswith((".cpp", ".h", ".hpp")) and not f.startswith("test_"):
                text = open(os.path.join(dirpath, f)).read()
                assert "ccoracle" not in text and "cco_" not in text, f


def test_struct_layouts_match_the_header(tmp_path):
    """ctypes mirrors of the header's structs have the C compiler's sizes and field offsets"""
    import subprocess
    fields = [f[0] for f in cc.PtxInfo._fields_]
    prog = ['#include <stdio.h>', '#include <stddef.h>', '#include "ccgpu.h"', "int main(void) {",
            '  printf("%zu\\n", sizeof(ccg_ptx_info));']
    prog += [f'  printf("%zu\\n", offsetof(ccg_ptx_info, {f}));' for f in fields]
    prog += ['  printf("%zu\\n", sizeof(ccg_stats));', "  return 0; }"]
    src = tmp_path / "layout.c"
    src.write_text("\n".join(prog))
    exe = tmp_path / "layout"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    out = [int(x) for x in subprocess.check_output([str(exe)]).split()]
    assert out[0] == ctypes.sizeof(cc.PtxInfo)
    assert out[1:1 + len(fields)] == [getattr(cc.PtxInfo, f).offset for f in fields]
    assert out[-1] == ctypes.sizeof(cc.Stats)
