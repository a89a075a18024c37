"""The C-ABI library loads and exports every symbol include/seqrush_b200.h declares; the ctypes
struct layouts match the header; without a CUDA device every entry point fails loudly (no CPU path)."""
import ctypes as C
import os
import re
import subprocess

import pytest

from conftest import ROOT

HEADER = os.path.join(ROOT, "include", "seqrush_b200.h")


def _declared():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(srgpu_[a-z0-9_]+)\s*\(", src)))


def test_exports_match_header(built):
    from seqrush_b200 import _lib
    names = _declared()
    assert sorted(_lib.EXPORTS) == names
    L = C.CDLL(_lib.LIB_PATH)
    for n in names:
        assert hasattr(L, n), n


def test_struct_layouts_match_header(built, tmp_path):
    from seqrush_b200 import _lib
    prog = tmp_path / "sz.c"
    prog.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "seqrush_b200.h"\nint main(){printf("%zu %zu %zu %zu %zu %zu %zu %zu\\n",'
                    'sizeof(srgpu_seqview),sizeof(srgpu_penalties),sizeof(srgpu_result),sizeof(srgpu_stats),sizeof(srgpu_region),'
                    'offsetof(srgpu_result,cigar_off),offsetof(srgpu_stats,align_ms),offsetof(srgpu_region,seq1_is_rc));return 0;}\n')
    exe = tmp_path / "sz"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), str(prog), "-o", str(exe)], check=True)
    got = [int(x) for x in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()]
    want = [C.sizeof(_lib.SeqView), C.sizeof(_lib.Penalties), C.sizeof(_lib.Result), C.sizeof(_lib.Stats),
            C.sizeof(_lib.Region), _lib.Result.cigar_off.offset, _lib.Stats.align_ms.offset, _lib.Region.seq1_is_rc.offset]
    assert got == want


def test_no_cpu_fallback(built):
    """Without a device srgpu_ctx_create must fail with SRGPU_E_CUDA, and the Python mirror must raise."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    from seqrush_b200 import _lib
    import seqrush_b200 as s
    h = C.c_void_p()
    assert _lib.lib().srgpu_ctx_create(0, None, C.byref(h)) == -1
    assert b"CUDA" in _lib.lib().srgpu_last_error() or b"device" in _lib.lib().srgpu_last_error()
    with pytest.raises(s.SrgpuError):
        s.Context(0)


def test_product_never_imports_oracle():
    """Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may touch oracle/."""
    pkg = os.path.join(ROOT, "seqrush_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dp, f), errors="replace").read()
                assert "_oracle" not in txt and "liboracle" not in txt and "libsrref" not in txt, f
