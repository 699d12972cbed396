"""CPU-only checks of the drop-in boundary: libthorcuda.so loads without a GPU, exports every function that
include/thorcuda.h and include/thorcuda_compat.h declare, the ctypes struct mirrors match the header's
sizes, and the product refuses to run (instead of falling back) when no sm_100 device is present."""
import ctypes as C
import os
import re
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
INC = os.path.join(ROOT, "include")
LIB = os.path.join(ROOT, "thor_b200", "libthorcuda.so")


def declared_functions(header):
    """names of the function prototypes of a C header (comments and typedef bodies stripped)"""
    src = open(os.path.join(INC, header)).read()
    src = re.sub(r"/\*.*?\*/", " ", src, flags=re.S)
    src = re.sub(r"//[^\n]*", " ", src)
    src = re.sub(r"typedef\s+struct\s*\{.*?\}\s*\w+\s*;", " ", src, flags=re.S)
    src = "\n".join(ln for ln in src.splitlines() if not ln.lstrip().startswith("#"))
    names = []
    for m in re.finditer(r"([A-Za-z_]\w*)\s*\(([^;{}()]|\([^()]*\))*\)\s*;", src):
        names.append(m.group(1))
    return sorted(set(names))


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(LIB):
        subprocess.run(["make", "-j8"], cwd=ROOT, check=True, capture_output=True)
    return C.CDLL(LIB)


def test_headers_declare_something():
    a, b = declared_functions("thorcuda.h"), declared_functions("thorcuda_compat.h")
    assert len(a) >= 40 and "thorcuda_loopfilter_frames" in a and "thorcuda_sad_table_build" in a
    # the reference's own kernel API: common_kernels.h:30-37, enc_kernels.h:32-37, transform.h:36-37
    assert b == sorted(["block_avg_simd", "sad_calc_simd_unaligned", "get_inter_prediction_luma_simd",
                        "get_inter_prediction_chroma_simd", "transform_simd", "inverse_transform_simd", "clpf_block4", "clpf_block8",
                        "sad_calc_simd", "ssd_calc_simd", "detect_clpf_simd", "sad_calc_fasthalf_simd", "sad_calc_fastquarter_simd",
                        "widesad_calc_simd", "transform", "inverse_transform"])


@pytest.mark.parametrize("header", ["thorcuda.h", "thorcuda_compat.h"])
def test_every_declared_symbol_is_exported(lib, header):
    missing = [n for n in declared_functions(header) if not hasattr(lib, n)]
    assert not missing, f"declared in include/{header} but not exported by libthorcuda.so: {missing}"


def test_struct_sizes_match_header():
    """compile a C probe against the header and compare with the numpy mirrors in thor_b200/lib.py"""
    sys.path.insert(0, ROOT)
    from thor_b200.lib import COST_RESULT, COST_TASK, PRED_TASK, TB_TASK
    from thor_b200.ops import LfJob
    probe = r'''
#include <stdio.h>
#include "thorcuda.h"
#include "thorcuda_compat.h"
int main(void){printf("%zu %zu %zu %zu %zu\n",sizeof(thorcuda_pred_task),sizeof(thorcuda_tb_task),sizeof(thorcuda_cost_task),
sizeof(thorcuda_cost_result),sizeof(thorcuda_lf_job));return 0;}
'''
    exe = os.path.join(ROOT, "build", "abi_probe")
    os.makedirs(os.path.dirname(exe), exist_ok=True)
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", INC, "-x", "c", "-", "-o", exe], input=probe.encode(), check=True)
    out = subprocess.run([exe], check=True, capture_output=True, text=True).stdout.split()
    assert [int(x) for x in out] == [PRED_TASK.itemsize, TB_TASK.itemsize, COST_TASK.itemsize, COST_RESULT.itemsize, C.sizeof(LfJob)]


def test_no_cpu_fallback_without_gpu(lib):
    """Without a CUDA device init fails with a message and every compute entry point aborts the process
    (fatalerror() convention), it never computes on the CPU.  Skipped where a GPU is visible."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is visible")
    lib.thorcuda_last_error.restype = C.c_char_p
    assert lib.thorcuda_init(0) != 0
    assert b"no CPU fallback" in lib.thorcuda_last_error()
    code = ("import ctypes,sys; l=ctypes.CDLL(%r); a=(ctypes.c_uint8*64)(); "
            "l.sad_calc_simd(a,a,8,8,8,8); print('computed')" % LIB)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True)
    assert r.returncode != 0 and "computed" not in r.stdout
    assert "Run-time error" in r.stderr


def test_product_does_not_reference_the_oracle():
    """oracle/ is test infrastructure: nothing under thor_b200/ or include/ may mention it."""
    bad = []
    for base in ("thor_b200", "include"):
        for dp, _, files in os.walk(os.path.join(ROOT, base)):
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".h")):
                    txt = open(os.path.join(dp, f), errors="ignore").read()
                    if re.search(r"oracle/|thor_oracle|libthorref|libthorpipe", txt):
                        bad.append(os.path.join(dp, f))
    assert not bad, bad
