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


@pytest.mark.parametrize("header", ["thorcuda.h", "thorcuda_compat.h", "thorcuda_enc.h"])
def test_every_declared_symbol_is_exported(lib, header):
    missing = [n for n in declared_functions(header) if not hasattr(lib, n)]
    assert not missing, f"declared in include/{header} but not exported by libthorcuda.so: {missing}"


def test_struct_sizes_match_header():
    """compile a C probe against the header and compare with the numpy mirrors in thor_b200/lib.py"""
    sys.path.insert(0, ROOT)
    from thor_b200.lib import COST_RESULT, COST_TASK, INTRA_TASK, PRED_TASK, SKIP_TASK, TB_TASK
    from thor_b200.ops import LfJob
    probe = r'''
#include <stdio.h>
#include "thorcuda.h"
#include "thorcuda_compat.h"
int main(void){printf("%zu %zu %zu %zu %zu %zu %zu\n",sizeof(thorcuda_pred_task),sizeof(thorcuda_tb_task),sizeof(thorcuda_cost_task),
sizeof(thorcuda_cost_result),sizeof(thorcuda_lf_job),sizeof(thorcuda_skip_task),sizeof(thorcuda_intra_task));return 0;}
'''
    exe = os.path.join(ROOT, "build", "abi_probe")
    os.makedirs(os.path.dirname(exe), exist_ok=True)
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", INC, "-x", "c", "-", "-o", exe], input=probe.encode(), check=True)
    out = subprocess.run([exe], check=True, capture_output=True, text=True).stdout.split()
    assert [int(x) for x in out] == [PRED_TASK.itemsize, TB_TASK.itemsize, COST_TASK.itemsize, COST_RESULT.itemsize, C.sizeof(LfJob), SKIP_TASK.itemsize, INTRA_TASK.itemsize]


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


def test_host_helpers_against_oracle(lib):
    """a20 and a19's floating-point step are pure host code in the library: clip_mv, the table index of a probe and
    the early-skip thresholds, checked against the oracle (itself pinned to the reference in test_oracle_vs_ref)."""
    import numpy as np
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from thorref import load_oracle
    sys.path.insert(0, ROOT)
    from thor_b200.ops import clip_mv, early_skip_thresholds, sad_table_index
    orc = load_oracle()
    orc.orc_early_skip_threshold.argtypes = [C.c_int, C.c_int, C.c_double]
    rng = np.random.default_rng(3)
    hit = 0
    for _ in range(3000):
        fw, fh = (416, 240) if rng.random() < 0.5 else (1920, 1080)
        size = int(rng.choice([8, 16, 32, 64]))
        ypos, xpos = int(rng.integers(0, fh // size)) * size, int(rng.integers(0, fw // size)) * size
        sign = int(rng.integers(0, 2))
        scale = 8 if rng.random() < 0.3 else 1
        mx, my = int(rng.integers(-900, 900)) * scale, int(rng.integers(-900, 900)) * scale
        ax, ay = C.c_int16(mx), C.c_int16(my)
        orc.orc_clip_mv(C.byref(ax), C.byref(ay), ypos, xpos, fw, fh, size, sign)
        assert clip_mv((mx, my), ypos, xpos, fw, fh, size, sign) == (ax.value, ay.value)
        # table index: the probe address the reference forms is ref + s*(mv.y>>2)*stride + s*(mv.x>>2)
        mv = (int(rng.integers(-90, 90)), int(rng.integers(-90, 90)))
        c = (int(rng.integers(-4, 5)), int(rng.integers(-4, 5)))
        s = -1 if sign else 1
        dx, dy = s * (mv[0] >> 2) - c[0], s * (mv[1] >> 2) - c[1]
        want = -1
        if -16 <= dx < 16 and -16 <= dy < 16 and xpos + size <= fw:
            want = (((ypos // size) * (fw // size) + xpos // size) * 32 + dy + 16) * 32 + dx + 16
            hit += 1
        assert sad_table_index(fw, size, xpos, ypos, mv, sign, c, 16) == want
    assert hit > 500
    for qp in range(0, 52):
        for size in (8, 16):
            for thr in (0.3, 0.8, 1.0, float(np.float32(1.3) * np.float32(0.8))):
                t32 = float(np.float32(thr))
                ty, tc = early_skip_thresholds(qp, size, t32)
                assert ty == orc.orc_early_skip_threshold(qp, size // 2, 0.5 * t32)
                qpc = orc.orc_chroma_qp(qp)
                assert tc == int(t32 * (float(1 << (16 + qpc // 6)) / [26214, 23302, 20560, 18396, 16384, 14564][qpc % 6]))
