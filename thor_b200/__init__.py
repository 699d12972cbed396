"""thor_b200 -- host-side binding of libthorcuda (hand-written CUDA for B200 / sm_100a).

The product is the C-ABI shared library ``thor_b200/libthorcuda.so`` (include/thorcuda.h).  This
package is the thin Python mirror used by the tests and the benchmark; it never computes pixels
itself and has no CPU fallback: if the library is missing, or no sm_100 device is present,
``init()`` raises.
"""
from .lib import (LF_CLPF, LF_DEBLOCK_UV, LF_DEBLOCK_Y, LF_MAKE_REF, LIST_COST, LIST_PRED, LIST_TB, COST_FASTHALF,  # noqa: F401
                  COST_FASTQUARTER, COST_SAD, COST_SSD, COST_WIDESAD, COST_TASK, COST_RESULT, PRED_TASK, TB_TASK, BLOCK_SYNTAX, NO_TB, Picture,
                  ThorCudaError, cdll, check, init, launch_count, lib_path, shutdown, sync)
from . import ops  # noqa: F401
