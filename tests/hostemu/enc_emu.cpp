// TEST INFRASTRUCTURE ONLY -- host emulation of libthorcuda's superblock encoder.
//
// thor_b200/csrc/enc/enc_core.h is written as uniform control flow + cooperative pixel loops; compiled with THOR_HOSTEMU
// the thread group is a single thread and barriers are no-ops, so the SAME source that nvcc turns into the sm_100a kernel
// runs here on the host, superblock after superblock in raster order.  integration/_build/Thorenc_emu links it behind the
// reference-side hook (integration/enc_gpu_hook.c, -DTHOR_EMU) so that the decision logic can be compared with the stock
// encoder byte for byte in a container without a GPU (tests/test_enc_emu.py).  Nothing under thor_b200/ uses this file:
// the library has no CPU path.
#define THOR_HOSTEMU 1
#include "../../thor_b200/csrc/enc/enc_core.h"

extern "C" void thor_emu_encode_frame(const thor_enc_frame *F) {
  using namespace thorenc;
  static SbMem S;
  Grp g;
  S.frame = *F;
  build_tables(g, S);
  const int nsbx = (F->width + 63) / 64, nsby = (F->height + 63) / 64;
  for (int k = 0; k < nsby; k++)
    for (int l = 0; l < nsbx; l++) {
      Enc E{g, S, S.frame, SbState(), F->nodes + (size_t)(k * nsbx + l) * THOR_ENC_NODES, F->coef + (k * nsbx + l), l * 64, k * 64};
      encode_superblock(E);
    }
}
