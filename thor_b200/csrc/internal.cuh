// libthorcuda internals: context, device pictures, scratch buffers, launch bookkeeping.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include "../../include/thorcuda.h"

namespace thor {

// Device view of a picture (geometry: reference create_yuv_frame, common/common_frame.c:324-351).
struct PicView {
  uint8_t *p[3];       // visible origin of Y,U,V
  int stride[3];       // bytes per row (U and V share stride_c)
  int w, h;            // luma size
  int pad[3];          // border width of each plane
};

struct Pic {
  bool used = false;
  int w = 0, h = 0, pad_y = 0, pad_c = 0, stride_y = 0, stride_c = 0, rows_y = 0, rows_c = 0;
  size_t bytes = 0;
  uint8_t *buf[2] = {nullptr, nullptr};   // buf[cur] holds the picture; the other is the filter target
  int cur = 0;
  PicView view(int which) const;
  PicView view() const { return view(cur); }
};

struct DevBuf {   // grow-only device scratch
  void *p = nullptr; size_t cap = 0;
  void *reserve(size_t bytes);
  void release();
};
struct PinBuf {   // grow-only pinned host scratch
  void *p = nullptr; size_t cap = 0;
  void *reserve(size_t bytes);
  void release();
};

struct Context {
  bool ready = false;
  int device = -1;
  int sm_count = 0;
  cudaStream_t stream = nullptr;
  uint64_t launches = 0;
  char err[512] = {0};
  Pic pics[256];
  // side information of the current frame, packed (2 bytes per 4x4 unit)
  DevBuf blockinfo_raw, blockinfo_packed;
  int bi_w = 0, bi_h = 0;
  DevBuf sb_bits, tasks, results, pool, cbp, ssd, refs;
  PinBuf h_tasks, h_misc;
  // sad tables: level 0..3 = block size 8,16,32,64
  DevBuf sadtab[4];
  size_t sadtab_bytes[4] = {0, 0, 0, 0};
  int sad_range = 0, sad_w = 0, sad_h = 0;
  DevBuf centres;
};

Context &ctx();
int fail(const char *fmt, ...);
Pic *get_pic(thorcuda_pic h);

#define THOR_CUDA(call)                                                                   \
  do {                                                                                    \
    cudaError_t e_ = (call);                                                              \
    if (e_ != cudaSuccess) return thor::fail("%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
  } while (0)

#define THOR_NEED_CTX()                                                                   \
  do { if (!thor::ctx().ready) thor::die("libthorcuda: thorcuda_init() was not called or failed (no CPU fallback)"); } while (0)

// counts the launch and checks the launch status
#define THOR_LAUNCHED()                                                                   \
  do { thor::ctx().launches++; THOR_CUDA(cudaGetLastError()); } while (0)

[[noreturn]] void die(const char *msg);   // fatalerror() convention: message on stderr, abort()

// reference tables living in constant memory (defined in tables.cu)
// DCT bases: c_dct4[k*4+n] ... c_dct32[k*32+n]; zig-zag scan positions; deblock thresholds.
}  // namespace thor
