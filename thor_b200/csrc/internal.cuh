// libthorcuda internals: context, device pictures, staging, task lists, launch bookkeeping.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>
#include "../../include/thorcuda.h"

namespace thor {

// Device view of a picture (geometry: reference create_yuv_frame, common/common_frame.c:324-351).
struct PicView {
  uint8_t *p[3];       // visible origin of Y,U,V
  int stride[3];       // bytes per row (U and V share stride_c)
  int w, h;            // luma size
  int pad[3];          // border width of each plane
  int reserved;
};                     // 64 bytes
static_assert(sizeof(PicView) == 64, "PicView is mirrored in a device table");

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

struct Pic {
  bool used = false;
  int w = 0, h = 0, pad_y = 0, pad_c = 0, stride_y = 0, stride_c = 0, rows_y = 0, rows_c = 0;
  size_t bytes = 0;
  uint8_t *buf[2] = {nullptr, nullptr};   // buf[cur] holds the picture; the other is the filter target
  int cur = 0;
  uint16_t *info = nullptr;               // packed side information (2 bytes per 4x4 luma unit)
  bool info_ok = false;
  uint32_t gen = 0;                       // bumped at every create: resident lists remember it (list_stale)
  PicView view(int which) const;
  PicView view() const { return view(cur); }
};

// A resident task list (thorcuda_list_create).
struct List {
  bool used = false;
  int kind = -1, ntasks = 0, npics = 0, ntiles = 0, nbig = 0;
  int cost_w = 0, cost_h = 0;      // cost lists: the largest block (sizes the kernel's window buffer)
  std::vector<int> pics;                  // handles
  std::vector<uint32_t> pic_gens;         // generation of each at list creation
  uint8_t *blob = nullptr;                // one device allocation: tasks | tiles | map
  size_t off_tiles = 0, off_map = 0, blob_bytes = 0;
  // TB lists
  size_t pool_entries = 0;
  int16_t *pool = nullptr;                // quantised coefficients
  uint8_t *cbp = nullptr;
  uint32_t *ssd = nullptr;
  uint32_t *bits = nullptr;               // write_coeff bit counts (bits.cu)
  int ntb[5] = {0, 0, 0, 0, 0};           // tasks per transform size 4,8,16,32,64 (tasks are sorted by size)
  int tb_first[5] = {0, 0, 0, 0, 0};
  uint32_t *order = nullptr;              // device: sorted position -> task index
  // COST lists
  thorcuda_cost_result *results = nullptr;   // the buffer the LAST cost call wrote (one of res_buf)
  thorcuda_cost_result *res_buf[2] = {nullptr, nullptr};
  int res_next = 0;
  // Host copies of a list's outputs run on the copy stream (d2h_late): they drain behind the big table fetches
  // without holding up the kernels queued after them.  Before anything overwrites or frees a buffer with a copy in
  // flight, the main stream is fenced (list_fence).  Cost results alternate between two buffers so that the usual
  // fasthalf -> fastquarter pair never needs the fence.
  bool late_pool = false, late_res[2] = {false, false};
  const void *tasks() const { return blob; }
  const void *tiles() const { return blob + off_tiles; }
  const uint16_t *map() const { return (const uint16_t *)(blob + off_map); }
};

struct StageSlot { PinBuf buf; cudaEvent_t ev = nullptr; bool pending = false; };

struct SadSlot {
  bool used = false;
  int w = 0, h = 0, range = 0;
  void *tab[4] = {nullptr, nullptr, nullptr, nullptr};   // levels 8,16,32,64
  size_t bytes[4] = {0, 0, 0, 0};
  size_t cap[4] = {0, 0, 0, 0};
};

struct Context {
  bool ready = false;
  int device = -1;
  int sm_count = 0;
  cudaStream_t stream = nullptr;
  // large device->host fetches (SAD tables) run on their own stream so that kernels queued after them keep the SMs
  // busy while the copy engine drains: see d2h_bulk()
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t ev_main = nullptr, ev_copy = nullptr;
  bool copy_pending = false;
  uint64_t launches = 0;
  char err[512] = {0};
  Pic pics[THORCUDA_MAX_PICS];
  PicView *d_pics = nullptr;              // device mirror of view() of every picture, indexed by handle
  List lists[256];
  StageSlot stage[8];
  int stage_next = 0;
  std::vector<std::pair<uintptr_t, size_t>> pinned;   // thorcuda_host_alloc ranges
  DevBuf raw_info, scratch, flush, lfjobs, sadjobs, centres;
  cudaEvent_t tm0[64], tm1[64];
  bool tm_used[64] = {false};
  cudaEvent_t fn_main[16], fn_copy[16];   // thorcuda_fence_record / _wait
  bool fn_used[16] = {false}, fn_made[16] = {false}, fn_copy_set[16] = {false};
  SadSlot sad[THORCUDA_SAD_SLOTS];
};

Context &ctx();
int fail(const char *fmt, ...);
Pic *get_pic(thorcuda_pic h);
List *get_list(thorcuda_list h);
[[noreturn]] void die(const char *msg);   // fatalerror() convention: message on stderr, abort()

// Queue a host->device copy of a caller buffer: through the pinned staging ring (the caller may reuse
// src on return) unless src lies in thorcuda_host_alloc memory (then it is the DMA source itself).
int h2d(void *dst, const void *src, size_t bytes);
// Pinned staging memory the caller fills itself, then queues with stage_commit_copy().
void *stage_acquire(size_t bytes);
int stage_commit_copy(void *dst, const void *staged, size_t bytes);
int d2h(void *dst, const void *src, size_t bytes);
// device->host copy on the copy stream, ordered after everything queued so far on the main stream; the host data is
// valid after thorcuda_sync().  Whoever overwrites or frees `src` later must call copies_fence() first.
int d2h_bulk(void *dst, const void *src, size_t bytes);
int copies_fence();    // make the main stream wait for the bulk copies queued so far
struct List;
int list_fence(List *L);
// non-zero (and an error) if a picture the list was validated against has been destroyed or recreated since: the kernels
// address pictures through the device table and never clamp, so such a list must not be launched
int list_stale(const List *L);
// transform.cu: the transform-block kernel over a sub-range of every size class of a TB list (nullptr: everything)
int run_tb_classes(List *L, int decode, bool want_ssd, const int *class_first, const int *class_count);   // ... if the list has output copies in flight
// refresh the device mirror entry of picture h
int sync_pic_view(thorcuda_pic h);
// one stand-alone forward / inverse transform on device buffers (transform.cu), queued on the stream
int tb_raw(int size, int fast, int inverse, const int16_t *d_in, int16_t *d_out);
void compat_reset();   // compat.cu: drop the per-block API's scratch pictures (thorcuda_shutdown)
void interp_reset();   // tinterp.cu: drop the interpolation scratch (thorcuda_shutdown)
void bits_reset();     // bits.cu: the scan tables are uploaded again after a shutdown
void enc_reset();      // encoder.cu: the frame encoder's device buffers

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

__device__ __forceinline__ int clampi(int v, int lo, int hi) { return min(max(v, lo), hi); }
__device__ __forceinline__ int clip255(int v) { return min(max(v, 0), 255); }

// reference common/common_block.c:78-83
static const uint8_t h_chroma_qp[52] = {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16, 17, 18, 19, 20, 21, 22, 23, 24, 25,
                                        26, 27, 28, 29, 29, 30, 31, 32, 33, 33, 34, 34, 35, 35, 36, 36, 37, 37, 38, 39, 40, 41, 42, 43, 44, 45};

}  // namespace thor
