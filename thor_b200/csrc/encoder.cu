// libthorcuda: the frame encoder -- every block decision of Thorenc's encode_frame on the device.
//
// Replaces the superblock loop of reference enc/encode_frame.c:112-148 and everything it drives in
// enc/encode_block.c:799-3033 (motion_estimate, search_*_prediction_params, encode_block with its trial encodes,
// check_early_skip_block, mode_decision_rdo, process_block) together with the neighbour state those read
// (find_block_contexts common/common_block.c:158-178, get_mv_pred / get_mv_skip / get_mv_merge
// common/inter_prediction.c:182-600) and the bit counts of write_block (enc/write_bits.c:364-650).  The device code is
// csrc/enc/enc_core.h; this file is the kernel around it and the C entry point.
//
// Mapping.  A frame's superblocks depend on each other exactly as SURVEY App. A.2 says: a block reads the decisions and
// the reconstruction of its left, up-left, up and up-right neighbours, so superblock (r, c) can start when (r, c-1) and
// (r-1, c+1) are finished -- a wavefront with a lag of two superblocks per row.  One CTA of 256 threads walks one
// superblock ROW; rows are handed out by a ticket (the row above is always already running, so the spin-wait below cannot
// deadlock however the CTAs are scheduled), a row publishes "superblocks finished" with a release store after a
// __threadfence() of all its threads, the row below polls that counter with acquire loads and reads what the upper row
// wrote (reconstruction, 16-byte neighbour units) with ld.global.cg, i.e. from L2.  Inside a superblock the CTA executes
// the reference's decision sequence with uniform control flow and spreads every pixel loop over its threads; all block
// state lives in ~100 KB of shared memory (original, prediction, trial and best reconstruction, coefficients, transform
// scratch, candidate lists).  At 1080p the wavefront keeps at most min(17, 30/2) = 15 CTAs busy: one frame cannot fill 148
// SMs; independent streams / segments are what fills the device (bench.py), exactly as SURVEY 8e lays out.
#include <string.h>
#include <algorithm>
#include "internal.cuh"
#include "enc/enc_core.h"

namespace thor {

using namespace thorenc;

__device__ __forceinline__ int ld_acquire(const int *p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release(int *p, int v) { asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }

// Scheduling.  Superblock (r, c) may start when (r, c-1) and (r-1, min(c+1, last)) are finished.  Nothing but `progress`
// carries over from one superblock to the next (the candidate lists and best_ref are reset per superblock,
// enc/encode_frame.c:117-121), so a CTA is not tied to a row: every CTA repeatedly claims the next superblock of the lowest
// row that is ready and not being worked on (one lock word per row).  With one CTA per row about half of each CTA's life
// is spent waiting for the row above while it holds an SM; pulling work lets fewer CTAs (about rows/2 + a few) finish the
// frame in the same number of wavefront steps, which is what lets twice as many independent frames share the device.
// The lowest unfinished row is always either ready or being worked on, so any number of resident CTAs >= 1 makes progress.
#ifndef THOR_ENC_MIN_CTAS
#define THOR_ENC_MIN_CTAS (NT <= 128 ? 2 : 1)
#endif
__global__ void __launch_bounds__(NT, THOR_ENC_MIN_CTAS) k_encode_frame(const thor_enc_frame F_arg, int *busy) {
  extern __shared__ __align__(16) uint8_t smem[];
  SbMem &S = *reinterpret_cast<SbMem *>(smem);
  __shared__ int s_row, s_col;
  Grp g;
  g.tid = threadIdx.x;
  g.red = S.red;
  g.flip = 0;
  for (int i = threadIdx.x; i < (int)(sizeof(thor_enc_frame) / 4); i += NT) reinterpret_cast<int *>(&S.frame)[i] = reinterpret_cast<const int *>(&F_arg)[i];
  __syncthreads();
  const thor_enc_frame &F = S.frame;
  const int nsbx = (F.width + 63) / 64, nsby = (F.height + 63) / 64;
  build_tables(g, S);
  for (;;) {
    if (threadIdx.x == 0) {
      PROF(PF_WAIT);
      int row = -1, col = 0;
      for (;;) {
        bool all_done = true;
        for (int r = 0; r < nsby && row < 0; r++) {
          const int done = ld_acquire(F.progress + r);
          if (done >= nsbx) continue;
          all_done = false;
          if (r > 0 && ld_acquire(F.progress + r - 1) < min(done + 2, nsbx)) continue;   // waits for the row above (a lower row may still be ready)
          if (atomicCAS(busy + r, 0, 1) == 0) {
            __threadfence();
            col = ld_acquire(F.progress + r);                // re-read under the lock
            if (col < nsbx && (r == 0 || ld_acquire(F.progress + r - 1) >= min(col + 2, nsbx))) row = r;
            else atomicExch(busy + r, 0);
          }
        }
        if (row >= 0 || all_done) break;
        __nanosleep(500);
      }
      s_row = row;
      s_col = col;
    }
    __syncthreads();
    const int row = s_row, l = s_col;
    if (row < 0) return;
    Enc E{g, S, F, SbState(), F.nodes + (size_t)(row * nsbx + l) * THOR_ENC_NODES, F.coef + (row * nsbx + l), l * 64, row * 64};
    encode_superblock(E);
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
      st_release(F.progress + row, l + 1);
      __threadfence();
      atomicExch(busy + row, 0);
    }
  }
}

// neighbour units -> the loop filter's packed side information (filters.cu: 2 bytes per 4x4 unit:
// bits 0-2 mode | 3-5 cbp y,u,v | 6-7 log2(size)-3 | 8 tb_split | 9-10 pb_part | 11 some |vector component| >= 4)
__global__ void k_units_to_info(const thor_enc_unit *__restrict__ units, uint16_t *__restrict__ info, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int4 raw = reinterpret_cast<const int4 *>(units)[i];
  thor_enc_unit u;
  *reinterpret_cast<int4 *>(&u) = raw;
  const int lg = u.size >= 64 ? 3 : u.size >= 32 ? 2 : u.size >= 16 ? 1 : 0;
  const bool big = abs((int)u.mv0x) >= 4 || abs((int)u.mv0y) >= 4 || abs((int)u.mv1x) >= 4 || abs((int)u.mv1y) >= 4;
  info[i] = (uint16_t)((u.mode & 7) | ((u.cbp & 7) << 3) | (lg << 6) | ((u.flags & 1) << 8) | (((u.flags >> 1) & 3) << 9) | (big ? 2048 : 0));
}

static DevBuf g_units, g_nodes, g_coef, g_sync;
static bool g_attr_set = false;
void enc_reset() { g_units.release(); g_nodes.release(); g_coef.release(); g_sync.release(); g_attr_set = false; }

static thor_enc_pic pic_of(const Pic *p, int frame_num) {
  const PicView v = p->view();
  thor_enc_pic o;
  o.y = v.p[0]; o.u = v.p[1]; o.v = v.p[2];
  o.stride_y = v.stride[0]; o.stride_c = v.stride[1];
  o.frame_num = frame_num; o.reserved = 0;
  return o;
}

}  // namespace thor

using namespace thor;

// cycle accounting of the superblock encoder (only in -DTHOR_ENC_PROF builds; otherwise all zero): out[PF_*], 32 entries
extern "C" int thorcuda_enc_profile(unsigned long long *out32) {
#ifdef THOR_ENC_PROF
  THOR_CUDA(cudaDeviceSynchronize());
  THOR_CUDA(cudaMemcpyFromSymbol(out32, thorenc::g_enc_prof, 32 * sizeof(unsigned long long)));
#else
  memset(out32, 0, 32 * sizeof(unsigned long long));
#endif
  return 0;
}

extern "C" int thorcuda_enc_frame(const thorcuda_enc_params *p, thor_enc_node *host_nodes, thor_enc_coef *host_coef) {
  THOR_NEED_CTX();
  Context &c = ctx();
  if (!p || !host_nodes || !host_coef) return fail("enc_frame: null argument");
  if (p->width <= 0 || p->height <= 0 || (p->width & 7) || (p->height & 7)) return fail("enc_frame: picture size %dx%d is not a multiple of 8", p->width, p->height);
  if (p->frame_type < 0 || p->frame_type > 2) return fail("enc_frame: frame type %d", p->frame_type);
  if (p->qp < 0 || p->qp > 51 || p->qp - p->max_delta_qp < 0 || p->qp + p->max_delta_qp > 51) return fail("enc_frame: qp %d (+-%d) outside 0..51", p->qp, p->max_delta_qp);
  if (p->max_delta_qp && p->delta_qp_step <= 0) return fail("enc_frame: delta_qp_step %d", p->delta_qp_step);
  if (p->num_ref < 0 || p->num_ref > THOR_ENC_MAX_REF || (p->frame_type != 0 && p->num_ref == 0)) return fail("enc_frame: %d reference pictures", p->num_ref);
  if (p->encoder_speed < 0 || p->encoder_speed > 2) return fail("enc_frame: encoder_speed %d", p->encoder_speed);
  Pic *org = get_pic(p->orig), *rec = get_pic(p->rec);
  if (!org || !rec) return fail("enc_frame: bad picture handle");
  if (org->w != p->width || org->h != p->height || rec->w != p->width || rec->h != p->height) return fail("enc_frame: orig / rec size differs from %dx%d", p->width, p->height);
  thor_enc_frame F;
  memset(&F, 0, sizeof(F));
  for (int r = 0; r < (p->frame_type == 0 ? 0 : p->num_ref); r++) {
    const Pic *rp = get_pic(p->ref[r]);
    if (!rp) return fail("enc_frame: bad reference handle %d", p->ref[r]);
    // clip_mv keeps a block within 80 samples of the picture; 64 + 3 + 3 more are read by the filters and widesad: the
    // reference relies on 96 / 48 samples of border (enc/mainenc.c:165) and so does this code
    if (rp->w != p->width || rp->h != p->height || rp->pad_y < 96 || rp->pad_c < 48) return fail("enc_frame: reference %d needs the picture size and a 96 / 48 border", r);
    F.ref[r] = pic_of(rp, p->ref_frame_num[r]);
  }
  F.width = p->width; F.height = p->height; F.frame_type = p->frame_type; F.qp = p->qp;
  F.num_ref = p->frame_type == 0 ? 0 : p->num_ref;
  F.interp_ref = p->interp_ref; F.num_intra_modes = p->num_intra_modes; F.frame_num = p->frame_num;
  F.encoder_speed = p->encoder_speed; F.enable_bipred = p->enable_bipred; F.enable_tb_split = p->enable_tb_split;
  F.enable_pb_split = p->enable_pb_split; F.intra_rdo = p->intra_rdo; F.max_delta_qp = p->max_delta_qp;
  F.delta_qp_step = p->delta_qp_step; F.use_block_contexts = p->use_block_contexts;
  F.early_skip_thr = p->early_skip_thr; F.lambda = p->lambda;
  F.orig = pic_of(org, p->frame_num);
  F.rec = pic_of(rec, p->frame_num);
  const int nsbx = (p->width + 63) / 64, nsby = (p->height + 63) / 64, nsb = nsbx * nsby;
  const size_t nunits = (size_t)(p->width / 4) * (p->height / 4);
  const size_t node_bytes = (size_t)nsb * THOR_ENC_NODES * sizeof(thor_enc_node), coef_bytes = (size_t)nsb * sizeof(thor_enc_coef);
  // the previous frame's result copies run on the main stream, so growing / reusing these buffers is ordered behind them
  F.units = (thor_enc_unit *)g_units.reserve(nunits * sizeof(thor_enc_unit));
  F.nodes = (thor_enc_node *)g_nodes.reserve(node_bytes);
  F.coef = (thor_enc_coef *)g_coef.reserve(coef_bytes);
  int *sync = (int *)g_sync.reserve((size_t)(2 * nsby) * sizeof(int));      // row locks | progress
  if (!F.units || !F.nodes || !F.coef || !sync) return fail("enc_frame: device allocation failed");
  F.progress = sync + nsby;
  THOR_CUDA(cudaMemsetAsync(F.units, 0, nunits * sizeof(thor_enc_unit), c.stream));   // encode_frame's memset of deblock_data (:74)
  THOR_CUDA(cudaMemsetAsync(F.nodes, 0, node_bytes, c.stream));
  THOR_CUDA(cudaMemsetAsync(sync, 0, (size_t)(2 * nsby) * sizeof(int), c.stream));
  if (!g_attr_set) {
    THOR_CUDA(cudaFuncSetAttribute(k_encode_frame, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(SbMem)));
    g_attr_set = true;
  }
  // CTAs per frame: by default as many as the wavefront can ever keep busy, min(rows, ceil(cols / 2)) -- the shortest frame
  // time; THORCUDA_ENC_CTAS lowers it when many frames share the device (bench.py: more streams per GPU at nearly the same
  // time per frame)
  int ctas = std::min(nsby, (nsbx + 1) / 2);
  if (const char *e = getenv("THORCUDA_ENC_CTAS")) ctas = std::max(1, std::min(atoi(e), nsby));
  k_encode_frame<<<ctas, NT, sizeof(SbMem), c.stream>>>(F, sync);
  THOR_LAUNCHED();
  if (!rec->info) THOR_CUDA(cudaMalloc((void **)&rec->info, nunits * 2));
  k_units_to_info<<<(unsigned)((nunits + 255) / 256), 256, 0, c.stream>>>(F.units, rec->info, (int)nunits);
  THOR_LAUNCHED();
  rec->info_ok = true;
  if (d2h(host_nodes, F.nodes, node_bytes)) return -1;
  if (d2h(host_coef, F.coef, coef_bytes)) return -1;
  return 0;
}
