// libthorcuda: batched intra prediction.
//
// Restates reference common/intra_prediction.c:39-389: the neighbour arrays of make_top_and_left (row above, column
// left, corner; 128 at picture edges; one extra sample when up-right / down-left is available, then replication to
// 2*size) and the ten modes of get_intra_prediction.  One warp per block: the neighbours are gathered into shared
// memory, smoothed as the mode asks ([1 2 1] over size or 2*size samples, [1 2 2 2 1] for planar), and every lane
// then produces 4 adjacent samples per step and stores them as one word.
//
// Blocks of one call must not depend on each other (the caller batches by wavefront: a block may enter a call once
// the blocks that own its neighbour samples have been reconstructed); neighbours are read before anything is
// written, so src == dst is allowed.
#include <string.h>
#include "internal.cuh"

namespace thor {

constexpr int IW = 4;                 // warps per CTA

struct IntraParams {
  PicView pics[8];
  const thorcuda_intra_task *tasks;
  int ntasks;
};

struct IntraSmem {
  uint8_t top[128], left[128];        // raw neighbours, 2 * size each
  uint8_t tf[128], lf[128];           // smoothed neighbours (uint8), or int16[64] each for the planar mode
};

__device__ __forceinline__ uint8_t smooth121(const uint8_t *in, int j, int len) {
  return (uint8_t)((in[max(j - 1, 0)] + 2 * in[j] + in[min(j + 1, len - 1)] + 2) >> 2);
}

__device__ __forceinline__ int avg_dn(int a, int b) { return (a + b) >> 1; }

__global__ void __launch_bounds__(IW * 32) k_intra(const IntraParams P) {
  __shared__ __align__(16) IntraSmem sm[IW];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ti = blockIdx.x * IW + warp;
  if (ti >= P.ntasks) return;
  const thorcuda_intra_task t = P.tasks[ti];
  IntraSmem &S = sm[warp];
  const int s = t.size, pl = t.plane, x = t.x, y = t.y, mode = t.mode;
  const PicView &vs = P.pics[t.src], &vd = P.pics[t.dst];
  const int sstride = vs.stride[pl];
  const uint8_t *src = vs.p[pl] + (size_t)y * sstride + x;
  const int toplen = (t.flags & 1) ? s + 1 : s, leftlen = (t.flags & 2) ? s + 1 : s;

  // ---- neighbours (intra_prediction.c:57-143) ----
  for (int k = lane; k < 2 * s; k += 32) {
    S.top[k] = y == 0 ? 128 : src[-sstride + min(k, toplen - 1)];
    S.left[k] = x == 0 ? 128 : src[(ptrdiff_t)min(k, leftlen - 1) * sstride - 1];
  }
  int tl;
  if (y == 0) tl = x == 0 ? 128 : src[-1];                       // = left[0]
  else tl = t.cb_x > 0 ? src[-sstride - 1] : src[-sstride];      // the corner only when the CODING block is off the left edge
  __syncwarp();

  // ---- smoothing the mode needs ----
  int tlf = 0, dc = 0;
  int16_t *tF = reinterpret_cast<int16_t *>(S.tf), *lF = reinterpret_cast<int16_t *>(S.lf);
  if (mode == 4 || mode == 7 || mode == 8) {
    for (int j = lane; j < s; j += 32) { S.tf[j] = smooth121(S.top, j, s); S.lf[j] = smooth121(S.left, j, s); }
    tlf = (2 * tl + S.left[0] + S.top[0] + 2) >> 2;
  } else if (mode == 5 || mode == 6) {
    for (int j = lane; j < 2 * s; j += 32) S.tf[j] = smooth121(S.top, j, 2 * s);
  } else if (mode == 9) {
    for (int j = lane; j < 2 * s; j += 32) S.lf[j] = smooth121(S.left, j, 2 * s);
  } else if (mode == 1) {
    for (int j = lane; j < s; j += 32) {
      const int a = max(j - 2, 0), b = max(j - 1, 0), c = min(j + 1, s - 1), d = min(j + 2, s - 1);
      tF[j] = (int16_t)(S.top[a] + 2 * S.top[b] + 2 * S.top[j] + 2 * S.top[c] + S.top[d]);
      lF[j] = (int16_t)(S.left[a] + 2 * S.left[b] + 2 * S.left[j] + 2 * S.left[c] + S.left[d]);
    }
    tlf = S.left[1] + 2 * S.left[0] + 2 * tl + 2 * S.top[0] + S.top[1];
  } else if (mode != 2 && mode != 3) {
    // DC (and unknown modes, :387): at a picture edge the side that exists counts twice (:367)
    const uint8_t *a = x != 0 ? S.left : S.top, *b = y != 0 ? S.top : S.left;
    int sum = 0;
    for (int j = lane; j < s; j += 32) sum += a[j] + b[j];
#pragma unroll
    for (int m = 16; m > 0; m >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, m);
    dc = (sum + s) / (2 * s);
  }
  __syncwarp();

  // ---- samples: 4 per lane and step ----
  const int qpr = s >> 2, dstride = vd.stride[pl];
  uint8_t *dst = vd.p[pl] + (size_t)y * dstride + x;
  for (int q = lane; q < s * qpr; q += 32) {
    const int i = q / qpr, j0 = 4 * (q - i * qpr);
    uint32_t out = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const int j = j0 + k;
      int v;
      switch (mode) {
        case 1: v = min(max(lF[i] + tF[j] - tlf + 4, 0) >> 3, 255); break;     // C division then clip: negatives clip to 0 either way
        case 2: v = S.left[i]; break;
        case 3: v = S.top[j]; break;
        case 4: { const int d = i - j; v = d > 0 ? S.lf[d - 1] : d == 0 ? tlf : S.tf[-d - 1]; break; }
        case 5: v = S.tf[i + j + 1]; break;
        case 6: { const int d = i + 2 * j; v = (d & 1) ? S.tf[(d + 1) >> 1] : avg_dn(S.tf[d >> 1], S.tf[(d >> 1) + 1]); break; }
        case 7: {
          const int d = i - 2 * j, h = (-d) >> 1;
          v = d > 1 ? S.lf[d - 2] : d == 1 ? tlf : d == 0 ? avg_dn(tlf, S.tf[0]) : (d & 1) ? S.tf[h] : avg_dn(S.tf[h], S.tf[h - 1]);
          break;
        }
        case 8: {
          const int d = 2 * i - j, h = d >> 1;
          v = d < -1 ? S.tf[-d - 2] : d == -1 ? tlf : d == 0 ? avg_dn(tlf, S.lf[0]) : (d & 1) ? S.lf[h] : avg_dn(S.lf[h], S.lf[h - 1]);
          break;
        }
        case 9: { const int d = 2 * i + j; v = (d & 1) ? S.lf[(d + 1) >> 1] : avg_dn(S.lf[d >> 1], S.lf[(d >> 1) + 1]); break; }
        default: v = dc; break;
      }
      out |= (uint32_t)v << (8 * k);
    }
    *reinterpret_cast<uint32_t *>(dst + (size_t)i * dstride + j0) = out;
  }
}

}  // namespace thor

using namespace thor;

extern "C" {

// validates the tasks, fills the picture views and uploads the tasks to the library's scratch; P.tasks then points at them
static int intra_prepare(const thorcuda_pic *pics, int npics, const thorcuda_intra_task *tasks, int ntasks, IntraParams &P) {
  Context &c = ctx();
  if (npics <= 0 || npics > 8) return fail("intra: npics %d out of range (1..8)", npics);
  memset(&P, 0, sizeof(P));
  const Pic *pp[8];
  for (int i = 0; i < npics; i++) {
    Pic *p = get_pic(pics[i]);
    if (!p) return fail("intra: bad picture handle %d", pics[i]);
    pp[i] = p;
    P.pics[i] = p->view();
  }
  for (int i = 0; i < ntasks; i++) {
    const thorcuda_intra_task &t = tasks[i];
    if (t.src >= npics || t.dst >= npics) return fail("intra: task %d names a picture outside pics[%d]", i, npics);
    const Pic *s = pp[t.src], *d = pp[t.dst];
    if (t.plane > 2 || s->w != d->w || s->h != d->h) return fail("intra: task %d: plane %d / picture size mismatch", i, t.plane);
    const int pw = t.plane ? s->w / 2 : s->w, ph = t.plane ? s->h / 2 : s->h;
    const bool size_ok = t.size == 4 || t.size == 8 || t.size == 16 || t.size == 32 || t.size == 64;
    if (!size_ok || (t.x & 3) || t.x + t.size > pw || t.y + t.size > ph || t.cb_x > t.x)
      return fail("intra: task %d has an invalid block (%d,%d size %d, plane %d)", i, t.x, t.y, t.size, t.plane);
    // the extra up-right / down-left sample must exist in the picture where it is read (a block in the first row /
    // column never reads it: the reference's sub-block rule sets the flag there too, :106-110)
    if ((t.flags & 1) && t.y > 0 && t.x + t.size >= pw) return fail("intra: task %d claims an up-right sample outside the picture", i);
    if ((t.flags & 2) && t.x > 0 && t.y + t.size >= ph) return fail("intra: task %d claims a down-left sample outside the picture", i);
  }
  const size_t tbytes = (size_t)ntasks * sizeof(thorcuda_intra_task);
  uint8_t *d = (uint8_t *)c.scratch.reserve(tbytes + 64);
  if (!d) return fail("intra: device allocation failed");
  if (h2d(d, tasks, tbytes)) return -1;
  P.tasks = (const thorcuda_intra_task *)d;
  P.ntasks = ntasks;
  return 0;
}

int thorcuda_intra_predict_blocks(const thorcuda_pic *pics, int npics, const thorcuda_intra_task *tasks, int ntasks) {
  THOR_NEED_CTX();
  if (ntasks <= 0) return 0;
  IntraParams P;
  if (intra_prepare(pics, npics, tasks, ntasks, P)) return -1;
  k_intra<<<(ntasks + IW - 1) / IW, IW * 32, 0, ctx().stream>>>(P);
  THOR_LAUNCHED();
  return 0;
}

int thorcuda_intra_reconstruct_levels(const thorcuda_pic *pics, int npics, const thorcuda_intra_task *tasks, const int *intra_first,
                                      const thorcuda_tb_task *tb, const int *tb_first, int nlevels, const int16_t *host_coeffq) {
  THOR_NEED_CTX();
  if (nlevels <= 0) return 0;
  if (!intra_first || !tb_first) return fail("intra levels: NULL level table");
  if (intra_first[0] != 0 || tb_first[0] != 0) return fail("intra levels: the level tables must start at 0");
  for (int l = 0; l < nlevels; l++)
    if (intra_first[l + 1] < intra_first[l] || tb_first[l + 1] < tb_first[l]) return fail("intra levels: level %d has a negative size", l);
  const int nintra = intra_first[nlevels], ntb = tb_first[nlevels];
  if (ntb > 0 && (!tb || !host_coeffq)) return fail("intra levels: transform blocks without tasks or coefficients");
  IntraParams P;
  memset(&P, 0, sizeof(P));
  if (nintra > 0 && intra_prepare(pics, npics, tasks, nintra, P)) return -1;
  // ONE resident transform-block list for the whole frame.  List creation sorts by size class but keeps the task order
  // inside a class, and the caller's tasks are in level order: a level is a contiguous range of every class.
  thorcuda_list hl = -1;
  List *L = nullptr;
  if (ntb > 0) {
    hl = thorcuda_list_create(THORCUDA_LIST_TB, pics, npics, tb, ntb);
    if (hl < 0) return -1;
    L = get_list(hl);
    if (h2d(L->pool, host_coeffq, L->pool_entries * sizeof(int16_t))) { thorcuda_list_destroy(hl); return -1; }
  }
  int done[5] = {0, 0, 0, 0, 0}, rc = 0;
  const thorcuda_intra_task *dtasks = P.tasks;
  for (int l = 0; l < nlevels && !rc; l++) {
    const int ni = intra_first[l + 1] - intra_first[l];
    if (ni > 0) {
      P.tasks = dtasks + intra_first[l];
      P.ntasks = ni;
      k_intra<<<(ni + IW - 1) / IW, IW * 32, 0, ctx().stream>>>(P);
      ctx().launches++;
      if (cudaGetLastError() != cudaSuccess) rc = fail("intra levels: launch failed");
    }
    if (L && tb_first[l + 1] > tb_first[l]) {
      int first[5], count[5] = {0, 0, 0, 0, 0};
      for (int i = tb_first[l]; i < tb_first[l + 1]; i++) {
        const int sz = tb[i].size;
        count[sz == 4 ? 0 : sz == 8 ? 1 : sz == 16 ? 2 : sz == 32 ? 3 : 4]++;
      }
      for (int k = 0; k < 5; k++) { first[k] = L->tb_first[k] + done[k]; done[k] += count[k]; }
      if (run_tb_classes(L, 1, false, first, count)) rc = -1;
    }
  }
  if (hl >= 0) thorcuda_list_destroy(hl);          // stream-ordered: the queued launches finish first
  return rc;
}

}  // extern "C"
