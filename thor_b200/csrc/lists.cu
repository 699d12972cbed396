// libthorcuda: resident task lists (validation, expansion, upload).
#include <string.h>
#include <algorithm>
#include "internal.cuh"
#include "lists.cuh"

using namespace thor;

namespace thor {

static int log2size(int s) { return s == 4 ? 0 : s == 8 ? 1 : s == 16 ? 2 : s == 32 ? 3 : s == 64 ? 4 : -1; }

static int check_pred(const Pic *const *P, const thorcuda_pic *handles, int npics, const thorcuda_pred_task *tasks, int ntasks, std::vector<PredTile> &tiles) {
  for (int i = 0; i < ntasks; i++) {
    const thorcuda_pred_task &t = tasks[i];
    if (t.dst >= npics) return fail("predict: task %d writes picture %d of %d", i, t.dst, npics);
    const Pic *dst = P[t.dst];
    if (t.w < 4 || t.h < 4 || t.w > 64 || t.h > 64 || (t.w & 3) || (t.h & 3) || t.x + t.w > dst->w || t.y + t.h > dst->h ||
        (t.x & 3) || (t.y & 3))
      return fail("predict: task %d has an invalid block (%d,%d %dx%d)", i, t.x, t.y, t.w, t.h);
    for (int pass = 0; pass < 2; pass++) {
      const int rix = pass ? t.ref1 : t.ref0;
      if (pass && rix < 0) break;
      if (rix < 0 || rix >= npics) return fail("predict: task %d references picture %d of %d", i, rix, npics);
      const Pic *r = P[rix];
      if (r->w != dst->w || r->h != dst->h) return fail("predict: task %d: reference size mismatch", i);
      if (rix == t.dst) return fail("predict: task %d predicts a picture from itself", i);
      const int sgn = pass ? t.sign1 : t.sign0;
      const int mx = sgn ? -(pass ? t.mv1x : t.mv0x) : (pass ? t.mv1x : t.mv0x);
      const int my = sgn ? -(pass ? t.mv1y : t.mv0y) : (pass ? t.mv1y : t.mv0y);
      const int lx = t.x + (mx >> 2), ly = t.y + (my >> 2);
      // filter reach: 2 left/up, 3 right/down (luma); the padded picture must contain the window
      if (lx - 2 < -r->pad_y || ly - 2 < -r->pad_y || lx + t.w + 3 > r->w + r->pad_y || ly + t.h + 3 > r->h + r->pad_y)
        return fail("predict: task %d reads outside the padded reference (mv %d,%d)", i, mx, my);
      const int cx = (t.x >> 1) + (mx >> 3), cy = (t.y >> 1) + (my >> 3);
      if (cx - 1 < -r->pad_c || cy - 1 < -r->pad_c || cx + t.w / 2 + 2 > r->w / 2 + r->pad_c || cy + t.h / 2 + 2 > r->h / 2 + r->pad_c)
        return fail("predict: task %d reads outside the padded chroma reference (mv %d,%d)", i, mx, my);
    }
    for (int ty = 0; ty < (t.h + 15) / 16; ty++)
      for (int tx = 0; tx < (t.w + 15) / 16; tx++) {
        PredTile q;
        memset(&q, 0, sizeof(q));
        q.x = (uint16_t)(t.x + 16 * tx); q.y = (uint16_t)(t.y + 16 * ty);
        q.tw = (uint8_t)std::min(16, t.w - 16 * tx); q.th = (uint8_t)std::min(16, t.h - 16 * ty);
        q.dst = (uint16_t)handles[t.dst]; q.ref0 = (uint16_t)handles[t.ref0]; q.ref1 = t.ref1 >= 0 ? (uint16_t)handles[t.ref1] : (uint16_t)0xFFFF;
        q.mv0x = (int16_t)(t.sign0 ? -t.mv0x : t.mv0x); q.mv0y = (int16_t)(t.sign0 ? -t.mv0y : t.mv0y);
        q.mv1x = (int16_t)(t.sign1 ? -t.mv1x : t.mv1x); q.mv1y = (int16_t)(t.sign1 ? -t.mv1y : t.mv1y);
        q.task = (uint32_t)i;
        tiles.push_back(q);
      }
  }
  return 0;
}

static int check_tb(const Pic *const *P, int npics, const thorcuda_tb_task *tasks, int ntasks, List &L, std::vector<uint32_t> &order) {
  size_t pool = 0;
  int cnt[5] = {0, 0, 0, 0, 0};
  for (int i = 0; i < ntasks; i++) {
    const thorcuda_tb_task &t = tasks[i];
    const int lg = log2size(t.size);
    if (lg < 0) return fail("tb: task %d has transform size %d", i, t.size);
    if (t.plane > 2 || t.qp > 51) return fail("tb: task %d has plane %d qp %d", i, t.plane, t.qp);
    if (t.org >= npics || t.pred >= npics || t.rec >= npics) return fail("tb: task %d names a picture outside pics[%d]", i, npics);
    const Pic *r = P[t.rec];
    const int pw = t.plane ? r->w / 2 : r->w, ph = t.plane ? r->h / 2 : r->h;
    if ((t.x & 3) || (t.y & 3) || t.x + t.size > pw || t.y + t.size > ph)
      return fail("tb: task %d block (%d,%d size %d) is misaligned or outside plane %d", i, t.x, t.y, t.size, t.plane);
    if (P[t.org]->w != r->w || P[t.org]->h != r->h || P[t.pred]->w != r->w || P[t.pred]->h != r->h)
      return fail("tb: task %d: picture size mismatch", i);
    if (t.size >= 8 && ((t.x | t.y) & 7)) return fail("tb: task %d: blocks of size >= 8 must lie on the 8-sample grid (%d,%d)", i, t.x, t.y);
    const int q = t.size < 16 ? t.size : 16;
    if ((t.coeff_offset & 7) && t.size > 4) return fail("tb: task %d: coeff_offset must be a multiple of 8 entries", i);
    if (t.coeff_offset & 3) return fail("tb: task %d: coeff_offset must be a multiple of 4 entries", i);
    pool = std::max(pool, (size_t)t.coeff_offset + q * q);
    cnt[lg]++;
  }
  int first = 0;
  for (int k = 0; k < 5; k++) { L.tb_first[k] = first; L.ntb[k] = cnt[k]; first += cnt[k]; }
  int pos[5];
  for (int k = 0; k < 5; k++) pos[k] = L.tb_first[k];
  order.resize(ntasks);
  for (int i = 0; i < ntasks; i++) order[pos[log2size(tasks[i].size)]++] = (uint32_t)i;
  L.pool_entries = (pool + 7) & ~(size_t)7;
  return 0;
}

static int check_cost(const Pic *const *P, int npics, const thorcuda_cost_task *tasks, int ntasks) {
  for (int i = 0; i < ntasks; i++) {
    const thorcuda_cost_task &t = tasks[i];
    if (t.org >= npics || t.ref >= npics) return fail("cost: task %d names a picture outside pics[%d]", i, npics);
    const Pic *o = P[t.org], *r = P[t.ref];
    if (o->w != r->w || o->h != r->h) return fail("cost: task %d: picture size mismatch", i);
    if (t.w < 4 || t.h < 4 || t.w > 64 || t.h > 64 || (t.w & 3) || (t.h & 3) || t.x + t.w > o->w || t.y + t.h > o->h)
      return fail("cost: task %d has an invalid block (%d,%d %dx%d)", i, t.x, t.y, t.w, t.h);
    if (t.x & 3) return fail("cost: task %d: block x %d is not on the 4-sample grid", i, t.x);
    if (t.hx < -2 || t.hx > 2 || t.hy < -2 || t.hy > 2 || (t.hx & 1) || (t.hy & 1)) return fail("cost: task %d: half-pel offset (%d,%d)", i, t.hx, t.hy);
    // widest reach of any kind: 3 columns (widesad), 2 rows/columns (fasthalf); the kernel stages whole words, i.e.
    // columns -4 .. w+3, so that is what must lie inside the padded picture
    const int lx = t.x + t.ix, ly = t.y + t.iy;
    if (lx - 4 < -r->pad_y || ly - 3 < -r->pad_y || lx + t.w + 4 > r->w + r->pad_y || ly + t.h + 3 > r->h + r->pad_y)
      return fail("cost: task %d reads outside the padded reference (%d,%d)", i, t.ix, t.iy);
  }
  return 0;
}

}  // namespace thor

extern "C" {

thorcuda_list thorcuda_list_create(int kind, const thorcuda_pic *pics, int npics, const void *tasks, int ntasks) {
  THOR_NEED_CTX();
  Context &c = ctx();
  if (kind < 0 || kind > 2) return fail("list: unknown kind %d", kind);
  if (npics <= 0 || npics > 128) return fail("list: npics %d out of range (1..128)", npics);
  if (ntasks <= 0 || !tasks) return fail("list: empty task list");
  const Pic *P[128];
  for (int i = 0; i < npics; i++) {
    P[i] = get_pic(pics[i]);
    if (!P[i]) return fail("list: bad picture handle %d", pics[i]);
  }
  int h = -1;
  for (int i = 0; i < 256; i++) if (!c.lists[i].used) { h = i; break; }
  if (h < 0) return fail("list: too many open lists");
  List &L = c.lists[h];
  L = List();
  L.kind = kind; L.ntasks = ntasks; L.npics = npics;
  L.pics.assign(pics, pics + npics);
  L.pic_gens.clear();
  for (int i = 0; i < npics; i++) L.pic_gens.push_back(get_pic(pics[i])->gen);

  std::vector<PredTile> tiles;
  std::vector<uint32_t> order;
  size_t task_size = 0;
  const void *second = nullptr;
  size_t second_bytes = 0;
  if (kind == THORCUDA_LIST_PRED) {
    if (check_pred(P, pics, npics, (const thorcuda_pred_task *)tasks, ntasks, tiles)) return -1;
    task_size = sizeof(thorcuda_pred_task);
    // tiles larger than 8x8 (one warp each) first, then the small ones (four to a warp); stable: neighbours stay together
    std::stable_partition(tiles.begin(), tiles.end(), [](const PredTile &q) { return q.tw > 8 || q.th > 8; });
    L.ntiles = (int)tiles.size();
    L.nbig = (int)std::count_if(tiles.begin(), tiles.end(), [](const PredTile &q) { return q.tw > 8 || q.th > 8; });
    second = tiles.data(); second_bytes = tiles.size() * sizeof(PredTile);
  } else if (kind == THORCUDA_LIST_TB) {
    if (check_tb(P, npics, (const thorcuda_tb_task *)tasks, ntasks, L, order)) return -1;
    task_size = sizeof(thorcuda_tb_task);
    second = order.data(); second_bytes = order.size() * sizeof(uint32_t);
  } else {
    if (check_cost(P, npics, (const thorcuda_cost_task *)tasks, ntasks)) return -1;
    task_size = sizeof(thorcuda_cost_task);
    for (int i = 0; i < ntasks; i++) {
      const thorcuda_cost_task &t = ((const thorcuda_cost_task *)tasks)[i];
      L.cost_w = std::max(L.cost_w, (int)t.w);
      L.cost_h = std::max(L.cost_h, (int)t.h);
    }
  }
  const size_t task_bytes = (size_t)ntasks * task_size;
  L.off_tiles = (task_bytes + 15) & ~(size_t)15;
  L.off_map = (L.off_tiles + second_bytes + 15) & ~(size_t)15;
  L.blob_bytes = L.off_map + sizeof(uint16_t) * npics;
  if (cudaMallocAsync((void **)&L.blob, L.blob_bytes, c.stream) != cudaSuccess) return fail("list: device allocation of %zu bytes failed", L.blob_bytes);
  uint8_t *st = (uint8_t *)stage_acquire(L.blob_bytes);
  if (!st) { cudaFreeAsync(L.blob, c.stream); return fail("list: staging allocation failed"); }
  memcpy(st, tasks, task_bytes);
  if (second_bytes) memcpy(st + L.off_tiles, second, second_bytes);
  uint16_t *map = (uint16_t *)(st + L.off_map);
  for (int i = 0; i < npics; i++) map[i] = (uint16_t)pics[i];
  if (stage_commit_copy(L.blob, st, L.blob_bytes)) { cudaFreeAsync(L.blob, c.stream); return -1; }
  cudaError_t e = cudaSuccess;
  if (kind == THORCUDA_LIST_TB) {
    L.order = (uint32_t *)(L.blob + L.off_tiles);
    e = cudaMallocAsync((void **)&L.pool, L.pool_entries * sizeof(int16_t) + 64, c.stream);
    if (e == cudaSuccess) e = cudaMallocAsync((void **)&L.cbp, (size_t)ntasks + 16, c.stream);
    if (e == cudaSuccess) e = cudaMallocAsync((void **)&L.ssd, (size_t)ntasks * 4 + 16, c.stream);
    if (e == cudaSuccess) e = cudaMallocAsync((void **)&L.bits, (size_t)ntasks * 4 + 16, c.stream);
    if (e == cudaSuccess) e = cudaMemsetAsync(L.pool, 0, L.pool_entries * sizeof(int16_t) + 64, c.stream);
  } else if (kind == THORCUDA_LIST_COST) {
    e = cudaMallocAsync((void **)&L.res_buf[0], (size_t)ntasks * sizeof(thorcuda_cost_result), c.stream);
    if (e == cudaSuccess) e = cudaMallocAsync((void **)&L.res_buf[1], (size_t)ntasks * sizeof(thorcuda_cost_result), c.stream);
    L.results = L.res_buf[0];
  }
  if (e != cudaSuccess) {
    cudaFreeAsync(L.blob, c.stream); if (L.pool) cudaFreeAsync(L.pool, c.stream); if (L.cbp) cudaFreeAsync(L.cbp, c.stream);
    if (L.ssd) cudaFreeAsync(L.ssd, c.stream);
    if (L.bits) cudaFreeAsync(L.bits, c.stream);
    for (int i = 0; i < 2; i++) if (L.res_buf[i]) cudaFreeAsync(L.res_buf[i], c.stream);
    L = List();
    return fail("list: device allocation failed (%s)", cudaGetErrorString(e));
  }
  L.used = true;
  return h;
}

void thorcuda_list_destroy(thorcuda_list h) {
  List *L = get_list(h);
  if (!L) return;
  list_fence(L);                              // output copies still in flight on the copy stream finish first
  cudaStream_t st = ctx().stream;             // stream-ordered: queued work that uses the list finishes first
  if (L->blob) cudaFreeAsync(L->blob, st);
  if (L->pool) cudaFreeAsync(L->pool, st);
  if (L->cbp) cudaFreeAsync(L->cbp, st);
  if (L->ssd) cudaFreeAsync(L->ssd, st);
  if (L->bits) cudaFreeAsync(L->bits, st);
  for (int i = 0; i < 2; i++) if (L->res_buf[i]) cudaFreeAsync(L->res_buf[i], st);
  *L = List();
}

}  // extern "C"

namespace thor {
int list_fence(List *L) {
  if (!L->late_pool && !L->late_res[0] && !L->late_res[1]) return 0;
  L->late_pool = L->late_res[0] = L->late_res[1] = false;
  return copies_fence();
}
}  // namespace thor

extern "C" {

size_t thorcuda_list_pool_entries(thorcuda_list h) {
  List *L = get_list(h);
  return L ? L->pool_entries : 0;
}

}  // extern "C"
