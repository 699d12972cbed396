// libthorcuda: context, error convention, picture pool, staging ring, host<->device picture copies.
#include <stdarg.h>
#include <string.h>
#include "internal.cuh"

namespace thor {

static Context g_ctx;
Context &ctx() { return g_ctx; }

int fail(const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_ctx.err, sizeof(g_ctx.err), fmt, ap);
  va_end(ap);
  return -1;
}

// Same observable behaviour as the reference's fatalerror() (common/global.h:38-44).
void die(const char *msg) {
  fprintf(stderr, "Run-time error...\n%s\n...now exiting to system...\n", msg);
  abort();
}

void *DevBuf::reserve(size_t bytes) {
  if (bytes <= cap && p) return p;
  if (p) { cudaStreamSynchronize(g_ctx.stream); cudaFree(p); }
  p = nullptr; cap = 0;
  size_t want = bytes + bytes / 4 + 256;
  if (cudaMalloc(&p, want) != cudaSuccess) { p = nullptr; return nullptr; }
  cap = want;
  return p;
}
void DevBuf::release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
void *PinBuf::reserve(size_t bytes) {
  if (bytes <= cap && p) return p;
  if (p) cudaFreeHost(p);
  p = nullptr; cap = 0;
  size_t want = bytes + bytes / 4 + 256;
  if (cudaMallocHost(&p, want) != cudaSuccess) { p = nullptr; return nullptr; }
  cap = want;
  return p;
}
void PinBuf::release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }

PicView Pic::view(int which) const {
  PicView v;
  uint8_t *b = buf[which];
  const size_t area_y = (size_t)rows_y * stride_y, area_c = (size_t)rows_c * stride_c;
  v.p[0] = b + (size_t)pad_y * stride_y + pad_y;
  v.p[1] = b + area_y + (size_t)pad_c * stride_c + pad_c;
  v.p[2] = b + area_y + area_c + (size_t)pad_c * stride_c + pad_c;
  v.stride[0] = stride_y; v.stride[1] = v.stride[2] = stride_c;
  v.w = w; v.h = h;
  v.pad[0] = pad_y; v.pad[1] = v.pad[2] = pad_c;
  v.reserved = 0;
  return v;
}

Pic *get_pic(thorcuda_pic h) {
  if (h < 0 || h >= THORCUDA_MAX_PICS || !g_ctx.pics[h].used) return nullptr;
  return &g_ctx.pics[h];
}
int list_stale(const List *L) {
  for (size_t i = 0; i < L->pics.size(); i++) {
    const Pic *p = get_pic(L->pics[i]);
    if (!p || p->gen != L->pic_gens[i])
      return fail("list: picture %d (handle %d) was destroyed or recreated after the list was made; recreate the list", (int)i, L->pics[i]);
  }
  return 0;
}

List *get_list(thorcuda_list h) {
  if (h < 0 || h >= 256 || !g_ctx.lists[h].used) return nullptr;
  return &g_ctx.lists[h];
}

static bool is_pinned(const void *p, size_t bytes) {
  const uintptr_t a = (uintptr_t)p;
  for (auto &r : g_ctx.pinned) if (a >= r.first && a + bytes <= r.first + r.second) return true;
  return false;
}

void *stage_acquire(size_t bytes) {
  Context &c = g_ctx;
  StageSlot &s = c.stage[c.stage_next];
  if (s.pending) { cudaEventSynchronize(s.ev); s.pending = false; }
  return s.buf.reserve(bytes);
}
int stage_commit_copy(void *dst, const void *staged, size_t bytes) {
  Context &c = g_ctx;
  StageSlot &s = c.stage[c.stage_next];
  THOR_CUDA(cudaMemcpyAsync(dst, staged, bytes, cudaMemcpyHostToDevice, c.stream));
  THOR_CUDA(cudaEventRecord(s.ev, c.stream));
  s.pending = true;
  c.stage_next = (c.stage_next + 1) % 8;
  return 0;
}
int h2d(void *dst, const void *src, size_t bytes) {
  if (!bytes) return 0;
  if (is_pinned(src, bytes)) {
    THOR_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, g_ctx.stream));
    return 0;
  }
  void *st = stage_acquire(bytes);
  if (!st) return fail("pinned staging allocation of %zu bytes failed", bytes);
  memcpy(st, src, bytes);
  return stage_commit_copy(dst, st, bytes);
}
int d2h(void *dst, const void *src, size_t bytes) {
  if (!bytes) return 0;
  THOR_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, g_ctx.stream));
  return 0;
}
int d2h_bulk(void *dst, const void *src, size_t bytes) {
  if (!bytes) return 0;
  Context &c = g_ctx;
  THOR_CUDA(cudaEventRecord(c.ev_main, c.stream));
  THOR_CUDA(cudaStreamWaitEvent(c.copy_stream, c.ev_main, 0));
  THOR_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, c.copy_stream));
  c.copy_pending = true;
  return 0;
}
int copies_fence() {
  Context &c = g_ctx;
  if (!c.copy_pending) return 0;
  THOR_CUDA(cudaEventRecord(c.ev_copy, c.copy_stream));
  THOR_CUDA(cudaStreamWaitEvent(c.stream, c.ev_copy, 0));
  return 0;
}
int sync_pic_view(thorcuda_pic h) {
  const PicView v = g_ctx.pics[h].view();
  return h2d(g_ctx.d_pics + h, &v, sizeof(v));
}

__global__ void k_fill(uint4 *p, size_t n, uint32_t v) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    p[i] = make_uint4(v, v, v, v);
}

// copy of the visible area of all three planes, 16 bytes per thread (widths are multiples of 8; the
// chroma tail of an odd multiple of 8 is handled with 4-byte accesses)
__global__ void k_copy_visible(PicView d, PicView s) {
  const int pl = blockIdx.z, w = pl ? s.w / 2 : s.w, h = pl ? s.h / 2 : s.h;
  const int y = blockIdx.y;
  if (y >= h) return;
  const uint8_t *sp = s.p[pl] + (size_t)y * s.stride[pl];
  uint8_t *dp = d.p[pl] + (size_t)y * d.stride[pl];
  for (int x = (blockIdx.x * blockDim.x + threadIdx.x) * 4; x < w; x += gridDim.x * blockDim.x * 4)
    *reinterpret_cast<uint32_t *>(dp + x) = *reinterpret_cast<const uint32_t *>(sp + x);
}

}  // namespace thor

using namespace thor;

extern "C" {

int thorcuda_init(int device) {
  Context &c = ctx();
  if (c.ready) return 0;
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0)
    return fail("libthorcuda: no CUDA device visible (there is no CPU fallback)");
  if (device < 0 || device >= n) return fail("libthorcuda: device %d out of range (%d visible)", device, n);
  cudaDeviceProp prop;
  THOR_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10)
    return fail("libthorcuda: device %d is sm_%d%d; this library is built for sm_100a (B200) only", device, prop.major, prop.minor);
  // THORCUDA_BLOCKING_SYNC=1: host threads sleep in thorcuda_sync() instead of spinning -- for many encoder / decoder processes
  // per host (one per stream, bench.py), whose frames take far longer on the device than a wake-up costs
  if (getenv("THORCUDA_BLOCKING_SYNC")) cudaSetDeviceFlags(cudaDeviceScheduleBlockingSync);
  THOR_CUDA(cudaSetDevice(device));
  THOR_CUDA(cudaStreamCreateWithFlags(&c.stream, cudaStreamNonBlocking));
  THOR_CUDA(cudaStreamCreateWithFlags(&c.copy_stream, cudaStreamNonBlocking));
  THOR_CUDA(cudaEventCreateWithFlags(&c.ev_main, cudaEventDisableTiming));
  THOR_CUDA(cudaEventCreateWithFlags(&c.ev_copy, cudaEventDisableTiming));
  c.copy_pending = false;
  {  // task lists come from the stream-ordered allocator; keep freed blocks cached instead of returning them to the OS
    cudaMemPool_t pool;
    THOR_CUDA(cudaDeviceGetDefaultMemPool(&pool, device));
    uint64_t keep = UINT64_MAX;
    THOR_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep));
  }
  THOR_CUDA(cudaMalloc((void **)&c.d_pics, sizeof(PicView) * THORCUDA_MAX_PICS));
  THOR_CUDA(cudaMemsetAsync(c.d_pics, 0, sizeof(PicView) * THORCUDA_MAX_PICS, c.stream));
  for (auto &s : c.stage) { THOR_CUDA(cudaEventCreateWithFlags(&s.ev, cudaEventDisableTiming)); s.pending = false; }
  for (int i = 0; i < 64; i++) { THOR_CUDA(cudaEventCreate(&c.tm0[i])); THOR_CUDA(cudaEventCreate(&c.tm1[i])); c.tm_used[i] = false; }
  c.stage_next = 0;
  c.device = device;
  c.sm_count = prop.multiProcessorCount;
  c.launches = 0;
  c.ready = true;
  return 0;
}

// ---- host helpers for table indexing (no device work) ----
void thorcuda_clip_mv(int16_t mv[2], int ypos, int xpos, int fwidth, int fheight, int size, int sign) {
  const int reach = 80;                        // PADDING_Y - 16 full pixels beyond each frame edge
  const int s = sign ? -1 : 1;
  int v[2] = {s * mv[0], s * mv[1]};           // x, y as the search applies them
  // the four tests run in the reference's order on the running value; C division truncates toward zero
  if (ypos + v[1] / 4 < -reach) v[1] = 4 * (-reach - ypos);
  if (ypos + v[1] / 4 + size > fheight + reach) v[1] = 4 * (fheight + reach - ypos - size);
  if (xpos + v[0] / 4 < -reach) v[0] = 4 * (-reach - xpos);
  if (xpos + v[0] / 4 > fwidth + reach) v[0] = 4 * (fwidth + reach - xpos - size);      // sic: the test has no "+ size"
  mv[0] = (int16_t)(s * v[0]);
  mv[1] = (int16_t)(s * v[1]);
}

int64_t thorcuda_sad_table_index(int fwidth, int bs, int xpos, int ypos, const int16_t mv[2], int sign, int centre_x, int centre_y, int range) {
  if (bs <= 0 || (xpos % bs) || (ypos % bs) || xpos + bs > fwidth) return -1;
  const int s = sign ? -1 : 1;
  const int dx = s * (mv[0] >> 2) - centre_x, dy = s * (mv[1] >> 2) - centre_y;   // the search reads ref + s*(mv.y>>2)*stride + s*(mv.x>>2)
  if (dx < -range || dx >= range || dy < -range || dy >= range) return -1;
  const int64_t block = (int64_t)(ypos / bs) * (fwidth / bs) + xpos / bs;
  return (block * 2 * range + (dy + range)) * 2 * range + (dx + range);
}

void thorcuda_shutdown(void) {
  Context &c = ctx();
  if (!c.ready) return;
  cudaStreamSynchronize(c.stream);
  cudaStreamSynchronize(c.copy_stream);
  compat_reset();
  interp_reset();
  bits_reset();
  enc_reset();
  for (int i = 0; i < 256; i++) if (c.lists[i].used) thorcuda_list_destroy(i);
  for (int i = 0; i < THORCUDA_MAX_PICS; i++) if (c.pics[i].used) thorcuda_pic_destroy(i);
  DevBuf *bufs[] = {&c.raw_info, &c.scratch, &c.flush, &c.lfjobs, &c.sadjobs, &c.centres};
  for (DevBuf *b : bufs) b->release();
  for (auto &s : c.sad) {
    for (int l = 0; l < 4; l++) if (s.tab[l]) cudaFree(s.tab[l]);
    s = SadSlot();
  }
  for (auto &s : c.stage) { s.buf.release(); cudaEventDestroy(s.ev); s.ev = nullptr; s.pending = false; }
  for (int i = 0; i < 64; i++) { cudaEventDestroy(c.tm0[i]); cudaEventDestroy(c.tm1[i]); }
  for (int i = 0; i < 16; i++) {
    if (c.fn_made[i]) { cudaEventDestroy(c.fn_main[i]); cudaEventDestroy(c.fn_copy[i]); }
    c.fn_made[i] = c.fn_used[i] = false;
  }
  cudaFree(c.d_pics); c.d_pics = nullptr;
  cudaStreamDestroy(c.stream);
  cudaStreamDestroy(c.copy_stream);
  cudaEventDestroy(c.ev_main); cudaEventDestroy(c.ev_copy);
  c.stream = c.copy_stream = nullptr;
  c.ev_main = c.ev_copy = nullptr;
  c.ready = false;
}

const char *thorcuda_last_error(void) { return ctx().err; }

int thorcuda_sync(void) {
  THOR_NEED_CTX();
  THOR_CUDA(cudaStreamSynchronize(ctx().stream));
  if (ctx().copy_pending) {
    THOR_CUDA(cudaStreamSynchronize(ctx().copy_stream));
    ctx().copy_pending = false;
  }
  return 0;
}

/* A marker behind everything queued so far -- kernels and copies on the library stream AND the bulk fetches on its copy
 * stream -- that can be waited for without waiting for what is queued later: a caller that works on several independent
 * groups of streams submits group B while group A's results are still travelling, and waits for A alone. */
int thorcuda_fence_record(void) {
  THOR_NEED_CTX();
  Context &c = ctx();
  for (int i = 0; i < 16; i++)
    if (!c.fn_used[i]) {
      if (!c.fn_made[i]) {
        if (cudaEventCreateWithFlags(&c.fn_main[i], cudaEventDisableTiming) != cudaSuccess ||
            cudaEventCreateWithFlags(&c.fn_copy[i], cudaEventDisableTiming) != cudaSuccess) { fail("fence: event creation failed"); return -1; }
        c.fn_made[i] = true;
      }
      if (cudaEventRecord(c.fn_main[i], c.stream) != cudaSuccess) { fail("fence: record failed"); return -1; }
      c.fn_copy_set[i] = c.copy_pending;
      if (c.copy_pending && cudaEventRecord(c.fn_copy[i], c.copy_stream) != cudaSuccess) { fail("fence: record failed"); return -1; }
      c.fn_used[i] = true;
      return i;
    }
  fail("fence: all 16 fences are in use");
  return -1;
}
int thorcuda_fence_wait(int id) {
  THOR_NEED_CTX();
  Context &c = ctx();
  if (id < 0 || id >= 16 || !c.fn_used[id]) return fail("fence: bad id %d", id);
  c.fn_used[id] = false;
  THOR_CUDA(cudaEventSynchronize(c.fn_main[id]));
  if (c.fn_copy_set[id]) THOR_CUDA(cudaEventSynchronize(c.fn_copy[id]));
  return 0;
}

void *thorcuda_stream(void) { return (void *)ctx().stream; }
uint64_t thorcuda_launch_count(void) { return ctx().launches; }

int thorcuda_timer_begin(void) {
  THOR_NEED_CTX();
  Context &c = ctx();
  for (int i = 0; i < 64; i++)
    if (!c.tm_used[i]) {
      c.tm_used[i] = true;
      if (cudaEventRecord(c.tm0[i], c.stream) != cudaSuccess) return -1;
      return i;
    }
  fail("timer: all 64 timers are open");
  return -1;
}
int thorcuda_timer_end(int id) {
  THOR_NEED_CTX();
  Context &c = ctx();
  if (id < 0 || id >= 64 || !c.tm_used[id]) return fail("timer: bad id %d", id);
  THOR_CUDA(cudaEventRecord(c.tm1[id], c.stream));
  return 0;
}
float thorcuda_timer_elapsed_ms(int id) {
  Context &c = ctx();
  if (id < 0 || id >= 64 || !c.tm_used[id]) return -1.f;
  float ms = -1.f;
  if (cudaEventSynchronize(c.tm1[id]) != cudaSuccess || cudaEventElapsedTime(&ms, c.tm0[id], c.tm1[id]) != cudaSuccess) ms = -1.f;
  c.tm_used[id] = false;
  return ms;
}

int thorcuda_flush_l2(size_t bytes) {
  THOR_NEED_CTX();
  Context &c = ctx();
  bytes = (bytes + 15) & ~(size_t)15;
  void *p = c.flush.reserve(bytes);
  if (!p) return fail("flush_l2: device allocation of %zu bytes failed", bytes);
  k_fill<<<c.sm_count * 8, 256, 0, c.stream>>>((uint4 *)p, bytes / 16, 0x5a5a5a5au);
  THOR_CUDA(cudaGetLastError());   // not counted in launch_count: it is measurement hygiene, not pipeline work
  return 0;
}

void *thorcuda_host_alloc(size_t bytes) {
  THOR_NEED_CTX();
  void *p = nullptr;
  if (cudaMallocHost(&p, bytes) != cudaSuccess) die("libthorcuda: pinned host allocation failed");
  ctx().pinned.push_back({(uintptr_t)p, bytes});
  return p;
}
int thorcuda_host_register(void *p, size_t bytes) {
  THOR_NEED_CTX();
  if (!p || !bytes) return fail("host_register: empty range");
  THOR_CUDA(cudaHostRegister(p, bytes, cudaHostRegisterDefault));
  ctx().pinned.push_back({(uintptr_t)p, bytes});
  return 0;
}
void thorcuda_host_unregister(void *p) {
  if (!p) return;
  auto &v = ctx().pinned;
  for (size_t i = 0; i < v.size(); i++) if (v[i].first == (uintptr_t)p) { v.erase(v.begin() + i); break; }
  cudaStreamSynchronize(ctx().stream);
  cudaStreamSynchronize(ctx().copy_stream);
  cudaHostUnregister(p);
}
void thorcuda_host_free(void *p) {
  if (!p) return;
  auto &v = ctx().pinned;
  for (size_t i = 0; i < v.size(); i++) if (v[i].first == (uintptr_t)p) { v.erase(v.begin() + i); break; }
  cudaStreamSynchronize(ctx().stream);
  cudaStreamSynchronize(ctx().copy_stream);      // a bulk fetch may still be writing into it
  cudaFreeHost(p);
}

thorcuda_pic thorcuda_pic_create(int width, int height, int pad_y, int pad_c) {
  THOR_NEED_CTX();
  Context &c = ctx();
  if (width <= 0 || height <= 0 || (width & 7) || (height & 7))   // enc/strings.c:437
    return fail("thorcuda_pic_create: width/height must be positive multiples of 8 (got %dx%d)", width, height);
  if (width > 16384 || height > 16384 || pad_y < 0 || pad_c < 0 || pad_y > 256 || pad_c > 256)
    return fail("thorcuda_pic_create: unsupported geometry");
  int h = -1;
  for (int i = 0; i < THORCUDA_MAX_PICS; i++) if (!c.pics[i].used) { h = i; break; }
  if (h < 0) return fail("thorcuda_pic_create: picture pool exhausted");
  Pic &p = c.pics[h];
  p = Pic();
  p.w = width; p.h = height; p.pad_y = pad_y; p.pad_c = pad_c;
  p.stride_y = (width + 2 * pad_y + 15) & ~15;
  p.stride_c = (width / 2 + 2 * pad_c + 15) & ~15;
  p.rows_y = height + 2 * pad_y;
  p.rows_c = height / 2 + 2 * pad_c;
  p.bytes = (size_t)p.rows_y * p.stride_y + 2 * (size_t)p.rows_c * p.stride_c;
  if (cudaMalloc((void **)&p.buf[0], p.bytes + 256) != cudaSuccess) return fail("thorcuda_pic_create: cudaMalloc(%zu) failed", p.bytes);
  cudaMemsetAsync(p.buf[0], 0, p.bytes + 256, c.stream);
  static uint32_t next_gen = 0;
  p.gen = ++next_gen;
  p.used = true;
  if (sync_pic_view(h)) return -1;
  return h;
}

void thorcuda_pic_destroy(thorcuda_pic h) {
  Pic *p = get_pic(h);
  if (!p) return;
  cudaStreamSynchronize(ctx().stream);
  for (int i = 0; i < 2; i++) if (p->buf[i]) cudaFree(p->buf[i]);
  if (p->info) cudaFree(p->info);
  *p = Pic();
  sync_pic_view(h);                       // the device table entry becomes all-zero: a stale list cannot reach freed memory through it
}

int thorcuda_pic_stride(thorcuda_pic h, int plane) {
  Pic *p = get_pic(h);
  if (!p) return fail("bad picture handle %d", h);
  return plane == 0 ? p->stride_y : p->stride_c;
}

void *thorcuda_pic_devptr(thorcuda_pic h, int plane) {
  Pic *p = get_pic(h);
  if (!p || plane < 0 || plane > 2) return nullptr;
  return p->view().p[plane];
}

static int copy_visible(thorcuda_pic h, uint8_t *const hp[3], int sy, int sc, bool to_device) {
  THOR_NEED_CTX();
  Pic *p = get_pic(h);
  if (!p) return fail("bad picture handle %d", h);
  PicView v = p->view();
  for (int i = 0; i < 3; i++) {
    if (!hp[i]) continue;
    const int w = i ? p->w / 2 : p->w, hh = i ? p->h / 2 : p->h, hs = i ? sc : sy;
    if (to_device) THOR_CUDA(cudaMemcpy2DAsync(v.p[i], v.stride[i], hp[i], hs, w, hh, cudaMemcpyHostToDevice, ctx().stream));
    else           THOR_CUDA(cudaMemcpy2DAsync(hp[i], hs, v.p[i], v.stride[i], w, hh, cudaMemcpyDeviceToHost, ctx().stream));
  }
  return 0;
}

int thorcuda_pic_upload(thorcuda_pic h, const uint8_t *y, const uint8_t *u, const uint8_t *v, int sy, int sc) {
  uint8_t *hp[3] = {(uint8_t *)y, (uint8_t *)u, (uint8_t *)v};
  return copy_visible(h, hp, sy, sc, true);
}
int thorcuda_pic_download(thorcuda_pic h, uint8_t *y, uint8_t *u, uint8_t *v, int sy, int sc) {
  uint8_t *hp[3] = {y, u, v};
  return copy_visible(h, hp, sy, sc, false);
}

static int copy_padded(thorcuda_pic h, uint8_t *const hp[3], bool to_device) {
  THOR_NEED_CTX();
  Pic *p = get_pic(h);
  if (!p) return fail("bad picture handle %d", h);
  uint8_t *b = p->buf[p->cur];
  const size_t ay = (size_t)p->rows_y * p->stride_y, ac = (size_t)p->rows_c * p->stride_c;
  uint8_t *dp[3] = {b, b + ay, b + ay + ac};
  const size_t n[3] = {ay, ac, ac};
  for (int i = 0; i < 3; i++) {
    if (!hp[i]) continue;
    if (to_device) THOR_CUDA(cudaMemcpyAsync(dp[i], hp[i], n[i], cudaMemcpyHostToDevice, ctx().stream));
    else           THOR_CUDA(cudaMemcpyAsync(hp[i], dp[i], n[i], cudaMemcpyDeviceToHost, ctx().stream));
  }
  return 0;
}
int thorcuda_pic_download_padded(thorcuda_pic h, uint8_t *y, uint8_t *u, uint8_t *v) {
  uint8_t *hp[3] = {y, u, v};
  return copy_padded(h, hp, false);
}
int thorcuda_pic_upload_padded(thorcuda_pic h, const uint8_t *y, const uint8_t *u, const uint8_t *v) {
  uint8_t *hp[3] = {(uint8_t *)y, (uint8_t *)u, (uint8_t *)v};
  return copy_padded(h, hp, true);
}

int thorcuda_pic_copy(thorcuda_pic hd, thorcuda_pic hs) {
  THOR_NEED_CTX();
  Pic *d = get_pic(hd), *s = get_pic(hs);
  if (!d || !s) return fail("pic_copy: bad handle");
  if (d->w != s->w || d->h != s->h) return fail("pic_copy: size mismatch");
  dim3 grid(4, s->h, 3);
  k_copy_visible<<<grid, 128, 0, ctx().stream>>>(d->view(), s->view());
  THOR_LAUNCHED();
  return 0;
}

}  // extern "C"
