// libthorcuda: context, error convention, picture pool, host<->device picture copies.
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
  if (bytes <= cap) return p;
  if (p) cudaFree(p);
  p = nullptr; cap = 0;
  size_t want = bytes + bytes / 4 + 256;
  if (cudaMalloc(&p, want) != cudaSuccess) { p = nullptr; return nullptr; }
  cap = want;
  return p;
}
void DevBuf::release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
void *PinBuf::reserve(size_t bytes) {
  if (bytes <= cap) return p;
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
  return v;
}

Pic *get_pic(thorcuda_pic h) {
  if (h < 0 || h >= 256 || !g_ctx.pics[h].used) return nullptr;
  return &g_ctx.pics[h];
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
  THOR_CUDA(cudaSetDevice(device));
  THOR_CUDA(cudaStreamCreateWithFlags(&c.stream, cudaStreamNonBlocking));
  c.device = device;
  c.sm_count = prop.multiProcessorCount;
  c.launches = 0;
  c.ready = true;
  return 0;
}

void thorcuda_shutdown(void) {
  Context &c = ctx();
  if (!c.ready) return;
  cudaStreamSynchronize(c.stream);
  for (int i = 0; i < 256; i++) if (c.pics[i].used) thorcuda_pic_destroy(i);
  DevBuf *bufs[] = {&c.blockinfo_raw, &c.blockinfo_packed, &c.sb_bits, &c.tasks, &c.results, &c.pool, &c.cbp, &c.ssd,
                    &c.refs, &c.sadtab[0], &c.sadtab[1], &c.sadtab[2], &c.sadtab[3], &c.centres};
  for (DevBuf *b : bufs) b->release();
  c.h_tasks.release(); c.h_misc.release();
  cudaStreamDestroy(c.stream);
  c.stream = nullptr;
  c.ready = false;
}

const char *thorcuda_last_error(void) { return ctx().err; }

int thorcuda_sync(void) {
  THOR_NEED_CTX();
  THOR_CUDA(cudaStreamSynchronize(ctx().stream));
  return 0;
}

void *thorcuda_stream(void) { return (void *)ctx().stream; }
uint64_t thorcuda_launch_count(void) { return ctx().launches; }

void *thorcuda_host_alloc(size_t bytes) {
  THOR_NEED_CTX();
  void *p = nullptr;
  if (cudaMallocHost(&p, bytes) != cudaSuccess) die("libthorcuda: pinned host allocation failed");
  return p;
}
void thorcuda_host_free(void *p) { if (p) cudaFreeHost(p); }

thorcuda_pic thorcuda_pic_create(int width, int height, int pad_y, int pad_c) {
  THOR_NEED_CTX();
  Context &c = ctx();
  if (width <= 0 || height <= 0 || (width & 7) || (height & 7))   // enc/strings.c:437
    return fail("thorcuda_pic_create: width/height must be positive multiples of 8 (got %dx%d)", width, height);
  int h = -1;
  for (int i = 0; i < 256; i++) if (!c.pics[i].used) { h = i; break; }
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
  p.used = true;
  return h;
}

void thorcuda_pic_destroy(thorcuda_pic h) {
  Pic *p = get_pic(h);
  if (!p) return;
  cudaStreamSynchronize(ctx().stream);
  for (int i = 0; i < 2; i++) if (p->buf[i]) cudaFree(p->buf[i]);
  *p = Pic();
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

}  // extern "C"
