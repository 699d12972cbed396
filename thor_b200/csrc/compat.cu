// libthorcuda: the reference's per-block kernel API (include/thorcuda_compat.h) on top of the device code.
//
// Each entry point copies ONE block's operands from the caller's host memory into three small scratch
// pictures that live in HBM, runs the SAME kernels the batched API uses (k_cost_batch, k_predict, the
// transform passes) on a one-task list, copies the result back and synchronises.  Only the host bytes the
// reference function itself reads are touched (reach of every function: see the comments below).
// CLPF / block average have no batched twin at block granularity and get two small kernels here.
#include <string.h>
#include "internal.cuh"
#include "../../include/thorcuda_compat.h"

namespace thor {

namespace {

struct Compat {
  bool ready = false;
  thorcuda_pic org = -1, ref = -1, dst = -1;   // 64x64 scratch pictures; ref carries a 16/8 sample border
  DevBuf d_in, d_out;
  PinBuf pin;
};
Compat g_compat;

constexpr int REF_PAD_Y = 16, REF_PAD_C = 8;

Compat &compat() {
  Compat &k = g_compat;
  if (k.ready && ctx().ready) return k;
  if (!ctx().ready) {
    const char *dev = getenv("THORCUDA_DEVICE");
    if (thorcuda_init(dev ? atoi(dev) : 0)) die(thorcuda_last_error());
  }
  k.org = thorcuda_pic_create(64, 64, 0, 0);
  k.ref = thorcuda_pic_create(64, 64, REF_PAD_Y, REF_PAD_C);
  k.dst = thorcuda_pic_create(64, 64, 0, 0);
  if (k.org < 0 || k.ref < 0 || k.dst < 0) die(thorcuda_last_error());
  k.d_in = DevBuf(); k.d_out = DevBuf(); k.pin = PinBuf();
  if (!k.d_in.reserve(64 * 64 * 2) || !k.d_out.reserve(64 * 64 * 2) || !k.pin.reserve(1 << 16)) die("libthorcuda: compat scratch allocation failed");
  k.ready = true;
  return k;
}

void must(int rc) { if (rc) die(thorcuda_last_error()); }
void must_cuda(cudaError_t e) { if (e != cudaSuccess) die(cudaGetErrorString(e)); }

// rows [y0,y1) x columns [x0,x1) relative to host pointer p (stride hs) -> the same window relative to the
// visible origin of plane `pl` of scratch picture h
void put_window(thorcuda_pic h, int pl, const uint8_t *p, int hs, int x0, int y0, int x1, int y1) {
  Pic *P = get_pic(h);
  const PicView v = P->view();
  must_cuda(cudaMemcpy2DAsync(v.p[pl] + (ptrdiff_t)y0 * v.stride[pl] + x0, v.stride[pl], p + (ptrdiff_t)y0 * hs + x0, hs, x1 - x0, y1 - y0,
                              cudaMemcpyHostToDevice, ctx().stream));
}
void get_window(thorcuda_pic h, int pl, uint8_t *p, int hs, int w, int hgt) {
  Pic *P = get_pic(h);
  const PicView v = P->view();
  must_cuda(cudaMemcpy2DAsync(p, hs, v.p[pl], v.stride[pl], w, hgt, cudaMemcpyDeviceToHost, ctx().stream));
}

void check_block(const char *fn, int w, int h, int lo, int hi, int mult) {
  if (w < lo || h < lo || w > hi || h > hi || (w % mult) || (h % mult)) {
    char msg[160];
    snprintf(msg, sizeof(msg), "libthorcuda: %s: unsupported block %dx%d", fn, w, h);
    die(msg);
  }
}

// one cost task on the scratch pictures; (l,r,u,d) = reach of the reference function around the block
thorcuda_cost_result cost1(const char *fn, int kind, const uint8_t *a, int as, const uint8_t *b, int bs, int w, int h, int hx, int hy,
                           int l, int r, int u, int d) {
  Compat &k = compat();
  check_block(fn, w, h, 4, 64, 4);
  put_window(k.org, 0, a, as, 0, 0, w, h);
  put_window(k.ref, 0, b, bs, -l, -u, w + r, h + d);
  thorcuda_cost_task t;
  memset(&t, 0, sizeof(t));
  t.w = (uint8_t)w; t.h = (uint8_t)h; t.org = 0; t.ref = 1; t.hx = (int8_t)hx; t.hy = (int8_t)hy;
  const thorcuda_pic pics[2] = {k.org, k.ref};
  thorcuda_cost_result *res = (thorcuda_cost_result *)k.pin.p;
  must(thorcuda_cost_batch(kind, pics, 2, &t, 1, res));
  must(thorcuda_sync());
  return *res;
}

// ---- CLPF of one size x size block whose (size+2)^2 neighbourhood (edges already replaced by the centre
// sample where the reference does so) is in `win`; optional SSD sums against org (detect_clpf) ----
__global__ void k_clpf_small(const uint8_t *__restrict__ win, uint8_t *__restrict__ out, const uint8_t *__restrict__ org, int size, int *sums) {
  const int i = threadIdx.x, n = size * size, p = size + 2;
  int s0 = 0, s1 = 0;
  if (i < n) {
    const int y = i / size, x = i - y * size;
    const uint8_t *c = win + (y + 1) * p + x + 1;
    const int X = c[0], A = c[-p], B = c[-1], C = c[1], D = c[p];
    const int F = X + ((A > X) + (B > X) + (C > X) + (D > X) > 2) - ((A < X) + (B < X) + (C < X) + (D < X) > 2);
    if (out) out[i] = (uint8_t)F;
    if (org) { const int O = org[i]; s0 = (O - X) * (O - X); s1 = (O - F) * (O - F); }
  }
  if (sums) {
#pragma unroll
    for (int m = 16; m > 0; m >>= 1) { s0 += __shfl_xor_sync(0xffffffffu, s0, m); s1 += __shfl_xor_sync(0xffffffffu, s1, m); }
    if ((i & 31) == 0) { atomicAdd(sums, s0); atomicAdd(sums + 1, s1); }
  }
}

__global__ void k_block_avg(const uint8_t *__restrict__ a, const uint8_t *__restrict__ b, uint8_t *__restrict__ o, int n) {
  for (int i = threadIdx.x; i < n; i += blockDim.x) o[i] = (uint8_t)((a[i] + b[i] + 1) >> 1);
}

// gather the clamped neighbourhood of the block at (x0,y0) on the host (clpf_block / detect_clpf edge rule:
// a neighbour outside [left,right] x [top,bottom] is replaced by the centre sample)
void gather_clpf(uint8_t *win, const uint8_t *src, int stride, int x0, int y0, int size, int left, int top, int right, int bottom) {
  const int p = size + 2;
  for (int y = -1; y <= size; y++)
    for (int x = -1; x <= size; x++) {
      if ((y < 0 || y >= size) && (x < 0 || x >= size)) { win[(y + 1) * p + x + 1] = 0; continue; }   // corners: unused
      int yy = y0 + y, xx = x0 + x;
      if (yy < top) yy = top; if (yy > bottom) yy = bottom;
      if (xx < left) xx = left; if (xx > right) xx = right;
      win[(y + 1) * p + x + 1] = src[(ptrdiff_t)yy * stride + xx];
    }
}

void clpf_block_n(const uint8_t *src, uint8_t *dst, int sstride, int dstride, int x0, int y0, int size, int width, int height) {
  Compat &k = compat();
  const int left = x0 & ~(dstride - 1), top = y0 & ~(dstride - 1);
  const int right = min(width - 1, left + dstride - 1), bottom = min(height - 1, top + dstride - 1);
  uint8_t *pin = (uint8_t *)k.pin.p;
  const int p = size + 2;
  gather_clpf(pin, src, sstride, x0, y0, size, left, top, right, bottom);
  uint8_t *d_win = (uint8_t *)k.d_in.p, *d_out = (uint8_t *)k.d_out.p;
  must_cuda(cudaMemcpyAsync(d_win, pin, p * p, cudaMemcpyHostToDevice, ctx().stream));
  k_clpf_small<<<1, 64, 0, ctx().stream>>>(d_win, d_out, nullptr, size, nullptr);
  ctx().launches++;
  must_cuda(cudaGetLastError());
  must_cuda(cudaMemcpyAsync(pin + 256, d_out, size * size, cudaMemcpyDeviceToHost, ctx().stream));
  must(thorcuda_sync());
  for (int y = 0; y < size; y++) memcpy(dst + (ptrdiff_t)(y0 + y - top) * dstride + (x0 - left), pin + 256 + y * size, size);
}

void transform_any(const int16_t *in, int16_t *out, int size, int fast, int inverse) {
  Compat &k = compat();
  if (size != 4 && size != 8 && size != 16 && size != 32 && size != 64) die("libthorcuda: transform: unsupported size");
  const int q = size < 16 ? size : 16;
  int16_t *pin = (int16_t *)k.pin.p;
  int16_t *d_in = (int16_t *)k.d_in.p, *d_out = (int16_t *)k.d_out.p;
  // forward reads the whole block; inverse reads only the low q x q corner (rows at stride size):
  // coefficients outside it are undefined in the reference and must not be touched
  if (!inverse) memcpy(pin, in, (size_t)size * size * 2);
  else for (int i = 0; i < q; i++) memcpy(pin + i * size, in + i * size, q * 2);
  must_cuda(cudaMemcpyAsync(d_in, pin, (size_t)size * size * 2, cudaMemcpyHostToDevice, ctx().stream));
  must(tb_raw(size, fast, inverse, d_in, d_out));
  must_cuda(cudaMemcpyAsync(pin + 4096, d_out, (size_t)size * size * 2, cudaMemcpyDeviceToHost, ctx().stream));
  must(thorcuda_sync());
  if (inverse) memcpy(out, pin + 4096, (size_t)size * size * 2);
  else for (int i = 0; i < q; i++) memcpy(out + i * size, pin + 4096 + i * size, q * 2);
}

}  // namespace
}  // namespace thor

using namespace thor;

namespace thor {
// called by thorcuda_shutdown(): the scratch pictures die with the picture pool
void compat_reset() {
  g_compat.d_in.release(); g_compat.d_out.release(); g_compat.pin.release();
  g_compat = Compat();
}
}  // namespace thor

extern "C" {

// ---- costs: reach = what the reference function reads around the w x h block at b ----
int sad_calc_simd(uint8_t *a, uint8_t *b, int astride, int bstride, int width, int height) {
  return (int)cost1("sad_calc_simd", THORCUDA_COST_SAD, a, astride, b, bstride, width, height, 0, 0, 0, 0, 0, 0).cost;
}
int sad_calc_simd_unaligned(uint8_t *a, uint8_t *b, int astride, int bstride, int width, int height) {
  // the reference's loop is only correct for widths up to 16 (common/common_kernels.c:111-119 advances the row
  // pointers once per 16-column step) and temporal interpolation never asks for more
  if (width > 16) die("libthorcuda: sad_calc_simd_unaligned: width > 16 is undefined in the reference");
  return (int)cost1("sad_calc_simd_unaligned", THORCUDA_COST_SAD, a, astride, b, bstride, width, height, 0, 0, 0, 0, 0, 0).cost;
}
int ssd_calc_simd(uint8_t *a, uint8_t *b, int astride, int bstride, int size) {
  return (int)cost1("ssd_calc_simd", THORCUDA_COST_SSD, a, astride, b, bstride, size, size, 0, 0, 0, 0, 0, 0).cost;
}
unsigned int widesad_calc_simd(uint8_t *a, uint8_t *b, int astride, int bstride, int width, int height, int *x) {
  // columns -3 .. w+2 (enc/encode_block.c:764-776)
  const thorcuda_cost_result r = cost1("widesad_calc_simd", THORCUDA_COST_WIDESAD, a, astride, b, bstride, width, height, 0, 0, 3, 3, 0, 0);
  *x = r.x;
  return r.cost;
}
unsigned int sad_calc_fasthalf_simd(const uint8_t *a, const uint8_t *b, int astride, int bstride, int width, int height, int *x, int *y) {
  // rows -2 .. h+1, columns -2 .. w+1 (enc/encode_block.c:508-553)
  const thorcuda_cost_result r = cost1("sad_calc_fasthalf_simd", THORCUDA_COST_FASTHALF, a, astride, b, bstride, width, height, 0, 0, 2, 2, 2, 2);
  *x = r.x; *y = r.y;
  return r.cost;
}
unsigned int sad_calc_fastquarter_simd(const uint8_t *o, const uint8_t *r, int os, int rs, int width, int height, int *x, int *y) {
  // rows -1 .. h, columns -1 .. w (enc/encode_block.c:616-700); (*x,*y) selects the formula set
  const thorcuda_cost_result c = cost1("sad_calc_fastquarter_simd", THORCUDA_COST_FASTQUARTER, o, os, r, rs, width, height, *x, *y, 1, 1, 1, 1);
  *x = c.x; *y = c.y;
  return c.cost;
}

// ---- inter prediction ----
void get_inter_prediction_luma_simd(int width, int height, int xoff, int yoff, unsigned char *THOR_RESTRICT qp, int qstride,
                                    const unsigned char *THOR_RESTRICT ip, int istride, int bipred) {
  Compat &k = compat();
  check_block("get_inter_prediction_luma_simd", width, height, 4, 64, 4);
  if (xoff < 0 || xoff > 3 || yoff < 0 || yoff > 3) die("libthorcuda: get_inter_prediction_luma_simd: offset out of range");
  // 6-tap reach: rows -2 .. h+2, columns -2 .. w+2 (common/inter_prediction.c:146-176); none for a phase of 0
  const int rx0 = xoff ? 2 : 0, rx1 = xoff ? 3 : 0, ry0 = yoff ? 2 : 0, ry1 = yoff ? 3 : 0;
  put_window(k.ref, 0, ip, istride, -rx0, -ry0, width + rx1, height + ry1);
  thorcuda_pred_task t;
  memset(&t, 0, sizeof(t));
  t.w = (uint8_t)width; t.h = (uint8_t)height; t.ref0 = 1; t.ref1 = -1; t.mv0x = (int16_t)xoff; t.mv0y = (int16_t)yoff; t.dst = 2;
  const thorcuda_pic pics[3] = {k.org, k.ref, k.dst};
  must(thorcuda_predict_blocks(pics, 3, &t, 1, bipred ? 1 : 0, 1));
  get_window(k.dst, 0, qp, qstride, width, height);
  must(thorcuda_sync());
}

void get_inter_prediction_chroma_simd(int width, int height, int xoff, int yoff, unsigned char *THOR_RESTRICT qp, int qstride,
                                      const unsigned char *THOR_RESTRICT ip, int istride) {
  Compat &k = compat();
  check_block("get_inter_prediction_chroma_simd", width, height, 2, 32, 2);
  if (xoff < 0 || xoff > 7 || yoff < 0 || yoff > 7) die("libthorcuda: get_inter_prediction_chroma_simd: offset out of range");
  // 4-tap reach: rows -1 .. h+1, columns -1 .. w+1 (common/inter_prediction.c:97-115)
  const int rx0 = xoff ? 1 : 0, rx1 = xoff ? 2 : 0, ry0 = yoff ? 1 : 0, ry1 = yoff ? 2 : 0;
  put_window(k.ref, 1, ip, istride, -rx0, -ry0, width + rx1, height + ry1);
  thorcuda_pred_task t;
  memset(&t, 0, sizeof(t));
  // the batched kernel derives the chroma block from the luma one: luma 2w x 2h, the mv is used at 1/8 chroma precision
  t.w = (uint8_t)(2 * width); t.h = (uint8_t)(2 * height); t.ref0 = 1; t.ref1 = -1; t.mv0x = (int16_t)xoff; t.mv0y = (int16_t)yoff; t.dst = 2;
  const thorcuda_pic pics[3] = {k.org, k.ref, k.dst};
  must(thorcuda_predict_blocks(pics, 3, &t, 1, 0, 2));
  get_window(k.dst, 1, qp, qstride, width, height);
  must(thorcuda_sync());
}

// ---- transforms ----
void transform_simd(const int16_t *block, int16_t *coeff, int size, int fast) { transform_any(block, coeff, size, fast, 0); }
void inverse_transform_simd(const int16_t *coeff, int16_t *block, int size) { transform_any(coeff, block, size, 0, 1); }
void transform(const int16_t *block, int16_t *coeff, int size, int fast) { transform_any(block, coeff, size, fast, 0); }
void inverse_transform(const int16_t *coeff, int16_t *block, int size) { transform_any(coeff, block, size, 0, 1); }

// ---- CLPF ----
void clpf_block4(const uint8_t *src, uint8_t *dst, int sstride, int dstride, int x0, int y0, int width, int height) {
  clpf_block_n(src, dst, sstride, dstride, x0, y0, 4, width, height);
}
void clpf_block8(const uint8_t *src, uint8_t *dst, int sstride, int dstride, int x0, int y0, int width, int height) {
  clpf_block_n(src, dst, sstride, dstride, x0, y0, 8, width, height);
}
void detect_clpf_simd(const uint8_t *rec, const uint8_t *org, int x0, int y0, int width, int height, int so, int stride,
                      int *sum0, int *sum1) {
  Compat &k = compat();
  const int left = x0 & ~63, top = y0 & ~63;                       // MAX_BLOCK_SIZE, enc/encode_block.c:3038-3041
  const int right = min(width - 1, left + 63), bottom = min(height - 1, top + 63);
  uint8_t *pin = (uint8_t *)k.pin.p;
  gather_clpf(pin, rec, stride, x0, y0, 8, left, top, right, bottom);
  for (int y = 0; y < 8; y++) memcpy(pin + 128 + 8 * y, org + (ptrdiff_t)(y0 + y) * so + x0, 8);
  memset(pin + 192, 0, 8);
  uint8_t *d = (uint8_t *)k.d_in.p;
  must_cuda(cudaMemcpyAsync(d, pin, 200, cudaMemcpyHostToDevice, ctx().stream));
  k_clpf_small<<<1, 64, 0, ctx().stream>>>(d, nullptr, d + 128, 8, (int *)(d + 192));
  ctx().launches++;
  must_cuda(cudaGetLastError());
  must_cuda(cudaMemcpyAsync(pin + 256, d + 192, 8, cudaMemcpyDeviceToHost, ctx().stream));
  must(thorcuda_sync());
  *sum0 += ((int *)(pin + 256))[0];
  *sum1 += ((int *)(pin + 256))[1];
}

// ---- rounding block average ----
void block_avg_simd(uint8_t *p, uint8_t *r0, uint8_t *r1, int sp, int s0, int s1, int width, int height) {
  Compat &k = compat();
  check_block("block_avg_simd", width, height, 2, 64, 2);
  uint8_t *pin = (uint8_t *)k.pin.p;
  const int n = width * height;
  for (int y = 0; y < height; y++) {
    memcpy(pin + y * width, r0 + (ptrdiff_t)y * s0, width);
    memcpy(pin + n + y * width, r1 + (ptrdiff_t)y * s1, width);
  }
  uint8_t *d = (uint8_t *)k.d_in.p, *o = (uint8_t *)k.d_out.p;
  must_cuda(cudaMemcpyAsync(d, pin, 2 * n, cudaMemcpyHostToDevice, ctx().stream));
  k_block_avg<<<1, 256, 0, ctx().stream>>>(d, d + n, o, n);
  ctx().launches++;
  must_cuda(cudaGetLastError());
  must_cuda(cudaMemcpyAsync(pin + 2 * n, o, n, cudaMemcpyDeviceToHost, ctx().stream));
  must(thorcuda_sync());
  for (int y = 0; y < height; y++) memcpy(p + (ptrdiff_t)y * sp, pin + 2 * n + y * width, width);
}

}  // extern "C"
