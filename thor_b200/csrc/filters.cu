// libthorcuda: in-loop filters.  One fused kernel per frame: deblock Y/U/V -> CLPF (with the encoder's
// per-superblock decision) -> write reconstructed picture and/or the interior of the next reference
// picture; a second small kernel replicates the borders.
//
// Restates (bit-exactly) reference common/common_frame.c:46-321 (deblock_frame_y/_uv), :485-557
// (clpf_frame), common/common_block.c:180-197 (clpf_block), enc/encode_frame.c:50-63 (clpf_decision),
// enc/encode_block.c:3036-3057 (detect_clpf), common/common_frame.c:405-483 (pad / create_reference).
//
// Work decomposition: one CTA per 64x64 superblock (SB).  The SB plus the halo the two deblocking
// passes need (8 rows above/below, 16 columns left/right so that loads stay 16-byte aligned) is staged
// in shared memory with 128-bit loads; both edge passes and CLPF run on the staged tile; results leave
// with 128-bit stores.  Every picture byte is read from HBM once (halo re-reads hit L2) and written once.
#include <string.h>
#include "internal.cuh"

namespace thor {

// ---- packed side information: 2 bytes per 4x4 luma unit --------------------------------------
// bits 0-2 mode | 3 cbp.y | 4 cbp.u | 5 cbp.v | 6-7 log2(size)-3 | 8 tb_split | 9-10 pb_part | 11 any |mv comp| >= 4
enum { BI_MODE = 7, BI_CBPY = 8, BI_CBPU = 16, BI_CBPV = 32, BI_TBSPLIT = 256, BI_MVLARGE = 2048 };
__device__ __forceinline__ int bi_mode(unsigned v) { return v & BI_MODE; }
__device__ __forceinline__ int bi_size(unsigned v) { return 8 << ((v >> 6) & 3); }
__device__ __forceinline__ int bi_part(unsigned v) { return (v >> 9) & 3; }

// deblock_data_t (common/types.h:127-135) viewed as 11 little-endian 32-bit words; packed on the host
// while the record is copied into the pinned staging buffer (22x fewer bytes cross PCIe).
static inline uint16_t pack_blockinfo(const int32_t *e) {
  const int mode = e[0], cy = e[1], cu = e[2], cv = e[3], sz = e[4] & 0xff, tb = (e[4] >> 8) & 0xff, part = e[5];
  auto big = [](int packed) {
    const int x = (int16_t)(packed & 0xffff), y = (int16_t)(packed >> 16);   // NEW_MV_TEST, common_frame.c:119-120
    return abs(x) >= 4 || abs(y) >= 4;
  };
  const int lg = sz >= 64 ? 3 : sz >= 32 ? 2 : sz >= 16 ? 1 : 0;
  const unsigned v = (mode & 7) | (cy ? BI_CBPY : 0) | (cu ? BI_CBPU : 0) | (cv ? BI_CBPV : 0) | (lg << 6) | (tb ? BI_TBSPLIT : 0) |
                     ((part & 3) << 9) | ((big(e[6]) || big(e[7])) ? BI_MVLARGE : 0);
  return (uint16_t)v;
}

constexpr int LF_THREADS = 256;
constexpr int LY_ROWS = 80, LY_PITCH = 96;   // luma tile: rows y0-8..y0+71, cols x0-16..x0+79
constexpr int LC_ROWS = 40, LC_PITCH = 64;   // chroma tile: rows cy0-4..cy0+35, cols cx0-16..cx0+47
constexpr int LI_ROWS = 20, LI_COLS = 24;    // side info units: rows (y0-8)/4.., cols (x0-16)/4..

struct LfParams {
  PicView src, dstA, dstB, org;     // dstB.p[0]==nullptr: no reference output; org.p[0]==nullptr: decoder mode
  const uint16_t *info;             // packed side info, bw units per row
  int bw;
  int beta, tc_y, tc_c;
  int stages;
  const uint8_t *bits_in;
  uint8_t *bits_out;
  int nsbx_full, nsby_full;         // floor(w/64), floor(h/64): SBs CLPF may touch (common_frame.c:496-497)
  PicView *self;                    // entry of rec in the device picture table: updated to dstA (the buffers flip)
};

// CLPF of 4 packed samples.  x = centre, a/b/c/d = up/left/right/down neighbours (already clamped).
// delta = (#nbr > x  > 2) - (#nbr < x > 2), per byte (common_block.c:186-194).
__device__ __forceinline__ uint32_t clpf4(uint32_t x, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  uint32_t r = 0;
#pragma unroll
  for (int k = 0; k < 4; k++) {
    const int X = (x >> (8 * k)) & 255, A = (a >> (8 * k)) & 255, B = (b >> (8 * k)) & 255, C = (c >> (8 * k)) & 255,
              D = (d >> (8 * k)) & 255;
    const int up = (A > X) + (B > X) + (C > X) + (D > X), dn = (A < X) + (B < X) + (C < X) + (D < X);
    r |= (uint32_t)((X + (up > 2) - (dn > 2)) & 255) << (8 * k);
  }
  return r;
}

__device__ __forceinline__ uint32_t ssd4(uint32_t o, uint32_t x) {
  uint32_t s = 0;
#pragma unroll
  for (int k = 0; k < 4; k++) {
    const int d = (int)((o >> (8 * k)) & 255) - (int)((x >> (8 * k)) & 255);
    s += d * d;
  }
  return s;
}

// CLPF of one 16-sample row chunk held in a tile of pitch P; (r,c16) position inside the SB-local
// window of extent n x n (64 luma / 32 chroma) whose origin is at tile[0].  Neighbours outside the
// window are replaced by the centre sample.
template <int P>
__device__ __forceinline__ uint4 clpf_chunk(const uint8_t *win, int r, int c, int n, uint4 &centre) {
  const uint8_t *row = win + r * P + c;
  const uint4 x = *reinterpret_cast<const uint4 *>(row);
  const uint4 up = r == 0 ? x : *reinterpret_cast<const uint4 *>(row - P);
  const uint4 dn = r == n - 1 ? x : *reinterpret_cast<const uint4 *>(row + P);
  const uint32_t lw = c == 0 ? (x.x & 255) : row[-1];              // sample left of the chunk
  const uint32_t rw = c + 16 == n ? (x.w >> 24) : row[16];         // sample right of the chunk
  // left neighbours: shift the row right by one byte; right neighbours: shift left by one byte
  uint4 l, rr;
  l.x = (x.x << 8) | lw;            l.y = __funnelshift_l(x.x, x.y, 8);
  l.z = __funnelshift_l(x.y, x.z, 8); l.w = __funnelshift_l(x.z, x.w, 8);
  rr.x = __funnelshift_r(x.x, x.y, 8); rr.y = __funnelshift_r(x.y, x.z, 8);
  rr.z = __funnelshift_r(x.z, x.w, 8); rr.w = (x.w >> 8) | (rw << 24);
  centre = x;
  uint4 f;
  f.x = clpf4(x.x, up.x, l.x, rr.x, dn.x);
  f.y = clpf4(x.y, up.y, l.y, rr.y, dn.y);
  f.z = clpf4(x.z, up.z, l.z, rr.z, dn.z);
  f.w = clpf4(x.w, up.w, l.w, rr.w, dn.w);
  return f;
}

__global__ void __launch_bounds__(LF_THREADS) k_loopfilter(const LfParams *__restrict__ jobs) {
  __shared__ LfParams P;
  __shared__ __align__(16) uint8_t sY[LY_ROWS * LY_PITCH];
  __shared__ __align__(16) uint8_t sC[2][LC_ROWS * LC_PITCH];
  __shared__ uint16_t sI[LI_ROWS * LI_COLS];
  __shared__ uint8_t sFlag[192];
  __shared__ int sRed[2][LF_THREADS / 32];
  __shared__ int sOn;

  const int tid = threadIdx.x;
  if (tid < (int)(sizeof(LfParams) / 4)) reinterpret_cast<uint32_t *>(&P)[tid] = reinterpret_cast<const uint32_t *>(jobs + blockIdx.z)[tid];
  __syncthreads();
  const int W = P.src.w, H = P.src.h;
  const int x0 = blockIdx.x * 64, y0 = blockIdx.y * 64;
  if (x0 >= W || y0 >= H) return;                 // frames of a batch may differ in size
  if (blockIdx.x == 0 && blockIdx.y == 0 && tid < (int)(sizeof(PicView) / 4))
    reinterpret_cast<uint32_t *>(P.self)[tid] = reinterpret_cast<const uint32_t *>(&P.dstA)[tid];
  const int tx0 = x0 - 16, ty0 = y0 - 8;           // luma tile origin in the frame
  const int cx0 = x0 >> 1, cy0 = y0 >> 1;
  const int tcx0 = cx0 - 16, tcy0 = cy0 - 4;       // chroma tile origin

  // ---------------- stage tile ----------------
  for (int i = tid; i < LY_ROWS * (LY_PITCH / 16); i += LF_THREADS) {
    const int r = i / (LY_PITCH / 16), c = i % (LY_PITCH / 16);
    const int y = ty0 + r, x = tx0 + c * 16;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (y >= 0 && y < H && x >= 0 && x < W) v = *reinterpret_cast<const uint4 *>(P.src.p[0] + (size_t)y * P.src.stride[0] + x);
    *reinterpret_cast<uint4 *>(&sY[r * LY_PITCH + c * 16]) = v;
  }
  for (int i = tid; i < 2 * LC_ROWS * (LC_PITCH / 16); i += LF_THREADS) {
    const int pl = i / (LC_ROWS * (LC_PITCH / 16)), j = i % (LC_ROWS * (LC_PITCH / 16));
    const int r = j / (LC_PITCH / 16), c = j % (LC_PITCH / 16);
    const int y = tcy0 + r, x = tcx0 + c * 16;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (y >= 0 && y < H / 2 && x >= 0 && x < W / 2)
      v = *reinterpret_cast<const uint4 *>(P.src.p[1 + pl] + (size_t)y * P.src.stride[1] + x);
    *reinterpret_cast<uint4 *>(&sC[pl][r * LC_PITCH + c * 16]) = v;
  }
  for (int i = tid; i < LI_ROWS * LI_COLS; i += LF_THREADS) {
    const int r = i / LI_COLS, c = i % LI_COLS;
    const int uy = (ty0 >> 2) + r, ux = (tx0 >> 2) + c;
    uint16_t v = 0;
    if (uy >= 0 && uy < (H >> 2) && ux >= 0 && ux < (W >> 2)) v = P.info[(size_t)uy * P.bw + ux];
    sI[i] = v;
  }
  __syncthreads();

  // side info of the unit containing frame position (x,y)
  auto unit = [&](int x, int y) -> unsigned { return sI[((y - ty0) >> 2) * LI_COLS + ((x - tx0) >> 2)]; };

  // ---------------- luma: vertical edges (filter across x) ----------------
  if (P.stages & THORCUDA_LF_DEBLOCK_Y) {
    // decision per (edge, 8-row block, 4-row segment): 9 x 10 x 2
    for (int i = tid; i < 180; i += LF_THREADS) {
      const int seg = i & 1, pair = i >> 1, e = pair % 9, b = pair / 9;
      const int x = x0 + 8 * e, y = ty0 + 8 * b;
      uint8_t on = 0;
      if (x > 0 && x < W && y >= 0 && y < H) {
        const uint8_t *s2 = &sY[(8 * b + 2) * LY_PITCH + (x - tx0)], *s5 = s2 + 3 * LY_PITCH;
        const int d = abs(s2[-2] - s2[-1]) + abs(s2[1] - s2[0]) + abs(s5[-2] - s5[-1]) + abs(s5[1] - s5[0]);
        const unsigned q = unit(x, y + 4 * seg), p = unit(x - 4, y + 4 * seg);
        int qs = bi_size(q);
        if (((q & BI_TBSPLIT) || bi_part(q) >= 2) && qs > 8) qs >>= 1;        // PART_VER(2) or PART_QUAD(3)
        const bool strong = ((p | q) & (BI_MVLARGE | BI_CBPY)) || bi_mode(p) == 1 || bi_mode(q) == 1;
        on = d < P.beta && (x & (qs - 1)) == 0 && strong;
      }
      sFlag[i] = on;
    }
    __syncthreads();
    for (int i = tid; i < 720; i += LF_THREADS) {
      const int ln = i & 7, pair = i >> 3, e = pair % 9, b = pair / 9;
      if (!sFlag[pair * 2 + (ln >> 2)]) continue;
      uint8_t *s = &sY[(8 * b + ln) * LY_PITCH + (x0 + 8 * e - tx0)];
      const int p1 = s[-2], p0 = s[-1], q0 = s[0], q1 = s[1];
      const int delta = clampi((18 * (q0 - p0) - 6 * (q1 - p1) + 16) >> 5, -P.tc_y, P.tc_y);
      s[-2] = (uint8_t)clampi(p1 + delta / 2, 0, 255);
      s[-1] = (uint8_t)clampi(p0 + delta, 0, 255);
      s[0] = (uint8_t)clampi(q0 - delta, 0, 255);
      s[1] = (uint8_t)clampi(q1 - delta / 2, 0, 255);
    }
    __syncthreads();
    // ---------------- luma: horizontal edges (filter across y) ----------------
    for (int i = tid; i < 144; i += LF_THREADS) {
      const int seg = i & 1, pair = i >> 1, g = pair & 7, f = pair >> 3;
      const int x = x0 + 8 * g, y = y0 + 8 * f;
      uint8_t on = 0;
      if (y > 0 && y < H && x < W) {
        const uint8_t *s2 = &sY[(y - ty0) * LY_PITCH + (x - tx0) + 2], *s5 = s2 + 3;
        const int d = abs(s2[-2 * LY_PITCH] - s2[-LY_PITCH]) + abs(s2[LY_PITCH] - s2[0]) +
                      abs(s5[-2 * LY_PITCH] - s5[-LY_PITCH]) + abs(s5[LY_PITCH] - s5[0]);
        const unsigned q = unit(x + 4 * seg, y), p = unit(x + 4 * seg, y - 4);
        int qs = bi_size(q);
        const int part = bi_part(q);
        if (((q & BI_TBSPLIT) || part == 1 || part == 3) && qs > 8) qs >>= 1;  // PART_HOR(1) or PART_QUAD(3)
        const bool strong = ((p | q) & (BI_MVLARGE | BI_CBPY)) || bi_mode(p) == 1 || bi_mode(q) == 1;
        on = d < P.beta && (y & (qs - 1)) == 0 && strong;
      }
      sFlag[i] = on;
    }
    __syncthreads();
    for (int i = tid; i < 576; i += LF_THREADS) {
      const int col = i & 63, f = i >> 6;
      if (!sFlag[((f << 3) + (col >> 3)) * 2 + ((col >> 2) & 1)]) continue;
      uint8_t *s = &sY[(y0 + 8 * f - ty0) * LY_PITCH + (x0 - tx0) + col];
      const int p1 = s[-2 * LY_PITCH], p0 = s[-LY_PITCH], q0 = s[0], q1 = s[LY_PITCH];
      const int delta = clampi((18 * (q0 - p0) - 6 * (q1 - p1) + 16) >> 5, -P.tc_y, P.tc_y);
      s[-2 * LY_PITCH] = (uint8_t)clampi(p1 + delta / 2, 0, 255);
      s[-LY_PITCH] = (uint8_t)clampi(p0 + delta, 0, 255);
      s[0] = (uint8_t)clampi(q0 - delta, 0, 255);
      s[LY_PITCH] = (uint8_t)clampi(q1 - delta / 2, 0, 255);
    }
  }

  // ---------------- chroma: both planes, vertical then horizontal edges ----------------
  if (P.stages & THORCUDA_LF_DEBLOCK_UV) {
    for (int i = tid; i < 2 * 360; i += LF_THREADS) {
      const int pl = i / 360, j = i % 360, ln = j & 3, pair = j >> 2, e = pair % 9, b = pair / 9;
      const int x = x0 + 8 * e, y = ty0 + 8 * b;          // luma coordinates of the 8x8 block edge
      if (!(x > 0 && x < W && y >= 0 && y < H)) continue;
      const unsigned q = unit(x, y), p = unit(x - 4, y);
      if ((x & (bi_size(q) - 1)) != 0 || !(bi_mode(p) == 1 || bi_mode(q) == 1)) continue;
      uint8_t *s = &sC[pl][(4 * b + ln) * LC_PITCH + ((x >> 1) - tcx0)];
      const int p1 = s[-2], p0 = s[-1], q0 = s[0], q1 = s[1];
      const int delta = clampi((4 * (q0 - p0) + (p1 - q1) + 4) >> 3, -P.tc_c, P.tc_c);
      s[-1] = (uint8_t)clampi(p0 + delta, 0, 255);
      s[0] = (uint8_t)clampi(q0 - delta, 0, 255);
    }
    __syncthreads();
    for (int i = tid; i < 2 * 288; i += LF_THREADS) {
      const int pl = i / 288, j = i % 288, col = j & 31, f = j >> 5;
      const int x = x0 + 8 * (col >> 2), y = y0 + 8 * f;
      if (!(y > 0 && y < H && x < W)) continue;
      const unsigned q = unit(x, y), p = unit(x, y - 4);
      if ((y & (bi_size(q) - 1)) != 0 || !(bi_mode(p) == 1 || bi_mode(q) == 1)) continue;
      uint8_t *s = &sC[pl][((y >> 1) - tcy0) * LC_PITCH + (cx0 - tcx0) + col];
      const int p1 = s[-2 * LC_PITCH], p0 = s[-LC_PITCH], q0 = s[0], q1 = s[LC_PITCH];
      const int delta = clampi((4 * (q0 - p0) + (p1 - q1) + 4) >> 3, -P.tc_c, P.tc_c);
      s[-LC_PITCH] = (uint8_t)clampi(p0 + delta, 0, 255);
      s[0] = (uint8_t)clampi(q0 - delta, 0, 255);
    }
  }
  __syncthreads();

  // ---------------- CLPF + output ----------------
  const uint8_t *winY = &sY[8 * LY_PITCH + 16];            // SB origin inside the luma tile
  const bool clpf_sb = (P.stages & THORCUDA_LF_CLPF) && (int)blockIdx.x < P.nsbx_full && (int)blockIdx.y < P.nsby_full;

  // this thread's luma chunk: row r, columns 16c..16c+15 of the SB
  const int r = tid >> 2, c = (tid & 3) * 16;
  uint4 xY, fY;
  fY = clpf_chunk<LY_PITCH>(winY, r, c, 64, xY);
  // side info of the two 8x8 blocks the chunk covers
  const unsigned ia = unit(x0 + c, y0 + r), ib = unit(x0 + c + 8, y0 + r);
  const bool ya = bi_mode(ia) != 3 && (ia & BI_CBPY), yb = bi_mode(ib) != 3 && (ib & BI_CBPY);

  int on = 0;
  if (clpf_sb) {
    // candidate test over the 64 8x8 blocks (first unit of each), common_frame.c:503-511
    int cand = 0;
    if (tid < 64) {
      const unsigned v = unit(x0 + (tid & 7) * 8, y0 + (tid >> 3) * 8);
      cand = bi_mode(v) != 3 && (v & (BI_CBPY | BI_CBPU | BI_CBPV));
    }
    cand = __syncthreads_or(cand);
    const int sb = blockIdx.y * P.nsbx_full + blockIdx.x;
    if (cand) {
      if (P.org.p[0]) {
        // encoder: SSD before/after over coded non-bipred luma blocks (encode_frame.c:50-63)
        int s0 = 0, s1 = 0;
        if (ya || yb) {
          const uint4 o = *reinterpret_cast<const uint4 *>(P.org.p[0] + (size_t)(y0 + r) * P.org.stride[0] + x0 + c);
          if (ya) { s0 += ssd4(o.x, xY.x) + ssd4(o.y, xY.y); s1 += ssd4(o.x, fY.x) + ssd4(o.y, fY.y); }
          if (yb) { s0 += ssd4(o.z, xY.z) + ssd4(o.w, xY.w); s1 += ssd4(o.z, fY.z) + ssd4(o.w, fY.w); }
        }
#pragma unroll
        for (int m = 16; m > 0; m >>= 1) { s0 += __shfl_xor_sync(0xffffffffu, s0, m); s1 += __shfl_xor_sync(0xffffffffu, s1, m); }
        if ((tid & 31) == 0) { sRed[0][tid >> 5] = s0; sRed[1][tid >> 5] = s1; }
        __syncthreads();
        if (tid == 0) {
          int a = 0, b = 0;
          for (int k = 0; k < LF_THREADS / 32; k++) { a += sRed[0][k]; b += sRed[1][k]; }
          sOn = b < a;
        }
        __syncthreads();
        on = sOn;
      } else {
        on = P.bits_in[sb] != 0;
      }
    }
    if (tid == 0 && P.bits_out) P.bits_out[sb] = cand ? (uint8_t)on : 0xFF;
  }

  // luma out
  {
    uint4 o = xY;
    if (on) {
      if (ya) { o.x = fY.x; o.y = fY.y; }
      if (yb) { o.z = fY.z; o.w = fY.w; }
    }
    const int y = y0 + r, x = x0 + c;
    if (y < H && x < W) {
      *reinterpret_cast<uint4 *>(P.dstA.p[0] + (size_t)y * P.dstA.stride[0] + x) = o;
      if (P.dstB.p[0]) *reinterpret_cast<uint4 *>(P.dstB.p[0] + (size_t)y * P.dstB.stride[0] + x) = o;
    }
  }
  // chroma out: 2 planes x 32 rows x 2 chunks = 128 threads
  if (tid < 128) {
    const int pl = tid >> 6, rr = (tid >> 1) & 31, cc = (tid & 1) * 16;
    const uint8_t *winC = &sC[pl][4 * LC_PITCH + 16];
    uint4 xC, fC;
    fC = clpf_chunk<LC_PITCH>(winC, rr, cc, 32, xC);
    uint4 o = xC;
    if (on) {
      const uint32_t bit = pl ? BI_CBPV : BI_CBPU;
      // four 8x8 luma blocks (4 chroma samples each) per 16-sample chunk
      const unsigned i0 = unit(x0 + 2 * cc, y0 + 2 * rr), i1 = unit(x0 + 2 * cc + 8, y0 + 2 * rr),
                     i2 = unit(x0 + 2 * cc + 16, y0 + 2 * rr), i3 = unit(x0 + 2 * cc + 24, y0 + 2 * rr);
      if (bi_mode(i0) != 3 && (i0 & bit)) o.x = fC.x;
      if (bi_mode(i1) != 3 && (i1 & bit)) o.y = fC.y;
      if (bi_mode(i2) != 3 && (i2 & bit)) o.z = fC.z;
      if (bi_mode(i3) != 3 && (i3 & bit)) o.w = fC.w;
    }
    const int y = cy0 + rr, x = cx0 + cc;
    if (y < H / 2 && x < W / 2) {
      *reinterpret_cast<uint4 *>(P.dstA.p[1 + pl] + (size_t)y * P.dstA.stride[1] + x) = o;
      if (P.dstB.p[0]) *reinterpret_cast<uint4 *>(P.dstB.p[1 + pl] + (size_t)y * P.dstB.stride[1] + x) = o;
    }
  }
}

// Border replication (pad_yuv_frame, common_frame.c:405-462).  One thread per 4 border bytes; only border bytes
// are enumerated: item < nA -> the top / bottom bands (2*pad full padded rows), else the left / right bands of
// the h picture rows.  grid.y = plane, grid.z = job; grid.x covers the largest plane of the batch.
__device__ __forceinline__ void pad_items(const PicView &v, int pl) {
  const int pad = v.pad[pl], w = pl ? v.w / 2 : v.w, h = pl ? v.h / 2 : v.h, stride = v.stride[pl];
  if (!pad) return;
  const int tw4 = (w + 2 * pad) >> 2, pw4 = pad >> 2;  // words per padded row / per side band (w, pad: multiples of 4)
  const int nA = 2 * pad * tw4, nB = h * 2 * pw4;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nA + nB; i += gridDim.x * blockDim.x) {
    int row, x;
    if (i < nA) {
      const int r = i / tw4;
      x = (i - r * tw4) * 4 - pad;
      row = r < pad ? r - pad : h + (r - pad);
    } else {
      const int j = i - nA, r = j / (2 * pw4), c = j - r * 2 * pw4;
      row = r;
      x = c < pw4 ? c * 4 - pad : w + (c - pw4) * 4;
    }
    const uint8_t *src = v.p[pl] + (ptrdiff_t)min(max(row, 0), h - 1) * stride;
    uint32_t val;
    if (x < 0) val = src[0] * 0x01010101u;
    else if (x >= w) val = src[w - 1] * 0x01010101u;
    else val = *reinterpret_cast<const uint32_t *>(src + x);
    *reinterpret_cast<uint32_t *>(v.p[pl] + (ptrdiff_t)row * stride + x) = val;
  }
}
__global__ void __launch_bounds__(256) k_pad(PicView v) { pad_items(v, blockIdx.y); }
__global__ void __launch_bounds__(256) k_pad_jobs(const LfParams *__restrict__ jobs) {
  const LfParams &J = jobs[blockIdx.z];
  if (!J.dstB.p[0]) return;
  pad_items(J.dstB, blockIdx.y);
}

}  // namespace thor

using namespace thor;

extern "C" {

int thorcuda_blockinfo_upload_signals(thorcuda_pic hrec, const void *deblock_data, uint8_t *sb_signals) {
  THOR_NEED_CTX();
  Pic *rec = get_pic(hrec);
  if (!rec) return fail("blockinfo: bad picture handle %d", hrec);
  if (!deblock_data) return fail("blockinfo: NULL side information");
  const int n = (rec->w / 4) * (rec->h / 4);
  if (!rec->info) THOR_CUDA(cudaMalloc((void **)&rec->info, (size_t)n * 2));
  uint16_t *st = (uint16_t *)stage_acquire((size_t)n * 2);
  if (!st) return fail("blockinfo: staging allocation failed");
  // 44 bytes in, 2 bytes out per 4x4 unit, on the calling thread: helper threads were measured and made the decoder
  // SLOWER (the reference clears this array at the start of every frame; once other cores have read it, that memset
  // pays for the cache lines: +0.7 ms per 1080p frame against 0.13 ms saved here)
  const int32_t *e = (const int32_t *)deblock_data;
  for (int i = 0; i < n; i++, e += 11) st[i] = pack_blockinfo(e);
  if (sb_signals) {
    // does clpf_frame ask its decision callback about a superblock (common/common_frame.c:507-521)?  -- any 8x8 block
    // that is not bi-predicted and has coded coefficients in some plane; read from the packed records
    const int upr = rec->w / 4, nh = rec->w / 64, nv = rec->h / 64;
    for (int k = 0; k < nv; k++)
      for (int l = 0; l < nh; l++) {
        int sig = 0;
        for (int m = 0; m < 16 && !sig; m += 2) {
          const uint16_t *row = st + (size_t)(k * 16 + m) * upr + l * 16;
          for (int q = 0; q < 16; q += 2) sig |= (row[q] & 7) != 3 /* MODE_BIPRED */ && (row[q] & (BI_CBPY | BI_CBPU | BI_CBPV));
        }
        sb_signals[k * nh + l] = (uint8_t)sig;
      }
  }
  if (stage_commit_copy(rec->info, st, (size_t)n * 2)) return -1;
  rec->info_ok = true;
  return 0;
}

int thorcuda_blockinfo_upload(thorcuda_pic hrec, const void *deblock_data) {
  return thorcuda_blockinfo_upload_signals(hrec, deblock_data, nullptr);
}

int thorcuda_pad_frame(thorcuda_pic h) {
  THOR_NEED_CTX();
  Pic *p = get_pic(h);
  if (!p) return fail("bad picture handle %d", h);
  if (p->pad_y == 0 && p->pad_c == 0) return 0;
  const int items = (2 * p->pad_y * (p->w + 2 * p->pad_y) + p->h * 2 * p->pad_y) / 4;
  dim3 grid((items + 255) / 256, 3, 1);
  k_pad<<<grid, 256, 0, ctx().stream>>>(p->view());
  THOR_LAUNCHED();
  return 0;
}

int thorcuda_loopfilter_frames(int n, const thorcuda_lf_job *jobs) {
  THOR_NEED_CTX();
  Context &c = ctx();
  static const uint8_t beta[52] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15,
                                   16, 17, 18, 20, 22, 24, 26, 28, 30, 32, 34, 36, 38, 40, 42, 44, 46, 48, 50, 52, 54, 56, 58, 60, 62, 64};
  static const uint8_t tc[56] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 1, 1, 1, 1, 1, 2,
                                 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 5, 5, 6, 6, 7, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13, 14, 14};
  if (n <= 0) return 0;
  if (n > 16384) return fail("loopfilter: too many frames in one batch");
  // validate, size the per-SB bit scratch
  size_t bits_total = 0;
  int maxw = 0, maxh = 0, max_ref_rows = 0, max_pad_items = 0;
  for (int i = 0; i < n; i++) {
    const thorcuda_lf_job &j = jobs[i];
    Pic *rec = get_pic(j.rec), *org = j.org >= 0 ? get_pic(j.org) : nullptr, *ref = j.ref_out >= 0 ? get_pic(j.ref_out) : nullptr;
    if (!rec) return fail("loopfilter: bad rec handle");
    if (j.org >= 0 && !org) return fail("loopfilter: bad org handle");
    if (j.ref_out >= 0 && !ref) return fail("loopfilter: bad ref handle");
    if (rec->pad_y || rec->pad_c) return fail("loopfilter: rec must be an unpadded picture");
    if (j.qp < 0 || j.qp > 51) return fail("loopfilter: qp %d out of range", j.qp);
    if ((j.stages & (THORCUDA_LF_DEBLOCK_Y | THORCUDA_LF_DEBLOCK_UV | THORCUDA_LF_CLPF)) && !rec->info_ok)
      return fail("loopfilter: no side information uploaded for picture %d", j.rec);
    if ((org && (org->w != rec->w || org->h != rec->h)) || (ref && (ref->w != rec->w || ref->h != rec->h)))
      return fail("loopfilter: picture size mismatch");
    if ((j.stages & THORCUDA_LF_CLPF) && !org && !j.sb_bits_in && (rec->w / 64) * (rec->h / 64) > 0)
      return fail("loopfilter: decoder-mode CLPF needs sb_bits_in");
    for (int k = 0; k < i; k++) if (jobs[k].rec == j.rec) return fail("loopfilter: picture %d appears twice in one batch", j.rec);
    if (!rec->buf[1 - rec->cur]) THOR_CUDA(cudaMalloc((void **)&rec->buf[1 - rec->cur], rec->bytes + 256));
    bits_total += 2 * (size_t)((rec->w / 64) * (rec->h / 64) + 1);
    maxw = max(maxw, rec->w); maxh = max(maxh, rec->h);
    if (ref && (j.stages & THORCUDA_LF_MAKE_REF)) {
      max_ref_rows = max(max_ref_rows, ref->rows_y);
      max_pad_items = max(max_pad_items, (2 * ref->pad_y * (ref->w + 2 * ref->pad_y) + ref->h * 2 * ref->pad_y) / 4);
    }
  }
  uint8_t *d_bits = (uint8_t *)c.scratch.reserve(bits_total);
  LfParams *d_jobs = (LfParams *)c.lfjobs.reserve(sizeof(LfParams) * n);
  uint8_t *st = (uint8_t *)stage_acquire(sizeof(LfParams) * n + bits_total);
  if (!d_bits || !d_jobs || !st) return fail("loopfilter: allocation failed");
  LfParams *hj = (LfParams *)st;
  uint8_t *hbits = st + sizeof(LfParams) * n;
  size_t off = 0;
  bool any_bits_in = false;
  for (int i = 0; i < n; i++) {
    const thorcuda_lf_job &j = jobs[i];
    Pic *rec = get_pic(j.rec), *org = j.org >= 0 ? get_pic(j.org) : nullptr, *ref = j.ref_out >= 0 ? get_pic(j.ref_out) : nullptr;
    const int nsbx = rec->w / 64, nsby = rec->h / 64, nsb = nsbx * nsby;
    LfParams &P = hj[i];
    memset(&P, 0, sizeof(P));
    P.src = rec->view(rec->cur);
    P.dstA = rec->view(1 - rec->cur);
    if (ref && (j.stages & THORCUDA_LF_MAKE_REF)) P.dstB = ref->view(); else { P.dstB = P.dstA; P.dstB.p[0] = nullptr; }
    if (org) P.org = org->view(); else { P.org = P.src; P.org.p[0] = nullptr; }
    P.info = rec->info;
    P.bw = rec->w / 4;
    P.beta = beta[j.qp]; P.tc_y = tc[j.qp]; P.tc_c = tc[h_chroma_qp[j.qp]];
    P.stages = j.stages;
    P.nsbx_full = nsbx; P.nsby_full = nsby;
    P.bits_in = nullptr; P.bits_out = nullptr;
    P.self = c.d_pics + j.rec;
    if ((j.stages & THORCUDA_LF_CLPF) && nsb > 0) {
      if (!org) {
        memcpy(hbits + off, j.sb_bits_in, nsb);
        P.bits_in = d_bits + off;
        any_bits_in = true;
      }
      P.bits_out = d_bits + off + nsb + 1;
    }
    off += 2 * (size_t)(nsb + 1);
  }
  // one staged copy carries the job table and (decoder mode) the signalled bits
  THOR_CUDA(cudaMemcpyAsync(d_jobs, hj, sizeof(LfParams) * n, cudaMemcpyHostToDevice, c.stream));
  if (any_bits_in) { if (stage_commit_copy(d_bits, hbits, bits_total)) return -1; }
  else if (stage_commit_copy(d_jobs, hj, 0)) return -1;
  dim3 grid((maxw + 63) / 64, (maxh + 63) / 64, n);
  k_loopfilter<<<grid, LF_THREADS, 0, c.stream>>>(d_jobs);
  THOR_LAUNCHED();
  off = 0;
  for (int i = 0; i < n; i++) {
    const thorcuda_lf_job &j = jobs[i];
    Pic *rec = get_pic(j.rec);
    const int nsb = (rec->w / 64) * (rec->h / 64);
    rec->cur ^= 1;                                   // the kernel updated the device picture table itself
    if ((j.stages & THORCUDA_LF_CLPF) && nsb > 0 && j.sb_bits_out)
      if (d2h(j.sb_bits_out, d_bits + off + nsb + 1, nsb)) return -1;
    off += 2 * (size_t)(nsb + 1);
  }
  if (max_ref_rows) {
    dim3 pg((max_pad_items + 255) / 256, 3, n);
    k_pad_jobs<<<pg, 256, 0, c.stream>>>(d_jobs);
    THOR_LAUNCHED();
  }
  return 0;
}

int thorcuda_loopfilter_frame(thorcuda_pic hrec, thorcuda_pic horg, thorcuda_pic href, int qp, int stages,
                              const uint8_t *sb_bits_in, uint8_t *sb_bits_out) {
  thorcuda_lf_job j;
  j.rec = hrec; j.org = horg; j.ref_out = href; j.qp = qp; j.stages = stages;
  j.sb_bits_in = sb_bits_in; j.sb_bits_out = sb_bits_out;
  return thorcuda_loopfilter_frames(1, &j);
}

}  // extern "C"
