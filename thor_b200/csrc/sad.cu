// libthorcuda: motion-search costs.
//
//  (1) k_sad_table : whole-frame integer-pel SAD tables for the 8/16/32/64 block levels in one pass
//                    (restates sad_calc, reference enc/encode_block.c:740-755, for every block and every
//                    displacement of the search window; larger levels are sums of the 8x8 level, which is
//                    exact because SAD is additive over disjoint sub-blocks).
//  (2) k_cost_batch: one warp per (block, candidate) task: sad_calc, widesad_calc (:757-780),
//                    sad_calc_fasthalf (:497-605), sad_calc_fastquarter (:609-738), ssd_calc (:782-797).
//
// Bound: integer ALU (VABSDIFF4.U8.ACC = 4 abs-diffs + accumulate per lane per instruction), not HBM.
#include <string.h>
#include "internal.cuh"

namespace thor {

struct SadJob {
  PicView org, ref;
  const int16_t *centres;       // device, (cx,cy) per SB, or nullptr
  uint16_t *t8, *t16;
  uint32_t *t32, *t64;
};

constexpr int SAD_THREADS = 256;

// acc + sum of |a.b[i] - b.b[i]| over the four bytes: ONE VABSDIFF4.U8.ACC (the intrinsic form
// __vsadu4(a,b)+acc is split by the compiler into VABSDIFF4 + IADD3, doubling the ALU-pipe work)
__device__ __forceinline__ uint32_t sad4(uint32_t a, uint32_t b, uint32_t acc) {
  uint32_t r;
  asm("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(acc));
  return r;
}

__device__ __forceinline__ void cp_async16(void *smem_dst, const void *gmem_src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait1() { asm volatile("cp.async.wait_group 1;" ::: "memory"); }

// Geometry of one launch configuration (search range R, a multiple of 4).
//   n x n displacements per block.  A warp pass covers UPW block columns x DXL consecutive horizontal
//   displacements (lane = u * DXL + d), so that a store instruction writes DXL consecutive table entries of a
//   block (one aligned 64-byte row for R = 16).
template <int R>
struct SadCfg {
  static constexpr int n = 2 * R, nn = n * n;   // displacements -R .. R-1 in both directions
  // 32 lanes per block column only for R = 16: at R = 32 the pending-half-sum rows would not fit in shared memory
  static constexpr int DXL = R == 16 ? 32 : (2 * R) % 16 == 0 ? 16 : 8;
  static constexpr int UPW = 32 / DXL;          // block columns per warp pass
  static constexpr int NB = 4 / UPW;            // column passes that cover a warp's 4 block columns
  static constexpr int NCH = 2 * R / DXL;       // displacement chunks
  static constexpr int win_rows = 64 + 2 * R;
  static constexpr int pitch = (64 + 2 * R + 15 + 15) & ~15;   // window rows start 16-byte aligned in global memory
  static constexpr int buf_bytes = 64 * 64 + win_rows * pitch; // one stage: original SB | reference window
  static constexpr int part_off = 2 * buf_bytes;               // uint32 [8 strips][nn]
  static constexpr int pair_off = part_off + 8 * nn * 4;       // uint16 [8 warps][nn] (UPW == 1 only)
  // Byte-phase copies of the window (R <= 16, where they fit beside two resident CTAs): copy k = the window shifted
  // left by k bytes, so that a lane finds the 4 samples at ANY column as one aligned word of copy (column & 3) and the
  // inner loop needs no funnel shifts (copy 0 is the staged window itself).  Copy k starts 32k bytes past a
  // 128-byte boundary: the 32 lanes of a pass (consecutive columns) then hit 32 different banks.
  static constexpr bool PH = R <= 16;
  static constexpr int win_bytes = (win_rows * pitch + 127) & ~127;
  static constexpr int ph_off = (pair_off + (UPW == 1 ? 8 * nn * 2 : 0) + 127) & ~127;
  static constexpr int smem_bytes = PH ? ph_off + 3 * win_bytes + 128 : ((ph_off + 15) & ~15);
  static_assert(buf_bytes % 128 == 0, "both stages must have the bank alignment the phase copies are laid out against");
};

// where one pass's results go
struct SadOut {
  uint16_t *t8T, *t8B, *t16;    // already offset to this lane's horizontal displacement
  uint32_t *part;               // this warp's 32x16 strip sums (+ displacement)
  uint16_t *pair;               // this warp's pending 16x16 half sums (+ displacement)
  bool okT, okB, ok16, lead;
  int b;                        // column pass index inside the warp's 4 block columns
};

// combine the 8x16 column sum v of this lane's block column into the 16x16 and 32x16 levels.
//   UPW 4 / 2: partner columns sit in other lanes of the same pass (shuffle);  UPW 1: they are other passes of
//   the SAME lane, so the running sums live in this warp's shared-memory rows (no synchronisation needed).
template <int UPW>
__device__ __forceinline__ void sad_reduce(uint32_t v, int e, const SadOut &O) {
  if (UPW == 4) {
    v += __shfl_xor_sync(0xffffffffu, v, 8);
    if (O.ok16) O.t16[e] = (uint16_t)v;
    v += __shfl_xor_sync(0xffffffffu, v, 16);
    if (O.lead) O.part[e] = v;
  } else if (UPW == 2) {
    v += __shfl_xor_sync(0xffffffffu, v, 16);
    if (O.ok16) O.t16[e] = (uint16_t)v;
    if (O.lead) O.part[e] = O.b ? O.part[e] + v : v;
  } else {
    if (!(O.b & 1)) {
      O.pair[e] = (uint16_t)v;
    } else {
      v += O.pair[e];
      if (O.ok16) O.t16[e] = (uint16_t)v;
      O.part[e] = O.b == 1 ? v : O.part[e] + v;
    }
  }
}

// One "pass": a lane owns the two vertically adjacent 8x8 blocks (block column bx, block rows 2p and 2p+1 of
// the SB) at ONE horizontal displacement and slides over all 2R vertical ones with 2 x 8 rotating
// accumulators.  The 8 reference rows that start / end the slide touch fewer displacement rows than the middle
// ones; FIRST / LAST skip those, so no abs-diff is computed that is not part of an output.
enum { BLK_FIRST = 0, BLK_MID = 1, BLK_LAST = 2 };

template <int MODE, int R, int UPW, bool PH>
__device__ __forceinline__ void sad_block(const uint32_t (&oT)[8][2], const uint32_t (&oB)[8][2], uint32_t (&aT)[8], uint32_t (&aB)[8],
                                          const uint8_t *col, int sh, const SadOut &O, int e0) {
  constexpr int n = SadCfg<R>::n, pitch = SadCfg<R>::pitch;
#pragma unroll
  for (int s = 0; s < (MODE == BLK_LAST ? 7 : 8); s++) {     // the last block: reference rows 2R .. 2R+6
    const uint32_t *rt = reinterpret_cast<const uint32_t *>(col + s * pitch);
    const uint32_t *rb = reinterpret_cast<const uint32_t *>(col + (s + 8) * pitch);
    uint32_t wT0, wT1, wB0, wB1;
    if (PH) {                                         // `col` points into the copy of this lane's byte phase
      wT0 = rt[0]; wT1 = rt[1]; wB0 = rb[0]; wB1 = rb[1];
    } else {
      const uint32_t t0 = rt[0], t1 = rt[1], t2 = rt[2], b0 = rb[0], b1 = rb[1], b2 = rb[2];
      wT0 = __funnelshift_r(t0, t1, sh); wT1 = __funnelshift_r(t1, t2, sh);
      wB0 = __funnelshift_r(b0, b1, sh); wB1 = __funnelshift_r(b1, b2, sh);
    }
#pragma unroll
    for (int r = 0; r < 8; r++) {
      if (MODE == BLK_FIRST && r > s) continue;      // displacement row j - r < 0
      if (MODE == BLK_LAST && r <= s) continue;      // displacement row j - r > 2R - 1
      const int k = (s - r) & 7;
      aT[k] = sad4(oT[r][0], wT0, aT[k]); aT[k] = sad4(oT[r][1], wT1, aT[k]);
      aB[k] = sad4(oB[r][0], wB0, aB[k]); aB[k] = sad4(oB[r][1], wB1, aB[k]);
    }
    if (MODE != BLK_FIRST || s == 7) {               // displacement row (s - 7) relative to this block is complete
      const int kc = (s + 1) & 7;
      const uint32_t vT = aT[kc], vB = aB[kc];
      const int e = e0 + (s - 7) * n;
      if (O.okT) O.t8T[e] = (uint16_t)vT;
      if (O.okB) O.t8B[e] = (uint16_t)vB;
      sad_reduce<UPW>(vT + vB, e, O);
      aT[kc] = 0; aB[kc] = 0;
    }
  }
}

struct SadTile {
  const uint8_t *sOrg, *sWin, *sPh;
  uint32_t *part;
  uint16_t *pair;
  uint16_t *t8, *t16;
  int W, H, x0, y0, o;
};

template <int R, int UPW>
__device__ __forceinline__ void sad_pass(const SadTile &C, int p, int bx, int dxi, int b, int warp, bool lead) {
  constexpr int n = SadCfg<R>::n, nn = SadCfg<R>::nn, pitch = SadCfg<R>::pitch;
  uint32_t oT[8][2], oB[8][2];
#pragma unroll
  for (int r = 0; r < 8; r++) {
    const uint32_t *a = reinterpret_cast<const uint32_t *>(C.sOrg + (16 * p + r) * 64 + bx * 8);
    const uint32_t *bb = reinterpret_cast<const uint32_t *>(C.sOrg + (16 * p + 8 + r) * 64 + bx * 8);
    oT[r][0] = a[0]; oT[r][1] = a[1]; oB[r][0] = bb[0]; oB[r][1] = bb[1];
  }
  // which outputs exist (frame edges) and where they go
  const int gx8 = C.x0 + bx * 8, gx16 = C.x0 + (bx & ~1) * 8, gyT = C.y0 + 16 * p, gyB = gyT + 8;
  const bool okx8 = gx8 + 8 <= C.W;
  const int bw8 = C.W >> 3, bw16 = C.W >> 4;
  SadOut O;
  O.okT = okx8 && gyT + 8 <= C.H; O.okB = okx8 && gyB + 8 <= C.H;
  // the lane that completes a 16x16 sum: the odd column when partners are other passes (UPW 1), the even one else
  O.ok16 = (UPW == 1 ? (bx & 1) : !(bx & 1)) && gx16 + 16 <= C.W && gyT + 16 <= C.H;
  O.lead = lead; O.b = b;
  O.t8T = C.t8 + ((size_t)(gyT >> 3) * bw8 + (gx8 >> 3)) * nn + dxi;
  O.t8B = C.t8 + ((size_t)(gyB >> 3) * bw8 + (gx8 >> 3)) * nn + dxi;
  O.t16 = C.t16 + ((size_t)(gyT >> 4) * bw16 + (gx16 >> 4)) * nn + dxi;
  O.part = C.part + (p * 2 + (bx >> 2)) * nn + dxi;
  O.pair = C.pair + warp * nn + dxi;
  const int off = bx * 8 + dxi + C.o;
  const int sh = (off & 3) * 8;
  const uint8_t *wbase = C.sWin;
  if (SadCfg<R>::PH && (off & 3)) wbase = C.sPh + (off & 3) * 32 + ((off & 3) - 1) * SadCfg<R>::win_bytes;
  const uint8_t *col = wbase + (16 * p) * pitch + (off & ~3);
  uint32_t aT[8], aB[8];
#pragma unroll
  for (int k = 0; k < 8; k++) { aT[k] = 0; aB[k] = 0; }
  sad_block<BLK_FIRST, R, UPW, SadCfg<R>::PH>(oT, oB, aT, aB, col, sh, O, 0);
  int e0 = 0;
#pragma unroll 1
  for (int jj = 8; jj < 2 * R; jj += 8) {
    col += 8 * pitch; e0 += 8 * n;
    sad_block<BLK_MID, R, UPW, SadCfg<R>::PH>(oT, oB, aT, aB, col, sh, O, e0);
  }
  col += 8 * pitch; e0 += 8 * n;
  sad_block<BLK_LAST, R, UPW, SadCfg<R>::PH>(oT, oB, aT, aB, col, sh, O, e0);
}

// Persistent kernel: gridDim.x CTAs walk the (job, SB row, SB column) tiles of the launch with a stride of
// gridDim.x.  Shared memory: two stages of {original SB 64x64 | reference window (64+2R rows)} filled by
// cp.async (the next tile's loads are in flight while the current tile is searched) | 32x16-strip sums
// [8][n*n] | pending 16x16 half sums [8 warps][n*n].
// Warp w owns block-row pair p = w & 3 and column half ch = w >> 2 (block columns 4ch..4ch+3): 32 equal passes
// per tile, 4 per warp.
template <int R>
__global__ void __launch_bounds__(SAD_THREADS, 2) k_sad_table(const SadJob *__restrict__ jobs, int gx, int gy, int njobs) {
  using Cfg = SadCfg<R>;
  constexpr int n = Cfg::n, nn = Cfg::nn, pitch = Cfg::pitch, win_rows = Cfg::win_rows;
  extern __shared__ __align__(128) uint8_t smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  uint32_t *part = reinterpret_cast<uint32_t *>(smem + Cfg::part_off);
  uint16_t *pair = reinterpret_cast<uint16_t *>(smem + Cfg::pair_off);
  const int per_job = gx * gy, total = per_job * njobs;

  struct Tile { int z, x0, y0, o, W, H; };
  // first valid tile at or after linear index L (frames of a batch may differ in size)
  auto find = [&](int L, Tile &t) -> int {
    for (; L < total; L += gridDim.x) {
      t.z = L / per_job;
      const int rem = L - t.z * per_job, sy = rem / gx, sx = rem - sy * gx;
      t.x0 = sx * 64; t.y0 = sy * 64;
      t.W = jobs[t.z].org.w; t.H = jobs[t.z].org.h;
      if (t.x0 < t.W && t.y0 < t.H) return L;
    }
    return total;
  };
  // queue the loads of a tile into stage `st`; returns the window's byte offset of displacement (-R) inside a row
  auto issue = [&](Tile &t, int st) {
    const SadJob &J = jobs[t.z];
    uint8_t *sOrg = smem + st * Cfg::buf_bytes, *sWin = sOrg + 64 * 64;
    int cx = 0, cy = 0;
    if (J.centres) {
      const int sb = (t.y0 >> 6) * gx + (t.x0 >> 6);       // host lays centres out with the launch's SB pitch
      cx = J.centres[2 * sb]; cy = J.centres[2 * sb + 1];
    }
    const int wx = t.x0 + cx - R, wy = t.y0 + cy - R;
    const int ax = wx & ~15;
    t.o = wx - ax;
    const int pad = J.ref.pad[0], stride = J.ref.stride[0];
    const uint8_t *rp = J.ref.p[0];
    // rows / 16-byte chunks outside the padded picture only feed blocks that are not output; clamp them to legal addresses
    const int xlo = -(pad & ~15), xhi = (t.W + pad - 1) & ~15;
    constexpr int cpr = pitch >> 4;
    for (int i = tid; i < win_rows * cpr; i += SAD_THREADS) {
      const int r = i / cpr, c = i - r * cpr;
      const int gyy = clampi(wy + r, -pad, t.H + pad - 1), gxx = clampi(ax + 16 * c, xlo, xhi);
      cp_async16(sWin + r * pitch + 16 * c, rp + (ptrdiff_t)gyy * stride + gxx);
    }
    const uint8_t *op = J.org.p[0];
    const int ostride = J.org.stride[0], oxhi = (t.W - 1) & ~15;
    for (int i = tid; i < 64 * 4; i += SAD_THREADS) {
      const int r = i >> 2, c = i & 3;
      const int gyy = min(t.y0 + r, t.H - 1), gxx = min(t.x0 + 16 * c, oxhi);
      cp_async16(sOrg + r * 64 + 16 * c, op + (size_t)gyy * ostride + gxx);
    }
  };

  Tile cur, nxt;
  int L = find(blockIdx.x, cur);
  int st = 0;
  if (L < total) issue(cur, 0);
  cp_async_commit();
  while (L < total) {
    const int Ln = find(L + gridDim.x, nxt);
    if (Ln < total) issue(nxt, st ^ 1);
    cp_async_commit();
    cp_async_wait1();                 // everything but the group just committed has landed
    __syncthreads();

    const SadJob &J = jobs[cur.z];
    SadTile C;
    C.sOrg = smem + st * Cfg::buf_bytes; C.sWin = C.sOrg + 64 * 64; C.sPh = smem + Cfg::ph_off;
    if (Cfg::PH) {
      // the three shifted copies of the window: 16 bytes in, 3 x 16 bytes out per item
      constexpr int cpr = pitch >> 4;
      uint8_t *sPh = smem + Cfg::ph_off;
      for (int i = tid; i < win_rows * cpr; i += SAD_THREADS) {
        const int r = i / cpr, c = i - r * cpr;
        const uint8_t *src = C.sWin + r * pitch + 16 * c;
        const uint4 v = *reinterpret_cast<const uint4 *>(src);
        const uint32_t nx = c + 1 < cpr ? *reinterpret_cast<const uint32_t *>(src + 16) : 0u;
#pragma unroll
        for (int k = 1; k < 4; k++) {
          uint4 o;
          o.x = __funnelshift_r(v.x, v.y, 8 * k); o.y = __funnelshift_r(v.y, v.z, 8 * k);
          o.z = __funnelshift_r(v.z, v.w, 8 * k); o.w = __funnelshift_r(v.w, nx, 8 * k);
          *reinterpret_cast<uint4 *>(sPh + 32 * k + (k - 1) * Cfg::win_bytes + r * pitch + 16 * c) = o;
        }
      }
      __syncthreads();
    }
    C.part = part; C.pair = pair; C.t8 = J.t8; C.t16 = J.t16;
    C.W = cur.W; C.H = cur.H; C.x0 = cur.x0; C.y0 = cur.y0; C.o = cur.o;
    {
      const int p = warp & 3, ch = warp >> 2, u = lane / Cfg::DXL, d = lane % Cfg::DXL;
#pragma unroll 1
      for (int c = 0; c < Cfg::NCH; c++)
#pragma unroll 1
        for (int b = 0; b < Cfg::NB; b++)
          sad_pass<R, Cfg::UPW>(C, p, 4 * ch + b * Cfg::UPW + u, c * Cfg::DXL + d, b, warp, u == 0);
    }
    __syncthreads();
    // 32x32 and 64x64 levels from the strip sums: coalesced 32-bit rows
    {
      const int W = cur.W, H = cur.H, x0 = cur.x0, y0 = cur.y0;
      const int bw32 = W >> 5, bw64 = W >> 6;
      const bool ok64 = x0 + 64 <= W && y0 + 64 <= H;
      uint32_t *d64 = J.t64 + ((size_t)(y0 >> 6) * bw64 + (x0 >> 6)) * nn;
      uint32_t *d32[4];
      bool ok32[4];
#pragma unroll
      for (int q = 0; q < 4; q++) {
        const int gxx = x0 + (q & 1) * 32, gyy = y0 + (q >> 1) * 32;
        ok32[q] = gxx + 32 <= W && gyy + 32 <= H;
        d32[q] = J.t32 + ((size_t)(gyy >> 5) * bw32 + (gxx >> 5)) * nn;
      }
      for (int i = tid; i < nn; i += SAD_THREADS) {
        uint32_t tot = 0;
#pragma unroll
        for (int q = 0; q < 4; q++) {   // 32x32 block q = (qy, qx): strips p = 2qy, 2qy+1 of column half qx
          const int qy = q >> 1, qx = q & 1;
          const uint32_t v = part[((2 * qy) * 2 + qx) * nn + i] + part[((2 * qy + 1) * 2 + qx) * nn + i];
          if (ok32[q]) d32[q][i] = v;
          tot += v;
        }
        if (ok64) d64[i] = tot;
      }
    }
    cur = nxt; L = Ln; st ^= 1;
    __syncthreads();                  // `part` and stage st are free again: the next iteration refills / rewrites them
  }
}

// --------------------------------------------------------------------------- cost batch ----
// One warp per (block, candidate) task.  The reference window (rows -3..h+2, columns -4..w+3 around the candidate
// position, so that the block's columns are word aligned) is staged in shared memory; a lane then works on 4
// horizontally adjacent samples at a time with byte-SIMD words: VAVG (rounding average = v64_avg_u8), VHADD
// (truncating average = v64_rdavg_u8) and VABSDIFF4 with accumulate, exactly the operations of the reference's
// SIMD kernels (enc/enc_kernels.c:162-648).  The kernel is compiled once per kind; the window buffer is sized by
// the largest block of the list (dynamic shared memory), so lists of small blocks run at full occupancy.
constexpr int CW = 8;                         // warps per CTA

struct CostParams {
  const PicView *pics;
  const uint16_t *map;
  const thorcuda_cost_task *tasks;
  thorcuda_cost_result *out;
  int ntasks;
  int pitch;                                  // window row pitch in words: (max w + 8) / 4
  int warp_words;                             // window words per warp: (max h + 6) * pitch
};

__device__ __forceinline__ uint32_t up4(uint32_t a, uint32_t b) { return __vavgu4(a, b); }    // (a+b+1)>>1 per byte
__device__ __forceinline__ uint32_t dn4(uint32_t a, uint32_t b) { return __vhaddu4(a, b); }   // (a+b)>>1 per byte

// the 4 samples starting K columns right of the quad, from the three window words around it
struct Row3 { uint32_t w0, w1, w2; };
template <int K>
__device__ __forceinline__ uint32_t col(const Row3 &r) {
  if (K < 0) return __funnelshift_r(r.w0, r.w1, 8 * (4 + K));
  if (K == 0) return r.w1;
  return __funnelshift_r(r.w1, r.w2, 8 * K);
}
__device__ __forceinline__ Row3 load_row(const uint32_t *p) { Row3 r; r.w0 = p[0]; r.w1 = p[1]; r.w2 = p[2]; return r; }

// bilinear diagonal half-pel samples (reference enc/encode_block.c:527-571: a fixed tree of rounding / truncating
// averages); R[0..4] = rows -2..+2; (QY,QX) = top-left sample of the 2x2 core relative to the quad
template <int QY, int QX>
__device__ __forceinline__ uint32_t diag_half(const Row3 (&R)[5]) {
  const uint32_t core = dn4(up4(col<QX>(R[QY + 2]), col<QX + 1>(R[QY + 2])), up4(col<QX>(R[QY + 3]), col<QX + 1>(R[QY + 3])));
  const uint32_t vert = dn4(up4(col<QX>(R[QY + 1]), col<QX>(R[QY + 4])), up4(col<QX + 1>(R[QY + 1]), col<QX + 1>(R[QY + 4])));
  const uint32_t hori = dn4(up4(col<QX - 1>(R[QY + 2]), col<QX + 2>(R[QY + 2])), up4(col<QX - 1>(R[QY + 3]), col<QX + 2>(R[QY + 3])));
  return dn4(dn4(vert, hori), core);
}

__device__ __forceinline__ uint32_t warp_sum(uint32_t v) {
#pragma unroll
  for (int m = 16; m > 0; m >>= 1) v += __shfl_xor_sync(0xffffffffu, v, m);
  return v;
}

// Sum each of s[0..7] over the warp with 9 shuffles instead of 40: every exchange halves the number of sums a lane
// still carries.  Afterwards lane l holds the warp total of s[k], k = 4 * bit4(l) + 2 * bit3(l) + bit2(l).
__device__ __forceinline__ uint32_t warp_sum8(const uint32_t (&s)[8], int lane) {
  uint32_t t[4], u[2];
  const bool b4 = lane & 16, b3 = lane & 8, b2 = lane & 4;
#pragma unroll
  for (int k = 0; k < 4; k++) t[k] = (b4 ? s[k + 4] : s[k]) + __shfl_xor_sync(0xffffffffu, b4 ? s[k] : s[k + 4], 16);
#pragma unroll
  for (int k = 0; k < 2; k++) u[k] = (b3 ? t[k + 2] : t[k]) + __shfl_xor_sync(0xffffffffu, b3 ? t[k] : t[k + 2], 8);
  uint32_t v = (b2 ? u[1] : u[0]) + __shfl_xor_sync(0xffffffffu, b2 ? u[0] : u[1], 4);
  v += __shfl_xor_sync(0xffffffffu, v, 2);
  v += __shfl_xor_sync(0xffffffffu, v, 1);
  return v;
}

// Stage rows x (4 * words) bytes starting at src (any alignment) into win, byte 0 of a row = src[0].  LPR lanes
// share a row (LPR > words: the lane after the last output word fetches the word the shift spills into); the loads
// of 8 row groups are issued before the first is used.
template <int LPR>
__device__ __forceinline__ void cost_stage(uint32_t *win, int pitch, const uint8_t *src, int stride, int rows, int words, int lane) {
  constexpr int RPP = 32 / LPR;                       // rows per pass
  const int o = (int)((uintptr_t)src & 3), sh = 8 * o;
  const uint8_t *base = src - o;
  const int wd = lane % LPR, r0 = lane / LPR;
  const bool fetch = 4 * wd < o + 4 * words;          // raw words that hold wanted bytes
  for (int rb = 0; rb < rows; rb += 8 * RPP) {
    uint32_t v[8];
#pragma unroll
    for (int k = 0; k < 8; k++) {
      const int r = rb + k * RPP + r0;
      v[k] = (fetch && r < rows) ? *reinterpret_cast<const uint32_t *>(base + (ptrdiff_t)r * stride + 4 * wd) : 0u;
    }
#pragma unroll
    for (int k = 0; k < 8; k++) {
      const int r = rb + k * RPP + r0;
      const uint32_t hi = __shfl_down_sync(0xffffffffu, v[k], 1);
      if (wd < words && r < rows) win[r * pitch + wd] = __funnelshift_r(v[k], hi, sh);
    }
  }
}

template <int KIND>
__global__ void __launch_bounds__(CW * 32, 4) k_cost_batch(const CostParams P) {
  extern __shared__ __align__(16) uint32_t sWin[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ti = blockIdx.x * CW + warp;
  if (ti >= P.ntasks) return;
  const thorcuda_cost_task t = P.tasks[ti];
  const PicView &org = P.pics[P.map[t.org]], &ref = P.pics[P.map[t.ref]];
  const int w = t.w, h = t.h;
  const int CPW = P.pitch;
  // stage reference rows -3..h+2, columns -4..w+3 around the candidate position; byte 0 of a row = column -4
  uint32_t *win = sWin + warp * P.warp_words;
  {
    const int rstride = ref.stride[0];
    const uint8_t *src = ref.p[0] + (ptrdiff_t)(t.y + t.iy - 3) * rstride + (t.x + t.ix - 4);
    const int words = (w + 8) >> 2;
    if (words < 8) cost_stage<8>(win, CPW, src, rstride, h + 6, words, lane);
    else if (words < 16) cost_stage<16>(win, CPW, src, rstride, h + 6, words, lane);
    else cost_stage<32>(win, CPW, src, rstride, h + 6, words, lane);
  }
  __syncwarp();
  const int ostride = org.stride[0];
  const uint8_t *a00 = org.p[0] + (size_t)t.y * ostride + t.x;
  const int wq = w >> 2, nquads = wq * h;
  const int wq_inv = (65536 + wq - 1) / wq;              // i / wq == (i * wq_inv) >> 16 for i < 64 * 16
  uint32_t s[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  const int sel = (t.hx & t.hy) ? 0 : (t.hx ? 1 : (t.hy ? 2 : 3));
  for (int i = lane; i < nquads; i += 32) {
    const int r = (i * wq_inv) >> 16, c4 = i - r * wq;
    const uint32_t A = *reinterpret_cast<const uint32_t *>(a00 + (size_t)r * ostride + 4 * c4);
    const uint32_t *p0 = win + (r + 3) * CPW + c4;       // words around the quad in its own row
    if (KIND == THORCUDA_COST_SAD) {
      s[0] = sad4(A, p0[1], s[0]);
    } else if (KIND == THORCUDA_COST_SSD) {
      const uint32_t d = __vabsdiffu4(A, p0[1]);
      s[0] = __dp4a(d, d, s[0]);
    } else if (KIND == THORCUDA_COST_WIDESAD) {
      const Row3 R = load_row(p0);
      s[0] = sad4(A, col<-3>(R), s[0]); s[1] = sad4(A, col<-1>(R), s[1]); s[2] = sad4(A, col<0>(R), s[2]);
      s[3] = sad4(A, col<1>(R), s[3]); s[4] = sad4(A, col<3>(R), s[4]);
    } else if (KIND == THORCUDA_COST_FASTHALF) {
      Row3 R[5];
#pragma unroll
      for (int y = 0; y < 5; y++) R[y] = load_row(p0 + (y - 2) * CPW);
      const uint32_t X = R[2].w1;
      // order: top down left right tl tr bl br
      s[0] = sad4(A, up4(X, R[1].w1), s[0]);
      s[1] = sad4(A, up4(X, R[3].w1), s[1]);
      s[2] = sad4(A, up4(col<-1>(R[2]), X), s[2]);
      s[3] = sad4(A, up4(X, col<1>(R[2])), s[3]);
      s[4] = sad4(A, diag_half<-1, -1>(R), s[4]);
      s[5] = sad4(A, diag_half<-1, 0>(R), s[5]);
      s[6] = sad4(A, diag_half<0, -1>(R), s[6]);
      s[7] = sad4(A, diag_half<0, 0>(R), s[7]);
    } else {
      // fastquarter (enc/encode_block.c:609-738): a = here, b = up, c = up-right, d = right, e = down-right,
      // f = down, g = down-left, hh = left; four formula sets chosen by the incoming half-pel offset
      const Row3 Ru = load_row(p0 - CPW), R0 = load_row(p0), Rd = load_row(p0 + CPW);
      const uint32_t a = R0.w1, b = Ru.w1, c2 = col<1>(Ru), d = col<1>(R0), e = col<1>(Rd), f = Rd.w1, g = col<-1>(Rd), hh = col<-1>(R0);
      const uint32_t ad = up4(a, d), de = up4(d, e), af = up4(a, f), fe = up4(f, e), dc = up4(d, c2), ab = up4(a, b),
                     ah = up4(a, hh), gf = up4(g, f);
      uint32_t v[8];   // tl top tr left right bl down br
      if (sel == 0)      { v[0]=dn4(ad,af); v[1]=dn4(de,a); v[2]=dn4(ad,de); v[3]=dn4(ad,f); v[4]=dn4(ad,e); v[5]=dn4(af,fe); v[6]=dn4(de,f); v[7]=dn4(de,fe); }
      else if (sel == 1) { v[0]=dn4(ad,ab); v[1]=dn4(dc,a); v[2]=dn4(ad,dc); v[3]=dn4(ad,a); v[4]=dn4(ad,d); v[5]=dn4(ad,af); v[6]=dn4(af,d); v[7]=dn4(ad,de); }
      else if (sel == 2) { v[0]=dn4(ah,af); v[1]=dn4(af,a); v[2]=dn4(ad,af); v[3]=dn4(gf,a); v[4]=dn4(ad,f); v[5]=dn4(af,gf); v[6]=dn4(af,f); v[7]=dn4(af,fe); }
      else               { v[0]=dn4(ah,ab); v[1]=dn4(ab,a); v[2]=dn4(ad,ab); v[3]=dn4(ah,a); v[4]=dn4(ad,a); v[5]=dn4(ah,af); v[6]=dn4(af,a); v[7]=dn4(af,ad); }
#pragma unroll
      for (int k = 0; k < 8; k++) s[k] = sad4(A, v[k], s[k]);
    }
  }
  thorcuda_cost_result res;
  res.x = 0; res.y = 0;
  if (KIND == THORCUDA_COST_SAD || KIND == THORCUDA_COST_SSD) {
    res.cost = warp_sum(s[0]);
  } else {
    // Every candidate position gets a rank = its place in the reference's comparison order; the reference keeps the
    // first strict minimum in that order, which is the minimum of (cost << 3 | rank) -- the packing its own SIMD
    // widesad uses (enc_kernels.c:93-97).  Costs stay below 2^21 (64 * 64 * 255).
    //   widesad     (:764-776)  order -3,-1,0,+1,+3
    //   fasthalf    (:577-603)  start top, then down,right,left,tl,tr,br,bl;  s = top down left right tl tr bl br
    //   fastquarter (:703-734)  start top, then tl,tr,left,right,bl,down,br;  s = tl top tr left right bl down br
    const int k = ((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1);
    if (KIND == THORCUDA_COST_WIDESAD) s[5] = s[6] = s[7] = 0x0fffffffu >> 5;    // never the minimum (summed over 32 lanes)
    const uint32_t total = warp_sum8(s, lane);
    const uint32_t rank_of = KIND == THORCUDA_COST_WIDESAD ? 0x76543210u : KIND == THORCUDA_COST_FASTHALF ? 0x67542310u : 0x76543201u;
    uint32_t key = (total << 3) | ((rank_of >> (4 * k)) & 7u);
    key = min(key, __shfl_xor_sync(0xffffffffu, key, 16));
    key = min(key, __shfl_xor_sync(0xffffffffu, key, 8));
    key = min(key, __shfl_xor_sync(0xffffffffu, key, 4));
    res.cost = key >> 3;
    const int rank = key & 7;
    // (x, y) of each rank, packed as 4-bit fields biased by 3
    if (KIND == THORCUDA_COST_WIDESAD) {
      res.x = (int16_t)((int)((0x64320u >> (4 * rank)) & 15u) - 3);                       // -3,-1,0,1,3
    } else if (KIND == THORCUDA_COST_FASTHALF) {
      res.x = (int16_t)((int)((0x15511533u >> (4 * rank)) & 15u) - 3);                    // 0,0,2,-2,-2,2,2,-2
      res.y = (int16_t)((int)((0x55113351u >> (4 * rank)) & 15u) - 3);                    // -2,2,0,0,-2,-2,2,2
    } else {
      res.x = (int16_t)((int)((0x43242423u >> (4 * rank)) & 15u) - 3);                    // 0,-1,1,-1,1,-1,0,1
      res.y = (int16_t)((int)((0x44433222u >> (4 * rank)) & 15u) - 3);                    // -1,-1,-1,0,0,1,1,1
    }
  }
  if (lane == 0) P.out[ti] = res;
}

template <int R>
static int launch_sad(int gx, int gy, int n, const SadJob *d_jobs) {
  static int ctas_per_sm = 0;
  constexpr size_t smem = SadCfg<R>::smem_bytes;
  if (!ctas_per_sm) {
    THOR_CUDA(cudaFuncSetAttribute(k_sad_table<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    THOR_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctas_per_sm, k_sad_table<R>, SAD_THREADS, smem));
    if (ctas_per_sm < 1) return fail("sad_table: kernel for range %d does not fit on an SM", R);
  }
  // persistent: as many CTAs as are resident at once (2 per SM up to R = 24) walk the tiles
  const int total = gx * gy * n;
  const int grid = total < ctas_per_sm * ctx().sm_count ? total : ctas_per_sm * ctx().sm_count;
  k_sad_table<R><<<grid, SAD_THREADS, smem, ctx().stream>>>(d_jobs, gx, gy, n);
  THOR_LAUNCHED();
  return 0;
}

// register-only VABSDIFF4.ACC loop for thorcuda_calibrate_absdiff
__global__ void __launch_bounds__(256) k_absdiff_peak(uint32_t *out, int iters) {
  uint32_t a[16], x = threadIdx.x * 0x01010101u + blockIdx.x, y = ~x;
#pragma unroll
  for (int k = 0; k < 16; k++) a[k] = k;
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int k = 0; k < 16; k++) a[k] = sad4(x, y + k, a[k]);
    x += 0x00010001u;
  }
  uint32_t r = 0;
#pragma unroll
  for (int k = 0; k < 16; k++) r ^= a[k];
  if (r == 0x12345678u) out[threadIdx.x & 15] = r;      // keeps the loop alive, practically never taken
}

}  // namespace thor

using namespace thor;

extern "C" {

size_t thorcuda_sad_table_bytes(int slot, int bs) {
  if (slot < 0 || slot >= THORCUDA_SAD_SLOTS) return 0;
  const int l = bs == 8 ? 0 : bs == 16 ? 1 : bs == 32 ? 2 : bs == 64 ? 3 : -1;
  if (l < 0 || !ctx().sad[slot].used) return 0;
  return ctx().sad[slot].bytes[l];
}

void *thorcuda_sad_table_devptr(int slot, int bs) {
  if (slot < 0 || slot >= THORCUDA_SAD_SLOTS) return nullptr;
  const int l = bs == 8 ? 0 : bs == 16 ? 1 : bs == 32 ? 2 : bs == 64 ? 3 : -1;
  if (l < 0 || !ctx().sad[slot].used) return nullptr;
  return ctx().sad[slot].tab[l];
}

int thorcuda_sad_table_fetch(int slot, int bs, void *host_dst, size_t bytes) {
  THOR_NEED_CTX();
  const size_t have = thorcuda_sad_table_bytes(slot, bs);
  if (!have) return fail("sad_table_fetch: slot %d has no %dx%d table", slot, bs, bs);
  if (bytes > have) return fail("sad_table_fetch: %zu bytes requested, table has %zu", bytes, have);
  return d2h_bulk(host_dst, thorcuda_sad_table_devptr(slot, bs), bytes);
}

int thorcuda_sad_table_build(int n, const int *slots, const thorcuda_pic *orgs, const thorcuda_pic *refs, int range,
                             const int16_t *const *centres) {
  THOR_NEED_CTX();
  Context &c = ctx();
  if (n <= 0) return 0;
  if (range < 4 || range > 32 || (range & 3)) return fail("sad_table: range must be 4,8,...,32 (got %d)", range);
  const int nn = (2 * range) * (2 * range);
  int maxw = 0, maxh = 0;
  size_t centre_entries = 0;
  for (int i = 0; i < n; i++) {
    Pic *o = get_pic(orgs[i]), *r = get_pic(refs[i]);
    if (!o || !r) return fail("sad_table: bad picture handle in pair %d", i);
    if (o->w != r->w || o->h != r->h) return fail("sad_table: pair %d: picture size mismatch", i);
    if (o->pad_y & 15) return fail("sad_table: pair %d: the original's border %d is not a multiple of 16 (its rows are fetched with 16-byte copies)", i, o->pad_y);
    if ((r->pad_y & 15) || r->pad_y < range) return fail("sad_table: pair %d: reference padding %d cannot hold range %d", i, r->pad_y, range);
    if (slots[i] < 0 || slots[i] >= THORCUDA_SAD_SLOTS) return fail("sad_table: slot %d out of range", slots[i]);
    for (int k = 0; k < i; k++) if (slots[k] == slots[i]) return fail("sad_table: slot %d used twice in one call", slots[i]);
    maxw = max(maxw, o->w); maxh = max(maxh, o->h);
  }
  const int gx = (maxw + 63) / 64, gy = (maxh + 63) / 64;
  for (int i = 0; i < n; i++) {
    if (!centres || !centres[i]) continue;
    Pic *o = get_pic(orgs[i]), *r = get_pic(refs[i]);
    const int nsbx = (o->w + 63) / 64, nsby = (o->h + 63) / 64;
    for (int sy = 0; sy < nsby; sy++)
      for (int sx = 0; sx < nsbx; sx++) {
        const int cx = centres[i][2 * (sy * nsbx + sx)], cy = centres[i][2 * (sy * nsbx + sx) + 1];
        const int x0 = sx * 64, y0 = sy * 64, x1 = min(x0 + 64, o->w), y1 = min(y0 + 64, o->h);
        if (x0 + cx - range < -r->pad_y || y0 + cy - range < -r->pad_y || x1 + cx + range - 1 > o->w + r->pad_y ||
            y1 + cy + range - 1 > o->h + r->pad_y)
          return fail("sad_table: pair %d SB (%d,%d): centre (%d,%d) +- %d leaves the padded reference", i, sx, sy, cx, cy, range);
      }
    centre_entries += (size_t)2 * gx * gy;
  }
  // fetches of the previous tables may still be draining on the copy stream: the build must not overwrite them
  if (copies_fence()) return -1;
  // (re)size the slots
  for (int i = 0; i < n; i++) {
    Pic *o = get_pic(orgs[i]);
    SadSlot &s = c.sad[slots[i]];
    for (int l = 0; l < 4; l++) {
      const int bs = 8 << l;
      const size_t bytes = (size_t)(o->w / bs) * (o->h / bs) * nn * (l < 2 ? 2 : 4);
      if (bytes > s.cap[l]) {
        if (s.tab[l]) { cudaStreamSynchronize(c.stream); cudaStreamSynchronize(c.copy_stream); cudaFree(s.tab[l]); s.tab[l] = nullptr; s.cap[l] = 0; }
        if (bytes && cudaMalloc(&s.tab[l], bytes + 64) != cudaSuccess) return fail("sad_table: device allocation of %zu bytes failed", bytes);
        s.cap[l] = bytes;
      }
      s.bytes[l] = bytes;
    }
    s.used = true; s.w = o->w; s.h = o->h; s.range = range;
  }
  SadJob *d_jobs = (SadJob *)c.sadjobs.reserve(sizeof(SadJob) * n);
  int16_t *d_centres = centre_entries ? (int16_t *)c.centres.reserve(centre_entries * 2) : nullptr;
  const size_t st_bytes = sizeof(SadJob) * n + centre_entries * 2;
  uint8_t *st = (uint8_t *)stage_acquire(st_bytes);
  if (!d_jobs || !st || (centre_entries && !d_centres)) return fail("sad_table: allocation failed");
  SadJob *hj = (SadJob *)st;
  int16_t *hc = (int16_t *)(st + sizeof(SadJob) * n);
  size_t coff = 0;
  for (int i = 0; i < n; i++) {
    Pic *o = get_pic(orgs[i]), *r = get_pic(refs[i]);
    SadSlot &s = c.sad[slots[i]];
    hj[i].org = o->view(); hj[i].ref = r->view();
    hj[i].t8 = (uint16_t *)s.tab[0]; hj[i].t16 = (uint16_t *)s.tab[1];
    hj[i].t32 = (uint32_t *)s.tab[2]; hj[i].t64 = (uint32_t *)s.tab[3];
    hj[i].centres = nullptr;
    if (centres && centres[i]) {
      // re-pitch to the launch grid (gx SBs per row) so the kernel can index with gridDim.x
      const int nsbx = (o->w + 63) / 64, nsby = (o->h + 63) / 64;
      memset(hc + coff, 0, (size_t)2 * gx * gy * 2);
      for (int sy = 0; sy < nsby; sy++) memcpy(hc + coff + (size_t)2 * sy * gx, centres[i] + (size_t)2 * sy * nsbx, (size_t)2 * nsbx * 2);
      hj[i].centres = d_centres + coff;
      coff += (size_t)2 * gx * gy;
    }
  }
  THOR_CUDA(cudaMemcpyAsync(d_jobs, hj, sizeof(SadJob) * n, cudaMemcpyHostToDevice, c.stream));
  if (stage_commit_copy(d_centres ? (void *)d_centres : (void *)d_jobs, hc, centre_entries * 2)) return -1;
  switch (range) {
    case 4: return launch_sad<4>(gx, gy, n, d_jobs);
    case 8: return launch_sad<8>(gx, gy, n, d_jobs);
    case 12: return launch_sad<12>(gx, gy, n, d_jobs);
    case 16: return launch_sad<16>(gx, gy, n, d_jobs);
    case 20: return launch_sad<20>(gx, gy, n, d_jobs);
    case 24: return launch_sad<24>(gx, gy, n, d_jobs);
    case 28: return launch_sad<28>(gx, gy, n, d_jobs);
    default: return launch_sad<32>(gx, gy, n, d_jobs);
  }
}

int thorcuda_cost_list(thorcuda_list hl, int kind, thorcuda_cost_result *host_results) {
  THOR_NEED_CTX();
  Context &c = ctx();
  List *L = get_list(hl);
  if (!L || L->kind != THORCUDA_LIST_COST) return fail("cost: not a cost list");
  if (list_stale(L)) return -1;
  if (kind < 0 || kind > THORCUDA_COST_SSD) return fail("cost: unknown kind %d", kind);
  CostParams P;
  P.pics = c.d_pics; P.map = L->map();
  P.tasks = (const thorcuda_cost_task *)L->tasks();
  // results alternate between two buffers: the copy of the previous call may still be in flight on the copy stream
  const int rb = L->res_next;
  if (L->late_res[rb] && list_fence(L)) return -1;
  L->results = L->res_buf[rb];
  L->res_next = rb ^ 1;
  P.out = L->results; P.ntasks = L->ntasks;
  P.pitch = (L->cost_w + 8) / 4;
  P.warp_words = (L->cost_h + 6) * P.pitch;
  const size_t smem = (size_t)CW * P.warp_words * 4;      // at most 8 x 70 x 18 words = 40 KB
  const unsigned grid = (unsigned)((L->ntasks + CW - 1) / CW);
  switch (kind) {
    case THORCUDA_COST_SAD: k_cost_batch<THORCUDA_COST_SAD><<<grid, CW * 32, smem, c.stream>>>(P); break;
    case THORCUDA_COST_WIDESAD: k_cost_batch<THORCUDA_COST_WIDESAD><<<grid, CW * 32, smem, c.stream>>>(P); break;
    case THORCUDA_COST_FASTHALF: k_cost_batch<THORCUDA_COST_FASTHALF><<<grid, CW * 32, smem, c.stream>>>(P); break;
    case THORCUDA_COST_FASTQUARTER: k_cost_batch<THORCUDA_COST_FASTQUARTER><<<grid, CW * 32, smem, c.stream>>>(P); break;
    default: k_cost_batch<THORCUDA_COST_SSD><<<grid, CW * 32, smem, c.stream>>>(P); break;
  }
  THOR_LAUNCHED();
  if (host_results) {
    L->late_res[rb] = true;
    return d2h_bulk(host_results, L->results, (size_t)L->ntasks * sizeof(thorcuda_cost_result));
  }
  return 0;
}

int thorcuda_cost_batch(int kind, const thorcuda_pic *pics, int npics, const thorcuda_cost_task *tasks, int ntasks,
                        thorcuda_cost_result *host_results) {
  if (ntasks <= 0) return 0;
  const thorcuda_list l = thorcuda_list_create(THORCUDA_LIST_COST, pics, npics, tasks, ntasks);
  if (l < 0) return -1;
  const int rc = thorcuda_cost_list(l, kind, host_results);
  thorcuda_list_destroy(l);
  return rc;
}

// Peak of the instruction the SAD search is made of, measured on the device it runs on ("calibrate peak with a
// register-only micro-kernel"): every thread keeps 16 independent VABSDIFF4.ACC chains busy, nothing touches memory.
// Returns abs-diffs per second (4 per instruction and lane), or a negative value on failure.  Synchronous.
double thorcuda_calibrate_absdiff(int iters) {
  THOR_NEED_CTX();
  Context &c = ctx();
  if (iters <= 0) iters = 4096;
  uint32_t *d = (uint32_t *)c.scratch.reserve(1 << 16);
  if (!d) return -1.0;
  const int ctas = c.sm_count * 8, threads = 256;
  cudaEvent_t e0, e1;
  if (cudaEventCreate(&e0) != cudaSuccess || cudaEventCreate(&e1) != cudaSuccess) return -1.0;
  k_absdiff_peak<<<ctas, threads, 0, c.stream>>>(d, iters / 8);       // warm-up
  cudaEventRecord(e0, c.stream);
  k_absdiff_peak<<<ctas, threads, 0, c.stream>>>(d, iters);
  cudaEventRecord(e1, c.stream);
  c.launches += 2;
  float ms = 0.f;
  const bool ok = cudaEventSynchronize(e1) == cudaSuccess && cudaEventElapsedTime(&ms, e0, e1) == cudaSuccess && cudaGetLastError() == cudaSuccess;
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  if (!ok || ms <= 0.f) return -1.0;
  return 4.0 * 16.0 * (double)iters * threads * ctas / (ms * 1e-3);
}

}  // extern "C"
