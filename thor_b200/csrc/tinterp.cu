// libthorcuda: temporal interpolation on the device.
//
// Restates reference common/temporal_interp.c:61-1053 (interpolate_frames; its USE_CHROMA is 0): a luma pyramid, one
// motion field per level estimated coarse to fine, a neighbour-merge pass per level, and the rounding average of the two
// displaced references.  The reference walks the 16x16 blocks of a level in raster order; each block only looks at its
// left, up-left, up and up-right neighbours, so the device runs the block ROWS concurrently as a wavefront:
//
//   k_halve     2x2 mean (double rounding) of both references into the next level, border replication fused in
//   k_estimate  one CTA per block row, rows handed out by a ticket so that a row's upper neighbour is always running;
//               a row waits (acquire load) until the row above is two blocks ahead.  Inside a block the independent
//               parts run side by side: warp 0 the skip test (four 8x8 SADs), warps 1..5 one candidate each, every
//               candidate refined speculatively by its cross search (4 lane groups = the 4 cross points, 8 lanes x 2
//               rows of a 16x16 SAD each); one thread then replays the reference's sequential selection on the results
//   k_merge     per 8x8 block the best of its own / neighbours' vectors (one warp per block, 8 lanes per candidate)
//   k_blend     8x8 luma + 4x4 chroma blocks: (a + b + 1) >> 1 of the displaced references, or a copy / clamped
//               fetch when a displaced block leaves the picture by more than half a block
//
// Vectors are 1/8 pel int16 pairs pointing into pic[1]; pic[0]'s vector is always the scaled one (-wt1/wt0), so only
// pic[1]'s field is stored.
#include <math.h>
#include <string.h>
#include "internal.cuh"

namespace thor {

constexpr int COST_MAX = 0x3fffffff;

struct LevelView {
  const uint8_t *y[2];      // pic[0], pic[1] luma origin
  int stride, w, h, pad;
  int bw, bh;               // field size in 8x8 units
  short2 *pre, *post;       // field after the block search / after the merge pass
  const short2 *guide;      // the coarser level's merged field (nullptr on the coarsest level)
  int guide_bw;
  int *progress;            // [bh/2] blocks finished per block row, + 1 ticket counter at [bh/2]
  int wt0, wt1;
};

// ---------------------------------------------------------------- vector arithmetic (temporal_interp.c:66-91) ----
__host__ __device__ __forceinline__ int scale_component(int v, int numer, int denom) {
  if (denom == 0) return 0;
  int prod = v * numer;
  if (denom < 0) { denom = -denom; prod = -prod; }
  return prod >= 0 ? (prod + denom / 2) / denom : -((-prod + denom / 2) / denom);
}
// the same rounding division for the device: denom = wt0 is a small positive frame distance, so |v * numer| + denom / 2
// stays below 2^24 and the quotient is exactly the high word of a multiplication by ceil(2^32 / denom) -- the integer
// division costs ~40 dependent instructions on the block search's critical path, four times per candidate
__device__ __forceinline__ int scale_exact(int v, int numer, int denom) {
  if (denom > 255 || denom < 2) return scale_component(v, numer, denom);
  const int prod = v * numer;
  const uint32_t magic = 0xffffffffu / (uint32_t)denom + 1;       // uniform: hoisted out of the loops by the compiler
  const int q = (int)__umulhi((uint32_t)abs(prod) + (uint32_t)(denom >> 1), magic);
  return prod >= 0 ? q : -q;
}
__device__ __forceinline__ short2 to_pic0(short2 v, int wt0, int wt1) {
  const int numer = -wt1, denom = wt0;
  if (numer == denom) return v;
  if (numer == -denom) return make_short2((short)-v.x, (short)-v.y);
  return make_short2((short)scale_exact(v.x, numer, denom), (short)scale_exact(v.y, numer, denom));
}
__device__ __forceinline__ int pel(int v) { return (v + 4) >> 3; }
__device__ __forceinline__ int l1(short2 a, short2 b) { return abs(a.x - b.x) + abs(a.y - b.y); }
__device__ __forceinline__ bool same(short2 a, short2 b) { return a.x == b.x && a.y == b.y; }

// ---------------------------------------------------------------- unaligned rows through aligned words ----
__device__ __forceinline__ void ldrow16(const uint8_t *p, uint32_t o[4]) {
  const uint32_t *q = reinterpret_cast<const uint32_t *>(reinterpret_cast<uintptr_t>(p) & ~(uintptr_t)3);
  const int sh = (int)(reinterpret_cast<uintptr_t>(p) & 3) * 8;
  const uint32_t a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2), d = __ldg(q + 3), e = __ldg(q + 4);
  o[0] = __funnelshift_r(a, b, sh); o[1] = __funnelshift_r(b, c, sh); o[2] = __funnelshift_r(c, d, sh); o[3] = __funnelshift_r(d, e, sh);
}
__device__ __forceinline__ void ldrow8(const uint8_t *p, uint32_t o[2]) {
  const uint32_t *q = reinterpret_cast<const uint32_t *>(reinterpret_cast<uintptr_t>(p) & ~(uintptr_t)3);
  const int sh = (int)(reinterpret_cast<uintptr_t>(p) & 3) * 8;
  const uint32_t a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2);
  o[0] = __funnelshift_r(a, b, sh); o[1] = __funnelshift_r(b, c, sh);
}
__device__ __forceinline__ uint32_t ldrow4(const uint8_t *p) {
  const uint32_t *q = reinterpret_cast<const uint32_t *>(reinterpret_cast<uintptr_t>(p) & ~(uintptr_t)3);
  const int sh = (int)(reinterpret_cast<uintptr_t>(p) & 3) * 8;
  return __funnelshift_r(__ldg(q), __ldg(q + 1), sh);
}

__device__ __forceinline__ bool inside(const LevelView &L, int x, int y, int size) {
  return x >= -L.pad && x + size <= L.w + L.pad && y >= -L.pad && y + size <= L.h + L.pad;
}

// One lane's share of the SAD between the blocks at (x0,y0) in pic[0] and (x1,y1) in pic[1]: rows r0 .. r0+nrows-1 of
// width W (8 or 16).  `in` = both blocks inside picture + pad; else every sample position is clamped (:507-520).
template <int W>
__device__ __forceinline__ uint32_t sad_rows(const LevelView &L, int x0, int y0, int x1, int y1, int r0, int nrows, bool in) {
  uint32_t s = 0;
  if (in) {
    for (int r = r0; r < r0 + nrows; r++) {
      const uint8_t *a = L.y[0] + (ptrdiff_t)(y0 + r) * L.stride + x0, *b = L.y[1] + (ptrdiff_t)(y1 + r) * L.stride + x1;
      uint32_t va[W / 4], vb[W / 4];
      if (W == 16) { ldrow16(a, va); ldrow16(b, vb); } else { ldrow8(a, va); ldrow8(b, vb); }
#pragma unroll
      for (int k = 0; k < W / 4; k++) s = __vsadu4(va[k], vb[k]) + s;
    }
  } else {
    const int wP = L.w + L.pad, hP = L.h + L.pad;
    for (int r = r0; r < r0 + nrows; r++) {
      const int ay = min(hP - 1, max(-L.pad, r + y0)), by = min(hP - 1, max(-L.pad, r + y1));
      for (int j = 0; j < W; j++) {
        const int ax = min(wP - 1, max(-L.pad, j + x0)), bx = min(wP - 1, max(-L.pad, j + x1));
        s += abs((int)L.y[1][(ptrdiff_t)by * L.stride + bx] - (int)L.y[0][(ptrdiff_t)ay * L.stride + ax]);
      }
    }
  }
  return s;
}

// ---------------------------------------------------------------- pyramid ----
struct HalveParams {
  const uint8_t *src[2];
  uint8_t *dst[2];          // origin of the visible picture; the border (pad on every side) is written too
  int sstride, dstride, wo, ho, pad;
};

// one thread = 4 adjacent output samples (border included: source coordinates clamp to the visible picture)
__global__ void __launch_bounds__(256) k_halve(const HalveParams P) {
  const int words = (P.wo + 2 * P.pad + 3) / 4, rows = P.ho + 2 * P.pad;   // a width that is not a multiple of 4 ends in a partial word (the stride has room)
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= words * rows) return;
  const int k = blockIdx.y, row = idx / words - P.pad, col = (idx % words) * 4 - P.pad;
  const uint8_t *s = P.src[k];
  const int i = min(max(row, 0), P.ho - 1);
  const uint8_t *a = s + (ptrdiff_t)(2 * i) * P.sstride, *b = a + P.sstride;
  uint32_t out = 0;
#pragma unroll
  for (int t = 0; t < 4; t++) {
    const int j = min(max(col + t, 0), P.wo - 1);
    const int v = (((a[2 * j] + b[2 * j] + 1) >> 1) + ((a[2 * j + 1] + b[2 * j + 1] + 1) >> 1)) >> 1;   // :164-165
    out |= (uint32_t)v << (8 * t);
  }
  *reinterpret_cast<uint32_t *>(P.dst[k] + (ptrdiff_t)row * P.dstride + col) = out;
}

// ---------------------------------------------------------------- block search (wavefront) ----
__device__ __forceinline__ int ld_acquire(const int *p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release(int *p, int v) { asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
// the field is written by other SMs during the launch: read it from L2
__device__ __forceinline__ short2 ldmv(const short2 *p) {
  const uint32_t v = __ldcg(reinterpret_cast<const uint32_t *>(p));
  return make_short2((short)(v & 0xffff), (short)(v >> 16));
}

struct Nbrs {              // the already decided neighbours of a 16x16 block (vectors into pic[1])
  short2 ur, up, ul, left;
  bool has_ur, has_up, has_left;
  int kind;                // which form of the smoothness term applies (:366-385)
};

__device__ __forceinline__ int field_cost(const Nbrs &n, short2 v, int lambda) {
  int d = 0;
  if (n.kind == 1) d = l1(v, n.ur) + l1(v, n.up) + l1(v, n.ul) + l1(v, n.left);
  else if (n.kind == 2) d = l1(v, n.left);
  else if (n.kind == 3) d = l1(v, n.ur) + l1(v, n.up);
  return (d * lambda) >> 7;
}

constexpr int EW = 6;      // warps per CTA: skip test + 5 candidates

// The wavefront hand-over.  A row publishes "blocks finished" with a release store after the vectors; the row below polls
// with a RELAXED gpu-scope load and then reads the vectors with relaxed gpu-scope loads (both are served by L2, and the
// vector load is only issued once the flag value has come back).  An acquire load would also work but invalidates the
// SM's L1 on every block, which turns each picture-row load of the search into an L2 round trip.
__device__ __forceinline__ int ld_flag(const int *p) {
  int v;
  asm volatile("ld.relaxed.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ short2 ld_mv_l2(const short2 *p) {
  uint32_t v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return make_short2((short)(v & 0xffff), (short)(v >> 16));
}
__device__ __forceinline__ void st_mv_l2(short2 *p, short2 v) {
  const uint32_t w = (uint32_t)(uint16_t)v.x | ((uint32_t)(uint16_t)v.y << 16);
  asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(w) : "memory");
}
__device__ __forceinline__ void prefetch_l1(const void *p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }

struct EstSmem {
  int row;
  int skip[2];               // [block parity]: results of a block are read while the next block is already being searched
  uint32_t init[2][5], fin[2][5];
  short2 mv[2][5];
  short2 skip_mv[2];
  short2 nxt_ur[2];          // up-right neighbour of the NEXT block, fetched by warp 0 while the candidates are searched
};

// cost of vector v for the 16x16 block at (xs, ys), spread over an 8-lane group (gl = lane in group); every lane of the
// group returns the total
__device__ __forceinline__ uint32_t block_cost(const LevelView &L, const Nbrs &nb, int lambda, int xs, int ys, short2 v, int gl) {
  const short2 v0 = to_pic0(v, L.wt0, L.wt1);
  const int x0 = xs + pel(v0.x), y0 = ys + pel(v0.y), x1 = xs + pel(v.x), y1 = ys + pel(v.y);
  const bool in = inside(L, x0, y0, 16) && inside(L, x1, y1, 16);
  uint32_t s = sad_rows<16>(L, x0, y0, x1, y1, 2 * gl, 2, in);
  s += __shfl_xor_sync(0xffffffffu, s, 1);
  s += __shfl_xor_sync(0xffffffffu, s, 2);
  s += __shfl_xor_sync(0xffffffffu, s, 4);
  return s + (uint32_t)field_cost(nb, v, lambda);
}

// block_cost of v (returned to every lane) and of the group's cross point t (returned to the group) in ONE round of
// loads: the candidate's own cost and its first cross iteration are independent, and the search is bound by the latency
// of that chain.  v is spread over all 32 lanes (row lane / 2, 8-sample half lane & 1), t over the group as in block_cost.
__device__ __forceinline__ void block_cost_pair(const LevelView &L, const Nbrs &nb, int lambda, int xs, int ys, short2 v, short2 t,
                                                int lane, int gl, uint32_t &cost_v, uint32_t &cost_t) {
  const short2 v0 = to_pic0(v, L.wt0, L.wt1), t0 = to_pic0(t, L.wt0, L.wt1);
  const int vx0 = xs + pel(v0.x), vy0 = ys + pel(v0.y), vx1 = xs + pel(v.x), vy1 = ys + pel(v.y);
  const int tx0 = xs + pel(t0.x), ty0 = ys + pel(t0.y), tx1 = xs + pel(t.x), ty1 = ys + pel(t.y);
  const bool vin = inside(L, vx0, vy0, 16) && inside(L, vx1, vy1, 16), tin = inside(L, tx0, ty0, 16) && inside(L, tx1, ty1, 16);
  const int h8 = 8 * (lane & 1);
  uint32_t sv = sad_rows<8>(L, vx0 + h8, vy0, vx1 + h8, vy1, lane >> 1, 1, vin);
  uint32_t st = sad_rows<16>(L, tx0, ty0, tx1, ty1, 2 * gl, 2, tin);
  sv += __shfl_xor_sync(0xffffffffu, sv, 1);
  st += __shfl_xor_sync(0xffffffffu, st, 1);
  sv += __shfl_xor_sync(0xffffffffu, sv, 2);
  st += __shfl_xor_sync(0xffffffffu, st, 2);
  sv += __shfl_xor_sync(0xffffffffu, sv, 4);
  st += __shfl_xor_sync(0xffffffffu, st, 4);
  sv += __shfl_xor_sync(0xffffffffu, sv, 8);
  sv += __shfl_xor_sync(0xffffffffu, sv, 16);
  cost_v = sv + (uint32_t)field_cost(nb, v, lambda);
  cost_t = st + (uint32_t)field_cost(nb, t, lambda);
}

template <bool GUIDED>
__global__ void __launch_bounds__(EW * 32) k_estimate(const LevelView L) {
  __shared__ EstSmem S;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, grp = lane >> 3, gl = lane & 7;
  const int nrows = L.bh / 2, ncols = L.bw / 2, bw = L.bw;
  if (threadIdx.x == 0) S.row = atomicAdd(L.progress + nrows, 1);      // rows in start order: the row above is always running
  __syncthreads();
  const int row = S.row, by = 2 * row;
  if (row >= nrows) return;
  const int lambda = GUIDED ? 750 : 3000;
  const short2 zero = make_short2(0, 0);
  const short2 *above = L.pre + (ptrdiff_t)(by - 2) * bw;                // the row of vectors above (by > 0 only)
  // neighbours of the first block: wait until the row above has finished two blocks
  if (threadIdx.x == 0) {
    short2 u = zero, r = zero;
    if (row > 0) {
      const int need = min(2, ncols);
      while (ld_flag(L.progress + row - 1) < need) __nanosleep(20);
      asm volatile("fence.acquire.gpu;" ::: "memory");                   // pairs with the row above's release store: its vectors are visible now
      u = ld_mv_l2(above);
      if (bw > 2) r = ld_mv_l2(above + 2);
    }
    S.nxt_ur[0] = u;
    S.nxt_ur[1] = r;
  }
  __syncthreads();
  short2 left = zero, ul = zero, up = S.nxt_ur[0], ur = S.nxt_ur[1];
  __syncthreads();                                                       // S.nxt_ur[1] is rewritten during block 0

  for (int col = 0; col < ncols; col++) {
    const int bx = 2 * col, xs = bx * 8, ys = by * 8, pos = by * bw + bx, par = col & 1;
    Nbrs nb;
    nb.has_ur = by > 0 && bx < bw - 2;
    nb.has_up = by > 0;
    nb.has_left = bx > 0;
    nb.ur = nb.has_ur ? ur : zero;
    nb.up = up;
    nb.ul = ul;
    nb.left = left;
    nb.kind = (bx == 0 && by == 0) ? 0 : (by > 0 && bx > 0 && bx < bw - 2) ? 1 : by == 0 ? 2 : bx == 0 ? 3 : 0;

    if (warp == 0) {
      // skip vector: the neighbour closest (L1) to the others, the last one on ties (:761-782, :820-832)
      short2 v[3];
      int n = 0;
      if (nb.has_ur) v[n++] = nb.ur;
      if (nb.has_left) v[n++] = nb.left;
      if (nb.has_up) v[n++] = nb.up;
      short2 sk = zero;
      if (n) {
        int best = 0, bc = COST_MAX;
        for (int j = 0; j < n; j++) {
          int c = 0;
          for (int i = 0; i < n; i++) c += l1(v[i], v[j]);
          if (c <= bc) { best = j; bc = c; }
        }
        sk = v[best];
      }
      // skip test (:525-581): the four 8x8 SADs of the block, one per lane group
      const short2 sk0 = to_pic0(sk, L.wt0, L.wt1);
      const int x0 = xs + pel(sk0.x), y0 = ys + pel(sk0.y), x1 = xs + pel(sk.x), y1 = ys + pel(sk.y);
      const bool in = inside(L, x0, y0, 16) && inside(L, x1, y1, 16);
      const int qx = (grp & 1) * 8, qy = (grp >> 1) * 8;
      uint32_t s = in ? sad_rows<8>(L, x0 + qx, y0 + qy, x1 + qx, y1 + qy, gl, 1, true) : 0;
      s += __shfl_xor_sync(0xffffffffu, s, 1);
      s += __shfl_xor_sync(0xffffffffu, s, 2);
      s += __shfl_xor_sync(0xffffffffu, s, 4);
      const bool ok = __all_sync(0xffffffffu, in && s <= 8 * 64);
      if (lane == 0) { S.skip[par] = ok; S.skip_mv[par] = sk; }
      // ---- look ahead, off the critical path of the candidate warps ----
      if (col + 1 < ncols) {
        // the picture rows the next block will most likely read (this block's skip vector, one block to the right):
        // lanes 0..15 pic[1], 16..31 pic[0]; only every eighth block actually starts a new 128-byte line
        if (in && x0 + 32 <= L.w + L.pad && x1 + 32 <= L.w + L.pad) {
          const int r = lane & 15;
          prefetch_l1(lane < 16 ? L.y[1] + (ptrdiff_t)(y1 + r) * L.stride + x1 + 31 : L.y[0] + (ptrdiff_t)(y0 + r) * L.stride + x0 + 31);
        }
        // the next block's up-right vector, once the row above has produced it
        if (lane == 0) {
          short2 r = zero;
          if (row > 0) {
            const int need = min(col + 3, ncols);
            while (ld_flag(L.progress + row - 1) < need) __nanosleep(20);
            asm volatile("fence.acquire.gpu;" ::: "memory");             // one fence per block, after the poll: the loads stay relaxed
            if (bx + 4 < bw) r = ld_mv_l2(above + bx + 4);
          }
          S.nxt_ur[par ^ 1] = r;
        }
      }
    }
    // candidate list (:303-351): zero, guide, up-right, left, up, without duplicates
    short2 cand[5];
    int nc = 0;
    cand[nc++] = zero;
    {
      short2 t[4];
      bool has[4];
      if (GUIDED) {
        const short2 g = L.guide[(by / 2) * L.guide_bw + bx / 2];
        t[0] = make_short2((short)(g.x << 1), (short)(g.y << 1));                // :266-267
      } else t[0] = zero;
      has[0] = GUIDED;
      t[1] = nb.ur; has[1] = nb.has_ur;
      t[2] = nb.left; has[2] = nb.has_left;
      t[3] = nb.up; has[3] = nb.has_up;
#pragma unroll
      for (int k = 0; k < 4; k++) {
        if (!has[k]) continue;
        bool dup = false;
#pragma unroll
        for (int i = 0; i < 5; i++) dup |= i < nc && same(cand[i], t[k]);
        if (!dup) {
#pragma unroll
          for (int i = 1; i < 5; i++) if (i == nc) cand[i] = t[k];
          nc++;
        }
      }
    }
    const int c = warp - 1;
    if (warp >= 1 && c < nc) {
      // this warp's candidate, refined on its own by the cross search (:674-712)
      short2 cur = zero;
      uint32_t my_init = 0;
#pragma unroll
      for (int i = 0; i < 5; i++) if (i == c) cur = cand[i];
      int shift = GUIDED ? 3 : 6, budget = GUIDED ? 8 : 64;
      uint32_t cost = 0, tc = 0;
      bool first = true;                     // the first cross iteration always runs: priced together with the candidate
      while (shift >= 3 && budget > 0) {
        const int off = 1 << shift;
        short2 t = cur;
        if (grp == 0) t.x = (short)(cur.x - off);
        else if (grp == 1) t.x = (short)(cur.x + off);
        else if (grp == 2) t.y = (short)(cur.y - off);
        else t.y = (short)(cur.y + off);
        if (first) {
          block_cost_pair(L, nb, lambda, xs, ys, cur, t, lane, gl, cost, tc);
          my_init = cost;
          first = false;
        } else tc = block_cost(L, nb, lambda, xs, ys, t, gl);
        // the reference takes the four points in order, each against the best so far: the smallest wins, the first on ties
        uint32_t bestc = cost;
        int bestk = -1;
#pragma unroll
        for (int k = 0; k < 4; k++) {
          const uint32_t ck = __shfl_sync(0xffffffffu, tc, 8 * k);
          if (ck < bestc) { bestc = ck; bestk = k; }
        }
        if (bestk >= 0) {
          cost = bestc;
          if (bestk == 0) cur.x = (short)(cur.x - off);
          else if (bestk == 1) cur.x = (short)(cur.x + off);
          else if (bestk == 2) cur.y = (short)(cur.y - off);
          else cur.y = (short)(cur.y + off);
        } else shift--;
        budget -= 4;
      }
      if (lane == 0) { S.init[par][c] = my_init; S.fin[par][c] = cost; S.mv[par][c] = cur; }
    }
    __syncthreads();                                                    // the only barrier of a block
    short2 best;
    if (S.skip[par]) best = S.skip_mv[par];
    else {
      // the reference's sequential selection (:685, :713-717) replayed on the speculative results
      uint32_t best_cost = COST_MAX;
      best = cand[0];
      for (int i = 0; i < nc; i++) {
        uint32_t cost = S.init[par][i];
        short2 mv = cand[0];
#pragma unroll
        for (int k = 0; k < 5; k++) if (k == i) mv = cand[k];
        if (((uint32_t)(4 + i) * cost) / 8 < best_cost) { cost = S.fin[par][i]; mv = S.mv[par][i]; }
        if (cost < best_cost) { best_cost = cost; best = mv; }
      }
    }
    if (threadIdx.x == EW * 32 - 1) {
      // publish: the four 8x8 entries of the block, then the progress counter (release orders it after them).  Done by a
      // thread of the last warp, so that warp 0 starts the next block's skip test at once.
      st_mv_l2(L.pre + pos, best);
      st_mv_l2(L.pre + pos + 1, best);
      st_mv_l2(L.pre + pos + bw, best);
      st_mv_l2(L.pre + pos + bw + 1, best);
      st_release(L.progress + row, col + 1);
    }
    // the next block's neighbours: its up-left is this block's up, its up is this block's up-right
    ul = up;
    up = nb.ur;
    ur = S.nxt_ur[par ^ 1];
    left = best;
  }
}

// ---------------------------------------------------------------- merge pass (:288-301, :727-759, :898-917) ----
// one warp per 8x8 block; lane group g evaluates candidate g (the fifth one in a second round)
constexpr int MW = 8;
__global__ void __launch_bounds__(MW * 32) k_merge(const LevelView L) {
  const int lane = threadIdx.x & 31, grp = lane >> 3, gl = lane & 7;
  const int pos = blockIdx.x * MW + (threadIdx.x >> 5);
  if (pos >= L.bw * L.bh) return;
  const int by = pos / L.bw, bx = pos - by * L.bw, off = (by & 1) ? 2 : 1;     // the distance follows the ROW parity in both directions
  short2 cand[5];
  int nc = 0;
  {
    short2 t[5];
    bool has[5];
    t[0] = L.pre[pos]; has[0] = true;
    has[1] = by - off >= 0; t[1] = has[1] ? L.pre[pos - off * L.bw] : t[0];
    has[2] = by + off < L.bh; t[2] = has[2] ? L.pre[pos + off * L.bw] : t[0];
    has[3] = bx - off >= 0; t[3] = has[3] ? L.pre[pos - off] : t[0];
    has[4] = bx + off < L.bw; t[4] = has[4] ? L.pre[pos + off] : t[0];
#pragma unroll
    for (int k = 0; k < 5; k++) {
      cand[k] = t[0];
      if (!has[k]) continue;
      bool dup = false;
#pragma unroll
      for (int i = 0; i < 5; i++) dup |= i < nc && i < k && same(cand[i], t[k]);
      if (!dup) {
#pragma unroll
        for (int i = 0; i < 5; i++) if (i == nc) cand[i] = t[k];
        nc++;
      }
    }
  }
  short2 best = cand[0];
  if (nc > 1) {
    uint32_t best_cost = COST_MAX;
    const int xs = bx * 8, ys = by * 8;
#pragma unroll
    for (int round = 0; round < 2; round++) {
      const int c = round * 4 + grp;
      if (round == 1 && nc <= 4) break;
      short2 v = cand[0];
#pragma unroll
      for (int i = 0; i < 5; i++) if (i == c) v = cand[i];
      const short2 v0 = to_pic0(v, L.wt0, L.wt1);
      const int x0 = xs + pel(v0.x), y0 = ys + pel(v0.y), x1 = xs + pel(v.x), y1 = ys + pel(v.y);
      const bool in = inside(L, x0, y0, 8) && inside(L, x1, y1, 8);
      uint32_t s = c < nc ? sad_rows<8>(L, x0, y0, x1, y1, gl, 1, in) : 0;
      s += __shfl_xor_sync(0xffffffffu, s, 1);
      s += __shfl_xor_sync(0xffffffffu, s, 2);
      s += __shfl_xor_sync(0xffffffffu, s, 4);
#pragma unroll
      for (int k = 0; k < 4; k++) {
        const uint32_t ck = __shfl_sync(0xffffffffu, s, 8 * k);
        const int ci = round * 4 + k;
        if (ci < nc && ck < best_cost) {
          best_cost = ck;
#pragma unroll
          for (int i = 0; i < 5; i++) if (i == ci) best = cand[i];
        }
      }
    }
  }
  if (lane == 0) L.post[pos] = best;
}

// ---------------------------------------------------------------- motion-compensated average (:387-441, :920-970) ----
struct BlendParams {
  const uint8_t *p0[3], *p1[3];   // pic[0], pic[1] plane origins
  uint8_t *out[3];
  int rstride[3], ostride[3];
  int w, h;                       // luma size
  int bw, bh;
  const short2 *mv;
  int wt0, wt1;
};

// 16 threads per 8x8 field entry: 8 luma rows, 4 U rows, 4 V rows
__global__ void __launch_bounds__(256) k_blend(const BlendParams P) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x, pos = t >> 4, k = t & 15;
  if (pos >= P.bw * P.bh) return;
  const int by = pos / P.bw, bx = pos - by * P.bw;
  const int pl = k < 8 ? 0 : k < 12 ? 1 : 2, i = k < 8 ? k : (k - 8) & 3;
  const int size = pl ? 4 : 8, pad = pl ? 2 : 4;
  const int wP = pl ? (P.w + 4) / 2 : P.w + 4, hP = pl ? (P.h + 4) / 2 : P.h + 4;
  const int vw = pl ? P.w / 2 : P.w, vh = pl ? P.h / 2 : P.h;
  const int xs = bx * size, ys = by * size;
  if (ys + i >= vh || xs >= vw) return;        // the overhang of the last block row / column is border, rebuilt by the caller's padding
  short2 v1 = P.mv[pos], v0;
  if (pl) { v1.x >>= 1; v1.y >>= 1; }          // :935-937
  v0 = to_pic0(v1, P.wt0, P.wt1);
  const int x0 = xs + pel(v0.x), y0 = ys + pel(v0.y), x1 = xs + pel(v1.x), y1 = ys + pel(v1.y);
  const bool in0 = x0 >= -pad && x0 + size <= wP && y0 >= -pad && y0 + size <= hP;
  const bool in1 = x1 >= -pad && x1 + size <= wP && y1 >= -pad && y1 + size <= hP;
  const uint8_t *r0 = P.p0[pl], *r1 = P.p1[pl];
  const int rs = P.rstride[pl];
  uint8_t *o = P.out[pl] + (ptrdiff_t)(ys + i) * P.ostride[pl] + xs;
  uint32_t a[2] = {0, 0}, b[2] = {0, 0}, res[2];
  const int nw = size / 4;
  if (in0 || in1) {
    if (in0) { if (pl) a[0] = ldrow4(r0 + (ptrdiff_t)(y0 + i) * rs + x0); else ldrow8(r0 + (ptrdiff_t)(y0 + i) * rs + x0, a); }
    if (in1) { if (pl) b[0] = ldrow4(r1 + (ptrdiff_t)(y1 + i) * rs + x1); else ldrow8(r1 + (ptrdiff_t)(y1 + i) * rs + x1, b); }
    for (int q = 0; q < 2; q++) res[q] = (in0 && in1) ? __vavgu4(a[q], b[q]) : in1 ? b[q] : a[q];
  } else {
    const int ay = min(hP - 1, max(-pad, i + y0)), cy = min(hP - 1, max(-pad, i + y1));
    res[0] = res[1] = 0;
    for (int j = 0; j < size; j++) {
      const int ax = min(wP - 1, max(-pad, j + x0)), cx = min(wP - 1, max(-pad, j + x1));
      const uint32_t v = (r0[(ptrdiff_t)ay * rs + ax] + r1[(ptrdiff_t)cy * rs + cx] + 1) >> 1;
      res[j >> 2] |= v << (8 * (j & 3));
    }
  }
  const int vis = min(size, vw - xs);          // visible samples of this row
  if (vis == size && (reinterpret_cast<uintptr_t>(o) & 3) == 0) {
    for (int q = 0; q < nw; q++) reinterpret_cast<uint32_t *>(o)[q] = res[q];
  } else {
    for (int j = 0; j < vis; j++) o[j] = (uint8_t)(res[j >> 2] >> (8 * (j & 3)));
  }
}

}  // namespace thor

using namespace thor;

namespace {

// Device scratch of the interpolation, kept between calls (grow-only): the down-scaled levels, the fields, the
// wavefront counters.
struct InterpState {
  DevBuf levels, fields, counters;
  int fw = 0, fh = 0, fbw = 0, fbh = 0;     // geometry of the last call (thorcuda_interp_field)
  short2 *field0 = nullptr;
} g_interp;

int interp_levels(int w, int h) {
  const int m = w < h ? w : h;
  const int l = (int)(log10((double)m) / log10(2.0) - 4.0);      // temporal_interp.c:977
  return l < 4 ? l : 4;
}

}  // namespace

namespace thor {
void interp_reset() {
  g_interp.levels.release();
  g_interp.fields.release();
  g_interp.counters.release();
  g_interp = InterpState();
}
}  // namespace thor

extern "C" {

int thorcuda_interp_levels(int width, int height) { return interp_levels(width, height); }

int thorcuda_interpolate_frames(thorcuda_pic dst_h, thorcuda_pic ref0_h, thorcuda_pic ref1_h, int ratio, int pos) {
  THOR_NEED_CTX();
  Context &c = ctx();
  Pic *dst = get_pic(dst_h), *r0 = get_pic(ref0_h), *r1 = get_pic(ref1_h);
  if (!dst || !r0 || !r1) return fail("interpolate: bad picture handle");
  if (dst == r0 || dst == r1) return fail("interpolate: the output must not be one of the references");
  if (r0->w != r1->w || r0->h != r1->h || r0->pad_y != r1->pad_y || r0->pad_c != r1->pad_c || dst->w != r0->w || dst->h != r0->h)
    return fail("interpolate: pictures differ in geometry");
  if (r0->pad_y < 4 || r0->pad_c < 2) return fail("interpolate: the references need a border of at least 4 (luma) / 2 (chroma) samples");
  if (ratio < 2 || pos < 1 || pos >= ratio) return fail("interpolate: position %d outside (0, %d)", pos, ratio);
  const int w = r0->w, h = r0->h, nlev = interp_levels(w, h);
  if (nlev < 1) return 0;                                        // the reference leaves the output untouched (:977, :1022)
  const int reversed = pos > ratio / 2;                          // :120-126
  const int wt0 = reversed ? pos : ratio - pos, wt1 = ratio - wt0;
  const PicView v0 = (reversed ? r1 : r0)->view(), v1 = (reversed ? r0 : r1)->view();

  // ---- layout of the scratch ----
  LevelView L[4];
  memset(L, 0, sizeof(L));
  size_t lev_bytes = 0, field_bytes = 0, cnt_ints = 0;
  size_t lev_off[4][2], pre_off[4], post_off[4], cnt_off[4];
  for (int l = 0; l < nlev; l++) {
    LevelView &q = L[l];
    q.w = w >> l;
    q.h = h >> l;
    q.bw = 2 * ((q.w + 15) / 16);                                // :97-99
    q.bh = 2 * ((q.h + 15) / 16);
    q.wt0 = wt0;
    q.wt1 = wt1;
    if (l == 0) { q.stride = v0.stride[0]; q.pad = r0->pad_y; }
    else {
      q.pad = 32;                                                // :1004-1005
      q.stride = (q.w + 64 + 3 + 15) & ~15;
      for (int k = 0; k < 2; k++) { lev_off[l][k] = lev_bytes; lev_bytes += (size_t)q.stride * (q.h + 64) + 256; }
    }
    pre_off[l] = field_bytes; field_bytes += (size_t)q.bw * q.bh * sizeof(short2);
    post_off[l] = field_bytes; field_bytes += (size_t)q.bw * q.bh * sizeof(short2);
    cnt_off[l] = cnt_ints; cnt_ints += q.bh / 2 + 1;
  }
  uint8_t *lev = (uint8_t *)g_interp.levels.reserve(lev_bytes + 256);
  uint8_t *fld = (uint8_t *)g_interp.fields.reserve(field_bytes);
  int *cnt = (int *)g_interp.counters.reserve(cnt_ints * sizeof(int));
  if (!lev || !fld || !cnt) return fail("interpolate: device allocation failed");
  THOR_CUDA(cudaMemsetAsync(cnt, 0, cnt_ints * sizeof(int), c.stream));
  for (int l = 0; l < nlev; l++) {
    LevelView &q = L[l];
    if (l == 0) { q.y[0] = v0.p[0]; q.y[1] = v1.p[0]; }
    else for (int k = 0; k < 2; k++) q.y[k] = lev + lev_off[l][k] + (size_t)32 * q.stride + 32;
    q.pre = (short2 *)(fld + pre_off[l]);
    q.post = (short2 *)(fld + post_off[l]);
    q.progress = cnt + cnt_off[l];
  }
  for (int l = 0; l + 1 < nlev; l++) { L[l].guide = L[l + 1].post; L[l].guide_bw = L[l + 1].bw; }

  // ---- pyramid ----
  for (int l = 1; l < nlev; l++) {
    HalveParams H;
    for (int k = 0; k < 2; k++) { H.src[k] = L[l - 1].y[k]; H.dst[k] = const_cast<uint8_t *>(L[l].y[k]); }
    H.sstride = L[l - 1].stride; H.dstride = L[l].stride; H.wo = L[l].w; H.ho = L[l].h; H.pad = 32;
    const int items = ((H.wo + 64 + 3) / 4) * (H.ho + 64);
    k_halve<<<dim3((items + 255) / 256, 2), 256, 0, c.stream>>>(H);
    THOR_LAUNCHED();
  }
  // ---- fields, coarse to fine ----
  for (int l = nlev - 1; l >= 0; l--) {
    if (l == nlev - 1) k_estimate<false><<<L[l].bh / 2, EW * 32, 0, c.stream>>>(L[l]);
    else k_estimate<true><<<L[l].bh / 2, EW * 32, 0, c.stream>>>(L[l]);
    THOR_LAUNCHED();
    k_merge<<<(L[l].bw * L[l].bh + MW - 1) / MW, MW * 32, 0, c.stream>>>(L[l]);
    THOR_LAUNCHED();
  }
  // ---- output ----
  BlendParams B;
  const PicView vd = dst->view();
  for (int pl = 0; pl < 3; pl++) {
    B.p0[pl] = v0.p[pl]; B.p1[pl] = v1.p[pl]; B.out[pl] = vd.p[pl];
    B.rstride[pl] = v0.stride[pl]; B.ostride[pl] = vd.stride[pl];
  }
  B.w = w; B.h = h; B.bw = L[0].bw; B.bh = L[0].bh; B.mv = L[0].post; B.wt0 = wt0; B.wt1 = wt1;
  k_blend<<<(B.bw * B.bh * 16 + 255) / 256, 256, 0, c.stream>>>(B);
  THOR_LAUNCHED();
  g_interp.fw = w; g_interp.fh = h; g_interp.fbw = B.bw; g_interp.fbh = B.bh; g_interp.field0 = L[0].post;
  return 0;
}

int thorcuda_interp_field(int16_t *mv1, int *bw, int *bh) {
  THOR_NEED_CTX();
  if (!g_interp.field0) return fail("interp_field: no interpolation has run");
  if (bw) *bw = g_interp.fbw;
  if (bh) *bh = g_interp.fbh;
  if (mv1) {
    if (d2h(mv1, g_interp.field0, (size_t)g_interp.fbw * g_interp.fbh * sizeof(short2))) return -1;
    THOR_CUDA(cudaStreamSynchronize(ctx().stream));
  }
  return 0;
}

}  // extern "C"
