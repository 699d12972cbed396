// Frame encoder of libthorcuda -- motion-search costs and the reference's search pattern (motion_estimate).
#pragma once
#include "enc_pred.h"

namespace thorenc {

// sad_calc (enc/encode_block.c:740-755) for K candidate positions at once: the width x height block `a` (shared, stride
// as) against ref + off[k] (reference picture, stride bs; off[] = S.me_off, written by the caller before the call).  Results in S.acc[0..K).  Ends with a
// barrier.  `wide`: each candidate is widesad_calc's five positions x + {-3,-1,0,1,3} (:757-780); results in
// S.acc[5k .. 5k+5).
//
// Device mapping: a candidate belongs to ONE WARP (or, when there are fewer candidates than warps, to 2 / 4 / 8 warps that
// split its rows): a lane walks 4-sample words of the block, the warp reduces once per candidate, nobody divides (block
// sides are powers of two) and the five wide positions come out of the same four aligned words.  Before round 2's last
// profile this function was a flat item loop over all candidates with four integer divisions and a shuffle reduction per
// word: 21 % of k_encode_frame's samples at LDB medium (profiles/r2_ncu_encode_frame_1080p.txt).
#ifndef THOR_HOSTEMU
TD uint32_t window_word(uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3, int off) {      // 4 bytes at byte offset off (0..9) of w0..w3
  const unsigned sh = (unsigned)(off & 3) * 8;
  return off < 4 ? __funnelshift_r(w0, w1, sh) : off < 8 ? __funnelshift_r(w1, w2, sh) : __funnelshift_r(w2, w3, sh);
}
#endif
TDN void sad_batch(Grp &g, SbMem &S, const uint8_t *a, int as, const uint8_t *ref, int bs, const int *off, int K, int width, int height, int wide) {
  PROF(PF_SAD);
  THOR_LOCAL(&g);
  THOR_SHS(&S); THOR_SHARED(a);
  const int per = wide ? 5 : 1, wq = width >> 2, items = height * wq;
  g.sync();                                                // readers of the previous batch's results are done; off[] is written
#ifdef THOR_HOSTEMU
  for (int k = 0; k < K; k++)
    for (int w = 0; w < per; w++) {
      const int dx = wide ? (w == 0 ? -3 : w == 1 ? -1 : w == 2 ? 0 : w == 3 ? 1 : 3) : 0;
      int v = 0;
      for (int it = 0; it < items; it++) {
        const int r = it / wq, c = (it % wq) * 4;
        v += (int)sad4(*reinterpret_cast<const uint32_t *>(a + r * as + c), ld_ro32u(ref + off[k] + dx + r * bs + c));
      }
      S.acc[k * per + w] = v;
    }
#else
  const int lane = g.tid & 31, warp = g.tid >> 5;
  const int lwq = 31 - __clz(wq);                          // width is 4, 8, 16, 32 or 64
  // warps per candidate: as many as keeps every warp busy, but no more than there are 32-word slices of the block
  int sp = 1;
  while (K * sp * 2 <= NW && sp * 2 * 32 <= items) sp *= 2;
  if (sp > 1) {
    TFOR(g, i, K * per) S.acc[i] = 0;
    g.sync();
  }
  const int chunk = items / sp;                            // items and sp are powers of two
  // A lane keeps its word column (32 is a multiple of the words per row) and steps down the rows with two pointer
  // increments; the reference rows share one byte alignment (the picture stride is a multiple of 16), so the aligned
  // word pointer and the funnel shift are set up once per candidate.
  const int c = (lane & (wq - 1)) << 2, rstep = 32 >> lwq;
  const int passes = chunk >= 32 ? chunk >> 5 : 1;
  const bool active = lane < chunk;
  for (int u = warp; u < K * sp; u += NW) {
    const int k = u / sp, part = u - k * sp;
    const int row0 = ((part * chunk) >> lwq) + (lane >> lwq);
    const uint8_t *pa = a + row0 * as + c;
    if (!wide) {
      const uintptr_t ad = reinterpret_cast<uintptr_t>(ref + off[k] + row0 * bs + c);
      const uint32_t *q = reinterpret_cast<const uint32_t *>(ad & ~(uintptr_t)3);
      const unsigned sh = (unsigned)(ad & 3) * 8;
      const int qstep = (rstep * bs) >> 2, astep = rstep * as;
      int v = 0;
      if (active)
        for (int j = 0; j < passes; j++) {
          v += (int)sad4(*reinterpret_cast<const uint32_t *>(pa), __funnelshift_r(__ldg(q), __ldg(q + 1), sh));
          pa += astep;
          q += qstep;
        }
      v = __reduce_add_sync(0xffffffffu, v);
      if (lane == 0) { if (sp > 1) atomicAdd(&S.acc[k], v); else S.acc[k] = v; }
    } else {
      const uintptr_t ad = reinterpret_cast<uintptr_t>(ref + off[k] + row0 * bs + c - 3);
      const uint32_t *q = reinterpret_cast<const uint32_t *>(ad & ~(uintptr_t)3);
      const int o0 = (int)(ad & 3);
      const int qstep = (rstep * bs) >> 2, astep = rstep * as;
      int v0 = 0, v1 = 0, v2 = 0, v3 = 0, v4 = 0;
      if (active)
        for (int j = 0; j < passes; j++) {
          const uint32_t A = *reinterpret_cast<const uint32_t *>(pa);
          const uint32_t w0 = __ldg(q), w1 = __ldg(q + 1), w2 = __ldg(q + 2), w3 = __ldg(q + 3);
          v0 += (int)sad4(A, window_word(w0, w1, w2, w3, o0));          // x - 3
          v1 += (int)sad4(A, window_word(w0, w1, w2, w3, o0 + 2));      // x - 1
          v2 += (int)sad4(A, window_word(w0, w1, w2, w3, o0 + 3));      // x
          v3 += (int)sad4(A, window_word(w0, w1, w2, w3, o0 + 4));      // x + 1
          v4 += (int)sad4(A, window_word(w0, w1, w2, w3, o0 + 6));      // x + 3
          pa += astep;
          q += qstep;
        }
      v0 = __reduce_add_sync(0xffffffffu, v0); v1 = __reduce_add_sync(0xffffffffu, v1); v2 = __reduce_add_sync(0xffffffffu, v2);
      v3 = __reduce_add_sync(0xffffffffu, v3); v4 = __reduce_add_sync(0xffffffffu, v4);
      if (lane < 5) {
        const int v = lane == 0 ? v0 : lane == 1 ? v1 : lane == 2 ? v2 : lane == 3 ? v3 : v4;
        if (sp > 1) atomicAdd(&S.acc[5 * k + lane], v); else S.acc[5 * k + lane] = v;
      }
    }
  }
#endif
  g.sync();
}

// the minimum of widesad_calc: first minimum in the order -3,-1,0,1,3 (enc/encode_block.c:764-776 = the packed minimum of
// enc/enc_kernels.c:94-97)
TD unsigned widesad_pick(const int *acc5, int *x) {
  THOR_SHARED(acc5);
  unsigned best = 1u << 31;
  int bx = 0;
  for (int k = 0; k < 5; k++) {
    const unsigned s = (unsigned)acc5[k];
    if (s < best) { best = s; bx = k == 0 ? -3 : k == 1 ? -1 : k == 2 ? 0 : k == 3 ? 1 : 3; }
  }
  *x = bx;
  return best;
}

TD int upavg(int a, int b) { return (a + b + 1) >> 1; }
TD int dnavg(int a, int b) { return (a + b) >> 1; }

// sad_calc_fasthalf (enc/encode_block.c:497-605): eight bilinear half-pel candidates around the integer position b.
TDN unsigned sad_fasthalf(Grp &g, const uint8_t *a, int as, const uint8_t *b, int bs, int width, int height, int *x, int *y) {
  PROF(PF_HALF);
  THOR_LOCAL(&g);
  THOR_SHARED(a);
  int s[8] = {0, 0, 0, 0, 0, 0, 0, 0};                     // top down left right tl tr bl br
  TFOR(g, i, width * height) {
    const int r = p2div(i, width), c = p2mod(i, width);
    const uint8_t *p = b + r * bs + c;
    const int A = a[r * as + c];
    // the neighbourhood rows -2..+2 (only the samples the trees use)
    const int m2m1 = ld_ro8(p - 2 * bs - 1), m2_0 = ld_ro8(p - 2 * bs), m2p1 = ld_ro8(p - 2 * bs + 1);
    const int m1m2 = ld_ro8(p - bs - 2), m1m1 = ld_ro8(p - bs - 1), m1_0 = ld_ro8(p - bs), m1p1 = ld_ro8(p - bs + 1), m1p2 = ld_ro8(p - bs + 2);
    const int z_m2 = ld_ro8(p - 2), z_m1 = ld_ro8(p - 1), z_0 = ld_ro8(p), z_p1 = ld_ro8(p + 1), z_p2 = ld_ro8(p + 2);
    const int p1m2 = ld_ro8(p + bs - 2), p1m1 = ld_ro8(p + bs - 1), p1_0 = ld_ro8(p + bs), p1p1 = ld_ro8(p + bs + 1), p1p2 = ld_ro8(p + bs + 2);
    const int p2m1 = ld_ro8(p + 2 * bs - 1), p2_0 = ld_ro8(p + 2 * bs), p2p1 = ld_ro8(p + 2 * bs + 1);
    int t1, t2, t3, t4, t5, t6, t7, t8;
    t1 = upavg(m1m1, m1_0); t2 = upavg(z_m1, z_0); t1 = dnavg(t1, t2);
    t3 = upavg(m2m1, p1m1); t4 = upavg(m2_0, p1_0); t3 = dnavg(t3, t4);
    t5 = upavg(m1m2, m1p1); t6 = upavg(z_m2, z_p1); t5 = dnavg(t5, t6);
    t5 = dnavg(t3, t5);
    const int ptl = dnavg(t5, t1);
    s[2] += iabs(A - t2);                                  // left
    t1 = upavg(m1_0, m1p1); t8 = upavg(z_0, z_p1); t1 = dnavg(t1, t8);
    t5 = upavg(m2p1, p1p1); t3 = dnavg(t4, t5);
    t4 = upavg(m1m1, m1p2); t7 = upavg(z_m1, z_p2); t5 = dnavg(t7, t4);
    t5 = dnavg(t3, t5);
    const int ptr = dnavg(t5, t1);
    s[3] += iabs(A - t8);                                  // right
    t1 = upavg(p1m1, p1_0); t3 = dnavg(t1, t2);
    t2 = upavg(m1m1, p2m1); t4 = upavg(m1_0, p2_0); t5 = dnavg(t4, t2);
    t1 = upavg(p1m2, p1p1); t2 = dnavg(t6, t1);
    t2 = dnavg(t5, t2);
    const int pbl = dnavg(t2, t3);
    t2 = upavg(p1_0, p1p1); t3 = dnavg(t8, t2);
    t5 = upavg(m1p1, p2p1); t6 = dnavg(t4, t5);
    t8 = upavg(p1m1, p1p2); t1 = dnavg(t7, t8);
    t2 = dnavg(t6, t1);
    const int pbr = dnavg(t2, t3);
    s[1] += iabs(A - upavg(z_0, p1_0));                    // down
    s[0] += iabs(A - upavg(z_0, m1_0));                    // top
    s[4] += iabs(A - ptl);
    s[5] += iabs(A - ptr);
    s[7] += iabs(A - pbr);
    s[6] += iabs(A - pbl);
  }
  g.sumk(s);
  unsigned best = (unsigned)s[0];
  int bx = 0, by = -2;
  if ((unsigned)s[1] < best) { best = (unsigned)s[1]; bx = 0; by = 2; }
  if ((unsigned)s[3] < best) { best = (unsigned)s[3]; bx = 2; by = 0; }
  if ((unsigned)s[2] < best) { best = (unsigned)s[2]; bx = -2; by = 0; }
  if ((unsigned)s[4] < best) { best = (unsigned)s[4]; bx = -2; by = -2; }
  if ((unsigned)s[5] < best) { best = (unsigned)s[5]; bx = 2; by = -2; }
  if ((unsigned)s[7] < best) { best = (unsigned)s[7]; bx = 2; by = 2; }
  if ((unsigned)s[6] < best) { best = (unsigned)s[6]; bx = -2; by = 2; }
  *x = bx; *y = by;
  return best;
}

// sad_calc_fastquarter (enc/encode_block.c:609-738): eight quarter-pel candidates; the incoming (*x, *y) select one of
// four formula sets.
TDN unsigned sad_fastquarter(Grp &g, const uint8_t *o, int os, const uint8_t *rf, int rs, int width, int height, int *x, int *y) {
  PROF(PF_QUARTER);
  THOR_LOCAL(&g);
  THOR_SHARED(o);
  const int sel = (*x & *y) ? 0 : (*x ? 1 : (*y ? 2 : 3));
  int s[8] = {0, 0, 0, 0, 0, 0, 0, 0};                     // tl top tr left right bl down br
  TFOR(g, i, width * height) {
    const int r0 = p2div(i, width), j = p2mod(i, width);
    const uint8_t *r = rf + r0 * rs + j;
    const int O = o[r0 * os + j];
    const int a = ld_ro8(r), b = ld_ro8(r - rs), c = ld_ro8(r - rs + 1), d = ld_ro8(r + 1), e = ld_ro8(r + rs + 1), f = ld_ro8(r + rs),
              gg = ld_ro8(r + rs - 1), hh = ld_ro8(r - 1);
    const int ad = upavg(a, d), de = upavg(d, e), af = upavg(a, f), fe = upavg(f, e), dc = upavg(d, c), ab = upavg(a, b), ah = upavg(a, hh),
              gf = upavg(gg, f);
    int v[8];
    if (sel == 0) { v[0] = dnavg(ad, af); v[1] = dnavg(de, a); v[2] = dnavg(ad, de); v[3] = dnavg(ad, f); v[4] = dnavg(ad, e); v[5] = dnavg(af, fe); v[6] = dnavg(de, f); v[7] = dnavg(de, fe); }
    else if (sel == 1) { v[0] = dnavg(ad, ab); v[1] = dnavg(dc, a); v[2] = dnavg(ad, dc); v[3] = dnavg(ad, a); v[4] = dnavg(ad, d); v[5] = dnavg(ad, af); v[6] = dnavg(af, d); v[7] = dnavg(ad, de); }
    else if (sel == 2) { v[0] = dnavg(ah, af); v[1] = dnavg(af, a); v[2] = dnavg(ad, af); v[3] = dnavg(gf, a); v[4] = dnavg(ad, f); v[5] = dnavg(af, gf); v[6] = dnavg(af, f); v[7] = dnavg(af, fe); }
    else { v[0] = dnavg(ah, ab); v[1] = dnavg(ab, a); v[2] = dnavg(ad, ab); v[3] = dnavg(ah, a); v[4] = dnavg(ad, a); v[5] = dnavg(ah, af); v[6] = dnavg(af, a); v[7] = dnavg(af, ad); }
#pragma unroll
    for (int k = 0; k < 8; k++) s[k] += iabs(O - v[k]);
  }
  g.sumk(s);
  unsigned best = (unsigned)s[1];
  int bx = 0, by = -1;
  if ((unsigned)s[0] < best) { best = (unsigned)s[0]; bx = -1; }
  if ((unsigned)s[2] < best) { best = (unsigned)s[2]; bx = 1; }
  if ((unsigned)s[3] < best) { best = (unsigned)s[3]; bx = -1; by = 0; }
  if ((unsigned)s[4] < best) { best = (unsigned)s[4]; bx = 1; by = 0; }
  if ((unsigned)s[5] < best) { best = (unsigned)s[5]; bx = -1; by = 1; }
  if ((unsigned)s[6] < best) { best = (unsigned)s[6]; bx = 0; by = 1; }
  if ((unsigned)s[7] < best) { best = (unsigned)s[7]; bx = 1; by = 1; }
  *x = bx; *y = by;
  return best;
}

// (unsigned)(lambda * bits + 0.5) of the vector cost (enc/encode_block.c:863)
TD unsigned mv_cost(double lambda, Mv mv, Mv mvp) { return (unsigned)dadd(dmul(lambda, (double)mv_bits(mv.y - mvp.y, mv.x - mvp.x)), 0.5); }

// add_mvcandidate (enc/encode_block.c:58-73)
TD void add_mvcandidate(Grp &g, SbMem &S, SbState &st, int ref_idx, Mv mv) {
  THOR_LOCAL(&st); THOR_LOCAL(&g);
  THOR_SHS(&S);
  const Mv imv = mkmv((mv.x + 2) >> 2, (mv.y + 2) >> 2);
  const uint64_t m = (uint64_t)1 << ((((int)imv.y << 3) ^ (int)imv.x) & 63);
  if (!(m & st.mvcand_mask[ref_idx])) {
    g.sync();                                              // readers of the list are done
    if (g.tid == 0) S.mvcand[ref_idx][st.mvcand_num[ref_idx]] = imv;
    g.sync();
    st.mvcand_num[ref_idx]++;
  }
  st.mvcand_mask[ref_idx] |= m;
}

// The searches below price up to 64 candidates and keep the first cheapest one (the reference's `if (sad < min_sad)` walk).
// On the device a lane owns candidates k = lane, lane + 32: it forms the vector, prices it, and the warp agrees on
// (smallest cost, then smallest k) with shuffles -- the same answer as the sequential walk, in a 25th of the instructions
// every warp used to replay.  The host emulation is one lane and walks all candidates in order.
#ifdef THOR_HOSTEMU
constexpr int ME_LANES = 1;
TD int me_lane(const Grp &) { return 0; }
TD void me_argmin(unsigned &, int &, Mv &) {}
#else
constexpr int ME_LANES = 32;
TD int me_lane(const Grp &g) { return g.tid & 31; }
TD void me_argmin(unsigned &sad, int &k, Mv &mv) {
#pragma unroll
  for (int o = 16; o; o >>= 1) {
    const unsigned s2 = __shfl_xor_sync(0xffffffffu, sad, o);
    const int k2 = __shfl_xor_sync(0xffffffffu, k, o);
    const int m2 = __shfl_xor_sync(0xffffffffu, ((int)(uint16_t)mv.x) | ((int)mv.y << 16), o);
    if (s2 < sad || (s2 == sad && k2 < k)) { sad = s2; k = k2; mv = mkmv((int16_t)(m2 & 0xffff), (int16_t)(m2 >> 16)); }
  }
}
#endif

// motion_estimate (enc/encode_block.c:830-1016).  orig: shared, stride `size` (the coding block's size); ref: block origin
// in the reference luma plane.  Returns min(cmin, min_sad) and the vector.
TDN unsigned motion_estimate(Grp &g, SbMem &S, const thor_enc_frame &F, const uint8_t *orig, const uint8_t *ref, int size, int stride_r, int width,
                             int height, Mv *mv_out, Mv mvc, Mv mvp, double lambda, int sign, int xpos, int ypos, const Mv *mvcand, int mvcand_num,
                             int enable_bipred) {
  PROF(PF_ME);
  THOR_LOCAL(&g);
  THOR_SHS(&S); THOR_SHARED(orig); THOR_SHARED(mvcand); THOR_SHF(&F);
  const int s = sign ? -1 : 1, speed = F.encoder_speed, fw = F.width, fh = F.height;
  const int lane = me_lane(g);
  unsigned min_sad = 1u << 31;
  Mv mv_opt = mkmv(0, 0), mv_ref = mkmv(((mvc.x + 2) >> 2) << 2, ((mvc.y + 2) >> 2) << 2);
  Mv mine[64 / ME_LANES];                                  // this lane's candidates of the current batch (k = lane, lane + 32)

  if ((size == 16 && enable_bipred) || speed == 0) {
    // telescope search: 5x5 grids of +-2 step, step = 32, 16, 8, 4 quarter-samples (:843-874); the centre only at step 32
    for (int step = 32; step >= 4; step >>= 1) {
      const int wide = step == 32 && size == 16 && speed < 2 && speed > 0;
      const int K = step < 32 ? 24 : 25;
      for (int k = lane, n = 0; k < K; k += ME_LANES, n++) {
        const int e = (step < 32 && k >= 12) ? k + 1 : k;
        const Mv cand = clip_mv(mkmv(mv_ref.x + (e % 5 - 2) * step, mv_ref.y + (e / 5 - 2) * step), ypos, xpos, fw, fh, size, sign);
        mine[n] = cand;
        S.me_off[k] = s * (cand.x >> 2) + s * (cand.y >> 2) * stride_r;      // every warp writes the same values
      }
      sad_batch(g, S, orig, size, ref, stride_r, S.me_off, K, width, height, wide);
      unsigned bsad = ~0u;
      int bk = 1 << 30;
      Mv bmv = mkmv(0, 0);
      for (int k = lane, n = 0; k < K; k += ME_LANES, n++) {
        Mv cand = mine[n];
        unsigned sad;
        if (wide) {
          int x;
          sad = widesad_pick(S.acc + 5 * k, &x);
          cand.x = (int16_t)(cand.x + (s * x << 2));
        } else sad = (unsigned)S.acc[k];
        sad += mv_cost(lambda, cand, mvp);
        if (sad < bsad) { bsad = sad; bk = k; bmv = cand; }
      }
      me_argmin(bsad, bk, bmv);
      if (bsad < min_sad) { min_sad = bsad; mv_opt = bmv; }
      mv_ref = mv_opt;
    }
  }

  // candidate search (:876-893)
  if (mvcand_num) {
    const int wide = size == 16;
    for (int k = lane, n = 0; k < mvcand_num; k += ME_LANES, n++) {
      const Mv cand = clip_mv(mkmv(mvcand[k].x << 2, mvcand[k].y << 2), ypos, xpos, fw, fh, size, sign);
      mine[n] = cand;
      S.me_off[k] = s * (cand.x >> 2) + s * (cand.y >> 2) * stride_r;
    }
    sad_batch(g, S, orig, size, ref, stride_r, S.me_off, mvcand_num, width, height, wide);
    unsigned bsad = ~0u;
    int bk = 1 << 30;
    Mv bmv = mkmv(0, 0);
    for (int k = lane, n = 0; k < mvcand_num; k += ME_LANES, n++) {
      Mv cand = mine[n];
      int x = 0;
      unsigned sad;
      if (wide) sad = widesad_pick(S.acc + 5 * k, &x);
      else sad = (unsigned)S.acc[k];
      cand.x = (int16_t)(cand.x + (s * x << 2));
      sad += mv_cost(lambda, cand, mvp);
      if (sad < bsad) { bsad = sad; bk = k; bmv = cand; }
    }
    me_argmin(bsad, bk, bmv);
    if (bsad < min_sad) { min_sad = bsad; mv_opt = bmv; }
  }
  mv_ref = mv_opt;

  // hexagon-like full-pel refinement (:895-929)
  {
    const int maxsteps = (size <= 16 || speed == 0) ? 6 : 0;
    int start = 0, end = 5;
    for (int step = 1; step < maxsteps; step++) {
      const int nd = (end - start + 6) % 6 + 1;            // the directions visited this round: start, start + 1, .. end (mod 6)
      for (int k = lane, n = 0; k < nd; k += ME_LANES, n++) {
        const int d = (start + k) % 6;
        const int diy = d == 0 ? 1 : d == 1 ? 2 : d == 2 ? 1 : d == 3 ? -1 : d == 4 ? -2 : -1;
        const int dix = d == 0 ? -1 : d == 1 ? 0 : d == 2 ? 1 : d == 3 ? 1 : d == 4 ? 0 : -1;
        const Mv cand = clip_mv(mkmv(mv_ref.x + diy * 4, mv_ref.y + dix * 4), ypos, xpos, fw, fh, size, sign);
        mine[n] = cand;
        S.me_off[k] = s * (cand.x >> 2) + s * (cand.y >> 2) * stride_r;
      }
      sad_batch(g, S, orig, size, ref, stride_r, S.me_off, nd, width, height, 0);
      unsigned bsad = ~0u;
      int bk = 1 << 30;
      Mv bmv = mkmv(0, 0);
      for (int k = lane, n = 0; k < nd; k += ME_LANES, n++) {
        const Mv cand = mine[n];
        const unsigned sad = (unsigned)S.acc[k] + mv_cost(lambda, cand, mvp);
        if (sad < bsad) { bsad = sad; bk = k; bmv = cand; }
      }
      me_argmin(bsad, bk, bmv);
      int best_dir = -1;
      if (bsad < min_sad) { min_sad = bsad; mv_opt = bmv; best_dir = (start + bk) % 6; }
      mv_ref = mv_opt;
      start = best_dir ? best_dir - 1 : 5;
      end = start + 2;
      end -= (end >= 6) * 6;
      if (best_dir < 0) break;
    }
  }

  int ydelta_hp = 0, xdelta_hp = 0, ydelta_qp = 0, xdelta_qp = 0;
  unsigned cmin = min_sad, sad;
  Mv cand;
  if (speed == 0) {
    // true interpolation at the 8 half-pel and then the 8 quarter-pel positions (:938-976)
    for (int pass = 0; pass < 2; pass++) {
      const int u = pass ? 1 : 2;
      for (int i = 1; i <= 8; i++) {
        const int m = i == 2 || i == 5 || i == 6 ? -u : i == 3 || i == 7 || i == 8 ? u : 0;        // hmpos / qmpos
        const int n = i == 1 || i == 5 || i == 7 ? -u : i == 4 || i == 6 || i == 8 ? u : 0;        // hnpos / qnpos
        cand = mkmv((pass ? mv_opt.x : mv_ref.x) + n, (pass ? mv_opt.y : mv_ref.y) + m);
        inter_luma(g, S, S.ipred, width, ref, stride_r, width, height, cand, sign, enable_bipred);
        int acc = 0;
        TFOR(g, j, width * height) acc += iabs((int)orig[p2div(j, width) * size + p2mod(j, width)] - (int)S.ipred[j]);
        sad = (unsigned)g.sum(acc);
        if (pass) sad += (unsigned)(int)dadd(dmul(lambda, (double)mv_bits(cand.y - mvp.y, cand.x - mvp.x)), 0.5);
        else sad += mv_cost(lambda, cand, mvp);
        if (sad < cmin) {
          cmin = sad;
          if (pass) { ydelta_qp = m; xdelta_qp = n; } else { ydelta_hp = m; xdelta_hp = n; }
        }
      }
      if (!pass) { mv_opt.x = (int16_t)(mv_opt.x + xdelta_hp); mv_opt.y = (int16_t)(mv_opt.y + ydelta_hp); }
    }
  } else {
    // bilinear approximations (:977-1008)
    mv_ref.x = (int16_t)(mv_ref.x * s);
    mv_ref.y = (int16_t)(mv_ref.y * s);
    int spx, spy;
    sad = sad_fasthalf(g, orig, size, ref + (mv_ref.x >> 2) + (mv_ref.y >> 2) * stride_r, stride_r, width, height, &spx, &spy);
    sad += (unsigned)dadd(dmul(lambda, (double)mv_bits(mv_ref.y + s * spy - mvp.y, mv_ref.x + s * spx - mvp.x)), 0.5);
    if (sad < cmin) { cmin = sad; xdelta_hp = s * spx; ydelta_hp = s * spy; }
    spx = xdelta_hp;
    spy = ydelta_hp;
    mv_ref.x = (int16_t)(mv_opt.x + s * spx);
    mv_ref.y = (int16_t)(mv_opt.y + s * spy);
    mv_opt.x = (int16_t)(mv_opt.x + xdelta_hp);
    mv_opt.y = (int16_t)(mv_opt.y + ydelta_hp);
    sad = sad_fastquarter(g, orig, size, ref + s * (mv_ref.x >> 2) + s * (mv_ref.y >> 2) * stride_r, stride_r, width, height, &spx, &spy);
    sad += (unsigned)(int)dadd(dmul(lambda, (double)mv_bits(mv_ref.y + s * spy - mvp.y, mv_ref.x + s * spx - mvp.x)), 0.5);
    if (sad < cmin) { cmin = sad; xdelta_qp = s * spx; ydelta_qp = s * spy; }
  }
  mv_opt.x = (int16_t)(mv_opt.x + xdelta_qp);
  mv_opt.y = (int16_t)(mv_opt.y + ydelta_qp);
  *mv_out = mv_opt;
  return cmin < min_sad ? cmin : min_sad;
}

// motion_estimate_bi (enc/encode_block.c:1102-1215): one vector used with opposite signs in two references (B frames at
// encoder_speed 0).  3x3 telescope from 8 samples down to a quarter sample around mvc with true interpolation of both
// references, then six extra candidates taken from -- and, as in the reference, WRITTEN INTO -- the candidate list: entries
// num..3 are zeroed, entry 4 becomes mvp, entry 5 zero (the list's length is left alone; later searches see the changes).
TDN unsigned motion_estimate_bi(Grp &g, SbMem &S, const thor_enc_frame &F, const uint8_t *orig, const uint8_t *ref0, const uint8_t *ref1, int size,
                                int stride_r, Mv *mv_out, Mv mvc, Mv mvp, double lambda, int sign, int xpos, int ypos, Mv *mvcand, int mvcand_num,
                                int enable_bipred) {
  THOR_SHS(&S);
  const int fw = F.width, fh = F.height, n = size * size;
  uint32_t min_sad = 1u << 31;
  Mv mv_opt = mkmv(0, 0), mv_ref = mkmv(((mvc.x + 2) >> 2) << 2, ((mvc.y + 2) >> 2) << 2);
  uint8_t *rf0 = S.p0.y, *rf1 = S.p1.y;
  int pass = 0;
  for (int step = 32;; ) {
    // pass 0: the telescope at this step; pass 1 (after step 0): the extra candidates
    const int ncand = pass ? 6 : 9;
    for (int c = 0; c < ncand; c++) {
      Mv cand;
      if (!pass) {
        const int k = (c / 3 - 1) * step, l = (c % 3 - 1) * step;
        if (step < 32 && k == 0 && l == 0) continue;
        if (step == 1) {
          const int vf = mv_ref.y & 3, hf = mv_ref.x & 3;
          bool skip;
          if (vf == 0 && hf == 0) skip = iabs(k) != iabs(l);
          else if (vf == 2 && hf == 2) skip = true;
          else skip = iabs(k) == iabs(l);
          if (skip) continue;
        }
        cand = mkmv(mv_ref.x + l, mv_ref.y + k);
      } else {
        cand = mvcand[c];
      }
      cand = clip_mv(cand, ypos, xpos, fw, fh, size, sign);
      inter_luma(g, S, rf0, size, ref0, stride_r, size, size, cand, sign, enable_bipred);
      cand = clip_mv(cand, ypos, xpos, fw, fh, size, 1 - sign);
      inter_luma(g, S, rf1, size, ref1, stride_r, size, size, cand, 1 - sign, enable_bipred);
      int acc = 0;
      TFOR(g, i, n / 4) acc += (int)sad4(reinterpret_cast<const uint32_t *>(orig)[i], havg4(reinterpret_cast<const uint32_t *>(rf0)[i], reinterpret_cast<const uint32_t *>(rf1)[i]));
      int sad = g.sum(acc);
      sad += (int)dadd(dmul(lambda, (double)mv_bits((int16_t)(cand.y - mvp.y), (int16_t)(cand.x - mvp.x))), 0.5);
      if ((uint32_t)sad < min_sad) { min_sad = (uint32_t)sad; mv_opt = cand; }
    }
    if (pass) break;
    mv_ref = mv_opt;
    step >>= 1;
    if (step == 0) {
      g.sync();
      if (g.tid == 0) {
        for (int idx = mvcand_num; idx < 4; idx++) mvcand[idx] = mkmv(0, 0);
        mvcand[4] = mvp;
        mvcand[5] = mkmv(0, 0);
      }
      g.sync();
      pass = 1;
    }
  }
  *mv_out = mv_opt;
  return min_sad;
}

}  // namespace thorenc
