// Frame encoder of libthorcuda -- residual, integer transforms, quantiser, reconstruction, distortion.
#pragma once
#include "enc_state.h"

namespace thorenc {

// sum of m[c] * x[c], c < n (n a multiple of 4; m and x 4-byte aligned): two dp2a per four terms instead of eight loads and
// four multiply-adds.  Exact: |x| < 2^15, |m| < 2^7, n <= 32.
TD int dot_i8_i16(const int8_t *m, const int16_t *x, int n) {
  int acc = 0;
#ifdef THOR_HOSTEMU
  for (int c = 0; c < n; c++) acc += m[c] * x[c];
#else
  const int *m4 = reinterpret_cast<const int *>(m), *x2 = reinterpret_cast<const int *>(x);
#pragma unroll 1
  for (int c = 0; c < (n >> 2); c++) {        // not unrolled: code size (instruction fetch) beats the few branches, measured
    const int mm = m4[c];
    acc = __dp2a_lo(x2[2 * c], mm, acc);
    acc = __dp2a_hi(x2[2 * c + 1], mm, acc);
  }
#endif
  return acc;
}

// the two matrix passes of transform() (common/transform.c:309-329) on S.resid (n x n, stride n): only the low
// keep x keep outputs are produced (keep = min(n, 16)), each pass rounds, shifts and stores int16.  Output: S.coeff,
// stride keep.  Ends with a barrier.
TDN void forward_core(Grp &g, SbMem &S, int n, int sh1, int sh2) {
  THOR_LOCAL(&g);
  THOR_SHS(&S);
  const int keep = imin(n, 16);
  const int16_t *in = S.resid;
  const int8_t *M = dct_matrix(S, n);
  TFOR(g, i, keep * n) {                                   // rows of the block -> horizontal frequency k
    const int r = p2div(i, keep), k = p2mod(i, keep);
    const int acc = dot_i8_i16(M + k * n, in + r * n, n);
    S.mid[k * 32 + r] = (int16_t)((acc + (1 << (sh1 - 1))) >> sh1);
  }
  g.sync();
  TFOR(g, i, keep * keep) {                                // vertical frequency v
    const int v = p2div(i, keep), k = p2mod(i, keep);
    const int acc = dot_i8_i16(M + v * n, S.mid + k * 32, n);
    S.coeff[v * keep + k] = (int16_t)((acc + (1 << (sh2 - 1))) >> sh2);
  }
  g.sync();
}

// transform() (common/transform.c:249-330) of org - pred for one size x size transform block.  The transform input is
// formed directly: the residual itself, its f x f box sums for `fast` blocks larger than 16 (:273-293), or its 2x2 sums
// for 64x64 (:295-307).  Output: S.coeff, the low q x q corner (q = min(size, 16)), stride q.  Ends with a barrier.
TDN void forward_tb(Grp &g, SbMem &S, const uint8_t *org, int ostride, const uint8_t *pred, int pstride, int size, int fast) {
  THOR_LOCAL(&g);
  THOR_SHS(&S); THOR_SHARED(org); THOR_SHARED(pred);
  int n = size, sh1 = ilog2((uint32_t)size), sh2 = sh1 + 5, f = 1;
  if (size > 16 && fast) { n = 16; f = size / 16; sh1 = size == 64 ? 8 : 6; sh2 = 9; }
  else if (size == 64) { n = 32; f = 2; sh1 = 7; sh2 = 10; }
  TFOR(g, i, n * n) {
    const int r = p2div(i, n), c = p2mod(i, n);
    int s = 0;
    for (int a = 0; a < f; a++)
      for (int b = 0; b < f; b++) s += (int)org[(r * f + a) * ostride + c * f + b] - (int)pred[(r * f + a) * pstride + c * f + b];
    S.resid[i] = (int16_t)s;
  }
  g.sync();
  forward_core(g, S, n, sh1, sh2);
}

// quantize() (enc/encode_block.c:75-178, rdoq == 0) of S.coeff; levels go to coeffq (stride `size`, low q x q corner).
// Returns cbp (uniform).  Ends with a barrier.
TDN int quantize_tb(Grp &g, SbMem &S, int16_t *coeffq, int qp, int size, int coeff_block_type) {
  THOR_LOCAL(&g);
  THOR_SHS(&S); THOR_SHARED(coeffq);
  const int intra = (coeff_block_type >> 1) & 1, chroma = coeff_block_type & 1;
  const int lg = ilog2((uint32_t)size), q = imin(size, 16), nq = q * q;
  const int scale = kQuant[qp % 6], shift = 21 - lg + qp / 6;
  const uint8_t *zz = zigzag(S, q);
  const int tail_off = (intra ? 38 : -26) * (1 << (shift - 8));
  const int off_small = (intra ? 102 : 51) << (shift - 8), off_big = (intra ? 115 : 90) << (shift - 8);
  int mylast = -1;
  TFOR(g, i, nq) {
    const int c = S.coeff[i], p = zz[i];
    S.lin[p] = (int16_t)c;
    if ((iabs(iabs(c) * scale + tail_off) >> shift) != 0) mylast = imax(mylast, p);
  }
  const int last = g.vmax(mylast);                         // last_pos (:95-104); barrier inside
  int mine = 0;
  TFOR(g, p, nq) {
    int l = 0;
    if (p <= last) {
      const int c = S.lin[p], mag = scale * iabs(c);
      const int off = ((mag >> shift) == 0 || chroma) ? off_small : off_big;
      l = (mag + off) >> shift;
      mine |= l != 0;
      if (c < 0) l = -l;
    }
    S.lev[p] = (int16_t)l;
  }
  if (g.tid < 8) S.fire[g.tid] = 0;
#ifdef THOR_HOSTEMU
  for (int k = 0; k < 8; k++) S.fire[k] = 0;
#endif
  const int cbp = g.any(mine);                             // barrier inside
  if (cbp) {
    // "RDOQ light" (:135-168): a level > 1 right behind two zeros is traded for a +-1.  Whether position p can fire at
    // all depends only on the ORIGINAL levels at p-2, p-1, p (the pass only ever turns zeros and large levels into +-1,
    // which cannot create a new zero-zero-large pattern); those positions are collected in parallel and walked in order.
    const int N = chroma ? last + 1 : nq;
    TFOR(g, p, N) {
      if (p >= 2 && S.lev[p - 2] == 0 && S.lev[p - 1] == 0 && iabs(S.lev[p]) > 1) {
#ifdef THOR_HOSTEMU
        S.fire[p >> 5] |= 1u << (p & 31);
#else
        atomicOr(&S.fire[p >> 5], 1u << (p & 31));
#endif
      }
    }
    g.sync();
    if (g.tid == 0) {
      const int thr = (73 * kDequant[qp % 6] << (qp / 6)) >> (4 + lg);
      for (int w = 0; w < 8; w++) {
        uint32_t m = S.fire[w];
        while (m) {
          const int b = ilog2(m & (0u - m));
          m &= m - 1;
          const int p = 32 * w + b;
          int ok = 1;
          if (p > 2 && iabs(S.lev[p - 3]) > 1) ok = 0;
          if (p > 3 && iabs(S.lev[p - 4]) > 1 && iabs(S.lev[p - 3]) > 0) ok = 0;
          if (p == 2 && (!chroma || last >= 6)) ok = 0;
          if (!(ok && S.lev[p - 2] == 0 && S.lev[p - 1] == 0 && iabs(S.lev[p]) > 1)) continue;
          const int k1 = iabs(S.lin[p]), k2 = iabs(S.lin[p - 1]), k3 = iabs(S.lin[p - 2]);
          if (k1 + imax(k2, k3) < thr) S.lev[p] = S.lin[p] < 0 ? -1 : 1;
          else if (k2 > k3) S.lev[p - 1] = S.lin[p - 1] < 0 ? -1 : 1;
          else S.lev[p - 2] = S.lin[p - 2] < 0 ? -1 : 1;
        }
      }
    }
    g.sync();
  }
  TFOR(g, i, nq) coeffq[(p2div(i, q)) * size + (p2mod(i, q))] = (int16_t)S.lev[zz[i]];
  g.sync();
  return cbp;
}

// dequantize (common/common_block.c:132-146) + inverse_transform (common/transform.c:432-518) + reconstruct_block
// (common/common_block.c:148-156).  coeffq: stride `size`; rec: stride rstride.  Ends with a barrier.
TDN void reconstruct_tb(Grp &g, SbMem &S, const int16_t *coeffq, const uint8_t *pred, int pstride, uint8_t *rec, int rstride, int qp, int size) {
  THOR_LOCAL(&g);
  THOR_SHS(&S); THOR_SHARED(coeffq); THOR_SHARED(pred); THOR_SHARED(rec);
  const int lg = ilog2((uint32_t)size), q = imin(size, 16), n = size == 64 ? 32 : size, keep = imin(n, 16);
  const int lshift = qp / 6, rshift = lg - 1, scale = kDequant[qp % 6], add = 1 << (rshift - 1);
  TFOR(g, i, q * q) {
    const int c = coeffq[(p2div(i, q)) * size + (p2mod(i, q))];
    S.rcoeff[i] = (int16_t)((((c * scale) << lshift) + add) >> rshift);
  }
  g.sync();
  const int8_t *M = dct_matrix(S, n);
  TFOR(g, i, keep * n) {                                   // column k of the coefficients -> spatial column
    const int r = p2div(i, keep), k = p2mod(i, keep);
    int acc = 0;
    for (int v = 0; v < keep; v++) acc += M[v * n + r] * S.rcoeff[v * q + k];
    S.mid[k * 32 + r] = (int16_t)iclip((acc + 64) >> 7, -32768, 32767);
  }
  g.sync();
  TFOR(g, i, n * n) {
    const int r = p2div(i, n), c = p2mod(i, n);
    int acc = 0;
    for (int k = 0; k < keep; k++) acc += M[k * n + c] * S.mid[k * 32 + r];
    S.rblock[i] = (int16_t)iclip((acc + 2048) >> 12, -32768, 32767);
  }
  g.sync();
  if (size == 64) {                                        // 2x nearest-neighbour up-sampling (:507-514)
    TFOR(g, i, 4096) {
      const int r = i >> 6, c = i & 63;
      rec[r * rstride + c] = (uint8_t)clip255(S.rblock[(r >> 1) * 32 + (c >> 1)] + pred[r * pstride + c]);
    }
  } else {
    TFOR(g, i, n * n) {
      const int r = p2div(i, n), c = p2mod(i, n);
      rec[r * rstride + c] = (uint8_t)clip255(S.rblock[i] + pred[r * pstride + c]);
    }
  }
  g.sync();
}

// one transform block of encode_and_reconstruct_block_inter / _intra (enc/encode_block.c:1398-1532): residual ->
// transform -> quantise -> (dequantise -> inverse transform -> reconstruct | copy the prediction).  Returns the cbp bit.
TDN int encode_tb(Grp &g, SbMem &S, const uint8_t *org, int ostride, const uint8_t *pred, int pstride, uint8_t *rec, int rstride,
                  int16_t *coeffq, int size, int qp, int coeff_type, int fast) {
  PROF(PF_TB);
  THOR_LOCAL(&g);
  THOR_SHS(&S); THOR_SHARED(org); THOR_SHARED(pred); THOR_SHARED(rec); THOR_SHARED(coeffq);
  forward_tb(g, S, org, ostride, pred, pstride, size, fast);
  const int cbp = quantize_tb(g, S, coeffq, qp, size, coeff_type);
  if (cbp) reconstruct_tb(g, S, coeffq, pred, pstride, rec, rstride, qp, size);
  else {
    TFOR(g, i, size * size) {
      const int r = p2div(i, size), c = p2mod(i, size);
      rec[r * rstride + c] = pred[r * pstride + c];
    }
    g.sync();
  }
  return cbp;
}

// ---- up to three transform blocks at once ------------------------------------------------------------------------
// The Y, U and V transform blocks of a coding block go through the same chain of short phases; done one after the other
// that is 3 x 14 barriers for at most a few hundred samples each.  encode_tb3 runs every phase over the samples of all
// (<= 3) blocks together -- same arithmetic per block as forward_tb / quantize_tb / reconstruct_tb above, slot t of the
// transform scratch per block -- so a trial encode costs one pass through the chain.  Returns the cbp bits (bit t).
struct TbDesc {
  const uint8_t *org; int ostride;
  const uint8_t *pred; int pstride;
  uint8_t *rec; int rstride;
  int16_t *coeffq;                    // stride `size`
  int size, qp, ctype, fast;
};
TD int tb_of(int i, int c0, int c1, int &j) {
  if (i < c0) { j = i; return 0; }
  if (i < c0 + c1) { j = i - c0; return 1; }
  j = i - c0 - c1;
  return 2;
}

// Per-block values are picked with two selects (T3) from arrays that are only indexed by constants, so they live in
// registers: indexing them -- or the descriptors -- by the run-time block number meant a local-memory load per use.
#define T3(a) (t == 0 ? a[0] : t == 1 ? a[1] : a[2])
TDN int encode_tb3(Grp &g, SbMem &S, const TbDesc *tb, int ntb) {
  PROF(PF_TB);
  THOR_SHS(&S);
  const uint8_t *t_org[3], *t_pred[3];
  uint8_t *t_rec[3];
  int16_t *t_cq[3];
  int t_os[3], t_ps[3], t_rs[3], t_size[3] = {0, 0, 0}, t_qp[3], t_ct[3];
#pragma unroll
  for (int t = 0; t < 3; t++) {
    const TbDesc &d = tb[t < ntb ? t : 0];
    t_org[t] = d.org; t_pred[t] = d.pred; t_rec[t] = d.rec; t_cq[t] = d.coeffq;
    t_os[t] = d.ostride; t_ps[t] = d.pstride; t_rs[t] = d.rstride; t_qp[t] = d.qp; t_ct[t] = d.ctype;
    if (t < ntb) t_size[t] = d.size;
  }
  // geometry per block: forward transform size n (after pooling by f), shifts, kept corner q, inverse size ni
  int n[3] = {0, 0, 0}, f[3] = {1, 1, 1}, sh1[3] = {0, 0, 0}, sh2[3] = {0, 0, 0}, q[3] = {0, 0, 0}, ni[3] = {0, 0, 0}, lg[3] = {0, 0, 0};
#pragma unroll
  for (int t = 0; t < 3; t++) {
    if (t >= ntb) continue;
    const int size = tb[t].size;
    lg[t] = ilog2((uint32_t)size);
    n[t] = size; sh1[t] = lg[t]; sh2[t] = lg[t] + 5;
    if (size > 16 && tb[t].fast) { n[t] = 16; f[t] = size / 16; sh1[t] = size == 64 ? 8 : 6; sh2[t] = 9; }
    else if (size == 64) { n[t] = 32; f[t] = 2; sh1[t] = 7; sh2[t] = 10; }
    q[t] = imin(size, 16);
    ni[t] = size == 64 ? 32 : size;
  }
  int c0, c1, tot, j;
  // 1. transform input
  c0 = n[0] * n[0]; c1 = n[1] * n[1]; tot = c0 + c1 + n[2] * n[2];
  TFOR(g, i, tot) {
    const int t = tb_of(i, c0, c1, j), nn = T3(n), ff = T3(f), r = p2div(j, nn), c = p2mod(j, nn);
    const uint8_t *o = T3(t_org) + (r * ff) * T3(t_os) + c * ff, *p = T3(t_pred) + (r * ff) * T3(t_ps) + c * ff;
    int sum = 0;
    for (int a = 0; a < ff; a++)
      for (int b = 0; b < ff; b++) sum += (int)o[a * T3(t_os) + b] - (int)p[a * T3(t_ps) + b];
    S.resid[t * 1024 + j] = (int16_t)sum;
  }
  g.sync();
  // 2. rows -> horizontal frequencies
  c0 = q[0] * n[0]; c1 = q[1] * n[1]; tot = c0 + c1 + q[2] * n[2];
  TFOR(g, i, tot) {
    const int t = tb_of(i, c0, c1, j), nn = T3(n), keep = T3(q), r = p2div(j, keep), k = p2mod(j, keep);
    const int8_t *M = dct_matrix(S, nn) + k * nn;
    const int16_t *in = S.resid + t * 1024 + r * nn;
    const int acc = dot_i8_i16(M, in, nn);
    S.mid[t * 512 + k * 32 + r] = (int16_t)((acc + (1 << (T3(sh1) - 1))) >> T3(sh1));
  }
  g.sync();
  // 3. vertical frequencies; 4. scan order + last significant position
  c0 = q[0] * q[0]; c1 = q[1] * q[1]; tot = c0 + c1 + q[2] * q[2];
  TFOR(g, i, tot) {
    const int t = tb_of(i, c0, c1, j), nn = T3(n), keep = T3(q), v = p2div(j, keep), k = p2mod(j, keep);
    const int8_t *M = dct_matrix(S, nn) + v * nn;
    const int16_t *m = S.mid + t * 512 + k * 32;
    const int acc = dot_i8_i16(M, m, nn);
    S.coeff[t * 256 + j] = (int16_t)((acc + (1 << (T3(sh2) - 1))) >> T3(sh2));
  }
  g.sync();
  int last[3] = {-1, -1, -1};
  TFOR(g, i, tot) {
    const int t = tb_of(i, c0, c1, j);
    const int intra = (T3(t_ct) >> 1) & 1, shift = 21 - T3(lg) + T3(t_qp) / 6, scale = kQuant[T3(t_qp) % 6];
    const int tail_off = (intra ? 38 : -26) * (1 << (shift - 8));
    const int c = S.coeff[t * 256 + j], p = zigzag(S, T3(q))[j];
    S.lin[t * 256 + p] = (int16_t)c;
    if ((iabs(iabs(c) * scale + tail_off) >> shift) != 0) { if (t == 0) last[0] = imax(last[0], p); else if (t == 1) last[1] = imax(last[1], p); else last[2] = imax(last[2], p); }
  }
  g.vmax3(last);                                                       // barrier inside
  // 5. levels with the dead-zone offsets (:106-131)
  int mine = 0;
  TFOR(g, i, tot) {
    const int t = tb_of(i, c0, c1, j);
    int l = 0;
    if (j <= T3(last)) {
      const int intra = (T3(t_ct) >> 1) & 1, chroma = T3(t_ct) & 1, shift = 21 - T3(lg) + T3(t_qp) / 6;
      const int c = S.lin[t * 256 + j], mag = kQuant[T3(t_qp) % 6] * iabs(c);
      const int off = (((mag >> shift) == 0 || chroma) ? (intra ? 102 : 51) : (intra ? 115 : 90)) << (shift - 8);
      l = (mag + off) >> shift;
      if (l) mine |= 1 << t;
      if (c < 0) l = -l;
    }
    S.lev[t * 256 + j] = (int16_t)l;
  }
  TFOR(g, i, 24) S.fire[i] = 0;
  const int cbp = g.orall(mine);                                       // barrier inside
  // 6. "RDOQ light" (:135-168), see quantize_tb: candidates found in parallel, walked in order by one thread per block
  if (cbp) {
    TFOR(g, i, tot) {
      const int t = tb_of(i, c0, c1, j);
      const int N = (T3(t_ct) & 1) ? T3(last) + 1 : T3(q) * T3(q);
      const int16_t *lev = S.lev + t * 256;
      if (((cbp >> t) & 1) && j >= 2 && j < N && lev[j - 2] == 0 && lev[j - 1] == 0 && iabs(lev[j]) > 1) {
#ifdef THOR_HOSTEMU
        S.fire[t * 8 + (j >> 5)] |= 1u << (j & 31);
#else
        atomicOr(&S.fire[t * 8 + (j >> 5)], 1u << (j & 31));
#endif
      }
    }
    g.sync();
    if ((g.tid & 31) == 0)
      for (int t = g.tid >> 5; t < ntb; t += (g.n + 31) >> 5) {
        if (!((cbp >> t) & 1)) continue;
        const int chroma = T3(t_ct) & 1, qp = T3(t_qp);
        const int thr = (73 * kDequant[qp % 6] << (qp / 6)) >> (4 + T3(lg));
        int16_t *lev = S.lev + t * 256;
        const int16_t *lin = S.lin + t * 256;
        for (int w = 0; w < 8; w++) {
          uint32_t m = S.fire[t * 8 + w];
          while (m) {
            const int b = ilog2(m & (0u - m));
            m &= m - 1;
            const int p = 32 * w + b;
            int ok = 1;
            if (p > 2 && iabs(lev[p - 3]) > 1) ok = 0;
            if (p > 3 && iabs(lev[p - 4]) > 1 && iabs(lev[p - 3]) > 0) ok = 0;
            if (p == 2 && (!chroma || T3(last) >= 6)) ok = 0;
            if (!(ok && lev[p - 2] == 0 && lev[p - 1] == 0 && iabs(lev[p]) > 1)) continue;
            const int k1 = iabs(lin[p]), k2 = iabs(lin[p - 1]), k3 = iabs(lin[p - 2]);
            if (k1 + imax(k2, k3) < thr) lev[p] = lin[p] < 0 ? -1 : 1;
            else if (k2 > k3) lev[p - 1] = lin[p - 1] < 0 ? -1 : 1;
            else lev[p - 2] = lin[p - 2] < 0 ? -1 : 1;
          }
        }
      }
    g.sync();
  }
  // 7. levels back to raster order (the block's coefficient array) and dequantised (common/common_block.c:132-146)
  TFOR(g, i, tot) {
    const int t = tb_of(i, c0, c1, j), qq = T3(q), qp = T3(t_qp), rshift = T3(lg) - 1;
    const int l = (int16_t)S.lev[t * 256 + zigzag(S, qq)[j]];
    T3(t_cq)[(p2div(j, qq)) * T3(t_size) + (p2mod(j, qq))] = (int16_t)l;
    S.rcoeff[t * 256 + j] = (int16_t)((((l * kDequant[qp % 6]) << (qp / 6)) + (1 << (rshift - 1))) >> rshift);
  }
  g.sync();
  if (cbp) {
    // 8. inverse transform (common/transform.c:432-518), blocks with levels only
    int k0 = ((cbp >> 0) & 1) ? q[0] * ni[0] : 0, k1 = ((cbp >> 1) & 1) ? q[1] * ni[1] : 0, k2 = ((cbp >> 2) & 1) ? q[2] * ni[2] : 0;
    TFOR(g, i, k0 + k1 + k2) {
      const int t = tb_of(i, k0, k1, j), nn = T3(ni), keep = T3(q), r = p2div(j, keep), k = p2mod(j, keep);
      const int8_t *M = dct_matrix(S, nn) + r;
      const int16_t *rc = S.rcoeff + t * 256 + k;
      int acc = 0;
      for (int v = 0; v < keep; v++) acc += M[v * nn] * rc[v * keep];
      S.mid[t * 512 + k * 32 + r] = (int16_t)iclip((acc + 64) >> 7, -32768, 32767);
    }
    g.sync();
    k0 = ((cbp >> 0) & 1) ? ni[0] * ni[0] : 0; k1 = ((cbp >> 1) & 1) ? ni[1] * ni[1] : 0; k2 = ((cbp >> 2) & 1) ? ni[2] * ni[2] : 0;
    TFOR(g, i, k0 + k1 + k2) {
      const int t = tb_of(i, k0, k1, j), nn = T3(ni), keep = T3(q), r = p2div(j, nn), c = p2mod(j, nn);
      const int8_t *M = dct_matrix(S, nn) + c;
      const int16_t *m = S.mid + t * 512 + r;
      int acc = 0;
      for (int k = 0; k < keep; k++) acc += M[k * nn] * m[k * 32];
      S.rblock[t * 1024 + j] = (int16_t)iclip((acc + 2048) >> 12, -32768, 32767);
    }
    g.sync();
  }
  // 9. reconstruction (common/common_block.c:148-156): prediction + residual where the block has levels, else the prediction
  c0 = tb[0].size * tb[0].size; c1 = ntb > 1 ? tb[1].size * tb[1].size : 0; tot = c0 + c1 + (ntb > 2 ? tb[2].size * tb[2].size : 0);
  TFOR(g, i, tot) {
    const int t = tb_of(i, c0, c1, j), size = T3(t_size), r = p2div(j, size), c = p2mod(j, size);
    int v = T3(t_pred)[r * T3(t_ps) + c];
    if ((cbp >> t) & 1) v = clip255(v + (size == 64 ? S.rblock[t * 1024 + (r >> 1) * 32 + (c >> 1)] : S.rblock[t * 1024 + j]));
    T3(t_rec)[r * T3(t_rs) + c] = (uint8_t)v;
  }
  g.sync();
  return cbp;
}

#undef T3

// ssd_calc (enc/encode_block.c:782-797) over the three planes of a block: cost_calc's distortion (:1218-1228)
TDN int ssd_yuv(Grp &g, const Blk &a, const Blk &b, int stride, int width, int height) {
  PROF(PF_SSD);
  THOR_SHARED(&a); THOR_SHARED(&b);
  THOR_LOCAL(&g);
  int s = 0;
  // four luma samples per item (rows are 4-byte aligned, widths multiples of 8)
  const int wq = width / 4;
  const Div dwq = mkdiv(wq);
  TFOR(g, i, wq * height) {
    const int o = dq(dwq, i) * stride + dr(dwq, i) * 4;
    const uint32_t x = *reinterpret_cast<const uint32_t *>(a.y + o), y = *reinterpret_cast<const uint32_t *>(b.y + o);
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const int d = (int)((x >> (8 * k)) & 255) - (int)((y >> (8 * k)) & 255);
      s += d * d;
    }
  }
  const int wc = width / 2, sc = stride / 2;
  const Div dwc = mkdiv(wc);
  TFOR(g, i, wc * (height / 2)) {
    const int o = dq(dwc, i) * sc + dr(dwc, i);
    const int du = (int)a.u[o] - (int)b.u[o], dv = (int)a.v[o] - (int)b.v[o];
    s += du * du + dv * dv;
  }
  return g.sum(s);
}

// cost_calc (enc/encode_block.c:1218-1228)
TD uint32_t rd_cost(int ssd, int nbits, double lambda) {
  uint32_t cost = (uint32_t)(ssd + (int32_t)dadd(dmul(lambda, (double)nbits), 0.5));
  if (cost > (1u << 30)) cost = 1u << 30;
  return cost;
}

}  // namespace thorenc
