// Frame encoder of libthorcuda -- per-superblock working memory, neighbour state and the bit-cost model.
#pragma once
#include "enc_common.h"

namespace thorenc {

// compact block of samples, stride = block size (the reference's yuv_block_t, enc/mainenc.h:90-95)
struct Blk { uint8_t y[4096]; uint8_t u[1024]; uint8_t v[1024]; };
// quantised coefficients in block_param_t's layout (common/types.h:164-166)
struct Coef { int16_t y[4096]; int16_t u[1024]; int16_t v[1024]; };

// block_param_t without the coefficient arrays
struct Param {
  int mode, intra_mode, skip_idx, pb_part;
  Mv mv0[4], mv1[4];
  int ref_idx0, ref_idx1, dir;
  int cbp_y, cbp_u, cbp_v;
  int tb_param, tb_split;
};

// what write_block reads of block_info_t besides the decision
struct BlockCtx {
  int size, ypos, xpos, bwidth, bheight;
  int ctx_split, ctx_cbp, ctx_index;      // block_context_t
  int num_skip_vec, num_merge_vec;
  InterPred skip_cand[2], merge_cand[2];  // MAX_NUM_SKIP = MAX_NUM_MERGE = 2 (LIMITED_SKIP, common/global.h:79-84)
  Mv mvp;
  int max_num_pb_part, max_num_tb_part, delta_qp;
};

struct SbMem {
  thor_enc_frame frame;              // the frame description: copied from the kernel parameters once per CTA
  Blk org, rec, best, pred, p0, p1;
  Coef tc, bc;                       // coefficients of the trial / of the best trial so far
  alignas(16) int16_t resid[3 * 1024];           // transform input (n <= 32 after pooling), slot t at t * 1024
  // transform scratch: three transform blocks (Y, U, V of a coding block) are in flight at once, slot t at t * stride
  alignas(16) int16_t mid[3 * 16 * 32];          // transform intermediate
  int16_t coeff[3 * 256];            // low q x q corner of the forward transform, stride q
  int16_t rcoeff[3 * 256];           // dequantised, stride q
  int16_t rblock[3 * 1024];          // inverse transform output (n <= 32), stride n
  int16_t lin[3 * 256], lev[3 * 256];// quantiser: coefficients / levels in scan order (levels < 2^15: |c| * 26214 >> 15)
  int16_t itmp[69 * 64];             // interpolation: first-pass intermediate (rows -2 .. h+2)
  alignas(16) int16_t scan[3 * 256]; // bit counter: levels in scan order (three planes at once; read 16 bytes at a time)
  int bits3[4];
  uint8_t ipred[4096];               // search: prediction under test
  uint8_t org8[4096];                // bi-pred search: 2*org - pred (search_bipred_prediction_params)
  uint8_t top[3][132], left[3][132]; // intra neighbours per plane
  uint8_t topleft[4];
  uint8_t tf[132], lf[132];          // smoothed neighbour rows
  alignas(16) int8_t dct[32 * 32 + 16 * 16 + 8 * 8 + 4 * 4];
  uint8_t zz[16 + 64 + 256];
  uint32_t fire[3 * 8];              // quantiser: positions the RDOQ-light pass has to look at
  int acc[64 * 5 + 8];               // SAD accumulators of a batch of candidates
  int me_off[64];                    // the batch's candidate offsets into the reference plane (motion_estimate)
  Mv mvcand[THOR_ENC_MAX_REF][64];   // frame_info.mvcand (enc/mainenc.h:124), reset per superblock
  // decision of a 16x16 block kept while its 8x8 children are tried (top-down split, enc/encode_block.c:2983-2995)
  uint8_t keep_y[256], keep_u[64], keep_v[64];
  int16_t keep_cy[256], keep_cu[64], keep_cv[64];
#ifndef THOR_HOSTEMU
  int red[2 * NW * 8];
#endif
};

// per-superblock scalar state of frame_info_t (enc/encode_frame.c:117-121)
struct SbState {
  int mvcand_num[THOR_ENC_MAX_REF];
  uint64_t mvcand_mask[THOR_ENC_MAX_REF];
  int best_ref;
};

// ---- tables built once per CTA ----------------------------------------------------------------------------------
TD const int8_t *dct_matrix(const SbMem &S, int n) { return S.dct + (n == 32 ? 0 : n == 16 ? 1024 : n == 8 ? 1280 : 1344); }
TD const uint8_t *zigzag(const SbMem &S, int q) { return S.zz + (q == 4 ? 0 : q == 8 ? 16 : 80); }

TDN void build_tables(Grp &g, SbMem &S) {
  THOR_LOCAL(&g);
  THOR_SHS(&S);
  // M_N[k][n] = C[(2n+1) * k * 32/N] with the symmetries of cos(pi a / 64): common/transform.c:41-245 (g1..g4mat_hevc)
  TFOR(g, i, 1360) {
    int N, base;
    if (i < 1024) { N = 32; base = 0; } else if (i < 1280) { N = 16; base = 1024; } else if (i < 1344) { N = 8; base = 1280; } else { N = 4; base = 1344; }
    const int k = (i - base) / N, n = (i - base) % N;
    int a = ((2 * n + 1) * k * (32 / N)) & 127;
    if (a > 64) a = 128 - a;
    S.dct[i] = (int8_t)(a > 32 ? -kCos32[64 - a] : kCos32[a]);
  }
  // classic zig-zag (common/common_block.c:38-73): serial per table, three tables
  TFOR(g, t, 3) {
    const int q = 4 << t, off = t == 0 ? 0 : t == 1 ? 16 : 80;
    int idx = 0;
    for (int s = 0; s <= 2 * q - 2; s++)
      for (int m = 0; m <= s; m++) {
        const int i = (s & 1) ? m : s - m, j = s - i;
        if (i < q && j < q) S.zz[off + i * q + j] = (uint8_t)idx++;
      }
  }
  g.sync();
}

// ---- neighbour availability (common/common_block.c:100-129) -----------------------------------------------------
TD int upright_available(int ypos, int xpos, int size, int width) {
  int a = (ypos > 0) && (xpos + size < width);
  if (size == 32 && (ypos % 64) == 32) a = 0;
  if (size == 16 && ((ypos % 32) == 16 || ((ypos % 64) == 32 && (xpos % 32) == 16))) a = 0;
  if (size == 8 && ((ypos % 16) == 8 || ((ypos % 32) == 16 && (xpos % 16) == 8) || ((ypos % 64) == 32 && (xpos % 32) == 24))) a = 0;
  return a;
}
TD int downleft_available(int ypos, int xpos, int size, int height) {
  int a = (xpos > 0) && (ypos + size < height);
  if (size == 64) a = 0;
  if (size == 32 && (ypos % 64) == 32) a = 0;
  if (size == 16 && ((ypos % 64) == 48 || ((ypos % 64) == 16 && (xpos % 32) == 16))) a = 0;
  if (size == 8 && ((ypos % 64) == 56 || ((ypos % 16) == 8 && (xpos % 16) == 8) || ((ypos % 64) == 24 && (xpos % 32) == 16))) a = 0;
  return a;
}

TD InterPred unit_pred(const thor_enc_unit *units, int idx) {
  const thor_enc_unit u = ld_unit(units + idx);
  InterPred p;
  p.mv0 = mkmv(u.mv0x, u.mv0y);
  p.mv1 = mkmv(u.mv1x, u.mv1y);
  p.ref_idx0 = u.ref_idx0;
  p.ref_idx1 = u.ref_idx1;
  p.dir = u.dir;
  return p;
}
TD InterPred zero_pred() {
  InterPred p;
  p.mv0 = mkmv(0, 0); p.mv1 = mkmv(0, 0);
  p.ref_idx0 = p.ref_idx1 = 0; p.dir = 0;
  return p;
}

// get_mv_pred (common/inter_prediction.c:182-294): median of three neighbour vectors chosen by availability
TDN Mv get_mv_pred(const thor_enc_frame &F, int ypos, int xpos, int size) {
  PROF(PF_NEIGH);
  THOR_SHF(&F);
  const int bs = size / 4, stride = F.width / 4, idx = (ypos / 4) * stride + xpos / 4;
  const int up0 = idx - stride, up1 = idx - stride + (bs - 1) / 2, up2 = idx - stride + bs - 1;
  const int left0 = idx - 1, left1 = idx + stride * ((bs - 1) / 2) - 1, left2 = idx + stride * (bs - 1) - 1;
  const int downleft = idx + stride * bs - 1, upright = idx - stride + bs, upleft = idx - stride - 1;
  const int U = ypos > 0, L = xpos > 0, UR = upright_available(ypos, xpos, size, F.width), DL = downleft_available(ypos, xpos, size, F.height);
  int ia = -1, ib = -1, ic = -1;
  if (U && !UR && !L && !DL) { ia = up0; ib = up1; ic = up2; }
  else if (U && UR && !L && !DL) { ia = up0; ib = up2; ic = upright; }
  else if (!U && !UR && L && !DL) { ia = left0; ib = left1; ic = left2; }
  else if (U && !UR && L && !DL) { ia = upleft; ib = up2; ic = left2; }
  else if (U && UR && L && !DL) { ia = up0; ib = upright; ic = left2; }
  else if (!U && !UR && L && DL) { ia = left0; ib = left2; ic = downleft; }
  else if (U && !UR && L && DL) { ia = up2; ib = left0; ic = downleft; }
  else if (U && UR && L && DL) { ia = up0; ib = upright; ic = left0; }
  Mv a = mkmv(0, 0), b = a, c = a;
  if (ia >= 0) { a = unit_pred(F.units, ia).mv0; b = unit_pred(F.units, ib).mv0; c = unit_pred(F.units, ic).mv0; }
  Mv p;
  p.x = (int16_t)(a.x < b.x ? imin(b.x, imax(a.x, c.x)) : imin(a.x, imax(b.x, c.x)));
  p.y = (int16_t)(a.y < b.y ? imin(b.y, imax(a.y, c.y)) : imin(a.y, imax(b.y, c.y)));
  return p;
}

// get_mv_skip / get_mv_merge (common/inter_prediction.c:296-600 with LIMITED_SKIP: the two functions are the same)
TDN int get_mv_skip(const thor_enc_frame &F, int ypos, int xpos, int size, InterPred *cand) {
  PROF(PF_NEIGH);
  THOR_SHF(&F);
  const int bs = size / 4, stride = F.width / 4, idx = (ypos / 4) * stride + xpos / 4;
  const int up0 = idx - stride, left0 = idx - 1, upright = idx - stride + bs;
  int up2 = idx - stride + bs - 1, left2 = idx + stride * (bs - 1) - 1;
  const int U = ypos > 0, L = xpos > 0, UR = upright_available(ypos, xpos, size, F.width);
  if (ypos + size > F.height) left2 = left0;
  if (xpos + size > F.width) up2 = up0;
  InterPred t0 = L ? unit_pred(F.units, left2) : zero_pred();
  InterPred t1 = UR ? unit_pred(F.units, upright) : U ? unit_pred(F.units, up2) : zero_pred();
  cand[0] = t0;
  const bool dup = mv_eq(t1.mv0, t0.mv0) && t1.ref_idx0 == t0.ref_idx0 && mv_eq(t1.mv1, t0.mv1) && t1.ref_idx1 == t0.ref_idx1 &&
                   (t1.dir == t0.dir || t1.dir == -1);
  if (dup) return 1;
  cand[1] = t1;
  return 2;
}

// find_block_contexts (common/common_block.c:158-178)
TD void find_block_contexts(const thor_enc_frame &F, int ypos, int xpos, int size, BlockCtx &B) {
  THOR_LOCAL(&B);
  if (ypos >= 8 && xpos >= 8 && ypos + size < F.height && xpos + size < F.width && F.use_block_contexts && size <= 64) {
    const int bs = F.width / 4, bindex = (ypos / 4) * bs + xpos / 4;
    const thor_enc_unit up = ld_unit(F.units + bindex - bs), lf = ld_unit(F.units + bindex - 1);
    B.ctx_split = (up.size < size) + (lf.size < size);
    B.ctx_cbp = ((up.cbp & 1) > 0) + ((lf.cbp & 1) > 0);
    B.ctx_index = 3 * B.ctx_split + (up.cbp != 0) + (lf.cbp != 0);
  } else {
    B.ctx_split = B.ctx_cbp = B.ctx_index = -1;
  }
}

// ---- bit-cost model: lengths of the reference's codes -----------------------------------------------------------
// quote_vlc for tables 0..5 (enc/putvlc.c:141-157) and table 10 (:210-213)
TD int vlc_len(int n, uint32_t cn) {
  const uint32_t lim = 6u << n;
  return cn < lim ? 1 + n + (int)(cn >> n) : 7 - n + 2 * ilog2(cn - lim + (1u << n));
}
TD int expgolomb_len(uint32_t cn) { return 1 + 2 * ilog2(cn + 1); }
// quote_mv_bits (enc/encode_block.c:799-814) = write_mv (enc/write_bits.c:50-69)
TD int mv_bits(int dy, int dx) {
  return expgolomb_len((uint32_t)(2 * iabs(dx) - (dx < 0))) + expgolomb_len((uint32_t)(2 * iabs(dy) - (dy < 0)));
}
TD int mvd_len(Mv mv, Mv p) { return mv_bits(mv.y - p.y, mv.x - p.x); }
TD int mv_list_len(const Mv *mv, Mv p, int part) {                          // write_bits.c:421-444, :468-490
  int n = mvd_len(mv[0], p);
  if (part == 1) n += mvd_len(mv[2], mv[0]);
  else if (part == 2) n += mvd_len(mv[1], mv[0]);
  else if (part == 3) n += mvd_len(mv[1], mv[0]) + mvd_len(mv[2], mv[0]) + mvd_len(mv[3], mv[0]);
  return n;
}
TD int idx_len(int nvec, int idx) { return nvec == 4 ? 2 : nvec == 3 ? (idx == 0 ? 1 : 2) : nvec == 2 ? 1 : 0; }   // :507-530
TD int cbp_code(int y, int u, int v, int ctx_cbp) {                          // :382, :561-566
  const int code = (int)((0x47362501u >> (4 * (y + 2 * u + 4 * v))) & 7);     // cbp_table {1,0,5,2,6,3,7,4}
  return ctx_cbp == 0 && code < 2 ? 1 - code : code;
}
// find_code (enc/write_bits.c:71-108) for a run-mode event
TD int find_code(int run, int level, int maxrun, bool chroma) {
  const int m = imax(maxrun, 4);
  const int index = run + (level > 1 ? m + 1 : 0);
  if (!chroma && index < 2) return index;
  if (index <= 4) return index + 1;
  if (index <= m) return index + 3;
  if (index == m + 1) return 6;
  if (index == m + 2) return 7;
  return index + 1;
}

// write_super_mode (enc/write_bits.c:268-362) with split_flag 0, as a length
TD int super_mode_len(const thor_enc_frame &F, const BlockCtx &B, int mode, int ref_idx0) {
  THOR_LOCAL(&B);
  if (F.frame_type == 0) return B.size > 8;
  const int bip = F.num_ref > 1 && F.enable_bipred, spl = B.size > 8, maxbit = 2 + F.num_ref + spl + bip;
  const bool demote = (B.ctx_index == 2 || B.ctx_index > 3) && spl;
  int code;
  if (F.interp_ref) {
    code = mode == 0 ? 0 : mode == 4 ? 2 : mode == 3 ? 3 : mode == 1 ? 4 : ref_idx0 > 0 ? 4 + ref_idx0 : 4 + F.num_ref;
    if (!bip && code > 3) code--;
    if (!spl && code > 1) code--;
    if (demote && code < 3) code = (code + 2) % 3;
  } else {
    code = mode == 0 ? 0 : mode == 4 ? 3 : mode == 3 ? 4 : mode == 1 ? 5 : ref_idx0 > 0 ? 5 + ref_idx0 : 2;
    if (!bip && code > 4) code--;
    if (!spl && code > 1) code--;
    if (demote && code < 4) code = (code + 3) & 3;
  }
  return code == maxbit ? maxbit : code + 1;
}

// the two-mode scan of write_coeff (:160-250) over levels in scan order, as a length
TDN int scan_bits(const int16_t *s, int N, int last_pos, bool chroma, bool intra, int size) {
  const bool small_chroma = chroma && size <= 8;
  int bits = 0, pos = 0;
  if (chroma) {
    bits = 1;
    if (last_pos == 0 && iabs(s[0]) == 1) return 2;
  }
  int vlc_adaptive = intra && !chroma;
  int level_mode = 1, level = 1;
  while (pos <= last_pos) {
    if (level_mode) {
      while (pos <= last_pos && level > 0) {
        level = iabs(s[pos]);
        bits += vlc_len(vlc_adaptive, (uint32_t)level) + (level > 0);
        if (!chroma) vlc_adaptive = level > 3;
        pos++;
      }
    }
    const int maxrun = N - pos - 1;
    int run = 0, c = 0;
    while (c == 0 && pos <= last_pos) {
      c = s[pos];
      if (c == 0) run++;
      else {
        level = iabs(c);
        const int cn = find_code(run, level, maxrun, chroma);
        bits += small_chroma ? expgolomb_len((uint32_t)cn) : (cn == 0 ? 2 : vlc_len(2, (uint32_t)cn + 1));
        bits += level > 1 ? vlc_len(0, (uint32_t)(2 * (level - 2) + (c < 0))) : 1;
        run = 0;
      }
      pos++;
      level_mode = level > 1;
    }
  }
  if (pos < N && level_mode) {
    level = iabs(s[pos]);
    bits += vlc_len(vlc_adaptive, (uint32_t)level) + (level > 0);
    pos++;
  }
  if (pos < N) bits += small_chroma ? 1 : chroma ? 2 : 3;   // end of block: table 0 cn 0 / 2 bits / table 2 cn 3
  return bits;
}

#ifndef THOR_HOSTEMU
// scan_bits by one warp, eight scan positions per lane.  The two-mode scan looks sequential but its state is a function of
// the levels before a position only:
//   - a position is coded in LEVEL mode iff the nearest earlier level that is not 1 is > 1 (or there is none): a zero
//     ends level mode, a level > 1 in run mode starts it, a level of 1 changes nothing;
//   - the adaptive table of a level-mode code is "the previous level-mode position held a level > 3";
//   - a run-mode code's run and maxrun count from the nearest earlier CODED position (level-mode position or non-zero).
// Three exclusive "latest index with property" scans give all of that; every position then prices itself.  (Checked
// against scan_bits on 200 000 random blocks as a Python model before it was written here; the streams of
// tests/test_gpu_encoder.py depend on every count.)  s: 16-byte aligned, N entries (a multiple of 8 or 16 = 4x4), zeros
// after last_pos.
TD int warp_latest(int local_before, int mine_incl, int lane) {       // exclusive scan of "latest key" across the warp
  int incl = mine_incl;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl = max(incl, t);
  }
  int excl = __shfl_up_sync(0xffffffffu, incl, 1);
  if (lane == 0) excl = -1;
  return max(local_before, excl);
}
TDN int scan_bits_warp(const int16_t *s, int N, int last_pos, bool chroma, bool intra, int size, int lane) {
  // a warp prices eight positions per lane in ~500 instructions whatever the block holds; the sequential walk costs ~25 per
  // position up to the last level, and most blocks end early (measured: always-parallel was 5 % faster per frame and 5 %
  // slower for 24 streams together)
#ifndef THOR_SCAN_MIN
#define THOR_SCAN_MIN 24
#endif
  if (last_pos < THOR_SCAN_MIN) return scan_bits(s, N, last_pos, chroma, intra, size);
  const bool small_chroma = chroma && size <= 8;
  int a[8];
  unsigned neg = 0;
  {
    uint4 v = make_uint4(0, 0, 0, 0);
    if (lane * 8 < N) v = *reinterpret_cast<const uint4 *>(s + lane * 8);
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 8; i++) {
      const int c = (int)(int16_t)(w[i >> 1] >> (16 * (i & 1)));
      a[i] = iabs(c);
      neg |= (unsigned)(c < 0) << i;
    }
  }
  const int p0 = lane * 8;
  // 1. nearest earlier level != 1: key = (index << 1) | (level > 1)
  int k1[8], run = -1;
#pragma unroll
  for (int i = 0; i < 8; i++) { k1[i] = run; if (a[i] != 1) run = ((p0 + i) << 1) | (a[i] > 1); }
  const int e1 = warp_latest(-1, run, lane);
  unsigned lvl = 0;                                                    // bit i: position p0 + i is coded in level mode
#pragma unroll
  for (int i = 0; i < 8; i++) { const int k = max(k1[i], e1); lvl |= (unsigned)(k < 0 || (k & 1)) << i; }
  // 2. nearest earlier level-mode position, 3. nearest earlier coded position
  int kl[8], kc[8], rl = -1, rc = -1;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    kl[i] = rl; kc[i] = rc;
    if ((lvl >> i) & 1) rl = p0 + i;
    if (((lvl >> i) & 1) || a[i] != 0) rc = p0 + i;
  }
  const int el = warp_latest(-1, rl, lane), ec = warp_latest(-1, rc, lane);
  // the adaptive-table flag needs the LEVEL at that position: fetch it from shared memory (any lane may own it)
  int bits = 0;
  const int P = last_pos + 1;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    const int p = p0 + i;
    if (p > P || p >= N) continue;
    const bool level_mode = (lvl >> i) & 1;
    if (level_mode) {
      int va = 0;
      if (!chroma) {
        const int ql = max(kl[i], el);
        va = ql < 0 ? (int)intra : (iabs(s[ql]) > 3);
      }
      bits += vlc_len(va, (uint32_t)a[i]) + (a[i] > 0);                // p == P: the extra level after the last one (a zero)
      if (p == P && p + 1 < N) bits += small_chroma ? 1 : chroma ? 2 : 3;
    } else if (p == P) {
      bits += small_chroma ? 1 : chroma ? 2 : 3;                       // end of block
    } else if (a[i] != 0) {
      const int c = max(kc[i], ec), start = c + 1;
      const int cn = find_code(p - start, a[i], N - start - 1, chroma);
      bits += small_chroma ? expgolomb_len((uint32_t)cn) : (cn == 0 ? 2 : vlc_len(2, (uint32_t)cn + 1));
      bits += a[i] > 1 ? vlc_len(0, (uint32_t)(2 * (a[i] - 2) + ((neg >> i) & 1))) : 1;
    }
  }
  return __reduce_add_sync(0xffffffffu, bits) + (chroma ? 1 : 0);
}
#endif

// write_coeff (enc/write_bits.c:110-253) as a length.  coeff: size x size block, stride `size`; the caller guarantees at
// least one non-zero level (cbp set).  Uses S.scan; every thread returns the same number.
TDN int coeff_bits(Grp &g, SbMem &S, const int16_t *coeff, int size, int type) {
  THOR_LOCAL(&g);
  THOR_SHS(&S); THOR_SHARED(coeff);
  const int q = imin(size, 16), N = q * q;
  const bool chroma = type & 1, intra = (type >> 1) & 1;
  const uint8_t *zz = zigzag(S, q);
  g.sync();                                      // earlier readers of S.scan are done
  int mylast = -1;
  TFOR(g, i, N) {
    const int16_t c = coeff[(p2div(i, q)) * size + (p2mod(i, q))];
    S.scan[zz[i]] = c;
    if (c) mylast = imax(mylast, (int)zz[i]);
  }
  const int last_pos = imax(g.vmax(mylast), 0);    // barrier inside
  g.sync();
#if defined(THOR_HOSTEMU) || defined(THOR_SEQ_SCAN)
  return scan_bits(S.scan, N, last_pos, chroma, intra, size);
#else
  return scan_bits_warp(S.scan, N, last_pos, chroma, intra, size, g.tid & 31);       // every warp computes the same number
#endif
}

// write_coeff for the Y, U and V blocks of an unsplit coding block at once (planes with their bit in `mask`): one scatter
// pass, then the three serial scans run side by side on three warps.  Returns the sum of the lengths.
TDN int coeff_bits3(Grp &g, SbMem &S, const int16_t *cy, const int16_t *cu, const int16_t *cv, int size, int ctype, int mask) {
  THOR_SHS(&S);
  const int qy = imin(size, 16), qc = imin(size / 2, 16);
  const int c0 = (mask & 1) ? qy * qy : 0, c1 = (mask & 2) ? qc * qc : 0, c2 = (mask & 4) ? qc * qc : 0;
  g.sync();                                      // earlier readers of S.scan are done
  int last[3] = {-1, -1, -1};
  TFOR(g, i, c0 + c1 + c2) {
    int j, t;
    if (i < c0) { j = i; t = 0; } else if (i < c0 + c1) { j = i - c0; t = 1; } else { j = i - c0 - c1; t = 2; }
    const int q = t ? qc : qy, sz = t ? size / 2 : size;
    const int16_t c = (t == 0 ? cy : t == 1 ? cu : cv)[(p2div(j, q)) * sz + (p2mod(j, q))];
    const int p = zigzag(S, q)[j];
    S.scan[t * 256 + p] = c;
    if (c) last[t] = imax(last[t], p);
  }
  g.vmax3(last);                                 // barrier inside
  g.sync();
  for (int t = g.tid >> 5; t < 3; t += (g.n + 31) >> 5)
    if ((mask >> t) & 1) {
      const int q = t ? qc : qy;
#if defined(THOR_HOSTEMU) || defined(THOR_SEQ_SCAN)
      const int b = scan_bits(S.scan + t * 256, q * q, imax(last[t], 0), t != 0, (ctype >> 1) & 1, t ? size / 2 : size);
#else
      const int b = scan_bits_warp(S.scan + t * 256, q * q, imax(last[t], 0), t != 0, (ctype >> 1) & 1, t ? size / 2 : size, g.tid & 31);
#endif
      if ((g.tid & 31) == 0) S.bits3[t] = b;
    }
  g.sync();
  return ((mask & 1) ? S.bits3[0] : 0) + ((mask & 2) ? S.bits3[1] : 0) + ((mask & 4) ? S.bits3[2] : 0);
}

// write_block (enc/write_bits.c:364-650) as a length
TDN int block_bits(Grp &g, SbMem &S, const thor_enc_frame &F, const BlockCtx &B, const Param &P, const Coef &C) {
  PROF(PF_BITS);
  THOR_LOCAL(&B); THOR_LOCAL(&P); THOR_LOCAL(&g);
  THOR_SHS(&S); THOR_SHF(&F);
  const int mode = P.mode, size = B.size, part = P.pb_part;
  int bits = super_mode_len(F, B, mode, P.ref_idx0);
  if (size == 64 && mode != 0 && F.max_delta_qp) bits += vlc_len(0, (uint32_t)iabs(B.delta_qp)) + (B.delta_qp != 0);
  if (mode == 1) {
    const int im = P.intra_mode;
    if (F.num_intra_modes <= 4) bits += 2;
    else if (F.num_intra_modes <= 8) bits += im == 0 || im == 2 || im == 3 ? 2 : im == 4 || im == 8 || im == 9 ? 4 : 5;
    else bits += im < 2 ? 3 : im < 4 ? 2 : im >= 8 ? 4 : 5;
  } else if (mode == 2) {
    if (B.max_num_pb_part > 1) bits += part == 0 ? 1 : part == 1 ? 2 : 3;
    bits += mv_list_len(P.mv0, B.mvp, part);
  } else if (mode == 3) {
    Mv p = B.mvp;
    if (part == 0) bits += mvd_len(P.mv0[0], p);
    if (F.frame_type == 2) p = P.mv0[0];
    bits += mv_list_len(P.mv1, p, part);
    if (F.frame_type == 1) {
      if (F.num_ref == 2) { const int c = 2 * P.ref_idx0 + P.ref_idx1; bits += c == 3 ? 3 : c + 1; }
      else bits += expgolomb_len((uint32_t)(4 * P.ref_idx0 + P.ref_idx1));
    }
  } else if (mode == 0) {
    bits += idx_len(B.num_skip_vec, P.skip_idx);
  } else {
    bits += idx_len(B.num_merge_vec, P.skip_idx);
  }
  if (mode == 0) return bits;
  const int cy = P.cbp_y, cu = P.cbp_u, cv = P.cbp_v, split = P.tb_split;
  const int ctype = (mode == 1) << 1;
  const int max_tb = mode == 2 ? (B.max_num_tb_part > 1 ? 2 : 1) : mode == 1 ? B.max_num_tb_part : 1;
  int code;
  if (max_tb > 1) {
    if (split) code = 2;
    else { code = cbp_code(cy, cu, cv, B.ctx_cbp); if (code > 1) code++; }
  } else if (mode == 4) {
    code = (int)((0x47362501u >> (4 * (cy + 2 * cu + 4 * cv))) & 7);
    code = code == 1 ? 7 : code > 1 ? code - 1 : code;
  } else code = cbp_code(cy, cu, cv, B.ctx_cbp);
  bits += vlc_len(0, (uint32_t)code);
  if (!split) {
    const int mask = (cy ? 1 : 0) | (cu ? 2 : 0) | (cv ? 4 : 0);
    if (mask) bits += coeff_bits3(g, S, C.y, C.u, C.v, size, ctype, mask);
  } else if (size > 8) {
    for (int k = 0; k < 4; k++) {
      const int y = (cy >> (3 - k)) & 1, u = (cu >> (3 - k)) & 1, v = (cv >> (3 - k)) & 1;
      bits += vlc_len(0, (uint32_t)cbp_code(y, u, v, B.ctx_cbp));
      if (y) bits += coeff_bits(g, S, C.y + k * (size / 2) * (size / 2), size / 2, ctype);
      if (u) bits += coeff_bits(g, S, C.u + k * (size / 4) * (size / 4), size / 4, ctype | 1);
      if (v) bits += coeff_bits(g, S, C.v + k * (size / 4) * (size / 4), size / 4, ctype | 1);
    }
  } else {
    for (int k = 0; k < 4; k++) {
      const int y = (cy >> (3 - k)) & 1;
      bits += 1;
      if (y) bits += coeff_bits(g, S, C.y + k * (size / 2) * (size / 2), size / 2, ctype);
    }
    const int c = cu + 2 * cv;
    bits += c == 0 ? 1 : c == 1 ? 2 : 3;
    if (cu) bits += coeff_bits(g, S, C.u, size / 2, ctype | 1);
    if (cv) bits += coeff_bits(g, S, C.v, size / 2, ctype | 1);
  }
  return bits;
}

}  // namespace thorenc
