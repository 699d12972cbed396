/* TEST INFRASTRUCTURE ONLY -- CPU oracle of the bit-cost model, first slice: the exact number of bits
 * the reference's entropy coder spends on a block's quantised coefficients and on a motion vector difference.
 * Restates   enc/putvlc.c:133-229   (quote_vlc: code lengths of the 14 VLC tables),
 *            enc/write_bits.c:50-69  (write_mv), :71-108 (find_code), :110-253 (write_coeff)
 * as a pure length computation (nothing is written anywhere).  Pinned against the compiled reference by
 * tests/test_oracle_vs_ref.py::test_coeff_bits_* (write_coeff into a real stream_t, bit position before / after).
 * Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may use this file. */
#include <stdlib.h>
#include "thor_oracle.h"

static int ilog2u(unsigned v) { int n = 0; while (v >>= 1) n++; return n; }

/* length of code number cn in table n; -1 where the reference calls fatalerror (putvlc.c:177,225) */
int orc_vlc_len(unsigned n, unsigned cn) {
  if (n <= 5) {                                   /* :141-157  Golomb-like prefix, exp-Golomb escape */
    const unsigned lim = 6u << n;
    if ((int)cn < (int)lim) return 1 + (int)n + (int)(cn >> n);
    return (6 - (int)n) + 1 + 2 * ilog2u(cn - lim + (1u << n));
  }
  if (n <= 7) return 1 + ((int)n - 4) + (int)(cn >> (n - 4));        /* :158-163 */
  if (n == 8) return cn == 0 ? 1 : cn <= 2 ? 2 : -1;                 /* :164-181 */
  if (n == 9) return cn == 0 ? 3 : cn <= 2 ? 4 : cn < 11 ? 5 : 5 + (int)((cn + 5) >> 4);   /* :182-209 */
  if (n == 10) return 1 + 2 * ilog2u(cn + 1);                        /* :210-213  exp-Golomb */
  if (n == 11) return cn < 2 ? (int)cn + 1 : (int)(cn / 2) + 3;      /* :214-217 */
  if (n == 12) return cn + 1 < 4 ? (int)cn + 1 : 4;                  /* :218-221 */
  if (n == 13) return cn + 1 < 6 ? (int)cn + 1 : 6;                  /* :222-225 */
  return -1;
}

/* write_mv (write_bits.c:50-69): two exp-Golomb codes of the signed differences */
int orc_mv_bits(int mvx, int mvy, int mvpx, int mvpy) {
  const int dx = mvx - mvpx, dy = mvy - mvpy;
  return orc_vlc_len(10, (unsigned)(2 * abs(dx) - (dx < 0))) + orc_vlc_len(10, (unsigned)(2 * abs(dy) - (dy < 0)));
}

/* find_code (write_bits.c:71-108): code number of the joint event (run, level > 1), or of the end of block */
int orc_find_code(int run, int level, int maxrun, int chroma, int eob) {
  const int m = maxrun > 4 ? maxrun : 4;
  const int index = run + (level > 1 ? m + 1 : 0);
  if (eob) return chroma ? 0 : 2;
  if (!chroma && index < 2) return index;
  if (index <= 4) return index + 1;
  if (index <= m) return index + 3;
  if (index == m + 1) return 6;
  if (index == m + 2) return 7;
  return index + 1;
}

/* bits of one run-mode / end-of-block code number (write_bits.c:196-206, :238-249) */
static int event_bits(int cn, int chroma, int size, int eob) {
  if (chroma && size <= 8) return orc_vlc_len(eob ? 0 : 10, (unsigned)cn);
  return cn == 0 ? 2 : orc_vlc_len(2, (unsigned)cn + 1);
}

/* write_coeff (write_bits.c:110-253) as a length: coeff = the block's quantised levels, row stride `stride` (the
 * reference passes stride == size; libthorcuda's pool uses stride == min(size,16)), type bit0 chroma, bit1 intra.
 * An all-zero block costs 0 bits here (the reference never calls write_coeff for one: cbp is 0, :146-147). */
int orc_coeff_bits(const int16_t *coeff, int stride, int size, int type) {
  const int q = size < 16 ? size : 16, n = q * q, chroma = type & 1;
  const int *zz = orc_zigzag(q);
  int16_t s[256];
  int last = -1, bits = 0;
  for (int i = 0; i < q; i++)
    for (int j = 0; j < q; j++) s[zz[i * q + j]] = coeff[i * stride + j];
  for (int p = 0; p < n; p++) if (s[p]) last = p;
  if (last < 0) return 0;
  if (chroma) {                                                 /* :147-158 */
    if (last == 0 && abs(s[0]) == 1) return 2;
    bits = 1;
  }
  int adaptive = ((type >> 1) & 1) && !chroma;                   /* table of the next level-mode symbol */
  int level_mode = 1, pos = 0, run = 0, maxrun = 0;
  while (pos <= last) {
    const int level = abs(s[pos]);
    if (level_mode) {                                            /* :166-178 */
      bits += orc_vlc_len((unsigned)adaptive, (unsigned)level) + (level > 0);
      if (!chroma) adaptive = level > 3;
      pos++;
      if (level == 0) { level_mode = 0; run = 0; maxrun = n - pos - 1; }
    } else if (level == 0) {                                     /* :184-190 */
      run++;
      pos++;
    } else {                                                     /* :191-218 */
      bits += event_bits(orc_find_code(run, level, maxrun, chroma, 0), chroma, size, 0);
      bits += level > 1 ? orc_vlc_len(0, (unsigned)(2 * (level - 2) + (s[pos] < 0))) : 1;
      pos++;
      level_mode = level > 1;
      run = 0;
      maxrun = n - pos - 1;
    }
  }
  if (pos < n && level_mode) {                                   /* :222-234  one more (zero) level before the EOB */
    bits += orc_vlc_len((unsigned)adaptive, 0);
    pos++;
  }
  if (pos < n) bits += event_bits(orc_find_code(0, 0, 0, chroma, 1), chroma, size, 1);   /* :237-250 */
  return bits;
}

/* ---- write_block as a length (enc/write_bits.c:268-650) ---- */
enum { FT_I = 0, FT_P = 1, FT_B = 2, M_SKIP = 0, M_INTRA = 1, M_INTER = 2, M_BIPRED = 3, M_MERGE = 4 };

/* write_super_mode with split_flag == 0 (:268-362): a unary code whose rank depends on what the frame offers */
static int super_mode_bits(const int32_t *d) {
  const int size = d[ORC_BS_SIZE], mode = d[ORC_BS_MODE], num_ref = d[ORC_BS_NUM_REF], r0 = d[ORC_BS_REF_IDX0];
  if (d[ORC_BS_FRAME_TYPE] == FT_I) return size > 8 ? 1 : 0;                       /* :356-360 */
  const int bipred_possible = num_ref > 1 && d[ORC_BS_ENABLE_BIPRED], split_possible = size > 8;
  const int maxbit = 2 + num_ref + split_possible + bipred_possible;
  const int demote_skip = (d[ORC_BS_CTX_INDEX] == 2 || d[ORC_BS_CTX_INDEX] > 3) && size > 8;
  int code;
  if (d[ORC_BS_INTERP_REF]) {                                                     /* :291-320 */
    code = mode == M_SKIP ? 0 : mode == M_MERGE ? 2 : mode == M_BIPRED ? 3 : mode == M_INTRA ? 4 : r0 > 0 ? 4 + r0 : 4 + num_ref;
    if (!bipred_possible && code > 3) code--;
    if (!split_possible && code > 1) code--;
    if (demote_skip && code < 3) code = (code + 2) % 3;
  } else {                                                                        /* :321-348 */
    code = mode == M_SKIP ? 0 : mode == M_MERGE ? 3 : mode == M_BIPRED ? 4 : mode == M_INTRA ? 5 : r0 > 0 ? 5 + r0 : 2;
    if (!bipred_possible && code > 4) code--;
    if (!split_possible && code > 1) code--;
    if (demote_skip && code < 4) code = (code + 3) % 4;
  }
  return code == maxbit ? maxbit : code + 1;                                      /* :350-353 */
}

static int pb_part_bits(int max_num_pb_part, int pb_part) {                       /* :410-420 */
  if (max_num_pb_part <= 1) return 0;
  return pb_part == 0 ? 1 : pb_part == 1 ? 2 : 3;
}
static int mvd_bits(const int32_t *mv, const int32_t *mvp) { return orc_mv_bits(mv[0], mv[1], mvp[0], mvp[1]); }
/* motion vectors of one list for a partition (:421-444, :468-490): the first against mvp, the others against the first */
static int mv_list_bits(const int32_t *mv, const int32_t *mvp, int pb_part) {
  int n = mvd_bits(mv, mvp);
  if (pb_part == 1) n += mvd_bits(mv + 4, mv);                /* HOR: vectors 0 and 2 */
  else if (pb_part == 2) n += mvd_bits(mv + 2, mv);           /* VER: vectors 0 and 1 */
  else if (pb_part == 3) n += mvd_bits(mv + 2, mv) + mvd_bits(mv + 4, mv) + mvd_bits(mv + 6, mv);
  return n;
}
static int idx_bits(int nvec, int idx) {                                          /* :507-530 */
  if (nvec == 4) return 2;
  if (nvec == 3) return idx == 0 ? 1 : 2;
  return nvec == 2 ? 1 : 0;
}
static const int cbp_rank[8] = {1, 0, 5, 2, 6, 3, 7, 4};                          /* :382 */
static int cbp_code(int y, int u, int v, int ctx_cbp) {                           /* :561-566 */
  int code = cbp_rank[y + 2 * u + 4 * v];
  if (ctx_cbp == 0 && code < 2) code = 1 - code;
  return code;
}

int orc_block_syntax_bits(const int32_t *d) {
  const int size = d[ORC_BS_SIZE], mode = d[ORC_BS_MODE], pb_part = d[ORC_BS_PB_PART], tb_split = d[ORC_BS_TB_SPLIT];
  const int cy = d[ORC_BS_CBP_Y], cu = d[ORC_BS_CBP_U], cv = d[ORC_BS_CBP_V], ctx_cbp = d[ORC_BS_CTX_CBP];
  int n = super_mode_bits(d);
  if (size == 64 && mode != M_SKIP && d[ORC_BS_MAX_DELTA_QP])                     /* :389-391, write_delta_qp :255-266 */
    n += orc_vlc_len(0, (unsigned)abs(d[ORC_BS_DELTA_QP])) + (d[ORC_BS_DELTA_QP] != 0);
  if (mode == M_INTRA) {                                                          /* :393-413 */
    const int nm = d[ORC_BS_NUM_INTRA_MODES], im = d[ORC_BS_INTRA_MODE];
    if (nm <= 4) n += 2;
    else if (nm <= 8) { static const int map[10] = {2, 8, 1, 0, 5, 9, 7, 6, 4, 3}, len[8] = {2, 2, 2, 4, 4, 4, 5, 5}; n += len[map[im]]; }
    else { static const int map[10] = {2, 3, 1, 0, 6, 9, 8, 7, 5, 4}, len[10] = {2, 2, 3, 3, 4, 4, 5, 5, 5, 5}; n += len[map[im]]; }
  } else if (mode == M_INTER) {
    n += pb_part_bits(d[ORC_BS_MAX_PB_PART], pb_part) + mv_list_bits(d + ORC_BS_MV0, d + ORC_BS_MVP, pb_part);
  } else if (mode == M_BIPRED) {                                                  /* :446-505, BIPRED_PART 0 */
    const int32_t *mvp = d + ORC_BS_MVP;
    if (pb_part == 0) n += mvd_bits(d + ORC_BS_MV0, mvp);
    if (d[ORC_BS_FRAME_TYPE] == FT_B) mvp = d + ORC_BS_MV0;
    n += mv_list_bits(d + ORC_BS_MV1, mvp, pb_part);
    if (d[ORC_BS_FRAME_TYPE] == FT_P) {
      if (d[ORC_BS_NUM_REF] == 2) { const int c = 2 * d[ORC_BS_REF_IDX0] + d[ORC_BS_REF_IDX1]; n += c == 3 ? 3 : c + 1; }
      else n += orc_vlc_len(10, (unsigned)(4 * d[ORC_BS_REF_IDX0] + d[ORC_BS_REF_IDX1]));
    }
  } else if (mode == M_SKIP) {
    n += idx_bits(d[ORC_BS_NUM_SKIP_VEC], d[ORC_BS_SKIP_IDX]);
  } else if (mode == M_MERGE) {
    n += idx_bits(d[ORC_BS_NUM_MERGE_VEC], d[ORC_BS_SKIP_IDX]);
  }
  if (mode == M_SKIP) return n;

  int max_tb = 1, code;                                                           /* :533-570 */
  if (mode == M_INTER) max_tb = d[ORC_BS_MAX_TB_PART] > 1 ? 2 : 1;
  else if (mode == M_INTRA) max_tb = d[ORC_BS_MAX_TB_PART];
  if (max_tb > 1) {
    if (tb_split) code = 2;
    else { code = cbp_code(cy, cu, cv, ctx_cbp); if (code > 1) code++; }
  } else if (mode == M_MERGE) {
    code = cbp_rank[cy + 2 * cu + 4 * cv];
    code = code == 1 ? 7 : code > 1 ? code - 1 : code;
  } else code = cbp_code(cy, cu, cv, ctx_cbp);
  n += orc_vlc_len(0, (unsigned)code);
  if (tb_split) {
    if (size > 8) {                                                               /* :585-612: a cbp code per quadrant */
      for (int i = 0; i < 4; i++)
        n += orc_vlc_len(0, (unsigned)cbp_code((cy >> (3 - i)) & 1, (cu >> (3 - i)) & 1, (cv >> (3 - i)) & 1, ctx_cbp));
    } else {                                                                      /* :613-641: 4 luma flags + chroma pair */
      const int c = cu + 2 * cv;
      n += 4 + (c == 0 ? 1 : c == 1 ? 2 : 3);
    }
  }
  return n;
}

int orc_block_bits(const int32_t *d, const int16_t *coeff_y, const int16_t *coeff_u, const int16_t *coeff_v) {
  const int size = d[ORC_BS_SIZE], mode = d[ORC_BS_MODE], type = (mode == M_INTRA) << 1;
  const int cy = d[ORC_BS_CBP_Y], cu = d[ORC_BS_CBP_U], cv = d[ORC_BS_CBP_V];
  int n = orc_block_syntax_bits(d);
  if (mode == M_SKIP) return n;
  if (!d[ORC_BS_TB_SPLIT]) {                                                      /* :572-582 */
    if (cy) n += orc_coeff_bits(coeff_y, size, size, type);
    if (cu) n += orc_coeff_bits(coeff_u, size / 2, size / 2, type | 1);
    if (cv) n += orc_coeff_bits(coeff_v, size / 2, size / 2, type | 1);
  } else if (size > 8) {
    const int h = size / 2, q = size / 4;
    for (int i = 0; i < 4; i++) {
      if ((cy >> (3 - i)) & 1) n += orc_coeff_bits(coeff_y + i * h * h, h, h, type);
      if ((cu >> (3 - i)) & 1) n += orc_coeff_bits(coeff_u + i * q * q, q, q, type | 1);
      if ((cv >> (3 - i)) & 1) n += orc_coeff_bits(coeff_v + i * q * q, q, q, type | 1);
    }
  } else {
    for (int i = 0; i < 4; i++)
      if ((cy >> (3 - i)) & 1) n += orc_coeff_bits(coeff_y + i * 16, 4, 4, type);
    if (cu) n += orc_coeff_bits(coeff_u, 4, 4, type | 1);
    if (cv) n += orc_coeff_bits(coeff_v, 4, 4, type | 1);
  }
  return n;
}
