// Frame encoder of libthorcuda -- the reference's block layer (enc/encode_block.c:1398-3033) for one superblock:
// trial encodes, early skip, motion / intra search, rate-distortion mode decision and the quad-tree recursion.
// One thread group (a CTA on the device) executes it with uniform control flow; see enc_common.h.
#pragma once
#include "enc_me.h"
#include "enc_xform.h"

namespace thorenc {

struct Enc {
  Grp &g;
  SbMem &S;
  const thor_enc_frame &F;
  SbState st;
  thor_enc_node *nodes;              // this superblock's 85 nodes (global)
  thor_enc_coef *coef;               // this superblock's coefficients (global)
  int sb_x, sb_y;                    // luma position of the superblock
};

TD const thor_enc_pic &ref_pic(const thor_enc_frame &F, int ref_idx) { return F.ref[ref_idx]; }

// z-order index (0..63) of the 8x8 unit at (y, x) inside its superblock
TD int zidx8(int ypos, int xpos) {
  const int by = (ypos & 63) >> 3, bx = (xpos & 63) >> 3;
  int z = 0;
  for (int b = 0; b < 3; b++) z |= (((bx >> b) & 1) << (2 * b)) | (((by >> b) & 1) << (2 * b + 1));
  return z;
}
TD int node_index(int size, int ypos, int xpos) {
  const int z = zidx8(ypos, xpos);
  return size == 64 ? 0 : size == 32 ? 1 + (z >> 4) : size == 16 ? 5 + (z >> 2) : 21 + z;
}

// copy_frame_to_block (enc/encode_block.c:1821-1838): the original samples of the block, compact, stride = size
TDN void load_org(Enc &E, const BlockCtx &B) {
  PROF(PF_LOAD_ORG);
  THOR_LOCAL(&E); THOR_LOCAL(&B);
  THOR_SHS(&E.S); THOR_SHF(&E.F);
  Grp &g = E.g;
  const thor_enc_pic &o = E.F.orig;
  const int size = B.size, wq = B.bwidth / 4;
  const Div dwq = mkdiv(wq);
  g.sync();
  TFOR(g, i, B.bheight * wq) {
    const int r = dq(dwq, i), c = dr(dwq, i) * 4;
    *reinterpret_cast<uint32_t *>(E.S.org.y + r * size + c) = ld_ro32u(o.y + (B.ypos + r) * o.stride_y + B.xpos + c);
  }
  const int wc = B.bwidth / 2, sc = size / 2;
  const Div dwc = mkdiv(wc);
  TFOR(g, i, (B.bheight / 2) * wc) {
    const int r = dq(dwc, i), c = dr(dwc, i);
    const int off = (B.ypos / 2 + r) * o.stride_c + B.xpos / 2 + c;
    E.S.org.u[r * sc + c] = ld_ro8(o.u + off);
    E.S.org.v[r * sc + c] = ld_ro8(o.v + off);
  }
  g.sync();
}

// make_top_and_left (common/intra_prediction.c:57-143), both branches.  rec_frame: the coding block's origin in the
// reconstructed plane; rblock: the sub-block's origin in the block's own compact reconstruction (tb_split only);
// (i, j): position of the sub-block inside the coding block; (ypos, xpos): plane position of the coding block.
TDN void make_top_and_left(Enc &E, int pl, const uint8_t *rec_frame, int fstride, const uint8_t *rblock, int rbstride, int i, int j, int ypos, int xpos,
                           int size, int cb_upright, int cb_downleft, int tb_split) {
  PROF(PF_PRED_INTRA);
  THOR_LOCAL(&E);
  THOR_SHS(&E.S); THOR_SHF(&E.F);
  Grp &g = E.g;
  uint8_t *top = E.S.top[pl], *left = E.S.left[pl];
  int downleft, upright;
  if (!tb_split) { downleft = cb_downleft; upright = cb_upright; }
  else { downleft = (j == 0 && (i == 0 || cb_downleft)); upright = (j == 0 || (i == 0 && cb_upright)); }
  const int leftlen = downleft ? size + 1 : size, toplen = upright ? size + 1 : size;
  const bool top_edge = ypos + i == 0, left_edge = xpos + j == 0;
  const bool top_in_block = tb_split && i > 0, left_in_block = tb_split && j > 0;
  g.sync();
  TFOR(g, k, 2 * size) {
    const int kt = imin(k, toplen - 1), kl = imin(k, leftlen - 1);
    top[k] = top_edge ? 128 : top_in_block ? rblock[-rbstride + kt] : ld_frame8(rec_frame - fstride + j + kt);
    left[k] = left_edge ? 128 : left_in_block ? rblock[kl * rbstride - 1] : ld_frame8(rec_frame + (i + kl) * fstride - 1);
  }
  g.sync();
  if (g.tid == 0) {
    int tl;
    if (top_edge) tl = left[0];
    else if (xpos <= 0) tl = top[0];
    else if (!top_in_block) tl = ld_frame8(rec_frame - fstride + j - 1);
    else tl = j > 0 ? rblock[-rbstride - 1] : ld_frame8(rec_frame + (i - 1) * fstride - 1);
    E.S.topleft[pl] = (uint8_t)tl;
  }
  g.sync();
}

// encode_and_reconstruct_block_intra (enc/encode_block.c:1398-1467) for one plane.  org: compact (stride size);
// pred / rec / coeffq: the plane's arrays of S.pred / S.rec / S.tc.  Returns cbp (a 4-bit mask under tb_split).
TDN int intra_plane(Enc &E, int pl, const uint8_t *org, const uint8_t *rec_frame, int fstride, int ypos, int xpos, int size, int qp, uint8_t *pred,
                    int16_t *coeffq, uint8_t *rec, int coeff_type, int tb_split, int intra_mode, int upright, int downleft) {
  THOR_LOCAL(&E);
  THOR_SHS(&E.S); THOR_SHF(&E.F); THOR_SHARED(org); THOR_SHARED(pred); THOR_SHARED(coeffq); THOR_SHARED(rec);
  Grp &g = E.g;
  SbMem &S = E.S;
  const int fast = E.F.encoder_speed > 1;
  if (!tb_split) {
    make_top_and_left(E, pl, rec_frame, fstride, nullptr, 0, 0, 0, ypos, xpos, size, upright, downleft, 0);
    intra_predict(g, S, pl, ypos, xpos, size, pred, intra_mode);
    return encode_tb(g, S, org, size, pred, size, rec, size, coeffq, size, qp, coeff_type, fast);
  }
  const int size2 = size / 2;
  int cbp = 0, index = 0;
  for (int i = 0; i < size; i += size2)
    for (int j = 0; j < size; j += size2) {
      make_top_and_left(E, pl, rec_frame, fstride, rec + i * size + j, size, i, j, ypos, xpos, size2, upright, downleft, 1);
      intra_predict(g, S, pl, ypos + i, xpos + j, size2, pred, intra_mode);     // pblock: compact size2 x size2
      const int bit = encode_tb(g, S, org + i * size + j, size, pred, size2, rec + i * size + j, size, coeffq + index, size2, qp, coeff_type, fast);
      cbp = (cbp << 1) + bit;
      index += size2 * size2;
    }
  return cbp;
}

// encode_and_reconstruct_block_inter (enc/encode_block.c:1469-1532) for one plane
TDN int inter_plane(Enc &E, const uint8_t *org, int size, int qp, const uint8_t *pred, int16_t *coeffq, uint8_t *rec, int coeff_type, int tb_split) {
  THOR_LOCAL(&E);
  THOR_SHS(&E.S); THOR_SHF(&E.F); THOR_SHARED(org); THOR_SHARED(pred); THOR_SHARED(coeffq); THOR_SHARED(rec);
  Grp &g = E.g;
  SbMem &S = E.S;
  const int speed = E.F.encoder_speed;
  if (!tb_split) return encode_tb(g, S, org, size, pred, size, rec, size, coeffq, size, qp, coeff_type, (size == 64 && speed > 0) || speed > 1);
  const int size2 = size / 2, fast = size == 64 || speed > 1;
  int cbp = 0, index = 0;
  for (int i = 0; i < size; i += size2)
    for (int j = 0; j < size; j += size2) {
      const int o = i * size + j;
      const int bit = encode_tb(g, S, org + o, size, pred + o, size, rec + o, size, coeffq + index, size2, qp, coeff_type, fast);
      cbp = (cbp << 1) + bit;
      index += size2 * size2;
    }
  return cbp;
}

// encode_block (enc/encode_block.c:1590-1800) as a trial: prediction, residual coding and reconstruction of the block with
// the parameters in P, into S.rec / S.tc; returns what write_block would emit (its length).  P.cbp_* / P.tb_split are set.
TDN int encode_block(Enc &E, const BlockCtx &B, Param &P) {
  THOR_LOCAL(&E); THOR_LOCAL(&B); THOR_LOCAL(&P);
  THOR_SHS(&E.S); THOR_SHF(&E.F);
  Grp &g = E.g;
  SbMem &S = E.S;
  const thor_enc_frame &F = E.F;
  const int size = B.size, sizeC = size / 2, yposC = B.ypos / 2, xposC = B.xpos / 2;
  const int mode = P.mode;
  const int qpY = F.qp + B.delta_qp, qpC = kChromaQp[qpY];
  const int tb_split = imax(0, P.tb_param), zero_block = P.tb_param == -1;
  const int ctype = (F.frame_type == 0) << 1;
  const int bip = F.enable_bipred;
  P.tb_split = tb_split;
  g.sync();
  if (mode == 1) {
    const thor_enc_pic &rec = F.rec;
    const int upright = upright_available(B.ypos, B.xpos, size, F.width), downleft = downleft_available(B.ypos, B.xpos, size, F.height);
    if (!tb_split) {
      // the three predictions only read neighbours outside the block: form them first, then code Y, U and V together
      make_top_and_left(E, 0, rec.y + B.ypos * rec.stride_y + B.xpos, rec.stride_y, nullptr, 0, 0, 0, B.ypos, B.xpos, size, upright, downleft, 0);
      intra_predict(g, S, 0, B.ypos, B.xpos, size, S.pred.y, P.intra_mode);
      make_top_and_left(E, 1, rec.u + yposC * rec.stride_c + xposC, rec.stride_c, nullptr, 0, 0, 0, yposC, xposC, sizeC, upright, downleft, 0);
      intra_predict(g, S, 1, yposC, xposC, sizeC, S.pred.u, P.intra_mode);
      make_top_and_left(E, 2, rec.v + yposC * rec.stride_c + xposC, rec.stride_c, nullptr, 0, 0, 0, yposC, xposC, sizeC, upright, downleft, 0);
      intra_predict(g, S, 2, yposC, xposC, sizeC, S.pred.v, P.intra_mode);
      const int fast = F.encoder_speed > 1;
      const TbDesc d[3] = {{S.org.y, size, S.pred.y, size, S.rec.y, size, S.tc.y, size, qpY, ctype, fast},
                           {S.org.u, sizeC, S.pred.u, sizeC, S.rec.u, sizeC, S.tc.u, sizeC, qpC, ctype | 1, fast},
                           {S.org.v, sizeC, S.pred.v, sizeC, S.rec.v, sizeC, S.tc.v, sizeC, qpC, ctype | 1, fast}};
      const int m = encode_tb3(g, S, d, 3);
      P.cbp_y = m & 1; P.cbp_u = (m >> 1) & 1; P.cbp_v = (m >> 2) & 1;
    } else {
      P.cbp_y = intra_plane(E, 0, S.org.y, rec.y + B.ypos * rec.stride_y + B.xpos, rec.stride_y, B.ypos, B.xpos, size, qpY, S.pred.y, S.tc.y, S.rec.y,
                            ctype, tb_split, P.intra_mode, upright, downleft);
      P.cbp_u = intra_plane(E, 1, S.org.u, rec.u + yposC * rec.stride_c + xposC, rec.stride_c, yposC, xposC, sizeC, qpC, S.pred.u, S.tc.u, S.rec.u,
                            ctype | 1, tb_split && size > 8, P.intra_mode, upright, downleft);
      P.cbp_v = intra_plane(E, 2, S.org.v, rec.v + yposC * rec.stride_c + xposC, rec.stride_c, yposC, xposC, sizeC, qpC, S.pred.v, S.tc.v, S.rec.v,
                            ctype | 1, tb_split && size > 8, P.intra_mode, upright, downleft);
    }
  } else {
    Blk &dst = mode == 0 ? S.rec : S.pred;
    const thor_enc_pic &r0 = ref_pic(F, P.ref_idx0);
    if (mode == 0 || mode == 4) {                                      // SKIP / MERGE (:1688-1732)
      if (P.dir == 2) {
        const thor_enc_pic &r1 = ref_pic(F, P.ref_idx1);
        inter_yuv(g, S, F, r0, S.p0, B, P.mv0, r0.frame_num >= F.rec.frame_num, bip, 0);
        inter_yuv(g, S, F, r1, S.p1, B, P.mv1, r1.frame_num >= F.rec.frame_num, bip, 0);
        average_blocks(g, dst, S.p0, S.p1, B);
      } else {
        inter_yuv(g, S, F, r0, dst, B, P.mv0, r0.frame_num > F.rec.frame_num, bip, 0);
      }
    } else if (mode == 2) {                                            // INTER (:1734-1739)
      inter_yuv(g, S, F, r0, dst, B, P.mv0, r0.frame_num > F.rec.frame_num, bip, F.enable_pb_split);
    } else {                                                           // BIPRED (:1740-1753)
      const thor_enc_pic &r1 = ref_pic(F, P.ref_idx1);
      inter_yuv(g, S, F, r0, S.p0, B, P.mv0, r0.frame_num >= F.rec.frame_num, bip, F.enable_pb_split);
      inter_yuv(g, S, F, r1, S.p1, B, P.mv1, r1.frame_num >= F.rec.frame_num, bip, F.enable_pb_split);
      average_blocks(g, dst, S.p0, S.p1, B);
    }
    if (mode == 0) {
      P.cbp_y = P.cbp_u = P.cbp_v = 0;
    } else if (zero_block) {
      TFOR(g, i, size * size / 4) reinterpret_cast<uint32_t *>(S.rec.y)[i] = reinterpret_cast<const uint32_t *>(S.pred.y)[i];
      TFOR(g, i, sizeC * sizeC / 4) {
        reinterpret_cast<uint32_t *>(S.rec.u)[i] = reinterpret_cast<const uint32_t *>(S.pred.u)[i];
        reinterpret_cast<uint32_t *>(S.rec.v)[i] = reinterpret_cast<const uint32_t *>(S.pred.v)[i];
      }
      g.sync();
      P.cbp_y = P.cbp_u = P.cbp_v = 0;
    } else {
      if (!tb_split) {
        const int speed = F.encoder_speed, fastC = speed > 1, fastY = (size == 64 && speed > 0) || speed > 1;
        const TbDesc d[3] = {{S.org.y, size, S.pred.y, size, S.rec.y, size, S.tc.y, size, qpY, ctype, fastY},
                             {S.org.u, sizeC, S.pred.u, sizeC, S.rec.u, sizeC, S.tc.u, sizeC, qpC, ctype | 1, fastC},
                             {S.org.v, sizeC, S.pred.v, sizeC, S.rec.v, sizeC, S.tc.v, sizeC, qpC, ctype | 1, fastC}};
        const int m = encode_tb3(g, S, d, 3);
        P.cbp_y = m & 1; P.cbp_u = (m >> 1) & 1; P.cbp_v = (m >> 2) & 1;
      } else {
        P.cbp_y = inter_plane(E, S.org.y, size, qpY, S.pred.y, S.tc.y, S.rec.y, ctype, tb_split);
        P.cbp_u = inter_plane(E, S.org.u, sizeC, qpC, S.pred.u, S.tc.u, S.rec.u, ctype | 1, tb_split && size > 8);
        P.cbp_v = inter_plane(E, S.org.v, sizeC, qpC, S.pred.v, S.tc.v, S.rec.v, ctype | 1, tb_split && size > 8);
      }
    }
  }
  return block_bits(g, S, F, B, P, S.tc);
}

// copy_best_parameters (enc/encode_block.c:1983-2045): the trial in S.rec / S.tc / P becomes the block's best so far
TDN void keep_best(Enc &E, const BlockCtx &B, const Param &P, Param &best) {
  PROF(PF_KEEP);
  THOR_LOCAL(&E); THOR_LOCAL(&B); THOR_LOCAL(&P); THOR_LOCAL(&best);
  THOR_SHS(&E.S); THOR_SHF(&E.F);
  Grp &g = E.g;
  SbMem &S = E.S;
  const int size = B.size, ny = size * size / 4, nc = ny / 4;
  g.sync();
  TFOR(g, i, ny) reinterpret_cast<uint32_t *>(S.best.y)[i] = reinterpret_cast<const uint32_t *>(S.rec.y)[i];
  TFOR(g, i, nc) {
    reinterpret_cast<uint32_t *>(S.best.u)[i] = reinterpret_cast<const uint32_t *>(S.rec.u)[i];
    reinterpret_cast<uint32_t *>(S.best.v)[i] = reinterpret_cast<const uint32_t *>(S.rec.v)[i];
  }
  if (P.cbp_y) TFOR(g, i, ny * 2) reinterpret_cast<uint32_t *>(S.bc.y)[i] = reinterpret_cast<const uint32_t *>(S.tc.y)[i];
  if (P.cbp_u) TFOR(g, i, nc * 2) reinterpret_cast<uint32_t *>(S.bc.u)[i] = reinterpret_cast<const uint32_t *>(S.tc.u)[i];
  if (P.cbp_v) TFOR(g, i, nc * 2) reinterpret_cast<uint32_t *>(S.bc.v)[i] = reinterpret_cast<const uint32_t *>(S.tc.v)[i];
  g.sync();
  best = P;
  if (P.mode == 0 || P.mode == 4) {
    const InterPred &c = P.mode == 0 ? B.skip_cand[P.skip_idx] : B.merge_cand[P.skip_idx];
    best.ref_idx0 = c.ref_idx0;
    best.ref_idx1 = c.ref_idx1;
    for (int i = 0; i < 4; i++) { best.mv0[i] = c.mv0; best.mv1[i] = c.mv1; }
    best.dir = c.dir;
  } else if (P.mode == 1) {
    best.ref_idx0 = best.ref_idx1 = 0;
    for (int i = 0; i < 4; i++) best.mv0[i] = best.mv1[i] = mkmv(0, 0);
    best.dir = -1;
  } else {
    best.dir = P.mode == 2 ? 0 : 2;
  }
}

// ---- early skip (enc/encode_block.c:2481-2785) -------------------------------------------------------------------
// check_early_skip_transform_coeff's threshold (:2491-2492): the only floating-point step, in double as the reference
TD int skip_threshold(int qp, int lg, double relative_threshold) {
  const int shift2 = 21 - lg + qp / 6;
  const double first_quantizer_level = ddiv((double)(1 << shift2), (double)kQuant[qp % 6]);
  return (int)dmul(relative_threshold, first_quantizer_level);
}

// check_early_skip_sub_block (:2507-2543): residual of a size0 x size0 luma sub-block -> 2x2 mean -> transform(size0/2)
// -> any |coefficient| > threshold.  org: stride ostride; pred: compact stride size0 (size0 >= 8).
TDN int skip_test_luma(Enc &E, const uint8_t *org, int ostride, int size0, int qp, const uint8_t *pred, float thr) {
  THOR_LOCAL(&E);
  THOR_SHS(&E.S); THOR_SHF(&E.F); THOR_SHARED(org); THOR_SHARED(pred);
  Grp &g = E.g;
  SbMem &S = E.S;
  const int n = size0 / 2, lg = ilog2((uint32_t)n);
  TFOR(g, i, n * n) {
    const int r = 2 * p2div(i, n), c = 2 * p2mod(i, n);
    const int s = ((int)org[r * ostride + c] - (int)pred[r * size0 + c]) + ((int)org[r * ostride + c + 1] - (int)pred[r * size0 + c + 1]) +
                  ((int)org[(r + 1) * ostride + c] - (int)pred[(r + 1) * size0 + c]) + ((int)org[(r + 1) * ostride + c + 1] - (int)pred[(r + 1) * size0 + c + 1]);
    S.resid[i] = (int16_t)((s + 2) >> 2);
  }
  g.sync();
  forward_core(g, S, n, lg, lg + 5);
  const int threshold = skip_threshold(qp, lg, dmul(0.5, (double)thr));
  int sig = 0;
  TFOR(g, i, n * n) sig |= iabs((int)S.coeff[i]) > threshold;
  return g.any(sig);
}

// check_early_skip_sub_blockC (:2545-2610).  The 8-wide form tests every column sum as int16; every other width runs the
// 4-wide form, which looks at the top-left 4 x 4 samples only (pairs of column sums) -- also for 16 x 16 sub-blocks.
TDN int skip_test_chroma(Enc &E, const uint8_t *org, int ostride, int size, int qp, const uint8_t *pred, float thr) {
  THOR_LOCAL(&E);
  THOR_SHS(&E.S); THOR_SHF(&E.F); THOR_SHARED(org); THOR_SHARED(pred);
  Grp &g = E.g;
  const int threshold = skip_threshold(qp, 5, (double)thr);
  int sig = 0;
  if (size == 8) {
    TFOR(g, j, 8) {
      int sum = 0;
      for (int i = 0; i < 8; i++) sum += (int)org[i * ostride + j] - (int)pred[i * 8 + j];
      sig |= (int16_t)sum > (int16_t)threshold;
    }
  } else {
    TFOR(g, j, 2) {
      int sum = 0;
      for (int i = 0; i < 4; i++)
        sum += (int16_t)((int)org[i * ostride + 2 * j] - (int)pred[i * size + 2 * j]) + (int16_t)((int)org[i * ostride + 2 * j + 1] - (int)pred[i * size + 2 * j + 1]);
      sig |= sum > threshold;
    }
  }
  return g.any(sig);
}

// check_early_skip_block (:2613-2741): 1 = no sub-block of the candidate's prediction leaves a significant residual
TDN int check_early_skip(Enc &E, const BlockCtx &B, const Param &P) {
  PROF(PF_EARLY);
  THOR_LOCAL(&E); THOR_LOCAL(&B); THOR_LOCAL(&P);
  THOR_SHS(&E.S); THOR_SHF(&E.F);
  Grp &g = E.g;
  SbMem &S = E.S;
  const thor_enc_frame &F = E.F;
  const int size = B.size, size0 = imin(size, 32), sc0 = size0 / 2;
  const int qpY = F.qp + B.delta_qp, qpC = kChromaQp[qpY];
  float thr = F.early_skip_thr;
  if (F.encoder_speed > 1 && size == 64) thr = (float)dmul(1.3, (double)thr);
  const int bip = F.enable_bipred;
  const thor_enc_pic &r0 = ref_pic(F, P.ref_idx0);
  const int stride_y = r0.stride_y, stride_c = r0.stride_c;
  uint8_t *pb = S.ipred, *pb0 = S.p0.y, *pb1 = S.p1.y;             // compact sub-block predictions
  Mv mv = P.mv0[0];
  for (int i = 0; i < size; i += size0)
    for (int j = 0; j < size; j += size0) {
      const int offy = (B.ypos + i) * stride_y + B.xpos + j, offc = (B.ypos + i) / 2 * stride_c + (B.xpos + j) / 2;
      const uint8_t *oy = S.org.y + size * i + j, *ou = S.org.u + (size / 2) * (i / 2) + j / 2, *ov = S.org.v + (size / 2) * (i / 2) + j / 2;
      if (P.dir == 2) {
        const thor_enc_pic &r1 = ref_pic(F, P.ref_idx1);
        const int sign0 = r0.frame_num >= F.frame_num, sign1 = r1.frame_num >= F.frame_num;
        inter_luma(g, S, pb0, size0, r0.y + offy, stride_y, size0, size0, clip_mv(P.mv0[0], B.ypos, B.xpos, F.width, F.height, size0, sign0), sign0, bip);
        inter_luma(g, S, pb1, size0, r1.y + offy, stride_y, size0, size0, clip_mv(P.mv1[0], B.ypos, B.xpos, F.width, F.height, size0, sign1), sign1, bip);
        TFOR(g, k, size0 * size0) pb[k] = (uint8_t)((pb0[k] + pb1[k]) >> 1);
        g.sync();
        if (skip_test_luma(E, oy, size, size0, qpY, pb, thr)) return 0;
        // chroma of a bi-predicted candidate uses the vectors as they are (not clipped): :2683-2704
        uint8_t *c0u = S.p0.u, *c0v = S.p0.v, *c1u = S.p1.u, *c1v = S.p1.v;
        inter_chroma(g, c0u, c0v, sc0, r0.u + offc, r0.v + offc, stride_c, sc0, sc0, P.mv0[0], sign0);
        inter_chroma(g, c1u, c1v, sc0, r1.u + offc, r1.v + offc, stride_c, sc0, sc0, P.mv1[0], sign1);
        TFOR(g, k, sc0 * sc0) { c0u[k] = (uint8_t)((c0u[k] + c1u[k]) >> 1); c0v[k] = (uint8_t)((c0v[k] + c1v[k]) >> 1); }
        g.sync();
        if (skip_test_chroma(E, ou, size / 2, sc0, qpC, c0u, thr)) return 0;
        if (skip_test_chroma(E, ov, size / 2, sc0, qpC, c0v, thr)) return 0;
      } else {
        const int sign = r0.frame_num > F.frame_num;
        mv = clip_mv(mv, B.ypos, B.xpos, F.width, F.height, size0, sign);
        inter_luma(g, S, pb, size0, r0.y + offy, stride_y, size0, size0, mv, sign, bip);
        if (skip_test_luma(E, oy, size, size0, qpY, pb, thr)) return 0;
        inter_chroma(g, S.p0.u, S.p0.v, sc0, r0.u + offc, r0.v + offc, stride_c, sc0, sc0, mv, sign);
        if (skip_test_chroma(E, ou, size / 2, sc0, qpC, S.p0.u, thr)) return 0;
        if (skip_test_chroma(E, ov, size / 2, sc0, qpC, S.p0.v, thr)) return 0;
      }
    }
  return 1;
}

// ---- searches ----------------------------------------------------------------------------------------------------
// search_intra_prediction_params (enc/encode_block.c:1230-1329): luma SAD of the intra modes in the reference's order
// DC, HOR, VER, PLANAR [, UPLEFT, UPRIGHT, UPUPRIGHT, UPUPLEFT, UPLEFTLEFT, DOWNLEFTLEFT]; first minimum wins.
TDN int search_intra(Enc &E, const BlockCtx &B, int num_intra_modes, int *intra_mode) {
  PROF(PF_SEARCH_INTRA);
  THOR_LOCAL(&E); THOR_LOCAL(&B);
  THOR_SHS(&E.S); THOR_SHF(&E.F);
  Grp &g = E.g;
  SbMem &S = E.S;
  const thor_enc_frame &F = E.F;
  const int size = B.size;
  const thor_enc_pic &rec = F.rec;
  make_top_and_left(E, 0, rec.y + B.ypos * rec.stride_y + B.xpos, rec.stride_y, nullptr, 0, 0, 0, B.ypos, B.xpos, size,
                    upright_available(B.ypos, B.xpos, size, F.width), downleft_available(B.ypos, B.xpos, size, F.height), 0);
  int min_sad = 1 << 30;
  *intra_mode = 0;
  const int n = num_intra_modes == 4 ? 4 : 10;
  for (int k = 0; k < n; k++) {
    const int mode = k == 0 ? 0 : k == 1 ? 2 : k == 2 ? 3 : k == 3 ? 1 : k == 4 ? 4 : k == 5 ? 5 : k == 6 ? 6 : k == 7 ? 7 : k == 8 ? 8 : 9;
    // the search's DC always averages left and top (:1250 tests "pos >= 0"), unlike get_intra_prediction at a picture edge
    intra_predict(g, S, 0, 1, 1, size, S.ipred, mode);
    int acc = 0;
    TFOR(g, i, size * size / 4) acc += (int)sad4(reinterpret_cast<const uint32_t *>(S.org.y)[i], reinterpret_cast<const uint32_t *>(S.ipred)[i]);
    const int sad = g.sum(acc);
    if (sad < min_sad) { min_sad = sad; *intra_mode = mode; }
  }
  return min_sad;
}

// search_inter_prediction_params (enc/encode_block.c:1331-1396): one motion_estimate per prediction block of the partition
TDN int search_inter(Enc &E, const uint8_t *org_y, const thor_enc_pic &ref, const BlockCtx &B, Mv mvc, Mv mvp, Mv *mv_arr, int part, double lambda,
                     int sign, int ref_idx, int enable_bipred) {
  THOR_LOCAL(&E); THOR_LOCAL(&B);
  THOR_SHS(&E.S); THOR_SHF(&E.F); THOR_SHARED(org_y);
  Grp &g = E.g;
  SbMem &S = E.S;
  const thor_enc_frame &F = E.F;
  const int size = B.size, rstride = ref.stride_y;
  const uint8_t *ref_y = ref.y + B.ypos * rstride + B.xpos;
  const Mv *mvcand = S.mvcand[ref_idx];
  const int ncand = E.st.mvcand_num[ref_idx];
  Mv mv, mvp2 = mvp;
  int sad = 0;
  if (part == 0) {
    sad += (int)motion_estimate(g, S, F, org_y, ref_y, size, rstride, size, size, &mv, mvc, mvp2, lambda, sign, B.xpos, B.ypos, mvcand, ncand, enable_bipred);
    mv_arr[0] = mv_arr[1] = mv_arr[2] = mv_arr[3] = mv;
  } else if (part == 1) {
    for (int index = 0; index < 4; index += 2) {
      const int py = index >> 1;
      sad += (int)motion_estimate(g, S, F, org_y + py * (size / 2) * size, ref_y + py * (size / 2) * rstride, size, rstride, size, size / 2, &mv, mvc, mvp2,
                                  lambda, sign, B.xpos, B.ypos, mvcand, ncand, enable_bipred);
      mv_arr[index] = mv_arr[index + 1] = mv;
      mvp2 = mv_arr[0];
    }
  } else if (part == 2) {
    for (int index = 0; index < 2; index++) {
      sad += (int)motion_estimate(g, S, F, org_y + index * (size / 2), ref_y + index * (size / 2), size, rstride, size / 2, size, &mv, mvc, mvp2, lambda,
                                  sign, B.xpos, B.ypos, mvcand, ncand, enable_bipred);
      mv_arr[index] = mv_arr[index + 2] = mv;
      mvp2 = mv_arr[0];
    }
  } else {
    for (int index = 0; index < 4; index++) {
      const int px = index & 1, py = index >> 1;
      sad += (int)motion_estimate(g, S, F, org_y + py * (size / 2) * size + px * (size / 2), ref_y + py * (size / 2) * rstride + px * (size / 2), size,
                                  rstride, size / 2, size / 2, &mv, mvc, mvp2, lambda, sign, B.xpos, B.ypos, mvcand, ncand, enable_bipred);
      mv_arr[index] = mv;
      mvp2 = mv_arr[0];
    }
  }
  return sad;
}

// search_bipred_prediction_params (enc/encode_block.c:2047-2202) with me_mode 0: iterative uni-directional search of one
// list against the target 2*org - prediction(other list).  (The me_mode 1 search -- mv0 = -mv1 on B frames at
// encoder_speed 0 -- is motion_estimate_bi in enc_me.h, called from mode_decision.)
TDN void search_bipred(Enc &E, const BlockCtx &B, int part, const Mv *mv_center, Mv mvp, int *ref_idx0, int *ref_idx1, Mv *mv_arr0, Mv *mv_arr1) {
  PROF(PF_BIPRED);
  THOR_LOCAL(&E); THOR_LOCAL(&B);
  THOR_SHS(&E.S); THOR_SHF(&E.F);
  Grp &g = E.g;
  SbMem &S = E.S;
  const thor_enc_frame &F = E.F;
  const int size = B.size;
  const double lambda = dsqrt(F.lambda);
  const int num_iter = F.encoder_speed == 0 ? 2 : 1;
  int min_ref_idx0 = (F.frame_type == 2 && F.interp_ref == 1) ? 1 : 0, min_ref_idx1 = 0;
  Mv min0[4], min1[4], mv_all[4];
  for (int i = 0; i < 4; i++) min0[i] = min1[i] = mvp;
  int min_sad = 1 << 30;
  for (int n = 0; n < num_iter; n++) {
    const int stop = part == 0 ? 0 : 1;
    for (int list = 1; list >= stop; list--) {
      const Mv mv = list ? min0[0] : min1[0];
      int ref_idx = list ? min_ref_idx0 : min_ref_idx1;
      {
        const thor_enc_pic &ref = ref_pic(F, ref_idx);
        inter_yuv(g, S, F, ref, S.pred, B, list ? min0 : min1, ref.frame_num > F.rec.frame_num, 1, 1);
      }
      TFOR(g, i, size * size) S.org8[i] = (uint8_t)clip255(2 * (int)S.org.y[i] - (int)S.pred.y[i]);
      g.sync();
      int ref_start, ref_end;
      if (F.frame_type == 1) { ref_start = 0; ref_end = F.num_ref - 1; }
      else { ref_start = ref_end = (list ? 1 : 0) + (F.interp_ref ? 1 : 0); }
      for (ref_idx = ref_start; ref_idx <= ref_end; ref_idx++) {
        const thor_enc_pic &ref = ref_pic(F, ref_idx);
        const int sign = ref.frame_num > F.rec.frame_num;
        const Mv mvp2 = (F.frame_type == 2 && list == 1) ? mv : mvp;
        const int sad = search_inter(E, S.org8, ref, B, mv_center[ref_idx], mvp2, mv_all, part, lambda, sign, ref_idx, 1);
        for (int i = 0; i < 4; i++) add_mvcandidate(g, S, E.st, ref_idx, mv_all[i]);
        if (sad < min_sad) {
          min_sad = sad;
          if (list) { min_ref_idx1 = ref_idx; for (int i = 0; i < 4; i++) min1[i] = mv_all[i]; }
          else { min_ref_idx0 = ref_idx; for (int i = 0; i < 4; i++) min0[i] = mv_all[i]; }
        }
      }
    }
  }
  *ref_idx0 = min_ref_idx0;
  *ref_idx1 = min_ref_idx1;
  for (int i = 0; i < 4; i++) { mv_arr0[i] = min0[i]; mv_arr1[i] = min1[i]; }
}

// ---- mode decision (enc/encode_block.c:2204-2479) ---------------------------------------------------------------
TD void try_cost(Enc &E, const BlockCtx &B, Param &P, Param &best, uint32_t &min_cost, int width, int height, uint32_t *cost_out) {
  THOR_LOCAL(&E); THOR_LOCAL(&B); THOR_LOCAL(&P); THOR_LOCAL(&best);
  THOR_SHS(&E.S); THOR_SHF(&E.F);
  const int nbits = encode_block(E, B, P);
  const uint32_t cost = rd_cost(ssd_yuv(E.g, E.S.org, E.S.rec, B.size, width, height), nbits, E.F.lambda);
  if (cost < min_cost) { min_cost = cost; keep_best(E, B, P, best); }
  if (cost_out) *cost_out = cost;
}
TD void param_from_cand(Param &P, const InterPred &c, int idx, int mode) {
  THOR_LOCAL(&P);
  P.mode = mode;
  P.skip_idx = idx;
  P.ref_idx0 = c.ref_idx0;
  P.ref_idx1 = c.ref_idx1;
  for (int i = 0; i < 4; i++) { P.mv0[i] = c.mv0; P.mv1[i] = c.mv1; }
  P.dir = c.dir;
}

TDN uint32_t mode_decision(Enc &E, BlockCtx &B, Param &best) {
  PROF(PF_MODE);
  THOR_LOCAL(&E); THOR_LOCAL(&B); THOR_LOCAL(&best);
  THOR_SHS(&E.S); THOR_SHF(&E.F);
  Grp &g = E.g;
  SbMem &S = E.S;
  const thor_enc_frame &F = E.F;
  const int size = B.size, speed = F.encoder_speed;
  const double lambda = F.lambda;
  int do_inter = 1, do_intra = 1;
  const int rectangular = B.bwidth != size || B.bheight != size;
  const int intra_inter_sad = speed > 0;
  uint32_t min_cost = 1u << 31, sad_intra = 1u << 31, sad_inter;
  Param P;
  P.mode = 0; P.intra_mode = 0; P.skip_idx = 0; P.pb_part = 0; P.ref_idx0 = P.ref_idx1 = 0; P.dir = 0;
  P.cbp_y = P.cbp_u = P.cbp_v = 0; P.tb_param = 0; P.tb_split = 0;
  for (int i = 0; i < 4; i++) P.mv0[i] = P.mv1[i] = mkmv(0, 0);

  if (F.frame_type != 0) {                                             // skip candidates (:2251-2279)
    P.tb_param = 0;
    P.pb_part = 0;
    for (int idx = 0; idx < B.num_skip_vec; idx++) {
      param_from_cand(P, B.skip_cand[idx], idx, 0);
      try_cost(E, B, P, best, min_cost, B.bwidth, B.bheight, nullptr);
    }
  }
  if (!rectangular && size <= 64) {
    if (F.frame_type != 0) {
      P.tb_param = 0;
      for (int idx = 0; idx < B.num_merge_vec; idx++) {                // merge candidates (:2283-2304)
        param_from_cand(P, B.merge_cand[idx], idx, 4);
        try_cost(E, B, P, best, min_cost, size, size, nullptr);
      }
      int intra_mode = 0;
      if (intra_inter_sad) {
        sad_intra = (uint32_t)search_intra(E, B, F.num_intra_modes, &intra_mode);
        sad_intra += (uint32_t)(int)dadd(dmul(dsqrt(lambda), 2.0), 0.5);
      }
      int min_idx, max_idx;
      if (E.st.best_ref < 0 || speed < 2 || F.enable_bipred) { min_idx = 0; max_idx = F.num_ref - 1; }
      else min_idx = max_idx = E.st.best_ref;
      int worst_cost = 0, best_cost = (int)(1u << 31);
      Mv mvp = mkmv(0, 0), mv_all[4][4], mv_center[THOR_ENC_MAX_REF];
      for (int ref_idx = min_idx; ref_idx <= max_idx; ref_idx++) {     // uni-directional inter (:2319-2375)
        const thor_enc_pic &ref = ref_pic(F, ref_idx);
        P.ref_idx0 = P.ref_idx1 = ref_idx;
        mvp = get_mv_pred(F, B.ypos, B.xpos, size);
        add_mvcandidate(g, S, E.st, ref_idx, mvp);
        B.mvp = mvp;
        const int sign = ref.frame_num >= F.rec.frame_num;
        mv_center[ref_idx] = mvp;
        sad_inter = 1u << 31;
        for (int part = 0; part < B.max_num_pb_part; part++) {
          const uint32_t sad = (uint32_t)search_inter(E, S.org.y, ref, B, mv_center[ref_idx], mvp, mv_all[part], part, dsqrt(lambda), sign, ref_idx, F.enable_bipred);
          for (int i = 0; i < 4; i++) add_mvcandidate(g, S, E.st, ref_idx, mv_all[part][i]);
          mv_center[ref_idx] = mv_all[0][0];
          sad_inter = sad_inter < sad ? sad_inter : sad;
        }
        if (intra_inter_sad) {
          do_inter = sad_inter < sad_intra;
          if (sad_inter < sad_intra) do_intra = 0;
        }
        if (do_inter) {
          for (int part = 0; part < B.max_num_pb_part; part++) {
            P.pb_part = part;
            for (int i = 0; i < 4; i++) P.mv0[i] = P.mv1[i] = mv_all[part][i];
            P.mode = 2;
            for (int tb_param = speed < 1 ? -1 : 0; tb_param <= B.max_num_tb_part - 1; tb_param++) {
              P.tb_param = tb_param;
              uint32_t cost;
              try_cost(E, B, P, best, min_cost, size, size, &cost);
              worst_cost = (uint32_t)worst_cost > cost ? worst_cost : (int)cost;
              best_cost = (uint32_t)best_cost < cost ? best_cost : (int)cost;
            }
          }
        }
      }
      // "one reference convincingly better": the reference's best_ref_idx is never anything but 0 (:2230, :2377)
      if (worst_cost && worst_cost * 3 > best_cost * 4) E.st.best_ref = 0;

      if (F.num_ref > 1 && F.enable_bipred && do_inter) {              // bi-prediction (:2380-2427)
        int r0, r1;
        Mv a0[4], a1[4];
        search_bipred(E, B, 0, mv_center, mvp, &r0, &r1, a0, a1);
        P.mode = 3;
        P.pb_part = 0;
        P.ref_idx0 = r0;
        P.ref_idx1 = r1;
        for (int i = 0; i < 4; i++) { P.mv0[i] = a0[i]; P.mv1[i] = a1[i]; }
        P.tb_param = 0;
        try_cost(E, B, P, best, min_cost, size, size, nullptr);
        if (F.frame_type == 2 && speed == 0) {
          // search_bipred_prediction_params with me_mode 1 (:2079-2110): mv0 = -mv1 through the two references' signs
          const int i0 = F.interp_ref ? 1 : 0, i1 = F.interp_ref ? 2 : 1;
          const thor_enc_pic &q0 = ref_pic(F, i0), &q1 = ref_pic(F, i1);
          const int off = B.ypos * q0.stride_y + B.xpos;
          Mv mv;
          motion_estimate_bi(g, S, F, S.org.y, q0.y + off, q1.y + off, size, q0.stride_y, &mv, mv_center[i0], mvp, dsqrt(lambda), 0, B.xpos, B.ypos,
                             S.mvcand[i0], E.st.mvcand_num[i0], 2);
          P.pb_part = 0;
          P.ref_idx0 = i0;
          P.ref_idx1 = i1;
          for (int i = 0; i < 4; i++) P.mv0[i] = P.mv1[i] = mv;
          P.tb_param = 0;
          try_cost(E, B, P, best, min_cost, size, size, nullptr);
        }
      }
    }
    if (do_intra) {                                                    // intra (:2431-2472)
      const int max_tb_param = B.max_num_tb_part - 1;
      int intra_mode = 0;
      P.mode = 1;
      if (F.intra_rdo) {
        uint32_t min_intra_cost = 1u << 31;
        for (int m = 0; m < F.num_intra_modes; m++) {
          P.intra_mode = m;
          for (int tb_param = 0; tb_param <= max_tb_param; tb_param++) {
            P.tb_param = tb_param;
            const int nbits = encode_block(E, B, P);
            const uint32_t cost = rd_cost(ssd_yuv(g, S.org, S.rec, size, size, size), nbits, lambda);
            if (cost < min_intra_cost) { min_intra_cost = cost; intra_mode = m; }
          }
        }
      } else {
        search_intra(E, B, F.num_intra_modes, &intra_mode);
      }
      P.intra_mode = intra_mode;
      for (int tb_param = 0; tb_param <= max_tb_param; tb_param++) {
        P.tb_param = tb_param;
        try_cost(E, B, P, best, min_cost, size, size, nullptr);
      }
    }
  }
  return min_cost;
}

// ---- the quad tree (enc/encode_block.c:2787-3033) ---------------------------------------------------------------
// What the final encode_block + copy_block_to_frame + copy_deblock_data (:1802-1819, :1947-1981) leave behind, plus the
// record the host's write_block needs: reconstruction -> picture, neighbour units, node, coefficients.
TDN void finalize_block(Enc &E, const BlockCtx &B, const Param &P) {
  PROF(PF_FINAL);
  THOR_LOCAL(&E); THOR_LOCAL(&B); THOR_LOCAL(&P);
  THOR_SHS(&E.S); THOR_SHF(&E.F);
  Grp &g = E.g;
  SbMem &S = E.S;
  const thor_enc_frame &F = E.F;
  const int size = B.size, z = zidx8(B.ypos, B.xpos);
  g.sync();
  {                                                                    // copy_block_to_frame
    const thor_enc_pic &rec = F.rec;
    const int wq = B.bwidth / 4;
    const Div dwq = mkdiv(wq);
    TFOR(g, i, B.bheight * wq) {
      const int r = dq(dwq, i), c = dr(dwq, i) * 4;
      *reinterpret_cast<uint32_t *>(rec.y + (B.ypos + r) * rec.stride_y + B.xpos + c) = *reinterpret_cast<const uint32_t *>(S.best.y + r * size + c);
    }
    const int wc = B.bwidth / 2, sc = size / 2;
    const Div dwc = mkdiv(wc);
    TFOR(g, i, (B.bheight / 2) * wc) {
      const int r = dq(dwc, i), c = dr(dwc, i), off = (B.ypos / 2 + r) * rec.stride_c + B.xpos / 2 + c;
      rec.u[off] = S.best.u[r * sc + c];
      rec.v[off] = S.best.v[r * sc + c];
    }
  }
  const int tb_split = imax(0, P.tb_param);
  {                                                                    // copy_deblock_data
    const int uw = B.bwidth / 4, div = size / 8, ustride = F.width / 4;
    const Div duw = mkdiv(uw);
    TFOR(g, i, (B.bheight / 4) * uw) {
      const int m = dq(duw, i), n = dr(duw, i), index = 2 * p2div(m, div) + p2div(n, div);
      thor_enc_unit u;
      u.mv0x = P.mv0[index].x; u.mv0y = P.mv0[index].y; u.mv1x = P.mv1[index].x; u.mv1y = P.mv1[index].y;
      u.mode = (uint8_t)P.mode;
      u.size = (uint8_t)size;
      u.cbp = tb_split ? 7 : (uint8_t)((P.cbp_y ? 1 : 0) | (P.cbp_u ? 2 : 0) | (P.cbp_v ? 4 : 0));   // :1785-1788
      u.flags = (uint8_t)(tb_split | ((P.mode == 2 ? P.pb_part : 0) << 1));
      u.ref_idx0 = (int8_t)P.ref_idx0; u.ref_idx1 = (int8_t)P.ref_idx1; u.dir = (int8_t)P.dir;
      u.pad = 0;
      st_unit(F.units + (B.ypos / 4 + m) * ustride + B.xpos / 4 + n, u);
    }
  }
  if (P.mode != 0) {                                                   // coefficients where write_block reads them
    const int ny = size * size / 2, nc = ny / 4;                       // 32-bit words
    if (P.cbp_y) TFOR(g, i, ny) reinterpret_cast<uint32_t *>(E.coef->y + 64 * z)[i] = reinterpret_cast<const uint32_t *>(S.bc.y)[i];
    if (P.cbp_u) TFOR(g, i, nc) reinterpret_cast<uint32_t *>(E.coef->u + 16 * z)[i] = reinterpret_cast<const uint32_t *>(S.bc.u)[i];
    if (P.cbp_v) TFOR(g, i, nc) reinterpret_cast<uint32_t *>(E.coef->v + 16 * z)[i] = reinterpret_cast<const uint32_t *>(S.bc.v)[i];
  }
  if (g.tid == 0) {
    thor_enc_node nd;
    nd.state = THOR_NODE_CODED;
    nd.mode = (uint8_t)P.mode; nd.intra_mode = (uint8_t)P.intra_mode; nd.skip_idx = (uint8_t)P.skip_idx; nd.pb_part = (uint8_t)P.pb_part;
    nd.tb_param = (int8_t)P.tb_param; nd.ref_idx0 = (int8_t)P.ref_idx0; nd.ref_idx1 = (int8_t)P.ref_idx1; nd.dir = (int8_t)P.dir;
    nd.cbp_y = (uint8_t)P.cbp_y; nd.cbp_u = (uint8_t)P.cbp_u; nd.cbp_v = (uint8_t)P.cbp_v;
    nd.delta_qp = (int8_t)B.delta_qp;
    nd.num_skip_vec = (uint8_t)B.num_skip_vec; nd.num_merge_vec = (uint8_t)B.num_merge_vec;
    nd.ctx_index = (int8_t)B.ctx_index; nd.ctx_cbp = (int8_t)B.ctx_cbp;
    nd.split_early = 0; nd.pad[0] = nd.pad[1] = 0;
    nd.mvp[0] = B.mvp.x; nd.mvp[1] = B.mvp.y;
    for (int i = 0; i < 4; i++) { nd.mv0[i][0] = P.mv0[i].x; nd.mv0[i][1] = P.mv0[i].y; nd.mv1[i][0] = P.mv1[i].x; nd.mv1[i][1] = P.mv1[i].y; }
    E.nodes[node_index(size, B.ypos, B.xpos)] = nd;
  }
  g.sync();
}

TD void mark_split(Enc &E, const BlockCtx &B) {
  THOR_LOCAL(&E); THOR_LOCAL(&B);
  THOR_SHS(&E.S); THOR_SHF(&E.F);
  if (E.g.tid == 0) {
    thor_enc_node &nd = E.nodes[node_index(B.size, B.ypos, B.xpos)];
    nd.state = THOR_NODE_SPLIT;
    nd.ctx_index = (int8_t)B.ctx_index;
    nd.delta_qp = (int8_t)B.delta_qp;
  }
}

// The parts of process_block that do not depend on the block size as a compile-time constant live in functions of their own:
// process_block is instantiated four times (64 / 32 / 16 / 8) and the kernel is bound by instruction fetch.
TDN void block_begin(Enc &E, BlockCtx &B, Param &best, int size, int ypos, int xpos, int qp) {
  THOR_LOCAL(&E); THOR_LOCAL(&B); THOR_LOCAL(&best);
  const thor_enc_frame &F = E.F;
  B.size = size; B.ypos = ypos; B.xpos = xpos;
  B.bwidth = imin(size, F.width - xpos);
  B.bheight = imin(size, F.height - ypos);
  B.max_num_tb_part = F.enable_tb_split == 1 ? 2 : 1;
  B.max_num_pb_part = F.enable_pb_split ? 4 : 1;
  B.delta_qp = qp - F.qp;
  B.mvp = mkmv(0, 0);
  B.num_skip_vec = B.num_merge_vec = 0;
  find_block_contexts(F, ypos, xpos, size, B);
  if (F.frame_type != 0) {
    B.num_skip_vec = get_mv_skip(F, ypos, xpos, size, B.skip_cand);
    B.num_merge_vec = get_mv_skip(F, ypos, xpos, size, B.merge_cand);  // get_mv_merge is the same function under LIMITED_SKIP
  }
  best.mode = 0; best.intra_mode = 0; best.skip_idx = 0; best.pb_part = 0; best.ref_idx0 = best.ref_idx1 = 0; best.dir = 0;
  best.cbp_y = best.cbp_u = best.cbp_v = 0; best.tb_param = 0; best.tb_split = 0;
  for (int i = 0; i < 4; i++) best.mv0[i] = best.mv1[i] = mkmv(0, 0);
}

// search_early_skip_candidates (:2743-2783): true when a skip candidate passed the test; the block is then finished
TDN bool early_skip_block(Enc &E, BlockCtx &B, Param &best, uint32_t *cost_out) {
  THOR_LOCAL(&E); THOR_LOCAL(&B); THOR_LOCAL(&best);
  load_org(E, B);
  uint32_t min_cost = 1u << 31;
  int early = 0;
  Param P = best;
  for (int idx = 0; idx < B.num_skip_vec; idx++) {
    param_from_cand(P, B.skip_cand[idx], idx, 0);
    P.tb_param = 0;
    if (check_early_skip(E, B, P)) {
      early = 1;
      try_cost(E, B, P, best, min_cost, B.size, B.size, nullptr);
    }
  }
  if (early) finalize_block(E, B, best);
  *cost_out = min_cost;
  return early != 0;
}

// the decision of a 16x16 block waits while its four children are tried (top-down split, enc/encode_block.c:2983-2995)
TDN void keep_16(Enc &E, int restore) {
  THOR_LOCAL(&E);
  Grp &g = E.g;
  SbMem &S = E.S;
  g.sync();
  if (!restore) {
    TFOR(g, i, 256) { S.keep_y[i] = S.best.y[i]; S.keep_cy[i] = S.bc.y[i]; }
    TFOR(g, i, 64) { S.keep_u[i] = S.best.u[i]; S.keep_v[i] = S.best.v[i]; S.keep_cu[i] = S.bc.u[i]; S.keep_cv[i] = S.bc.v[i]; }
  } else {
    TFOR(g, i, 256) { S.best.y[i] = S.keep_y[i]; S.bc.y[i] = S.keep_cy[i]; }
    TFOR(g, i, 64) { S.best.u[i] = S.keep_u[i]; S.best.v[i] = S.keep_v[i]; S.bc.u[i] = S.keep_cu[i]; S.bc.v[i] = S.keep_cv[i]; }
  }
  g.sync();
}

// mode_decision of a whole or picture-edge block and, if it beats the smaller blocks, its final encode
TDN uint32_t decide_block(Enc &E, BlockCtx &B, Param &best) {
  THOR_LOCAL(&E); THOR_LOCAL(&B); THOR_LOCAL(&best);
  load_org(E, B);
  return mode_decision(E, B, best);
}
TDN void split_mark(Enc &E, const BlockCtx &B) { mark_split(E, B); }

template <int SIZE>
TDN int process_block(Enc &E, int ypos, int xpos, int qp) {
  THOR_LOCAL(&E);
  THOR_SHS(&E.S); THOR_SHF(&E.F);
  const thor_enc_frame &F = E.F;
  if (ypos >= F.height || xpos >= F.width) return 0;
  const int encode_this_size = ypos + SIZE <= F.height && xpos + SIZE <= F.width;
  const int encode_smaller_size = SIZE > 8 * ((encode_this_size && F.frame_type != 0 && F.encoder_speed > 0) ? 2 : 1);
  const int top_down = !encode_smaller_size && SIZE > 8;
  const int encode_rectangular_size = !encode_this_size && F.frame_type != 0;
  if (!encode_this_size && !encode_smaller_size) return 0;
  uint32_t cost_small = 1u << 28, cost = 1u << 28;

  BlockCtx B;
  Param best;
  block_begin(E, B, best, SIZE, ypos, xpos, qp);

  if (encode_this_size && F.frame_type != 0 && F.early_skip_thr > 0.0f) {
    uint32_t min_cost;
    if (early_skip_block(E, B, best, &min_cost)) return (int)min_cost;
  }

  if (encode_smaller_size) {
    split_mark(E, B);
    if constexpr (SIZE > 8) {
      constexpr int NS = SIZE / 2;
      cost_small = 0;
      cost_small += (uint32_t)process_block<NS>(E, ypos, xpos, qp);
      cost_small += (uint32_t)process_block<NS>(E, ypos + NS, xpos, qp);
      cost_small += (uint32_t)process_block<NS>(E, ypos, xpos + NS, qp);
      cost_small += (uint32_t)process_block<NS>(E, ypos + NS, xpos + NS, qp);
    }
  }

  if (encode_this_size) {
    cost = decide_block(E, B, best);
    const int me_threshold = SIZE * SIZE * kIq8x8[qp] / 8;
    bool kept = false;
    if (top_down && cost > (uint32_t)me_threshold) {
      if constexpr (SIZE == 16) {
        keep_16(E, 0);
        kept = true;
        cost_small = 0;
        cost_small += (uint32_t)process_block<8>(E, ypos, xpos, qp);
        cost_small += (uint32_t)process_block<8>(E, ypos + 8, xpos, qp);
        cost_small += (uint32_t)process_block<8>(E, ypos, xpos + 8, qp);
        cost_small += (uint32_t)process_block<8>(E, ypos + 8, xpos + 8, qp);
      }
    }
    if (cost <= cost_small) {
      if (kept) keep_16(E, 1);
      finalize_block(E, B, best);
    } else {
      split_mark(E, B);
    }
  } else if (encode_rectangular_size) {
    cost = decide_block(E, B, best);
    if (cost <= cost_small) finalize_block(E, B, best);
  }
  return (int)(cost < cost_small ? cost : cost_small);
}

// one superblock of encode_frame's loop (enc/encode_frame.c:112-148), including the per-superblock delta-qp search
TDN void encode_superblock(Enc &E) {
  PROF(PF_SB);
  THOR_LOCAL(&E);
  THOR_SHS(&E.S); THOR_SHF(&E.F);
  const thor_enc_frame &F = E.F;
  for (int r = 0; r < THOR_ENC_MAX_REF; r++) { E.st.mvcand_num[r] = 0; E.st.mvcand_mask[r] = 0; }
  E.st.best_ref = -1;
  const int qp = F.qp;
  if (F.max_delta_qp) {
    int min_cost = 1 << 30, best_qp = qp;
    for (int qp0 = qp - F.max_delta_qp; qp0 <= qp + F.max_delta_qp; qp0 += F.delta_qp_step) {
      const int cost = process_block<64>(E, E.sb_y, E.sb_x, qp0);
      if (cost < min_cost) { min_cost = cost; best_qp = qp0; }
    }
    process_block<64>(E, E.sb_y, E.sb_x, best_qp);
  } else {
    process_block<64>(E, E.sb_y, E.sb_x, qp);
  }
}

}  // namespace thorenc
