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
  Grp &g = E.g;
  const thor_enc_pic &o = E.F.orig;
  const int size = B.size, wq = B.bwidth / 4;
  g.sync();
  TFOR(g, i, B.bheight * wq) {
    const int r = i / wq, c = (i % wq) * 4;
    *reinterpret_cast<uint32_t *>(E.S.org.y + r * size + c) = ld_ro32u(o.y + (B.ypos + r) * o.stride_y + B.xpos + c);
  }
  const int wc = B.bwidth / 2, sc = size / 2;
  TFOR(g, i, (B.bheight / 2) * wc) {
    const int r = i / wc, c = i % wc;
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
