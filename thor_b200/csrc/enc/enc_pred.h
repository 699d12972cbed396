// Frame encoder of libthorcuda -- inter and intra prediction of one block into shared memory.
#pragma once
#include "enc_state.h"

namespace thorenc {

// clip_mv (enc/encode_block.c:816-828): keeps the block within 80 samples of the picture; the right-edge test omits
// "+ size" and "/ 4" truncates toward zero, exactly as there.
TD Mv clip_mv(Mv mv, int ypos, int xpos, int fwidth, int fheight, int size, int sign) {
  const int ext = 96 - 16;
  int mvy = sign ? -mv.y : mv.y, mvx = sign ? -mv.x : mv.x;
  if (ypos + mvy / 4 < -ext) mvy = 4 * (-ext - ypos);
  if (ypos + mvy / 4 + size > fheight + ext) mvy = 4 * (fheight + ext - ypos - size);
  if (xpos + mvx / 4 < -ext) mvx = 4 * (-ext - xpos);
  if (xpos + mvx / 4 > fwidth + ext) mvx = 4 * (fwidth + ext - xpos - size);
  return mkmv(sign ? -mvx : mvx, sign ? -mvy : mvy);
}

// get_inter_prediction_luma (common/inter_prediction.c:120-180).  ref: block origin in the padded reference plane.
// Writes width x height samples at dst (shared, stride pstride, 4-byte aligned rows).  Uses S.itmp.  Ends with a barrier.
TDN void inter_luma(Grp &g, SbMem &S, uint8_t *dst, int pstride, const uint8_t *ref, int stride, int width, int height, Mv mv, int sign,
                    int bipred) {
  THOR_LOCAL(&g);
  THOR_SHS(&S); THOR_SHARED(dst);
  const int mx = sign ? -mv.x : mv.x, my = sign ? -mv.y : mv.y;
  const int fx = mx & 3, fy = my & 3;
  const uint8_t *src = ref + (my >> 2) * stride + (mx >> 2);
  const int wq = width >> 2;                              // 4-sample groups per row (width is a multiple of 4)
  const Div dwq = mkdiv(wq);
  if (!fx && !fy) {
    TFOR(g, i, height * wq) {
      const int r = dq(dwq, i), c = dr(dwq, i) * 4;
      *reinterpret_cast<uint32_t *>(dst + r * pstride + c) = ld_ro32u(src + r * stride + c);
    }
    g.sync();
    return;
  }
  if (fx == 2 && fy == 2) {
    // centre position: [0 1 1 0; 1 2 2 1; 1 2 2 1; 0 1 1 0] / 16 (:146-160)
    TFOR(g, i, height * wq) {
      const int r = dq(dwq, i), c = dr(dwq, i) * 4;
      int acc[4] = {0, 0, 0, 0};
#pragma unroll
      for (int dy = 0; dy < 4; dy++) {
        const uint8_t *p = src + (r + dy - 1) * stride + c - 1;
        const uint32_t w0 = ld_ro32u(p), w1 = ld_ro32u(p + 4);
        int b[7];
#pragma unroll
        for (int k = 0; k < 4; k++) b[k] = (w0 >> (8 * k)) & 255;
#pragma unroll
        for (int k = 0; k < 3; k++) b[4 + k] = (w1 >> (8 * k)) & 255;
        const int wy = (dy == 0 || dy == 3) ? 0 : 1;      // rows 0,3: [0 1 1 0]; rows 1,2: [1 2 2 1]
#pragma unroll
        for (int k = 0; k < 4; k++) acc[k] += wy ? (b[k] + 2 * b[k + 1] + 2 * b[k + 2] + b[k + 3]) : (b[k + 1] + b[k + 2]);
      }
      uint32_t out = 0;
#pragma unroll
      for (int k = 0; k < 4; k++) out |= (uint32_t)clip255((acc[k] + 8) >> 4) << (8 * k);
      *reinterpret_cast<uint32_t *>(dst + r * pstride + c) = out;
    }
    g.sync();
    return;
  }
  const int8_t *kx = bipred ? kLumaBi[fx] : kLumaUni[fx], *ky = bipred ? kLumaBi[fy] : kLumaUni[fy];
  const int k0 = kx[0], k1 = kx[1], k2 = kx[2], k3 = kx[3], k4 = kx[4], k5 = kx[5];
  // pass 1: horizontal 6-tap at full precision for rows -2 .. height+2 (the reference filters vertically first; without
  // intermediate rounding the order is immaterial)
  const int rows = height + 5;
  TFOR(g, i, rows * wq) {
    const int r = dq(dwq, i), c = dr(dwq, i) * 4;
    const uint8_t *p = src + (r - 2) * stride + c - 2;
    const uint32_t w0 = ld_ro32u(p), w1 = ld_ro32u(p + 4), w2 = ld_ro32u(p + 8);
    int b[9];
#pragma unroll
    for (int k = 0; k < 4; k++) { b[k] = (w0 >> (8 * k)) & 255; b[4 + k] = (w1 >> (8 * k)) & 255; }
    b[8] = w2 & 255;
#pragma unroll
    for (int k = 0; k < 4; k++)
      S.itmp[r * 64 + c + k] = (int16_t)(k0 * b[k] + k1 * b[k + 1] + k2 * b[k + 2] + k3 * b[k + 3] + k4 * b[k + 4] + k5 * b[k + 5]);
  }
  g.sync();
  const int v0 = ky[0], v1 = ky[1], v2 = ky[2], v3 = ky[3], v4 = ky[4], v5 = ky[5];
  TFOR(g, i, height * wq) {
    const int r = dq(dwq, i), c = dr(dwq, i) * 4;
    uint32_t out = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const int16_t *t = S.itmp + r * 64 + c + k;
      const int acc = v0 * t[0] + v1 * t[64] + v2 * t[128] + v3 * t[192] + v4 * t[256] + v5 * t[320];
      out |= (uint32_t)clip255((acc + 2048) >> 12) << (8 * k);
    }
    *reinterpret_cast<uint32_t *>(dst + r * pstride + c) = out;
  }
  g.sync();
}

// get_inter_prediction_chroma (common/inter_prediction.c:72-118): both chroma planes of one block in one pass.
// width, height: chroma size (>= 2).  Ends with a barrier.
TDN void inter_chroma(Grp &g, uint8_t *dst_u, uint8_t *dst_v, int pstride, const uint8_t *ref_u, const uint8_t *ref_v, int stride, int width,
                      int height, Mv mv, int sign) {
  THOR_LOCAL(&g);
  THOR_SHARED(dst_u); THOR_SHARED(dst_v);
  const int mx = sign ? -mv.x : mv.x, my = sign ? -mv.y : mv.y;
  const int fx = mx & 7, fy = my & 7;
  const int off = (my >> 3) * stride + (mx >> 3);
  const int n = width * height;
  const Div dw = mkdiv(width);
  if (!fx && !fy) {
    TFOR(g, i, 2 * n) {
      const int pl = i >= n, j = pl ? i - n : i, r = dq(dw, j), c = dr(dw, j);
      (pl ? dst_v : dst_u)[r * pstride + c] = ld_ro8((pl ? ref_v : ref_u) + off + r * stride + c);
    }
    g.sync();
    return;
  }
  const int8_t *kx = kChroma[fx], *ky = kChroma[fy];
  const int h0 = kx[0], h1 = kx[1], h2 = kx[2], h3 = kx[3];
  const int v0 = ky[0], v1 = ky[1], v2 = ky[2], v3 = ky[3];
  TFOR(g, i, 2 * n) {
    const int pl = i >= n, j = pl ? i - n : i, r = dq(dw, j), c = dr(dw, j);
    const uint8_t *p = (pl ? ref_v : ref_u) + off + (r - 1) * stride + c - 1;
    int acc = 0;
#pragma unroll
    for (int m = 0; m < 4; m++) {
      const uint8_t *q = p + m * stride;
      const int row = h0 * ld_ro8(q) + h1 * ld_ro8(q + 1) + h2 * ld_ro8(q + 2) + h3 * ld_ro8(q + 3);
      acc += (m == 0 ? v0 : m == 1 ? v1 : m == 2 ? v2 : v3) * row;
    }
    (pl ? dst_v : dst_u)[r * pstride + c] = (uint8_t)clip255((acc + 2048) >> 12);
  }
  g.sync();
}

// get_inter_prediction_yuv (enc/encode_block.c:1534-1567): the prediction of a whole coding block from one reference.
// `split` is the reference's argument (0: one vector for the block; 1: one per quadrant).
TDN void inter_yuv(Grp &g, SbMem &S, const thor_enc_frame &F, const thor_enc_pic &ref, Blk &dst, const BlockCtx &B, const Mv *mv_arr, int sign,
                   int bipred, int split) {
  PROF(PF_PRED_INTER);
  THOR_SHARED(&dst);
  THOR_LOCAL(&B); THOR_LOCAL(&g);
  THOR_SHS(&S); THOR_SHF(&F);
  const int div = split + 1, bwidth = B.bwidth / div, bheight = B.bheight / div, pstride = B.size;
  const int yC = B.ypos / 2, xC = B.xpos / 2;
  const uint8_t *ref_y = ref.y + B.ypos * ref.stride_y + B.xpos;
  const uint8_t *ref_u = ref.u + yC * ref.stride_c + xC, *ref_v = ref.v + yC * ref.stride_c + xC;
  // quadrants that share one vector are predicted in one go (the result does not depend on how the block is cut)
  const bool same = !split || (mv_eq(mv_arr[0], mv_arr[1]) && mv_eq(mv_arr[0], mv_arr[2]) && mv_eq(mv_arr[0], mv_arr[3]));
  if (same) {
    const Mv mv = clip_mv(mv_arr[0], B.ypos, B.xpos, F.width, F.height, B.size, sign);
    inter_luma(g, S, dst.y, pstride, ref_y, ref.stride_y, B.bwidth, B.bheight, mv, sign, bipred);
    inter_chroma(g, dst.u, dst.v, pstride / 2, ref_u, ref_v, ref.stride_c, B.bwidth / 2, B.bheight / 2, mv, sign);
    return;
  }
  for (int index = 0; index < 4; index++) {
    const int idx = index & 1, idy = index >> 1;
    const int offpY = idy * bheight * pstride + idx * bwidth, offpC = idy * bheight * pstride / 4 + idx * bwidth / 2;
    const int offrY = idy * bheight * ref.stride_y + idx * bwidth, offrC = idy * bheight * ref.stride_c / 2 + idx * bwidth / 2;
    const Mv mv = clip_mv(mv_arr[index], B.ypos, B.xpos, F.width, F.height, B.size, sign);
    inter_luma(g, S, dst.y + offpY, pstride, ref_y + offrY, ref.stride_y, bwidth, bheight, mv, sign, bipred);
    inter_chroma(g, dst.u + offpC, dst.v + offpC, pstride / 2, ref_u + offrC, ref_v + offrC, ref.stride_c, bwidth / 2, bheight / 2, mv, sign);
  }
}

// average_blocks_all (enc/encode_block.c:1569-1588): truncating mean of two predictions
TDN void average_blocks(Grp &g, Blk &dst, const Blk &a, const Blk &b, const BlockCtx &B) {
  THOR_SHARED(&dst); THOR_SHARED(&a); THOR_SHARED(&b);
  THOR_LOCAL(&B); THOR_LOCAL(&g);
  const int size = B.size, wq = B.bwidth / 4, ny = B.bheight * wq;
  const Div dwq = mkdiv(wq);
  TFOR(g, i, ny) {
    const int o = dq(dwq, i) * size + dr(dwq, i) * 4;
    *reinterpret_cast<uint32_t *>(dst.y + o) = havg4(*reinterpret_cast<const uint32_t *>(a.y + o), *reinterpret_cast<const uint32_t *>(b.y + o));
  }
  const int wc = B.bwidth / 2, nc = (B.bheight / 2) * wc, sc = size / 2;
  const Div dwc = mkdiv(wc);
  TFOR(g, i, nc) {
    const int o = dq(dwc, i) * sc + dr(dwc, i);
    dst.u[o] = (uint8_t)((a.u[o] + b.u[o]) >> 1);
    dst.v[o] = (uint8_t)((a.v[o] + b.v[o]) >> 1);
  }
  g.sync();
}

// ---- intra prediction (common/intra_prediction.c:39-389) --------------------------------------------------------
// make_top_and_left for the non-split case (:72-99): neighbours of the size x size block at (x, y) of plane `pl` from the
// reconstructed picture.  top/left hold 2*size samples.  Ends with a barrier.
TDN void intra_neighbours(Grp &g, SbMem &S, const uint8_t *rec, int stride, int x, int y, int size, int upright, int downleft, int pl) {
  THOR_LOCAL(&g);
  THOR_SHS(&S);
  uint8_t *top = S.top[pl], *left = S.left[pl];
  const int toplen = upright ? size + 1 : size, leftlen = downleft ? size + 1 : size;
  const uint8_t *p = rec + y * stride + x;
  TFOR(g, k, 2 * size) {
    top[k] = y == 0 ? 128 : ld_frame8(p - stride + imin(k, toplen - 1));
    left[k] = x == 0 ? 128 : ld_frame8(p + imin(k, leftlen - 1) * stride - 1);
  }
  g.sync();
  if (g.tid == 0) {
    uint8_t tl;
    if (y == 0) tl = left[0];
    else tl = x > 0 ? ld_frame8(p - stride - 1) : top[0];
    S.topleft[pl] = tl;
  }
  g.sync();
}

// [1 2 1] / 4 with the end samples repeated (:39-48)
TD void smooth121(Grp &g, const uint8_t *in, uint8_t *out, int len) {
  THOR_LOCAL(&g);
  THOR_SHARED(in); THOR_SHARED(out);
  TFOR(g, j, len) out[j] = (uint8_t)((in[imax(j - 1, 0)] + 2 * in[j] + in[imin(j + 1, len - 1)] + 2) >> 2);
}

// get_intra_prediction (:363-389) with the ten mode functions (:145-361); pblock: size x size, stride size.
// Ends with a barrier.
TDN void intra_predict(Grp &g, SbMem &S, int pl, int ypos, int xpos, int size, uint8_t *pblock, int mode) {
  PROF(PF_PRED_INTRA);
  THOR_LOCAL(&g);
  THOR_SHS(&S); THOR_SHARED(pblock);
  const uint8_t *top = S.top[pl], *left = S.left[pl];
  const int top_left = S.topleft[pl];
  const int n = size * size, sh = ilog2((uint32_t)size);
  switch (mode) {
    case 2:                                                        // MODE_HOR
      TFOR(g, i, n) pblock[i] = left[i >> sh];
      break;
    case 3:                                                        // MODE_VER
      TFOR(g, i, n) pblock[i] = top[i & (size - 1)];
      break;
    case 1: {                                                      // MODE_PLANAR: [1 2 2 2 1] smoothing, plane sum, C division
      int16_t *tF = S.itmp, *lF = S.itmp + 64;
      TFOR(g, j, size) {
        const int a = imax(j - 2, 0), b = imax(j - 1, 0), c = imin(j + 1, size - 1), d = imin(j + 2, size - 1);
        tF[j] = (int16_t)(top[a] + 2 * top[b] + 2 * top[j] + 2 * top[c] + top[d]);
        lF[j] = (int16_t)(left[a] + 2 * left[b] + 2 * left[j] + 2 * left[c] + left[d]);
      }
      g.sync();
      const int tlF = left[1] + 2 * left[0] + 2 * top_left + 2 * top[0] + top[1];
      TFOR(g, i, n) pblock[i] = (uint8_t)clip255((lF[i >> sh] + tF[i & (size - 1)] - tlF + 4) / 8);
      break;
    }
    case 4: case 7: case 8: {                                      // UPLEFT, UPUPLEFT, UPLEFTLEFT
      smooth121(g, left, S.lf, size);
      smooth121(g, top, S.tf, size);
      g.sync();
      const uint8_t *lf = S.lf, *tf = S.tf;
      const int tlf = (2 * top_left + left[0] + top[0] + 2) >> 2;
      TFOR(g, idx, n) {
        const int i = idx >> sh, j = idx & (size - 1);
        int v;
        if (mode == 4) {
          const int d = i - j;
          v = d > 0 ? lf[d - 1] : d == 0 ? tlf : tf[-d - 1];
        } else if (mode == 7) {
          const int d = i - 2 * j;
          if (d > 1) v = lf[d - 2];
          else if (d == 1) v = tlf;
          else if (d == 0) v = (tlf + tf[0]) >> 1;
          else v = (d & 1) ? tf[(-d) / 2] : (tf[(-d) / 2] + tf[(-d) / 2 - 1]) >> 1;
        } else {
          const int d = 2 * i - j;
          if (d < -1) v = tf[-d - 2];
          else if (d == -1) v = tlf;
          else if (d == 0) v = (tlf + lf[0]) >> 1;
          else v = (d & 1) ? lf[d / 2] : (lf[d / 2] + lf[d / 2 - 1]) >> 1;
        }
        pblock[idx] = (uint8_t)v;
      }
      break;
    }
    case 5: case 6: {                                              // UPRIGHT, UPUPRIGHT
      smooth121(g, top, S.tf, 2 * size);
      g.sync();
      const uint8_t *tf = S.tf;
      TFOR(g, idx, n) {
        const int i = idx >> sh, j = idx & (size - 1);
        if (mode == 5) pblock[idx] = tf[i + j + 1];
        else {
          const int d = i + 2 * j;
          pblock[idx] = (uint8_t)((d & 1) ? tf[(d + 1) / 2] : (tf[d / 2] + tf[d / 2 + 1]) >> 1);
        }
      }
      break;
    }
    case 9: {                                                      // DOWNLEFTLEFT
      smooth121(g, left, S.lf, 2 * size);
      g.sync();
      const uint8_t *lf = S.lf;
      TFOR(g, idx, n) {
        const int i = idx >> sh, j = idx & (size - 1);
        const int d = 2 * i + j;
        pblock[idx] = (uint8_t)((d & 1) ? lf[(d + 1) / 2] : (lf[d / 2] + lf[d / 2 + 1]) >> 1);
      }
      break;
    }
    default: {                                                     // MODE_DC: the available side counts twice at a picture edge
      const uint8_t *a = xpos != 0 ? left : top, *b = ypos != 0 ? top : left;
      THOR_SHARED(a); THOR_SHARED(b);
      int sum = 0;
      for (int j = 0; j < size; j++) sum += a[j] + b[j];           // uniform: every thread adds the same samples
      const int dc = (sum + size) / (2 * size);
      TFOR(g, i, n) pblock[i] = (uint8_t)dc;
      break;
    }
  }
  g.sync();
}

}  // namespace thorenc
