/* Reference-side binding of libthorcuda's fused frame tail for the Thor ENCODER.
 *
 * encode_frame() calls
 *     thorcuda_enc_tail(encoder_info, stream, qp)     instead of deblock_frame_y / deblock_frame_uv / clpf_frame
 *     thorcuda_enc_make_reference(ref0, rec)          instead of create_reference_frame
 * (integration/apply_hooks.py substitutes them in a scratch copy at build time).  The encoder-side CLPF decision
 * -- per superblock, SSD(filtered) < SSD(unfiltered) over the coded non-bi-predicted 8x8 luma blocks,
 * enc/encode_frame.c:50-63 -- is taken on the device against the original frame; the hook writes the resulting bits
 * into the stream in the order clpf_frame's callback would have (raster order of the signalling superblocks).
 * Compiled against the reference's own headers where they lie. */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "global.h"
#include "mainenc.h"
#include "putbits.h"
#include "thorcuda.h"

static thorcuda_pic dev_rec = -1, dev_org = -1, dev_ref = -1;
static uint8_t *sb_bits;

static void need(int rc) {
  if (rc) fatalerror((char *)thorcuda_last_error());
}

void thorcuda_enc_tail(encoder_info_t *ei, stream_t *stream, int qp) {
  yuv_frame_t *rec = ei->rec, *org = ei->orig;
  if (dev_rec < 0) {
    const char *dev = getenv("THORCUDA_DEVICE");
    need(thorcuda_init(dev ? atoi(dev) : 0));
    dev_rec = thorcuda_pic_create(rec->width, rec->height, 0, 0);
    dev_org = thorcuda_pic_create(rec->width, rec->height, 0, 0);
    dev_ref = thorcuda_pic_create(rec->width, rec->height, PADDING_Y, PADDING_Y / 2);
    if (dev_rec < 0 || dev_org < 0 || dev_ref < 0) fatalerror((char *)thorcuda_last_error());
    sb_bits = (uint8_t *)thorcuda_host_alloc((size_t)(rec->width / MAX_BLOCK_SIZE + 1) * (rec->height / MAX_BLOCK_SIZE + 1));
  }
  need(thorcuda_pic_upload(dev_rec, rec->y, rec->u, rec->v, rec->stride_y, rec->stride_c));
  need(thorcuda_pic_upload(dev_org, org->y, org->u, org->v, org->stride_y, org->stride_c));
  need(thorcuda_blockinfo_upload(dev_rec, ei->deblock_data));
  int stages = THORCUDA_LF_MAKE_REF;
  if (ei->params->deblocking) stages |= THORCUDA_LF_DEBLOCK_Y | THORCUDA_LF_DEBLOCK_UV;
  if (ei->params->clpf) stages |= THORCUDA_LF_CLPF;
  need(thorcuda_loopfilter_frame(dev_rec, dev_org, dev_ref, qp, stages, NULL, sb_bits));
  need(thorcuda_pic_download(dev_rec, rec->y, rec->u, rec->v, rec->stride_y, rec->stride_c));
  need(thorcuda_sync());
  if (ei->params->clpf) {
    putbits(1, 1, stream);            /* frame-level enable */
    putbits(1, 0, stream);            /* "every superblock signals its own bit" */
    const int n = (rec->width / MAX_BLOCK_SIZE) * (rec->height / MAX_BLOCK_SIZE);
    for (int i = 0; i < n; i++)
      if (sb_bits[i] != 0xFF) putbits(1, sb_bits[i], stream);
  }
}

void thorcuda_enc_make_reference(yuv_frame_t *ref, yuv_frame_t *rec) {
  ref->frame_num = rec->frame_num;
  need(thorcuda_pic_download_padded(dev_ref, ref->y - ref->pad_ver_y * ref->stride_y - ref->pad_hor_y,
                                    ref->u - ref->pad_ver_c * ref->stride_c - ref->pad_hor_c,
                                    ref->v - ref->pad_ver_c * ref->stride_c - ref->pad_hor_c));
  need(thorcuda_sync());
}
