/*
 * thorcuda_compat.h -- the reference's own per-block kernel API, served by libthorcuda.
 *
 * Same symbol names and signatures as awakecoding/thor's
 *     common/common_kernels.h:30-37   enc/enc_kernels.h:32-37   common/transform.h:36-37
 * so that Thorenc / Thordec link against libthorcuda.so INSTEAD of common_kernels.o, enc_kernels.o and
 * transform.o with no source change (INTEGRATION.md "link-time replacement"; the reference dispatches to
 * these symbols whenever its global `use_simd` is set, common/simd.h:134-140).
 *
 * All pointers are HOST pointers, exactly as the reference passes them (into malloc'ed frames or
 * alloca'ed compact blocks).  Every call is synchronous: the operands are copied to the B200, ONE block is
 * processed by the same device code the batched API of thorcuda.h runs, the result is copied back, and the
 * call returns after the stream has drained.  They exist for API fidelity and kernel-level parity tests;
 * the performance path is the batched API.  There is no CPU fallback: the first call creates the CUDA
 * context on device $THORCUDA_DEVICE (default 0); without an sm_100 device it aborts like the
 * reference's fatalerror() (common/global.h:38-44).
 *
 * Operand ranges are the ones the reference's dispatchers guarantee: luma prediction
 * width/height 4..64 (multiples of 4), (xoff,yoff) in 0..3; chroma width/height 4..32 (even), offsets 0..7;
 * SAD/SSD/half/quarter blocks 8..64 (multiples of 4; 4 is accepted too); widesad any such block (the
 * reference only calls it for 16x16); transforms 4,8,16,32,64.  Anything else aborts.
 */
#ifndef THORCUDA_COMPAT_H
#define THORCUDA_COMPAT_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#define THOR_RESTRICT __restrict__
#else
#define THOR_RESTRICT restrict
#endif

/* ---- common/common_kernels.h ------------------------------------------------------------------ */
/* :30  rounding average (r0+r1+1)>>1 of two blocks (temporal interpolation, common/temporal_interp.c:404-411) */
void block_avg_simd(uint8_t *p, uint8_t *r0, uint8_t *r1, int sp, int s0, int s1, int width, int height);
/* :31  SAD with unaligned operands, width <= 16 (common/temporal_interp.c:470-471,562-563; the reference's own
 *      loop is wrong for wider blocks, common/common_kernels.c:111-119, so they are rejected) */
int sad_calc_simd_unaligned(uint8_t *a, uint8_t *b, int astride, int bstride, int width, int height);
/* :32  quarter-pel luma prediction; ip already includes the integer MV; (xoff,yoff) != (0,0)
 *      (dispatcher: get_inter_prediction_luma, common/inter_prediction.c:120-180) */
void get_inter_prediction_luma_simd(int width, int height, int xoff, int yoff, unsigned char *THOR_RESTRICT qp, int qstride,
                                    const unsigned char *THOR_RESTRICT ip, int istride, int bipred);
/* :33  eighth-pel chroma prediction (dispatcher: get_inter_prediction_chroma, common/inter_prediction.c:72-118) */
void get_inter_prediction_chroma_simd(int width, int height, int xoff, int yoff, unsigned char *THOR_RESTRICT qp, int qstride,
                                      const unsigned char *THOR_RESTRICT ip, int istride);
/* :34-35  forward / inverse integer DCT of one block (dispatchers: transform/inverse_transform, common/transform.c:249,488) */
void transform_simd(const int16_t *block, int16_t *coeff, int size, int fast);
void inverse_transform_simd(const int16_t *coeff, int16_t *block, int size);
/* :36-37  CLPF of one 4x4 / 8x8 block at (x0,y0) of plane `src`, written into the SB-local buffer `dst`
 *      whose stride doubles as the SB size (clpf_block, common/common_block.c:180-197).  Contract as in the
 *      reference: the block's superblock lies completely inside the plane (clpf_frame, common_frame.c:496-497) */
void clpf_block4(const uint8_t *src, uint8_t *dst, int sstride, int dstride, int x0, int y0, int width, int height);
void clpf_block8(const uint8_t *src, uint8_t *dst, int sstride, int dstride, int x0, int y0, int width, int height);

/* ---- enc/enc_kernels.h ---------------------------------------------------------------------- */
/* :32  sad_calc (enc/encode_block.c:740-755) */
int sad_calc_simd(uint8_t *a, uint8_t *b, int astride, int bstride, int width, int height);
/* :33  ssd_calc (enc/encode_block.c:782-797) */
int ssd_calc_simd(uint8_t *a, uint8_t *b, int astride, int bstride, int size);
/* :34  CLPF decision SSDs of one 8x8 luma block, ACCUMULATED into *sum0 (unfiltered), *sum1 (filtered)
 *      (detect_clpf, enc/encode_block.c:3036-3057) */
void detect_clpf_simd(const uint8_t *rec, const uint8_t *org, int x0, int y0, int width, int height, int so, int stride,
                      int *sum0, int *sum1);
/* :35  8 bilinear half-pel SADs around b, returns the minimum and its (x,y) in {-2,0,2}^2
 *      (sad_calc_fasthalf, enc/encode_block.c:497-605).  The reference DEFINES this with a ninth,
 *      never-written parameter `unsigned int *maxsad` (enc/enc_kernels.c:162) although the header declares
 *      eight; under the C ABI an unused trailing argument is harmless, so eight are declared here too. */
unsigned int sad_calc_fasthalf_simd(const uint8_t *a, const uint8_t *b, int astride, int bstride, int width, int height,
                                    int *x, int *y);
/* :36  8 quarter-pel SADs around the half-pel point (*x,*y) (in/out) (sad_calc_fastquarter, enc/encode_block.c:609-738) */
unsigned int sad_calc_fastquarter_simd(const uint8_t *o, const uint8_t *r, int os, int rs, int width, int height, int *x, int *y);
/* :37  SAD at x offsets {-3,-1,0,+1,+3}: returns the minimum, *x its offset (widesad_calc, enc/encode_block.c:757-780) */
unsigned int widesad_calc_simd(uint8_t *a, uint8_t *b, int astride, int bstride, int width, int height, int *x);

/* ---- common/transform.h ---------------------------------------------------------------------- */
/* :36-37  the dispatchers themselves, so transform.o can be dropped from the link as well */
void transform(const int16_t *block, int16_t *coeff, int size, int fast);
void inverse_transform(const int16_t *coeff, int16_t *block, int size);

#ifdef __cplusplus
}
#endif
#endif
