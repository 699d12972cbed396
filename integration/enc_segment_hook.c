/* Closed-GOP segment encoding for the reference's sequence loop.
 *
 * main() of enc/mainenc.c (a scratch copy patched by integration/apply_hooks.py, mode "mainseg"/"mainenc") takes two
 * extra arguments that the reference's own parser never sees:
 *
 *     -seg_start A   first frame (display number) this process emits; must be an intra frame of the stream
 *     -seg_end   B   first frame it no longer emits; an intra frame too, or absent / -1 for "to the end of the stream"
 *
 * The process runs the WHOLE sequence loop (enc/mainenc.c:222-591) so that num_encoded_frames, coded_phase, b_level, the
 * frame QP, num_ref, ref_array, last_PorI_frame, last_intra_frame_num and the PPP tail rule take exactly the values of
 * the serial run -- none of that arithmetic is restated here -- but
 *   - frames outside [A, B] are not encoded: only the reference ring slides (thor_seg_skip_frame) so that ring indices
 *     and the frame numbers the random-access stripping rule reads (enc/mainenc.c:472-486) line up;
 *   - frame B, the next segment's intra frame, IS encoded (its leading B frames, display numbers below B, belong to this
 *     segment and predict from it) but its bits and its reconstruction are dropped: the next segment emits them;
 *   - the sequence header (enc/mainenc.c:196-212) leaves with the first frame of the stream, i.e. only when A == 0.
 * Frames keep their absolute numbers (enc/encode_frame.c:110 writes frame_num), which is why -skip cannot do this job
 * (enc/mainenc.c:232 rebases the numbering) and is refused together with -seg_start.
 *
 * The .bit a segment writes is a sequence of length-prefixed frame chunks (flush_all_bits, enc/putbits.c:57-95) in coding
 * order; thor_b200/segments.py joins the chunks of all segments by coding index into the serial stream.  The
 * reconstruction file holds display frames A .. B-1, so plain concatenation joins it.
 * Without -seg_start every function here answers "yes" and the program is the stock encoder. */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "global.h"
#include "mainenc.h"
#include "putbits.h"

static int seg_on, seg_a, seg_b = -1;

static int int_arg(const char *flag, const char *s) {
  char *end;
  long v = strtol(s, &end, 10);
  if (*s == 0 || *end != 0 || v < -1 || v > 0x3fffffff) {
    fprintf(stderr, "%s: '%s' is not a frame number\n", flag, s);
    exit(2);
  }
  return (int)v;
}

/* strips -seg_start / -seg_end from argv before parse_config_params (enc/strings.c) rejects them */
void thor_seg_args(int *argc, char **argv) {
  int n = 1;
  for (int i = 1; i < *argc; i++) {
    const int is_a = !strcmp(argv[i], "-seg_start"), is_b = !strcmp(argv[i], "-seg_end");
    if (is_a || is_b) {
      if (i + 1 >= *argc) {
        fprintf(stderr, "%s needs a frame number\n", argv[i]);
        exit(2);
      }
      const int v = int_arg(argv[i], argv[i + 1]);
      if (is_a) { seg_a = v; seg_on = 1; } else seg_b = v;
      i++;
    } else
      argv[n++] = argv[i];
  }
  *argc = n;
  if (seg_b >= 0 && !seg_on) { seg_on = 1; seg_a = 0; }
}

void thor_seg_check(const enc_params *p) {
  if (!seg_on) return;
  const char *why = NULL;
  if (seg_a < 0) why = "-seg_start must not be negative";
  else if (p->skip != 0) why = "-skip rebases the frame numbers (enc/mainenc.c:232): not allowed with -seg_start";
  else if (p->intra_period <= 0 || p->num_reorder_pics == 0) why = "only streams with -intra_period and reordering have closed-GOP segments";
  else if (seg_a % p->intra_period) why = "-seg_start is not an intra frame of this stream";
  else if (seg_b >= 0 && (seg_b % p->intra_period || seg_b <= seg_a)) why = "-seg_end is not a later intra frame of this stream";
  else if (p->intra_period % (p->num_reorder_pics + 1)) why = "-intra_period is not a multiple of the sub-GOP: intra frames would not be anchors";
  if (why) {
    fprintf(stderr, "segment encoding refused: %s\n", why);
    exit(2);
  }
  fprintf(stderr, "segment: emitting frames %d .. %s\n", seg_a, seg_b < 0 ? "end" : "(seg_end - 1)");
}

int thor_seg_encodes(int frame_num) { return !seg_on || (frame_num >= seg_a && (seg_b < 0 || frame_num <= seg_b)); }
int thor_seg_emits(int frame_num) { return !seg_on || (frame_num >= seg_a && (seg_b < 0 || frame_num < seg_b)); }
/* the reconstruction writer starts at display frame A (last_frame_output, enc/mainenc.c:81) */
int thor_seg_first_output(void) { return seg_on ? seg_a - 1 : -1; }

/* What encode_frame() leaves behind for the NEXT frames when nobody will ever look at this frame's pixels: the sliding
 * window moved by one (enc/encode_frame.c:165-177) and the new head carrying the frame's number.  The device twins of the
 * ring (enc_gpu_hook.c) are keyed by these yuv_frame_t, so they slide with it. */
void thor_seg_skip_frame(encoder_info_t *e) {
  yuv_frame_t *head = e->ref[MAX_REF_FRAMES - 1];
  for (int r = MAX_REF_FRAMES - 1; r > 0; r--) e->ref[r] = e->ref[r - 1];
  e->ref[0] = head;
  head->frame_num = e->rec->frame_num;
}

/* flush_all_bits for an emitted frame; otherwise the frame's bits (and, before frame A, the sequence header) are dropped */
void thor_seg_flush(stream_t *str, FILE *out, int frame_num) {
  if (thor_seg_emits(frame_num)) {
    flush_all_bits(str, out);
    return;
  }
  str->bytepos = 0;
  str->bitbuf = 0;
  str->bitrest = 32;
}
