/*
 * sk_oracle.h — CPU oracle for the Selkies pre-encode hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under selkies_gstreamer_b200/ may include,
 * link or load this; only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs do, and there only as the checker or
 * as the timed host baseline.
 *
 * PARITY UNPINNED: the arithmetic lives in GStreamer 1.24.12 (pinned by the
 * reference at addons/gstreamer/Dockerfile:85 and addons/conda/build/meta.yaml:6),
 * which is neither vendored under /root/reference nor installable here.  This
 * file restates the published algorithm of its generic videoconvert /
 * videoscale path (see sk_oracle.c for the per-function upstream names) and is
 * anchored on the reference's call sites (src/selkies/legacy/gstwebrtc_app.py:611-617
 * and siblings) and on the known-answer vectors of SURVEY.md Appendix A.7.
 */
#ifndef SK_ORACLE_H
#define SK_ORACLE_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { SKO_IN_BGRX = 0, SKO_IN_RGBX = 1 };            /* BGRA / RGBA share the unpack */
enum { SKO_OUT_NV12 = 0, SKO_OUT_I420 = 1 };
enum { SKO_MATRIX_BT601 = 0, SKO_MATRIX_BT709 = 1, SKO_MATRIX_BT2020 = 2 };
enum { SKO_SITE_NONE = 0, SKO_SITE_H_COSITED = 1 };
/* GST_LQ   : upstream resampler as recalled (filter stretched by 1/scale when
 *            shrinking, truncated to 2 taps, renormalised), 6-bit vertical taps,
 *            edge-aligned 8-bit horizontal accumulator.
 * PLAIN_LQ : same integer kernels, but textbook centre-aligned bilinear
 *            vertical weights (no stretch). */
enum { SKO_SCALE_GST_LQ = 0, SKO_SCALE_PLAIN_LQ = 1 };

/* Parity switches — one bit per recollection of upstream that could be wrong (SURVEY.md A.8 risk
 * list); 0 = the restatement as written.  The product carries the same bits (sk_config.parity_switches,
 * include/selkies_preencode.h) so that real GStreamer vectors, once somebody can produce them
 * (tools/make_gst_golden.py), either turn parity green or name the bit to flip. */
enum {
    SKO_PS_MATRIX_ROUND   = 1 << 0,  /* R8: the 8-bit matrix adds a +128 rounding term before >>8           */
    SKO_PS_COEFF_HALF_UP  = 1 << 1,  /* R8: coefficients rounded half away from zero instead of rint()      */
    SKO_PS_CS_EDGE_KEEP   = 1 << 2,  /* R3: the cosited filter leaves the first and last sample unfiltered  */
    SKO_PS_CHROMA_V_FIRST = 1 << 3,  /* R4: vertical average first, horizontal filter second                */
    SKO_PS_CHROMA_BOX4    = 1 << 4,  /* R4: site none: one rounding, (a+b+c+d+2)>>2, not cascaded averages  */
    SKO_PS_HINC_PLAIN     = 1 << 5,  /* R6: horizontal increment ((in-1)<<16)/(out-1) without the "-1"      */
    SKO_PS_V_FIRST_ON_TIE = 1 << 6   /* R7: equal cost of the two pass orders goes to vertical first        */
};

typedef struct sko_config {
    int in_w, in_h;
    int in_stride;      /* bytes; >= 4*in_w */
    int in_fmt;
    int out_w, out_h;
    int out_fmt;
    int matrix;
    int chroma_site;
    int scale_variant;
    int n_threads;      /* row bands, as videoconvert n-threads; <=1 means serial */
    int switches;       /* SKO_PS_* */
} sko_config;

/* Size of the packed output frame in GstVideoInfo default layout. */
size_t sko_out_size(int out_w, int out_h, int out_fmt);
/* Plane geometry of that layout: offsets/strides for Y, U(or UV), V. */
void sko_out_layout(int out_w, int out_h, int out_fmt,
                    size_t offset[3], int stride[3]);

/* The 3x4 integer matrix (8 fractional bits) for [Y,U,V] rows x [R,G,B,offset]. */
void sko_matrix_coeffs(int matrix, int coeffs[3][4]);
void sko_matrix_coeffs_sw(int matrix, int switches, int coeffs[3][4]);

/* Vertical / generic 2-tap tables of the resampler: for each of out_size
 * outputs the first source index and the s16 second tap at `precision` bits. */
void sko_resampler_2tap(int in_size, int out_size, int variant, int precision,
                        int32_t *offset, int16_t *tap1);
/* Horizontal edge-aligned increment (16.16). */
int32_t sko_h_increment(int in_w, int out_w);
int32_t sko_h_increment_sw(int in_w, int out_w, int switches);

/* Convert (+scale) one frame. Returns 0 or a negative error. */
int sko_convert(const sko_config *cfg, const uint8_t *in, uint8_t *out);

/* Per-16x16 damage map on the OUTPUT frame. prev==NULL → all 1.
 * map has ceil(h/16) rows of ceil(w/16) bytes. Returns damaged count. */
int sko_damage(int out_w, int out_h, int out_fmt,
               const uint8_t *cur, const uint8_t *prev, uint8_t *map);

/* Temporal damage policy (definition owned by this project, shaped by the knobs at
 * src/selkies/selkies.py:2949-2951).  state: one uint32 per macroblock, in/out
 * (streak | hot<<8 | still<<16 | dirty<<24); policy: one byte per macroblock, out
 * (1 damaged, 2 hot, 4 paint-over). */
void sko_policy_update(int n_blocks, const uint8_t *map, uint32_t *state, uint8_t *policy,
                       int threshold, int duration, int trigger);

/* Chroma filters exposed for known-answer tests: operate on AYUV lines in place. */
void sko_chroma_down_h2(uint8_t *ayuv, int w);
void sko_chroma_down_h2_cs(uint8_t *ayuv, int w);
void sko_chroma_down_v2(uint8_t *ayuv0, const uint8_t *ayuv1, int w);

#ifdef __cplusplus
}
#endif
#endif
