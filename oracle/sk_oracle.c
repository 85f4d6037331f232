/*
 * sk_oracle.c — CPU restatement of the reference's pre-encode stage.
 *
 * TEST INFRASTRUCTURE ONLY (see sk_oracle.h).  PARITY UNPINNED: GStreamer
 * 1.24.12 is a non-vendored dependency of the reference; this restates its
 * generic converter chain for BGRx/RGBx -> I420/NV12, 8 bit:
 *
 *     unpack -> [2-tap scale, picture not growing] -> 8-bit matrix table ->
 *     [2-tap scale, picture growing] -> chroma h/v down-sample -> pack
 *
 * Reference call sites that select this chain (element names only; the
 * reference holds no pixel arithmetic of its own):
 *     src/selkies/legacy/gstwebrtc_app.py:611-617   videoconvert -> NV12 (x264enc)
 *     src/selkies/legacy/gstwebrtc_app.py:643-649   videoconvert -> I420 (openh264enc)
 *     ... :669-675, :686-692, :725-731, :742-748, :767-773 (I420 family)
 *     src/selkies/legacy/gstwebrtc_app.py:263-284   cudaupload+cudaconvert -> NV12
 *
 * Each function names the upstream routine it follows (gst-plugins-base
 * 1.24.12, gst-libs/gst/video/...).
 */
#include "sk_oracle.h"

#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

#define RU2(x) (((x) + 1) & ~1)
#define RU4(x) (((x) + 3) & ~3)

/* ------------------------------------------------------------------ layout */

/* video-info.c: fill_planes() for I420 and NV12. */
void sko_out_layout(int w, int h, int fmt, size_t offset[3], int stride[3])
{
    int h2 = RU2(h);
    if (fmt == SKO_OUT_I420) {
        stride[0] = RU4(w);
        stride[1] = RU4(RU2(w) / 2);
        stride[2] = stride[1];
        offset[0] = 0;
        offset[1] = (size_t)stride[0] * h2;
        offset[2] = offset[1] + (size_t)stride[1] * (h2 / 2);
    } else {
        stride[0] = RU4(w);
        stride[1] = stride[0];
        stride[2] = 0;
        offset[0] = 0;
        offset[1] = (size_t)stride[0] * h2;
        offset[2] = 0;
    }
}

size_t sko_out_size(int w, int h, int fmt)
{
    size_t off[3];
    int st[3];
    int h2 = RU2(h);
    sko_out_layout(w, h, fmt, off, st);
    if (fmt == SKO_OUT_I420)
        return off[2] + (size_t)st[2] * (h2 / 2);
    return off[1] + (size_t)st[1] * (h2 / 2);
}

/* ------------------------------------------------------------------ matrix */

/* video-color.c: gst_video_color_matrix_get_Kr_Kb(). */
static void kr_kb(int matrix, double *kr, double *kb)
{
    switch (matrix) {
    case SKO_MATRIX_BT709:  *kr = 0.2126; *kb = 0.0722; break;
    case SKO_MATRIX_BT2020: *kr = 0.2627; *kb = 0.0593; break;
    default:                *kr = 0.299;  *kb = 0.114;  break;
    }
}

/*
 * video-converter.c: compute_matrix_to_RGB() (RGB in: offset 0, scale 1/255),
 * color_matrix_RGB_to_YCbCr(), compute_matrix_to_YUV() (range 16-235: scale
 * 219/224/224, offset 16/128/128), prepare_matrix() (x256) and
 * color_matrix_convert() (rint per entry).
 */
void sko_matrix_coeffs(int matrix, int c[3][4])
{
    sko_matrix_coeffs_sw(matrix, 0, c);
}

void sko_matrix_coeffs_sw(int matrix, int switches, int c[3][4])
{
    double kr, kb, kg, x, m[3][4];
    const double in_scale = 1 / ((float)255);
    const double sc[3] = {(float)219, (float)224, (float)224};
    const double of[3] = {16, 128, 128};
    int i, j;

    kr_kb(matrix, &kr, &kb);
    kg = 1.0 - kr - kb;
    m[0][0] = kr;  m[0][1] = kg;  m[0][2] = kb;
    x = 1 / (2 * (1 - kb));
    m[1][0] = -x * kr;  m[1][1] = -x * kg;  m[1][2] = x * (1 - kb);
    x = 1 / (2 * (1 - kr));
    m[2][0] = x * (1 - kr);  m[2][1] = -x * kg;  m[2][2] = -x * kb;
    for (i = 0; i < 3; i++) {
        for (j = 0; j < 3; j++)
            m[i][j] = m[i][j] * in_scale * sc[i] * 256.0;
        m[i][3] = of[i] * 256.0;
        for (j = 0; j < 4; j++)
            c[i][j] = (switches & SKO_PS_COEFF_HALF_UP) ? (int)round(m[i][j]) : (int)rint(m[i][j]);
        if (switches & SKO_PS_MATRIX_ROUND)
            c[i][3] += 128;          /* R8: round to nearest at the >>8 instead of flooring */
    }
}

typedef struct {
    int64_t t_r[256], t_g[256], t_b[256], t_c;
} matrix_tables;

/* video-converter.c: videoconvert_convert_init_tables() — three 16-bit lanes
 * accumulated in one int64 per channel value. */
static void init_tables(matrix_tables *t, int matrix, int switches)
{
    int c[3][4], i, j;
    sko_matrix_coeffs_sw(matrix, switches, c);
    for (i = 0; i < 256; i++) {
        int64_t r = 0, g = 0, b = 0;
        for (j = 0; j < 3; j++) {
            r = (r << 16) + (int64_t)c[j][0] * i;
            g = (g << 16) + (int64_t)c[j][1] * i;
            b = (b << 16) + (int64_t)c[j][2] * i;
        }
        t->t_r[i] = r;  t->t_g[i] = g;  t->t_b[i] = b;
    }
    t->t_c = ((int64_t)c[0][3] << 32) + ((int64_t)c[1][3] << 16) + (int64_t)c[2][3];
}

/* video-converter.c: video_converter_matrix8_table() — ARGB line to AYUV in place. */
static void matrix8_table(const matrix_tables *t, uint8_t *p, int w)
{
    int i;
    for (i = 0; i < w * 4; i += 4) {
        int64_t x = t->t_r[p[i + 1]] + t->t_g[p[i + 2]] + t->t_b[p[i + 3]] + t->t_c;
        p[i + 1] = (uint8_t)(x >> (32 + 8));
        p[i + 2] = (uint8_t)(x >> (16 + 8));
        p[i + 3] = (uint8_t)(x >> 8);
    }
}

/* ------------------------------------------------------------------ unpack */

/* video-format.c: unpack_BGRA() (BGRx shares it) / unpack_RGBA(): to A,R,G,B. */
static void unpack_line(int fmt, const uint8_t *s, uint8_t *d, int w)
{
    int i;
    if (fmt == SKO_IN_BGRX) {
        for (i = 0; i < w; i++) {
            d[4 * i + 0] = s[4 * i + 3];
            d[4 * i + 1] = s[4 * i + 2];
            d[4 * i + 2] = s[4 * i + 1];
            d[4 * i + 3] = s[4 * i + 0];
        }
    } else {
        for (i = 0; i < w; i++) {
            d[4 * i + 0] = s[4 * i + 3];
            d[4 * i + 1] = s[4 * i + 0];
            d[4 * i + 2] = s[4 * i + 1];
            d[4 * i + 3] = s[4 * i + 2];
        }
    }
}

/* ------------------------------------------------------------------ chroma */

/* video-chroma.c / video-orc.orc: video_orc_chroma_down_h2_u8 — avgub of each
 * pixel pair's U,V into the even pixel. */
void sko_chroma_down_h2(uint8_t *p, int w)
{
    int i;
    for (i = 0; i + 1 < w; i += 2) {
        p[4 * i + 2] = (uint8_t)((p[4 * i + 2] + p[4 * i + 6] + 1) >> 1);
        p[4 * i + 3] = (uint8_t)((p[4 * i + 3] + p[4 * i + 7] + 1) >> 1);
    }
}

#define F31(a, b)     (uint8_t)((3 * (a) + (b) + 2) >> 2)
#define F121(a, b, c) (uint8_t)(((a) + 2 * (b) + (c) + 2) >> 2)

/* video-chroma.c: video_chroma_down_h2_cs_u8 — [3 1] at x=0, [1 2 1] at even x,
 * [1 3] closing tap; odd samples stay original while the loop runs. */
static void chroma_down_h2_cs_sw(uint8_t *p, int w, int keep_edges)
{
    int i, k;
    if (w < 2)
        return;
    if (!keep_edges)
        for (k = 2; k <= 3; k++)
            p[k] = F31(p[k], p[4 + k]);
    for (i = 2; i < w - 2; i += 2)
        for (k = 2; k <= 3; k++)
            p[4 * i + k] = F121(p[4 * (i - 1) + k], p[4 * i + k], p[4 * (i + 1) + k]);
    if (i < w && !keep_edges)
        for (k = 2; k <= 3; k++)
            p[4 * i + k] = F31(p[4 * i + k], p[4 * (i - 1) + k]);
}

void sko_chroma_down_h2_cs(uint8_t *p, int w)
{
    chroma_down_h2_cs_sw(p, w, 0);
}

/* R4 alternative for site none: both pairs of both lines in one rounding. */
static void chroma_down_box4(uint8_t *l0, const uint8_t *l1, int w)
{
    int i, k;
    for (i = 0; i + 1 < w; i += 2)
        for (k = 2; k <= 3; k++)
            l0[4 * i + k] = (uint8_t)((l0[4 * i + k] + l0[4 * i + 4 + k] + l1[4 * i + k] + l1[4 * i + 4 + k] + 2) >> 2);
}

/* video-orc.orc: video_orc_chroma_down_v2_u8 — avgub of U,V of two lines into
 * the first. */
void sko_chroma_down_v2(uint8_t *l0, const uint8_t *l1, int w)
{
    int i;
    for (i = 0; i < w; i++) {
        l0[4 * i + 2] = (uint8_t)((l0[4 * i + 2] + l1[4 * i + 2] + 1) >> 1);
        l0[4 * i + 3] = (uint8_t)((l0[4 * i + 3] + l1[4 * i + 3] + 1) >> 1);
    }
}

/* ------------------------------------------------------------------- scale */

/* video-scaler.c: gst_video_scaler_new() — scale->inc for the 2-tap h path. */
int32_t sko_h_increment(int in_w, int out_w)
{
    return sko_h_increment_sw(in_w, out_w, 0);
}

int32_t sko_h_increment_sw(int in_w, int out_w, int switches)
{
    if (out_w == 1)
        return 0;
    return (int32_t)((((int64_t)in_w - 1) << 16) / (out_w - 1) - ((switches & SKO_PS_HINC_PLAIN) ? 0 : 1));
}

/* video-orc.orc: video_orc_resample_h_2tap_4u8_lq == ldreslinl d, s, 0, inc. */
static void hscale_2tap_lq(const uint8_t *s, uint8_t *d, int32_t inc, int out_w)
{
    int i, k;
    for (i = 0; i < out_w; i++) {
        int32_t acc = i * inc;
        int x0 = acc >> 16, f = (acc >> 8) & 0xff;
        for (k = 0; k < 4; k++)
            d[4 * i + k] = (uint8_t)((s[4 * x0 + k] * (256 - f) + s[4 * x0 + 4 + k] * f) >> 8);
    }
}

/* video-orc.orc: video_orc_resample_v_2tap_u8_lq. */
static void vscale_2tap_lq(const uint8_t *s1, const uint8_t *s2, uint8_t *d, int16_t p1, int n)
{
    int i;
    for (i = 0; i < n; i++) {
        int16_t a = s1[i];
        int16_t t = (int16_t)(s2[i] - a);
        t = (int16_t)(t * p1);
        t = (int16_t)(t + 32);
        t = (int16_t)(t >> 6);
        a = (int16_t)(a + t);
        d[i] = (uint8_t)(a < 0 ? 0 : (a > 255 ? 255 : a));
    }
}

/* video-resampler.c: get_linear_tap(). */
static double linear_tap(double fx, int l, int xi, double x)
{
    double a = fabs(x - (xi + l)) * fx;
    return a < 1.0 ? 1.0 - a : 0.0;
}

/* video-scaler.c / video-resampler.c: the DC-preserving bias search that turns
 * double taps into integers summing to 1<<precision. */
static void taps_to_int(const double *tmp, int n, int precision, int *out)
{
    double mult = (double)(1 << precision);
    double lo = 0.0, hi = 1.0, off = 0.5;
    int i, j;
    for (i = 0; i < 64; i++) {
        int sum = 0;
        for (j = 0; j < n; j++)
            sum += (int)floor(off + tmp[j] * mult);
        if (sum == (1 << precision))
            break;
        if (lo == hi)
            break;
        if (sum < mult) {
            if (off > lo) lo = off;
            off += (hi - lo) / 2;
        } else {
            if (off < hi) hi = off;
            off -= (hi - lo) / 2;
        }
    }
    for (j = 0; j < n; j++)
        out[j] = (int)floor(off + tmp[j] * mult);
}

/* video-resampler.c: gst_video_resampler_init() + resampler_calculate_taps()
 * for METHOD_LINEAR with max-taps 2 (videoscale method=bilinear). */
void sko_resampler_2tap(int in_size, int out_size, int variant, int precision,
                        int32_t *offset, int16_t *tap1)
{
    const int max_taps = 2;
    double scale = (double)in_size / (double)out_size;
    double fx = (variant == SKO_SCALE_GST_LQ && scale > 1.0) ? 1.0 / scale : 1.0;
    int j, l;
    for (j = 0; j < out_size; j++) {
        double taps[2], weight = 0.0, ox, x;
        int xi, it[2];
        ox = (0.5 + (double)j) / out_size;
        x = ox * (double)in_size - 0.5;
        if (x < 0) x = 0;
        if (x > in_size - 1) x = in_size - 1;
        xi = (int)floor(x);           /* tap_offs = (2-1)/2 = 0 */
        for (l = 0; l < max_taps; l++) {
            taps[l] = linear_tap(fx, l, xi, x);
            weight += taps[l];
        }
        for (l = 0; l < max_taps; l++)
            taps[l] /= weight;
        if (xi > in_size - max_taps) {   /* last source line: fold right tap in */
            taps[1] = taps[0] + taps[1];
            taps[0] = 0;
            xi = in_size - max_taps;
        }
        if (xi < 0) {
            taps[0] += taps[1];
            taps[1] = 0;
            xi = 0;
        }
        taps_to_int(taps, max_taps, precision, it);
        offset[j] = xi;
        tap1[j] = (int16_t)it[1];
    }
}

/* ----------------------------------------------------------------- convert */

typedef struct {
    const sko_config *cfg;
    const uint8_t *in;
    uint8_t *out;
    matrix_tables tab;
    int scaling, h_first, scale_after;
    int32_t h_inc;
    int32_t *v_off;
    int16_t *v_tap;
    size_t off[3];
    int stride[3];
} job;

typedef struct {
    job *j;
    int y0, y1;     /* output line band, y0 even */
} band;

/* One source line as the scaler sees it: unpacked ARGB when the scaler runs in front of
 * the matrix (fewer or equal output pixels), already converted AYUV when it runs behind it
 * (video-converter.c: chain_scale() is called twice, in front of chain_convert() without and
 * behind it with "force"; the first call only acts when out pixels <= in pixels). */
static void source_line(const job *j, int row, uint8_t *dst)
{
    const sko_config *c = j->cfg;
    unpack_line(c->in_fmt, j->in + (size_t)row * c->in_stride, dst, c->in_w);
    if (j->scale_after)
        matrix8_table(&j->tab, dst, c->in_w);
}

/* One converted AYUV output line (video-converter.c: the line-cache chain up to
 * and including chain_convert, plus the second chain_scale when enlarging). tmp holds 3
 * lines of max(in_w,out_w)*4 bytes. */
static void make_ayuv_line(const job *j, int y, uint8_t *dst, uint8_t *tmp)
{
    const sko_config *c = j->cfg;
    if (!j->scaling) {
        unpack_line(c->in_fmt, j->in + (size_t)y * c->in_stride, dst, c->in_w);
    } else {
        int wmax = c->in_w > c->out_w ? c->in_w : c->out_w;
        uint8_t *a = tmp, *b = tmp + (size_t)wmax * 4, *u = tmp + (size_t)wmax * 8;
        int o = (c->in_h == c->out_h) ? y : j->v_off[y];
        int16_t p1 = (c->in_h == c->out_h) ? 0 : j->v_tap[y];
        int o2 = (o + 1 < c->in_h) ? o + 1 : o;
        if (j->h_first) {
            source_line(j, o, u);
            if (c->in_w != c->out_w) hscale_2tap_lq(u, a, j->h_inc, c->out_w);
            else memcpy(a, u, (size_t)c->in_w * 4);
            if (c->in_h != c->out_h) {
                source_line(j, o2, u);
                if (c->in_w != c->out_w) hscale_2tap_lq(u, b, j->h_inc, c->out_w);
                else memcpy(b, u, (size_t)c->in_w * 4);
                vscale_2tap_lq(a, b, dst, p1, c->out_w * 4);
            } else {
                memcpy(dst, a, (size_t)c->out_w * 4);
            }
        } else {
            source_line(j, o, a);
            if (c->in_h != c->out_h) {
                source_line(j, o2, b);
                vscale_2tap_lq(a, b, u, p1, c->in_w * 4);
            } else {
                memcpy(u, a, (size_t)c->in_w * 4);
            }
            if (c->in_w != c->out_w) hscale_2tap_lq(u, dst, j->h_inc, c->out_w);
            else memcpy(dst, u, (size_t)c->in_w * 4);
        }
        if (j->scale_after)
            return;
    }
    matrix8_table(&j->tab, dst, c->out_w);
}

/* video-format.c: pack_I420()/pack_NV12(): Y from every line, U/V from even
 * pixels of even lines only. */
static void pack_pair(const job *j, int y, const uint8_t *l0, const uint8_t *l1)
{
    const sko_config *c = j->cfg;
    uint8_t *dy0 = j->out + j->off[0] + (size_t)y * j->stride[0];
    uint8_t *dy1 = dy0 + j->stride[0];
    int i, w = c->out_w;
    for (i = 0; i < w; i++)
        dy0[i] = l0[4 * i + 1];
    if (l1)
        for (i = 0; i < w; i++)
            dy1[i] = l1[4 * i + 1];
    if (c->out_fmt == SKO_OUT_I420) {
        uint8_t *du = j->out + j->off[1] + (size_t)(y / 2) * j->stride[1];
        uint8_t *dv = j->out + j->off[2] + (size_t)(y / 2) * j->stride[2];
        for (i = 0; i < w / 2; i++) {
            du[i] = l0[8 * i + 2];
            dv[i] = l0[8 * i + 3];
        }
    } else {
        uint8_t *duv = j->out + j->off[1] + (size_t)(y / 2) * j->stride[1];
        for (i = 0; i < w / 2; i++) {
            duv[2 * i + 0] = l0[8 * i + 2];
            duv[2 * i + 1] = l0[8 * i + 3];
        }
    }
}

/* video-converter.c: video_converter_generic() over one band of lines, with
 * chain_downsample (video-chroma.c: video_chroma_down_v2 → h on both lines,
 * then v) feeding chain_pack. */
static void *band_main(void *arg)
{
    band *b = (band *)arg;
    job *j = b->j;
    const sko_config *c = j->cfg;
    int wmax = c->in_w > c->out_w ? c->in_w : c->out_w;
    /* + one pixel: with SKO_PS_HINC_PLAIN the last h tap pair may start on the last source pixel (weight 0) */
    uint8_t *l0 = (uint8_t *)calloc((size_t)wmax * 4 * 5 + 16, 1);
    uint8_t *l1 = l0 + (size_t)wmax * 4;
    uint8_t *tmp = l1 + (size_t)wmax * 4;
    int y;
    for (y = b->y0; y < b->y1; y += 2) {
        int have2 = (y + 1 < c->out_h);
        make_ayuv_line(j, y, l0, tmp);
        if (have2)
            make_ayuv_line(j, y + 1, l1, tmp);
        {
            const int keep = (c->switches & SKO_PS_CS_EDGE_KEEP) != 0;
            const int cs = c->chroma_site == SKO_SITE_H_COSITED;
            if (!cs && have2 && (c->switches & SKO_PS_CHROMA_BOX4)) {
                chroma_down_box4(l0, l1, c->out_w);                    /* R4: one rounding */
            } else if (have2 && (c->switches & SKO_PS_CHROMA_V_FIRST)) {
                sko_chroma_down_v2(l0, l1, c->out_w);                   /* R4: vertical first ... */
                if (cs) chroma_down_h2_cs_sw(l0, c->out_w, keep);       /* ... then horizontal */
                else sko_chroma_down_h2(l0, c->out_w);
            } else {
                if (cs) {
                    chroma_down_h2_cs_sw(l0, c->out_w, keep);
                    if (have2) chroma_down_h2_cs_sw(l1, c->out_w, keep);
                } else {
                    sko_chroma_down_h2(l0, c->out_w);
                    if (have2) sko_chroma_down_h2(l1, c->out_w);
                }
                if (have2)
                    sko_chroma_down_v2(l0, l1, c->out_w);
            }
        }
        pack_pair(j, y, l0, have2 ? l1 : NULL);
    }
    free(l0);
    return NULL;
}

int sko_convert(const sko_config *cfg, const uint8_t *in, uint8_t *out)
{
    job j;
    int nt, i, pairs, per;
    band *bands;
    pthread_t *th;

    if (!cfg || !in || !out) return -1;
    if (cfg->in_w < 2 || cfg->in_h < 2 || cfg->out_w < 2 || cfg->out_h < 2) return -2;
    if ((cfg->in_w | cfg->in_h | cfg->out_w | cfg->out_h) & 1) return -2;
    if (cfg->in_stride < cfg->in_w * 4) return -3;

    memset(&j, 0, sizeof j);
    j.cfg = cfg;  j.in = in;  j.out = out;
    init_tables(&j.tab, cfg->matrix, cfg->switches);
    sko_out_layout(cfg->out_w, cfg->out_h, cfg->out_fmt, j.off, j.stride);
    j.scaling = (cfg->in_w != cfg->out_w) || (cfg->in_h != cfg->out_h);
    /* video-converter.c chain_scale(): in front of the matrix while the picture does not
     * grow, behind it (on AYUV) when it does. */
    j.scale_after = (int64_t)cfg->out_w * cfg->out_h > (int64_t)cfg->in_w * cfg->in_h;
    /* chain_scale(): h first when out_w*in_h <= in_w*out_h. */
    j.h_first = ((int64_t)cfg->out_w * cfg->in_h <= (int64_t)cfg->in_w * cfg->out_h);
    if ((cfg->switches & SKO_PS_V_FIRST_ON_TIE) && (int64_t)cfg->out_w * cfg->in_h == (int64_t)cfg->in_w * cfg->out_h)
        j.h_first = 0;                                                  /* R7 */
    if (j.scaling) {
        j.h_inc = sko_h_increment_sw(cfg->in_w, cfg->out_w, cfg->switches);
        j.v_off = (int32_t *)malloc(sizeof(int32_t) * cfg->out_h);
        j.v_tap = (int16_t *)malloc(sizeof(int16_t) * cfg->out_h);
        sko_resampler_2tap(cfg->in_h, cfg->out_h, cfg->scale_variant, 6, j.v_off, j.v_tap);
    }

    nt = cfg->n_threads < 1 ? 1 : cfg->n_threads;
    pairs = (cfg->out_h + 1) / 2;
    if (nt > pairs) nt = pairs;
    per = (pairs + nt - 1) / nt;
    bands = (band *)calloc(nt, sizeof(band));
    th = (pthread_t *)calloc(nt, sizeof(pthread_t));
    for (i = 0; i < nt; i++) {
        bands[i].j = &j;
        bands[i].y0 = 2 * per * i;
        bands[i].y1 = 2 * per * (i + 1);
        if (bands[i].y1 > cfg->out_h) bands[i].y1 = cfg->out_h;
        if (bands[i].y0 > cfg->out_h) bands[i].y0 = cfg->out_h;
    }
    for (i = 1; i < nt; i++)
        pthread_create(&th[i], NULL, band_main, &bands[i]);
    band_main(&bands[0]);
    for (i = 1; i < nt; i++)
        pthread_join(th[i], NULL);
    free(bands);  free(th);  free(j.v_off);  free(j.v_tap);
    return 0;
}

/* ------------------------------------------------------------------ damage */

/* SURVEY.md Appendix A.6 (definition owned by this project; the reference only
 * shows the policy knobs at src/selkies/selkies.py:2949-2951). */
int sko_damage(int w, int h, int fmt, const uint8_t *cur, const uint8_t *prev, uint8_t *map)
{
    size_t off[3];
    int st[3], mbw = (w + 15) / 16, mbh = (h + 15) / 16, bx, by, x, y, count = 0;
    sko_out_layout(w, h, fmt, off, st);
    for (by = 0; by < mbh; by++) {
        for (bx = 0; bx < mbw; bx++) {
            int d = 0;
            if (!prev) {
                d = 1;
            } else {
                int x1 = bx * 16 + 16 > w ? w : bx * 16 + 16;
                int y1 = by * 16 + 16 > h ? h : by * 16 + 16;
                for (y = by * 16; y < y1 && !d; y++)
                    for (x = bx * 16; x < x1; x++)
                        if (cur[off[0] + (size_t)y * st[0] + x] != prev[off[0] + (size_t)y * st[0] + x]) { d = 1; break; }
                for (y = by * 8; y < (y1 + 1) / 2 && !d; y++)
                    for (x = bx * 8; x < (x1 + 1) / 2; x++) {
                        if (fmt == SKO_OUT_I420) {
                            size_t u = off[1] + (size_t)y * st[1] + x, v = off[2] + (size_t)y * st[2] + x;
                            if (cur[u] != prev[u] || cur[v] != prev[v]) { d = 1; break; }
                        } else {
                            size_t uv = off[1] + (size_t)y * st[1] + 2 * x;
                            if (cur[uv] != prev[uv] || cur[uv + 1] != prev[uv + 1]) { d = 1; break; }
                        }
                    }
            }
            map[by * mbw + bx] = (uint8_t)d;
            count += d;
        }
    }
    return count;
}

/* ------------------------------------------------------------------ policy */

void sko_policy_update(int n, const uint8_t *map, uint32_t *state, uint8_t *policy,
                       int threshold, int duration, int trigger)
{
    int i;
    for (i = 0; i < n; i++) {
        uint32_t st = state[i];
        uint32_t streak = st & 0xff, hot = (st >> 8) & 0xff, still = (st >> 16) & 0xff, dirty = (st >> 24) & 1;
        uint32_t pol = 0;
        if (map[i]) {
            streak = streak < 255 ? streak + 1 : 255;
            still = 0;
            dirty = 1;
            if (streak >= (uint32_t)threshold || hot > 0)
                hot = (uint32_t)duration;
            pol = 1;
        } else {
            streak = 0;
            still = still < 255 ? still + 1 : 255;
            if (hot > 0)
                hot--;
            if (dirty && hot == 0 && still >= (uint32_t)trigger) {
                pol = 4;
                dirty = 0;
            }
        }
        if (hot > 0)
            pol |= 2;
        state[i] = streak | (hot << 8) | (still << 16) | (dirty << 24);
        policy[i] = (uint8_t)pol;
    }
}
