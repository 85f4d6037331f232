// sk_host_math.h — host-side arithmetic shared by the session runtime (sk_session.cu) and the
// allocation-free entry point (sk_noalloc.cu): integer colour matrix, dp4a coefficient packing,
// output layout and the 2-tap resampler table.  Header-only, no allocation (callers own the arrays).
#pragma once
#include "../../include/selkies_preencode.h"
#include "sk_kernels.h"

#include <cmath>
#include <cstdint>

namespace skhost {

inline int ru(int v, int a) { return (v + a - 1) / a * a; }

// Integer colour matrix at 8 fractional bits: full-range RGB → limited-range Y'CbCr (scale
// 219/224/224 over 255), each entry rounded to nearest-even — the derivation GStreamer's converter
// performs in double before its 8-bit table path.  c[row Y,U,V][col R,G,B].
inline void matrix_coeffs(int matrix, int c[3][3], unsigned switches = 0)
{
    double kr = 0.299, kb = 0.114;
    if (matrix == SK_MATRIX_BT709) { kr = 0.2126; kb = 0.0722; }
    else if (matrix == SK_MATRIX_BT2020) { kr = 0.2627; kb = 0.0593; }
    const double kg = 1.0 - kr - kb;
    const double inv255 = 1 / ((float)255);
    const double ys = 219.0 * inv255 * 256.0, cs = 224.0 * inv255 * 256.0;
    const double ub = 1 / (2 * (1 - kb)), vr = 1 / (2 * (1 - kr));
    const double m[3][3] = {
        {kr * ys, kg * ys, kb * ys},
        {-ub * kr * cs, -ub * kg * cs, ub * (1 - kb) * cs},
        {vr * (1 - kr) * cs, -vr * kg * cs, -vr * kb * cs}};
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++)
            c[i][j] = (switches & SK_PS_COEFF_HALF_UP) ? (int)std::round(m[i][j]) : (int)std::nearbyint(m[i][j]);
}

// c = {R, G, B}; byte k of the result multiplies byte k of the pixel word (dp4a operand).
inline uint32_t pack_coeffs(const int c[3], int in_fmt)
{
    const uint32_t r = (uint32_t)(c[0] & 0xff), g = (uint32_t)(c[1] & 0xff), b = (uint32_t)(c[2] & 0xff);
    if (in_fmt == SK_IN_BGRX) return b | (g << 8) | (r << 16);
    return r | (g << 8) | (b << 16);
}

// 2-tap vertical table of the linear resampler (videoscale method=bilinear): centre-aligned source
// position, triangle filter (stretched by out/in when shrinking in the GST_LQ variant), two taps
// renormalised, 6-bit integer taps found by a bias search so that they sum to 64.
// off[j] = first source row, tap[j] = weight of the second row, j < out_size.
inline void vertical_taps(int in_size, int out_size, int variant, int32_t *off, int16_t *tap)
{
    const double ratio = (double)in_size / out_size;
    const double stretch = (variant == SK_SCALE_VARIANT_GST_LQ && ratio > 1.0) ? 1.0 / ratio : 1.0;
    for (int j = 0; j < out_size; j++) {
        double pos = (j + 0.5) / out_size * in_size - 0.5;
        pos = std::fmin(std::fmax(pos, 0.0), (double)(in_size - 1));
        int first = (int)std::floor(pos);
        double w[2];
        for (int l = 0; l < 2; l++) {
            const double a = std::fabs(pos - (first + l)) * stretch;
            w[l] = a < 1.0 ? 1.0 - a : 0.0;
        }
        const double sum = w[0] + w[1];
        w[0] /= sum;  w[1] /= sum;
        if (first > in_size - 2) { w[1] += w[0]; w[0] = 0.0; first = in_size - 2; }
        double lo = 0.0, hi = 1.0, bias = 0.5;
        for (int it = 0; it < 64; it++) {
            const int s = (int)std::floor(bias + w[0] * 64.0) + (int)std::floor(bias + w[1] * 64.0);
            if (s == 64 || lo == hi) break;
            if (s < 64) { if (bias > lo) lo = bias; bias += (hi - lo) / 2; }
            else        { if (bias < hi) hi = bias; bias -= (hi - lo) / 2; }
        }
        off[j] = first;
        tap[j] = (int16_t)std::floor(bias + w[1] * 64.0);
    }
}

// Everything of SkGeom that follows from the configuration alone (GstVideoInfo default layout,
// matrix, scale increments).  in_stride and the tiled-kernel fields are the caller's business.
// Returns the size in bytes of one packed output frame.
inline size_t fill_geometry(SkGeom &g, int in_w, int in_h, int out_w, int out_h, int in_format, int out_format,
                            int matrix, int chroma_site, bool damage, unsigned switches = 0)
{
    g.in_w = in_w;  g.in_h = in_h;  g.out_w = out_w;  g.out_h = out_h;
    const int h2 = ru(out_h, 2);
    g.y_stride = ru(out_w, 4);
    size_t frame_bytes;
    if (out_format == SK_OUT_I420) {
        g.c_stride = ru(ru(out_w, 2) / 2, 4);
        g.off_c0 = (uint32_t)g.y_stride * h2;
        g.off_c1 = g.off_c0 + (uint32_t)g.c_stride * (h2 / 2);
        frame_bytes = (size_t)g.off_c1 + (size_t)g.c_stride * (h2 / 2);
    } else {
        g.c_stride = g.y_stride;
        g.off_c0 = (uint32_t)g.y_stride * h2;
        g.off_c1 = 0;
        frame_bytes = (size_t)g.off_c0 + (size_t)g.c_stride * (h2 / 2);
    }
    g.mb_w = (out_w + 15) / 16;
    g.mb_h = (out_h + 15) / 16;
    int m[3][3];
    matrix_coeffs(matrix, m, switches);
    g.cy = pack_coeffs(m[0], in_format);
    g.cu = pack_coeffs(m[1], in_format);
    g.cv = pack_coeffs(m[2], in_format);
    g.out_fmt = out_format;
    g.site = chroma_site;
    g.damage = damage ? 1 : 0;
    g.scale_h = in_w != out_w;
    g.scale_v = in_h != out_h;
    // upstream chain_scale(): in front of the matrix while out pixels <= in pixels, behind it otherwise
    g.scale_after = ((int64_t)out_w * out_h > (int64_t)in_w * in_h) ? 1 : 0;
    g.h_inc = out_w > 1 ? (int32_t)((((int64_t)in_w - 1) << 16) / (out_w - 1) - ((switches & SK_PS_HINC_PLAIN) ? 0 : 1)) : 0;
    g.h_first = ((int64_t)out_w * in_h <= (int64_t)in_w * out_h) ? 1 : 0;
    if ((switches & SK_PS_V_FIRST_ON_TIE) && (int64_t)out_w * in_h == (int64_t)in_w * out_h) g.h_first = 0;
    // what the generic kernel has to honour per pixel
    g.switches = switches & (SK_PS_MATRIX_ROUND | SK_PS_CS_EDGE_KEEP | SK_PS_CHROMA_V_FIRST | SK_PS_CHROMA_BOX4);
    g.y_off = 4096u + ((switches & SK_PS_MATRIX_ROUND) ? 128u : 0u);
    g.c_off = 32768u + ((switches & SK_PS_MATRIX_ROUND) ? 128u : 0u);
    return frame_bytes;
}

}  // namespace skhost
