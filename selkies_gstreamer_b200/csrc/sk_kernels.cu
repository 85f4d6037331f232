// sk_kernels.cu — fused pre-encode kernels for sm_100a (B200).
//
// One pass per frame replaces what the reference's pipeline runs between
// ximagesrc and the encoder (src/selkies/legacy/gstwebrtc_app.py:611-617
// videoconvert → NV12/I420; :263-284 cudaupload+cudaconvert → NV12):
//
//   BGRx/RGBx read (128-bit coalesced) → integer 3x4 matrix (dp4a, 8 fractional
//   bits, no rounding term) → 4:2:0 chroma (h filter with a warp-shuffle halo,
//   then v average) → Y/UV stores in GstVideoInfo layout → compare against the
//   previous output frame (kept resident in HBM) → per-16x16 damage flags by
//   warp ballot + shared-memory OR, damaged count by one atomic per block.
//
// HBM-bound integer work: no tensor cores on purpose.  Algorithmic traffic is
// 4 B/px in + 1.5 B/px out + 1.5 B/px previous frame (+1 B per macroblock).
#include "sk_kernels.h"
#include "../../include/selkies_preencode.h"     // SK_PS_* switch bits

namespace {

constexpr int kThreads = 256;         // 8 warps per CTA
constexpr uint32_t kFull = 0xffffffffu;

__device__ __forceinline__ uint32_t dp4a_uu(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t d;
    asm("dp4a.u32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ uint32_t dp4a_us(uint32_t a, uint32_t b, int c)
{
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return (uint32_t)d;
}

// ---- loads -----------------------------------------------------------------
template <int HINT> __device__ __forceinline__ uint4 ld16(const void *p)
{
    if (HINT & 1) return __ldcs(reinterpret_cast<const uint4 *>(p));
    return __ldg(reinterpret_cast<const uint4 *>(p));
}
template <int HINT> __device__ __forceinline__ uint2 ld8(const void *p)
{
    if (HINT & 1) return __ldcs(reinterpret_cast<const uint2 *>(p));
    return __ldg(reinterpret_cast<const uint2 *>(p));
}
template <int HINT> __device__ __forceinline__ uint32_t ld4(const void *p)
{
    if (HINT & 1) return __ldcs(reinterpret_cast<const uint32_t *>(p));
    return __ldg(reinterpret_cast<const uint32_t *>(p));
}
template <int HINT> __device__ __forceinline__ uint32_t ld2(const void *p)
{
    if (HINT & 1) return __ldcs(reinterpret_cast<const unsigned short *>(p));
    return __ldg(reinterpret_cast<const unsigned short *>(p));
}
__device__ __forceinline__ uint32_t ld1(const void *p)
{
    return __ldg(reinterpret_cast<const unsigned char *>(p));
}

// nv = number of valid pixels of this quad: 4, 2 or 0 (W is even).
template <int HINT> __device__ __forceinline__ uint4 load_quad(const uint8_t *p, int nv)
{
    uint4 v = make_uint4(0, 0, 0, 0);
    if (nv == 4) {
        v = ld16<HINT>(p);
    } else if (nv == 2) {
        uint2 t = ld8<HINT>(p);
        v.x = t.x;  v.y = t.y;
    }
    return v;
}
template <int HINT> __device__ __forceinline__ uint32_t load_word(const uint8_t *p, int nv)
{
    if (nv == 4) return ld4<HINT>(p);
    if (nv == 2) return ld2<HINT>(p);
    return 0;
}
// HINT bit 1 (values 2, 3): streaming (evict-first) stores for the interior path
template <int HINT> __device__ __forceinline__ void st4(uint8_t *p, uint32_t w)
{
    if (HINT & 2) __stcs(reinterpret_cast<uint32_t *>(p), w);
    else *reinterpret_cast<uint32_t *>(p) = w;
}
__device__ __forceinline__ void store_word(uint8_t *p, uint32_t w, int nv)
{
    if (nv == 4) *reinterpret_cast<uint32_t *>(p) = w;
    else if (nv == 2) *reinterpret_cast<unsigned short *>(p) = (unsigned short)w;
}

// ---- arithmetic -------------------------------------------------------------
// Four pixels → packed Y word and four (U | V<<16) lane pairs.
__device__ __forceinline__ void matrix_quad(const uint4 &q, const SkGeom &g, uint32_t &yw, uint32_t uv[4])
{
    const uint32_t p[4] = {q.x, q.y, q.z, q.w};
    uint32_t ay[4];
#pragma unroll
    for (int i = 0; i < 4; i++) {
        ay[i] = dp4a_uu(p[i], g.cy, 4096u);
        uint32_t au = dp4a_us(p[i], g.cu, 32768);
        uint32_t av = dp4a_us(p[i], g.cv, 32768);
        uv[i] = __byte_perm(au, av, 0x7531);   // (au>>8)&0xff | ((av>>8)&0xff)<<16 (sums < 65536)
    }
    uint32_t t01 = __byte_perm(ay[0], ay[1], 0x0051);
    uint32_t t23 = __byte_perm(ay[2], ay[3], 0x0051);
    yw = __byte_perm(t01, t23, 0x5410);
}

// Same with the accumulator offsets taken from the launch geometry (parity-switch path only: the
// fast kernels keep them as immediates).
__device__ __forceinline__ void matrix_quad_sw(const uint4 &q, const SkGeom &g, uint32_t &yw, uint32_t uv[4])
{
    const uint32_t p[4] = {q.x, q.y, q.z, q.w};
    uint32_t ay[4];
#pragma unroll
    for (int i = 0; i < 4; i++) {
        ay[i] = dp4a_uu(p[i], g.cy, g.y_off);
        uint32_t au = dp4a_us(p[i], g.cu, (int)g.c_off);
        uint32_t av = dp4a_us(p[i], g.cv, (int)g.c_off);
        uv[i] = __byte_perm(au, av, 0x7531);
    }
    uint32_t t01 = __byte_perm(ay[0], ay[1], 0x0051);
    uint32_t t23 = __byte_perm(ay[2], ay[3], 0x0051);
    yw = __byte_perm(t01, t23, 0x5410);
}

// Horizontal 4:2:0 filter on 16-bit lanes; x = first pixel of the quad.
template <int SITE>
__device__ __forceinline__ void chroma_h(const uint32_t uv[4], uint32_t left, int x, int w,
                                         uint32_t &h0, uint32_t &h1)
{
    if (SITE == 0) {
        h0 = ((uv[0] + uv[1] + 0x00010001u) >> 1) & 0x00ff00ffu;
        h1 = ((uv[2] + uv[3] + 0x00010001u) >> 1) & 0x00ff00ffu;
    } else {
        const bool closing = w > 2;
        uint32_t l0 = (x == 0) ? uv[0] : left;                      // [3 1] opening tap
        uint32_t r0 = (closing && x == w - 2) ? uv[0] : uv[1];      // [1 3] closing tap (partial quad)
        uint32_t r1 = (closing && x + 2 == w - 2) ? uv[2] : uv[3];  // [1 3] closing tap (full quad)
        h0 = ((l0 + 2u * uv[0] + r0 + 0x00020002u) >> 2) & 0x00ff00ffu;
        h1 = ((uv[1] + 2u * uv[2] + r1 + 0x00020002u) >> 2) & 0x00ff00ffu;
    }
}

// Parity-switch variant: SK_PS_CS_EDGE_KEEP leaves the first and the last chroma sample of a row unfiltered.
template <int SITE>
__device__ __forceinline__ void chroma_h_sw(const uint32_t uv[4], uint32_t left, int x, int w, bool keep_edges,
                                            uint32_t &h0, uint32_t &h1)
{
    chroma_h<SITE>(uv, left, x, w, h0, h1);
    if (SITE == 1 && keep_edges) {
        if (x == 0 || x == w - 2) h0 = uv[0];
        if (x + 2 == w - 2) h1 = uv[2];
    }
}

// Vertical average of two filtered rows and byte gather.
// NV12 → U0 V0 U1 V1 ; I420 → U0 U1 V0 V1 (low half to the U plane, high half to V).
template <int FMT>
__device__ __forceinline__ uint32_t chroma_v(uint32_t a0, uint32_t a1, uint32_t b0, uint32_t b1)
{
    uint32_t v0 = (a0 + b0 + 0x00010001u) >> 1;
    uint32_t v1 = (a1 + b1 + 0x00010001u) >> 1;
    return __byte_perm(v0, v1, FMT == 0 ? 0x6420 : 0x6240);
}

// Block-level tail of the damage reduction: one atomic per CTA into the running
// count, then a wrapping ticket; the last CTA of this frame publishes the total
// and leaves both scratch words at zero for the next launch.
__device__ __forceinline__ void finish_count(const SkFrame &f, uint32_t n)
{
    if (n) atomicAdd(&f.scratch[0], n);
    __threadfence();
    const uint32_t nblk = gridDim.x * gridDim.y;
    uint32_t t = atomicInc(&f.scratch[1], nblk - 1);
    if (t == nblk - 1) {
        __threadfence();
        *f.count_out = atomicExch(&f.scratch[0], 0u);
    }
}

// Temporal policy of one macroblock: consecutive-damage streak, "hot" countdown
// while a block keeps changing, still-frame counter that triggers one paint-over. Flags:
// 1 = damaged this frame, 2 = hot (persistently damaged), 4 = paint over now.
__device__ __forceinline__ void policy_update(const SkFrame &f, const SkGeom &g, size_t idx, bool damaged, uint32_t st)
{
    if (f.state == nullptr) return;
    uint32_t streak = st & 0xffu, hot = (st >> 8) & 0xffu, still = (st >> 16) & 0xffu, dirty = (st >> 24) & 1u;
    uint32_t pol = 0;
    if (damaged) {
        streak = min(streak + 1u, 255u);
        still = 0;
        dirty = 1;
        if (streak >= (uint32_t)g.pol_threshold || hot > 0) hot = (uint32_t)g.pol_duration;
        pol = 1;
    } else {
        streak = 0;
        still = min(still + 1u, 255u);
        if (hot > 0) hot--;
        if (dirty && hot == 0 && still >= (uint32_t)g.pol_trigger) {
            pol = 4;
            dirty = 0;
        }
    }
    if (hot > 0) pol |= 2;
    f.state[idx] = streak | (hot << 8) | (still << 16) | (dirty << 24);
    f.policy[idx] = (uint8_t)pol;
}
// The state word of the macroblock a thread will publish, fetched at kernel entry so that its
// latency is hidden behind the pixel work instead of sitting on the CTA's tail.
__device__ __forceinline__ uint32_t policy_prefetch(const SkFrame &f, const SkGeom &g, bool mine, int bx, int by)
{
    if (!mine || f.state == nullptr || bx >= g.mb_w || by >= g.mb_h) return 0;
    return f.state[(size_t)by * g.mb_w + bx];
}

// ---- bulk-copy (TMA engine) helpers: 1-D cp.async.bulk global -> shared, completion on an mbarrier
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
// 2-D tiled tensor-map copy global -> shared (c0 = element column, c1 = row), completion on an mbarrier
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap *map, int c0, int c1, uint32_t bar)
{
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(bar) : "memory");
}
// the same box, only as far as L2 (no shared-memory destination, no completion to wait for)
__device__ __forceinline__ void tma_prefetch_2d(const CUtensorMap *map, int c0, int c1)
{
    asm volatile("cp.async.bulk.prefetch.tensor.2d.L2.global.tile [%0, {%1, %2}];"
                 ::"l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    uint32_t done = 0;
    for (int spin = 0; spin < (1 << 24) && !done; spin++) {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                     "selp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    }
    if (!done) __trap();        // never hang the GPU on a protocol error
}

__device__ __forceinline__ uint32_t lds32(uint32_t addr)
{
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ void sts32(uint32_t addr, uint32_t v)
{
    asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}

// ---- no-scale kernel ----------------------------------------------------------
// CTA = 8 warps.  A warp covers 128*QX pixels of 2*RP rows; lane l owns the
// quads at x = bx0 + q*128 + 4*l.  The CTA covers 16*RP rows = RP macroblock
// rows, so a macroblock is 4 adjacent lanes of 8/RP adjacent warps.
//
// FULL = the CTA's whole footprint lies inside the frame: every access is an
// unconditional, fully vectorised instruction.  Edge CTAs take the guarded path.
template <int RP, int QX, int FMT, int SITE, bool DAMAGE, int HINT, bool FULL>
__device__ __forceinline__ void convert_tile(const SkGeom &g, const SkFrame &f, const int bx0, const int y0,
                                             const int lane, const bool have_prev,
                                             uint32_t (&diff)[QX], int (&nvq)[QX],
                                             const uint8_t *tile = nullptr, uint32_t bar = 0, int warp = 0)
{
    constexpr int ROWS = 2 * RP;
    constexpr bool BULK = HINT == 4;            // input rows arrive in shared memory by cp.async.bulk
    constexpr int PITCH = 512 * QX + 16;        // bytes per staged row: 16-byte halo slot + the CTA's pixels
    const int w = g.out_w, h = g.out_h;
    const int x0 = bx0 + lane * 4;

    uint4 px[ROWS][QX];
    uint32_t halo[ROWS];
    uint32_t py[ROWS][QX];
    uint32_t pc[RP][QX];

#pragma unroll
    for (int q = 0; q < QX; q++) {
        const int nv = w - (x0 + q * 128);
        nvq[q] = FULL ? 4 : (nv >= 4 ? 4 : (nv > 0 ? 2 : 0));
        diff[q] = 0;
    }

    // 32-bit byte offsets (largest frame: 8192*8192*4 B input, 100 MB output)
    const uint32_t in_off = (uint32_t)y0 * (uint32_t)g.in_stride + 4u * (uint32_t)x0;
    const uint32_t y_off = (uint32_t)y0 * (uint32_t)g.y_stride + (uint32_t)x0;
    const uint32_t c_off = (uint32_t)(y0 >> 1) * (uint32_t)g.c_stride + (FMT == 0 ? (uint32_t)x0 : (uint32_t)(x0 >> 1));
    const uint8_t *in = f.in + in_off;
    const bool want_halo = SITE == 1 && lane == 0 && bx0 > 0;

    // ---- phase 1: issue every load of this thread up front; the narrow previous-frame loads go
    //      first so that ptxas keeps all of them ahead of the first dependent instruction
    if (DAMAGE) {
        const uint8_t *pyb = f.prev + y_off;
#pragma unroll
        for (int r = 0; r < ROWS; r++) {
            const bool rv = have_prev && (FULL || (y0 + r) < h);
            const uint8_t *row = pyb + (uint32_t)r * (uint32_t)g.y_stride;
#pragma unroll
            for (int q = 0; q < QX; q++) {
                if (FULL) py[r][q] = rv ? ld4<HINT>(row + q * 128) : 0u;
                else py[r][q] = load_word<HINT>(row + q * 128, rv ? nvq[q] : 0);
            }
        }
#pragma unroll
        for (int rp = 0; rp < RP; rp++) {
            const bool rv = have_prev && (FULL || (y0 + 2 * rp) < h);
            const uint32_t o = c_off + (uint32_t)rp * (uint32_t)g.c_stride;
#pragma unroll
            for (int q = 0; q < QX; q++) {
                const int nv = rv ? nvq[q] : 0;
                if (FMT == 0) {
                    const uint8_t *p = f.prev + g.off_c0 + o + q * 128;
                    if (FULL) pc[rp][q] = rv ? ld4<HINT>(p) : 0u;
                    else pc[rp][q] = load_word<HINT>(p, nv);
                } else {
                    const uint8_t *pu = f.prev + g.off_c0 + o + q * 64;
                    const uint8_t *pv = f.prev + g.off_c1 + o + q * 64;
                    uint32_t a = 0, b = 0;
                    if (nv == 4) { a = ld2<HINT>(pu); b = ld2<HINT>(pv); }
                    else if (nv == 2) { a = ld1(pu); b = ld1(pv); }
                    pc[rp][q] = a | (b << 16);
                }
            }
        }
    }

    if (BULK && FULL) {
        mbar_wait(bar, 0);
#pragma unroll
        for (int r = 0; r < ROWS; r++) {
            const uint8_t *trow = tile + (warp * ROWS + r) * PITCH;
#pragma unroll
            for (int q = 0; q < QX; q++)
                px[r][q] = *reinterpret_cast<const uint4 *>(trow + 16 + lane * 16 + q * 512);
            halo[r] = want_halo ? *reinterpret_cast<const uint32_t *>(trow + 12) : 0u;
        }
    } else {
#pragma unroll
        for (int r = 0; r < ROWS; r++) {
            const bool rv = FULL || (y0 + r) < h;
            const uint8_t *row = in + (uint32_t)r * (uint32_t)g.in_stride;
#pragma unroll
            for (int q = 0; q < QX; q++) {
                if (FULL) px[r][q] = ld16<HINT>(row + q * 512);
                else px[r][q] = load_quad<HINT>(row + q * 512, rv ? nvq[q] : 0);
            }
            halo[r] = 0;
            if (want_halo && rv) halo[r] = ld4<0>(row - 4);   // lane 0: pixel left of the warp
        }
    }
    // ---- phase 2: arithmetic, stores, compare
    uint8_t *oy = f.out + y_off;
#pragma unroll
    for (int rp = 0; rp < RP; rp++) {
        uint32_t hh[2][QX][2];
#pragma unroll
        for (int rr = 0; rr < 2; rr++) {
            const int r = 2 * rp + rr;
            const bool rv = FULL || (y0 + r) < h;
            uint32_t uvq[QX][4];
            uint32_t yw[QX];
#pragma unroll
            for (int q = 0; q < QX; q++)
                matrix_quad(px[r][q], g, yw[q], uvq[q]);
            uint32_t halo_uv = 0;
            if (SITE == 1) {
                uint32_t au = dp4a_us(halo[r], g.cu, 32768);
                uint32_t av = dp4a_us(halo[r], g.cv, 32768);
                halo_uv = __byte_perm(au, av, 0x7531);
            }
            uint8_t *orow = oy + (uint32_t)r * (uint32_t)g.y_stride;
#pragma unroll
            for (int q = 0; q < QX; q++) {
                const int x = x0 + q * 128;
                uint32_t left = 0;
                if (SITE == 1) {
                    left = __shfl_up_sync(kFull, uvq[q][3], 1);
                    uint32_t wrap = (q > 0) ? __shfl_sync(kFull, uvq[q > 0 ? q - 1 : 0][3], 31) : halo_uv;
                    if (lane == 0) left = wrap;
                }
                chroma_h<SITE>(uvq[q], left, x, w, hh[rr][q][0], hh[rr][q][1]);
                if (FULL) st4<HINT>(orow + q * 128, yw[q]);
                else if (rv) store_word(orow + q * 128, yw[q], nvq[q]);
                if (DAMAGE && rv) {
                    if (FULL) diff[q] |= yw[q] ^ py[r][q];
                    else diff[q] |= (yw[q] ^ py[r][q]) & (nvq[q] == 4 ? 0xffffffffu : 0x0000ffffu);
                }
            }
        }
        const bool rv = FULL || (y0 + 2 * rp) < h;
        const uint32_t o = c_off + (uint32_t)rp * (uint32_t)g.c_stride;
#pragma unroll
        for (int q = 0; q < QX; q++) {
            uint32_t cw = chroma_v<FMT>(hh[0][q][0], hh[0][q][1], hh[1][q][0], hh[1][q][1]);
            const int nv = rv ? nvq[q] : 0;
            if (FMT == 0) {
                uint8_t *p = f.out + g.off_c0 + o + q * 128;
                if (FULL) st4<HINT>(p, cw);
                else store_word(p, cw, nv);
                if (DAMAGE && rv) diff[q] |= (cw ^ pc[rp][q]) & ((FULL || nvq[q] == 4) ? 0xffffffffu : 0x0000ffffu);
            } else {
                uint8_t *pu = f.out + g.off_c0 + o + q * 64;
                uint8_t *pv = f.out + g.off_c1 + o + q * 64;
                if (nv == 4) {
                    *reinterpret_cast<unsigned short *>(pu) = (unsigned short)cw;
                    *reinterpret_cast<unsigned short *>(pv) = (unsigned short)(cw >> 16);
                } else if (nv == 2) {
                    *pu = (uint8_t)cw;
                    *pv = (uint8_t)(cw >> 16);
                }
                if (DAMAGE && rv) diff[q] |= (cw ^ pc[rp][q]) & ((FULL || nvq[q] == 4) ? 0xffffffffu : 0x00ff00ffu);
            }
        }
    }
}

// Register budget: RP*QX <= 2 must stay at 5 CTAs/SM (<= 48 registers) — the sweep in profiles/ shows the
// occupancy cliff costs ~12 % of the HBM rate.
template <int RP, int QX, int FMT, int SITE, bool DAMAGE, int HINT, bool POLICY>
__global__ void __launch_bounds__(kThreads, (RP * QX <= 2) ? 5 : 1)
sk_convert_kernel(const __grid_constant__ SkGeom g, const __grid_constant__ SkBatch batch)
{
    constexpr int ROWS = 2 * RP;
    constexpr int NMB = RP * QX * 8;      // macroblocks per CTA
    // Launches alternate the traversal direction (g.reverse): what the previous launch wrote last —
    // the "previous frame" of the sessions / rows it finished with — is still in the 126 MB L2 and is
    // read first by this one.
    const unsigned bix = g.reverse ? gridDim.x - 1 - blockIdx.x : blockIdx.x;
    const unsigned biy = g.reverse ? gridDim.y - 1 - blockIdx.y : blockIdx.y;
    const unsigned biz = g.reverse ? gridDim.z - 1 - blockIdx.z : blockIdx.z;
    const SkFrame &f = batch.f[biz];
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int bx0 = bix * (128 * QX);
    const int by0 = biy * (16 * RP);
    const int y0 = by0 + warp * ROWS;
    const bool have_prev = DAMAGE && (f.prev != nullptr);
    const bool full = (bx0 + 128 * QX <= g.out_w) && (by0 + 16 * RP <= g.out_h);
    if (DAMAGE && blockIdx.x == 0 && blockIdx.y == 0 && (int)threadIdx.x < f.count_zero_n) f.count_zero[threadIdx.x] = 0u;

    uint32_t pol_state = 0;
    if (DAMAGE && POLICY) {
        const int t = threadIdx.x;
        pol_state = policy_prefetch(f, g, t < NMB, bix * (8 * QX) + ((t % (QX * 8)) / 8) * 8 + t % 8,
                                    biy * RP + t / (QX * 8));
    }
    uint32_t diff[QX];
    int nvq[QX];
    constexpr bool BULK = HINT == 4;
    constexpr int PITCH = 512 * QX + 16;
    __shared__ __align__(16) uint8_t s_tile[BULK ? 16 * RP * PITCH : 16];
    __shared__ __align__(8) unsigned long long s_bar;
    const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&s_bar);
    if (BULK && full) {
        // The TMA engine copies the CTA's 16*RP input rows (plus the 16-byte halo slot on the left)
        // into shared memory; one lane per row issues, lane 0 arms the barrier with the byte total.
        if (threadIdx.x == 0) mbar_init(bar, 1);
        __syncthreads();
        if (warp == 0) {
            const uint32_t lead = bx0 > 0 ? 16u : 0u;
            const uint32_t row_bytes = 512u * QX + lead;
            if (lane == 0) mbar_expect_tx(bar, row_bytes * 16u * RP);
            __syncwarp();
            if (lane < 16 * RP) {
                const uint8_t *src = f.in + (size_t)(by0 + lane) * (size_t)g.in_stride + 4 * (size_t)bx0 - lead;
                bulk_g2s((uint32_t)__cvta_generic_to_shared(s_tile) + lane * PITCH + 16 - lead, src, row_bytes, bar);
            }
        }
    }
    if (full) convert_tile<RP, QX, FMT, SITE, DAMAGE, HINT, true>(g, f, bx0, y0, lane, have_prev, diff, nvq, s_tile, bar, warp);
    else      convert_tile<RP, QX, FMT, SITE, DAMAGE, HINT, false>(g, f, bx0, y0, lane, have_prev, diff, nvq);

    // ---- phase 3: per-macroblock flags
    if (DAMAGE) {
        __shared__ uint32_t s_mask[8][QX];
        const bool rows_valid = y0 < g.out_h;
#pragma unroll
        for (int q = 0; q < QX; q++) {
            bool d = rows_valid && nvq[q] > 0 && (diff[q] != 0 || !have_prev);
            uint32_t m = __ballot_sync(kFull, d);
            if (lane == 0) s_mask[warp][q] = m;
        }
        __syncthreads();
        constexpr int WPM = 8 / RP;                 // warps per macroblock row
        bool hit = false;
        if (threadIdx.x < NMB) {
            const int t = threadIdx.x;
            const int mbrow = t / (QX * 8);
            const int q = (t % (QX * 8)) / 8;
            const int m = t % 8;
            uint32_t acc = 0;
#pragma unroll
            for (int k = 0; k < WPM; k++)
                acc |= s_mask[mbrow * WPM + k][q];
            const bool flag = ((acc >> (4 * m)) & 0xfu) != 0;
            const int bx = bix * (8 * QX) + q * 8 + m;
            const int by = biy * RP + mbrow;
            if (bx < g.mb_w && by < g.mb_h) {
                f.map[(size_t)by * g.mb_w + bx] = flag ? 1 : 0;
                if (POLICY) policy_update(f, g, (size_t)by * g.mb_w + bx, flag, pol_state);
                hit = flag;
            }
        }
        // Block count by the barrier itself, then ONE fire-and-forget atomic (RED) per CTA into a word the
        // host zeroed before the launch.  (A returning atomic here — the ticket scheme of finish_count —
        // kept every CTA's registers allocated ~1 us longer and cost 4.7 % of the kernel rate.)
        const int total = __syncthreads_count(hit);
        if (threadIdx.x == 0 && total != 0 && f.count_out != nullptr) atomicAdd(f.count_out, (uint32_t)total);
    }
}

// ---- warp-tile variant of the no-scale kernel ---------------------------------------------
// One warp owns a whole 256 x 16 tile (16 macroblocks) and walks its 8 row pairs in a loop, so the
// damage flags never leave the warp: no shared memory, no __syncthreads, and a CTA lives 8x longer.
template <int FMT, int SITE, bool DAMAGE, bool POLICY>
__global__ void __launch_bounds__(kThreads, 5)
sk_convert_warp_kernel(const __grid_constant__ SkGeom g, const __grid_constant__ SkBatch batch)
{
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int tiles_x = (g.out_w + 255) / 256;
    const int tiles = tiles_x * ((g.out_h + 15) / 16);
    int wid = blockIdx.x * 8 + warp;                    // this warp's tile within the frame
    if (wid >= tiles) return;                           // no block-wide barrier below
    const unsigned biz = g.reverse ? gridDim.z - 1 - blockIdx.z : blockIdx.z;
    if (g.reverse) wid = tiles - 1 - wid;
    const SkFrame &f = batch.f[biz];
    const int tx = wid % tiles_x, ty = wid / tiles_x;
    const int bx0 = tx * 256, by0 = ty * 16;
    const bool have_prev = DAMAGE && (f.prev != nullptr);
    const bool full = (bx0 + 256 <= g.out_w) && (by0 + 16 <= g.out_h);

    if (DAMAGE && blockIdx.x == 0 && (int)threadIdx.x < f.count_zero_n) f.count_zero[threadIdx.x] = 0u;
    uint32_t pol_state = 0;
    if (DAMAGE && POLICY) pol_state = policy_prefetch(f, g, lane < 16, tx * 16 + lane, ty);

    uint32_t acc[2] = {0, 0};
    int nvq[2] = {0, 0};
#pragma unroll 1
    for (int it = 0; it < 8; it++) {
        uint32_t d[2];
        const int y0 = by0 + 2 * it;
        if (full) convert_tile<1, 2, FMT, SITE, DAMAGE, 0, true>(g, f, bx0, y0, lane, have_prev, d, nvq);
        else      convert_tile<1, 2, FMT, SITE, DAMAGE, 0, false>(g, f, bx0, y0, lane, have_prev, d, nvq);
        acc[0] |= d[0];
        acc[1] |= d[1];
    }
    if (DAMAGE) {
        // 4 adjacent lanes = one macroblock; lanes 0..7 publish the blocks of q = 0, lanes 8..15 those of q = 1
        const uint32_t m0 = __ballot_sync(kFull, nvq[0] > 0 && (acc[0] != 0 || !have_prev));
        const uint32_t m1 = __ballot_sync(kFull, nvq[1] > 0 && (acc[1] != 0 || !have_prev));
        bool hit = false;
        if (lane < 16) {
            const uint32_t m = lane < 8 ? m0 : m1;
            const bool flag = ((m >> (4 * (lane & 7))) & 0xfu) != 0;
            const int bx = tx * 16 + lane;
            if (bx < g.mb_w && ty < g.mb_h) {
                f.map[(size_t)ty * g.mb_w + bx] = flag ? 1 : 0;
                if (POLICY) policy_update(f, g, (size_t)ty * g.mb_w + bx, flag, pol_state);
                hit = flag;
            }
        }
        const uint32_t votes = __ballot_sync(kFull, hit);
        if (lane == 0 && votes != 0 && f.count_out != nullptr) atomicAdd(f.count_out, (uint32_t)__popc(votes));
    }
}

// ---- scale kernel ---------------------------------------------------------------
// 2-tap shrink in the RGB domain before the matrix (order as upstream's
// chain_scale: h first when out_w*in_h <= in_w*out_h), then the same
// matrix/chroma/damage tail.  Thread = 4 output pixels x 2 rows.
// 2-tap blends on the dot-product unit: the two pixels' bytes are interleaved per channel
// ([aB bB aG bG], [aR bR . .]) and dp2a multiplies each byte pair by the 16-bit weight pair.
// Both weight pairs sum to 256, so every result is byte 1 of its accumulator (X is dropped: the
// matrix ignores it).
__device__ __forceinline__ uint32_t blend2(uint32_t a, uint32_t b, uint32_t wpair, uint32_t round)
{
    const uint32_t t0 = __byte_perm(a, b, 0x5140);     // aB bB aG bG
    const uint32_t t1 = __byte_perm(a, b, 0x7362);     // aR bR aX bX
    const uint32_t rb = __dp2a_lo(wpair, t0, round);
    const uint32_t rg = __dp2a_hi(wpair, t0, round);
    const uint32_t rr = __dp2a_lo(wpair, t1, round);
    return __byte_perm(__byte_perm(rb, rg, 0x0051), rr, 0x7510);
}
// per byte (a*(256-f) + b*f) >> 8
__device__ __forceinline__ uint32_t lerp_h2(uint32_t a, uint32_t b, uint32_t f)
{
    return blend2(a, b, (256u - f) | (f << 16), 0u);
}
// per byte (a*(64-p) + b*p + 32) >> 6  ==  (a*(256-4p) + b*4p + 128) >> 8   (0 <= p <= 64)
__device__ __forceinline__ uint32_t lerp_v2(uint32_t a, uint32_t b, uint32_t p)
{
    return blend2(a, b, (256u - 4u * p) | ((4u * p) << 16), 128u);
}

__device__ __forceinline__ uint32_t lerp_h(uint32_t a, uint32_t b, uint32_t f) { return lerp_h2(a, b, f); }
__device__ __forceinline__ uint32_t lerp_v(uint32_t a, uint32_t b, uint32_t p) { return lerp_v2(a, b, p); }

// One source pixel as the scaler sees it.  While the picture does not grow the scaler runs in front of the
// matrix, on the raw B,G,R,X bytes; when it grows (g.scale_after) upstream scales behind the matrix, so each
// source pixel is converted first and the 2-tap blends run on its Y,U,V bytes.
__device__ __forceinline__ uint32_t to_yuv_word(uint32_t px, const SkGeom &g)
{
    const uint32_t ay = dp4a_uu(px, g.cy, g.y_off);
    const uint32_t au = dp4a_us(px, g.cu, (int)g.c_off);
    const uint32_t av = dp4a_us(px, g.cv, (int)g.c_off);
    return __byte_perm(__byte_perm(ay, au, 0x0051), av, 0x7510);      // Y | U<<8 | V<<16 (sums < 65536)
}
__device__ __forceinline__ uint32_t source_px(const uint8_t *p, const SkGeom &g)
{
    const uint32_t v = ld4<0>(p);
    return g.scale_after ? to_yuv_word(v, g) : v;
}

__device__ __forceinline__ uint32_t sample_px(const uint8_t *in, const SkGeom &g, int x, int o, int o2, uint32_t p)
{
    int x0 = x;
    uint32_t fr = 0;
    if (g.scale_h) {
        int32_t acc = x * g.h_inc;
        x0 = acc >> 16;
        fr = (uint32_t)(acc >> 8) & 0xffu;
    }
    const uint8_t *r0 = in + (size_t)o * g.in_stride + 4 * (size_t)x0;
    const uint8_t *r1 = in + (size_t)o2 * g.in_stride + 4 * (size_t)x0;
    uint32_t a0 = source_px(r0, g);
    uint32_t b0 = g.scale_v ? source_px(r1, g) : 0;
    if (!g.scale_h)
        return g.scale_v ? lerp_v(a0, b0, p) : a0;
    // (SK_PS_HINC_PLAIN can put the last output pixel exactly on the last source pixel: its second tap has
    //  weight 0 and must not be fetched from beyond the row)
    const int next = (x0 + 1 < g.in_w) ? 4 : 0;
    uint32_t a1 = source_px(r0 + next, g);
    uint32_t b1 = g.scale_v ? source_px(r1 + next, g) : 0;
    if (!g.scale_v)
        return lerp_h(a0, a1, fr);
    if (g.h_first)
        return lerp_v(lerp_h(a0, a1, fr), lerp_h(b0, b1, fr), p);
    return lerp_h(lerp_v(a0, b0, p), lerp_v(a1, b1, p), fr);
}

// Four already converted pixels (Y | U<<8 | V<<16) -> packed Y word and the four (U | V<<16) pairs.
__device__ __forceinline__ void split_yuv_quad(const uint4 &q, uint32_t &yw, uint32_t uv[4])
{
    yw = __byte_perm(__byte_perm(q.x, q.y, 0x0040), __byte_perm(q.z, q.w, 0x0040), 0x5410);
    uv[0] = __byte_perm(q.x, 0u, 0x4241);
    uv[1] = __byte_perm(q.y, 0u, 0x4241);
    uv[2] = __byte_perm(q.z, 0u, 0x4241);
    uv[3] = __byte_perm(q.w, 0u, 0x4241);
}

// SW = true: the parity-switch instantiation (g.switches, g.y_off / g.c_off honoured); SW = false: the
// arithmetic as restated, untouched.
template <int FMT, int SITE, bool DAMAGE, bool SW = false>
__global__ void __launch_bounds__(kThreads)
sk_scale_kernel(const __grid_constant__ SkGeom g, const __grid_constant__ SkBatch batch)
{
    const SkFrame &f = batch.f[blockIdx.z];
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int bx0 = blockIdx.x * 128;
    const int y0 = (blockIdx.y * 8 + warp) * 2;
    const int w = g.out_w, h = g.out_h;
    const int x = bx0 + lane * 4;
    const int nvx = (w - x) >= 4 ? 4 : ((w - x) > 0 ? 2 : 0);
    const bool rows_valid = y0 < h;
    const int nv = rows_valid ? nvx : 0;
    const bool have_prev = DAMAGE && (f.prev != nullptr);
    if (DAMAGE && blockIdx.x == 0 && blockIdx.y == 0 && (int)threadIdx.x < f.count_zero_n) f.count_zero[threadIdx.x] = 0u;

    uint32_t hh[2][2];
    uint32_t uvr[2][4] = {}, leftr[2] = {};      // SW only
    uint32_t diff = 0;
#pragma unroll
    for (int rr = 0; rr < 2; rr++) {
        const int y = y0 + rr;
        int o = y, o2 = y;
        uint32_t p = 0;
        if (rows_valid && g.scale_v) {
            o = f.v_off[y];
            p = (uint32_t)f.v_tap[y];
            o2 = min(o + 1, g.in_h - 1);
        }
        uint4 q = make_uint4(0, 0, 0, 0);
        if (nv >= 2) {
            q.x = sample_px(f.in, g, x, o, o2, p);
            q.y = sample_px(f.in, g, x + 1, o, o2, p);
        }
        if (nv == 4) {
            q.z = sample_px(f.in, g, x + 2, o, o2, p);
            q.w = sample_px(f.in, g, x + 3, o, o2, p);
        }
        uint32_t yw, uv[4];
        if (g.scale_after) split_yuv_quad(q, yw, uv);
        else if (SW) matrix_quad_sw(q, g, yw, uv);
        else matrix_quad(q, g, yw, uv);
        uint32_t left = 0;
        if (SITE == 1) {
            left = __shfl_up_sync(kFull, uv[3], 1);
            if (lane == 0) {
                uint32_t hp = (bx0 > 0 && rows_valid) ? sample_px(f.in, g, bx0 - 1, o, o2, p) : 0;
                if (g.scale_after) {
                    left = __byte_perm(hp, 0u, 0x4241);
                } else {
                    uint32_t au = dp4a_us(hp, g.cu, SW ? (int)g.c_off : 32768);
                    uint32_t av = dp4a_us(hp, g.cv, SW ? (int)g.c_off : 32768);
                    left = __byte_perm(au, av, 0x7531);
                }
            }
        }
        if (SW) {
            // keep the unfiltered samples: the switch paths below combine the two rows differently
#pragma unroll
            for (int i = 0; i < 4; i++) uvr[rr][i] = uv[i];
            leftr[rr] = left;
            chroma_h_sw<SITE>(uv, left, x, w, (g.switches & SK_PS_CS_EDGE_KEEP) != 0, hh[rr][0], hh[rr][1]);
        } else {
            chroma_h<SITE>(uv, left, x, w, hh[rr][0], hh[rr][1]);
        }
        store_word(f.out + (size_t)y * g.y_stride + x, yw, nv);
        if (DAMAGE) {
            uint32_t pyw = load_word<0>(f.prev + (size_t)y * g.y_stride + x, have_prev ? nv : 0);
            diff |= (yw ^ pyw) & (nvx == 4 ? 0xffffffffu : 0x0000ffffu);
        }
    }
    {
        const int yc = y0 >> 1;
        uint32_t cw = chroma_v<FMT>(hh[0][0], hh[0][1], hh[1][0], hh[1][1]);
        if (SW) {
            constexpr uint32_t kGather = FMT == 0 ? 0x6420 : 0x6240;
            if (SITE == 0 && (g.switches & SK_PS_CHROMA_BOX4)) {
                // R4 alternative: the four samples of a 2x2 quad in one rounding
                const uint32_t b0 = ((uvr[0][0] + uvr[0][1] + uvr[1][0] + uvr[1][1] + 0x00020002u) >> 2) & 0x00ff00ffu;
                const uint32_t b1 = ((uvr[0][2] + uvr[0][3] + uvr[1][2] + uvr[1][3] + 0x00020002u) >> 2) & 0x00ff00ffu;
                cw = __byte_perm(b0, b1, kGather);
            } else if (g.switches & SK_PS_CHROMA_V_FIRST) {
                // R4 alternative: average the two rows first, filter horizontally second
                uint32_t va[4], h0, h1;
#pragma unroll
                for (int i = 0; i < 4; i++) va[i] = ((uvr[0][i] + uvr[1][i] + 0x00010001u) >> 1) & 0x00ff00ffu;
                const uint32_t vl = ((leftr[0] + leftr[1] + 0x00010001u) >> 1) & 0x00ff00ffu;
                chroma_h_sw<SITE>(va, vl, x, w, (g.switches & SK_PS_CS_EDGE_KEEP) != 0, h0, h1);
                cw = __byte_perm(h0, h1, kGather);
            }
        }
        if (FMT == 0) {
            store_word(f.out + g.off_c0 + (size_t)yc * g.c_stride + x, cw, nv);
            if (DAMAGE) {
                uint32_t pcw = load_word<0>(f.prev + g.off_c0 + (size_t)yc * g.c_stride + x, have_prev ? nv : 0);
                diff |= (cw ^ pcw) & (nvx == 4 ? 0xffffffffu : 0x0000ffffu);
            }
        } else {
            uint8_t *pu = f.out + g.off_c0 + (size_t)yc * g.c_stride + (x >> 1);
            uint8_t *pv = f.out + g.off_c1 + (size_t)yc * g.c_stride + (x >> 1);
            if (nv == 4) {
                *reinterpret_cast<unsigned short *>(pu) = (unsigned short)cw;
                *reinterpret_cast<unsigned short *>(pv) = (unsigned short)(cw >> 16);
            } else if (nv == 2) {
                *pu = (uint8_t)cw;
                *pv = (uint8_t)(cw >> 16);
            }
            if (DAMAGE) {
                uint32_t a = 0, b = 0;
                const uint8_t *qu = f.prev + g.off_c0 + (size_t)yc * g.c_stride + (x >> 1);
                const uint8_t *qv = f.prev + g.off_c1 + (size_t)yc * g.c_stride + (x >> 1);
                if (have_prev && nv == 4) { a = ld2<0>(qu); b = ld2<0>(qv); }
                else if (have_prev && nv == 2) { a = ld1(qu); b = ld1(qv); }
                diff |= (cw ^ (a | (b << 16))) & (nvx == 4 ? 0xffffffffu : 0x00ff00ffu);
            }
        }
    }
    if (DAMAGE) {
        __shared__ uint32_t s_mask[8];
        bool d = rows_valid && nvx > 0 && (diff != 0 || !have_prev);
        uint32_t m = __ballot_sync(kFull, d);
        if (lane == 0) s_mask[warp] = m;
        __syncthreads();
        if (warp == 0) {
            bool flag = false, valid = false;
            if (lane < 8) {
                uint32_t acc = 0;
#pragma unroll
                for (int k = 0; k < 8; k++) acc |= s_mask[k];
                flag = ((acc >> (4 * lane)) & 0xfu) != 0;
                int bx = blockIdx.x * 8 + lane, by = blockIdx.y;
                valid = bx < g.mb_w && by < g.mb_h;
                if (valid) {
                    f.map[(size_t)by * g.mb_w + bx] = flag ? 1 : 0;
                    policy_update(f, g, (size_t)by * g.mb_w + bx, flag, f.state ? f.state[(size_t)by * g.mb_w + bx] : 0u);
                }
            }
            uint32_t votes = __ballot_sync(kFull, valid && flag);
            if (lane == 0) {
                if (f.scratch != nullptr) finish_count(f, (uint32_t)__popc(votes));      // allocation-free entry point
                else if (votes != 0 && f.count_out != nullptr) atomicAdd(f.count_out, (uint32_t)__popc(votes));
            }
        }
    }
}

// ---- tiled scale kernel (fast path) -------------------------------------------------
// Used when the horizontal pass comes first (or only one axis changes) and the source
// rows of an output tile fit the shared-memory window (kScaleMaxRows).  Per CTA tile of
// 128 x 32 (or 128 x 16) output pixels:
//   -  the raw source window arrives by ONE 2-D tensor-map copy (TMA); meanwhile the CTA asks for its
//      successor's window to be brought into L2;
//   A  every window row is h-scaled ONCE into shared memory (in place over the raw row when it is long
//      enough); lanes own adjacent output columns, so the gathers are (nearly) bank-conflict free;
//   B  each thread takes 4 adjacent output pixels x 2 rows with 128-bit LDS, applies the
//      vertical 2-tap, then the same matrix / chroma / store / damage tail as above (once per 16 rows).
constexpr int kScaleRowWords = 132;     // 128 columns + halo column (index 128), 16-byte aligned rows

// All frames of a launch share the geometry, hence the vertical table: any frame's copy serves.
__device__ __forceinline__ uint32_t batch_vpack(const SkFrame &f, int y) { return f.v_pack[y].x; }

// TR = output rows per CTA tile (16 or 32): a 32-row tile runs pass B twice on one staged window, which
// halves the per-thread fixed cost (geometry, staging wait, damage tail) per pixel and shrinks the window
// overlap (44 instead of 2 x 23 source rows at 4:3).
//
// RS = words per staged raw row when the session rounded the window width up to one of the compiled classes
// (in-place layout only), 0 = taken from the geometry at run time.  A compile-time row stride turns every
// row step of pass A into an immediate offset of the LDS / STS instructions.
// AFTER = the picture grows: the window's pixels are converted as they are picked up in pass A, both 2-tap
// passes then blend Y,U,V bytes and pass B only has to split them (upstream scales behind the matrix then).
template <int FMT, int SITE, bool DAMAGE, bool SH, bool SV, int RPW, int TR, int RS, bool AFTER, bool FULL>
__device__ __forceinline__ void scale_tile(const SkGeom &g, const SkFrame &f, uint32_t *Sb, const int Ss_rt, uint32_t *RAW, uint32_t *HALO,
                                           const int bx0, const int by0, const int lane, const int warp,
                                           const int sy0, const int sy1,
                                           const bool have_prev, uint32_t (&diff)[TR / 16], int &nvx_out,
                                           bool (&rows_valid_out)[TR / 16],
                                           const uint32_t bar, const CUtensorMap *tmap)
{
    constexpr int NG = TR / 16;          // 16-row groups (= macroblock rows) per tile
    const int w = g.out_w, h = g.out_h;

    // f.v_pack[y] = {first source row, dp2a weight pair of the two vertical taps}; the second row is always
    // first + 1 (vertical_taps clamps first to in_h - 2), and the table is padded to an even length.
    const uint2 *vp = f.v_pack;
    const int nr = sy1 - sy0 + 1;           // source-row window [sy0, sy1]: fetched by the caller, first thing
    // the warp that owns the halo column (see pass A) needs the vertical taps of output row by0 + lane
    uint2 hvp = make_uint2(0u, 0u);
    if (SITE == 1 && SV && bx0 > 0 && warp == 7 && lane < TR) hvp = vp[min(by0 + lane, h - 1)];

    // Previous-frame words of pass B do not depend on shared memory: fetch them first so their
    // latency hides behind pass A.
    const int x = bx0 + lane * 4;
    const int nvx = FULL ? 4 : ((w - x) >= 4 ? 4 : ((w - x) > 0 ? 2 : 0));
    nvx_out = nvx;
    // (I420: the previous U and V half-words stay in separate registers until pass B compares them, so that
    //  nothing here waits for a global load before the staging copy has been issued)
    uint32_t pyw[NG][2], pcw[NG], pcv[NG];
    int vo[NG][2];
    uint32_t vt[NG][2];
#pragma unroll
    for (int gi = 0; gi < NG; gi++) {
        const int y0 = by0 + 16 * gi + 2 * warp;
        const bool rows_valid = FULL || y0 < h;
        const int nv = rows_valid ? nvx : 0;
        rows_valid_out[gi] = rows_valid;
        diff[gi] = 0;
        const uint32_t y_off = (uint32_t)y0 * (uint32_t)g.y_stride + (uint32_t)x;
        const uint32_t c_off = (uint32_t)(y0 >> 1) * (uint32_t)g.c_stride + (FMT == 0 ? (uint32_t)x : (uint32_t)(x >> 1));
        pyw[gi][0] = pyw[gi][1] = pcw[gi] = pcv[gi] = 0;
        // vertical taps of this thread's two rows: fetched now, used after the staging wait
        vo[gi][0] = vo[gi][1] = 0;
        vt[gi][0] = vt[gi][1] = 0;
        if (SV && rows_valid) {                 // y0 is even: both rows' entries in one 16-byte load
            const uint4 t = *reinterpret_cast<const uint4 *>(vp + y0);
            vo[gi][0] = (int)t.x;  vt[gi][0] = t.y;
            vo[gi][1] = (int)t.z;  vt[gi][1] = t.w;
        }
        if (DAMAGE && have_prev) {
            if (FULL) {
                pyw[gi][0] = ld4<0>(f.prev + y_off);
                pyw[gi][1] = ld4<0>(f.prev + y_off + g.y_stride);
                if (FMT == 0) pcw[gi] = ld4<0>(f.prev + g.off_c0 + c_off);
                else { pcw[gi] = ld2<0>(f.prev + g.off_c0 + c_off); pcv[gi] = ld2<0>(f.prev + g.off_c1 + c_off); }
            } else {
                pyw[gi][0] = load_word<0>(f.prev + y_off, nv);
                pyw[gi][1] = load_word<0>(f.prev + y_off + g.y_stride, nv);
                if (FMT == 0) {
                    pcw[gi] = load_word<0>(f.prev + g.off_c0 + c_off, nv);
                } else {
                    uint32_t a = 0, b = 0;
                    if (nv == 4) { a = ld2<0>(f.prev + g.off_c0 + c_off); b = ld2<0>(f.prev + g.off_c1 + c_off); }
                    else if (nv == 2) { a = ld1(f.prev + g.off_c0 + c_off); b = ld1(f.prev + g.off_c1 + c_off); }
                    pcw[gi] = a;  pcv[gi] = b;
                }
            }
        }
    }

    // ---- stage: the raw source window of this tile (rows sy0.., 16-byte aligned columns cx0..) comes
    //      global -> shared through ONE 2-D tensor-map copy issued by one thread of the CTA
    //      (cp.async.bulk.tensor.2d, UTMALDG in SASS; completion on the mbarrier).  The box is the
    //      largest window any tile of this geometry needs; what falls outside the frame arrives as zeros,
    //      so edge tiles need no special path.  No thread computes a load address.
    const int xo_first = bx0 > 0 ? bx0 - 1 : 0;                       // halo column included
    const int xo_last = FULL ? bx0 + 127 : min(bx0 + 127, w - 1);
    int xs_first = xo_first;
    if (SH) xs_first = (xo_first * g.h_inc) >> 16;
    const int cx0 = xs_first & ~3;                                    // 16-byte aligned start
    const uint32_t RSB = 4u * (uint32_t)(RS ? RS : g.scale_raw_stride);   // bytes per raw row
    const int Ss = RS ? RS : Ss_rt;
    if (threadIdx.x == 0) {
        mbar_expect_tx(bar, RSB * (uint32_t)g.scale_rows);
        tma_load_2d((uint32_t)__cvta_generic_to_shared(RAW), tmap, cx0 >> 1, sy0, bar);
        // While this window is on its way: ask for the window of the tile that will take this CTA's place
        // (g.scale_prefetch = CTAs resident on the whole GPU, tiles are handed out x-fastest, then y, then
        // frame) to be brought into L2, so that its own staging copy later finds it there instead of in HBM.
        if (g.scale_prefetch > 0) {
            const uint32_t per_frame = gridDim.x * gridDim.y;
            const uint32_t lin = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z) + (uint32_t)g.scale_prefetch;
            if (lin < per_frame * gridDim.z) {
                const uint32_t fz = lin / per_frame, rem = lin - fz * per_frame;
                const uint32_t ty = rem / gridDim.x, tx = rem - ty * gridDim.x;
                int px = tx > 0 ? (int)tx * 128 - 1 : 0;
                if (SH) px = (px * g.h_inc) >> 16;
                const int py = SV ? (int)batch_vpack(f, (int)ty * TR) : (int)ty * TR;
                tma_prefetch_2d(tmap + ((int)fz - (int)blockIdx.z), (px & ~3) >> 1, py);
            }
        }
    }
    // Pass A's per-thread set-up does not need the window: it runs while the copy is in flight.
    // Column k of this lane is output pixel bx0 + lane + 32k: byte offset of its first source pixel inside a
    // raw row, and the dp2a weight pair of its two horizontal taps.
    uint32_t rel4[4], fr[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int xo = bx0 + lane + 32 * k;
        const int xc = (FULL || xo < w) ? xo : xo_last;
        rel4[k] = 4u * (uint32_t)(xc - cx0);
        fr[k] = 0;
        if (SH) {
            const int32_t acc = xc * g.h_inc;
            rel4[k] = 4u * (uint32_t)((acc >> 16) - cx0);
            const uint32_t fq = (uint32_t)(acc >> 8) & 0xffu;
            fr[k] = (256u - fq) | (fq << 16);          // the dp2a weight pair, built once per column
        }
    }
    // Shared-space addresses by hand, one register per column (and one for the stores) that already holds
    // this warp's first row; the empty asm keeps the compiler from re-deriving them row by row.  With a
    // compile-time stride (RS) every further row is an immediate offset, otherwise one add per register.
    const uint32_t raw_s = (uint32_t)__cvta_generic_to_shared(RAW) + (uint32_t)warp * RSB;
    uint32_t ak[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        ak[k] = raw_s + rel4[k];
        asm volatile("" : "+r"(ak[k]));
    }
    uint32_t os = (uint32_t)__cvta_generic_to_shared(Sb) + 4u * (uint32_t)(lane + warp * Ss);
    asm volatile("" : "+r"(os));
#pragma unroll
    for (int k = 0; k < 4; k++) asm volatile("" : "+r"(fr[k]));

    // One warp polls the mbarrier; the other seven sleep at the CTA barrier instead of spending issue
    // slots on try_wait loops (this kernel is issue-bound, not HBM-bound).
    if (warp == 0) mbar_wait(bar, 0);
    __syncthreads();

    // ---- pass A: h-scale (or copy) the window once into the second shared buffer; lanes own
    //      adjacent output columns, so the gathers are (nearly) bank-conflict free
    {
        const uint8_t *rawb = reinterpret_cast<const uint8_t *>(RAW);
        // The halo column (output pixel bx0-1, needed by the cosited chroma filter of the tile's first
        // quad): ONE warp h-scales it for every window row, lane = row, instead of every warp spending
        // a one-lane instruction stream per row on it.  Warp 7 owns the fewest window rows.
        if (SITE == 1 && bx0 > 0 && warp == 7) {
            uint32_t hrel4 = 4u * (uint32_t)(xo_first - cx0), hf = 0;
            if (SH) {
                const int32_t acc = xo_first * g.h_inc;
                hrel4 = 4u * (uint32_t)((acc >> 16) - cx0);
                hf = (uint32_t)(acc >> 8) & 0xffu;
            }
#pragma unroll 1
            for (int r = lane; r < nr; r += 32) {
                const uint8_t *R = rawb + (uint32_t)r * RSB;
                uint32_t ha = *reinterpret_cast<const uint32_t *>(R + hrel4);
                uint32_t hb = SH ? *reinterpret_cast<const uint32_t *>(R + hrel4 + 4) : 0u;
                if (AFTER) { ha = to_yuv_word(ha, g);  hb = SH ? to_yuv_word(hb, g) : 0u; }
                HALO[r] = SH ? lerp_h2(ha, hb, hf) : ha;
            }
            __syncwarp();
            // ... and finishes it: lane = output row of the tile, vertical 2-tap + the two chroma dot
            // products, so that pass B only picks one word up per row.
            if (lane < TR) {
                const int ra = SV ? (int)hvp.x - sy0 : min(lane, nr - 1);
                uint32_t hp = HALO[ra];
                if (SV) hp = blend2(hp, HALO[ra + 1], hvp.y, 128u);
                HALO[kScaleMaxRows + lane] = AFTER ? __byte_perm(hp, 0u, 0x4241)
                                        : __byte_perm(dp4a_us(hp, g.cu, 32768), dp4a_us(hp, g.cv, 32768), 0x7531);
            }
        }
#pragma unroll
        for (int i = 0; i < RPW; i++) {
            const uint32_t ioff = RS ? (uint32_t)i * 8u * 4u * (uint32_t)RS : 0u;      // same stride for loads and stores
            if (warp + 8 * i < nr) {                      // warp-uniform
                uint32_t a[4], b[4];
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    a[k] = lds32(ak[k] + ioff);
                    b[k] = SH ? lds32(ak[k] + ioff + 4u) : 0u;
                    if (AFTER) { a[k] = to_yuv_word(a[k], g);  b[k] = SH ? to_yuv_word(b[k], g) : 0u; }
                }
                // in-place mode: the row's results overwrite the raw row they were read from, so every
                // lane's reads of this row must precede any lane's writes
                __syncwarp();
#pragma unroll
                for (int k = 0; k < 4; k++) sts32(os + ioff + 128u * k, SH ? blend2(a[k], b[k], fr[k], 0u) : a[k]);
            }
            if (!RS) {
#pragma unroll
                for (int k = 0; k < 4; k++) ak[k] += 8u * RSB;
                os += 32u * (uint32_t)Ss;
            }
        }
    }
    __syncthreads();

    // ---- pass B: vertical 2-tap + matrix + chroma + stores + compare, once per 16-row group
#pragma unroll
    for (int gi = 0; gi < NG; gi++) {
        const int y0 = by0 + 16 * gi + 2 * warp;
        const bool rows_valid = FULL || y0 < h;
        const int nv = rows_valid ? nvx : 0;
        const uint32_t y_off = (uint32_t)y0 * (uint32_t)g.y_stride + (uint32_t)x;
        const uint32_t c_off = (uint32_t)(y0 >> 1) * (uint32_t)g.c_stride + (FMT == 0 ? (uint32_t)x : (uint32_t)(x >> 1));
        uint32_t hh[2][2];
        uint32_t d = 0;
#pragma unroll
        for (int rr = 0; rr < 2; rr++) {
            const int y = y0 + rr;
            uint4 q = make_uint4(0, 0, 0, 0);
            if (rows_valid) {
                const int ra = SV ? vo[gi][rr] - sy0 : y - sy0;
                const uint32_t *Sq = Sb + ra * Ss + 4 * lane;
                q = *reinterpret_cast<const uint4 *>(Sq);
                if (SV) {
                    const uint4 q2 = *reinterpret_cast<const uint4 *>(Sq + Ss);
                    const uint32_t wp = vt[gi][rr];
                    q.x = blend2(q.x, q2.x, wp, 128u);
                    q.y = blend2(q.y, q2.y, wp, 128u);
                    q.z = blend2(q.z, q2.z, wp, 128u);
                    q.w = blend2(q.w, q2.w, wp, 128u);
                }
            }
            uint32_t yw, uv[4];
            if (AFTER) split_yuv_quad(q, yw, uv);
            else matrix_quad(q, g, yw, uv);
            uint32_t left = 0;
            if (SITE == 1) {
                left = __shfl_up_sync(kFull, uv[3], 1);
                if (lane == 0) left = HALO[kScaleMaxRows + 16 * gi + 2 * warp + rr];     // unused when bx0 == 0
            }
            chroma_h<SITE>(uv, left, x, w, hh[rr][0], hh[rr][1]);
            uint8_t *oy = f.out + y_off + (uint32_t)rr * (uint32_t)g.y_stride;
            if (FULL) *reinterpret_cast<uint32_t *>(oy) = yw;
            else store_word(oy, yw, nv);
            if (DAMAGE) d |= (yw ^ pyw[gi][rr]) & ((FULL || nvx == 4) ? 0xffffffffu : 0x0000ffffu);
        }
        uint32_t cw = chroma_v<FMT>(hh[0][0], hh[0][1], hh[1][0], hh[1][1]);
        if (FMT == 0) {
            if (FULL) *reinterpret_cast<uint32_t *>(f.out + g.off_c0 + c_off) = cw;
            else store_word(f.out + g.off_c0 + c_off, cw, nv);
            if (DAMAGE) d |= (cw ^ pcw[gi]) & ((FULL || nvx == 4) ? 0xffffffffu : 0x0000ffffu);
        } else {
            uint8_t *pu = f.out + g.off_c0 + c_off;
            uint8_t *pv = f.out + g.off_c1 + c_off;
            if (nv == 4) {
                *reinterpret_cast<unsigned short *>(pu) = (unsigned short)cw;
                *reinterpret_cast<unsigned short *>(pv) = (unsigned short)(cw >> 16);
            } else if (nv == 2) {
                *pu = (uint8_t)cw;
                *pv = (uint8_t)(cw >> 16);
            }
            if (DAMAGE) d |= (((cw & 0xffffu) ^ pcw[gi]) | ((cw >> 16) ^ pcv[gi])) & ((FULL || nvx == 4) ? 0xffffu : 0x00ffu);
        }
        diff[gi] = d;
    }
}

// RPW = window rows per warp (3: window <= 24 rows, 5: <= 40, 7: <= 56 = kScaleMaxRows); TR = output rows per tile.
template <int FMT, int SITE, bool DAMAGE, bool SH, bool SV, int RPW, int TR, int RS, bool AFTER>
#ifndef SK_SCALE_CTAS32
#define SK_SCALE_CTAS32 6
#endif
#ifndef SK_SCALE_CTAS16
#define SK_SCALE_CTAS16 6
#endif
// (6 CTAs per SM = 40 registers: +1-1.5 % on every shrink ratio measured; the growing-picture instantiation
//  converts in pass A as well and is 4 % faster with the 48 registers of 5 CTAs per SM)
__global__ void __launch_bounds__(kThreads, RPW == 3 ? 6 : (AFTER ? 5 : (TR == 32 ? SK_SCALE_CTAS32 : SK_SCALE_CTAS16)))
sk_scale_tiled_kernel(const __grid_constant__ SkGeom g, const __grid_constant__ SkBatch batch,
                      const __grid_constant__ SkTmaps maps)
{
    constexpr int NG = TR / 16;
    // dynamic shared memory: the raw window first (a TMA destination: 128-byte aligned), then the h-scaled rows
    extern __shared__ __align__(128) uint32_t sk_smem[];
    uint32_t *RAW = sk_smem;
    // The h-scaled rows live in their own buffer behind the window — or, when a raw row is long enough
    // (g.scale_inplace), IN the raw row they came from, 8 words in (the halo column's sources are the first
    // pixels of the raw row and stay intact): half the shared memory, more CTAs per SM.
    uint32_t *Sb = (RS || g.scale_inplace) ? RAW + 8 : sk_smem + (size_t)g.scale_rows * g.scale_raw_stride;
    const int Ss = g.scale_inplace ? g.scale_raw_stride : kScaleRowWords;
    // h-scaled halo column (output pixel bx0-1), one entry per window row; then its chroma pair per output row
    __shared__ uint32_t s_halo[kScaleMaxRows + 32];
    const SkFrame &f = batch.f[blockIdx.z];
    const int lane = threadIdx.x & 31;
    // the shuffle tells the compiler that the warp index is warp-uniform: row addressing and the row-count
    // branches of pass A then run on the uniform datapath instead of per lane
    const int warp = __shfl_sync(kFull, (int)(threadIdx.x >> 5), 0);
    const int bx0 = blockIdx.x * 128;
    const int by0 = blockIdx.y * TR;
    const bool have_prev = DAMAGE && (f.prev != nullptr);
    const bool full = (bx0 + 128 <= g.out_w) && (by0 + TR <= g.out_h);

    // The tile's source-row window decides where the staging copy starts: these two loads go out before
    // anything else so that their latency overlaps the set-up below.
    int sy0 = by0, sy1 = min(by0 + TR - 1, g.out_h - 1);
    if (SV) {
        const int ylast = sy1;
        sy0 = (int)f.v_pack[by0].x;
        sy1 = (int)f.v_pack[ylast].x + 1;
    }
    __shared__ __align__(8) unsigned long long s_bar;
    const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&s_bar);
    if (threadIdx.x == 0) mbar_init(bar, 1);
    if (DAMAGE && blockIdx.x == 0 && blockIdx.y == 0 && (int)threadIdx.x < f.count_zero_n) f.count_zero[threadIdx.x] = 0u;
    if (warp == 0) __syncwarp();        // only warp 0 ever touches the mbarrier: no CTA barrier needed here
    // thread t < 8*NG publishes macroblock (column t % 8, row t / 8) of this tile
    const int mb_x = blockIdx.x * 8 + (threadIdx.x & 7), mb_y = blockIdx.y * NG + (threadIdx.x >> 3);
    const uint32_t pol_state = DAMAGE ? policy_prefetch(f, g, threadIdx.x < 8 * NG, mb_x, mb_y) : 0u;
    uint32_t diff[NG];
    int nvx = 0;
    bool rows_valid[NG];
    const CUtensorMap *tmap = &maps.m[blockIdx.z];
    if (full) scale_tile<FMT, SITE, DAMAGE, SH, SV, RPW, TR, RS, AFTER, true>(g, f, Sb, Ss, RAW, s_halo, bx0, by0, lane, warp, sy0, sy1, have_prev, diff, nvx, rows_valid, bar, tmap);
    else      scale_tile<FMT, SITE, DAMAGE, SH, SV, RPW, TR, RS, AFTER, false>(g, f, Sb, Ss, RAW, s_halo, bx0, by0, lane, warp, sy0, sy1, have_prev, diff, nvx, rows_valid, bar, tmap);

    if (DAMAGE) {
        __shared__ uint32_t s_mask[NG][8];
#pragma unroll
        for (int gi = 0; gi < NG; gi++) {
            const bool d = rows_valid[gi] && nvx > 0 && (diff[gi] != 0 || !have_prev);
            const uint32_t m = __ballot_sync(kFull, d);
            if (lane == 0) s_mask[gi][warp] = m;
        }
        __syncthreads();
        bool hit = false;
        if (threadIdx.x < 8 * NG) {
            const int gi = threadIdx.x >> 3;
            uint32_t acc = 0;
#pragma unroll
            for (int k = 0; k < 8; k++) acc |= s_mask[gi][k];
            const bool flag = ((acc >> (4 * (threadIdx.x & 7))) & 0xfu) != 0;
            if (mb_x < g.mb_w && mb_y < g.mb_h) {
                f.map[(size_t)mb_y * g.mb_w + mb_x] = flag ? 1 : 0;
                policy_update(f, g, (size_t)mb_y * g.mb_w + mb_x, flag, pol_state);
                hit = flag;
            }
        }
        // Block count by the barrier itself, then ONE fire-and-forget atomic (RED) per CTA into a word the
        // previous use of the ring entry zeroed.  (A returning atomic here — the ticket scheme of finish_count —
        // kept every CTA's registers allocated ~1 us longer and cost 4.7 % of the kernel rate.)
        const int total = __syncthreads_count(hit);
        if (threadIdx.x == 0 && total != 0 && f.count_out != nullptr) atomicAdd(f.count_out, (uint32_t)total);
    }
}

// ---- damage-gated read-back ------------------------------------------------------------
// System-memory consumers need the converted frame on the host, but on a desktop most of it did not
// change.  Instead of a full-frame D2H the macroblocks whose map byte is set are written straight
// into a persistent frame in mapped pinned host memory (zero-copy stores over PCIe): thread = one
// macroblock column of one macroblock row, so adjacent damaged blocks become contiguous 16-byte-per-lane
// warp stores.  The host frame of ring entry k was last refreshed n_ring frames ago: `stale[j]` keeps,
// per host frame j, the union of the maps it has missed; entry k copies map|stale[k] and clears it,
// the others accumulate this frame's map.
template <int FMT, bool VEC>
__global__ void __launch_bounds__(kThreads)
sk_gather_kernel(const __grid_constant__ SkGeom g, const __grid_constant__ SkGather a)
{
    const int bx = blockIdx.x * kThreads + threadIdx.x;
    const int by = blockIdx.y;
    if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0 && a.blocks_zero != nullptr) *a.blocks_zero = 0u;
    bool need = false;
    if (bx < g.mb_w) {
        const size_t idx = (size_t)by * g.mb_w + bx;
        const uint8_t m = a.map ? a.map[idx] : (uint8_t)1;
        const size_t plane = (size_t)g.mb_w * g.mb_h;
        need = (m | a.stale[(size_t)a.ring_idx * plane + idx]) != 0;
        for (int j = 0; j < a.n_ring; j++) {
            uint8_t *sp = a.stale + (size_t)j * plane + idx;
            if (j == a.ring_idx) *sp = 0;
            else if (m) *sp = 1;
        }
    }
    if (need) {
        const int y0 = by * 16;
        const int rows = min(16, g.out_h - y0);              // out_h is even
        const int wpx = min(16, g.out_w - bx * 16);          // valid pixels of this block (even)
        // luma: `rows` segments of up to 16 bytes
        const size_t yo = (size_t)y0 * g.y_stride + (size_t)bx * 16;
        if (VEC && wpx == 16) {
#pragma unroll 4
            for (int r = 0; r < rows; r++)
                *reinterpret_cast<uint4 *>(a.dst + yo + (size_t)r * g.y_stride) =
                    *reinterpret_cast<const uint4 *>(a.src + yo + (size_t)r * g.y_stride);
        } else {
            const int nw = (wpx + 3) >> 2;                   // whole words; the row stride is a multiple of 4
            for (int r = 0; r < rows; r++)
                for (int k = 0; k < nw; k++)
                    *reinterpret_cast<uint32_t *>(a.dst + yo + (size_t)r * g.y_stride + 4 * k) =
                        *reinterpret_cast<const uint32_t *>(a.src + yo + (size_t)r * g.y_stride + 4 * k);
        }
        // chroma: rows/2 segments; NV12 = up to 16 interleaved bytes, I420 = up to 8 bytes in each plane
        const int crow = rows >> 1;
        if (FMT == 0) {
            const size_t co = g.off_c0 + (size_t)(y0 >> 1) * g.c_stride + (size_t)bx * 16;
            if (VEC && wpx == 16) {
#pragma unroll 4
                for (int r = 0; r < crow; r++)
                    *reinterpret_cast<uint4 *>(a.dst + co + (size_t)r * g.c_stride) =
                        *reinterpret_cast<const uint4 *>(a.src + co + (size_t)r * g.c_stride);
            } else {
                const int nw = (wpx + 3) >> 2;
                for (int r = 0; r < crow; r++)
                    for (int k = 0; k < nw; k++)
                        *reinterpret_cast<uint32_t *>(a.dst + co + (size_t)r * g.c_stride + 4 * k) =
                            *reinterpret_cast<const uint32_t *>(a.src + co + (size_t)r * g.c_stride + 4 * k);
            }
        } else {
            const int nw = (wpx / 2 + 3) >> 2;
            for (int pl = 0; pl < 2; pl++) {
                const size_t co = (pl ? g.off_c1 : g.off_c0) + (size_t)(y0 >> 1) * g.c_stride + (size_t)bx * 8;
                for (int r = 0; r < crow; r++)
                    for (int k = 0; k < nw; k++)
                        *reinterpret_cast<uint32_t *>(a.dst + co + (size_t)r * g.c_stride + 4 * k) =
                            *reinterpret_cast<const uint32_t *>(a.src + co + (size_t)r * g.c_stride + 4 * k);
            }
        }
    }
    // number of macroblocks written, one RED per CTA
    const int total = __syncthreads_count(need);
    if (threadIdx.x == 0 && total != 0 && a.blocks_out != nullptr) atomicAdd(a.blocks_out, (uint32_t)total);
}

template <int RP, int QX, int HINT>
cudaError_t launch_convert_shape(const SkGeom &g, const SkBatch &b, int n, cudaStream_t st)
{
    dim3 grid((g.out_w + 128 * QX - 1) / (128 * QX), (g.out_h + 16 * RP - 1) / (16 * RP), n);
    dim3 block(kThreads);
    const bool policy = g.damage != 0 && b.f[0].state != nullptr;
#define SK_CASE(FMT, SITE, DMG)                                                                    \
    if (g.out_fmt == FMT && g.site == SITE && (g.damage != 0) == DMG) {                             \
        if (DMG && policy) sk_convert_kernel<RP, QX, FMT, SITE, DMG, HINT, DMG><<<grid, block, 0, st>>>(g, b);   \
        else sk_convert_kernel<RP, QX, FMT, SITE, DMG, HINT, false><<<grid, block, 0, st>>>(g, b);  \
        return cudaGetLastError();                                                                \
    }
    SK_CASE(0, 0, true) SK_CASE(0, 1, true) SK_CASE(1, 0, true) SK_CASE(1, 1, true)
    SK_CASE(0, 0, false) SK_CASE(0, 1, false) SK_CASE(1, 0, false) SK_CASE(1, 1, false)
#undef SK_CASE
    return cudaErrorInvalidValue;
}

#ifdef SK_TUNING
cudaError_t launch_convert_warp(const SkGeom &g, const SkBatch &b, int n, cudaStream_t st)
{
    const int tiles = ((g.out_w + 255) / 256) * ((g.out_h + 15) / 16);
    dim3 grid((tiles + 7) / 8, 1, n);
    dim3 block(kThreads);
    const bool policy = g.damage != 0 && b.f[0].state != nullptr;
#define SK_CASE(FMT, SITE, DMG)                                                                        \
    if (g.out_fmt == FMT && g.site == SITE && (g.damage != 0) == DMG) {                                 \
        if (DMG && policy) sk_convert_warp_kernel<FMT, SITE, DMG, DMG><<<grid, block, 0, st>>>(g, b);   \
        else sk_convert_warp_kernel<FMT, SITE, DMG, false><<<grid, block, 0, st>>>(g, b);               \
        return cudaGetLastError();                                                                    \
    }
    SK_CASE(0, 0, true) SK_CASE(0, 1, true) SK_CASE(1, 0, true) SK_CASE(1, 1, true)
    SK_CASE(0, 0, false) SK_CASE(0, 1, false) SK_CASE(1, 0, false) SK_CASE(1, 1, false)
#undef SK_CASE
    return cudaErrorInvalidValue;
}
#endif  // SK_TUNING

}  // namespace

cudaError_t sk_launch_convert(const SkGeom &g, const SkBatch &b, int n, cudaStream_t st, int variant)
{
    // The release library carries ONE shape: RP=1, QX=2, plain 128-bit __ldg loads — the winner of the
    // sweep in profiles/ (r1_default_ncu_summary.md).  The losing shapes and the cp.async.bulk (TMA engine)
    // staging variant are compiled only into the tuning build (make tuning, -DSK_TUNING).
#ifdef SK_TUNING
    switch (variant) {
    default:
    case 0: return launch_convert_shape<1, 2, 0>(g, b, n, st);
    case 15: return launch_convert_shape<1, 1, 0>(g, b, n, st);
    case 1: return launch_convert_shape<2, 1, 0>(g, b, n, st);
    case 2: return launch_convert_shape<1, 2, 0>(g, b, n, st);
    case 3: return launch_convert_shape<2, 2, 0>(g, b, n, st);
    case 4: return launch_convert_shape<1, 1, 1>(g, b, n, st);
    case 5: return launch_convert_shape<2, 1, 1>(g, b, n, st);
    case 6: return launch_convert_shape<1, 2, 1>(g, b, n, st);
    case 7: return launch_convert_shape<2, 2, 1>(g, b, n, st);
    case 8: return launch_convert_shape<4, 1, 0>(g, b, n, st);
    case 9: return launch_convert_shape<4, 2, 0>(g, b, n, st);
    case 10: return launch_convert_shape<2, 4, 0>(g, b, n, st);
    case 11: return launch_convert_shape<1, 4, 0>(g, b, n, st);
    case 12: return launch_convert_warp(g, b, n, st);             // one warp per 256x16 tile, no shared memory / barriers
    case 13: return launch_convert_shape<1, 2, 3>(g, b, n, st);   // default shape, streaming loads + stores
    case 14: return launch_convert_shape<1, 2, 4>(g, b, n, st);   // default shape, input rows staged by cp.async.bulk (TMA engine)
    }
#else
    (void)variant;
    return launch_convert_shape<1, 2, 0>(g, b, n, st);
#endif
}

// One dispatch for both uses of the tiled kernel's template table: launch it (resident == nullptr), or
// report how many of its CTAs fit one SM for this geometry's shared-memory footprint.
static cudaError_t tiled_dispatch(const SkGeom &g, const SkBatch *b, const SkTmaps *maps, int n, cudaStream_t st, int *resident)
{
    const int tr = g.scale_tile_rows;                      // 16 or 32, chosen by the session runtime
    dim3 grid((g.out_w + 127) / 128, (g.out_h + tr - 1) / tr, n);
    dim3 block(kThreads);
    const size_t smem = (size_t)g.scale_rows * ((g.scale_inplace ? 0 : kScaleRowWords) + (size_t)g.scale_raw_stride) * sizeof(uint32_t);
#define SK_TILED_LAUNCH(FMT, SITE, DMG, SH, SV, RPW, TR, RS, AFTER)                                   \
        {                                                                                             \
            auto kern = sk_scale_tiled_kernel<FMT, SITE, DMG, SH, SV, RPW, TR, RS, AFTER>;            \
            if (smem > 48 * 1024)                                                                     \
                cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);   \
            if (resident) return cudaOccupancyMaxActiveBlocksPerMultiprocessor(resident, kern, kThreads, smem); \
            kern<<<grid, block, smem, st>>>(g, *b, *maps);                                            \
        }
    // compiled row-stride classes (SkGeom::scale_raw_stride was rounded up to one by the session runtime)
    const int rs = (g.scale_inplace && g.scale_h) ? g.scale_raw_stride : 0;
#define SK_TILED(FMT, SITE, DMG, SH, SV)                                                              \
    if (g.out_fmt == FMT && g.site == SITE && (g.damage != 0) == DMG && (g.scale_h != 0) == SH &&     \
        (g.scale_v != 0) == SV) {                                                                     \
        /* Specialised instantiations exist for the usual case only, both axes scaled (BOTH): the     \
           compile-time row strides, and the growing picture (AFTER; 32-row tiles, or 16-row tiles    \
           of <= 24 window rows - the session runtime sends everything else to the generic kernel). */ \
        constexpr bool BOTH = SH && SV;                                                               \
        if (BOTH && g.scale_after) {                                                                  \
            if (tr == 32) SK_TILED_LAUNCH(FMT, SITE, DMG, SH, SV, 7, 32, 0, BOTH)                     \
            else SK_TILED_LAUNCH(FMT, SITE, DMG, SH, SV, 3, 16, 0, BOTH)                              \
        }                                                                                             \
        else if (tr == 32 && BOTH && rs == kScaleStride32) SK_TILED_LAUNCH(FMT, SITE, DMG, SH, SV, 7, 32, (BOTH ? kScaleStride32 : 0), false) \
        else if (tr == 32 && BOTH && rs == kScaleStride16n) SK_TILED_LAUNCH(FMT, SITE, DMG, SH, SV, 7, 32, (BOTH ? kScaleStride16n : 0), false) \
        else if (tr == 32) SK_TILED_LAUNCH(FMT, SITE, DMG, SH, SV, 7, 32, 0, false)                   \
        else if (g.scale_rows <= 24) SK_TILED_LAUNCH(FMT, SITE, DMG, SH, SV, 3, 16, 0, false)         \
        else if (BOTH && rs == kScaleStride16) SK_TILED_LAUNCH(FMT, SITE, DMG, SH, SV, 5, 16, (BOTH ? kScaleStride16 : 0), false) \
        else if (BOTH && rs == kScaleStride16n) SK_TILED_LAUNCH(FMT, SITE, DMG, SH, SV, 5, 16, (BOTH ? kScaleStride16n : 0), false) \
        else SK_TILED_LAUNCH(FMT, SITE, DMG, SH, SV, 5, 16, 0, false)                                 \
        return cudaGetLastError();                                                                    \
    }
#define SK_TILED_MODES(FMT, SITE, DMG) \
    SK_TILED(FMT, SITE, DMG, true, true) SK_TILED(FMT, SITE, DMG, true, false) SK_TILED(FMT, SITE, DMG, false, true)
    SK_TILED_MODES(0, 0, true) SK_TILED_MODES(0, 1, true) SK_TILED_MODES(1, 0, true) SK_TILED_MODES(1, 1, true)
    SK_TILED_MODES(0, 0, false) SK_TILED_MODES(0, 1, false) SK_TILED_MODES(1, 0, false) SK_TILED_MODES(1, 1, false)
#undef SK_TILED_MODES
#undef SK_TILED
#undef SK_TILED_LAUNCH
    return cudaErrorInvalidValue;
}

int sk_scale_resident_ctas(const SkGeom &g)
{
    if (!g.scale_fast) return 0;
    int per_sm = 0, dev = 0, sms = 0;
    if (tiled_dispatch(g, nullptr, nullptr, 1, nullptr, &per_sm) != cudaSuccess) { cudaGetLastError(); return 0; }
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
    return per_sm * sms;
}

cudaError_t sk_launch_scale(const SkGeom &g, const SkBatch &b, const SkTmaps *maps, int n, cudaStream_t st, int /*variant*/)
{
    if (g.scale_fast && maps != nullptr) return tiled_dispatch(g, &b, maps, n, st, nullptr);
    dim3 grid((g.out_w + 127) / 128, (g.out_h + 15) / 16, n);
    dim3 block(kThreads);
#define SK_CASE(FMT, SITE, DMG)                                                              \
    if (g.out_fmt == FMT && g.site == SITE && (g.damage != 0) == DMG) {                       \
        if (g.switches) sk_scale_kernel<FMT, SITE, DMG, true><<<grid, block, 0, st>>>(g, b); \
        else sk_scale_kernel<FMT, SITE, DMG, false><<<grid, block, 0, st>>>(g, b);           \
        return cudaGetLastError();                                                          \
    }
    SK_CASE(0, 0, true) SK_CASE(0, 1, true) SK_CASE(1, 0, true) SK_CASE(1, 1, true)
    SK_CASE(0, 0, false) SK_CASE(0, 1, false) SK_CASE(1, 0, false) SK_CASE(1, 1, false)
#undef SK_CASE
    return cudaErrorInvalidValue;
}

cudaError_t sk_launch_gather(const SkGeom &g, const SkGather &a, cudaStream_t st)
{
    dim3 grid((g.mb_w + kThreads - 1) / kThreads, g.mb_h, 1);
    dim3 block(kThreads);
    // 16-byte stores need 16-byte aligned rows: every plane stride a multiple of 16 (all BASELINE widths)
    const bool vec = (g.y_stride % 16 == 0) && (g.c_stride % 16 == 0) && (g.off_c0 % 16 == 0);
    if (g.out_fmt == 0) {
        if (vec) sk_gather_kernel<0, true><<<grid, block, 0, st>>>(g, a);
        else sk_gather_kernel<0, false><<<grid, block, 0, st>>>(g, a);
    } else {
        if (g.y_stride % 16 == 0) sk_gather_kernel<1, true><<<grid, block, 0, st>>>(g, a);   // VEC: luma only
        else sk_gather_kernel<1, false><<<grid, block, 0, st>>>(g, a);
    }
    return cudaGetLastError();
}

// Load every instantiation of the generic scale kernel into the current device's context now, so
// that no later launch triggers a lazy module load (which allocates on the device).
cudaError_t sk_preload_scale_generic()
{
    cudaFuncAttributes fa;
    cudaError_t e;
#define SK_CASE(FMT, SITE, DMG) \
    if ((e = cudaFuncGetAttributes(&fa, sk_scale_kernel<FMT, SITE, DMG, false>)) != cudaSuccess) return e;
    SK_CASE(0, 0, true) SK_CASE(0, 1, true) SK_CASE(1, 0, true) SK_CASE(1, 1, true)
    SK_CASE(0, 0, false) SK_CASE(0, 1, false) SK_CASE(1, 0, false) SK_CASE(1, 1, false)
#undef SK_CASE
    return cudaSuccess;
}
