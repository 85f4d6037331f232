// sk_kernels.h — internal launch interface between the session runtime and the
// sm_100a kernels.  Not part of the C ABI.
#pragma once
#include <cuda.h>             // CUtensorMap (type only: the driver entry point is looked up at run time)
#include <cuda_runtime.h>
#include <stdint.h>

constexpr int SK_MAX_BATCH = 32;
// Row strides (words) the tiled scale kernel is also compiled for as constants: in-place windows of 32-row
// tiles up to 192 words (shrink ratios up to ~1.4) or 208 words (3:2), of 16-row tiles 208 or 272 words (2:1).
constexpr int kScaleMaxRows = 56;     // tallest staged window (8 warps x 7 rows)
constexpr int kScaleStride32 = 192;
constexpr int kScaleStride16n = 208;
constexpr int kScaleStride16 = 272;

// Geometry + arithmetic constants shared by every frame of one launch.
struct SkGeom {
    int in_w, in_h;
    int out_w, out_h;
    int in_stride;            // bytes, multiple of 16
    int y_stride, c_stride;   // output plane strides (GstVideoInfo default layout)
    uint32_t off_c0, off_c1;  // byte offset of U (or UV) plane and of V plane
    int mb_w, mb_h;
    uint32_t cy, cu, cv;      // dp4a-packed coefficients in the input's byte order
    int out_fmt;              // SK_OUT_*
    int site;                 // SK_SITE_*
    int damage;               // 0 = convert only
    // scale path
    int32_t h_inc;            // 16.16 edge-aligned increment
    int h_first;              // order of the two 2-tap passes
    int scale_h, scale_v;     // which axes actually change size
    int scale_after;          // 1: the picture grows, the scaler runs behind the matrix (generic kernel only)
    int scale_fast;           // 1: the shared-memory tiled kernel applies (set by the session runtime)
    int scale_tile_rows;      // tiled kernel: output rows per CTA tile, 16 or 32
    int scale_inplace;        // tiled kernel: h-scaled rows overwrite their raw rows (no second shared buffer)
    int scale_rows;           // tiled kernel: most source rows any output tile needs
    int scale_raw_stride;     // tiled kernel: words per staged raw row (widest tile window, 16-byte chunks)
    int scale_prefetch;       // tiled kernel: CTAs resident on the GPU = how far ahead a CTA asks for a window in L2 (0: off)
    // temporal damage policy (per-macroblock saturating counters updated in the same kernel)
    int pol_threshold, pol_duration, pol_trigger;
    int reverse;              // 1: walk sessions / tiles back to front (alternates per launch, for L2 reuse)
    // parity switches that change per-pixel arithmetic (SK_PS_*): honoured by the generic kernel only
    uint32_t switches;
    uint32_t y_off, c_off;    // matrix accumulator offsets: 4096 / 32768, +128 with SK_PS_MATRIX_ROUND
};

// Per-session pointers of one launch (all device memory).
struct SkFrame {
    const uint8_t *in;        // BGRx/RGBx rows, in_stride apart
    uint8_t *out;             // packed output frame being written
    const uint8_t *prev;      // previous packed output frame, or nullptr (all damaged)
    uint8_t *map;             // [mb_h][mb_w]
    uint32_t *count_out;      // damaged-block count of this frame: CTAs add to it, it must start at zero
    uint32_t *count_zero;     // the word the NEXT use of this ring entry will add to: zeroed by this launch, so
                              // that no memset has to be queued in front of any launch (nullptr: nothing to zero)
    int count_zero_n;         // words to zero there: 1, or a whole set when this frame leads a batch
    uint32_t *scratch;        // [0] running count, [1] CTA ticket; both return to 0 at kernel end
    const int32_t *v_off;     // scale: first source row per output row
    const int16_t *v_tap;     // scale: second vertical tap (6-bit) per output row
    const uint2 *v_pack;      // tiled scale kernel: {v_off, (256-4*tap) | (4*tap)<<16} per output row, even length
    uint32_t *state;          // per-macroblock policy state (streak | hot<<8 | still<<16 | dirty<<24), or nullptr
    uint8_t *policy;          // per-macroblock policy flags of this frame (SK_POLICY_*), or nullptr
};

struct SkBatch {
    SkFrame f[SK_MAX_BATCH];
};

// Damage-gated read-back of one frame (sk_gather_kernel).
struct SkGather {
    const uint8_t *src;       // packed output frame in HBM
    uint8_t *dst;             // persistent frame in mapped pinned host memory (device-visible address)
    const uint8_t *map;       // this frame's damage map [mb_h][mb_w], or nullptr (copy everything)
    uint8_t *stale;           // [n_ring][mb_h][mb_w]: blocks each host frame has missed so far
    int n_ring, ring_idx;
    uint32_t *blocks_out;     // += number of macroblocks written (starts at zero)
    uint32_t *blocks_zero;    // the word the next use of this ring entry adds to: zeroed here
};
cudaError_t sk_launch_gather(const SkGeom &g, const SkGather &a, cudaStream_t st);
cudaError_t sk_preload_scale_generic();

// One 2-D tensor map per frame of a launch (tiled scale kernel): the frame's input as
// [in_h rows][in_w/2 elements of 8 bytes], box = the raw source window of one output tile.  The TMA
// engine then stages a tile's whole window with ONE instruction (zero-filling what lies outside the frame).
struct SkTmaps {
    CUtensorMap m[SK_MAX_BATCH];
};

// variant: 0 = default. Returns the launch error (no sync).
cudaError_t sk_launch_convert(const SkGeom &g, const SkBatch &b, int n, cudaStream_t st, int variant);
// maps: one tensor map per frame when g.scale_fast (tiled kernel), else may be nullptr.
cudaError_t sk_launch_scale(const SkGeom &g, const SkBatch &b, const SkTmaps *maps, int n, cudaStream_t st, int variant);
// CTAs of the tiled scale kernel this geometry selects that are resident on the current device at once
// (occupancy x SM count); 0 when the tiled kernel does not apply.
int sk_scale_resident_ctas(const SkGeom &g);
