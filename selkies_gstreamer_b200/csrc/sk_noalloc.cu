// sk_noalloc.cu — sk_scale_frame_noalloc(): the one entry point of this library that
// performs no allocation of any kind on the GPU.
//
// Who needs it.  Everything else in this library hangs off a session, and
// sk_session_create() allocates (device frames, pinned staging, streams, events).
// Two kinds of callers of this stage cannot do that at the moment they need a frame
// converted:
//   * a streaming host that already runs many desktop sessions on the GPU:
//     cudaMalloc/cudaFree are device-wide synchronisation points, so allocating while
//     the other sessions stream makes every one of them miss a frame.  Such a host
//     wants to convert e.g. the first frame after a resize, a thumbnail or a
//     screenshot on buffers it already owns, without creating a session;
//   * a host whose GPU memory is handed out by someone else's pool (an encoder's
//     buffer pool, a framework's caching allocator) and is fully reserved: a raw
//     cudaMalloc there can fail or fragment the pool, and the host's rule is simply
//     "library code does not allocate".
//
// How.  It is the generic shrink+convert+damage kernel (sk_scale_kernel: every output pixel gathers
// its own taps; it also handles equal sizes — not the tiled kernel a session runs, which needs
// per-session set-up) on caller-owned input/output buffers.  The kernel's scratch — the
// reduction words of the damage count and the per-row vertical tap tables — lives in
// one file-scope __device__ array of fixed size instead of cudaMalloc'd memory.  Every
// GPU that loads the library carries that array (about 26 KB) until the process exits.
#include "../../include/selkies_preencode.h"
#include "sk_kernels.h"
#include "sk_host_math.h"

#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>

static constexpr int kMaxRows = SK_NOALLOC_MAX_OUT_HEIGHT;
static constexpr size_t kCounterBytes = 256;                        // [0] running count, [1] CTA ticket
static constexpr size_t kOffBytes = sizeof(int32_t) * kMaxRows;     // first source row per output row
static constexpr size_t kTapBytes = sizeof(int16_t) * kMaxRows;     // second vertical tap per output row
static constexpr size_t kScratchBytes = (kCounterBytes + kOffBytes + kTapBytes + 255) / 256 * 256;
static constexpr int kMaxDevices = 16;

// The scratch itself.  In host code this name is only a handle for the runtime, never a
// device address: the address each GPU gave it comes from cudaGetSymbolAddress below.
static __device__ __align__(256) unsigned char g_noalloc_scratch[kScratchBytes];

namespace {

struct PerDevice {
    unsigned char *base = nullptr;      // device address of g_noalloc_scratch on this GPU
    bool ready = false;
};
PerDevice g_dev[kMaxDevices];
int32_t g_host_off[kMaxRows];           // host staging of the tap tables (static: nothing is allocated)
int16_t g_host_tap[kMaxRows];
std::mutex g_mu;

int fail(char *errbuf, size_t n, int code, const char *fmt, ...)
{
    if (errbuf && n) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(errbuf, n, fmt, ap);
        va_end(ap);
    }
    return code;
}

}  // namespace

extern "C" int sk_scale_frame_noalloc(const sk_config *cfg, const void *d_in, size_t in_stride,
                                      void *d_out, const void *d_prev, uint8_t *d_map, uint32_t *d_count,
                                      void *cuda_stream, char *errbuf, size_t errbuf_len)
{
    if (errbuf && errbuf_len) errbuf[0] = 0;
    if (!cfg || !d_in || !d_out) return fail(errbuf, errbuf_len, SK_ERR_ARG, "cfg, d_in and d_out must not be NULL");
    if (cfg->struct_size != sizeof(sk_config))
        return fail(errbuf, errbuf_len, SK_ERR_ARG, "sk_config.struct_size does not match this library");
    const int in_w = cfg->in_width, in_h = cfg->in_height;
    const int out_w = cfg->out_width ? cfg->out_width : in_w;
    const int out_h = cfg->out_height ? cfg->out_height : in_h;
    auto bad_dim = [](int v) { return v < 2 || v > 8192 || (v & 1); };
    if (bad_dim(in_w) || bad_dim(in_h) || bad_dim(out_w) || bad_dim(out_h))
        return fail(errbuf, errbuf_len, SK_ERR_ARG, "widths and heights must be even and in [2, 8192]");
    if (out_h > kMaxRows)
        return fail(errbuf, errbuf_len, SK_ERR_ARG,
                    "out_height %d exceeds SK_NOALLOC_MAX_OUT_HEIGHT (%d): the fixed scratch array holds one tap "
                    "entry per output row; use a session (sk_session_create) for taller output", out_h, kMaxRows);
    if ((int64_t)out_w * out_h > (int64_t)in_w * in_h)
        return fail(errbuf, errbuf_len, SK_ERR_UNSUPPORTED, "enlarging is not supported");
    if (cfg->in_format != SK_IN_BGRX && cfg->in_format != SK_IN_RGBX) return fail(errbuf, errbuf_len, SK_ERR_ARG, "unknown in_format");
    if (cfg->out_format != SK_OUT_NV12 && cfg->out_format != SK_OUT_I420) return fail(errbuf, errbuf_len, SK_ERR_ARG, "unknown out_format");
    if (cfg->matrix < SK_MATRIX_BT601 || cfg->matrix > SK_MATRIX_BT2020) return fail(errbuf, errbuf_len, SK_ERR_ARG, "unknown matrix");
    if (cfg->chroma_site != SK_SITE_NONE && cfg->chroma_site != SK_SITE_H_COSITED) return fail(errbuf, errbuf_len, SK_ERR_ARG, "unknown chroma_site");
    if (cfg->scale_variant != SK_SCALE_VARIANT_GST_LQ && cfg->scale_variant != SK_SCALE_VARIANT_PLAIN_LQ)
        return fail(errbuf, errbuf_len, SK_ERR_ARG, "unknown scale_variant");
    const bool damage = !(cfg->flags & SK_FLAG_NO_DAMAGE);
    if (damage && (!d_map || !d_count)) return fail(errbuf, errbuf_len, SK_ERR_ARG, "d_map and d_count are required unless SK_FLAG_NO_DAMAGE is set");
    if (in_stride == 0) in_stride = (size_t)4 * in_w;
    if (in_stride < (size_t)4 * in_w) return fail(errbuf, errbuf_len, SK_ERR_ARG, "in_stride smaller than a row");
    if ((uint64_t)in_stride * (uint64_t)in_h >= (1ull << 32))
        return fail(errbuf, errbuf_len, SK_ERR_ARG, "in_stride * in_height must stay below 4 GiB (32-bit offsets in the kernel)");
    if (((uintptr_t)d_in & 3) || (in_stride & 3) || ((uintptr_t)d_out & 3) || ((uintptr_t)d_prev & 3) || ((uintptr_t)d_count & 3))
        return fail(errbuf, errbuf_len, SK_ERR_ALIGN, "d_in, in_stride, d_out, d_prev and d_count must be 4-byte aligned");

    int dev = cfg->device_id;
    cudaError_t e;
    if (dev < 0) {
        if ((e = cudaGetDevice(&dev)) != cudaSuccess) return fail(errbuf, errbuf_len, SK_ERR_CUDA, "cudaGetDevice: %s", cudaGetErrorString(e));
    } else if ((e = cudaSetDevice(dev)) != cudaSuccess) {
        return fail(errbuf, errbuf_len, SK_ERR_CUDA, "cudaSetDevice(%d): %s", dev, cudaGetErrorString(e));
    }
    if (dev >= kMaxDevices) return fail(errbuf, errbuf_len, SK_ERR_ARG, "device index %d beyond the %d this entry point tracks", dev, kMaxDevices);
    cudaStream_t st = (cudaStream_t)cuda_stream;

    std::lock_guard<std::mutex> lock(g_mu);
    PerDevice &pd = g_dev[dev];
    if (!pd.ready) {
        // Set-up, once per GPU, with that GPU current: ask the runtime where this GPU keeps the array.
        void *p = nullptr;
        if ((e = cudaGetSymbolAddress(&p, g_noalloc_scratch)) != cudaSuccess || !p)
            return fail(errbuf, errbuf_len, SK_ERR_CUDA, "cudaGetSymbolAddress(scratch) on device %d: %s", dev, cudaGetErrorString(e));
        // ... and have every instantiation of the kernel loaded into this GPU's context now: a lazy
        // code load on some later call (another format or chroma site) would allocate device memory.
        if ((e = sk_preload_scale_generic()) != cudaSuccess)
            return fail(errbuf, errbuf_len, SK_ERR_CUDA, "loading the scale kernels on device %d: %s", dev, cudaGetErrorString(e));
        pd.base = static_cast<unsigned char *>(p);
        pd.ready = true;
    }
    uint32_t *d_counters = reinterpret_cast<uint32_t *>(pd.base);
    int32_t *d_off = reinterpret_cast<int32_t *>(pd.base + kCounterBytes);
    int16_t *d_tap = reinterpret_cast<int16_t *>(pd.base + kCounterBytes + kOffBytes);

    // The array still holds whatever the previous call left: the reduction words must start at
    // zero, and the tap tables are rewritten for this call's geometry — both on the caller's stream.
    if ((e = cudaMemsetAsync(d_counters, 0, kCounterBytes, st)) != cudaSuccess)
        return fail(errbuf, errbuf_len, SK_ERR_CUDA, "cudaMemsetAsync(scratch counters): %s", cudaGetErrorString(e));
    if (in_h != out_h) {
        skhost::vertical_taps(in_h, out_h, cfg->scale_variant, g_host_off, g_host_tap);
        if ((e = cudaMemcpyAsync(d_off, g_host_off, sizeof(int32_t) * out_h, cudaMemcpyHostToDevice, st)) != cudaSuccess ||
            (e = cudaMemcpyAsync(d_tap, g_host_tap, sizeof(int16_t) * out_h, cudaMemcpyHostToDevice, st)) != cudaSuccess)
            return fail(errbuf, errbuf_len, SK_ERR_CUDA, "cudaMemcpyAsync(tap tables): %s", cudaGetErrorString(e));
    }

    SkGeom g{};
    if (cfg->parity_switches & ~0x7fu) return fail(errbuf, errbuf_len, SK_ERR_ARG, "unknown parity_switches bit");
    skhost::fill_geometry(g, in_w, in_h, out_w, out_h, cfg->in_format, cfg->out_format, cfg->matrix,
                          cfg->chroma_site, damage, cfg->parity_switches);
    g.in_stride = (int)in_stride;

    SkBatch b{};
    b.f[0].in = static_cast<const uint8_t *>(d_in);
    b.f[0].out = static_cast<uint8_t *>(d_out);
    b.f[0].prev = static_cast<const uint8_t *>(d_prev);
    b.f[0].map = d_map;
    b.f[0].count_out = d_count;
    b.f[0].count_zero = nullptr;        // the ticket scheme below publishes the total; nothing to pre-zero
    b.f[0].count_zero_n = 0;
    b.f[0].scratch = d_counters;
    b.f[0].v_off = d_off;
    b.f[0].v_tap = d_tap;
    if ((e = sk_launch_scale(g, b, nullptr, 1, st, 0)) != cudaSuccess)
        return fail(errbuf, errbuf_len, SK_ERR_CUDA, "kernel launch: %s", cudaGetErrorString(e));
    return SK_OK;
}
