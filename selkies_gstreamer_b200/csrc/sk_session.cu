// sk_session.cu — session runtime and C ABI (include/selkies_preencode.h).
//
// A session is what one instance of the reference's convert stage is
// (src/selkies/legacy/gstwebrtc_app.py:611-617 / :263-284): fixed caps, one
// frame in → one frame out, plus the state this project adds (previous output
// frame and damage map).  Everything a session will ever need is allocated at
// create time; the per-frame path only enqueues copies, one kernel and events.
//
//   copy stream   : pinned staging slot  --cudaMemcpyAsync-->  device input slot
//   compute stream: fused kernel (reads input slot + previous output, writes
//                   next ring entry + damage map + count)
//   d2h stream    : map+count (and the frame with SK_FLAG_HOST_OUTPUT) to pinned host
#include "../../include/selkies_preencode.h"
#include "sk_kernels.h"
#include "sk_host_math.h"

#include <cudaTypedefs.h>     // PFN_cuTensorMapEncodeTiled: looked up through the runtime, libcuda is not linked

#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>

static_assert(sizeof(sk_result) == 144 && sizeof(sk_config) == 76, "ABI structs changed: bump SK_ABI_VERSION and the bindings");

namespace {

constexpr int kMaxInflight = 8;
// Header in front of the maps of a meta block: two alternating sets of count words (one u32 per batch
// member each) and two alternating gathered-block words.  A use of the block adds into one set and zeroes
// the other inside the same kernel, so no memset is ever queued in front of a launch: the previous use
// of the same block — the only reader of the set being zeroed — has finished its read-back by then
// (that is the ring-entry dependency every submission waits on anyway).
constexpr size_t kCountHdr = 512;
constexpr int kCountSet = 32;           // words per set (SK_MAX_BATCH)
constexpr int kGatherWord = 64;         // first of the two gathered-block words
std::atomic<uint64_t> g_launches{0};
std::atomic<int> g_variant{0};         // tuning build only (sk_tuning_set_variant); the release library has one shape
std::atomic<int> g_snake{1};            // alternate traversal direction per launch (tuning build: 102 off / 103 on)
thread_local std::string g_create_error;

using skhost::ru;

struct Slot {          // one staging slot (caller-chosen index)
    uint8_t *h_in = nullptr;
    uint8_t *d_in = nullptr;
    cudaEvent_t consumed = nullptr;   // recorded after the kernel that read d_in
    bool consumed_valid = false;
};

struct TmapSlot {      // a cached 2-D tensor map of one input frame address (tiled scale kernel)
    const void *ptr = nullptr;
    size_t stride = 0;
    CUtensorMap map;
};
constexpr int kTmapCache = 8;

struct Ring {          // one output ring entry (submission order)
    uint8_t *d_out = nullptr;
    uint8_t *h_out = nullptr;
    uint8_t *d_meta = nullptr;        // map bytes, policy bytes, then the u32 count
    uint8_t *h_meta = nullptr;
    uint8_t *d_meta_view = nullptr;   // where this frame's maps actually are: own buffers, or a slice of
    uint8_t *h_meta_view = nullptr;   // the batch leader's contiguous block (one D2H for the whole batch)
    cudaEvent_t h2d_start = nullptr, h2d_done = nullptr;
    cudaEvent_t k_start = nullptr, k_done = nullptr;
    cudaEvent_t d2h_start = nullptr, done = nullptr;
    cudaEvent_t wait_on = nullptr;    // event sk_wait synchronises on (own `done` or a batch leader's)
    cudaEvent_t kernel_done = nullptr; // event after the kernel that wrote d_out (own `k_done` or a batch leader's)
    cudaEvent_t consumer_done = nullptr; // recorded by sk_release_device on the consumer's stream
    bool consumer_pending = false;
    uint64_t frame_id = 0;
    bool valid = false;
    bool had_h2d = false;
    bool used = false;                // `done` has been recorded at least once
    bool waited = false;              // the host has synchronised on wait_on since: nothing on the GPU still uses the entry
    uint8_t *h_out_user = nullptr;    // caller's output buffer of this frame (sk_submit_io), else nullptr
    uint8_t *h_out_dev = nullptr;     // device-visible address of h_out (mapped; gated read-back)
    bool gated = false;               // this frame came back through the gather kernel
    unsigned use = 0;                 // how often this entry's own meta block has been used (picks the count set)
    const uint32_t *h_blocks = nullptr; // gathered-block count of this frame in pinned memory
    const uint32_t *h_count = nullptr; // this frame's damaged-block count in pinned memory
    const uint32_t *d_count = nullptr; // ... and where the kernel adds it up in HBM
};

}  // namespace

struct sk_session {
    sk_config cfg{};
    int dev = 0;
    int n = 2;
    SkGeom geom{};
    bool scaling = false;
    bool generic_only = false;    // route every frame through sk_scale_kernel (parity switches)
    size_t in_bytes_host = 0;     // in_stride * in_h
    int d_in_stride = 0;
    size_t frame_bytes = 0;
    size_t map_bytes = 0, meta_bytes = 0;
    size_t plane_offset[3]{};
    int plane_stride[3]{};
    Slot slots[kMaxInflight];
    Ring ring[kMaxInflight];
    uint32_t *d_state = nullptr;      // per-macroblock temporal policy state
    uint8_t *d_stale = nullptr;       // gated read-back: [n][mb_h*mb_w] blocks each host frame has missed
    bool bmeta_ready = false;         // leader: d_bmeta/h_bmeta allocated for every ring entry
    unsigned bmeta_use[kMaxInflight] = {};  // leader: uses of each batch block (picks the count set)
    uint8_t *d_bmeta[kMaxInflight] = {};   // leader only: maps of all batch members, contiguous per ring entry
    uint8_t *h_bmeta[kMaxInflight] = {};
    int32_t *d_voff = nullptr;
    int16_t *d_vtap = nullptr;
    uint2 *d_vpack = nullptr;
    TmapSlot tmaps[kTmapCache];
    int tmap_next = 0;
    cudaStream_t s_copy = nullptr, s_comp = nullptr, s_d2h = nullptr;
    cudaEvent_t mark[2]{};
    uint64_t seq = 0;
    int prev_ring = -1;           // -1: no predecessor (all damaged)
    cudaStream_t last_kernel_stream = nullptr;
    cudaEvent_t last_kernel_event = nullptr;
    float last_kernel_us = 0.f;
    uint32_t last_count = 0;
    std::string err;
};

namespace {

int fail(sk_session *s, int code, const std::string &msg)
{
    if (s) s->err = msg; else g_create_error = msg;
    return code;
}

#define SK_CUDA(s, call)                                                                     \
    do {                                                                                     \
        cudaError_t e_ = (call);                                                             \
        if (e_ != cudaSuccess)                                                               \
            return fail((s), SK_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); \
    } while (0)

void free_session(sk_session *s)
{
    if (!s) return;
    cudaSetDevice(s->dev);
    if (s->s_copy) cudaStreamSynchronize(s->s_copy);
    if (s->s_comp) cudaStreamSynchronize(s->s_comp);
    if (s->s_d2h) cudaStreamSynchronize(s->s_d2h);
    for (int i = 0; i < kMaxInflight; i++) {
        Slot &sl = s->slots[i];
        if (sl.h_in) cudaFreeHost(sl.h_in);
        if (sl.d_in) cudaFree(sl.d_in);
        if (sl.consumed) cudaEventDestroy(sl.consumed);
        Ring &r = s->ring[i];
        if (r.d_out) cudaFree(r.d_out);
        if (r.h_out) cudaFreeHost(r.h_out);
        if (r.d_meta) cudaFree(r.d_meta);
        if (r.h_meta) cudaFreeHost(r.h_meta);
        cudaEvent_t evs[] = {r.h2d_start, r.h2d_done, r.k_start, r.k_done, r.d2h_start, r.done, r.consumer_done};
        for (cudaEvent_t e : evs) if (e) cudaEventDestroy(e);
    }
    if (s->d_state) cudaFree(s->d_state);
    if (s->d_stale) cudaFree(s->d_stale);
    for (int i = 0; i < kMaxInflight; i++) {
        if (s->d_bmeta[i]) cudaFree(s->d_bmeta[i]);
        if (s->h_bmeta[i]) cudaFreeHost(s->h_bmeta[i]);
    }
    if (s->d_voff) cudaFree(s->d_voff);
    if (s->d_vtap) cudaFree(s->d_vtap);
    if (s->d_vpack) cudaFree(s->d_vpack);
    for (cudaEvent_t e : s->mark) if (e) cudaEventDestroy(e);
    if (s->s_copy) cudaStreamDestroy(s->s_copy);
    if (s->s_comp) cudaStreamDestroy(s->s_comp);
    if (s->s_d2h) cudaStreamDestroy(s->s_d2h);
    delete s;
}

int validate(const sk_config &c, std::string &why)
{
    auto bad_dim = [](int v) { return v < 2 || v > 8192 || (v & 1); };
    if (bad_dim(c.in_width) || bad_dim(c.in_height)) { why = "input width/height must be even and in [2, 8192]"; return SK_ERR_ARG; }
    if (bad_dim(c.out_width) || bad_dim(c.out_height)) { why = "output width/height must be even and in [2, 8192]"; return SK_ERR_ARG; }
    if (c.in_stride < 4 * c.in_width) { why = "in_stride smaller than 4*in_width"; return SK_ERR_ARG; }
    // the kernels address the input with 32-bit byte offsets
    if ((uint64_t)ru(c.in_stride, 16) * (uint64_t)c.in_height >= (1ull << 32)) { why = "in_stride * in_height must stay below 4 GiB"; return SK_ERR_ARG; }
    if (c.in_format != SK_IN_BGRX && c.in_format != SK_IN_RGBX) { why = "unknown in_format"; return SK_ERR_ARG; }
    if (c.out_format != SK_OUT_NV12 && c.out_format != SK_OUT_I420) { why = "unknown out_format"; return SK_ERR_ARG; }
    if (c.matrix < SK_MATRIX_BT601 || c.matrix > SK_MATRIX_BT2020) { why = "unknown matrix"; return SK_ERR_ARG; }
    if (c.chroma_site != SK_SITE_NONE && c.chroma_site != SK_SITE_H_COSITED) { why = "unknown chroma_site"; return SK_ERR_ARG; }
    if (c.n_inflight < 2 || c.n_inflight > kMaxInflight) { why = "n_inflight must be in [2, 8]"; return SK_ERR_ARG; }
    auto bad_knob = [](int v) { return v < 1 || v > 255; };
    if (bad_knob(c.damage_block_threshold) || bad_knob(c.damage_block_duration) || bad_knob(c.paint_over_trigger_frames)) {
        why = "damage_block_threshold, damage_block_duration and paint_over_trigger_frames must be in [1, 255]";
        return SK_ERR_ARG;
    }
    if (c.parity_switches & ~0x7fu) { why = "unknown parity_switches bit"; return SK_ERR_ARG; }
    if ((c.flags & SK_FLAG_GATED_READBACK) && (!(c.flags & SK_FLAG_HOST_OUTPUT) || (c.flags & SK_FLAG_NO_DAMAGE))) {
        why = "SK_FLAG_GATED_READBACK needs SK_FLAG_HOST_OUTPUT and the damage map";
        return SK_ERR_ARG;
    }
    const bool scaling = c.in_width != c.out_width || c.in_height != c.out_height;
    if (scaling) {
        if (c.scale_method != SK_SCALE_BILINEAR) { why = "only scale_method bilinear is supported"; return SK_ERR_UNSUPPORTED; }
        if (c.scale_variant != SK_SCALE_VARIANT_GST_LQ && c.scale_variant != SK_SCALE_VARIANT_PLAIN_LQ) { why = "unknown scale_variant"; return SK_ERR_ARG; }
    }
    return SK_OK;
}

int find_ring(sk_session *s, uint64_t frame_id)
{
    for (int i = 0; i < s->n; i++)
        if (s->ring[i].valid && s->ring[i].frame_id == frame_id) return i;
    return -1;
}

// Fill the launch descriptor of the next frame of `s` and advance its ring.
// `in` is the device input pointer. Returns the ring index.
int stage_frame(sk_session *s, const uint8_t *in, uint64_t frame_id, SkFrame &f)
{
    const int ri = (int)(s->seq % (uint64_t)s->n);
    Ring &r = s->ring[ri];
    f.in = in;
    f.out = r.d_out;
    f.prev = (s->prev_ring >= 0) ? s->ring[s->prev_ring].d_out : nullptr;
    r.d_meta_view = r.d_meta + kCountHdr;
    r.h_meta_view = r.h_meta + kCountHdr;
    const unsigned set = r.use & 1u;                           // provisional: a batch on the leader's block overrides it
    r.h_count = reinterpret_cast<const uint32_t *>(r.h_meta) + set * kCountSet;
    r.d_count = reinterpret_cast<const uint32_t *>(r.d_meta) + set * kCountSet;
    f.map = r.d_meta_view;
    f.count_out = reinterpret_cast<uint32_t *>(r.d_meta) + set * kCountSet;
    f.count_zero = reinterpret_cast<uint32_t *>(r.d_meta) + (set ^ 1u) * kCountSet;
    f.count_zero_n = 1;
    f.state = s->d_state;                       // nullptr unless SK_FLAG_POLICY
    f.policy = r.d_meta_view + s->map_bytes;
    f.scratch = nullptr;
    f.v_off = s->d_voff;
    f.v_tap = s->d_vtap;
    f.v_pack = s->d_vpack;
    r.frame_id = frame_id;
    r.valid = true;
    r.waited = false;
    r.gated = false;
    r.h_out_user = nullptr;

    s->prev_ring = ri;
    s->seq++;
    return ri;
}

// cuTensorMapEncodeTiled through the runtime's driver entry point lookup (once per process).
PFN_cuTensorMapEncodeTiled tensor_map_encoder()
{
    static PFN_cuTensorMapEncodeTiled fn = []() -> PFN_cuTensorMapEncodeTiled {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess) {
            cudaGetLastError();
            return nullptr;
        }
        return reinterpret_cast<PFN_cuTensorMapEncodeTiled>(p);
    }();
    return fn;
}

// The tensor map the tiled scale kernel stages its source windows with, for the frame at `dptr`:
// [in_h][in_w/2] elements of 8 bytes (two pixels), rows `stride` bytes apart, box = the widest / tallest
// window any output tile needs (g.scale_raw_stride words x g.scale_rows).  Encoding is a pure host
// computation (~1 us); maps are cached per input address since sessions cycle through a few buffers.
const CUtensorMap *input_tensor_map(sk_session *s, const void *dptr, size_t stride)
{
    for (TmapSlot &t : s->tmaps)
        if (t.ptr == dptr && t.stride == stride) return &t.map;
    PFN_cuTensorMapEncodeTiled enc = tensor_map_encoder();
    if (!enc) return nullptr;
    TmapSlot &t = s->tmaps[s->tmap_next];
    const cuuint64_t dims[2] = {(cuuint64_t)s->geom.in_w / 2, (cuuint64_t)s->geom.in_h};
    const cuuint64_t strides[1] = {(cuuint64_t)stride};
    const cuuint32_t box[2] = {(cuuint32_t)s->geom.scale_raw_stride / 2, (cuuint32_t)s->geom.scale_rows};
    const cuuint32_t estr[2] = {1, 1};
    if (enc(&t.map, CU_TENSOR_MAP_DATA_TYPE_UINT64, 2, const_cast<void *>(dptr), dims, strides, box, estr,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) {
        t.ptr = nullptr;
        return nullptr;
    }
    t.ptr = dptr;
    t.stride = stride;
    s->tmap_next = (s->tmap_next + 1) % kTmapCache;
    return &t.map;
}

// ss[i] / b.f[i].in: the frames of this launch.  in_stride comes from g_in.
cudaError_t launch(sk_session *const *ss, const SkGeom &g_in, const SkBatch &b, int n, cudaStream_t st)
{
    const sk_session *s = ss[0];
    g_launches.fetch_add(1, std::memory_order_relaxed);
    if (s->scaling || s->generic_only) {
        if (!g_in.scale_fast) return sk_launch_scale(g_in, b, nullptr, n, st, g_variant.load());
        SkTmaps maps;
        for (int i = 0; i < n; i++) {
            const CUtensorMap *m = input_tensor_map(ss[i], b.f[i].in, (size_t)g_in.in_stride);
            if (!m) return cudaErrorInvalidValue;
            maps.m[i] = *m;
        }
        return sk_launch_scale(g_in, b, &maps, n, st, g_variant.load());
    }
    SkGeom g = g_in;
    g.reverse = (int)(s->seq & 1) && g_snake.load();     // seq was advanced by stage_frame: alternates per frame
    return sk_launch_convert(g, b, n, st, g_variant.load());
}

// What must have finished before `stream` may run the kernel that overwrites the ring entry `s`
// will use next and reads the session's previous output.
int wait_for_ring_entry(sk_session *err_to, sk_session *s, cudaStream_t stream, cudaEvent_t &already_waited)
{
    Ring &r = s->ring[s->seq % (uint64_t)s->n];
    // its read-back may still be running (unless the host has already waited for it)
    if (r.used && !r.waited && r.wait_on != already_waited) {
        SK_CUDA(err_to, cudaStreamWaitEvent(stream, r.wait_on, 0));
        already_waited = r.wait_on;
    }
    // a device-side consumer may still be reading it (sk_release_device)
    if (r.consumer_pending) {
        SK_CUDA(err_to, cudaStreamWaitEvent(stream, r.consumer_done, 0));
        r.consumer_pending = false;
    }
    // the session's previous kernel may have run on another stream (a batch leader's)
    if (s->last_kernel_stream && s->last_kernel_stream != stream)
        SK_CUDA(err_to, cudaStreamWaitEvent(stream, s->last_kernel_event, 0));
    return SK_OK;
}

// H2D of one captured frame into the device copy of `slot`, on the session's copy stream.
// Pinned / registered memory is DMA'd asynchronously; pageable memory is staged by the driver before
// the call returns.  Records the ring entry's h2d_done.
int enqueue_h2d(sk_session *err_to, sk_session *s, int slot, const void *src, size_t stride)
{
    const bool timing = s->cfg.flags & SK_FLAG_TIMING;
    const size_t row = (size_t)4 * s->cfg.in_width;
    Slot &sl = s->slots[slot];
    Ring &r = s->ring[s->seq % (uint64_t)s->n];
    // The device copy of this slot may still be read by an earlier kernel.
    if (sl.consumed_valid) SK_CUDA(err_to, cudaStreamWaitEvent(s->s_copy, sl.consumed, 0));
    if (timing) SK_CUDA(err_to, cudaEventRecord(r.h2d_start, s->s_copy));
    if (stride == (size_t)s->d_in_stride) {
        SK_CUDA(err_to, cudaMemcpyAsync(sl.d_in, src, stride * (size_t)(s->cfg.in_height - 1) + row,
                                        cudaMemcpyHostToDevice, s->s_copy));
    } else {
        SK_CUDA(err_to, cudaMemcpy2DAsync(sl.d_in, s->d_in_stride, src, stride, row, s->cfg.in_height,
                                          cudaMemcpyHostToDevice, s->s_copy));
    }
    SK_CUDA(err_to, cudaEventRecord(r.h2d_done, s->s_copy));
    return SK_OK;
}

// Read-back of one ring entry on stream `st` (which already waits for the kernel): maps + counts,
// then the frame — whole (copy engine) into the caller's buffer or the pinned ring, or only the
// macroblocks that changed (gather kernel into the mapped persistent frame).
int enqueue_frame_readback(sk_session *err_to, sk_session *s, int ri, cudaStream_t st, void *host_out, bool own_meta)
{
    Ring &r = s->ring[ri];
    const bool damage = !(s->cfg.flags & SK_FLAG_NO_DAMAGE);
    const bool host_out_flag = s->cfg.flags & SK_FLAG_HOST_OUTPUT;
    r.h_out_user = static_cast<uint8_t *>(host_out);
    if (s->d_stale && !host_out) {
        SkGather a{};
        a.src = r.d_out;
        a.dst = r.h_out_dev;
        a.map = r.d_meta_view;
        a.stale = s->d_stale;
        a.n_ring = s->n;
        a.ring_idx = ri;
        a.blocks_out = reinterpret_cast<uint32_t *>(r.d_meta) + kGatherWord + (r.use & 1u);
        a.blocks_zero = reinterpret_cast<uint32_t *>(r.d_meta) + kGatherWord + ((r.use & 1u) ^ 1u);
        r.h_blocks = reinterpret_cast<const uint32_t *>(r.h_meta) + kGatherWord + (r.use & 1u);
        g_launches.fetch_add(1, std::memory_order_relaxed);
        SK_CUDA(err_to, sk_launch_gather(s->geom, a, st));
        r.gated = true;
    }
    if (own_meta) r.use++;             // the entry's own block has now been used once more
    if (damage && own_meta)
        SK_CUDA(err_to, cudaMemcpyAsync(r.h_meta, r.d_meta, kCountHdr + (s->d_state ? s->meta_bytes : s->map_bytes),
                                        cudaMemcpyDeviceToHost, st));
    if (host_out)
        SK_CUDA(err_to, cudaMemcpyAsync(host_out, r.d_out, s->frame_bytes, cudaMemcpyDeviceToHost, st));
    else if (host_out_flag && !r.gated)
        SK_CUDA(err_to, cudaMemcpyAsync(r.h_out, r.d_out, s->frame_bytes, cudaMemcpyDeviceToHost, st));
    return SK_OK;
}

int submit_common(sk_session *s, const SkGeom &g, const uint8_t *d_in, int slot, uint64_t frame_id, void *host_out)
{
    const bool timing = s->cfg.flags & SK_FLAG_TIMING;
    const int ri_next = (int)(s->seq % (uint64_t)s->n);
    Ring &r = s->ring[ri_next];
    cudaEvent_t waited = nullptr;
    int rc = wait_for_ring_entry(s, s, s->s_comp, waited);
    if (rc != SK_OK) return rc;
    r.had_h2d = slot >= 0;
    if (slot >= 0) SK_CUDA(s, cudaStreamWaitEvent(s->s_comp, r.h2d_done, 0));

    SkBatch b;
    const int ri = stage_frame(s, d_in, frame_id, b.f[0]);
    if (timing) SK_CUDA(s, cudaEventRecord(r.k_start, s->s_comp));
    sk_session *one[1] = {s};
    cudaError_t le = launch(one, g, b, 1, s->s_comp);
    if (le != cudaSuccess) {
        s->prev_ring = -1;          // the ring entry was never written: the next frame has no predecessor
        r.valid = false;
        cudaGetLastError();
        cudaMemset(r.d_meta, 0, kCountHdr);     // the launch that would have zeroed the next count word never ran
        return fail(s, SK_ERR_CUDA, std::string("kernel launch: ") + cudaGetErrorString(le));
    }
    SK_CUDA(s, cudaEventRecord(r.k_done, s->s_comp));
    r.kernel_done = r.k_done;
    s->last_kernel_stream = s->s_comp;
    s->last_kernel_event = r.k_done;
    if (slot >= 0) {
        Slot &sl = s->slots[slot];
        SK_CUDA(s, cudaEventRecord(sl.consumed, s->s_comp));
        sl.consumed_valid = true;
    }
    SK_CUDA(s, cudaStreamWaitEvent(s->s_d2h, r.k_done, 0));
    if (timing) SK_CUDA(s, cudaEventRecord(r.d2h_start, s->s_d2h));
    rc = enqueue_frame_readback(s, s, ri, s->s_d2h, host_out, true);
    if (rc != SK_OK) return rc;
    SK_CUDA(s, cudaEventRecord(r.done, s->s_d2h));
    r.wait_on = r.done;
    r.used = true;
    return SK_OK;
}

int check_slot(sk_session *s, int slot)
{
    if (slot < 0 || slot >= s->n) return fail(s, SK_ERR_ARG, "slot out of range");
    return SK_OK;
}

// Members of one launch must agree on everything the kernel is specialised or parameterised on.
int validate_batch(sk_session **ss, int n)
{
    sk_session *L = ss[0];
    if (n > SK_MAX_BATCH) return fail(L, SK_ERR_ARG, "batch larger than SK_MAX_BATCH (32)");
    for (int i = 0; i < n; i++) {
        sk_session *s = ss[i];
        if (!s) return fail(L, SK_ERR_ARG, "null session in batch");
        const sk_config &a = L->cfg, &b = s->cfg;
        if (s->dev != L->dev || a.in_width != b.in_width || a.in_height != b.in_height ||
            a.out_width != b.out_width || a.out_height != b.out_height || a.in_format != b.in_format ||
            a.out_format != b.out_format || a.matrix != b.matrix || a.chroma_site != b.chroma_site ||
            a.scale_variant != b.scale_variant ||
            ((a.flags ^ b.flags) & (SK_FLAG_NO_DAMAGE | SK_FLAG_POLICY | SK_FLAG_SCALE_GENERIC)) ||
            a.damage_block_threshold != b.damage_block_threshold || a.damage_block_duration != b.damage_block_duration ||
            a.paint_over_trigger_frames != b.paint_over_trigger_frames || a.n_inflight != b.n_inflight ||
            a.parity_switches != b.parity_switches ||
            s->d_in_stride != L->d_in_stride)
            return fail(L, SK_ERR_STATE, "sessions of one batch must share device and configuration");
        if (s->cfg.flags & SK_FLAG_GATED_READBACK)
            return fail(L, SK_ERR_UNSUPPORTED, "SK_FLAG_GATED_READBACK sessions cannot be batched");
        for (int j = 0; j < i; j++)
            if (ss[j] == s) return fail(L, SK_ERR_STATE, "a session appears twice in one batch");
    }
    return SK_OK;
}

// The leader's contiguous map blocks (one per ring entry), all or nothing.
int prepare_bmeta(sk_session *L)
{
    if (L->bmeta_ready || (L->cfg.flags & SK_FLAG_NO_DAMAGE)) return SK_OK;
    const size_t bytes = kCountHdr + (size_t)SK_MAX_BATCH * L->meta_bytes;
    for (int k = 0; k < L->n; k++) {
        cudaError_t e = cudaSuccess;
        if (!L->d_bmeta[k]) e = cudaMalloc((void **)&L->d_bmeta[k], bytes);
        if (e == cudaSuccess) e = cudaMemset(L->d_bmeta[k], 0, kCountHdr);
        if (e == cudaSuccess && !L->h_bmeta[k]) e = cudaHostAlloc((void **)&L->h_bmeta[k], bytes, cudaHostAllocDefault);
        if (e != cudaSuccess) {
            cudaGetLastError();
            for (int j = 0; j <= k; j++) {
                if (L->d_bmeta[j]) { cudaFree(L->d_bmeta[j]); L->d_bmeta[j] = nullptr; }
                if (L->h_bmeta[j]) { cudaFreeHost(L->h_bmeta[j]); L->h_bmeta[j] = nullptr; }
            }
            return fail(L, SK_ERR_NOMEM, std::string("batch map block: ") + cudaGetErrorString(e));
        }
    }
    L->bmeta_ready = true;
    return SK_OK;
}

// A batch step failed after some members' rings were advanced: their next frame has no predecessor.
void abandon_batch(sk_session **ss, int n)
{
    cudaGetLastError();
    for (int i = 0; i < n; i++) {
        ss[i]->prev_ring = -1;
        // count words must start at zero again: the launch that would have zeroed them never ran
        for (int k = 0; k < ss[i]->n; k++) {
            cudaMemset(ss[i]->ring[k].d_meta, 0, kCountHdr);
            if (ss[i]->d_bmeta[k]) cudaMemset(ss[i]->d_bmeta[k], 0, kCountHdr);
        }
        for (int k = 0; k < ss[i]->n; k++)
            if (ss[i]->ring[k].valid && !ss[i]->ring[k].used) ss[i]->ring[k].valid = false;
    }
}

}  // namespace

extern "C" {

int sk_version(void) { return SK_ABI_VERSION; }

uint64_t sk_launch_count(void) { return g_launches.load(); }

#ifdef SK_TUNING
// Tuning build only (make tuning): process-wide kernel-shape selector for A/B runs; not in the public header.
int sk_tuning_set_variant(int variant)
{
    if (variant == 102 || variant == 103) {
        g_snake.store(variant == 103 ? 1 : 0);
        return SK_OK;
    }
    if (variant < 0 || variant > 15) return SK_ERR_ARG;
    g_variant.store(variant);
    return SK_OK;
}
#endif

int sk_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

const char *sk_last_error(sk_session *s)
{
    return s ? s->err.c_str() : g_create_error.c_str();
}

int sk_session_create(const sk_config *cfg_in, sk_session **out)
{
    if (!cfg_in || !out) return fail(nullptr, SK_ERR_ARG, "null argument");
    *out = nullptr;
    if (cfg_in->struct_size != sizeof(sk_config))
        return fail(nullptr, SK_ERR_ARG, "sk_config.struct_size does not match this library");
    sk_config c = *cfg_in;
    if (c.out_width == 0) c.out_width = c.in_width;
    if (c.out_height == 0) c.out_height = c.in_height;
    if (c.in_stride == 0) c.in_stride = 4 * c.in_width;
    if (c.n_inflight == 0) c.n_inflight = 2;
    if (c.scale_method == 0) c.scale_method = SK_SCALE_BILINEAR;
    if (c.device_id < 0) c.device_id = 0;
    if (c.damage_block_threshold == 0) c.damage_block_threshold = 10;       // selkies.py:2950
    if (c.damage_block_duration == 0) c.damage_block_duration = 20;         // selkies.py:2951
    if (c.paint_over_trigger_frames == 0) c.paint_over_trigger_frames = 15; // selkies.py:2949
    std::string why;
    int rc = validate(c, why);
    if (rc != SK_OK) return fail(nullptr, rc, why);

    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(nullptr, SK_ERR_CUDA, std::string("no CUDA device: ") + cudaGetErrorString(e) +
                                              " (this library has no CPU path)");
    }
    if (c.device_id >= ndev) return fail(nullptr, SK_ERR_ARG, "cuda-device-id out of range");

    sk_session *s = new (std::nothrow) sk_session();
    if (!s) return fail(nullptr, SK_ERR_NOMEM, "out of host memory");
    s->cfg = c;
    s->dev = c.device_id;
    s->n = c.n_inflight;
    s->scaling = c.in_width != c.out_width || c.in_height != c.out_height;
    // Per-pixel parity switches are implemented by the generic kernel only (it converts equal sizes too).
    s->generic_only = (c.parity_switches & (SK_PS_MATRIX_ROUND | SK_PS_CS_EDGE_KEEP | SK_PS_CHROMA_V_FIRST | SK_PS_CHROMA_BOX4)) != 0;

#define SK_TRY(call)                                                                           \
    do {                                                                                       \
        cudaError_t e_ = (call);                                                               \
        if (e_ != cudaSuccess) {                                                               \
            g_create_error = std::string(#call) + ": " + cudaGetErrorString(e_);               \
            free_session(s);                                                                   \
            return SK_ERR_CUDA;                                                                \
        }                                                                                      \
    } while (0)

    SK_TRY(cudaSetDevice(s->dev));
    // geometry (GstVideoInfo default layout)
    SkGeom &g = s->geom;
    s->frame_bytes = skhost::fill_geometry(g, c.in_width, c.in_height, c.out_width, c.out_height, c.in_format,
                                           c.out_format, c.matrix, c.chroma_site, !(c.flags & SK_FLAG_NO_DAMAGE),
                                           c.parity_switches);
    s->plane_offset[0] = 0;  s->plane_offset[1] = g.off_c0;  s->plane_offset[2] = g.off_c1;
    s->plane_stride[0] = g.y_stride;  s->plane_stride[1] = g.c_stride;
    s->plane_stride[2] = (c.out_format == SK_OUT_I420) ? g.c_stride : 0;
    g.pol_threshold = c.damage_block_threshold;
    g.pol_duration = c.damage_block_duration;
    g.pol_trigger = c.paint_over_trigger_frames;
    s->d_in_stride = ru(c.in_stride, 16);
    g.in_stride = s->d_in_stride;
    s->in_bytes_host = (size_t)c.in_stride * c.in_height;
    s->map_bytes = (size_t)ru(g.mb_w * g.mb_h, 4);
    s->meta_bytes = 2 * s->map_bytes;
    const bool gated = c.flags & SK_FLAG_GATED_READBACK;

    SK_TRY(cudaStreamCreateWithFlags(&s->s_copy, cudaStreamNonBlocking));
    SK_TRY(cudaStreamCreateWithFlags(&s->s_comp, cudaStreamNonBlocking));
    SK_TRY(cudaStreamCreateWithFlags(&s->s_d2h, cudaStreamNonBlocking));
    const unsigned evflags = (c.flags & SK_FLAG_TIMING) ? cudaEventDefault : cudaEventDisableTiming;
    // Events a host thread waits on put that thread to sleep instead of spinning, unless asked otherwise.
    const unsigned waitflags = evflags | ((c.flags & SK_FLAG_SPIN_WAIT) ? 0u : (unsigned)cudaEventBlockingSync);
    for (int i = 0; i < s->n; i++) {
        Slot &sl = s->slots[i];
        SK_TRY(cudaHostAlloc((void **)&sl.h_in, s->in_bytes_host, cudaHostAllocDefault));
        SK_TRY(cudaMalloc((void **)&sl.d_in, (size_t)s->d_in_stride * c.in_height + 16));
        SK_TRY(cudaEventCreateWithFlags(&sl.consumed, cudaEventDisableTiming));
        Ring &r = s->ring[i];
        SK_TRY(cudaMalloc((void **)&r.d_out, s->frame_bytes));
        SK_TRY(cudaMemset(r.d_out, 0, s->frame_bytes));
        SK_TRY(cudaMalloc((void **)&r.d_meta, kCountHdr + s->meta_bytes));
        SK_TRY(cudaMemset(r.d_meta, 0, kCountHdr + s->meta_bytes));
        SK_TRY(cudaHostAlloc((void **)&r.h_meta, kCountHdr + s->meta_bytes, cudaHostAllocDefault));
        std::memset(r.h_meta, 0, kCountHdr + s->meta_bytes);
        if (c.flags & SK_FLAG_HOST_OUTPUT) {
            SK_TRY(cudaHostAlloc((void **)&r.h_out, s->frame_bytes, gated ? cudaHostAllocMapped : cudaHostAllocDefault));
            if (gated) {
                std::memset(r.h_out, 0, s->frame_bytes);
                SK_TRY(cudaHostGetDevicePointer((void **)&r.h_out_dev, r.h_out, 0));
            }
        }
        cudaEvent_t *evs[] = {&r.h2d_start, &r.h2d_done, &r.k_start, &r.k_done, &r.d2h_start};
        for (cudaEvent_t *ev : evs) SK_TRY(cudaEventCreateWithFlags(ev, evflags));
        SK_TRY(cudaEventCreateWithFlags(&r.done, waitflags));
        SK_TRY(cudaEventCreateWithFlags(&r.consumer_done, cudaEventDisableTiming));
    }
    if ((c.flags & SK_FLAG_POLICY) && !(c.flags & SK_FLAG_NO_DAMAGE)) {
        SK_TRY(cudaMalloc((void **)&s->d_state, s->map_bytes * sizeof(uint32_t)));
        SK_TRY(cudaMemset(s->d_state, 0, s->map_bytes * sizeof(uint32_t)));
    }
    if (gated) {
        // every host frame starts out missing every block
        SK_TRY(cudaMalloc((void **)&s->d_stale, (size_t)s->n * g.mb_w * g.mb_h));
        SK_TRY(cudaMemset(s->d_stale, 1, (size_t)s->n * g.mb_w * g.mb_h));
    }
    for (cudaEvent_t &ev : s->mark) SK_TRY(cudaEventCreate(&ev));
    if (s->scaling) {
        std::vector<int32_t> off;
        std::vector<int16_t> tap;
        off.resize(g.out_h);
        tap.resize(g.out_h);
        skhost::vertical_taps(g.in_h, g.out_h, c.scale_variant, off.data(), tap.data());
        // The tiled kernel needs the h pass first (or a single axis) and every output tile's source rows
        // inside its shared-memory window.  Window height for 16-row and for 32-row tiles:
        auto window_rows = [&](int tr) {
            if (!g.scale_v) return tr;
            int most = 0;
            for (int by0 = 0; by0 < g.out_h; by0 += tr) {
                const int ylast = (by0 + tr - 1 < g.out_h) ? by0 + tr - 1 : g.out_h - 1;
                int last = off[ylast] + 1;
                if (last > g.in_h - 1) last = g.in_h - 1;
                const int rows = last - off[by0] + 1;
                if (rows > most) most = rows;
            }
            return most;
        };
        int max_rows = window_rows(16);
        const int max_rows32 = window_rows(32);
        // widest raw window (in 16-byte chunks) any 128-column output tile stages
        int max_chunks = 0;
        for (int bx0 = 0; bx0 < g.out_w; bx0 += 128) {
            const int xo_first = bx0 > 0 ? bx0 - 1 : 0;
            const int xo_last = (bx0 + 127 < g.out_w) ? bx0 + 127 : g.out_w - 1;
            int xs_first = xo_first, xs_last = xo_last;
            if (g.scale_h) {
                xs_first = (int)(((int64_t)xo_first * g.h_inc) >> 16);
                xs_last = (int)(((int64_t)xo_last * g.h_inc) >> 16) + 1;
            }
            const int chunks = ((xs_last - (xs_first & ~3)) >> 2) + 1;
            if (chunks > max_chunks) max_chunks = chunks;
        }
        const int need_stride = 4 * max_chunks;
        const bool order_ok = g.h_first || !(g.scale_h && g.scale_v);
        // in-place h-scaled rows need 8 + 128 words inside a raw row (sk_kernels.cu), and a horizontal pass
        g.scale_inplace = (g.scale_h && need_stride >= 140 && !std::getenv("SELKIES_PREENCODE_SCALE_SEPARATE")) ? 1 : 0;
        // In-place windows are widened to the row stride the kernels were compiled for (kScaleStride32 /
        // kScaleStride16) when they fit one: the TMA box simply reads a few columns more, and every row step
        // of the kernel becomes an immediate offset.
        auto stride_for = [&](bool t32, int rows) {
            if (!g.scale_inplace || !(g.scale_h && g.scale_v)) return need_stride;     // classes: both axes only
            if (t32) return need_stride <= kScaleStride32 ? kScaleStride32 : (need_stride <= kScaleStride16n ? kScaleStride16n : need_stride);
            if (rows <= 24) return need_stride;
            return need_stride <= kScaleStride16n ? kScaleStride16n : (need_stride <= kScaleStride16 ? kScaleStride16 : need_stride);
        };
        // 32-row tiles (pass B twice per staged window) when the taller window still lets 4 CTAs share an
        // SM; SELKIES_PREENCODE_SCALE_TILE=16|32 overrides the choice for A/B runs and tests.
        const size_t sep_words = g.scale_inplace ? 0 : 132;
        const size_t smem32 = (size_t)max_rows32 * (sep_words + (size_t)stride_for(true, max_rows32)) * 4;
        bool tile32 = max_rows32 <= kScaleMaxRows && smem32 <= 56 * 1024 && g.out_h >= 64;
        if (const char *e = std::getenv("SELKIES_PREENCODE_SCALE_TILE")) {
            if (std::atoi(e) == 16) tile32 = false;
            else if (std::atoi(e) == 32) tile32 = max_rows32 <= kScaleMaxRows && smem32 <= 96 * 1024;
        }
        if (tile32) max_rows = max_rows32;
        g.scale_raw_stride = stride_for(tile32, max_rows);
        const size_t row_words = sep_words + (size_t)g.scale_raw_stride;
        g.scale_tile_rows = tile32 ? 32 : 16;
        g.scale_rows = max_rows;
        const size_t smem = (size_t)max_rows * row_words * 4;
        // (a growing picture is scaled behind the matrix: the tiled kernel is compiled for that with 32-row
        //  tiles and with 16-row tiles of up to 24 window rows only)
        g.scale_fast = (order_ok && max_rows <= kScaleMaxRows && max_chunks <= 96 && smem <= 96 * 1024 &&
                        (tile32 || max_rows <= (g.scale_after ? 24 : 40)) &&
                        (!g.scale_after || (g.scale_h && g.scale_v)) &&
                        !(c.flags & SK_FLAG_SCALE_GENERIC) && !s->generic_only) ? 1 : 0;
        // how far ahead a CTA asks for a window to be brought into L2: the CTAs resident on this GPU
        g.scale_prefetch = std::getenv("SELKIES_PREENCODE_SCALE_NO_PREFETCH") ? 0 : sk_scale_resident_ctas(g);
        SK_TRY(cudaMalloc((void **)&s->d_voff, off.size() * sizeof(int32_t)));
        SK_TRY(cudaMalloc((void **)&s->d_vtap, tap.size() * sizeof(int16_t)));
        SK_TRY(cudaMemcpy(s->d_voff, off.data(), off.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
        SK_TRY(cudaMemcpy(s->d_vtap, tap.data(), tap.size() * sizeof(int16_t), cudaMemcpyHostToDevice));
        if (g.scale_fast && g.scale_v) {
            // the tiled kernel's copy of the table: row and ready-made dp2a weight pair side by side, padded
            // to an even length so that rows (y, y+1) always load as one 16-byte word
            std::vector<uint2> pack(((size_t)g.out_h + 1) & ~(size_t)1);
            for (size_t j = 0; j < pack.size(); j++) {
                const size_t k = j < (size_t)g.out_h ? j : (size_t)g.out_h - 1;
                const uint32_t p4 = 4u * (uint32_t)tap[k];
                pack[j] = make_uint2((uint32_t)off[k], (256u - p4) | (p4 << 16));
            }
            SK_TRY(cudaMalloc((void **)&s->d_vpack, pack.size() * sizeof(uint2)));
            SK_TRY(cudaMemcpy(s->d_vpack, pack.data(), pack.size() * sizeof(uint2), cudaMemcpyHostToDevice));
        }
    }
    SK_TRY(cudaDeviceSynchronize());
#undef SK_TRY
    *out = s;
    return SK_OK;
}

int sk_session_destroy(sk_session *s)
{
    if (!s) return SK_ERR_ARG;
    free_session(s);
    return SK_OK;
}

int sk_get_input_staging(sk_session *s, int slot, void **host_ptr, size_t *bytes)
{
    if (!s || !host_ptr) return SK_ERR_ARG;
    if (check_slot(s, slot) != SK_OK) return SK_ERR_ARG;
    *host_ptr = s->slots[slot].h_in;
    if (bytes) *bytes = s->in_bytes_host;
    return SK_OK;
}

int sk_get_device_input(sk_session *s, int slot, void **dptr, size_t *stride)
{
    if (!s || !dptr) return SK_ERR_ARG;
    if (check_slot(s, slot) != SK_OK) return SK_ERR_ARG;
    *dptr = s->slots[slot].d_in;
    if (stride) *stride = (size_t)s->d_in_stride;
    return SK_OK;
}

int sk_upload(sk_session *s, int slot, const void *src, size_t src_stride)
{
    if (!s || !src) return SK_ERR_ARG;
    if (check_slot(s, slot) != SK_OK) return SK_ERR_ARG;
    const size_t row = (size_t)4 * s->cfg.in_width;
    if (src_stride == 0) src_stride = row;
    if (src_stride < row) return fail(s, SK_ERR_ARG, "src_stride smaller than a row");
    uint8_t *dst = s->slots[slot].h_in;
    if (src_stride == (size_t)s->cfg.in_stride) {
        std::memcpy(dst, src, s->in_bytes_host - (src_stride - row));
    } else {
        for (int y = 0; y < s->cfg.in_height; y++)
            std::memcpy(dst + (size_t)y * s->cfg.in_stride, (const uint8_t *)src + (size_t)y * src_stride, row);
    }
    return SK_OK;
}

int sk_submit(sk_session *s, int slot, uint64_t frame_id)
{
    if (!s) return SK_ERR_ARG;
    if (check_slot(s, slot) != SK_OK) return SK_ERR_ARG;
    SK_CUDA(s, cudaSetDevice(s->dev));
    int rc = enqueue_h2d(s, s, slot, s->slots[slot].h_in, (size_t)s->cfg.in_stride);
    if (rc != SK_OK) return rc;
    return submit_common(s, s->geom, s->slots[slot].d_in, slot, frame_id, nullptr);
}

int sk_submit_io(sk_session *s, int slot, const void *host_in, size_t in_stride, void *host_out, uint64_t frame_id)
{
    if (!s || !host_in) return SK_ERR_ARG;
    if (check_slot(s, slot) != SK_OK) return SK_ERR_ARG;
    const size_t row = (size_t)4 * s->cfg.in_width;
    if (in_stride == 0) in_stride = row;
    if (in_stride < row) return fail(s, SK_ERR_ARG, "stride smaller than a row");
    if (host_out && s->d_stale)
        return fail(s, SK_ERR_UNSUPPORTED, "a SK_FLAG_GATED_READBACK session keeps its own persistent host frames: host_out must be NULL");
    SK_CUDA(s, cudaSetDevice(s->dev));
    // Straight from the caller's buffer into the device slot: no trip through the pinned staging.
    // Either way the buffer must stay valid until sk_wait(frame_id) returns.
    int rc = enqueue_h2d(s, s, slot, host_in, in_stride);
    if (rc != SK_OK) return rc;
    return submit_common(s, s->geom, s->slots[slot].d_in, slot, frame_id, host_out);
}

int sk_submit_host(sk_session *s, int slot, const void *host_ptr, size_t stride, uint64_t frame_id)
{
    return sk_submit_io(s, slot, host_ptr, stride, nullptr, frame_id);
}

int sk_host_register(void *ptr, size_t bytes)
{
    if (!ptr || !bytes) return SK_ERR_ARG;
    cudaError_t e = cudaHostRegister(ptr, bytes, cudaHostRegisterDefault);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(nullptr, SK_ERR_CUDA, std::string("cudaHostRegister: ") + cudaGetErrorString(e));
    }
    return SK_OK;
}

int sk_host_unregister(void *ptr)
{
    if (!ptr) return SK_ERR_ARG;
    cudaError_t e = cudaHostUnregister(ptr);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(nullptr, SK_ERR_CUDA, std::string("cudaHostUnregister: ") + cudaGetErrorString(e));
    }
    return SK_OK;
}

int sk_submit_device(sk_session *s, const void *dptr, size_t stride, uint64_t frame_id)
{
    if (!s || !dptr) return SK_ERR_ARG;
    if (stride == 0) stride = (size_t)s->d_in_stride;
    if (((uintptr_t)dptr & 15) || (stride & 15)) return fail(s, SK_ERR_ALIGN, "device frame pointer and stride must be 16-byte aligned");
    if (stride < (size_t)4 * s->cfg.in_width) return fail(s, SK_ERR_ARG, "stride smaller than a row");
    if ((uint64_t)stride * (uint64_t)s->cfg.in_height >= (1ull << 32)) return fail(s, SK_ERR_ARG, "stride * in_height must stay below 4 GiB");
    SK_CUDA(s, cudaSetDevice(s->dev));
    SkGeom g = s->geom;
    g.in_stride = (int)stride;
    return submit_common(s, g, (const uint8_t *)dptr, -1, frame_id, nullptr);
}

int sk_batch_prepare(sk_session **ss, int n)
{
    if (!ss || n < 1 || !ss[0]) return SK_ERR_ARG;
    int rc = validate_batch(ss, n);
    if (rc != SK_OK) return rc;
    SK_CUDA(ss[0], cudaSetDevice(ss[0]->dev));
    return prepare_bmeta(ss[0]);
}

int sk_batch_submit(sk_session **ss, int n, const void *const *dptrs, size_t stride, uint64_t frame_id)
{
    if (!ss || !dptrs || n < 1) return SK_ERR_ARG;
    sk_session *L = ss[0];
    if (!L) return SK_ERR_ARG;
    // ---- everything that can be refused or has to be allocated comes before any ring is advanced
    int rc = validate_batch(ss, n);
    if (rc != SK_OK) return rc;
    if (stride == 0) stride = (size_t)L->d_in_stride;
    if (stride & 15) return fail(L, SK_ERR_ALIGN, "stride must be 16-byte aligned");
    if (stride < (size_t)4 * L->cfg.in_width) return fail(L, SK_ERR_ARG, "stride smaller than a row");
    if ((uint64_t)stride * (uint64_t)L->cfg.in_height >= (1ull << 32)) return fail(L, SK_ERR_ARG, "stride * in_height must stay below 4 GiB");
    for (int i = 0; i < n; i++) {
        if (!dptrs[i]) return fail(L, SK_ERR_ARG, "null frame in batch");
        if ((uintptr_t)dptrs[i] & 15) return fail(L, SK_ERR_ALIGN, "device frame pointers must be 16-byte aligned");
    }
    SK_CUDA(L, cudaSetDevice(L->dev));
    if ((rc = prepare_bmeta(L)) != SK_OK) return rc;     // no-op after sk_batch_prepare / the first batch

    const bool timing = L->cfg.flags & SK_FLAG_TIMING;
    const bool damage = !(L->cfg.flags & SK_FLAG_NO_DAMAGE);
    SkGeom g = L->geom;
    g.in_stride = (int)stride;
    SkBatch b;
    int ris[SK_MAX_BATCH];
    cudaEvent_t waited = nullptr;      // members of a steady batch share one event: wait on it once
    for (int i = 0; i < n; i++) {
        sk_session *s = ss[i];
        if ((rc = wait_for_ring_entry(L, s, L->s_comp, waited)) != SK_OK) { abandon_batch(ss, i); return rc; }
        s->ring[s->seq % (uint64_t)s->n].had_h2d = false;
        ris[i] = stage_frame(s, (const uint8_t *)dptrs[i], frame_id, b.f[i]);
    }
    const int lri = ris[0];
    Ring &lr = L->ring[lri];
    if (damage) {
        // All members' maps land in one block owned by the leader, so the read-back is ONE copy per
        // step instead of one small copy per session.
        for (int i = 0; i < n; i++) {
            Ring &r = ss[i]->ring[ris[i]];
            r.d_meta_view = L->d_bmeta[lri] + kCountHdr + (size_t)i * L->meta_bytes;
            r.h_meta_view = L->h_bmeta[lri] + kCountHdr + (size_t)i * L->meta_bytes;
            const unsigned set = L->bmeta_use[lri] & 1u;
            r.h_count = reinterpret_cast<const uint32_t *>(L->h_bmeta[lri]) + set * kCountSet + i;
            r.d_count = reinterpret_cast<const uint32_t *>(L->d_bmeta[lri]) + set * kCountSet + i;
            b.f[i].count_out = reinterpret_cast<uint32_t *>(L->d_bmeta[lri]) + set * kCountSet + i;
            // the other set is zeroed for the block's next use — all of it, by the first member's frame:
            // the next batch may have more members than this one
            b.f[i].count_zero = i == 0 ? reinterpret_cast<uint32_t *>(L->d_bmeta[lri]) + (set ^ 1u) * kCountSet : nullptr;
            b.f[i].count_zero_n = i == 0 ? kCountSet : 0;
            b.f[i].map = r.d_meta_view;
            b.f[i].policy = r.d_meta_view + L->map_bytes;
        }
    }
#define SK_BATCH_CUDA(call)                                                                          \
    do {                                                                                             \
        cudaError_t e_ = (call);                                                                     \
        if (e_ != cudaSuccess) {                                                                     \
            abandon_batch(ss, n);                                                                    \
            return fail(L, SK_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_));         \
        }                                                                                            \
    } while (0)
    if (damage) L->bmeta_use[lri]++;
    if (timing) SK_BATCH_CUDA(cudaEventRecord(lr.k_start, L->s_comp));
    SK_BATCH_CUDA(launch(ss, g, b, n, L->s_comp));
    SK_BATCH_CUDA(cudaEventRecord(lr.k_done, L->s_comp));
    // Read-back of every member rides on the leader's d2h stream; one event closes the batch.
    SK_BATCH_CUDA(cudaStreamWaitEvent(L->s_d2h, lr.k_done, 0));
    if (timing) SK_BATCH_CUDA(cudaEventRecord(lr.d2h_start, L->s_d2h));
    if (damage)
        SK_BATCH_CUDA(cudaMemcpyAsync(L->h_bmeta[lri], L->d_bmeta[lri], kCountHdr + (size_t)n * L->meta_bytes,
                                      cudaMemcpyDeviceToHost, L->s_d2h));
    for (int i = 0; i < n; i++) {
        if ((rc = enqueue_frame_readback(L, ss[i], ris[i], L->s_d2h, nullptr, false)) != SK_OK) { abandon_batch(ss, n); return rc; }
    }
    SK_BATCH_CUDA(cudaEventRecord(lr.done, L->s_d2h));
    for (int i = 0; i < n; i++) {
        sk_session *s = ss[i];
        Ring &r = s->ring[ris[i]];
        r.wait_on = lr.done;
        r.kernel_done = lr.k_done;
        r.used = true;
        s->last_kernel_stream = L->s_comp;
        s->last_kernel_event = lr.k_done;
    }
    return SK_OK;
}

int sk_batch_submit_host(sk_session **ss, int n, const void *const *host_in, size_t in_stride,
                         void *const *host_out, uint64_t frame_id)
{
    if (!ss || !host_in || n < 1) return SK_ERR_ARG;
    sk_session *L = ss[0];
    if (!L) return SK_ERR_ARG;
    int rc = validate_batch(ss, n);
    if (rc != SK_OK) return rc;
    const size_t row = (size_t)4 * L->cfg.in_width;
    if (in_stride == 0) in_stride = row;
    if (in_stride < row) return fail(L, SK_ERR_ARG, "stride smaller than a row");
    for (int i = 0; i < n; i++)
        if (!host_in[i]) return fail(L, SK_ERR_ARG, "null frame in batch");
    SK_CUDA(L, cudaSetDevice(L->dev));

    const bool timing = L->cfg.flags & SK_FLAG_TIMING;
    SkBatch b;
    int ris[SK_MAX_BATCH];
    int slots[SK_MAX_BATCH];
    cudaEvent_t waited = nullptr;
    for (int i = 0; i < n; i++) {
        sk_session *s = ss[i];
        // each member's frame goes up on its own copy stream, into the device slot that pairs with
        // the ring entry it is about to fill; the leader's compute stream waits for every upload
        slots[i] = (int)(s->seq % (uint64_t)s->n);
        Ring &r = s->ring[slots[i]];
        if ((rc = enqueue_h2d(L, s, slots[i], host_in[i], in_stride)) != SK_OK) { abandon_batch(ss, i); return rc; }
        if ((rc = wait_for_ring_entry(L, s, L->s_comp, waited)) != SK_OK) { abandon_batch(ss, i); return rc; }
        cudaError_t e = cudaStreamWaitEvent(L->s_comp, r.h2d_done, 0);
        if (e != cudaSuccess) { abandon_batch(ss, i); return fail(L, SK_ERR_CUDA, std::string("cudaStreamWaitEvent: ") + cudaGetErrorString(e)); }
        r.had_h2d = true;
        ris[i] = stage_frame(s, s->slots[slots[i]].d_in, frame_id, b.f[i]);
    }
    Ring &lr = L->ring[ris[0]];
    if (timing) SK_BATCH_CUDA(cudaEventRecord(lr.k_start, L->s_comp));
    SK_BATCH_CUDA(launch(ss, L->geom, b, n, L->s_comp));
    SK_BATCH_CUDA(cudaEventRecord(lr.k_done, L->s_comp));
    for (int i = 0; i < n; i++) {
        Slot &sl = ss[i]->slots[slots[i]];
        SK_BATCH_CUDA(cudaEventRecord(sl.consumed, L->s_comp));
        sl.consumed_valid = true;
    }
    SK_BATCH_CUDA(cudaStreamWaitEvent(L->s_d2h, lr.k_done, 0));
    if (timing) SK_BATCH_CUDA(cudaEventRecord(lr.d2h_start, L->s_d2h));
    for (int i = 0; i < n; i++) {
        void *ho = host_out ? host_out[i] : nullptr;
        if ((rc = enqueue_frame_readback(L, ss[i], ris[i], L->s_d2h, ho, true)) != SK_OK) { abandon_batch(ss, n); return rc; }
    }
    SK_BATCH_CUDA(cudaEventRecord(lr.done, L->s_d2h));
#undef SK_BATCH_CUDA
    for (int i = 0; i < n; i++) {
        sk_session *s = ss[i];
        Ring &r = s->ring[ris[i]];
        r.wait_on = lr.done;
        r.kernel_done = lr.k_done;
        r.used = true;
        s->last_kernel_stream = L->s_comp;
        s->last_kernel_event = lr.k_done;
    }
    return SK_OK;
}

namespace {
void describe(sk_session *s, Ring &r, uint64_t frame_id, sk_result *out, bool host_side)
{
    std::memset(out, 0, sizeof(*out));
    out->frame_id = frame_id;
    out->dev_frame = r.d_out;
    out->frame_bytes = s->frame_bytes;
    for (int i = 0; i < 3; i++) {
        out->plane_offset[i] = s->plane_offset[i];
        out->plane_stride[i] = s->plane_stride[i];
    }
    out->mb_w = s->geom.mb_w;
    out->mb_h = s->geom.mb_h;
    const bool damage = !(s->cfg.flags & SK_FLAG_NO_DAMAGE);
    if (damage) {
        out->dev_damage_map = r.d_meta_view;
        out->dev_damaged_count = r.d_count;
        if (s->d_state) out->dev_policy_map = r.d_meta_view + s->map_bytes;
    }
    if (!host_side) return;
    if (r.h_out_user) out->host_frame = r.h_out_user;
    else if (s->cfg.flags & SK_FLAG_HOST_OUTPUT) out->host_frame = r.h_out;
    if (out->host_frame)
        out->readback_bytes = r.gated ? (uint64_t)*r.h_blocks * 384u : s->frame_bytes;
    if (damage) {
        out->host_damage_map = r.h_meta_view;
        out->damaged_count = *r.h_count;
        if (s->d_state) out->host_policy_map = r.h_meta_view + s->map_bytes;
    }
}
}  // namespace

int sk_wait(sk_session *s, uint64_t frame_id, sk_result *out)
{
    if (!s) return SK_ERR_ARG;
    const int ri = find_ring(s, frame_id);
    if (ri < 0) return fail(s, SK_ERR_STATE, "unknown frame id (never submitted, or its ring entry was reused)");
    Ring &r = s->ring[ri];
    SK_CUDA(s, cudaSetDevice(s->dev));
    SK_CUDA(s, cudaEventSynchronize(r.wait_on));
    // The host has seen this frame complete: later submissions need not chain on its events again (a
    // batch leader that owned them may even be gone by then).
    r.waited = true;
    if (s->last_kernel_event == r.kernel_done) s->last_kernel_stream = nullptr;
    const bool timing = s->cfg.flags & SK_FLAG_TIMING;
    const bool own_events = r.wait_on == r.done;
    if (timing && own_events) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, r.k_start, r.k_done) == cudaSuccess) s->last_kernel_us = ms * 1e3f;
        else cudaGetLastError();
    }
    if (out) {
        describe(s, r, frame_id, out, true);
        if (timing && own_events) {
            float ms = 0.f;
            out->kernel_us = s->last_kernel_us;
            if (r.had_h2d && cudaEventElapsedTime(&ms, r.h2d_start, r.h2d_done) == cudaSuccess) out->h2d_us = ms * 1e3f;
            if (cudaEventElapsedTime(&ms, r.d2h_start, r.done) == cudaSuccess) out->d2h_us = ms * 1e3f;
            cudaGetLastError();
        }
    }
    return SK_OK;
}

int sk_read_frame(sk_session *s, uint64_t frame_id, void *host_out)
{
    if (!s || !host_out) return SK_ERR_ARG;
    const int ri = find_ring(s, frame_id);
    if (ri < 0) return fail(s, SK_ERR_STATE, "unknown frame id (never submitted, or its ring entry was reused)");
    Ring &r = s->ring[ri];
    if (!r.kernel_done) return fail(s, SK_ERR_STATE, "frame has no kernel queued");
    SK_CUDA(s, cudaSetDevice(s->dev));
    SK_CUDA(s, cudaStreamWaitEvent(s->s_d2h, r.kernel_done, 0));
    SK_CUDA(s, cudaMemcpyAsync(host_out, r.d_out, s->frame_bytes, cudaMemcpyDeviceToHost, s->s_d2h));
    SK_CUDA(s, cudaStreamSynchronize(s->s_d2h));
    return SK_OK;
}

int sk_wait_device(sk_session *s, uint64_t frame_id, void *cuda_stream, sk_result *out)
{
    if (!s) return SK_ERR_ARG;
    const int ri = find_ring(s, frame_id);
    if (ri < 0) return fail(s, SK_ERR_STATE, "unknown frame id (never submitted, or its ring entry was reused)");
    Ring &r = s->ring[ri];
    if (!r.kernel_done) return fail(s, SK_ERR_STATE, "frame has no kernel queued");
    SK_CUDA(s, cudaSetDevice(s->dev));
    SK_CUDA(s, cudaStreamWaitEvent((cudaStream_t)cuda_stream, r.kernel_done, 0));
    if (out) describe(s, r, frame_id, out, false);
    return SK_OK;
}

int sk_release_device(sk_session *s, uint64_t frame_id, void *cuda_stream)
{
    if (!s) return SK_ERR_ARG;
    const int ri = find_ring(s, frame_id);
    if (ri < 0) return fail(s, SK_ERR_STATE, "unknown frame id (never submitted, or its ring entry was reused)");
    Ring &r = s->ring[ri];
    SK_CUDA(s, cudaSetDevice(s->dev));
    SK_CUDA(s, cudaEventRecord(r.consumer_done, (cudaStream_t)cuda_stream));
    r.consumer_pending = true;
    return SK_OK;
}

int sk_batch_wait(sk_session **ss, int n, uint64_t frame_id, uint32_t *damaged_counts)
{
    if (!ss || n < 1 || !ss[0]) return SK_ERR_ARG;
    for (int i = 0; i < n; i++) {
        if (!ss[i]) return fail(ss[0], SK_ERR_ARG, "null session in batch");
        if (i == 0 || !damaged_counts) {
            sk_result r;
            int rc = sk_wait(ss[i], frame_id, &r);
            if (rc != SK_OK) {
                if (i > 0) ss[0]->err = ss[i]->err;
                return rc;
            }
            if (damaged_counts) damaged_counts[i] = r.damaged_count;
            continue;
        }
        // members of one batch share the leader's completion event: after the leader's wait only the
        // host-side sum is left (a member submitted outside the batch falls back to a full sk_wait)
        sk_session *s = ss[i];
        const int ri = find_ring(s, frame_id);
        if (ri < 0) return fail(ss[0], SK_ERR_STATE, "unknown frame id in batch");
        Ring &r = s->ring[ri];
        const int lri = find_ring(ss[0], frame_id);
        if (lri < 0 || r.wait_on != ss[0]->ring[lri].wait_on || (s->cfg.flags & SK_FLAG_NO_DAMAGE)) {
            sk_result res;
            int rc = sk_wait(s, frame_id, &res);
            if (rc != SK_OK) { ss[0]->err = s->err; return rc; }
            damaged_counts[i] = res.damaged_count;
            continue;
        }
        r.waited = true;
        if (s->last_kernel_event == r.kernel_done) s->last_kernel_stream = nullptr;
        damaged_counts[i] = *r.h_count;
    }
    return SK_OK;
}

int sk_last_kernel_us(sk_session *s, float *us)
{
    if (!s || !us) return SK_ERR_ARG;
    if (!(s->cfg.flags & SK_FLAG_TIMING)) return fail(s, SK_ERR_STATE, "session was created without SK_FLAG_TIMING");
    *us = s->last_kernel_us;
    return SK_OK;
}

int sk_marker_record(sk_session *s, int which)
{
    if (!s || which < 0 || which > 1) return SK_ERR_ARG;
    SK_CUDA(s, cudaSetDevice(s->dev));
    SK_CUDA(s, cudaEventRecord(s->mark[which], s->s_comp));
    return SK_OK;
}

int sk_marker_elapsed_us(sk_session *s, float *us)
{
    if (!s || !us) return SK_ERR_ARG;
    SK_CUDA(s, cudaSetDevice(s->dev));
    SK_CUDA(s, cudaEventSynchronize(s->mark[1]));
    float ms = 0.f;
    SK_CUDA(s, cudaEventElapsedTime(&ms, s->mark[0], s->mark[1]));
    *us = ms * 1e3f;
    return SK_OK;
}

int sk_pcie_probe(int device_id, size_t bytes, int iters, int mode, double *h2d_gbs, double *d2h_gbs)
{
    if (bytes == 0 || iters < 1 || mode < 0 || mode > 2) return SK_ERR_ARG;
    constexpr size_t kRot = 8;
    if (h2d_gbs) *h2d_gbs = 0.0;
    if (d2h_gbs) *d2h_gbs = 0.0;
    if (device_id < 0) device_id = 0;
    void *h[2] = {nullptr, nullptr}, *d[2] = {nullptr, nullptr};
    cudaStream_t st[2] = {nullptr, nullptr};
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    cudaError_t e = cudaSetDevice(device_id);
    for (int i = 0; i < 2 && e == cudaSuccess; i++) {
        // kRot distinct host buffers, used in turn: a single buffer copied over and over can be served from
        // the host's last-level cache, which flatters the rate real frames (a new buffer each) ever see
        e = cudaHostAlloc(&h[i], bytes * kRot, cudaHostAllocDefault);
        if (e == cudaSuccess) { std::memset(h[i], 0x5a, bytes * kRot); e = cudaMalloc(&d[i], bytes); }
        if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&st[i], cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaEventCreate(&ev[2 * i]);
        if (e == cudaSuccess) e = cudaEventCreate(&ev[2 * i + 1]);
    }
    const bool up = mode != 1, down = mode != 0;
    if (e == cudaSuccess) {
        // one untimed pass each way, then the timed loops (both started before either is waited for)
        if (up) e = cudaMemcpyAsync(d[0], h[0], bytes, cudaMemcpyHostToDevice, st[0]);
        if (e == cudaSuccess && down) e = cudaMemcpyAsync(h[1], d[1], bytes, cudaMemcpyDeviceToHost, st[1]);
        if (e == cudaSuccess) e = cudaDeviceSynchronize();
    }
    if (e == cudaSuccess) {
        if (up) cudaEventRecord(ev[0], st[0]);
        if (down) cudaEventRecord(ev[2], st[1]);
        for (int k = 0; k < iters && e == cudaSuccess; k++) {
            const size_t o = ((size_t)k % kRot) * bytes;
            if (up) e = cudaMemcpyAsync(d[0], static_cast<char *>(h[0]) + o, bytes, cudaMemcpyHostToDevice, st[0]);
            if (e == cudaSuccess && down)
                e = cudaMemcpyAsync(static_cast<char *>(h[1]) + o, d[1], bytes, cudaMemcpyDeviceToHost, st[1]);
        }
        if (up) cudaEventRecord(ev[1], st[0]);
        if (down) cudaEventRecord(ev[3], st[1]);
        if (e == cudaSuccess) e = cudaDeviceSynchronize();
    }
    if (e == cudaSuccess) {
        float ms = 0.f;
        if (up && cudaEventElapsedTime(&ms, ev[0], ev[1]) == cudaSuccess && ms > 0.f && h2d_gbs)
            *h2d_gbs = (double)bytes * iters / (ms * 1e-3) / 1e9;
        if (down && cudaEventElapsedTime(&ms, ev[2], ev[3]) == cudaSuccess && ms > 0.f && d2h_gbs)
            *d2h_gbs = (double)bytes * iters / (ms * 1e-3) / 1e9;
    }
    for (int i = 0; i < 2; i++) {
        if (h[i]) cudaFreeHost(h[i]);
        if (d[i]) cudaFree(d[i]);
        if (st[i]) cudaStreamDestroy(st[i]);
    }
    for (cudaEvent_t v : ev) if (v) cudaEventDestroy(v);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(nullptr, SK_ERR_CUDA, std::string("sk_pcie_probe: ") + cudaGetErrorString(e));
    }
    return SK_OK;
}

int sk_reset_history(sk_session *s)
{
    if (!s) return SK_ERR_ARG;
    s->prev_ring = -1;
    return SK_OK;
}

}  // extern "C"
