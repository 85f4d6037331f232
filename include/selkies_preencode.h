/*
 * selkies_preencode.h — C ABI of the B200 pre-encode stage.
 *
 * One shared library (libselkies_preencode.so, hand-written sm_100a CUDA + a
 * small C++ runtime) replaces the elements the reference instantiates between
 * ximagesrc and the video encoder:
 *
 *   videoconvert (+ capsfilter NV12 / I420)   src/selkies/legacy/gstwebrtc_app.py:611-617,
 *                                             :643-649, :669-675, :686-692, :725-731, :742-748, :767-773
 *   cudaupload + cudaconvert (+ NV12 caps)    src/selkies/legacy/gstwebrtc_app.py:263-284, :370-380, :427-437
 *   (would-be) videoscale method=bilinear     no call site; resolution comes from
 *                                             src/selkies/legacy/resize.py:15-29
 *
 * It is bound from Python with ctypes (selkies_gstreamer_b200/_abi.py); the
 * GstBase.BaseTransform element in selkies_gstreamer_b200/gst_element.py is the
 * drop-in for the factories above.  Plain C types only; no exception crosses;
 * every function returns an int status (0 = SK_OK, negative = error) unless
 * stated otherwise.  There is no CPU fallback: without a CUDA device every
 * entry point that touches a frame fails with SK_ERR_CUDA.
 *
 * Ownership: the library owns all device memory, pinned staging, streams and
 * events of a session.  Callers own their frame buffers and only borrow the
 * pointers returned in sk_result until the slot / ring entry is reused.
 * Threading: one producer thread per session (GStreamer's streaming thread);
 * different sessions are independent.
 */
#ifndef SELKIES_PREENCODE_H
#define SELKIES_PREENCODE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SK_ABI_VERSION 3

/* status codes */
enum {
    SK_OK = 0,
    SK_ERR_ARG = -1,          /* bad pointer / enum / geometry (odd or out-of-range size)      */
    SK_ERR_CUDA = -2,         /* CUDA runtime error, incl. "no device"; see sk_last_error()    */
    SK_ERR_ALIGN = -3,        /* device input pointer / stride not 16-byte aligned             */
    SK_ERR_STATE = -4,        /* slot busy, unknown frame id, session mismatch in a batch      */
    SK_ERR_UNSUPPORTED = -5,  /* e.g. a scale_method other than bilinear */
    SK_ERR_NOMEM = -6
};

/* sink caps formats that ximagesrc can negotiate (gstwebrtc_app.py:210-255; docs/component.md:336) */
enum { SK_IN_BGRX = 0, SK_IN_RGBX = 1, SK_IN_BGRA = 0, SK_IN_RGBA = 1 };
/* src caps formats forced by the capsfilter after the stage (gstwebrtc_app.py:615 NV12, :647 I420) */
enum { SK_OUT_NV12 = 0, SK_OUT_I420 = 1 };
/* colorimetry of the src caps; GStreamer's default by output height is applied by the Python host */
enum { SK_MATRIX_BT601 = 0, SK_MATRIX_BT709 = 1, SK_MATRIX_BT2020 = 2 };
/* chroma-site of the src caps */
enum { SK_SITE_NONE = 0, SK_SITE_H_COSITED = 1 };
/* videoscale "method": only bilinear (2-tap) is required */
enum { SK_SCALE_BILINEAR = 1 };
/* resampler restatement variant ("scale") */
enum { SK_SCALE_VARIANT_GST_LQ = 0, SK_SCALE_VARIANT_PLAIN_LQ = 1 };

/* sk_config.flags */
enum {
    SK_FLAG_NO_DAMAGE   = 1 << 0,  /* convert-only: no previous-frame read, no map                   */
    SK_FLAG_HOST_OUTPUT = 1 << 1,  /* copy the converted frame back to pinned host memory each frame */
    SK_FLAG_TIMING      = 1 << 2,  /* record CUDA events; sk_result.*_us are filled                  */
    SK_FLAG_POLICY      = 1 << 3,  /* keep per-macroblock temporal counters in the kernel and return
                                      the SK_POLICY_* map (hot / paint-over), see below              */
    SK_FLAG_SCALE_GENERIC = 1 << 4, /* scale with the gather kernel even where the shared-memory tiled
                                      kernel applies (A/B measurements, tests); results are identical  */
    SK_FLAG_SPIN_WAIT   = 1 << 5,  /* sk_wait busy-polls (lowest latency, burns one host core).  Default:
                                      the waiting thread sleeps (cudaEventBlockingSync) — this stage exists
                                      to take load OFF the host cores (docs/design.md:33)               */
    SK_FLAG_GATED_READBACK = 1 << 6 /* with SK_FLAG_HOST_OUTPUT: sk_result.host_frame is a persistent pinned
                                      frame per ring entry into which only the macroblocks that changed
                                      since that copy was last refreshed are written (zero-copy stores
                                      from the GPU) — a mostly static desktop (BASELINE.json config 5)
                                      sends a few KB back instead of the whole frame.  Not with batches
                                      or a caller-supplied output buffer.                               */
};

/* sk_config.parity_switches — one bit per recollection of upstream GStreamer that this restatement
 * could have wrong.  0 = the arithmetic as restated, and the fast kernels.
 * GStreamer 1.24.12 (the reference's pin, addons/gstreamer/Dockerfile:85-93) cannot be run where this
 * was built; whoever can, produces golden vectors with tools/make_gst_golden.py and
 * tests/test_gst_golden.py then either passes or names the bit that makes it pass.  Bits 0,2,3,4 change
 * per-pixel arithmetic and route the session through the generic kernel (slower, same traffic) until
 * the fast kernels are updated to the confirmed behaviour; bits 1,5,6 only change host-derived
 * constants and keep the fast kernels. */
enum {
    SK_PS_MATRIX_ROUND   = 1 << 0,  /* R8: +128 rounding term before the >>8 of the 8-bit matrix           */
    SK_PS_COEFF_HALF_UP  = 1 << 1,  /* R8: coefficients rounded half away from zero instead of rint()      */
    SK_PS_CS_EDGE_KEEP   = 1 << 2,  /* R3: cosited filter leaves the first and last chroma sample as is    */
    SK_PS_CHROMA_V_FIRST = 1 << 3,  /* R4: vertical average first, horizontal filter second                */
    SK_PS_CHROMA_BOX4    = 1 << 4,  /* R4: chroma-site none: (a+b+c+d+2)>>2 in one rounding                */
    SK_PS_HINC_PLAIN     = 1 << 5,  /* R6: horizontal increment without the "-1"                           */
    SK_PS_V_FIRST_ON_TIE = 1 << 6   /* R7: equal cost of the two scale orders goes to vertical first       */
};

typedef struct sk_session sk_session;

/*
 * Replaces: the GObject properties + negotiated caps of the stage
 * (cuda-device-id gstwebrtc_app.py:264-272; n-threads/qos :612-613; caps :614-617).
 */
typedef struct sk_config {
    uint32_t struct_size;     /* = sizeof(sk_config); ABI versioning                              */
    int32_t  device_id;       /* cuda-device-id; -1 = device 0                                     */
    int32_t  in_width, in_height;
    int32_t  in_stride;       /* bytes per captured row; 0 = 4*in_width                            */
    int32_t  in_format;       /* SK_IN_*                                                           */
    int32_t  out_width, out_height;   /* 0 = same as input; smaller or larger (a picture that grows is
                                         scaled behind the matrix, as upstream's chain_scale() does)  */
    int32_t  out_format;      /* SK_OUT_*                                                          */
    int32_t  matrix;          /* SK_MATRIX_*                                                       */
    int32_t  chroma_site;     /* SK_SITE_*                                                         */
    int32_t  scale_method;    /* SK_SCALE_BILINEAR (ignored when sizes are equal)                  */
    int32_t  scale_variant;   /* SK_SCALE_VARIANT_*                                                */
    int32_t  n_inflight;      /* staging slots and output ring depth, 2..8 (0 = 2)                 */
    uint32_t flags;           /* SK_FLAG_*                                                         */
    /* Temporal damage policy, the knobs the reference hands to its capture module
     * (src/selkies/selkies.py:2949-2951); 0 = that default; each 1..255.                         */
    int32_t  damage_block_threshold;     /* consecutive damaged frames that make a block "hot" (10) */
    int32_t  damage_block_duration;      /* frames a hot block stays hot after its last change (20) */
    int32_t  paint_over_trigger_frames;  /* still frames after which a changed block is repainted (15) */
    uint32_t parity_switches;            /* SK_PS_*; 0 = as restated                                    */
} sk_config;

/* Per-macroblock policy flags (sk_result.host_policy_map), updated inside the convert kernel. */
enum {
    SK_POLICY_DAMAGED    = 1,  /* differs from the previous frame (same as the damage map)          */
    SK_POLICY_HOT        = 2,  /* has kept changing: treat as motion, hold quality, no paint-over   */
    SK_POLICY_PAINT_OVER = 4   /* went still long enough: send one high-quality refresh now         */
};

/* What one converted frame looks like to the caller (GstVideoInfo default layout). */
typedef struct sk_result {
    uint64_t frame_id;
    const uint8_t *dev_frame;     /* packed output frame in HBM                                   */
    const uint8_t *host_frame;    /* pinned copy when SK_FLAG_HOST_OUTPUT, else NULL              */
    size_t   frame_bytes;
    size_t   plane_offset[3];     /* Y, U (or interleaved UV), V (0 for NV12)                     */
    int32_t  plane_stride[3];
    const uint8_t *host_damage_map; /* pinned, uint8[mb_h][mb_w], 1 = damaged; NULL if NO_DAMAGE  */
    const uint8_t *dev_damage_map;
    int32_t  mb_w, mb_h;
    uint32_t damaged_count;       /* exact sum of the map (reduced on the GPU with the map)        */
    float    h2d_us, kernel_us, d2h_us;   /* only with SK_FLAG_TIMING                             */
    const uint8_t *host_policy_map; /* pinned, uint8[mb_h][mb_w] of SK_POLICY_* bits; NULL without SK_FLAG_POLICY */
    const uint8_t *dev_policy_map;
    const uint32_t *dev_damaged_count; /* the count word in HBM (what a device-side consumer reads)    */
    uint64_t readback_bytes;      /* frame bytes that crossed PCIe for host_frame: frame_bytes, or with
                                     SK_FLAG_GATED_READBACK 384 B per macroblock actually written     */
} sk_result;

/* Replaces: Gst.ElementFactory.make("videoconvert"/"cudaconvert") + set_caps
 * (gstwebrtc_app.py:611, :270). Allocates everything the session will ever use. */
int sk_session_create(const sk_config *cfg, sk_session **out);

/* Replaces: element finalize at pipeline teardown (gstwebrtc_app.py:1727-1742). */
int sk_session_destroy(sk_session *s);

/* Pinned, library-owned staging for slot in [0, n_inflight): the capture side
 * writes the BGRx frame here (replaces cudaupload's GstBuffer input,
 * gstwebrtc_app.py:263-265).  Rows are in_stride bytes apart. */
int sk_get_input_staging(sk_session *s, int slot, void **host_ptr, size_t *bytes);

/* Device-side copy of staging slot `slot` (valid after an sk_submit of that
 * slot has completed): lets a caller re-submit a resident frame with
 * sk_submit_device / sk_batch_submit without another upload. */
int sk_get_device_input(sk_session *s, int slot, void **dptr, size_t *stride);

/* Copy a caller-owned (pageable or pinned) frame into staging slot `slot`
 * on the calling thread (replaces the GstBuffer map + upload of cudaupload). */
int sk_upload(sk_session *s, int slot, const void *src, size_t src_stride);

/* Asynchronous: H2D of staging slot on the copy stream, then the fused
 * convert[+scale]+damage kernel on the compute stream, then D2H of the map
 * (and of the frame with SK_FLAG_HOST_OUTPUT).  Replaces one
 * GstBaseTransform::transform call of the stage (upstream), as driven by the
 * pipeline assembled at gstwebrtc_app.py:941-994. */
int sk_submit(sk_session *s, int slot, uint64_t frame_id);

/* Same, but the H2D copy reads the caller's own frame buffer (rows `stride` apart, 0 = tight)
 * instead of the staging slot — the GstBuffer of ximagesrc, or its XShm segment
 * (gstwebrtc_app.py:210-241) — saving the host-side memcpy of sk_upload.  `slot` only names the
 * device-side input slot.  With memory pinned by sk_host_register (or cudaHostAlloc) the copy is
 * an asynchronous DMA; pageable memory is staged by the driver.  The buffer must stay valid and
 * unchanged until sk_wait(frame_id) returns. */
int sk_submit_host(sk_session *s, int slot, const void *host_ptr, size_t stride, uint64_t frame_id);

/* Same as sk_submit_host, and the converted frame is copied back INTO THE CALLER'S buffer
 * `host_out` (frame_bytes of it; NULL = as configured by SK_FLAG_HOST_OUTPUT) instead of the
 * library's pinned ring: the GstBuffer a GstBaseTransform hands to transform() as `outbuf`
 * (upstream; the reference's pipeline links the stage's src pad to the encoder at
 * gstwebrtc_app.py:991-994).  sk_result.host_frame is then host_out.  Pin recycled pool buffers
 * once with sk_host_register so both copies are asynchronous DMA without a host-side memcpy. */
int sk_submit_io(sk_session *s, int slot, const void *host_in, size_t in_stride, void *host_out,
                 uint64_t frame_id);

/* Pin / unpin a caller-owned capture buffer (e.g. the XShm segment) for sk_submit_host.
 * Pinning costs milliseconds: do it once per buffer, not per frame. */
int sk_host_register(void *ptr, size_t bytes);
int sk_host_unregister(void *ptr);

/* Same, for a frame that is already in HBM (video/x-raw(memory:CUDAMemory)
 * input, gstwebrtc_app.py:281-284, and the device-resident benchmark).
 * dptr and stride must be 16-byte aligned. */
int sk_submit_device(sk_session *s, const void *dptr, size_t stride, uint64_t frame_id);

/* Block until frame_id is complete and describe it. */
int sk_wait(sk_session *s, uint64_t frame_id, sk_result *out);

/* Copy the converted frame of frame_id from HBM into `host_out` (frame_bytes) and block until it is
 * there — for sessions without SK_FLAG_HOST_OUTPUT whose host occasionally does need the pixels
 * (a screenshot, a checker), while the ring entry is still valid.  The reference has no
 * counterpart: its stage always hands the whole frame downstream (gstwebrtc_app.py:991-994). */
int sk_read_frame(sk_session *s, uint64_t frame_id, void *host_out);

/* Device-side completion for consumers that stay on the GPU — the reference's primary branch:
 * `video/x-raw(memory:CUDAMemory),format=NV12` into nvh264enc / nvh265enc / nvav1enc
 * (gstwebrtc_app.py:281-284, :286-367, :370-380, :427-437).  Makes `cuda_stream` (a cudaStream_t of
 * the session's device) wait for the kernel that produced frame_id; the host does not block and no
 * frame D2H is needed (create the session without SK_FLAG_HOST_OUTPUT).  `out` receives the device
 * pointers (dev_frame, dev_damage_map, dev_policy_map, dev_damaged_count, layout); its host_*
 * fields are NULL and damaged_count is 0 — use sk_wait for the host copies of map and count. */
int sk_wait_device(sk_session *s, uint64_t frame_id, void *cuda_stream, sk_result *out);

/* The consumer's half of that contract: call after queueing the last read of frame_id's device
 * buffers on `cuda_stream`.  The ring entry is not overwritten (n_inflight frames later) before the
 * work queued on cuda_stream up to this point has finished.  Without this call the ring depth alone
 * protects the consumer, as for sk_wait. */
int sk_release_device(sk_session *s, uint64_t frame_id, void *cuda_stream);

/* Next frame is treated as having no predecessor: map = all ones.  Replaces the
 * renegotiation after a resize (src/selkies/legacy/webrtc_input.py:672-678 →
 * resize.py:75-117) or a keyframe request. */
int sk_reset_history(sk_session *s);

/* One launch over n sessions of identical configuration on the same device
 * (throughput mode: independent desktop sessions, gstwebrtc_app.py:264-272
 * pins each to a GPU).  dptrs[i] is session i's frame in HBM (stride as in
 * sk_submit_device); all get frame_id.  Completion via sk_wait on each (or sk_batch_wait).
 * The members share the first session's streams and completion event: destroy the sessions of a
 * batch together, after the last batch has been waited for. */
int sk_batch_submit(sk_session **sessions, int n, const void *const *dptrs,
                    size_t stride, uint64_t frame_id);

/* Allocate the batch bookkeeping of a group led by sessions[0] (one contiguous map block per ring
 * entry) ahead of time, so that no later sk_batch_submit* allocates on the frame path.  Optional:
 * the first sk_batch_submit of a leader does the same, before anything is queued. */
int sk_batch_prepare(sk_session **sessions, int n);

/* sk_batch_submit for frames that are still in HOST memory — what N stage instances of one process
 * hand over when their transform() calls arrive together (the reference builds several apps per
 * process, src/selkies/legacy/webrtc.py:684-685; stage call sites gstwebrtc_app.py:611-617):
 * every member's frame goes up on that member's own copy stream (host_in[i], rows in_stride apart,
 * 0 = tight; pinned or registered memory is asynchronous DMA), then ONE launch converts all of them,
 * then the frames come back — into host_out[i] when host_out and host_out[i] are non-NULL, else as
 * the member's SK_FLAG_HOST_OUTPUT says.  Membership may change from call to call.  Each member's
 * maps land in its own meta buffers.  Completion via sk_wait on each member (or sk_batch_wait). */
int sk_batch_submit_host(sk_session **sessions, int n, const void *const *host_in, size_t in_stride,
                         void *const *host_out, uint64_t frame_id);

/* Wait for frame_id on every session of a batch; damaged_counts (may be NULL)
 * receives one exact count per session.  Maps stay readable through sk_wait. */
int sk_batch_wait(sk_session **sessions, int n, uint64_t frame_id, uint32_t *damaged_counts);

/* Device-side time of the last `sk_batch_submit`/`sk_submit*` kernel launch on
 * this session in microseconds (needs SK_FLAG_TIMING), for roofline accounting. */
int sk_last_kernel_us(sk_session *s, float *us);

/* Measurement helpers: record a CUDA event (which = 0 start, 1 stop) on the
 * session's compute stream, and read the device time between them.  Used by
 * bench.py to time K steps on the stream the kernels run on.  (The reference
 * has no timing hooks for this stage; its only knob is GST_TRACERS.) */
int sk_marker_record(sk_session *s, int which);
int sk_marker_elapsed_us(sk_session *s, float *us);

/* Measurement helper: the bare PCIe rate of this process's GPU `device_id`, with nothing of the
 * session machinery involved — `iters` back-to-back cudaMemcpyAsync of `bytes` between eight pinned
 * host buffers used in turn (so the host's cache cannot stand in for its memory) and one device
 * buffer, timed with CUDA events.  mode 0: host->device only,
 * 1: device->host only, 2: both directions at once on two streams.  The rates (GB/s, 0 when the
 * direction was not exercised) are what the frame path's copies are judged against
 * (north_star: "the H2D copy is reported separately against PCIe bandwidth"; the reference has
 * no such probe).  Allocates and frees its own buffers: not for the frame path. */
int sk_pcie_probe(int device_id, size_t bytes, int iters, int mode, double *h2d_gbs, double *d2h_gbs);

/*
 * Convert (and shrink) ONE frame on buffers the caller owns, without allocating anything.
 * This entry point has no counterpart in the reference.
 *
 * It exists for callers that must not allocate GPU memory when they need a frame converted:
 * a host that is already streaming many desktop sessions on the GPU (cudaMalloc/cudaFree
 * stall the whole device, so every running session would drop a frame), or a host whose
 * GPU memory belongs to somebody else's pool (an encoder's buffer pool, a framework's
 * caching allocator) and is fully handed out.  No cudaMalloc, cudaFree or other allocating
 * call is made here, on any path.
 *
 * The kernel is the library's generic scale kernel (every output pixel gathers its own taps
 * through L1; equal sizes are allowed and then it simply converts) — NOT the shared-memory tiled
 * kernel a session normally runs, which needs per-session set-up.  Output is bit-identical to a
 * session's; it is 1.6-2.7x slower per frame (profiles/r2_noalloc_timing.md).  Its scratch — the
 * two reduction words of the damage count and one vertical tap entry per output row — lives in a
 * fixed array inside the library image, about 26 KB on every GPU that loads the library, for the
 * life of the process.  Hence the limit: out_height <= SK_NOALLOC_MAX_OUT_HEIGHT, anything taller
 * is refused with SK_ERR_ARG and a message in errbuf (use a session for it).  Enlarging (more
 * output than input pixels) is refused with SK_ERR_UNSUPPORTED here; sessions do it.
 *
 *   cfg        geometry, formats, matrix, chroma_site, scale_variant, SK_FLAG_NO_DAMAGE as
 *              for sk_session_create; device_id < 0 means the calling thread's current
 *              device, otherwise that device is made current.  n_inflight is ignored.
 *   d_in       device pointer to the BGRx/RGBx frame, rows in_stride bytes apart (0 = 4*width)
 *   d_out      device pointer to frame_bytes of output (GstVideoInfo default layout)
 *   d_prev     device pointer to the previous output frame, or NULL (map = all ones)
 *   d_map      device pointer to ceil(out_h/16)*ceil(out_w/16) bytes
 *   d_count    device pointer to one uint32 (the damaged-block count)
 *   cuda_stream  a cudaStream_t; the work is queued there and the call returns at once
 *   errbuf     optional buffer for a human-readable error, errbuf_len bytes
 * All device pointers and in_stride must be 4-byte aligned.
 *
 * The first call on each GPU looks up where that GPU keeps the array (cudaGetSymbolAddress) and
 * loads all eight instantiations of the kernel into that GPU's context, so that later calls with
 * another format / chroma site do not trigger a lazy code load; make that first call before
 * starting any stream capture.  When in_height != out_height the tap entries are uploaded from a
 * static host table that every call rewrites.
 * All calls share the one array, which still holds what the previous call left there;
 * each call re-zeroes the reduction words and rewrites the tap entries on cuda_stream
 * before its kernel.  Two calls whose GPU work can overlap (different streams, or
 * different host threads) therefore race on the array: serialise them.
 */
#define SK_NOALLOC_MAX_OUT_HEIGHT 4320
int sk_scale_frame_noalloc(const sk_config *cfg, const void *d_in, size_t in_stride,
                           void *d_out, const void *d_prev, uint8_t *d_map, uint32_t *d_count,
                           void *cuda_stream, char *errbuf, size_t errbuf_len);

/* Number of kernel launches issued by this library in this process so far. */
uint64_t sk_launch_count(void);

/* Human-readable description of the last error on this session (or of the
 * last failed sk_session_create when s == NULL).  Owned by the library. */
const char *sk_last_error(sk_session *s);

/* SK_ABI_VERSION of the loaded library. */
int sk_version(void);

/* Number of CUDA devices visible (0 on a CPU-only host; never fails). */
int sk_device_count(void);

#ifdef __cplusplus
}
#endif
#endif /* SELKIES_PREENCODE_H */
