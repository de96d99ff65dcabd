/* vepb200.h — C ABI of the B200 frame path (libvepb200.so).
 *
 * Drop-in boundary for ONE path of tangtang888/video-edge-ai-proxy: the conversion of a decoded
 * YUV 4:2:0 frame to packed BGR24 and its packing into the `VideoFrame` message, i.e. reference
 * python/read_image.py:94-119.  The reference has no FFI seam there (it calls PyAV, which calls
 * libswscale); each entry point below cites the reference lines whose work it takes over.
 * Everything is plain C: pointers, sizes, ints.  No torch / CUDA types appear in a signature
 * (CUDA streams travel as `void*`).
 *
 * Conventions: every function returns 0 on success or a negative vep_status; the text of the
 * last error of the calling thread is available from vep_last_error().  Nothing here calls
 * abort()/exit().  There is no CPU fallback: without a usable CUDA device vep_create fails.
 */
#ifndef VEPB200_H
#define VEPB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VEP_ABI_VERSION 1
#define VEP_MAX_TICKS_IN_FLIGHT 4

typedef enum {
    VEP_OK = 0,
    VEP_ERR_INVALID_ARG = -1,   /* null pointer, bad id, bad enum */
    VEP_ERR_GEOMETRY = -2,      /* odd width/height, stride < row bytes, nv12 with height < 8 */
    VEP_ERR_CUDA = -3,          /* a CUDA runtime call failed; see vep_last_error() */
    VEP_ERR_NOMEM = -4,
    VEP_ERR_STATE = -5,         /* call out of order (e.g. result before any tick) */
    VEP_ERR_NO_FRAME = -6,      /* nothing converted yet for this stream */
    VEP_ERR_BUSY = -7           /* too many ticks in flight / slot still in use */
} vep_status;

/* Pixel formats a decoder hands over (PyAV `frame.format.name` at read_image.py:94). */
typedef enum {
    VEP_FMT_YUV420P = 0,   /* planar, limited (MPEG) range  — swscale coefficient set 1 */
    VEP_FMT_YUVJ420P = 1,  /* planar, full (JPEG) range     — swscale coefficient set 2 */
    VEP_FMT_NV12 = 2,      /* Y plane + interleaved UV, limited range */
    VEP_FMT_NV21 = 3,      /* as NV12 with V first in each chroma pair */
    VEP_FMT_YUV422P = 4,   /* planar 4:2:2, limited range: chroma planes are (W/2) x H */
    VEP_FMT_YUVJ422P = 5,  /* planar 4:2:2, full range (what MJPEG cameras decode to) */
    VEP_FMT_YUV420P10LE = 6, /* planar 4:2:0, 10-bit samples (<= 1023) in little-endian 16-bit containers, limited range
                               (HEVC Main10); strides in BYTES; height >= 8 */
    VEP_FMT_YUV444P = 7,   /* planar 4:4:4, limited range: chroma planes are W x H */
    VEP_FMT_YUVJ444P = 8,  /* planar 4:4:4, full range */
    VEP_FMT_GRAY8 = 9      /* luma only (monochrome decode): plane[0] only, B = G = R = Y as libswscale does */
} vep_pixfmt;

/* One frame of a batched conversion, all pointers in DEVICE memory.
 * Takes the place of one `frame.to_ndarray(format='bgr24')` call (read_image.py:94):
 * plane[]/stride[] are the AVFrame planes and linesizes, dst is the (H, W, 3) uint8 result
 * (dst_stride = 3*width for the contiguous array PyAV returns after its padding strip).
 * Any even geometry and any alignment is accepted and gives the same bytes.  Fastest: 16-byte aligned plane and dst
 * pointers and strides that are multiples of 16 (planar 8-bit chroma: 8), any even width (what an AVFrame's padded
 * linesizes give, and what the host pipeline stages).  8-bit 4:2:0 / 4:2:2 / NV12 / NV21 frames of any other alignment
 * (tight rows of a width like 854, odd offsets) keep the same tiling with loads and stores assembled from the aligned
 * 16-byte chunks around them (about half the speed); 4:4:4, gray and 10-bit frames that are not aligned take a scalar
 * path.  Loads are always whole aligned 16-byte (8-byte) chunks that contain at least one visible sample of the row, so
 * up to 15 bytes beside a row may be READ but never a byte of an unmapped page; nothing outside the visible 3*width
 * bytes of an output row is ever written. */
typedef struct {
    const uint8_t *plane[3]; /* Y,U,V  (nv12/nv21: Y,UV,NULL) */
    int32_t stride[3];       /* bytes per row of each plane */
    uint8_t *dst;            /* B,G,R interleaved */
    int32_t dst_stride;      /* bytes per output row, >= 3*width */
    int32_t width, height;   /* both even; nv12/nv21 need height >= 8 */
    int32_t fmt;             /* vep_pixfmt */
    int32_t convert;         /* 0: the decode gate skipped this frame (read_image.py:80), no work */
} vep_frame_desc;

/* Per-frame scalars copied into the VideoFrame message (read_image.py:101-111). */
typedef struct {
    int64_t timestamp;   /* field 4: frame.time*time_base.denominator, else wall-clock ms (:89-91) */
    int64_t pts;         /* field 6 */
    int64_t dts;         /* field 7 */
    int64_t packet;      /* field 13: index of the packet inside its GOP (:107) */
    int64_t keyframe;    /* field 14: running GOP counter (:108) */
    double time_base;    /* field 10 */
    int32_t is_keyframe; /* field 5 — of the packet last popped from the queue (:110) */
    int32_t is_corrupt;  /* field 9 (:111) */
    char frame_type[8];  /* field 8: pict_type.name, NUL-terminated ("I","P","B",…) */
} vep_frame_meta;

typedef struct vep_ctx vep_ctx;   /* one per GPU; owns streams, staging, device buffers */
typedef struct vep_plan vep_plan; /* a validated, device-resident batch of frame descriptors */

/* ---- library ------------------------------------------------------------------------------ */
int vep_abi_version(void);
const char *vep_last_error(void);
const char *vep_status_name(int status);

/* ---- context (replaces "one container per camera", rtsp_process_manager.go:106-111, with one
 *      context per GPU serving many camera streams) ------------------------------------------- */
int vep_create(int device, vep_ctx **out);
void vep_destroy(vep_ctx *ctx);

/* ---- device-resident batched conversion (read_image.py:94 for N frames at once) -------------
 * Descriptors hold device pointers (e.g. torch tensors' data_ptr()).  One kernel launch converts
 * every descriptor with convert != 0.  `cuda_stream` is a cudaStream_t (NULL = default stream).
 * vep_convert_batch validates, uploads the table and launches; a plan does the first two once
 * so a fixed camera set costs one launch per tick. */
int vep_convert_batch(vep_ctx *ctx, const vep_frame_desc *descs, int n, void *cuda_stream);
/* The same with an explicit launch order (VEP_ORDER_*, below); vep_convert_batch is VEP_ORDER_STRICT. */
int vep_convert_batch_ordered(vep_ctx *ctx, const vep_frame_desc *descs, int n, void *cuda_stream, int order);
int vep_plan_create(vep_ctx *ctx, const vep_frame_desc *descs, int n, vep_plan **out);
int vep_plan_launch(vep_plan *plan, void *cuda_stream);
/* Same launch, but the caller asserts that it neither reads nor overwrites anything the PREVIOUS
 * kernel launch on this stream writes (e.g. consecutive ticks over disjoint buffer sets).  The GPU may
 * then start this launch while the previous one is still draining; ordering against everything
 * else on the stream (copies, events, other kernels) is unchanged.  Results are identical. */
int vep_plan_launch_overlapped(vep_plan *plan, void *cuda_stream);
/* The general form.  `order` says what this launch must be ordered against with respect to the PREVIOUS KERNEL
 * on the stream (copies, events and everything else keep ordinary stream order in all three):
 *   VEP_ORDER_STRICT  everything (== vep_plan_launch): safe whatever the previous kernel did;
 *   VEP_ORDER_WRITES  only this launch's stores wait for the previous kernel: valid when that kernel did not
 *                     PRODUCE this launch's input planes (inputs that came from cudaMemcpy or from older work are
 *                     fine) — consecutive ticks may then re-use one output buffer and the last one still wins;
 *   VEP_ORDER_NONE    nothing (== vep_plan_launch_overlapped). */
#define VEP_ORDER_STRICT 0
#define VEP_ORDER_WRITES 1
#define VEP_ORDER_NONE 2
int vep_plan_launch_ordered(vep_plan *plan, void *cuda_stream, int order);
/* plans[0..n) launched back to back on one stream with one call (all of one device): what a native producer
 * converting tick after tick does; saves the per-call overhead of n separate vep_plan_launch_ordered calls.
 * event_before / event_after: optional cudaEvent_t (NULL: none) recorded on the stream directly before the first
 * and directly after the last launch, for timing exactly these launches without host latency in between. */
int vep_plan_launch_sequence(vep_plan *const *plans, int n, void *cuda_stream, int order, void *event_before, void *event_after);
/* frames with convert!=0, CTAs per launch, algorithmic bytes (read + written: 4.5*W*H per converted 8-bit 4:2:0
 * frame, 5 for 4:2:2, 6 for 4:4:4 and 10-bit, 4 for gray) */
int vep_plan_info(const vep_plan *plan, int *n_converted, int *n_ctas, int64_t *alg_bytes);
void vep_plan_destroy(vep_plan *plan);

/* No counterpart in the reference (it never touches a GPU): the same batched conversion for callers
 * that are not allowed to allocate device memory — a process on a GPU slice or in a container whose
 * memory was handed out in full at start-up, a plugin inside a host application whose own pool already
 * owns almost all of HBM (a raw cudaMalloc can fail there although the pool has room), or a
 * real-time loop that cannot take cudaMalloc's device-wide synchronisation mid-stream.  Needs no
 * vep_ctx; works on the CURRENT device and on the caller's stream; calls no cudaMalloc/cudaFree or
 * any other allocating routine.  Its descriptor tables live in a fixed array inside the library's
 * CUDA module (545,024 bytes on every GPU that loads the library, until the process exits), hence the
 * limits: at most VEP_NOALLOC_MAX_FRAMES converted frames and VEP_NOALLOC_MAX_CTAS thread blocks per
 * call.  On the vector path (8-bit 4:2:0 planar / NV12 / NV21 with aligned planes) one block covers
 * 8 x 512 px x 2 rows: a 1080p frame takes 270 blocks, 4K takes 1080; every other frame (4:2:2, 4:4:4,
 * gray, 10-bit, unaligned planes) runs the scalar path at 1024 2x2 blocks per thread block: 507 for
 * 1080p.  Larger batches are refused with VEP_ERR_INVALID_ARG and the text says which limit — split them.
 * The first call on each device also resolves the array's address on that device; make that first
 * call before any stream capture is begun.  All calls share the one array per device and each call
 * rewrites it: calls whose GPU work can overlap (different streams, different threads) race and
 * must be serialised by the caller; calls queued on one stream are safe.  The tables are uploaded from
 * ordinary (pageable) host memory inside the library, so a call may hold the calling thread until work
 * queued earlier on the same stream has drained — keep that in mind in a loop with a deadline.  Outputs
 * are byte-identical to vep_convert_batch. */
#define VEP_NOALLOC_MAX_FRAMES 256
#define VEP_NOALLOC_MAX_CTAS 65536
int vep_convert_batch_noalloc(const vep_frame_desc *descs, int n, void *cuda_stream);

/* ---- host pipeline: camera streams, pinned staging, tick (read_image.py:87-121) --------------
 * stream_open declares a camera (device_id in the reference) with a fixed decoded geometry.
 * `ring_depth` mirrors --memory_buffer / XADD MAXLEN (read_image.py:121): a result stays valid
 * until `ring_depth` newer frames of that stream have been converted (ring_depth + VEP_MAX_TICKS_IN_FLIGHT + 2
 * slots are kept per stream, each with pinned input, pinned output and device buffers). */
int vep_stream_open(vep_ctx *ctx, int stream_id, int fmt, int width, int height, int ring_depth);
/* Several cameras at once, their slots sharing one pinned and one device allocation per direction, laid out
 * [slot index][camera in the order given].  Everything else is as if each had been opened alone; the difference is
 * what a tick costs: cameras that are neighbours in `specs` and stage their frames in lock step (one frame each per
 * tick, the usual fleet) are copied in and out with ONE cudaMemcpyAsync per direction for the whole run instead of one
 * per camera (about +12 % end to end on one B200 for 16 x 1080p).  Cameras that skip a tick simply split the run.
 * A member is closed with vep_stream_close; the memory is released with the last member. */
typedef struct { int32_t stream_id, fmt, width, height; } vep_stream_spec;
int vep_stream_open_many(vep_ctx *ctx, const vep_stream_spec *specs, int n, int ring_depth);
/* The most frames one such merged copy may carry (0, the default: no limit; 1: one copy per frame as for cameras
 * opened singly).  A run longer than the limit is cut into several copies, each on its first camera's streams. */
int vep_set_copy_run_limit(vep_ctx *ctx, int max_frames_per_copy);
/* Page-lock / release a range of host memory the caller owns (cudaHostRegister, usable from every GPU), e.g. the
 * arena a decoder's get_buffer callback hands frames out of: once at start-up, not per frame (registering costs about
 * a millisecond per 10 MB).  Unregister before freeing the memory. */
int vep_host_register(void *ptr, size_t bytes);
int vep_host_unregister(void *ptr);
/* Not while ticks are in flight (VEP_ERR_BUSY) and not while another thread is inside a call for this stream. */
int vep_stream_close(vep_ctx *ctx, int stream_id);
/* Helper threads for the staging copy (default 0).  One core copies 10-16 GB/s into pinned memory, a PCIe Gen5
 * link takes 55.  vep_submit: frames of 4 MB and more (4K) are cut into pieces copied by the submitting thread and
 * the helpers; smaller frames are copied by the submitting thread alone (one producer thread per camera already
 * covers them, as the reference's one process per camera does).  vep_submit_batch: all pieces of all frames.
 * Change the count only while no vep_submit / vep_submit_batch call is in progress. */
int vep_set_staging_threads(vep_ctx *ctx, int n);

/* Hand one decoded frame (HOST pointers: AVFrame planes + linesizes) to the library.  `convert` is
 *   VEP_SUBMIT_SKIP   (0) a gate-skipped frame: counted, never copied (read_image.py:80);
 *   VEP_SUBMIT_COPY   (1) the planes are copied into pinned staging before the call returns, so the caller's frame
 *                         may be freed at once (what PyAV's pageable frames need);
 *   VEP_SUBMIT_PINNED (2) no CPU copy: the planes already lie in page-locked host memory (a decoder buffer pool
 *                         registered once with vep_host_register, or any cudaHostAlloc'ed / cudaHostRegister'ed memory) and
 *                         the tick's H2D copy reads them in place, any linesize.  The caller keeps them valid and unchanged
 *                         until the tick that takes the frame has been waited for (vep_tick / vep_tick_wait*), i.e. it
 *                         holds its AVFrame reference that long.  Refused with VEP_ERR_INVALID_ARG when a plane is not
 *                         page-locked.  On a host whose memory system is shared by several GPUs' DMA traffic this is the
 *                         difference between ~0.77 and ~0.98 of the box's DMA ceiling.
 * Callable concurrently for different streams; one producer per stream. */
#define VEP_SUBMIT_SKIP 0
#define VEP_SUBMIT_COPY 1
#define VEP_SUBMIT_PINNED 2
int vep_submit(vep_ctx *ctx, int stream_id, const uint8_t *const plane[3], const int32_t stride[3],
               const vep_frame_meta *meta, int convert);
/* The frames of one tick with ONE call (one item per camera, at most one frame per stream and call): the staging
 * copies of all items are cut into ~1 MB pieces and shared between the calling thread and the context's helper
 * threads (vep_set_staging_threads), so a producer that gathers many cameras pays one call per tick and the copy
 * runs on several cores.  Items with convert == 0 are counted like in vep_submit.  Stops at the first item that
 * cannot be staged (its status is returned, *n_done = items staged and committed before it). */
typedef struct {
    int32_t stream_id;
    int32_t convert;           /* VEP_SUBMIT_SKIP / _COPY / _PINNED, as for vep_submit */
    const uint8_t *plane[3];   /* HOST pointers, as for vep_submit */
    int32_t stride[3];
    int32_t reserved;
    vep_frame_meta meta;
} vep_submit_item;
int vep_submit_batch(vep_ctx *ctx, const vep_submit_item *items, int n, int *n_done);
/* Zero-copy variant: get pointers into the next pinned staging slot (a decoder can write its
 * planes there directly), then commit it.  stride_out is the row pitch of each plane in the slot: the row bytes
 * when they are a multiple of 16, else rounded up to 64. */
int vep_acquire_input(vep_ctx *ctx, int stream_id, uint8_t *plane_out[3], int32_t stride_out[3]);
int vep_commit_input(vep_ctx *ctx, int stream_id, const vep_frame_meta *meta);

/* One tick: for every stream with a committed frame — H2D on that camera's CUDA stream, ONE
 * batched kernel launch over all of them, D2H into the pinned output ring.  vep_tick is
 * begin+wait.  Up to VEP_MAX_TICKS_IN_FLIGHT ticks may be in flight (begin, begin, wait, …) so the
 * staging copy and H2D of later ticks overlap the D2H of earlier ones (PCIe is full duplex). */
int vep_tick_begin(vep_ctx *ctx, int *n_frames);
int vep_tick_wait(vep_ctx *ctx);   /* waits for the OLDEST tick in flight */
int vep_tick(vep_ctx *ctx, int *n_frames);   /* begin + wait for THAT tick (and any older one) */
/* Ticketed form for several ticking threads or sparse ticks: begin hands out a ticket (ticks that found nothing
 * to convert return the ticket of the newest real tick, 0 if there never was one); wait_ticket returns once that
 * tick and all older ones have landed, whoever else has begun or waited for ticks meanwhile. */
int vep_tick_begin_ticket(vep_ctx *ctx, int *n_frames, uint64_t *ticket);
int vep_tick_wait_ticket(vep_ctx *ctx, uint64_t ticket);

/* Newest converted frame of a stream: pointer to H*W*3 contiguous BGR bytes in pinned host
 * memory (what `np.ndarray.tobytes(img)` yields at read_image.py:97) and shape = {H, W, 3}
 * (img.shape, :95). */
int vep_result(vep_ctx *ctx, int stream_id, const uint8_t **bgr, size_t *len, int32_t shape[3]);
/* Same frame as the complete serialized VideoFrame (vf.SerializeToString(), read_image.py:119;
 * schema proto/video_streaming.proto:78-93) — header and trailer are written around the payload
 * in the same pinned buffer, so the pointer can go to XADD (:121) without another copy. */
int vep_result_packed(vep_ctx *ctx, int stream_id, const uint8_t **msg, size_t *len);
/* Older results of the same stream: age 0 is the newest, age k the k-th before it (a catch-up
 * burst, read_image.py:81-85, converts several frames of one camera in one tick).  Guaranteed for
 * age < ring_depth; frames of the tick just waited for are all still there (up to ring_depth+3)
 * until the stream is fed again. */
int vep_result_packed_at(vep_ctx *ctx, int stream_id, int age, const uint8_t **msg, size_t *len);
int vep_result_at(vep_ctx *ctx, int stream_id, int age, const uint8_t **bgr, size_t *len, int32_t shape[3]);
/* Device-side view of the same result for GPU consumers (zero-copy; valid like vep_result).  On the device a row
 * occupies *row_pitch bytes: 3*width when that is a multiple of 16 (every common resolution), else 3*width rounded
 * up to 64 (the copy to the host removes the padding).  vep_result_device serves the unpadded case only and returns
 * VEP_ERR_GEOMETRY otherwise. */
int vep_result_device_pitched(vep_ctx *ctx, int stream_id, const uint8_t **bgr_dev, int32_t shape[3], int32_t *row_pitch);
int vep_result_device(vep_ctx *ctx, int stream_id, const uint8_t **bgr_dev, int32_t shape[3]);

/* ---- VideoFrame wire framing alone (read_image.py:99-119 minus the payload copy) -------------
 * Writes the bytes that precede and follow `data` in the serialized message.  prefix needs
 * >= 32 bytes, suffix >= 192 bytes.  shape is {height, width, 3}. */
int vep_videoframe_frame(const vep_frame_meta *meta, int32_t width, int32_t height,
                         uint8_t *prefix, int32_t *prefix_len, uint8_t *suffix,
                         int32_t *suffix_len);

/* ---- counters (the reference only prints) -------------------------------------- */
typedef struct {
    int64_t frames_offered;   /* submits incl. gate-skipped */
    int64_t frames_converted;
    int64_t ticks;
    int64_t kernel_launches;
    int64_t h2d_bytes, d2h_bytes;
    float last_kernel_ms;     /* device time of the most recent tick's kernel (CUDA events) */
    int32_t reserved;
    int64_t copies_in, copies_out;   /* H2D / D2H copy operations issued by ticks (one per run of neighbouring slots) */
} vep_stats;
int vep_get_stats(vep_ctx *ctx, vep_stats *out);

#ifdef __cplusplus
}
#endif
#endif /* VEPB200_H */
