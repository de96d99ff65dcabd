// vep_abi.cu — host side of libvepb200.so: the C ABI declared in include/vepb200.h.
//
// Replaces, for many cameras at once, the inner loop of reference python/read_image.py:87-121:
// pinned staging of decoded planes, H2D on per-camera CUDA streams, ONE batched conversion
// launch per tick (vep_kernels.cuh), D2H into a pinned output ring whose slots already carry
// the VideoFrame wire framing (proto/video_streaming.proto:78-93) around the BGR payload.
// No CPU conversion path exists in this library.
#include "vep_kernels.cuh"
#include "../../include/vepb200.h"

#include <immintrin.h>

#include <algorithm>
#include <atomic>
#include <cstdarg>
#include <cstddef>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <condition_variable>
#include <deque>
#include <functional>
#include <mutex>
#include <new>
#include <shared_mutex>
#include <thread>
#include <unordered_map>
#include <vector>

static_assert(sizeof(vep_frame_desc) == 72 && offsetof(vep_frame_desc, dst) == 40 && offsetof(vep_frame_desc, width) == 52,
              "vep_frame_desc layout is part of the ABI (mirrored in _abi.py)");
static_assert(sizeof(vep_frame_meta) == 64 && offsetof(vep_frame_meta, frame_type) == 56, "vep_frame_meta layout is part of the ABI");
static_assert(sizeof(vep_submit_item) == 112 && offsetof(vep_submit_item, plane) == 8 && offsetof(vep_submit_item, meta) == 48, "vep_submit_item layout is part of the ABI");
static_assert(sizeof(vep::FrameDev) == 80 && offsetof(vep::FrameDev, w) == 48 && offsetof(vep::FrameDev, item_begin) == 64 &&
              sizeof(vep::CtaWork) == 8, "device table layout (the kernels read 16-byte groups of FrameDev)");

using vep::CtaWork;
using vep::FrameDev;

// ------------------------------------------------------------------------------------------------
// errors
// ------------------------------------------------------------------------------------------------
static thread_local char g_err[512] = "";

static int fail(int code, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
    return code;
}
#define CUDA_TRY(expr)                                                                       \
    do {                                                                                     \
        cudaError_t e_ = (expr);                                                             \
        if (e_ != cudaSuccess)                                                               \
            return fail(VEP_ERR_CUDA, "%s -> %s (%s:%d)", #expr, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

extern "C" int vep_abi_version(void) { return VEP_ABI_VERSION; }
extern "C" const char *vep_last_error(void) { return g_err; }
extern "C" const char *vep_status_name(int s) {
    switch (s) {
        case VEP_OK: return "VEP_OK";
        case VEP_ERR_INVALID_ARG: return "VEP_ERR_INVALID_ARG";
        case VEP_ERR_GEOMETRY: return "VEP_ERR_GEOMETRY";
        case VEP_ERR_CUDA: return "VEP_ERR_CUDA";
        case VEP_ERR_NOMEM: return "VEP_ERR_NOMEM";
        case VEP_ERR_STATE: return "VEP_ERR_STATE";
        case VEP_ERR_NO_FRAME: return "VEP_ERR_NO_FRAME";
        case VEP_ERR_BUSY: return "VEP_ERR_BUSY";
        default: return "VEP_ERR_UNKNOWN";
    }
}

// ------------------------------------------------------------------------------------------------
// descriptor validation -> device tables
// ------------------------------------------------------------------------------------------------
static constexpr int kMaxDim = 16384;

struct Tables {
    std::vector<FrameDev> frames;   // n_converted entries + 1 sentinel (item_begin = total_items)
    std::vector<CtaWork> work;      // per-CTA table of the one-shot kernel
    int64_t alg_bytes = 0;
    int n_converted = 0;
    int total_items = 0;            // work items over all frames (0: nothing to launch)
    std::vector<int32_t> tick_cta_begin;   // tick kernel: first CTA of each frame, then the grid size
};

static inline bool aligned(const void *p, size_t a) { return (reinterpret_cast<uintptr_t>(p) % a) == 0; }

static int build_tables(const vep_frame_desc *d, int n, Tables &t) {
    constexpr int rpt = vep::kTickRpt;
    if (n < 0 || (n > 0 && !d)) return fail(VEP_ERR_INVALID_ARG, "descs is null or n < 0");
    t.frames.clear();
    t.work.clear();
    t.alg_bytes = 0;
    t.n_converted = 0;
    t.total_items = 0;
    t.tick_cta_begin.clear();
    int tick_ctas = 0;
    // the 4:4:4 vector path exists only in the format-specialised tick kernels: usable when the launch is single-format
    int only_fmt = -2;
    for (int i = 0; i < n; ++i) if (d[i].convert) only_fmt = (only_fmt == -2 || only_fmt == d[i].fmt) ? d[i].fmt : -1;
    const bool homogeneous = only_fmt >= 0;
    for (int i = 0; i < n; ++i) {
        const vep_frame_desc &s = d[i];
        if (!s.convert) continue;
        if (s.fmt < VEP_FMT_YUV420P || s.fmt > VEP_FMT_GRAY8) return fail(VEP_ERR_INVALID_ARG, "desc %d: unknown fmt %d", i, s.fmt);
        const int w = s.width, h = s.height;
        if (w <= 0 || h <= 0 || w > kMaxDim || h > kMaxDim || (w & 1) || (h & 1))
            return fail(VEP_ERR_GEOMETRY, "desc %d: %dx%d — 4:2:0 needs even, positive dimensions <= %d", i, w, h, kMaxDim);
        const bool nv12 = vep::fmt_is_semiplanar(s.fmt);
        const int bps = vep::fmt_bytes_per_sample(s.fmt);
        if ((nv12 || bps == 2) && h < 8) return fail(VEP_ERR_GEOMETRY, "desc %d: nv12/nv21/10-bit height %d < 8 (swscale changes its vertical chroma filter there)", i, h);
        const bool gray = s.fmt == VEP_FMT_GRAY8;
        if (!s.plane[0] || (!gray && (!s.plane[1] || (!nv12 && !s.plane[2]))) || !s.dst) return fail(VEP_ERR_INVALID_ARG, "desc %d: null plane or dst", i);
        const int cw = ((nv12 || vep::fmt_is_444(s.fmt)) ? w : w / 2) * bps;
        if (s.stride[0] < w * bps || (!gray && (s.stride[1] < cw || (!nv12 && s.stride[2] < cw))) || s.dst_stride < 3 * w)
            return fail(VEP_ERR_GEOMETRY, "desc %d: a stride is smaller than its row", i);
        if (bps == 2 && ((s.stride[0] | s.stride[1] | s.stride[2]) & 1 || !aligned(s.plane[0], 2) || !aligned(s.plane[1], 2) || !aligned(s.plane[2], 2)))
            return fail(VEP_ERR_GEOMETRY, "desc %d: 16-bit samples need 2-byte aligned planes and even strides", i);
        FrameDev f{};
        f.y = s.plane[0]; f.c0 = gray ? s.plane[0] : s.plane[1]; f.c1 = (nv12 || gray) ? nullptr : s.plane[2]; f.dst = s.dst;
        f.sy = s.stride[0]; f.sc0 = gray ? s.stride[0] : s.stride[1]; f.sc1 = (nv12 || gray) ? 0 : s.stride[2]; f.sd = s.dst_stride;
        f.w = w; f.h = h; f.fmt = s.fmt;
        // Vector path: every plane and the output 16-byte aligned (planar 8-bit chroma: 8) with strides that are
        // multiples of the same.  Any even width qualifies for 8-bit samples: a stride that is a multiple of 16 and
        // >= the row already contains the row rounded up to 16 bytes, which is all the last lane of a row reads.
        const size_t ca = (nv12 || bps == 2 || vep::fmt_is_444(s.fmt)) ? 16 : 8;
        const bool vec = (bps == 1 || w % 16 == 0) && aligned(f.y, 16) && f.sy % 16 == 0 && aligned(f.c0, ca) && f.sc0 % ca == 0 &&
                         (nv12 || gray || (aligned(f.c1, ca) && f.sc1 % ca == 0)) && aligned(f.dst, 16) && f.sd % 16 == 0;
        // 8-bit 4:2:0 / 4:2:2 / NV12 / NV21 keep the vector tiling whatever the alignment (tight rows of an odd width, odd
        // offsets): PATH_UNALIGNED assembles its loads and stores from the aligned chunks around them.  The other formats
        // fall back to the scalar path when they are not aligned.
        const bool any_align = bps == 1 && !gray && !vep::fmt_is_444(s.fmt);
        f.fast = vec ? vep::PATH_VECTOR : any_align ? vep::PATH_UNALIGNED : vep::PATH_GENERIC;
        if (vep::fmt_vector_only_when_specialised(s.fmt))
            f.fast = (f.fast != vep::PATH_GENERIC && homogeneous) ? vep::PATH_VECTOR : vep::PATH_GENERIC;
        // only the tick kernel has a vector path for 4:2:2, 4:4:4, gray and for 10-bit samples
        f.fast_tbl = (vep::fmt_is_422(s.fmt) || vep::fmt_vector_only_when_specialised(s.fmt) || bps == 2) ? vep::PATH_GENERIC : f.fast;
        f.segs = (w + vep::kSegPx - 1) / vep::kSegPx;
        f.segs_magic = f.segs == 1 ? 0u : (uint32_t)((((uint64_t)1 << 32) + f.segs - 1) / f.segs);
        f.item_begin = t.total_items;
        const int fi = (int)t.frames.size();
        t.frames.push_back(f);
        int items;
        t.tick_cta_begin.push_back(tick_ctas);
        if (f.fast != vep::PATH_GENERIC) tick_ctas += (f.segs * ((h / 2 + rpt - 1) / rpt) + vep::kWarpsPerCta - 1) / vep::kWarpsPerCta;
        else tick_ctas += ((w / 2) * (h / 2) + vep::kGenericBlocksPerCta - 1) / vep::kGenericBlocksPerCta;
        if (f.fast_tbl != vep::PATH_GENERIC) {
            items = f.segs * (h / 2);
            for (int it = 0; it < items; it += vep::kWarpsPerCta) t.work.push_back(CtaWork{fi, it});
        } else {
            const int blocks = (w / 2) * (h / 2);
            items = (blocks + vep::kGenericBlocksPerCta - 1) / vep::kGenericBlocksPerCta;
            for (int b = 0; b < blocks; b += vep::kGenericBlocksPerCta) t.work.push_back(CtaWork{fi, b});
        }
        if ((int64_t)t.total_items + items > (int64_t)1 << 30) return fail(VEP_ERR_INVALID_ARG, "batch too large");
        t.total_items += items;
        t.alg_bytes += gray ? (int64_t)w * h * 4 : vep::fmt_is_422(s.fmt) ? (int64_t)w * h * 5 : (bps == 2 || vep::fmt_is_444(s.fmt)) ? (int64_t)w * h * 6
                                              : (int64_t)w * h * 9 / 2;   // read: gray 1, 4:2:0 1.5, 4:2:2 2, 10-bit / 4:4:4 3 B/px
        t.n_converted++;
    }
    FrameDev sentinel{};
    sentinel.item_begin = t.total_items;
    t.frames.push_back(sentinel);
    t.tick_cta_begin.push_back(tick_ctas);
    return VEP_OK;
}

// ------------------------------------------------------------------------------------------------
// context
// ------------------------------------------------------------------------------------------------
struct TableBuf {           // one upload buffer of the per-call path
    void *h = nullptr;      // pinned
    void *d = nullptr;      // device
    size_t cap = 0;
    cudaEvent_t used = nullptr;
    bool pending = false;
};

enum SlotState { SLOT_FREE = 0, SLOT_FILLING, SLOT_SUBMITTED, SLOT_INFLIGHT, SLOT_READY };

static constexpr size_t kHdrRoom = 64;      // wire prefix is written right-aligned in front of the payload
static constexpr size_t kTrailerRoom = 192;

struct Slot {
    uint8_t *h_in = nullptr, *h_out = nullptr, *d_in = nullptr, *d_out = nullptr;
    vep_frame_meta meta{};
    int state = SLOT_FREE;
    size_t msg_off = 0, msg_len = 0;
    cudaEvent_t done = nullptr;
    cudaEvent_t done_ref = nullptr;   // the event that covers this slot's copy out (its own, or that of the merged copy it rode in)
    // VEP_SUBMIT_PINNED: the frame stays in the caller's page-locked planes until the tick copies it to the device
    bool external = false;
    const uint8_t *ext_plane[3] = {nullptr, nullptr, nullptr};
    int32_t ext_stride[3] = {0, 0, 0};
};

struct StreamGroup {
    uint8_t *h_in = nullptr, *h_out = nullptr, *d_in = nullptr, *d_out = nullptr;   // [slot index][member] each
    size_t in_stride = 0, out_stride = 0;                                           // bytes per slot index
    std::atomic<int> refs{0};
};

struct Stream {
    int id = 0, fmt = 0, w = 0, h = 0, ring = 1;
    // staged planes (pinned and device alike) have rows padded to pitch_y / pitch_c so that every geometry meets the
    // vector path's alignment; the device output has pitch_out (>= 3w, multiple of 16), the pinned output is tight
    int pitch_y = 0, pitch_c = 0, pitch_out = 0;
    size_t off_c0 = 0, off_c1 = 0, in_bytes = 0, out_bytes = 0, dev_out_bytes = 0, in_slot = 0, out_slot = 0, dev_out_slot = 0;
    std::vector<Slot> slots;
    int wr = 0;             // next slot to fill
    int newest = -1;        // newest READY slot
    int n_ready = 0;
    StreamGroup *grp = nullptr;   // owner of the slot memory (streams opened together share one allocation per kind)
    int member = 0;               // position inside the group: slot k of member m+1 follows slot k of member m in memory
    bool padded_out = false;      // device output rows are padded (3w not a multiple of 16): own device buffer, 2-D copy out
    uint8_t *d_out_own = nullptr;
    cudaStream_t cs_in = nullptr, cs_out = nullptr;
    cudaEvent_t ev_in = nullptr;
    std::mutex mu;
};

struct SlotRef { Stream *s; int slot; };

struct TickRec {
    std::vector<SlotRef> slots;
    cudaEvent_t k0 = nullptr, k1 = nullptr;
    bool active = false;
    uint64_t ticket = 0;
};

// Helper threads for the staging copy of large frames (vep_submit): a 4K frame is 12-25 MB, one core copies
// ~10-16 GB/s, a PCIe link takes 55 -- so frames above kBandSplitBytes are copied as row bands by the submitting
// thread plus these workers.  They sleep on a condition variable otherwise.
struct StagePool {
    std::vector<std::thread> threads;
    std::mutex mu;
    std::condition_variable cv;
    std::deque<std::function<void()>> q;
    bool stop = false;
    void start(int n) {
        for (int i = 0; i < n; ++i)
            threads.emplace_back([this] {
                for (;;) {
                    std::function<void()> job;
                    {
                        std::unique_lock<std::mutex> lk(mu);
                        cv.wait(lk, [this] { return stop || !q.empty(); });
                        if (q.empty()) return;
                        job = std::move(q.front());
                        q.pop_front();
                    }
                    job();
                }
            });
    }
    void post(std::function<void()> job) {
        { std::lock_guard<std::mutex> lk(mu); q.push_back(std::move(job)); }
        cv.notify_one();
    }
    void shutdown() {
        { std::lock_guard<std::mutex> lk(mu); stop = true; }
        cv.notify_all();
        for (auto &t : threads) t.join();
        threads.clear();
    }
};
static constexpr size_t kBandSplitBytes = (size_t)4 << 20;

struct vep_ctx {
    int device = 0;
    cudaStream_t main = nullptr;
    TableBuf tb[4];
    int tb_next = 0;
    std::shared_mutex smu;                          // guards the streams map (shared: lookups, exclusive: open/close)
    std::mutex mu;                                  // guards ticks, table buffers, stats, scratch
    std::unordered_map<int, Stream *> streams;
    TickRec ticks[VEP_MAX_TICKS_IN_FLIGHT];
    int tick_head = 0, ticks_in_flight = 0;         // oldest in flight = tick_head
    uint64_t next_ticket = 1, retired_ticket = 0;   // tickets are issued in order and retire in order
    vep_stats stats{};
    Tables scratch;
    std::vector<uint8_t> tick_blob;
    StagePool pool;
    int stage_threads = 0;                          // helpers besides the submitting thread
    int run_limit = 0;                              // most frames one merged copy may carry (0: no limit)
};

struct vep_plan {
    vep_ctx *ctx = nullptr;
    FrameDev *d_frames = nullptr;
    CtaWork *d_work = nullptr;
    int n_ctas = 0, n_converted = 0, total_items = 0;
    std::vector<uint8_t> tick_blob, tick_blob_overlap, tick_blob_writes;   // TickParams<16|128>, wait_prior = 1 | 0 | 2, passed by value
    int tick_cap = 0, tick_grid = 0;
    int64_t alg_bytes = 0;
};

struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int dev) { cudaGetDevice(&prev); if (prev != dev) cudaSetDevice(dev); else prev = -1; }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

// ---- tick kernel: descriptor table travels as a kernel parameter -------------------------------------
static constexpr int kTickCapSmall = 16, kTickCapLarge = 128;
static constexpr int kOnlyFmtUnaligned = 0x100;   // flag in TickParams::only_fmt (host side only)

template <int NF>
static void fill_tick(const Tables &t, std::vector<uint8_t> &blob, int wait_prior) {
    blob.assign(sizeof(vep::TickParams<NF>), 0);
    auto *P = reinterpret_cast<vep::TickParams<NF> *>(blob.data());
    const int n = t.n_converted;
    P->n_frames = n;
    P->wait_prior = wait_prior;
    int per = n ? t.tick_cta_begin[1] - t.tick_cta_begin[0] : 0;
    for (int i = 1; i < n && per > 0; ++i) if (t.tick_cta_begin[i + 1] - t.tick_cta_begin[i] != per) per = 0;
    P->uniform_ctas = (n > 0 && n <= 65535) ? per : 0;
    int only = n ? t.frames[0].fmt : -1;
    for (int i = 1; i < n; ++i) if (t.frames[i].fmt != only) only = -1;
    bool any_unaligned = false;
    for (int i = 0; i < n; ++i) any_unaligned |= t.frames[i].fast == vep::PATH_UNALIGNED;
    // host-side hint: all frames share one pixel format -> format-specialised kernel; bit 8: some frame needs the
    // any-alignment loads and stores -> the flavour of that kernel which carries them
    P->only_fmt = only < 0 ? -1 : (only | (any_unaligned ? kOnlyFmtUnaligned : 0));
    for (int i = 0; i <= n; ++i) P->cta_begin[i] = t.tick_cta_begin[i];
    for (int i = 0; i < n; ++i) P->frames[i] = t.frames[i];
}
// TickParams::wait_prior for a launch order of the ABI
static inline int wait_prior_of(int order) { return order == VEP_ORDER_NONE ? 0 : order == VEP_ORDER_WRITES ? 2 : 1; }
// returns capacity used (0: does not fit -> caller falls back to the table-driven kernel)
static int make_tick_blob(const Tables &t, std::vector<uint8_t> &blob, int *grid, int order = VEP_ORDER_STRICT) {
    const int n = t.n_converted;
    *grid = n ? t.tick_cta_begin[n] : 0;
    if (n <= kTickCapSmall) { fill_tick<kTickCapSmall>(t, blob, wait_prior_of(order)); return kTickCapSmall; }
    if (n <= kTickCapLarge) { fill_tick<kTickCapLarge>(t, blob, wait_prior_of(order)); return kTickCapLarge; }
    return 0;
}
// single-format launches run the instantiation in which the other formats' code is compiled out
template <int NF>
static const void *tick_kernel(int only_fmt) {
    constexpr int R = vep::kTickRpt;
    if (only_fmt >= 0 && (only_fmt & kOnlyFmtUnaligned)) {
        switch (only_fmt & 0xff) {
            case VEP_FMT_YUV420P: return (const void *)vep::yuv_to_bgr24_tick<NF, R, 4, vep::FMT_YUV420P, true>;
            case VEP_FMT_YUVJ420P: return (const void *)vep::yuv_to_bgr24_tick<NF, R, 4, vep::FMT_YUVJ420P, true>;
            case VEP_FMT_NV12: return (const void *)vep::yuv_to_bgr24_tick<NF, R, 4, vep::FMT_NV12, true>;
            case VEP_FMT_NV21: return (const void *)vep::yuv_to_bgr24_tick<NF, R, 4, vep::FMT_NV21, true>;
            case VEP_FMT_YUV422P: return (const void *)vep::yuv_to_bgr24_tick<NF, R, 4, vep::FMT_YUV422P, true>;
            case VEP_FMT_YUVJ422P: return (const void *)vep::yuv_to_bgr24_tick<NF, R, 4, vep::FMT_YUVJ422P, true>;
            default: return (const void *)vep::yuv_to_bgr24_tick<NF, R, 4>;   // cannot happen: other formats are never PATH_UNALIGNED
        }
    }
    switch (only_fmt) {
        case VEP_FMT_YUV420P: return (const void *)vep::yuv_to_bgr24_tick<NF, R, 4, vep::FMT_YUV420P>;
        case VEP_FMT_YUVJ420P: return (const void *)vep::yuv_to_bgr24_tick<NF, R, 4, vep::FMT_YUVJ420P>;
        case VEP_FMT_NV12: return (const void *)vep::yuv_to_bgr24_tick<NF, R, 4, vep::FMT_NV12>;
        case VEP_FMT_NV21: return (const void *)vep::yuv_to_bgr24_tick<NF, R, 4, vep::FMT_NV21>;
        case VEP_FMT_YUV422P: return (const void *)vep::yuv_to_bgr24_tick<NF, R, 4, vep::FMT_YUV422P>;
        case VEP_FMT_YUVJ422P: return (const void *)vep::yuv_to_bgr24_tick<NF, R, 4, vep::FMT_YUVJ422P>;
        case VEP_FMT_YUV420P10LE: return (const void *)vep::yuv_to_bgr24_tick<NF, R, 3, vep::FMT_YUV420P10>;   // 80 registers: no spills
        case VEP_FMT_YUV444P: return (const void *)vep::yuv_to_bgr24_tick<NF, R, 3, vep::FMT_YUV444P>;
        case VEP_FMT_YUVJ444P: return (const void *)vep::yuv_to_bgr24_tick<NF, R, 3, vep::FMT_YUVJ444P>;
        case VEP_FMT_GRAY8: return (const void *)vep::yuv_to_bgr24_tick<NF, R, 4, vep::FMT_GRAY8>;
        default: return (const void *)vep::yuv_to_bgr24_tick<NF, R, 4>;
    }
}
static int launch_tick(int cap, const std::vector<uint8_t> &blob, int grid, cudaStream_t st) {
    if (grid <= 0) return VEP_OK;
    const int32_t *hdr = reinterpret_cast<const int32_t *>(blob.data());   // n_frames, wait_prior, uniform_ctas, only_fmt
    const void *fn = cap == kTickCapSmall ? tick_kernel<kTickCapSmall>(hdr[3]) : tick_kernel<kTickCapLarge>(hdr[3]);
    void *args[1] = {const_cast<uint8_t *>(blob.data())};
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = hdr[2] > 0 ? dim3(hdr[2], hdr[0]) : dim3(grid);
    cfg.blockDim = dim3(vep::kThreads);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;   // programmatic dependent launch (see the kernel)
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    CUDA_TRY(cudaLaunchKernelExC(&cfg, fn, args));
    return VEP_OK;
}

static int launch_tables(const FrameDev *d_frames, const CtaWork *d_work, int n_ctas, int total_items, cudaStream_t st) {
    if (total_items <= 0) return VEP_OK;
    vep::yuv_to_bgr24_batched<<<n_ctas, vep::kThreads, 0, st>>>(d_frames, d_work);
    CUDA_TRY(cudaGetLastError());
    return VEP_OK;
}

extern "C" int vep_create(int device, vep_ctx **out) {
    if (!out) return fail(VEP_ERR_INVALID_ARG, "out is null");
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail(VEP_ERR_CUDA, "no CUDA device (%s); this library has no CPU path", e == cudaSuccess ? "count 0" : cudaGetErrorString(e));
    if (device < 0 || device >= ndev) return fail(VEP_ERR_INVALID_ARG, "device %d out of range [0,%d)", device, ndev);
    cudaDeviceProp prop;
    CUDA_TRY(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10) return fail(VEP_ERR_CUDA, "device %d is sm_%d%d; this build carries sm_100a code only", device, prop.major, prop.minor);
    DeviceGuard g(device);
    vep_ctx *c = new (std::nothrow) vep_ctx();
    if (!c) return fail(VEP_ERR_NOMEM, "out of host memory");
    c->device = device;
    CUDA_TRY(cudaStreamCreateWithFlags(&c->main, cudaStreamNonBlocking));
    for (auto &t : c->tb) CUDA_TRY(cudaEventCreateWithFlags(&t.used, cudaEventDisableTiming));
    for (auto &t : c->ticks) { CUDA_TRY(cudaEventCreate(&t.k0)); CUDA_TRY(cudaEventCreate(&t.k1)); }
    *out = c;
    return VEP_OK;
}

extern "C" int vep_host_register(void *ptr, size_t bytes) {
    if (!ptr || !bytes) return fail(VEP_ERR_INVALID_ARG, "null pointer or zero length");
    CUDA_TRY(cudaHostRegister(ptr, bytes, cudaHostRegisterPortable));
    return VEP_OK;
}
extern "C" int vep_host_unregister(void *ptr) {
    if (!ptr) return fail(VEP_ERR_INVALID_ARG, "null pointer");
    CUDA_TRY(cudaHostUnregister(ptr));
    return VEP_OK;
}

extern "C" int vep_set_copy_run_limit(vep_ctx *c, int max_frames_per_copy) {
    if (!c) return fail(VEP_ERR_INVALID_ARG, "ctx is null");
    if (max_frames_per_copy < 0) return fail(VEP_ERR_INVALID_ARG, "limit %d < 0", max_frames_per_copy);
    std::lock_guard<std::mutex> lk(c->mu);
    c->run_limit = max_frames_per_copy;
    return VEP_OK;
}

extern "C" int vep_set_staging_threads(vep_ctx *c, int n) {
    if (!c) return fail(VEP_ERR_INVALID_ARG, "ctx is null");
    if (n < 0 || n > 64) return fail(VEP_ERR_INVALID_ARG, "staging threads %d outside [0,64]", n);
    std::lock_guard<std::mutex> lk(c->mu);
    c->pool.shutdown();
    c->pool.stop = false;
    c->stage_threads = n;
    if (n) c->pool.start(n);
    return VEP_OK;
}

static void free_group(StreamGroup *g) {
    if (!g) return;
    if (g->h_in) cudaFreeHost(g->h_in);
    if (g->h_out) cudaFreeHost(g->h_out);
    if (g->d_in) cudaFree(g->d_in);
    if (g->d_out) cudaFree(g->d_out);
    delete g;
}
static void free_stream(Stream *s) {
    if (!s) return;
    for (auto &sl : s->slots) if (sl.done) cudaEventDestroy(sl.done);
    if (s->cs_in) { cudaStreamSynchronize(s->cs_in); cudaStreamDestroy(s->cs_in); }
    if (s->cs_out) { cudaStreamSynchronize(s->cs_out); cudaStreamDestroy(s->cs_out); }
    if (s->ev_in) cudaEventDestroy(s->ev_in);
    if (s->d_out_own) cudaFree(s->d_out_own);
    if (s->grp && --s->grp->refs == 0) free_group(s->grp);   // the slot memory goes with the last member
    delete s;
}

extern "C" void vep_destroy(vep_ctx *c) {
    if (!c) return;
    DeviceGuard g(c->device);
    c->pool.shutdown();
    if (c->main) cudaStreamSynchronize(c->main);
    for (auto &kv : c->streams) free_stream(kv.second);
    for (auto &t : c->tb) { if (t.h) cudaFreeHost(t.h); if (t.d) cudaFree(t.d); if (t.used) cudaEventDestroy(t.used); }
    for (auto &t : c->ticks) { if (t.k0) cudaEventDestroy(t.k0); if (t.k1) cudaEventDestroy(t.k1); }
    if (c->main) cudaStreamDestroy(c->main);
    delete c;
}

// launch `t` on `st`: up to 128 frames ride in the kernel parameters, larger batches go through the upload ring
static int upload_and_launch(vep_ctx *c, const Tables &t, cudaStream_t st, int order) {
    if (t.total_items == 0) return VEP_OK;
    int grid = 0;
    const int cap = make_tick_blob(t, c->tick_blob, &grid, order);
    if (cap) {   // cudaLaunchKernel copies the parameter block before returning: the blob is reusable at once
        int rc = launch_tick(cap, c->tick_blob, grid, st);
        if (rc) return rc;
        c->stats.kernel_launches++;
        return VEP_OK;
    }
    TableBuf &b = c->tb[c->tb_next];
    c->tb_next = (c->tb_next + 1) % 4;
    if (b.pending) { CUDA_TRY(cudaEventSynchronize(b.used)); b.pending = false; }
    const size_t fb = (t.frames.size() * sizeof(FrameDev) + 255) / 256 * 256;
    const size_t need = fb + t.work.size() * sizeof(CtaWork);
    if (need > b.cap) {
        if (b.h) cudaFreeHost(b.h);
        if (b.d) cudaFree(b.d);
        b.h = b.d = nullptr;
        b.cap = 0;
        const size_t cap = std::max(need * 2, (size_t)1 << 16);
        CUDA_TRY(cudaHostAlloc(&b.h, cap, cudaHostAllocDefault));
        CUDA_TRY(cudaMalloc(&b.d, cap));
        b.cap = cap;
    }
    memcpy(b.h, t.frames.data(), t.frames.size() * sizeof(FrameDev));
    memcpy((char *)b.h + fb, t.work.data(), t.work.size() * sizeof(CtaWork));
    CUDA_TRY(cudaMemcpyAsync(b.d, b.h, need, cudaMemcpyHostToDevice, st));
    int rc = launch_tables((const FrameDev *)b.d, (const CtaWork *)((char *)b.d + fb), (int)t.work.size(), t.total_items, st);
    if (rc) return rc;
    CUDA_TRY(cudaEventRecord(b.used, st));
    b.pending = true;
    c->stats.kernel_launches++;
    return VEP_OK;
}

extern "C" int vep_convert_batch_ordered(vep_ctx *c, const vep_frame_desc *descs, int n, void *cuda_stream, int order) {
    if (!c) return fail(VEP_ERR_INVALID_ARG, "ctx is null");
    if (order < VEP_ORDER_STRICT || order > VEP_ORDER_NONE) return fail(VEP_ERR_INVALID_ARG, "unknown launch order %d", order);
    DeviceGuard g(c->device);
    std::lock_guard<std::mutex> lk(c->mu);
    int rc = build_tables(descs, n, c->scratch);
    if (rc) return rc;
    rc = upload_and_launch(c, c->scratch, (cudaStream_t)cuda_stream, order);
    if (rc) return rc;
    c->stats.frames_converted += c->scratch.n_converted;
    return VEP_OK;
}
extern "C" int vep_convert_batch(vep_ctx *c, const vep_frame_desc *descs, int n, void *cuda_stream) {
    return vep_convert_batch_ordered(c, descs, n, cuda_stream, VEP_ORDER_STRICT);
}

extern "C" int vep_plan_create(vep_ctx *c, const vep_frame_desc *descs, int n, vep_plan **out) {
    if (!c || !out) return fail(VEP_ERR_INVALID_ARG, "ctx or out is null");
    *out = nullptr;
    DeviceGuard g(c->device);
    Tables t;
    int rc = build_tables(descs, n, t);
    if (rc) return rc;
    vep_plan *p = new (std::nothrow) vep_plan();
    if (!p) return fail(VEP_ERR_NOMEM, "out of host memory");
    p->ctx = c;
    p->total_items = t.total_items;
    p->n_converted = t.n_converted;
    p->tick_cap = make_tick_blob(t, p->tick_blob, &p->tick_grid);
    if (p->tick_cap) {
        p->n_ctas = p->tick_grid;
        p->tick_blob_overlap = p->tick_blob;
        reinterpret_cast<int32_t *>(p->tick_blob_overlap.data())[1] = wait_prior_of(VEP_ORDER_NONE);   // TickParams::wait_prior
        p->tick_blob_writes = p->tick_blob;
        reinterpret_cast<int32_t *>(p->tick_blob_writes.data())[1] = wait_prior_of(VEP_ORDER_WRITES);
    } else {
        p->n_ctas = (int)t.work.size();
    }
    p->alg_bytes = t.alg_bytes;
    if (t.total_items && !p->tick_cap) {   // more than 128 frames: the table-driven kernel reads its tables from device memory
        cudaError_t e = cudaMalloc(&p->d_frames, t.frames.size() * sizeof(FrameDev));
        if (e == cudaSuccess) e = cudaMalloc(&p->d_work, t.work.size() * sizeof(CtaWork));
        if (e == cudaSuccess) e = cudaMemcpy(p->d_frames, t.frames.data(), t.frames.size() * sizeof(FrameDev), cudaMemcpyHostToDevice);
        if (e == cudaSuccess) e = cudaMemcpy(p->d_work, t.work.data(), t.work.size() * sizeof(CtaWork), cudaMemcpyHostToDevice);
        if (e != cudaSuccess) {
            if (p->d_frames) cudaFree(p->d_frames);
            if (p->d_work) cudaFree(p->d_work);
            delete p;
            return fail(VEP_ERR_CUDA, "plan upload failed: %s", cudaGetErrorString(e));
        }
    }
    *out = p;
    return VEP_OK;
}

static int plan_enqueue(vep_plan *p, cudaStream_t st, int order) {
    const std::vector<uint8_t> &blob = order == VEP_ORDER_NONE ? p->tick_blob_overlap : order == VEP_ORDER_WRITES ? p->tick_blob_writes : p->tick_blob;
    return p->tick_cap ? launch_tick(p->tick_cap, blob, p->tick_grid, st) : launch_tables(p->d_frames, p->d_work, p->n_ctas, p->total_items, st);
}
static int plan_launch(vep_plan *p, void *cuda_stream, int order) {
    if (!p) return fail(VEP_ERR_INVALID_ARG, "plan is null");
    if (order < VEP_ORDER_STRICT || order > VEP_ORDER_NONE) return fail(VEP_ERR_INVALID_ARG, "unknown launch order %d", order);
    DeviceGuard g(p->ctx->device);
    int rc = plan_enqueue(p, (cudaStream_t)cuda_stream, order);
    if (rc) return rc;
    __atomic_add_fetch(&p->ctx->stats.kernel_launches, 1, __ATOMIC_RELAXED);
    __atomic_add_fetch(&p->ctx->stats.frames_converted, (int64_t)p->n_converted, __ATOMIC_RELAXED);
    return VEP_OK;
}
extern "C" int vep_plan_launch(vep_plan *p, void *cuda_stream) { return plan_launch(p, cuda_stream, VEP_ORDER_STRICT); }
extern "C" int vep_plan_launch_overlapped(vep_plan *p, void *cuda_stream) { return plan_launch(p, cuda_stream, VEP_ORDER_NONE); }
extern "C" int vep_plan_launch_ordered(vep_plan *p, void *cuda_stream, int order) { return plan_launch(p, cuda_stream, order); }

// A run of launches issued from native code in one call: what a C producer that converts tick after tick does, and
// what the benchmark uses so that the Python call overhead per launch stays out of a short timed region.
extern "C" int vep_plan_launch_sequence(vep_plan *const *plans, int n, void *cuda_stream, int order, void *event_before, void *event_after) {
    if (n < 0 || (n > 0 && !plans)) return fail(VEP_ERR_INVALID_ARG, "plans is null or n < 0");
    if (order < VEP_ORDER_STRICT || order > VEP_ORDER_NONE) return fail(VEP_ERR_INVALID_ARG, "unknown launch order %d", order);
    for (int i = 0; i < n; ++i) if (!plans[i]) return fail(VEP_ERR_INVALID_ARG, "plans[%d] is null", i);
    if (n == 0) {
        if (event_before) CUDA_TRY(cudaEventRecord((cudaEvent_t)event_before, (cudaStream_t)cuda_stream));
        if (event_after) CUDA_TRY(cudaEventRecord((cudaEvent_t)event_after, (cudaStream_t)cuda_stream));
        return VEP_OK;
    }
    DeviceGuard g(plans[0]->ctx->device);
    if (event_before) CUDA_TRY(cudaEventRecord((cudaEvent_t)event_before, (cudaStream_t)cuda_stream));
    for (int i = 0; i < n; ++i) {
        vep_plan *p = plans[i];
        if (p->ctx->device != plans[0]->ctx->device) return fail(VEP_ERR_INVALID_ARG, "plans[%d] belongs to another device", i);
        int rc = plan_enqueue(p, (cudaStream_t)cuda_stream, order);
        if (rc) return rc;
        __atomic_add_fetch(&p->ctx->stats.kernel_launches, 1, __ATOMIC_RELAXED);
        __atomic_add_fetch(&p->ctx->stats.frames_converted, (int64_t)p->n_converted, __ATOMIC_RELAXED);
    }
    if (event_after) CUDA_TRY(cudaEventRecord((cudaEvent_t)event_after, (cudaStream_t)cuda_stream));
    return VEP_OK;
}

extern "C" int vep_plan_info(const vep_plan *p, int *n_converted, int *n_ctas, int64_t *alg_bytes) {
    if (!p) return fail(VEP_ERR_INVALID_ARG, "plan is null");
    if (n_converted) *n_converted = p->n_converted;
    if (n_ctas) *n_ctas = p->n_ctas;
    if (alg_bytes) *alg_bytes = p->alg_bytes;
    return VEP_OK;
}

extern "C" void vep_plan_destroy(vep_plan *p) {
    if (!p) return;
    DeviceGuard g(p->ctx->device);
    if (p->d_frames) cudaFree(p->d_frames);
    if (p->d_work) cudaFree(p->d_work);
    delete p;
}

// ------------------------------------------------------------------------------------------------
// VideoFrame wire framing (proto3; proto/video_streaming.proto:67-93, filled as read_image.py:99-117)
// ------------------------------------------------------------------------------------------------
static inline uint8_t *put_varint(uint8_t *p, uint64_t v) {
    while (v >= 0x80) { *p++ = (uint8_t)(v | 0x80); v >>= 7; }
    *p++ = (uint8_t)v;
    return p;
}
static inline uint8_t *put_int(uint8_t *p, int field, int64_t v) {   // int64 / bool, default omitted
    if (v == 0) return p;
    p = put_varint(p, (uint64_t)field << 3);
    return put_varint(p, (uint64_t)v);   // negative values: 10-byte two's complement varint
}

extern "C" int vep_videoframe_frame(const vep_frame_meta *m, int32_t width, int32_t height, uint8_t *prefix,
                                    int32_t *prefix_len, uint8_t *suffix, int32_t *suffix_len) {
    if (!m || !prefix || !prefix_len || !suffix || !suffix_len) return fail(VEP_ERR_INVALID_ARG, "null argument");
    if (width <= 0 || height <= 0) return fail(VEP_ERR_GEOMETRY, "bad geometry %dx%d", width, height);
    uint8_t *p = prefix;
    p = put_int(p, 1, width);
    p = put_int(p, 2, height);
    *p++ = (3 << 3) | 2;   // data: length-delimited, payload follows
    p = put_varint(p, (uint64_t)width * height * 3);
    *prefix_len = (int32_t)(p - prefix);

    uint8_t *s = suffix;
    s = put_int(s, 4, m->timestamp);
    s = put_int(s, 5, m->is_keyframe ? 1 : 0);
    s = put_int(s, 6, m->pts);
    s = put_int(s, 7, m->dts);
    const size_t ft = strnlen(m->frame_type, sizeof m->frame_type);
    if (ft) { *s++ = (8 << 3) | 2; *s++ = (uint8_t)ft; memcpy(s, m->frame_type, ft); s += ft; }
    s = put_int(s, 9, m->is_corrupt ? 1 : 0);
    uint64_t tb;
    memcpy(&tb, &m->time_base, 8);
    if (tb != 0) { *s++ = (10 << 3) | 1; memcpy(s, &tb, 8); s += 8; }   // little-endian host (x86-64)
    // shape = ShapeProto{ dim: [ {size=H,name="0"}, {size=W,name="1"}, {size=3,name="2"} ] }
    uint8_t shp[64], *q = shp;
    const int64_t dims[3] = {height, width, 3};
    for (int i = 0; i < 3; ++i) {
        uint8_t dim[16], *r = dim;
        r = put_int(r, 1, dims[i]);
        *r++ = (2 << 3) | 2; *r++ = 1; *r++ = (uint8_t)('0' + i);
        *q++ = (2 << 3) | 2; *q++ = (uint8_t)(r - dim);
        memcpy(q, dim, r - dim); q += r - dim;
    }
    *s++ = (11 << 3) | 2; *s++ = (uint8_t)(q - shp);
    memcpy(s, shp, q - shp); s += q - shp;
    s = put_int(s, 13, m->packet);
    s = put_int(s, 14, m->keyframe);
    *suffix_len = (int32_t)(s - suffix);
    return VEP_OK;
}

// ------------------------------------------------------------------------------------------------
// camera streams, staging, ticks
// ------------------------------------------------------------------------------------------------
static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

static Stream *find_stream(vep_ctx *c, int id) {
    std::shared_lock<std::shared_mutex> lk(c->smu);
    auto it = c->streams.find(id);
    return it == c->streams.end() ? nullptr : it->second;
}

// row pitch of a staged plane / of the device output: tight when that already is a multiple of 16 bytes (every
// common resolution), else padded to 64 so that widths like 854 or 1366 keep the kernels' vector path
static inline int pitch_of(size_t row_bytes) { return (int)(row_bytes % 16 == 0 ? row_bytes : align_up(row_bytes, 64)); }

static int check_stream_spec(int fmt, int w, int h) {
    if (fmt < VEP_FMT_YUV420P || fmt > VEP_FMT_GRAY8) return fail(VEP_ERR_INVALID_ARG, "unknown fmt %d", fmt);
    if (w <= 0 || h <= 0 || w > kMaxDim || h > kMaxDim || (w & 1) || (h & 1)) return fail(VEP_ERR_GEOMETRY, "%dx%d: need even, positive dimensions", w, h);
    if ((vep::fmt_is_semiplanar(fmt) || vep::fmt_bytes_per_sample(fmt) == 2) && h < 8) return fail(VEP_ERR_GEOMETRY, "nv12/nv21/10-bit height %d < 8", h);
    return VEP_OK;
}

// geometry of one stream's slots
static void layout_stream(Stream *s, int stream_id, int fmt, int w, int h, int ring_depth) {
    s->id = stream_id; s->fmt = fmt; s->w = w; s->h = h; s->ring = ring_depth;
    const size_t bps = vep::fmt_bytes_per_sample(fmt);
    const bool semi = vep::fmt_is_semiplanar(fmt), gray = fmt == VEP_FMT_GRAY8;
    const size_t crow = ((semi || vep::fmt_is_444(fmt)) ? (size_t)w : (size_t)w / 2) * bps;   // bytes per chroma row
    const size_t crows = (vep::fmt_is_422(fmt) || vep::fmt_is_444(fmt)) ? (size_t)h : (size_t)h / 2;
    s->pitch_y = pitch_of((size_t)w * bps);
    s->pitch_c = gray ? 0 : pitch_of(crow);
    s->pitch_out = pitch_of((size_t)w * 3);
    const size_t ysz = (size_t)s->pitch_y * h, csz = (size_t)s->pitch_c * crows;
    s->off_c0 = align_up(ysz, 256);
    if (gray) { s->off_c0 = s->off_c1 = 0; s->in_bytes = ysz; }
    else if (semi) { s->off_c1 = 0; s->in_bytes = s->off_c0 + csz; }
    else { s->off_c1 = s->off_c0 + align_up(csz, 256); s->in_bytes = s->off_c1 + csz; }
    s->out_bytes = (size_t)w * h * 3;
    s->dev_out_bytes = (size_t)s->pitch_out * h;
    s->in_slot = align_up(s->in_bytes, 4096);
    s->out_slot = align_up(kHdrRoom + s->out_bytes + kTrailerRoom, 4096);
    s->dev_out_slot = align_up(s->dev_out_bytes, 4096);
    s->padded_out = (size_t)s->pitch_out != (size_t)w * 3;
}

// Open n camera streams whose slots share one allocation per kind, laid out [slot index][stream]: slot k of
// consecutive streams is contiguous in pinned memory and, identically, in device memory.  A tick in which such
// neighbours each stage a frame in the same slot index (a fleet of cameras ticking in lock step) then moves all of
// them with ONE copy per direction instead of one per camera (tick_enqueue); measured on one B200 the 16 x 1080p
// tick's DMA ceiling is 79 GB/s with one copy per direction against 70 GB/s with 16 + 16 (profiles/r2_pcie_probe_*).
extern "C" int vep_stream_open_many(vep_ctx *c, const vep_stream_spec *specs, int n, int ring_depth) {
    if (!c) return fail(VEP_ERR_INVALID_ARG, "ctx is null");
    if (n <= 0 || !specs) return fail(VEP_ERR_INVALID_ARG, "specs is null or n <= 0");
    if (ring_depth < 1) ring_depth = 1;
    if (ring_depth > 64) return fail(VEP_ERR_INVALID_ARG, "ring_depth %d > 64", ring_depth);
    for (int i = 0; i < n; ++i) {
        int rc = check_stream_spec(specs[i].fmt, specs[i].width, specs[i].height);
        if (rc) return rc;
        for (int j = 0; j < i; ++j) if (specs[j].stream_id == specs[i].stream_id) return fail(VEP_ERR_INVALID_ARG, "stream %d listed twice", specs[i].stream_id);
    }
    DeviceGuard g(c->device);
    {
        std::shared_lock<std::shared_mutex> lk(c->smu);
        for (int i = 0; i < n; ++i) if (c->streams.count(specs[i].stream_id)) return fail(VEP_ERR_STATE, "stream %d already open", specs[i].stream_id);
    }
    const int ns = ring_depth + VEP_MAX_TICKS_IN_FLIGHT + 2;   // ring + ticks in flight + one filling + slack
    StreamGroup *grp = new (std::nothrow) StreamGroup();
    if (!grp) return fail(VEP_ERR_NOMEM, "out of host memory");
    std::vector<Stream *> made;
    auto undo = [&](int code, const char *what, cudaError_t e) {
        for (Stream *s : made) free_stream(s);
        if (made.empty()) free_group(grp);
        return fail(code, "stream_open: %s%s%s", what, e != cudaSuccess ? ": " : "", e != cudaSuccess ? cudaGetErrorString(e) : "");
    };
    std::vector<size_t> in_off(n), out_off(n);
    bool any_padded_rows = false;
    for (int i = 0; i < n; ++i) {
        Stream *s = new (std::nothrow) Stream();
        if (!s) return undo(VEP_ERR_NOMEM, "out of host memory", cudaSuccess);
        layout_stream(s, specs[i].stream_id, specs[i].fmt, specs[i].width, specs[i].height, ring_depth);
        s->grp = grp; s->member = i;
        grp->refs++;
        made.push_back(s);
        in_off[i] = grp->in_stride; grp->in_stride += s->in_slot;
        out_off[i] = grp->out_stride; grp->out_stride += s->out_slot;
        any_padded_rows |= s->pitch_y != (int)(s->w * vep::fmt_bytes_per_sample(s->fmt));
    }
    // Input staging is written by the CPU (streaming stores) and read by the DMA engine.  Write-combined pinned
    // memory measured the same end to end as ordinary pinned memory on the B200 host (round 1), and a caller that
    // reads back a slot obtained from vep_acquire_input would crawl on it, so ordinary pinned memory is used.
    cudaError_t e = cudaHostAlloc((void **)&grp->h_out, grp->out_stride * ns, cudaHostAllocDefault);
    if (e == cudaSuccess) e = cudaHostAlloc((void **)&grp->h_in, grp->in_stride * ns, cudaHostAllocDefault);
    if (e == cudaSuccess) e = cudaMalloc((void **)&grp->d_in, grp->in_stride * ns);
    if (e == cudaSuccess) e = cudaMalloc((void **)&grp->d_out, grp->out_stride * ns);
    // padded rows: the padding is read by the last lane of a row (never used); give it defined contents once
    if (e == cudaSuccess && any_padded_rows) memset(grp->h_in, 0, grp->in_stride * ns);
    for (int i = 0; i < n && e == cudaSuccess; ++i) {
        Stream *s = made[i];
        s->slots.resize(ns);
        // device rows of the output are padded for this geometry: it cannot share the mirrored layout, own buffer
        if (s->padded_out) e = cudaMalloc((void **)&s->d_out_own, s->dev_out_slot * ns);
        if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&s->cs_in, cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&s->cs_out, cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&s->ev_in, cudaEventDisableTiming);
        for (int k = 0; k < ns && e == cudaSuccess; ++k) {
            Slot &sl = s->slots[k];
            sl.h_in = grp->h_in + grp->in_stride * k + in_off[i];
            sl.d_in = grp->d_in + grp->in_stride * k + in_off[i];
            sl.h_out = grp->h_out + grp->out_stride * k + out_off[i];
            // mirrored layout: the device payload sits at the same offset as the pinned payload (behind the header room)
            sl.d_out = s->padded_out ? s->d_out_own + s->dev_out_slot * k : grp->d_out + grp->out_stride * k + out_off[i] + kHdrRoom;
            e = cudaEventCreateWithFlags(&sl.done, cudaEventDisableTiming);
        }
    }
    if (e != cudaSuccess) return undo(e == cudaErrorMemoryAllocation ? VEP_ERR_NOMEM : VEP_ERR_CUDA, "allocation failed", e);
    std::unique_lock<std::shared_mutex> lk(c->smu);
    for (int i = 0; i < n; ++i)
        if (c->streams.count(specs[i].stream_id)) {
            lk.unlock();
            return undo(VEP_ERR_STATE, "a stream of the group was opened meanwhile", cudaSuccess);
        }
    for (Stream *s : made) c->streams[s->id] = s;
    return VEP_OK;
}

extern "C" int vep_stream_open(vep_ctx *c, int stream_id, int fmt, int w, int h, int ring_depth) {
    const vep_stream_spec one = {stream_id, fmt, w, h};
    return vep_stream_open_many(c, &one, 1, ring_depth);
}

extern "C" int vep_stream_close(vep_ctx *c, int stream_id) {
    if (!c) return fail(VEP_ERR_INVALID_ARG, "ctx is null");
    DeviceGuard g(c->device);
    Stream *s = nullptr;
    {
        std::lock_guard<std::mutex> tk(c->mu);
        if (c->ticks_in_flight) return fail(VEP_ERR_BUSY, "ticks in flight; call vep_tick_wait first");
        std::unique_lock<std::shared_mutex> lk(c->smu);
        auto it = c->streams.find(stream_id);
        if (it == c->streams.end()) return fail(VEP_ERR_INVALID_ARG, "stream %d not open", stream_id);
        s = it->second;
        c->streams.erase(it);
    }
    free_stream(s);
    return VEP_OK;
}

// pick the slot a new frame goes into; caller holds s->mu
static int claim_slot(Stream *s) {
    Slot &sl = s->slots[s->wr];
    if (sl.state == SLOT_FILLING) return s->wr;   // re-acquire without commit
    if (sl.state == SLOT_SUBMITTED || sl.state == SLOT_INFLIGHT) return -1;
    if (sl.state == SLOT_READY) {
        // READY slots are reused oldest-first; the ring newest ones must survive
        if (s->n_ready <= s->ring) return -1;
        s->n_ready--;
    }
    sl.state = SLOT_FILLING;
    return s->wr;
}

static void plane_ptrs(const Stream *s, uint8_t *base, uint8_t *pl[3], int32_t st[3]) {
    const bool semi = vep::fmt_is_semiplanar(s->fmt);
    pl[0] = base; pl[1] = base + s->off_c0; pl[2] = semi ? nullptr : base + s->off_c1;
    st[0] = s->pitch_y; st[1] = s->pitch_c; st[2] = semi ? 0 : s->pitch_c;
    if (s->fmt == VEP_FMT_GRAY8) { pl[1] = pl[2] = nullptr; st[1] = st[2] = 0; }
}

extern "C" int vep_acquire_input(vep_ctx *c, int stream_id, uint8_t *plane_out[3], int32_t stride_out[3]) {
    if (!c || !plane_out || !stride_out) return fail(VEP_ERR_INVALID_ARG, "null argument");
    Stream *s = find_stream(c, stream_id);
    if (!s) return fail(VEP_ERR_INVALID_ARG, "stream %d not open", stream_id);
    std::lock_guard<std::mutex> lk(s->mu);
    const int i = claim_slot(s);
    if (i < 0) return fail(VEP_ERR_BUSY, "stream %d: no free staging slot (tick more often or raise ring_depth)", stream_id);
    s->slots[i].external = false;
    plane_ptrs(s, s->slots[i].h_in, plane_out, stride_out);
    return VEP_OK;
}

extern "C" int vep_commit_input(vep_ctx *c, int stream_id, const vep_frame_meta *meta) {
    if (!c) return fail(VEP_ERR_INVALID_ARG, "ctx is null");
    Stream *s = find_stream(c, stream_id);
    if (!s) return fail(VEP_ERR_INVALID_ARG, "stream %d not open", stream_id);
    std::lock_guard<std::mutex> lk(s->mu);
    Slot &sl = s->slots[s->wr];
    if (sl.state != SLOT_FILLING) return fail(VEP_ERR_STATE, "stream %d: commit without acquire", stream_id);
    if (meta) sl.meta = *meta; else memset(&sl.meta, 0, sizeof sl.meta);
    sl.state = SLOT_SUBMITTED;
    s->wr = (s->wr + 1) % (int)s->slots.size();
    __atomic_add_fetch(&c->stats.frames_offered, 1, __ATOMIC_RELAXED);
    return VEP_OK;
}

// Staging copy into pinned memory.  The destination is read next by the GPU's DMA engine, never by
// this core, so it is written with non-temporal stores: no read-for-ownership traffic and no dirty
// lines left in the producer core's cache for the H2D copy to snoop (measured on the B200 host:
// H2D of freshly memcpy'd pinned data ran ~2x slower than from cold pinned memory).
__attribute__((target("avx2"))) static void stream_copy_avx2(uint8_t *dst, const uint8_t *src, size_t n) {
    size_t head = (32 - (reinterpret_cast<uintptr_t>(dst) & 31)) & 31;
    if (head > n) head = n;
    memcpy(dst, src, head);
    dst += head; src += head; n -= head;
    size_t i = 0;
    for (; i + 128 <= n; i += 128) {
        const __m256i a = _mm256_loadu_si256((const __m256i *)(src + i)), b = _mm256_loadu_si256((const __m256i *)(src + i + 32));
        const __m256i c = _mm256_loadu_si256((const __m256i *)(src + i + 64)), d = _mm256_loadu_si256((const __m256i *)(src + i + 96));
        _mm256_stream_si256((__m256i *)(dst + i), a); _mm256_stream_si256((__m256i *)(dst + i + 32), b);
        _mm256_stream_si256((__m256i *)(dst + i + 64), c); _mm256_stream_si256((__m256i *)(dst + i + 96), d);
    }
    for (; i + 32 <= n; i += 32) _mm256_stream_si256((__m256i *)(dst + i), _mm256_loadu_si256((const __m256i *)(src + i)));
    memcpy(dst + i, src + i, n - i);
}
static void stream_copy(uint8_t *dst, const uint8_t *src, size_t n) {
    static const bool avx2 = __builtin_cpu_supports("avx2");
    if (avx2 && n >= 4096) stream_copy_avx2(dst, src, n); else memcpy(dst, src, n);
}
static void copy_rows(uint8_t *dst, int dstride, const uint8_t *src, int sstride, int row_bytes, int r0, int r1) {
    if (dstride == sstride && sstride == row_bytes) stream_copy(dst + (size_t)r0 * dstride, src + (size_t)r0 * sstride, (size_t)row_bytes * (r1 - r0));
    else for (int r = r0; r < r1; ++r) stream_copy(dst + (size_t)r * dstride, src + (size_t)r * sstride, row_bytes);
    _mm_sfence();   // order the streaming stores before the commit that makes the slot visible
}

struct PlaneCopy { uint8_t *dst; int dstride; const uint8_t *src; int sstride, row_bytes, rows; };
struct CopyJob { PlaneCopy p; int r0, r1; };

// Run the copy jobs on the calling thread plus the context's helper threads: every thread takes the next job off a
// shared counter until none is left.  Returns when all jobs are done.
static void run_copy_jobs(vep_ctx *c, const std::vector<CopyJob> &jobs) {
    auto one = [&jobs](int i) { const CopyJob &j = jobs[i]; copy_rows(j.p.dst, j.p.dstride, j.p.src, j.p.sstride, j.p.row_bytes, j.r0, j.r1); };
    const int n = (int)jobs.size();
    const int helpers = std::min(c->stage_threads, n - 1);
    if (helpers <= 0) {
        for (int i = 0; i < n; ++i) one(i);
        return;
    }
    struct Shared { std::atomic<int> next{0}; std::mutex mu; std::condition_variable cv; int left; } sh;
    sh.left = helpers;
    auto drain = [&sh, &one, n] { for (int i; (i = sh.next.fetch_add(1, std::memory_order_relaxed)) < n;) one(i); };
    for (int k = 0; k < helpers; ++k)
        c->pool.post([&sh, drain] {
            drain();
            std::lock_guard<std::mutex> lk(sh.mu);
            if (--sh.left == 0) sh.cv.notify_one();
        });
    drain();
    std::unique_lock<std::mutex> lk(sh.mu);
    sh.cv.wait(lk, [&sh] { return sh.left == 0; });
}

// cut the planes of one frame into jobs of about kJobBytes (whole rows)
static constexpr size_t kJobBytes = (size_t)1 << 20;
static void add_plane_jobs(std::vector<CopyJob> &jobs, const PlaneCopy &p) {
    const int rows_per = std::max(1, (int)(kJobBytes / (size_t)std::max(p.row_bytes, 1)));
    for (int r = 0; r < p.rows; r += rows_per) jobs.push_back(CopyJob{p, r, std::min(p.rows, r + rows_per)});
}

// page-locked (cudaHostAlloc'ed or cudaHostRegister'ed) host memory?  first and last byte of the range are asked for
static bool range_is_pinned(const uint8_t *p, size_t bytes) {
    for (const uint8_t *q : {p, p + (bytes ? bytes - 1 : 0)}) {
        cudaPointerAttributes a{};
        if (cudaPointerGetAttributes(&a, q) != cudaSuccess) { cudaGetLastError(); return false; }
        if (a.type != cudaMemoryTypeHost) return false;
    }
    return true;
}

// validate one frame against its stream and claim its staging slot; fills the plane copies (mode VEP_SUBMIT_COPY) or
// notes the caller's planes in the slot (VEP_SUBMIT_PINNED: *np = 0, nothing to copy)
static int prepare_submit(vep_ctx *c, Stream *s, const uint8_t *const plane[3], const int32_t stride[3], PlaneCopy pc[3], int *np, int mode = VEP_SUBMIT_COPY) {
    const bool semi = vep::fmt_is_semiplanar(s->fmt), gray = s->fmt == VEP_FMT_GRAY8;
    if (!plane || !stride || !plane[0] || (!gray && (!plane[1] || (!semi && !plane[2])))) return fail(VEP_ERR_INVALID_ARG, "null plane");
    const int bps = vep::fmt_bytes_per_sample(s->fmt);
    const int cw = ((semi || vep::fmt_is_444(s->fmt)) ? s->w : s->w / 2) * bps;      // chroma row bytes
    const int crows = (vep::fmt_is_422(s->fmt) || vep::fmt_is_444(s->fmt)) ? s->h : s->h / 2;
    if (stride[0] < s->w * bps || (!gray && (stride[1] < cw || (!semi && stride[2] < cw)))) return fail(VEP_ERR_GEOMETRY, "stride smaller than row");
    if (mode != VEP_SUBMIT_COPY && mode != VEP_SUBMIT_PINNED) return fail(VEP_ERR_INVALID_ARG, "unknown submit mode %d", mode);
    if (mode == VEP_SUBMIT_PINNED) {
        const int np_ = gray ? 1 : semi ? 2 : 3;
        for (int i = 0; i < np_; ++i) {
            const size_t rb = i == 0 ? (size_t)s->w * bps : (size_t)cw, rows = i == 0 ? (size_t)s->h : (size_t)crows;
            if (!range_is_pinned(plane[i], (rows - 1) * (size_t)stride[i] + rb))
                return fail(VEP_ERR_INVALID_ARG, "stream %d: plane %d is not in page-locked host memory (vep_host_register it, or submit with VEP_SUBMIT_COPY)", s->id, i);
        }
    }
    uint8_t *pl[3];
    int32_t st[3];
    {
        std::lock_guard<std::mutex> lk(s->mu);
        const int i = claim_slot(s);
        if (i < 0) return fail(VEP_ERR_BUSY, "stream %d: no free staging slot (tick more often or raise ring_depth)", s->id);
        Slot &sl = s->slots[i];
        plane_ptrs(s, sl.h_in, pl, st);
        sl.external = mode == VEP_SUBMIT_PINNED;
        for (int k = 0; k < 3; ++k) { sl.ext_plane[k] = sl.external ? plane[k] : nullptr; sl.ext_stride[k] = sl.external ? stride[k] : 0; }
    }
    if (mode == VEP_SUBMIT_PINNED) { *np = 0; return VEP_OK; }
    int n = 0;
    pc[n++] = PlaneCopy{pl[0], st[0], plane[0], stride[0], s->w * bps, s->h};
    if (!gray) pc[n++] = PlaneCopy{pl[1], st[1], plane[1], stride[1], cw, crows};
    if (!semi && !gray) pc[n++] = PlaneCopy{pl[2], st[2], plane[2], stride[2], cw, crows};
    *np = n;
    return VEP_OK;
}

extern "C" int vep_submit(vep_ctx *c, int stream_id, const uint8_t *const plane[3], const int32_t stride[3],
                          const vep_frame_meta *meta, int convert) {
    if (!c) return fail(VEP_ERR_INVALID_ARG, "ctx is null");
    Stream *s = find_stream(c, stream_id);
    if (!s) return fail(VEP_ERR_INVALID_ARG, "stream %d not open", stream_id);
    if (!convert) {   // gate said no (read_image.py:80): counted, nothing copied
        __atomic_add_fetch(&c->stats.frames_offered, 1, __ATOMIC_RELAXED);
        return VEP_OK;
    }
    PlaneCopy pc[3];
    int np = 0;
    int rc = prepare_submit(c, s, plane, stride, pc, &np, convert);
    if (rc) return rc;
    // the copy runs outside the lock (one producer per stream); a large frame is shared with the helper threads
    size_t total = 0;
    for (int i = 0; i < np; ++i) total += (size_t)pc[i].row_bytes * pc[i].rows;
    if (c->stage_threads == 0 || total < kBandSplitBytes) {
        for (int i = 0; i < np; ++i) copy_rows(pc[i].dst, pc[i].dstride, pc[i].src, pc[i].sstride, pc[i].row_bytes, 0, pc[i].rows);
    } else {
        std::vector<CopyJob> jobs;
        for (int i = 0; i < np; ++i) add_plane_jobs(jobs, pc[i]);
        run_copy_jobs(c, jobs);
    }
    return vep_commit_input(c, stream_id, meta);
}

// All frames of a tick with one call: the copies of ALL items are cut into ~1 MB jobs and spread over the calling
// thread and the helper threads, then every item is committed.  Stops at the first item that cannot be staged.
extern "C" int vep_submit_batch(vep_ctx *c, const vep_submit_item *items, int n, int *n_done) {
    if (n_done) *n_done = 0;
    if (!c) return fail(VEP_ERR_INVALID_ARG, "ctx is null");
    if (n < 0 || (n > 0 && !items)) return fail(VEP_ERR_INVALID_ARG, "items is null or n < 0");
    std::vector<CopyJob> jobs;
    int ok = 0, rc = VEP_OK;
    for (; ok < n; ++ok) {
        const vep_submit_item &it = items[ok];
        Stream *s = find_stream(c, it.stream_id);
        if (!s) { rc = fail(VEP_ERR_INVALID_ARG, "item %d: stream %d not open", ok, it.stream_id); break; }
        if (!it.convert) continue;
        PlaneCopy pc[3];
        int np = 0;
        rc = prepare_submit(c, s, it.plane, it.stride, pc, &np, it.convert);
        if (rc) break;
        for (int i = 0; i < np; ++i) add_plane_jobs(jobs, pc[i]);
    }
    run_copy_jobs(c, jobs);
    for (int i = 0; i < ok; ++i) {
        if (!items[i].convert) { __atomic_add_fetch(&c->stats.frames_offered, 1, __ATOMIC_RELAXED); continue; }
        const int rc2 = vep_commit_input(c, items[i].stream_id, &items[i].meta);
        if (rc2 && !rc) rc = rc2;
    }
    if (n_done) *n_done = ok;
    return rc;
}

static void write_framing(Stream *s, Slot &sl) {
    uint8_t pre[32], suf[kTrailerRoom];
    int32_t pl = 0, sl_len = 0;
    vep_videoframe_frame(&sl.meta, s->w, s->h, pre, &pl, suf, &sl_len);
    memcpy(sl.h_out + kHdrRoom - pl, pre, pl);
    memcpy(sl.h_out + kHdrRoom + s->out_bytes, suf, sl_len);
    sl.msg_off = kHdrRoom - pl;
    sl.msg_len = (size_t)pl + s->out_bytes + sl_len;
}

static int tick_enqueue(vep_ctx *c, TickRec &t, int *n_frames);

extern "C" int vep_tick_begin_ticket(vep_ctx *c, int *n_frames, uint64_t *ticket) {
    if (!c) return fail(VEP_ERR_INVALID_ARG, "ctx is null");
    if (n_frames) *n_frames = 0;
    if (ticket) *ticket = 0;
    DeviceGuard g(c->device);
    std::lock_guard<std::mutex> lk(c->mu);
    if (c->ticks_in_flight >= VEP_MAX_TICKS_IN_FLIGHT) return fail(VEP_ERR_BUSY, "%d ticks already in flight; call vep_tick_wait", VEP_MAX_TICKS_IN_FLIGHT);
    TickRec &t = c->ticks[(c->tick_head + c->ticks_in_flight) % VEP_MAX_TICKS_IN_FLIGHT];
    t.slots.clear();
    int nf = 0;
    const int rc = tick_enqueue(c, t, &nf);
    if (rc != VEP_OK) {
        // a CUDA call failed half-way: the frames of this tick are dropped (the reference drops a frame on any
        // exception, read_image.py:130-131) and their slots are released so the cameras are not wedged
        cudaDeviceSynchronize();     // copies of a run are queued on its first camera's streams: wait for all of them
        for (auto &r : t.slots) {
            std::lock_guard<std::mutex> sl_lk(r.s->mu);
            r.s->slots[r.slot].state = SLOT_FREE;
        }
        t.slots.clear();
        return rc;
    }
    if (n_frames) *n_frames = nf;
    // an empty tick starts nothing; its ticket is that of the newest real tick, so waiting for it still means
    // "everything begun before this call has landed"
    if (ticket) *ticket = nf ? t.ticket : c->next_ticket - 1;
    return VEP_OK;
}
extern "C" int vep_tick_begin(vep_ctx *c, int *n_frames) { return vep_tick_begin_ticket(c, n_frames, nullptr); }

// VEP_SUBMIT_PINNED frames: plane p goes over as one linear copy when its linesize is a multiple of 16 and its rows,
// padding included, fit the plane's region of the slot; the kernel is then told the caller's linesize
static void external_plane_geometry(const Stream *s, int p, size_t *row_bytes, size_t *rows, size_t *room) {
    const int bps = vep::fmt_bytes_per_sample(s->fmt);
    const bool semi = vep::fmt_is_semiplanar(s->fmt);
    const size_t crow = ((semi || vep::fmt_is_444(s->fmt)) ? (size_t)s->w : (size_t)s->w / 2) * bps;
    const size_t crows = (vep::fmt_is_422(s->fmt) || vep::fmt_is_444(s->fmt)) ? (size_t)s->h : (size_t)s->h / 2;
    const size_t region_end[3] = {s->fmt == VEP_FMT_GRAY8 ? s->in_slot : s->off_c0, semi ? s->in_slot : s->off_c1, s->in_slot};
    const size_t region_begin[3] = {0, s->off_c0, s->off_c1};
    *row_bytes = p == 0 ? (size_t)s->w * bps : crow;
    *rows = p == 0 ? (size_t)s->h : crows;
    *room = region_end[p] - region_begin[p];
}
static bool external_plane_is_linear(const Stream *s, const Slot &sl, int p) {
    size_t rb, rows, room;
    external_plane_geometry(s, p, &rb, &rows, &room);
    const size_t ls = (size_t)sl.ext_stride[p];
    return ls % 16 == 0 && (rows - 1) * ls + rb <= room;
}
// Straight from the caller's page-locked planes: linear where possible (decoders pad linesizes to 32/64 bytes, so
// this is the usual case); any other linesize is re-pitched by the copy engine (row-wise DMA: slow for narrow rows,
// but correct for every layout).
static int enqueue_external_h2d(vep_ctx *c, Stream *s, Slot &sl) {
    uint8_t *pl[3];
    int32_t st[3];
    plane_ptrs(s, sl.d_in, pl, st);
    for (int p = 0; p < 3; ++p) {
        if (!pl[p]) continue;
        size_t rb, rows, room;
        external_plane_geometry(s, p, &rb, &rows, &room);
        const size_t ls = (size_t)sl.ext_stride[p];
        if (external_plane_is_linear(s, sl, p)) {
            const size_t span = (rows - 1) * ls + rb;
            CUDA_TRY(cudaMemcpyAsync(pl[p], sl.ext_plane[p], span, cudaMemcpyHostToDevice, s->cs_in));
            c->stats.h2d_bytes += (int64_t)span;
        } else {
            CUDA_TRY(cudaMemcpy2DAsync(pl[p], st[p], sl.ext_plane[p], ls, rb, rows, cudaMemcpyHostToDevice, s->cs_in));
            c->stats.h2d_bytes += (int64_t)(rb * rows);
        }
    }
    return VEP_OK;
}

static int tick_enqueue(vep_ctx *c, TickRec &t, int *n_frames) {
    std::vector<vep_frame_desc> descs;
    std::shared_lock<std::shared_mutex> map_lk(c->smu);
    // 1. claim every committed frame (oldest first per camera) and describe it for the kernel
    for (auto &kv : c->streams) {
        Stream *s = kv.second;
        std::lock_guard<std::mutex> sl_lk(s->mu);
        const int ns = (int)s->slots.size();
        // slots are committed in ring order ending just before wr; walk them oldest first
        for (int k = ns; k >= 1; --k) {
            const int i = ((s->wr - k) % ns + ns) % ns;
            Slot &sl = s->slots[i];
            if (sl.state != SLOT_SUBMITTED) continue;
            vep_frame_desc d{};
            uint8_t *pl[3];
            plane_ptrs(s, sl.d_in, pl, d.stride);
            d.plane[0] = pl[0]; d.plane[1] = pl[1]; d.plane[2] = pl[2];
            if (sl.external)      // the kernel reads rows with the caller's linesize where the plane travels as one linear copy
                for (int p = 0; p < 3; ++p) if (pl[p] && external_plane_is_linear(s, sl, p)) d.stride[p] = sl.ext_stride[p];
            d.dst = sl.d_out; d.dst_stride = s->pitch_out; d.width = s->w; d.height = s->h; d.fmt = s->fmt; d.convert = 1;
            descs.push_back(d);
            sl.state = SLOT_INFLIGHT;
            t.slots.push_back(SlotRef{s, i});
        }
    }
    if (t.slots.empty()) return VEP_OK;
    // Copies are issued per RUN: frames whose slots are neighbours in memory (same group, same slot index, consecutive
    // members — a fleet ticking in lock step) travel as one copy per direction; everything else one copy per frame.
    std::vector<int> order(t.slots.size());
    for (size_t i = 0; i < order.size(); ++i) order[i] = (int)i;
    std::sort(order.begin(), order.end(), [&t](int a, int b) {
        const SlotRef &x = t.slots[a], &y = t.slots[b];
        if (x.s->grp != y.s->grp) return x.s->grp < y.s->grp;
        if (x.slot != y.slot) return x.slot < y.slot;
        return x.s->member < y.s->member;
    });
    struct Run { int first, last; };   // positions in `order`
    std::vector<Run> in_runs, out_runs;
    auto neighbours = [&t, &order](int a, int b) {
        const SlotRef &x = t.slots[order[a]], &y = t.slots[order[b]];
        return x.s->grp == y.s->grp && x.slot == y.slot && y.s->member == x.s->member + 1;
    };
    for (int i = 0; i < (int)order.size(); ++i) {
        const SlotRef &r = t.slots[order[i]];
        const bool ext = r.s->slots[r.slot].external;
        const int lim = c->run_limit > 0 ? c->run_limit : (int)order.size();
        if (!in_runs.empty() && !ext && neighbours(i - 1, i) && !t.slots[order[i - 1]].s->slots[t.slots[order[i - 1]].slot].external &&
            in_runs.back().last == i - 1 && i - in_runs.back().first < lim) in_runs.back().last = i;
        else in_runs.push_back(Run{i, i});
        if (!out_runs.empty() && !r.s->padded_out && neighbours(i - 1, i) && !t.slots[order[i - 1]].s->padded_out &&
            out_runs.back().last == i - 1 && i - out_runs.back().first < lim) out_runs.back().last = i;
        else out_runs.push_back(Run{i, i});
    }
    // 2. H2D, one copy per run on the input stream of the run's first camera
    for (const Run &run : in_runs) {
        const SlotRef &a = t.slots[order[run.first]], &b = t.slots[order[run.last]];
        Stream *s = a.s;
        Slot &sa = a.s->slots[a.slot], &sb = b.s->slots[b.slot];
        if (!sa.external) {
            const size_t bytes = (size_t)(sb.h_in - sa.h_in) + b.s->in_bytes;
            CUDA_TRY(cudaMemcpyAsync(sa.d_in, sa.h_in, bytes, cudaMemcpyHostToDevice, s->cs_in));
            c->stats.h2d_bytes += (int64_t)bytes;
        } else {
            int rc = enqueue_external_h2d(c, s, sa);
            if (rc) return rc;
        }
        CUDA_TRY(cudaEventRecord(s->ev_in, s->cs_in));
        CUDA_TRY(cudaStreamWaitEvent(c->main, s->ev_in, 0));
    }
    // 3. one batched launch for the whole tick.  Its inputs come from the copies above and its outputs are slots no
    //    earlier tick still owns, so it needs no ordering against the previous tick's kernel (VEP_ORDER_NONE): with
    //    programmatic dependent launch it may start while that kernel drains.
    int rc = build_tables(descs.data(), (int)descs.size(), c->scratch);
    if (rc) return rc;
    CUDA_TRY(cudaEventRecord(t.k0, c->main));
    rc = upload_and_launch(c, c->scratch, c->main, VEP_ORDER_NONE);
    if (rc) return rc;
    CUDA_TRY(cudaEventRecord(t.k1, c->main));
    // 4. D2H, one copy per run on the output stream of the run's first camera: payloads land behind their wire prefixes
    //    (a merged copy also sweeps the header/trailer room between payloads: the framing is written when the tick
    //    retires).  A padded device pitch (width*3 not a multiple of 16) is removed by the copy engine on the way out.
    for (const Run &run : out_runs) {
        const SlotRef &a = t.slots[order[run.first]], &b = t.slots[order[run.last]];
        Stream *s = a.s;
        Slot &sa = a.s->slots[a.slot], &sb = b.s->slots[b.slot];
        CUDA_TRY(cudaStreamWaitEvent(s->cs_out, t.k1, 0));
        if (!s->padded_out) {
            const size_t bytes = (size_t)(sb.h_out - sa.h_out) + b.s->out_bytes;
            CUDA_TRY(cudaMemcpyAsync(sa.h_out + kHdrRoom, sa.d_out, bytes, cudaMemcpyDeviceToHost, s->cs_out));
        } else {
            CUDA_TRY(cudaMemcpy2DAsync(sa.h_out + kHdrRoom, (size_t)s->w * 3, sa.d_out, s->pitch_out, (size_t)s->w * 3, s->h, cudaMemcpyDeviceToHost, s->cs_out));
        }
        CUDA_TRY(cudaEventRecord(sa.done, s->cs_out));
        for (int i = run.first; i <= run.last; ++i) {
            const SlotRef &r = t.slots[order[i]];
            r.s->slots[r.slot].done_ref = sa.done;
            c->stats.d2h_bytes += (int64_t)r.s->out_bytes;
        }
    }
    t.active = true;
    t.ticket = c->next_ticket++;
    c->ticks_in_flight++;
    c->stats.ticks++;
    c->stats.copies_in += (int64_t)in_runs.size();
    c->stats.copies_out += (int64_t)out_runs.size();
    c->stats.frames_converted += (int64_t)t.slots.size();
    if (n_frames) *n_frames = (int)t.slots.size();
    return VEP_OK;
}

// Retire ticks, oldest first: only the oldest one (only_oldest), or all up to and including ticket `upto`.  The event
// waits run without the context lock, so producers keep staging and other threads keep beginning ticks meanwhile;
// when two threads wait for the same tick, the one that finds it already retired simply moves on.
static int retire_ticks(vep_ctx *c, uint64_t upto, bool only_oldest) {
    DeviceGuard g(c->device);
    for (;;) {
        std::vector<cudaEvent_t> evs;
        uint64_t id = 0;
        {
            std::lock_guard<std::mutex> lk(c->mu);
            if (!c->ticks_in_flight) return VEP_OK;
            TickRec &t = c->ticks[c->tick_head];
            if (!only_oldest && t.ticket > upto) return VEP_OK;
            id = t.ticket;
            for (auto &r : t.slots) evs.push_back(r.s->slots[r.slot].done_ref);
        }
        for (cudaEvent_t e : evs) CUDA_TRY(cudaEventSynchronize(e));
        {
            std::lock_guard<std::mutex> lk(c->mu);
            TickRec &t = c->ticks[c->tick_head];
            if (c->ticks_in_flight && t.ticket == id) {     // else another waiter retired it meanwhile
                float ms = 0.f;
                if (cudaEventElapsedTime(&ms, t.k0, t.k1) == cudaSuccess) c->stats.last_kernel_ms = ms;
                for (auto &r : t.slots) {
                    write_framing(r.s, r.s->slots[r.slot]);      // after the copy out: a merged copy sweeps the framing room
                    std::lock_guard<std::mutex> sl_lk(r.s->mu);
                    r.s->slots[r.slot].state = SLOT_READY;
                    r.s->newest = r.slot;
                    r.s->n_ready++;
                }
                t.slots.clear();
                t.active = false;
                c->retired_ticket = id;
                c->tick_head = (c->tick_head + 1) % VEP_MAX_TICKS_IN_FLIGHT;
                c->ticks_in_flight--;
            }
        }
        if (only_oldest) return VEP_OK;
    }
}

extern "C" int vep_tick_wait(vep_ctx *c) {
    if (!c) return fail(VEP_ERR_INVALID_ARG, "ctx is null");
    return retire_ticks(c, 0, true);
}
extern "C" int vep_tick_wait_ticket(vep_ctx *c, uint64_t ticket) {
    if (!c) return fail(VEP_ERR_INVALID_ARG, "ctx is null");
    return retire_ticks(c, ticket, false);
}

extern "C" int vep_tick(vep_ctx *c, int *n_frames) {
    uint64_t ticket = 0;
    int rc = vep_tick_begin_ticket(c, n_frames, &ticket);
    if (rc) return rc;
    return vep_tick_wait_ticket(c, ticket);   // waits for the tick it began (and any older one), whoever else ticks
}

static int ready_slot(vep_ctx *c, int stream_id, int age, Stream **so, Slot **slo) {
    if (!c) return fail(VEP_ERR_INVALID_ARG, "ctx is null");
    Stream *s = find_stream(c, stream_id);
    if (!s) return fail(VEP_ERR_INVALID_ARG, "stream %d not open", stream_id);
    std::lock_guard<std::mutex> lk(s->mu);
    if (age < 0) return fail(VEP_ERR_INVALID_ARG, "age %d < 0", age);
    if (s->newest < 0 || age >= s->n_ready)
        return fail(VEP_ERR_NO_FRAME, "stream %d: no converted frame of age %d", stream_id, age);
    const int ns = (int)s->slots.size();
    Slot &sl = s->slots[((s->newest - age) % ns + ns) % ns];
    if (sl.state != SLOT_READY) return fail(VEP_ERR_NO_FRAME, "stream %d: frame of age %d was recycled", stream_id, age);
    *so = s;
    *slo = &sl;
    return VEP_OK;
}

extern "C" int vep_result_at(vep_ctx *c, int stream_id, int age, const uint8_t **bgr, size_t *len, int32_t shape[3]) {
    Stream *s = nullptr; Slot *sl = nullptr;
    int rc = ready_slot(c, stream_id, age, &s, &sl);
    if (rc) return rc;
    if (bgr) *bgr = sl->h_out + kHdrRoom;
    if (len) *len = s->out_bytes;
    if (shape) { shape[0] = s->h; shape[1] = s->w; shape[2] = 3; }
    return VEP_OK;
}
extern "C" int vep_result(vep_ctx *c, int stream_id, const uint8_t **bgr, size_t *len, int32_t shape[3]) {
    return vep_result_at(c, stream_id, 0, bgr, len, shape);
}

extern "C" int vep_result_packed_at(vep_ctx *c, int stream_id, int age, const uint8_t **msg, size_t *len) {
    Stream *s = nullptr; Slot *sl = nullptr;
    int rc = ready_slot(c, stream_id, age, &s, &sl);
    if (rc) return rc;
    if (msg) *msg = sl->h_out + sl->msg_off;
    if (len) *len = sl->msg_len;
    return VEP_OK;
}
extern "C" int vep_result_packed(vep_ctx *c, int stream_id, const uint8_t **msg, size_t *len) {
    return vep_result_packed_at(c, stream_id, 0, msg, len);
}

extern "C" int vep_result_device_pitched(vep_ctx *c, int stream_id, const uint8_t **bgr_dev, int32_t shape[3], int32_t *row_pitch) {
    Stream *s = nullptr; Slot *sl = nullptr;
    int rc = ready_slot(c, stream_id, 0, &s, &sl);
    if (rc) return rc;
    if (bgr_dev) *bgr_dev = sl->d_out;
    if (shape) { shape[0] = s->h; shape[1] = s->w; shape[2] = 3; }
    if (row_pitch) *row_pitch = s->pitch_out;
    return VEP_OK;
}
extern "C" int vep_result_device(vep_ctx *c, int stream_id, const uint8_t **bgr_dev, int32_t shape[3]) {
    int32_t pitch = 0, shp[3] = {0, 0, 0};
    int rc = vep_result_device_pitched(c, stream_id, bgr_dev, shp, &pitch);
    if (rc) return rc;
    if (pitch != 3 * shp[1])
        return fail(VEP_ERR_GEOMETRY, "stream %d: device rows are padded to %d bytes (width*3 = %d is not a multiple of 16); use vep_result_device_pitched",
                    stream_id, pitch, 3 * shp[1]);
    if (shape) { shape[0] = shp[0]; shape[1] = shp[1]; shape[2] = shp[2]; }
    return VEP_OK;
}

extern "C" int vep_get_stats(vep_ctx *c, vep_stats *out) {
    if (!c || !out) return fail(VEP_ERR_INVALID_ARG, "null argument");
    std::lock_guard<std::mutex> lk(c->mu);
    *out = c->stats;
    return VEP_OK;
}

// ------------------------------------------------------------------------------------------------
// vep_convert_batch_noalloc — batched conversion for callers that must not allocate device memory.
//
// Who needs this: a producer that runs with a fixed device-memory budget and may not call
// cudaMalloc once it is streaming — a GPU slice or container whose memory was handed out in full
// at start-up, a host application (an inference server, a torch process) whose own pool already
// holds nearly all of HBM so that a plugin's cudaMalloc can fail although the pool has room, or
// a real-time loop that cannot afford cudaMalloc's device-wide synchronisation in the middle of
// a tick.  The other entry points keep their descriptor tables in cudaMalloc'ed buffers (plans,
// the per-context upload ring); this one keeps them in the fixed array below, which the CUDA
// module carries itself.
//
// The kernel is the table-driven vep::yuv_to_bgr24_batched; only the home of its two tables
// (FrameDev[], CtaWork[]) differs.
// ------------------------------------------------------------------------------------------------
static constexpr size_t kNoallocFramesBytes = (size_t)(VEP_NOALLOC_MAX_FRAMES + 1) * sizeof(FrameDev);
static constexpr size_t kNoallocWorkOffset = (kNoallocFramesBytes + 255) / 256 * 256;
static constexpr size_t kNoallocScratchBytes = kNoallocWorkOffset + (size_t)VEP_NOALLOC_MAX_CTAS * sizeof(CtaWork);

// Device-resident scratch: one copy per GPU that loads this library, held until the process exits
// (545,024 bytes).  In HOST code this name is only a handle for cudaGetSymbolAddress — it is not a
// device address and must never be handed to a kernel or a memcpy as one.
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoallocScratchBytes];

static std::mutex g_noalloc_mu;
static void *g_noalloc_dev_ptr[64] = {};          // per device ordinal: address of g_noalloc_scratch on that GPU
// Fixed host staging (no heap use on this path).  These two arrays are PAGEABLE on purpose: cudaMemcpyAsync from
// pageable memory returns only after the runtime has taken its own copy of the source (for larger tables it first
// waits for earlier work on the stream), which is what makes rewriting them on the next call safe.  Do not
// cudaHostRegister them: from pinned memory the copy would read them later, and an event would have to guard re-use.
static FrameDev g_noalloc_h_frames[VEP_NOALLOC_MAX_FRAMES + 1];
static CtaWork g_noalloc_h_work[VEP_NOALLOC_MAX_CTAS];

// Validation + table row for one descriptor; mirrors build_tables() without touching the heap.
static int noalloc_frame(const vep_frame_desc &s, int i, FrameDev &f) {
    if (s.fmt < VEP_FMT_YUV420P || s.fmt > VEP_FMT_GRAY8) return fail(VEP_ERR_INVALID_ARG, "desc %d: unknown fmt %d", i, s.fmt);
    const int w = s.width, h = s.height;
    if (w <= 0 || h <= 0 || w > kMaxDim || h > kMaxDim || (w & 1) || (h & 1))
        return fail(VEP_ERR_GEOMETRY, "desc %d: %dx%d — 4:2:0 needs even, positive dimensions <= %d", i, w, h, kMaxDim);
    const bool nv12 = vep::fmt_is_semiplanar(s.fmt);
    const int bps = vep::fmt_bytes_per_sample(s.fmt);
    if ((nv12 || bps == 2) && h < 8) return fail(VEP_ERR_GEOMETRY, "desc %d: nv12/nv21/10-bit height %d < 8", i, h);
    const bool gray = s.fmt == VEP_FMT_GRAY8;
    if (!s.plane[0] || (!gray && (!s.plane[1] || (!nv12 && !s.plane[2]))) || !s.dst) return fail(VEP_ERR_INVALID_ARG, "desc %d: null plane or dst", i);
    const int cw = ((nv12 || vep::fmt_is_444(s.fmt)) ? w : w / 2) * bps;
    if (s.stride[0] < w * bps || (!gray && (s.stride[1] < cw || (!nv12 && s.stride[2] < cw))) || s.dst_stride < 3 * w)
        return fail(VEP_ERR_GEOMETRY, "desc %d: a stride is smaller than its row", i);
    if (bps == 2 && ((s.stride[0] | s.stride[1] | s.stride[2]) & 1 || !aligned(s.plane[0], 2) || !aligned(s.plane[1], 2) || !aligned(s.plane[2], 2)))
        return fail(VEP_ERR_GEOMETRY, "desc %d: 16-bit samples need 2-byte aligned planes and even strides", i);
    f = FrameDev{};
    f.y = s.plane[0]; f.c0 = gray ? s.plane[0] : s.plane[1]; f.c1 = (nv12 || gray) ? nullptr : s.plane[2]; f.dst = s.dst;
    f.sy = s.stride[0]; f.sc0 = gray ? s.stride[0] : s.stride[1]; f.sc1 = (nv12 || gray) ? 0 : s.stride[2]; f.sd = s.dst_stride;
    f.w = w; f.h = h; f.fmt = s.fmt;
    const size_t ca = nv12 ? 16 : 8;
    const bool vec = aligned(f.y, 16) && f.sy % 16 == 0 && aligned(f.c0, ca) && f.sc0 % ca == 0 &&
                     (nv12 || gray || (aligned(f.c1, 8) && f.sc1 % 8 == 0)) && aligned(f.dst, 16) && f.sd % 16 == 0;
    f.fast = vec ? vep::PATH_VECTOR : vep::PATH_UNALIGNED;   // the table-driven kernel: aligned, or assembled from aligned chunks
    if (vep::fmt_is_422(s.fmt) || vep::fmt_vector_only_when_specialised(s.fmt) || bps == 2) f.fast = vep::PATH_GENERIC;   // ... 4:2:2 / 4:4:4 / gray / 10-bit on the scalar one
    f.fast_tbl = f.fast;
    f.segs = (w + vep::kSegPx - 1) / vep::kSegPx;
    return VEP_OK;
}

extern "C" int vep_convert_batch_noalloc(const vep_frame_desc *descs, int n, void *cuda_stream) {
    if (n < 0 || (n > 0 && !descs)) return fail(VEP_ERR_INVALID_ARG, "descs is null or n < 0");
    int dev = -1;
    CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return fail(VEP_ERR_INVALID_ARG, "device ordinal %d not supported", dev);
    std::lock_guard<std::mutex> lk(g_noalloc_mu);

    // ---- set-up, once per device, with that device current: resolve the scratch array's address there ----
    if (!g_noalloc_dev_ptr[dev]) {
        cudaDeviceProp prop;
        CUDA_TRY(cudaGetDeviceProperties(&prop, dev));
        if (prop.major < 10) return fail(VEP_ERR_CUDA, "device %d is sm_%d%d; this build carries sm_100a code only", dev, prop.major, prop.minor);
        void *ptr = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&ptr, g_noalloc_scratch);
        if (e != cudaSuccess || !ptr) return fail(VEP_ERR_CUDA, "cudaGetSymbolAddress(scratch) on device %d -> %s", dev, cudaGetErrorString(e));
        g_noalloc_dev_ptr[dev] = ptr;
    }
    uint8_t *scratch = static_cast<uint8_t *>(g_noalloc_dev_ptr[dev]);

    // ---- tables into the fixed host staging, refusing anything beyond the stated maxima ----
    int nf = 0, nw = 0;
    for (int i = 0; i < n; ++i) {
        if (!descs[i].convert) continue;
        if (nf >= VEP_NOALLOC_MAX_FRAMES)
            return fail(VEP_ERR_INVALID_ARG, "more than VEP_NOALLOC_MAX_FRAMES (%d) frames to convert in one call", VEP_NOALLOC_MAX_FRAMES);
        FrameDev &f = g_noalloc_h_frames[nf];
        int rc = noalloc_frame(descs[i], i, f);
        if (rc) return rc;
        const int units = f.fast != vep::PATH_GENERIC ? f.segs * (f.h / 2) : (f.w / 2) * (f.h / 2);
        const int per_cta = f.fast != vep::PATH_GENERIC ? vep::kWarpsPerCta : vep::kGenericBlocksPerCta;
        const int ctas = (units + per_cta - 1) / per_cta;
        if (nw + ctas > VEP_NOALLOC_MAX_CTAS)
            return fail(VEP_ERR_INVALID_ARG, "batch needs more than VEP_NOALLOC_MAX_CTAS (%d) thread blocks (about %d 1080p frames on the vector path, %d on the scalar one); split the call",
                        VEP_NOALLOC_MAX_CTAS, VEP_NOALLOC_MAX_CTAS / 270, VEP_NOALLOC_MAX_CTAS / 507);
        for (int k = 0; k < ctas; ++k) g_noalloc_h_work[nw++] = CtaWork{nf, k * per_cta};
        ++nf;
    }
    if (nf == 0) return VEP_OK;

    // ---- both tables are rewritten in full for the entries this launch reads; nothing in the scratch has to
    //      start at zero, and stale rows of an earlier, larger call are never indexed ----
    cudaStream_t st = (cudaStream_t)cuda_stream;
    CUDA_TRY(cudaMemcpyAsync(scratch, g_noalloc_h_frames, (size_t)nf * sizeof(FrameDev), cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaMemcpyAsync(scratch + kNoallocWorkOffset, g_noalloc_h_work, (size_t)nw * sizeof(CtaWork), cudaMemcpyHostToDevice, st));
    vep::yuv_to_bgr24_batched<<<nw, vep::kThreads, 0, st>>>(reinterpret_cast<const FrameDev *>(scratch),
                                                            reinterpret_cast<const CtaWork *>(scratch + kNoallocWorkOffset));
    CUDA_TRY(cudaGetLastError());
    return VEP_OK;
}
