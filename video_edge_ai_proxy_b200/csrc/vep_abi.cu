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
#include <cstdarg>
#include <cstddef>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <unordered_map>
#include <vector>

static_assert(sizeof(vep_frame_desc) == 72 && offsetof(vep_frame_desc, dst) == 40 && offsetof(vep_frame_desc, width) == 52,
              "vep_frame_desc layout is part of the ABI (mirrored in _abi.py)");
static_assert(sizeof(vep_frame_meta) == 64 && offsetof(vep_frame_meta, frame_type) == 56, "vep_frame_meta layout is part of the ABI");
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
    int total_items = 0;            // item space of the persistent kernel
    std::vector<int32_t> tick_cta_begin;   // tick kernel: first CTA of each frame, then the grid size
};

static inline bool aligned(const void *p, size_t a) { return (reinterpret_cast<uintptr_t>(p) % a) == 0; }

static int build_tables(const vep_frame_desc *d, int n, Tables &t, int rpt = 1, bool vec444_ok = false) {
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
        const size_t ca = (nv12 || bps == 2 || vep::fmt_is_444(s.fmt)) ? 16 : 8;
        const bool vec = (w % 16 == 0) && aligned(f.y, 16) && f.sy % 16 == 0 && aligned(f.c0, ca) && f.sc0 % ca == 0 &&
                         (nv12 || gray || (aligned(f.c1, ca) && f.sc1 % ca == 0)) && aligned(f.dst, 16) && f.sd % 16 == 0;
        const bool bulk = vec && bps == 1 && (w % 32 == 0) && aligned(f.c0, 16) && f.sc0 % 16 == 0 && (nv12 || (aligned(f.c1, 16) && f.sc1 % 16 == 0));
        f.fast = bulk ? vep::PATH_BULK : vec ? vep::PATH_VECTOR : vep::PATH_GENERIC;
        if (vep::fmt_vector_only_when_specialised(s.fmt))
            f.fast = (f.fast != vep::PATH_GENERIC && homogeneous && vec444_ok) ? vep::PATH_VECTOR : vep::PATH_GENERIC;
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
};

struct Stream {
    int id = 0, fmt = 0, w = 0, h = 0, ring = 1;
    size_t off_c0 = 0, off_c1 = 0, in_bytes = 0, out_bytes = 0, in_slot = 0, out_slot = 0;
    std::vector<Slot> slots;
    int wr = 0;             // next slot to fill
    int newest = -1;        // newest READY slot
    int n_ready = 0;
    uint8_t *h_base = nullptr, *h_in_base = nullptr, *d_base = nullptr;   // pinned out slots, pinned in slots, device slots
    cudaStream_t cs_in = nullptr, cs_out = nullptr;
    cudaEvent_t ev_in = nullptr;
    std::mutex mu;
};

struct SlotRef { Stream *s; int slot; };

struct TickRec {
    std::vector<SlotRef> slots;
    cudaEvent_t k0 = nullptr, k1 = nullptr;
    bool active = false;
};

struct vep_ctx {
    int device = 0;
    int kernel = 2;              // 2: tick kernel, table in kernel params (default); 1: table-driven one-shot
                                 // (VEP_KERNEL=oneshot; also the fallback beyond 128 frames); 0: VEP_KERNEL=persistent
    int rpt = 2, minb = 4;       // tick kernel variant: row pairs per lane, min CTAs/SM (VEP_RPT, VEP_MINB)
    int pdl = 1;                 // programmatic dependent launch attribute on tick launches (VEP_PDL=0 disables)
    int specialize = 1;          // use the per-format instantiation for single-format launches (VEP_SPECIALIZE=0 disables)
    int store_mode = 0;          // one-shot kernel only (VEP_STORE_MODE)
    int resident_warps = 0;      // persistent kernel: warps the whole GPU keeps resident
    cudaStream_t main = nullptr;
    TableBuf tb[4];
    int tb_next = 0;
    std::mutex mu;                                  // guards streams map, ticks, table buffers
    std::unordered_map<int, Stream *> streams;
    TickRec ticks[VEP_MAX_TICKS_IN_FLIGHT];
    int tick_head = 0, ticks_in_flight = 0;         // oldest in flight = tick_head
    vep_stats stats{};
    Tables scratch;
    std::vector<uint8_t> tick_blob;
};

// 4:4:4 frames may take the vector path only where a format-specialised tick kernel will run them
static inline bool vec444_ok(const vep_ctx *c) { return c->kernel == 2 && c->specialize && c->rpt == 2 && c->minb == 4; }

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

// grid of the persistent kernel: every warp gets the same number of items (no partial last round)
static int persistent_ctas(const vep_ctx *c, int total_items) {
    const int rounds = (total_items + c->resident_warps - 1) / c->resident_warps;
    const int warps = (total_items + rounds - 1) / rounds;
    return (warps + vep::kPWarps - 1) / vep::kPWarps;
}

// ---- tick kernel: descriptor table travels as a kernel parameter -------------------------------------
static constexpr int kTickCapSmall = 16, kTickCapLarge = 128;

template <int NF>
static void fill_tick(const Tables &t, std::vector<uint8_t> &blob) {
    blob.assign(sizeof(vep::TickParams<NF>), 0);
    auto *P = reinterpret_cast<vep::TickParams<NF> *>(blob.data());
    const int n = t.n_converted;
    P->n_frames = n;
    P->wait_prior = 1;
    int per = n ? t.tick_cta_begin[1] - t.tick_cta_begin[0] : 0;
    for (int i = 1; i < n && per > 0; ++i) if (t.tick_cta_begin[i + 1] - t.tick_cta_begin[i] != per) per = 0;
    P->uniform_ctas = (n > 0 && n <= 65535) ? per : 0;
    int only = n ? t.frames[0].fmt : -1;
    for (int i = 1; i < n; ++i) if (t.frames[i].fmt != only) only = -1;
    P->only_fmt = only;        // host-side hint: all frames share one pixel format -> format-specialised kernel
    for (int i = 0; i <= n; ++i) P->cta_begin[i] = t.tick_cta_begin[i];
    for (int i = 0; i < n; ++i) P->frames[i] = t.frames[i];
}
// returns capacity used (0: does not fit -> caller falls back to the table-driven kernel)
static int make_tick_blob(const Tables &t, std::vector<uint8_t> &blob, int *grid) {
    const int n = t.n_converted;
    *grid = n ? t.tick_cta_begin[n] : 0;
    if (n <= kTickCapSmall) { fill_tick<kTickCapSmall>(t, blob); return kTickCapSmall; }
    if (n <= kTickCapLarge) { fill_tick<kTickCapLarge>(t, blob); return kTickCapLarge; }
    return 0;
}
template <int NF>
static const void *tick_kernel(int rpt, int minb, int only_fmt) {
    if (rpt == 2 && minb == 4) {   // the default variant also exists specialised per pixel format
        switch (only_fmt) {
            case VEP_FMT_YUV420P: return (const void *)vep::yuv_to_bgr24_tick<NF, 2, 4, vep::FMT_YUV420P>;
            case VEP_FMT_YUVJ420P: return (const void *)vep::yuv_to_bgr24_tick<NF, 2, 4, vep::FMT_YUVJ420P>;
            case VEP_FMT_NV12: return (const void *)vep::yuv_to_bgr24_tick<NF, 2, 4, vep::FMT_NV12>;
            case VEP_FMT_NV21: return (const void *)vep::yuv_to_bgr24_tick<NF, 2, 4, vep::FMT_NV21>;
            case VEP_FMT_YUV422P: return (const void *)vep::yuv_to_bgr24_tick<NF, 2, 4, vep::FMT_YUV422P>;
            case VEP_FMT_YUVJ422P: return (const void *)vep::yuv_to_bgr24_tick<NF, 2, 4, vep::FMT_YUVJ422P>;
            case VEP_FMT_YUV420P10LE: return (const void *)vep::yuv_to_bgr24_tick<NF, 2, 3, vep::FMT_YUV420P10>;   // 80 registers: no spills
            case VEP_FMT_YUV444P: return (const void *)vep::yuv_to_bgr24_tick<NF, 2, 3, vep::FMT_YUV444P>;
            case VEP_FMT_YUVJ444P: return (const void *)vep::yuv_to_bgr24_tick<NF, 2, 3, vep::FMT_YUVJ444P>;
            case VEP_FMT_GRAY8: return (const void *)vep::yuv_to_bgr24_tick<NF, 2, 4, vep::FMT_GRAY8>;
            default: break;
        }
    }
    if (rpt == 2) return minb >= 4 ? (const void *)vep::yuv_to_bgr24_tick<NF, 2, 4> : (const void *)vep::yuv_to_bgr24_tick<NF, 2, 3>;
    if (minb >= 8) return (const void *)vep::yuv_to_bgr24_tick<NF, 1, 8>;
    if (minb >= 6) return (const void *)vep::yuv_to_bgr24_tick<NF, 1, 6>;
    return (const void *)vep::yuv_to_bgr24_tick<NF, 1, 5>;
}
static int launch_tick(const vep_ctx *c, int cap, const std::vector<uint8_t> &blob, int grid, cudaStream_t st) {
    if (grid <= 0) return VEP_OK;
    const int32_t *hdr0 = reinterpret_cast<const int32_t *>(blob.data());   // n_frames, wait_prior, uniform_ctas, only_fmt
    const int only = c->specialize ? hdr0[3] : -1;
    const void *fn = cap == kTickCapSmall ? tick_kernel<kTickCapSmall>(c->rpt, c->minb, only) : tick_kernel<kTickCapLarge>(c->rpt, c->minb, only);
    void *args[1] = {const_cast<uint8_t *>(blob.data())};
    cudaLaunchConfig_t cfg{};
    const int32_t *hdr = reinterpret_cast<const int32_t *>(blob.data());   // n_frames, wait_prior, uniform_ctas
    cfg.gridDim = hdr[2] > 0 ? dim3(hdr[2], hdr[0]) : dim3(grid);
    cfg.blockDim = dim3(vep::kThreads);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = c->pdl ? 1 : 0;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    CUDA_TRY(cudaLaunchKernelExC(&cfg, fn, args));
    return VEP_OK;
}

static int launch_tables(const vep_ctx *c, const FrameDev *d_frames, const CtaWork *d_work, int n_ctas, int total_items, cudaStream_t st) {
    if (total_items <= 0) return VEP_OK;
    if (c->kernel == 0) {
        vep::yuv_to_bgr24_persistent<<<persistent_ctas(c, total_items), vep::kPThreads, vep::kPSmemBytes, st>>>(d_frames, total_items);
    } else {
        switch (c->store_mode) {
            case 1: vep::yuv_to_bgr24_batched<1><<<n_ctas, vep::kThreads, 0, st>>>(d_frames, d_work); break;
            case 2: vep::yuv_to_bgr24_batched<2><<<n_ctas, vep::kThreads, 0, st>>>(d_frames, d_work); break;
            default: vep::yuv_to_bgr24_batched<0><<<n_ctas, vep::kThreads, 0, st>>>(d_frames, d_work); break;
        }
    }
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
    const char *sm = getenv("VEP_STORE_MODE");
    c->store_mode = sm ? atoi(sm) : 0;
    if (c->store_mode < 0 || c->store_mode > 2) c->store_mode = 0;
    const char *kn = getenv("VEP_KERNEL");
    c->kernel = !kn ? 2 : strcmp(kn, "persistent") == 0 ? 0 : strcmp(kn, "oneshot") == 0 ? 1 : 2;
    const char *er = getenv("VEP_RPT"), *em = getenv("VEP_MINB");
    c->rpt = (er && atoi(er) == 1) ? 1 : 2;
    c->minb = em ? atoi(em) : (c->rpt == 2 ? 4 : 6);
    const char *es = getenv("VEP_SPECIALIZE");
    c->specialize = (es && atoi(es) == 0) ? 0 : 1;
    const char *ep = getenv("VEP_PDL");
    c->pdl = (ep && atoi(ep) == 0) ? 0 : 1;
    CUDA_TRY(cudaFuncSetAttribute(vep::yuv_to_bgr24_persistent, cudaFuncAttributeMaxDynamicSharedMemorySize, vep::kPSmemBytes));
    CUDA_TRY(cudaFuncSetAttribute(vep::yuv_to_bgr24_persistent, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    int per_sm = 0;
    CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, vep::yuv_to_bgr24_persistent, vep::kPThreads, vep::kPSmemBytes));
    if (per_sm < 1) return fail(VEP_ERR_CUDA, "persistent kernel does not fit on an SM");
    const char *occ = getenv("VEP_CTAS_PER_SM");
    if (occ && atoi(occ) >= 1 && atoi(occ) < per_sm) per_sm = atoi(occ);
    c->resident_warps = per_sm * prop.multiProcessorCount * vep::kPWarps;
    CUDA_TRY(cudaStreamCreateWithFlags(&c->main, cudaStreamNonBlocking));
    for (auto &t : c->tb) CUDA_TRY(cudaEventCreateWithFlags(&t.used, cudaEventDisableTiming));
    for (auto &t : c->ticks) { CUDA_TRY(cudaEventCreate(&t.k0)); CUDA_TRY(cudaEventCreate(&t.k1)); }
    *out = c;
    return VEP_OK;
}

static void free_stream(Stream *s) {
    if (!s) return;
    for (auto &sl : s->slots) if (sl.done) cudaEventDestroy(sl.done);
    if (s->cs_in) { cudaStreamSynchronize(s->cs_in); cudaStreamDestroy(s->cs_in); }
    if (s->cs_out) { cudaStreamSynchronize(s->cs_out); cudaStreamDestroy(s->cs_out); }
    if (s->ev_in) cudaEventDestroy(s->ev_in);
    if (s->h_base) cudaFreeHost(s->h_base);
    if (s->h_in_base) cudaFreeHost(s->h_in_base);
    if (s->d_base) cudaFree(s->d_base);
    delete s;
}

extern "C" void vep_destroy(vep_ctx *c) {
    if (!c) return;
    DeviceGuard g(c->device);
    if (c->main) cudaStreamSynchronize(c->main);
    for (auto &kv : c->streams) free_stream(kv.second);
    for (auto &t : c->tb) { if (t.h) cudaFreeHost(t.h); if (t.d) cudaFree(t.d); if (t.used) cudaEventDestroy(t.used); }
    for (auto &t : c->ticks) { if (t.k0) cudaEventDestroy(t.k0); if (t.k1) cudaEventDestroy(t.k1); }
    if (c->main) cudaStreamDestroy(c->main);
    delete c;
}

// upload `t` through the next ring buffer and launch on `st`
static int upload_and_launch(vep_ctx *c, const Tables &t, cudaStream_t st) {
    if (t.total_items == 0) return VEP_OK;
    if (c->kernel == 2) {
        int grid = 0;
        const int cap = make_tick_blob(t, c->tick_blob, &grid);
        if (cap) {   // cudaLaunchKernel copies the parameter block before returning: the blob is reusable at once
            int rc = launch_tick(c, cap, c->tick_blob, grid, st);
            if (rc) return rc;
            c->stats.kernel_launches++;
            return VEP_OK;
        }
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
    int rc = launch_tables(c, (const FrameDev *)b.d, (const CtaWork *)((char *)b.d + fb), (int)t.work.size(), t.total_items, st);
    if (rc) return rc;
    CUDA_TRY(cudaEventRecord(b.used, st));
    b.pending = true;
    c->stats.kernel_launches++;
    return VEP_OK;
}

extern "C" int vep_convert_batch(vep_ctx *c, const vep_frame_desc *descs, int n, void *cuda_stream) {
    if (!c) return fail(VEP_ERR_INVALID_ARG, "ctx is null");
    DeviceGuard g(c->device);
    std::lock_guard<std::mutex> lk(c->mu);
    int rc = build_tables(descs, n, c->scratch, c->rpt, vec444_ok(c));
    if (rc) return rc;
    rc = upload_and_launch(c, c->scratch, (cudaStream_t)cuda_stream);
    if (rc) return rc;
    c->stats.frames_converted += c->scratch.n_converted;
    return VEP_OK;
}

extern "C" int vep_plan_create(vep_ctx *c, const vep_frame_desc *descs, int n, vep_plan **out) {
    if (!c || !out) return fail(VEP_ERR_INVALID_ARG, "ctx or out is null");
    *out = nullptr;
    DeviceGuard g(c->device);
    Tables t;
    int rc = build_tables(descs, n, t, c->rpt, vec444_ok(c));
    if (rc) return rc;
    vep_plan *p = new (std::nothrow) vep_plan();
    if (!p) return fail(VEP_ERR_NOMEM, "out of host memory");
    p->ctx = c;
    p->n_ctas = c->kernel == 0 ? (t.total_items ? persistent_ctas(c, t.total_items) : 0) : (int)t.work.size();
    p->total_items = t.total_items;
    p->n_converted = t.n_converted;
    if (c->kernel == 2) {
        p->tick_cap = make_tick_blob(t, p->tick_blob, &p->tick_grid);
        if (p->tick_cap) {
            p->n_ctas = p->tick_grid;
            p->tick_blob_overlap = p->tick_blob;
            reinterpret_cast<int32_t *>(p->tick_blob_overlap.data())[1] = 0;   // TickParams::wait_prior
            p->tick_blob_writes = p->tick_blob;
            reinterpret_cast<int32_t *>(p->tick_blob_writes.data())[1] = 2;
        }
        else p->n_ctas = (int)t.work.size();
    }
    p->alg_bytes = t.alg_bytes;
    if (t.total_items) {
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

static int plan_launch(vep_plan *p, void *cuda_stream, int order) {
    if (!p) return fail(VEP_ERR_INVALID_ARG, "plan is null");
    if (order < VEP_ORDER_STRICT || order > VEP_ORDER_NONE) return fail(VEP_ERR_INVALID_ARG, "unknown launch order %d", order);
    DeviceGuard g(p->ctx->device);
    const std::vector<uint8_t> &blob = order == VEP_ORDER_NONE ? p->tick_blob_overlap : order == VEP_ORDER_WRITES ? p->tick_blob_writes : p->tick_blob;
    int rc = p->tick_cap ? launch_tick(p->ctx, p->tick_cap, blob, p->tick_grid, (cudaStream_t)cuda_stream)
                         : launch_tables(p->ctx, p->d_frames, p->d_work, p->n_ctas, p->total_items, (cudaStream_t)cuda_stream);
    if (rc) return rc;
    p->ctx->stats.kernel_launches++;
    p->ctx->stats.frames_converted += p->n_converted;
    return VEP_OK;
}
extern "C" int vep_plan_launch(vep_plan *p, void *cuda_stream) { return plan_launch(p, cuda_stream, VEP_ORDER_STRICT); }
extern "C" int vep_plan_launch_overlapped(vep_plan *p, void *cuda_stream) { return plan_launch(p, cuda_stream, VEP_ORDER_NONE); }
extern "C" int vep_plan_launch_ordered(vep_plan *p, void *cuda_stream, int order) { return plan_launch(p, cuda_stream, order); }

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
    std::lock_guard<std::mutex> lk(c->mu);
    auto it = c->streams.find(id);
    return it == c->streams.end() ? nullptr : it->second;
}

extern "C" int vep_stream_open(vep_ctx *c, int stream_id, int fmt, int w, int h, int ring_depth) {
    if (!c) return fail(VEP_ERR_INVALID_ARG, "ctx is null");
    if (fmt < VEP_FMT_YUV420P || fmt > VEP_FMT_GRAY8) return fail(VEP_ERR_INVALID_ARG, "unknown fmt %d", fmt);
    if (w <= 0 || h <= 0 || w > kMaxDim || h > kMaxDim || (w & 1) || (h & 1)) return fail(VEP_ERR_GEOMETRY, "%dx%d: need even, positive dimensions", w, h);
    if ((vep::fmt_is_semiplanar(fmt) || vep::fmt_bytes_per_sample(fmt) == 2) && h < 8) return fail(VEP_ERR_GEOMETRY, "nv12/nv21/10-bit height %d < 8", h);
    if (ring_depth < 1) ring_depth = 1;
    if (ring_depth > 64) return fail(VEP_ERR_INVALID_ARG, "ring_depth %d > 64", ring_depth);
    DeviceGuard g(c->device);
    {
        std::lock_guard<std::mutex> lk(c->mu);
        if (c->streams.count(stream_id)) return fail(VEP_ERR_STATE, "stream %d already open", stream_id);
    }
    Stream *s = new (std::nothrow) Stream();
    if (!s) return fail(VEP_ERR_NOMEM, "out of host memory");
    s->id = stream_id; s->fmt = fmt; s->w = w; s->h = h; s->ring = ring_depth;
    const size_t bps = vep::fmt_bytes_per_sample(fmt);
    const size_t ysz = (size_t)w * h * bps;       // bytes of the luma plane (chroma planes scale with it)
    s->off_c0 = align_up(ysz, 256);
    const size_t csz = vep::fmt_is_444(fmt) ? ysz : vep::fmt_is_422(fmt) ? ysz / 2 : ysz / 4;   // bytes of one planar chroma plane
    if (fmt == VEP_FMT_GRAY8) { s->off_c0 = s->off_c1 = 0; s->in_bytes = ysz; }
    else if (vep::fmt_is_semiplanar(fmt)) { s->off_c1 = 0; s->in_bytes = s->off_c0 + ysz / 2; }
    else { s->off_c1 = s->off_c0 + align_up(csz, 256); s->in_bytes = s->off_c1 + csz; }
    s->out_bytes = (size_t)w * h * 3;
    s->in_slot = align_up(s->in_bytes, 4096);
    s->out_slot = align_up(kHdrRoom + s->out_bytes + kTrailerRoom, 4096);
    const int ns = ring_depth + VEP_MAX_TICKS_IN_FLIGHT + 2;   // ring + ticks in flight + one filling + slack
    s->slots.resize(ns);
    // input staging is written by the CPU (streaming stores) and read by the DMA engine; write-combined pinned
    // memory (VEP_WC=1) measured the same end to end as ordinary pinned memory on the B200 host, so it stays
    // opt-in (a caller that reads back a slot obtained from vep_acquire_input would crawl on WC memory)
    static const bool wc = getenv("VEP_WC") && atoi(getenv("VEP_WC")) == 1;
    cudaError_t e = cudaHostAlloc((void **)&s->h_base, s->out_slot * ns, cudaHostAllocDefault);
    if (e == cudaSuccess) e = cudaHostAlloc((void **)&s->h_in_base, s->in_slot * ns, wc ? cudaHostAllocWriteCombined : cudaHostAllocDefault);
    if (e == cudaSuccess) e = cudaMalloc((void **)&s->d_base, (s->in_slot + s->out_slot) * ns);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&s->cs_in, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&s->cs_out, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&s->ev_in, cudaEventDisableTiming);
    for (int i = 0; i < ns && e == cudaSuccess; ++i) {
        Slot &sl = s->slots[i];
        sl.h_in = s->h_in_base + s->in_slot * i;
        sl.h_out = s->h_base + s->out_slot * i;
        sl.d_in = s->d_base + (s->in_slot + s->out_slot) * i;
        sl.d_out = sl.d_in + s->in_slot;
        e = cudaEventCreateWithFlags(&sl.done, cudaEventDisableTiming);
    }
    if (e != cudaSuccess) {
        free_stream(s);
        return fail(e == cudaErrorMemoryAllocation ? VEP_ERR_NOMEM : VEP_ERR_CUDA, "stream_open: %s", cudaGetErrorString(e));
    }
    std::lock_guard<std::mutex> lk(c->mu);
    c->streams[stream_id] = s;
    return VEP_OK;
}

extern "C" int vep_stream_close(vep_ctx *c, int stream_id) {
    if (!c) return fail(VEP_ERR_INVALID_ARG, "ctx is null");
    DeviceGuard g(c->device);
    Stream *s = nullptr;
    {
        std::lock_guard<std::mutex> lk(c->mu);
        if (c->ticks_in_flight) return fail(VEP_ERR_BUSY, "ticks in flight; call vep_tick_wait first");
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

static void plane_ptrs(Stream *s, uint8_t *base, uint8_t *pl[3], int32_t st[3]) {
    const bool semi = vep::fmt_is_semiplanar(s->fmt);
    const int bps = vep::fmt_bytes_per_sample(s->fmt);
    const int cwp = vep::fmt_is_444(s->fmt) ? s->w : s->w / 2;     // planar chroma samples per row
    pl[0] = base; pl[1] = base + s->off_c0; pl[2] = semi ? nullptr : base + s->off_c1;
    st[0] = s->w * bps; st[1] = semi ? s->w : cwp * bps; st[2] = semi ? 0 : cwp * bps;
    if (s->fmt == VEP_FMT_GRAY8) { pl[1] = pl[2] = nullptr; st[1] = st[2] = 0; }
}

extern "C" int vep_acquire_input(vep_ctx *c, int stream_id, uint8_t *plane_out[3], int32_t stride_out[3]) {
    if (!c || !plane_out || !stride_out) return fail(VEP_ERR_INVALID_ARG, "null argument");
    Stream *s = find_stream(c, stream_id);
    if (!s) return fail(VEP_ERR_INVALID_ARG, "stream %d not open", stream_id);
    std::lock_guard<std::mutex> lk(s->mu);
    const int i = claim_slot(s);
    if (i < 0) return fail(VEP_ERR_BUSY, "stream %d: no free staging slot (tick more often or raise ring_depth)", stream_id);
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
static void copy_plane(uint8_t *dst, int dstride, const uint8_t *src, int sstride, int row_bytes, int rows) {
    if (dstride == sstride && sstride == row_bytes) stream_copy(dst, src, (size_t)row_bytes * rows);
    else for (int r = 0; r < rows; ++r) stream_copy(dst + (size_t)r * dstride, src + (size_t)r * sstride, row_bytes);
    _mm_sfence();   // order the streaming stores before the commit that makes the slot visible
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
    const bool semi = vep::fmt_is_semiplanar(s->fmt), gray = s->fmt == VEP_FMT_GRAY8;
    if (!plane || !stride || !plane[0] || (!gray && (!plane[1] || (!semi && !plane[2])))) return fail(VEP_ERR_INVALID_ARG, "null plane");
    const int bps = vep::fmt_bytes_per_sample(s->fmt);
    const int cw = ((semi || vep::fmt_is_444(s->fmt)) ? s->w : s->w / 2) * bps;      // chroma row bytes
    const int crows = (vep::fmt_is_422(s->fmt) || vep::fmt_is_444(s->fmt)) ? s->h : s->h / 2;
    if (stride[0] < s->w * bps || (!gray && (stride[1] < cw || (!semi && stride[2] < cw)))) return fail(VEP_ERR_GEOMETRY, "stride smaller than row");
    uint8_t *pl[3];
    int32_t st[3];
    {
        std::lock_guard<std::mutex> lk(s->mu);
        const int i = claim_slot(s);
        if (i < 0) return fail(VEP_ERR_BUSY, "stream %d: no free staging slot (tick more often or raise ring_depth)", stream_id);
        plane_ptrs(s, s->slots[i].h_in, pl, st);
    }
    // the copy runs outside the lock: one producer per stream
    copy_plane(pl[0], st[0], plane[0], stride[0], s->w * bps, s->h);
    if (!gray) copy_plane(pl[1], st[1], plane[1], stride[1], cw, crows);
    if (!semi && !gray) copy_plane(pl[2], st[2], plane[2], stride[2], cw, crows);
    return vep_commit_input(c, stream_id, meta);
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

extern "C" int vep_tick_begin(vep_ctx *c, int *n_frames) {
    if (!c) return fail(VEP_ERR_INVALID_ARG, "ctx is null");
    if (n_frames) *n_frames = 0;
    DeviceGuard g(c->device);
    std::lock_guard<std::mutex> lk(c->mu);
    if (c->ticks_in_flight >= VEP_MAX_TICKS_IN_FLIGHT) return fail(VEP_ERR_BUSY, "%d ticks already in flight; call vep_tick_wait", VEP_MAX_TICKS_IN_FLIGHT);
    TickRec &t = c->ticks[(c->tick_head + c->ticks_in_flight) % VEP_MAX_TICKS_IN_FLIGHT];
    t.slots.clear();
    const int rc = tick_enqueue(c, t, n_frames);
    if (rc != VEP_OK) {
        // a CUDA call failed half-way: the frames of this tick are dropped (the reference drops a frame on any
        // exception, read_image.py:130-131) and their slots are released so the cameras are not wedged
        for (auto &r : t.slots) {
            cudaStreamSynchronize(r.s->cs_in);
            cudaStreamSynchronize(r.s->cs_out);
            std::lock_guard<std::mutex> sl_lk(r.s->mu);
            r.s->slots[r.slot].state = SLOT_FREE;
        }
        t.slots.clear();
    }
    return rc;
}

static int tick_enqueue(vep_ctx *c, TickRec &t, int *n_frames) {
    std::vector<vep_frame_desc> descs;
    std::vector<Stream *> involved;
    // 1. H2D of every committed frame on its camera's input stream
    for (auto &kv : c->streams) {
        Stream *s = kv.second;
        std::lock_guard<std::mutex> sl_lk(s->mu);
        bool any = false;
        const int ns = (int)s->slots.size();
        // slots are committed in ring order ending just before wr; walk them oldest first
        for (int k = ns; k >= 1; --k) {
            const int i = ((s->wr - k) % ns + ns) % ns;
            Slot &sl = s->slots[i];
            if (sl.state != SLOT_SUBMITTED) continue;
            CUDA_TRY(cudaMemcpyAsync(sl.d_in, sl.h_in, s->in_bytes, cudaMemcpyHostToDevice, s->cs_in));
            c->stats.h2d_bytes += (int64_t)s->in_bytes;
            vep_frame_desc d{};
            const bool semi = vep::fmt_is_semiplanar(s->fmt);
            const int bps = vep::fmt_bytes_per_sample(s->fmt);
            d.plane[0] = sl.d_in; d.plane[1] = sl.d_in + s->off_c0; d.plane[2] = semi ? nullptr : sl.d_in + s->off_c1;
            const int cwp = vep::fmt_is_444(s->fmt) ? s->w : s->w / 2;
            d.stride[0] = s->w * bps; d.stride[1] = semi ? s->w : cwp * bps; d.stride[2] = semi ? 0 : cwp * bps;
            if (s->fmt == VEP_FMT_GRAY8) { d.plane[1] = d.plane[2] = nullptr; d.stride[1] = d.stride[2] = 0; }
            d.dst = sl.d_out; d.dst_stride = 3 * s->w; d.width = s->w; d.height = s->h; d.fmt = s->fmt; d.convert = 1;
            descs.push_back(d);
            write_framing(s, sl);
            sl.state = SLOT_INFLIGHT;
            t.slots.push_back(SlotRef{s, i});
            any = true;
        }
        if (any) {
            CUDA_TRY(cudaEventRecord(s->ev_in, s->cs_in));
            CUDA_TRY(cudaStreamWaitEvent(c->main, s->ev_in, 0));
            involved.push_back(s);
        }
    }
    if (t.slots.empty()) return VEP_OK;
    // 2. one batched launch for the whole tick
    int rc = build_tables(descs.data(), (int)descs.size(), c->scratch, c->rpt, vec444_ok(c));
    if (rc) return rc;
    CUDA_TRY(cudaEventRecord(t.k0, c->main));
    rc = upload_and_launch(c, c->scratch, c->main);
    if (rc) return rc;
    CUDA_TRY(cudaEventRecord(t.k1, c->main));
    // 3. D2H of each result on its camera's output stream, payload lands behind the wire prefix
    for (Stream *s : involved) CUDA_TRY(cudaStreamWaitEvent(s->cs_out, t.k1, 0));
    for (auto &r : t.slots) {
        Slot &sl = r.s->slots[r.slot];
        CUDA_TRY(cudaMemcpyAsync(sl.h_out + kHdrRoom, sl.d_out, r.s->out_bytes, cudaMemcpyDeviceToHost, r.s->cs_out));
        CUDA_TRY(cudaEventRecord(sl.done, r.s->cs_out));
        c->stats.d2h_bytes += (int64_t)r.s->out_bytes;
    }
    t.active = true;
    c->ticks_in_flight++;
    c->stats.ticks++;
    c->stats.frames_converted += (int64_t)t.slots.size();
    if (n_frames) *n_frames = (int)t.slots.size();
    return VEP_OK;
}

extern "C" int vep_tick_wait(vep_ctx *c) {
    if (!c) return fail(VEP_ERR_INVALID_ARG, "ctx is null");
    DeviceGuard g(c->device);
    std::lock_guard<std::mutex> lk(c->mu);
    if (!c->ticks_in_flight) return VEP_OK;
    TickRec &t = c->ticks[c->tick_head];
    for (auto &r : t.slots) CUDA_TRY(cudaEventSynchronize(r.s->slots[r.slot].done));
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, t.k0, t.k1) == cudaSuccess) c->stats.last_kernel_ms = ms;
    for (auto &r : t.slots) {
        std::lock_guard<std::mutex> sl_lk(r.s->mu);
        r.s->slots[r.slot].state = SLOT_READY;
        r.s->newest = r.slot;
        r.s->n_ready++;
    }
    t.slots.clear();
    t.active = false;
    c->tick_head = (c->tick_head + 1) % VEP_MAX_TICKS_IN_FLIGHT;
    c->ticks_in_flight--;
    return VEP_OK;
}

extern "C" int vep_tick(vep_ctx *c, int *n_frames) {
    int rc = vep_tick_begin(c, n_frames);
    if (rc) return rc;
    return vep_tick_wait(c);
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

extern "C" int vep_result_device(vep_ctx *c, int stream_id, const uint8_t **bgr_dev, int32_t shape[3]) {
    Stream *s = nullptr; Slot *sl = nullptr;
    int rc = ready_slot(c, stream_id, 0, &s, &sl);
    if (rc) return rc;
    if (bgr_dev) *bgr_dev = sl->d_out;
    if (shape) { shape[0] = s->h; shape[1] = s->w; shape[2] = 3; }
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
// (about 532 KB).  In HOST code this name is only a handle for cudaGetSymbolAddress — it is not a
// device address and must never be handed to a kernel or a memcpy as one.
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoallocScratchBytes];

static std::mutex g_noalloc_mu;
static void *g_noalloc_dev_ptr[64] = {};          // per device ordinal: address of g_noalloc_scratch on that GPU
static FrameDev g_noalloc_h_frames[VEP_NOALLOC_MAX_FRAMES + 1];   // fixed host staging (no heap use on this path)
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
    const bool vec = (w % 16 == 0) && aligned(f.y, 16) && f.sy % 16 == 0 && aligned(f.c0, ca) && f.sc0 % ca == 0 &&
                     (nv12 || gray || (aligned(f.c1, 8) && f.sc1 % 8 == 0)) && aligned(f.dst, 16) && f.sd % 16 == 0;
    f.fast = vec ? vep::PATH_VECTOR : vep::PATH_GENERIC;   // the table-driven kernel knows these two paths
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
            return fail(VEP_ERR_INVALID_ARG, "batch needs more than VEP_NOALLOC_MAX_CTAS (%d) thread blocks (about %d 1080p frames); split the call",
                        VEP_NOALLOC_MAX_CTAS, VEP_NOALLOC_MAX_CTAS / 270);
        for (int k = 0; k < ctas; ++k) g_noalloc_h_work[nw++] = CtaWork{nf, k * per_cta};
        ++nf;
    }
    if (nf == 0) return VEP_OK;

    // ---- both tables are rewritten in full for the entries this launch reads; nothing in the scratch has to
    //      start at zero, and stale rows of an earlier, larger call are never indexed ----
    cudaStream_t st = (cudaStream_t)cuda_stream;
    CUDA_TRY(cudaMemcpyAsync(scratch, g_noalloc_h_frames, (size_t)nf * sizeof(FrameDev), cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaMemcpyAsync(scratch + kNoallocWorkOffset, g_noalloc_h_work, (size_t)nw * sizeof(CtaWork), cudaMemcpyHostToDevice, st));
    vep::yuv_to_bgr24_batched<0><<<nw, vep::kThreads, 0, st>>>(reinterpret_cast<const FrameDev *>(scratch),
                                                               reinterpret_cast<const CtaWork *>(scratch + kNoallocWorkOffset));
    CUDA_TRY(cudaGetLastError());
    return VEP_OK;
}
