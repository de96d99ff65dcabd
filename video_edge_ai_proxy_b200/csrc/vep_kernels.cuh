// vep_kernels.cuh — sm_100a kernels: batched YUV -> BGR24 (planar 4:2:0 / 4:2:2 / 4:4:4 limited and full
// range, NV12 / NV21, 10-bit 4:2:0).
//
// Takes over the arithmetic of reference python/read_image.py:94 (`frame.to_ndarray('bgr24')`,
// i.e. PyAV -> libswscale `sws_scale`).  Results are bit-exact with libswscale's default x86
// path; the closed forms are restated in oracle/formulas.py and DESIGN.md.
//
// Shape of the work (HBM-bound byte arithmetic, no tensor cores): one launch covers every frame
// of a tick.  Each warp converts "items" = 2 luma rows x 512 px.  A lane owns 16 px of both rows: two
// 16-byte Y loads, one 8-byte U and one 8-byte V load (NV12: two 16-byte UV loads), so every chroma
// sample is fetched once for its 2x2 luma block.  The 96 output bytes per lane are re-ordered through a
// per-warp shared-memory tile so that global stores are 16 bytes per lane and contiguous across
// the warp (3-byte pixels would otherwise scatter 48-byte strided pieces).
//
// Widths that are not a multiple of 16 take the same path when the rows are padded to 16-byte strides: the last
// lane of a row then loads its 16 bytes partly from the row padding (read, never written) and its partial
// output chunk leaves as 2-byte stores.
//
// Two kernels, in file order: yuv_to_bgr24_batched (descriptor tables in device memory: more than 128 frames
// per launch, and vep_convert_batch_noalloc) and yuv_to_bgr24_tick (the DEFAULT: descriptor table in kernel
// parameters, programmatic dependent launch, per-format instantiations).  DESIGN.md section 4 has the numbers,
// including those of the variants that were measured and removed (TMA bulk-copy persistent kernel, TMA and
// strided store modes).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace vep {

#ifndef VEP_WARPS_PER_CTA
#define VEP_WARPS_PER_CTA 8
#endif
constexpr int kWarpsPerCta = VEP_WARPS_PER_CTA;   // 8 measured best (4: same overlapped, strict order within 1 %)
constexpr int kThreads = kWarpsPerCta * 32;
constexpr int kSegPx = 512;                       // pixels per warp item row (32 lanes x 16 px)
constexpr int kSegBytes = kSegPx * 3;             // 1536 output bytes per row of an item
constexpr int kGenericBlocksPerCta = 1024;        // 2x2 blocks per CTA on the generic path
constexpr int kTickRpt = 2;                       // tick kernel: vertically adjacent row pairs per lane (1 measured the same overlapped, 2 % slower in strict order)

enum : int { FMT_YUV420P = 0, FMT_YUVJ420P = 1, FMT_NV12 = 2, FMT_NV21 = 3, FMT_YUV422P = 4, FMT_YUVJ422P = 5, FMT_YUV420P10 = 6, FMT_YUV444P = 7, FMT_YUVJ444P = 8, FMT_GRAY8 = 9 };
__host__ __device__ constexpr bool fmt_is_semiplanar(int f) { return f == FMT_NV12 || f == FMT_NV21; }
__host__ __device__ constexpr bool fmt_is_422(int f) { return f == FMT_YUV422P || f == FMT_YUVJ422P; }
__host__ __device__ constexpr bool fmt_is_444(int f) { return f == FMT_YUV444P || f == FMT_YUVJ444P; }
// formats whose vector path is compiled into their own instantiation of the tick kernel only
__host__ __device__ constexpr bool fmt_vector_only_when_specialised(int f) { return fmt_is_444(f) || f == FMT_GRAY8; }
__host__ __device__ constexpr bool fmt_is_full_range(int f) { return f == FMT_YUVJ420P || f == FMT_YUVJ422P || f == FMT_YUVJ444P; }
__host__ __device__ constexpr int fmt_bytes_per_sample(int f) { return f == FMT_YUV420P10 ? 2 : 1; }

enum : int { PATH_GENERIC = 0, PATH_VECTOR = 1, PATH_UNALIGNED = 2 };

struct __align__(16) FrameDev {   // 80 bytes, one per converted frame (+1 sentinel), lives in device memory
    const uint8_t *y, *c0, *c1;   // Y, U|UV, V
    uint8_t *dst;
    int32_t sy, sc0, sc1, sd;     // strides in bytes
    int32_t w, h;
    int32_t fmt;
    int32_t fast;                 // PATH_*: 0 scalar (any even geometry, any alignment), 1 vector (16-byte aligned
                                  //         planes and strides; any even width), 2 the vector tiling over planes, strides
                                  //         and output of ANY alignment (8-bit 4:2:0 / 4:2:2 / NV12 / NV21): loads and
                                  //         stores are assembled from the aligned 16-byte chunks around them
    int32_t item_begin;           // first work item of this frame in the launch-wide item space (sentinel: total)
    int32_t segs;                 // 512-px segments per row pair
    uint32_t segs_magic;          // ceil(2^32 / segs): rp = umulhi(local, magic) for local < 2^20
    int32_t fast_tbl;             // PATH_* as seen by the table-driven kernel (it has a vector path for 8-bit 4:2:0 only)
};
struct CtaWork {                  // 8 bytes per CTA
    int32_t frame;                // index into the FrameDev table
    int32_t first;                // first warp item (fast path) / first 2x2 block (generic path)
};

// ---------------------------------------------------------------------------------------------
// Fixed-point constants.  swscale x86 path (formula A):
//   y  = (((Y<<3) - yOff) * yC) >> 16,   cb = (((u<<3)-1024) * ub) >> 16, ...
// Each product is folded into one IMAD  t = sample * (8*coef) + add  whose UPPER 16 bits are the
// floor-shifted term.  Biases ride in `add` so that halves can be combined as packed s16x2:
//   y term carries -1024, cb/cr carry +1024, cgu/cgv carry +512 each (sum +1024).
// ---------------------------------------------------------------------------------------------
template <bool FULL> struct KA;
template <> struct KA<false> { static constexpr int yC = 9539, vr = 13075, ub = 16525, ug = -3209, vg = -6660, yOff = 128; };
template <> struct KA<true>  { static constexpr int yC = 8192, vr = 11485, ub = 14516, ug = -2819, vg = -5850, yOff = 0; };

template <bool FULL> struct KM {
    using K = KA<FULL>;
    static constexpr int y_mul = 8 * K::yC,  y_add = -K::yOff * K::yC - (1024 << 16);
    static constexpr int b_mul = 8 * K::ub,  b_add = -1024 * K::ub + (1024 << 16);
    static constexpr int r_mul = 8 * K::vr,  r_add = -1024 * K::vr + (1024 << 16);
    static constexpr int gu_mul = 8 * K::ug, gu_add = -1024 * K::ug + (512 << 16);
    static constexpr int gv_mul = 8 * K::vg, gv_add = -1024 * K::vg + (512 << 16);
};

// swscale plain-C table path (formula B, limited range) — used by NV12's last two rows.
struct KB {
    static constexpr int cy = 76309, oy = 1048576, crv = 89830, cbu = 113537, cgu = -22049, cgv = -45756;
    static constexpr int base = -(384 << 16) - oy + 326 * cy + 0x8000;
    static constexpr int sub_r = crv >> 9, sub_b = cbu >> 9;
    static constexpr int sub_gu = -((-cgu + 511) >> 9), sub_gv = -((-cgv + 511) >> 9);  // floor
};

__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) { return __byte_perm(a, b, sel); }
__device__ __forceinline__ uint32_t byte_of(uint32_t w, int k) { return __byte_perm(w, 0, 0x4440 | k); }
__device__ __forceinline__ uint32_t word_of(const uint4 &v, int i) { return i == 0 ? v.x : i == 1 ? v.y : i == 2 ? v.z : v.w; }

__device__ __forceinline__ uint4 ld_stream_v4(const uint8_t *p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ uint2 ld_stream_v2(const uint8_t *p) {
    uint2 r;
    asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p));
    return r;
}
__device__ __forceinline__ void st_stream_v4(uint8_t *p, const uint4 &v) {
    asm volatile("st.global.cs.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// ---- 16 / 8 bytes from ANY address (PATH_UNALIGNED).  The aligned chunk that holds the first byte is loaded, and the
// next one only if one of the n bytes the lane needs lies in it: every chunk touched contains a byte of the plane, so no
// load can reach an unmapped page.  Rows of a warp are 16 (8) bytes apart per lane, so `a` is warp-uniform. ----
__device__ __forceinline__ uint4 shift_chunks16(const uint4 &lo, const uint4 &hi, uint32_t a) {   // bytes a .. a+15 of lo:hi, 0 < a < 16
    const uint32_t sh = (a & 3u) * 8u;
    switch (a >> 2) {
        case 0: return make_uint4(__funnelshift_r(lo.x, lo.y, sh), __funnelshift_r(lo.y, lo.z, sh), __funnelshift_r(lo.z, lo.w, sh), __funnelshift_r(lo.w, hi.x, sh));
        case 1: return make_uint4(__funnelshift_r(lo.y, lo.z, sh), __funnelshift_r(lo.z, lo.w, sh), __funnelshift_r(lo.w, hi.x, sh), __funnelshift_r(hi.x, hi.y, sh));
        case 2: return make_uint4(__funnelshift_r(lo.z, lo.w, sh), __funnelshift_r(lo.w, hi.x, sh), __funnelshift_r(hi.x, hi.y, sh), __funnelshift_r(hi.y, hi.z, sh));
        default: return make_uint4(__funnelshift_r(lo.w, hi.x, sh), __funnelshift_r(hi.x, hi.y, sh), __funnelshift_r(hi.y, hi.z, sh), __funnelshift_r(hi.z, hi.w, sh));
    }
}
__device__ __forceinline__ uint4 ldg16_any(const uint8_t *p, int n) {
    const uint32_t a = (uint32_t)(reinterpret_cast<uintptr_t>(p) & 15u);
    const uint8_t *q = p - a;
    const uint4 lo = __ldg(reinterpret_cast<const uint4 *>(q));      // cached: the neighbouring lane needs this chunk too
    if (a == 0) return lo;
    uint4 hi = make_uint4(0, 0, 0, 0);
    if ((int)a + n > 16) hi = __ldg(reinterpret_cast<const uint4 *>(q + 16));
    return shift_chunks16(lo, hi, a);
}
__device__ __forceinline__ uint2 ldg8_any(const uint8_t *p, int n) {
    const uint32_t a = (uint32_t)(reinterpret_cast<uintptr_t>(p) & 7u);
    const uint8_t *q = p - a;
    const uint2 lo = __ldg(reinterpret_cast<const uint2 *>(q));
    if (a == 0) return lo;
    uint2 hi = make_uint2(0, 0);
    if ((int)a + n > 8) hi = __ldg(reinterpret_cast<const uint2 *>(q + 8));
    const uint32_t sh = (a & 3u) * 8u;
    return (a >> 2) == 0 ? make_uint2(__funnelshift_r(lo.x, lo.y, sh), __funnelshift_r(lo.y, hi.x, sh))
                         : make_uint2(__funnelshift_r(lo.y, hi.x, sh), __funnelshift_r(hi.x, hi.y, sh));
}

// packed saturating add: per 16-bit lane max(min(a + b, 255), 0)  (VIADDMNMX.S16x2.RELU on sm_90+)
__device__ __forceinline__ uint32_t add_sat_u8_x2(uint32_t a, uint32_t b) {
    return __viaddmin_s16x2_relu(a, b, 0x00ff00ffu);
}

struct ChromaPair { uint32_t cb2, cg2, cr2; };   // each: the term (+bias) replicated in both s16 lanes

// chroma terms from raw IMAD results (upper halves hold cb+1024, cr+1024, cgu+512, cgv+512)
__device__ __forceinline__ ChromaPair chroma_pack(int tb, int tr, int tgu, int tgv) {
    ChromaPair c;
    c.cb2 = prmt((uint32_t)tb, (uint32_t)tb, 0x3232);
    c.cr2 = prmt((uint32_t)tr, (uint32_t)tr, 0x3232);
    const uint32_t tg = ((uint32_t)tgu & 0xffff0000u) + (uint32_t)tgv;   // upper half: cgu+cgv+1024, no carry in
    c.cg2 = prmt(tg, tg, 0x3232);
    return c;
}
template <bool FULL> __device__ __forceinline__ ChromaPair chroma_a(uint32_t u, uint32_t v) {
    using M = KM<FULL>;   // u, v are plain samples (<<3 folded into the multiplier)
    return chroma_pack((int)u * M::b_mul + M::b_add, (int)v * M::r_mul + M::r_add,
                       (int)u * M::gu_mul + M::gu_add, (int)v * M::gv_mul + M::gv_add);
}
// NV12 even rows: su = u_up + u_cur, "(su<<2) - 1024" -> half the multiplier
__device__ __forceinline__ ChromaPair chroma_a_sum2(uint32_t su, uint32_t sv) {
    using M = KM<false>;
    return chroma_pack((int)su * (M::b_mul / 2) + M::b_add, (int)sv * (M::r_mul / 2) + M::r_add,
                       (int)su * (M::gu_mul / 2) + M::gu_add, (int)sv * (M::gv_mul / 2) + M::gv_add);
}

// channel pairs [c0,0,c1,0] of pixels (0,1) and (2,3) -> the 12 bytes B0 G0 R0 B1 | G1 R1 B2 G2 | R2 B3 G3 R3
__device__ __forceinline__ void interleave_bgr(uint32_t bp, uint32_t gp, uint32_t rp, uint32_t bq, uint32_t gq, uint32_t rq,
                                               uint32_t &w0, uint32_t &w1, uint32_t &w2) {
    const uint32_t bgp = gp * 256u + bp;   // [B0 G0 B1 G1]
    const uint32_t bgq = gq * 256u + bq;   // [B2 G2 B3 G3]
    w0 = prmt(bgp, rp, 0x2410);            // B0 G0 R0 B1
    const uint32_t x = prmt(bgp, rp, 0x0063);   // G1 R1 . .
    w1 = prmt(x, bgq, 0x5410);             // G1 R1 B2 G2
    w2 = prmt(bgq, rq, 0x6324);            // R2 B3 G3 R3
}

// four luma IMAD results (upper halves = y - 1024) + the two chroma pairs covering them -> 12 output bytes
__device__ __forceinline__ void pack4_a(int t0, int t1, int t2, int t3, const ChromaPair &P, const ChromaPair &Q,
                                        uint32_t &w0, uint32_t &w1, uint32_t &w2) {
    const uint32_t yp = prmt((uint32_t)t0, (uint32_t)t1, 0x7632);   // [y0-1024, y1-1024] as s16x2
    const uint32_t yq = prmt((uint32_t)t2, (uint32_t)t3, 0x7632);
    const uint32_t bp = add_sat_u8_x2(yp, P.cb2), gp = add_sat_u8_x2(yp, P.cg2), rp = add_sat_u8_x2(yp, P.cr2);
    const uint32_t bq = add_sat_u8_x2(yq, Q.cb2), gq = add_sat_u8_x2(yq, Q.cg2), rq = add_sat_u8_x2(yq, Q.cr2);
    interleave_bgr(bp, gp, rp, bq, gq, rq, w0, w1, w2);
}

// 4 luma samples (one 32-bit word) + the two chroma pairs covering them -> 12 output bytes
template <bool FULL>
__device__ __forceinline__ void luma4_a(uint32_t yw, const ChromaPair &P, const ChromaPair &Q,
                                        uint32_t &w0, uint32_t &w1, uint32_t &w2) {
    using M = KM<FULL>;
    const int t0 = (int)byte_of(yw, 0) * M::y_mul + M::y_add;
    const int t1 = (int)byte_of(yw, 1) * M::y_mul + M::y_add;
    const int t2 = (int)byte_of(yw, 2) * M::y_mul + M::y_add;
    const int t3 = (int)(yw >> 24) * M::y_mul + M::y_add;
    pack4_a(t0, t1, t2, t3, P, Q, w0, w1, w2);
}

// formula B; kb/kg/kr = base + d*cy for one chroma column pair
struct ChromaB { int kb, kg, kr; };
__device__ __forceinline__ ChromaB chroma_b(int u, int v) {
    ChromaB c;
    const int db = ((u * KB::cbu) >> 16) - KB::sub_b;
    const int dr = ((v * KB::crv) >> 16) - KB::sub_r;
    const int dg = ((u * KB::cgu) >> 16) - KB::sub_gu + ((v * KB::cgv) >> 16) - KB::sub_gv;
    c.kb = KB::base + db * KB::cy;
    c.kg = KB::base + dg * KB::cy;
    c.kr = KB::base + dr * KB::cy;
    return c;
}
__device__ __forceinline__ uint32_t clamp_pair_hi(int a, int b) {   // [sat_u8(a>>16), sat_u8(b>>16)] as s16x2
    return __vimin_s16x2_relu(prmt((uint32_t)a, (uint32_t)b, 0x7632), 0x00ff00ffu);
}
// y0..y3 are luma * cy (formula B), P/Q the chroma offsets of the two column pairs
__device__ __forceinline__ void pack4_b(int y0, int y1, int y2, int y3, const ChromaB &P, const ChromaB &Q,
                                        uint32_t &w0, uint32_t &w1, uint32_t &w2) {
    const uint32_t bp = clamp_pair_hi(y0 + P.kb, y1 + P.kb), gp = clamp_pair_hi(y0 + P.kg, y1 + P.kg),
                   rp = clamp_pair_hi(y0 + P.kr, y1 + P.kr);
    const uint32_t bq = clamp_pair_hi(y2 + Q.kb, y3 + Q.kb), gq = clamp_pair_hi(y2 + Q.kg, y3 + Q.kg),
                   rq = clamp_pair_hi(y2 + Q.kr, y3 + Q.kr);
    const uint32_t bgp = gp * 256u + bp, bgq = gq * 256u + bq;
    w0 = prmt(bgp, rp, 0x2410);
    const uint32_t x = prmt(bgp, rp, 0x0063);
    w1 = prmt(x, bgq, 0x5410);
    w2 = prmt(bgq, rq, 0x6324);
}
__device__ __forceinline__ void luma4_b(uint32_t yw, const ChromaB &P, const ChromaB &Q, uint32_t &w0, uint32_t &w1, uint32_t &w2) {
    pack4_b((int)byte_of(yw, 0) * KB::cy, (int)byte_of(yw, 1) * KB::cy, (int)byte_of(yw, 2) * KB::cy, (int)(yw >> 24) * KB::cy, P, Q, w0, w1, w2);
}

// ---- yuv420p10le: 10-bit samples in 16-bit containers.  Same stages with 11-bit terms: luma (Y<<1), odd rows (u<<1),
//      even rows (u_above + u); rows H-2/H-1 use formula B on y = (Y+2)>>2, u = min((blend+8192)>>14, 255). ----
struct P10 {   // multipliers: the 8-bit constants carry "<<3", 10-bit samples need "<<1" (a quarter), sums need "<<0"
    using M = KM<false>;
    static constexpr int y_mul = M::y_mul / 4, b_mul = M::b_mul / 4, r_mul = M::r_mul / 4, gu_mul = M::gu_mul / 4, gv_mul = M::gv_mul / 4;
};
__device__ __forceinline__ ChromaPair chroma_p10(uint32_t u, uint32_t v, int shift /*0: (u<<1), 1: sum of two rows*/) {
    using M = KM<false>;
    const int bm = P10::b_mul >> shift, rm = P10::r_mul >> shift, gum = P10::gu_mul / (1 << shift), gvm = P10::gv_mul / (1 << shift);
    return chroma_pack((int)u * bm + M::b_add, (int)v * rm + M::r_add, (int)u * gum + M::gu_add, (int)v * gvm + M::gv_add);
}
// one group of 4 px x 2 rows: ya/yb = two 32-bit words each (4 samples), cu/cv/pu/pv = one word each (2 samples)
__device__ __forceinline__ void group_p10(uint32_t ya0, uint32_t ya1, uint32_t yb0, uint32_t yb1, uint32_t cu, uint32_t cv,
                                          uint32_t pu, uint32_t pv, bool tail, uint32_t *o0, uint32_t *o1) {
    using M = KM<false>;
    const uint32_t u0 = cu & 0xffffu, u1 = cu >> 16, v0 = cv & 0xffffu, v1 = cv >> 16;
    const uint32_t a0 = pu & 0xffffu, a1 = pu >> 16, b0 = pv & 0xffffu, b1 = pv >> 16;     // row above
    const int ye0 = ya0 & 0xffffu, ye1 = ya0 >> 16, ye2 = ya1 & 0xffffu, ye3 = ya1 >> 16;
    const int yo0 = yb0 & 0xffffu, yo1 = yb0 >> 16, yo2 = yb1 & 0xffffu, yo3 = yb1 >> 16;
    if (!tail) {
        const ChromaPair Pe = chroma_p10(a0 + u0, b0 + v0, 1), Qe = chroma_p10(a1 + u1, b1 + v1, 1);
        const ChromaPair Po = chroma_p10(u0, v0, 0), Qo = chroma_p10(u1, v1, 0);
        pack4_a(ye0 * P10::y_mul + M::y_add, ye1 * P10::y_mul + M::y_add, ye2 * P10::y_mul + M::y_add, ye3 * P10::y_mul + M::y_add, Pe, Qe, o0[0], o0[1], o0[2]);
        pack4_a(yo0 * P10::y_mul + M::y_add, yo1 * P10::y_mul + M::y_add, yo2 * P10::y_mul + M::y_add, yo3 * P10::y_mul + M::y_add, Po, Qo, o1[0], o1[1], o1[2]);
    } else {
        const ChromaB Pe = chroma_b(min((int)(a0 * 1024u + u0 * 3072u + 8192u) >> 14, 255), min((int)(b0 * 1024u + v0 * 3072u + 8192u) >> 14, 255));
        const ChromaB Qe = chroma_b(min((int)(a1 * 1024u + u1 * 3072u + 8192u) >> 14, 255), min((int)(b1 * 1024u + v1 * 3072u + 8192u) >> 14, 255));
        const ChromaB Po = chroma_b(min((int)(u0 + 2) >> 2, 255), min((int)(v0 + 2) >> 2, 255)), Qo = chroma_b(min((int)(u1 + 2) >> 2, 255), min((int)(v1 + 2) >> 2, 255));
        pack4_b(((ye0 + 2) >> 2) * KB::cy, ((ye1 + 2) >> 2) * KB::cy, ((ye2 + 2) >> 2) * KB::cy, ((ye3 + 2) >> 2) * KB::cy, Pe, Qe, o0[0], o0[1], o0[2]);
        pack4_b(((yo0 + 2) >> 2) * KB::cy, ((yo1 + 2) >> 2) * KB::cy, ((yo2 + 2) >> 2) * KB::cy, ((yo3 + 2) >> 2) * KB::cy, Po, Qo, o1[0], o1[1], o1[2]);
    }
}
// one lane's 16 px x 2 rows: ya/yb two uint4 each (16 samples), cu/cv/pu/pv one uint4 each (8 samples)
__device__ __forceinline__ void compute_p10(const uint4 &ya_lo, const uint4 &ya_hi, const uint4 &yb_lo, const uint4 &yb_hi, const uint4 &cu,
                                            const uint4 &cv, const uint4 &pu, const uint4 &pv, bool tail, uint32_t (&o0)[12], uint32_t (&o1)[12]) {
#pragma unroll
    for (int g = 0; g < 4; ++g) {
        const uint4 &a = g < 2 ? ya_lo : ya_hi, &b = g < 2 ? yb_lo : yb_hi;
        const int h = (g & 1) * 2;
        group_p10(word_of(a, h), word_of(a, h + 1), word_of(b, h), word_of(b, h + 1), word_of(cu, g), word_of(cv, g), word_of(pu, g), word_of(pv, g),
                  tail, &o0[3 * g], &o1[3 * g]);
    }
}



// ---- 4:4:4, one row of 16 px: swscale's full-chroma C stage (one rounding per channel, wrapping 32-bit sums) ----
template <bool FULL>
__device__ __forceinline__ void compute_444_row(const uint4 &y, const uint4 &u, const uint4 &v, uint32_t (&o)[12]) {
    using K = KA<FULL>;
    constexpr uint32_t yC9 = (uint32_t)K::yC << 9;
    constexpr uint32_t kB = (uint32_t)(-(K::yOff << 6) * K::yC + (1 << 21)) - 128u * ((uint32_t)K::ub << 9);
    constexpr uint32_t kR = (uint32_t)(-(K::yOff << 6) * K::yC + (1 << 21)) - 128u * ((uint32_t)K::vr << 9);
    constexpr uint32_t kG = (uint32_t)(-(K::yOff << 6) * K::yC + (1 << 21)) - 128u * ((uint32_t)(K::ug << 9)) - 128u * ((uint32_t)(K::vg << 9));
#pragma unroll
    for (int g = 0; g < 4; ++g) {
        const uint32_t yw = word_of(y, g), uw = word_of(u, g), vw = word_of(v, g);
        int b[4], gg[4], r[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const uint32_t Y = i == 3 ? yw >> 24 : byte_of(yw, i), U = i == 3 ? uw >> 24 : byte_of(uw, i), V = i == 3 ? vw >> 24 : byte_of(vw, i);
            const uint32_t yt = Y * yC9;
            b[i] = (int)(yt + U * ((uint32_t)K::ub << 9) + kB) >> 22;
            r[i] = (int)(yt + V * ((uint32_t)K::vr << 9) + kR) >> 22;
            gg[i] = (int)(yt + U * (uint32_t)(K::ug << 9) + V * (uint32_t)(K::vg << 9) + kG) >> 22;
        }
        auto pair = [](int a, int c) { return __vimin_s16x2_relu(prmt((uint32_t)a, (uint32_t)c, 0x5410), 0x00ff00ffu); };   // values fit s16: |x>>22| <= 512
        interleave_bgr(pair(b[0], b[1]), pair(gg[0], gg[1]), pair(r[0], r[1]), pair(b[2], b[3]), pair(gg[2], gg[3]), pair(r[2], r[3]),
                       o[3 * g], o[3 * g + 1], o[3 * g + 2]);
    }
}

// ---- one lane's 16 px x 2 rows, planar: uu/vv hold 8 chroma samples each ---------------------------
template <bool FULL>
__device__ __forceinline__ void compute_planar(const uint4 &ya, const uint4 &yb, const uint2 &uu, const uint2 &vv,
                                               uint32_t (&o0)[12], uint32_t (&o1)[12]) {
#pragma unroll
    for (int g = 0; g < 4; ++g) {
        const uint32_t uw = g < 2 ? uu.x : uu.y, vw = g < 2 ? vv.x : vv.y;
        const int b = (g & 1) * 2;
        const ChromaPair P = chroma_a<FULL>(byte_of(uw, b), byte_of(vw, b));
        const ChromaPair Q = chroma_a<FULL>(b == 2 ? (uw >> 24) : byte_of(uw, 1), b == 2 ? (vw >> 24) : byte_of(vw, 1));
        luma4_a<FULL>(word_of(ya, g), P, Q, o0[3 * g], o0[3 * g + 1], o0[3 * g + 2]);
        luma4_a<FULL>(word_of(yb, g), P, Q, o1[3 * g], o1[3 * g + 1], o1[3 * g + 2]);
    }
}
// 4:2:2: each luma row has its own chroma row (no vertical sharing); one row at a time
template <bool FULL>
__device__ __forceinline__ void compute_planar_row(const uint4 &y, const uint2 &uu, const uint2 &vv, uint32_t (&o)[12]) {
#pragma unroll
    for (int g = 0; g < 4; ++g) {
        const uint32_t uw = g < 2 ? uu.x : uu.y, vw = g < 2 ? vv.x : vv.y;
        const int b = (g & 1) * 2;
        const ChromaPair P = chroma_a<FULL>(byte_of(uw, b), byte_of(vw, b));
        const ChromaPair Q = chroma_a<FULL>(b == 2 ? (uw >> 24) : byte_of(uw, 1), b == 2 ? (vw >> 24) : byte_of(vw, 1));
        luma4_a<FULL>(word_of(y, g), P, Q, o[3 * g], o[3 * g + 1], o[3 * g + 2]);
    }
}
__device__ __forceinline__ uint32_t swap_pairs(uint32_t w) { return prmt(w, 0, 0x2301); }   // V,U -> U,V in both pairs
__device__ __forceinline__ uint4 swap_pairs(const uint4 &v) { return make_uint4(swap_pairs(v.x), swap_pairs(v.y), swap_pairs(v.z), swap_pairs(v.w)); }

template <bool FULL>
__device__ __forceinline__ void lane_planar(const FrameDev &f, int rp, int x, bool un, uint32_t (&o0)[12], uint32_t (&o1)[12]) {
    const uint8_t *py = f.y + (size_t)(2 * rp) * f.sy + x, *pu = f.c0 + (size_t)rp * f.sc0 + (x >> 1), *pv = f.c1 + (size_t)rp * f.sc1 + (x >> 1);
    const int nb = min(16, f.w - x);                 // visible luma bytes of this lane
    const uint4 ya = un ? ldg16_any(py, nb) : ld_stream_v4(py);
    const uint4 yb = un ? ldg16_any(py + f.sy, nb) : ld_stream_v4(py + f.sy);
    const uint2 uu = un ? ldg8_any(pu, nb >> 1) : ld_stream_v2(pu);
    const uint2 vv = un ? ldg8_any(pv, nb >> 1) : ld_stream_v2(pv);
    compute_planar<FULL>(ya, yb, uu, vv, o0, o1);
}

// ---- one lane's 16 px x 2 rows, NV12 (formula C): cu = this chroma row, pu = the row above ---------
__device__ __forceinline__ void compute_nv12(const uint4 &ya, const uint4 &yb, const uint4 &cu, const uint4 &pu, bool tail,
                                             uint32_t (&o0)[12], uint32_t (&o1)[12]) {
#pragma unroll
    for (int g = 0; g < 4; ++g) {
        const uint32_t c = word_of(cu, g), p = word_of(pu, g);
        const uint32_t u0 = byte_of(c, 0), v0 = byte_of(c, 1), u1 = byte_of(c, 2), v1 = c >> 24;
        const uint32_t pu0 = byte_of(p, 0), pv0 = byte_of(p, 1), pu1 = byte_of(p, 2), pv1 = p >> 24;
        if (!tail) {
            const ChromaPair Pe = chroma_a_sum2(pu0 + u0, pv0 + v0), Qe = chroma_a_sum2(pu1 + u1, pv1 + v1);
            const ChromaPair Po = chroma_a<false>(u0, v0), Qo = chroma_a<false>(u1, v1);
            luma4_a<false>(word_of(ya, g), Pe, Qe, o0[3 * g], o0[3 * g + 1], o0[3 * g + 2]);
            luma4_a<false>(word_of(yb, g), Po, Qo, o1[3 * g], o1[3 * g + 1], o1[3 * g + 2]);
        } else {
            // row H-2: 12-bit blend 1/4 above + 3/4 current (rounded); row H-1: current row as is
            const ChromaB Pe = chroma_b((int)(pu0 * 1024u + u0 * 3072u + 2048u) >> 12, (int)(pv0 * 1024u + v0 * 3072u + 2048u) >> 12);
            const ChromaB Qe = chroma_b((int)(pu1 * 1024u + u1 * 3072u + 2048u) >> 12, (int)(pv1 * 1024u + v1 * 3072u + 2048u) >> 12);
            const ChromaB Po = chroma_b((int)u0, (int)v0), Qo = chroma_b((int)u1, (int)v1);
            luma4_b(word_of(ya, g), Pe, Qe, o0[3 * g], o0[3 * g + 1], o0[3 * g + 2]);
            luma4_b(word_of(yb, g), Po, Qo, o1[3 * g], o1[3 * g + 1], o1[3 * g + 2]);
        }
    }
}
__device__ __forceinline__ void lane_nv12(const FrameDev &f, int rp, int x, bool un, uint32_t (&o0)[12], uint32_t (&o1)[12]) {
    const uint8_t *py = f.y + (size_t)(2 * rp) * f.sy + x;
    const int nb = min(16, f.w - x);
    const uint4 ya = un ? ldg16_any(py, nb) : ld_stream_v4(py);
    const uint4 yb = un ? ldg16_any(py + f.sy, nb) : ld_stream_v4(py + f.sy);
    const int up = rp > 0 ? rp - 1 : 0;
    uint4 cu = un ? ldg16_any(f.c0 + (size_t)rp * f.sc0 + x, nb) : __ldg(reinterpret_cast<const uint4 *>(f.c0 + (size_t)rp * f.sc0 + x));   // u0 v0 u1 v1 ...
    uint4 pu = un ? ldg16_any(f.c0 + (size_t)up * f.sc0 + x, nb) : __ldg(reinterpret_cast<const uint4 *>(f.c0 + (size_t)up * f.sc0 + x));   // row above (L2/L1 hit)
    if (f.fmt == FMT_NV21) { cu = swap_pairs(cu); pu = swap_pairs(pu); }
    compute_nv12(ya, yb, cu, pu, rp == (f.h >> 1) - 1, o0, o1);   // rows H-2, H-1: swscale leaves its SIMD output stage
}

// ---- generic path: one 2x2 block per step, scalar, any even geometry / alignment ----------------
__device__ __forceinline__ uint8_t sat8(int v) { return (uint8_t)min(max(v, 0), 255); }

template <bool FULL>
__device__ __forceinline__ void px_a(int Y, int Uw, int Vw, uint8_t *d) {
    using K = KA<FULL>;
    const int y = (((Y << 3) - K::yOff) * K::yC) >> 16;
    const int cb = (Uw * K::ub) >> 16, cr = (Vw * K::vr) >> 16;
    const int cg = ((Uw * K::ug) >> 16) + ((Vw * K::vg) >> 16);
    d[0] = sat8(y + cb); d[1] = sat8(y + cg); d[2] = sat8(y + cr);
}
__device__ __forceinline__ void px_b(int Y, int u, int v, uint8_t *d) {
    const ChromaB c = chroma_b(u, v);
    d[0] = sat8((Y * KB::cy + c.kb) >> 16); d[1] = sat8((Y * KB::cy + c.kg) >> 16); d[2] = sat8((Y * KB::cy + c.kr) >> 16);
}

__device__ __forceinline__ void px_a10(int Y, int Uw, int Vw, uint8_t *d) {   // 11-bit terms: Y = sample<<1, Uw/Vw already "- 1024"
    using K = KA<false>;
    const int y = ((Y - K::yOff) * K::yC) >> 16;
    const int cb = (Uw * K::ub) >> 16, cr = (Vw * K::vr) >> 16;
    const int cg = ((Uw * K::ug) >> 16) + ((Vw * K::vg) >> 16);
    d[0] = sat8(y + cb); d[1] = sat8(y + cg); d[2] = sat8(y + cr);
}
__device__ __forceinline__ void generic_block_p10(const FrameDev &f, int bx, int by) {
    const uint16_t *y0 = reinterpret_cast<const uint16_t *>(f.y + (size_t)(2 * by) * f.sy) + 2 * bx;
    const uint16_t *y1 = reinterpret_cast<const uint16_t *>(f.y + (size_t)(2 * by + 1) * f.sy) + 2 * bx;
    uint8_t *d0 = f.dst + (size_t)(2 * by) * f.sd + 6 * bx, *d1 = d0 + f.sd;
    const int up = by > 0 ? by - 1 : 0;
    const int u = reinterpret_cast<const uint16_t *>(f.c0 + (size_t)by * f.sc0)[bx], v = reinterpret_cast<const uint16_t *>(f.c1 + (size_t)by * f.sc1)[bx];
    const int pu = reinterpret_cast<const uint16_t *>(f.c0 + (size_t)up * f.sc0)[bx], pv = reinterpret_cast<const uint16_t *>(f.c1 + (size_t)up * f.sc1)[bx];
    if (by != (f.h >> 1) - 1) {
        px_a10(y0[0] << 1, pu + u - 1024, pv + v - 1024, d0); px_a10(y0[1] << 1, pu + u - 1024, pv + v - 1024, d0 + 3);
        px_a10(y1[0] << 1, (u << 1) - 1024, (v << 1) - 1024, d1); px_a10(y1[1] << 1, (u << 1) - 1024, (v << 1) - 1024, d1 + 3);
    } else {
        const int ue = min((pu * 1024 + u * 3072 + 8192) >> 14, 255), ve = min((pv * 1024 + v * 3072 + 8192) >> 14, 255);
        const int uo = min((u + 2) >> 2, 255), vo = min((v + 2) >> 2, 255);
        px_b((y0[0] + 2) >> 2, ue, ve, d0); px_b((y0[1] + 2) >> 2, ue, ve, d0 + 3);
        px_b((y1[0] + 2) >> 2, uo, vo, d1); px_b((y1[1] + 2) >> 2, uo, vo, d1 + 3);
    }
}

// 4:4:4: swscale's plain-C "full chroma" stage — one rounding per channel, wrapping 32-bit arithmetic
// (the limited-range blue term can pass 2^31 and then clips to 0, as in the library), every pixel its own chroma
template <bool FULL>
__device__ __forceinline__ void px_444(uint32_t Y, uint32_t U, uint32_t V, uint8_t *d) {
    using K = KA<FULL>;
    constexpr uint32_t yC9 = (uint32_t)K::yC << 9, kY = (uint32_t)(-(K::yOff << 6) * K::yC + (1 << 21));
    const uint32_t yt = Y * yC9 + kY;
    const int u9 = ((int)U - 128) << 9, v9 = ((int)V - 128) << 9;
    const int b = (int)(yt + (uint32_t)(u9 * K::ub)), g = (int)(yt + (uint32_t)(u9 * K::ug) + (uint32_t)(v9 * K::vg)), r = (int)(yt + (uint32_t)(v9 * K::vr));
    d[0] = sat8(b >> 22); d[1] = sat8(g >> 22); d[2] = sat8(r >> 22);
}
__device__ __forceinline__ void generic_block_444(const FrameDev &f, int bx, int by) {
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        const uint8_t *y = f.y + (size_t)(2 * by + r) * f.sy + 2 * bx, *u = f.c0 + (size_t)(2 * by + r) * f.sc0 + 2 * bx,
                      *v = f.c1 + (size_t)(2 * by + r) * f.sc1 + 2 * bx;
        uint8_t *d = f.dst + (size_t)(2 * by + r) * f.sd + 6 * bx;
        if (f.fmt == FMT_YUVJ444P) { px_444<true>(y[0], u[0], v[0], d); px_444<true>(y[1], u[1], v[1], d + 3); }
        else { px_444<false>(y[0], u[0], v[0], d); px_444<false>(y[1], u[1], v[1], d + 3); }
    }
}

__device__ __forceinline__ void generic_block(const FrameDev &f, int bx, int by) {
    if (f.fmt == FMT_YUV420P10) { generic_block_p10(f, bx, by); return; }
    if (fmt_is_444(f.fmt)) { generic_block_444(f, bx, by); return; }
    if (f.fmt == FMT_GRAY8) {   // monochrome decode: swscale replicates the (full-range) luma byte, B = G = R = Y
#pragma unroll
        for (int r = 0; r < 2; ++r) {
            const uint8_t *y = f.y + (size_t)(2 * by + r) * f.sy + 2 * bx;
            uint8_t *d = f.dst + (size_t)(2 * by + r) * f.sd + 6 * bx;
            d[0] = d[1] = d[2] = y[0]; d[3] = d[4] = d[5] = y[1];
        }
        return;
    }
    const uint8_t *y0 = f.y + (size_t)(2 * by) * f.sy + 2 * bx, *y1 = y0 + f.sy;
    uint8_t *d0 = f.dst + (size_t)(2 * by) * f.sd + 6 * bx, *d1 = d0 + f.sd;
    if (!fmt_is_semiplanar(f.fmt)) {
        // 4:2:0: one chroma row per row pair; 4:2:2: chroma row j for luma row j
        const int r0 = fmt_is_422(f.fmt) ? 2 * by : by, r1 = fmt_is_422(f.fmt) ? 2 * by + 1 : by;
        const int Uw = ((int)f.c0[(size_t)r0 * f.sc0 + bx] << 3) - 1024, Vw = ((int)f.c1[(size_t)r0 * f.sc1 + bx] << 3) - 1024;
        const int Ux = ((int)f.c0[(size_t)r1 * f.sc0 + bx] << 3) - 1024, Vx = ((int)f.c1[(size_t)r1 * f.sc1 + bx] << 3) - 1024;
        if (fmt_is_full_range(f.fmt)) {
            px_a<true>(y0[0], Uw, Vw, d0); px_a<true>(y0[1], Uw, Vw, d0 + 3);
            px_a<true>(y1[0], Ux, Vx, d1); px_a<true>(y1[1], Ux, Vx, d1 + 3);
        } else {
            px_a<false>(y0[0], Uw, Vw, d0); px_a<false>(y0[1], Uw, Vw, d0 + 3);
            px_a<false>(y1[0], Ux, Vx, d1); px_a<false>(y1[1], Ux, Vx, d1 + 3);
        }
        return;
    }
    const int up = by > 0 ? by - 1 : 0;
    const uint8_t *c = f.c0 + (size_t)by * f.sc0 + 2 * bx, *p = f.c0 + (size_t)up * f.sc0 + 2 * bx;
    const int iu = f.fmt == FMT_NV21 ? 1 : 0;
    const int u = c[iu], v = c[1 - iu], pu = p[iu], pv = p[1 - iu];
    if (by != (f.h >> 1) - 1) {
        const int Ue = ((pu + u) << 2) - 1024, Ve = ((pv + v) << 2) - 1024;
        const int Uo = (u << 3) - 1024, Vo = (v << 3) - 1024;
        px_a<false>(y0[0], Ue, Ve, d0); px_a<false>(y0[1], Ue, Ve, d0 + 3);
        px_a<false>(y1[0], Uo, Vo, d1); px_a<false>(y1[1], Uo, Vo, d1 + 3);
    } else {
        const int ue = (pu * 1024 + u * 3072 + 2048) >> 12, ve = (pv * 1024 + v * 3072 + 2048) >> 12;
        px_b(y0[0], ue, ve, d0); px_b(y0[1], ue, ve, d0 + 3);
        px_b(y1[0], u, v, d1);   px_b(y1[1], u, v, d1 + 3);
    }
}

// ---- store of one row segment from the per-warp tile ---------------------------------------------
// t: the segment's bytes in shared memory (lane l parked its 48 bytes at l*48), g: 16-byte aligned global
// address of the segment's first byte, seg_bytes: visible bytes (3 * pixels, even).  16 bytes per lane,
// contiguous across the warp; a last partial chunk (width not a multiple of 16) leaves as 2-byte stores so
// that nothing beyond the visible pixels is written (swscale overwrites the row padding; we must not).
__device__ __forceinline__ void store_segment(uint8_t *g, const uint8_t *t, int seg_bytes, int lane) {
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const int off = (k * 32 + lane) * 16;
        if (off + 16 <= seg_bytes) {
            st_stream_v4(g + off, *reinterpret_cast<const uint4 *>(t + off));
        } else if (off < seg_bytes) {
            for (int b = off; b < seg_bytes; b += 2)
                *reinterpret_cast<uint16_t *>(g + b) = *reinterpret_cast<const uint16_t *>(t + b);
        }
    }
}

// The same for a row that starts at ANY address (PATH_UNALIGNED): the aligned 16-byte chunks of the OUTPUT are assembled
// from the two tile chunks each straddles (tile rows are 16-byte aligned and 1536 bytes long: the second chunk of a
// fully covered output chunk always lies inside the row); the partial chunks at both ends leave as byte stores.
__device__ __forceinline__ void store_segment_any(uint8_t *g, const uint8_t *t, int seg_bytes, int lane) {
    const uint32_t a = (uint32_t)(reinterpret_cast<uintptr_t>(g) & 15u);
    if (a == 0) { store_segment(g, t, seg_bytes, lane); return; }   // warp-uniform
    uint8_t *G = g - a;                                              // first aligned chunk that holds a byte of the row
    const int nch = ((int)a + seg_bytes + 15) >> 4;
    for (int k = lane; k < nch; k += 32) {
        const int o = 16 * k - (int)a;                               // tile offset of the chunk's first byte
        if (o >= 0 && o + 16 <= seg_bytes) {
            const int r = o & 15;                                    // = 16 - a
            const uint4 lo = *reinterpret_cast<const uint4 *>(t + (o - r)), hi = *reinterpret_cast<const uint4 *>(t + (o - r) + 16);
            st_stream_v4(G + 16 * k, shift_chunks16(lo, hi, (uint32_t)r));
        }
    }
    // the partial chunks at the two ends of the row: one byte per lane, one store request each
    const int head = min(16 - (int)a, seg_bytes);                    // tile bytes [0, head) share the first output chunk
    if (lane < head) g[lane] = t[lane];
    const int tail0 = ((int)a + seg_bytes) & ~15;                    // row-relative (from G) start of the last chunk
    const int tb = tail0 - (int)a + lane;                            // tile byte this lane would carry
    if (tail0 > 0 && tb >= head && tb < seg_bytes && (((int)a + seg_bytes) & 15) != 0) g[tb] = t[tb];
}

// ---- the table-driven kernel -----------------------------------------------------------------------
// Descriptor tables in device memory (FrameDev[], CtaWork[]): the path for more than 128 frames per launch
// and for vep_convert_batch_noalloc.  One warp item = 512 px x 2 luma rows; shared-memory re-order, then
// 16 B/lane warp-contiguous st.global.
__global__ void __launch_bounds__(kThreads)
yuv_to_bgr24_batched(const FrameDev *__restrict__ frames, const CtaWork *__restrict__ work) {
    __shared__ __align__(128) uint8_t tile[kWarpsPerCta][2][kSegBytes];
    const CtaWork cw = work[blockIdx.x];
    const FrameDev f = frames[cw.frame];

    if (f.fast_tbl == PATH_GENERIC) {
        const int bw = f.w >> 1, nblk = bw * (f.h >> 1);
        for (int q = cw.first + threadIdx.x; q < min(cw.first + kGenericBlocksPerCta, nblk); q += kThreads)
            generic_block(f, q % bw, q / bw);
        return;
    }

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int segs = (f.w + kSegPx - 1) / kSegPx;
    const int item = cw.first + warp;
    if (item >= segs * (f.h >> 1)) return;   // warp-uniform
    const int rp = item / segs, seg = item - rp * segs;
    const int x = seg * kSegPx + lane * 16;

    uint32_t o0[12], o1[12];
    const bool active = x < f.w;             // the last active lane of a row may cover fewer than 16 visible pixels
    const bool un = f.fast_tbl == PATH_UNALIGNED;
    if (active) {
        if (f.fmt == FMT_YUV420P) lane_planar<false>(f, rp, x, un, o0, o1);
        else if (f.fmt == FMT_YUVJ420P) lane_planar<true>(f, rp, x, un, o0, o1);
        else lane_nv12(f, rp, x, un, o0, o1);   // nv12 / nv21 (other formats never reach here: fast_tbl is generic)
    }
    uint8_t *g0 = f.dst + (size_t)(2 * rp) * f.sd + (size_t)seg * kSegBytes;
    const int seg_bytes = min(kSegPx, f.w - seg * kSegPx) * 3;

    uint8_t *t0 = tile[warp][0], *t1 = tile[warp][1];
    if (active) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {   // 48 B lane stride: conflict-free per quarter warp
            *reinterpret_cast<uint4 *>(t0 + lane * 48 + k * 16) = make_uint4(o0[4 * k], o0[4 * k + 1], o0[4 * k + 2], o0[4 * k + 3]);
            *reinterpret_cast<uint4 *>(t1 + lane * 48 + k * 16) = make_uint4(o1[4 * k], o1[4 * k + 1], o1[4 * k + 2], o1[4 * k + 3]);
        }
    }
    __syncwarp();
    if (un) { store_segment_any(g0, t0, seg_bytes, lane); store_segment_any(g0 + f.sd, t1, seg_bytes, lane); }
    else { store_segment(g0, t0, seg_bytes, lane); store_segment(g0 + f.sd, t1, seg_bytes, lane); }
}


// ================================================================================================
// Tick kernel (THE DEFAULT): descriptor table in the kernel-parameter constant bank.
//
// Same tiling as the table-driven kernel above (one warp = 512 px x 2 luma rows per row pair,
// shared-memory re-order, 16-byte warp-contiguous stores) but nothing on a CTA's critical path
// waits for global memory before its pixel loads are issued: the frame descriptors and the
// CTA->frame prefix arrive with the launch itself (__grid_constant__), the CTA finds its frame with
// independent constant-bank reads, and row pair / segment come from a multiply-high.  With RPT = 2
// a lane converts two vertically adjacent row pairs: all eight loads are in flight before the
// first multiply, and NV12 re-uses the first pair's chroma row as the "row above" of the second.
// ================================================================================================
template <int NF>
struct TickParams {
    int32_t n_frames;
    int32_t wait_prior;      // 1: griddepcontrol.wait before touching memory (ordinary stream order);
                             // 2: wait only before the first global store (the previous kernel did not produce our inputs);
                             // 0: the caller declared this launch independent of the previous kernel
    int32_t uniform_ctas;    // > 0: every frame has this many CTAs and the grid is (uniform_ctas, n_frames):
                             //      blockIdx.y is the frame, no search (homogeneous camera fleets)
    int32_t only_fmt;        // host-side only: the single pixel format of the launch, or -1 (selects the instantiation)
    int32_t cta_begin[(NF + 1 + 3) / 4 * 4];   // cta_begin[f] = first CTA of frame f; [n_frames] = grid size
    FrameDev frames[NF];
};

// ONLY >= 0: every frame of the launch has pixel format ONLY (a homogeneous fleet, the common case) and the other
// formats' code is compiled out: a third of the instruction footprint, fewer instruction-cache misses.
// UN: the instantiation also carries the any-alignment loads and stores (PATH_UNALIGNED frames); the all-format kernel
// (ONLY < 0) always does, the per-format kernels exist in both flavours so that the aligned ones stay lean.
template <int NF, int RPT, int MINB, int ONLY = -1, bool UN = false>
__global__ void __launch_bounds__(kThreads, MINB)
yuv_to_bgr24_tick(const __grid_constant__ TickParams<NF> p) {
    __shared__ __align__(128) uint8_t tile[kWarpsPerCta][2][kSegBytes];
    // Programmatic dependent launch: let the next launch on this stream start filling SMs as this
    // one drains (a 16-frame tick lasts ~27 us, of which ~4 us are ramp and tail); unless the caller
    // vouched for independence, wait here until the previous kernel's writes are complete.
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    if (p.wait_prior == 1) asm volatile("griddepcontrol.wait;" ::: "memory");
    int cta = blockIdx.x, fi = 0;
    if (p.uniform_ctas > 0) {
        fi = blockIdx.y;
        cta += p.cta_begin[fi];
    } else {   // mixed fleet: binary search over the CTA prefix in the constant bank (uniform datapath)
        int lo = 0, hi = p.n_frames - 1;
        while (lo < hi) { const int mid = (lo + hi + 1) >> 1; if (cta >= p.cta_begin[mid]) lo = mid; else hi = mid - 1; }
        fi = lo;
    }
    const FrameDev &f = p.frames[fi];
    const int fmt = ONLY >= 0 ? ONLY : f.fmt;
    const int first = (cta - p.cta_begin[fi]) * (f.fast == PATH_GENERIC ? kGenericBlocksPerCta : kWarpsPerCta);

    if (f.fast == PATH_GENERIC) {
        if (p.wait_prior == 2) asm volatile("griddepcontrol.wait;" ::: "memory");   // the scalar path stores as it goes
        const int bw = f.w >> 1, nblk = bw * (f.h >> 1);
        for (int q = first + threadIdx.x; q < min(first + kGenericBlocksPerCta, nblk); q += kThreads)
            generic_block(f, q % bw, q / bw);
        return;
    }

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int segs = f.segs, hp = f.h >> 1;
    const int groups = (hp + RPT - 1) / RPT;               // row-pair groups per segment column
    const int item = first + warp;
    if (item >= segs * groups) return;                      // warp-uniform
    const int grp = f.segs_magic ? (int)__umulhi((uint32_t)item, f.segs_magic) : item;
    const int seg = item - grp * segs;
    const int rp0 = grp * RPT;
    const int x = seg * kSegPx + lane * 16;
    const bool active = x < f.w;
    const int seg_bytes = min(kSegPx, f.w - seg * kSegPx) * 3;   // visible bytes of this segment's rows (even)
    uint8_t *t0 = tile[warp][0], *t1 = tile[warp][1];

    // ---- all loads first ----
    uint4 ya[RPT], yb[RPT], ca[RPT];   // ca: planar {U.lo,U.hi,V.lo,V.hi}; nv12: the 16 interleaved chroma bytes
    uint4 pu = make_uint4(0, 0, 0, 0);
    const bool nv12 = fmt_is_semiplanar(fmt);
    const bool p10 = fmt == FMT_YUV420P10;               // 16-bit containers: loaded per row pair below, not here
    const int cshift = fmt_is_422(fmt) ? 1 : 0;           // 4:2:2: chroma row of luma row j is row j, not j/2
    // planes, strides or output of any alignment (the host sends launches with such frames to an instantiation that
    // carries this code)
    const bool un = (ONLY < 0 || UN) && f.fast == PATH_UNALIGNED;
    const int nb = min(16, f.w - x);                      // visible luma bytes of this lane
    if (active && !p10 && !fmt_vector_only_when_specialised(fmt)) {
#pragma unroll
        for (int r = 0; r < RPT; ++r) {
            const int rp = min(rp0 + r, hp - 1);            // clamp: an odd tail group re-reads the last pair (not stored)
            const uint8_t *py = f.y + (size_t)(2 * rp) * f.sy + x;
            ya[r] = un ? ldg16_any(py, nb) : ld_stream_v4(py);
            yb[r] = un ? ldg16_any(py + f.sy, nb) : ld_stream_v4(py + f.sy);
            if (nv12) {
                const uint8_t *pc = f.c0 + (size_t)rp * f.sc0 + x;
                ca[r] = un ? ldg16_any(pc, nb) : ld_stream_v4(pc);
            } else {
                const uint8_t *pcu = f.c0 + (size_t)(rp << cshift) * f.sc0 + (x >> 1), *pcv = f.c1 + (size_t)(rp << cshift) * f.sc1 + (x >> 1);
                const uint2 uu = un ? ldg8_any(pcu, nb >> 1) : ld_stream_v2(pcu);
                const uint2 vv = un ? ldg8_any(pcv, nb >> 1) : ld_stream_v2(pcv);
                ca[r] = make_uint4(uu.x, uu.y, vv.x, vv.y);
            }
        }
        if (nv12) {
            const uint8_t *pp = f.c0 + (size_t)(rp0 > 0 ? rp0 - 1 : 0) * f.sc0 + x;
            pu = un ? ldg16_any(pp, nb) : __ldg(reinterpret_cast<const uint4 *>(pp));
        }
        if (fmt == FMT_NV21) {                            // V first: swap each pair once, then it is NV12
            pu = swap_pairs(pu);
#pragma unroll
            for (int r = 0; r < RPT; ++r) ca[r] = swap_pairs(ca[r]);
        }
    }

#pragma unroll
    for (int r = 0; r < RPT; ++r) {
        const int rp = rp0 + r;
        if (rp >= hp) break;                                // warp-uniform
        if (ONLY >= 0 && active && fmt == FMT_GRAY8) {   // B = G = R = Y: three byte permutes per four pixels
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                const uint4 yy = ld_stream_v4(f.y + (size_t)(2 * rp + q) * f.sy + x);
                uint8_t *t = q ? t1 : t0;
                uint32_t o[12];
#pragma unroll
                for (int g = 0; g < 4; ++g) {
                    const uint32_t yw = word_of(yy, g);
                    o[3 * g] = prmt(yw, yw, 0x1000); o[3 * g + 1] = prmt(yw, yw, 0x2211); o[3 * g + 2] = prmt(yw, yw, 0x3332);
                }
#pragma unroll
                for (int k = 0; k < 3; ++k) *reinterpret_cast<uint4 *>(t + lane * 48 + k * 16) = make_uint4(o[4 * k], o[4 * k + 1], o[4 * k + 2], o[4 * k + 3]);
            }
        } else if (ONLY >= 0 && active && fmt_is_444(fmt)) {   // compiled into the 4:4:4 instantiations only (register pressure)
            // 4:4:4: three full-resolution planes, one row at a time (park it, then the next)
            uint32_t o[12];
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                const size_t row = (size_t)(2 * rp + q);
                const uint4 yy = ld_stream_v4(f.y + row * f.sy + x), uu4 = ld_stream_v4(f.c0 + row * f.sc0 + x), vv4 = ld_stream_v4(f.c1 + row * f.sc1 + x);
                if (fmt == FMT_YUV444P) compute_444_row<false>(yy, uu4, vv4, o);
                else compute_444_row<true>(yy, uu4, vv4, o);
                uint8_t *t = q ? t1 : t0;
#pragma unroll
                for (int k = 0; k < 3; ++k) *reinterpret_cast<uint4 *>(t + lane * 48 + k * 16) = make_uint4(o[4 * k], o[4 * k + 1], o[4 * k + 2], o[4 * k + 3]);
            }
        } else if (active && p10) {
            // 10-bit: 32 B of luma per row and lane, 16 B per chroma plane; the row above feeds the even row
            const uint8_t *yr = f.y + (size_t)(2 * rp) * f.sy + 2 * x;
            const uint4 a_lo = ld_stream_v4(yr), a_hi = ld_stream_v4(yr + 16), b_lo = ld_stream_v4(yr + f.sy), b_hi = ld_stream_v4(yr + f.sy + 16);
            const int upr = rp > 0 ? rp - 1 : 0;
            const uint4 cu = ld_stream_v4(f.c0 + (size_t)rp * f.sc0 + x), cv = ld_stream_v4(f.c1 + (size_t)rp * f.sc1 + x);
            const uint4 pu10 = __ldg(reinterpret_cast<const uint4 *>(f.c0 + (size_t)upr * f.sc0 + x));
            const uint4 pv10 = __ldg(reinterpret_cast<const uint4 *>(f.c1 + (size_t)upr * f.sc1 + x));
            uint32_t o0[12], o1[12];
            compute_p10(a_lo, a_hi, b_lo, b_hi, cu, cv, pu10, pv10, rp == hp - 1, o0, o1);
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                *reinterpret_cast<uint4 *>(t0 + lane * 48 + k * 16) = make_uint4(o0[4 * k], o0[4 * k + 1], o0[4 * k + 2], o0[4 * k + 3]);
                *reinterpret_cast<uint4 *>(t1 + lane * 48 + k * 16) = make_uint4(o1[4 * k], o1[4 * k + 1], o1[4 * k + 2], o1[4 * k + 3]);
            }
        } else if (active && cshift) {
            // 4:2:2: the odd luma row has a chroma row of its own.  One row at a time (park it, then the next)
            // keeps this secondary format from raising the register count of the 4:2:0 path.
            uint32_t o[12];
            if (fmt == FMT_YUV422P) compute_planar_row<false>(ya[r], make_uint2(ca[r].x, ca[r].y), make_uint2(ca[r].z, ca[r].w), o);
            else compute_planar_row<true>(ya[r], make_uint2(ca[r].x, ca[r].y), make_uint2(ca[r].z, ca[r].w), o);
#pragma unroll
            for (int k = 0; k < 3; ++k) *reinterpret_cast<uint4 *>(t0 + lane * 48 + k * 16) = make_uint4(o[4 * k], o[4 * k + 1], o[4 * k + 2], o[4 * k + 3]);
            const uint8_t *pub = f.c0 + (size_t)(2 * rp + 1) * f.sc0 + (x >> 1), *pvb = f.c1 + (size_t)(2 * rp + 1) * f.sc1 + (x >> 1);
            const uint2 ub = un ? ldg8_any(pub, nb >> 1) : ld_stream_v2(pub);
            const uint2 vb = un ? ldg8_any(pvb, nb >> 1) : ld_stream_v2(pvb);
            if (fmt == FMT_YUV422P) compute_planar_row<false>(yb[r], ub, vb, o);
            else compute_planar_row<true>(yb[r], ub, vb, o);
#pragma unroll
            for (int k = 0; k < 3; ++k) *reinterpret_cast<uint4 *>(t1 + lane * 48 + k * 16) = make_uint4(o[4 * k], o[4 * k + 1], o[4 * k + 2], o[4 * k + 3]);
        } else if (active) {
            uint32_t o0[12], o1[12];
            if (fmt == FMT_YUV420P) compute_planar<false>(ya[r], yb[r], make_uint2(ca[r].x, ca[r].y), make_uint2(ca[r].z, ca[r].w), o0, o1);
            else if (fmt == FMT_YUVJ420P) compute_planar<true>(ya[r], yb[r], make_uint2(ca[r].x, ca[r].y), make_uint2(ca[r].z, ca[r].w), o0, o1);
            else compute_nv12(ya[r], yb[r], ca[r], r == 0 ? pu : ca[r > 0 ? r - 1 : 0], rp == hp - 1, o0, o1);
#pragma unroll
            for (int k = 0; k < 3; ++k) {   // 48 B lane stride: conflict-free per quarter warp
                *reinterpret_cast<uint4 *>(t0 + lane * 48 + k * 16) = make_uint4(o0[4 * k], o0[4 * k + 1], o0[4 * k + 2], o0[4 * k + 3]);
                *reinterpret_cast<uint4 *>(t1 + lane * 48 + k * 16) = make_uint4(o1[4 * k], o1[4 * k + 1], o1[4 * k + 2], o1[4 * k + 3]);
            }
        }
        __syncwarp();
        // "ordered writes": loads and arithmetic of the first row pair overlapped the previous kernel's tail; its
        // writes must be complete before ours land (same output buffer re-used by consecutive launches)
        if (p.wait_prior == 2 && r == 0) asm volatile("griddepcontrol.wait;" ::: "memory");
        uint8_t *g0 = f.dst + (size_t)(2 * rp) * f.sd + (size_t)seg * kSegBytes;
        if (un) { store_segment_any(g0, t0, seg_bytes, lane); store_segment_any(g0 + f.sd, t1, seg_bytes, lane); }
        else { store_segment(g0, t0, seg_bytes, lane); store_segment(g0 + f.sd, t1, seg_bytes, lane); }
        if (RPT > 1) __syncwarp();                          // tile is re-used by the next row pair
    }
}

}  // namespace vep
