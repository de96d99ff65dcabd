/* TEST INFRASTRUCTURE ONLY -- plain-C restatement of the arithmetic behind the reference's
 * `frame.to_ndarray(format='bgr24')` (reference python/read_image.py:94), which PyAV av=8.0.2
 * (reference python/environment.yml:10) delegates to libswscale.  libswscale is not in the
 * reference tree; the closed forms below are the ones in oracle/formulas.py (formula A, B, C)
 * and are pinned against the live libswscale 9.1.100 of this image by tests/test_oracle.py.
 * Parity status: unpinned by the reference's own tests (it has none for this path); pinned by
 * the live library and tests/golden/.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load the library
 * built from this file (oracle/Makefile -> oracle/_build/libyuv2bgr_port.so).
 */
#include <stddef.h>
#include <stdint.h>

typedef struct { int yC, vr, ub, ug, vg, yOff; } simd_k;
/* formula A coefficient sets (oracle/formulas.py: coeffs_simd) */
static const simd_k K_LIMITED = { 9539, 13075, 16525, -3209, -6660, 128 };
static const simd_k K_FULL    = { 8192, 11485, 14516, -2819, -5850, 0 };
/* formula B, limited range (oracle/formulas.py: _coeffs_c_path) */
enum { C_CY = 76309, C_OY = 1048576, C_CRV = 89830, C_CBU = 113537, C_CGU = -22049, C_CGV = -45756 };

static inline uint8_t sat_u8(int v) { return (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v)); }
/* arithmetic (floor) shift on a possibly negative product, independent of the compiler */
static inline int asr16(int v) { return v >= 0 ? (v >> 16) : -((-v + 65535) >> 16); }

/* SIMD output stage for one row: Uw/Vw per chroma column already in ((u<<3)-1024) form */
static void row_simd(const uint8_t *y, const int *Uw, const int *Vw, uint8_t *dst, int w,
                     const simd_k *k)
{
    for (int x = 0; x < w; x++) {
        int uw = Uw[x >> 1], vw = Vw[x >> 1];
        int cb = asr16(uw * k->ub);
        int cr = asr16(vw * k->vr);
        int cg = asr16(uw * k->ug) + asr16(vw * k->vg);
        int yy = asr16((((int)y[x] << 3) - k->yOff) * k->yC);
        dst[3 * x + 0] = sat_u8(yy + cb);
        dst[3 * x + 1] = sat_u8(yy + cg);
        dst[3 * x + 2] = sat_u8(yy + cr);
    }
}

/* plain-C table stage (formula B) for one row on per-column chroma values u, v */
static void row_cpath(const uint8_t *y, const int *u, const int *v, uint8_t *dst, int w)
{
    const int base = -(384 << 16) - C_OY + 326 * C_CY + 0x8000;
    for (int x = 0; x < w; x++) {
        int uu = u[x >> 1], vv = v[x >> 1];
        int dr = asr16(vv * C_CRV) - (C_CRV >> 9);
        int db = asr16(uu * C_CBU) - (C_CBU >> 9);
        int dg = asr16(uu * C_CGU) - (-((-C_CGU + 511) >> 9)) + asr16(vv * C_CGV) - (-((-C_CGV + 511) >> 9));
        int Y = y[x];
        dst[3 * x + 0] = sat_u8(asr16(base + (Y + db) * C_CY));
        dst[3 * x + 1] = sat_u8(asr16(base + (Y + dg) * C_CY));
        dst[3 * x + 2] = sat_u8(asr16(base + (Y + dr) * C_CY));
    }
}

#define MAXW 8192

/* Formula A.  Returns 0, or -1 on unsupported geometry. */
int port_yuv420p_to_bgr24(const uint8_t *Y, int sy, const uint8_t *U, int su, const uint8_t *V,
                          int sv, int w, int h, int full_range, uint8_t *dst, int sd)
{
    if (w <= 0 || h <= 0 || (w & 1) || (h & 1) || w > MAXW) return -1;
    const simd_k *k = full_range ? &K_FULL : &K_LIMITED;
    int Uw[MAXW / 2], Vw[MAXW / 2];
    for (int j = 0; j < h; j++) {
        const uint8_t *u = U + (size_t)(j >> 1) * su, *v = V + (size_t)(j >> 1) * sv;
        for (int i = 0; i < w / 2; i++) { Uw[i] = (u[i] << 3) - 1024; Vw[i] = (v[i] << 3) - 1024; }
        row_simd(Y + (size_t)j * sy, Uw, Vw, dst + (size_t)j * sd, w, k);
    }
    return 0;
}

/* 4:2:2 planar: luma row j uses chroma row j (oracle/formulas.py: yuv422p_to_bgr24). */
int port_yuv422p_to_bgr24(const uint8_t *Y, int sy, const uint8_t *U, int su, const uint8_t *V,
                          int sv, int w, int h, int full_range, uint8_t *dst, int sd)
{
    if (w <= 0 || h <= 0 || (w & 1) || (h & 1) || w > MAXW) return -1;
    const simd_k *k = full_range ? &K_FULL : &K_LIMITED;
    int Uw[MAXW / 2], Vw[MAXW / 2];
    for (int j = 0; j < h; j++) {
        const uint8_t *u = U + (size_t)j * su, *v = V + (size_t)j * sv;
        for (int i = 0; i < w / 2; i++) { Uw[i] = (u[i] << 3) - 1024; Vw[i] = (v[i] << 3) - 1024; }
        row_simd(Y + (size_t)j * sy, Uw, Vw, dst + (size_t)j * sd, w, k);
    }
    return 0;
}

/* Formula C (limited range).  H >= 8.  vu != 0: NV21 byte order (V first). */
int port_nv12_to_bgr24(const uint8_t *Y, int sy, const uint8_t *UV, int suv, int w, int h,
                       uint8_t *dst, int sd, int vu)
{
    if (w <= 0 || h < 8 || (w & 1) || (h & 1) || w > MAXW) return -1;
    int a0[MAXW / 2], a1[MAXW / 2];
    const int hc = h / 2;
    const int iu = vu ? 1 : 0, iv = vu ? 0 : 1;
    for (int j = 0; j < h; j++) {
        const int k = j >> 1;
        const uint8_t *y = Y + (size_t)j * sy;
        uint8_t *d = dst + (size_t)j * sd;
        if (j < h - 2) {
            const uint8_t *cur = UV + (size_t)k * suv;
            const uint8_t *up = UV + (size_t)(k > 0 ? k - 1 : 0) * suv;
            for (int i = 0; i < w / 2; i++) {
                if (j & 1) {
                    a0[i] = (cur[2 * i + iu] << 3) - 1024;
                    a1[i] = (cur[2 * i + iv] << 3) - 1024;
                } else {
                    a0[i] = ((up[2 * i + iu] + cur[2 * i + iu]) << 2) - 1024;
                    a1[i] = ((up[2 * i + iv] + cur[2 * i + iv]) << 2) - 1024;
                }
            }
            row_simd(y, a0, a1, d, w, &K_LIMITED);
        } else {
            int k0, k1, a;
            if (j & 1) { k0 = k; k1 = k + 1 < hc ? k + 1 : hc - 1; a = 1024; }
            else       { k0 = k - 1; k1 = k; a = 3072; }
            if (k0 == k1) a = 0;
            const uint8_t *r0 = UV + (size_t)k0 * suv, *r1 = UV + (size_t)k1 * suv;
            for (int i = 0; i < w / 2; i++) {
                a0[i] = (r0[2 * i + iu] * (4096 - a) + r1[2 * i + iu] * a + 2048) >> 12;
                a1[i] = (r0[2 * i + iv] * (4096 - a) + r1[2 * i + iv] * a + 2048) >> 12;
            }
            row_cpath(y, a0, a1, d, w);
        }
    }
    return 0;
}

/* yuv420p10le (oracle/formulas.py: yuv420p10_to_bgr24).  Planes hold little-endian 16-bit samples <= 1023;
 * strides in bytes.  H >= 8. */
static void row_simd11(const uint16_t *y, const int *Uw, const int *Vw, uint8_t *dst, int w)
{
    const simd_k *k = &K_LIMITED;
    for (int x = 0; x < w; x++) {
        int uw = Uw[x >> 1], vw = Vw[x >> 1];
        int cb = asr16(uw * k->ub), cr = asr16(vw * k->vr);
        int cg = asr16(uw * k->ug) + asr16(vw * k->vg);
        int yy = asr16((((int)y[x] << 1) - k->yOff) * k->yC);
        dst[3 * x + 0] = sat_u8(yy + cb);
        dst[3 * x + 1] = sat_u8(yy + cg);
        dst[3 * x + 2] = sat_u8(yy + cr);
    }
}

int port_yuv420p10_to_bgr24(const uint8_t *Y, int sy, const uint8_t *U, int su, const uint8_t *V, int sv,
                            int w, int h, uint8_t *dst, int sd)
{
    if (w <= 0 || h < 8 || (w & 1) || (h & 1) || w > MAXW) return -1;
    int a0[MAXW / 2], a1[MAXW / 2];
    uint8_t y8[MAXW];
    const int hc = h / 2;
    for (int j = 0; j < h; j++) {
        const int k = j >> 1;
        const uint16_t *y = (const uint16_t *)(Y + (size_t)j * sy);
        uint8_t *d = dst + (size_t)j * sd;
        if (j < h - 2) {
            const uint16_t *cu = (const uint16_t *)(U + (size_t)k * su), *cv = (const uint16_t *)(V + (size_t)k * sv);
            const uint16_t *pu = (const uint16_t *)(U + (size_t)(k > 0 ? k - 1 : 0) * su), *pv = (const uint16_t *)(V + (size_t)(k > 0 ? k - 1 : 0) * sv);
            for (int i = 0; i < w / 2; i++) {
                if (j & 1) { a0[i] = (cu[i] << 1) - 1024; a1[i] = (cv[i] << 1) - 1024; }
                else       { a0[i] = pu[i] + cu[i] - 1024; a1[i] = pv[i] + cv[i] - 1024; }
            }
            row_simd11(y, a0, a1, d, w);
        } else {
            int k0, k1, a;
            if (j & 1) { k0 = k; k1 = k + 1 < hc ? k + 1 : hc - 1; a = 1024; }
            else       { k0 = k - 1; k1 = k; a = 3072; }
            if (k0 == k1) a = 0;
            const uint16_t *u0 = (const uint16_t *)(U + (size_t)k0 * su), *u1 = (const uint16_t *)(U + (size_t)k1 * su);
            const uint16_t *v0 = (const uint16_t *)(V + (size_t)k0 * sv), *v1 = (const uint16_t *)(V + (size_t)k1 * sv);
            for (int i = 0; i < w / 2; i++) {
                a0[i] = (u0[i] * (4096 - a) + u1[i] * a + 8192) >> 14; if (a0[i] > 255) a0[i] = 255;
                a1[i] = (v0[i] * (4096 - a) + v1[i] * a + 8192) >> 14; if (a1[i] > 255) a1[i] = 255;
            }
            /* formula B takes 8-bit luma; (Y+2)>>2 can reach 256, which row_cpath's closed form handles */
            {
                const int base = -(384 << 16) - C_OY + 326 * C_CY + 0x8000;
                (void)y8;
                for (int x = 0; x < w; x++) {
                    int uu = a0[x >> 1], vv = a1[x >> 1];
                    int dr = asr16(vv * C_CRV) - (C_CRV >> 9);
                    int db = asr16(uu * C_CBU) - (C_CBU >> 9);
                    int dg = asr16(uu * C_CGU) - (-((-C_CGU + 511) >> 9)) + asr16(vv * C_CGV) - (-((-C_CGV + 511) >> 9));
                    int Yv = (y[x] + 2) >> 2;
                    d[3 * x + 0] = sat_u8(asr16(base + (Yv + db) * C_CY));
                    d[3 * x + 1] = sat_u8(asr16(base + (Yv + dg) * C_CY));
                    d[3 * x + 2] = sat_u8(asr16(base + (Yv + dr) * C_CY));
                }
            }
        }
    }
    return 0;
}

/* 4:4:4 planar (oracle/formulas.py: yuv444p_to_bgr24): swscale's plain-C full-chroma stage, wrapping 32-bit sums. */
int port_yuv444p_to_bgr24(const uint8_t *Y, int sy, const uint8_t *U, int su, const uint8_t *V, int sv,
                          int w, int h, int full_range, uint8_t *dst, int sd)
{
    if (w <= 0 || h <= 0 || (w & 1) || (h & 1)) return -1;
    const simd_k *k = full_range ? &K_FULL : &K_LIMITED;
    for (int j = 0; j < h; j++) {
        const uint8_t *y = Y + (size_t)j * sy, *u = U + (size_t)j * su, *v = V + (size_t)j * sv;
        uint8_t *d = dst + (size_t)j * sd;
        for (int x = 0; x < w; x++) {
            uint32_t yt = (uint32_t)((((int)y[x] << 9) - (k->yOff << 6)) * k->yC + (1 << 21));
            int u9 = ((int)u[x] - 128) << 9, v9 = ((int)v[x] - 128) << 9;
            int32_t b = (int32_t)(yt + (uint32_t)(u9 * k->ub));
            int32_t g = (int32_t)(yt + (uint32_t)(u9 * k->ug) + (uint32_t)(v9 * k->vg));
            int32_t r = (int32_t)(yt + (uint32_t)(v9 * k->vr));
            d[3 * x + 0] = sat_u8(b >= 0 ? b >> 22 : -1);
            d[3 * x + 1] = sat_u8(g >= 0 ? g >> 22 : -1);
            d[3 * x + 2] = sat_u8(r >= 0 ? r >> 22 : -1);
        }
    }
    return 0;
}
