/* Pure-C caller of libvepb200.so: no Python, no torch.  Opens camera streams of every pixel format,
 * submits synthetic frames with decoder-style padded linesizes, runs one tick, and compares every
 * output byte with the plain-C oracle (oracle/_build/libyuv2bgr_port.so).  Also checks the packed
 * VideoFrame framing against vep_videoframe_frame.  Exit code 0 = all good.
 *
 * Built and run by tests/test_gpu_c_abi.py:
 *   gcc -std=c99 -I include tests/c/abi_roundtrip.c -L video_edge_ai_proxy_b200 -lvepb200 -L oracle/_build -lyuv2bgr_port
 */
#include "vepb200.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* oracle (test infrastructure) */
int port_yuv420p_to_bgr24(const uint8_t *, int, const uint8_t *, int, const uint8_t *, int, int, int, int, uint8_t *, int);
int port_yuv422p_to_bgr24(const uint8_t *, int, const uint8_t *, int, const uint8_t *, int, int, int, int, uint8_t *, int);
int port_yuv444p_to_bgr24(const uint8_t *, int, const uint8_t *, int, const uint8_t *, int, int, int, int, uint8_t *, int);
int port_yuv420p10_to_bgr24(const uint8_t *, int, const uint8_t *, int, const uint8_t *, int, int, int, uint8_t *, int);
int port_nv12_to_bgr24(const uint8_t *, int, const uint8_t *, int, int, int, uint8_t *, int, int);

static uint32_t rng_state = 12345u;
static uint32_t rnd(void) { rng_state = rng_state * 1664525u + 1013904223u; return rng_state >> 8; }

#define CHECK(x) do { int rc_ = (x); if (rc_ != 0) { fprintf(stderr, "%s -> %d (%s): %s\n", #x, rc_, vep_status_name(rc_), vep_last_error()); return 1; } } while (0)

typedef struct { int fmt, w, h; } cam_t;

int main(void)
{
    const cam_t cams[] = { {VEP_FMT_YUV420P, 1920, 1080}, {VEP_FMT_YUVJ420P, 640, 480}, {VEP_FMT_NV12, 1280, 720}, {VEP_FMT_NV21, 640, 480},
                           {VEP_FMT_YUV422P, 640, 480}, {VEP_FMT_YUVJ422P, 1280, 720}, {VEP_FMT_YUV420P10LE, 640, 480},
                           {VEP_FMT_YUV444P, 640, 480}, {VEP_FMT_YUVJ444P, 50, 6}, {VEP_FMT_YUV420P, 50, 6} };
    const int n = (int)(sizeof cams / sizeof cams[0]);
    uint8_t *planes[10][3]; int strides[10][3]; int rows[10][3], rowbytes[10][3];
    vep_ctx *ctx = NULL;
    CHECK(vep_create(0, &ctx));
    for (int c = 0; c < n; ++c) {
        const int fmt = cams[c].fmt, w = cams[c].w, h = cams[c].h;
        const int bps = fmt == VEP_FMT_YUV420P10LE ? 2 : 1;
        const int semi = fmt == VEP_FMT_NV12 || fmt == VEP_FMT_NV21;
        const int f444 = fmt == VEP_FMT_YUV444P || fmt == VEP_FMT_YUVJ444P, f422 = fmt == VEP_FMT_YUV422P || fmt == VEP_FMT_YUVJ422P;
        rowbytes[c][0] = w * bps; rows[c][0] = h;
        rowbytes[c][1] = (semi || f444 ? w : w / 2) * bps; rows[c][1] = (f444 || f422) ? h : h / 2;
        rowbytes[c][2] = semi ? 0 : rowbytes[c][1]; rows[c][2] = semi ? 0 : rows[c][1];
        for (int p = 0; p < 3; ++p) {
            planes[c][p] = NULL; strides[c][p] = 0;
            if (!rows[c][p]) continue;
            strides[c][p] = rowbytes[c][p] + 32;                       /* decoder-style linesize padding */
            planes[c][p] = malloc((size_t)strides[c][p] * rows[c][p]);
            for (int r = 0; r < rows[c][p]; ++r)
                for (int x = 0; x < strides[c][p]; ++x) {
                    uint8_t v = (uint8_t)rnd();
                    if (bps == 2 && (x & 1)) v &= 3;                   /* high byte of a 10-bit sample */
                    planes[c][p][(size_t)r * strides[c][p] + x] = v;
                }
        }
        CHECK(vep_stream_open(ctx, 100 + c, fmt, w, h, 1));
        vep_frame_meta m; memset(&m, 0, sizeof m);
        m.timestamp = 1000 + c; m.pts = c; m.dts = c; m.packet = c; m.keyframe = 1; m.time_base = 1.0 / 90000; m.is_keyframe = 1;
        strcpy(m.frame_type, "I");
        const uint8_t *pl[3] = { planes[c][0], planes[c][1], planes[c][2] };
        CHECK(vep_submit(ctx, 100 + c, pl, strides[c], &m, 1));
    }
    int converted = 0;
    CHECK(vep_tick(ctx, &converted));
    if (converted != n) { fprintf(stderr, "converted %d of %d\n", converted, n); return 1; }
    for (int c = 0; c < n; ++c) {
        const int fmt = cams[c].fmt, w = cams[c].w, h = cams[c].h;
        const uint8_t *bgr = NULL; size_t len = 0; int32_t shape[3];
        CHECK(vep_result(ctx, 100 + c, &bgr, &len, shape));
        if (len != (size_t)w * h * 3 || shape[0] != h || shape[1] != w || shape[2] != 3) { fprintf(stderr, "cam %d: bad shape\n", c); return 1; }
        uint8_t *want = malloc(len);
        int rc = -1;
        switch (fmt) {
            case VEP_FMT_YUV420P: case VEP_FMT_YUVJ420P:
                rc = port_yuv420p_to_bgr24(planes[c][0], strides[c][0], planes[c][1], strides[c][1], planes[c][2], strides[c][2], w, h, fmt == VEP_FMT_YUVJ420P, want, w * 3); break;
            case VEP_FMT_YUV422P: case VEP_FMT_YUVJ422P:
                rc = port_yuv422p_to_bgr24(planes[c][0], strides[c][0], planes[c][1], strides[c][1], planes[c][2], strides[c][2], w, h, fmt == VEP_FMT_YUVJ422P, want, w * 3); break;
            case VEP_FMT_YUV444P: case VEP_FMT_YUVJ444P:
                rc = port_yuv444p_to_bgr24(planes[c][0], strides[c][0], planes[c][1], strides[c][1], planes[c][2], strides[c][2], w, h, fmt == VEP_FMT_YUVJ444P, want, w * 3); break;
            case VEP_FMT_YUV420P10LE:
                rc = port_yuv420p10_to_bgr24(planes[c][0], strides[c][0], planes[c][1], strides[c][1], planes[c][2], strides[c][2], w, h, want, w * 3); break;
            case VEP_FMT_NV12: case VEP_FMT_NV21:
                rc = port_nv12_to_bgr24(planes[c][0], strides[c][0], planes[c][1], strides[c][1], w, h, want, w * 3, fmt == VEP_FMT_NV21); break;
        }
        if (rc != 0) { fprintf(stderr, "cam %d: oracle refused\n", c); return 1; }
        if (memcmp(bgr, want, len) != 0) { fprintf(stderr, "cam %d (fmt %d %dx%d): GPU bytes differ from the oracle\n", c, fmt, w, h); return 1; }
        /* packed message = prefix | payload | suffix */
        const uint8_t *msg = NULL; size_t mlen = 0;
        CHECK(vep_result_packed(ctx, 100 + c, &msg, &mlen));
        vep_frame_meta m; memset(&m, 0, sizeof m);
        m.timestamp = 1000 + c; m.pts = c; m.dts = c; m.packet = c; m.keyframe = 1; m.time_base = 1.0 / 90000; m.is_keyframe = 1;
        strcpy(m.frame_type, "I");
        uint8_t pre[32], suf[192]; int32_t pl_ = 0, sl_ = 0;
        CHECK(vep_videoframe_frame(&m, w, h, pre, &pl_, suf, &sl_));
        if (mlen != (size_t)pl_ + len + sl_ || memcmp(msg, pre, pl_) || memcmp(msg + pl_, want, len) || memcmp(msg + pl_ + len, suf, sl_)) {
            fprintf(stderr, "cam %d: packed message differs\n", c); return 1;
        }
        free(want);
    }
    vep_stats st;
    CHECK(vep_get_stats(ctx, &st));
    if (st.kernel_launches != 1 || st.frames_converted != n) { fprintf(stderr, "stats: %lld launches, %lld frames\n", (long long)st.kernel_launches, (long long)st.frames_converted); return 1; }
    /* errors are codes, not aborts */
    if (vep_stream_open(ctx, 999, VEP_FMT_NV12, 640, 6, 1) != VEP_ERR_GEOMETRY) { fprintf(stderr, "nv12 height 6 was not refused\n"); return 1; }
    if (vep_stream_open(ctx, 100, VEP_FMT_YUV420P, 64, 64, 1) != VEP_ERR_STATE) { fprintf(stderr, "duplicate stream id was not refused\n"); return 1; }
    vep_destroy(ctx);
    printf("C_ABI_ROUNDTRIP_OK %d cameras, 1 launch\n", n);
    return 0;
}
