/* Pure-C producer over the round-2 entry points of libvepb200.so: a fleet opened as one group
 * (vep_stream_open_many), whole ticks staged with one call (vep_submit_batch + helper threads), ticketed ticks
 * in flight, frames read in place from a registered arena (vep_host_register + VEP_SUBMIT_PINNED), and the
 * decoder-writes-into-the-slot path (vep_acquire_input / vep_commit_input).  Every result is compared with the
 * plain-C oracle; the copy counters prove that a lock-step fleet moves with one copy per direction per tick.
 * Exit code 0 = all good.  Built and run by tests/test_gpu_c_abi.py.
 */
#include "vepb200.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>

int port_yuv420p_to_bgr24(const uint8_t *, int, const uint8_t *, int, const uint8_t *, int, int, int, int, uint8_t *, int);

#define CHECK(x) do { int rc_ = (x); if (rc_ != 0) { fprintf(stderr, "%s -> %d (%s): %s\n", #x, rc_, vep_status_name(rc_), vep_last_error()); return 1; } } while (0)
#define NCAM 4
#define W 1280
#define H 720

static uint32_t rng_state = 777u;
static uint8_t rnd8(void) { rng_state = rng_state * 1664525u + 1013904223u; return (uint8_t)(rng_state >> 16); }

typedef struct { uint8_t *y, *u, *v; int sy, sc; } frame_t;

/* a "decoded frame" with decoder-style padded linesizes inside `arena` (64-byte aligned planes) */
static size_t carve(uint8_t *arena, size_t off, frame_t *f)
{
    f->sy = W + 64; f->sc = W / 2 + 64;
    f->y = arena + off; off += (size_t)f->sy * H;
    f->u = arena + off; off += (size_t)f->sc * (H / 2);
    f->v = arena + off; off += (size_t)f->sc * (H / 2);
    return (off + 63) / 64 * 64;
}
static void fill(frame_t *f)
{
    for (size_t i = 0; i < (size_t)f->sy * H; ++i) f->y[i] = rnd8();
    for (size_t i = 0; i < (size_t)f->sc * (H / 2); ++i) { f->u[i] = rnd8(); f->v[i] = rnd8(); }
}
static int same_as_oracle(vep_ctx *ctx, int id, const frame_t *f, uint8_t *want)
{
    const uint8_t *bgr = NULL; size_t len = 0; int32_t shape[3];
    if (vep_result(ctx, id, &bgr, &len, shape) != 0 || len != (size_t)W * H * 3) return 0;
    if (port_yuv420p_to_bgr24(f->y, f->sy, f->u, f->sc, f->v, f->sc, W, H, 0, want, W * 3) != 0) return 0;
    return memcmp(bgr, want, len) == 0;
}

int main(void)
{
    vep_ctx *ctx = NULL;
    CHECK(vep_create(0, &ctx));
    vep_stream_spec specs[NCAM];
    for (int c = 0; c < NCAM; ++c) { specs[c].stream_id = 10 + c; specs[c].fmt = VEP_FMT_YUV420P; specs[c].width = W; specs[c].height = H; }
    CHECK(vep_stream_open_many(ctx, specs, NCAM, 2));
    CHECK(vep_set_staging_threads(ctx, 2));

    /* the producer's own frame pool: 3 generations x NCAM frames, page-locked once */
    const size_t per_frame = ((size_t)(W + 64) * H + 2 * (size_t)(W / 2 + 64) * (H / 2) + 256);
    const size_t arena_bytes = per_frame * NCAM * 3 + 4096;
    uint8_t *raw = malloc(arena_bytes + 64), *arena = (uint8_t *)(((uintptr_t)raw + 63) & ~(uintptr_t)63);
    frame_t fr[3][NCAM];
    size_t off = 0;
    for (int g = 0; g < 3; ++g) for (int c = 0; c < NCAM; ++c) { off = carve(arena, off, &fr[g][c]); fill(&fr[g][c]); }
    uint8_t *want = malloc((size_t)W * H * 3);
    vep_stats s0, s1;

    /* 1. pageable frames, one vep_submit_batch per tick, two ticks in flight, waited for by ticket */
    CHECK(vep_get_stats(ctx, &s0));
    uint64_t ticket[2];
    for (int g = 0; g < 2; ++g) {
        vep_submit_item items[NCAM];
        memset(items, 0, sizeof items);
        for (int c = 0; c < NCAM; ++c) {
            items[c].stream_id = 10 + c; items[c].convert = VEP_SUBMIT_COPY;
            items[c].plane[0] = fr[g][c].y; items[c].plane[1] = fr[g][c].u; items[c].plane[2] = fr[g][c].v;
            items[c].stride[0] = fr[g][c].sy; items[c].stride[1] = items[c].stride[2] = fr[g][c].sc;
            items[c].meta.timestamp = 100 * g + c; strcpy(items[c].meta.frame_type, "P");
        }
        int done = 0, nconv = 0;
        CHECK(vep_submit_batch(ctx, items, NCAM, &done));
        if (done != NCAM) { fprintf(stderr, "batch staged %d of %d\n", done, NCAM); return 1; }
        CHECK(vep_tick_begin_ticket(ctx, &nconv, &ticket[g]));
        if (nconv != NCAM) { fprintf(stderr, "tick %d converted %d\n", g, nconv); return 1; }
    }
    CHECK(vep_tick_wait_ticket(ctx, ticket[1]));           /* both ticks have landed */
    CHECK(vep_get_stats(ctx, &s1));
    if (s1.copies_in - s0.copies_in != 2 || s1.copies_out - s0.copies_out != 2) {
        fprintf(stderr, "copies: %lld in, %lld out for 2 ticks of a lock-step group\n", (long long)(s1.copies_in - s0.copies_in), (long long)(s1.copies_out - s0.copies_out));
        return 1;
    }
    for (int c = 0; c < NCAM; ++c) if (!same_as_oracle(ctx, 10 + c, &fr[1][c], want)) { fprintf(stderr, "batch: camera %d differs\n", c); return 1; }

    /* 2. the same pool page-locked: frames are read in place, no CPU copy; pageable memory is refused in this mode */
    const uint8_t *pl0[3] = { fr[2][0].y, fr[2][0].u, fr[2][0].v };
    const int32_t st0[3] = { fr[2][0].sy, fr[2][0].sc, fr[2][0].sc };
    if (vep_submit(ctx, 10, pl0, st0, NULL, VEP_SUBMIT_PINNED) != VEP_ERR_INVALID_ARG) { fprintf(stderr, "pageable planes were not refused in pinned mode\n"); return 1; }
    CHECK(vep_host_register(arena, arena_bytes));
    for (int c = 0; c < NCAM; ++c) {
        const uint8_t *pl[3] = { fr[2][c].y, fr[2][c].u, fr[2][c].v };
        const int32_t st[3] = { fr[2][c].sy, fr[2][c].sc, fr[2][c].sc };
        CHECK(vep_submit(ctx, 10 + c, pl, st, NULL, VEP_SUBMIT_PINNED));
    }
    int nconv = 0;
    CHECK(vep_tick(ctx, &nconv));
    for (int c = 0; c < NCAM; ++c) if (!same_as_oracle(ctx, 10 + c, &fr[2][c], want)) { fprintf(stderr, "pinned: camera %d differs\n", c); return 1; }
    CHECK(vep_host_unregister(arena));

    /* 3. the decoder writes into the slot the library hands out */
    for (int c = 0; c < NCAM; ++c) {
        uint8_t *pl[3]; int32_t st[3];
        CHECK(vep_acquire_input(ctx, 10 + c, pl, st));
        for (int r = 0; r < H; ++r) memcpy(pl[0] + (size_t)r * st[0], fr[0][c].y + (size_t)r * fr[0][c].sy, W);
        for (int r = 0; r < H / 2; ++r) {
            memcpy(pl[1] + (size_t)r * st[1], fr[0][c].u + (size_t)r * fr[0][c].sc, W / 2);
            memcpy(pl[2] + (size_t)r * st[2], fr[0][c].v + (size_t)r * fr[0][c].sc, W / 2);
        }
        CHECK(vep_commit_input(ctx, 10 + c, NULL));
    }
    CHECK(vep_tick(ctx, &nconv));
    for (int c = 0; c < NCAM; ++c) if (!same_as_oracle(ctx, 10 + c, &fr[0][c], want)) { fprintf(stderr, "slot: camera %d differs\n", c); return 1; }
    /* ring_depth 2: the previous frame of a camera is still there */
    const uint8_t *older = NULL; size_t len = 0; int32_t shape[3];
    CHECK(vep_result_at(ctx, 10, 1, &older, &len, shape));
    if (port_yuv420p_to_bgr24(fr[2][0].y, fr[2][0].sy, fr[2][0].u, fr[2][0].sc, fr[2][0].v, fr[2][0].sc, W, H, 0, want, W * 3) != 0 || memcmp(older, want, len)) {
        fprintf(stderr, "age 1 of camera 0 differs\n"); return 1;
    }
    for (int c = 0; c < NCAM; ++c) CHECK(vep_stream_close(ctx, 10 + c));
    vep_destroy(ctx);
    free(want); free(raw);
    printf("C_ABI_BATCH_OK %d cameras: batch submit, tickets, pinned pool, decoder-written slots\n", NCAM);
    return 0;
}
