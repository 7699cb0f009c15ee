/*
 * CPU oracle for the segmentation hot path -- plain C restatement.
 *
 * TEST INFRASTRUCTURE, NOT PRODUCT.  Linked/loaded only by tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / `--impl reference` legs (as the checker or the timed CPU
 * baseline).  The product library (libssb200.so) never links or calls it.
 *
 * Parity: these functions restate oracle/segpath.py one for one and are checked against it
 * and against tests/golden/ in tests/test_oracle.py.  Pinned / unpinned status is the same
 * as stated in the header of oracle/segpath.py: pre-process and foreground mask are pinned by
 * fixtures produced by the reference's own code; argmax + per-label sums are a north-star
 * restatement ("parity unpinned" by the reference, anchored on numpy semantics).
 *
 * Build: oracle/build_oracle.sh  (gcc -O3 -shared -fPIC -pthread) -> oracle/_build/liboracle.so
 */
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* fp16 bits -> float, exact (no hardware half support assumed). */
static float half_bits_to_float(uint16_t h) {
    uint32_t sign = (uint32_t)(h & 0x8000u) << 16;
    uint32_t exp = (h >> 10) & 0x1fu;
    uint32_t man = h & 0x3ffu;
    uint32_t bits;
    if (exp == 0) {
        if (man == 0) {
            bits = sign;
        } else { /* subnormal: renormalise */
            int e = -1;
            do { man <<= 1; ++e; } while (!(man & 0x400u));
            bits = sign | ((uint32_t)(127 - 15 - e) << 23) | ((man & 0x3ffu) << 13);
        }
    } else if (exp == 31) {
        bits = sign | 0x7f800000u | (man << 13);
    } else {
        bits = sign | ((exp + 112u) << 23) | (man << 13);
    }
    float f;
    memcpy(&f, &bits, 4);
    return f;
}

/* value * 2^24 of a finite fp16 as an exact integer (oracle/segpath.py winning_logit_fixed). */
static int64_t half_bits_to_fixed24(uint16_t h) {
    int64_t exp = (h >> 10) & 0x1f, man = h & 0x3ff, v;
    if (exp == 0) v = man;                       /* man * 2^-24 * 2^24 */
    else v = (man | 0x400) << (exp - 1);         /* (1.man) * 2^(exp-15) * 2^24 = (man|0x400) << (exp-1) */
    return (h & 0x8000u) ? -v : v;
}

/* One frame: first-index argmax over C (np.argmax) and per-label {n, sum x, sum y, sum q}
 * over [:roi_h,:roi_w]  (arithmetic of sem_seg_server.py:102-122 restated per label). */
static void frame_argmax_stats_u8(const uint8_t *q, int H, int W, int C, int roi_h, int roi_w,
                                  uint8_t *labels, uint64_t *stats) {
    memset(stats, 0, (size_t)C * 4 * sizeof(uint64_t));
    for (int y = 0; y < H; ++y) {
        for (int x = 0; x < W; ++x) {
            const uint8_t *p = q + ((size_t)y * W + x) * C;
            int best = 0;
            uint8_t bv = p[0];
            for (int c = 1; c < C; ++c)
                if (p[c] > bv) { bv = p[c]; best = c; }
            if (labels) labels[(size_t)y * W + x] = (uint8_t)best;
            if (y < roi_h && x < roi_w) {
                uint64_t *s = stats + (size_t)best * 4;
                s[0] += 1; s[1] += (uint64_t)x; s[2] += (uint64_t)y; s[3] += bv;
            }
        }
    }
}

static void frame_argmax_stats_f16(const uint16_t *q, int H, int W, int C, int roi_h, int roi_w,
                                   uint8_t *labels, uint64_t *stats) {
    memset(stats, 0, (size_t)C * 4 * sizeof(uint64_t));
    for (int y = 0; y < H; ++y) {
        for (int x = 0; x < W; ++x) {
            const uint16_t *p = q + ((size_t)y * W + x) * C;
            int best = 0;
            float bv = half_bits_to_float(p[0]);
            for (int c = 1; c < C; ++c) {
                float v = half_bits_to_float(p[c]);
                if (v > bv) { bv = v; best = c; }   /* strict >: first index wins, -0 == +0 */
            }
            if (labels) labels[(size_t)y * W + x] = (uint8_t)best;
            if (y < roi_h && x < roi_w) {
                uint64_t *s = stats + (size_t)best * 4;
                s[0] += 1; s[1] += (uint64_t)x; s[2] += (uint64_t)y;
                s[3] += (uint64_t)half_bits_to_fixed24(p[best]);
            }
        }
    }
}

/* sem_seg_server.py:150-161 with host-built NEAREST tables: BGR->RGB gather + zero pad. */
static void frame_preprocess(const uint8_t *bgr, int Wc, const int32_t *xtab, const int32_t *ytab,
                             int new_w, int new_h, uint8_t *out, int H, int W) {
    memset(out, 0, (size_t)H * W * 3);
    for (int y = 0; y < new_h; ++y) {
        const uint8_t *row = bgr + (size_t)ytab[y] * Wc * 3;
        uint8_t *o = out + (size_t)y * W * 3;
        for (int x = 0; x < new_w; ++x) {
            const uint8_t *s = row + (size_t)xtab[x] * 3;
            o[3 * x + 0] = s[2]; o[3 * x + 1] = s[1]; o[3 * x + 2] = s[0];
        }
    }
}

static int reflect101(int i, int n) {
    if (n == 1) return 0;
    while (i < 0 || i >= n) i = i < 0 ? -i : 2 * (n - 1) - i;
    return i;
}

/* sem_seg_server.py:173 + :79-85 as integers: colormap -> 3x3 box (REFLECT_101) ->
 * round(s/9) -> fixed-point luma with swapped channel roles -> > 127. */
static void frame_fg_mask(const uint8_t *lab, int h, int w, int ld, const uint8_t *cmap /*[256][3]*/,
                          uint8_t *mask, int ldm) {
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) {
            int s[3] = {0, 0, 0};
            for (int dy = -1; dy <= 1; ++dy)
                for (int dx = -1; dx <= 1; ++dx) {
                    const uint8_t *c = cmap + 3 * lab[(size_t)reflect101(y + dy, h) * ld + reflect101(x + dx, w)];
                    s[0] += c[0]; s[1] += c[1]; s[2] += c[2];
                }
            int b0 = (2 * s[0] + 9) / 18, b1 = (2 * s[1] + 9) / 18, b2 = (2 * s[2] + 9) / 18;
            int gray = (b0 * 1868 + b1 * 9617 + b2 * 4899 + 8192) >> 14;
            mask[(size_t)y * ldm + x] = gray > 127 ? 255 : 0;
        }
}

/* ---- frame-sharded thread pool (frames are independent: sem_seg_server.py:145-195) ----
 * Workers are created once and kept (a timed CPU baseline should not pay pthread_create per batch);
 * frames are handed out one at a time from a shared counter, the calling thread works too. */
typedef struct {
    int kind;            /* 0 u8 argmax, 1 f16 argmax, 2 preprocess */
    const void *in; void *out; uint64_t *stats;
    int B, H, W, C, roi_h, roi_w;
    int Hc, Wc, new_w, new_h; const int32_t *xtab, *ytab; int64_t in_stride;
} job_t;

static void do_frame(const job_t *j, int b) {
    size_t px = (size_t)j->H * j->W;
    if (j->kind == 3) {   /* copy chunk b of 64 KB: whichever thread takes it touches the destination pages first */
        size_t off = (size_t)b << 16, n = (size_t)j->in_stride - off;
        memcpy((uint8_t *)j->out + off, (const uint8_t *)j->in + off, n < 65536 ? n : 65536);
        return;
    }
    if (j->kind == 0)
        frame_argmax_stats_u8((const uint8_t *)j->in + (size_t)b * px * j->C, j->H, j->W, j->C, j->roi_h,
                              j->roi_w, j->out ? (uint8_t *)j->out + (size_t)b * px : NULL,
                              j->stats + (size_t)b * j->C * 4);
    else if (j->kind == 1)
        frame_argmax_stats_f16((const uint16_t *)j->in + (size_t)b * px * j->C, j->H, j->W, j->C, j->roi_h,
                               j->roi_w, j->out ? (uint8_t *)j->out + (size_t)b * px : NULL,
                               j->stats + (size_t)b * j->C * 4);
    else
        frame_preprocess((const uint8_t *)j->in + (size_t)b * j->in_stride, j->Wc, j->xtab, j->ytab,
                         j->new_w, j->new_h, (uint8_t *)j->out + (size_t)b * px * 3, j->H, j->W);
}

#define POOL_MAX 512
static struct {
    pthread_mutex_t mu;
    pthread_cond_t work, done;
    pthread_t tid[POOL_MAX];
    int n_workers;          /* threads created so far */
    int helpers;            /* workers invited to the current job: ids < helpers */
    unsigned generation;    /* bumped per job */
    int running;            /* invited workers that have not finished the current job yet */
    const job_t *job;
    int next_frame;         /* next frame to hand out (atomic) */
} g_pool = {PTHREAD_MUTEX_INITIALIZER, PTHREAD_COND_INITIALIZER, PTHREAD_COND_INITIALIZER, {0}, 0, 0, 0, 0, NULL, 0};
static pthread_mutex_t g_call_mu = PTHREAD_MUTEX_INITIALIZER;   /* one sharded call at a time */

static void drain(const job_t *j) {
    for (;;) {
        int b = __atomic_fetch_add(&g_pool.next_frame, 1, __ATOMIC_RELAXED);
        if (b >= j->B) return;
        do_frame(j, b);
    }
}

static void *pool_worker(void *arg) {
    const int id = (int)(intptr_t)arg;
    unsigned seen = 0;
    pthread_mutex_lock(&g_pool.mu);
    for (;;) {
        while (g_pool.generation == seen || id >= g_pool.helpers) {
            if (g_pool.generation != seen) seen = g_pool.generation;   /* not invited to this one */
            pthread_cond_wait(&g_pool.work, &g_pool.mu);
        }
        seen = g_pool.generation;
        const job_t *j = g_pool.job;
        pthread_mutex_unlock(&g_pool.mu);
        drain(j);
        pthread_mutex_lock(&g_pool.mu);
        if (--g_pool.running == 0) pthread_cond_signal(&g_pool.done);
    }
    return NULL;
}

static int run_sharded(job_t *j, int B, int threads) {
    j->B = B;
    if (threads > B) threads = B;
    if (threads > POOL_MAX + 1) threads = POOL_MAX + 1;
    if (threads <= 1) {
        for (int b = 0; b < B; ++b) do_frame(j, b);
        return 0;
    }
    pthread_mutex_lock(&g_call_mu);
    pthread_mutex_lock(&g_pool.mu);
    while (g_pool.n_workers < threads - 1) {
        if (pthread_create(&g_pool.tid[g_pool.n_workers], NULL, pool_worker, (void *)(intptr_t)g_pool.n_workers)) break;
        pthread_detach(g_pool.tid[g_pool.n_workers]);
        ++g_pool.n_workers;
    }
    g_pool.helpers = threads - 1 < g_pool.n_workers ? threads - 1 : g_pool.n_workers;
    g_pool.running = g_pool.helpers;
    g_pool.job = j;
    __atomic_store_n(&g_pool.next_frame, 0, __ATOMIC_RELAXED);
    ++g_pool.generation;
    pthread_cond_broadcast(&g_pool.work);
    pthread_mutex_unlock(&g_pool.mu);
    drain(j);
    pthread_mutex_lock(&g_pool.mu);
    while (g_pool.running > 0) pthread_cond_wait(&g_pool.done, &g_pool.mu);
    pthread_mutex_unlock(&g_pool.mu);
    pthread_mutex_unlock(&g_call_mu);
    return 0;
}

int oracle_argmax_stats_u8(const uint8_t *logits, int B, int H, int W, int C, int roi_h, int roi_w,
                           uint8_t *labels_or_null, uint64_t *stats, int threads) {
    job_t j; memset(&j, 0, sizeof j);
    j.kind = 0; j.in = logits; j.out = labels_or_null; j.stats = stats;
    j.H = H; j.W = W; j.C = C; j.roi_h = roi_h; j.roi_w = roi_w;
    return run_sharded(&j, B, threads);
}

int oracle_argmax_stats_f16(const uint16_t *logits, int B, int H, int W, int C, int roi_h, int roi_w,
                            uint8_t *labels_or_null, uint64_t *stats, int threads) {
    job_t j; memset(&j, 0, sizeof j);
    j.kind = 1; j.in = logits; j.out = labels_or_null; j.stats = stats;
    j.H = H; j.W = W; j.C = C; j.roi_h = roi_h; j.roi_w = roi_w;
    return run_sharded(&j, B, threads);
}

int oracle_preprocess_nearest_u8(const uint8_t *frames_bgr, int B, int Hc, int Wc, int64_t frame_stride,
                                 const int32_t *xtab, const int32_t *ytab, int new_w, int new_h,
                                 uint8_t *out_rgb, int H, int W, int threads) {
    job_t j; memset(&j, 0, sizeof j);
    j.kind = 2; j.in = frames_bgr; j.out = out_rgb; j.H = H; j.W = W;
    j.Hc = Hc; j.Wc = Wc; j.new_w = new_w; j.new_h = new_h; j.xtab = xtab; j.ytab = ytab;
    j.in_stride = frame_stride;
    return run_sharded(&j, B, threads);
}

int oracle_fg_mask_u8(const uint8_t *labels, int B, int H, int W, int roi_h, int roi_w,
                      const uint8_t *colormap, uint8_t *mask) {
    for (int b = 0; b < B; ++b) {
        uint8_t *m = mask + (size_t)b * H * W;
        memset(m, 0, (size_t)H * W);
        frame_fg_mask(labels + (size_t)b * H * W, roi_h, roi_w, W, colormap, m, W);
    }
    return 0;
}

/* Copies `bytes` from src to dst in 64 KB chunks handed out to the pooled threads.  With dst freshly allocated (pages
 * not yet touched) its pages end up spread over the memory nodes the threads run on, instead of all on the node of the
 * thread that generated the data: the timed CPU baseline then reads at the machine's average bandwidth, whichever
 * node the generating thread happened to sit on (the figure was bimodal, +-15 %, without this). */
int oracle_interleaved_copy(uint8_t *dst, const uint8_t *src, int64_t bytes, int threads) {
    job_t j; memset(&j, 0, sizeof j);
    j.kind = 3; j.in = src; j.out = dst; j.in_stride = bytes;
    return run_sharded(&j, (int)((bytes + 65535) >> 16), threads);
}
