/*
 * ebcc_codec.c -- host side of the B200-native EBCC codec (plain C).  Mirrors the driver of the
 * reference (src/ebcc_codec.c) at its API boundary -- validation, env vars, 48-byte "EBCC" frame
 * header, 80-byte "EBCK" chunk container, buffer ownership -- and hands every numerical stage to
 * the CUDA layer through include/ebcc_cuda.h.  There is no CPU implementation of any stage here:
 * without a CUDA device every encode/decode call logs a fatal error and returns 0.
 */
#include "../../include/ebcc_codec.h"
#include "../../include/ebcc_cuda.h"
#include <assert.h>
#include <math.h>
#include <pthread.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <unistd.h>

/* ---- logging: levels as rxi log.c (reference src/log/log.h), EBCC_LOG_LEVEL (src/ebcc_codec.c:430-448) */
enum { LOG_TRACE, LOG_DEBUG, LOG_INFO, LOG_WARN, LOG_ERROR, LOG_FATAL };
static int g_log_level = LOG_WARN;
static const char *const k_level_names[] = {"TRACE", "DEBUG", "INFO", "WARN", "ERROR", "FATAL"};

static void ebcc_log(int level, const char *fmt, ...)
{
    if (level < g_log_level) return;
    va_list ap;
    fprintf(stderr, "%-5s ebcc_b200: ", k_level_names[level]);
    va_start(ap, fmt);
    vfprintf(stderr, fmt, ap);
    va_end(ap);
    fputc('\n', stderr);
}

void log_set_level_from_env(void)
{
    const char *env = getenv("EBCC_LOG_LEVEL");
    g_log_level = LOG_WARN;
    if (env) {
        char *end;
        long v = strtol(env, &end, 10);
        if (*end != '\0') ebcc_log(LOG_WARN, "Ignore log level: %s, should be in [0, 5]", env);
        else g_log_level = (int)v;
    }
}

static const char *const k_residual_names[] = {"NONE", "MAX_ERROR", "RELATIVE_ERROR"};

void print_config(codec_config_t *config)
{
    ebcc_log(LOG_INFO, "dimensions:\t(%lu, %lu, %lu)", config->dims[0], config->dims[1], config->dims[2]);
    ebcc_log(LOG_INFO, "chunk dimensions:\t(%lu, %lu, %lu)", config->chunk_dims[0], config->chunk_dims[1], config->chunk_dims[2]);
    ebcc_log(LOG_INFO, "base_cr:\t%f", config->base_cr);
    if ((unsigned)config->residual_compression_type <= RELATIVE_ERROR)
        ebcc_log(LOG_INFO, "residual type:\t%s", k_residual_names[config->residual_compression_type]);
    if (config->residual_compression_type == MAX_ERROR) ebcc_log(LOG_INFO, "max error:\t%f", config->error);
    if (config->residual_compression_type == RELATIVE_ERROR) ebcc_log(LOG_INFO, "relative error:\t%f", config->error);
}

void free_buffer(void *buffer) { if (buffer) free(buffer); }

/* ---- serialised headers (reference src/ebcc_codec.c:190-217) -------------------------------------- */
typedef struct {
    uint8_t magic[4];
    uint8_t version;
    uint8_t flags;
    uint16_t reserved;
    uint32_t minval_bits;
    uint32_t maxval_bits;
    uint64_t coeffs_size;
    uint32_t residual_minval_bits;
    uint32_t residual_maxval_bits;
    uint64_t compressed_size;
    uint64_t tail_size;
} ebcc_header_t;

typedef struct {
    uint8_t magic[4];
    uint32_t version;
    uint32_t ndims;
    uint32_t reserved;
    uint64_t dims[NDIMS];
    uint64_t chunk_dims[NDIMS];
    uint64_t num_chunks;
    uint64_t chunk_size;
} ebcc_chunking_header_t;

_Static_assert(sizeof(ebcc_header_t) == 48, "EBCC header must be 48 bytes");
_Static_assert(sizeof(ebcc_chunking_header_t) == 80, "EBCK header must be 80 bytes");

static uint32_t float_to_bits(float v) { uint32_t b; memcpy(&b, &v, 4); return b; }
static float bits_to_float(uint32_t b) { float v; memcpy(&v, &b, 4); return v; }

static int checked_mul(size_t a, size_t b, size_t *out)
{
    if (a != 0 && b > SIZE_MAX / a) return 0;
    *out = a * b;
    return 1;
}

/* reference dims_are_valid (src/ebcc_codec.c:286-297) */
static int dims_are_valid(const size_t dims[NDIMS])
{
    size_t h = 1;
    for (size_t i = 0; i < NDIMS - 1; ++i)
        if (dims[i] == 0 || !checked_mul(h, dims[i], &h)) return 0;
    return h >= EBCC_MIN_INTERNAL_IMAGE_DIM && h <= EBCC_MAX_INTERNAL_IMAGE_DIM &&
           dims[NDIMS - 1] >= EBCC_MIN_INTERNAL_IMAGE_DIM && dims[NDIMS - 1] <= EBCC_MAX_INTERNAL_IMAGE_DIM;
}

/* ---- process-wide pool of device pipelines (context + stream each), created lazily and re-created after fork.
 * Every encode / decode call (and every worker of a multi-batch call) owns one pipeline while it runs, so callers on
 * several threads -- or several batches of one call -- are in flight on the device at the same time: the kernels of
 * one batch fill the issue slots the latency-bound tier-1 coder of another leaves idle.  EBCC_PIPELINES (1..32,
 * default 16) bounds how many PER GPU.
 *
 * The pool spans GPUs: frames are independent (the reference's frame loop, src/ebcc_codec.c:1007-1046), so the batches of
 * one call are dealt to the pipelines of every device of EBCC_CUDA_DEVICES ("0,1,3", or "all"); slot k lives on device
 * k mod n, hence the first n workers of a call land on n different GPUs.  No data travels between GPUs: each batch is
 * uploaded to, coded on and downloaded from its own device, the per-frame sizes meet on the host where the EBCK container
 * is laid out by a prefix sum.  EBCC_CUDA_DEVICE (one ordinal) keeps a process on one GPU, which is what a launcher with
 * one process per GPU sets; with neither variable every visible GPU is used. */
/* More than 8 streams need more than the default 8 hardware queues (16 pipelines on 8: one 768-frame call 9.5 instead of
 * 12.4 GB/s).  The driver reads the variable when CUDA starts in the process, and the CUDA runtime linked into this library
 * starts it from a constructor of its own when the library is loaded -- so this one has to run first.  Too late, and
 * harmless, if the application has started CUDA already; an existing value is respected. */
__attribute__((constructor(101))) static void ebcc_more_hardware_queues(void) { setenv("CUDA_DEVICE_MAX_CONNECTIONS", "32", 0); }

#define EBCC_MAX_PIPELINES 32
/* 16 batches of 24 frames in flight per GPU is where a B200 stops getting faster on 721 x 1440 frames (about 3.2 GB of work
 * buffers per pipeline, allocated when a pipeline is first used: a call of one batch builds one) */
#define EBCC_DEFAULT_PIPELINES 16
#define EBCC_DEFAULT_BATCH_FRAMES 24
#define EBCC_MAX_DEVICES 16
#define EBCC_MAX_SLOTS (EBCC_MAX_PIPELINES * EBCC_MAX_DEVICES)
static pthread_mutex_t g_mu = PTHREAD_MUTEX_INITIALIZER;
static pthread_cond_t g_cv = PTHREAD_COND_INITIALIZER;
static ebcc_cuda_ctx *g_pool[EBCC_MAX_SLOTS];
static int g_busy[EBCC_MAX_SLOTS];
static int g_expected;     /* workers the running multi-batch calls are about to start (they may not have a pipeline yet) */
static pid_t g_ctx_pid = 0;
static int g_devs[EBCC_MAX_DEVICES], g_ndev = 0;   /* device ordinals the pool spans; resolved once per process */

/* caller holds g_mu */
static void resolve_devices(void)
{
    if (g_ndev > 0) return;
    const char *list = getenv("EBCC_CUDA_DEVICES"), *one = getenv("EBCC_CUDA_DEVICE");
    if (list && *list && strcmp(list, "all") != 0) {
        const char *p = list;
        while (*p && g_ndev < EBCC_MAX_DEVICES) {
            char *end = NULL;
            const long v = strtol(p, &end, 10);
            if (end == p) break;
            if (v >= 0) g_devs[g_ndev++] = (int)v;
            p = end;
            while (*p == ',' || *p == ' ') ++p;
        }
    } else if (!list && one) {
        g_devs[g_ndev++] = atoi(one);
    } else {
        int n = ebcc_cuda_device_count();
        if (n > EBCC_MAX_DEVICES) n = EBCC_MAX_DEVICES;
        for (int d = 0; d < n; ++d) g_devs[g_ndev++] = d;
    }
    if (g_ndev == 0) g_devs[g_ndev++] = 0;           /* no device: ctx creation fails loudly below */
}

static int pipelines_per_device(void)
{
    const char *env = getenv("EBCC_PIPELINES");
    long v = env ? atol(env) : EBCC_DEFAULT_PIPELINES;
    return v < 1 ? 1 : (v > EBCC_MAX_PIPELINES ? EBCC_MAX_PIPELINES : (int)v);
}

/* pipelines over all devices of the pool */
static int n_pipelines(void)
{
    pthread_mutex_lock(&g_mu);
    resolve_devices();
    const int nd = g_ndev;
    pthread_mutex_unlock(&g_mu);
    return pipelines_per_device() * nd;
}

/* blocks until a pipeline is free; NULL when no CUDA device is usable */
static ebcc_cuda_ctx *acquire_ctx(void)
{
    pthread_mutex_lock(&g_mu);
    const pid_t me = getpid();
    if (g_ctx_pid != me) {               /* contexts inherited through fork are abandoned, not reused */
        for (int k = 0; k < EBCC_MAX_SLOTS; ++k) { g_pool[k] = NULL; g_busy[k] = 0; }
        g_ctx_pid = me;
    }
    resolve_devices();
    const int np = pipelines_per_device() * g_ndev;
    for (;;) {
        for (int k = 0; k < np; ++k) {
            if (g_busy[k]) continue;
            g_busy[k] = 1;
            if (!g_pool[k]) {
                /* created outside the lock (the slot is already ours): callers that start together build their pipelines
                 * side by side instead of queueing behind one another's cudaMalloc / pinned allocations */
                const int dev = g_devs[k % g_ndev];
                pthread_mutex_unlock(&g_mu);
                ebcc_cuda_ctx *c = ebcc_cuda_ctx_create(dev);
                pthread_mutex_lock(&g_mu);
                if (!c) {
                    g_busy[k] = 0;
                    pthread_cond_signal(&g_cv);
                    pthread_mutex_unlock(&g_mu);
                    ebcc_log(LOG_FATAL, "no usable CUDA device: this build of EBCC has no CPU code path");
                    return NULL;
                }
                g_pool[k] = c;
            }
            ebcc_cuda_ctx *ctx = g_pool[k];
            pthread_mutex_unlock(&g_mu);
            return ctx;
        }
        pthread_cond_wait(&g_cv, &g_mu);
    }
}

/* tells the device layer how crowded ITS GPU is: the largest number of pipelines seen at work there (held right now, or
 * about to be by the multi-batch calls in flight) during the last two seconds.  The memory matters: callers on several
 * threads start and finish at slightly different times, and a call that sized its kernels for an empty GPU while fifteen
 * others are about to start would hog the SMs (measured: 22.6 -> 5 GB/s end to end without it). */
static int g_crowd_max[EBCC_MAX_DEVICES];
static double g_crowd_t[EBCC_MAX_DEVICES];

static void hint_crowd(ebcc_cuda_ctx *ctx)
{
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    const double now = (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
    pthread_mutex_lock(&g_mu);
    const int nd = g_ndev > 0 ? g_ndev : 1;
    int slot = 0;
    for (int k = 0; k < EBCC_MAX_SLOTS; ++k) if (g_pool[k] == ctx) { slot = k; break; }
    const int d = slot % nd;
    int busy = 0;
    for (int k = d; k < EBCC_MAX_SLOTS; k += nd) busy += g_busy[k];
    const int expected = (g_expected + nd - 1) / nd;             /* the workers of a call spread evenly over the devices */
    if (expected > busy) busy = expected;
    if (busy >= g_crowd_max[d] || now - g_crowd_t[d] > 2.0) { g_crowd_max[d] = busy; g_crowd_t[d] = now; }
    busy = g_crowd_max[d];
    pthread_mutex_unlock(&g_mu);
    ebcc_cuda_set_crowd(ctx, busy);
}

static void expect_workers(int delta)
{
    pthread_mutex_lock(&g_mu);
    g_expected += delta;
    if (g_expected < 0) g_expected = 0;
    pthread_mutex_unlock(&g_mu);
}

static void release_ctx(ebcc_cuda_ctx *ctx)
{
    pthread_mutex_lock(&g_mu);
    for (int k = 0; k < EBCC_MAX_SLOTS; ++k) if (g_pool[k] == ctx) g_busy[k] = 0;
    pthread_cond_broadcast(&g_cv);
    pthread_mutex_unlock(&g_mu);
}

static size_t max_batch_frames(void)
{
    const char *env = getenv("EBCC_BATCH_FRAMES");
    long v = env ? atol(env) : EBCC_DEFAULT_BATCH_FRAMES;
    return v > 0 ? (size_t)v : EBCC_DEFAULT_BATCH_FRAMES;
}

/* ---- frame assembly (reference src/ebcc_codec.c:870-907) ------------------------------------------- */
typedef struct {
    int const_field;
    float minval, maxval;
    float residual_minval, residual_maxval;
    size_t coeffs_size;
    const uint8_t *zstd; size_t zstd_size;
    const uint8_t *j2k; size_t j2k_size;
    size_t tot_size;
} frame_parts_t;

static size_t assemble_frame(const frame_parts_t *p, uint8_t **out)
{
    const size_t codec_size = p->const_field ? sizeof(uint64_t) : p->j2k_size;
    const size_t out_size = sizeof(ebcc_header_t) + p->zstd_size + codec_size;
    uint8_t *buf = (uint8_t *)malloc(out_size);
    if (!buf) return 0;
    ebcc_header_t h;
    memset(&h, 0, sizeof(h));
    memcpy(h.magic, EBCC_HEADER_MAGIC, 4);
    h.version = EBCC_HEADER_VERSION;
    if (p->const_field) h.flags |= EBCC_HEADER_FLAG_CONST_FIELD;
    h.minval_bits = float_to_bits(p->minval);
    h.maxval_bits = float_to_bits(p->maxval);
    h.coeffs_size = (uint64_t)p->coeffs_size;
    h.residual_minval_bits = float_to_bits(p->residual_minval);
    h.residual_maxval_bits = float_to_bits(p->residual_maxval);
    h.compressed_size = (uint64_t)p->zstd_size;
    h.tail_size = (uint64_t)codec_size;
    uint8_t *it = buf;
    memcpy(it, &h, sizeof(h)); it += sizeof(h);
    if (p->zstd_size) { memcpy(it, p->zstd, p->zstd_size); it += p->zstd_size; }
    if (p->const_field) { uint64_t n = (uint64_t)p->tot_size; memcpy(it, &n, sizeof(n)); }
    else memcpy(it, p->j2k, p->j2k_size);
    *out = buf;
    return out_size;
}

/* ---- optional wall-clock breakdown of the device calls (EBCC_PROFILE=1), printed to stderr ---------- */
#include <time.h>
enum { TM_ANALYZE, TM_PROBE, TM_RESID_ENC, TM_RESID_PROBE, TM_PACK, TM_DOWNLOAD, TM_COUNT };
static const char *const k_tm_names[TM_COUNT] = {"upload+analyze", "rate probes", "spiht encode", "spiht probes", "zstd pack", "download"};
/* per thread: every worker of a multi-batch call (and every concurrent caller) times its own batches */
static _Thread_local double g_tm[TM_COUNT]; static _Thread_local int g_tm_calls[TM_COUNT]; static _Thread_local int g_profile = 0;
static _Thread_local unsigned long g_probe_needed, g_probe_run;   /* frame-probes the searches asked for / the lock-step rounds executed */
static double now_ms(void) { struct timespec t; clock_gettime(CLOCK_MONOTONIC, &t); return t.tv_sec * 1e3 + t.tv_nsec * 1e-6; }
#define TIMED(slot, expr) ({ double t0_ = g_profile ? now_ms() : 0; int r_ = (expr); if (g_profile) { g_tm[slot] += now_ms() - t0_; g_tm_calls[slot]++; } r_; })

/* ---- compression-ratio search: error_bound_j2k_compression (reference src/ebcc_codec.c:545-596) as a
 * resumable state machine, so that a whole batch of frames advances in lock step, one device probe
 * (ebcc_cuda_rate_probe) per round.  Pure control flow over the quantiles the device returns. */
enum { PH_LO, PH_LO_WAIT, PH_HI, PH_HI_WAIT, PH_BIS, PH_BIS_WAIT, PH_DONE };
typedef struct {
    int phase;
    float cr_lo, cr_hi, cur, pending, result;
    double q, q_prev, target;
} crsearch_t;

static void crs_init(crsearch_t *s, float start_cr, double q0, double target)
{
    s->phase = PH_LO; s->cr_lo = s->cr_hi = start_cr; s->cur = start_cr; s->q = s->q_prev = q0; s->target = target;
    s->pending = s->result = start_cr;
}

/* returns 1 when a probe at s->pending is needed (the caller stores its quantile in s->q), 0 when done */
static int crs_advance(crsearch_t *s)
{
    const double eps = 1e-8;
    for (;;) {
        switch (s->phase) {
        case PH_LO:
            if (s->q < s->target && s->cr_lo >= 1. / 2) { s->cr_lo /= 2; s->pending = s->cr_lo; s->phase = PH_LO_WAIT; return 1; }
            s->q = s->q_prev; s->phase = PH_HI; break;
        case PH_LO_WAIT: s->phase = PH_LO; break;
        case PH_HI:
            if (s->q >= s->target && s->cr_hi <= 1000) { s->cr_hi *= 2; s->pending = s->cr_hi; s->phase = PH_HI_WAIT; return 1; }
            if (s->q >= s->target) { s->result = s->cr_hi; s->phase = PH_DONE; return 0; }
            s->q = s->q_prev; s->phase = PH_BIS; break;
        case PH_HI_WAIT: s->phase = PH_HI; break;
        case PH_BIS:
            if ((fabs(s->q - s->target) > eps || s->q == 1.0) && s->cr_hi - s->cr_lo > 1.) {
                s->cur = (s->cr_lo + s->cr_hi) / 2; s->pending = s->cur; s->phase = PH_BIS_WAIT; return 1;
            }
            s->result = s->cr_lo; s->phase = PH_DONE; return 0;
        case PH_BIS_WAIT:
            if (s->q < s->target) s->cr_hi = s->cur; else s->cr_lo = s->cur;
            s->phase = PH_BIS; break;
        default: return 0;
        }
    }
}

#define MEMO_MAX 96
typedef struct { int n; float cr[MEMO_MAX]; unsigned long long n_over[MEMO_MAX]; } memo_t;

static int memo_find(const memo_t *m, float cr, unsigned long long *n_over)
{
    for (int i = 0; i < m->n; ++i) if (m->cr[i] == cr) { *n_over = m->n_over[i]; return 1; }
    return 0;
}
static void memo_put(memo_t *m, float cr, unsigned long long n_over)
{
    if (m->n < MEMO_MAX) { m->cr[m->n] = cr; m->n_over[m->n] = n_over; m->n++; }
}

/* runs the search for every frame with run[i] != 0; result_cr[i] receives the ratio the reference would settle on */
static int search_batch(ebcc_cuda_ctx *ctx, size_t nb, const int *run, const float *start_cr, const double *q0, double target,
                        const float *err_target, size_t tot_size, memo_t *memo, float *result_cr, float *last_cr, double cap_quantile)
{
    crsearch_t *st = (crsearch_t *)calloc(nb, sizeof(*st));
    int *need = (int *)calloc(nb, sizeof(int));
    /* Every decision of crs_advance is a comparison of q = 1 - n_over / N with a quantile target 1 - e (q < target,
     * |q - target| > 1e-8, q == 1), here and in the later search that reuses the memo with a smaller e.  All of them have
     * the same outcome for every count >= (e + 1e-8) N + 2, so the device may stop counting there (and stop looking at
     * the frame).  Trace logging prints the exact quantile of every probe like the reference: no cap then. */
    const unsigned long long cap = (g_log_level <= LOG_TRACE || !(cap_quantile >= 0.0) || !(cap_quantile <= 1.0)) ? 0ull
                                   : (unsigned long long)floor((cap_quantile + 1e-8) * (double)tot_size) + 2ull;
    ebcc_probe_result *pr = (ebcc_probe_result *)calloc(nb, sizeof(*pr));
    int rc = 0;
    for (size_t i = 0; i < nb; ++i) { if (run[i]) crs_init(&st[i], start_cr[i], q0[i], target); else st[i].phase = PH_DONE; result_cr[i] = start_cr[i]; }
    for (;;) {
        int any_dev = 0, any = 0;
        for (size_t i = 0; i < nb; ++i) {
            need[i] = 0;
            while (st[i].phase != PH_DONE) {
                if (!crs_advance(&st[i])) { result_cr[i] = st[i].result; break; }
                any = 1;
                unsigned long long n;
                if (memo_find(&memo[i], st[i].pending, &n)) { st[i].q = 1. - ((double)n / tot_size); continue; }
                need[i] = 1; any_dev = 1; break;
            }
        }
        if (!any_dev) { if (!any) break; else continue; }
        for (size_t i = 0; i < nb; ++i) if (need[i]) { last_cr[i] = st[i].pending; g_probe_needed++; }
        g_probe_run += nb;
        if (TIMED(TM_PROBE, ebcc_cuda_rate_probe_capped(ctx, nb, last_cr, err_target, cap, pr))) { rc = -1; break; }
        for (size_t i = 0; i < nb; ++i) {
            if (!need[i]) continue;
            memo_put(&memo[i], st[i].pending, pr[i].n_over);
            st[i].q = 1. - ((double)pr[i].n_over / tot_size);
            ebcc_log(LOG_TRACE, "frame %lu: cr %f, 1-error_target_quantile: %.1e, jp2_length: %lu", i, st[i].pending, 1 - st[i].q, pr[i].stream_bytes);
        }
    }
    free(st); free(need); free(pr);
    return rc;
}

/* ---- batched encode: one pass of the reference's ebcc_encode (src/ebcc_codec.c:607-918) over nb frames -------- */
static int encode_chunk(ebcc_cuda_ctx *ctx, const float *frames, size_t nb, size_t base, const codec_config_t *config,
                        uint8_t **outs, size_t *sizes, size_t *done)
{
    /* dims[0] tiles of dims[1] x dims[2] per image (reference j2k_encode_internal, src/ebcc_codec.c:105-123) */
    const size_t T = config->dims[0], H = config->dims[1], W = config->dims[2], N = T * H * W;
    const residual_t mode = config->residual_compression_type;
    int rc = -1;
    double base_error_quantile = 1e-6;
    const char *env_q = getenv("EBCC_INIT_BASE_ERROR_QUANTILE");
    const int pure_disabled = getenv("EBCC_DISABLE_PURE_BASE_COMPRESSION_FALLBACK") != NULL;
    const int consistency_disabled = getenv("EBCC_DISABLE_PURE_BASE_COMPRESSION_FALLBACK_CONSISTENCY") != NULL;
    const int mean_adjust_disabled = getenv("EBCC_DISABLE_MEAN_ADJUSTMENT") != NULL;
    if (env_q) base_error_quantile = strtod(env_q, NULL);
    const double base_quantile_target = 1 - base_error_quantile;

    const uint8_t **streams = (const uint8_t **)calloc(nb, sizeof(*streams));
    size_t *ssz = (size_t *)calloc(nb, sizeof(size_t));
    float *mn = (float *)calloc(nb, sizeof(float)), *mx = (float *)calloc(nb, sizeof(float));
    int *fl = (int *)calloc(nb, sizeof(int));
    float *cr = (float *)calloc(nb, sizeof(float)), *cr1 = (float *)calloc(nb, sizeof(float)), *cr2 = (float *)calloc(nb, sizeof(float));
    float *tgt = (float *)calloc(nb, sizeof(float)), *start = (float *)calloc(nb, sizeof(float));
    double *q0 = (double *)calloc(nb, sizeof(double)), *mean = (double *)calloc(nb, sizeof(double));
    int *run = (int *)calloc(nb, sizeof(int)), *resid = (int *)calloc(nb, sizeof(int)), *pure_required = (int *)calloc(nb, sizeof(int));
    memo_t *memo = (memo_t *)calloc(nb, sizeof(memo_t));
    ebcc_probe_result *p0 = (ebcc_probe_result *)calloc(nb, sizeof(*p0)), *p1 = (ebcc_probe_result *)calloc(nb, sizeof(*p1)),
                      *p2 = (ebcc_probe_result *)calloc(nb, sizeof(*p2));
    uint8_t **j2k1 = (uint8_t **)calloc(nb, sizeof(*j2k1));
    size_t *len1 = (size_t *)calloc(nb, sizeof(size_t)), *coeffs_size = (size_t *)calloc(nb, sizeof(size_t)), *tb = (size_t *)calloc(nb, sizeof(size_t));
    const uint8_t **zs = (const uint8_t **)calloc(nb, sizeof(*zs));
    uint8_t **zcopy = (uint8_t **)calloc(nb, sizeof(*zcopy));
    size_t *zn = (size_t *)calloc(nb, sizeof(size_t));
    float *perr = (float *)calloc(nb, sizeof(float)), *best_err = (float *)calloc(nb, sizeof(float));
    double *psum = (double *)calloc(nb, sizeof(double)), *thi = (double *)calloc(nb, sizeof(double)), *tlo = (double *)calloc(nb, sizeof(double)),
           *best_trunc = (double *)calloc(nb, sizeof(double));
    int *use_pure = (int *)calloc(nb, sizeof(int));

    g_profile = getenv("EBCC_PROFILE") != NULL;
    const double t_call0 = g_profile ? now_ms() : 0;
    if (g_profile) { memset(g_tm, 0, sizeof(g_tm)); memset(g_tm_calls, 0, sizeof(g_tm_calls)); g_probe_needed = g_probe_run = 0; }
    hint_crowd(ctx);
    if (ebcc_cuda_reserve_tiled(ctx, T, H, W, nb)) { rc = -2; goto out; }     /* -2: the caller retries with fewer frames per batch */
    if (TIMED(TM_ANALYZE, ebcc_cuda_upload(ctx, frames, nb) || ebcc_cuda_analyze(ctx, nb, mode != NONE) || ebcc_cuda_sync(ctx))) goto out;
    for (size_t i = 0; i < nb; ++i) { cr[i] = config->base_cr; tgt[i] = 0; }

    if (mode == NONE) {
        if (ebcc_cuda_rate_probe_streams_only(ctx, nb, cr)) goto out;
    } else {
        /* first encode + decode at base_cr (reference :693-712); targets need min/max, so fetch the frame statistics first */
        if (ebcc_cuda_frame_stats(ctx, nb, mn, mx, fl)) goto out;
        for (size_t i = 0; i < nb; ++i) {
            float t = config->error;
            if (mode == RELATIVE_ERROR) t *= (mx[i] - mn[i]);       /* get_data_range, reference :724-726 */
            tgt[i] = t;
        }
        if (TIMED(TM_PROBE, ebcc_cuda_rate_probe(ctx, nb, cr, tgt, 0, p0))) goto out;
        for (size_t i = 0; i < nb; ++i) {
            run[i] = !(fl[i] & (EBCC_FRAME_CONST | EBCC_FRAME_NANINF));
            memo_put(&memo[i], cr[i], p0[i].n_over);
            q0[i] = 1. - ((double)p0[i].n_over / N);
            start[i] = config->base_cr;
            mean[i] = p0[i].sum_err / N;                            /* get_mean_error after the first decode, :709 */
        }
        if (search_batch(ctx, nb, run, start, q0, base_quantile_target, tgt, N, memo, cr1, cr, base_error_quantile)) goto out;
        if (TIMED(TM_PROBE, ebcc_cuda_rate_probe(ctx, nb, cr1, tgt, 1, p1))) goto out;
        for (size_t i = 0; i < nb; ++i) {
            const double q = 1. - ((double)p1[i].n_over / N);
            if (run[i] && q < base_quantile_target)      /* reference :591-593 */
                ebcc_log(LOG_WARN, "Could not reach error target quantile of (1-%.2e) (1-%.2e instead).", 1 - base_quantile_target, 1 - q);
        }
    }
    if (TIMED(TM_DOWNLOAD, ebcc_cuda_download(ctx, nb, streams, ssz, mn, mx, fl))) goto out;
    for (size_t i = 0; i < nb; ++i) {
        if (fl[i] & EBCC_FRAME_NANINF) {
            /* reference check_nan_inf: log_fatal + exit(1) (src/ebcc_codec.c:598-605) */
            ebcc_log(LOG_FATAL, "NaN or Inf found in data (frame %lu)", base + i);
            exit(1);
        }
    }

    if (mode != NONE) {
        const int pure_done = base_quantile_target == 1.0;
        int any_resid = 0;
        for (size_t i = 0; i < nb; ++i) {
            if (!run[i]) continue;
            len1[i] = ssz[i];
            j2k1[i] = (uint8_t *)streams[i];      /* stays valid across the one later download of this batch (pure base-layer fallback) */
            const float cur_max_error = fmaxf(fabsf(p1[i].rmin), fabsf(p1[i].rmax));
            resid[i] = !(cur_max_error <= tgt[i]);
            best_err[i] = -1;
            tb[i] = len1[i] * 8;
            if (resid[i]) any_resid = 1;
            ebcc_log(LOG_TRACE, "frame %lu: base cr %f, jp2_length %lu, max error %f, target %f", base + i, cr1[i], len1[i], cur_max_error, tgt[i]);
        }
        if (any_resid) {
            /* SPIHT encode of the normalised residual + full decode check (reference :744-763) */
            if (TIMED(TM_RESID_ENC, ebcc_cuda_residual_encode(ctx, nb, resid, tb, coeffs_size))) goto out;
            for (size_t i = 0; i < nb; ++i) tb[i] = coeffs_size[i] * 8;
            if (TIMED(TM_RESID_PROBE, ebcc_cuda_residual_probe(ctx, nb, resid, tb, perr, psum))) goto out;
            for (size_t i = 0; i < nb; ++i) {
                if (!resid[i]) continue;
                if (perr[i] > tgt[i]) {
                    ebcc_log(LOG_INFO, "Could not reach error target of %f (%f instead). Retry with pure base compression.", tgt[i], perr[i]);
                    /* like the reference (:752-758) the full-length residual stays in place: it is packed and written
                     * unless the pure base-layer fallback below replaces it (which it does whenever it runs) */
                    resid[i] = 0; pure_required[i] = 1;
                } else {
                    best_err[i] = perr[i]; mean[i] = psum[i] / N;
                    thi[i] = (double)coeffs_size[i] * 8; tlo[i] = 112.0; best_trunc[i] = thi[i];
                }
            }
            /* truncation bisection (reference :766-796), all frames in lock step */
            for (;;) {
                int any = 0;
                for (size_t i = 0; i < nb; ++i) {
                    tb[i] = 0;
                    if (!resid[i]) continue;
                    if (((tgt[i] - best_err[i]) / tgt[i] > 1e-8) && (thi[i] - tlo[i] > 8 * 4)) {
                        tb[i] = ((size_t)ceill((thi[i] + tlo[i]) / 2 / 8)) * 8;
                        any = 1;
                    }
                }
                if (!any) break;
                int *act = (int *)calloc(nb, sizeof(int));
                for (size_t i = 0; i < nb; ++i) act[i] = tb[i] != 0;
                int prc = TIMED(TM_RESID_PROBE, ebcc_cuda_residual_probe(ctx, nb, act, tb, perr, psum));
                free(act);
                if (prc) goto out;
                for (size_t i = 0; i < nb; ++i) {
                    if (!tb[i]) continue;
                    if (perr[i] > tgt[i]) tlo[i] = (double)tb[i];
                    else {
                        thi[i] = (double)tb[i];
                        if (perr[i] >= best_err[i]) { best_err[i] = perr[i]; best_trunc[i] = (double)tb[i]; mean[i] = psum[i] / N; }
                    }
                    ebcc_log(LOG_TRACE, "frame %lu: trunc_lo %.1f trunc_hi %.1f bytes %lu max error %f", base + i, tlo[i], thi[i], tb[i] / 8, perr[i]);
                }
            }
            for (size_t i = 0; i < nb; ++i) {
                if (resid[i]) coeffs_size[i] = (size_t)(best_trunc[i] / 8.);
                if (!resid[i] && !pure_required[i]) coeffs_size[i] = 0;
                if (coeffs_size[i] <= 16) coeffs_size[i] = 0;
            }
            /* zstd stage on the device (reference :813-817) */
            if (TIMED(TM_PACK, ebcc_cuda_residual_pack(ctx, nb, coeffs_size, zs, zn))) goto out;
            for (size_t i = 0; i < nb; ++i) {
                if (!zn[i]) continue;
                zcopy[i] = (uint8_t *)malloc(zn[i]);
                memcpy(zcopy[i], zs[i], zn[i]);
            }
        }
        /* pure base-layer fallback (reference :819-854) */
        if (!pure_done && !pure_disabled) {
            for (size_t i = 0; i < nb; ++i) {
                start[i] = consistency_disabled ? cr1[i] : config->base_cr;
                unsigned long long n = 0;
                memo_find(&memo[i], start[i], &n);
                q0[i] = 1. - ((double)n / N);
            }
            if (search_batch(ctx, nb, run, start, q0, 1.0, tgt, N, memo, cr2, cr, base_error_quantile)) goto out;
            if (TIMED(TM_PROBE, ebcc_cuda_rate_probe(ctx, nb, cr2, tgt, 1, p2))) goto out;
            if (TIMED(TM_DOWNLOAD, ebcc_cuda_download(ctx, nb, streams, ssz, NULL, NULL, NULL))) goto out;
            for (size_t i = 0; i < nb; ++i) {
                if (!run[i]) continue;
                if (ssz[i] < zn[i] + len1[i] || pure_required[i]) {
                    if (ssz[i] < zn[i] + len1[i])
                        ebcc_log(LOG_INFO, "Pure base compression (%lu) is better than base (%lu) + residual (%lu) compression (sum: %lu)",
                                 ssz[i], len1[i], zn[i], zn[i] + len1[i]);
                    use_pure[i] = 1; mean[i] = p2[i].sum_err / N; zn[i] = 0; coeffs_size[i] = 0;
                }
            }
        }
    }

    for (size_t i = 0; i < nb; ++i) {
        frame_parts_t p;
        memset(&p, 0, sizeof(p));
        p.const_field = (fl[i] & EBCC_FRAME_CONST) != 0;
        p.minval = mn[i]; p.maxval = mx[i];
        p.tot_size = N;
        if (mode != NONE && run[i]) {
            p.residual_minval = p1[i].rmin; p.residual_maxval = p1[i].rmax;
            p.coeffs_size = coeffs_size[i];
            p.zstd = zcopy[i]; p.zstd_size = zn[i];
            if (use_pure[i]) { p.j2k = streams[i]; p.j2k_size = ssz[i]; }
            else { p.j2k = j2k1[i]; p.j2k_size = len1[i]; }
            ebcc_log(LOG_INFO, "Mean of compression error: %e", mean[i]);
            if (!mean_adjust_disabled && fabs(mean[i]) > 1e-18) { p.minval += mean[i]; p.maxval += mean[i]; }
        } else {
            p.j2k = streams[i]; p.j2k_size = ssz[i];
        }
        if (fl[i] & EBCC_FRAME_OVERFLOW) { ebcc_log(LOG_FATAL, "tier-1 work buffer overflow in frame %lu: no stream written", base + i); continue; }
        if (!p.const_field && p.j2k_size == 0) { ebcc_log(LOG_FATAL, "empty codestream for frame %lu", base + i); continue; }
        sizes[base + i] = assemble_frame(&p, &outs[base + i]);
        if (sizes[base + i]) {
            (*done)++;
            ebcc_log(LOG_INFO, "coeffs_size: %lu, compressed_size: %lu, jp2_length: %lu, compression ratio: %f", p.coeffs_size, p.zstd_size,
                     p.const_field ? sizeof(uint64_t) : p.j2k_size, (double)(N * sizeof(float)) / sizes[base + i]);
        }
    }
    rc = 0;
    if (g_profile) {
        fprintf(stderr, "[ebcc profile] %lu frames:", nb);
        for (int k = 0; k < TM_COUNT; ++k) fprintf(stderr, " %s %.1f ms/%d;", k_tm_names[k], g_tm[k], g_tm_calls[k]);
        double dev = 0; for (int k = 0; k < TM_COUNT; ++k) dev += g_tm[k];
        fprintf(stderr, " host side (copies, assembly) %.1f ms of %.1f;", now_ms() - t_call0 - dev, now_ms() - t_call0);
        fprintf(stderr, " search rounds: %lu frame-probes needed of %lu run;", g_probe_needed, g_probe_run);
        fputc('\n', stderr);
    }
out:
    for (size_t i = 0; i < nb; ++i) free(zcopy[i]);
    free(streams); free(ssz); free(mn); free(mx); free(fl); free(cr); free(cr1); free(cr2); free(tgt); free(start); free(q0); free(mean);
    free(run); free(resid); free(pure_required); free(memo); free(p0); free(p1); free(p2); free(j2k1); free(len1); free(coeffs_size); free(tb);
    free(zs); free(zcopy); free(zn); free(perr); free(best_err); free(psum); free(thi); free(tlo); free(best_trunc); free(use_pure);
    return rc;
}

/* batches of one call are handed to up to EBCC_PIPELINES workers, each on its own device pipeline */
typedef struct {
    const float *frames; size_t nframes; const codec_config_t *config; uint8_t **outs; size_t *sizes;
    size_t N, maxb, next, done;
    int failed;
    pthread_mutex_t mu;
} enc_job_t;

/* one batch; when the device (or pinned host) memory for nb frames cannot be had, the batch is split in halves */
static int encode_range(ebcc_cuda_ctx *ctx, enc_job_t *j, size_t base, size_t nb, size_t *done)
{
    const int rc = encode_chunk(ctx, j->frames + base * j->N, nb, base, j->config, j->outs, j->sizes, done);
    if (rc != -2 || nb < 2) return rc;
    ebcc_log(LOG_WARN, "work buffers for %lu frames could not be allocated: retrying with %lu", nb, nb / 2);
    pthread_mutex_lock(&j->mu);
    if (j->maxb > nb / 2) j->maxb = nb / 2;          /* later batches of this call start smaller right away */
    pthread_mutex_unlock(&j->mu);
    const int r1 = encode_range(ctx, j, base, nb / 2, done);
    if (r1) return r1;
    return encode_range(ctx, j, base + nb / 2, nb - nb / 2, done);
}

static void *enc_worker(void *arg)
{
    enc_job_t *j = (enc_job_t *)arg;
    ebcc_cuda_ctx *ctx = acquire_ctx();
    if (!ctx) { pthread_mutex_lock(&j->mu); j->failed = 1; pthread_mutex_unlock(&j->mu); return NULL; }
    for (;;) {
        pthread_mutex_lock(&j->mu);
        const size_t base = j->next;
        const int stop = j->failed || base >= j->nframes;
        const size_t maxb = j->maxb;
        if (!stop) j->next = base + maxb;
        pthread_mutex_unlock(&j->mu);
        if (stop) break;
        const size_t nb = j->nframes - base < maxb ? j->nframes - base : maxb;
        size_t done = 0;
        const int rc = encode_range(ctx, j, base, nb, &done);
        pthread_mutex_lock(&j->mu);
        j->done += done;
        if (rc) j->failed = 1;
        pthread_mutex_unlock(&j->mu);
        if (rc) { ebcc_log(LOG_FATAL, "CUDA encode of frames %lu..%lu failed", base, base + nb - 1); break; }
    }
    release_ctx(ctx);
    return NULL;
}

size_t ebcc_encode_batch(const float *frames, size_t nframes, const codec_config_t *config, uint8_t **outs, size_t *sizes)
{
    log_set_level_from_env();
    for (size_t i = 0; i < nframes; ++i) { outs[i] = NULL; sizes[i] = 0; }
    if (!dims_are_valid(config->dims)) {
        ebcc_log(LOG_FATAL, "Invalid EBCC dimensions: product(dims[0..%d]) and dims[%d] must be between %d and %d",
                 NDIMS - 2, NDIMS - 1, EBCC_MIN_INTERNAL_IMAGE_DIM, EBCC_MAX_INTERNAL_IMAGE_DIM);
        return 0;
    }
    if ((unsigned)config->residual_compression_type > RELATIVE_ERROR) {
        ebcc_log(LOG_FATAL, "unknown residual compression type %d", (int)config->residual_compression_type);
        return 0;
    }
    enc_job_t job;
    job.frames = frames; job.nframes = nframes; job.config = config; job.outs = outs; job.sizes = sizes;
    job.N = config->dims[0] * config->dims[1] * config->dims[2]; job.maxb = max_batch_frames(); job.next = 0; job.done = 0; job.failed = 0;
    pthread_mutex_init(&job.mu, NULL);
    const size_t nbatches = (nframes + job.maxb - 1) / job.maxb;
    int nworkers = n_pipelines();
    if ((size_t)nworkers > nbatches) nworkers = (int)nbatches;
    pthread_t th[EBCC_MAX_SLOTS];
    int started = 0;
    expect_workers(nworkers);
    for (int k = 1; k < nworkers; ++k) if (pthread_create(&th[started], NULL, enc_worker, &job) == 0) started++;
    enc_worker(&job);                      /* the calling thread is worker 0 */
    for (int k = 0; k < started; ++k) pthread_join(th[k], NULL);
    expect_workers(-nworkers);
    pthread_mutex_destroy(&job.mu);
    return job.done;
}

size_t ebcc_encode(float *data, codec_config_t *config, uint8_t **out_buffer)
{
    size_t size = 0;
    *out_buffer = NULL;
    log_set_level_from_env();
    print_config(config);
    if (ebcc_encode_batch(data, 1, config, out_buffer, &size) != 1) return 0;
    return size;
}

/* Decoded frames the library allocates itself (the caller frees them with free_buffer = free, like the reference's).  A fresh
 * 4 MB malloc is faulted in page by page inside the device-to-host copy; a 2 MB-aligned block that asks for transparent huge
 * pages takes two faults instead of a thousand where the kernel grants them (it is a hint: plain pages otherwise). */
#include <sys/mman.h>
static void *frame_alloc(size_t bytes)
{
    const size_t huge = (size_t)2 << 20;
    if (bytes >= huge) {
        void *p = NULL;
        if (posix_memalign(&p, huge, bytes) == 0 && p) {
#ifdef MADV_HUGEPAGE
            (void)madvise(p, bytes & ~(huge - 1), MADV_HUGEPAGE);
#endif
            return p;
        }
    }
    return malloc(bytes);
}

/* ---- batched decode (reference ebcc_decode, src/ebcc_codec.c:1215-1320) --------------------------------- */
typedef struct {
    int ok, const_field;
    float minval, maxval, rmin, rmax;
    size_t coeffs_size, zstd_size, tail_size, tot_size;
    const uint8_t *zstd, *tail;
    size_t H, W, T;              /* tile height, width, tiles per image */
} parsed_stream_t;

static uint32_t be32(const uint8_t *p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }

static int legacy_take(const uint8_t **it, const uint8_t *end, void *dst, size_t nb)
{
    if ((size_t)(end - *it) < nb) return 0;
    memcpy(dst, *it, nb);
    *it += nb;
    return 1;
}

/* header-less streams of older EBCC releases (reference ebcc_decode_legacy, src/ebcc_codec.c:1147-1213):
 * f32 min, f32 max, u64 coeffs_size, f32 rmin, f32 rmax, u64 zstd_size, zstd payload, then u64 count (constant
 * field, min == max) or the JPEG 2000 codestream */
static int parse_legacy_prefix(const uint8_t *data, size_t n, parsed_stream_t *ps)
{
    const uint8_t *it = data, *end = data + n;
    uint64_t cs = 0, zs = 0;
    if (!legacy_take(&it, end, &ps->minval, 4) || !legacy_take(&it, end, &ps->maxval, 4) || !legacy_take(&it, end, &cs, 8) ||
        !legacy_take(&it, end, &ps->rmin, 4) || !legacy_take(&it, end, &ps->rmax, 4) || !legacy_take(&it, end, &zs, 8)) {
        ebcc_log(LOG_FATAL, "Invalid legacy encoded data: truncated header");
        return 0;
    }
    if ((uint64_t)(end - it) < zs) { ebcc_log(LOG_FATAL, "Invalid legacy encoded data: truncated residual payload"); return 0; }
    ps->coeffs_size = (size_t)cs; ps->zstd_size = (size_t)zs;
    ps->zstd = it; it += zs;
    ps->tail = it; ps->tail_size = (size_t)(end - it);
    ps->const_field = ps->minval == ps->maxval;
    if (ps->const_field && ps->tail_size < sizeof(uint64_t)) { ebcc_log(LOG_FATAL, "Invalid legacy encoded data: missing const-field length"); return 0; }
    if (ps->const_field) ps->tail_size = sizeof(uint64_t);
    return 1;
}

static int parse_stream(const uint8_t *data, size_t n, parsed_stream_t *ps)
{
    memset(ps, 0, sizeof(*ps));
    if (n < sizeof(ebcc_header_t) || memcmp(data, EBCC_HEADER_MAGIC, 4) != 0) {
        if (!parse_legacy_prefix(data, n, ps)) return 0;
    } else {
        ebcc_header_t h;
        memcpy(&h, data, sizeof(h));
        if (h.version != EBCC_HEADER_VERSION) { ebcc_log(LOG_FATAL, "Unsupported EBCC header version: %u", h.version); return 0; }
        size_t used = sizeof(h);
        if (h.compressed_size > n - used) { ebcc_log(LOG_FATAL, "Invalid encoded data: truncated payload"); return 0; }
        used += (size_t)h.compressed_size;
        if (h.tail_size > n - used) { ebcc_log(LOG_FATAL, "Invalid encoded data: truncated payload"); return 0; }
        used += (size_t)h.tail_size;
        /* bytes beyond header + payload are tolerated like the reference does (its iterator check, :1314, only
         * compares what it consumed with what the header announced): padded buffers, over-reported nbytes */
        ps->minval = bits_to_float(h.minval_bits); ps->maxval = bits_to_float(h.maxval_bits);
        ps->rmin = bits_to_float(h.residual_minval_bits); ps->rmax = bits_to_float(h.residual_maxval_bits);
        ps->coeffs_size = (size_t)h.coeffs_size; ps->zstd_size = (size_t)h.compressed_size; ps->tail_size = (size_t)h.tail_size;
        ps->zstd = data + sizeof(h); ps->tail = ps->zstd + ps->zstd_size;
        ps->const_field = (h.flags & EBCC_HEADER_FLAG_CONST_FIELD) != 0;
    }
    if (ps->const_field) {
        if (ps->tail_size != sizeof(uint64_t)) { ebcc_log(LOG_FATAL, "Invalid encoded data: const-field payload must contain uint64_t length"); return 0; }
        uint64_t t; memcpy(&t, ps->tail, sizeof(t));
        /* the largest frame the codec can write is 2047 x 2047 samples: anything else is a crafted or damaged stream
         * (the count feeds a malloc and a fill loop; 2^62+1 would wrap the byte count) */
        if (t == 0 || t > (uint64_t)EBCC_MAX_INTERNAL_IMAGE_DIM * EBCC_MAX_INTERNAL_IMAGE_DIM) {
            ebcc_log(LOG_FATAL, "Invalid encoded data: const-field size %llu out of range", (unsigned long long)t); return 0;
        }
        ps->tot_size = (size_t)t;
        if (ps->zstd_size > 0 && ps->coeffs_size > 0) { ebcc_log(LOG_FATAL, "Invalid encoded data: residual data cannot be applied to const field"); return 0; }
    } else {
        /* image size from the SIZ marker: SOC(2) SIZ(2) Lsiz(2) Rsiz(2) Xsiz Ysiz XOsiz YOsiz XTsiz YTsiz */
        if (ps->tail_size < 40 || ps->tail[0] != 0xff || ps->tail[1] != 0x4f || ps->tail[2] != 0xff || ps->tail[3] != 0x51) {
            ebcc_log(LOG_FATAL, "Invalid encoded data: no JPEG 2000 codestream"); return 0;
        }
        const size_t X = be32(ps->tail + 8), Y = be32(ps->tail + 12), XT = be32(ps->tail + 24), YT = be32(ps->tail + 28);
        /* one column of tiles, all of the same height (what j2k_encode_internal writes for dims[0] > 1) */
        if (be32(ps->tail + 16) || be32(ps->tail + 20) || be32(ps->tail + 32) || be32(ps->tail + 36) || X != XT || YT == 0 || Y % YT != 0) {
            ebcc_log(LOG_FATAL, "Invalid encoded data: unsupported JPEG 2000 tiling"); return 0;
        }
        ps->W = X; ps->H = YT; ps->T = Y / YT; ps->tot_size = X * Y;
        if (X < EBCC_MIN_INTERNAL_IMAGE_DIM || X > EBCC_MAX_INTERNAL_IMAGE_DIM || YT < EBCC_MIN_INTERNAL_IMAGE_DIM || Y > EBCC_MAX_INTERNAL_IMAGE_DIM) {
            ebcc_log(LOG_FATAL, "Invalid encoded data: image size %lu x %lu out of range", Y, X); return 0;
        }
    }
    ps->ok = 1;
    return 1;
}

typedef struct {
    parsed_stream_t *ps; float **outs; size_t *counts; int *owned;
    size_t nframes, maxb, ngroups, next, done;
    size_t *starts, *ends;
    pthread_mutex_t mu;
} dec_job_t;

static void *dec_worker(void *arg)
{
    dec_job_t *jb = (dec_job_t *)arg;
    const size_t maxb = jb->maxb;
    parsed_stream_t *ps = jb->ps;
    const uint8_t **js = (const uint8_t **)malloc(sizeof(*js) * maxb);
    size_t *jn = (size_t *)malloc(sizeof(size_t) * maxb), *idx = (size_t *)malloc(sizeof(size_t) * maxb);
    float *mn = (float *)malloc(sizeof(float) * maxb), *mx = (float *)malloc(sizeof(float) * maxb);
    float **dst = (float **)malloc(sizeof(float *) * maxb);
    ebcc_residual_in *rs = (ebcc_residual_in *)calloc(maxb, sizeof(*rs));
    ebcc_cuda_ctx *ctx = NULL;
    for (;;) {
        pthread_mutex_lock(&jb->mu);
        const size_t g = jb->next < jb->ngroups ? jb->next++ : (size_t)-1;
        pthread_mutex_unlock(&jb->mu);
        if (g == (size_t)-1) break;
        const size_t i = jb->starts[g], H = ps[i].H, W = ps[i].W, T = ps[i].T;
        size_t nb = 0;
        for (size_t j = i; j < jb->ends[g]; ++j) {
            if (!ps[j].ok || ps[j].const_field) continue;
            if (!jb->outs[j]) { jb->outs[j] = (float *)frame_alloc(T * H * W * sizeof(float)); jb->owned[j] = 1; }
            if (!jb->outs[j]) { ps[j].ok = 0; continue; }
            js[nb] = ps[j].tail; jn[nb] = ps[j].tail_size; mn[nb] = ps[j].minval; mx[nb] = ps[j].maxval; dst[nb] = jb->outs[j]; idx[nb] = j;
            memset(&rs[nb], 0, sizeof(rs[nb]));
            if (ps[j].zstd_size > 0 && ps[j].coeffs_size > 0) {
                rs[nb].zstd = ps[j].zstd; rs[nb].zstd_size = ps[j].zstd_size; rs[nb].coeffs_size = ps[j].coeffs_size;
                rs[nb].rmin = ps[j].rmin; rs[nb].rmax = ps[j].rmax;
            }
            nb++;
        }
        if (!nb) continue;
        if (!ctx) ctx = acquire_ctx();
        if (ctx) hint_crowd(ctx);
        const int rc = ctx ? ebcc_cuda_decode_batch_tiled(ctx, js, jn, nb, T, H, W, mn, mx, rs, dst) : -1;
        size_t ok = 0;
        for (size_t k = 0; k < nb; ++k) {
            if (rc == 0) { jb->counts[idx[k]] = T * H * W; ok++; }
            else if (jb->owned[idx[k]]) { free(jb->outs[idx[k]]); jb->outs[idx[k]] = NULL; }
        }
        if (rc) ebcc_log(LOG_FATAL, "CUDA decode of %lu frame(s) failed", nb);
        pthread_mutex_lock(&jb->mu);
        jb->done += ok;
        pthread_mutex_unlock(&jb->mu);
    }
    if (ctx) release_ctx(ctx);
    free(js); free(jn); free(idx); free(mn); free(mx); free(dst); free(rs);
    return NULL;
}

/* const_inplace: constant fields are written into a caller-provided outs[i] (batch API, host or device memory); 0 =
 * the reference's ebcc_decode, whose const-field branch always returns a fresh buffer (src/ebcc_codec.c:1278) while its
 * JPEG 2000 branch writes into *out_buffer when that is non-NULL on entry (j2k_decode_internal, :1126-1128).
 * expect_count != 0: every stream must announce exactly that many samples (chunk containers). */
static size_t decode_batch_impl(uint8_t *const *streams, const size_t *stream_sizes, size_t nframes, float **outs, size_t *counts,
                                int const_inplace, size_t expect_count)
{
    log_set_level_from_env();
    size_t done = 0;
    parsed_stream_t *ps = (parsed_stream_t *)calloc(nframes ? nframes : 1, sizeof(*ps));
    int *owned = (int *)calloc(nframes ? nframes : 1, sizeof(int));
    for (size_t i = 0; i < nframes; ++i) {
        counts[i] = 0;
        if (!parse_stream(streams[i], stream_sizes[i], &ps[i])) continue;
        if (expect_count && ps[i].tot_size != expect_count) {
            /* chunk containers know the size of every chunk: refuse before anything is allocated or decoded */
            ebcc_log(LOG_FATAL, "chunk %lu holds %lu values, expected %lu", i, ps[i].tot_size, expect_count);
            ps[i].ok = 0; continue;
        }
        if (ps[i].const_field) {
            /* constant field: the stream carries one value and a count (reference :1265-1281) */
            float *o = (float *)malloc(ps[i].tot_size * sizeof(float));
            if (!o) { ps[i].ok = 0; continue; }
            for (size_t k = 0; k < ps[i].tot_size; ++k) o[k] = ps[i].minval;
            if (outs[i] && const_inplace) {
                /* caller-provided destination (host or device memory): written in place like every other frame */
                const int rc = ebcc_cuda_copy_out(outs[i], o, ps[i].tot_size * sizeof(float));
                free(o);
                if (rc) { ps[i].ok = 0; continue; }
            } else outs[i] = o;
            counts[i] = ps[i].tot_size; done++;
        }
    }
    /* runs of frames with the same geometry, at most EBCC_BATCH_FRAMES each, become the jobs of the workers */
    dec_job_t job;
    memset(&job, 0, sizeof(job));
    job.ps = ps; job.outs = outs; job.counts = counts; job.owned = owned; job.nframes = nframes; job.maxb = max_batch_frames();
    job.starts = (size_t *)malloc(sizeof(size_t) * (nframes ? nframes : 1));
    job.ends = (size_t *)malloc(sizeof(size_t) * (nframes ? nframes : 1));
    for (size_t i = 0; i < nframes;) {
        if (!ps[i].ok || ps[i].const_field) { ++i; continue; }
        size_t nb = 0, j = i;
        for (; j < nframes && nb < job.maxb; ++j) {
            if (!ps[j].ok || ps[j].const_field) continue;
            if (ps[j].H != ps[i].H || ps[j].W != ps[i].W || ps[j].T != ps[i].T) break;
            nb++;
        }
        job.starts[job.ngroups] = i; job.ends[job.ngroups] = j; job.ngroups++;
        i = j;
    }
    pthread_mutex_init(&job.mu, NULL);
    int nworkers = n_pipelines();
    if ((size_t)nworkers > job.ngroups) nworkers = (int)job.ngroups;
    pthread_t th[EBCC_MAX_SLOTS];
    int started = 0;
    expect_workers(nworkers);
    for (int k = 1; k < nworkers; ++k) if (pthread_create(&th[started], NULL, dec_worker, &job) == 0) started++;
    if (job.ngroups) dec_worker(&job);
    for (int k = 0; k < started; ++k) pthread_join(th[k], NULL);
    expect_workers(-nworkers);
    pthread_mutex_destroy(&job.mu);
    done += job.done;
    free(job.starts); free(job.ends); free(ps); free(owned);
    return done;
}

size_t ebcc_decode_batch(uint8_t *const *streams, const size_t *stream_sizes, size_t nframes, float **outs, size_t *counts)
{
    return decode_batch_impl(streams, stream_sizes, nframes, outs, counts, 1, 0);
}

size_t ebcc_decode(uint8_t *data, size_t data_size, float **out_buffer)
{
    /* Reference contract (src/ebcc_codec.c:1215-1320, :1126-1128): when *out_buffer is non-NULL on entry the frame is
     * decoded INTO that buffer (the zarr codec's decode(buf, out=arr) relies on it, ebcc/zarr_filter.py:63-72) and the
     * pointer is left unchanged; when it is NULL a fresh malloc'ed buffer is returned.  Constant-field streams always
     * come back in a fresh buffer (:1278). */
    size_t count = 0;
    uint8_t *s[1] = {data};
    float *o = *out_buffer;
    if (decode_batch_impl(s, &data_size, 1, &o, &count, 0, 0) != 1) return 0;
    *out_buffer = o;
    return count;
}

/* ---- N-D chunk container "EBCK" (reference src/ebcc_codec.c:920-1090, :1322-1449) --------------------- */
static size_t ceil_div(size_t a, size_t b) { return a / b + (a % b != 0); }

static size_t flat_offset(const size_t idx[NDIMS], const size_t dims[NDIMS])
{
    size_t o = idx[0];
    for (size_t i = 1; i < NDIMS; ++i) o = o * dims[i] + idx[i];
    return o;
}

static void chunk_origin(size_t linear, const size_t counts[NDIMS], const size_t cdims[NDIMS], size_t origin[NDIMS])
{
    for (size_t d = NDIMS; d-- > 0;) { origin[d] = (linear % counts[d]) * cdims[d]; linear /= counts[d]; }
}

/* edge chunks are padded by replicating the last valid sample along each axis (reference :339-351) */
static void gather_chunk_padded(const float *data, float *chunk, const size_t dims[NDIMS], const size_t origin[NDIMS],
                                const size_t cdims[NDIMS])
{
    size_t idx[NDIMS];
    for (size_t a = 0; a < cdims[0]; ++a) {
        idx[0] = origin[0] + a; if (idx[0] > dims[0] - 1) idx[0] = dims[0] - 1;
        for (size_t b = 0; b < cdims[1]; ++b) {
            idx[1] = origin[1] + b; if (idx[1] > dims[1] - 1) idx[1] = dims[1] - 1;
            float *dst = chunk + (a * cdims[1] + b) * cdims[2];
            for (size_t c = 0; c < cdims[2]; ++c) {
                idx[2] = origin[2] + c; if (idx[2] > dims[2] - 1) idx[2] = dims[2] - 1;
                dst[c] = data[flat_offset(idx, dims)];
            }
        }
    }
}

static void scatter_chunk_cropped(const float *chunk, float *data, const size_t dims[NDIMS], const size_t origin[NDIMS],
                                  const size_t cdims[NDIMS])
{
    size_t idx[NDIMS];
    for (size_t a = 0; a < cdims[0]; ++a) {
        idx[0] = origin[0] + a; if (idx[0] >= dims[0]) break;
        for (size_t b = 0; b < cdims[1]; ++b) {
            idx[1] = origin[1] + b; if (idx[1] >= dims[1]) break;
            const float *src = chunk + (a * cdims[1] + b) * cdims[2];
            for (size_t c = 0; c < cdims[2]; ++c) {
                idx[2] = origin[2] + c; if (idx[2] >= dims[2]) break;
                data[flat_offset(idx, dims)] = src[c];
            }
        }
    }
}

size_t ebcc_encode_chunking(float *data, codec_config_t *config, uint8_t **out_buffer)
{
    log_set_level_from_env();
    *out_buffer = NULL;
    size_t cdims[NDIMS];
    int all_zero = 1;
    for (size_t i = 0; i < NDIMS; ++i) { cdims[i] = config->chunk_dims[i]; if (cdims[i]) all_zero = 0; }
    if (all_zero) for (size_t i = 0; i < NDIMS; ++i) cdims[i] = config->dims[i];
    if (!dims_are_valid(cdims)) {
        ebcc_log(LOG_FATAL, "Invalid chunking dimensions: product(chunk_dims[0..%d]) and chunk_dims[%d] must be between %d and %d",
                 NDIMS - 2, NDIMS - 1, EBCC_MIN_INTERNAL_IMAGE_DIM, EBCC_MAX_INTERNAL_IMAGE_DIM);
        return 0;
    }
    size_t counts[NDIMS], chunk_size = 1, num_chunks = 1;
    for (size_t i = 0; i < NDIMS; ++i) {
        if (config->dims[i] == 0 || cdims[i] == 0) {
            ebcc_log(LOG_FATAL, "Invalid chunking dimensions: dims and chunk_dims must be non-zero");
            return 0;
        }
        counts[i] = ceil_div(config->dims[i], cdims[i]);
        if (!checked_mul(chunk_size, cdims[i], &chunk_size) || !checked_mul(num_chunks, counts[i], &num_chunks)) {
            ebcc_log(LOG_FATAL, "Invalid chunking dimensions: size overflow");
            return 0;
        }
    }
    /* gather every chunk (padded where it overhangs) into one contiguous batch for the device */
    size_t total = 0;
    if (!checked_mul(chunk_size, num_chunks, &total) || total > SIZE_MAX / sizeof(float)) {
        ebcc_log(LOG_FATAL, "Invalid chunking dimensions: padded size overflow");
        return 0;
    }
    int slabs = cdims[1] == config->dims[1] && cdims[2] == config->dims[2] && config->dims[0] % cdims[0] == 0;
    float *batch = data;
    if (!slabs) {
        batch = (float *)malloc(total * sizeof(float));
        if (!batch) { ebcc_log(LOG_FATAL, "Failed to allocate chunk buffer"); return 0; }
        for (size_t k = 0; k < num_chunks; ++k) {
            size_t origin[NDIMS];
            chunk_origin(k, counts, cdims, origin);
            gather_chunk_padded(data, batch + k * chunk_size, config->dims, origin, cdims);
        }
    }
    codec_config_t cc = *config;
    for (size_t i = 0; i < NDIMS; ++i) { cc.dims[i] = cdims[i]; cc.chunk_dims[i] = 0; }
    uint8_t **outs = (uint8_t **)calloc(num_chunks, sizeof(*outs));
    size_t *sizes = (size_t *)calloc(num_chunks, sizeof(size_t));
    size_t ok = ebcc_encode_batch(batch, num_chunks, &cc, outs, sizes);
    if (!slabs) free(batch);
    size_t out_size = sizeof(ebcc_chunking_header_t);
    for (size_t k = 0; k < num_chunks; ++k) out_size += sizeof(uint64_t) + sizes[k];
    uint8_t *out = NULL;
    if (ok == num_chunks) out = (uint8_t *)malloc(out_size);
    if (!out) {
        if (ok != num_chunks) ebcc_log(LOG_FATAL, "Failed to encode %lu of %lu chunks", num_chunks - ok, num_chunks);
        for (size_t k = 0; k < num_chunks; ++k) free(outs[k]);
        free(outs); free(sizes);
        return 0;
    }
    ebcc_chunking_header_t h;
    memset(&h, 0, sizeof(h));
    memcpy(h.magic, EBCC_CHUNKING_HEADER_MAGIC, 4);
    h.version = EBCC_CHUNKING_HEADER_VERSION;
    h.ndims = NDIMS;
    for (size_t i = 0; i < NDIMS; ++i) { h.dims[i] = config->dims[i]; h.chunk_dims[i] = cdims[i]; }
    h.num_chunks = num_chunks; h.chunk_size = chunk_size;
    uint8_t *it = out;
    memcpy(it, &h, sizeof(h)); it += sizeof(h);
    for (size_t k = 0; k < num_chunks; ++k) {
        uint64_t n = sizes[k];
        memcpy(it, &n, sizeof(n)); it += sizeof(n);
        memcpy(it, outs[k], sizes[k]); it += sizes[k];
        free(outs[k]);
    }
    free(outs); free(sizes);
    *out_buffer = out;
    return out_size;
}

size_t ebcc_encode_chunking_compat(float *data, codec_config_t *config, uint8_t **out_buffer)
{
    log_set_level_from_env();
    codec_config_t cc = *config;
    int all_zero = 1;
    for (size_t i = 0; i < NDIMS; ++i) if (cc.chunk_dims[i]) { all_zero = 0; break; }
    if (all_zero) {
        cc.chunk_dims[0] = 1;
        cc.chunk_dims[1] = cc.dims[1] > EBCC_MAX_INTERNAL_IMAGE_DIM ? 1024 : cc.dims[1];
        cc.chunk_dims[2] = cc.dims[2] > EBCC_MAX_INTERNAL_IMAGE_DIM ? 1024 : cc.dims[2];
        ebcc_log(LOG_INFO, "ebcc_encode_chunking_compat chunk dimensions: (%lu, %lu, %lu)", cc.chunk_dims[0], cc.chunk_dims[1], cc.chunk_dims[2]);
    }
    if (cc.residual_compression_type == RELATIVE_ERROR) {
        /* global-range relative error becomes an absolute bound (reference :1077-1087); this is
         * configuration arithmetic over the caller's array, not a codec stage */
        size_t total = 1;
        for (size_t i = 0; i < NDIMS; ++i)
            if (!checked_mul(total, cc.dims[i], &total)) total = 0;
        if (total == 0) { ebcc_log(LOG_FATAL, "Invalid EBCC dimensions: size overflow or zero-sized data"); return 0; }
        float lo = data[0], hi = data[0];
        for (size_t i = 0; i < total; ++i) {
            if (isnan(data[i]) || isinf(data[i])) { ebcc_log(LOG_FATAL, "NaN or Inf found in data at index %lu", i); exit(1); }
            if (data[i] > hi) hi = data[i];
            if (data[i] < lo) lo = data[i];
        }
        cc.error *= hi - lo;
        cc.residual_compression_type = MAX_ERROR;
    }
    return ebcc_encode_chunking(data, &cc, out_buffer);
}

size_t ebcc_decode_chunking(uint8_t *data, size_t data_size, float **out_buffer)
{
    log_set_level_from_env();
    if (data_size < 4 || memcmp(data, EBCC_CHUNKING_HEADER_MAGIC, 4) != 0)
        return ebcc_decode(data, data_size, out_buffer);       /* plain EBCC stream (reference :1326-1329) */
    ebcc_chunking_header_t h;
    if (data_size < sizeof(h)) { ebcc_log(LOG_FATAL, "Invalid chunked EBCC header: truncated"); return 0; }
    memcpy(&h, data, sizeof(h));
    if (h.version != EBCC_CHUNKING_HEADER_VERSION || h.ndims != NDIMS) {
        ebcc_log(LOG_FATAL, "Invalid chunked EBCC header: unsupported version %u or ndims %u", h.version, h.ndims);
        return 0;
    }
    size_t dims[NDIMS], cdims[NDIMS], counts[NDIMS], total = 1, chunk_size = 1, num_chunks = 1;
    for (size_t i = 0; i < NDIMS; ++i) {
        dims[i] = (size_t)h.dims[i]; cdims[i] = (size_t)h.chunk_dims[i];
        if (dims[i] == 0 || cdims[i] == 0) { ebcc_log(LOG_FATAL, "Invalid chunked EBCC header: zero dimension"); return 0; }
        counts[i] = ceil_div(dims[i], cdims[i]);
        if (!checked_mul(total, dims[i], &total) || !checked_mul(chunk_size, cdims[i], &chunk_size) ||
            !checked_mul(num_chunks, counts[i], &num_chunks)) {
            ebcc_log(LOG_FATAL, "Invalid chunked EBCC header: size overflow");
            return 0;
        }
    }
    if (num_chunks != h.num_chunks || chunk_size != h.chunk_size || !dims_are_valid(cdims)) {
        ebcc_log(LOG_FATAL, "Invalid chunked EBCC header: inconsistent chunk geometry");
        return 0;
    }
    /* sequential scan of the length prefixes, then one batched decode */
    uint8_t **streams = (uint8_t **)calloc(num_chunks, sizeof(*streams));
    size_t *ssz = (size_t *)calloc(num_chunks, sizeof(size_t));
    size_t *counts_out = (size_t *)calloc(num_chunks, sizeof(size_t));
    float **outs = (float **)calloc(num_chunks, sizeof(*outs));
    size_t off = sizeof(h), ok = 1;
    for (size_t k = 0; k < num_chunks && ok; ++k) {
        uint64_t n;
        if (data_size - off < sizeof(n)) { ok = 0; break; }
        memcpy(&n, data + off, sizeof(n)); off += sizeof(n);
        if (n > data_size - off) { ok = 0; break; }
        streams[k] = data + off; ssz[k] = (size_t)n; off += (size_t)n;
    }
    float *out = NULL;
    size_t ret = 0;
    if (!ok) ebcc_log(LOG_FATAL, "Invalid chunked EBCC stream: truncated chunk table");
    else if (decode_batch_impl(streams, ssz, num_chunks, outs, counts_out, 1, chunk_size) == num_chunks) {
        out = (float *)malloc(total * sizeof(float));
        for (size_t k = 0; k < num_chunks && out; ++k) {
            if (counts_out[k] != chunk_size) { ebcc_log(LOG_FATAL, "chunk %lu decoded to %lu values, expected %lu", k, counts_out[k], chunk_size); free(out); out = NULL; break; }
            size_t origin[NDIMS];
            chunk_origin(k, counts, cdims, origin);
            scatter_chunk_cropped(outs[k], out, dims, origin, cdims);
        }
        if (out) ret = total;
    }
    for (size_t k = 0; k < num_chunks; ++k) free(outs[k]);
    free(outs); free(counts_out); free(streams); free(ssz);
    *out_buffer = out;
    return ret;
}
