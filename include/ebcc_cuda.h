/*
 * ebcc_cuda.h -- thin C ABI between the host C codec (ebcc_b200/csrc/ebcc_codec.c) and the CUDA
 * device layer (ebcc_b200/csrc/ebcc_cuda.cu, sm_100a).  Plain pointers and sizes only.
 *
 * Every entry point replaces a CPU stage of the reference:
 *   ebcc_cuda_j2k_encode_batch  <- findMinMaxf + u16 scaling + j2k_encode_internal
 *                                  (src/ebcc_codec.c:676-693, :105-180 -> OpenJPEG opj_encode)
 * All functions return 0 on success and a negative value on failure (after logging to stderr);
 * there is no CPU fallback: without a CUDA device every call fails.
 */
#ifndef EBCC_CUDA_H
#define EBCC_CUDA_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct ebcc_cuda_ctx ebcc_cuda_ctx;

#define EBCC_FRAME_CONST 1      /* min == max: no codestream produced          */
#define EBCC_FRAME_NANINF 2     /* NaN or Inf in the frame (reference exit(1)) */
#define EBCC_FRAME_OVERFLOW 4   /* a code-block outgrew its work-buffer slice (cannot happen within the stated bounds): no stream */

/* stage names for ebcc_cuda_stage_ms */
enum {
    EBCC_ST_H2D = 0, EBCC_ST_MINMAX, EBCC_ST_DWT, EBCC_ST_T1_CM, EBCC_ST_T1_MQ, EBCC_ST_PCRD_T2,
    EBCC_ST_BODIES, EBCC_ST_D2H, EBCC_ST_COUNT
};

/* kernel classes of ebcc_cuda_kernel_timing_report */
enum {
    EBCC_KT_MINMAX = 0, EBCC_KT_DWT, EBCC_KT_T1_CM, EBCC_KT_T1_MQ, EBCC_KT_PCRD_T2, EBCC_KT_BODIES, EBCC_KT_T1_DEC, EBCC_KT_RECON,
    EBCC_KT_IDWT, EBCC_KT_FINALIZE, EBCC_KT_SP_FWD, EBCC_KT_SP_ENCODE, EBCC_KT_SP_PROBE, EBCC_KT_ZSTD_ENC, EBCC_KT_T2_PARSE,
    EBCC_KT_ZSTD_DEC, EBCC_KT_SP_DECODE, EBCC_KT_SP_INV, EBCC_KT_COUNT
};
#define EBCC_KT_NAMES "minmax,dwt,t1_cm,t1_mq,pcrd_t2,bodies,t1_dec,recon,idwt,finalize,sp_fwd,sp_encode,sp_probe,zstd_enc,t2_parse,zstd_dec,sp_decode,sp_inv"

int ebcc_cuda_device_count(void);

/* process-wide: enable != 0 clears the accumulators and brackets every launch group of every pipeline with CUDA events on
 * its stream; the report returns, per kernel class, the summed device milliseconds and the number of launch groups timed
 * since then (read back at the pipelines' synchronisation points).  Returns EBCC_KT_COUNT. */
void ebcc_cuda_kernel_timing(int enable);
int ebcc_cuda_kernel_timing_report(double *ms, unsigned long long *groups, int n);

/* one context = one device + one stream + reusable work buffers; not thread-safe (callers lock) */
ebcc_cuda_ctx *ebcc_cuda_ctx_create(int device);
void ebcc_cuda_ctx_destroy(ebcc_cuda_ctx *ctx);

/* size the work buffers for batches of up to nframes frames of height x width samples */
int ebcc_cuda_reserve(ebcc_cuda_ctx *ctx, size_t height, size_t width, size_t nframes);

/* multi-tile internal images (dims[0] > 1; j2k_encode_internal with n_tiles > 1, src/ebcc_codec.c:105-180): every
 * frame of the batch is an image of ntiles tiles of height x width stacked vertically, (ntiles * height) x width
 * samples; upload / analyze / rate_probe / residual_* / download then work on whole images, each tile being coded
 * with the geometry of its canvas position.  ntiles == 1 is ebcc_cuda_reserve. */
int ebcc_cuda_reserve_tiled(ebcc_cuda_ctx *ctx, size_t ntiles, size_t height, size_t width, size_t nframes);

/* host -> device copy of nframes frames (float32, C-contiguous) into the context's input buffer */
int ebcc_cuda_upload(ebcc_cuda_ctx *ctx, const float *frames, size_t nframes);

/* device pass over the resident batch: min/max, scaling, DWT, tier-1, PCRD, tier-2.
 * base_cr is EBCC's compression ratio (OpenJPEG rate = base_cr / 2).  Asynchronous. */
int ebcc_cuda_j2k_encode_resident(ebcc_cuda_ctx *ctx, size_t nframes, float base_cr);

/* The encoder in two stages, for the compression-ratio search of MAX_ERROR / RELATIVE_ERROR
 * (error_bound_j2k_compression, src/ebcc_codec.c:545-596): analyze = everything independent of the ratio
 * (min/max, scaling, DWT, tier-1), run once; rate_probe = PCRD + tier-2 at cr[i] for frame i followed by a
 * real decode of the truncated code-block segments and the error statistics against the original frame
 * (emulate_j2k_compression + get_error_target_quantile / get_max_error / get_mean_error / findMinMaxf on the
 * residual, src/ebcc_codec.c:464-543).  with_streams != 0 also assembles the codestreams for ebcc_cuda_download. */
typedef struct {
    unsigned long long n_over;   /* samples with |d - decoded| > targets[i] */
    float max_err;               /* max |d - decoded| */
    float rmin, rmax;            /* min / max of d - decoded */
    double sum_err;              /* sum of (d - decoded), double accumulation */
    size_t stream_bytes;         /* codestream length at this ratio */
} ebcc_probe_result;
/* for_search != 0: the analysis also leaves behind what the probes of the compression-ratio search need -- the
 * tier-1 recording (decoded value + pass of significance per sample, straight from the encoder's context modelling)
 * and, per code-block and pass, whether the segment cut at that pass decodes like the complete one */
int ebcc_cuda_analyze(ebcc_cuda_ctx *ctx, size_t nframes, int for_search);
int ebcc_cuda_analyze_resident(ebcc_cuda_ctx *ctx, size_t nframes);   /* = ebcc_cuda_analyze(ctx, nframes, 1) */
int ebcc_cuda_rate_probe(ebcc_cuda_ctx *ctx, size_t nframes, const float *cr, const float *targets, int with_streams,
                         ebcc_probe_result *res);

/* a probe of the search itself: the host only compares n_over with the quantile target, so the count has to be exact only
 * below cap; a frame that reaches cap violations is not examined further (its n_over is some value >= cap, its other
 * statistics cover the examined part only).  cap == 0: ebcc_cuda_rate_probe(with_streams = 0). */
int ebcc_cuda_rate_probe_capped(ebcc_cuda_ctx *ctx, size_t nframes, const float *cr, const float *targets, unsigned long long cap,
                                ebcc_probe_result *res);
int ebcc_cuda_rate_probe_streams_only(ebcc_cuda_ctx *ctx, size_t nframes, const float *cr);
int ebcc_cuda_frame_stats(ebcc_cuda_ctx *ctx, size_t nframes, float *minv, float *maxv, int *flags);

/* No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops).  The min / max / flags stage of the encoder
 * (what ebcc_cuda_frame_stats reports after an analysis pass) on nframes device-resident frames of samples_per_frame
 * floats each, on the CURRENT device, enqueued on `stream` (a cudaStream_t, NULL = default stream); the results land in
 * the caller's DEVICE arrays d_minv / d_maxv / d_flags once the stream has run (the call does not synchronise).
 * Allocates nothing: the per-frame reduction records live in a static __device__ array of
 * EBCC_NOALLOC_MAX_FRAMES entries (128 KiB per GPU, held until exit); more frames per call are rejected with -1.
 * All calls share that array: calls that overlap on the GPU race, so order them on one stream or with events. */
#define EBCC_NOALLOC_MAX_FRAMES 4096
int ebcc_cuda_frame_stats_noalloc(const float *d_frames, size_t nframes, size_t samples_per_frame,
                                  float *d_minv, float *d_maxv, int *d_flags, void *stream);

/* Residual layer on the resident batch (src/ebcc_codec.c:744-817): SPIHT encode of the normalised residual of the
 * last rate probe with a bit budget (spiht_encode), probes of the truncation search (spiht_decode of a prefix +
 * get_max_error / get_mean_error), and the zstd stage.  active[i] == 0 skips frame i. */
int ebcc_cuda_residual_encode(ebcc_cuda_ctx *ctx, size_t nframes, const int *active, const size_t *trunc_bits, size_t *coeffs_size);
int ebcc_cuda_residual_probe(ebcc_cuda_ctx *ctx, size_t nframes, const int *active, const size_t *num_bits, float *max_err, double *sum_err);
int ebcc_cuda_residual_pack(ebcc_cuda_ctx *ctx, size_t nframes, const size_t *coeffs_size, const uint8_t **frames, size_t *sizes);

/* device -> host: waits for the pass, then returns per-frame codestream sizes, min, max, flags;
 * streams[i] receives a pointer into a context-owned pinned buffer; two buffers take turns, so the pointers stay valid
 * until the download after the next one on this context */
int ebcc_cuda_download(ebcc_cuda_ctx *ctx, size_t nframes, const uint8_t **streams, size_t *sizes,
                       float *minv, float *maxv, int *flags);

/* convenience: reserve + upload + encode + download */
int ebcc_cuda_j2k_encode_batch(ebcc_cuda_ctx *ctx, const float *frames, size_t nframes, size_t height,
                               size_t width, float base_cr, const uint8_t **streams, size_t *sizes,
                               float *minv, float *maxv, int *flags);

/* residual layer of one frame as stored in its EBCC stream (src/ebcc_codec.c:1294-1312); zstd_size == 0: none */
typedef struct {
    const uint8_t *zstd;
    size_t zstd_size;
    size_t coeffs_size;      /* SPIHT bytes after zstd decompression */
    float rmin, rmax;        /* residual range used to de-normalise */
} ebcc_residual_in;

/* replaces j2k_decode_internal + ZSTD_decompress + spiht_decode + the residual add of ebcc_decode
 * (src/ebcc_codec.c:1092-1136, :1294-1312): nframes J2K codestreams of the same H x W geometry are
 * decoded on the device and written to outs[i] (host, H*W floats each). */
int ebcc_cuda_decode_batch(ebcc_cuda_ctx *ctx, const uint8_t *const *j2k, const size_t *j2k_sizes, size_t nframes,
                           size_t height, size_t width, const float *minv, const float *maxv,
                           const ebcc_residual_in *res, float *const *outs);

/* the same for multi-tile codestreams: ntiles tiles of height x width per image, outs[i] holds ntiles * height * width floats */
int ebcc_cuda_decode_batch_tiled(ebcc_cuda_ctx *ctx, const uint8_t *const *j2k, const size_t *j2k_sizes, size_t nframes,
                                 size_t ntiles, size_t height, size_t width, const float *minv, const float *maxv,
                                 const ebcc_residual_in *res, float *const *outs);

/* hint from the host codec: how many pipelines (contexts) are at work on this device right now.  The serial tier-1
 * coders step more code-blocks per warp when the device is shared (fewer instructions) and fewer when a call has the
 * GPU to itself (shorter chains); results do not depend on it. */
void ebcc_cuda_set_crowd(ebcc_cuda_ctx *ctx, int pipelines_at_work);

/* copies bytes from host memory to dst, which may be host or device memory (caller-provided decode destinations) */
int ebcc_cuda_copy_out(void *dst, const void *src, size_t bytes);

/* profiling: when enabled, CUDA events bracket every stage on the context's stream */
void ebcc_cuda_set_profiling(ebcc_cuda_ctx *ctx, int enabled);
/* milliseconds spent in a stage during the last pass (valid after ebcc_cuda_sync) */
float ebcc_cuda_stage_ms(ebcc_cuda_ctx *ctx, int stage);
int ebcc_cuda_sync(ebcc_cuda_ctx *ctx);
/* kernels launched by every context of the process so far */
unsigned long long ebcc_cuda_total_launches(void);
/* number of kernel launches issued by this context so far */
unsigned long long ebcc_cuda_launch_count(ebcc_cuda_ctx *ctx);

/* bench helper: `warmup` untimed + `steps` timed device passes over the resident batch, L2 flushed
 * (flush_bytes written) between passes outside the timed region; CUDA-event times on the context's
 * stream.  total_ms: sum over timed passes; stage_ms_sum[EBCC_ST_COUNT]: per-stage sums. */
int ebcc_cuda_bench_resident(ebcc_cuda_ctx *ctx, size_t nframes, float base_cr, int steps, int warmup,
                             size_t flush_bytes, float *total_ms, float *stage_ms_sum);

/* Same measurement with several resident batches in flight: step i runs on ctxs[i % nctx], each context on its own
 * stream (the per-batch hot path is unchanged; kernels of different batches overlap).  The flush buffer is written
 * before every step inside the timed region.  total_ms: device time of all `steps` timed steps. */
int ebcc_cuda_bench_resident_pipelined(ebcc_cuda_ctx **ctxs, int nctx, size_t nframes, float base_cr, int steps, int warmup,
                                       size_t flush_bytes, float *total_ms);

/* measurement tap: EBCOT symbols (one byte each, k_t1_cm -> k_t1_mq) and MQ-coded bytes tier-1 produced for the first
 * nframes frames of the resident batch -- the algorithmic traffic of the tier-1 kernels */
int ebcc_cuda_t1_totals(ebcc_cuda_ctx *ctx, size_t nframes, unsigned long long *symbols, unsigned long long *mq_bytes);

/* debug/test taps on the resident batch (device -> host copies of intermediates) */
int ebcc_cuda_debug_coef(ebcc_cuda_ctx *ctx, size_t frame, float *dst);
int ebcc_cuda_debug_check_cuts(ebcc_cuda_ctx *ctx, size_t nframes, unsigned long long counts[6]);
int ebcc_cuda_debug_spiht_stream(ebcc_cuda_ctx *ctx, size_t frame, uint8_t *dst, size_t cap, size_t *len);
/* one zstd frame through the device decompressor alone (tests: frames written by the reference's ZSTD_compress) */
int ebcc_cuda_debug_zstd_decompress(int device, const uint8_t *frame, size_t n, uint8_t *out, size_t cap, size_t *out_len);
int ebcc_cuda_debug_cblk(ebcc_cuda_ctx *ctx, size_t frame, size_t cblk, int *numbps, int *totalpasses,
                         int *rates, double *dists, int *numpasses);

#ifdef __cplusplus
}
#endif
#endif
