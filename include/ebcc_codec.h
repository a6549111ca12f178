/*
 * ebcc_codec.h -- public C API of the B200-native EBCC codec.  Drop-in for the reference's
 * src/ebcc_codec.h:16-49: same symbols, argument meaning, ownership and return conventions, so
 * that h5py (filter 308), the ctypes zarr codec (ebcc/zarr_filter.py) and tests/test_c_api.py of
 * the reference run unchanged against ebcc_b200/libh5z_ebcc.so.
 *
 *   ebcc_encode / ebcc_decode           reference src/ebcc_codec.c:607 / :1215
 *   ebcc_encode_chunking[_compat]       reference src/ebcc_codec.c:920 / :1054
 *   ebcc_decode_chunking                reference src/ebcc_codec.c:1322
 *   free_buffer, print_config,
 *   log_set_level_from_env              reference src/ebcc_codec.c:1451, :414, :430
 *   populate_config, H5PLget_plugin_*   reference src/h5z_ebcc.c:38, :27-28
 *
 * Additive (not in the reference): ebcc_encode_batch / ebcc_decode_batch expose the batched device
 * pipeline; each frame's stream is identical to a standalone ebcc_encode of that frame.
 */
#ifndef EBCC_CODEC_H
#define EBCC_CODEC_H

#include <stddef.h>
#include <stdint.h>

#ifndef EBCC_API
#define EBCC_API __attribute__((visibility("default")))
#endif

#define NDIMS 3
#define EBCC_MIN_INTERNAL_IMAGE_DIM 32
#define EBCC_MAX_INTERNAL_IMAGE_DIM 2047
#define EBCC_VERSION_MAJOR 0
#define EBCC_VERSION_MINOR 1
#define EBCC_HEADER_VERSION 1
#define EBCC_HEADER_FLAG_CONST_FIELD 0x01
#define EBCC_HEADER_MAGIC "EBCC"
#define EBCC_CHUNKING_HEADER_VERSION 1
#define EBCC_CHUNKING_HEADER_MAGIC "EBCK"

#ifdef __cplusplus
extern "C" {
#endif

typedef enum { NONE, MAX_ERROR, RELATIVE_ERROR } residual_t;

typedef struct {
    size_t dims[NDIMS];
    float base_cr;
    residual_t residual_compression_type;
    float residual_cr;
    float error;
    size_t chunk_dims[NDIMS];
} codec_config_t;

EBCC_API size_t ebcc_encode(float *data, codec_config_t *config, uint8_t **out_buffer);
EBCC_API size_t ebcc_decode(uint8_t *data, size_t data_size, float **out_buffer);
EBCC_API size_t ebcc_encode_chunking(float *data, codec_config_t *config, uint8_t **out_buffer);
EBCC_API size_t ebcc_encode_chunking_compat(float *data, codec_config_t *config, uint8_t **out_buffer);
EBCC_API size_t ebcc_decode_chunking(uint8_t *data, size_t data_size, float **out_buffer);
EBCC_API void free_buffer(void *buffer);
EBCC_API void print_config(codec_config_t *config);
EBCC_API void log_set_level_from_env(void);
EBCC_API void populate_config(codec_config_t *config, size_t cd_nelmts, const unsigned int cd_values[], size_t buf_size);

/* Batched entry points.  frames: nframes contiguous frames of prod(config->dims) floats each; a host pointer or a
 * CUDA device pointer (unified virtual addressing: frames produced on the GPU are compressed without a host round
 * trip; the work producing them must have completed, the codec runs on its own streams).  The same holds for
 * ebcc_encode's data and for caller-provided outs[i] of ebcc_decode_batch.
 * outs[i] receives a malloc'ed stream (free with free_buffer), sizes[i] its length (0 = failure).
 * Returns the number of frames encoded successfully. */
EBCC_API size_t ebcc_encode_batch(const float *frames, size_t nframes, const codec_config_t *config,
                                  uint8_t **outs, size_t *sizes);
/* streams[i]/stream_sizes[i]: nframes EBCC streams; outs[i] receives a malloc'ed float array (or is
 * written in place when non-NULL on entry), counts[i] the element count (0 = failure). */
EBCC_API size_t ebcc_decode_batch(uint8_t *const *streams, const size_t *stream_sizes, size_t nframes,
                                  float **outs, size_t *counts);

#ifdef __cplusplus
}
#endif
#endif
