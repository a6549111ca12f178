/*
 * h5z_ebcc.c -- HDF5 dynamically-loaded filter 308 over the B200 codec.  Same plugin surface as the
 * reference's src/h5z_ebcc.c (H5PLget_plugin_type/info :27-28, populate_config :38-93, filter
 * callback :124-148), compiled without HDF5 headers like the reference (src/hdf5_stub.h:29-52).
 */
#include "../../include/ebcc_codec.h"
#include <assert.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* the handful of HDF5 plugin types the loader needs (values fixed by the HDF5 ABI) */
typedef int H5Z_filter_t;
typedef size_t (*H5Z_func_t)(unsigned int flags, size_t cd_nelmts, const unsigned int cd_values[], size_t nbytes,
                             size_t *buf_size, void **buf);
typedef struct {
    int version;
    H5Z_filter_t id;
    unsigned encoder_present;
    unsigned decoder_present;
    const char *name;
    void *can_apply;
    void *set_local;
    H5Z_func_t filter;
} H5Z_class2_t;
typedef enum { H5PL_TYPE_ERROR = -1, H5PL_TYPE_FILTER = 0, H5PL_TYPE_VOL = 1, H5PL_TYPE_VFD = 2, H5PL_TYPE_NONE = 3 } H5PL_type_t;
#define H5Z_CLASS_T_VERS 1
#define H5Z_FLAG_REVERSE 0x0100
#define H5Z_FILTER_EBCC 308

static size_t H5Z_filter_ebcc(unsigned int flags, size_t cd_nelmts, const unsigned int cd_values[], size_t nbytes,
                              size_t *buf_size, void **buf);

const H5Z_class2_t H5Z_EBCC[1] = {{H5Z_CLASS_T_VERS, (H5Z_filter_t)H5Z_FILTER_EBCC, 1, 1, "HDF5 EBCC filter L&L", NULL, NULL,
                                   (H5Z_func_t)H5Z_filter_ebcc}};

EBCC_API H5PL_type_t H5PLget_plugin_type(void) { return H5PL_TYPE_FILTER; }
EBCC_API const void *H5PLget_plugin_info(void) { return H5Z_EBCC; }

static float bits_as_float(unsigned int v) { float f; memcpy(&f, &v, sizeof f); return f; }

#define PC_FATAL(...) do { fprintf(stderr, "FATAL ebcc_b200: " __VA_ARGS__); fputc('\n', stderr); exit(1); } while (0)

/* cd_values = [tile_height, tile_width, f32bits(base_cr), mode, f32bits(error)] */
void populate_config(codec_config_t *config, size_t cd_nelmts, const unsigned int cd_values[], size_t buf_size)
{
    if (cd_nelmts < 4) PC_FATAL("EBCC filter requires at least 4 configuration values, got %lu", cd_nelmts);
    for (size_t i = 0; i < NDIMS; ++i) config->chunk_dims[i] = 0;
    const size_t th = cd_values[0], tw = cd_values[1];
    if (th < EBCC_MIN_INTERNAL_IMAGE_DIM || tw < EBCC_MIN_INTERNAL_IMAGE_DIM || th > EBCC_MAX_INTERNAL_IMAGE_DIM ||
        tw > EBCC_MAX_INTERNAL_IMAGE_DIM)
        PC_FATAL("Tile size %lu x %lu is invalid, each dimension must be between %d and %d", th, tw,
                 EBCC_MIN_INTERNAL_IMAGE_DIM, EBCC_MAX_INTERNAL_IMAGE_DIM);
    const size_t tile = th * tw, nfloats = buf_size / sizeof(float);
    if (nfloats < tile) PC_FATAL("Buffer size %lu is smaller than the tile size %lu x %lu = %lu", nfloats, th, tw, tile);
    if (nfloats % tile != 0) PC_FATAL("Buffer size %lu is not divisible by the tile size %lu x %lu = %lu", nfloats, th, tw, tile);
    config->dims[0] = nfloats / tile; config->dims[1] = th; config->dims[2] = tw;
    if (config->dims[0] > EBCC_MAX_INTERNAL_IMAGE_DIM / config->dims[1])
        PC_FATAL("Flattened EBCC image height %lu x %lu exceeds the limit of %d", config->dims[0], config->dims[1],
                 EBCC_MAX_INTERNAL_IMAGE_DIM);
    config->base_cr = bits_as_float(cd_values[2]);
    config->residual_compression_type = (residual_t)cd_values[3];
    switch (config->residual_compression_type) {
    case NONE: assert(cd_nelmts == 4); break;
    case MAX_ERROR:
    case RELATIVE_ERROR: assert(cd_nelmts == 5); config->error = bits_as_float(cd_values[4]); break;
    }
}

static size_t H5Z_filter_ebcc(unsigned int flags, size_t cd_nelmts, const unsigned int cd_values[], size_t nbytes,
                              size_t *buf_size, void **buf)
{
    if (flags & H5Z_FLAG_REVERSE) {
        float *out = NULL;
        /* like the reference (:129-134) the decode path reports the ELEMENT count */
        *buf_size = ebcc_decode((uint8_t *)*buf, nbytes, &out);
        free_buffer(*buf);
        *buf = out;
        return *buf_size;
    }
    codec_config_t config;
    memset(&config, 0, sizeof(config));
    populate_config(&config, cd_nelmts, cd_values, *buf_size);
    uint8_t *out = NULL;
    *buf_size = ebcc_encode((float *)*buf, &config, &out);
    free_buffer(*buf);
    *buf = out;
    return *buf_size;
}
