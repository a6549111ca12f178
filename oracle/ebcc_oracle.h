/*
 * ebcc_oracle.h -- CPU restatement of the EBCC per-frame hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is linked into, imported by or
 * executed from the product library (ebcc_b200/).  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may use it, and only as the checker.
 *
 * Parity pinning: every function here is checked byte-for-byte (integers, bytes) or
 * bit-for-bit (floats) against oracle/_ref/libh5z_ebcc_ref.so, which is the UNMODIFIED
 * reference compiled by oracle/Makefile from /root/reference/src (OpenJPEG 2.5.3 @210a8a5,
 * zstd 1.5.6 @794ea1b, EBCC @e836fbd); see tests/test_oracle_vs_ref.py and tests/golden/.
 *
 * Plain sequential C, written from the reference's behaviour (file:line cited per function),
 * not copied from it.
 */
#ifndef EBCC_ORACLE_H
#define EBCC_ORACLE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORC_MAXPASSES 100
#define ORC_NRES 6

typedef struct {
    int resno, orient;       /* orient: 0 LL, 1 HL, 2 LH, 3 HH (OPJ bandno) */
    int x0, y0, x1, y1;      /* band rectangle in band coordinates */
    int offx, offy;          /* top-left of the band inside the Mallat-layout tile buffer */
    int level;               /* numresolutions-1-resno */
    float stepsize;          /* encoder-side Delta_b (with sub-band gain) */
    float stepsize_dec;      /* decoder-side Delta_b (log2_gain = 0, OPJ two_invK quirk) */
    int numbps;              /* Mb */
    int expn, mant;
    int cw, ch;              /* code-block grid */
    int first_cblk;          /* index of the first code-block of this band */
} orc_band_t;

typedef struct {
    int band;
    int x0, y0, x1, y1;      /* in band coordinates */
    int tx, ty;              /* top-left inside the tile buffer */
    int w, h;
    int numbps;              /* magnitude bit-planes actually present */
    int totalpasses;
    int rate[ORC_MAXPASSES];        /* cumulative bytes after each pass (after OPJ's fix-ups) */
    int len[ORC_MAXPASSES];
    double distortiondec[ORC_MAXPASSES]; /* cumulative */
    uint8_t *data;           /* MQ bytes, capacity 74 + w*h*4 like OPJ */
    int datalen;             /* total MQ bytes produced */
    int numpasses;           /* passes selected by PCRD (single layer) */
    int prev_numpasses;      /* layer->numpasses memory of opj_tcd_makelayer */
} orc_cblk_t;

typedef struct {
    int w, h;
    int y0;                  /* canvas row of the tile's first line (tile t of a multi-tile image: t * h) */
    int cas_v[ORC_NRES];     /* row parity of each resolution's canvas origin (dwt.c:1986-1987) */
    int nbands, ncblks;
    int res_w[ORC_NRES], res_h[ORC_NRES];
    orc_band_t bands[3 * ORC_NRES - 2];
    orc_cblk_t *cblks;
    float *coef;             /* DWT coefficients, Mallat layout, w*h */
} orc_tile_t;

/* --- driver-level scalar stages (src/ebcc_codec.c) --- */
void orc_minmax(const float *d, size_t n, float *mn, float *mx);
void orc_scale_u16(const float *d, size_t n, float mn, float mx, uint16_t *out);

/* --- JPEG 2000 encoder as EBCC configures OpenJPEG --- */
orc_tile_t *orc_tile_create(int h, int w);
/* tile of h x w lines whose first line is canvas row y0 (tcd.c:961-1264 with a non-zero tile origin) */
orc_tile_t *orc_tile_create_at(int h, int w, int y0);
void orc_tile_destroy(orc_tile_t *t);
void orc_dwt97_forward(orc_tile_t *t, const uint16_t *img);   /* DC shift + 5-level 9/7 */
void orc_t1_encode_all(orc_tile_t *t);                         /* quantise + EBCOT tier-1 */
double orc_rate_allocate(orc_tile_t *t, float base_cr, int main_header_bytes); /* PCRD; returns threshold */
double orc_rate_allocate_tiled(orc_tile_t *t, float base_cr, int main_header_bytes, int ntiles);
size_t orc_t2_encode(orc_tile_t *t, uint8_t *dst, size_t cap, int dry_run);
size_t orc_main_header(int h, int w, uint8_t *dst);
/* whole J2K encode: returns malloc'ed codestream */
size_t orc_j2k_encode(const uint16_t *img, int h, int w, float base_cr, uint8_t **out);
/* multi-tile image (j2k_encode_internal with n_tiles > 1, src/ebcc_codec.c:105-180): ntiles tiles of h x w stacked
 * vertically, img holds ntiles * h * w samples; main header | one tile-part per tile | EOC */
size_t orc_main_header_tiled(int himg, int h, int w, uint8_t *dst);
size_t orc_j2k_encode_tiled(const uint16_t *img, int ntiles, int h, int w, float base_cr, uint8_t **out);

/* test helpers */
void orc_get_luts(uint8_t *zc /*2048*/, uint8_t *sc /*256*/, uint8_t *spb /*256*/,
                  int16_t *nm_sig, int16_t *nm_sig0, int16_t *nm_ref, int16_t *nm_ref0 /*128 each*/);
void orc_get_mq_table(uint16_t *qe, uint8_t *nmps, uint8_t *nlps, uint8_t *sw /*47 each*/);

#ifdef __cplusplus
}
#endif
#endif
