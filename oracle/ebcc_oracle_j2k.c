/*
 * ebcc_oracle_j2k.c -- CPU restatement of the JPEG 2000 encoder exactly as EBCC drives
 * OpenJPEG 2.5.3 (1 component, 16-bit unsigned, irreversible 9/7, 6 resolutions, 64x64
 * code-blocks, LRCP, 1 layer, 1 precinct per resolution, cblksty 0, no SOP/EPH).
 *
 * TEST INFRASTRUCTURE ONLY (see ebcc_oracle.h).  Pinned byte-for-byte against
 * oracle/_ref (j2k_encode_internal, src/ebcc_codec.c:105-180) by tests/test_oracle_vs_ref.py.
 *
 * Compile with -ffp-contract=off: OpenJPEG on x86-64 never fuses multiply-add.
 */
#include "ebcc_oracle.h"
#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------------------------ */
/* small integer helpers (opj_intmath.h semantics)                                       */
static int ceildivpow2(int a, int b) { return (int)(((int64_t)a + ((int64_t)1 << b) - 1) >> b); }
static int floorlog2(int a) { int l = 0; while (a > 1) { a >>= 1; l++; } return l; }
static int imin(int a, int b) { return a < b ? a : b; }
static int imax(int a, int b) { return a > b ? a : b; }

/* ------------------------------------------------------------------------------------ */
/* driver scalar stages: src/ebcc_codec.c:515-533 (findMinMaxf), :686-689 (u16 scaling)   */
void orc_minmax(const float *d, size_t n, float *mn, float *mx)
{
    float lo = d[0], hi = d[0];
    for (size_t i = 0; i < n; ++i) {
        if (d[i] < lo) lo = d[i];
        if (d[i] > hi) hi = d[i];
    }
    *mn = lo; *mx = hi;
}

void orc_scale_u16(const float *d, size_t n, float mn, float mx, uint16_t *out)
{
    /* ((d-min)/(max-min)) * (uint16_t)-1 : f32 sub, f32 div, f32 mul by 65535.0f, C truncation */
    const float range = mx - mn;
    for (size_t i = 0; i < n; ++i) {
        float v = ((d[i] - mn) / range) * 65535.0f;
        out[i] = (uint16_t)v;
    }
}

/* ------------------------------------------------------------------------------------ */
/* geometry + quantisation step sizes                                                   */
/* 9/7 sub-band norms (dwt.c:212-217); indexed [orient][level]                           */
static const double norms97[4][10] = {
    {1.000, 1.965, 4.177, 8.403, 16.90, 33.84, 67.69, 135.3, 270.6, 540.9},
    {2.022, 3.989, 8.355, 17.04, 34.27, 68.63, 137.3, 274.6, 549.0, 0},
    {2.022, 3.989, 8.355, 17.04, 34.27, 68.63, 137.3, 274.6, 549.0, 0},
    {2.080, 3.865, 8.307, 17.18, 34.71, 69.59, 139.3, 278.6, 557.2, 0}
};

static double getnorm97(int level, int orient)
{
    if (orient == 0 && level >= 10) level = 9;
    else if (orient > 0 && level >= 9) level = 8;
    return norms97[orient][level];
}

orc_tile_t *orc_tile_create(int h, int w) { return orc_tile_create_at(h, w, 0); }

/* tcd.c:961-1264 for one tile whose first line is canvas row y0 (x origin 0): resolution and band rectangles are
 * ceil-divisions of the canvas coordinates y0 and y0 + h (tcd.c:961-1055); the code-block grid is anchored at
 * multiples of 64 of the band's canvas rows (tcd.c:1122-1230), so the first block row of a band may be partial */
orc_tile_t *orc_tile_create_at(int h, int w, int y0)
{
    orc_tile_t *t = (orc_tile_t *)calloc(1, sizeof(*t));
    t->w = w; t->h = h; t->y0 = y0;
    for (int r = 0; r < ORC_NRES; ++r) {
        int lvl = ORC_NRES - 1 - r;
        t->res_w[r] = ceildivpow2(w, lvl);
        t->res_h[r] = ceildivpow2(y0 + h, lvl) - ceildivpow2(y0, lvl);
        t->cas_v[r] = ceildivpow2(y0, lvl) & 1;
    }
    int nb = 0, ncb = 0;
    for (int r = 0; r < ORC_NRES; ++r) {
        int lvl = ORC_NRES - 1 - r;
        int nbands = r == 0 ? 1 : 3;
        for (int bi = 0; bi < nbands; ++bi) {
            orc_band_t *b = &t->bands[nb];
            int orient = r == 0 ? 0 : bi + 1;
            b->resno = r; b->orient = orient; b->level = lvl;
            if (r == 0) {
                b->x0 = 0; b->y0 = ceildivpow2(y0, lvl);
                b->x1 = ceildivpow2(w, lvl); b->y1 = ceildivpow2(y0 + h, lvl);
                b->offx = 0; b->offy = 0;
            } else {
                int x0b = orient & 1, y0b = orient >> 1;
                b->x0 = 0; b->y0 = ceildivpow2(y0 - (y0b << lvl), lvl + 1);
                b->x1 = ceildivpow2(w - (x0b << lvl), lvl + 1);
                b->y1 = ceildivpow2(y0 + h - (y0b << lvl), lvl + 1);
                b->offx = x0b ? t->res_w[r - 1] : 0;
                b->offy = y0b ? t->res_h[r - 1] : 0;
            }
            /* dwt.c:2176-2198 + 1302-1311: explicit step sizes, prec 16, gain 0 for 9/7 */
            {
                double stepsize = 1.0 / getnorm97(lvl, orient);
                int si = (int)floor(stepsize * 8192.0);
                int p = floorlog2(si) - 13;
                int n = 11 - floorlog2(si);
                b->mant = (n < 0 ? si >> -n : si << n) & 0x7ff;
                b->expn = 16 - p;
            }
            /* tcd.c:1069-1089 */
            {
                int log2_gain = orient == 0 ? 0 : (orient == 3 ? 2 : 1);
                b->stepsize = (float)((1.0 + b->mant / 2048.0) * pow(2.0, (16 + log2_gain) - b->expn));
                b->stepsize_dec = (float)((1.0 + b->mant / 2048.0) * pow(2.0, 16 - b->expn));
                b->numbps = b->expn + 2 - 1;
            }
            int bw = b->x1 - b->x0, bh = b->y1 - b->y0;
            b->cw = bw > 0 && bh > 0 ? (bw + 63) / 64 : 0;
            b->ch = bw > 0 && bh > 0 ? (b->y1 + 63) / 64 - b->y0 / 64 : 0;
            b->first_cblk = ncb;
            ncb += b->cw * b->ch;
            nb++;
        }
    }
    t->nbands = nb; t->ncblks = ncb;
    t->cblks = (orc_cblk_t *)calloc((size_t)ncb, sizeof(orc_cblk_t));
    for (int bi = 0; bi < nb; ++bi) {
        orc_band_t *b = &t->bands[bi];
        for (int cy = 0; cy < b->ch; ++cy)
            for (int cx = 0; cx < b->cw; ++cx) {
                orc_cblk_t *c = &t->cblks[b->first_cblk + cy * b->cw + cx];
                c->band = bi;
                c->x0 = cx * 64; c->y0 = imax((b->y0 / 64 + cy) * 64, b->y0);
                c->x1 = imin(c->x0 + 64, b->x1); c->y1 = imin((b->y0 / 64 + cy + 1) * 64, b->y1);
                c->w = c->x1 - c->x0; c->h = c->y1 - c->y0;
                c->tx = b->offx + c->x0; c->ty = b->offy + (c->y0 - b->y0);
                c->data = (uint8_t *)calloc((size_t)(74 + c->w * c->h * 4), 1);
            }
    }
    t->coef = (float *)malloc(sizeof(float) * (size_t)w * h);
    return t;
}

void orc_tile_destroy(orc_tile_t *t)
{
    if (!t) return;
    for (int i = 0; i < t->ncblks; ++i) free(t->cblks[i].data);
    free(t->cblks); free(t->coef); free(t);
}

/* ------------------------------------------------------------------------------------ */
/* forward 9/7: dwt.c:111-117 constants, :1222-1300 lifting, :1930-2108 level loop        */
static const float kAlpha = -1.586134342f;
static const float kBeta = -0.052980118f;
static const float kGamma = 0.882911075f;
static const float kDelta = 0.443506852f;
static const float kK = 1.230174105f;
static const float kInvK = (float)(1.0 / 1.230174105);

/* one lifting step over an interleaved line starting on an even sample (cas = 0).
 * tgt_first = index of first updated sample (1 for d-steps, 0 for s-steps) */
static void lift_d(float *x, int sn, int dn, float c)
{
    int m = imin(dn, sn - 1);
    for (int i = 0; i < m; ++i) x[2 * i + 1] += (x[2 * i] + x[2 * i + 2]) * c;
    if (m < dn) x[2 * m + 1] += (2 * x[2 * m]) * c;
}

static void lift_s(float *x, int sn, int dn, float c)
{
    int m = imin(sn, dn);
    if (m > 0) {
        x[0] += (x[1] + x[1]) * c;
        for (int i = 1; i < m; ++i) x[2 * i] += (x[2 * i - 1] + x[2 * i + 1]) * c;
    }
    if (m < sn) x[2 * m] += (2 * x[2 * m - 1]) * c;
}

static void fwd97_line(float *x, int n)
{
    if (n == 1) return;
    int sn = (n + 1) >> 1, dn = n - sn;
    lift_d(x, sn, dn, kAlpha);
    lift_s(x, sn, dn, kBeta);
    lift_d(x, sn, dn, kGamma);
    lift_s(x, sn, dn, kDelta);
    for (int i = 0; i < sn; ++i) x[2 * i] *= kInvK;
    for (int i = 0; i < dn; ++i) x[2 * i + 1] *= kK;
}

/* the same line when its first sample has odd parity on the canvas (cas = 1; opj_dwt_encode_1_real with a = 1, b = 0,
 * dwt.c:1251-1300, and opj_dwt_encode_and_deinterleave_v_real :1860-1924): the even-indexed samples are the dn
 * high-pass ones, the odd-indexed the sn low-pass ones, and the lifting steps swap shapes.  lift_d / lift_s are
 * "update the odd samples from their even neighbours" / "update the even samples from their odd neighbours" for a line
 * with (first argument) even-indexed and (second argument) odd-indexed samples. */
static void fwd97_line_odd(float *x, int n, int sn)
{
    if (n == 1) return;
    int dn = n - sn;
    lift_s(x, dn, sn, kAlpha);
    lift_d(x, dn, sn, kBeta);
    lift_s(x, dn, sn, kGamma);
    lift_d(x, dn, sn, kDelta);
    for (int i = 0; i < dn; ++i) x[2 * i] *= kK;
    for (int i = 0; i < sn; ++i) x[2 * i + 1] *= kInvK;
}

void orc_dwt97_forward(orc_tile_t *t, const uint16_t *img)
{
    const int w = t->w, h = t->h;
    float *a = t->coef;
    /* tcd.c:2469-2508 DC level shift: (float)(pix - 32768) */
    for (size_t i = 0; i < (size_t)w * h; ++i) a[i] = (float)((int)img[i] - 32768);
    float *line = (float *)malloc(sizeof(float) * (size_t)(w > h ? w : h));
    for (int r = ORC_NRES - 1; r >= 1; --r) {
        int rw = t->res_w[r], rh = t->res_h[r];
        int sn;
        /* vertical first */
        sn = t->res_h[r - 1];
        if (rh > 1)
            for (int x = 0; x < rw; ++x) {
                const int cas = t->cas_v[r];
                for (int y = 0; y < rh; ++y) line[y] = a[(size_t)y * w + x];
                if (cas == 0) fwd97_line(line, rh); else fwd97_line_odd(line, rh, sn);
                for (int i = 0; i < sn; ++i) a[(size_t)i * w + x] = line[2 * i + cas];
                for (int i = 0; i < rh - sn; ++i) a[(size_t)(sn + i) * w + x] = line[2 * i + 1 - cas];
            }
        /* then horizontal */
        sn = t->res_w[r - 1];
        if (rw > 1)
            for (int y = 0; y < rh; ++y) {
                float *row = a + (size_t)y * w;
                memcpy(line, row, sizeof(float) * (size_t)rw);
                fwd97_line(line, rw);
                for (int i = 0; i < sn; ++i) row[i] = line[2 * i];
                for (int i = 0; i < rw - sn; ++i) row[sn + i] = line[2 * i + 1];
            }
    }
    free(line);
}

/* ------------------------------------------------------------------------------------ */
/* MQ coder (ISO/IEC 15444-1 Annex C; mqc.c:61-157 table, :188-227 init/flush, :492-523   */
/* byteout, mqc_inl.h:210-260 encode)                                                    */
typedef struct { uint16_t qe; uint8_t nmps, nlps, sw; } mq_state_t;
static const mq_state_t mq_tab[47] = {
    {0x5601, 1, 1, 1},   {0x3401, 2, 6, 0},   {0x1801, 3, 9, 0},   {0x0AC1, 4, 12, 0},
    {0x0521, 5, 29, 0},  {0x0221, 38, 33, 0}, {0x5601, 7, 6, 1},   {0x5401, 8, 14, 0},
    {0x4801, 9, 14, 0},  {0x3801, 10, 14, 0}, {0x3001, 11, 17, 0}, {0x2401, 12, 18, 0},
    {0x1C01, 13, 20, 0}, {0x1601, 29, 21, 0}, {0x5601, 15, 14, 1}, {0x5401, 16, 14, 0},
    {0x5101, 17, 15, 0}, {0x4801, 18, 16, 0}, {0x3801, 19, 17, 0}, {0x3401, 20, 18, 0},
    {0x3001, 21, 19, 0}, {0x2801, 22, 19, 0}, {0x2401, 23, 20, 0}, {0x2201, 24, 21, 0},
    {0x1C01, 25, 22, 0}, {0x1801, 26, 23, 0}, {0x1601, 27, 24, 0}, {0x1401, 28, 25, 0},
    {0x1201, 29, 26, 0}, {0x1101, 30, 27, 0}, {0x0AC1, 31, 28, 0}, {0x09C1, 32, 29, 0},
    {0x08A1, 33, 30, 0}, {0x0521, 34, 31, 0}, {0x0441, 35, 32, 0}, {0x02A1, 36, 33, 0},
    {0x0221, 37, 34, 0}, {0x0141, 38, 35, 0}, {0x0111, 39, 36, 0}, {0x0085, 40, 37, 0},
    {0x0049, 41, 38, 0}, {0x0025, 42, 39, 0}, {0x0015, 43, 40, 0}, {0x0009, 44, 41, 0},
    {0x0005, 45, 42, 0}, {0x0001, 45, 43, 0}, {0x5601, 46, 46, 0}
};

void orc_get_mq_table(uint16_t *qe, uint8_t *nmps, uint8_t *nlps, uint8_t *sw)
{
    for (int i = 0; i < 47; ++i) { qe[i] = mq_tab[i].qe; nmps[i] = mq_tab[i].nmps; nlps[i] = mq_tab[i].nlps; sw[i] = mq_tab[i].sw; }
}

#define CTX_ZC 0
#define CTX_SC 9
#define CTX_MAG 14
#define CTX_AGG 17
#define CTX_UNI 18
#define NCTX 19

typedef struct {
    uint32_t a, c; int ct;
    uint8_t *start, *bp;       /* bp starts at start-1 (a zero guard byte precedes start) */
    uint8_t idx[NCTX], mps[NCTX];
} mq_t;

static void mq_init(mq_t *q, uint8_t *start)
{
    memset(q->idx, 0, sizeof q->idx); memset(q->mps, 0, sizeof q->mps);
    q->idx[CTX_UNI] = 46; q->idx[CTX_AGG] = 3; q->idx[CTX_ZC] = 4;
    q->a = 0x8000; q->c = 0; q->ct = 12; q->start = start; q->bp = start - 1;
}

static void mq_byteout(mq_t *q)
{
    if (*q->bp == 0xff) {
        q->bp++; *q->bp = (uint8_t)(q->c >> 20); q->c &= 0xfffff; q->ct = 7;
    } else if ((q->c & 0x8000000) == 0) {
        q->bp++; *q->bp = (uint8_t)(q->c >> 19); q->c &= 0x7ffff; q->ct = 8;
    } else {
        (*q->bp)++;
        if (*q->bp == 0xff) {
            q->c &= 0x7ffffff;
            q->bp++; *q->bp = (uint8_t)(q->c >> 20); q->c &= 0xfffff; q->ct = 7;
        } else {
            q->bp++; *q->bp = (uint8_t)(q->c >> 19); q->c &= 0x7ffff; q->ct = 8;
        }
    }
}

static void mq_renorm(mq_t *q)
{
    do {
        q->a <<= 1; q->c <<= 1; q->ct--;
        if (q->ct == 0) mq_byteout(q);
    } while ((q->a & 0x8000) == 0);
}

static void mq_encode(mq_t *q, int ctx, int d)
{
    const mq_state_t *s = &mq_tab[q->idx[ctx]];
    uint32_t qe = s->qe;
    if (q->mps[ctx] == d) {
        q->a -= qe;
        if ((q->a & 0x8000) == 0) {
            if (q->a < qe) q->a = qe; else q->c += qe;
            q->idx[ctx] = s->nmps;
            mq_renorm(q);
        } else {
            q->c += qe;
        }
    } else {
        q->a -= qe;
        if (q->a < qe) q->c += qe; else q->a = qe;
        if (s->sw) q->mps[ctx] ^= 1;
        q->idx[ctx] = s->nlps;
        mq_renorm(q);
    }
}

static void mq_flush(mq_t *q)
{
    uint32_t tempc = q->c + q->a;
    q->c |= 0xffff;
    if (q->c >= tempc) q->c -= 0x8000;
    q->c <<= q->ct; mq_byteout(q);
    q->c <<= q->ct; mq_byteout(q);
    if (*q->bp != 0xff) q->bp++;
}

static int mq_numbytes(const mq_t *q) { return (int)(q->bp - q->start); }

/* ------------------------------------------------------------------------------------ */
/* EBCOT context formation (Annex D; t1_generate_luts.c restated as functions)            */
static int zc_ctx(int orient, int h, int v, int d)
{
    if (orient == 1) { int t = h; h = v; v = t; }      /* HL band swaps roles */
    if (orient != 3) {
        if (h == 0) { if (v == 0) return d == 0 ? 0 : (d == 1 ? 1 : 2); return v == 1 ? 3 : 4; }
        if (h == 1) { if (v == 0) return d == 0 ? 5 : 6; return 7; }
        return 8;
    } else {
        int hv = h + v;
        if (d == 0) return hv == 0 ? 0 : (hv == 1 ? 1 : 2);
        if (d == 1) return hv == 0 ? 3 : (hv == 1 ? 4 : 5);
        if (d == 2) return hv == 0 ? 6 : 7;
        return 8;
    }
}

/* contribution of a neighbour pair: +1 if any significant positive, -1 if any significant negative */
static int sc_contrib(int sig_a, int neg_a, int sig_b, int neg_b)
{
    int pos = imin((sig_a && !neg_a) + (sig_b && !neg_b), 1);
    int neg = imin((sig_a && neg_a) + (sig_b && neg_b), 1);
    return pos - neg;
}

static void sc_ctx(int hc, int vc, int *ctx, int *spb)
{
    int n;
    if (hc == 0 && vc == 0) *spb = 0; else *spb = !(hc > 0 || (hc == 0 && vc > 0));
    if (hc < 0) { hc = -hc; vc = -vc; }
    if (hc == 0) n = vc == 0 ? 0 : 1;
    else n = vc == -1 ? 2 : (vc == 0 ? 3 : 4);
    *ctx = CTX_SC + n;
}

/* distortion-estimate tables (t1_generate_luts.c:259-290) */
static int16_t nm_sig[128], nm_sig0[128], nm_ref[128], nm_ref0[128];
static int nm_ready = 0;
static void nm_init(void)
{
    if (nm_ready) return;
    for (int i = 0; i < 128; ++i) {
        double t = i / 64.0, u = t, v = t - 1.5;
        nm_sig[i] = (int16_t)imax(0, (int)(floor((u * u - v * v) * 64.0 + 0.5) / 64.0 * 8192.0));
        nm_sig0[i] = (int16_t)imax(0, (int)(floor((u * u) * 64.0 + 0.5) / 64.0 * 8192.0));
        u = t - 1.0;
        v = (i & 64) ? t - 1.5 : t - 0.5;
        nm_ref[i] = (int16_t)imax(0, (int)(floor((u * u - v * v) * 64.0 + 0.5) / 64.0 * 8192.0));
        nm_ref0[i] = (int16_t)imax(0, (int)(floor((u * u) * 64.0 + 0.5) / 64.0 * 8192.0));
    }
    nm_ready = 1;
}
static int nmsedec_sig(uint32_t x, int bpno) { return bpno > 0 ? nm_sig[(x >> bpno) & 127] : nm_sig0[x & 127]; }
static int nmsedec_ref(uint32_t x, int bpno) { return bpno > 0 ? nm_ref[(x >> bpno) & 127] : nm_ref0[x & 127]; }

/* expose tables in OPJ's LUT layout so a test can diff them against t1_luts.h */
void orc_get_luts(uint8_t *zc, uint8_t *sc, uint8_t *spb, int16_t *a, int16_t *b, int16_t *c, int16_t *d)
{
    nm_init();
    for (int orient = 0; orient < 4; ++orient)
        for (int f = 0; f < 512; ++f) {
            /* bit i of f = T1_SIGMA_i: 0 NW,1 N,2 NE,3 W,4 THIS,5 E,6 SW,7 S,8 SE */
            int h = ((f >> 3) & 1) + ((f >> 5) & 1), v = ((f >> 1) & 1) + ((f >> 7) & 1);
            int dg = (f & 1) + ((f >> 2) & 1) + ((f >> 6) & 1) + ((f >> 8) & 1);
            zc[(orient << 9) | f] = (uint8_t)(CTX_ZC + zc_ctx(orient, h, v, dg));
        }
    for (int f = 0; f < 256; ++f) {
        /* bits: 0 SGN_W,1 SIG_N,2 SGN_E,3 SIG_W,4 SGN_N,5 SIG_E,6 SGN_S,7 SIG_S */
        int hc = sc_contrib((f >> 5) & 1, (f >> 2) & 1, (f >> 3) & 1, f & 1);
        int vc = sc_contrib((f >> 1) & 1, (f >> 4) & 1, (f >> 7) & 1, (f >> 6) & 1);
        int ctx, s; sc_ctx(hc, vc, &ctx, &s);
        sc[f] = (uint8_t)ctx; spb[f] = (uint8_t)s;
    }
    memcpy(a, nm_sig, sizeof nm_sig); memcpy(b, nm_sig0, sizeof nm_sig0);
    memcpy(c, nm_ref, sizeof nm_ref); memcpy(d, nm_ref0, sizeof nm_ref0);
}

/* ------------------------------------------------------------------------------------ */
/* tier-1 block coder: t1.c:2206-2449 (quantise), :2552-2741 (pass loop), :479-580 sigpass, */
/* :814-922 refpass, :1163-1256 clnpass, :1416-1449 weighted distortion                  */
#define BS 66  /* 64 + 1 border each side */
typedef struct {
    int w, h, orient;
    uint32_t mag[64 * 64];     /* |q| with 6 fractional bits */
    uint8_t neg[64 * 64];
    uint8_t sig[BS * BS], sgn[BS * BS], pi[BS * BS], mu[BS * BS];
    mq_t mq;
    int nmsedec;
} t1_t;

#define AT(x, y) (((y) + 1) * BS + (x) + 1)

static int any_neighbour(const t1_t *t, int p)
{
    const uint8_t *s = t->sig;
    return s[p - 1] | s[p + 1] | s[p - BS] | s[p + BS] | s[p - BS - 1] | s[p - BS + 1] | s[p + BS - 1] | s[p + BS + 1];
}

static int zc_of(const t1_t *t, int p)
{
    const uint8_t *s = t->sig;
    int h = s[p - 1] + s[p + 1], v = s[p - BS] + s[p + BS];
    int d = s[p - BS - 1] + s[p - BS + 1] + s[p + BS - 1] + s[p + BS + 1];
    return CTX_ZC + zc_ctx(t->orient, h, v, d);
}

static void code_sign_and_mark(t1_t *t, int x, int y, int bpno)
{
    int p = AT(x, y);
    const uint8_t *s = t->sig, *g = t->sgn;
    int hc = sc_contrib(s[p + 1], g[p + 1], s[p - 1], g[p - 1]);
    int vc = sc_contrib(s[p - BS], g[p - BS], s[p + BS], g[p + BS]);
    int ctx, spb; sc_ctx(hc, vc, &ctx, &spb);
    int neg = t->neg[y * 64 + x];
    t->nmsedec += nmsedec_sig(t->mag[y * 64 + x], bpno);
    mq_encode(&t->mq, ctx, neg ^ spb);
    t->sig[p] = 1; t->sgn[p] = (uint8_t)neg;
}

static void enc_sigpass(t1_t *t, int bpno)
{
    const uint32_t one = 1u << (bpno + 6);
    t->nmsedec = 0;
    for (int k = 0; k < t->h; k += 4)
        for (int x = 0; x < t->w; ++x)
            for (int y = k; y < imin(k + 4, t->h); ++y) {
                int p = AT(x, y);
                if (!t->sig[p] && !t->pi[p] && any_neighbour(t, p)) {
                    int v = (t->mag[y * 64 + x] & one) ? 1 : 0;
                    mq_encode(&t->mq, zc_of(t, p), v);
                    if (v) code_sign_and_mark(t, x, y, bpno);
                    t->pi[p] = 1;
                }
            }
}

static void enc_refpass(t1_t *t, int bpno)
{
    const uint32_t one = 1u << (bpno + 6);
    t->nmsedec = 0;
    for (int k = 0; k < t->h; k += 4)
        for (int x = 0; x < t->w; ++x)
            for (int y = k; y < imin(k + 4, t->h); ++y) {
                int p = AT(x, y);
                if (t->sig[p] && !t->pi[p]) {
                    int ctx = t->mu[p] ? CTX_MAG + 2 : (any_neighbour(t, p) ? CTX_MAG + 1 : CTX_MAG);
                    uint32_t a = t->mag[y * 64 + x];
                    t->nmsedec += nmsedec_ref(a, bpno);
                    mq_encode(&t->mq, ctx, (a & one) ? 1 : 0);
                    t->mu[p] = 1;
                }
            }
}

static void enc_clnpass(t1_t *t, int bpno)
{
    const uint32_t one = 1u << (bpno + 6);
    t->nmsedec = 0;
    for (int k = 0; k < t->h; k += 4) {
        int full = k + 4 <= t->h;
        int lim = imin(4, t->h - k);
        for (int x = 0; x < t->w; ++x) {
            int runlen = 0, agg = 0;
            if (full) {
                agg = 1;
                for (int j = 0; j < 4; ++j) {
                    int p = AT(x, k + j);
                    if (t->sig[p] || t->pi[p] || any_neighbour(t, p)) { agg = 0; break; }
                }
            }
            if (agg) {
                for (runlen = 0; runlen < 4; ++runlen)
                    if (t->mag[(k + runlen) * 64 + x] & one) break;
                mq_encode(&t->mq, CTX_AGG, runlen != 4);
                if (runlen == 4) continue;
                mq_encode(&t->mq, CTX_UNI, runlen >> 1);
                mq_encode(&t->mq, CTX_UNI, runlen & 1);
            }
            for (int ci = runlen; ci < lim; ++ci) {
                int y = k + ci, p = AT(x, y);
                if (agg && ci == runlen) {
                    code_sign_and_mark(t, x, y, bpno);
                } else if (!t->sig[p] && !t->pi[p]) {
                    int v = (t->mag[y * 64 + x] & one) ? 1 : 0;
                    mq_encode(&t->mq, zc_of(t, p), v);
                    if (v) code_sign_and_mark(t, x, y, bpno);
                }
                t->pi[p] = 0;
            }
        }
    }
}

static double wmsedec(int nmsedec, int level, int orient, int bpno, double stepsize)
{
    const int log2_gain = orient == 0 ? 0 : (orient == 3 ? 2 : 1);
    double w1 = 1, w2 = getnorm97(level, orient), w;
    stepsize /= (1 << log2_gain);
    w = w1 * w2 * stepsize * (1 << bpno);
    w *= w * nmsedec / 8192.0;
    return w;
}

static void t1_encode_cblk(orc_tile_t *tile, orc_cblk_t *cb, t1_t *t)
{
    const orc_band_t *b = &tile->bands[cb->band];
    nm_init();
    memset(t->sig, 0, sizeof t->sig); memset(t->sgn, 0, sizeof t->sgn);
    memset(t->pi, 0, sizeof t->pi); memset(t->mu, 0, sizeof t->mu);
    t->w = cb->w; t->h = cb->h; t->orient = b->orient;
    /* quantise: lrintf((coef / stepsize) * 64), t1.c:2397-2425 */
    int max = 0;
    for (int y = 0; y < cb->h; ++y)
        for (int x = 0; x < cb->w; ++x) {
            float c = tile->coef[(size_t)(cb->ty + y) * tile->w + cb->tx + x];
            int q = (int)lrintf((c / b->stepsize) * 64.0f);
            int a = q < 0 ? -q : q;
            t->mag[y * 64 + x] = (uint32_t)a; t->neg[y * 64 + x] = q < 0;
            if (a > max) max = a;
        }
    cb->numbps = max ? (floorlog2(max) + 1) - 6 : 0;
    cb->totalpasses = 0; cb->datalen = 0;
    memset(cb->data, 0, (size_t)(74 + cb->w * cb->h * 4));
    if (cb->numbps <= 0) return;

    /* data[0] is OPJ's fake leading byte; MQ bytes start at data+1 */
    mq_init(&t->mq, cb->data + 1);
    double cum = 0.0;
    int bpno = cb->numbps - 1, passtype = 2, passno;
    for (passno = 0; bpno >= 0; ++passno) {
        switch (passtype) {
        case 0: enc_sigpass(t, bpno); break;
        case 1: enc_refpass(t, bpno); break;
        default: enc_clnpass(t, bpno); break;
        }
        cum += wmsedec(t->nmsedec, b->level, b->orient, bpno, (double)b->stepsize);
        cb->distortiondec[passno] = cum;
        if (passtype == 2 && bpno == 0) {      /* only the very last pass terminates (cblksty 0) */
            mq_flush(&t->mq);
            cb->rate[passno] = mq_numbytes(&t->mq);
        } else {
            cb->rate[passno] = mq_numbytes(&t->mq) + 3;
        }
        if (++passtype == 3) { passtype = 0; bpno--; }
    }
    cb->totalpasses = passno;
    int last = mq_numbytes(&t->mq);
    cb->datalen = last;
    for (passno = cb->totalpasses; passno > 0;) {
        --passno;
        if (cb->rate[passno] > last) cb->rate[passno] = last; else last = cb->rate[passno];
    }
    for (passno = 0; passno < cb->totalpasses; ++passno) {
        if (cb->data[1 + cb->rate[passno] - 1] == 0xFF) cb->rate[passno]--;
        cb->len[passno] = cb->rate[passno] - (passno == 0 ? 0 : cb->rate[passno - 1]);
    }
}

void orc_t1_encode_all(orc_tile_t *t)
{
    t1_t *t1 = (t1_t *)malloc(sizeof(t1_t));
    for (int i = 0; i < t->ncblks; ++i) t1_encode_cblk(t, &t->cblks[i], t1);
    free(t1);
}

/* ------------------------------------------------------------------------------------ */
/* tier-2: bit writer with 0xFF stuffing (bio.c:75-84,149-211), tag trees (tgt.c:48-305),  */
/* packet headers + bodies (t2.c:662-1006), LRCP with one precinct per resolution          */
typedef struct { uint8_t *start, *bp, *end; uint32_t buf; int ct; int dry; } bio_t;

static void bio_init(bio_t *b, uint8_t *p, size_t len, int dry)
{ b->start = b->bp = p; b->end = p + len; b->buf = 0; b->ct = 8; b->dry = dry; }

static int bio_byteout(bio_t *b)
{
    b->buf = (b->buf << 8) & 0xffff;
    b->ct = b->buf == 0xff00 ? 7 : 8;
    if (b->bp >= b->end) return 0;
    if (!b->dry) *b->bp = (uint8_t)(b->buf >> 8);
    b->bp++;
    return 1;
}
static void bio_putbit(bio_t *b, uint32_t v)
{
    if (b->ct == 0) bio_byteout(b);
    b->ct--;
    b->buf |= v << b->ct;
}
static void bio_write(bio_t *b, uint32_t v, int n) { for (int i = n - 1; i >= 0; --i) bio_putbit(b, (v >> i) & 1); }
static int bio_flush(bio_t *b)
{
    if (!bio_byteout(b)) return 0;
    if (b->ct == 7 && !bio_byteout(b)) return 0;
    return 1;
}

typedef struct { int value, low, known, parent; } tg_node_t;
typedef struct { int nnodes; tg_node_t *n; } tgt_t;

static tgt_t *tgt_create(int lw, int lh)
{
    int nplh[32], nplv[32], nl = 0, total = 0, n;
    nplh[0] = lw; nplv[0] = lh;
    do {
        n = nplh[nl] * nplv[nl];
        nplh[nl + 1] = (nplh[nl] + 1) / 2; nplv[nl + 1] = (nplv[nl] + 1) / 2;
        total += n; ++nl;
    } while (n > 1);
    tgt_t *t = (tgt_t *)malloc(sizeof(*t));
    t->nnodes = total; t->n = (tg_node_t *)calloc((size_t)total, sizeof(tg_node_t));
    int base = 0;
    for (int l = 0; l < nl; ++l) {
        int cnt = nplh[l] * nplv[l], nbase = base + cnt;
        for (int y = 0; y < nplv[l]; ++y)
            for (int x = 0; x < nplh[l]; ++x)
                t->n[base + y * nplh[l] + x].parent = l == nl - 1 ? -1 : nbase + (y / 2) * nplh[l + 1] + x / 2;
        base = nbase;
    }
    for (int i = 0; i < total; ++i) { t->n[i].value = 999; t->n[i].low = 0; t->n[i].known = 0; }
    return t;
}
static void tgt_free(tgt_t *t) { if (t) { free(t->n); free(t); } }
static void tgt_setvalue(tgt_t *t, int leaf, int value)
{
    int i = leaf;
    while (i >= 0 && t->n[i].value > value) { t->n[i].value = value; i = t->n[i].parent; }
}
static void tgt_encode(bio_t *bio, tgt_t *t, int leaf, int threshold)
{
    int stk[32], sp = 0, i = leaf, low = 0;
    while (t->n[i].parent >= 0) { stk[sp++] = i; i = t->n[i].parent; }
    for (;;) {
        tg_node_t *nd = &t->n[i];
        if (low > nd->low) nd->low = low; else low = nd->low;
        while (low < threshold) {
            if (low >= nd->value) {
                if (!nd->known) { bio_putbit(bio, 1); nd->known = 1; }
                break;
            }
            bio_putbit(bio, 0); ++low;
        }
        nd->low = low;
        if (sp == 0) break;
        i = stk[--sp];
    }
}

static void put_numpasses(bio_t *b, int n)
{
    if (n == 1) bio_putbit(b, 0);
    else if (n == 2) bio_write(b, 2, 2);
    else if (n <= 5) bio_write(b, 0xc | (uint32_t)(n - 3), 4);
    else if (n <= 36) bio_write(b, 0x1e0 | (uint32_t)(n - 6), 9);
    else if (n <= 164) bio_write(b, 0xff80 | (uint32_t)(n - 37), 16);
}

/* returns bytes produced, or (size_t)-1 if the packets do not fit in cap */
size_t orc_t2_encode(orc_tile_t *t, uint8_t *dst, size_t cap, int dry)
{
    uint8_t *c = dst;
    size_t length = cap;
    for (int r = 0; r < ORC_NRES; ++r) {
        bio_t bio;
        bio_init(&bio, c, length, dry);
        bio_putbit(&bio, 1);                      /* non-empty packet */
        for (int bi = 0; bi < t->nbands; ++bi) {
            orc_band_t *b = &t->bands[bi];
            if (b->resno != r || b->cw * b->ch == 0) continue;
            int nblk = b->cw * b->ch;
            tgt_t *incl = tgt_create(b->cw, b->ch), *imsb = tgt_create(b->cw, b->ch);
            for (int k = 0; k < nblk; ++k) tgt_setvalue(imsb, k, b->numbps - t->cblks[b->first_cblk + k].numbps);
            for (int k = 0; k < nblk; ++k) if (t->cblks[b->first_cblk + k].numpasses) tgt_setvalue(incl, k, 0);
            for (int k = 0; k < nblk; ++k) {
                orc_cblk_t *cb = &t->cblks[b->first_cblk + k];
                tgt_encode(&bio, incl, k, 1);
                if (!cb->numpasses) continue;
                tgt_encode(&bio, imsb, k, 999);
                put_numpasses(&bio, cb->numpasses);
                /* single codeword segment: only the last pass of the block terminates */
                int len = cb->rate[cb->numpasses - 1];
                int numlenbits = 3;
                int increment = imax(0, floorlog2(len) + 1 - (numlenbits + floorlog2(cb->numpasses)));
                for (int i = 0; i < increment; ++i) bio_putbit(&bio, 1);
                bio_putbit(&bio, 0);
                numlenbits += increment;
                bio_write(&bio, (uint32_t)len, numlenbits + floorlog2(cb->numpasses));
            }
            tgt_free(incl); tgt_free(imsb);
        }
        if (!bio_flush(&bio)) return (size_t)-1;
        size_t nb = (size_t)(bio.bp - bio.start);
        c += nb; length -= nb;
        for (int bi = 0; bi < t->nbands; ++bi) {
            orc_band_t *b = &t->bands[bi];
            if (b->resno != r) continue;
            for (int k = 0; k < b->cw * b->ch; ++k) {
                orc_cblk_t *cb = &t->cblks[b->first_cblk + k];
                if (!cb->numpasses) continue;
                size_t len = (size_t)cb->rate[cb->numpasses - 1];
                if (len > length) return (size_t)-1;
                if (!dry) memcpy(c, cb->data + 1, len);
                c += len; length -= len;
            }
        }
    }
    return (size_t)(c - dst);
}

/* ------------------------------------------------------------------------------------ */
/* PCRD: tcd.c:248-353 (makelayer), :465-712 (rateallocate); budget j2k.c:5320-5430       */
static int makelayer(orc_tile_t *t, double thresh)
{
    int same = 1;
    for (int i = 0; i < t->ncblks; ++i) {
        orc_cblk_t *cb = &t->cblks[i];
        int n = 0;
        if (thresh < 0) {
            n = cb->totalpasses;
        } else {
            for (int p = 0; p < cb->totalpasses; ++p) {
                uint32_t dr; double dd;
                if (n == 0) { dr = (uint32_t)cb->rate[p]; dd = cb->distortiondec[p]; }
                else { dr = (uint32_t)(cb->rate[p] - cb->rate[n - 1]); dd = cb->distortiondec[p] - cb->distortiondec[n - 1]; }
                if (!dr) { if (dd != 0) n = p + 1; continue; }
                if (thresh - (dd / dr) < DBL_EPSILON) n = p + 1;
            }
        }
        if (cb->prev_numpasses != n) { same = 0; cb->prev_numpasses = n; }
        cb->numpasses = n;
    }
    return same;
}

/* per-tile byte budget exactly as opj_j2k_update_rates computes it in f32 */
static float tile_budget(int h, int w, float base_cr, int main_header_bytes, int ntiles)
{
    float rate = base_cr / 2;                       /* src/ebcc_codec.c:116 */
    if (rate <= 1.0) return 0.0f;                   /* j2k.c:8159-8162 force lossless */
    float r = (float)(((double)16 * (uint32_t)w * (uint32_t)h) / (rate * (float)8)) - 0.0f;
    r -= (float)main_header_bytes / (float)(uint32_t)ntiles;   /* l_sot_remove = stream position / (th * tw), j2k.c:5351 */
    if (r < 30.0f) r = 30.0f;
    return r;
}

double orc_rate_allocate(orc_tile_t *t, float base_cr, int main_header_bytes)
{
    return orc_rate_allocate_tiled(t, base_cr, main_header_bytes, 1);
}

/* opj_tcd_rateallocate runs per tile; opj_j2k_update_rates gives every tile the budget of its own area minus its
 * share of the main header (j2k.c:5361-5408) */
double orc_rate_allocate_tiled(orc_tile_t *t, float base_cr, int main_header_bytes, int ntiles)
{
    double min = DBL_MAX, max = 0;
    for (int i = 0; i < t->ncblks; ++i) {
        orc_cblk_t *cb = &t->cblks[i];
        cb->prev_numpasses = 0; cb->numpasses = 0;
        for (int p = 0; p < cb->totalpasses; ++p) {
            int dr; double dd;
            if (p == 0) { dr = cb->rate[p]; dd = cb->distortiondec[p]; }
            else { dr = cb->rate[p] - cb->rate[p - 1]; dd = cb->distortiondec[p] - cb->distortiondec[p - 1]; }
            if (dr == 0) continue;
            double s = dd / dr;
            if (s < min) min = s;
            if (s > max) max = s;
        }
    }
    float rates0 = tile_budget(t->h, t->w, base_cr, main_header_bytes, ntiles);
    /* tile buffer: j2k.c:5439-5470 (1.4*bits/8 + 500 + specific headers), minus SOT(12), SOD(2)+EOC room(2) */
    uint64_t tile_size = (uint64_t)((double)((uint64_t)t->w * t->h * 16) * 1.4 / 8) + 500;
    uint32_t len = (uint32_t)tile_size - 12 - 4;
    double goodthresh;
    if (rates0 > 0.0f) {
        uint32_t maxlen = (uint32_t)ceil(rates0); if (maxlen > len) maxlen = len;
        double lo = min, hi = max, thresh = 0, stable = 0;
        int last_ok = 0;
        for (int i = 0; i < 128; ++i) {
            double nt = (lo + hi) / 2;
            if (fabs(nt - thresh) <= 0.5 * 1e-5 * thresh) break;
            thresh = nt;
            int same = makelayer(t, thresh) && i != 0;
            int fail;
            if (same) fail = !last_ok;
            else fail = orc_t2_encode(t, NULL, maxlen, 1) == (size_t)-1;
            if (fail) { last_ok = 0; lo = thresh; continue; }
            last_ok = 1; hi = thresh; stable = thresh;
        }
        goodthresh = stable == 0 ? thresh : stable;
    } else {
        goodthresh = -1;
    }
    makelayer(t, goodthresh);
    return goodthresh;
}

/* ------------------------------------------------------------------------------------ */
/* codestream markers: j2k.c:13080-13133 (SOC SIZ COD QCD COM), :4355 SOT, :4835 SOD, :5194 EOC */
static uint8_t *put16(uint8_t *p, uint32_t v) { p[0] = (uint8_t)(v >> 8); p[1] = (uint8_t)v; return p + 2; }
static uint8_t *put32(uint8_t *p, uint32_t v) { p = put16(p, v >> 16); return put16(p, v & 0xffff); }

size_t orc_main_header(int h, int w, uint8_t *dst) { return orc_main_header_tiled(h, h, w, dst); }

/* SIZ carries the image size (Xsiz, Ysiz) and the tile size (XTsiz, YTsiz); everything else is the single-tile header */
size_t orc_main_header_tiled(int himg, int h, int w, uint8_t *dst)
{
    uint8_t *p = dst;
    orc_tile_t *t = orc_tile_create(h, w);
    p = put16(p, 0xFF4F);
    p = put16(p, 0xFF51); p = put16(p, 41); p = put16(p, 0);
    p = put32(p, (uint32_t)w); p = put32(p, (uint32_t)himg); p = put32(p, 0); p = put32(p, 0);
    p = put32(p, (uint32_t)w); p = put32(p, (uint32_t)h); p = put32(p, 0); p = put32(p, 0);
    p = put16(p, 1); *p++ = 15; *p++ = 1; *p++ = 1;
    p = put16(p, 0xFF52); p = put16(p, 12); *p++ = 0; *p++ = 0; p = put16(p, 1); *p++ = 0;
    *p++ = 5; *p++ = 4; *p++ = 4; *p++ = 0; *p++ = 0;
    p = put16(p, 0xFF5C); p = put16(p, 3 + 2 * t->nbands); *p++ = (2 << 5) | 2;
    for (int i = 0; i < t->nbands; ++i) p = put16(p, (uint32_t)((t->bands[i].expn << 11) | t->bands[i].mant));
    static const char com[] = "Created by OpenJPEG version 2.5.3";
    p = put16(p, 0xFF64); p = put16(p, (uint32_t)(4 + sizeof(com) - 1)); p = put16(p, 1);
    memcpy(p, com, sizeof(com) - 1); p += sizeof(com) - 1;
    orc_tile_destroy(t);
    return (size_t)(p - dst);
}

size_t orc_j2k_encode(const uint16_t *img, int h, int w, float base_cr, uint8_t **out)
{
    orc_tile_t *t = orc_tile_create(h, w);
    orc_dwt97_forward(t, img);
    orc_t1_encode_all(t);
    size_t cap = (size_t)((double)((uint64_t)w * h * 16) * 1.4 / 8) + 1024;
    uint8_t *buf = (uint8_t *)malloc(cap);
    size_t hdr = orc_main_header(h, w, buf);
    orc_rate_allocate(t, base_cr, (int)hdr);
    uint8_t *sot = buf + hdr;
    put16(sot, 0xFF90); put16(sot + 2, 10); put16(sot + 4, 0); sot[10] = 0; sot[11] = 1;
    put16(sot + 12, 0xFF93);
    size_t body = orc_t2_encode(t, sot + 14, cap - hdr - 14 - 2, 0);
    if (body == (size_t)-1) body = 0;
    put32(sot + 6, (uint32_t)(body + 14));
    put16(sot + 14 + body, 0xFFD9);
    *out = buf;
    orc_tile_destroy(t);
    return hdr + 14 + body + 2;
}

/* opj_write_tile per tile then opj_end_compress (src/ebcc_codec.c:166-176): each tile-part is SOT (Isot = tile index,
 * Psot, TPsot 0, TNsot 1) SOD packets, coded with the geometry of the tile's canvas position and its own byte budget */
size_t orc_j2k_encode_tiled(const uint16_t *img, int ntiles, int h, int w, float base_cr, uint8_t **out)
{
    if (ntiles <= 1) return orc_j2k_encode(img, h, w, base_cr, out);
    const size_t tile_cap = (size_t)((double)((uint64_t)w * h * 16) * 1.4 / 8) + 1024;
    uint8_t *buf = (uint8_t *)malloc(256 + tile_cap * (size_t)ntiles);
    size_t pos = orc_main_header_tiled(ntiles * h, h, w, buf);
    const int hdr = (int)pos;
    for (int ti = 0; ti < ntiles; ++ti) {
        orc_tile_t *t = orc_tile_create_at(h, w, ti * h);
        orc_dwt97_forward(t, img + (size_t)ti * h * w);
        orc_t1_encode_all(t);
        orc_rate_allocate_tiled(t, base_cr, hdr, ntiles);
        uint8_t *sot = buf + pos;
        put16(sot, 0xFF90); put16(sot + 2, 10); put16(sot + 4, (uint32_t)ti); sot[10] = 0; sot[11] = 1;
        put16(sot + 12, 0xFF93);
        size_t body = orc_t2_encode(t, sot + 14, tile_cap - 14 - 2, 0);
        if (body == (size_t)-1) body = 0;
        put32(sot + 6, (uint32_t)(body + 14));
        pos += 14 + body;
        orc_tile_destroy(t);
    }
    put16(buf + pos, 0xFFD9);
    *out = buf;
    return pos + 2;
}
