// ebcc_cuda.cu -- device layer of the EBCC hot path for NVIDIA B200 (sm_100a) behind the C ABI of
// include/ebcc_cuda.h.  One context = one GPU, one stream, one set of work buffers sized for a
// batch of frames; a batch flows  H2D -> min/max -> (scale + DC shift + DWT) -> tier-1 context
// modelling -> MQ coding -> PCRD + tier-2 -> compaction -> D2H  without host round trips.
#include "common.cuh"
#include "k_dwt.cuh"
#include "k_t1.cuh"
#include "k_t2.cuh"
#include "k_dec.cuh"
#include "k_idwt.cuh"
#include "k_idwt_fused.cuh"
#include "k_spiht.cuh"
#include "k_zstd.cuh"
#include "../../include/ebcc_cuda.h"
#include <math.h>
#include <stdio.h>
#include <time.h>
#include <stdlib.h>
#include <string.h>

#define CK(call)                                                                                     \
    do {                                                                                             \
        cudaError_t e_ = (call);                                                                     \
        if (e_ != cudaSuccess) {                                                                     \
            fprintf(stderr, "[ebcc-cuda] %s failed at %s:%d: %s\n", #call, __FILE__, __LINE__,       \
                    cudaGetErrorString(e_));                                                         \
            (void)cudaGetLastError();   /* a failed call must not poison the next check */            \
            return -1;                                                                               \
        }                                                                                            \
    } while (0)

// ---- per-kernel-class device times (bench.py's roofline; off by default) -------------------------------------------
// When switched on (ebcc_cuda_kernel_timing(1)) every launch group below is bracketed by two CUDA events on the stream it
// is launched on; the pairs are read back at the pipeline's next synchronisation point and added up process-wide, so
// the figures describe the kernels in the configuration that is actually running (all pipelines in flight).
#include <pthread.h>
static int g_kt_on = 0;
static pthread_mutex_t g_kt_mu = PTHREAD_MUTEX_INITIALIZER;
static double g_kt_ms[EBCC_KT_COUNT];
static unsigned long long g_kt_n[EBCC_KT_COUNT];
#define KT_RING 192
#define UPLOAD_SLOTS 4
#define UPLOAD_SLOT_BYTES ((size_t)8 << 20)

// ---- compaction of the per-frame codestreams into one contiguous buffer --------------------------
__global__ void k_offsets(const ProbeOut *__restrict__ pout, unsigned long long *offs, int nframes)
{
    // single thread: nframes is small (<= a few thousand) and this is 8 bytes per frame
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        unsigned long long o = 0;
        for (int i = 0; i < nframes; ++i) { offs[i] = o; o += (unsigned long long)((pout[i].stream_bytes + 15) & ~15); }
        offs[nframes] = o;
    }
}

__global__ void k_compact(const unsigned char *__restrict__ out_pool, const ProbeOut *__restrict__ pout,
                          const unsigned long long *__restrict__ offs, unsigned char *packed, size_t out_stride)
{
    const int frame = blockIdx.y;
    const int n16 = (pout[frame].stream_bytes + 15) >> 4;
    const uint4 *src = (const uint4 *)(out_pool + (size_t)frame * out_stride);
    uint4 *dst = (uint4 *)(packed + offs[frame]);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += gridDim.x * blockDim.x) dst[i] = src[i];
}

// ---- multi-tile codestream assembly ---------------------------------------------------------------------------------
// Every tile context writes "main header | SOT(Isot = t) SOD packets | EOC" for its tile; the image's codestream is
// the main header, the tile-parts in tile order and one EOC (opj_j2k_write_tile per tile + opj_end_compress,
// src/ebcc_codec.c:166-176).  grid = (ntiles, nframes).
#define EBCC_MAX_TILES 64
struct TileSrc { const unsigned char *out; const ProbeOut *pout; size_t stride; };

__global__ void k_assemble_tiles(const TileSrc *__restrict__ src, int ntiles, int hl, unsigned char *out, size_t out_stride, ProbeOut *pout)
{
    const int t = blockIdx.x, frame = blockIdx.y;
    const int sb = src[t].pout[frame].stream_bytes;
    if (sb == 0) {                                   // constant image: no codestream
        if (t == 0 && threadIdx.x == 0) { pout[frame].stream_bytes = 0; pout[frame].total_bytes = 0; }
        return;
    }
    int off = hl;
    for (int k = 0; k < t; ++k) off += src[k].pout[frame].stream_bytes - hl - 2;
    const int len = sb - hl - 2;
    const unsigned char *s = src[t].out + (size_t)frame * src[t].stride;
    unsigned char *o = out + (size_t)frame * out_stride;
    for (int i = threadIdx.x; i < len; i += blockDim.x) o[off + i] = s[hl + i];
    if (t == 0) for (int i = threadIdx.x; i < hl; i += blockDim.x) o[i] = s[i];
    if (t == ntiles - 1 && threadIdx.x == 0) {
        o[off + len] = 0xff; o[off + len + 1] = 0xd9;
        pout[frame].stream_bytes = off + len + 2;
        pout[frame].total_bytes = off + len - hl;
        pout[frame].iters = 0; pout[frame].thresh = 0.0;
    }
}

// ---- host-side geometry (tcd.c:961-1264) ---------------------------------------------------------------------------
static int h_ceildivpow2(int a, int b) { return (int)(((long long)a + (1ll << b) - 1) >> b); }
static int h_floorlog2(int a) { int l = 0; while (a > 1) { a >>= 1; l++; } return l; }

static const double h_norms97[4][10] = {
    {1.000, 1.965, 4.177, 8.403, 16.90, 33.84, 67.69, 135.3, 270.6, 540.9},
    {2.022, 3.989, 8.355, 17.04, 34.27, 68.63, 137.3, 274.6, 549.0, 0},
    {2.022, 3.989, 8.355, 17.04, 34.27, 68.63, 137.3, 274.6, 549.0, 0},
    {2.080, 3.865, 8.307, 17.18, 34.71, 69.59, 139.3, 278.6, 557.2, 0}};

static void put16(unsigned char *&p, unsigned v) { *p++ = (unsigned char)(v >> 8); *p++ = (unsigned char)v; }
static void put32(unsigned char *&p, unsigned v) { put16(p, v >> 16); put16(p, v & 0xffff); }

// Geometry of one tile of H x W whose first line is canvas row y0 (tile tile_index of ntiles; the image is Himg lines
// high).  Tiles below the first one start at a non-zero canvas row: resolutions and bands are the canvas rectangles of
// tcd.c:961-1055 (ceil-divisions of y0 and y0 + H), the row parity of a resolution's origin decides which lines are
// low-pass (dwt.c:1986-1987), and the code-block grid stays anchored to multiples of 64 of the band's canvas
// coordinates (tcd.c:1122-1230), so the first block row of a band can be a partial one.
static void build_plan(int H, int W, int y0, int tile_index, int ntiles, int Himg, Plan *pl, CblkDesc **cblks_out)
{
    memset(pl, 0, sizeof(*pl));
    pl->H = H; pl->W = W; pl->y0 = y0; pl->tile_index = tile_index; pl->ntiles = ntiles;
    for (int r = 0; r < J2K_NRES; ++r) {
        const int lvl = J2K_NRES - 1 - r, ry0 = h_ceildivpow2(y0, lvl);
        pl->res_w[r] = h_ceildivpow2(W, lvl);
        pl->res_h[r] = h_ceildivpow2(y0 + H, lvl) - ry0;
        pl->cas_v[r] = ry0 & 1;
    }
    int nb = 0, ncb = 0, nnodes = 0;
    int band_y0[J2K_MAXBANDS];                       // canvas row of each band's first line
    for (int r = 0; r < J2K_NRES; ++r) {
        const int lvl = J2K_NRES - 1 - r;
        pl->res_first[r] = ncb;
        for (int bi = 0; bi < (r == 0 ? 1 : 3); ++bi) {
            BandDesc *b = &pl->bands[nb];
            const int orient = r == 0 ? 0 : bi + 1;
            b->resno = r; b->orient = orient; b->level = lvl;
            if (r == 0) {
                b->bw = h_ceildivpow2(W, lvl); b->bh = pl->res_h[0];
                band_y0[nb] = h_ceildivpow2(y0, lvl);
            } else {
                const int x0b = orient & 1, y0b = orient >> 1;
                b->bw = h_ceildivpow2(W - (x0b << lvl), lvl + 1);
                band_y0[nb] = h_ceildivpow2(y0 - (y0b << lvl), lvl + 1);
                b->bh = h_ceildivpow2(y0 + H - (y0b << lvl), lvl + 1) - band_y0[nb];
                b->offx = x0b ? pl->res_w[r - 1] : 0;
                b->offy = y0b ? pl->res_h[r - 1] : 0;
            }
            // explicit quantisation step (dwt.c:2176-2198, 1302-1311), band parameters (tcd.c:1069-1089)
            int lv = lvl;
            if (orient == 0 && lv >= 10) lv = 9; else if (orient > 0 && lv >= 9) lv = 8;
            const double norm = h_norms97[orient][lv];
            const int si = (int)floor((1.0 / norm) * 8192.0);
            const int p = h_floorlog2(si) - 13, n = 11 - h_floorlog2(si);
            b->mant = (n < 0 ? si >> -n : si << n) & 0x7ff;
            b->expn = 16 - p;
            const int log2_gain = orient == 0 ? 0 : (orient == 3 ? 2 : 1);
            b->stepsize = (float)((1.0 + b->mant / 2048.0) * pow(2.0, (16 + log2_gain) - b->expn));
            b->stepsize_dec = (float)((1.0 + b->mant / 2048.0) * pow(2.0, 16 - b->expn));
            b->numbps = b->expn + 2 - 1;
            {
                double st = (double)b->stepsize;
                st /= (1 << log2_gain);
                b->wnorm = 1.0 * norm * st;
            }
            const bool empty = b->bw <= 0 || b->bh <= 0;
            b->cw = empty ? 0 : (b->bw + 63) / 64;
            b->ch = empty ? 0 : (band_y0[nb] + b->bh + 63) / 64 - band_y0[nb] / 64;
            b->first_cblk = ncb;
            ncb += b->cw * b->ch;
            // tag tree shape
            b->tt_base = nnodes;
            int lw = b->cw, lh = b->ch, nl = 0, off = 0;
            if (!empty) {
                int nn;
                do {
                    b->tt_w[nl] = lw; b->tt_h[nl] = lh; b->tt_off[nl] = off;
                    nn = lw * lh; off += nn;
                    lw = (lw + 1) / 2; lh = (lh + 1) / 2; ++nl;
                } while (nn > 1);
            }
            b->tt_nlev = nl; nnodes += off;
            nb++;
        }
        pl->res_count[r] = ncb - pl->res_first[r];
    }
    pl->nbands = nb; pl->ncblks = ncb; pl->nnodes = nnodes;
    CblkDesc *cb = (CblkDesc *)calloc((size_t)ncb, sizeof(CblkDesc));
    size_t sym_off = 0, mq_off = 0;
    for (int bi = 0; bi < nb; ++bi) {
        const BandDesc *b = &pl->bands[bi];
        for (int gy = 0; gy < b->ch; ++gy)
            for (int gx = 0; gx < b->cw; ++gx) {
                CblkDesc *c = &cb[b->first_cblk + gy * b->cw + gx];
                c->band = bi; c->gx = gx; c->gy = gy; c->orient = b->orient;
                c->w = (gx * 64 + 64 <= b->bw ? 64 : b->bw - gx * 64);
                // rows of the band relative to its first line: grid row gy of a grid anchored at canvas multiples of 64
                const int by0 = band_y0[bi], gr0 = (by0 / 64 + gy) * 64 - by0;
                const int r0 = gr0 < 0 ? 0 : gr0, r1 = gr0 + 64 < b->bh ? gr0 + 64 : b->bh;
                c->h = r1 - r0;
                c->tx = b->offx + gx * 64; c->ty = b->offy + r0;
                // symbols: <= 1.5 per sample per bit-plane (zero coding / refinement / run mode) + one sign
                size_t cap = (size_t)c->w * c->h * (3 * (size_t)b->numbps / 2 + 2) + 32;
                cap = (cap + 15) & ~(size_t)15;
                c->sym_off = (unsigned)sym_off; c->sym_cap = (unsigned)cap; sym_off += cap;
                size_t mcap = ((size_t)80 + (size_t)c->w * c->h * 4 + 15) & ~(size_t)15;
                c->mq_off = (unsigned)mq_off; c->mq_cap = (unsigned)mcap; mq_off += mcap;
            }
    }
    pl->sym_stride = (sym_off + 255) & ~(size_t)255;
    pl->mq_stride = (mq_off + 255) & ~(size_t)255;
    // main header: SOC SIZ COD QCD COM (j2k.c:13080-13133); 135 bytes for one component
    unsigned char *p = pl->main_header;
    put16(p, 0xFF4F);
    put16(p, 0xFF51); put16(p, 41); put16(p, 0);
    put32(p, (unsigned)W); put32(p, (unsigned)Himg); put32(p, 0); put32(p, 0);
    put32(p, (unsigned)W); put32(p, (unsigned)H); put32(p, 0); put32(p, 0);
    put16(p, 1); *p++ = 15; *p++ = 1; *p++ = 1;
    put16(p, 0xFF52); put16(p, 12); *p++ = 0; *p++ = 0; put16(p, 1); *p++ = 0;
    *p++ = 5; *p++ = 4; *p++ = 4; *p++ = 0; *p++ = 0;
    put16(p, 0xFF5C); put16(p, 3 + 2 * (unsigned)nb); *p++ = (2 << 5) | 2;
    for (int i = 0; i < nb; ++i) put16(p, (unsigned)((pl->bands[i].expn << 11) | pl->bands[i].mant));
    static const char com[] = "Created by OpenJPEG version 2.5.3";
    put16(p, 0xFF64); put16(p, (unsigned)(4 + sizeof(com) - 1)); put16(p, 1);
    memcpy(p, com, sizeof(com) - 1); p += sizeof(com) - 1;
    pl->main_header_len = (int)(p - pl->main_header);
    // room for every tile-part of a multi-tile image when this plan is the image-level one (Himg == H, ntiles > 1)
    size_t os = (size_t)((double)((unsigned long long)W * H * 16) * 1.4 / 8) + 500 + 256 + (size_t)pl->main_header_len + 520 * (size_t)ntiles;
    pl->out_stride = (os + 255) & ~(size_t)255;
    *cblks_out = cb;
}

// per-tile byte budget in f32 exactly as opj_j2k_update_rates (j2k.c:5376-5408); < 0 => all passes
static int tile_maxlen(int H, int W, float base_cr, int main_header_bytes, int ntiles)
{
    float rate = base_cr / 2;                            // src/ebcc_codec.c:116
    if (rate <= 1.0) return -1;                          // j2k.c:8159-8162
    float r = (float)(((double)16 * (unsigned)W * (unsigned)H) / (rate * (float)8)) - 0.0f;
    r -= (float)main_header_bytes / (float)(unsigned)ntiles;     // l_sot_remove, j2k.c:5351-5352
    if (r < 30.0f) r = 30.0f;
    unsigned long long tile_size = (unsigned long long)((double)((unsigned long long)W * H * 16) * 1.4 / 8) + 500;
    unsigned len = (unsigned)tile_size - 12 - 4;
    unsigned ml = (unsigned)ceil(r);
    return (int)(ml < len ? ml : len);
}

static void fill_luts(T1Luts *l, MqTable *t)
{
    for (int i = 0; i < 128; ++i) {
        double tt = i / 64.0, u = tt, v = tt - 1.5;
        int a = (int)(floor((u * u - v * v) * 64.0 + 0.5) / 64.0 * 8192.0);
        int b = (int)(floor((u * u) * 64.0 + 0.5) / 64.0 * 8192.0);
        l->sig[i] = (short)(a > 0 ? a : 0); l->sig0[i] = (short)(b > 0 ? b : 0);
        u = tt - 1.0; v = (i & 64) ? tt - 1.5 : tt - 0.5;
        a = (int)(floor((u * u - v * v) * 64.0 + 0.5) / 64.0 * 8192.0);
        b = (int)(floor((u * u) * 64.0 + 0.5) / 64.0 * 8192.0);
        l->ref[i] = (short)(a > 0 ? a : 0); l->ref0[i] = (short)(b > 0 ? b : 0);
    }
    // MQ probability state machine, ISO/IEC 15444-1 table C.2: Qe, NMPS, NLPS, SWITCH
    static const unsigned short qe[47] = {
        0x5601, 0x3401, 0x1801, 0x0AC1, 0x0521, 0x0221, 0x5601, 0x5401, 0x4801, 0x3801, 0x3001, 0x2401,
        0x1C01, 0x1601, 0x5601, 0x5401, 0x5101, 0x4801, 0x3801, 0x3401, 0x3001, 0x2801, 0x2401, 0x2201,
        0x1C01, 0x1801, 0x1601, 0x1401, 0x1201, 0x1101, 0x0AC1, 0x09C1, 0x08A1, 0x0521, 0x0441, 0x02A1,
        0x0221, 0x0141, 0x0111, 0x0085, 0x0049, 0x0025, 0x0015, 0x0009, 0x0005, 0x0001, 0x5601};
    static const unsigned char nmps[47] = {1, 2, 3, 4, 5, 38, 7, 8, 9, 10, 11, 12, 13, 29, 15, 16, 17, 18, 19, 20, 21, 22, 23, 24,
                                           25, 26, 27, 28, 29, 30, 31, 32, 33, 34, 35, 36, 37, 38, 39, 40, 41, 42, 43, 44, 45, 45, 46};
    static const unsigned char nlps[47] = {1, 6, 9, 12, 29, 33, 6, 14, 14, 14, 17, 18, 20, 21, 14, 14, 15, 16, 17, 18, 19, 19, 20, 21,
                                           22, 23, 24, 25, 26, 27, 28, 29, 30, 31, 32, 33, 34, 35, 36, 37, 38, 39, 40, 41, 42, 43, 46};
    for (int i = 0; i < 47; ++i) {
        unsigned sw = (i == 0 || i == 6 || i == 14) ? 1u : 0u;
        t->e[i] = qe[i] | ((unsigned)nmps[i] << 16) | ((unsigned)nlps[i] << 22) | (sw << 28);
    }
}

static void fill_dec_luts(DecLuts *d, const MqTable *t)
{
    for (int o = 0; o < 4; ++o)
        for (int nb = 0; nb < 512; ++nb) {
            const int h = ((nb >> 3) & 1) + ((nb >> 5) & 1), v = ((nb >> 1) & 1) + ((nb >> 7) & 1);
            const int dg = (nb & 1) + ((nb >> 2) & 1) + ((nb >> 6) & 1) + ((nb >> 8) & 1);
            d->zc[o * 512 + nb] = (unsigned char)t1_zc_ctx(o, h, v, dg);
        }
    for (int st = 0; st < 47; ++st)
        for (int mps = 0; mps < 2; ++mps) {
            const unsigned e = t->e[st], qe = e & 0xffffu, nm = (e >> 16) & 63u, nl = (e >> 22) & 63u, sw = (e >> 28) & 1u;
            d->mq[st * 2 + mps] = qe | ((nm * 2 + (unsigned)mps) << 16) | ((nl * 2 + ((unsigned)mps ^ sw)) << 23) | ((unsigned)mps << 30);
        }
}

// sum of the errors of a frame as a double (fixed-point accumulator of the kernels, see ERR_FX_BITS)
static double es_sum(const ErrStat &e) { return ldexp((double)e.sum_fx, e.sum_exp - ERR_FX_BITS); }

// ---- context ---------------------------------------------------------------------------------------
struct ebcc_cuda_ctx {
    int device;
    cudaStream_t stream;
    Plan plan;                 // host copy
    CblkDesc *h_cblks;
    Plan *d_plan;
    CblkDesc *d_cblks;
    int *d_order;              // code-block indices by descending area: long decode chains start first
    BandDesc *d_bands;
    T1Luts *d_luts;
    MqTable *d_mqtab;
    size_t cap_frames;
    float *d_in, *d_coef;
    FrameStat *d_stat;
    unsigned char *d_sym, *d_mq, *d_out, *d_packed;
    CblkOut *d_cbout;
    int *d_maxlen, *d_body_off;
    ProbeOut *d_pout;
    PcrdCache *d_pcache;       // [B] threshold -> packet bytes evaluations of the frame's probes since the analysis (k_pcrd_t2)
    unsigned long long *d_offs;
    // decoder side
    DecLuts *d_declut;
    unsigned char *d_streams;  // [B][out_stride] codestreams to decode
    int *d_slens, *d_status;
    DecJob *d_jobs;            // [B][ncblks]
    uint2 *d_rjobs;            // [B][ncblks] compact per-block records of a probe (k_jobs_from_enc -> k_idwt_tile)
    int *d_iplane;             // [B][H][W] tier-1 decoder output (integer, Mallat layout); the recording of the CR search
    int *d_scratch;            // [B][H][W] per-probe real decodes of blocks the recording cannot serve
    unsigned char *d_spass;    // [B][H][W] pass in which each sample became significant (recording)
    DecState *d_dstate;        // [B][ncblks]
    T1CutEntry *d_cut_list;    // unproven truncation points of the batch and their pre-decoded blocks (k_t1_dec_cuts)
    int *d_cut_count, *d_cut_pool, cut_cap;
    unsigned char *d_ctxsnap;  // [B * ncblks][T1_SNAP_PLANES][T1_SNAP_BYTES] context states at clean-up pass ends (k_t1_mq -> k_t1_dec_cuts)
    int *h_cut_count;          // pinned copy of the batch's number of unproven cuts
    int cut_known, cut_overflow;   // read back after the analysis: when every unproven cut sits in the pool no probe launches the tier-1 decoder
    unsigned *d_decsave;       // [B][ncblks][DEC_SAVE_WORDS] decoder state of each block's recording (resumed by later probes)
    int use_recording;         // 0: every probe decodes every block for real (EBCC_NO_RECORDING=1)
    int rec_from_encoder;      // recording + cut verdicts come from the encoder (k_t1_cm / k_t1_mq); 0: from k_t1_dec (EBCC_DECODER_RECORDING=1)
    int for_search;            // the running analysis feeds a compression-ratio search
    float *d_dcoef;            // [B][H][W] dequantised coefficients / IDWT work plane
    char *h_up[4]; cudaEvent_t ev_up[4];   // pinned staging ring for uploads from pageable memory (ebcc_cuda_upload)
    float *d_dcoef2;           // [B][H][W] second work plane: the fused IDWT levels ping-pong between the two
    int skip_dec;              // this probe's decoded frames are not needed (search probe: only the statistics are)
    unsigned long long probe_cap;   // search probe: violation counts are only needed up to this value (0: exact), see k_idwt_tile
    float *d_dec;              // [B][H][W] decoded frames
    ErrStat *d_es;
    float *d_targets;
    ErrStat *h_es;
    int *h_status, *h_slens;
    float *h_targets;
    unsigned char *h_streams;  // pinned staging for codestream upload
    // residual layer (SPIHT + zstd)
    SpGeom sg;
    SpBufs sp;
    float *d_sp_img0, *d_sp_img1;   // [B][Np] data / temp planes of the SPIHT wavelet transform
    SpInfo *d_sp_info, *h_sp_info;
    unsigned *d_sp_P, *h_sp_P, *d_sp_budget, *h_sp_budget;
    int *d_sp_active, *h_sp_active, *d_step_tab;
    float *d_sp_rmm, *h_sp_rmm;
    ErrStat *d_es2, *h_es2;
    unsigned char *d_zout, *h_zout; // [B][zstride] zstd frames
    size_t zstride;
    unsigned *d_zlen, *h_zlen, *d_zsrc_len, *h_zsrc_len;
    unsigned char *d_zin, *h_zin;   // [B][zstride] zstd frames to decode (decode path)
    unsigned char *d_zblk;
    unsigned *d_zblen;
    int zmaxchunks;
    ZDecWork *d_zdec_work;
    unsigned *d_zcap, *h_zcap;
    // pinned host mirrors
    ProbeOut *h_pout;
    FrameStat *h_stat;
    unsigned long long *h_offs;
    int *h_maxlen;
    unsigned char *h_packed;   // pinned destination of the current download; two buffers take turns (h_packed_alt), so the
    size_t h_packed_cap;       // streams of the previous download stay valid while the next one lands
    unsigned char *h_packed_alt;
    size_t h_packed_alt_cap;
    float *h_stage;            // pinned staging for pageable inputs
    size_t h_stage_cap;
    int profiling;
    cudaEvent_t ev[EBCC_ST_COUNT + 1];
    cudaEvent_t ev_sync;       // blocking-sync event: a host thread waiting for its pipeline sleeps instead of spinning
    int spin_sync;
    float stage_ms[EBCC_ST_COUNT];
    int ev_valid;
    unsigned long long launches;
    // multi-tile images (dims[0] > 1, src/ebcc_codec.c:105-180): this context holds the image-level data (input,
    // min/max, decoded image, residual layer, assembled codestreams); tile t of every image of the batch is coded by
    // tile[t], a light context with the tile's own geometry that shares this context's stream and look-up tables
    cudaEvent_t kt_ev[2 * KT_RING];   // kernel-class timing: event pairs recorded since the last synchronisation
    int kt_id[KT_RING], kt_n, kt_have;
    // The long serial kernels (MQ coder, decode of the unproven cuts) run on a second stream of the LOWEST priority,
    // everything else on `stream`, which has the highest: with a dozen pipelines in flight the shared memory of every SM is
    // full of resident MQ CTAs, and without priorities the short kernels of the other pipelines (9/7 passes, context
    // modelling, rate allocation, probes) queue behind MQ CTAs that are still waiting for a place themselves.
    cudaStream_t stream_lo;
    cudaEvent_t ev_fork, ev_join;
    int prio_split;
    int crowd;                 // pipelines at work on this device (hint from the host codec), see mq_lanes()
    int env_mq_lanes, env_dec_lanes, env_no_recording, env_dec_recording, env_debug_lanes, env_t1rec_ctas, env_dwt_cw;   // tuning / debugging variables, read once per host call (read_env)
    int env_fused_idwt, env_fused_dwt, env_sp_serial, env_no_pcrd_cache, env_cuts_from_zero, env_all_unproven;
    int ntiles;
    ebcc_cuda_ctx *tile[EBCC_MAX_TILES];
    ebcc_cuda_ctx *parent;
    TileSrc *d_tile_src;
};

static void read_env(ebcc_cuda_ctx *c);
static ebcc_cuda_ctx *ctx_root_mut(ebcc_cuda_ctx *c) { return c->parent ? c->parent : c; }

// NVTX: every launch group is also an NVTX range named after its kernel class ("ebcc/dwt", "ebcc/t1_cm", ...), so that a
// timeline (Nsight Systems, ncu --nvtx) shows the stages of the hot path; header-only NVTX 3 costs a null-pointer test when
// no tool is attached.  Switch off with -DEBCC_NO_NVTX.
#if !defined(EBCC_CUEMU) && !defined(EBCC_NO_NVTX)
#include <nvtx3/nvToolsExt.h>
static const char *const k_kt_nvtx[EBCC_KT_COUNT] = {
    "ebcc/minmax", "ebcc/dwt", "ebcc/t1_cm", "ebcc/t1_mq", "ebcc/pcrd_t2", "ebcc/bodies", "ebcc/t1_dec", "ebcc/recon", "ebcc/idwt", "ebcc/finalize",
    "ebcc/sp_fwd", "ebcc/sp_encode", "ebcc/sp_probe", "ebcc/zstd_enc", "ebcc/t2_parse", "ebcc/zstd_dec", "ebcc/sp_decode", "ebcc/sp_inv"};
#define EBCC_NVTX_PUSH(id) nvtxRangePushA(k_kt_nvtx[id])
#define EBCC_NVTX_POP() nvtxRangePop()
#else
#define EBCC_NVTX_PUSH(id) ((void)0)
#define EBCC_NVTX_POP() ((void)0)
#endif

// scope of one launch group of kernel class `id`
struct KT {
    ebcc_cuda_ctx *r; int slot; cudaStream_t st;
    KT(ebcc_cuda_ctx *c, int id, cudaStream_t on = NULL) : r(NULL), slot(-1), st(on)
    {
        EBCC_NVTX_PUSH(id);
        if (!g_kt_on) return;
        r = ctx_root_mut(c);
        if (!st) st = r->stream;
        if (!r->kt_have) { for (int i = 0; i < 2 * KT_RING; ++i) cudaEventCreate(&r->kt_ev[i]); r->kt_have = 1; }
        if (r->kt_n >= KT_RING) { r = NULL; return; }
        slot = r->kt_n++; r->kt_id[slot] = id;
        cudaEventRecord(r->kt_ev[2 * slot], st);
    }
    ~KT() { if (r) cudaEventRecord(r->kt_ev[2 * slot + 1], st); EBCC_NVTX_POP(); }
};

// after a synchronisation of the pipeline's stream: every recorded pair has completed
static void kt_harvest(ebcc_cuda_ctx *c)
{
    ebcc_cuda_ctx *r = ctx_root_mut(c);
    if (!r->kt_n) return;
    pthread_mutex_lock(&g_kt_mu);
    for (int i = 0; i < r->kt_n; ++i) {
        float ms = 0;
        if (cudaEventElapsedTime(&ms, r->kt_ev[2 * i], r->kt_ev[2 * i + 1]) == cudaSuccess) { g_kt_ms[r->kt_id[i]] += ms; g_kt_n[r->kt_id[i]]++; }
        else (void)cudaGetLastError();
    }
    pthread_mutex_unlock(&g_kt_mu);
    r->kt_n = 0;
}

extern "C" void ebcc_cuda_kernel_timing(int enable)
{
    pthread_mutex_lock(&g_kt_mu);
    for (int i = 0; i < EBCC_KT_COUNT; ++i) { g_kt_ms[i] = 0; g_kt_n[i] = 0; }
    g_kt_on = enable;
    pthread_mutex_unlock(&g_kt_mu);
}

extern "C" int ebcc_cuda_kernel_timing_report(double *ms, unsigned long long *groups, int n)
{
    pthread_mutex_lock(&g_kt_mu);
    for (int i = 0; i < n && i < EBCC_KT_COUNT; ++i) { ms[i] = g_kt_ms[i]; groups[i] = g_kt_n[i]; }
    pthread_mutex_unlock(&g_kt_mu);
    return EBCC_KT_COUNT;
}

// Wait for the pipeline's stream.  Several pipelines (and, on a multi-GPU box, several ranks) share the host cores, and an
// error-bounded batch waits some forty times (one wait per probe).  EBCC_SPIN_SYNC=1: spin (lowest latency, one core per
// waiting thread); =0: sleep on a blocking-sync event (no CPU at all, but the wake-up costs most of a millisecond under load:
// one 768-frame call 10.0 instead of 12.4 GB/s); unset: poll the event, spinning briefly and then sleeping 30 us between
// looks -- close to the spin wait's latency at a few per cent of a core, whoever else runs on the box.
static cudaError_t ctx_sync(ebcc_cuda_ctx *c)
{
    cudaError_t e;
    if (c->spin_sync == 1) e = cudaStreamSynchronize(c->stream);
    else {
        e = cudaEventRecord(c->ev_sync, c->stream);
        if (e == cudaSuccess && c->spin_sync == 0) e = cudaEventSynchronize(c->ev_sync);
        else if (e == cudaSuccess) {
            int looks = 0;
            while ((e = cudaEventQuery(c->ev_sync)) == cudaErrorNotReady)
                if (++looks > 64) { struct timespec ts = {0, 30000}; nanosleep(&ts, NULL); }
            const cudaError_t le = cudaGetLastError();          // the query's "not ready" must not linger as the last error
            if (e == cudaSuccess && le != cudaSuccess && le != cudaErrorNotReady) e = le;
        }
    }
    if (e == cudaSuccess) kt_harvest(c);
    return e;
}


extern "C" int ebcc_cuda_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}

extern "C" ebcc_cuda_ctx *ebcc_cuda_ctx_create(int device)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0 || device >= n) {
        fprintf(stderr, "[ebcc-cuda] no CUDA device available (requested %d of %d): the EBCC B200 codec has no CPU path\n", device, n);
        return NULL;
    }
    if (cudaSetDevice(device) != cudaSuccess) return NULL;
    ebcc_cuda_ctx *c = (ebcc_cuda_ctx *)calloc(1, sizeof(*c));
    c->device = device; c->ntiles = 1;
    {
        int least = 0, greatest = 0;
        cudaDeviceGetStreamPriorityRange(&least, &greatest);
        const char *ps = getenv("EBCC_PRIO_SPLIT");
        c->prio_split = !(ps && ps[0] == '0') && least != greatest;
        if (cudaStreamCreateWithPriority(&c->stream, cudaStreamNonBlocking, c->prio_split ? greatest : 0) != cudaSuccess) { free(c); return NULL; }
        if (c->prio_split) {
            if (cudaStreamCreateWithPriority(&c->stream_lo, cudaStreamNonBlocking, least) != cudaSuccess) { cudaStreamDestroy(c->stream); free(c); return NULL; }
            cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming);
            cudaEventCreateWithFlags(&c->ev_join, cudaEventDisableTiming);
        }
    }
    for (int i = 0; i <= EBCC_ST_COUNT; ++i) cudaEventCreate(&c->ev[i]);
    cudaEventCreateWithFlags(&c->ev_sync, cudaEventBlockingSync | cudaEventDisableTiming);
    { const char *sp = getenv("EBCC_SPIN_SYNC"); c->spin_sync = !sp ? 2 : (sp[0] == '1' ? 1 : 0); }      // 2: poll (ctx_sync)
    read_env(c);
    T1Luts luts; MqTable tab;
    fill_luts(&luts, &tab);
    if (cudaMalloc((void **)&c->d_luts, sizeof(luts)) != cudaSuccess || cudaMalloc((void **)&c->d_mqtab, sizeof(tab)) != cudaSuccess) { free(c); return NULL; }
    cudaMemcpy(c->d_luts, &luts, sizeof(luts), cudaMemcpyHostToDevice);
    cudaMemcpy(c->d_mqtab, &tab, sizeof(tab), cudaMemcpyHostToDevice);
    DecLuts *dl = (DecLuts *)malloc(sizeof(DecLuts));
    fill_dec_luts(dl, &tab);
    if (cudaMalloc((void **)&c->d_declut, sizeof(DecLuts)) != cudaSuccess) { free(dl); free(c); return NULL; }
    cudaMemcpy(c->d_declut, dl, sizeof(DecLuts), cudaMemcpyHostToDevice);
    free(dl);
    {
        // SPIHT top bit-plane for max = 2^k exactly as the reference's libm evaluates floor(log(max) / log(2.0))
        // (spiht_re.c:60); configuration arithmetic, done once on the host with the same libm call
        int tab[32];
        for (int k = 0; k < 32; ++k) tab[k] = (int)floor(log(ldexp(1.0, k)) / log(2.0));
        if (cudaMalloc((void **)&c->d_step_tab, sizeof(tab)) != cudaSuccess) { free(c); return NULL; }
        cudaMemcpy(c->d_step_tab, tab, sizeof(tab), cudaMemcpyHostToDevice);
    }
    return c;
}

// light context of one tile: shares the parent's stream and tables, owns its geometry and work buffers
static ebcc_cuda_ctx *ctx_create_tile(ebcc_cuda_ctx *parent)
{
    ebcc_cuda_ctx *c = (ebcc_cuda_ctx *)calloc(1, sizeof(*c));
    if (!c) return NULL;
    c->device = parent->device; c->stream = parent->stream; c->parent = parent; c->spin_sync = parent->spin_sync;
    c->d_luts = parent->d_luts; c->d_mqtab = parent->d_mqtab; c->d_declut = parent->d_declut; c->d_step_tab = parent->d_step_tab;
    for (int i = 0; i <= EBCC_ST_COUNT; ++i) cudaEventCreate(&c->ev[i]);
    cudaEventCreateWithFlags(&c->ev_sync, cudaEventBlockingSync | cudaEventDisableTiming);
    return c;
}

static void free_buffers(ebcc_cuda_ctx *c);
static void drop_tiles(ebcc_cuda_ctx *c)
{
    for (int t = 0; t < EBCC_MAX_TILES; ++t) {
        ebcc_cuda_ctx *s = c->tile[t];
        if (!s) continue;
        free_buffers(s);
        for (int i = 0; i <= EBCC_ST_COUNT; ++i) cudaEventDestroy(s->ev[i]);
        cudaEventDestroy(s->ev_sync);
        free(s);
        c->tile[t] = NULL;
    }
    cudaFree(c->d_tile_src); c->d_tile_src = NULL;
    c->ntiles = 1;
}

static void free_buffers(ebcc_cuda_ctx *c)
{
    cudaFree(c->d_plan); cudaFree(c->d_cblks); cudaFree(c->d_order); c->d_order = NULL; cudaFree(c->d_in); cudaFree(c->d_coef); cudaFree(c->d_stat);
    cudaFree(c->d_sym); cudaFree(c->d_mq); cudaFree(c->d_out); cudaFree(c->d_packed); cudaFree(c->d_cbout);
    cudaFree(c->d_maxlen); cudaFree(c->d_body_off); cudaFree(c->d_pout); cudaFree(c->d_pcache); c->d_pcache = NULL; cudaFree(c->d_offs);
    cudaFreeHost(c->h_pout); cudaFreeHost(c->h_stat); cudaFreeHost(c->h_offs); cudaFreeHost(c->h_maxlen);
    cudaFreeHost(c->h_packed); cudaFreeHost(c->h_packed_alt); cudaFreeHost(c->h_stage);
    cudaFree(c->d_streams); cudaFree(c->d_slens); cudaFree(c->d_status); cudaFree(c->d_jobs); cudaFree(c->d_rjobs); c->d_rjobs = NULL; cudaFree(c->d_iplane); cudaFree(c->d_scratch); cudaFree(c->d_spass); cudaFree(c->d_dstate); cudaFree(c->d_decsave);
    cudaFree(c->d_cut_list); cudaFree(c->d_cut_count); cudaFree(c->d_cut_pool); cudaFree(c->d_ctxsnap); c->d_ctxsnap = NULL; c->d_cut_list = NULL; c->d_cut_count = NULL; c->d_cut_pool = NULL;
    cudaFreeHost(c->h_cut_count); c->h_cut_count = NULL;
    c->d_scratch = NULL; c->d_spass = NULL; c->d_dstate = NULL; c->d_decsave = NULL;
    cudaFree(c->d_dcoef); cudaFree(c->d_dcoef2); c->d_dcoef2 = NULL; cudaFree(c->d_dec); cudaFree(c->d_es); cudaFree(c->d_targets);
    cudaFreeHost(c->h_es); cudaFreeHost(c->h_status); cudaFreeHost(c->h_slens); cudaFreeHost(c->h_targets); cudaFreeHost(c->h_streams);
    c->d_streams = NULL; c->d_slens = c->d_status = NULL; c->d_jobs = NULL; c->d_iplane = NULL; c->d_dcoef = c->d_dec = NULL;
    c->d_es = NULL; c->d_targets = NULL; c->h_es = NULL; c->h_status = c->h_slens = NULL; c->h_targets = NULL; c->h_streams = NULL;
    cudaFree(c->sp.coef); cudaFree(c->sp.D); cudaFree(c->sp.L); cudaFree(c->sp.lip0); cudaFree(c->sp.lip1); cudaFree(c->sp.lsp);
    cudaFree(c->sp.lis0); cudaFree(c->sp.lis1); cudaFree(c->sp.sign_pos); cudaFree(c->sp.lsp_idx); cudaFree(c->sp.bits);
    cudaFree(c->d_sp_img0); cudaFree(c->d_sp_img1); cudaFree(c->d_sp_info); cudaFree(c->d_sp_P); cudaFree(c->d_sp_budget);
    cudaFree(c->d_sp_active); cudaFree(c->d_sp_rmm); cudaFree(c->d_es2); cudaFree(c->d_zout); cudaFree(c->d_zlen); cudaFree(c->d_zsrc_len);
    cudaFree(c->d_zin); cudaFree(c->d_zblk); cudaFree(c->d_zblen); c->d_zblk = NULL; c->d_zblen = NULL; cudaFree(c->d_zdec_work); cudaFree(c->d_zcap); cudaFreeHost(c->h_zcap);
    c->d_zdec_work = NULL; c->d_zcap = NULL; c->h_zcap = NULL;
    cudaFreeHost(c->h_sp_info); cudaFreeHost(c->h_sp_P); cudaFreeHost(c->h_sp_budget); cudaFreeHost(c->h_sp_active);
    cudaFreeHost(c->h_sp_rmm); cudaFreeHost(c->h_es2); cudaFreeHost(c->h_zout); cudaFreeHost(c->h_zlen); cudaFreeHost(c->h_zsrc_len);
    cudaFreeHost(c->h_zin);
    memset(&c->sp, 0, sizeof(c->sp));
    c->d_sp_img0 = c->d_sp_img1 = NULL; c->d_sp_info = NULL; c->d_sp_P = c->d_sp_budget = NULL; c->d_sp_active = NULL; c->d_sp_rmm = NULL;
    c->d_es2 = NULL; c->d_zout = NULL; c->d_zlen = c->d_zsrc_len = NULL; c->d_zin = NULL;
    c->h_sp_info = NULL; c->h_sp_P = c->h_sp_budget = NULL; c->h_sp_active = NULL; c->h_sp_rmm = NULL; c->h_es2 = NULL; c->h_zout = NULL;
    c->h_zlen = c->h_zsrc_len = NULL; c->h_zin = NULL;
    free(c->h_cblks);
    c->d_plan = NULL; c->d_cblks = NULL; c->d_in = c->d_coef = NULL; c->d_stat = NULL;
    c->d_sym = c->d_mq = c->d_out = c->d_packed = NULL; c->d_cbout = NULL; c->d_maxlen = c->d_body_off = NULL;
    c->d_pout = NULL; c->d_offs = NULL; c->h_pout = NULL; c->h_stat = NULL; c->h_offs = NULL; c->h_maxlen = NULL;
    c->h_packed = NULL; c->h_packed_cap = 0; c->h_packed_alt = NULL; c->h_packed_alt_cap = 0; c->h_stage = NULL; c->h_stage_cap = 0; c->h_cblks = NULL;
    c->cap_frames = 0;
}

extern "C" void ebcc_cuda_ctx_destroy(ebcc_cuda_ctx *c)
{
    if (!c) return;
    cudaSetDevice(c->device);
    ctx_sync(c);
    drop_tiles(c);
    free_buffers(c);
    cudaFree(c->d_luts); cudaFree(c->d_mqtab); cudaFree(c->d_declut); cudaFree(c->d_step_tab);
    for (int i = 0; i <= EBCC_ST_COUNT; ++i) cudaEventDestroy(c->ev[i]);
    if (c->kt_have) for (int i = 0; i < 2 * KT_RING; ++i) cudaEventDestroy(c->kt_ev[i]);
    for (int k = 0; k < UPLOAD_SLOTS; ++k) if (c->h_up[k]) { cudaFreeHost(c->h_up[k]); cudaEventDestroy(c->ev_up[k]); }
    if (c->prio_split) { cudaStreamDestroy(c->stream_lo); cudaEventDestroy(c->ev_fork); cudaEventDestroy(c->ev_join); }
    cudaStreamDestroy(c->stream);
    free(c);
}

// work buffers of one context for nframes tiles of H x W at canvas row y0 (tile_index of ntiles, image height Himg);
// light: no residual-layer buffers (tile contexts of a multi-tile image)
static int reserve_alloc(ebcc_cuda_ctx *c, size_t H, size_t W, size_t nframes, int y0, int tile_index, int ntiles, int Himg, int light);
static int reserve_impl(ebcc_cuda_ctx *c, size_t H, size_t W, size_t nframes, int y0, int tile_index, int ntiles, int Himg, int light)
{
    const int rc = reserve_alloc(c, H, W, nframes, y0, tile_index, ntiles, Himg, light);
    if (rc && c) {            // out of device or pinned memory half way: give everything back, the caller may retry with fewer frames
        cudaSetDevice(c->device);
        free_buffers(c);
        (void)cudaGetLastError();
    }
    return rc;
}

static int reserve_alloc(ebcc_cuda_ctx *c, size_t H, size_t W, size_t nframes, int y0, int tile_index, int ntiles, int Himg, int light)
{
    if (!c || H < 32 || W < 32 || H > 2047 || W > 2047 || nframes == 0) return -1;
    CK(cudaSetDevice(c->device));
    if (c->cap_frames >= nframes && c->plan.H == (int)H && c->plan.W == (int)W && c->plan.y0 == y0 && c->plan.tile_index == tile_index &&
        c->plan.ntiles == ntiles) return 0;
    CK(ctx_sync(c));
    free_buffers(c);
    build_plan((int)H, (int)W, y0, tile_index, ntiles, Himg, &c->plan, &c->h_cblks);
    const Plan *pl = &c->plan;
    const size_t B = nframes, N = H * W;
    CK(cudaMalloc((void **)&c->d_plan, sizeof(Plan)));
    CK(cudaMalloc((void **)&c->d_cblks, sizeof(CblkDesc) * pl->ncblks));
    CK(cudaMemcpy(c->d_plan, pl, sizeof(Plan), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(c->d_cblks, c->h_cblks, sizeof(CblkDesc) * pl->ncblks, cudaMemcpyHostToDevice));
    {
        // longest-processing-time-first launch order for the one-block-per-warp decoder: when the grid does not fit
        // the GPU in one wave, the blocks left for the tail are the small ones
        int *ord = (int *)malloc(sizeof(int) * pl->ncblks);
        for (int i = 0; i < pl->ncblks; ++i) ord[i] = i;
        for (int i = 1; i < pl->ncblks; ++i) {              // insertion sort, stable, by area descending
            const int v = ord[i], a = c->h_cblks[v].w * c->h_cblks[v].h;
            int j = i - 1;
            while (j >= 0 && c->h_cblks[ord[j]].w * c->h_cblks[ord[j]].h < a) { ord[j + 1] = ord[j]; --j; }
            ord[j + 1] = v;
        }
        CK(cudaMalloc((void **)&c->d_order, sizeof(int) * pl->ncblks));
        CK(cudaMemcpy(c->d_order, ord, sizeof(int) * pl->ncblks, cudaMemcpyHostToDevice));
        free(ord);
    }
    c->d_bands = (BandDesc *)((char *)c->d_plan + offsetof(Plan, bands));
    CK(cudaMalloc((void **)&c->d_in, B * N * sizeof(float)));
    CK(cudaMalloc((void **)&c->d_coef, B * N * sizeof(float)));
    CK(cudaMalloc((void **)&c->d_stat, B * sizeof(FrameStat)));
    CK(cudaMalloc((void **)&c->d_sym, B * pl->sym_stride));
    CK(cudaMalloc((void **)&c->d_mq, B * pl->mq_stride));
    CK(cudaMalloc((void **)&c->d_out, B * pl->out_stride));
    CK(cudaMalloc((void **)&c->d_packed, B * pl->out_stride));
    CK(cudaMalloc((void **)&c->d_cbout, B * pl->ncblks * sizeof(CblkOut)));
    CK(cudaMalloc((void **)&c->d_maxlen, B * sizeof(int)));
    CK(cudaMalloc((void **)&c->d_body_off, B * pl->ncblks * sizeof(int)));
    CK(cudaMalloc((void **)&c->d_pout, B * sizeof(ProbeOut)));
    CK(cudaMalloc((void **)&c->d_pcache, B * sizeof(PcrdCache)));
    CK(cudaMalloc((void **)&c->d_offs, (B + 1) * sizeof(unsigned long long)));
    CK(cudaMallocHost((void **)&c->h_pout, B * sizeof(ProbeOut)));
    CK(cudaMallocHost((void **)&c->h_stat, B * sizeof(FrameStat)));
    CK(cudaMallocHost((void **)&c->h_offs, (B + 1) * sizeof(unsigned long long)));
    CK(cudaMallocHost((void **)&c->h_maxlen, B * sizeof(int)));
    CK(cudaMalloc((void **)&c->d_streams, B * pl->out_stride));
    CK(cudaMalloc((void **)&c->d_slens, B * sizeof(int)));
    CK(cudaMalloc((void **)&c->d_status, B * sizeof(int)));
    CK(cudaMalloc((void **)&c->d_jobs, B * pl->ncblks * sizeof(DecJob)));
    CK(cudaMalloc((void **)&c->d_rjobs, B * pl->ncblks * sizeof(uint2)));
    CK(cudaMalloc((void **)&c->d_iplane, B * N * sizeof(int)));
    CK(cudaMalloc((void **)&c->d_scratch, B * N * sizeof(int)));
    CK(cudaMalloc((void **)&c->d_spass, B * N));
    CK(cudaMalloc((void **)&c->d_dstate, B * pl->ncblks * sizeof(DecState)));
    // about 13 unproven cuts per 721 x 1440 frame (1 in 1000 of ~13 400 truncation points): room for several times that
    c->cut_cap = (int)(B * (size_t)(pl->ncblks / 3 + 8) + 256);
    CK(cudaMalloc((void **)&c->d_cut_list, (size_t)c->cut_cap * sizeof(T1CutEntry)));
    CK(cudaMalloc((void **)&c->d_cut_count, sizeof(int)));
    CK(cudaMallocHost((void **)&c->h_cut_count, sizeof(int)));
    CK(cudaMalloc((void **)&c->d_cut_pool, (size_t)c->cut_cap * T1_CUT_WORDS * sizeof(int)));
    CK(cudaMalloc((void **)&c->d_ctxsnap, B * (size_t)pl->ncblks * T1_SNAP_PLANES * T1_SNAP_BYTES));
    CK(cudaMalloc((void **)&c->d_decsave, B * pl->ncblks * (size_t)DEC_SAVE_WORDS * sizeof(unsigned)));
    CK(cudaMalloc((void **)&c->d_dcoef, B * N * sizeof(float)));
    CK(cudaMalloc((void **)&c->d_dcoef2, B * N * sizeof(float)));
    CK(cudaMalloc((void **)&c->d_dec, B * N * sizeof(float)));
    CK(cudaMalloc((void **)&c->d_es, B * sizeof(ErrStat)));
    CK(cudaMalloc((void **)&c->d_targets, B * sizeof(float)));
    CK(cudaMallocHost((void **)&c->h_es, B * sizeof(ErrStat)));
    CK(cudaMallocHost((void **)&c->h_status, B * sizeof(int)));
    CK(cudaMallocHost((void **)&c->h_slens, B * sizeof(int)));
    CK(cudaMallocHost((void **)&c->h_targets, B * sizeof(float)));
    CK(cudaMallocHost((void **)&c->h_streams, B * pl->out_stride));
    if (!light) {   // residual layer
        SpGeom &g = c->sg;
        g.W = (int)W; g.H = (int)H; g.Wp = ((int)W + 15) & ~15; g.Hp = ((int)H + 15) & ~15; g.lx = g.Wp >> SP_STAGES; g.ly = g.Hp >> SP_STAGES;
        const size_t Np = (size_t)g.Wp * g.Hp;
        c->sp.bits_words = pl->out_stride / 4 + 64;
        c->zstride = pl->out_stride + 1024;
        CK(cudaMalloc((void **)&c->sp.coef, B * Np * 4)); CK(cudaMalloc((void **)&c->sp.D, B * Np * 4)); CK(cudaMalloc((void **)&c->sp.L, B * Np * 4));
        CK(cudaMalloc((void **)&c->sp.lip0, B * Np * 4)); CK(cudaMalloc((void **)&c->sp.lip1, B * Np * 4)); CK(cudaMalloc((void **)&c->sp.lsp, B * Np * 4));
        CK(cudaMalloc((void **)&c->sp.lis0, B * Np * 8)); CK(cudaMalloc((void **)&c->sp.lis1, B * Np * 8));
        CK(cudaMalloc((void **)&c->sp.sign_pos, B * Np * 4)); CK(cudaMalloc((void **)&c->sp.lsp_idx, B * Np * 4));
        CK(cudaMalloc((void **)&c->sp.bits, B * c->sp.bits_words * 4));
        CK(cudaMalloc((void **)&c->d_sp_img0, B * Np * 4)); CK(cudaMalloc((void **)&c->d_sp_img1, B * Np * 4));
        CK(cudaMalloc((void **)&c->d_sp_info, B * sizeof(SpInfo))); CK(cudaMalloc((void **)&c->d_sp_P, B * 4)); CK(cudaMalloc((void **)&c->d_sp_budget, B * 4));
        CK(cudaMalloc((void **)&c->d_sp_active, B * 4)); CK(cudaMalloc((void **)&c->d_sp_rmm, B * 8)); CK(cudaMalloc((void **)&c->d_es2, B * sizeof(ErrStat)));
        CK(cudaMalloc((void **)&c->d_zout, B * c->zstride)); CK(cudaMalloc((void **)&c->d_zlen, B * 4)); CK(cudaMalloc((void **)&c->d_zsrc_len, B * 4));
        CK(cudaMalloc((void **)&c->d_zin, B * c->zstride));
        c->zmaxchunks = (int)((c->sp.bits_words * 4 + Z_CHUNK - 1) / Z_CHUNK);
        CK(cudaFuncSetAttribute(k_zstd_compress_chunks, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(ZEncWork)));
        CK(cudaMalloc((void **)&c->d_zblk, B * c->zmaxchunks * (size_t)Z_BLK_STRIDE)); CK(cudaMalloc((void **)&c->d_zblen, B * c->zmaxchunks * 4)); CK(cudaMalloc((void **)&c->d_zdec_work, B * sizeof(ZDecWork)));
        CK(cudaMalloc((void **)&c->d_zcap, B * 4)); CK(cudaMallocHost((void **)&c->h_zcap, B * 4));
        CK(cudaMallocHost((void **)&c->h_sp_info, B * sizeof(SpInfo))); CK(cudaMallocHost((void **)&c->h_sp_P, B * 4)); CK(cudaMallocHost((void **)&c->h_sp_budget, B * 4));
        CK(cudaMallocHost((void **)&c->h_sp_active, B * 4)); CK(cudaMallocHost((void **)&c->h_sp_rmm, B * 8)); CK(cudaMallocHost((void **)&c->h_es2, B * sizeof(ErrStat)));
        CK(cudaMallocHost((void **)&c->h_zout, B * c->zstride)); CK(cudaMallocHost((void **)&c->h_zlen, B * 4)); CK(cudaMallocHost((void **)&c->h_zsrc_len, B * 4));
        CK(cudaMallocHost((void **)&c->h_zin, B * c->zstride));
        CK(cudaFuncSetAttribute(k_sp_dwt_row, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        CK(cudaFuncSetAttribute(k_sp_dwt_col, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        CK(cudaFuncSetAttribute(k_sp_idwt_row, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        CK(cudaFuncSetAttribute(k_sp_idwt_col, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    }
    c->cap_frames = nframes;
    // large dynamic shared memory opt-ins
    CK(cudaFuncSetAttribute(k_t1_dec<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)DEC_SMEM_BYTES(1)));
    CK(cudaFuncSetAttribute(k_t1_dec<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)DEC_SMEM_BYTES(2)));
    CK(cudaFuncSetAttribute(k_t1_dec<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)DEC_SMEM_BYTES(4)));
    CK(cudaFuncSetAttribute(k_t1_dec<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)DEC_SMEM_BYTES(8)));
    // four CTAs of the tier-1 decoder per SM need the largest shared-memory carve-out
    CK(cudaFuncSetAttribute(k_t1_dec<1>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    CK(cudaFuncSetAttribute(k_t1_dec<2>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    CK(cudaFuncSetAttribute(k_t1_dec<4>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    CK(cudaFuncSetAttribute(k_t1_dec<8>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    CK(cudaFuncSetAttribute(k_idwt_v, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    CK(cudaFuncSetAttribute(k_idwt_h, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    CK(cudaFuncSetAttribute(k_t2_parse, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    CK(cudaFuncSetAttribute(k_dwt_v<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    CK(cudaFuncSetAttribute(k_dwt_v<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    CK(cudaFuncSetAttribute(k_dwt_h, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    CK(cudaFuncSetAttribute(k_pcrd_t2, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    return 0;
}

extern "C" int ebcc_cuda_reserve_tiled(ebcc_cuda_ctx *c, size_t ntiles, size_t H, size_t W, size_t nframes)
{
    if (!c || c->parent || ntiles == 0 || ntiles > EBCC_MAX_TILES || ntiles * H > 2047) return -1;
    CK(cudaSetDevice(c->device));
    if (ntiles == 1) {
        if (c->ntiles > 1) { CK(ctx_sync(c)); drop_tiles(c); }
        c->ntiles = 1;
        return reserve_impl(c, H, W, nframes, 0, 0, 1, (int)H, 0);
    }
    const int T = (int)ntiles, Himg = (int)(ntiles * H);
    if (c->ntiles != T) { CK(ctx_sync(c)); drop_tiles(c); }
    // image-level context: input, statistics, decoded image, residual layer and the assembled codestreams
    if (reserve_impl(c, (size_t)Himg, W, nframes, 0, 0, T, Himg, 0)) return -1;
    c->ntiles = T;
    TileSrc src[EBCC_MAX_TILES];
    for (int t = 0; t < T; ++t) {
        if (!c->tile[t] && !(c->tile[t] = ctx_create_tile(c))) return -1;
        if (reserve_impl(c->tile[t], H, W, nframes, t * (int)H, t, T, Himg, 1)) return -1;
        src[t].out = c->tile[t]->d_out; src[t].pout = c->tile[t]->d_pout; src[t].stride = c->tile[t]->plan.out_stride;
    }
    if (!c->d_tile_src) CK(cudaMalloc((void **)&c->d_tile_src, sizeof(TileSrc) * EBCC_MAX_TILES));
    CK(cudaMemcpyAsync(c->d_tile_src, src, sizeof(TileSrc) * T, cudaMemcpyHostToDevice, c->stream));
    CK(ctx_sync(c));
    return 0;
}

extern "C" int ebcc_cuda_reserve(ebcc_cuda_ctx *c, size_t H, size_t W, size_t nframes)
{
    return ebcc_cuda_reserve_tiled(c, 1, H, W, nframes);
}

static void mark(ebcc_cuda_ctx *c, int i)
{
    if (c->profiling) cudaEventRecord(c->ev[i], c->stream);
}

extern "C" int ebcc_cuda_upload(ebcc_cuda_ctx *c, const float *frames, size_t nframes)
{
    if (!c || nframes > c->cap_frames) return -1;
    CK(cudaSetDevice(c->device));
    const size_t bytes = nframes * (size_t)c->plan.H * c->plan.W * sizeof(float);
    mark(c, EBCC_ST_H2D);
#ifndef EBCC_CUEMU
    // Pageable caller memory (what HDF5, zarr and numpy hand over): cudaMemcpyAsync would stage it through the driver's own
    // bounce buffer, synchronously and at a fraction of the link rate.  The pipeline stages it itself through a small ring
    // of pinned slots: this thread copies slice k + 1 into a slot while the copy engine moves slice k, and the threads of
    // the other pipelines do the same side by side, so together they fill the link.
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, frames) != cudaSuccess) { (void)cudaGetLastError(); at.type = cudaMemoryTypeUnregistered; }
    if (at.type == cudaMemoryTypeUnregistered && bytes > 0) {
        if (!c->h_up[0]) {
            for (int k = 0; k < UPLOAD_SLOTS; ++k) {
                CK(cudaMallocHost((void **)&c->h_up[k], UPLOAD_SLOT_BYTES));
                CK(cudaEventCreateWithFlags(&c->ev_up[k], cudaEventDisableTiming));
            }
        }
        const char *src = (const char *)frames;
        size_t off = 0;
        for (int k = 0; off < bytes; ++k) {
            const int slot = k % UPLOAD_SLOTS;
            const size_t n = bytes - off < UPLOAD_SLOT_BYTES ? bytes - off : UPLOAD_SLOT_BYTES;
            if (k >= UPLOAD_SLOTS) CK(cudaEventSynchronize(c->ev_up[slot]));      // the slot's previous slice has left
            memcpy(c->h_up[slot], src + off, n);
            CK(cudaMemcpyAsync((char *)c->d_in + off, c->h_up[slot], n, cudaMemcpyHostToDevice, c->stream));
            CK(cudaEventRecord(c->ev_up[slot], c->stream));
            off += n;
        }
        return 0;
    }
#endif
    // cudaMemcpyDefault: frames may be pinned host memory or a device buffer of any GPU of the process (unified virtual
    // addressing), e.g. a tensor produced on the device -- then no H2D copy happens at all
    CK(cudaMemcpyAsync(c->d_in, frames, bytes, cudaMemcpyDefault, c->stream));
    return 0;
}

static int launch_dwt(ebcc_cuda_ctx *c, int B)
{
    const Plan *pl = &c->plan;
    const int W = pl->W, H = pl->H;
    KT kt_(c, EBCC_KT_DWT);
    if (ctx_root_mut(c)->env_fused_dwt) {
        // one launch per level: vertical + horizontal analysis of a tile in shared memory; the LL band travels through the
        // decoder's two work planes (idle during the analysis), the high bands go straight to their place
        for (int r = J2K_NRES - 1; r >= 1; --r) {
            DwtFusedArgs a;
            const bool first = r == J2K_NRES - 1;
            a.in = first ? c->d_in : (((J2K_NRES - 1 - r) & 1) ? c->d_dcoef : c->d_dcoef2);
            a.ll_out = r == 1 ? c->d_coef : (((J2K_NRES - 1 - r) & 1) ? c->d_dcoef2 : c->d_dcoef);
            a.coef = c->d_coef; a.stat = c->d_stat; a.W = W; a.H = H;
            a.rw = pl->res_w[r]; a.rh = pl->res_h[r]; a.snw = pl->res_w[r - 1]; a.snh = pl->res_h[r - 1];
            const dim3 grid((unsigned)((a.rw + FT_TW - 1) / FT_TW), (unsigned)((a.rh + FT_TH - 1) / FT_TH), (unsigned)B);
            void (*kern)(DwtFusedArgs) = pl->cas_v[r] == 0 ? (first ? k_dwt_tile<true, 0> : k_dwt_tile<false, 0>)
                                                           : (first ? k_dwt_tile<true, 1> : k_dwt_tile<false, 1>);
            EBCC_LAUNCH(kern, grid, dim3(FT_THREADS), 0, c->stream, a);
            c->launches++;
        }
        return 0;
    }
    for (int r = J2K_NRES - 1; r >= 1; --r) {
        const int rw = pl->res_w[r], rh = pl->res_h[r];
        // vertical: strips of CW columns, full height in shared memory
        int CW = ctx_root_mut(c)->env_dwt_cw;
        while ((size_t)rh * CW * sizeof(float) > 96 * 1024 && CW > 4) CW >>= 1;
        dim3 bv((unsigned)CW, (unsigned)(512 / CW));
        dim3 gv((unsigned)((rw + CW - 1) / CW), (unsigned)B);
        size_t sv = (size_t)rh * CW * sizeof(float);
        if (r == J2K_NRES - 1)
            EBCC_LAUNCH(k_dwt_v<true>, gv, bv, sv, c->stream, c->d_in, c->d_coef, c->d_stat, W, H, rw, rh, pl->res_h[r - 1], pl->cas_v[r]);
        else
            EBCC_LAUNCH(k_dwt_v<false>, gv, bv, sv, c->stream, c->d_in, c->d_coef, c->d_stat, W, H, rw, rh, pl->res_h[r - 1], pl->cas_v[r]);
        // horizontal: RH full rows per CTA
        int RH = 4;
        while ((size_t)rw * RH * sizeof(float) > 64 * 1024 && RH > 1) RH >>= 1;
        dim3 bh(128, (unsigned)RH);
        dim3 gh((unsigned)((rh + RH - 1) / RH), (unsigned)B);
        EBCC_LAUNCH(k_dwt_h, gh, bh, (size_t)rw * RH * sizeof(float), c->stream, c->d_coef, W, H, rw, rh, pl->res_w[r - 1]);
        c->launches += 2;
    }
    return 0;
}

// Code-blocks per warp of the two serial tier-1 coders.  A warp that steps L blocks in lock step issues about 1/L of the
// instructions but every block's chain gets longer (lanes diverge), so the best L depends on what else runs: alone on
// the GPU a short chain wins (measured on B200, k_t1_mq alone: 1 frame L=1 5.0 ms / L=32 12.2 ms; 24 frames L=4 8.4 ms /
// L=32 12.6 ms), with several pipelines in flight the issue slots are the scarce resource and L = 32 wins.
// EBCC_MQ_LANES / EBCC_DEC_LANES override.
static const ebcc_cuda_ctx *ctx_root(const ebcc_cuda_ctx *c) { return c->parent ? c->parent : c; }
static int ctx_crowd(const ebcc_cuda_ctx *c) { const ebcc_cuda_ctx *r = ctx_root(c); return r->crowd > 0 ? r->crowd : 1; }
// the environment is read when a context is created and at every ebcc_cuda_set_crowd (once per host-level call), not per launch
static void read_env(ebcc_cuda_ctx *c)
{
    const char *e = getenv("EBCC_MQ_LANES");
    c->env_mq_lanes = e ? atoi(e) : 0;
    e = getenv("EBCC_DEC_LANES");
    c->env_dec_lanes = e ? atoi(e) : 0;
    e = getenv("EBCC_NO_RECORDING");                // debugging aid: every probe decodes every block for real
    c->env_no_recording = e && e[0] == '1';
    e = getenv("EBCC_DECODER_RECORDING");           // cross-check aid: recordings made by the tier-1 decoder (round-1 path)
    c->env_dec_recording = e && e[0] == '1';
    c->env_debug_lanes = getenv("EBCC_DEBUG_LANES") != NULL;
    e = getenv("EBCC_T1REC_CTAS");                  // tuning: CTAs per SM the recording variant of k_t1_cm is compiled for (2: 64 registers; 3, the default: 40, +2.8 % on the pipelined step)
    c->env_t1rec_ctas = e ? atoi(e) : 3;
    e = getenv("EBCC_FUSED_IDWT");                  // 0: the unfused inverse transform (k_dequant / k_recon_dequant, k_idwt_h / k_idwt_v, k_finalize)
    c->env_fused_idwt = !(e && e[0] == '0');
    c->env_cuts_from_zero = getenv("EBCC_CUTS_FROM_PASS0") != NULL;   // cross-check aid: every unproven cut decoded from pass 0
    c->env_all_unproven = getenv("EBCC_DEBUG_ALL_CUTS_UNPROVEN") != NULL;   // test aid: every cut goes to the pool (as far as it holds)
    c->env_no_pcrd_cache = getenv("EBCC_NO_PCRD_CACHE") != NULL;   // cross-check aid: every probe bisects from scratch
    e = getenv("EBCC_SP_DECODE_SERIAL");            // cross-check aid: the one-thread SPIHT list decoder
    c->env_sp_serial = e && e[0] == '1';
    e = getenv("EBCC_FUSED_DWT");                   // 0: the unfused forward transform (k_dwt_v / k_dwt_h)
    c->env_fused_dwt = !(e && e[0] == '0');
    e = getenv("EBCC_DWT_CW");                      // tuning: columns per CTA of the vertical 9/7 passes (shared memory = height x columns x 4 bytes)
    c->env_dwt_cw = e ? atoi(e) : 32;
    if (c->env_dwt_cw != 4 && c->env_dwt_cw != 8 && c->env_dwt_cw != 16 && c->env_dwt_cw != 32) c->env_dwt_cw = 32;
}
// measurement aid (EBCC_DEBUG_REPEAT=cm|mq|cuts|idwt|dwt): the named launch group runs twice (same results), which shows
// what one more of it costs a step with every pipeline in flight
static int dbg_repeat(const char *cls)
{
    static const char *e = getenv("EBCC_DEBUG_REPEAT");
    return (e && !strcmp(e, cls)) ? 2 : 1;
}
static int mq_lanes(const ebcc_cuda_ctx *c, long long jobs)
{
    { const int l = ctx_root(c)->env_mq_lanes; if (l == 1 || l == 2 || l == 4 || l == 8 || l == 16 || l == 32) return l; }
    const int crowd = ctx_crowd(c);
    if (crowd >= 3) return MQ_DEFAULT_LANES;
    if (crowd == 2) return 8;
    return jobs <= 1600 ? 1 : (jobs <= 3600 ? 2 : 4);
}
static int dec_lanes(const ebcc_cuda_ctx *c)
{
    { const int l = ctx_root(c)->env_dec_lanes; if (l == 1 || l == 2 || l == 4 || l == 8) return l; }
    return ctx_crowd(c) >= 3 ? DEC_DEFAULT_LANES : 1;
}

extern "C" void ebcc_cuda_set_crowd(ebcc_cuda_ctx *c, int pipelines_at_work) { if (c) { c->crowd = pipelines_at_work; read_env(c); } }

static int launch_minmax(ebcc_cuda_ctx *c, int B)
{
    const int N = c->plan.H * c->plan.W;
    KT kt_(c, EBCC_KT_MINMAX);
    EBCC_LAUNCH(k_stat_init, dim3((unsigned)((B + 127) / 128)), dim3(128), 0, c->stream, c->d_stat, B);
    // enough CTAs to keep every SM loading: 74 per frame for a batch, 296 for a handful of frames
    EBCC_LAUNCH(k_minmax, dim3(B >= 8 ? 74u : 296u, (unsigned)B), dim3(256), 0, c->stream, c->d_in, c->d_stat, N);
    EBCC_LAUNCH(k_minmax_final, dim3((unsigned)((B + 127) / 128)), dim3(128), 0, c->stream, c->d_in, c->d_stat, N, B);
    c->launches += 3;
    return 0;
}

static int analyze_impl(ebcc_cuda_ctx *c, size_t nframes, int with_minmax, int for_search);

// multi-tile image: min/max over the whole image (src/ebcc_codec.c:676-689 scales the image, not the tile), then
// every tile context transforms and codes its tile of every image
static int analyze_tiled(ebcc_cuda_ctx *c, size_t nframes, int for_search)
{
    const int B = (int)nframes, T = c->ntiles;
    mark(c, EBCC_ST_MINMAX);
    if (launch_minmax(c, B)) return -1;
    mark(c, EBCC_ST_DWT);
    for (int t = 0; t < T; ++t) {
        ebcc_cuda_ctx *s = c->tile[t];
        const size_t Nt = (size_t)s->plan.H * s->plan.W;
        CK(cudaMemcpyAsync(s->d_stat, c->d_stat, sizeof(FrameStat) * B, cudaMemcpyDeviceToDevice, c->stream));
        CK(cudaMemcpy2DAsync(s->d_in, Nt * sizeof(float), c->d_in + (size_t)t * Nt, (size_t)T * Nt * sizeof(float), Nt * sizeof(float),
                             (size_t)B, cudaMemcpyDeviceToDevice, c->stream));
        if (analyze_impl(s, nframes, 0, for_search)) return -1;
    }
    mark(c, EBCC_ST_T1_CM);
    mark(c, EBCC_ST_T1_MQ);
    return 0;
}

// stage 1 of the encoder: everything that does not depend on the compression ratio
extern "C" int ebcc_cuda_analyze(ebcc_cuda_ctx *c, size_t nframes, int for_search)
{
    if (!c || nframes == 0 || nframes > c->cap_frames) return -1;
    CK(cudaSetDevice(c->device));
    if (c->ntiles > 1) return analyze_tiled(c, nframes, for_search);
    return analyze_impl(c, nframes, 1, for_search);
}

extern "C" int ebcc_cuda_analyze_resident(ebcc_cuda_ctx *c, size_t nframes) { return ebcc_cuda_analyze(c, nframes, 1); }

static int analyze_impl(ebcc_cuda_ctx *c, size_t nframes, int with_minmax, int for_search)
{
    const Plan *pl = &c->plan;
    const int B = (int)nframes;
    c->for_search = for_search;
    CK(cudaMemsetAsync(c->d_pcache, 0, (size_t)B * sizeof(PcrdCache), c->stream));     // new frames: nothing evaluated yet
    c->use_recording = !ctx_root(c)->env_no_recording;
    c->rec_from_encoder = for_search && c->use_recording && !ctx_root(c)->env_dec_recording;
    if (for_search && !c->rec_from_encoder) {
        const long long nb = (long long)pl->ncblks * B;
        EBCC_LAUNCH(k_dstate_reset, dim3((unsigned)((nb + 255) / 256)), dim3(256), 0, c->stream, c->d_dstate, nb);
        c->launches++;
    }
    mark(c, EBCC_ST_MINMAX);
    if (with_minmax && launch_minmax(c, B)) return -1;

    mark(c, EBCC_ST_DWT);
    for (int rep = 0; rep < dbg_repeat("dwt"); ++rep) if (launch_dwt(c, B)) return -1;

    mark(c, EBCC_ST_T1_CM);
    KT *kt_cm = new KT(c, EBCC_KT_T1_CM);
    if (c->rec_from_encoder && ctx_root(c)->env_t1rec_ctas == 3)
        EBCC_LAUNCH((k_t1_cm<true, 3>), dim3((unsigned)pl->ncblks, (unsigned)B), dim3(T1_THREADS), 0, c->stream,
                    c->d_coef, c->d_cblks, c->d_bands, c->d_luts, (const unsigned char *)c->d_declut, c->d_cbout, c->d_sym, pl->W, pl->H, pl->ncblks,
                    pl->sym_stride, c->d_iplane, c->d_spass, c->d_stat);
    else if (c->rec_from_encoder)
        for (int rep = 0; rep < dbg_repeat("cm"); ++rep)
        EBCC_LAUNCH((k_t1_cm<true, 2>), dim3((unsigned)pl->ncblks, (unsigned)B), dim3(T1_THREADS), 0, c->stream,
                    c->d_coef, c->d_cblks, c->d_bands, c->d_luts, (const unsigned char *)c->d_declut, c->d_cbout, c->d_sym, pl->W, pl->H, pl->ncblks,
                    pl->sym_stride, c->d_iplane, c->d_spass, c->d_stat);
    else
        EBCC_LAUNCH((k_t1_cm<false, T1_MIN_CTAS>), dim3((unsigned)pl->ncblks, (unsigned)B), dim3(T1_THREADS), 0, c->stream,
                    c->d_coef, c->d_cblks, c->d_bands, c->d_luts, (const unsigned char *)c->d_declut, c->d_cbout, c->d_sym, pl->W, pl->H, pl->ncblks,
                    pl->sym_stride, (int *)NULL, (unsigned char *)NULL, c->d_stat);
    delete kt_cm;
    mark(c, EBCC_ST_T1_MQ);
    ebcc_cuda_ctx *root = ctx_root_mut(c);
    cudaStream_t ls = c->stream;                   // stream of the long serial kernels
    if (root->prio_split) {
        CK(cudaEventRecord(root->ev_fork, c->stream));
        CK(cudaStreamWaitEvent(root->stream_lo, root->ev_fork, 0));
        ls = root->stream_lo;
    }
    {
        KT kt_(c, EBCC_KT_T1_MQ, ls);
        const long long jobs = (long long)pl->ncblks * B;
        const int lanes = mq_lanes(c, jobs);                     // code-blocks per warp: 1, 2, 4, 8, 16 or 32
        if (ctx_root(c)->env_debug_lanes) fprintf(stderr, "[ebcc-cuda] mq lanes %d (crowd %d, jobs %lld)\n", lanes, ctx_crowd(c), jobs);
        const long long per_cta = (long long)MQ_WARPS(lanes) * lanes;
        const dim3 grid((unsigned)((jobs + per_cta - 1) / per_cta)), block(MQ_THREADS(lanes));
#define MQ_GO(LL) EBCC_LAUNCH(k_t1_mq<LL>, grid, block, 0, ls, c->d_cblks, c->d_cbout, c->d_sym, c->d_mq, \
                              c->d_mqtab, pl->ncblks, B, pl->sym_stride, pl->mq_stride, c->rec_from_encoder ? c->d_dstate : (T1CutState *)NULL, c->d_stat, \
                              (c->rec_from_encoder && !ctx_root(c)->env_cuts_from_zero) ? c->d_ctxsnap : (unsigned char *)NULL, ctx_root(c)->env_all_unproven)
        for (int rep = 0; rep < dbg_repeat("mq"); ++rep) {
        if (lanes == 32) MQ_GO(32); else if (lanes == 16) MQ_GO(16); else if (lanes == 8) MQ_GO(8); else if (lanes == 4) MQ_GO(4);
        else if (lanes == 2) MQ_GO(2); else MQ_GO(1);
        }
#undef MQ_GO
    }
    c->launches += 2;
    if (c->rec_from_encoder) {
        // every unproven truncation point of the batch is decoded now, all of them side by side (see k_t1_dec_cuts)
        KT kt_(c, EBCC_KT_T1_DEC, ls);
        const long long jobs = (long long)pl->ncblks * B;
        CK(cudaMemsetAsync(c->d_cut_count, 0, sizeof(int), ls));
        // entries a full pool leaves unwritten must read as invalid (k_cut_collect reserves before it checks the room)
        CK(cudaMemsetAsync(c->d_cut_list, 0xff, (size_t)c->cut_cap * sizeof(T1CutEntry), ls));
        EBCC_LAUNCH(k_cut_collect, dim3((unsigned)((jobs + 255) / 256)), dim3(256), 0, ls, (const CblkOut *)c->d_cbout, c->d_dstate, c->d_cut_list,
                    c->d_cut_count, c->cut_cap, jobs);
        for (int rep = 0; rep < dbg_repeat("cuts"); ++rep)
        EBCC_LAUNCH(k_t1_dec_cuts, dim3((unsigned)((c->cut_cap + CUT_WPC - 1) / CUT_WPC)), dim3(CUT_WPC * 32), 0, ls, c->d_cblks, (const CblkOut *)c->d_cbout,
                    (const T1CutEntry *)c->d_cut_list, (const int *)c->d_cut_count, c->cut_cap, (const unsigned char *)c->d_mq, pl->mq_stride, c->d_declut,
                    c->d_cut_pool, pl->ncblks, (const int *)c->d_iplane, (const unsigned char *)c->d_spass,
                    ctx_root(c)->env_cuts_from_zero ? (const unsigned char *)NULL : (const unsigned char *)c->d_ctxsnap, pl->W, pl->H, B);
        CK(cudaMemcpyAsync(c->h_cut_count, c->d_cut_count, sizeof(int), cudaMemcpyDeviceToHost, ls));
        c->launches += 2;
    }
    if (root->prio_split) {
        CK(cudaEventRecord(root->ev_join, ls));
        CK(cudaStreamWaitEvent(c->stream, root->ev_join, 0));
    }
    c->cut_known = 0; c->cut_overflow = 1;
    CK(cudaGetLastError());
    return 0;
}

// stage 2: PCRD + tier-2 for one compression ratio per frame; with_bodies also assembles the codestreams
static int launch_rate_stage(ebcc_cuda_ctx *c, int B, const float *cr, int with_bodies)
{
    const Plan *pl = &c->plan;
    if (c->ntiles > 1) {
        // per-tile rate allocation (opj_j2k_update_rates splits the budget per tile, j2k.c:5361-5408), then the
        // tile-parts are gathered behind one main header
        mark(c, EBCC_ST_PCRD_T2);
        for (int t = 0; t < c->ntiles; ++t) if (launch_rate_stage(c->tile[t], B, cr, with_bodies)) return -1;
        mark(c, EBCC_ST_BODIES);
        if (with_bodies) {
            EBCC_LAUNCH(k_assemble_tiles, dim3((unsigned)c->ntiles, (unsigned)B), dim3(256), 0, c->stream, c->d_tile_src, c->ntiles,
                        c->tile[0]->plan.main_header_len, c->d_out, pl->out_stride, c->d_pout);
            EBCC_LAUNCH(k_offsets, dim3(1), dim3(32), 0, c->stream, c->d_pout, c->d_offs, B);
            EBCC_LAUNCH(k_compact, dim3(8, (unsigned)B), dim3(256), 0, c->stream, c->d_out, c->d_pout, c->d_offs, c->d_packed, pl->out_stride);
            c->launches += 3;
        }
        mark(c, EBCC_ST_D2H);
        CK(cudaGetLastError());
        return 0;
    }
    for (int i = 0; i < B; ++i) c->h_maxlen[i] = tile_maxlen(pl->H, pl->W, cr[i], pl->main_header_len, pl->ntiles);
    CK(cudaMemcpyAsync(c->d_maxlen, c->h_maxlen, sizeof(int) * B, cudaMemcpyHostToDevice, c->stream));
    mark(c, EBCC_ST_PCRD_T2);
    {
        KT kt_(c, EBCC_KT_PCRD_T2);
        EBCC_LAUNCH(k_pcrd_t2, dim3((unsigned)B), dim3(T2_THREADS), t2_smem_bytes(pl->ncblks, pl->nnodes), c->stream,
                    c->d_plan, c->d_cblks, c->d_cbout, c->d_maxlen, c->d_out, c->d_body_off, c->d_pout, c->d_stat, ctx_root(c)->env_no_pcrd_cache ? (PcrdCache *)NULL : c->d_pcache);
    }
    c->launches += 1;
    mark(c, EBCC_ST_BODIES);
    if (with_bodies) {
        KT kt_(c, EBCC_KT_BODIES);
        EBCC_LAUNCH(k_t2_bodies, dim3((unsigned)pl->ncblks, (unsigned)B), dim3(128), 0, c->stream,
                    c->d_plan, c->d_cblks, c->d_cbout, c->d_mq, c->d_body_off, c->d_out, c->d_pout);
        EBCC_LAUNCH(k_offsets, dim3(1), dim3(32), 0, c->stream, c->d_pout, c->d_offs, B);
        EBCC_LAUNCH(k_compact, dim3(8, (unsigned)B), dim3(256), 0, c->stream, c->d_out, c->d_pout, c->d_offs, c->d_packed, pl->out_stride);
        c->launches += 3;
    }
    mark(c, EBCC_ST_D2H);
    CK(cudaGetLastError());
    return 0;
}

extern "C" int ebcc_cuda_j2k_encode_resident(ebcc_cuda_ctx *c, size_t nframes, float base_cr)
{
    if (ebcc_cuda_analyze(c, nframes, 0)) return -1;
    float *cr = (float *)malloc(sizeof(float) * nframes);
    for (size_t i = 0; i < nframes; ++i) cr[i] = base_cr;
    int rc = launch_rate_stage(c, (int)nframes, cr, 1);
    free(cr);
    return rc;
}

// ---- decoder back half on the resident batch: tier-1 decode -> dequant -> IDWT -> float frames ---------
// jobs/pool describe where every code-block's MQ segment lives; if orig != NULL the error statistics of the
// reconstruction against orig (per-frame targets in d_targets) are accumulated into d_es.
static int run_decode_stage(ebcc_cuda_ctx *c, int B, const unsigned char *pool, size_t pool_stride, const float *orig)
{
    const Plan *pl = &c->plan;
    const int W = pl->W, H = pl->H, N = W * H;
    const long long njobs = (long long)pl->ncblks * B;
    const bool probe = orig != NULL;         // probe of the CR search: jobs come from the encoder, the recording is used
    // a probe whose batch has every unproven cut in the pool has nothing to decode: recording, verdicts and pool serve it
    if (!(probe && c->rec_from_encoder && c->use_recording && c->cut_known && !c->cut_overflow)) {
        KT kt_(c, EBCC_KT_T1_DEC);
        const int lanes = dec_lanes(c);                         // code-blocks per warp of the tier-1 decoder: 1, 2, 4 or 8
        const long long per_cta = (long long)DEC_WPC(lanes) * lanes;
        const dim3 grid((unsigned)((njobs + per_cta - 1) / per_cta)), block((unsigned)(DEC_WPC(lanes) * 32));
#define DEC_GO(LL) EBCC_LAUNCH(k_t1_dec<LL>, grid, block, DEC_SMEM_BYTES(LL), c->stream, c->d_cblks, c->d_order, c->d_jobs, pool, pool_stride, c->d_declut, \
                               c->d_iplane, probe ? c->d_spass : (unsigned char *)NULL, c->d_scratch,                                  \
                               probe ? (const CblkOut *)c->d_cbout : (const CblkOut *)NULL, probe ? c->d_dstate : (DecState *)NULL,   \
                               probe ? c->d_decsave : (unsigned *)NULL, W, H, pl->ncblks, B)
        if (lanes == 8) DEC_GO(8); else if (lanes == 4) DEC_GO(4); else if (lanes == 2) DEC_GO(2); else DEC_GO(1);
#undef DEC_GO
    }
    // The fused path serves decodes and every probe the recording, the verdicts and the pool of pre-decoded cuts answer by
    // themselves (no tier-1 decoder ran above); the rare probe that had to decode blocks into the scratch plane takes the
    // unfused kernels below.
    const bool no_t1_decode = probe && c->rec_from_encoder && c->use_recording && c->cut_known && !c->cut_overflow;
    if (ctx_root(c)->env_fused_idwt && (!probe || no_t1_decode)) {
        KT kt_iw(c, EBCC_KT_IDWT);
        // (a probe's statistics were reset by k_jobs_from_enc)
        for (int r = 1; r < J2K_NRES; ++r) {
            IdwtFusedArgs a;
            a.plan = c->d_plan; a.cblks = c->d_cblks; a.iplane = c->d_iplane; a.spass = c->d_spass;
            a.cut_pool = (const int *)c->d_cut_pool; a.rjobs = c->d_rjobs;
            a.ll_in = (r & 1) ? c->d_dcoef2 : c->d_dcoef;
            a.out = (r & 1) ? c->d_dcoef : c->d_dcoef2;
            a.r = r; a.rw = pl->res_w[r]; a.rh = pl->res_h[r]; a.snw = pl->res_w[r - 1]; a.snh = pl->res_h[r - 1];
            a.stat = c->d_stat; a.orig = orig; a.targets = orig ? c->d_targets : (const float *)NULL; a.es = c->d_es;
            a.cap = (probe && c->skip_dec) ? ctx_root(c)->probe_cap : 0ull;
            const dim3 grid((unsigned)B, (unsigned)(((a.rw + FT_TW - 1) / FT_TW) * ((a.rh + FT_TH - 1) / FT_TH)));
            const bool fin = r == J2K_NRES - 1;
            if (fin) a.out = (probe && c->skip_dec) ? (float *)NULL : c->d_dec;
            void (*kern)(IdwtFusedArgs);
            if (pl->cas_v[r] == 0)
                kern = probe ? (fin ? k_idwt_tile<IDF_SRC_RECON, true, 0> : k_idwt_tile<IDF_SRC_RECON, false, 0>)
                             : (fin ? k_idwt_tile<IDF_SRC_IPLANE, true, 0> : k_idwt_tile<IDF_SRC_IPLANE, false, 0>);
            else
                kern = probe ? (fin ? k_idwt_tile<IDF_SRC_RECON, true, 1> : k_idwt_tile<IDF_SRC_RECON, false, 1>)
                             : (fin ? k_idwt_tile<IDF_SRC_IPLANE, true, 1> : k_idwt_tile<IDF_SRC_IPLANE, false, 1>);
            EBCC_LAUNCH(kern, grid, dim3(FT_THREADS), 0, c->stream, a);
            c->launches++;
        }
        CK(cudaGetLastError());
        return 0;
    }
    KT *kt_rc = new KT(c, EBCC_KT_RECON);
    if (probe)
        EBCC_LAUNCH(k_recon_dequant, dim3(148 * 2, (unsigned)B), dim3(256), 0, c->stream, c->d_plan, c->d_cbout, c->d_dstate, c->d_jobs,
                    c->d_iplane, c->d_spass, c->d_scratch, c->d_dcoef, c->d_cblks, (const int *)c->d_cut_pool);
    else
        EBCC_LAUNCH(k_dequant, dim3(148 * 2, (unsigned)B), dim3(256), 0, c->stream, c->d_iplane, c->d_dcoef, c->d_plan, B);
    delete kt_rc;
    c->launches += 2;
    KT *kt_iw = new KT(c, EBCC_KT_IDWT);
    for (int r = 1; r < J2K_NRES; ++r) {
        const int rw = pl->res_w[r], rh = pl->res_h[r];
        int RH = 4;
        while ((size_t)rw * RH * sizeof(float) > 64 * 1024 && RH > 1) RH >>= 1;
        EBCC_LAUNCH(k_idwt_h, dim3((unsigned)((rh + RH - 1) / RH), (unsigned)B), dim3(128, (unsigned)RH),
                    (size_t)rw * RH * sizeof(float), c->stream, c->d_dcoef, W, H, rw, rh, pl->res_w[r - 1]);
        int CW = ctx_root(c)->env_dwt_cw;
        while ((size_t)rh * CW * sizeof(float) > 96 * 1024 && CW > 4) CW >>= 1;
        EBCC_LAUNCH(k_idwt_v, dim3((unsigned)((rw + CW - 1) / CW), (unsigned)B), dim3((unsigned)CW, (unsigned)(512 / CW)),
                    (size_t)rh * CW * sizeof(float), c->stream, c->d_dcoef, W, H, rw, rh, pl->res_h[r - 1], pl->cas_v[r]);
        c->launches += 2;
    }
    delete kt_iw;
    KT kt_fin(c, EBCC_KT_FINALIZE);
    if (orig) { EBCC_LAUNCH(k_errstat_init, dim3((unsigned)((B + 127) / 128)), dim3(128), 0, c->stream, c->d_es, B); c->launches++; }
    EBCC_LAUNCH(k_finalize, dim3(148 * 2, (unsigned)B), dim3(256), 0, c->stream, c->d_dcoef, c->d_dec, c->d_stat, orig,
                orig ? c->d_targets : (const float *)NULL, c->d_es, N);
    c->launches++;
    CK(cudaGetLastError());
    return 0;
}


// One probe of the compression-ratio search (emulate_j2k_compression, src/ebcc_codec.c:535-543) for every frame of
// the resident batch: rate allocation at cr[i], real tier-1 decode of the truncated segments, IDWT, and the error
// statistics against the original frame.  Synchronous; res[i] is filled for every frame.
// device work of one probe on one (tile) context; the statistics land in the context's pinned h_es / h_pout
static int probe_launch(ebcc_cuda_ctx *c, int B, const float *targets)
{
    const Plan *pl = &c->plan;
    if (c->rec_from_encoder && !c->cut_known) {
        // first probe after the analysis: how many unproven cuts did the batch have?  (the host codec has synchronised
        // already, the wait is over before it starts)
        CK(ctx_sync(c));
        c->cut_overflow = *c->h_cut_count > c->cut_cap;
        c->cut_known = 1;
    }
    for (int i = 0; i < B; ++i) c->h_targets[i] = targets[i];
    CK(cudaMemcpyAsync(c->d_targets, c->h_targets, sizeof(float) * B, cudaMemcpyHostToDevice, c->stream));
    const long long njobs = (long long)pl->ncblks * B;
    for (int rep = 0; rep < dbg_repeat("idwt"); ++rep) {
    EBCC_LAUNCH(k_jobs_from_enc, dim3((unsigned)((njobs + 255) / 256)), dim3(256), 0, c->stream, c->d_cblks, c->d_cbout, c->d_dstate, c->d_jobs,
                pl->ncblks, B, c->use_recording, c->d_rjobs, c->d_es);
    c->launches++;
    if (run_decode_stage(c, B, c->d_mq, pl->mq_stride, c->d_in)) return -1;
    }
    CK(cudaMemcpyAsync(c->h_es, c->d_es, sizeof(ErrStat) * B, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaMemcpyAsync(c->h_pout, c->d_pout, sizeof(ProbeOut) * B, cudaMemcpyDeviceToHost, c->stream));
    return 0;
}

// tile t's decoded lines of every image -> their place in the image-level plane
static int gather_tiles(ebcc_cuda_ctx *c, int B)
{
    const int T = c->ntiles;
    for (int t = 0; t < T; ++t) {
        ebcc_cuda_ctx *s = c->tile[t];
        const size_t Nt = (size_t)s->plan.H * s->plan.W;
        CK(cudaMemcpy2DAsync(c->d_dec + (size_t)t * Nt, (size_t)T * Nt * sizeof(float), s->d_dec, Nt * sizeof(float), Nt * sizeof(float),
                             (size_t)B, cudaMemcpyDeviceToDevice, c->stream));
    }
    return 0;
}

// search probe: like ebcc_cuda_rate_probe(with_streams = 0), but n_over only has to be exact below cap -- a frame that
// reaches cap violations is not examined any further and reports some count >= cap; max_err / rmin / rmax / sum_err of
// such a frame are those of the part examined
extern "C" int ebcc_cuda_rate_probe(ebcc_cuda_ctx *c, size_t nframes, const float *cr, const float *targets, int with_streams, ebcc_probe_result *res);
extern "C" int ebcc_cuda_rate_probe_capped(ebcc_cuda_ctx *c, size_t nframes, const float *cr, const float *targets, unsigned long long cap,
                                           ebcc_probe_result *res)
{
    if (!c) return -1;
    c->probe_cap = cap;
    const int rc = ebcc_cuda_rate_probe(c, nframes, cr, targets, 0, res);
    c->probe_cap = 0;
    return rc;
}

extern "C" int ebcc_cuda_rate_probe(ebcc_cuda_ctx *c, size_t nframes, const float *cr, const float *targets,
                                    int with_streams, ebcc_probe_result *res)
{
    if (!c || nframes == 0 || nframes > c->cap_frames) return -1;
    CK(cudaSetDevice(c->device));
    const int B = (int)nframes;
    if (launch_rate_stage(c, B, cr, with_streams)) return -1;
    c->skip_dec = !with_streams && c->ntiles == 1;      // the decoded frames of a search probe are never read, its statistics are
    if (c->ntiles > 1) {
        const int T = c->ntiles, hl = c->tile[0]->plan.main_header_len;
        for (int t = 0; t < T; ++t) if (probe_launch(c->tile[t], B, targets)) return -1;
        if (gather_tiles(c, B)) return -1;
        CK(ctx_sync(c));
        // image-level statistics: the samples of an image are the samples of its tiles
        for (int i = 0; i < B; ++i) {
            ErrStat e = c->tile[0]->h_es[i];
            int bytes = c->tile[0]->h_pout[i].stream_bytes;
            for (int t = 1; t < T; ++t) {
                const ErrStat &x = c->tile[t]->h_es[i];
                e.n_over += x.n_over; e.sum_fx += x.sum_fx;       // same image range => same fixed-point unit in every tile
                if (x.max_err_bits > e.max_err_bits) e.max_err_bits = x.max_err_bits;
                if (x.rmin_key < e.rmin_key) e.rmin_key = x.rmin_key;
                if (x.rmax_key > e.rmax_key) e.rmax_key = x.rmax_key;
                if (bytes) bytes += c->tile[t]->h_pout[i].stream_bytes - hl - 2;
            }
            c->h_es[i] = e;
            c->h_pout[i].stream_bytes = bytes;
        }
    } else {
        if (probe_launch(c, B, targets)) return -1;
        CK(ctx_sync(c));
    }
    for (int i = 0; i < B; ++i) {
        const ErrStat &e = c->h_es[i];
        res[i].n_over = e.n_over;
        memcpy(&res[i].max_err, &e.max_err_bits, 4);
        unsigned k = e.rmin_key; k = (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k; memcpy(&res[i].rmin, &k, 4);
        k = e.rmax_key; k = (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k; memcpy(&res[i].rmax, &k, 4);
        res[i].sum_err = es_sum(e);
        res[i].stream_bytes = (size_t)c->h_pout[i].stream_bytes;
    }
    return 0;
}


// ---- residual layer ------------------------------------------------------------------------------------------
static int sp_col_cw(int ny) { int cw = 32; while ((size_t)2 * ny * cw * sizeof(float) > 160 * 1024 && cw > 1) cw >>= 1; return cw; }

// SPIHT-encodes the normalised residual (original - reconstruction of the LAST rate probe) of every active frame with
// a bit budget of trunc_bits[i] (spiht_encode, src/spiht/spiht_re.c:432-475; call site src/ebcc_codec.c:744-748).
// The residual range is the one the last ebcc_cuda_rate_probe reported.  coeffs_size[i] receives the stream length.
extern "C" int ebcc_cuda_residual_encode(ebcc_cuda_ctx *c, size_t nframes, const int *active, const size_t *trunc_bits, size_t *coeffs_size)
{
    if (!c || nframes == 0 || nframes > c->cap_frames) return -1;
    CK(cudaSetDevice(c->device));
    const int B = (int)nframes;
    const SpGeom g = c->sg;
    const size_t Np = (size_t)g.Wp * g.Hp;
    for (int i = 0; i < B; ++i) {
        const ErrStat &e = c->h_es[i];
        unsigned k = e.rmin_key; k = (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k; memcpy(&c->h_sp_rmm[2 * i], &k, 4);
        k = e.rmax_key; k = (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k; memcpy(&c->h_sp_rmm[2 * i + 1], &k, 4);
        c->h_sp_active[i] = active[i] ? 1 : 0;
        size_t tb = trunc_bits[i];
        if (tb + 128 + 64 > c->sp.bits_words * 32) tb = c->sp.bits_words * 32 - 192;
        c->h_sp_budget[i] = (unsigned)tb;
    }
    CK(cudaMemcpyAsync(c->d_sp_rmm, c->h_sp_rmm, 8 * B, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(c->d_sp_active, c->h_sp_active, 4 * B, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(c->d_sp_budget, c->h_sp_budget, 4 * B, cudaMemcpyHostToDevice, c->stream));
    EBCC_LAUNCH(k_sp_init, dim3((unsigned)((B + 127) / 128)), dim3(128), 0, c->stream, c->d_sp_info, c->d_sp_rmm, c->d_sp_budget, c->d_sp_active, B);
    CK(cudaMemsetAsync(c->sp.sign_pos, 0, (size_t)B * Np * 4, c->stream));
    CK(cudaMemsetAsync(c->sp.bits, 0, (size_t)B * c->sp.bits_words * 4, c->stream));
    CK(cudaMemsetAsync(c->sp.D, 0, (size_t)B * Np * 4, c->stream));
    CK(cudaMemsetAsync(c->sp.L, 0, (size_t)B * Np * 4, c->stream));
    const dim3 ge(148 * 2, (unsigned)B);
    KT *kt_fw = new KT(c, EBCC_KT_SP_FWD);
    EBCC_LAUNCH(k_sp_load, ge, dim3(256), 0, c->stream, c->d_in, c->d_dec, c->d_sp_img0, c->d_sp_info, g);
    EBCC_LAUNCH(k_sp_subdc, ge, dim3(256), 0, c->stream, c->d_sp_img0, c->d_sp_info, g);
    c->launches += 3;
    for (int l = 0; l < SP_STAGES; ++l) {
        const int nx = g.Wp >> l, ny = g.Hp >> l;
        EBCC_LAUNCH(k_sp_dwt_row, dim3((unsigned)ny, (unsigned)B), dim3(256), (size_t)2 * nx * sizeof(float), c->stream,
                    c->d_sp_img0, c->d_sp_img1, c->d_sp_info, g, nx);
        const int cw = sp_col_cw(ny);
        EBCC_LAUNCH(k_sp_dwt_col, dim3((unsigned)((nx + cw - 1) / cw), (unsigned)B), dim3((unsigned)cw, (unsigned)(512 / cw)),
                    (size_t)2 * ny * cw * sizeof(float), c->stream, c->d_sp_img1, c->d_sp_img0, c->d_sp_info, g, nx, ny);
        c->launches += 2;
    }
    EBCC_LAUNCH(k_sp_normalize, ge, dim3(256), 0, c->stream, c->d_sp_img0, c->sp.coef, c->d_sp_info, g);
    EBCC_LAUNCH(k_sp_desc, dim3(148, (unsigned)B), dim3(256), 0, c->stream, c->sp.coef, c->sp.D, c->sp.L, c->d_sp_info, g, g.Wp / 4, g.Hp / 4, g.Wp / 2, g.Hp / 2);
    EBCC_LAUNCH(k_sp_desc, dim3(148, (unsigned)B), dim3(256), 0, c->stream, c->sp.coef, c->sp.D, c->sp.L, c->d_sp_info, g, g.Wp / 8, g.Hp / 8, g.Wp / 4, g.Hp / 4);
    EBCC_LAUNCH(k_sp_desc, dim3(148, (unsigned)B), dim3(256), 0, c->stream, c->sp.coef, c->sp.D, c->sp.L, c->d_sp_info, g, 0, 0, g.Wp / 8, g.Hp / 8);
    delete kt_fw;
    { KT kt_(c, EBCC_KT_SP_ENCODE);
    EBCC_LAUNCH(k_sp_encode, dim3((unsigned)B), dim3(SP_THREADS), 0, c->stream, c->sp, c->d_sp_info, g, c->d_step_tab); }
    c->launches += 5;
    CK(cudaMemcpyAsync(c->h_sp_info, c->d_sp_info, sizeof(SpInfo) * B, cudaMemcpyDeviceToHost, c->stream));
    CK(ctx_sync(c));
    CK(cudaGetLastError());
    for (int i = 0; i < B; ++i) if (active[i]) coeffs_size[i] = c->h_sp_info[i].nbytes;
    return 0;
}

// inverse SPIHT transform of the plane in d_sp_img0 for every frame with P != 0 (idwt2full, dwt.h:305-317)
static int sp_launch_inverse(ebcc_cuda_ctx *c, int B)
{
    const SpGeom g = c->sg;
    for (int l = SP_STAGES - 1; l >= 0; --l) {
        const int nx = g.Wp >> l, ny = g.Hp >> l;
        const int cw = sp_col_cw(ny);
        EBCC_LAUNCH(k_sp_idwt_col, dim3((unsigned)((nx + cw - 1) / cw), (unsigned)B), dim3((unsigned)cw, (unsigned)(512 / cw)),
                    (size_t)2 * ny * cw * sizeof(float), c->stream, c->d_sp_img0, c->d_sp_img1, c->d_sp_P, g, nx, ny);
        EBCC_LAUNCH(k_sp_idwt_row, dim3((unsigned)ny, (unsigned)B), dim3(256), (size_t)2 * nx * sizeof(float), c->stream,
                    c->d_sp_img1, c->d_sp_img0, c->d_sp_P, g, nx);
        c->launches += 2;
    }
    return 0;
}

// One probe of the truncation search (src/ebcc_codec.c:749-757, :775-780): what spiht_decode(buf, num_bits/8, ..., num_bits)
// would reconstruct from the first num_bits[i] bits of frame i's stream, added to the base layer and compared with the
// original frame.  Frames with active[i] == 0 are skipped.
extern "C" int ebcc_cuda_residual_probe(ebcc_cuda_ctx *c, size_t nframes, const int *active, const size_t *num_bits, float *max_err, double *sum_err)
{
    if (!c || nframes == 0 || nframes > c->cap_frames) return -1;
    CK(cudaSetDevice(c->device));
    const int B = (int)nframes;
    const SpGeom g = c->sg;
    for (int i = 0; i < B; ++i) {
        unsigned P = 0;
        if (active[i]) {
            size_t nb = num_bits[i];
            const size_t bits0 = (size_t)c->h_sp_budget[i] + 128;
            if (nb > bits0) nb = bits0;
            P = nb > 128 ? (unsigned)(nb - 128 + 1) : 1u;
        }
        c->h_sp_P[i] = P;
    }
    CK(cudaMemcpyAsync(c->d_sp_P, c->h_sp_P, 4 * B, cudaMemcpyHostToDevice, c->stream));
    const dim3 ge(148 * 2, (unsigned)B);
    KT *kt_pr = new KT(c, EBCC_KT_SP_PROBE);
    EBCC_LAUNCH(k_sp_recon, ge, dim3(256), 0, c->stream, c->sp, c->d_sp_info, c->d_sp_P, c->d_sp_img0, g);
    c->launches++;
    if (sp_launch_inverse(c, B)) return -1;
    EBCC_LAUNCH(k_errstat_init, dim3((unsigned)((B + 127) / 128)), dim3(128), 0, c->stream, c->d_es2, B);
    EBCC_LAUNCH(k_sp_errstat, ge, dim3(256), 0, c->stream, c->d_sp_img0, c->d_in, c->d_dec, c->d_sp_info, c->d_sp_P, c->d_es2, g);
    c->launches += 2;
    delete kt_pr;
    CK(cudaMemcpyAsync(c->h_es2, c->d_es2, sizeof(ErrStat) * B, cudaMemcpyDeviceToHost, c->stream));
    CK(ctx_sync(c));
    CK(cudaGetLastError());
    for (int i = 0; i < B; ++i) {
        if (!active[i]) continue;
        memcpy(&max_err[i], &c->h_es2[i].max_err_bits, 4);
        sum_err[i] = es_sum(c->h_es2[i]);
    }
    return 0;
}

// zstd stage (ZSTD_compress call site src/ebcc_codec.c:813-817): compresses the first coeffs_size[i] bytes of frame i's
// SPIHT stream into a zstd frame on the device.  frames[i] points into a context-owned pinned buffer.
extern "C" int ebcc_cuda_residual_pack(ebcc_cuda_ctx *c, size_t nframes, const size_t *coeffs_size, const uint8_t **frames, size_t *sizes)
{
    if (!c || nframes == 0 || nframes > c->cap_frames) return -1;
    CK(cudaSetDevice(c->device));
    const int B = (int)nframes;
    int any = 0;
    for (int i = 0; i < B; ++i) { c->h_zsrc_len[i] = (unsigned)coeffs_size[i]; if (coeffs_size[i]) any = 1; frames[i] = NULL; sizes[i] = 0; }
    if (!any) return 0;
    CK(cudaMemcpyAsync(c->d_zsrc_len, c->h_zsrc_len, 4 * B, cudaMemcpyHostToDevice, c->stream));
    size_t maxn = 0;
    for (int i = 0; i < B; ++i) if (coeffs_size[i] > maxn) maxn = coeffs_size[i];
    const int nch = (int)((maxn + Z_CHUNK - 1) / Z_CHUNK);
    KT *kt_z = new KT(c, EBCC_KT_ZSTD_ENC);
    EBCC_LAUNCH(k_zstd_compress_chunks, dim3((unsigned)nch, (unsigned)B), dim3(ZC_THREADS), sizeof(ZEncWork), c->stream, (const unsigned char *)c->sp.bits,
                c->sp.bits_words * 4, c->d_zsrc_len, c->d_zblk, c->d_zblen, c->zmaxchunks);
    EBCC_LAUNCH(k_zstd_assemble, dim3((unsigned)B), dim3(256), 0, c->stream, c->d_zblk, c->d_zblen, c->zmaxchunks, c->d_zsrc_len, c->d_zout,
                c->zstride, c->d_zlen);
    delete kt_z;
    c->launches++;
    c->launches++;
    CK(cudaMemcpyAsync(c->h_zlen, c->d_zlen, 4 * B, cudaMemcpyDeviceToHost, c->stream));
    CK(ctx_sync(c));
    for (int i = 0; i < B; ++i)
        if (c->h_zlen[i]) CK(cudaMemcpyAsync(c->h_zout + (size_t)i * c->zstride, c->d_zout + (size_t)i * c->zstride, c->h_zlen[i], cudaMemcpyDeviceToHost, c->stream));
    CK(ctx_sync(c));
    CK(cudaGetLastError());
    for (int i = 0; i < B; ++i) { frames[i] = c->h_zout + (size_t)i * c->zstride; sizes[i] = c->h_zlen[i]; }
    return 0;
}

// NONE mode: rate allocation + codestream assembly only
extern "C" int ebcc_cuda_rate_probe_streams_only(ebcc_cuda_ctx *c, size_t nframes, const float *cr)
{
    if (!c || nframes == 0 || nframes > c->cap_frames) return -1;
    CK(cudaSetDevice(c->device));
    return launch_rate_stage(c, (int)nframes, cr, 1);
}

extern "C" int ebcc_cuda_frame_stats(ebcc_cuda_ctx *c, size_t nframes, float *minv, float *maxv, int *flags)
{
    if (!c || nframes == 0 || nframes > c->cap_frames) return -1;
    CK(cudaSetDevice(c->device));
    CK(cudaMemcpyAsync(c->h_stat, c->d_stat, sizeof(FrameStat) * nframes, cudaMemcpyDeviceToHost, c->stream));
    CK(ctx_sync(c));
    for (size_t i = 0; i < nframes; ++i) { minv[i] = c->h_stat[i].minv; maxv[i] = c->h_stat[i].maxv; flags[i] = c->h_stat[i].flags; }
    return 0;
}

__global__ void k_set_stat(FrameStat *stat, const float *mm, int n)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { stat[i].minv = mm[2 * i]; stat[i].maxv = mm[2 * i + 1]; stat[i].flags = 0; }
}

// decoded frames of the resident batch -> the callers' buffers.  Pinned host memory and device memory take an asynchronous
// copy each; pageable host memory (malloc'ed buffers, numpy arrays) is served through the pinned ring of ebcc_cuda_upload,
// the other way round: the copy engine fills slot k + 1 while this thread copies slot k to its destination.
static int download_frames(ebcc_cuda_ctx *c, float *const *outs, int B, size_t frame_bytes)
{
#ifndef EBCC_CUEMU
    int npage = 0;
    for (int i = 0; i < B; ++i) {
        cudaPointerAttributes at;
        if (cudaPointerGetAttributes(&at, outs[i]) != cudaSuccess) { (void)cudaGetLastError(); at.type = cudaMemoryTypeUnregistered; }
        c->h_targets[i] = at.type == cudaMemoryTypeUnregistered ? 1.0f : 0.0f;      // (h_targets is idle during a decode)
        npage += at.type == cudaMemoryTypeUnregistered;
    }
    for (int i = 0; i < B; ++i)
        if (c->h_targets[i] == 0.0f)
            CK(cudaMemcpyAsync(outs[i], (const char *)c->d_dec + (size_t)i * frame_bytes, frame_bytes, cudaMemcpyDefault, c->stream));
    if (npage == 0) return 0;
    if (!c->h_up[0]) {
        for (int k = 0; k < UPLOAD_SLOTS; ++k) {
            CK(cudaMallocHost((void **)&c->h_up[k], UPLOAD_SLOT_BYTES));
            CK(cudaEventCreateWithFlags(&c->ev_up[k], cudaEventDisableTiming));
        }
    }
    // slices of at most one slot, frame by frame; `issued` runs UPLOAD_SLOTS - 1 slices ahead of `done`
    struct Slice { char *dst; size_t off, n; };
    const size_t per = (frame_bytes + UPLOAD_SLOT_BYTES - 1) / UPLOAD_SLOT_BYTES;
    const size_t total = (size_t)npage * per;
    size_t issued = 0, done = 0; int fi = -1; size_t si = per;       // iteration state of `issued`
    int fd = -1; size_t sd = per;                                    // ... and of `done`
    auto next = [&](int &f, size_t &sl, Slice &out) {
        if (++sl >= per) { sl = 0; do { ++f; } while (f < B && c->h_targets[f] == 0.0f); }
        out.off = sl * UPLOAD_SLOT_BYTES;
        out.n = frame_bytes - out.off < UPLOAD_SLOT_BYTES ? frame_bytes - out.off : UPLOAD_SLOT_BYTES;
        out.dst = (char *)outs[f] + out.off;
        out.off += (size_t)f * frame_bytes;
    };
    while (done < total) {
        while (issued < total && issued < done + UPLOAD_SLOTS) {
            Slice sl; next(fi, si, sl);
            const int slot = (int)(issued % UPLOAD_SLOTS);
            CK(cudaMemcpyAsync(c->h_up[slot], (const char *)c->d_dec + sl.off, sl.n, cudaMemcpyDeviceToHost, c->stream));
            CK(cudaEventRecord(c->ev_up[slot], c->stream));
            ++issued;
        }
        Slice sl; next(fd, sd, sl);
        const int slot = (int)(done % UPLOAD_SLOTS);
        CK(cudaEventSynchronize(c->ev_up[slot]));
        memcpy(sl.dst, c->h_up[slot], sl.n);
        ++done;
    }
    return 0;
#else
    for (int i = 0; i < B; ++i)
        CK(cudaMemcpyAsync(outs[i], (const char *)c->d_dec + (size_t)i * frame_bytes, frame_bytes, cudaMemcpyDefault, c->stream));
    return 0;
#endif
}

extern "C" int ebcc_cuda_decode_batch_tiled(ebcc_cuda_ctx *c, const uint8_t *const *streams, const size_t *sizes, size_t nframes,
                                            size_t ntiles, size_t H, size_t W, const float *minv, const float *maxv,
                                            const ebcc_residual_in *res, float *const *outs)
{
    if (ebcc_cuda_reserve_tiled(c, ntiles, H, W, nframes)) return -1;
    int any_res = 0;
    for (size_t i = 0; res && i < nframes; ++i)
        if (res[i].zstd_size) {
            any_res = 1;
            if (res[i].zstd_size + 16 > c->zstride || res[i].coeffs_size + 16 > c->sp.bits_words * 4) {
                fprintf(stderr, "[ebcc-cuda] residual payload of frame %zu too large\n", i); return -1;
            }
        }
    const Plan *pl = &c->plan;
    const int B = (int)nframes, T = c->ntiles;
    const size_t N = (size_t)T * H * W;
    for (int i = 0; i < B; ++i) {
        if (sizes[i] > pl->out_stride) { fprintf(stderr, "[ebcc-cuda] codestream %d too large (%zu bytes)\n", i, sizes[i]); return -1; }
        memcpy(c->h_streams + (size_t)i * pl->out_stride, streams[i], sizes[i]);
        c->h_slens[i] = (int)sizes[i];
    }
    // min / max travel in the (pinned) target array: 2 floats per frame would overflow it, so use h_es as scratch
    float *mm = (float *)c->h_es;
    for (int i = 0; i < B; ++i) { mm[2 * i] = minv[i]; mm[2 * i + 1] = maxv[i]; }
    for (int i = 0; i < B; ++i)
        CK(cudaMemcpyAsync(c->d_streams + (size_t)i * pl->out_stride, c->h_streams + (size_t)i * pl->out_stride, sizes[i],
                           cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(c->d_slens, c->h_slens, sizeof(int) * B, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(c->d_es, mm, sizeof(float) * 2 * B, cudaMemcpyHostToDevice, c->stream));
    EBCC_LAUNCH(k_set_stat, dim3((unsigned)((B + 127) / 128)), dim3(128), 0, c->stream, c->d_stat, (const float *)c->d_es, B);
    c->launches++;
    if (T == 1) {
        {
            KT kt_(c, EBCC_KT_T2_PARSE);
            EBCC_LAUNCH(k_t2_parse, dim3((unsigned)B), dim3(32), (size_t)pl->nnodes * 4 * sizeof(int), c->stream,
                        c->d_plan, c->d_cblks, c->d_streams, pl->out_stride, c->d_slens, c->d_jobs, c->d_status);
        }
        c->launches++;
        if (run_decode_stage(c, B, c->d_streams, pl->out_stride, NULL)) return -1;
    } else {
        // every tile context finds its own tile-part in the image's codestream and decodes it
        for (int t = 0; t < T; ++t) {
            ebcc_cuda_ctx *s = c->tile[t];
            CK(cudaMemcpyAsync(s->d_stat, c->d_stat, sizeof(FrameStat) * B, cudaMemcpyDeviceToDevice, c->stream));
            EBCC_LAUNCH(k_t2_parse, dim3((unsigned)B), dim3(32), (size_t)s->plan.nnodes * 4 * sizeof(int), c->stream,
                        s->d_plan, s->d_cblks, c->d_streams, pl->out_stride, c->d_slens, s->d_jobs, s->d_status);
            s->launches++;
            if (run_decode_stage(s, B, c->d_streams, pl->out_stride, NULL)) return -1;
            CK(cudaMemcpyAsync(s->h_status, s->d_status, sizeof(int) * B, cudaMemcpyDeviceToHost, c->stream));
        }
        if (gather_tiles(c, B)) return -1;
    }
    if (any_res) {
        // residual layer: zstd frame -> SPIHT stream -> coefficient plane -> inverse transform -> add (src/ebcc_codec.c:1294-1312)
        const SpGeom g = c->sg;
        for (int i = 0; i < B; ++i) {
            const int on = res[i].zstd_size != 0;
            c->h_sp_active[i] = on; c->h_sp_budget[i] = 0; c->h_sp_P[i] = on ? 1u : 0u;
            c->h_sp_rmm[2 * i] = res[i].rmin; c->h_sp_rmm[2 * i + 1] = res[i].rmax;
            c->h_zsrc_len[i] = (unsigned)res[i].zstd_size; c->h_zcap[i] = (unsigned)res[i].coeffs_size;
            if (on) {
                memcpy(c->h_zin + (size_t)i * c->zstride, res[i].zstd, res[i].zstd_size);
                CK(cudaMemcpyAsync(c->d_zin + (size_t)i * c->zstride, c->h_zin + (size_t)i * c->zstride, res[i].zstd_size, cudaMemcpyHostToDevice, c->stream));
            }
        }
        CK(cudaMemcpyAsync(c->d_sp_rmm, c->h_sp_rmm, 8 * B, cudaMemcpyHostToDevice, c->stream));
        CK(cudaMemcpyAsync(c->d_sp_active, c->h_sp_active, 4 * B, cudaMemcpyHostToDevice, c->stream));
        CK(cudaMemcpyAsync(c->d_sp_budget, c->h_sp_budget, 4 * B, cudaMemcpyHostToDevice, c->stream));
        CK(cudaMemcpyAsync(c->d_sp_P, c->h_sp_P, 4 * B, cudaMemcpyHostToDevice, c->stream));
        CK(cudaMemcpyAsync(c->d_zsrc_len, c->h_zsrc_len, 4 * B, cudaMemcpyHostToDevice, c->stream));
        CK(cudaMemcpyAsync(c->d_zcap, c->h_zcap, 4 * B, cudaMemcpyHostToDevice, c->stream));
        CK(cudaMemsetAsync(c->sp.bits, 0, (size_t)B * c->sp.bits_words * 4, c->stream));
        EBCC_LAUNCH(k_sp_init, dim3((unsigned)((B + 127) / 128)), dim3(128), 0, c->stream, c->d_sp_info, c->d_sp_rmm, c->d_sp_budget, c->d_sp_active, B);
        {
            KT kt_(c, EBCC_KT_ZSTD_DEC);
            EBCC_LAUNCH(k_zstd_decompress, dim3((unsigned)B), dim3(32), 0, c->stream, c->d_zin, c->zstride, c->d_zsrc_len, (unsigned char *)c->sp.bits,
                        c->sp.bits_words * 4, c->d_zcap, c->d_zlen, c->d_zdec_work);
        }
        {
            KT kt_(c, EBCC_KT_SP_DECODE);
            if (ctx_root(c)->env_sp_serial)
                EBCC_LAUNCH(k_sp_decode_serial, dim3((unsigned)B), dim3(32), 0, c->stream, (const unsigned char *)c->sp.bits, c->sp.bits_words * 4, (const int *)c->d_zcap, c->sp,
                            c->d_sp_info, c->d_sp_img0, g, c->d_maxlen);
            else
                EBCC_LAUNCH(k_sp_decode, dim3((unsigned)B), dim3(SPD_THREADS), 0, c->stream, (const unsigned char *)c->sp.bits, c->sp.bits_words * 4, (const int *)c->d_zcap, c->sp,
                            c->d_sp_info, c->d_sp_img0, g, c->d_maxlen);
        }
        c->launches += 3;
        {
            KT kt_(c, EBCC_KT_SP_INV);
            if (sp_launch_inverse(c, B)) return -1;
            EBCC_LAUNCH(k_sp_add_residual, dim3(148 * 2, (unsigned)B), dim3(256), 0, c->stream, c->d_sp_img0, c->d_dec, c->d_sp_info, g);
        }
        c->launches++;
        CK(cudaMemcpyAsync(c->h_zlen, c->d_zlen, 4 * B, cudaMemcpyDeviceToHost, c->stream));
        CK(cudaMemcpyAsync(c->h_maxlen, c->d_maxlen, 4 * B, cudaMemcpyDeviceToHost, c->stream));
    }
    if (T == 1) CK(cudaMemcpyAsync(c->h_status, c->d_status, sizeof(int) * B, cudaMemcpyDeviceToHost, c->stream));
    if (download_frames(c, outs, B, (size_t)N * sizeof(float))) return -1;    // host (pinned or pageable) or device destinations
    CK(ctx_sync(c));
    int bad = 0;
    for (int t = 0; T > 1 && t < T; ++t)
        for (int i = 0; i < B; ++i) {
            if (t == 0) c->h_status[i] = PARSE_OK;
            if (c->tile[t]->h_status[i] != PARSE_OK) c->h_status[i] = c->tile[t]->h_status[i];
        }
    for (int i = 0; i < B; ++i) if (c->h_status[i] != PARSE_OK) { fprintf(stderr, "[ebcc-cuda] malformed JPEG 2000 codestream in frame %d\n", i); bad = 1; }
    for (int i = 0; any_res && i < B; ++i)
        if (res[i].zstd_size && (c->h_zlen[i] == 0xffffffffu || c->h_maxlen[i] != 0)) {
            fprintf(stderr, "[ebcc-cuda] malformed residual layer (%s) in frame %d\n", c->h_zlen[i] == 0xffffffffu ? "zstd frame" : "SPIHT header", i); bad = 1;
        }
    return bad ? -2 : 0;
}

extern "C" int ebcc_cuda_decode_batch(ebcc_cuda_ctx *c, const uint8_t *const *streams, const size_t *sizes, size_t nframes,
                                      size_t H, size_t W, const float *minv, const float *maxv, const ebcc_residual_in *res,
                                      float *const *outs)
{
    return ebcc_cuda_decode_batch_tiled(c, streams, sizes, nframes, 1, H, W, minv, maxv, res, outs);
}

extern "C" int ebcc_cuda_download(ebcc_cuda_ctx *c, size_t nframes, const uint8_t **streams, size_t *sizes,
                                  float *minv, float *maxv, int *flags)
{
    if (!c || nframes > c->cap_frames) return -1;
    CK(cudaSetDevice(c->device));
    const int B = (int)nframes;
    CK(cudaMemcpyAsync(c->h_pout, c->d_pout, sizeof(ProbeOut) * B, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaMemcpyAsync(c->h_stat, c->d_stat, sizeof(FrameStat) * B, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaMemcpyAsync(c->h_offs, c->d_offs, sizeof(unsigned long long) * (B + 1), cudaMemcpyDeviceToHost, c->stream));
    CK(ctx_sync(c));
    const size_t total = (size_t)c->h_offs[B];
    { unsigned char *t = c->h_packed; c->h_packed = c->h_packed_alt; c->h_packed_alt = t;
      const size_t n = c->h_packed_cap; c->h_packed_cap = c->h_packed_alt_cap; c->h_packed_alt_cap = n; }
    if (total > c->h_packed_cap) {
        cudaFreeHost(c->h_packed);
        c->h_packed_cap = total + total / 4 + 4096;
        CK(cudaMallocHost((void **)&c->h_packed, c->h_packed_cap));
    }
    if (total) CK(cudaMemcpyAsync(c->h_packed, c->d_packed, total, cudaMemcpyDeviceToHost, c->stream));
    if (c->profiling) cudaEventRecord(c->ev[EBCC_ST_COUNT], c->stream);
    CK(ctx_sync(c));
    for (int i = 0; i < B; ++i) {
        if (streams) streams[i] = c->h_packed + c->h_offs[i];
        if (sizes) sizes[i] = (size_t)c->h_pout[i].stream_bytes;
        if (minv) minv[i] = c->h_stat[i].minv;
        if (maxv) maxv[i] = c->h_stat[i].maxv;
        if (flags) flags[i] = c->h_stat[i].flags;
    }
    if (c->profiling) {
        for (int s = 0; s < EBCC_ST_COUNT; ++s) {
            float ms = 0;
            cudaEventElapsedTime(&ms, c->ev[s], c->ev[s + 1]);
            c->stage_ms[s] = ms;
        }
    }
    return 0;
}

extern "C" int ebcc_cuda_j2k_encode_batch(ebcc_cuda_ctx *c, const float *frames, size_t nframes, size_t H, size_t W,
                                          float base_cr, const uint8_t **streams, size_t *sizes, float *minv,
                                          float *maxv, int *flags)
{
    if (ebcc_cuda_reserve(c, H, W, nframes)) return -1;
    if (ebcc_cuda_upload(c, frames, nframes)) return -1;
    if (ebcc_cuda_j2k_encode_resident(c, nframes, base_cr)) return -1;
    return ebcc_cuda_download(c, nframes, streams, sizes, minv, maxv, flags);
}

// bytes from host memory to a caller-provided destination that may be host or device memory
extern "C" int ebcc_cuda_copy_out(void *dst, const void *src, size_t bytes)
{
    CK(cudaMemcpy(dst, src, bytes, cudaMemcpyDefault));
    return 0;
}

extern "C" void ebcc_cuda_set_profiling(ebcc_cuda_ctx *c, int enabled) { if (c) c->profiling = enabled; }
extern "C" float ebcc_cuda_stage_ms(ebcc_cuda_ctx *c, int stage) { return (c && stage >= 0 && stage < EBCC_ST_COUNT) ? c->stage_ms[stage] : -1.0f; }
extern "C" int ebcc_cuda_sync(ebcc_cuda_ctx *c) { if (!c) return -1; CK(cudaSetDevice(c->device)); CK(ctx_sync(c)); return 0; }
extern "C" unsigned long long ebcc_cuda_launch_count(ebcc_cuda_ctx *c) { return c ? c->launches : 0; }
extern "C" unsigned long long ebcc_cuda_total_launches(void) { return __atomic_load_n(&g_ebcc_launches, __ATOMIC_RELAXED); }

extern "C" int ebcc_cuda_debug_coef(ebcc_cuda_ctx *c, size_t frame, float *dst)
{
    if (!c || frame >= c->cap_frames) return -1;
    const size_t N = (size_t)c->plan.H * c->plan.W;
    CK(ctx_sync(c));
    CK(cudaMemcpy(dst, c->d_coef + frame * N, N * sizeof(float), cudaMemcpyDeviceToHost));
    return 0;
}

extern "C" int ebcc_cuda_debug_cblk(ebcc_cuda_ctx *c, size_t frame, size_t cblk, int *numbps, int *totalpasses,
                                    int *rates, double *dists, int *numpasses)
{
    if (!c || frame >= c->cap_frames || cblk >= (size_t)c->plan.ncblks) return -1;
    CblkOut o;
    CK(ctx_sync(c));
    CK(cudaMemcpy(&o, c->d_cbout + frame * c->plan.ncblks + cblk, sizeof(o), cudaMemcpyDeviceToHost));
    *numbps = o.numbps; *totalpasses = o.totalpasses; *numpasses = o.numpasses;
    for (int i = 0; i < o.totalpasses && i < J2K_MAXPASSES; ++i) { rates[i] = o.rate[i]; dists[i] = o.dist[i]; }
    return 0;
}

// measurement tap: what tier-1 moved for the resident batch -- EBCOT symbols (1 byte each, written by k_t1_cm, read by
// k_t1_mq) and MQ-coded bytes (written by k_t1_mq), summed over all code-blocks of the first nframes frames
extern "C" int ebcc_cuda_t1_totals(ebcc_cuda_ctx *c, size_t nframes, unsigned long long *symbols, unsigned long long *mq_bytes)
{
    if (!c || nframes > c->cap_frames) return -1;
    const size_t n = nframes * (size_t)c->plan.ncblks;
    CblkOut *h = (CblkOut *)malloc(n * sizeof(CblkOut));
    if (!h) return -1;
    CK(ctx_sync(c));
    CK(cudaMemcpy(h, c->d_cbout, n * sizeof(CblkOut), cudaMemcpyDeviceToHost));
    unsigned long long s = 0, b = 0;
    for (size_t i = 0; i < n; ++i)
        if (h[i].totalpasses > 0) { s += h[i].nsym_end[h[i].totalpasses - 1]; b += (unsigned long long)h[i].datalen; }
    free(h);
    *symbols = s; *mq_bytes = b;
    return 0;
}

// test tap: after ebcc_cuda_analyze(for_search = 1), checks EVERY truncation point of EVERY code-block of the first nframes
// frames: the block is decoded for real from the segment cut at rate[np - 1] and compared with what the recording +
// verdict would reconstruct.  counts[0] cuts checked, [1] flagged unproven, [2] flagged proven but different from the real
// decode (must be 0), [3] flagged unproven although equal, [4] unproven cuts found pre-decoded in the pool, [5] pool entries
// that differ from the real decode (must be 0).  Leaves the PCRD allocation of the context clobbered.
extern "C" int ebcc_cuda_debug_check_cuts(ebcc_cuda_ctx *c, size_t nframes, unsigned long long *counts)
{
    if (!c || nframes == 0 || nframes > c->cap_frames || c->ntiles > 1 || !c->for_search) return -1;
    CK(cudaSetDevice(c->device));
    const Plan *pl = &c->plan;
    const int B = (int)nframes;
    const long long njobs = (long long)pl->ncblks * B;
    unsigned long long *d_cnt = NULL;
    CK(cudaMalloc((void **)&d_cnt, 6 * sizeof(unsigned long long)));
    CK(cudaMemsetAsync(d_cnt, 0, 6 * sizeof(unsigned long long), c->stream));
    unsigned long long prev = 0;
    for (int np = 1; np < J2K_MAXPASSES; ++np) {
        EBCC_LAUNCH(k_dbg_force_np, dim3((unsigned)((njobs + 255) / 256)), dim3(256), 0, c->stream, c->d_cbout, njobs, np);
        EBCC_LAUNCH(k_jobs_from_enc, dim3((unsigned)((njobs + 255) / 256)), dim3(256), 0, c->stream, c->d_cblks, c->d_cbout, c->d_dstate, c->d_jobs,
                    pl->ncblks, B, 0, (uint2 *)NULL, (ErrStat *)NULL);
        const long long per_cta = (long long)DEC_WPC(1);
        EBCC_LAUNCH(k_t1_dec<1>, dim3((unsigned)((njobs + per_cta - 1) / per_cta)), dim3((unsigned)(DEC_WPC(1) * 32)), DEC_SMEM_BYTES(1), c->stream, c->d_cblks, c->d_order,
                    c->d_jobs, c->d_mq, pl->mq_stride, c->d_declut, c->d_iplane, (unsigned char *)NULL, c->d_scratch, (const CblkOut *)c->d_cbout,
                    c->d_dstate, c->d_decsave, pl->W, pl->H, pl->ncblks, B);
        EBCC_LAUNCH(k_dbg_compare, dim3((unsigned)pl->ncblks, (unsigned)B), dim3(128), 0, c->stream, c->d_cblks, c->d_cbout, c->d_dstate, c->d_iplane,
                    c->d_spass, c->d_scratch, d_cnt, np, pl->W, pl->H, pl->ncblks, (const int *)c->d_cut_pool);
        CK(cudaMemcpyAsync(counts, d_cnt, 6 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
        CK(ctx_sync(c));
        if (counts[0] == prev) break;          // no block has this many passes
        prev = counts[0];
    }
    cudaFree(d_cnt);
    CK(cudaGetLastError());
    return 0;
}

// test tap: SPIHT bitstream of a frame after ebcc_cuda_residual_encode
extern "C" int ebcc_cuda_debug_spiht_stream(ebcc_cuda_ctx *c, size_t frame, uint8_t *dst, size_t cap, size_t *len)
{
    if (!c || frame >= c->cap_frames) return -1;
    CK(ctx_sync(c));
    const size_t n = c->h_sp_info[frame].nbytes;
    *len = n;
    if (n > cap) return -1;
    CK(cudaMemcpy(dst, (const unsigned char *)c->sp.bits + frame * c->sp.bits_words * 4, n, cudaMemcpyDeviceToHost));
    return 0;
}

// ---- device-side timing helper for bench.py -------------------------------------------------------------
// Runs `warmup` untimed and `steps` timed passes over the resident batch.  Between passes a scratch buffer
// larger than L2 is overwritten (outside the timed events) so every pass starts with a cold L2.  Times come
// from CUDA events recorded on the context's own stream.
extern "C" int ebcc_cuda_bench_resident(ebcc_cuda_ctx *c, size_t nframes, float base_cr, int steps, int warmup,
                                        size_t flush_bytes, float *total_ms, float *stage_ms_sum)
{
    if (!c || steps <= 0) return -1;
    CK(cudaSetDevice(c->device));
    void *flush = NULL;
    if (flush_bytes) CK(cudaMalloc(&flush, flush_bytes));
    const int prev = c->profiling;
    c->profiling = 1;
    *total_ms = 0;
    for (int s = 0; s < EBCC_ST_COUNT; ++s) stage_ms_sum[s] = 0;
    for (int it = 0; it < warmup + steps; ++it) {
        if (flush) CK(cudaMemsetAsync(flush, it & 0xff, flush_bytes, c->stream));
        if (ebcc_cuda_j2k_encode_resident(c, nframes, base_cr)) { cudaFree(flush); return -1; }
        CK(ctx_sync(c));
        if (it >= warmup) {
            float ms = 0;
            CK(cudaEventElapsedTime(&ms, c->ev[EBCC_ST_MINMAX], c->ev[EBCC_ST_D2H]));
            *total_ms += ms;
            for (int s = EBCC_ST_MINMAX; s < EBCC_ST_D2H; ++s) {
                CK(cudaEventElapsedTime(&ms, c->ev[s], c->ev[s + 1]));
                stage_ms_sum[s] += ms;
            }
        }
    }
    c->profiling = prev;
    if (flush) cudaFree(flush);
    return 0;
}

// Several resident batches in flight: step it runs on ctxs[it % nctx], every context on its own stream, so that the
// tier-1 context-modelling kernel of one batch overlaps the (latency-bound) MQ kernel of another.  All launches are
// issued by the calling thread; the timed region is bracketed by events on a timing stream that every context
// stream waits on / is waited for.  flush_bytes > 0 writes a scratch buffer of that size before every step (inside
// the timed region).  total_ms: device time of the `steps` timed steps.
extern "C" int ebcc_cuda_bench_resident_pipelined(ebcc_cuda_ctx **ctxs, int nctx, size_t nframes, float base_cr, int steps, int warmup,
                                                  size_t flush_bytes, float *total_ms)
{
    if (!ctxs || nctx <= 0 || nctx > 16 || steps <= 0) return -1;
    CK(cudaSetDevice(ctxs[0]->device));
    for (int k = 0; k < nctx; ++k) ctxs[k]->crowd = nctx;
    void *flush[16] = {0};
    cudaStream_t ts;
    cudaEvent_t e0, e1, done[16];
    CK(cudaStreamCreateWithFlags(&ts, cudaStreamNonBlocking));
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    for (int k = 0; k < nctx; ++k) {
        CK(cudaEventCreateWithFlags(&done[k], cudaEventDisableTiming));
        if (flush_bytes) CK(cudaMalloc(&flush[k], flush_bytes));
    }
    int rc = 0;
    for (int phase = 0; phase < 2 && !rc; ++phase) {
        const int n = phase == 0 ? warmup : steps;
        if (phase == 1) {
            CK(cudaEventRecord(e0, ts));
            for (int k = 0; k < nctx; ++k) CK(cudaStreamWaitEvent(ctxs[k]->stream, e0, 0));
        }
        for (int it = 0; it < n && !rc; ++it) {
            ebcc_cuda_ctx *c = ctxs[it % nctx];
            if (flush[it % nctx]) CK(cudaMemsetAsync(flush[it % nctx], it & 0xff, flush_bytes, c->stream));
            rc = ebcc_cuda_j2k_encode_resident(c, nframes, base_cr);
        }
        for (int k = 0; k < nctx; ++k) { CK(cudaEventRecord(done[k], ctxs[k]->stream)); CK(cudaStreamWaitEvent(ts, done[k], 0)); }
        if (phase == 1) CK(cudaEventRecord(e1, ts));
        CK(cudaStreamSynchronize(ts));
    }
    if (!rc) CK(cudaEventElapsedTime(total_ms, e0, e1));
    for (int k = 0; k < nctx; ++k) { cudaEventDestroy(done[k]); if (flush[k]) cudaFree(flush[k]); }
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaStreamDestroy(ts);
    return rc;
}

// test tap: one zstd frame through the device decompressor (k_zstd_decompress), nothing else of the codec involved.
// Returns 0 and *out_len = bytes produced, or -2 when the decoder rejects the frame.  Allocates its own scratch.
extern "C" int ebcc_cuda_debug_zstd_decompress(int device, const uint8_t *frame, size_t n, uint8_t *out, size_t cap, size_t *out_len)
{
    if (!frame || !out || !out_len || n == 0 || n > 0x7fffffffu || cap > 0x7fffffffu) return -1;
    CK(cudaSetDevice(device));
    unsigned char *d_in = NULL, *d_out = NULL; unsigned *d_len = NULL; ZDecWork *d_work = NULL;
    const size_t stride = (n + 64 + 15) & ~(size_t)15;
    CK(cudaMalloc((void **)&d_in, stride)); CK(cudaMalloc((void **)&d_out, cap + 64)); CK(cudaMalloc((void **)&d_len, 3 * sizeof(unsigned)));
    CK(cudaMalloc((void **)&d_work, sizeof(ZDecWork)));
    CK(cudaMemset(d_in, 0, stride));
    CK(cudaMemcpy(d_in, frame, n, cudaMemcpyHostToDevice));
    const unsigned h[3] = {(unsigned)n, (unsigned)cap, 0u};
    CK(cudaMemcpy(d_len, h, sizeof(h), cudaMemcpyHostToDevice));
    EBCC_LAUNCH(k_zstd_decompress, dim3(1), dim3(32), 0, (cudaStream_t)0, (const unsigned char *)d_in, stride, (const unsigned *)d_len, d_out, cap + 64,
                (const unsigned *)(d_len + 1), d_len + 2, d_work);
    CK(cudaDeviceSynchronize());
    unsigned got = 0;
    CK(cudaMemcpy(&got, d_len + 2, sizeof(got), cudaMemcpyDeviceToHost));
    int rc = 0;
    if (got == 0xffffffffu || got > cap) rc = -2;
    else { CK(cudaMemcpy(out, d_out, got, cudaMemcpyDeviceToHost)); *out_len = got; }
    cudaFree(d_in); cudaFree(d_out); cudaFree(d_len); cudaFree(d_work);
    return rc;
}

// ---- allocation-free entry point (no counterpart in the reference) -----------
// Frame statistics (findMinMaxf + check_nan_inf) of device-resident frames for callers that must not allocate: code
// under CUDA graph capture, tight serving loops.  The FrameStat records k_minmax reduces into with atomics -- the scratch
// memory a context normally holds in its cudaMalloc'ed d_stat -- live in a file-scope __device__ array instead, which
// every GPU that loads the library keeps until exit (EBCC_NOALLOC_MAX_FRAMES x 32 bytes = 128 KiB).  In host code the
// array's name is only a stand-in for the symbol: the device address comes from cudaGetSymbolAddress, taken once per
// device with that device current.  All calls share the array, so calls that overlap on the GPU (different streams
// without ordering) race; k_stat_init clears the records a call uses before the reduction touches them.
static __device__ __align__(256) FrameStat g_noalloc_stat[EBCC_NOALLOC_MAX_FRAMES];
#define NOALLOC_MAX_DEVICES 64
static FrameStat *g_noalloc_ptr[NOALLOC_MAX_DEVICES];
static pthread_mutex_t g_noalloc_mu = PTHREAD_MUTEX_INITIALIZER;

__global__ void k_stat_export(const FrameStat *__restrict__ stat, float *minv, float *maxv, int *flags, int nframes)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < nframes) { minv[i] = stat[i].minv; maxv[i] = stat[i].maxv; flags[i] = stat[i].flags; }
}

extern "C" int ebcc_cuda_frame_stats_noalloc(const float *d_frames, size_t nframes, size_t samples_per_frame,
                                             float *d_minv, float *d_maxv, int *d_flags, void *stream)
{
    if (!d_frames || !d_minv || !d_maxv || !d_flags || nframes == 0 || samples_per_frame == 0) {
        fprintf(stderr, "[ebcc-cuda] ebcc_cuda_frame_stats_noalloc: null pointer or empty batch\n");
        return -1;
    }
    if (nframes > EBCC_NOALLOC_MAX_FRAMES || samples_per_frame > 0x7fffffffull) {
        fprintf(stderr, "[ebcc-cuda] ebcc_cuda_frame_stats_noalloc: %zu frames of %zu samples exceed the static scratch "
                        "(at most %d frames of at most 2^31-1 samples per call)\n", nframes, samples_per_frame, EBCC_NOALLOC_MAX_FRAMES);
        return -1;
    }
    int dev = -1;
    CK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) { fprintf(stderr, "[ebcc-cuda] ebcc_cuda_frame_stats_noalloc: device %d out of range\n", dev); return -1; }
    pthread_mutex_lock(&g_noalloc_mu);
    FrameStat *stat = g_noalloc_ptr[dev];
    if (!stat) {                                       // set-up, once per device, with that device current
        cudaError_t e = cudaGetSymbolAddress((void **)&stat, g_noalloc_stat);
        if (e != cudaSuccess || !stat) {
            pthread_mutex_unlock(&g_noalloc_mu);
            fprintf(stderr, "[ebcc-cuda] cudaGetSymbolAddress(g_noalloc_stat) failed on device %d: %s\n", dev, cudaGetErrorString(e));
            (void)cudaGetLastError();
            return -1;
        }
        g_noalloc_ptr[dev] = stat;
    }
    pthread_mutex_unlock(&g_noalloc_mu);
    cudaStream_t s = (cudaStream_t)stream;
    const int B = (int)nframes, N = (int)samples_per_frame;
    const dim3 gsmall((unsigned)((B + 127) / 128));
    EBCC_LAUNCH(k_stat_init, gsmall, dim3(128), 0, s, stat, B);               // the shared records start from scratch
    EBCC_LAUNCH(k_minmax, dim3(B >= 8 ? 74u : 296u, (unsigned)B), dim3(256), 0, s, d_frames, stat, N);
    EBCC_LAUNCH(k_minmax_final, gsmall, dim3(128), 0, s, d_frames, stat, N, B);
    EBCC_LAUNCH(k_stat_export, gsmall, dim3(128), 0, s, stat, d_minv, d_maxv, d_flags, B);
    CK(cudaGetLastError());
    return 0;
}
