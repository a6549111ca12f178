// k_dec.cuh -- JPEG 2000 decoder front half: tier-2 packet-header parse and EBCOT tier-1 decode.
//
//   k_t2_parse : one CTA per frame, lane 0 walks the codestream (main header check, SOT/SOD, the
//                six LRCP packets of the single tile: tag trees, pass counts, segment lengths;
//                t2.c:1050-1330, tgt.c:307-344, bio.c) and emits one DecJob per code-block.
//   k_t1_dec   : one warp per code-block, lane 0 runs the serial MQ decoder (mqc_inl.h:100-175,
//                mqc.c:420-460) and the three coding passes (t1.c:412-478, 766-812, 1118-1414) on a
//                packed neighbourhood-flag array in shared memory; sample values are accumulated
//                straight into the frame's integer coefficient plane with fire-and-forget atomics
//                (all updates of one sample commute), so no per-block data buffer is needed.
//
// The same k_t1_dec serves ebcc_decode (jobs from k_t2_parse) and the compression-ratio search of
// ebcc_encode (jobs built from the encoder's own PCRD result by k_jobs_from_enc), which therefore
// sees exactly what a real decoder sees -- including the effect of truncated MQ segments.
#pragma once
#include "common.cuh"
#include "k_t1.cuh"
#include "k_t2.cuh"

struct DecJob { int np; int numbps; unsigned off; int len; };

// decoder-side LUTs: zero-coding context by (orient, 9-bit neighbourhood), filled by the host with
// the same Annex-D rule the encoder evaluates arithmetically (t1_zc_ctx)
struct DecLuts {
    unsigned char zc[4 * 512];
    unsigned mq[94];           // per (state, mps): qe | next-if-mps << 16 | next-if-lps << 24
};

#define DEC_FW 66
#define DEC_FWORDS (18 * DEC_FW)
#define DEC_WARPS 4

// flag word of one stripe column: bits 0-17 significance of the 3x6 neighbourhood
// (bit (r+1)*3 + (dx+1), r = -1..4, dx = -1..1), bits 18-23 sign of own column rows -1..4,
// bits 24-27 pi (coded in the significance pass), bits 28-31 mu (already refined)
#define DF_SIG(ci) (1u << ((ci) * 3 + 4))
#define DF_NB(ci) (0x1efu << ((ci) * 3))
#define DF_CHI(r) (1u << (18 + (r) + 1))
#define DF_PI(ci) (1u << (24 + (ci)))
#define DF_MU(ci) (1u << (28 + (ci)))

struct MqDec {
    unsigned a, c, ct;
    int bp, len;
    unsigned cur;                  // byte at bp
    const unsigned char *data;
};

__device__ __forceinline__ unsigned mqd_byte(const MqDec &q, int i)
{
    return (i >= 0 && i < q.len) ? (unsigned)q.data[i] : 0xffu;
}

__device__ __forceinline__ void mqd_bytein(MqDec &q)
{
    const unsigned nx = mqd_byte(q, q.bp + 1);
    if (q.cur == 0xffu) {
        if (nx > 0x8fu) { q.c += 0xff00u; q.ct = 8; }
        else { q.bp++; q.cur = nx; q.c += nx << 9; q.ct = 7; }
    } else { q.bp++; q.cur = nx; q.c += nx << 8; q.ct = 8; }
}

__device__ __forceinline__ void mqd_init(MqDec &q, const unsigned char *data, int len)
{
    q.data = data; q.len = len; q.bp = 0;
    q.cur = mqd_byte(q, 0);
    q.c = q.cur << 16;
    mqd_bytein(q);
    q.c <<= 7; q.ct -= 7; q.a = 0x8000u;
}

__device__ __forceinline__ void mqd_renorm(MqDec &q)
{
    do {
        if (q.ct == 0) mqd_bytein(q);
        int k = __clz((int)q.a) - 16;          // shifts still needed
        if ((unsigned)k > q.ct) k = (int)q.ct;
        q.a <<= k; q.c <<= k; q.ct -= (unsigned)k;
    } while (q.a < 0x8000u);
}

__device__ __forceinline__ unsigned mqd_decode(MqDec &q, unsigned char *ctx, const unsigned *tab)
{
    const unsigned st = *ctx, e = tab[st], qe = e & 0xffffu, mps = st & 1u;
    unsigned d;
    q.a -= qe;
    if ((q.c >> 16) < qe) {
        if (q.a < qe) { d = mps; *ctx = (unsigned char)((e >> 16) & 0xffu); }
        else { d = mps ^ 1u; *ctx = (unsigned char)(e >> 24); }
        q.a = qe;
        mqd_renorm(q);
    } else {
        q.c -= qe << 16;
        if ((q.a & 0x8000u) == 0) {
            if (q.a < qe) { d = mps ^ 1u; *ctx = (unsigned char)(e >> 24); }
            else { d = mps; *ctx = (unsigned char)((e >> 16) & 0xffu); }
            mqd_renorm(q);
        } else d = mps;
    }
    return d;
}

struct T1DecState {
    unsigned *F;                 // flag words of this block
    unsigned char *cx;           // 19 context states
    const unsigned char *zc;     // zero-coding LUT of this orientation
    const unsigned *tab;
    int *plane;                  // top-left of the block inside the integer coefficient plane
    int W;
};

// sample (stripe word index p, row ci) becomes significant with sign s: update the neighbourhood bits
__device__ __forceinline__ void t1d_set_sig(T1DecState &S, int p, int ci, unsigned s, unsigned &f)
{
    f |= DF_SIG(ci) | (s ? DF_CHI(ci) : 0u);
    S.F[p - 1] |= 1u << ((ci + 1) * 3 + 2);
    S.F[p + 1] |= 1u << ((ci + 1) * 3 + 0);
    if (ci == 0) {
        S.F[p - DEC_FW] |= (1u << 16) | (s ? DF_CHI(4) : 0u);
        S.F[p - DEC_FW - 1] |= 1u << 17;
        S.F[p - DEC_FW + 1] |= 1u << 15;
    }
    if (ci == 3) {
        S.F[p + DEC_FW] |= (1u << 1) | (s ? DF_CHI(-1) : 0u);
        S.F[p + DEC_FW - 1] |= 1u << 2;
        S.F[p + DEC_FW + 1] |= 1u << 0;
    }
}

// decode the sign of (p, ci); returns 1 for negative
__device__ __forceinline__ unsigned t1d_sign(T1DecState &S, MqDec &q, int p, int ci, unsigned f)
{
    const unsigned fl = S.F[p - 1], fr = S.F[p + 1];
    const int sw = (f >> ((ci + 1) * 3)) & 1, se = (f >> ((ci + 1) * 3 + 2)) & 1;
    const int sn = (f >> (ci * 3 + 1)) & 1, ss = (f >> ((ci + 2) * 3 + 1)) & 1;
    const int nw = (fl >> (18 + ci + 1)) & 1, ne = (fr >> (18 + ci + 1)) & 1;
    const int nn = (f >> (18 + ci)) & 1, ns = (f >> (18 + ci + 2)) & 1;
    const int hc = t1_contrib(se, ne, sw, nw), vc = t1_contrib(sn, nn, ss, ns);
    const int e = t1_sc_lut(hc, vc);
    const unsigned v = mqd_decode(q, S.cx + CTX_SC + (e & 7), S.tab);
    return v ^ (unsigned)(e >> 3);
}

__device__ void t1d_clnpass(T1DecState &S, MqDec &q, int w, int h, int bpno)
{
    const int oneplushalf = 3 << bpno;
    for (int s = 0, y0 = 0; y0 < h; ++s, y0 += 4) {
        const int lim = min(4, h - y0);
        for (int x = 0; x < w; ++x) {
            const int p = (s + 1) * DEC_FW + x + 1;
            unsigned f = S.F[p];
            int first = 0, partial = 0;
            if (lim == 4 && f == 0) {
                if (!mqd_decode(q, S.cx + CTX_AGG, S.tab)) continue;
                unsigned r = mqd_decode(q, S.cx + CTX_UNI, S.tab);
                r = (r << 1) | mqd_decode(q, S.cx + CTX_UNI, S.tab);
                first = (int)r; partial = 1;
            }
            for (int ci = first; ci < lim; ++ci) {
                if (!partial && (f & (DF_SIG(ci) | DF_PI(ci)))) continue;
                if (!partial) {
                    const unsigned ctx = S.zc[(f >> (ci * 3)) & 0x1ffu];
                    if (!mqd_decode(q, S.cx + CTX_ZC + ctx, S.tab)) continue;
                }
                partial = 0;
                const unsigned sg = t1d_sign(S, q, p, ci, f);
                atomicAdd(S.plane + (size_t)(y0 + ci) * S.W + x, sg ? -oneplushalf : oneplushalf);
                t1d_set_sig(S, p, ci, sg, f);
            }
            S.F[p] = f & ~(0xfu << 24);
        }
    }
}

__device__ void t1d_sigpass(T1DecState &S, MqDec &q, int w, int h, int bpno)
{
    const int oneplushalf = 3 << bpno;
    for (int s = 0, y0 = 0; y0 < h; ++s, y0 += 4) {
        const int lim = min(4, h - y0);
        for (int x = 0; x < w; ++x) {
            const int p = (s + 1) * DEC_FW + x + 1;
            unsigned f = S.F[p];
            if (f == 0) continue;
            for (int ci = 0; ci < lim; ++ci) {
                if ((f & (DF_SIG(ci) | DF_PI(ci))) == 0 && (f & DF_NB(ci)) != 0) {
                    const unsigned ctx = S.zc[(f >> (ci * 3)) & 0x1ffu];
                    if (mqd_decode(q, S.cx + CTX_ZC + ctx, S.tab)) {
                        const unsigned sg = t1d_sign(S, q, p, ci, f);
                        atomicAdd(S.plane + (size_t)(y0 + ci) * S.W + x, sg ? -oneplushalf : oneplushalf);
                        t1d_set_sig(S, p, ci, sg, f);
                    }
                    f |= DF_PI(ci);
                }
            }
            S.F[p] = f;
        }
    }
}

__device__ void t1d_refpass(T1DecState &S, MqDec &q, int w, int h, int bpno)
{
    const int poshalf = 1 << bpno;
    for (int s = 0, y0 = 0; y0 < h; ++s, y0 += 4) {
        const int lim = min(4, h - y0);
        for (int x = 0; x < w; ++x) {
            const int p = (s + 1) * DEC_FW + x + 1;
            unsigned f = S.F[p];
            if (f == 0) continue;
            for (int ci = 0; ci < lim; ++ci) {
                if ((f & (DF_SIG(ci) | DF_PI(ci))) == DF_SIG(ci)) {
                    const int ctx = (f & DF_MU(ci)) ? CTX_MAG + 2 : ((f & DF_NB(ci)) ? CTX_MAG + 1 : CTX_MAG);
                    const unsigned v = mqd_decode(q, S.cx + ctx, S.tab);
                    const unsigned neg = (f >> (18 + ci + 1)) & 1u;
                    atomicAdd(S.plane + (size_t)(y0 + ci) * S.W + x, (v ^ neg) ? poshalf : -poshalf);
                    f |= DF_MU(ci);
                }
            }
            S.F[p] = f;
        }
    }
}

// grid = ceil(ncblks * nframes / DEC_WARPS) CTAs of DEC_WARPS warps; plane must be zero on entry
__global__ void __launch_bounds__(DEC_WARPS * 32)
k_t1_dec(const CblkDesc *__restrict__ cblks, const DecJob *__restrict__ jobs, const unsigned char *__restrict__ pool,
         size_t pool_stride, const DecLuts *__restrict__ g_luts, int *iplane, int W, int H, int ncblks, int nframes)
{
    __shared__ unsigned F[DEC_WARPS][DEC_FWORDS];
    __shared__ unsigned char cxs[DEC_WARPS][32];
    __shared__ unsigned tab[94];
    __shared__ unsigned char zc[4 * 512];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int i = tid; i < 94; i += blockDim.x) tab[i] = g_luts->mq[i];
    for (int i = tid; i < 4 * 512; i += blockDim.x) zc[i] = g_luts->zc[i];
    const long long job = (long long)blockIdx.x * DEC_WARPS + warp;
    const bool live = job < (long long)ncblks * nframes;
    int frame = 0, cbi = 0;
    DecJob jb; jb.np = 0; jb.numbps = 0; jb.off = 0; jb.len = 0;
    if (live) {
        frame = (int)(job / ncblks); cbi = (int)(job - (long long)frame * ncblks);
        jb = jobs[job];
    }
    const bool work = live && jb.np > 0 && jb.numbps > 0;
    if (work) for (int i = lane; i < DEC_FWORDS; i += 32) F[warp][i] = 0;
    __syncthreads();
    if (!work || lane != 0) return;

    const CblkDesc cb = cblks[cbi];
    unsigned char *cx = cxs[warp];
    for (int i = 0; i < 19; ++i) cx[i] = 0;
    cx[CTX_UNI] = 46 << 1; cx[CTX_AGG] = 3 << 1; cx[CTX_ZC] = 4 << 1;
    T1DecState S;
    S.F = F[warp]; S.cx = cx; S.zc = zc + cb.orient * 512; S.tab = tab; S.W = W;
    S.plane = iplane + (size_t)frame * H * W + (size_t)cb.ty * W + cb.tx;
    MqDec q;
    mqd_init(q, pool + (size_t)frame * pool_stride + jb.off, jb.len);
    int bpno = jb.numbps - 1, passtype = 2;
    for (int p = 0; p < jb.np && bpno >= 0; ++p) {
        if (passtype == 0) t1d_sigpass(S, q, cb.w, cb.h, bpno);
        else if (passtype == 1) t1d_refpass(S, q, cb.w, cb.h, bpno);
        else t1d_clnpass(S, q, cb.w, cb.h, bpno);
        if (++passtype == 3) { passtype = 0; bpno--; }
    }
}

// decode jobs for the encoder's own allocation (compression-ratio search): block k of a frame is
// decoded from its MQ pool segment truncated at rate[numpasses-1]
__global__ void k_jobs_from_enc(const CblkDesc *__restrict__ cblks, const CblkOut *__restrict__ cbout, DecJob *jobs,
                                int ncblks, int nframes)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)ncblks * nframes) return;
    const int k = (int)(i % ncblks);
    const CblkOut &c = cbout[i];
    DecJob j;
    j.np = c.numpasses; j.numbps = c.numbps; j.off = cblks[k].mq_off;
    j.len = c.numpasses > 0 ? c.rate[c.numpasses - 1] : 0;
    jobs[i] = j;
}

// ---- tier-2 parse ---------------------------------------------------------------------------------
struct BitIn {
    const unsigned char *p; int pos, end; unsigned buf; int ct; int err;
};
__device__ __forceinline__ void bi_bytein(BitIn &b)
{
    b.buf = (b.buf << 8) & 0xffffu;
    b.ct = b.buf == 0xff00u ? 7 : 8;
    if (b.pos < b.end) b.buf |= b.p[b.pos++]; else b.err = 1;
}
__device__ __forceinline__ unsigned bi_bit(BitIn &b)
{
    if (b.ct == 0) bi_bytein(b);
    b.ct--;
    return (b.buf >> b.ct) & 1u;
}
__device__ __forceinline__ unsigned bi_read(BitIn &b, int n)
{
    unsigned v = 0;
    for (int i = n - 1; i >= 0; --i) v |= bi_bit(b) << i;
    return v;
}

struct TgtState { int *val, *low; };

__device__ int tgt_decode(BitIn &b, const BandDesc &bd, TgtState t, int x, int y, int threshold)
{
    int low = 0, node = 0;
    for (int l = bd.tt_nlev - 1; l >= 0; --l) {
        node = bd.tt_base + bd.tt_off[l] + (y >> l) * bd.tt_w[l] + (x >> l);
        if (low > t.low[node]) t.low[node] = low; else low = t.low[node];
        while (low < threshold && low < t.val[node]) {
            if (bi_bit(b)) t.val[node] = low; else ++low;
            if (b.err) return 0;
        }
        t.low[node] = low;
    }
    return t.val[node] < threshold;
}

#define PARSE_OK 0
#define PARSE_BAD 1

// grid = nframes, block = 32.  streams: [frame][stream_stride] bytes, lens[frame] valid bytes.
// jobs[frame][ncblks]; job.off is relative to the frame's stream.  status[frame] = PARSE_*.
__global__ void k_t2_parse(const Plan *__restrict__ plan, const CblkDesc *__restrict__ cblks,
                           const unsigned char *__restrict__ streams, size_t stream_stride,
                           const int *__restrict__ lens, DecJob *jobs, int *status)
{
    EBCC_DYN_SMEM(int, sm);
    const int frame = blockIdx.x, ncblks = plan->ncblks, nnodes = plan->nnodes;
    DecJob *jb = jobs + (size_t)frame * ncblks;
    for (int k = threadIdx.x; k < ncblks; k += blockDim.x) { DecJob z; z.np = 0; z.numbps = 0; z.off = 0; z.len = 0; jb[k] = z; }
    TgtState incl, imsb;
    incl.val = sm; incl.low = sm + nnodes; imsb.val = sm + 2 * nnodes; imsb.low = sm + 3 * nnodes;
    for (int i = threadIdx.x; i < nnodes; i += blockDim.x) { incl.val[i] = 999; incl.low[i] = 0; imsb.val[i] = 999; imsb.low[i] = 0; }
    __syncthreads();
    if (threadIdx.x != 0) return;
    const unsigned char *s = streams + (size_t)frame * stream_stride;
    const int n = lens[frame];
    status[frame] = PARSE_BAD;
    if (n <= 0) { status[frame] = PARSE_OK; return; }         // nothing to decode (constant field)
    // main header: SOC SIZ COD QCD must be what this codec writes for (H, W)
    const int fixed = plan->main_header_len - 39;
    if (n < fixed + 14) return;
    for (int i = 0; i < fixed; ++i) if (s[i] != plan->main_header[i]) return;
    int pos = fixed;
    for (;;) {                                                  // skip COM and any other main-header segment
        if (pos + 4 > n || s[pos] != 0xff) return;
        if (s[pos + 1] == 0x90) break;
        pos += 2 + ((s[pos + 2] << 8) | s[pos + 3]);
    }
    if (pos + 14 > n) return;
    const unsigned psot = ((unsigned)s[pos + 6] << 24) | ((unsigned)s[pos + 7] << 16) | ((unsigned)s[pos + 8] << 8) | s[pos + 9];
    if (s[pos + 4] != 0 || s[pos + 5] != 0) return;             // single tile
    int tp_end = psot ? pos + (int)psot : n - 2;
    if (tp_end > n) return;
    pos += 12;
    if (s[pos] != 0xff || s[pos + 1] != 0x93) return;
    pos += 2;
    for (int r = 0; r < J2K_NRES; ++r) {
        BitIn b; b.p = s; b.pos = pos; b.end = tp_end; b.buf = 0; b.ct = 0; b.err = 0;
        const int f0 = plan->res_first[r], cnt = plan->res_count[r];
        if (!bi_bit(b)) {
            if (b.err) return;
            if ((b.buf & 0xffu) == 0xffu) bi_bytein(b);
            pos = b.pos;
            continue;
        }
        for (int k = f0; k < f0 + cnt; ++k) {
            const CblkDesc &cb = cblks[k];
            const BandDesc &bd = plan->bands[cb.band];
            if (!tgt_decode(b, bd, incl, cb.gx, cb.gy, 1)) { if (b.err) return; continue; }
            int i = 0;
            while (!tgt_decode(b, bd, imsb, cb.gx, cb.gy, i)) { if (b.err || i > 64) return; ++i; }
            DecJob j;
            j.numbps = bd.numbps + 1 - i;
            int np;
            if (!bi_bit(b)) np = 1;
            else if (!bi_bit(b)) np = 2;
            else { unsigned v = bi_read(b, 2); if (v != 3) np = 3 + (int)v; else { v = bi_read(b, 5); if (v != 31) np = 6 + (int)v; else np = 37 + (int)bi_read(b, 7); } }
            int inc = 0;
            while (bi_bit(b)) { if (b.err || ++inc > 32) return; }
            const int nbits = 3 + inc + t2_floorlog2(np);
            if (nbits > 32) return;
            j.np = np; j.len = (int)bi_read(b, nbits); j.off = 0;
            if (b.err || j.len < 0) return;
            jb[k] = j;
        }
        if ((b.buf & 0xffu) == 0xffu) bi_bytein(b);              // opj_bio_inalign
        if (b.err) return;
        pos = b.pos;
        for (int k = f0; k < f0 + cnt; ++k) {
            if (jb[k].np == 0) continue;
            if (jb[k].len > tp_end - pos) return;
            jb[k].off = (unsigned)pos; pos += jb[k].len;
        }
    }
    status[frame] = PARSE_OK;
}
