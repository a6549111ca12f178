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

// mode 4 = the cut was decoded ahead of time into the pool of unproven cuts, entry `pad` (k_t1_dec_cuts)
// mode: 0 = nothing to do, 1 = decode into the recording plane (iplane + pass-of-significance plane), continuing
//       the block's saved decoder state if it has one,
//       2 = decode into the scratch plane (a truncation the recording cannot reproduce, see DecState),
//       3 = the scratch plane already holds exactly this decode (same block, same pass count)
struct DecJob { int np; int numbps; unsigned off; int len; int mode; int pad; };

// T1CutState (common.cuh): what the recording of a block holds.  Two producers: the encoder itself (k_t1_cm writes the
// recording, k_t1_mq the verdicts -- the product path) and, for cross-checking, this decoder reading the COMPLETE
// segment with truncation shadows (mode 1 jobs, EBCC_DECODER_RECORDING=1).
typedef T1CutState DecState;

// per-block save area of a recording (32-bit words): flag words, then the serial decoder's registers
#define DEC_SAVE_MQ DEC_FWORDS_C
#define DEC_SAVE_WORDS 1280

// decoder-side LUTs: zero-coding context by (orient, 9-bit neighbourhood), filled by the host with
// the same Annex-D rule the encoder evaluates arithmetically (t1_zc_ctx)
struct DecLuts {
    unsigned char zc[4 * 512];
    unsigned mq[94];           // per (state, mps) pair: qe | pair after an MPS << 16 | pair after an LPS << 23 | mps << 30
};

#define DEC_FW 66
#define DEC_FWORDS (18 * DEC_FW)
#define DEC_FWORDS_C (18 * 66)
#define DEC_DEFAULT_LANES 2
#define DEC_NSHADOW 3

// flag word of one stripe column: bits 0-17 significance of the 3x6 neighbourhood
// (bit (r+1)*3 + (dx+1), r = -1..4, dx = -1..1), bits 18-23 sign of own column rows -1..4,
// bits 24-27 pi (coded in the significance pass), bits 28-31 mu (already refined)
#define DF_SIG(ci) (1u << ((ci) * 3 + 4))
#define DF_NB(ci) (0x1efu << ((ci) * 3))
#define DF_CHI(r) (1u << (18 + (r) + 1))
#define DF_PI(ci) (1u << (24 + (ci)))
#define DF_MU(ci) (1u << (28 + (ci)))

// byte-consuming half of the MQ decoder (C register, bit counter, read position) over a segment of len bytes
// followed by the synthetic 0xFF 0xFF terminator of opj_mqc_init_dec
struct MqCore { unsigned c, ct; int bp, len; unsigned cur; };

__device__ __forceinline__ unsigned mqc_byte(const MqCore &k, const unsigned char *data, int i)
{
    return (i >= 0 && i < k.len) ? (unsigned)data[i] : 0xffu;
}
__device__ __forceinline__ void mqc_bytein(MqCore &k, const unsigned char *data)
{
    const unsigned nx = mqc_byte(k, data, k.bp + 1);
    if (k.cur == 0xffu) {
        if (nx > 0x8fu) { k.c += 0xff00u; k.ct = 8; }
        else { k.bp++; k.cur = nx; k.c += nx << 9; k.ct = 7; }
    } else { k.bp++; k.cur = nx; k.c += nx << 8; k.ct = 8; }
}
// n renormalisation shifts
__device__ __forceinline__ void mqc_shift(MqCore &k, const unsigned char *data, int n)
{
    while (n > 0) {
        if (k.ct == 0) mqc_bytein(k, data);
        const int s = n < (int)k.ct ? n : (int)k.ct;
        k.c <<= s; k.ct -= (unsigned)s; n -= s;
    }
}

// Truncation shadows (compression-ratio search only), one set per warp in shared memory: shadow j replays the
// byte-consuming half of the decoder on the segment cut at trunc[j] = rate[pass + j]; it forks from the main decoder the
// first time a byte at or beyond the cut is read.  Everything that touches them is out of line, so the per-symbol code
// of the (eight times inlined) decoder stays small: the kernel is bound by instruction fetch otherwise.
struct MqShadows {
    MqCore sh[DEC_NSHADOW];
    int trunc[DEC_NSHADOW];        // cut length of slot j, < 0: slot unused
    unsigned sh_bad;               // bit j: slot j diverged (or could not be verified)
};

struct MqDec {
    unsigned a;
    MqCore m;                      // the decoder proper
    const unsigned char *data;
    MqShadows *sw;                 // nullptr: plain decode
    unsigned sh_on;                // bit j: slot j forked
    int fork_at;                   // smallest armed cut that has not forked yet (INT_MAX: none)
    int maxread;                   // highest byte index the main decoder has looked at
};

__device__ __forceinline__ int mqd_next_fork(const MqShadows *sw, unsigned sh_on)
{
    int f = 0x7fffffff;
#pragma unroll
    for (int j = 0; j < DEC_NSHADOW; ++j)
        if (sw->trunc[j] >= 0 && !((sh_on >> j) & 1u) && sw->trunc[j] < f) f = sw->trunc[j];
    return f;
}

// the main decoder (state m, about to read byte idx) reaches one or more cuts: fork their shadows.
// nrem: renormalisation shifts of the current symbol still to be applied (a shadow forked in the middle of a
// renormalisation has to finish it on its own).  Returns the new sh_on.
__device__ __noinline__ unsigned mqd_fork(MqShadows *sw, unsigned sh_on, MqCore m, const unsigned char *data, int idx, int nrem)
{
#pragma unroll
    for (int j = 0; j < DEC_NSHADOW; ++j)
        if (sw->trunc[j] >= 0 && !((sh_on >> j) & 1u) && idx >= sw->trunc[j]) {
            MqCore k = m;
            k.len = sw->trunc[j]; sh_on |= 1u << j;
            // the shadow's current byte may already lie beyond its cut (a cut inside the decoder's look-ahead)
            if (k.bp >= sw->trunc[j]) sw->sh_bad |= 1u << j;
            mqc_bytein(k, data);
            mqc_shift(k, data, nrem);
            sw->sh[j] = k;
        }
    return sh_on;
}

// one decision of the main decoder replayed on every forked shadow
__device__ __noinline__ void mqd_shadow_step(MqShadows *sw, unsigned sh_on, unsigned qe, unsigned a0, unsigned mps, unsigned d,
                                             int nshift, const unsigned char *data)
{
#pragma unroll
    for (int j = 0; j < DEC_NSHADOW; ++j)
        if ((sh_on >> j) & 1u) {
            MqCore k = sw->sh[j];
            unsigned ds;
            if ((k.c >> 16) < qe) ds = (a0 < qe) ? mps : (mps ^ 1u);
            else { k.c -= qe << 16; ds = ((a0 & 0x8000u) == 0) ? ((a0 < qe) ? (mps ^ 1u) : mps) : mps; }
            if (ds != d) sw->sh_bad |= 1u << j;
            else if (!((sw->sh_bad >> j) & 1u)) mqc_shift(k, data, nshift);
            sw->sh[j] = k;
        }
}

// end of pass p: slot 0 holds the verdict of "stop after pass p" (not forked => the decoder never looked beyond the
// cut); retire it, shift the others down, arm a new last slot with cut tn (< 0: none).  Returns the verdict (1 = bad).
__device__ __noinline__ unsigned mqd_retire(MqShadows *sw, unsigned &sh_on, int tn, int maxread)
{
    const unsigned bad = sw->sh_bad & 1u;
#pragma unroll
    for (int j = 0; j + 1 < DEC_NSHADOW; ++j) { sw->sh[j] = sw->sh[j + 1]; sw->trunc[j] = sw->trunc[j + 1]; }
    sh_on >>= 1; sw->sh_bad >>= 1;
    sw->trunc[DEC_NSHADOW - 1] = tn;
    if (tn >= 0 && tn <= maxread) { sh_on |= 1u << (DEC_NSHADOW - 1); sw->sh_bad |= 1u << (DEC_NSHADOW - 1); }
    return bad;
}

__device__ __forceinline__ void mqd_main_bytein(MqDec &q, int nrem)
{
    const int idx = q.m.bp + 1;
    if (idx > q.maxread) q.maxread = idx;
    if (idx >= q.fork_at) {
        q.sh_on = mqd_fork(q.sw, q.sh_on, q.m, q.data, idx, nrem);
        q.fork_at = mqd_next_fork(q.sw, q.sh_on);
    }
    mqc_bytein(q.m, q.data);
}

__device__ __forceinline__ void mqd_init(MqDec &q, const unsigned char *data, int len, MqShadows *sw)
{
    q.data = data; q.m.len = len; q.m.bp = 0;
    q.m.cur = mqc_byte(q.m, data, 0);
    q.m.c = q.m.cur << 16;
    q.m.ct = 0;
    q.sw = sw; q.sh_on = 0; q.fork_at = 0x7fffffff; q.maxread = 1;
    if (sw) {
        sw->sh_bad = 0;
#pragma unroll
        for (int j = 0; j < DEC_NSHADOW; ++j) sw->trunc[j] = -1;
    }
    mqc_bytein(q.m, data);
    q.m.c <<= 7; q.m.ct -= 7; q.a = 0x8000u;
}

// a context is a copy of its pair's table word, so one shared-memory load brings qe, mps and both successors
// SH: truncation shadows may be armed (the cross-check recordings of the compression-ratio search); a plain decode
// (ebcc_decode, the pool of unproven cuts) compiles without every shadow test
template <bool SH>
__device__ __forceinline__ unsigned mqd_decode(MqDec &q, unsigned *ctx, mq_saddr tab)
{
    // Branch-free decision (the serial chain pays for every resolved branch): with lz = "code value in the lower
    // sub-interval" and sw = "sub-intervals exchanged" (a - qe < qe), the decoded bit is mps ^ (lz != sw), the new
    // interval is qe (lz) or a - qe, and a renormalisation happens exactly when that interval lost its top bit --
    // mqc_inl.h:100-175 (opj_mqc_mpsexchange / lpsexchange / decode) folded into selects.
    const unsigned e = *ctx, qe = e & 0xffffu, mps = e >> 30;
    const unsigned a0 = q.a - qe;
    const bool lz = (q.m.c >> 16) < qe;
    const unsigned flip = (unsigned)(lz != (a0 < qe));
    const unsigned d = mps ^ flip;
    const unsigned a1 = lz ? qe : a0;
    q.m.c -= lz ? 0u : (qe << 16);
    const int nshift = __clz((int)a1) - 16;                  // 0 <=> no renormalisation (a1 keeps bit 15)
    if (SH && q.sh_on) mqd_shadow_step(q.sw, q.sh_on, qe, a0, mps, d, nshift, q.data);
    q.a = a1 << nshift;
    if (nshift) {
        // the table by its 32-bit shared address held in a register: the compiler otherwise re-derives the shared window
        // base (S2UR SR_CgaCtaId ...) in front of this load, on the chain of the next decision with the same context
        *ctx = mq_ld32(tab + (mq_saddr)(((flip ? (e >> 23) : (e >> 16)) & 127u) * 4u));
        int n = nshift;
        if (!SH && n <= (int)q.m.ct) { q.m.c <<= n; q.m.ct -= (unsigned)n; }      // the usual case: the bits are in the register
        else
            while (n > 0) {
                // main decoder: shifts, with fork checks at every byte read when shadows ride along
                if (q.m.ct == 0) { if (SH) mqd_main_bytein(q, n); else mqc_bytein(q.m, q.data); }
                const int s = n < (int)q.m.ct ? n : (int)q.m.ct;
                q.m.c <<= s; q.m.ct -= (unsigned)s; n -= s;
            }
    }
    return d;
}

struct T1DecState {
    unsigned *F;                 // flag words of this block
    unsigned char *F2;           // candidate bytes of this block (sig4 | nb4 << 4)
    unsigned *cx;                // 19 contexts (table words)
    mq_saddr zc;                 // zero-coding LUT of this orientation (shared address)
    mq_saddr tab;                // the MQ state table (shared address)
    int *plane;                  // top-left of the block inside the integer coefficient plane
    unsigned char *spass;        // top-left of the block inside the pass-of-significance plane, or nullptr
    int W, pass;
};

// Second word per stripe column (one byte), kept next to the flag word so that a pass finds the rows it has to visit with
// two independent loads and three logic operations instead of testing the four rows one by one:
//   bits 0-3 sig4 (row significant), 4-7 nb4 (row has a significant neighbour); pi4 (row coded in this plane's
//   significance pass) comes from the flag word's top byte.  All bits zero <=> flag word zero.  One byte, not two, per
//   column: with 5 bytes per column eight blocks take 47.5 KB and FOUR CTAs fit an SM (32 serial chains instead of 24).
#define D2_SIG(f2) ((f2) & 0xfu)
#define D2_NB(f2) (((f2) >> 4) & 0xfu)
#define DF_PI4(f) (((f) >> 24) & 0xfu)
__device__ __forceinline__ unsigned d2_from_flags(unsigned f)
{
    unsigned r = 0;
    for (int ci = 0; ci < 4; ++ci) {
        if (f & DF_SIG(ci)) r |= 1u << ci;
        if (f & DF_NB(ci)) r |= 0x10u << ci;
    }
    return r;
}

// sample (stripe word index p, row ci) becomes significant with sign s: update the neighbourhood bits
__device__ __forceinline__ void t1d_set_sig(T1DecState &S, int p, int ci, unsigned s, unsigned &f, unsigned &f2)
{
    f |= DF_SIG(ci) | (s ? DF_CHI(ci) : 0u);
    f2 |= (1u << ci) | ((((5u << ci) >> 1) & 0xfu) << 4);            // own column: rows ci - 1 and ci + 1 gain a neighbour
    const unsigned char side = (unsigned char)((((7u << ci) >> 1) & 0xfu) << 4);   // left / right column: rows ci - 1 .. ci + 1
    S.F[p - 1] |= 1u << ((ci + 1) * 3 + 2);
    S.F[p + 1] |= 1u << ((ci + 1) * 3 + 0);
    S.F2[p - 1] |= side;
    S.F2[p + 1] |= side;
    if (ci == 0) {
        S.F[p - DEC_FW] |= (1u << 16) | (s ? DF_CHI(4) : 0u);
        S.F[p - DEC_FW - 1] |= 1u << 17;
        S.F[p - DEC_FW + 1] |= 1u << 15;
        S.F2[p - DEC_FW] |= 0x80; S.F2[p - DEC_FW - 1] |= 0x80; S.F2[p - DEC_FW + 1] |= 0x80;      // row 3 of the stripe above
    }
    if (ci == 3) {
        S.F[p + DEC_FW] |= (1u << 1) | (s ? DF_CHI(-1) : 0u);
        S.F[p + DEC_FW - 1] |= 1u << 2;
        S.F[p + DEC_FW + 1] |= 1u << 0;
        S.F2[p + DEC_FW] |= 0x10; S.F2[p + DEC_FW - 1] |= 0x10; S.F2[p + DEC_FW + 1] |= 0x10;      // row 0 of the stripe below
    }
}

// decode the sign of (p, ci); returns 1 for negative
template <bool SH>
__device__ __forceinline__ unsigned t1d_sign(T1DecState &S, MqDec &q, int p, int ci, unsigned f)
{
    const unsigned fl = S.F[p - 1], fr = S.F[p + 1];
    const int sw = (f >> ((ci + 1) * 3)) & 1, se = (f >> ((ci + 1) * 3 + 2)) & 1;
    const int sn = (f >> (ci * 3 + 1)) & 1, ss = (f >> ((ci + 2) * 3 + 1)) & 1;
    const int nw = (fl >> (18 + ci + 1)) & 1, ne = (fr >> (18 + ci + 1)) & 1;
    const int nn = (f >> (18 + ci)) & 1, ns = (f >> (18 + ci + 2)) & 1;
    const int hc = t1_contrib(se, ne, sw, nw), vc = t1_contrib(sn, nn, ss, ns);
    const int e = t1_sc_lut(hc, vc);
    const unsigned v = mqd_decode<SH>(q, S.cx + CTX_SC + (e & 7), S.tab);
    return v ^ (unsigned)(e >> 3);
}

// o = offset of the sample inside the plane (the passes keep a per-stripe row base and add x and ci * W: one multiply-add
// per sample instead of a 64-bit address chain)
__device__ __forceinline__ void t1d_new_sig(T1DecState &S, int o, unsigned sg, int oneplushalf)
{
    atomicAdd(S.plane + o, sg ? -oneplushalf : oneplushalf);
    if (S.spass) S.spass[o] = (unsigned char)S.pass;
}

template <bool SH>
__device__ void t1d_clnpass(T1DecState &S, MqDec &q, int w, int h, int bpno)
{
    const int oneplushalf = 3 << bpno;
    for (int s = 0, y0 = 0; y0 < h; ++s, y0 += 4) {
        const int lim = min(4, h - y0);
        const unsigned limmask = (1u << lim) - 1u;
        const int rowo = y0 * S.W;
        for (int x = 0; x < w; ++x) {
            const int p = (s + 1) * DEC_FW + x + 1;
            unsigned f2 = S.F2[p];
            unsigned f = S.F[p];
            unsigned cand = ~(D2_SIG(f2) | DF_PI4(f)) & limmask;            // rows neither significant nor coded in the significance pass
            if (lim == 4 && f2 == 0) {                                      // flag word all zero: run-length mode
                if (!mqd_decode<SH>(q, S.cx + CTX_AGG, S.tab)) continue;
                unsigned r = mqd_decode<SH>(q, S.cx + CTX_UNI, S.tab);
                r = (r << 1) | mqd_decode<SH>(q, S.cx + CTX_UNI, S.tab);
                f = 0;
                const unsigned sg = t1d_sign<SH>(S, q, p, (int)r, f);
                t1d_new_sig(S, rowo + x + (int)r * S.W, sg, oneplushalf);
                t1d_set_sig(S, p, (int)r, sg, f, f2);
                cand &= ~((2u << r) - 1u);                                   // the rows behind it are coded normally
            } else {
                if (cand == 0) {                                            // nothing to code; the pass still ends the plane for this column
                    if (f & (0xfu << 24)) S.F[p] = f & ~(0xfu << 24);
                    continue;
                }
            }
            while (cand) {
                const int ci = __ffs((int)cand) - 1;
                cand &= cand - 1u;
                const unsigned ctx = mq_ld8(S.zc + (mq_saddr)((f >> (ci * 3)) & 0x1ffu));
                if (!mqd_decode<SH>(q, S.cx + CTX_ZC + ctx, S.tab)) continue;
                const unsigned sg = t1d_sign<SH>(S, q, p, ci, f);
                t1d_new_sig(S, rowo + x + ci * S.W, sg, oneplushalf);
                t1d_set_sig(S, p, ci, sg, f, f2);
            }
            S.F[p] = f & ~(0xfu << 24);
            S.F2[p] = (unsigned char)f2;
        }
    }
}

template <bool SH>
__device__ void t1d_sigpass(T1DecState &S, MqDec &q, int w, int h, int bpno)
{
    const int oneplushalf = 3 << bpno;
    for (int s = 0, y0 = 0; y0 < h; ++s, y0 += 4) {
        const int lim = min(4, h - y0);
        const unsigned limmask = (1u << lim) - 1u;
        const int rowo = y0 * S.W;
        for (int x = 0; x < w; ++x) {
            const int p = (s + 1) * DEC_FW + x + 1;
            unsigned f2 = S.F2[p];
            unsigned f = S.F[p];
            unsigned cand = D2_NB(f2) & ~(D2_SIG(f2) | DF_PI4(f)) & limmask;   // insignificant rows with a significant neighbour
            if (cand == 0) continue;
            do {
                const int ci = __ffs((int)cand) - 1;
                const unsigned ctx = mq_ld8(S.zc + (mq_saddr)((f >> (ci * 3)) & 0x1ffu));
                if (mqd_decode<SH>(q, S.cx + CTX_ZC + ctx, S.tab)) {
                    const unsigned sg = t1d_sign<SH>(S, q, p, ci, f);
                    t1d_new_sig(S, rowo + x + ci * S.W, sg, oneplushalf);
                    t1d_set_sig(S, p, ci, sg, f, f2);
                }
                f |= DF_PI(ci);
                // a row that just turned significant makes the rows behind it candidates
                cand = D2_NB(f2) & ~(D2_SIG(f2) | DF_PI4(f)) & limmask & ~((2u << ci) - 1u);
            } while (cand);
            S.F[p] = f; S.F2[p] = (unsigned char)f2;
        }
    }
}

template <bool SH>
__device__ void t1d_refpass(T1DecState &S, MqDec &q, int w, int h, int bpno)
{
    const int poshalf = 1 << bpno;
    for (int s = 0, y0 = 0; y0 < h; ++s, y0 += 4) {
        const int lim = min(4, h - y0);
        const unsigned limmask = (1u << lim) - 1u;
        const int rowo = y0 * S.W;
        for (int x = 0; x < w; ++x) {
            const int p = (s + 1) * DEC_FW + x + 1;
            const unsigned f2 = S.F2[p];
            unsigned f = S.F[p];
            unsigned cand = D2_SIG(f2) & ~DF_PI4(f) & limmask;              // significant before this plane
            if (cand == 0) continue;
            do {
                const int ci = __ffs((int)cand) - 1;
                cand &= cand - 1u;
                const int ctx = (f & DF_MU(ci)) ? CTX_MAG + 2 : ((f & DF_NB(ci)) ? CTX_MAG + 1 : CTX_MAG);
                const unsigned v = mqd_decode<SH>(q, S.cx + ctx, S.tab);
                const unsigned neg = (f >> (18 + ci + 1)) & 1u;
                atomicAdd(S.plane + (rowo + x + ci * S.W), (v ^ neg) ? poshalf : -poshalf);
                f |= DF_MU(ci);
            } while (cand);
            S.F[p] = f;
        }
    }
}

// coding passes p0 .. p1-1 of one block (pass 0 = cleanup of the top bit-plane, then sig / ref / cleanup per plane).
// rates != nullptr: truncation shadows for the cuts rates[0 .. nrates-1]; the verdict of cut p lands in ds.bad when
// pass p ends.
template <bool SH>
__device__ void t1d_passes(T1DecState &S, MqDec &q, int w, int h, int numbps, int p0, int p1,
                           const int *rates, int nrates, DecState &ds)
{
    for (int p = p0; p < p1; ++p) {
        const int passtype = p == 0 ? 2 : (p - 1) % 3, bpno = p == 0 ? numbps - 1 : numbps - 2 - (p - 1) / 3;
        if (bpno < 0) break;
        S.pass = p;
        if (passtype == 0) t1d_sigpass<SH>(S, q, w, h, bpno);
        else if (passtype == 1) t1d_refpass<SH>(S, q, w, h, bpno);
        else t1d_clnpass<SH>(S, q, w, h, bpno);
        if (rates) {
            const int pn = p + DEC_NSHADOW;
            if (mqd_retire(q.sw, q.sh_on, pn < nrates ? rates[pn] : -1, q.maxread) && p < 128) ds.bad[p >> 5] |= 1u << (p & 31);
            q.fork_at = mqd_next_fork(q.sw, q.sh_on);
        }
    }
}

// serial decoder registers <-> save area
__device__ __noinline__ void mqd_save(const MqDec &q, const unsigned *cx, unsigned *o)
{
    int k = 0;
    o[k++] = q.a; o[k++] = q.m.c; o[k++] = q.m.ct; o[k++] = (unsigned)q.m.bp; o[k++] = (unsigned)q.m.len; o[k++] = q.m.cur;
    for (int j = 0; j < DEC_NSHADOW; ++j) {
        const MqCore &h = q.sw->sh[j];
        o[k++] = h.c; o[k++] = h.ct; o[k++] = (unsigned)h.bp; o[k++] = (unsigned)h.len; o[k++] = h.cur;
        o[k++] = (unsigned)q.sw->trunc[j];
    }
    o[k++] = q.sh_on; o[k++] = q.sw->sh_bad; o[k++] = (unsigned)q.maxread;
    for (int i = 0; i < 19; ++i) o[k++] = cx[i];
}
__device__ __noinline__ void mqd_load(MqDec &q, unsigned *cx, const unsigned *o, MqShadows *sw)
{
    int k = 0;
    q.a = o[k++]; q.m.c = o[k++]; q.m.ct = o[k++]; q.m.bp = (int)o[k++]; q.m.len = (int)o[k++]; q.m.cur = o[k++];
    q.sw = sw;
    for (int j = 0; j < DEC_NSHADOW; ++j) {
        MqCore &h = sw->sh[j];
        h.c = o[k++]; h.ct = o[k++]; h.bp = (int)o[k++]; h.len = (int)o[k++]; h.cur = o[k++];
        sw->trunc[j] = (int)o[k++];
    }
    q.sh_on = o[k++]; sw->sh_bad = o[k++]; q.maxread = (int)o[k++];
    q.fork_at = mqd_next_fork(sw, q.sh_on);
    for (int i = 0; i < 19; ++i) cx[i] = o[k++];
}

// One decode job as the kernel sees it (every lane can rebuild any job of its warp from the job index).
struct DecView {
    int cbi, frame, mode;
    size_t bi, o;              // block index in the [frame][cblk] tables; offset of the block's top-left sample in a plane
    DecJob jb;
    CblkDesc cb;
};

__device__ __forceinline__ bool dec_view(DecView &v, long long job, const CblkDesc *cblks, const int *order, const DecJob *jobs,
                                         int W, int H, int ncblks, int nframes)
{
    v.mode = 0;
    if (job >= (long long)ncblks * nframes) return false;
    // frame varies fastest so that the blocks of a warp (and of neighbouring warps) are of similar length
    const int slot = (int)(job / nframes);
    v.frame = (int)(job - (long long)slot * nframes);
    v.cbi = order[slot];                       // largest blocks first
    v.bi = (size_t)v.frame * ncblks + v.cbi;
    v.jb = jobs[v.bi];
    v.mode = (v.jb.mode == 1 || v.jb.mode == 2) ? v.jb.mode : 0;
    v.cb = cblks[v.cbi];
    v.o = (size_t)v.frame * H * W + (size_t)v.cb.ty * W + v.cb.tx;
    return v.mode != 0;
}

// Warps per CTA for L blocks per warp: 8 blocks (39 KB of flag words) per CTA from L = 2 on.
#define DEC_WPC(L) ((L) == 1 ? 4 : 8 / (L))
#define DEC_SMEM_BYTES(L) ((size_t)(DEC_WPC(L) * (L)) * DEC_FWORDS * (sizeof(unsigned) + sizeof(unsigned char)))

// grid = ceil(ncblks * nframes / (DEC_WPC(L) * L)) CTAs.  A warp owns L code-blocks; lanes 0 .. L-1 decode one each
// (SIMT: they share every instruction they have in common, which trades a longer serial chain per block for fewer
// issued instructions -- worth it when other batches are in flight), all 32 lanes move state.
//   plain decode (ebcc_decode; cbout == nullptr): mode-1 jobs from k_t2_parse, decoded into iplane.
//   CR search (cbout, dstate, save given): mode 1 continues the block's recording from the complete segment up to
//   jb.np passes with truncation shadows; if the cut this probe needs turns out not to be reproducible the lane
//   decodes the truncated segment for real into the scratch plane right away (job becomes mode 2).  Mode 2 does only
//   that.  The block region of a plane is cleared before a decode that starts at pass 0.
template <int L>
__global__ void __launch_bounds__(DEC_WPC(L) * 32)
k_t1_dec(const CblkDesc *__restrict__ cblks, const int *__restrict__ order, DecJob *jobs, const unsigned char *__restrict__ pool,
         size_t pool_stride, const DecLuts *__restrict__ g_luts, int *iplane, unsigned char *spass, int *scratch,
         const CblkOut *__restrict__ cbout, DecState *dstate, unsigned *save, int W, int H, int ncblks, int nframes)
{
    constexpr int NB = DEC_WPC(L) * L;          // blocks per CTA
    // flag words + candidate bytes of the CTA's blocks: dynamic shared memory (DEC_SMEM_BYTES(L): 47.5 KB for eight blocks)
    EBCC_DYN_SMEM(unsigned, dynsm);
    unsigned (*F)[DEC_FWORDS] = (unsigned (*)[DEC_FWORDS])dynsm;
    unsigned char (*F2)[DEC_FWORDS] = (unsigned char (*)[DEC_FWORDS])(dynsm + NB * DEC_FWORDS);
    __shared__ unsigned cxs[NB][20];
    __shared__ MqShadows shw[NB];
    __shared__ int need2[NB];
    __shared__ unsigned tab[94];
    __shared__ unsigned char zc[4 * 512];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int i = tid; i < 94; i += blockDim.x) tab[i] = g_luts->mq[i];
    for (int i = tid; i < 4 * 512; i += blockDim.x) zc[i] = g_luts->zc[i];
    __syncthreads();
    const long long job0 = ((long long)blockIdx.x * DEC_WPC(L) + warp) * L;
    const bool search = cbout != nullptr && dstate != nullptr && save != nullptr;

    // ---- all lanes: bring every block of the warp to its starting state ----
    for (int sb = 0; sb < L; ++sb) {
        DecView v;
        if (!dec_view(v, job0 + sb, cblks, order, jobs, W, H, ncblks, nframes)) continue;
        unsigned *Fb = F[warp * L + sb];
        const bool resume = v.mode == 1 && search && dstate[v.bi].np_dec > 0;
        if (resume) {
            const unsigned *sv = save + v.bi * DEC_SAVE_WORDS;
            for (int i = lane; i < DEC_FWORDS; i += 32) { Fb[i] = sv[i]; F2[warp * L + sb][i] = (unsigned char)d2_from_flags(sv[i]); }
        } else {
            for (int i = lane; i < DEC_FWORDS; i += 32) { Fb[i] = 0; F2[warp * L + sb][i] = 0; }
            int *dst = (v.mode == 1 ? iplane : scratch) + v.o;
            unsigned char *sp = (v.mode == 1 && spass) ? spass + v.o : nullptr;
            for (int i = lane; i < v.cb.w * v.cb.h; i += 32) {
                const int y = i / v.cb.w, x = i - y * v.cb.w;
                dst[(size_t)y * W + x] = 0;
                if (sp) sp[(size_t)y * W + x] = 255;
            }
        }
    }
    if (lane < L) need2[warp * L + lane] = 0;
    __syncwarp();

    // ---- lanes 0 .. L-1: decode ----
    DecView v;
    const bool mine = lane < L && dec_view(v, job0 + lane, cblks, order, jobs, W, H, ncblks, nframes);
    const int blk = warp * L + (lane < L ? lane : 0);
    T1DecState S;
    S.F = F[blk]; S.F2 = F2[blk]; S.cx = cxs[blk]; S.tab = MQ_OPAQUE(MQ_SADDR(tab)); S.W = W; S.pass = 0;
    DecState ds; ds.np_dec = 0; ds.scratch_np = 0; ds.bad[0] = ds.bad[1] = ds.bad[2] = ds.bad[3] = 0; ds.first_entry = -1; ds.pad = 0;
    if (mine) {
        S.zc = MQ_OPAQUE(MQ_SADDR(zc + v.cb.orient * 512));
        if (search) ds = dstate[v.bi];
        const unsigned char *seg = pool + (size_t)v.frame * pool_stride + v.jb.off;
        unsigned *cx = cxs[blk];
        if (v.mode == 1) {
            unsigned *sv = search ? save + v.bi * DEC_SAVE_WORDS : nullptr;
            const bool resume = search && ds.np_dec > 0;
            if (v.jb.np > 0 && v.jb.numbps > 0) {
                S.plane = iplane + v.o; S.spass = spass ? spass + v.o : nullptr;
                MqDec q;
                const int *rates = search ? cbout[v.bi].rate : nullptr;
                const int nrates = search ? cbout[v.bi].totalpasses : 0;
                if (resume) { mqd_load(q, cx, sv + DEC_SAVE_MQ, &shw[blk]); q.data = seg; }
                else {
                    for (int i = 0; i < 19; ++i) cx[i] = tab[0];
                    cx[CTX_UNI] = tab[46 << 1]; cx[CTX_AGG] = tab[3 << 1]; cx[CTX_ZC] = tab[4 << 1];
                    mqd_init(q, seg, v.jb.len, rates ? &shw[blk] : (MqShadows *)nullptr);
                    if (rates) {
                        for (int j = 0; j < DEC_NSHADOW; ++j) {
                            const int tn = j < nrates ? rates[j] : -1;
                            shw[blk].trunc[j] = tn;
                            // cuts inside the two bytes the decoder reads while initialising cannot fork any more
                            if (tn >= 0 && tn <= q.maxread) { q.sh_on |= 1u << j; shw[blk].sh_bad |= 1u << j; }
                        }
                        q.fork_at = mqd_next_fork(&shw[blk], q.sh_on);
                    }
                }
                if (rates) t1d_passes<true>(S, q, v.cb.w, v.cb.h, v.jb.numbps, ds.np_dec, v.jb.np, rates, nrates, ds);
                else t1d_passes<false>(S, q, v.cb.w, v.cb.h, v.jb.numbps, ds.np_dec, v.jb.np, rates, nrates, ds);
                ds.np_dec = v.jb.np;
                if (search) mqd_save(q, cx, sv + DEC_SAVE_MQ);
            }
            if (search) {
                // the cut this probe needs: rate[np - 1]
                const int pc = v.jb.np - 1;
                int need = v.jb.np > 0 && (pc >= 128 || ((ds.bad[pc >> 5] >> (pc & 31)) & 1u));
                if (need && ds.scratch_np == v.jb.np) { need = 0; jobs[v.bi].mode = 3; }
                need2[blk] = need ? 1 : 0;
                if (!need) dstate[v.bi] = ds;
            }
        } else need2[blk] = 1;
    }
    __syncwarp();

    // ---- all lanes: save the recordings' flag words; clear what a real truncated decode needs ----
    for (int sb = 0; sb < L; ++sb) {
        DecView u;
        if (!dec_view(u, job0 + sb, cblks, order, jobs, W, H, ncblks, nframes)) continue;
        unsigned *Fb = F[warp * L + sb];
        if (u.mode == 1 && search) {
            unsigned *sv = save + u.bi * DEC_SAVE_WORDS;
            for (int i = lane; i < DEC_FWORDS; i += 32) sv[i] = Fb[i];
        }
        if (need2[warp * L + sb] && (u.mode == 2 ? false : true)) {
            // recording just made, cut not reproducible: the scratch decode starts from a clean block
            for (int i = lane; i < DEC_FWORDS; i += 32) { Fb[i] = 0; F2[warp * L + sb][i] = 0; }
            for (int i = lane; i < u.cb.w * u.cb.h; i += 32) {
                const int y = i / u.cb.w, x = i - y * u.cb.w;
                scratch[u.o + (size_t)y * W + x] = 0;
            }
        }
    }
    __syncwarp();

    // ---- lanes 0 .. L-1: real decode of the segment cut at rate[np - 1] into the scratch plane ----
    if (mine && need2[blk]) {
        if (v.jb.np > 0 && v.jb.numbps > 0) {
            unsigned *cx = cxs[blk];
            for (int i = 0; i < 19; ++i) cx[i] = tab[0];
            cx[CTX_UNI] = tab[46 << 1]; cx[CTX_AGG] = tab[3 << 1]; cx[CTX_ZC] = tab[4 << 1];
            S.plane = scratch + v.o; S.spass = nullptr;
            MqDec q;
            mqd_init(q, pool + (size_t)v.frame * pool_stride + v.jb.off, search ? cbout[v.bi].rate[v.jb.np - 1] : v.jb.len, nullptr);
            DecState dummy = ds;
            t1d_passes<false>(S, q, v.cb.w, v.cb.h, v.jb.numbps, 0, v.jb.np, nullptr, 0, dummy);
        }
        if (search) { ds.scratch_np = v.jb.np; dstate[v.bi] = ds; }
        if (v.jb.mode == 1) jobs[v.bi].mode = 2;
    }
}

// ---- unproven cuts, decoded ahead of the search --------------------------------------------------------------------------
// About one truncation point in a thousand does not decode like the complete segment (mq_cut_proven).  A probe that
// picks such a point needs the real thing, and one block is a serial chain of several milliseconds: decoded on demand
// it would set the pace of every probe.  So right after tier-1 every unproven (block, pass count) of the batch becomes
// an entry of a list and ONE launch decodes them all side by side, each into its own 64 x 64 slot of a pool.
__global__ void k_cut_collect(const CblkOut *__restrict__ cbout, T1CutState *dstate, T1CutEntry *list, int *count, int cap, long long njobs)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= njobs) return;
    T1CutState &d = dstate[i];
    const int n = __popc(d.bad[0]) + __popc(d.bad[1]) + __popc(d.bad[2]) + __popc(d.bad[3]);
    if (n == 0) return;
    const int first = atomicAdd(count, n);
    if (first + n > cap) return;                  // pool full: these cuts are decoded when a probe asks (first_entry stays < 0)
    d.first_entry = first;
    int e = first;
    for (int p = 0; p < 128 && p < cbout[i].totalpasses; ++p)
        if ((d.bad[p >> 5] >> (p & 31)) & 1u) { list[e].bi = (unsigned)i; list[e].np = p + 1; ++e; }
}

// value of a sample after the first np coding passes, from its recording (value v after npd >= np passes, pass of
// significance ps)
__device__ __forceinline__ int recon_value(int v, int ps, int np, int npd, int numbps)
{
    if (np == npd || v == 0) return v;
    if (ps >= np) return 0;
    const int kk = ps == 0 ? 0 : (ps + 2) / 3, kr = np / 3;
    const int b = numbps - 1 - (kr > kk ? kr : kk);
    const int a = v < 0 ? -v : v;
    const int mag = ((((a >> 1) >> b) << b) << 1) + (1 << b);
    return v < 0 ? -mag : mag;
}

#define CUT_WPC 4

// "Does the segment cut at R bytes decode like the complete one up to the end of the pass whose coder snapshot is (a, c, ct,
// pend, pos)?" for ANY R (mq_cut_proven, k_t1.cuh, answers it for the pass's own rate, at most two bytes behind the pending
// one).  Same criterion: the value the decoder reads -- final bytes, then one-bits -- must not exceed L + a.  Bytes whose
// weight falls below 2^-8 of a's LSB only add a fraction in (0, 1] of that unit, so with true bytes left over the test is
// A + 1 <= U for the integer part A.
__device__ __forceinline__ bool mq_cut_covers(unsigned a, unsigned c, int ct, unsigned pend, int pos, int R, const unsigned char *v)
{
    if (R - 1 < pos) return false;
    const int s = 27 - ct + 8;
    unsigned long long X = 0;
    int off = 0, last = 0, i = pos;
    for (; i < R && s - off >= 0; ++i) {
        const unsigned vi = i < 0 ? 0u : (unsigned)v[i];
        X += (unsigned long long)vi << (s - off);
        last = off;
        off += vi == 0xffu ? 7 : 8;
    }
    X += i < R ? 1ull : (1ull << (s - last));
    const unsigned long long U = ((((unsigned long long)pend << (27 - ct)) + c + a) << 8);
    return X <= U;
}

// One entry = (block, np): the block decoded from its segment cut at rate[np - 1], np passes, into the entry's pool slot.
// Decoding from pass 0 would be a chain over every decision of the block so far.  Instead the entry starts at the
// bit-plane boundary below its last pass (pass p0 = 1 + 3 j, a significance pass), because everything a decoder holds there
// is known without decoding:
//   * sample state: the encoder's recording (value after all passes, pass of first significance) gives significance, sign,
//     "already refined" and the value after p0 passes (recon_value) of every sample; the flag words follow from those;
//   * contexts: k_t1_mq saved the 19 context states at the end of every clean-up pass (ctxsnap);
//   * MQ registers: a is the coder's a at the end of pass p0 - 1 (CblkOut::mqs).  The decoder's C register is linear in the
//     bytes read and in the subtracted qe's: C = (bits read so far, as the byte-reading half alone shifts them in) minus
//     (the coder's low bound << 16), modulo 2^32; only the low 16 bits of the coder's c matter.  The byte-reading half is
//     replayed over N shifts, N = 12 - ct + sum over the bytes before the pending one of (7 after an 0xFF, else 8).
// Valid iff the cut segment decodes like the complete one up to the end of pass p0 - 1 (mq_cut_covers); otherwise, and for
// entries whose last pass is pass 0, the chain starts at pass 0 as before.
__global__ void __launch_bounds__(CUT_WPC * 32)
k_t1_dec_cuts(const CblkDesc *__restrict__ cblks, const CblkOut *__restrict__ cbout, const T1CutEntry *__restrict__ list, const int *__restrict__ count,
              int cap, const unsigned char *__restrict__ pool, size_t pool_stride, const DecLuts *__restrict__ g_luts, int *cut_pool, int ncblks,
              const int *__restrict__ rec_val, const unsigned char *__restrict__ rec_pass, const unsigned char *__restrict__ ctxsnap, int W, int H, int nframes)
{
    __shared__ unsigned F[CUT_WPC][DEC_FWORDS];
    __shared__ unsigned char F2[CUT_WPC][DEC_FWORDS];
    __shared__ unsigned cxs[CUT_WPC][20];
    __shared__ unsigned tab[94];
    __shared__ unsigned char zc[4 * 512];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int n = min(*count, cap);
    if (blockIdx.x * CUT_WPC >= n) return;
    for (int i = tid; i < 94; i += blockDim.x) tab[i] = g_luts->mq[i];
    for (int i = tid; i < 4 * 512; i += blockDim.x) zc[i] = g_luts->zc[i];
    __syncthreads();
    const int e = blockIdx.x * CUT_WPC + warp;
    if (e >= n) return;
    const T1CutEntry en = list[e];
    if (en.bi >= (unsigned)ncblks * (unsigned)nframes) return;      // slot reserved by a block that found the pool full
    const int k = (int)(en.bi % (unsigned)ncblks), frame = (int)(en.bi / (unsigned)ncblks);
    const CblkDesc cb = cblks[k];
    const CblkOut &co = cbout[en.bi];
    if (en.np > co.totalpasses) return;
    const unsigned char *seg = pool + (size_t)frame * pool_stride + cb.mq_off;
    int *dst = cut_pool + (size_t)e * T1_CUT_WORDS;
    unsigned *Fw = F[warp];
    for (int i = lane; i < DEC_FWORDS; i += 32) Fw[i] = 0;
    // start at a bit-plane boundary?
    const int plast = en.np - 1, R = en.np > 0 ? co.rate[en.np - 1] : 0;
    int p0 = 0;
    if (ctxsnap != nullptr && rec_val != nullptr && plast >= 1 && co.numbps > 0 && (plast - 1) / 3 < T1_SNAP_PLANES) {
        const int q = ((plast - 1) / 3) * 3;                     // the clean-up pass in front of the plane of the last pass
        const unsigned w0 = co.mqs[q][0], w1 = co.mqs[q][1];
        if (mq_cut_covers(w1 & 0xffffu, w0 & 0x0fffffffu, (int)(w0 >> 28), (w1 >> 16) & 0xffu, (int)co.mqs[q][2], R, seg)) p0 = q + 1;
    }
    __syncwarp();
    if (p0 == 0) {
        for (int i = lane; i < cb.h * J2K_CBLK; i += 32) dst[i] = 0;
    } else {
        // sample state after p0 passes from the recording: values into the slot, flag words by OR-ing what t1d_set_sig would
        // have set sample by sample (significance of the 3 x 6 neighbourhood, signs, mu)
        const int jp = (p0 - 1) / 3;                              // plane index of pass p0
        const size_t base = (size_t)frame * W * H + (size_t)cb.ty * W + cb.tx;
        for (int i = lane; i < cb.h * J2K_CBLK; i += 32) {
            const int y = i / J2K_CBLK, x = i - y * J2K_CBLK;
            int val = 0;
            if (x < cb.w) {
                const int v = rec_val[base + (size_t)y * W + x], ps = rec_pass[base + (size_t)y * W + x];
                if (v != 0 && ps < p0) {
                    val = recon_value(v, ps, p0, co.totalpasses, co.numbps);
                    const unsigned sg = v < 0 ? 1u : 0u;
                    const int js = ps == 0 ? -1 : (ps - 1) / 3;      // plane of first significance; first refined in plane js + 1
                    const int ci = y & 3, pw = ((y >> 2) + 1) * DEC_FW + x + 1;
                    atomicOr(&Fw[pw], DF_SIG(ci) | (sg ? DF_CHI(ci) : 0u) | (js <= jp - 2 ? DF_MU(ci) : 0u));
                    atomicOr(&Fw[pw - 1], 1u << ((ci + 1) * 3 + 2));
                    atomicOr(&Fw[pw + 1], 1u << ((ci + 1) * 3 + 0));
                    if (ci == 0) {
                        atomicOr(&Fw[pw - DEC_FW], (1u << 16) | (sg ? DF_CHI(4) : 0u));
                        atomicOr(&Fw[pw - DEC_FW - 1], 1u << 17);
                        atomicOr(&Fw[pw - DEC_FW + 1], 1u << 15);
                    }
                    if (ci == 3) {
                        atomicOr(&Fw[pw + DEC_FW], (1u << 1) | (sg ? DF_CHI(-1) : 0u));
                        atomicOr(&Fw[pw + DEC_FW - 1], 1u << 2);
                        atomicOr(&Fw[pw + DEC_FW + 1], 1u << 0);
                    }
                }
            }
            dst[i] = val;
        }
    }
    __syncwarp();
    for (int i = lane; i < DEC_FWORDS; i += 32) F2[warp][i] = (unsigned char)d2_from_flags(Fw[i]);
    __syncwarp();
    if (lane != 0) return;
    if (en.np <= 0 || co.numbps <= 0) return;
    unsigned *cx = cxs[warp];
    T1DecState S;
    S.F = Fw; S.F2 = F2[warp]; S.cx = cx; S.tab = MQ_OPAQUE(MQ_SADDR(tab)); S.W = J2K_CBLK; S.pass = 0; S.zc = MQ_OPAQUE(MQ_SADDR(zc + cb.orient * 512));
    S.plane = dst; S.spass = nullptr;
    MqDec q;
    mqd_init(q, seg, R, nullptr);
    if (p0 == 0) {
        for (int i = 0; i < 19; ++i) cx[i] = tab[0];
        cx[CTX_UNI] = tab[46 << 1]; cx[CTX_AGG] = tab[3 << 1]; cx[CTX_ZC] = tab[4 << 1];
    } else {
        const int qp = p0 - 1;
        const unsigned char *sn = ctxsnap + ((size_t)en.bi * T1_SNAP_PLANES + qp / 3) * T1_SNAP_BYTES;
        for (int i = 0; i < 19; ++i) cx[i] = tab[sn[i] < 94 ? sn[i] : 0];
        const unsigned w0 = co.mqs[qp][0], w1 = co.mqs[qp][1];
        const int ct_e = (int)(w0 >> 28), pos = (int)co.mqs[qp][2];
        int N = 12 - ct_e;
        for (int i = -1; i < pos; ++i) N += (i >= 0 && seg[i] == 0xffu) ? 7 : 8;
        mqc_shift(q.m, seg, N);
        q.m.c -= (w0 & 0xffffu) << 16;
        q.a = w1 & 0xffffu;
    }
    DecState dummy; dummy.np_dec = 0; dummy.scratch_np = 0; dummy.bad[0] = dummy.bad[1] = dummy.bad[2] = dummy.bad[3] = 0; dummy.first_entry = -1; dummy.pad = 0;
    t1d_passes<false>(S, q, cb.w, cb.h, co.numbps, p0, en.np, nullptr, 0, dummy);
}

// decode jobs for the encoder's own allocation (compression-ratio search).  Block k is recorded (mode 1) when the
// allocation needs more passes than its recording holds, served from the recording (mode 0) when the recording can
// reproduce "first numpasses passes", and decoded for real into the scratch plane (mode 2) otherwise.
__global__ void k_jobs_from_enc(const CblkDesc *__restrict__ cblks, const CblkOut *__restrict__ cbout, const DecState *__restrict__ dstate,
                                DecJob *jobs, int ncblks, int nframes, int use_recording, uint2 *rjobs, ErrStat *es_reset)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    // the probe's statistics start from scratch (k_errstat_init folded in: one launch less per probe)
    if (es_reset && i < nframes) { es_reset[i].n_over = 0; es_reset[i].max_err_bits = 0; es_reset[i].rmin_key = 0xffffffffu; es_reset[i].rmax_key = 0; es_reset[i].sum_exp = 0; es_reset[i].sum_fx = 0; }
    if (i >= (long long)ncblks * nframes) return;
    const int k = (int)(i % ncblks);
    const CblkOut &c = cbout[i];
    DecJob j;
    const int np = c.numpasses;
    j.np = np; j.numbps = c.numbps; j.off = cblks[k].mq_off; j.pad = 0;
    j.len = np > 0 ? c.rate[np - 1] : 0;
    if (!use_recording) j.mode = 2;
    else if (np == 0) j.mode = 0;
    else {
        const DecState &d = dstate[i];
        if (np > d.np_dec) { j.mode = 1; j.len = c.datalen; }          // record (on) from the complete segment
        else if ((np - 1) < 128 && !((d.bad[(np - 1) >> 5] >> ((np - 1) & 31)) & 1u)) j.mode = 0;
        else if (d.first_entry >= 0) {
            // pre-decoded by k_t1_dec_cuts: entry = first + number of unproven cuts of this block below this one
            int rank = 0;
            for (int w = 0; w < ((np - 1) >> 5); ++w) rank += __popc(d.bad[w]);
            rank += __popc(d.bad[(np - 1) >> 5] & ((1u << ((np - 1) & 31)) - 1u));
            j.mode = 4; j.pad = d.first_entry + rank;
        }
        else j.mode = d.scratch_np == np ? 3 : 2;
    }
    jobs[i] = j;
    // compact record for the fused inverse transform (k_idwt_tile): mode | np << 8 | recorded passes << 16 | bit-planes << 24, pool entry
    if (rjobs) rjobs[i] = make_uint2((unsigned)j.mode | ((unsigned)np << 8) | ((unsigned)(use_recording ? dstate[i].np_dec : 0) << 16) | ((unsigned)c.numbps << 24),
                                     (unsigned)j.pad);
}

__global__ void k_dstate_reset(DecState *d, long long n)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { d[i].np_dec = 0; d[i].scratch_np = 0; d[i].bad[0] = d[i].bad[1] = d[i].bad[2] = d[i].bad[3] = 0; d[i].first_entry = -1; d[i].pad = 0; }
}

// Probe reconstruction: integer tier-1 output of "the first numpasses passes" for every sample, taken from the
// recording (value after np_dec passes + pass of significance) or from the scratch plane (mode 2 blocks), then
// dequantised: (float)v * (0.5f * stepsize_dec)  (t1.c:1853-1865).
//   decoder units: |v| = 2*M + 2^b, M = magnitude bits decoded down to plane b, 2^b the half-bit bias
//   (t1.c:582-1140: oneplushalf on significance, +-poshalf per refinement)

__global__ void k_recon_dequant(const Plan *__restrict__ plan, const CblkOut *__restrict__ cbout, const DecState *__restrict__ dstate,
                                const DecJob *__restrict__ jobs, const int *__restrict__ iplane, const unsigned char *__restrict__ spass,
                                const int *__restrict__ scratch, float *coef, const CblkDesc *__restrict__ cblks, const int *__restrict__ cut_pool)
{
    const int W = plan->W, H = plan->H, ncblks = plan->ncblks;
    const size_t n = (size_t)W * H;
    const int frame = blockIdx.y;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const int y = (int)(i / W), x = (int)(i - (size_t)y * W);
        int r = 0;
        while (r < J2K_NRES - 1 && (x >= plan->res_w[r] || y >= plan->res_h[r])) ++r;
        int band = 0;
        if (r > 0) {
            const int o = (x >= plan->res_w[r - 1] ? 1 : 0) | (y >= plan->res_h[r - 1] ? 2 : 0);
            band = 1 + 3 * (r - 1) + (o - 1);
        }
        const BandDesc &bd = plan->bands[band];
        int gy = (y - bd.offy) >> 6, k = bd.first_cblk + gy * bd.cw + ((x - bd.offx) >> 6);
        // tiles below the first one: the block grid is anchored on the canvas, the band's first block row may be a partial one
        while (gy + 1 < bd.ch && y >= cblks[k].ty + cblks[k].h) { ++gy; k += bd.cw; }
        const size_t bi = (size_t)frame * ncblks + k, si = (size_t)frame * n + i;
        int v;
        const int md = jobs[bi].mode;
        if (md == 4) v = cut_pool[(size_t)jobs[bi].pad * T1_CUT_WORDS + (size_t)(y - cblks[k].ty) * J2K_CBLK + (x - cblks[k].tx)];
        else if (md >= 2) v = scratch[si];
        else {
            const int np = cbout[bi].numpasses;
            v = np > 0 ? recon_value(iplane[si], spass[si], np, dstate[bi].np_dec, cbout[bi].numbps) : 0;   // a block that was never included has no recording
        }
        coef[si] = __fmul_rn((float)v, __fmul_rn(0.5f, bd.stepsize_dec));
    }
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
    for (int k = threadIdx.x; k < ncblks; k += blockDim.x) { DecJob z; z.np = 0; z.numbps = 0; z.off = 0; z.len = 0; z.mode = 1; z.pad = 0; jb[k] = z; }
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
    // tile-parts follow one another (SOT .. data, Psot bytes each); find the one of this plan's tile
    int tp_end;
    for (;;) {
        if (pos + 14 > n || s[pos] != 0xff || s[pos + 1] != 0x90) return;
        const unsigned psot = ((unsigned)s[pos + 6] << 24) | ((unsigned)s[pos + 7] << 16) | ((unsigned)s[pos + 8] << 8) | s[pos + 9];
        const int isot = (s[pos + 4] << 8) | s[pos + 5];
        tp_end = psot ? pos + (int)psot : n - 2;
        if (psot != 0 && psot < 14) return;
        if (tp_end > n || tp_end < pos) return;
        if (isot == plan->tile_index) break;
        if (!psot) return;
        pos = tp_end;
    }
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
            j.np = np; j.len = (int)bi_read(b, nbits); j.off = 0; j.mode = 1; j.pad = 0;
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

// ---- test taps: every truncation point of every code-block checked against a real decode ------------------------------
__global__ void k_dbg_force_np(CblkOut *cbout, long long n, int np)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) cbout[i].numpasses = cbout[i].totalpasses >= np ? np : 0;
}

// grid (ncblks, nframes): block k of a frame, cut after np passes: recording vs the real truncated decode in `scratch`.
// cnt[0] cuts checked, [1] flagged unproven, [2] flagged proven but DIFFERENT (must stay 0), [3] flagged unproven but equal,
// [4] unproven cuts found in the pre-decoded pool, [5] pool entries that differ from the real decode (must stay 0)
__global__ void k_dbg_compare(const CblkDesc *__restrict__ cblks, const CblkOut *__restrict__ cbout, const DecState *__restrict__ dstate,
                              const int *__restrict__ iplane, const unsigned char *__restrict__ spass, const int *__restrict__ scratch,
                              unsigned long long *cnt, int np, int W, int H, int ncblks, const int *__restrict__ cut_pool)
{
    const int k = blockIdx.x, frame = blockIdx.y;
    const size_t bi = (size_t)frame * ncblks + k;
    if (cbout[bi].numpasses != np || np <= 0) return;
    const CblkDesc cb = cblks[k];
    const int npd = dstate[bi].np_dec, numbps = cbout[bi].numbps;
    int diff = 0, pdiff = 0;
    const int bad = (np - 1) >= 128 || ((dstate[bi].bad[(np - 1) >> 5] >> ((np - 1) & 31)) & 1u);
    int entry = -1;
    if (bad && np - 1 < 128 && dstate[bi].first_entry >= 0) {
        int rank = 0;
        for (int w = 0; w < ((np - 1) >> 5); ++w) rank += __popc(dstate[bi].bad[w]);
        rank += __popc(dstate[bi].bad[(np - 1) >> 5] & ((1u << ((np - 1) & 31)) - 1u));
        entry = dstate[bi].first_entry + rank;
    }
    for (int i = threadIdx.x; i < cb.w * cb.h; i += blockDim.x) {
        const int y = i / cb.w, x = i - y * cb.w;
        const size_t si = (size_t)frame * W * H + (size_t)(cb.ty + y) * W + cb.tx + x;
        if (recon_value(iplane[si], spass[si], np, npd, numbps) != scratch[si]) diff = 1;
        if (entry >= 0 && cut_pool[(size_t)entry * T1_CUT_WORDS + (size_t)y * J2K_CBLK + x] != scratch[si]) pdiff = 1;
    }
    diff = __syncthreads_or(diff);
    pdiff = __syncthreads_or(pdiff);
    if (threadIdx.x == 0) {
        if (entry >= 0) { atomicAdd(&cnt[4], 1ull); if (pdiff) atomicAdd(&cnt[5], 1ull); }
        atomicAdd(&cnt[0], 1ull);
        if (bad) atomicAdd(&cnt[1], 1ull);
        if (!bad && diff) atomicAdd(&cnt[2], 1ull);
        if (bad && !diff) atomicAdd(&cnt[3], 1ull);
    }
}
