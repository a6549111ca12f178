// k_t1.cuh -- EBCOT tier-1 encoder, split in two kernels:
//
//   k_t1_cm : one CTA per 64x64 code-block.  Quantises the block (t1.c:2397-2425), then runs the
//             bit-plane passes (cleanup / significance propagation / magnitude refinement) in
//             PARALLEL over stripe columns and emits the (context, decision) symbol stream in
//             scan order through a CTA-wide prefix sum.  Every stripe column owns one flag word
//             holding the significance of its whole 3x6 neighbourhood (kept current by atomicOr
//             scatter), so a context is one shift + mask + table look-up.  The scan-order
//             dependence of the significance state is resolved analytically: a neighbour counts
//             as significant if it was before the pass, or becomes so in this pass AND precedes
//             the sample in scan order.  For the cleanup pass "becomes significant" is known a
//             priori (insignificant, not visited, bit set); for the significance-propagation pass
//             it is the least fixed point of a monotone map, reached by cheap sweeps that emit
//             nothing, followed by one emitting sweep.
//   k_t1_mq : one code-block per warp.  The warp stages the symbol stream through shared memory, lane 0
//             runs the serial MQ arithmetic coder (mqc.c / mqc_inl.h) over it, pass rates as
//             opj_t1_encode_cblk computes them (t1.c:2655-2723).
//
// Integer/latency bound; bit-exact with opj_t1_encode_cblk by construction.
#pragma once
#include "common.cuh"

#ifndef T1_THREADS
#define T1_THREADS 512
#endif
#define T1_CPT (1024 / T1_THREADS)  // stripe columns per thread
#define T1_SW 66                    // state row stride: 64 + border
#define T1_NSTATE (18 * T1_SW)      // 16 stripes + border
// private symbol slot of a thread: 3 passes x T1_CPT columns x 10 symbols, padded to an odd number of words so that
// the 32 lanes of a warp hit 32 different banks
#define T1_SLOT (((30 * T1_CPT + 3) / 4 | 1) * 4)

#define CTX_ZC 0
#define CTX_SC 9
#define CTX_MAG 14
#define CTX_AGG 17
#define CTX_UNI 18

// distortion LUTs (t1_generate_luts.c:259-290), filled by the host at start-up
struct T1Luts { short sig[128], sig0[128], ref[128], ref0[128]; };

// zero-coding context from neighbour counts (Annex D table D.1; t1_generate_luts.c:48-128)
__host__ __device__ __forceinline__ int t1_zc_ctx(int orient, int h, int v, int d)
{
    if (orient == 1) { int t = h; h = v; v = t; }
    if (orient != 3) {
        if (h == 0) { if (v == 0) return d == 0 ? 0 : (d == 1 ? 1 : 2); return v == 1 ? 3 : 4; }
        if (h == 1) { if (v == 0) return d == 0 ? 5 : 6; return 7; }
        return 8;
    }
    int hv = h + v;
    if (d == 0) return hv == 0 ? 0 : (hv == 1 ? 1 : 2);
    if (d == 1) return hv == 0 ? 3 : (hv == 1 ? 4 : 5);
    if (d == 2) return hv == 0 ? 6 : 7;
    return 8;
}

// sign context + xor bit from horizontal / vertical contributions in {-1,0,1}
// entry (hc+1)*3+(vc+1) packs n (3 bits) | spb << 3
__device__ __forceinline__ int t1_sc_lut(int hc, int vc)
{
    const unsigned long long tab = 0x4ull | (0x3ull << 4) | (0x2ull << 8) | (0x1ull << 12) | (0x0ull << 16) |
                                   (0x1ull << 20) | (0x2ull << 24) | (0x3ull << 28) | (0x4ull << 32) |
                                   (0x8ull << 0) | (0x8ull << 4) | (0x8ull << 8) | (0x8ull << 12);
    return (int)((tab >> (4 * ((hc + 1) * 3 + (vc + 1)))) & 0xf);
}
__device__ __forceinline__ int t1_contrib(int sa, int na, int sb, int nb)
{
    int pos = min(1, (sa & (na ^ 1)) + (sb & (nb ^ 1)));
    int neg = min(1, (sa & na) + (sb & nb));
    return pos - neg;
}

// symbols of one coding pass go straight into the thread's private slot of shared memory as ctx << 2 | decision
struct T1Sym { unsigned char *p; int cnt; };
__device__ __forceinline__ void t1_put(T1Sym &s, int ctx, int d)
{
    s.p[s.cnt++] = (unsigned char)((ctx << 2) | d);
}

__device__ __forceinline__ int t1_nm_sig(const T1Luts *l, unsigned x, int bpno)
{ return bpno > 0 ? l->sig[(x >> bpno) & 127] : l->sig0[x & 127]; }
__device__ __forceinline__ int t1_nm_ref(const T1Luts *l, unsigned x, int bpno)
{ return bpno > 0 ? l->ref[(x >> bpno) & 127] : l->ref0[x & 127]; }

// Flag word of one stripe column (the layout of OPJ's opj_flag_t and of the decoder's DF_* word): bits 0-17
// significance of the 3x6 neighbourhood (bit (r+1)*3 + (dx+1), r = -1..4, dx = -1..1; the column's own rows are
// dx = 0, r = 0..3), bits 18-21 the own rows again as a nibble, bits 24-27 pi (coded in the significance pass),
// bits 28-31 mu (already refined).
#define CF_W 66
#define CF_WORDS (18 * CF_W)
#define CF_SIG(ci) (1u << ((ci) * 3 + 4))
#define CF_NB(ci) (0x1efu << ((ci) * 3))
#define CF_PI(ci) (1u << (24 + (ci)))
#define CF_MU(ci) (1u << (28 + (ci)))
// neighbours of a column that precede ALL of its rows in scan order: column to the west rows -1..3, row -1 of the
// own and the eastern column.  (Own rows above the current one are tracked in registers while a column is coded.)
#define CF_EARLY 0x124fu

// bits 18-21 repeat the significance of the column's own rows 0..3 as a nibble
__device__ __forceinline__ unsigned cf_own4(unsigned f) { return (f >> 18) & 15u; }

// rows b4 of column p become significant: set the bits every neighbouring word sees
__device__ __forceinline__ void cf_scatter(unsigned *N, int p, unsigned b4)
{
    unsigned own = 0, we = 0;
#pragma unroll
    for (int ci = 0; ci < 4; ++ci)
        if ((b4 >> ci) & 1u) { own |= CF_SIG(ci); we |= 1u << ((ci + 1) * 3); }
    atomicOr(&N[p], own | (b4 << 18));
    atomicOr(&N[p - 1], we << 2);
    atomicOr(&N[p + 1], we);
    if (b4 & 1u) { atomicOr(&N[p - CF_W], 1u << 16); atomicOr(&N[p - CF_W - 1], 1u << 17); atomicOr(&N[p - CF_W + 1], 1u << 15); }
    if (b4 & 8u) { atomicOr(&N[p + CF_W], 1u << 1); atomicOr(&N[p + CF_W - 1], 1u << 2); atomicOr(&N[p + CF_W + 1], 1u << 0); }
}

// sign coding of row j of column p; t = time-correct significance window, static signs in SG (rows -1..4 of a column
// at bits 0..5)
__device__ __forceinline__ void cf_sign(T1Sym &s, unsigned t, const unsigned *SG, int p, int j, int neg)
{
    const unsigned gc = SG[p], gw = SG[p - 1], ge = SG[p + 1];
    const int hc = t1_contrib((t >> ((j + 1) * 3 + 2)) & 1, (ge >> (j + 1)) & 1, (t >> ((j + 1) * 3)) & 1, (gw >> (j + 1)) & 1);
    const int vc = t1_contrib((t >> (j * 3 + 1)) & 1, (gc >> j) & 1, (t >> ((j + 2) * 3 + 1)) & 1, (gc >> (j + 2)) & 1);
    const int e = t1_sc_lut(hc, vc);
    t1_put(s, CTX_SC + (e & 7), neg ^ (e >> 3));
}

// One column of a significance-propagation pass (t1.c:363-410, 479-580).  Returns the rows that become significant;
// EMIT adds the symbols, the rows coded (pi) and the distortion estimate.
template <bool EMIT>
__device__ __forceinline__ unsigned cf_sigpass_col(unsigned f, unsigned n, const unsigned *mag, unsigned one, int lim, unsigned sgn4,
                                                    const unsigned char *zc, const unsigned *SG, int p, const T1Luts *luts, int bpno,
                                                    T1Sym &sy, unsigned &coded, int &nmsedec)
{
    unsigned t = f | (n & CF_EARLY), fresh = 0;
    if ((t & 0x3ffffu) == 0) return 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        if (j >= lim) break;
        if (!(f & CF_SIG(j)) && (t & CF_NB(j))) {
            const int v = (mag[j] & one) ? 1 : 0;
            if (EMIT) { t1_put(sy, CTX_ZC + zc[(t >> (j * 3)) & 0x1ffu], v); coded |= 1u << j; }
            if (v) {
                if (EMIT) { cf_sign(sy, t, SG, p, j, (sgn4 >> j) & 1); nmsedec += t1_nm_sig(luts, mag[j], bpno); }
                fresh |= 1u << j;
                t |= CF_SIG(j);
            }
        }
    }
    return fresh;
}

// rows b4 of the stripe column with flag-word index p became significant in coding pass `pass`
// (recp = the block's top-left sample in the pass plane: one pointer lives through the kernel instead of frame, H, ty, tx)
__device__ __forceinline__ void t1_rec_pass(unsigned char *recp, int W, int p, unsigned b4, int pass)
{
    const int s = p / CF_W - 1, x = p - (s + 1) * CF_W - 1;
    unsigned char *d = recp + (4 * s) * W + x;
#pragma unroll
    for (int j = 0; j < 4; ++j)
        if ((b4 >> j) & 1u) d[j * W] = (unsigned char)pass;
}

// 3 CTAs per SM caps the kernel at 40 registers: room for the long-lived MQ CTAs of other batches in flight
#ifndef T1_MIN_CTAS
#define T1_MIN_CTAS 3
#endif
template <bool REC, int MINB>      // MINB CTAs per SM: 3 caps the kernel at 40 registers, 2 at 64 (the recording variant then keeps eight hoisted
                                   // store addresses in local memory, which costs less than the third CTA brings with other batches in flight)
__global__ void __launch_bounds__(T1_THREADS, MINB)
k_t1_cm(const float *__restrict__ coef, const CblkDesc *__restrict__ cblks, const BandDesc *__restrict__ bands,
        const T1Luts *__restrict__ g_luts, const unsigned char *__restrict__ zc_lut, CblkOut *cbout, unsigned char *sym_pool,
        int W, int H, int ncblks, size_t sym_stride, int *rec_val, unsigned char *rec_pass, FrameStat *stat)
{
    // REC: rec_val / rec_pass (compression-ratio search): the *recording* a tier-1 decoder of the complete
    // segment would leave behind -- per sample the decoded value after all passes, sign * (2 M + 1) in the decoder's
    // half units (t1.c:582-1140), and the coding pass in which the sample became significant.  The encoder knows both:
    // k_recon_dequant rebuilds from them what "decode the first n passes" yields for any n, without decoding anything.
    __shared__ unsigned F[CF_WORDS];        // flag words
    __shared__ unsigned N[CF_WORDS];        // significance gained in the running pass, same neighbourhood layout
    __shared__ unsigned SG[CF_WORDS];       // static: sign of rows -1..4 of the column (bits 0..5)
    __shared__ unsigned char priv[T1_THREADS * T1_SLOT];   // per thread: 3 passes x T1_CPT columns x 10 symbols (+ pad)
    __shared__ unsigned char zc[512];
    __shared__ T1Luts luts;
    __shared__ unsigned long long warp_sum[T1_THREADS / 32];
    __shared__ int s_max, s_nmsedec[3];

    const int tid = threadIdx.x;
    const int cbi = blockIdx.x, frame = blockIdx.y;
    const CblkDesc cb = cblks[cbi];
    const BandDesc *bd = &bands[cb.band];
    const int bw = cb.w, bh = cb.h, orient = cb.orient;
    const int nstripes = (bh + 3) >> 2, nsc = nstripes * bw;
    CblkOut *out = cbout + (size_t)frame * ncblks + cbi;
    unsigned char *sym = sym_pool + (size_t)frame * sym_stride + cb.sym_off;

    for (int i = tid; i < CF_WORDS; i += T1_THREADS) { F[i] = 0; N[i] = 0; SG[i] = 0; }
    for (int i = tid; i < (int)(sizeof(T1Luts) / 2); i += T1_THREADS) ((short *)&luts)[i] = ((const short *)g_luts)[i];
    for (int i = tid; i < 512; i += T1_THREADS) zc[i] = zc_lut[orient * 512 + i];
    if (tid == 0) { s_max = 0; s_nmsedec[0] = s_nmsedec[1] = s_nmsedec[2] = 0; }
    __syncthreads();

    // ---- quantise: lrintf((coef / stepsize) * 64) ------------------------------------------------
    unsigned mag[T1_CPT][4];
    int sidx[T1_CPT], lim[T1_CPT];
    unsigned sgn4[T1_CPT];
    const float stepsize = bd->stepsize;
    const float *src = coef + (size_t)frame * H * W + (size_t)cb.ty * W + cb.tx;
    unsigned char *recp = REC ? rec_pass + (size_t)frame * H * W + (size_t)cb.ty * W + cb.tx : nullptr;
    int *recv = REC ? rec_val + (size_t)frame * H * W + (size_t)cb.ty * W + cb.tx : nullptr;
    int mymax = 0;
#pragma unroll
    for (int k = 0; k < T1_CPT; ++k) {
        int q = tid * T1_CPT + k;
        sidx[k] = -1; lim[k] = 0; sgn4[k] = 0;
#pragma unroll
        for (int j = 0; j < 4; ++j) mag[k][j] = 0;
        if (q < nsc) {
            int s = q / bw, x = q - s * bw;
            sidx[k] = (s + 1) * CF_W + x + 1;
            lim[k] = min(4, bh - 4 * s);
#pragma unroll
            for (int j = 0; j < 4; ++j)
                if (j < lim[k]) {
                    float c = src[(size_t)(4 * s + j) * W + x];
                    int v = __float2int_rn(__fmul_rn(__fdiv_rn(c, stepsize), 64.0f));
                    int a = v < 0 ? -v : v;
                    mag[k][j] = (unsigned)a;
                    if (v < 0) sgn4[k] |= 1u << j;
                    mymax = max(mymax, a);
                    if (REC) {
                        const int m = a >> 6;
                        recv[(4 * s + j) * W + x] = m ? (v < 0 ? -(2 * m + 1) : 2 * m + 1) : 0;
                    }
                }
            N[sidx[k]] = sgn4[k];          // staging for SG
        }
    }
    for (int o = 16; o > 0; o >>= 1) mymax = max(mymax, __shfl_xor_sync(0xffffffffu, mymax, o));
    if ((tid & 31) == 0) atomicMax(&s_max, mymax);
    __syncthreads();
#pragma unroll
    for (int k = 0; k < T1_CPT; ++k)
        if (sidx[k] >= 0) SG[sidx[k]] = ((N[sidx[k] - CF_W] >> 3) & 1u) | (sgn4[k] << 1) | ((N[sidx[k] + CF_W] & 1u) << 5);
    const int maxv = s_max;
    const int numbps = maxv ? (32 - __clz(maxv)) - 6 : 0;     // floorlog2(max)+1-T1_NMSEDEC_FRACBITS
    if (tid == 0) { out->numbps = numbps; out->totalpasses = 0; out->datalen = 0; out->numpasses = 0; }
    if (numbps <= 0) return;
    __syncthreads();
#pragma unroll
    for (int k = 0; k < T1_CPT; ++k) if (sidx[k] >= 0) N[sidx[k]] = 0;
    __syncthreads();

    // ---- bit-planes.  The (up to) three passes of a plane run back to back, each thread writing its symbols into
    // its private slot; ONE packed prefix sum per plane then places the three symbol runs in the stream. -----------
    unsigned total_syms = 0;
    int passno = 0;
    unsigned char *slot = priv + tid * T1_SLOT;
    for (int bpno = numbps - 1; bpno >= 0; --bpno) {
        const bool first = bpno == numbps - 1;          // the top plane has a clean-up pass only
        const unsigned one = 1u << (bpno + 6);
        T1Sym sq[3];
        int nm[3] = {0, 0, 0};
#pragma unroll
        for (int q = 0; q < 3; ++q) { sq[q].p = slot + q * (10 * T1_CPT); sq[q].cnt = 0; }

        if (!first) {
            // ---- significance propagation: least fixed point of "becomes significant", then one emitting sweep ----
            // todo[k]: insignificant rows whose bit is set in this plane and that are not fresh yet -- the only rows a
            // sweep can still change; columns without any skip the sweeps altogether
            unsigned acc[T1_CPT], todo[T1_CPT];
#pragma unroll
            for (int k = 0; k < T1_CPT; ++k) {
                acc[k] = 0; todo[k] = 0;
                if (sidx[k] < 0) continue;
                unsigned bits = 0;
#pragma unroll
                for (int j = 0; j < 4; ++j) bits |= ((mag[k][j] & one) ? 1u : 0u) << j;
                todo[k] = ~cf_own4(F[sidx[k]]) & bits & ((1u << lim[k]) - 1u);
            }
            for (;;) {
                int changed = 0;
#pragma unroll
                for (int k = 0; k < T1_CPT; ++k) {
                    if (!todo[k]) continue;
                    const int p = sidx[k];
                    const unsigned f = F[p];
                    unsigned dummy_c = 0; int dummy_n = 0;
                    const unsigned fr = cf_sigpass_col<false>(f, N[p], mag[k], one, lim[k], sgn4[k], zc, SG, p, &luts, bpno, sq[0], dummy_c, dummy_n);
                    const unsigned delta = fr & ~acc[k];
                    if (delta) { cf_scatter(N, p, delta); acc[k] |= delta; todo[k] &= ~delta; changed = 1; }
                }
                if (!__syncthreads_or(changed)) break;
            }
#pragma unroll
            for (int k = 0; k < T1_CPT; ++k) {
                if (sidx[k] < 0) continue;
                const int p = sidx[k];
                const unsigned f = F[p];
                unsigned coded = 0;
                if ((~cf_own4(f) & ((1u << lim[k]) - 1u)) != 0) {
                    const unsigned fr = cf_sigpass_col<true>(f, N[p], mag[k], one, lim[k], sgn4[k], zc, SG, p, &luts, bpno, sq[0], coded, nm[0]);
                    if (REC && fr) t1_rec_pass(recp, W, p, fr, passno);
                }
                // state: the word absorbs what its neighbourhood gained, rows coded here are marked pi
                const unsigned w = f | N[p] | (coded << 24);
                F[p] = w;
                N[p] = 0;
                // ---- magnitude refinement (t1.c:743-765, 814-922): needs nothing but the own word ----
                const unsigned sig = cf_own4(w), pi = (w >> 24) & 15u;
                if ((sig & ~pi & ((1u << lim[k]) - 1u)) == 0) continue;
                unsigned mu = 0;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    if (j >= lim[k]) break;
                    if (((sig >> j) & 1) && !((pi >> j) & 1)) {
                        const int ctx = (w & CF_MU(j)) ? CTX_MAG + 2 : ((w & CF_NB(j)) ? CTX_MAG + 1 : CTX_MAG);
                        nm[1] += t1_nm_ref(&luts, mag[k][j], bpno);
                        t1_put(sq[1], ctx, (mag[k][j] & one) ? 1 : 0);
                        mu |= 1u << j;
                    }
                }
                F[p] = w | (mu << 28);
            }
            __syncthreads();            // every N word is zero again before the clean-up pass scatters into it
        }

        {
            // ---- cleanup with run-length mode (t1.c:1051-1116, 1163-1256): what becomes significant is known up front ----
            unsigned fresh[T1_CPT];
#pragma unroll
            for (int k = 0; k < T1_CPT; ++k) {
                fresh[k] = 0;
                if (sidx[k] < 0) continue;
                const unsigned f = F[sidx[k]];
                const unsigned cand = ~cf_own4(f) & ~(f >> 24) & ((1u << lim[k]) - 1u);
                if (!cand) continue;
                unsigned bits = 0;
#pragma unroll
                for (int j = 0; j < 4; ++j) bits |= ((mag[k][j] & one) ? 1u : 0u) << j;
                fresh[k] = cand & bits;
                if (fresh[k]) {
                    cf_scatter(N, sidx[k], fresh[k]);
                    if (REC) t1_rec_pass(recp, W, sidx[k], fresh[k], first ? passno : passno + 2);
                }
            }
            __syncthreads();
#pragma unroll
            for (int k = 0; k < T1_CPT; ++k) {
                if (sidx[k] < 0) continue;
                const int p = sidx[k];
                const unsigned f = F[p];
                const unsigned sig = cf_own4(f), pi = (f >> 24) & 15u, rows = (1u << lim[k]) - 1u;
                if (((sig | pi) & rows) != rows) {
                    unsigned t = f | (N[p] & CF_EARLY);
                    int runlen = 0, agg = 0;
                    if (lim[k] == 4 && (t & 0x0f03ffffu) == 0) {
                        agg = 1;
                        runlen = fresh[k] ? (__ffs((int)fresh[k]) - 1) : 4;
                        t1_put(sq[2], CTX_AGG, runlen != 4);
                        if (runlen != 4) { t1_put(sq[2], CTX_UNI, runlen >> 1); t1_put(sq[2], CTX_UNI, runlen & 1); }
                    }
                    if (runlen < 4) {
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            if (j >= lim[k]) break;
                            if (j < runlen) continue;
                            int v;
                            if (agg && j == runlen) v = 1;
                            else if (!(((sig | pi) >> j) & 1)) {
                                v = (mag[k][j] & one) ? 1 : 0;
                                t1_put(sq[2], CTX_ZC + zc[(t >> (j * 3)) & 0x1ffu], v);
                            } else v = 0;
                            if (v) {
                                cf_sign(sq[2], t, SG, p, j, (sgn4[k] >> j) & 1);
                                nm[2] += t1_nm_sig(&luts, mag[k][j], bpno);
                                t |= CF_SIG(j);
                            }
                        }
                    }
                }
                F[p] = (f | N[p]) & ~(0xfu << 24);                // new significance, clear pi
                N[p] = 0;
            }
        }

        // ---- one CTA-wide exclusive scan of the three symbol counts (21-bit fields of a 64-bit word) ----
        const unsigned long long mine = (unsigned long long)sq[0].cnt | ((unsigned long long)sq[1].cnt << 21) | ((unsigned long long)sq[2].cnt << 42);
        unsigned long long incl = mine;
        for (int o = 1; o < 32; o <<= 1) { unsigned long long t = __shfl_up_sync(0xffffffffu, incl, o); if ((tid & 31) >= o) incl += t; }
        if ((tid & 31) == 31) warp_sum[tid >> 5] = incl;
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            int v = nm[q];
            for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            if ((tid & 31) == 0 && v) atomicAdd(&s_nmsedec[q], v);
        }
        __syncthreads();
        unsigned long long excl = incl - mine, total3;
        {
            const int nw = T1_THREADS / 32;
            const unsigned long long wv = (tid & 31) < nw ? warp_sum[tid & 31] : 0ull;
            unsigned long long ex = wv;
            for (int o = 1; o < 32; o <<= 1) { unsigned long long t = __shfl_up_sync(0xffffffffu, ex, o); if ((tid & 31) >= o) ex += t; }
            excl += __shfl_sync(0xffffffffu, ex - wv, tid >> 5);
            total3 = __shfl_sync(0xffffffffu, ex, nw - 1);
        }
        // ---- the runs leave for the stream: pass q starts where pass q-1 ended ----
        unsigned base = total_syms;
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            const unsigned tot = (unsigned)(total3 >> (21 * q)) & 0x1fffffu, off = (unsigned)(excl >> (21 * q)) & 0x1fffffu;
            unsigned char *dst = sym + base + off;
            // the slice holds the worst case (1.5 symbols per sample and plane + one sign, build_plan); checked all the same
            const int room = (int)cb.sym_cap - (int)(base + off), cn = min(sq[q].cnt, max(room, 0));
            for (int i = 0; i < cn; ++i) dst[i] = sq[q].p[i];
            base += tot;
            if (tid == 0 && (q == 2 || !first)) {
                const int pn = first ? passno : passno + q;
                out->nsym_end[pn] = base; out->nmsedec[pn] = s_nmsedec[q];
            }
        }
        total_syms = base;
        passno += first ? 1 : 3;
        __syncthreads();
        if (tid < 3) s_nmsedec[tid] = 0;
    }

    // ---- distortion per pass: opj_t1_getwmsedec, accumulated in pass order (t1.c:1416-1449, 2655) --
    if (tid == 0) {
        if (total_syms > cb.sym_cap) atomicOr(&stat[frame].flags, STAT_OVERFLOW);
        out->totalpasses = passno;
        const double wn = bd->wnorm;
        double cum = 0.0;
        for (int p = 0; p < passno; ++p) {
            int bp = p == 0 ? numbps - 1 : numbps - 2 - (p - 1) / 3;
            double w = __dmul_rn(wn, (double)(1 << bp));
            w = __dmul_rn(w, __ddiv_rn(__dmul_rn(w, (double)out->nmsedec[p]), 8192.0));
            cum = __dadd_rn(cum, w);
            out->dist[p] = cum;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// MQ coder.  State table (Annex C table C.2; mqc.c:61-157) arrives packed as qe | nmps<<16 | nlps<<22 | sw<<28
// and is expanded to one word per (state, mps) pair, index p = 2 * state + mps:
//   word(p) = qe << 16 | p << 3 | mps
// Shared memory holds, per pair, the words of its two successors {after an MPS, after an LPS}.  A context is
// a 16-byte record {word, word after MPS, word after LPS}: a state transition is a register select, and the
// successors of the new state are fetched off the critical path.  The symbol byte is ctx << 2 | decision.
struct MqTable { unsigned e[47]; };

#define MQ_WARPS(L) ((L) >= 32 ? 2 : 4)       // warps per CTA (static shared memory stays below 48 KB)
#define MQ_THREADS(L) (MQ_WARPS(L) * 32)
#define MQ_DEFAULT_LANES 32

// shared-memory accesses of the serial chain by 32-bit shared address: one LDS / STS each, no generic
// address arithmetic (the host-side emulation build uses plain pointers)
#ifdef EBCC_CUEMU
typedef unsigned long long mq_saddr;
#define MQ_SADDR(p) ((mq_saddr)(size_t)(p))
#define MQ_OPAQUE(a) (a)
__device__ __forceinline__ unsigned mq_ld8(mq_saddr a) { return *(const unsigned char *)(size_t)a; }
__device__ __forceinline__ unsigned mq_ld32(mq_saddr a) { return *(const unsigned *)(size_t)a; }
__device__ __forceinline__ uint2 mq_ld64(mq_saddr a) { return *(const uint2 *)(size_t)a; }
__device__ __forceinline__ uint4 mq_ld128(mq_saddr a) { return *(const uint4 *)(size_t)a; }
__device__ __forceinline__ void mq_st128(mq_saddr a, unsigned x, unsigned y, unsigned z) { *(uint4 *)(size_t)a = make_uint4(x, y, z, 0); }
#else
typedef unsigned mq_saddr;
#define MQ_SADDR(p) ((mq_saddr)__cvta_generic_to_shared(p))
// keeps an address in a register instead of re-deriving the shared window base at every use
__device__ __forceinline__ unsigned mq_opaque(unsigned a) { unsigned r; asm volatile("mov.u32 %0, %1;" : "=r"(r) : "r"(a)); return r; }
#define MQ_OPAQUE(a) mq_opaque(a)
__device__ __forceinline__ unsigned mq_ld8(mq_saddr a) { unsigned v; asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(a)); return v; }
__device__ __forceinline__ unsigned mq_ld32(mq_saddr a) { unsigned v; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a)); return v; }
__device__ __forceinline__ uint2 mq_ld64(mq_saddr a) { uint2 v; asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(a)); return v; }
__device__ __forceinline__ uint4 mq_ld128(mq_saddr a) { uint4 v; asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a)); return v; }
__device__ __forceinline__ void mq_st128(mq_saddr a, unsigned x, unsigned y, unsigned z) { asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" :: "r"(a), "r"(x), "r"(y), "r"(z), "r"(0u)); }
#endif

struct MqEnc {
    unsigned a, c;
    int ct;
    int pos;                  // index of the pending (not yet stored) byte, -1 = OPJ's fake byte
    int cap;                  // bytes the block's slice of the MQ pool holds
    unsigned pend;
    unsigned char *dst;
};

__device__ __forceinline__ void mq_advance(MqEnc &q, unsigned nb)
{
    if (q.pos >= 0 && q.pos < q.cap) q.dst[q.pos] = (unsigned char)q.pend;
    q.pos++; q.pend = nb;
}

__device__ __forceinline__ void mq_byteout(MqEnc &q)
{
    if (q.pend == 0xff) {
        mq_advance(q, (q.c >> 20) & 0xff); q.c &= 0xfffff; q.ct = 7;
    } else if ((q.c & 0x8000000) == 0) {
        mq_advance(q, (q.c >> 19) & 0xff); q.c &= 0x7ffff; q.ct = 8;
    } else {
        q.pend++;
        if (q.pend == 0xff) {
            q.c &= 0x7ffffff;
            mq_advance(q, (q.c >> 20) & 0xff); q.c &= 0xfffff; q.ct = 7;
        } else {
            mq_advance(q, (q.c >> 19) & 0xff); q.c &= 0x7ffff; q.ct = 8;
        }
    }
}

// Is "decode passes 0..p from the segment cut at R bytes" the same as decoding them from the complete segment?
// OPJ cuts a non-terminated pass at numbytes + 3 (t1.c:2673-2683), which is usually but not always enough; the
// compression-ratio search has to know, because its probes must see what a real decoder sees.  The MQ decoder is an
// exact interval decoder: it reproduces the symbols up to the end of pass p if and only if the value it reads lies in
// the coder's interval [L, L + a) at that point.  Beyond the cut it reads the synthetic 0xFF 0xFF marker of
// opj_mqc_init_dec, i.e. one-bits for ever (mqc_inl.h:77-99: c += 0xff00, ct = 8).  In units of a's LSB, relative to
// everything before the coder's pending byte (index pos, value pend, may still take a carry):
//     L = pend * 2^s + c,  s = 27 - ct        (bit 27 - ct of c is the carry into the pending byte, byteout mqc.c:492-523)
//     X = sum_{i = pos}^{R-1} v[i] * 2^(s - off_i)  +  2^(s - off_{R-1})      (final bytes, then the ones)
//         off_pos = 0, off_{i+1} = off_i + (v[i] == 0xff ? 7 : 8)               (bit stuffing)
// and the cut is safe iff X <= L + a (X minus "an infinitesimal" is what the decoder sees).  Everything is scaled by
// 2^8 so that the exponents stay non-negative (R - 1 - pos <= 2: at most two whole bytes lie behind the pending one).
__device__ __forceinline__ bool mq_cut_proven(unsigned a, unsigned c, int ct, unsigned pend, int pos, int R, const unsigned char *v)
{
    if (R - 1 < pos || R - 1 - pos > 2) return false;       // cannot happen (rate >= numbytes + 1); a real decode settles it
    const int s = 27 - ct + 8;
    unsigned long long X = 0;
    int off = 0, last = 0;
    for (int i = pos; i < R; ++i) {
        const unsigned vi = i < 0 ? 0u : (unsigned)v[i];      // index -1: OPJ's fake byte in front of the segment, never carried into
        X += (unsigned long long)vi << (s - off);
        last = off;
        off += vi == 0xffu ? 7 : 8;
    }
    X += 1ull << (s - last);
    const unsigned long long U = ((((unsigned long long)pend << (27 - ct)) + c + a) << 8);
    return X <= U;
}

// end of coding pass p: segment length as opj_t1_encode_cblk records it (t1.c:2655-2699)
__device__ __forceinline__ void mq_pass_end(MqEnc &q, CblkOut *out, int p, int npasses, bool snap)
{
    if (snap) { out->mqs[p][0] = q.c | ((unsigned)q.ct << 28); out->mqs[p][1] = q.a | (q.pend << 16); out->mqs[p][2] = (unsigned)q.pos; }
    if (p == npasses - 1) {
        // opj_mqc_flush (mqc.c:212-227)
        unsigned tempc = q.c + q.a;
        q.c |= 0xffffu;
        if (q.c >= tempc) q.c -= 0x8000u;
        q.c <<= q.ct; mq_byteout(q);
        q.c <<= q.ct; mq_byteout(q);
        if (q.pend != 0xff) mq_advance(q, 0);
        out->rate[p] = q.pos;
    } else {
        out->rate[p] = q.pos + 3;
    }
}

// L code-blocks per warp (L = 1, 2, 4, ... 32).  The 32 / L lanes of a group move their block's symbol stream
// HBM -> registers -> shared memory in refills of up to 512 bytes (the next refill is in flight while the current one
// is coded); the first lane of each group runs the serial MQ chain out of shared memory.  The coding lanes
// advance in lock step: every stretch of symbols up to the nearest pass end of any of them has a uniform
// trip count, the common step (MPS, interval still normalised) is shared by all lanes, and the state
// update + renormalisation is one predicated region so that only byte output really diverges.
#ifndef MQ_MIN_CTAS
#define MQ_MIN_CTAS 1
#endif
template <int L>
__global__ void __launch_bounds__(MQ_THREADS(L), MQ_MIN_CTAS)
k_t1_mq(const CblkDesc *__restrict__ cblks, CblkOut *cbout, const unsigned char *__restrict__ sym_pool,
        unsigned char *mq_pool, const MqTable *__restrict__ g_tab, int ncblks, int nframes,
        size_t sym_stride, size_t mq_stride, T1CutState *cuts, FrameStat *stat, unsigned char *ctxsnap, int dbg_all_unproven)
{
    // cuts != nullptr (compression-ratio search): the coder state at every pass end is kept and, once the segment is
    // complete, every pass gets its verdict "cut at rate[p] decodes like the complete segment" (mq_cut_proven)
    const bool snap = cuts != nullptr;
    constexpr int G = 32 / L;                     // lanes per code-block
    constexpr int R = L < 8 ? L : 8;              // 16-byte loads per lane and refill
    constexpr int CH = 16 * G * R;                // symbol bytes per refill and block: 512 (L <= 8), 256, 128
    constexpr unsigned CM = L == 1 ? 0x1u : (L == 2 ? 0x00010001u : (L == 4 ? 0x01010101u : (L == 8 ? 0x11111111u :
                            (L == 16 ? 0x55555555u : 0xffffffffu))));
    __shared__ __align__(16) uint2 tabn[96];      // pair -> {word after MPS, word after LPS}
    __shared__ __align__(16) uint4 ctxs[MQ_WARPS(L) * L][20];
    __shared__ __align__(16) unsigned char stage[MQ_WARPS(L) * L][CH + 16];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    auto word = [&](unsigned pr) { return ((g_tab->e[pr >> 1] & 0xffffu) << 16) | (pr << 3) | (pr & 1u); };
    auto succ = [&](unsigned pr) {
        const unsigned e = g_tab->e[pr >> 1], m = pr & 1u;
        return make_uint2(word(((e >> 16) & 63u) * 2u + m), word(((e >> 22) & 63u) * 2u + (m ^ ((e >> 28) & 1u))));
    };
    for (int i = tid; i < 94; i += MQ_THREADS(L)) tabn[i] = succ((unsigned)i);
    __syncthreads();
    const int g = lane / G, gl = lane - g * G, blk = wid * L + g;
    const bool coder = gl == 0;
    const long long job = ((long long)blockIdx.x * MQ_WARPS(L) + wid) * L + g;
    const bool valid = job < (long long)ncblks * nframes;
    // frame varies fastest: the lanes of a warp code the same code-block position of different frames, whose
    // symbol counts are similar
    const int cbi = valid ? (int)(job / nframes) : 0, frame = valid ? (int)(job - (long long)cbi * nframes) : 0;
    const CblkDesc cb = cblks[cbi];
    CblkOut *out = cbout + (size_t)frame * ncblks + cbi;
    const int npasses = valid ? max(out->totalpasses, 0) : 0;
    const unsigned total = npasses > 0 ? out->nsym_end[npasses - 1] : 0u;
    unsigned maxtotal = total;
    for (int o = 16; o > 0; o >>= 1) maxtotal = max(maxtotal, __shfl_xor_sync(0xffffffffu, maxtotal, o));
    const unsigned char *sym = sym_pool + (size_t)frame * sym_stride + cb.sym_off;

    uint4 *cx = ctxs[blk];
    for (int k = gl; k < 19; k += G) {
        const unsigned pr = k == CTX_UNI ? 92u : (k == CTX_AGG ? 6u : (k == CTX_ZC ? 8u : 0u));
        const uint2 nx = tabn[pr];
        cx[k] = make_uint4(word(pr), nx.x, nx.y, 0);
    }
    for (int k = gl; k < 16; k += G) stage[blk][CH + k] = 0;       // look-ahead slack behind the refill

    MqEnc q;
    q.a = 0x8000; q.c = 0; q.ct = 12; q.pos = -1; q.pend = 0; q.cap = (int)cb.mq_cap;
    q.dst = mq_pool + (size_t)frame * mq_stride + cb.mq_off;

    unsigned char *stg = stage[blk];
    const mq_saddr sba = MQ_SADDR(stg), cxa = MQ_SADDR(cx), taba = MQ_OPAQUE(MQ_SADDR(tabn));
    unsigned i = 0, end = npasses > 0 ? out->nsym_end[0] : 0xffffffffu;
    int p = 0;
    // context states at the end of clean-up pass pp (passes 0, 3, 6, ...), see T1_SNAP_PLANES
    auto save_ctx = [&](int pp) {
        if (!snap || ctxsnap == nullptr || pp % 3 != 0 || pp / 3 >= T1_SNAP_PLANES) return;
        unsigned char *d = ctxsnap + (((size_t)frame * ncblks + cbi) * T1_SNAP_PLANES + pp / 3) * T1_SNAP_BYTES;
        for (int k = 0; k < 19; ++k) d[k] = (unsigned char)((mq_ld32(cxa + (mq_saddr)(k * 16)) >> 3) & 127u);
    };
    uint4 pre[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const unsigned o = (gl + r * G) * 16u;
        pre[r] = o < total ? *(const uint4 *)(sym + o) : make_uint4(0, 0, 0, 0);
    }
    for (unsigned cbeg = 0; cbeg < maxtotal; cbeg += CH) {
        __syncwarp();
#pragma unroll
        for (int r = 0; r < R; ++r) *(uint4 *)(stg + (gl + r * G) * 16) = pre[r];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const unsigned o = cbeg + CH + (gl + r * G) * 16u;
            if (o < total) pre[r] = *(const uint4 *)(sym + o);
        }
        __syncwarp();
        if (!coder) continue;
        const unsigned cend = min(cbeg + CH, total);
        for (;;) {
            const bool act = i < cend;
            const unsigned m = __ballot_sync(CM, act);
            if (m == 0) break;
            // symbols every active lane can code before one of them reaches a pass end (or the refill ends)
            unsigned n = act ? min(end, cend) - i : 0xffffffffu;
            for (int o = G; o < 32; o <<= 1) n = min(n, __shfl_xor_sync(CM, n, o));
            if (!act) continue;
            // software pipeline: symbol k + 2 and the context record of symbol k + 1 are in flight while
            // symbol k is coded; a record changed by symbol k is forwarded when symbol k + 1 uses it too
            mq_saddr sp = sba + (i - cbeg);
            unsigned s0 = mq_ld8(sp), s1 = mq_ld8(sp + 1);
            mq_saddr ca = cxa + ((s0 & 0xfcu) << 2);
            uint4 r = mq_ld128(ca);
#pragma unroll 2
            for (unsigned k = 0; k < n; ++k) {
                const mq_saddr can = cxa + ((s1 & 0xfcu) << 2);
                uint4 rn = mq_ld128(can);
                ++sp;
                const unsigned s2 = mq_ld8(sp + 1);
                const unsigned qe = r.x >> 16, lps = (s0 ^ r.x) & 1u;
                q.a -= qe; q.c += qe;                     // the MPS / no-renormalisation outcome
                if (lps | (~q.a & 0x8000u)) {
                    // MPS: a < qe ? a = qe : c += qe;  LPS: a < qe ? c += qe : a = qe   (conditional exchange)
                    if ((unsigned)(q.a < qe) != lps) { q.a = qe; q.c -= qe; }
                    const unsigned w = lps ? r.z : r.y;
                    const uint2 nx = mq_ld64(taba + (w & 0x3f8u));
                    // renormalise: nsh shifts bring bit 15 of a back up; a byte leaves every time ct reaches 0
                    int nsh = __clz((int)q.a) - 16;
                    q.a <<= nsh;
                    while (nsh >= q.ct) { q.c <<= q.ct; nsh -= q.ct; mq_byteout(q); }
                    q.c <<= nsh; q.ct -= nsh;
                    mq_st128(ca, w, nx.x, nx.y);
                    if (can == ca) { rn.x = w; rn.y = nx.x; rn.z = nx.y; }
                }
                s0 = s1; s1 = s2; r = rn; ca = can;
            }
            i += n;
            if (i == end) {
                mq_pass_end(q, out, p, npasses, snap);
                save_ctx(p);
                end = ++p < npasses ? out->nsym_end[p] : 0xffffffffu;
            }
        }
    }
    if (coder && valid && snap && npasses <= 0) {
        T1CutState &cs = cuts[(size_t)frame * ncblks + cbi];
        cs.np_dec = 0; cs.scratch_np = 0; cs.bad[0] = cs.bad[1] = cs.bad[2] = cs.bad[3] = 0; cs.first_entry = -1; cs.pad = 0;
    }
    if (!coder || npasses <= 0) return;
    for (; p < npasses; ++p) { mq_pass_end(q, out, p, npasses, snap); save_ctx(p); }
    // final pending byte (beyond numbytes) is never part of the stream
    int last = q.pos;
    if (last > q.cap) { atomicOr(&stat[frame].flags, STAT_OVERFLOW); last = q.cap; }
    out->datalen = last;
    // monotone fix-up, then strip a trailing 0xFF from each pass (t1.c:2700-2723)
    for (int pp = npasses - 1; pp >= 0; --pp) {
        int r = out->rate[pp];
        if (r > last) out->rate[pp] = last; else last = r;
    }
    for (int pp = 0; pp < npasses; ++pp) {
        int r = out->rate[pp];
        if (r > 0 && q.dst[r - 1] == 0xff) out->rate[pp] = r - 1;
    }
    if (snap) {
        unsigned bad[4] = {0, 0, 0, 0};
        for (int pp = 0; pp < npasses; ++pp) {
            const unsigned w0 = out->mqs[pp][0], w1 = out->mqs[pp][1];
            const bool ok = !dbg_all_unproven && pp < 128 && mq_cut_proven(w1 & 0xffffu, w0 & 0x0fffffffu, (int)(w0 >> 28), (w1 >> 16) & 0xffu, (int)out->mqs[pp][2],
                                                      out->rate[pp], q.dst);
            if (!ok && pp < 128) bad[pp >> 5] |= 1u << (pp & 31);
        }
        T1CutState &cs = cuts[(size_t)frame * ncblks + cbi];
        cs.np_dec = npasses; cs.scratch_np = 0; cs.first_entry = -1; cs.pad = 0;
        cs.bad[0] = bad[0]; cs.bad[1] = bad[1]; cs.bad[2] = bad[2]; cs.bad[3] = bad[3];
    }
}
