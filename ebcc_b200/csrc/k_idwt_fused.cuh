// k_idwt_fused.cuh -- the 9/7 transform of the JPEG 2000 layer, one launch per resolution level and direction-fused:
// a CTA holds a tile of the level in shared memory and runs BOTH 1-D passes on it, the lifting itself in registers.
//
//   k_idwt_tile : inverse transform (OpenJPEG opj_dwt_decode_tile_97, dwt.c:3512-3733, 1-D kernel opj_v8dwt_decode
//                 :3353-3428): per level every row is synthesised horizontally first, then every column vertically.
//                 The stage in front is fused into the loads (dequantisation of the tier-1 output, t1.c:1853-1865, or
//                 the probe's reconstruction from the encoder's recording) and, at the last level, the stage behind it
//                 into the stores (DC shift / clamp / float rescale, tcd.c:2320-2340 + src/ebcc_codec.c:1129-1131, and
//                 the error statistics of the compression-ratio search, src/ebcc_codec.c:464-543).
//   k_dwt_tile  : forward transform (opj_dwt_encode_real, dwt.c:1930-2108, lifting :1222-1300): vertical pass, then
//                 horizontal pass; the first level reads the raw frame (u16 scaling + DC shift on the fly), the high
//                 bands go straight to their Mallat place, the LL band to a work plane the next level reads.
//
// Arithmetic and operation order are those of the whole-line kernels (k_dwt.cuh, k_idwt.cuh): every lifting update is
// x += (l + r) * c with separate IEEE add / mul / add.  A sample of a level's output depends on the level's input within
// +-4 positions in each direction, so a tile of TW x TH outputs is computed from (TW + 8) x (TH + 8) inputs and its own
// samples go through exactly the operations the whole-line kernels apply to them.  Image borders: OpenJPEG's edge rules
// ("x[1] twice" at the start of a line, (c + c) * x or (2 * x) * c at its end) are what the generic update computes on the
// whole-sample symmetric extension x[-k] = x[k], x[n-1+k] = x[n-1-k] -- float addition commutes and (2c) * x, (2x) * c round
// the same real product -- so the loads mirror and the compute phases carry no border cases.
//
// Phases of the inverse kernel (256 threads, tile = 120 x 32 outputs):
//   1 load    : one warp per tile row; source -> shared memory, columns de-interleaved [low | high], horizontal scaling folded in
//   2 rows    : one thread per (row, 8 outputs): 16 inputs by four 16-byte loads, 22 lifting updates in registers, outputs to a
//               second tile in natural column order
//   3 columns : one thread per (column, 16 output rows): 24 inputs, 38 lifting updates in registers, results straight to
//               global memory (coalesced across the threads of a warp), through the finalisation at the last level
// The forward kernel mirrors it (load natural, columns, rows, coalesced de-interleaved store).
//
// HBM traffic per level of rw x rh samples: inverse -- read 3/4 of the inputs from the source planes (5 bytes per sample for
// a probe: value + pass of significance; 4 for a decode), 1/4 from the previous level's output, write the outputs once (the
// last level writes the decoded frame, or nothing but statistics for a search probe, and reads the original frame);
// forward -- read the inputs once, write the outputs once.  Levels ping-pong between two work planes because a CTA must
// not overwrite inputs another CTA still reads.
#pragma once
#include "common.cuh"
#include "k_idwt.cuh"
#include "k_dec.cuh"

#define FT_TW 120                        // outputs per tile row
#define FT_TH 32                         // output rows per tile
#define FT_HALO 4
#define FT_COLS (FT_TW + 2 * FT_HALO)    // 128 loaded columns
#define FT_ROWS (FT_TH + 2 * FT_HALO)    // 40 loaded rows
#define FT_HALF (FT_COLS / 2)            // 64 low-pass (and high-pass) columns per row
#define FT_HOFF 80                       // first high-pass column of a de-interleaved row; 80 % 32 == 16: (de-)interleaving accesses of a warp hit 32 banks
#define FT_P1 144                        // floats per de-interleaved row (16-byte aligned halves)
#define FT_P2 128                        // floats per natural-order row
#define FT_SEGS (FT_TW / 8)              // 8-output segments per row
#define FT_THREADS 256

enum { IDF_SRC_IPLANE = 0, IDF_SRC_RECON = 1 };

// compact per-block record of a probe (written by k_jobs_from_enc next to the DecJob): what the reconstruction of a sample needs
//   x = mode | passes of the allocation << 8 | passes the recording holds << 16 | bit-planes << 24,   y = entry in the pool of pre-decoded cuts
struct IdwtFusedArgs {
    const Plan *plan; const CblkDesc *cblks;
    const int *iplane;                   // tier-1 output / recording: value per sample
    const unsigned char *spass;          // recording: pass of significance per sample          (probe)
    const int *cut_pool; const uint2 *rjobs;                                                 // (probe)
    const float *ll_in;                  // previous level's output (LL of this level), row stride W; unused at level 1
    float *out;                          // this level's output, row stride W (last level: decoded frame, may be NULL)
    int r, rw, rh, snw, snh;             // level, its size, low-pass counts
    const FrameStat *stat; const float *orig; const float *targets; ErrStat *es;      // last level only
    unsigned long long cap;              // last level of a search probe: a frame whose violation count has reached cap is not looked at any further (0: exact counts)
};

// whole-sample symmetric extension of a line of n samples
__device__ __forceinline__ int ft_reflect(int p, int n)
{
    if (p >= 0 && p < n) return p;
    if (n < 2) return 0;
    const int per = 2 * n - 2;
    p %= per;
    if (p < 0) p += per;
    return p < n ? p : per - p;
}

// the 4 lifting steps on a register window of 8 low-pass + 8 high-pass samples of a row (position of L[i]: 2i, of H[i]: 2i + 1);
// afterwards L[2..5] and H[2..5] are final.  Inverse: low, high, low, high with -delta, -gamma, -beta, -alpha; forward: high, low,
// high, low with alpha, beta, gamma, delta.
#define FT_UPD(x, l, r, c) x = __fadd_rn(x, __fmul_rn(__fadd_rn(l, r), c))
template <bool INV>
__device__ __forceinline__ void ft_row_window(float (&L)[8], float (&H)[8])
{
    if (INV) {
#pragma unroll
        for (int i = 1; i <= 7; ++i) FT_UPD(L[i], H[i - 1], H[i], -DWT_DELTA);
#pragma unroll
        for (int i = 1; i <= 6; ++i) FT_UPD(H[i], L[i], L[i + 1], -DWT_GAMMA);
#pragma unroll
        for (int i = 2; i <= 6; ++i) FT_UPD(L[i], H[i - 1], H[i], -DWT_BETA);
#pragma unroll
        for (int i = 2; i <= 5; ++i) FT_UPD(H[i], L[i], L[i + 1], -DWT_ALPHA);
    } else {
#pragma unroll
        for (int i = 0; i <= 6; ++i) FT_UPD(H[i], L[i], L[i + 1], DWT_ALPHA);
#pragma unroll
        for (int i = 1; i <= 6; ++i) FT_UPD(L[i], H[i - 1], H[i], DWT_BETA);
#pragma unroll
        for (int i = 1; i <= 5; ++i) FT_UPD(H[i], L[i], L[i + 1], DWT_GAMMA);
#pragma unroll
        for (int i = 2; i <= 5; ++i) FT_UPD(L[i], H[i - 1], H[i], DWT_DELTA);
    }
}
// the 4 lifting steps on a register window of 24 consecutive samples of a column (local index parity = parity of the canvas
// row); afterwards v[4..19] are final.  Step k updates the positions of parity P0 + k within [k + 1, 22 - k].
template <int P0>
__device__ __forceinline__ void ft_col_window(float (&v)[24], float c0, float c1, float c2, float c3)
{
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const float c = k == 0 ? c0 : (k == 1 ? c1 : (k == 2 ? c2 : c3));
        const int p = (P0 + k) & 1, lo = k + 1, hi = 22 - k;
        const int first = lo + ((lo ^ p) & 1);
#pragma unroll
        for (int i = first; i <= hi; i += 2) FT_UPD(v[i], v[i - 1], v[i + 1], c);
    }
}

// j2k_pixel_to_float (k_idwt.cuh) with cheaper instructions and the same bits:
//   * lrintf + 32768 clamped to [0, 65535]: the float -> int conversion saturates, so clamp(t, -32768, 32767) + 32768 in 32-bit
//     integers equals the 64-bit form for every input (NaN converts to 0 in both);
//   * p / 65535 in f32: the double product p * (1 / 65535) rounded to f32 is the correctly rounded quotient -- p < 2^16, so the
//     exact quotient is at least 2^-41 (relative) away from any f32 rounding boundary unless it is exact, far more than the
//     2^-52 the double arithmetic can be off -- hence identical to __fdiv_rn.
__device__ __forceinline__ float ft_pixel_to_float(float v, float mn, float range)
{
    const int p = min(max(__float2int_rn(v), -32768), 32767) + 32768;
    const float q = (float)__dmul_rn((double)p, 1.0 / 65535.0);
    return __fadd_rn(__fmul_rn(q, range), mn);
}

// ---- inverse -----------------------------------------------------------------------------------------------------------------
// grid = (frames, tiles of the level), block = FT_THREADS: the frame index varies fastest in CTA dispatch order, so at any time
// every frame has only a few dozen tiles in flight and the early exit of capped search probes (below) finds the counts of the
// tiles before it.  CAS = row parity of the level's canvas origin.
template <int SRC, bool FINAL, int CAS>
__global__ void __launch_bounds__(FT_THREADS) k_idwt_tile(IdwtFusedArgs a)
{
    __shared__ __align__(16) float s1[FT_ROWS * FT_P1];
    __shared__ __align__(16) float s2[FT_ROWS * FT_P2];
    const Plan *plan = a.plan;
    const int W = plan->W, frame = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const size_t fbase = (size_t)frame * W * plan->H;
    const int ntx = (a.rw + FT_TW - 1) / FT_TW, tyi = (int)blockIdx.y / ntx, txi = (int)blockIdx.y - tyi * ntx;
    const int X0 = txi * FT_TW, Y0 = tyi * FT_TH;
    const bool hx = a.rw > 1, vx = a.rh > 1;                 // a line of one sample is neither transformed nor scaled
    const int ncblks = plan->ncblks;
    if (FINAL && a.cap) {
        // A search probe only has to tell "violations >= cap" from an exact smaller count (the host's decisions are
        // comparisons of the count with the quantile target): once the tiles finished so far have seen cap violations the
        // rest of the frame is skipped.  Most probes of a search fail, and fail in the first tiles.
        __shared__ int s_skip;
        if (tid == 0) s_skip = *(volatile unsigned long long *)&a.es[frame].n_over >= a.cap;
        __syncthreads();
        if (s_skip) return;
    }

    // ---- 1: load (+ dequantise / reconstruct), de-interleaved columns, horizontal scaling folded in ----
    // Per tile row (one warp) two groups of samples: low-pass columns (slots 0..63) and high-pass columns (64..127), each from one
    // band or from the previous level's output.  All loads of a row are issued before the first use (12 independent
    // requests per thread in flight), the arithmetic follows.
    // the Mallat column of each of this thread's four slots does not depend on the row
    int mxs[4];
#pragma unroll
    for (int it = 0; it < 4; ++it) {
        const int high = it >> 1, k = lane + 32 * (it & 1);          // warp-uniform half
        const int X = ft_reflect(X0 - FT_HALO + 2 * k + high, a.rw);
        mxs[it] = ((hx && high) ? a.snw : 0) + (X >> 1);             // (the mirror keeps the parity; a 1-sample line has no high-pass part)
    }
    for (int lr = warp; lr < FT_ROWS; lr += FT_THREADS / 32) {
        const int Y = ft_reflect(Y0 - FT_HALO + lr, a.rh);
        const int ylow = vx ? ((Y & 1) == CAS) : (a.snh > 0);
        const int my = (ylow ? 0 : a.snh) + (Y >> 1);                    // Mallat row
        const size_t rowbase = fbase + (size_t)my * W;
        int qv[4], kbs[4]; unsigned psv[4]; uint2 rjv[4]; float lv[4], stepv[2];
        const bool ll_row = ylow && a.r > 1;                             // this row's low-pass columns come from the previous level
        // band of each half, and (probe) the code-block row of this Mallat row inside it
        int kb0[2], offx[2];
#pragma unroll
        for (int hf = 0; hf < 2; ++hf) {
            const int o = (hx ? hf : 0) | (ylow ? 0 : 2);
            const BandDesc &bd = plan->bands[o == 0 ? 0 : 1 + 3 * (a.r - 1) + (o - 1)];
            stepv[hf] = __fmul_rn(0.5f, bd.stepsize_dec);
            kb0[hf] = 0; offx[hf] = bd.offx;
            if (SRC == IDF_SRC_RECON && !(hf == 0 && ll_row)) {
                int gy = (my - bd.offy) >> 6, kb = bd.first_cblk + gy * bd.cw;
                // tiles below the first one: the block grid is anchored on the canvas, the band's first block row may be a partial one
                while (gy + 1 < bd.ch && my >= a.cblks[kb].ty + a.cblks[kb].h) { ++gy; kb += bd.cw; }
                kb0[hf] = kb;
            }
        }
#pragma unroll
        for (int it = 0; it < 4; ++it) {
            const int xhigh = hx ? (it >> 1) : 0;
            const int mx = mxs[it];
            lv[it] = 0.0f; qv[it] = 0; psv[it] = 0; rjv[it] = make_uint2(0u, 0u); kbs[it] = 0;
            if (!xhigh && ll_row) lv[it] = a.ll_in[rowbase + mx];
            else {
                qv[it] = a.iplane[rowbase + mx];
                if (SRC == IDF_SRC_RECON) {
                    psv[it] = a.spass[rowbase + mx];
                    kbs[it] = (xhigh ? kb0[1] : kb0[0]) + ((mx - (xhigh ? offx[1] : offx[0])) >> 6);
                    rjv[it] = a.rjobs[(size_t)frame * ncblks + kbs[it]];
                }
            }
        }
#pragma unroll
        for (int it = 0; it < 4; ++it) {
            const int high = it >> 1, k = lane + 32 * (it & 1);
            const int xhigh = hx ? high : 0;
            float v;
            if (!xhigh && ll_row) v = lv[it];
            else {
                int q = qv[it];
                if (SRC == IDF_SRC_RECON) {
                    const uint2 rj = rjv[it];
                    const int mode = (int)(rj.x & 0xffu), np = (int)((rj.x >> 8) & 0xffu);
                    if (mode == 4) q = a.cut_pool[(size_t)rj.y * T1_CUT_WORDS + (size_t)(my - a.cblks[kbs[it]].ty) * J2K_CBLK + (mxs[it] - a.cblks[kbs[it]].tx)];
                    else q = np > 0 ? recon_value(q, (int)psv[it], np, (int)((rj.x >> 16) & 0xffu), (int)(rj.x >> 24)) : 0;
                }
                v = __fmul_rn((float)q, xhigh ? stepv[1] : stepv[0]);
            }
            if (hx) v = __fmul_rn(v, xhigh ? IDWT_TWO_INVK : DWT_K);
            s1[lr * FT_P1 + (high ? FT_HOFF : 0) + k] = v;
        }
    }
    __syncthreads();

    // ---- 2: rows ----
    for (int u = tid; u < FT_ROWS * FT_SEGS; u += FT_THREADS) {
        const int row = u / FT_SEGS, j = u - row * FT_SEGS;
        float L[8], H[8];
        const float4 *pl = (const float4 *)(s1 + row * FT_P1 + 4 * j), *ph = (const float4 *)(s1 + row * FT_P1 + FT_HOFF + 4 * j);
        const float4 l0 = pl[0], l1 = pl[1], h0 = ph[0], h1 = ph[1];
        L[0] = l0.x; L[1] = l0.y; L[2] = l0.z; L[3] = l0.w; L[4] = l1.x; L[5] = l1.y; L[6] = l1.z; L[7] = l1.w;
        H[0] = h0.x; H[1] = h0.y; H[2] = h0.z; H[3] = h0.w; H[4] = h1.x; H[5] = h1.y; H[6] = h1.z; H[7] = h1.w;
        if (hx) ft_row_window<true>(L, H);
        float4 *po = (float4 *)(s2 + row * FT_P2 + 8 * j);
        po[0] = make_float4(L[2], H[2], L[3], H[3]);
        po[1] = make_float4(L[4], H[4], L[5], H[5]);
    }
    __syncthreads();

    // ---- 3: columns, straight to global memory ----
    float mn = 0.0f, range = 0.0f, target = 0.0f;
    double fxs = 0.0;
    unsigned over = 0; float mxe = 0.0f; unsigned kmin = 0xffffffffu, kmax = 0; long long sum = 0; int ex = 0;
    if (FINAL) {
        mn = a.stat[frame].minv; range = __fsub_rn(a.stat[frame].maxv, mn);
        target = a.targets ? a.targets[frame] : 0.0f;
        ex = err_fx_exp(range);
        fxs = __longlong_as_double((long long)(1023 + ERR_FX_BITS - ex) << 52);      // 2^(ERR_FX_BITS - ex): the fixed-point unit of the error sum
    }
    if (tid < 2 * FT_TW) {
        const int h = tid >= FT_TW, c = tid - (h ? FT_TW : 0);
        const int X = X0 + c;
        float v[24], og[16];
        if (FINAL && a.orig) {                                           // the original frame's samples travel while the columns are lifted
            const float *op = a.orig + fbase + (size_t)(Y0 + 16 * h) * W + X;
            if (X0 + FT_TW <= a.rw && Y0 + FT_TH <= a.rh) {
#pragma unroll
                for (int i = 0; i < 16; ++i) og[i] = op[(size_t)i * W];
            } else {
#pragma unroll
                for (int i = 0; i < 16; ++i) og[i] = (X < a.rw && Y0 + 16 * h + i < a.rh) ? op[(size_t)i * W] : 0.0f;
            }
        }
#pragma unroll
        for (int i = 0; i < 24; ++i) v[i] = s2[(16 * h + i) * FT_P2 + c];
        if (vx) {
#pragma unroll
            for (int i = 0; i < 24; ++i) v[i] = __fmul_rn(v[i], ((i & 1) == CAS) ? DWT_K : IDWT_TWO_INVK);
            ft_col_window<CAS>(v, -DWT_DELTA, -DWT_GAMMA, -DWT_BETA, -DWT_ALPHA);
        }
        if (X < a.rw) {
#pragma unroll
            for (int i = 4; i < 20; ++i) {
                const int Y = Y0 + 16 * h + i - 4;
                if (Y < a.rh) {
                    const size_t gi = fbase + (size_t)Y * W + X;
                    if (!FINAL) a.out[gi] = v[i];
                    else {
                        const float f = ft_pixel_to_float(v[i], mn, range);
                        if (a.out) a.out[gi] = f;
                        if (a.orig) {
                            const float rr = __fsub_rn(og[i - 4], f), e = fabsf(rr);
                            if (e > target) over++;
                            mxe = fmaxf(mxe, e);
                            const unsigned key = f32_order_key(rr);
                            kmin = min(kmin, key); kmax = max(kmax, key);
                            sum += __double2ll_rn(__dmul_rn((double)rr, fxs));       // = err_fx(rr, ex): scaling by a power of two is exact
                        }
                    }
                }
            }
        }
    }
    if (FINAL && a.orig) {
        for (int o = 16; o > 0; o >>= 1) {
            over += __shfl_xor_sync(0xffffffffu, over, o);
            mxe = fmaxf(mxe, __shfl_xor_sync(0xffffffffu, mxe, o));
            kmin = min(kmin, __shfl_xor_sync(0xffffffffu, kmin, o));
            kmax = max(kmax, __shfl_xor_sync(0xffffffffu, kmax, o));
            sum += __shfl_xor_sync(0xffffffffu, sum, o);
        }
        if (lane == 0) {
            if (over) atomicAdd(&a.es[frame].n_over, (unsigned long long)over);
            atomicMax(&a.es[frame].max_err_bits, __float_as_uint(mxe));
            atomicMin(&a.es[frame].rmin_key, kmin);
            atomicMax(&a.es[frame].rmax_key, kmax);
            atomicAdd((unsigned long long *)&a.es[frame].sum_fx, (unsigned long long)sum);
        }
        if (blockIdx.y == 0 && tid == 0) a.es[frame].sum_exp = ex;
    }
}

// scale_shift (k_dwt.cuh) without the f32 division: q = fl32(x / range) equals fl32((double)x * (1 / (double)range)) for all
// f32 x, range -- a quotient of two 24-bit significands is either a rounding boundary of f32 exactly or at least 2^-49
// (relative) away from one, and reciprocal + product in double are off by less than 2^-51 -- so the conversion to f32 rounds
// the same way (checked on 4e8 random pairs besides).  Three FP64-pipe instructions instead of the division subroutine,
// which was 53 % of the stall samples of the first level.
__device__ __forceinline__ float ft_scale_shift(float d, float mn, double rcp_range)
{
    const float q = (float)__dmul_rn((double)__fsub_rn(d, mn), rcp_range);
    const int u = (int)(__float2uint_rz(__fmul_rn(q, 65535.0f)) & 0xffffu);
    return (float)(u - 32768);
}

// ---- forward -----------------------------------------------------------------------------------------------------------------
struct DwtFusedArgs {
    const float *in;                     // level's input, row stride W: raw frame (first level) or the previous level's LL plane
    float *ll_out;                       // LL of this level, row stride W (level 1: the coefficient plane itself)
    float *coef;                         // Mallat-layout coefficient plane [frame][H][W]
    const FrameStat *stat;
    int W, H, rw, rh, snw, snh;
};

template <bool FIRST, int CAS>
__global__ void __launch_bounds__(FT_THREADS) k_dwt_tile(DwtFusedArgs a)
{
    __shared__ __align__(16) float s1[FT_TH * FT_P1];
    __shared__ __align__(16) float s2[FT_ROWS * FT_P2];
    const int W = a.W, frame = blockIdx.z, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const size_t fbase = (size_t)frame * W * a.H;
    const int X0 = blockIdx.x * FT_TW, Y0 = blockIdx.y * FT_TH;
    const bool hx = a.rw > 1, vx = a.rh > 1;
    float mn = 0.0f; double rcp_range = 0.0;
    if (FIRST) { mn = a.stat[frame].minv; rcp_range = __ddiv_rn(1.0, (double)__fsub_rn(a.stat[frame].maxv, mn)); }

    // ---- 1: load, natural order ----
    for (int lr = warp; lr < FT_ROWS; lr += FT_THREADS / 32) {
        const size_t rowbase = fbase + (size_t)ft_reflect(Y0 - FT_HALO + lr, a.rh) * W;
#pragma unroll
        for (int it = 0; it < 4; ++it) {
            const int lx = lane + 32 * it;
            float v = a.in[rowbase + ft_reflect(X0 - FT_HALO + lx, a.rw)];
            if (FIRST) v = ft_scale_shift(v, mn, rcp_range);
            s2[lr * FT_P2 + lx] = v;
        }
    }
    __syncthreads();

    // ---- 2: columns; the tile's own rows go on, de-interleaved by column, scaled ----
    {
        const int h = tid >> 7, c = tid & 127;
        float v[24];
#pragma unroll
        for (int i = 0; i < 24; ++i) v[i] = s2[(16 * h + i) * FT_P2 + c];
        if (vx) {
            ft_col_window<CAS ^ 1>(v, DWT_ALPHA, DWT_BETA, DWT_GAMMA, DWT_DELTA);
#pragma unroll
            for (int i = 4; i < 20; ++i) v[i] = __fmul_rn(v[i], ((i & 1) == CAS) ? DWT_INVK : DWT_K);
        }
        float *dst = s1 + (16 * h) * FT_P1 + ((c & 1) ? FT_HOFF : 0) + (c >> 1);
#pragma unroll
        for (int i = 4; i < 20; ++i) dst[(i - 4) * FT_P1] = v[i];
    }
    __syncthreads();

    // ---- 3: rows; results staged [low | high] per row in s2 ----
    for (int u = tid; u < FT_TH * FT_SEGS; u += FT_THREADS) {
        const int row = u / FT_SEGS, j = u - row * FT_SEGS;
        float L[8], H[8];
        const float4 *pl = (const float4 *)(s1 + row * FT_P1 + 4 * j), *ph = (const float4 *)(s1 + row * FT_P1 + FT_HOFF + 4 * j);
        const float4 l0 = pl[0], l1 = pl[1], h0 = ph[0], h1 = ph[1];
        L[0] = l0.x; L[1] = l0.y; L[2] = l0.z; L[3] = l0.w; L[4] = l1.x; L[5] = l1.y; L[6] = l1.z; L[7] = l1.w;
        H[0] = h0.x; H[1] = h0.y; H[2] = h0.z; H[3] = h0.w; H[4] = h1.x; H[5] = h1.y; H[6] = h1.z; H[7] = h1.w;
        if (hx) {
            ft_row_window<false>(L, H);
#pragma unroll
            for (int i = 2; i <= 5; ++i) { L[i] = __fmul_rn(L[i], DWT_INVK); H[i] = __fmul_rn(H[i], DWT_K); }
        }
        *(float4 *)(s2 + row * FT_P2 + 4 * j) = make_float4(L[2], L[3], L[4], L[5]);
        *(float4 *)(s2 + row * FT_P2 + FT_HALF + 4 * j) = make_float4(H[2], H[3], H[4], H[5]);
    }
    __syncthreads();

    // ---- 4: store: (X, Y) -> Mallat column (X odd ? snw : 0) + X / 2, row (Y high-pass ? snh : 0) + Y / 2 ----
    const int tw = min(FT_TW, a.rw - X0), th = min(FT_TH, a.rh - Y0);
    const int nlo = hx ? (tw + 1) >> 1 : tw, nhi = hx ? tw >> 1 : 0;       // X0 is even
    for (int yy = warp; yy < th; yy += FT_THREADS / 32) {
        const int Y = Y0 + yy;
        const int ylow = vx ? ((Y & 1) == CAS) : (a.snh > 0);
        const size_t rowbase = fbase + (size_t)((ylow ? 0 : a.snh) + (Y >> 1)) * W;
#pragma unroll
        for (int it = 0; it < 4; ++it) {
            const int high = it >> 1, k = lane + 32 * (it & 1);
            if (k >= (high ? nhi : nlo)) continue;
            float *dst = (!high && ylow) ? a.ll_out : a.coef;
            dst[rowbase + (high ? a.snw : 0) + (X0 >> 1) + k] = s2[yy * FT_P2 + (high ? FT_HALF : 0) + k];
        }
    }
}
