// k_idwt.cuh -- decoder back half: dequantisation, inverse 9/7 DWT, DC shift / clamp / float rescale,
// and the error statistics of the compression-ratio search.
//
// Bit-exact with OpenJPEG's opj_dwt_decode_tile_97 (dwt.c:3512-3733; 1-D kernel opj_v8dwt_decode
// :3353-3428 with its historic two_invK constant) and opj_tcd_dc_level_shift_decode (tcd.c:2265-2357):
// per level horizontal pass over every row first, then vertical pass; each lifting update is
// x += (l + r) * c with separate IEEE add / mul / add (no FMA), edge samples use (c + c) * neighbour.
//
// HBM-bound.  Algorithmic bytes per frame (N samples): k_dequant 4N + 4N; each level l reads and writes
// 4N/4^(5-l) twice (h, v); k_finalize 4N + 4N (+ 4N when it also compares against the original).
#pragma once
#include "common.cuh"
#include "k_dwt.cuh"

#define IDWT_TWO_INVK 1.625732422f

// integer tier-1 output -> float coefficient: (float)v * (0.5f * stepsize_dec)   (t1.c:1853-1865)
__global__ void k_dequant(const int *__restrict__ iplane, float *coef, const Plan *__restrict__ plan, int nframes)
{
    const int W = plan->W, H = plan->H;
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
        const float step = __fmul_rn(0.5f, plan->bands[band].stepsize_dec);
        coef[(size_t)frame * n + i] = __fmul_rn((float)iplane[(size_t)frame * n + i], step);
    }
    (void)nframes;
}

// x holds one interleaved line (element stride es): evens = low-pass, odds = high-pass
__device__ __forceinline__ void ilift_scale(float *x, int es, int n, int i0, int istr, int cas = 0)
{
    for (int i = i0; i < n; i += istr) x[i * es] = __fmul_rn(x[i * es], ((i ^ cas) & 1) ? IDWT_TWO_INVK : DWT_K);
}
// even[i] += (odd[i-1] + odd[i]) * c ; even[0] uses odd[0] twice ; trailing even of an odd-length line: (c+c) * odd[i-1]
__device__ __forceinline__ void ilift_s(float *x, int es, int sn, int dn, float c, int i0, int istr)
{
    const int m = min(sn, dn);
    for (int i = i0; i < sn; i += istr) {
        float v;
        if (i < m) v = __fmul_rn(__fadd_rn(x[(i == 0 ? 1 : 2 * i - 1) * es], x[(2 * i + 1) * es]), c);
        else v = __fmul_rn(__fadd_rn(c, c), x[(2 * i - 1) * es]);
        x[(2 * i) * es] = __fadd_rn(x[(2 * i) * es], v);
    }
}
// odd[i] += (even[i] + even[i+1]) * c ; trailing odd of an even-length line: (c+c) * even[i]
__device__ __forceinline__ void ilift_d(float *x, int es, int sn, int dn, float c, int i0, int istr)
{
    const int m = min(dn, sn - 1);
    for (int i = i0; i < dn; i += istr) {
        float v;
        if (i < m) v = __fmul_rn(__fadd_rn(x[(2 * i) * es], x[(2 * i + 2) * es]), c);
        else v = __fmul_rn(__fadd_rn(c, c), x[(2 * i) * es]);
        x[(2 * i + 1) * es] = __fadd_rn(x[(2 * i + 1) * es], v);
    }
}

// horizontal synthesis of RH rows; block = (TX, RH); grid = (ceil(rh/RH), B)
__global__ void k_idwt_h(float *coef, int W, int H, int rw, int rh, int sn)
{
    EBCC_DYN_SMEM(float, x);
    const int tx = threadIdx.x, TX = blockDim.x, r = threadIdx.y;
    const int frame = blockIdx.y, row = blockIdx.x * blockDim.y + r;
    const bool active = row < rh;
    float *line = x + (size_t)r * rw;
    float *g = coef + (size_t)frame * H * W + (size_t)row * W;
    const int dn = rw - sn;
    if (active) for (int j = tx; j < rw; j += TX) line[(j < sn) ? 2 * j : 2 * (j - sn) + 1] = g[j];
    __syncthreads();
    if (dn > 0 || sn > 1) {
        if (active) ilift_scale(line, 1, rw, tx, TX);
        __syncthreads();
        if (active) ilift_s(line, 1, sn, dn, -DWT_DELTA, tx, TX);
        __syncthreads();
        if (active) ilift_d(line, 1, sn, dn, -DWT_GAMMA, tx, TX);
        __syncthreads();
        if (active) ilift_s(line, 1, sn, dn, -DWT_BETA, tx, TX);
        __syncthreads();
        if (active) ilift_d(line, 1, sn, dn, -DWT_ALPHA, tx, TX);
        __syncthreads();
    }
    if (active) for (int j = tx; j < rw; j += TX) g[j] = line[j];
}

// vertical synthesis of a strip of CW columns; block = (CW, TY); grid = (ceil(rw/CW), B).  cas = row parity of the
// resolution's canvas origin: with cas 1 the low-pass lines are the odd ones and the steps swap roles
// (opj_v8dwt_decode with a = 1, b = 0, dwt.c:3353-3428; interleave dwt.c:3575-3600).
__global__ void k_idwt_v(float *coef, int W, int H, int rw, int rh, int sn, int cas)
{
    EBCC_DYN_SMEM(float, x);
    const int CW = blockDim.x, c = threadIdx.x, ty = threadIdx.y, TY = blockDim.y;
    const int frame = blockIdx.y, col = blockIdx.x * CW + c;
    const bool active = col < rw;
    const size_t base = (size_t)frame * H * W;
    const int dn = rh - sn;
    if (active)
        for (int i = ty; i < rh; i += TY) x[((i < sn) ? 2 * i + cas : 2 * (i - sn) + 1 - cas) * CW + c] = coef[base + (size_t)i * W + col];
    __syncthreads();
    float *xc = x + c;
    if (cas == 0 ? (dn > 0 || sn > 1) : (sn > 0 || dn > 1)) {
        if (active) ilift_scale(xc, CW, rh, ty, TY, cas);
        __syncthreads();
        if (cas == 0) {
            if (active) ilift_s(xc, CW, sn, dn, -DWT_DELTA, ty, TY);
            __syncthreads();
            if (active) ilift_d(xc, CW, sn, dn, -DWT_GAMMA, ty, TY);
            __syncthreads();
            if (active) ilift_s(xc, CW, sn, dn, -DWT_BETA, ty, TY);
            __syncthreads();
            if (active) ilift_d(xc, CW, sn, dn, -DWT_ALPHA, ty, TY);
        } else {
            if (active) ilift_d(xc, CW, dn, sn, -DWT_DELTA, ty, TY);
            __syncthreads();
            if (active) ilift_s(xc, CW, dn, sn, -DWT_GAMMA, ty, TY);
            __syncthreads();
            if (active) ilift_d(xc, CW, dn, sn, -DWT_BETA, ty, TY);
            __syncthreads();
            if (active) ilift_s(xc, CW, dn, sn, -DWT_ALPHA, ty, TY);
        }
        __syncthreads();
    }
    if (active) for (int i = ty; i < rh; i += TY) coef[base + (size_t)i * W + col] = x[i * CW + c];
}

// reconstructed sample -> float: lrintf + 32768, clamp to [0, 65535] (tcd.c:2320-2340), then
// ((float)p / 65535) * (max - min) + min   (src/ebcc_codec.c:1129-1131)
__device__ __forceinline__ float j2k_pixel_to_float(float v, float mn, float range)
{
    int p;
    if (v > 2147483647.0f) p = 65535;
    else if (v < -2147483648.0f) p = 0;
    else {
        long long t = (long long)__float2int_rn(v) + 32768;
        p = t < 0 ? 0 : (t > 65535 ? 65535 : (int)t);
    }
    return __fadd_rn(__fmul_rn(__fdiv_rn((float)p, 65535.0f), range), mn);
}

// (struct ErrStat: common.cuh)
// The mean error feeds the header's min / max (src/ebcc_codec.c:863-868), so the sum must not depend on the order in
// which CTAs and warps finish: every term is rounded to a fixed-point integer (units of 2^(e - 39), |term| < 2^e,
// at most 2^22 terms) and the integers are added with 64-bit atomics -- exact, hence order-free; the rounding is
// 2^-39 of the value range per term, far below one ulp of the f32 it ends up in.
#define ERR_FX_BITS 39
__device__ __forceinline__ long long err_fx(float r, int e) { return __double2ll_rn(ldexp((double)r, ERR_FX_BITS - e)); }
__host__ __device__ __forceinline__ int err_fx_exp(float bound) { int e = 0; (void)frexpf(bound, &e); return e; }

__global__ void k_errstat_init(ErrStat *es, int n)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { es[i].n_over = 0; es[i].max_err_bits = 0; es[i].rmin_key = 0xffffffffu; es[i].rmax_key = 0; es[i].sum_exp = 0; es[i].sum_fx = 0; }
}

__device__ __forceinline__ float key_to_f32(unsigned k)
{
    return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

// out = decoded frame; if orig != nullptr also accumulates ErrStat against it with the frame's target
__global__ void k_finalize(const float *__restrict__ coef, float *out, const FrameStat *__restrict__ stat,
                           const float *__restrict__ orig, const float *__restrict__ targets, ErrStat *es, int n)
{
    const int frame = blockIdx.y;
    const float mn = stat[frame].minv, range = __fsub_rn(stat[frame].maxv, mn);
    const float target = targets ? targets[frame] : 0.0f;
    unsigned long long over = 0;
    float mx = 0.0f;
    unsigned kmin = 0xffffffffu, kmax = 0;
    long long sum = 0;
    const int ex = err_fx_exp(range);         // |orig - decoded| <= max - min < 2^ex
    const size_t base = (size_t)frame * n;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float v = j2k_pixel_to_float(coef[base + i], mn, range);
        out[base + i] = v;
        if (orig) {
            const float r = __fsub_rn(orig[base + i], v), e = fabsf(r);
            if (e > target) over++;
            mx = fmaxf(mx, e);
            const unsigned k = f32_order_key(r);
            kmin = min(kmin, k); kmax = max(kmax, k);
            sum += err_fx(r, ex);
        }
    }
    if (!orig) return;
    for (int o = 16; o > 0; o >>= 1) {
        over += __shfl_xor_sync(0xffffffffu, over, o);
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        kmin = min(kmin, __shfl_xor_sync(0xffffffffu, kmin, o));
        kmax = max(kmax, __shfl_xor_sync(0xffffffffu, kmax, o));
        sum += __shfl_xor_sync(0xffffffffu, sum, o);
    }
    if ((threadIdx.x & 31) == 0) {
        if (over) atomicAdd(&es[frame].n_over, over);
        atomicMax(&es[frame].max_err_bits, __float_as_uint(mx));
        atomicMin(&es[frame].rmin_key, kmin);
        atomicMax(&es[frame].rmax_key, kmax);
        atomicAdd((unsigned long long *)&es[frame].sum_fx, (unsigned long long)sum);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) es[frame].sum_exp = ex;
}
