// k_dwt.cuh -- min/max reduction, u16 scaling + DC shift, forward 9/7 DWT.
//
// Bit-exact with OpenJPEG's opj_dwt_encode_real (dwt.c:1930-2108, lifting :1222-1300): every
// lifting update is (a + b) * c followed by +=, never contracted into an FMA (explicit
// __fadd_rn/__fmul_rn), vertical pass before horizontal pass at each level.
//
// HBM-bound stages.  Algorithmic bytes per frame (N = H*W samples):
//   k_minmax       : read 4N
//   k_dwt_v<FIRST> : read 4N (raw f32) + write 4N;   level l>1 passes read+write 4N/4^(l-1) each
//   k_dwt_h        : read + write 4N/4^(l-1)
#pragma once
#include "common.cuh"

// ---- ordered keys: monotone map float -> u32 with -0.0 == +0.0 --------------------------------
__device__ __forceinline__ unsigned f32_order_key(float f)
{
    unsigned b = __float_as_uint(f);
    if ((b << 1) == 0) b = 0;                         // canonical zero
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}

// findMinMaxf semantics: the FIRST element attaining the extremum wins (matters only for +-0).
// A thread walks its elements in ascending index order and replaces its candidate only on a strict improvement of the
// 32-bit order key, so it keeps the first occurrence; threads are merged with 64-bit keys
// key_min = (order << 32) | idx  -> atomicMin ; key_max = (order << 32) | ~idx -> atomicMax.
// 16-byte loads when the frame is 16-byte aligned (HBM-bound: 4N bytes read, nothing written).
struct MinMaxAcc { unsigned kmin, kmax, imin, imax; int bad; };
__device__ __forceinline__ void minmax_take(MinMaxAcc &a, float v, unsigned i)
{
    const unsigned b = __float_as_uint(v);
    if ((b & 0x7f800000u) == 0x7f800000u) a.bad = 1;
    const unsigned k = f32_order_key(v);
    if (k < a.kmin) { a.kmin = k; a.imin = i; }
    if (k > a.kmax) { a.kmax = k; a.imax = i; }
}

__global__ void __launch_bounds__(256) k_minmax(const float *__restrict__ in, FrameStat *stat, int n)
{
    const int frame = blockIdx.y;
    const float *p = in + (size_t)frame * n;
    // "no element yet": the keys below are neutral for atomicMin / atomicMax once widened
    MinMaxAcc a; a.kmin = 0xffffffffu; a.imin = 0xffffffffu; a.kmax = 0u; a.imax = 0xffffffffu; a.bad = 0;
    const int gtid = blockIdx.x * blockDim.x + threadIdx.x, stride = gridDim.x * blockDim.x;
    const int n4 = (((size_t)p & 15) == 0) ? n >> 2 : 0;
    const float4 *p4 = (const float4 *)p;
    int j = gtid;
    for (; j + stride < n4; j += 2 * stride) {          // two independent 16-byte loads in flight per thread
        const float4 u = p4[j], v = p4[j + stride];
        const unsigned i0 = 4u * (unsigned)j, i1 = 4u * (unsigned)(j + stride);
        minmax_take(a, u.x, i0); minmax_take(a, u.y, i0 + 1); minmax_take(a, u.z, i0 + 2); minmax_take(a, u.w, i0 + 3);
        minmax_take(a, v.x, i1); minmax_take(a, v.y, i1 + 1); minmax_take(a, v.z, i1 + 2); minmax_take(a, v.w, i1 + 3);
    }
    if (j < n4) {
        const float4 u = p4[j];
        const unsigned i0 = 4u * (unsigned)j;
        minmax_take(a, u.x, i0); minmax_take(a, u.y, i0 + 1); minmax_take(a, u.z, i0 + 2); minmax_take(a, u.w, i0 + 3);
    }
    for (int i = 4 * n4 + gtid; i < n; i += stride) minmax_take(a, p[i], (unsigned)i);   // tail / unaligned frame
    unsigned long long kmin = ((unsigned long long)a.kmin << 32) | a.imin;
    unsigned long long kmax = ((unsigned long long)a.kmax << 32) | (unsigned)(~a.imax);
    int bad = a.bad;
    for (int o = 16; o > 0; o >>= 1) {
        unsigned long long x = __shfl_xor_sync(0xffffffffu, kmin, o);
        unsigned long long y = __shfl_xor_sync(0xffffffffu, kmax, o);
        int c = __shfl_xor_sync(0xffffffffu, bad, o);
        kmin = x < kmin ? x : kmin; kmax = y > kmax ? y : kmax; bad |= c;
    }
    if ((threadIdx.x & 31) == 0) {
        atomicMin(&stat[frame].minkey, kmin);
        atomicMax(&stat[frame].maxkey, kmax);
        if (bad) atomicOr(&stat[frame].flags, STAT_NANINF);
    }
}

__global__ void k_stat_init(FrameStat *stat, int nframes)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < nframes) { stat[i].minkey = ~0ull; stat[i].maxkey = 0ull; stat[i].flags = 0; stat[i].minv = 0; stat[i].maxv = 0; stat[i].pad = 0; }
}

__global__ void k_minmax_final(const float *__restrict__ in, FrameStat *stat, int n, int nframes)
{
    int f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= nframes) return;
    unsigned imin = (unsigned)(stat[f].minkey & 0xffffffffu);
    unsigned imax = ~(unsigned)(stat[f].maxkey & 0xffffffffu);
    float mn = in[(size_t)f * n + imin], mx = in[(size_t)f * n + imax];
    stat[f].minv = mn; stat[f].maxv = mx;
    if (mn == mx) stat[f].flags |= STAT_CONST;
}

// ---- 9/7 lifting on interleaved lines held in shared memory ----------------------------------
#define DWT_ALPHA (-1.586134342f)
#define DWT_BETA (-0.052980118f)
#define DWT_GAMMA (0.882911075f)
#define DWT_DELTA (0.443506852f)
#define DWT_K (1.230174105f)
#define DWT_INVK ((float)(1.0 / 1.230174105))

// The lifting helpers are written for an interleaved line with ne even-indexed and no odd-indexed samples
// (parameters sn = ne, dn = no).  For a line whose origin has even parity (cas 0) the even samples are the low-pass
// ones; for odd parity (cas 1, rows of a tile that starts at an odd canvas row of this resolution) they are the
// high-pass ones and the four steps swap roles (opj_dwt_encode_1_real with a = 1, b = 0, dwt.c:1251-1300).
// d[i] += (s[i] + s[i+1]) * c ; the last d of an even-length line mirrors: (2 * s[i]) * c
__device__ __forceinline__ void lift_d(float *x, int es, int sn, int dn, float c, int i0, int istr)
{
    const int m = min(dn, sn - 1);
    for (int i = i0; i < dn; i += istr) {
        float a = x[(2 * i) * es];
        float v = (i < m) ? __fmul_rn(__fadd_rn(a, x[(2 * i + 2) * es]), c) : __fmul_rn(__fmul_rn(2.0f, a), c);
        x[(2 * i + 1) * es] = __fadd_rn(x[(2 * i + 1) * es], v);
    }
}

// s[i] += (d[i-1] + d[i]) * c ; s[0] uses d[0] twice ; the last s of an odd-length line: (2 * d[i-1]) * c
__device__ __forceinline__ void lift_s(float *x, int es, int sn, int dn, float c, int i0, int istr)
{
    const int m = min(sn, dn);
    for (int i = i0; i < sn; i += istr) {
        float v;
        if (i < m) {
            float l = x[(i == 0 ? 1 : 2 * i - 1) * es];
            v = __fmul_rn(__fadd_rn(l, x[(2 * i + 1) * es]), c);
        } else {
            v = __fmul_rn(__fmul_rn(2.0f, x[(2 * i - 1) * es]), c);
        }
        x[(2 * i) * es] = __fadd_rn(x[(2 * i) * es], v);
    }
}

__device__ __forceinline__ void lift_scale(float *x, int es, int n, int i0, int istr, int cas = 0)
{
    for (int i = i0; i < n; i += istr)
        x[i * es] = __fmul_rn(x[i * es], ((i ^ cas) & 1) ? DWT_K : DWT_INVK);
}

// the raw sample as OpenJPEG sees it: u16 scaling (src/ebcc_codec.c:686-689) then DC shift (tcd.c:2469-2508)
__device__ __forceinline__ float scale_shift(float d, float mn, float range)
{
    float v = __fmul_rn(__fdiv_rn(__fsub_rn(d, mn), range), 65535.0f);
    int u = (int)(__float2uint_rz(v) & 0xffffu);
    return (float)(u - 32768);
}

// Vertical pass over a strip of CW columns x rh rows.  block = (CW, TY); grid = (ceil(rw/CW), B).
// FIRST: source is the raw frame (scaled on the fly), else the coefficient buffer (in place).
template <bool FIRST>
__global__ void k_dwt_v(const float *__restrict__ in, float *coef, const FrameStat *__restrict__ stat,
                        int W, int H, int rw, int rh, int sn, int cas)
{
    EBCC_DYN_SMEM(float, x);
    const int CW = blockDim.x, c = threadIdx.x, ty = threadIdx.y, TY = blockDim.y;
    const int frame = blockIdx.y, col = blockIdx.x * CW + c;
    const bool active = col < rw;
    const size_t base = (size_t)frame * H * W;
    if (active) {
        if (FIRST) {
            const float mn = stat[frame].minv, range = __fsub_rn(stat[frame].maxv, mn);
            for (int i = ty; i < rh; i += TY) x[i * CW + c] = scale_shift(in[base + (size_t)i * W + col], mn, range);
        } else {
            for (int i = ty; i < rh; i += TY) x[i * CW + c] = coef[base + (size_t)i * W + col];
        }
    }
    __syncthreads();
    if (rh > 1) {
        const int dn = rh - sn;
        float *xc = x + c;
        if (cas == 0) {
            if (active) lift_d(xc, CW, sn, dn, DWT_ALPHA, ty, TY);
            __syncthreads();
            if (active) lift_s(xc, CW, sn, dn, DWT_BETA, ty, TY);
            __syncthreads();
            if (active) lift_d(xc, CW, sn, dn, DWT_GAMMA, ty, TY);
            __syncthreads();
            if (active) lift_s(xc, CW, sn, dn, DWT_DELTA, ty, TY);
        } else {
            if (active) lift_s(xc, CW, dn, sn, DWT_ALPHA, ty, TY);
            __syncthreads();
            if (active) lift_d(xc, CW, dn, sn, DWT_BETA, ty, TY);
            __syncthreads();
            if (active) lift_s(xc, CW, dn, sn, DWT_GAMMA, ty, TY);
            __syncthreads();
            if (active) lift_d(xc, CW, dn, sn, DWT_DELTA, ty, TY);
        }
        __syncthreads();
        if (active) lift_scale(xc, CW, rh, ty, TY, cas);
        __syncthreads();
    }
    if (active) {
        if (rh > 1) {
            for (int i = ty; i < rh; i += TY) {
                int dst = ((i ^ cas) & 1) ? sn + (i >> 1) : (i >> 1);
                coef[base + (size_t)dst * W + col] = x[i * CW + c];
            }
        } else if (FIRST) {
            coef[base + col] = x[c];
        }
    }
}

// Horizontal pass over RH rows x rw columns.  block = (TX, RH); grid = (ceil(rh/RH), B).
__global__ void k_dwt_h(float *coef, int W, int H, int rw, int rh, int sn)
{
    EBCC_DYN_SMEM(float, x);
    const int tx = threadIdx.x, TX = blockDim.x, r = threadIdx.y;
    const int frame = blockIdx.y, row = blockIdx.x * blockDim.y + r;
    const bool active = row < rh;
    float *line = x + (size_t)r * rw;
    float *g = coef + (size_t)frame * H * W + (size_t)row * W;
    if (active) for (int i = tx; i < rw; i += TX) line[i] = g[i];
    __syncthreads();
    const int dn = rw - sn;
    if (active) lift_d(line, 1, sn, dn, DWT_ALPHA, tx, TX);
    __syncthreads();
    if (active) lift_s(line, 1, sn, dn, DWT_BETA, tx, TX);
    __syncthreads();
    if (active) lift_d(line, 1, sn, dn, DWT_GAMMA, tx, TX);
    __syncthreads();
    if (active) lift_s(line, 1, sn, dn, DWT_DELTA, tx, TX);
    __syncthreads();
    if (active) lift_scale(line, 1, rw, tx, TX);
    __syncthreads();
    if (active)
        for (int j = tx; j < rw; j += TX) g[j] = (j < sn) ? line[2 * j] : line[2 * (j - sn) + 1];
}
