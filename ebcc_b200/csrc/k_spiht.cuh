// k_spiht.cuh -- residual layer: SPIHT wavelet coder of the reference (src/spiht/spiht_re.c, dwt.h, bitio.h)
// re-designed for the GPU.
//
// Forward path (per frame, all HBM-bound elementwise / separable kernels except the coder):
//   k_sp_load      residual = d - decoded, normalise to [0,1], x255, mirror-pad to a multiple of 16
//                  (src/ebcc_codec.c:730-746, dwt.h:41-78) + double sum for the DC term
//   k_sp_subdc     dc = floor(mean) (dwt.h:319-334)
//   k_sp_dwt_row / k_sp_dwt_col   3-level 9/7 with the reference's own boundary rules and xi scaling
//                  (dwt.h:87-112, 142-167), same float operation order, no FMA
//   k_sp_normalize truncation toward zero -> int32 coefficients + max magnitude (dwt.h:355-368)
//   k_sp_desc      per node: max magnitude over all descendants (D) and over grand-descendants (L),
//                  i.e. the significance tests of is_significant_set_A/B (spiht_re.c:160-206) in O(1)
//   k_sp_encode    one CTA per frame.  The list-driven coder (spiht_re.c:208-317) is data-parallel once the
//                  emission order is fixed: every LIP / LIS / LSP entry's bits depend only on coefficient
//                  magnitudes, and its position in the stream is an exclusive prefix sum over the entries
//                  before it.  Entries appended to the LIS while it is scanned are processed as successive
//                  wavefronts (the reference visits them in the same pass, in append order).  Besides the
//                  bitstream the kernel records, per coefficient, the stream position of its sign bit and its
//                  LSP index, and per bit-plane the base position of the refinement pass.
//
// Truncation search (src/ebcc_codec.c:766-796): decoding a prefix of P payload bits equals "apply every sign /
// refinement event whose position is <= P", so a probe is k_sp_recon (elementwise) + inverse DWT + error
// statistics; no serial decode is needed on the encode side.
//
// Decode path (ebcc_decode): k_sp_decode is the serial list decoder (one thread per frame; stored residual
// streams are a few KB), followed by the same inverse DWT.
#pragma once
#include "common.cuh"
#include "k_idwt.cuh"

#define SP_STAGES 3
#define SP_ALPHA (-1.586134342f)
#define SP_BETA (-0.05298011854f)
#define SP_GAMMA (0.8829110762f)
#define SP_DELTA (0.44355068522f)
#define SP_XI (1.149604398f)
#define SP_THREADS 1024
#define SP_NOPOS 0xffffffffu

struct SpGeom { int W, H, Wp, Hp, lx, ly; };

struct SpInfo {                 // per frame
    long long sum;              // sum of the padded image in units of 2^-30 (for the DC term): integer, hence order-free
    float rmin, rmax;           // residual range
    int dc;                     // floor(mean), 0..255
    int maxabs;                 // max |coefficient|
    int step0;                  // top bit-plane
    unsigned budget;            // bit budget of the coder (trunc_bits)
    unsigned emitted;           // payload bits emitted (<= budget + 1)
    unsigned nbytes;            // bitstream bytes
    unsigned ref_base[32];      // payload position (0-based) of the first refinement bit of each plane
    unsigned lsp_before[32];    // LSP length when the plane's refinement pass starts (entries of earlier planes)
    int active;
};

// ---- forward transform ----------------------------------------------------------------------------------
__global__ void k_sp_init(SpInfo *info, const float *rmm, const unsigned *budget, const int *active, int n)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    SpInfo &s = info[i];
    s.sum = 0; s.rmin = rmm[2 * i]; s.rmax = rmm[2 * i + 1]; s.dc = 0; s.maxabs = 0; s.step0 = 0;
    s.budget = budget[i]; s.emitted = 0; s.nbytes = 0; s.active = active[i];
    for (int p = 0; p < 32; ++p) { s.ref_base[p] = SP_NOPOS; s.lsp_before[p] = 0; }
}

__global__ void k_sp_load(const float *__restrict__ orig, const float *__restrict__ dec, float *img, SpInfo *info, SpGeom g)
{
    const int frame = blockIdx.y;
    if (!info[frame].active) return;
    const float rmin = info[frame].rmin, range = __fsub_rn(info[frame].rmax, rmin);
    const size_t n = (size_t)g.Wp * g.Hp, nsrc = (size_t)g.W * g.H;
    long long sum = 0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const int y = (int)(i / g.Wp), x = (int)(i - (size_t)y * g.Wp);
        float v = 0.0f;
        if (!(y >= g.H && x >= g.W)) {
            const int sy = y < g.H ? y : 2 * g.H - y - 1, sx = x < g.W ? x : 2 * g.W - x - 1;
            const size_t si = (size_t)frame * nsrc + (size_t)sy * g.W + sx;
            const float r = __fsub_rn(orig[si], dec[si]);
            v = __fmul_rn(__fdiv_rn(__fsub_rn(r, rmin), range), 255.0f);
        }
        img[(size_t)frame * n + i] = v;
        sum += __double2ll_rn((double)v * 1073741824.0);        // v in [0, 255]: < 2^38 per term, < 2^61 per frame
    }
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if ((threadIdx.x & 31) == 0) atomicAdd((unsigned long long *)&info[frame].sum, (unsigned long long)sum);
}

__global__ void k_sp_subdc(float *img, SpInfo *info, SpGeom g)
{
    const int frame = blockIdx.y;
    if (!info[frame].active) return;
    const size_t n = (size_t)g.Wp * g.Hp;
    // floor(mean) (dwt.h:319-334); the terms are non-negative, integer division is the floor
    const long long dc = info[frame].sum / ((long long)n << 30);
    const float dcf = (float)dc;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        img[(size_t)frame * n + i] = __fsub_rn(img[(size_t)frame * n + i], dcf);
    if (blockIdx.x == 0 && threadIdx.x == 0) info[frame].dc = (int)dc;
}

// one line of n items held interleaved in x (stride es); result de-interleaved into lo[] / hi[] (stride os)
__device__ __forceinline__ void sp_fwd_line(const float *x, int es, float *lo, float *hi, int os, int n, int t0, int ts, float *h, float *l)
{
    // h[], l[]: scratch of n/2 floats each (stride 1 per thread group is handled by the caller via pointers)
    const int m = n >> 1;
    for (int i = t0; i < m; i += ts) {
        float v;
        if (i < m - 1) v = __fadd_rn(x[(2 * i + 1) * es], __fmul_rn(SP_ALPHA, __fadd_rn(x[(2 * i) * es], x[(2 * i + 2) * es])));
        else v = __fadd_rn(x[(n - 1) * es], __fmul_rn(__fmul_rn(2.0f, SP_ALPHA), x[(n - 2) * es]));
        h[i] = v;
    }
    __syncthreads();
    for (int i = t0; i < m; i += ts) {
        const float s = i == 0 ? __fadd_rn(h[0], h[1]) : __fadd_rn(h[i], h[i - 1]);
        l[i] = __fadd_rn(x[(2 * i) * es], __fmul_rn(SP_BETA, s));
    }
    __syncthreads();
    for (int i = t0; i < m; i += ts) {
        const float s = i < m - 1 ? __fadd_rn(l[i], l[i + 1]) : __fadd_rn(l[m - 1], l[m - 2]);
        h[i] = __fadd_rn(h[i], __fmul_rn(SP_GAMMA, s));
    }
    __syncthreads();
    for (int i = t0; i < m; i += ts) {
        const float s = i == 0 ? __fadd_rn(h[0], h[1]) : __fadd_rn(h[i], h[i - 1]);
        l[i] = __fadd_rn(l[i], __fmul_rn(SP_DELTA, s));
    }
    __syncthreads();
    for (int i = t0; i < m; i += ts) {
        lo[i * os] = __fmul_rn(l[i], SP_XI);
        hi[i * os] = __fdiv_rn(h[i], SP_XI);
    }
}

// rows: src -> dst (different buffers), one row per CTA; block = 256; grid = (rows, B)
__global__ void k_sp_dwt_row(const float *__restrict__ src, float *dst, const SpInfo *__restrict__ info, SpGeom g, int nx)
{
    EBCC_DYN_SMEM(float, sm);
    const int frame = blockIdx.y, row = blockIdx.x;
    if (!info[frame].active) return;
    const size_t off = (size_t)frame * g.Wp * g.Hp + (size_t)row * g.Wp;
    float *x = sm, *h = sm + nx, *l = h + nx / 2;
    for (int i = threadIdx.x; i < nx; i += blockDim.x) x[i] = src[off + i];
    __syncthreads();
    sp_fwd_line(x, 1, dst + off, dst + off + nx / 2, 1, nx, threadIdx.x, blockDim.x, h, l);
}

// columns: src -> dst, CW columns per CTA; block = (CW, TY); grid = (ceil(nx/CW), B)
__global__ void k_sp_dwt_col(const float *__restrict__ src, float *dst, const SpInfo *__restrict__ info, SpGeom g, int nx, int ny)
{
    EBCC_DYN_SMEM(float, sm);
    const int frame = blockIdx.y;
    if (!info[frame].active) return;
    const int CW = blockDim.x, c = threadIdx.x, col = blockIdx.x * CW + c;
    const bool act = col < nx;
    const size_t base = (size_t)frame * g.Wp * g.Hp;
    float *x = sm, *h = sm + (size_t)ny * CW, *l = h + (size_t)(ny / 2) * CW;
    if (act) for (int i = threadIdx.y; i < ny; i += blockDim.y) x[i * CW + c] = src[base + (size_t)i * g.Wp + col];
    __syncthreads();
    // same lifting as sp_fwd_line, column-strided; inactive columns still take part in the barriers
    const int m = ny >> 1, t0 = threadIdx.y, ts = blockDim.y;
    float *xc = x + c, *hc = h + c, *lc = l + c;
    if (act) for (int i = t0; i < m; i += ts) {
        float v;
        if (i < m - 1) v = __fadd_rn(xc[(2 * i + 1) * CW], __fmul_rn(SP_ALPHA, __fadd_rn(xc[(2 * i) * CW], xc[(2 * i + 2) * CW])));
        else v = __fadd_rn(xc[(ny - 1) * CW], __fmul_rn(__fmul_rn(2.0f, SP_ALPHA), xc[(ny - 2) * CW]));
        hc[i * CW] = v;
    }
    __syncthreads();
    if (act) for (int i = t0; i < m; i += ts) {
        const float s = i == 0 ? __fadd_rn(hc[0], hc[CW]) : __fadd_rn(hc[i * CW], hc[(i - 1) * CW]);
        lc[i * CW] = __fadd_rn(xc[(2 * i) * CW], __fmul_rn(SP_BETA, s));
    }
    __syncthreads();
    if (act) for (int i = t0; i < m; i += ts) {
        const float s = i < m - 1 ? __fadd_rn(lc[i * CW], lc[(i + 1) * CW]) : __fadd_rn(lc[(m - 1) * CW], lc[(m - 2) * CW]);
        hc[i * CW] = __fadd_rn(hc[i * CW], __fmul_rn(SP_GAMMA, s));
    }
    __syncthreads();
    if (act) for (int i = t0; i < m; i += ts) {
        const float s = i == 0 ? __fadd_rn(hc[0], hc[CW]) : __fadd_rn(hc[i * CW], hc[(i - 1) * CW]);
        lc[i * CW] = __fadd_rn(lc[i * CW], __fmul_rn(SP_DELTA, s));
    }
    __syncthreads();
    if (act) for (int i = t0; i < m; i += ts) {
        dst[base + (size_t)i * g.Wp + col] = __fmul_rn(lc[i * CW], SP_XI);
        dst[base + (size_t)(m + i) * g.Wp + col] = __fdiv_rn(hc[i * CW], SP_XI);
    }
}

// the parts of a level's source plane that the row pass does not rewrite must still be visible in the other
// buffer: rows write temp[0..ny)[0..nx), columns write data[0..ny)[0..nx) -- exactly the same region, so the
// ping-pong needs no extra copies.

__global__ void k_sp_normalize(const float *__restrict__ img, int *coef, SpInfo *info, SpGeom g)
{
    const int frame = blockIdx.y;
    if (!info[frame].active) return;
    const size_t n = (size_t)g.Wp * g.Hp;
    int mx = 0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const int v = __float2int_rz(img[(size_t)frame * n + i]);
        coef[(size_t)frame * n + i] = v;
        mx = max(mx, v < 0 ? -v : v);
    }
    for (int o = 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if ((threadIdx.x & 31) == 0 && mx) atomicMax(&info[frame].maxabs, mx);
}

// ---- tree helpers (get_successor, spiht_re.c:127-158) ------------------------------------------------------
__device__ __forceinline__ bool sp_succ(const SpGeom &g, int x, int y, int &sx, int &sy)
{
    if (x < g.lx && y < g.ly) {
        sx = (x & 1) ? x + g.lx - 1 : x;
        sy = (y & 1) ? y + g.ly - 1 : y;
        return !(sx == x && sy == y);
    }
    sx = 2 * x; sy = 2 * y;
    return !(sx >= g.Wp || sy >= g.Hp);
}

// D[i] = max |C| over all descendants, L[i] = max over descendants that are not children.
// One launch per generation, finest first: nodes are the pixels of [0,X1)x[0,Y1) outside [0,X0)x[0,Y0).
__global__ void k_sp_desc(const int *__restrict__ coef, int *D, int *L, const SpInfo *__restrict__ info, SpGeom g,
                          int X0, int Y0, int X1, int Y1)
{
    const int frame = blockIdx.y;
    if (!info[frame].active) return;
    const size_t fo = (size_t)frame * g.Wp * g.Hp;
    const int n = X1 * Y1;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int y = i / X1, x = i - y * X1;
        if (x < X0 && y < Y0) continue;
        int sx, sy, d = 0, l = 0;
        if (sp_succ(g, x, y, sx, sy)) {
            for (int k = 0; k < 4; ++k) {
                const size_t c = fo + (size_t)(sy + (k >> 1)) * g.Wp + sx + (k & 1);
                const int v = coef[c], a = v < 0 ? -v : v, dc = D[c];
                d = max(d, max(a, dc)); l = max(l, dc);
            }
        }
        D[fo + (size_t)y * g.Wp + x] = d; L[fo + (size_t)y * g.Wp + x] = l;
    }
}

// ---- the coder -------------------------------------------------------------------------------------------
struct SpBufs {
    int *coef, *D, *L;               // [B][Np]
    unsigned *lip0, *lip1, *lsp;     // [B][Np]
    int *lis0, *lis1;                // [B][2*Np]
    unsigned *sign_pos, *lsp_idx;    // [B][Np]   sign_pos: 1-based payload index of the sign bit, 0 = never coded
    unsigned *bits;                  // [B][bits_words]  bitstream, MSB-first bytes
    size_t bits_words;
};

__device__ __forceinline__ void sp_setbit(unsigned *bits, unsigned abs_bit)
{
    atomicOr(&bits[abs_bit >> 5], 1u << ((((abs_bit >> 3) & 3u) << 3) + (7u - (abs_bit & 7u))));
}

// CTA-wide exclusive scan of one 64-bit value per thread (four packed 16-bit counters); returns the exclusive
// prefix, *total receives the CTA sum.  scratch: 33 x u64 in shared memory.
__device__ __forceinline__ unsigned long long sp_scan4(unsigned long long v, unsigned long long *scratch, unsigned long long *total)
{
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    unsigned long long incl = v;
    for (int o = 1; o < 32; o <<= 1) { unsigned long long t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
    if (lane == 31) scratch[w] = incl;
    __syncthreads();
    if (w == 0) {
        unsigned long long s = scratch[lane], si = s;
        for (int o = 1; o < 32; o <<= 1) { unsigned long long t = __shfl_up_sync(0xffffffffu, si, o); if (lane >= o) si += t; }
        scratch[lane] = si - s;
        if (lane == 31) scratch[32] = si;
    }
    __syncthreads();
    const unsigned long long r = scratch[w] + incl - v;
    *total = scratch[32];
    __syncthreads();
    return r;
}

#define SP_HDR_BITS 120u

__global__ void __launch_bounds__(SP_THREADS)
k_sp_encode(SpBufs b, SpInfo *info, SpGeom g, const int *__restrict__ step_of_pow2)
{
    __shared__ unsigned long long scratch[33];
    __shared__ unsigned s_bitpos, s_lip_n, s_lsp_n, s_lis_tail, s_prev_lsp;
    const int frame = blockIdx.x, tid = threadIdx.x;
    SpInfo &inf = info[frame];
    if (!inf.active) return;
    const size_t Np = (size_t)g.Wp * g.Hp, fo = (size_t)frame * Np;
    const int *C = b.coef + fo, *D = b.D + fo, *Lm = b.L + fo;
    unsigned *lip[2] = {b.lip0 + fo, b.lip1 + fo};
    int *lis[2] = {b.lis0 + 2 * fo, b.lis1 + 2 * fo};
    unsigned *lsp = b.lsp + fo, *sign_pos = b.sign_pos + fo, *lsp_idx = b.lsp_idx + fo;
    unsigned *bits = b.bits + (size_t)frame * b.bits_words;
    const unsigned budget = inf.budget;                 // bits 0..budget (0-based) are emitted, later ones dropped

    // step = floor(log(max) / log(2)) with max >= 2 (spiht_re.c:52-60); exact powers of two go through the
    // host-computed table so that libm's rounding of the quotient is reproduced
    int mxa = max(inf.maxabs, 2);
    int step0 = 31 - __clz(mxa);
    if ((mxa & (mxa - 1)) == 0) step0 = step_of_pow2[step0];

    // initial lists (spiht_re.c:64-77)
    const int nll = g.lx * g.ly;
    for (int i = tid; i < nll; i += SP_THREADS) {
        const int y = i / g.lx, x = i - y * g.lx;
        lip[0][i] = (unsigned)(x + y * g.Wp);
    }
    if (tid == 0) {
        // LIS: LL pixels with an odd coordinate, row-major; serial fill keeps the order (nll <= 65k, once per frame)
        unsigned n = 0;
        for (int y = 0; y < g.ly; ++y)
            for (int x = 0; x < g.lx; ++x)
                if ((x & 1) || (y & 1)) lis[0][n++] = x + y * g.Wp + 1;
        s_lis_tail = n; s_bitpos = 0; s_lip_n = (unsigned)nll; s_lsp_n = 0; s_prev_lsp = 0;
        // header: 'I' 'M' 'S', stages(6), size_x(12), size_y(12), extra_x(10), extra_y(10), colour(1), bits0(29), dc(8), step(8)
        unsigned pos = 0;
        unsigned char *bytes = (unsigned char *)bits;
        const unsigned long long fields[11][2] = {{'I', 8}, {'M', 8}, {'S', 8}, {SP_STAGES, 6}, {(unsigned long long)g.W, 12},
            {(unsigned long long)g.H, 12}, {(unsigned long long)(g.Wp - g.W), 10}, {(unsigned long long)(g.Hp - g.H), 10}, {0, 1},
            {(unsigned long long)budget + 128ull, 29}, {(unsigned long long)inf.dc, 8}};
        for (int f = 0; f < 11; ++f)
            for (int k = (int)fields[f][1] - 1; k >= 0; --k) { if ((fields[f][0] >> k) & 1) bytes[pos >> 3] |= (unsigned char)(0x80u >> (pos & 7)); pos++; }
        for (int k = 7; k >= 0; --k) { if ((step0 >> k) & 1) bytes[pos >> 3] |= (unsigned char)(0x80u >> (pos & 7)); pos++; }
        inf.step0 = step0;
    }
    __syncthreads();
    int cur_lip = 0, cur_lis = 0;
    bool stop = false;

    for (int step = step0; step >= 0 && !stop; --step) {
        const int thr = 1 << step;
        // ---- LIP pass ----
        {
            const unsigned n = s_lip_n, lsp_base = s_lsp_n;
            unsigned bitpos = s_bitpos, surv = 0, nlsp = 0;
            __syncthreads();
            for (unsigned base = 0; base < n; base += SP_THREADS) {
                const unsigned i = base + tid;
                unsigned pix = 0; int v = 0; unsigned sig = 0, valid = i < n;
                if (valid) { pix = lip[cur_lip][i]; v = C[pix]; sig = (v < 0 ? -v : v) >= thr; }
                unsigned long long cnt = valid ? ((unsigned long long)(1 + sig) | ((unsigned long long)sig << 16) | ((unsigned long long)(sig ^ 1u) << 32)) : 0ull;
                unsigned long long tot, ex = sp_scan4(cnt, scratch, &tot);
                if (valid) {
                    const unsigned p = bitpos + (unsigned)(ex & 0xffff);
                    if (sig) {
                        if (p <= budget) sp_setbit(bits, SP_HDR_BITS + p);
                        if (v < 0 && p + 1 <= budget) sp_setbit(bits, SP_HDR_BITS + p + 1);
                        sign_pos[pix] = p + 2;
                        const unsigned li = lsp_base + nlsp + (unsigned)((ex >> 16) & 0xffff);
                        lsp[li] = pix; lsp_idx[pix] = li;
                    } else lip[cur_lip ^ 1][surv + (unsigned)((ex >> 32) & 0xffff)] = pix;
                }
                bitpos += (unsigned)(tot & 0xffff); nlsp += (unsigned)((tot >> 16) & 0xffff); surv += (unsigned)((tot >> 32) & 0xffff);
                if (bitpos > budget) { stop = true; break; }
            }
            __syncthreads();
            if (tid == 0) { s_bitpos = bitpos; s_lip_n = surv; s_lsp_n = lsp_base + nlsp; }
            cur_lip ^= 1;
            __syncthreads();
            if (stop) break;
        }
        // ---- LIS pass, wave by wave ----
        {
            unsigned wb = 0, we = s_lis_tail, tail = s_lis_tail;
            unsigned bitpos = s_bitpos, lip_n = s_lip_n, lsp_n = s_lsp_n;
            int *ls = lis[cur_lis];
            unsigned *lp = lip[cur_lip];
            __syncthreads();
            while (wb < we && !stop) {
                for (unsigned base = wb; base < we; base += SP_THREADS) {
                    const unsigned i = base + tid;
                    const bool valid = i < we;
                    int item = 0, x = 0, y = 0, sx = 0, sy = 0;
                    unsigned sig = 0, isA = 0, csig = 0, cneg = 0, nbits = 0, nlsp = 0, nlip = 0, nlis = 0;
                    if (valid) {
                        item = ls[i];
                        isA = item > 0;
                        const int pix = isA ? item - 1 : -item - 1;
                        y = pix / g.Wp; x = pix - y * g.Wp;
                        nbits = 1;
                        if (isA) {
                            sig = D[pix] >= thr;
                            if (sig) {
                                sp_succ(g, x, y, sx, sy);
                                for (int k = 0; k < 4; ++k) {
                                    const int cv = C[(sy + (k >> 1)) * g.Wp + sx + (k & 1)];
                                    if ((cv < 0 ? -cv : cv) >= thr) { csig |= 1u << k; if (cv < 0) cneg |= 1u << k; }
                                }
                                nlsp = (unsigned)__popc(csig); nlip = 4 - nlsp; nbits += 4 + nlsp;
                                int gx, gy;
                                nlis = sp_succ(g, sx, sy, gx, gy) ? 1u : 0u;
                            }
                        } else {
                            sig = Lm[pix] >= thr;
                            if (sig) { sp_succ(g, x, y, sx, sy); nlis = 4; }
                        }
                    }
                    unsigned long long cnt = (unsigned long long)nbits | ((unsigned long long)nlsp << 16) | ((unsigned long long)nlip << 32) | ((unsigned long long)nlis << 48);
                    unsigned long long tot, ex = sp_scan4(cnt, scratch, &tot);
                    if (valid && sig) {
                        unsigned p = bitpos + (unsigned)(ex & 0xffff);
                        if (p <= budget) sp_setbit(bits, SP_HDR_BITS + p);
                        p++;
                        unsigned ol = lsp_n + (unsigned)((ex >> 16) & 0xffff), oi = lip_n + (unsigned)((ex >> 32) & 0xffff);
                        unsigned ot = tail + (unsigned)((ex >> 48) & 0xffff);
                        if (isA) {
                            for (int k = 0; k < 4; ++k) {
                                const unsigned cp = (unsigned)((sy + (k >> 1)) * g.Wp + sx + (k & 1));
                                if ((csig >> k) & 1) {
                                    if (p <= budget) sp_setbit(bits, SP_HDR_BITS + p);
                                    if (((cneg >> k) & 1) && p + 1 <= budget) sp_setbit(bits, SP_HDR_BITS + p + 1);
                                    sign_pos[cp] = p + 2;
                                    lsp[ol] = cp; lsp_idx[cp] = ol; ol++;
                                    p += 2;
                                } else { lp[oi++] = cp; p += 1; }
                            }
                            if (nlis) ls[ot] = -(x + y * g.Wp + 1);
                        } else {
                            for (int k = 0; k < 4; ++k) ls[ot + k] = (sy + (k >> 1)) * g.Wp + sx + (k & 1) + 1;
                        }
                        ls[i] = 0;                                   // removed
                    }
                    bitpos += (unsigned)(tot & 0xffff); lsp_n += (unsigned)((tot >> 16) & 0xffff);
                    lip_n += (unsigned)((tot >> 32) & 0xffff); tail += (unsigned)((tot >> 48) & 0xffff);
                    if (bitpos > budget) { stop = true; break; }
                }
                wb = we; we = tail;
                __syncthreads();
            }
            // consolidate (ml_consolidate): surviving items keep their order
            unsigned kept = 0;
            if (!stop) {
                int *ld = lis[cur_lis ^ 1];
                for (unsigned base = 0; base < tail; base += SP_THREADS) {
                    const unsigned i = base + tid;
                    const int item = i < tail ? ls[i] : 0;
                    unsigned long long tot, ex = sp_scan4(item != 0 ? 1ull : 0ull, scratch, &tot);
                    if (item != 0) ld[kept + (unsigned)ex] = item;
                    kept += (unsigned)tot;
                }
                cur_lis ^= 1;
            }
            __syncthreads();
            if (tid == 0) { s_bitpos = bitpos; s_lip_n = lip_n; s_lsp_n = lsp_n; s_lis_tail = kept; }
            __syncthreads();
            if (stop) break;
        }
        // ---- refinement pass: one bit per LSP entry of earlier planes, positions are closed-form ----
        {
            const unsigned bitpos = s_bitpos, nprev = s_prev_lsp;
            for (unsigned i = tid; i < nprev; i += SP_THREADS) {
                const unsigned p = bitpos + i;
                if (p > budget) break;
                const int v = C[lsp[i]];
                if (((v < 0 ? -v : v) >> step) & 1) sp_setbit(bits, SP_HDR_BITS + p);
            }
            __syncthreads();
            if (tid == 0) {
                inf.ref_base[step] = bitpos; inf.lsp_before[step] = nprev;
                s_bitpos = bitpos + nprev; s_prev_lsp = s_lsp_n;
            }
            __syncthreads();
            if (s_bitpos > budget) stop = true;
        }
    }
    __syncthreads();
    if (tid == 0) {
        const unsigned total = s_bitpos;
        const unsigned emitted = total > budget + 1u ? budget + 1u : total;
        inf.emitted = emitted;
        inf.nbytes = (SP_HDR_BITS + emitted + 7u) >> 3;
    }
}

// ---- reconstruction of a bitstream prefix from the recorded event positions ----------------------------------
// P = number of payload bits the reference decoder would consume (spiht_decode: min(num_bits, bits0) - 128 + 1)
__global__ void k_sp_recon(SpBufs b, const SpInfo *__restrict__ info, const unsigned *__restrict__ Pbits, float *img, SpGeom g)
{
    const int frame = blockIdx.y;
    const unsigned P = Pbits[frame];
    if (P == 0) return;
    const SpInfo &inf = info[frame];
    const size_t n = (size_t)g.Wp * g.Hp, fo = (size_t)frame * n;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const int v = b.coef[fo + i];
        const unsigned sp = b.sign_pos[fo + i];
        float out = 0.0f;
        if (v != 0 && sp != 0 && sp <= P) {
            const int a = v < 0 ? -v : v, s = 31 - __clz(a);
            const unsigned li = b.lsp_idx[fo + i];
            int mag = 1 << s;
            for (int p = s - 1; p >= 0; --p) {
                const unsigned rb = inf.ref_base[p];
                if (rb == SP_NOPOS || li >= inf.lsp_before[p] || rb + li + 1u > P) break;
                mag |= a & (1 << p);
            }
            out = (float)(v < 0 ? -mag : mag);
        }
        img[fo + i] = out;
    }
}

// ---- inverse transform (dwt.h:114-140, 169-272, 282-317): columns first, then rows -----------------------------
// columns: src(data) -> dst(temp); block = (CW, TY); grid = (ceil(nx/CW), B)
__global__ void k_sp_idwt_col(const float *__restrict__ src, float *dst, const unsigned *__restrict__ Pbits, SpGeom g, int nx, int ny)
{
    EBCC_DYN_SMEM(float, sm);
    const int frame = blockIdx.y;
    if (Pbits[frame] == 0) return;
    const int CW = blockDim.x, c = threadIdx.x, col = blockIdx.x * CW + c;
    const bool act = col < nx;
    const size_t base = (size_t)frame * g.Wp * g.Hp;
    const int m = ny >> 1, t0 = threadIdx.y, ts = blockDim.y;
    float *lo = sm + c, *hi = sm + (size_t)m * CW + c, *t = sm + (size_t)ny * CW + c;     // lo[i*CW], hi[i*CW], t[i*CW]
    if (act) for (int i = t0; i < m; i += ts) {
        lo[i * CW] = __fdiv_rn(src[base + (size_t)i * g.Wp + col], SP_XI);
        hi[i * CW] = __fmul_rn(src[base + (size_t)(m + i) * g.Wp + col], SP_XI);
    }
    __syncthreads();
    if (act) for (int i = t0; i < m; i += ts) {
        const float s = i == 0 ? __fadd_rn(hi[0], hi[CW]) : __fadd_rn(hi[i * CW], hi[(i - 1) * CW]);
        lo[i * CW] = __fsub_rn(lo[i * CW], __fmul_rn(SP_DELTA, s));
    }
    __syncthreads();
    if (act) for (int i = t0; i < m; i += ts) {
        const float s = i < m - 1 ? __fadd_rn(lo[i * CW], lo[(i + 1) * CW]) : __fadd_rn(lo[(m - 1) * CW], lo[(m - 2) * CW]);
        hi[i * CW] = __fsub_rn(hi[i * CW], __fmul_rn(SP_GAMMA, s));
    }
    __syncthreads();
    if (act) for (int i = t0; i < m; i += ts) {
        const float s = i == 0 ? __fadd_rn(hi[0], hi[CW]) : __fadd_rn(hi[i * CW], hi[(i - 1) * CW]);
        t[(2 * i) * CW] = __fsub_rn(lo[i * CW], __fmul_rn(SP_BETA, s));
    }
    __syncthreads();
    if (act) for (int i = t0; i < m; i += ts) {
        float v;
        if (i < m - 1) v = __fsub_rn(hi[i * CW], __fmul_rn(SP_ALPHA, __fadd_rn(t[(2 * i) * CW], t[(2 * i + 2) * CW])));
        else v = __fsub_rn(hi[i * CW], __fmul_rn(__fmul_rn(2.0f, SP_ALPHA), t[(ny - 2) * CW]));
        t[(2 * i + 1) * CW] = v;
    }
    __syncthreads();
    if (act) for (int i = t0; i < ny; i += ts) dst[base + (size_t)i * g.Wp + col] = t[i * CW];
}

// rows: src(temp) -> dst(data); one row per CTA
__global__ void k_sp_idwt_row(const float *__restrict__ src, float *dst, const unsigned *__restrict__ Pbits, SpGeom g, int nx)
{
    EBCC_DYN_SMEM(float, sm);
    const int frame = blockIdx.y, row = blockIdx.x;
    if (Pbits[frame] == 0) return;
    const size_t off = (size_t)frame * g.Wp * g.Hp + (size_t)row * g.Wp;
    const int m = nx >> 1, t0 = threadIdx.x, ts = blockDim.x;
    float *lo = sm, *hi = sm + m, *t = sm + nx;
    for (int i = t0; i < m; i += ts) {
        lo[i] = __fdiv_rn(src[off + i], SP_XI);
        hi[i] = __fmul_rn(src[off + m + i], SP_XI);
    }
    __syncthreads();
    for (int i = t0; i < m; i += ts) {
        const float s = i == 0 ? __fadd_rn(hi[0], hi[1]) : __fadd_rn(hi[i], hi[i - 1]);
        lo[i] = __fsub_rn(lo[i], __fmul_rn(SP_DELTA, s));
    }
    __syncthreads();
    for (int i = t0; i < m; i += ts) {
        const float s = i < m - 1 ? __fadd_rn(lo[i], lo[i + 1]) : __fadd_rn(lo[m - 1], lo[m - 2]);
        hi[i] = __fsub_rn(hi[i], __fmul_rn(SP_GAMMA, s));
    }
    __syncthreads();
    for (int i = t0; i < m; i += ts) {
        const float s = i == 0 ? __fadd_rn(hi[0], hi[1]) : __fadd_rn(hi[i], hi[i - 1]);
        t[2 * i] = __fsub_rn(lo[i], __fmul_rn(SP_BETA, s));
    }
    __syncthreads();
    for (int i = t0; i < m; i += ts) {
        float v;
        if (i < m - 1) v = __fsub_rn(hi[i], __fmul_rn(SP_ALPHA, __fadd_rn(t[2 * i], t[2 * i + 2])));
        else v = __fsub_rn(hi[i], __fmul_rn(__fmul_rn(2.0f, SP_ALPHA), t[nx - 2]));
        t[2 * i + 1] = v;
    }
    __syncthreads();
    for (int i = t0; i < nx; i += ts) dst[off + i] = t[i];
}

// add_dc + /255 + de-normalisation (dwt.h:336-353, spiht_re.c:511-515, src/ebcc_codec.c:751-753)
__device__ __forceinline__ float sp_pixel_to_residual(float v, float dc, float rmin, float range)
{
    float t = floorf(__fadd_rn(v, dc));
    t = t > 255.0f ? 255.0f : (t < 0.0f ? 0.0f : t);
    return __fadd_rn(__fmul_rn(__fdiv_rn(t, 255.0f), range), rmin);
}

// error statistics of (decoded + residual) against the original: get_max_error / get_mean_error
__global__ void k_sp_errstat(const float *__restrict__ img, const float *__restrict__ orig, const float *__restrict__ dec,
                             const SpInfo *__restrict__ info, const unsigned *__restrict__ Pbits, ErrStat *es, SpGeom g)
{
    const int frame = blockIdx.y;
    if (Pbits[frame] == 0) return;
    const float rmin = info[frame].rmin, range = __fsub_rn(info[frame].rmax, rmin), dc = (float)info[frame].dc;
    const size_t n = (size_t)g.W * g.H, np = (size_t)g.Wp * g.Hp;
    float mx = 0.0f; long long sum = 0;
    // |orig - dec| and the de-normalised residual both lie within the residual range: |e| <= 2 max(|rmin|, |rmax|) < 2^ex
    const int ex = err_fx_exp(__fmul_rn(2.0f, fmaxf(fabsf(rmin), fabsf(info[frame].rmax))));
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const int y = (int)(i / g.W), x = (int)(i - (size_t)y * g.W);
        const float r = sp_pixel_to_residual(img[(size_t)frame * np + (size_t)y * g.Wp + x], dc, rmin, range);
        const float e = __fsub_rn(orig[(size_t)frame * n + i], __fadd_rn(dec[(size_t)frame * n + i], r));
        mx = fmaxf(mx, fabsf(e)); sum += err_fx(e, ex);
    }
    for (int o = 16; o > 0; o >>= 1) { mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o)); sum += __shfl_xor_sync(0xffffffffu, sum, o); }
    if ((threadIdx.x & 31) == 0) { atomicMax(&es[frame].max_err_bits, __float_as_uint(mx)); atomicAdd((unsigned long long *)&es[frame].sum_fx, (unsigned long long)sum); }
    if (blockIdx.x == 0 && threadIdx.x == 0) es[frame].sum_exp = ex;
}

// decode path: out += residual
__global__ void k_sp_add_residual(const float *__restrict__ img, float *out, const SpInfo *__restrict__ info, SpGeom g)
{
    const int frame = blockIdx.y;
    if (!info[frame].active) return;
    const float rmin = info[frame].rmin, range = __fsub_rn(info[frame].rmax, rmin), dc = (float)info[frame].dc;
    const size_t n = (size_t)g.W * g.H, np = (size_t)g.Wp * g.Hp;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const int y = (int)(i / g.W), x = (int)(i - (size_t)y * g.W);
        const float r = sp_pixel_to_residual(img[(size_t)frame * np + (size_t)y * g.Wp + x], dc, rmin, range);
        out[(size_t)frame * n + i] = __fadd_rn(out[(size_t)frame * n + i], r);
    }
}

// ---- warp-cooperative list decoder for stored streams (spiht_re.c:319-430, 477-520) ----------------------------------
// One CTA of SPD_THREADS per frame; warp 0 walks the lists, all warps share the passes that need no parsing.  The list
// coder is serial only in WHERE an entry's bits sit: an entry of the LIP takes 1 or 2 bits, one of the LIS 1 to 9,
// depending on the bits themselves.  So warp 0 takes a list 32 entries at a time: a short warp-uniform loop over the
// chunk yields each entry's bit offset -- it works on a register window of the stream, skips a whole run of
// insignificant entries (one 0 bit each, whatever their kind) with one count-leading-zeros, and spends instructions
// only on significant ones; then the lanes pick their own bits out of the window (shuffles, no memory access), apply
// them to their own pixel and append to the lists at positions given by a warp prefix sum.  List loads (the next chunk
// is prefetched), pixel updates and appends of 32 entries travel together instead of one dependent global access after
// the other.  Entries appended to the LIS while it is scanned form the next wave, as in k_sp_encode.  The refinement
// pass is one bit per entry at a closed-form position: the whole CTA.  Truncated streams: the reference checks its bit
// budget after every bit; an action is applied iff the check before it passed, which is a comparison of the action's own
// bit count with the budget, and the decode ends with the chunk in which the budget ran out.
#define SPD_THREADS 128
struct SpWarpBits {
    const unsigned *words; int nw;       // the stream as 32-bit words (bytes in stream order), zero beyond nw
    int k0; unsigned held;               // lane l holds word k0 + l, MSB-first
};
__device__ __forceinline__ unsigned spw_load(const unsigned *words, int nw, int k)
{
    return (k >= 0 && k < nw) ? __byte_perm(words[k], 0u, 0x0123) : 0u;
}
// all lanes: make sure the words of a chunk that starts at bit pos (at most 9 words + look-ahead) are held
__device__ __forceinline__ void spw_seek(SpWarpBits &b, unsigned pos, int lane)
{
    const int k = (int)(pos >> 5);
    if (k < b.k0 || k + 14 > b.k0 + 32) { b.k0 = k; b.held = spw_load(b.words, b.nw, k + lane); }
}
// the 32 bits starting at pos, picked out of the held words (pos may differ per lane)
__device__ __forceinline__ unsigned spw_bits32(const SpWarpBits &b, unsigned pos)
{
    const int k = ((int)(pos >> 5) - b.k0) & 31;
    const unsigned hi = __shfl_sync(0xffffffffu, b.held, k), lo = __shfl_sync(0xffffffffu, b.held, (k + 1) & 31);
    return __funnelshift_l(lo, hi, pos & 31u);
}
// per thread, any position: straight from memory (header, refinement pass)
__device__ __forceinline__ unsigned spw_own(const SpWarpBits &b, unsigned pos)
{
    const int k = (int)(pos >> 5);
    return __funnelshift_l(spw_load(b.words, b.nw, k + 1), spw_load(b.words, b.nw, k), pos & 31u);
}

// grid = nframes, block = SPD_THREADS.  streams: [frame][stride]; lens[frame] = coeffs_size bytes (0 = none).
// img receives the integer-valued coefficient plane (float), geometry must match g.  status[frame] != 0 on malformed input.
__global__ void __launch_bounds__(SPD_THREADS)
k_sp_decode(const unsigned char *__restrict__ streams, size_t stride, const int *__restrict__ lens, SpBufs b,
            SpInfo *info, float *img, SpGeom g, int *status)
{
    __shared__ unsigned s_pos, s_cnt, s_lsp_prev, s_stop;
    const int frame = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned lt = (1u << lane) - 1u;
    SpInfo &inf = info[frame];
    const size_t Np = (size_t)g.Wp * g.Hp, fo = (size_t)frame * Np;
    if (lens[frame] <= 0) { if (tid == 0) { inf.active = 0; status[frame] = 0; } return; }
    float *data = img + fo;
    for (size_t i = tid; i < Np; i += SPD_THREADS) data[i] = 0.0f;
    if (tid == 0) { inf.active = 1; status[frame] = 1; }
    SpWarpBits wb;
    wb.words = (const unsigned *)(streams + (size_t)frame * stride); wb.nw = (lens[frame] + 3) >> 2; wb.k0 = -1000; wb.held = 0;
    // header (every thread reads it): 'I' 'M' 'S', stages(6), size_x(12), size_y(12), extra_x(10), extra_y(10), colour(1), bits0(29), dc(8), step(8)
    unsigned hp = 0;
    auto hdr = [&](int n) { const unsigned v = spw_own(wb, hp) >> (32 - n); hp += (unsigned)n; return v; };
    if (hdr(8) != 'I' || hdr(8) != 'M' || hdr(8) != 'S') return;
    const unsigned stages = hdr(6), sx_ = hdr(12), sy_ = hdr(12), ex_ = hdr(10), ey_ = hdr(10);
    hdr(1);
    const unsigned bits0 = hdr(29);
    if (stages != SP_STAGES || (int)sx_ != g.W || (int)sy_ != g.H || (int)(sx_ + ex_) != g.Wp || (int)(sy_ + ey_) != g.Hp) return;
    unsigned num_bits = (unsigned)lens[frame] * 8u;
    if (num_bits > bits0) num_bits = bits0;
    if (num_bits <= 128) return;
    num_bits -= 128;
    const int dc = (int)hdr(8), step0 = (int)hdr(8);
    if (tid == 0) inf.dc = dc;
    if (step0 > 30) return;
    unsigned *lip = b.lip0 + fo, *lsp = b.lsp + fo;
    int *lis = b.lis0 + 2 * fo;
    // initial lists (spiht_re.c:64-77): LIP = the LL band row-major, LIS = its pixels with an odd coordinate, same order
    const int half = g.lx >> 1;
    for (int i = tid; i < g.lx * g.ly; i += SPD_THREADS) {
        const int y = i / g.lx, x = i - y * g.lx;
        lip[i] = (unsigned)(x + y * g.Wp);
        if ((x & 1) || (y & 1)) {
            const int before = ((y + 1) >> 1) * half + (y >> 1) * g.lx;           // entries of the rows above: even rows hold lx / 2, odd rows lx
            lis[before + ((y & 1) ? x : (x >> 1))] = x + y * g.Wp + 1;
        }
    }
    unsigned lip_n = (unsigned)(g.lx * g.ly), lsp_n = 0, lis_n = (unsigned)(((g.ly + 1) >> 1) * half + (g.ly >> 1) * g.lx);
    unsigned pos = hp, cnt = 0;            // absolute bit position; payload bits consumed so far (warp 0 keeps them, shared memory hands them on)
    if (tid == 0) { status[frame] = 0; s_stop = 0; }
    __syncthreads();
    for (int step = step0; step >= 0; --step) {
        const float one = (float)(1 << step);
        if (warp == 0) {
            const unsigned lsp_prev = lsp_n;
            bool stop = false;
            // ---- LIP pass: 1 bit, + 1 sign bit when significant ----
            unsigned keep = 0;
            unsigned nxt = lane < (int)lip_n ? lip[lane] : 0u;
            for (unsigned base = 0; base < lip_n && !stop; base += 32) {
                const unsigned n = min(32u, lip_n - base);
                const bool valid = lane < (int)n;
                const unsigned pix = nxt;
                nxt = base + 32 + lane < lip_n ? lip[base + 32 + lane] : 0u;        // next chunk travels while this one is parsed
                spw_seek(wb, pos, lane);
                unsigned off = 0, myoff = 0, e = 0;
                while (e < n) {
                    const unsigned z = min((unsigned)__clz((int)spw_bits32(wb, pos + off)), n - e);   // run of insignificant entries
                    if (lane >= (int)e && lane < (int)(e + z)) myoff = off + ((unsigned)lane - e);
                    e += z; off += z;
                    if (e < n && z < 32u) { if (lane == (int)e) myoff = off; e += 1; off += 2; }   // a significant one: bit + sign
                }
                const unsigned own = spw_bits32(wb, pos + myoff);
                const bool sig = valid && (own >> 31), applied = valid && cnt + myoff + 1u <= num_bits;
                if (sig && applied) data[pix] = ((own >> 30) & 1u) ? -one : one;
                const unsigned bs = __ballot_sync(0xffffffffu, sig && applied), bk = __ballot_sync(0xffffffffu, valid && !sig);
                if (sig && applied) lsp[lsp_n + __popc(bs & lt)] = pix;
                if (valid && !sig) lip[keep + __popc(bk & lt)] = pix;
                lsp_n += __popc(bs); keep += __popc(bk);
                pos += off; cnt += off;
                if (cnt > num_bits) stop = true;
            }
            lip_n = keep;
            // ---- LIS pass, wave by wave ----
            unsigned wbeg = 0, wend = lis_n, tail = lis_n;
            while (wbeg < wend && !stop) {
                int nxi = wbeg + lane < wend ? lis[wbeg + lane] : 0;
                for (unsigned base = wbeg; base < wend && !stop; base += 32) {
                    const unsigned n = min(32u, wend - base);
                    const bool valid = lane < (int)n;
                    const int item = nxi;
                    nxi = base + 32 + lane < wend ? lis[base + 32 + lane] : 0;
                    const bool isA = item > 0;
                    const int pix = isA ? item - 1 : -item - 1;
                    const unsigned amask = __ballot_sync(0xffffffffu, isA);
                    spw_seek(wb, pos, lane);
                    unsigned off = 0, myoff = 0, e = 0;
                    while (e < n) {
                        const unsigned z = min((unsigned)__clz((int)spw_bits32(wb, pos + off)), n - e);
                        if (lane >= (int)e && lane < (int)(e + z)) myoff = off + ((unsigned)lane - e);
                        e += z; off += z;
                        if (e < n && z < 32u) {
                            unsigned len = 1;
                            if ((amask >> e) & 1u) {                    // children: significance bit, + sign when set
                                unsigned c = spw_bits32(wb, pos + off + 1u);
                                for (int k = 0; k < 4; ++k) { const unsigned cs = c >> 31; c <<= 1u + cs; len += 1u + cs; }
                            }
                            if (lane == (int)e) myoff = off;
                            e += 1; off += len;
                        }
                    }
                    const unsigned own = spw_bits32(wb, pos + myoff);
                    const bool sig = valid && (own >> 31) && cnt + myoff + 1u <= num_bits;       // entries beyond the budget do nothing
                    int sx = 0, sy = 0;
                    unsigned nlsp = 0, nlip = 0, nlis = 0, csig = 0, cneg = 0, capp = 0;
                    if (sig) {
                        const int y = pix / g.Wp, x = pix - y * g.Wp;
                        sp_succ(g, x, y, sx, sy);
                        if (isA) {
                            unsigned p = 1;                                   // next bit of this entry
                            for (int k = 0; k < 4; ++k) {
                                const unsigned cs = (own >> (31u - p)) & 1u;
                                const bool ok = cnt + myoff + p + 1u <= num_bits;      // the child's significance bit is inside the budget
                                ++p;
                                if (ok) capp |= 1u << k;
                                if (cs) { csig |= 1u << k; if ((own >> (31u - p)) & 1u) cneg |= 1u << k; ++p; }
                            }
                            nlsp = (unsigned)__popc(csig & capp); nlip = (unsigned)__popc(~csig & capp & 15u);
                            int gx, gy;
                            nlis = sp_succ(g, sx, sy, gx, gy) ? 1u : 0u;
                        } else nlis = 4;
                    }
                    // exclusive prefix sums of the three append counts (packed 10 bits each: at most 32 * 4 = 128 per chunk)
                    unsigned pk = nlsp | (nlip << 10) | (nlis << 20), inc = pk;
                    for (int o = 1; o < 32; o <<= 1) { const unsigned t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
                    const unsigned tot = __shfl_sync(0xffffffffu, inc, 31), ex = inc - pk;
                    if (sig) {
                        unsigned ol = lsp_n + (ex & 1023u), oi = lip_n + ((ex >> 10) & 1023u), ot = tail + (ex >> 20);
                        if (isA) {
                            for (int k = 0; k < 4; ++k) {
                                if (!((capp >> k) & 1u)) continue;
                                const unsigned cp = (unsigned)((sy + (k >> 1)) * g.Wp + sx + (k & 1));
                                if ((csig >> k) & 1u) { lsp[ol++] = cp; data[cp] = ((cneg >> k) & 1u) ? -one : one; }
                                else lip[oi++] = cp;
                            }
                            if (nlis) lis[ot] = -(pix + 1);
                        } else {
                            for (int k = 0; k < 4; ++k) lis[ot + k] = (sy + (k >> 1)) * g.Wp + sx + (k & 1) + 1;
                        }
                        lis[base + lane] = 0;                             // removed
                    }
                    lsp_n += tot & 1023u; lip_n += (tot >> 10) & 1023u; tail += tot >> 20;
                    pos += off; cnt += off;
                    if (cnt > num_bits) stop = true;
                }
                wbeg = wend; wend = tail;
                __syncwarp();
            }
            // consolidate: surviving items keep their order
            if (!stop) {
                keep = 0;
                for (unsigned base = 0; base < tail; base += 32) {
                    const int item = base + lane < tail ? lis[base + lane] : 0;
                    const unsigned bk = __ballot_sync(0xffffffffu, item != 0);
                    if (item != 0) lis[keep + __popc(bk & lt)] = item;
                    keep += __popc(bk);
                }
                lis_n = keep;
            }
            if (lane == 0) { s_pos = pos; s_cnt = cnt; s_lsp_prev = lsp_prev; s_stop = stop ? 1u : 0u; }
        }
        __syncthreads();
        if (s_stop) return;
        // ---- refinement pass: one bit per entry the LSP held before this plane; the whole CTA ----
        const unsigned rpos = s_pos, rcnt = s_cnt, lsp_prev = s_lsp_prev;
        for (unsigned i = tid; i < lsp_prev; i += SPD_THREADS) {
            if (rcnt + i <= num_bits) {                               // applied iff the check after the previous bit passed
                const unsigned pix = lsp[i];
                const int vi = (int)data[pix];
                if (spw_own(wb, rpos + i) >> 31) data[pix] = (float)(vi >= 0 ? (vi | (1 << step)) : -((-vi) | (1 << step)));
                else data[pix] = (float)(vi & ~(1 << step));
            }
        }
        pos = rpos + lsp_prev; cnt = rcnt + lsp_prev;
        if (cnt > num_bits) return;
        __syncthreads();
    }
}

// ---- reference-shaped serial decoder, kept for cross-checks (EBCC_SP_DECODE_SERIAL=1) ------------------------------------
// ---- serial decoder for stored streams (spiht_re.c:319-430, 477-520) ---------------------------------------------
struct SpBitIn { const unsigned char *p; unsigned nbits_avail; unsigned pos; };
__device__ __forceinline__ unsigned spb_get(SpBitIn &b)
{
    unsigned v = 0;
    if (b.pos < b.nbits_avail) v = (b.p[b.pos >> 3] >> (7 - (b.pos & 7))) & 1u;
    b.pos++;
    return v;
}
__device__ __forceinline__ unsigned spb_gets(SpBitIn &b, int n) { unsigned v = 0; for (int i = 0; i < n; ++i) v = (v << 1) | spb_get(b); return v; }

// grid = nframes, block = 32 (lane 0 decodes).  streams: [frame][stride]; lens[frame] = coeffs_size bytes (0 = none).
// img receives the integer-valued coefficient plane (float), geometry must match g.  status[frame] != 0 on malformed input.
__global__ void k_sp_decode_serial(const unsigned char *__restrict__ streams, size_t stride, const int *__restrict__ lens, SpBufs b,
                                   SpInfo *info, float *img, SpGeom g, int *status)
{
    const int frame = blockIdx.x;
    SpInfo &inf = info[frame];
    const size_t Np = (size_t)g.Wp * g.Hp, fo = (size_t)frame * Np;
    if (lens[frame] <= 0) { if (threadIdx.x == 0) { inf.active = 0; status[frame] = 0; } return; }
    float *data = img + fo;
    for (size_t i = threadIdx.x; i < Np; i += blockDim.x) data[i] = 0.0f;
    __syncthreads();
    if (threadIdx.x != 0) return;
    inf.active = 1; status[frame] = 1;
    SpBitIn bi; bi.p = streams + (size_t)frame * stride; bi.nbits_avail = (unsigned)lens[frame] * 8u; bi.pos = 0;
    if (spb_gets(bi, 8) != 'I' || spb_gets(bi, 8) != 'M' || spb_gets(bi, 8) != 'S') return;
    const unsigned stages = spb_gets(bi, 6), sx_ = spb_gets(bi, 12), sy_ = spb_gets(bi, 12), ex_ = spb_gets(bi, 10), ey_ = spb_gets(bi, 10);
    spb_get(bi);
    const unsigned bits0 = spb_gets(bi, 29);
    if (stages != SP_STAGES || (int)sx_ != g.W || (int)sy_ != g.H || (int)(sx_ + ex_) != g.Wp || (int)(sy_ + ey_) != g.Hp) return;
    unsigned num_bits = (unsigned)lens[frame] * 8u;
    if (num_bits > bits0) num_bits = bits0;
    if (num_bits <= 128) return;
    num_bits -= 128;
    inf.dc = (int)spb_gets(bi, 8);
    int step0 = (int)spb_gets(bi, 8);
    if (step0 > 30) return;
    unsigned *lip = b.lip0 + fo, *lsp = b.lsp + fo;
    int *lis = b.lis0 + 2 * fo;
    unsigned lip_n = 0, lsp_n = 0, lis_n = 0, cnt = 0;
    for (int y = 0; y < g.ly; ++y)
        for (int x = 0; x < g.lx; ++x) {
            lip[lip_n++] = (unsigned)(x + y * g.Wp);
            if ((x & 1) || (y & 1)) lis[lis_n++] = x + y * g.Wp + 1;
        }
    status[frame] = 0;
#define SP_COUNT() do { if (++cnt > num_bits) return; } while (0)
    for (int step = step0; step >= 0; --step) {
        const float one = (float)(1 << step);
        unsigned keep = 0;
        const unsigned lsp_prev = lsp_n;
        for (unsigned i = 0; i < lip_n; ++i) {
            const unsigned pix = lip[i];
            const unsigned sig = spb_get(bi);
            SP_COUNT();
            if (sig) { lsp[lsp_n++] = pix; data[pix] = spb_get(bi) ? -one : one; SP_COUNT(); }
            else lip[keep++] = pix;
        }
        lip_n = keep;
        unsigned tail = lis_n;
        for (unsigned i = 0; i < tail; ++i) {
            const int item = lis[i];
            const int pix = item > 0 ? item - 1 : -item - 1;
            const int y = pix / g.Wp, x = pix - y * g.Wp;
            int sx, sy;
            const unsigned sig = spb_get(bi);
            SP_COUNT();
            if (!sig) continue;
            sp_succ(g, x, y, sx, sy);
            if (item > 0) {
                for (int k = 0; k < 4; ++k) {
                    const unsigned cp = (unsigned)((sy + (k >> 1)) * g.Wp + sx + (k & 1));
                    const unsigned cs = spb_get(bi);
                    SP_COUNT();
                    if (cs) { lsp[lsp_n++] = cp; data[cp] = spb_get(bi) ? -one : one; SP_COUNT(); }
                    else lip[lip_n++] = cp;
                }
                int gx, gy;
                if (sp_succ(g, sx, sy, gx, gy)) lis[tail++] = -(pix + 1);
            } else {
                for (int k = 0; k < 4; ++k) lis[tail++] = (sy + (k >> 1)) * g.Wp + sx + (k & 1) + 1;
            }
            lis[i] = 0;
        }
        keep = 0;
        for (unsigned i = 0; i < tail; ++i) if (lis[i] != 0) lis[keep++] = lis[i];
        lis_n = keep;
        for (unsigned i = 0; i < lsp_prev; ++i) {
            const unsigned pix = lsp[i];
            const int vi = (int)data[pix];
            if (spb_get(bi)) data[pix] = (float)(vi >= 0 ? (vi | (1 << step)) : -((-vi) | (1 << step)));
            else data[pix] = (float)(vi & ~(1 << step));
            SP_COUNT();
        }
    }
#undef SP_COUNT
}
