// common.cuh -- shared declarations of the device layer (sm_100a).
//
// Data layout in HBM for a batch of B frames of H x W float32 (one JPEG 2000 tile each):
//   in     [B][H][W] f32   input frames (H2D copy of the caller's buffer)
//   coef   [B][H][W] f32   9/7 coefficients, Mallat layout, transformed in place level by level
//   stat   [B]             min / max / flags per frame
//   sym    [B][sym_stride] u8  EBCOT symbol stream per code-block: (ctx << 2 | bit), scan order
//   cbout  [B][ncblks]     per code-block pass table (symbol counts, nmsedec, rates, distortion)
//   mq     [B][mq_stride]  u8  MQ-coded bytes per code-block
//   out    [B][out_stride] u8  finished J2K codestreams
#pragma once
#ifdef EBCC_CUEMU
#include "cuemu.h"
#else
#include <cuda_runtime.h>
#include <stdint.h>
// every kernel launch of the library goes through this macro and is counted process-wide (ebcc_cuda_total_launches)
static unsigned long long g_ebcc_launches = 0;
#define EBCC_LAUNCH(kern, grid, block, smem, stream, ...) \
    (__atomic_fetch_add(&g_ebcc_launches, 1ull, __ATOMIC_RELAXED), kern<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__))
#define EBCC_DYN_SMEM(type, name) extern __shared__ __align__(16) unsigned char name##_raw_[]; type *name = (type *)name##_raw_
#endif
#include <stddef.h>

#define J2K_NRES 6
#define J2K_MAXBANDS 16
#define J2K_MAXPASSES 100
#define J2K_CBLK 64

// frame statistics written by k_minmax / k_minmax_final (reference: findMinMaxf + check_nan_inf,
// src/ebcc_codec.c:515-533, :598-605)
struct FrameStat {
    float minv, maxv;
    int flags;              // bit0: constant field, bit1: NaN/Inf present
    int pad;
    unsigned long long minkey, maxkey;
};
#define STAT_CONST 1
#define STAT_NANINF 2
#define STAT_OVERFLOW 4    // a code-block outgrew its slice of the symbol or MQ pool: the frame fails loudly

// per-frame error statistics of one reconstruction against the original frame
struct ErrStat {
    unsigned long long n_over;     // samples with |d - rec| > target           (get_error_target_quantile)
    unsigned max_err_bits;         // max |d - rec| as float bits                 (get_max_error)
    unsigned rmin_key, rmax_key;   // ordered keys of min / max of (d - rec)      (findMinMaxf on the residual)
    int sum_exp;                   // sum_fx counts units of 2^(sum_exp - ERR_FX_BITS)
    long long sum_fx;              // sum of (d - rec), fixed point               (get_mean_error)
};

struct BandDesc {
    int resno, orient, level;
    int offx, offy, bw, bh;      // position / size inside the Mallat-layout tile buffer
    int cw, ch, first_cblk;
    int numbps, expn, mant;
    float stepsize;              // encoder Delta_b (tcd.c:1069-1084)
    float stepsize_dec;          // decoder Delta_b (log2_gain 0)
    double wnorm;                // norm(level,orient) * stepsize / 2^log2_gain   (t1.c:1416-1449)
    // tag-tree shape of the band's single precinct (tgt.c:48-150): level 0 = leaves
    int tt_nlev, tt_base;        // number of levels, first node index inside the frame's node arrays
    int tt_off[8], tt_w[8], tt_h[8];
};

struct CblkDesc {
    int band;
    int tx, ty, w, h;            // rectangle inside the tile buffer
    int gx, gy;                  // position in the band's code-block grid
    int orient;
    unsigned sym_off, sym_cap;   // byte offset / capacity inside the frame's symbol pool
    unsigned mq_off, mq_cap;     // byte offset / capacity inside the frame's MQ pool
};

// per code-block result of tier-1 (opj_tcd_cblk_enc_t + opj_tcd_pass_t, tcd.h)
struct CblkOut {
    int numbps;
    int totalpasses;
    int datalen;
    int numpasses;               // passes selected by PCRD
    unsigned nsym_end[J2K_MAXPASSES];   // cumulative symbol count at the end of each pass
    int nmsedec[J2K_MAXPASSES];
    int rate[J2K_MAXPASSES];            // cumulative bytes (after OPJ's fix-ups)
    double dist[J2K_MAXPASSES];         // cumulative distortion decrease
    unsigned mqs[J2K_MAXPASSES][3];     // MQ coder at the end of each pass (compression-ratio search): c | ct << 28, a | pending byte << 16, index of the pending byte
};

// What the compression-ratio search knows about a code-block's truncation points.  The recording (value after all
// passes + pass of significance per sample, written by k_t1_cm) covers np_dec passes.  Bit p of bad[] is set when
// "decode only passes 0..p from the segment truncated at rate[p]" is NOT proven identical to the first p+1 passes of
// a decode of the complete segment (OPJ truncates non-terminated MQ segments at numbytes+3, which is not always
// enough; mq_cut_proven in k_t1.cuh).  scratch_np: pass count of the real truncated decode the scratch plane
// currently holds for this block (0 = none).
// first_entry: index of the block's first entry in the pool of pre-decoded unproven cuts (k_t1_dec_cuts), its entries
// follow in pass order; < 0: not pooled, such a cut is decoded when a probe asks for it.
struct T1CutState { int np_dec; int scratch_np; unsigned bad[4]; int first_entry; int pad; };
struct T1CutEntry { unsigned bi; int np; };        // block (frame * ncblks + cblk), passes to decode from the segment cut at rate[np - 1]
#define T1_CUT_WORDS (J2K_CBLK * J2K_CBLK)         // ints per pool entry: the decoded block, row stride J2K_CBLK
// MQ context states (pair index = state * 2 + mps, one byte each) at the end of every clean-up pass of a block: what a decode
// that starts at a bit-plane boundary needs besides the coder registers in CblkOut::mqs (k_t1_mq writes, k_t1_dec_cuts reads)
#define T1_SNAP_PLANES ((J2K_MAXPASSES + 2) / 3)
#define T1_SNAP_BYTES 20

struct Plan {
    int H, W, nbands, ncblks;
    // tile of a multi-tile image (src/ebcc_codec.c:105-123: dims[0] tiles of dims[1] x dims[2] stacked vertically):
    // canvas row of the tile's first line, its index (Isot) and the tile count; a single-tile image is {0, 0, 1}
    int y0, tile_index, ntiles;
    int cas_v[J2K_NRES];              // row parity of each resolution's origin on the canvas (dwt.c:1986-1987, 3575)
    int res_w[J2K_NRES], res_h[J2K_NRES];
    BandDesc bands[J2K_MAXBANDS];
    int nnodes;                       // tag-tree nodes over all bands
    int res_first[J2K_NRES], res_count[J2K_NRES];   // code-block index range of each resolution's packet
    size_t sym_stride, mq_stride, out_stride;
    int main_header_len;
    unsigned char main_header[160];
};
