// k_zstd.cuh -- the zstd stage of the residual layer as device kernels (format: RFC 8878).
//
// The reference calls ZSTD_compress(level 22) on the truncated SPIHT byte string and ZSTD_decompress on the way
// back (src/ebcc_codec.c:813-817, :1301).  The contract is the FORMAT, not the bytes: k_zstd_compress must emit a
// frame any conforming decoder accepts, k_zstd_decompress must accept any frame the reference's level-22 encoder
// can produce for these payloads (raw / RLE / compressed blocks, Huffman literals with FSE-compressed or direct
// weights, 1 or 4 streams, treeless literals, predefined / RLE / FSE / repeat sequence tables, repeat offsets).
//
// Payloads are a few KB (the SPIHT prefix kept by the truncation search), so both kernels run one frame per CTA
// with a single working thread; entropy coding of such short strings is a serial dependency chain.
//
// Compressor: byte runs (offset-1 matches, which is what dominates SPIHT output: long zero runs in the
// significance passes) become sequences with repeat-offset code 1 and predefined LL/ML tables, everything else
// goes to Huffman-coded literals (length-limited canonical code, FSE-compressed weight table).
#pragma once
#include "common.cuh"

#define ZC_THREADS 32
#define Z_MAGIC 0xFD2FB528u
#define Z_BLOCK_MAX (128 * 1024)
#define Z_CHUNK (16 * 1024)          // block size of the compressor: one block per CTA, blocks of a frame in parallel
#define Z_HUF_MAXLOG 11

// ---- shared tables ----------------------------------------------------------------------------------------
__device__ const short z_ll_default[36] = {4, 3, 2, 2, 2, 2, 2, 2, 2, 2, 2, 2, 2, 1, 1, 1, 2, 2, 2, 2, 2, 2, 2, 2, 2, 3, 2, 1, 1, 1, 1, 1, -1, -1, -1, -1};
__device__ const short z_ml_default[53] = {1, 4, 3, 2, 2, 2, 2, 2, 2, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, -1, -1, -1, -1, -1, -1, -1};
__device__ const short z_of_default[29] = {1, 1, 1, 1, 1, 1, 2, 2, 2, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, -1, -1, -1, -1, -1};
__device__ const unsigned z_ll_base[36] = {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16, 18, 20, 22, 24, 28, 32, 40, 48, 64, 128, 256, 512, 1024, 2048, 4096, 8192, 16384, 32768, 65536};
__device__ const unsigned char z_ll_bits[36] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 3, 3, 4, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16};
__device__ const unsigned z_ml_base[53] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16, 17, 18, 19, 20, 21, 22, 23, 24, 25, 26, 27, 28, 29, 30, 31, 32, 33, 34, 35, 37, 39, 41, 43, 47, 51, 59, 67, 83, 99, 131, 259, 515, 1027, 2051, 4099, 8195, 16387, 32771, 65539};
__device__ const unsigned char z_ml_bits[53] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 3, 3, 4, 4, 5, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16};

struct ZFseEntry { unsigned short base; unsigned char sym, nbits; };

__device__ __forceinline__ int z_highbit(unsigned v) { return 31 - __clz((int)v); }

// FSE decoding table from normalised counts (RFC 8878 4.1.1); norm[s] == -1: "less than one"
__device__ void z_fse_build(ZFseEntry *tab, int log, const short *norm, int nsym)
{
    const int size = 1 << log, mask = size - 1;
    unsigned short next[256];
    int high = size - 1;
    for (int s = 0; s < nsym; ++s) {
        if (norm[s] == -1) { tab[high--].sym = (unsigned char)s; next[s] = 1; }
        else next[s] = (unsigned short)norm[s];
    }
    const int step = (size >> 1) + (size >> 3) + 3;
    int pos = 0;
    for (int s = 0; s < nsym; ++s)
        for (int i = 0; i < norm[s]; ++i) {
            tab[pos].sym = (unsigned char)s;
            pos = (pos + step) & mask;
            while (pos > high) pos = (pos + step) & mask;
        }
    for (int u = 0; u < size; ++u) {
        const int s = tab[u].sym;
        const unsigned ns = next[s]++;
        const int nb = log - z_highbit(ns);
        tab[u].nbits = (unsigned char)nb;
        tab[u].base = (unsigned short)((ns << nb) - size);
    }
}

// ---- forward (LSB-first) bit reader used by table descriptions ---------------------------------------------
struct ZFwdBits { const unsigned char *p; int n; long long pos; };
__device__ __forceinline__ unsigned zf_peek(const ZFwdBits &b, int nb)
{
    unsigned long long v = 0;
    const long long byte = b.pos >> 3;
    for (int i = 0; i < 5; ++i) { long long k = byte + i; if (k < b.n) v |= (unsigned long long)b.p[k] << (8 * i); }
    return (unsigned)((v >> (b.pos & 7)) & ((1ull << nb) - 1));
}

// reads an FSE table description; returns bytes consumed or -1
__device__ int z_read_ncount(const unsigned char *src, int n, short *norm, int *nsym_io, int *log_out, int maxlog)
{
    if (n < 1) return -1;
    ZFwdBits b; b.p = src; b.n = n; b.pos = 0;
    const int log = (int)zf_peek(b, 4) + 5; b.pos += 4;
    if (log > maxlog) return -1;
    int remaining = (1 << log) + 1, threshold = 1 << log, nbits = log + 1, sym = 0;
    const int maxsym = *nsym_io;
    bool prev0 = false;
    while (remaining > 1 && sym < maxsym) {
        if (prev0) {
            for (;;) {
                const unsigned r = zf_peek(b, 2); b.pos += 2;
                for (unsigned k = 0; k < r && sym < maxsym; ++k) norm[sym++] = 0;
                if (r != 3) break;
            }
            prev0 = false;
            if (sym >= maxsym) break;
        }
        const int maxv = (2 * threshold - 1) - remaining;
        int count;
        const unsigned lowbits = zf_peek(b, nbits - 1);
        if ((int)lowbits < maxv) { count = (int)lowbits; b.pos += nbits - 1; }
        else {
            count = (int)zf_peek(b, nbits);
            if (count >= threshold) count -= maxv;
            b.pos += nbits;
        }
        count--;
        remaining -= count < 0 ? -count : count;
        norm[sym++] = (short)count;
        prev0 = count == 0;
        while (remaining < threshold && threshold > 1) { nbits--; threshold >>= 1; }
        if ((b.pos >> 3) > n) return -1;
    }
    if (remaining != 1) return -1;
    *nsym_io = sym; *log_out = log;
    return (int)((b.pos + 7) >> 3);
}

// ---- backward bit reader (FSE / Huffman streams) -----------------------------------------------------------
struct ZBackBits { const unsigned char *p; long long bitpos; int err; };   // bitpos = number of unread bits below the marker
__device__ int zb_init(ZBackBits &b, const unsigned char *p, int n)
{
    b.p = p; b.err = 0; b.bitpos = 0;
    if (n <= 0 || p[n - 1] == 0) { b.err = 1; return -1; }
    b.bitpos = (long long)(n - 1) * 8 + z_highbit(p[n - 1]);
    return 0;
}
// read nb bits (nb <= 24) just below the current position; bits beyond the start of the stream read as 0
__device__ __forceinline__ unsigned zb_read(ZBackBits &b, int nb)
{
    if (nb == 0) return 0;
    b.bitpos -= nb;
    long long start = b.bitpos;
    unsigned long long v = 0;
    int shift = 0;
    if (start < 0) { shift = (int)(-start); start = 0; if (shift >= nb) return 0; }
    const long long byte = start >> 3;
    for (int i = 0; i < 5; ++i) v |= (unsigned long long)b.p[byte + i < 0 ? 0 : byte + i] << (8 * i);   // caller guarantees 4 readable bytes of slack
    v >>= (start & 7);
    return (unsigned)((v & ((1ull << (nb - shift)) - 1)) << shift);
}

// ---- Huffman -------------------------------------------------------------------------------------------
struct ZHufDec { unsigned char sym[1 << Z_HUF_MAXLOG], nb[1 << Z_HUF_MAXLOG]; int log; int valid; };

__device__ int z_huf_build_dec(ZHufDec *h, const unsigned char *weights, int nw)
{
    // last weight is implicit
    unsigned total = 0;
    for (int i = 0; i < nw; ++i) { if (weights[i] > Z_HUF_MAXLOG) return -1; if (weights[i]) total += 1u << (weights[i] - 1); }
    if (total == 0) return -1;
    const int log = z_highbit(total) + 1;
    if (log > Z_HUF_MAXLOG) return -1;
    const unsigned rest = (1u << log) - total;
    if (rest == 0 || (rest & (rest - 1))) return -1;
    unsigned char w[256];
    for (int i = 0; i < nw; ++i) w[i] = weights[i];
    w[nw] = (unsigned char)(z_highbit(rest) + 1);
    const int n = nw + 1;
    // canonical order: ascending weight, then symbol; codes fill the table from index 0 with the longest codes
    unsigned pos = 0;
    for (int wt = 1; wt <= log; ++wt) {
        const unsigned span = 1u << (wt - 1);
        for (int s = 0; s < n; ++s)
            if (w[s] == wt) {
                for (unsigned k = 0; k < span; ++k) { h->sym[pos + k] = (unsigned char)s; h->nb[pos + k] = (unsigned char)(log + 1 - wt); }
                pos += span;
            }
    }
    if (pos != (1u << log)) return -1;
    h->log = log; h->valid = 1;
    return 0;
}

// one Huffman stream of nsym symbols
__device__ int z_huf_dec_stream(const ZHufDec *h, const unsigned char *src, int n, unsigned char *dst, int nsym)
{
    ZBackBits b;
    if (zb_init(b, src, n)) return -1;
    const int log = h->log;
    for (int i = 0; i < nsym; ++i) {
        // peek log bits
        long long save = b.bitpos;
        const unsigned idx = zb_read(b, log);
        b.bitpos = save - h->nb[idx];
        dst[i] = h->sym[idx];
    }
    return b.bitpos == 0 ? 0 : (b.bitpos < 0 ? -1 : -1);
}

// Huffman tree description -> decoding table; returns bytes consumed or -1
__device__ int z_huf_read_tree(ZHufDec *h, const unsigned char *src, int n, ZFseEntry *scratch)
{
    if (n < 1) return -1;
    unsigned char weights[256];
    int nw = 0, used;
    const int hb = src[0];
    if (hb >= 128) {
        nw = hb - 127;
        used = 1 + (nw + 1) / 2;
        if (used > n) return -1;
        for (int i = 0; i < nw; ++i) weights[i] = (i & 1) ? (src[1 + i / 2] & 15) : (src[1 + i / 2] >> 4);
    } else {
        used = 1 + hb;
        if (used > n || hb < 1) return -1;
        short norm[16]; int nsym = 13, log;
        const int hl = z_read_ncount(src + 1, hb, norm, &nsym, &log, 6);
        if (hl < 0) return -1;
        z_fse_build(scratch, log, norm, nsym);
        ZBackBits b;
        if (zb_init(b, src + 1 + hl, hb - hl)) return -1;
        unsigned s1 = zb_read(b, log), s2 = zb_read(b, log);
        for (;;) {
            if (nw >= 254) return -1;
            weights[nw++] = scratch[s1].sym;
            { const int nb = scratch[s1].nbits; if (b.bitpos < nb) { weights[nw++] = scratch[s2].sym; break; } s1 = scratch[s1].base + zb_read(b, nb); }
            if (nw >= 254) return -1;
            weights[nw++] = scratch[s2].sym;
            { const int nb = scratch[s2].nbits; if (b.bitpos < nb) { weights[nw++] = scratch[s1].sym; break; } s2 = scratch[s2].base + zb_read(b, nb); }
        }
    }
    if (z_huf_build_dec(h, weights, nw)) return -1;
    return used;
}

// ---- decompressor ---------------------------------------------------------------------------------------
struct ZDecWork {
    ZHufDec huf;
    ZFseEntry ll[512], of[256], ml[512], tmp[64];
    int ll_log, of_log, ml_log, tables_valid;
    unsigned rep[3];
    unsigned char lit[Z_BLOCK_MAX + 32];
};

__device__ int z_seq_table(int mode, const unsigned char *src, int n, ZFseEntry *tab, int *log, const short *defnorm, int defsyms, int deflog,
                           int maxsyms, int maxlog, int have_prev)
{
    if (mode == 0) { z_fse_build(tab, deflog, defnorm, defsyms); *log = deflog; return 0; }
    if (mode == 1) { if (n < 1) return -1; tab[0].sym = src[0]; tab[0].nbits = 0; tab[0].base = 0; *log = 0; return 1; }
    if (mode == 2) {
        short norm[64]; int nsym = maxsyms, lg;
        const int used = z_read_ncount(src, n, norm, &nsym, &lg, maxlog);
        if (used < 0) return -1;
        z_fse_build(tab, lg, norm, nsym); *log = lg;
        return used;
    }
    return have_prev ? 0 : -1;
}

__device__ int z_decode_block(ZDecWork *w, const unsigned char *src, int n, unsigned char *dst, int dst_pos, int dst_cap, int *produced)
{
    if (n < 1) return -1;
    // ---- literals section ----
    const int ltype = src[0] & 3, sf = (src[0] >> 2) & 3;
    int pos = 0, regen = 0, csize = 0, streams = 1;
    const unsigned char *lit = w->lit;
    if (ltype < 2) {
        if (sf == 0 || sf == 2) { regen = src[0] >> 3; pos = 1; }
        else if (sf == 1) { if (n < 2) return -1; regen = (src[0] >> 4) | (src[1] << 4); pos = 2; }
        else { if (n < 3) return -1; regen = (src[0] >> 4) | (src[1] << 4) | (src[2] << 12); pos = 3; }
        if (regen > Z_BLOCK_MAX) return -1;
        if (ltype == 0) { if (pos + regen > n) return -1; lit = src + pos; pos += regen; }
        else { if (pos + 1 > n) return -1; for (int i = 0; i < regen; ++i) w->lit[i] = src[pos]; pos += 1; }
    } else {
        if (n < 5) return -1;
        const unsigned long long hv = (unsigned long long)src[0] | ((unsigned long long)src[1] << 8) | ((unsigned long long)src[2] << 16) |
                                      ((unsigned long long)src[3] << 24) | ((unsigned long long)src[4] << 32);
        if (sf == 0 || sf == 1) { regen = (int)((hv >> 4) & 0x3ff); csize = (int)((hv >> 14) & 0x3ff); pos = 3; streams = sf == 0 ? 1 : 4; }
        else if (sf == 2) { regen = (int)((hv >> 4) & 0x3fff); csize = (int)((hv >> 18) & 0x3fff); pos = 4; streams = 4; }
        else { regen = (int)((hv >> 4) & 0x3ffff); csize = (int)((hv >> 22) & 0x3ffff); pos = 5; streams = 4; }
        if (regen > Z_BLOCK_MAX || pos + csize > n) return -1;
        const unsigned char *p = src + pos;
        int left = csize;
        if (ltype == 2) {
            const int used = z_huf_read_tree(&w->huf, p, left, w->tmp);
            if (used < 0) return -1;
            p += used; left -= used;
        } else if (!w->huf.valid) return -1;
        if (streams == 1) { if (z_huf_dec_stream(&w->huf, p, left, w->lit, regen)) return -1; }
        else {
            if (left < 6) return -1;
            const int s1 = p[0] | (p[1] << 8), s2 = p[2] | (p[3] << 8), s3 = p[4] | (p[5] << 8);
            const int s4 = left - 6 - s1 - s2 - s3, seg = (regen + 3) / 4;
            if (s4 < 0 || regen < 3 * seg) return -1;
            p += 6;
            if (z_huf_dec_stream(&w->huf, p, s1, w->lit, seg)) return -1;
            if (z_huf_dec_stream(&w->huf, p + s1, s2, w->lit + seg, seg)) return -1;
            if (z_huf_dec_stream(&w->huf, p + s1 + s2, s3, w->lit + 2 * seg, seg)) return -1;
            if (z_huf_dec_stream(&w->huf, p + s1 + s2 + s3, s4, w->lit + 3 * seg, regen - 3 * seg)) return -1;
        }
        pos += csize;
    }
    // ---- sequences section ----
    if (pos >= n) return -1;
    int nseq = src[pos++];
    if (nseq >= 128) {
        if (nseq == 255) { if (pos + 2 > n) return -1; nseq = src[pos] + (src[pos + 1] << 8) + 0x7f00; pos += 2; }
        else { if (pos + 1 > n) return -1; nseq = ((nseq - 128) << 8) + src[pos]; pos += 1; }
    }
    int out = dst_pos, lp = 0;
    if (nseq > 0) {
        if (pos >= n) return -1;
        const int modes = src[pos++];
        int used = z_seq_table((modes >> 6) & 3, src + pos, n - pos, w->ll, &w->ll_log, z_ll_default, 36, 6, 36, 9, w->tables_valid);
        if (used < 0) return -1; pos += used;
        used = z_seq_table((modes >> 4) & 3, src + pos, n - pos, w->of, &w->of_log, z_of_default, 29, 5, 32, 8, w->tables_valid);
        if (used < 0) return -1; pos += used;
        used = z_seq_table((modes >> 2) & 3, src + pos, n - pos, w->ml, &w->ml_log, z_ml_default, 53, 6, 53, 9, w->tables_valid);
        if (used < 0) return -1; pos += used;
        w->tables_valid = 1;
        ZBackBits b;
        if (zb_init(b, src + pos, n - pos)) return -1;
        unsigned sl = zb_read(b, w->ll_log), so = zb_read(b, w->of_log), sm = zb_read(b, w->ml_log);
        for (int i = 0; i < nseq; ++i) {
            const int oc = w->of[so].sym, mc = w->ml[sm].sym, lc = w->ll[sl].sym;
            if (oc > 31 || mc > 52 || lc > 35) return -1;
            unsigned ov = (1u << oc) + (oc > 24 ? (zb_read(b, oc - 24) << 24) + zb_read(b, 24) : zb_read(b, oc));
            const unsigned mlen = z_ml_base[mc] + (z_ml_bits[mc] ? zb_read(b, z_ml_bits[mc]) : 0u);
            const unsigned llen = z_ll_base[lc] + (z_ll_bits[lc] ? zb_read(b, z_ll_bits[lc]) : 0u);
            unsigned offset;
            if (ov > 3) { offset = ov - 3; w->rep[2] = w->rep[1]; w->rep[1] = w->rep[0]; w->rep[0] = offset; }
            else {
                unsigned idx = ov - 1 + (llen == 0 ? 1u : 0u);
                if (idx == 0) offset = w->rep[0];
                else {
                    offset = idx < 3 ? w->rep[idx] : w->rep[0] - 1;
                    if (offset == 0) return -1;
                    if (idx > 1) w->rep[2] = w->rep[1];
                    w->rep[1] = w->rep[0]; w->rep[0] = offset;
                }
            }
            if (i + 1 < nseq) {
                sl = w->ll[sl].base + zb_read(b, w->ll[sl].nbits);
                sm = w->ml[sm].base + zb_read(b, w->ml[sm].nbits);
                so = w->of[so].base + zb_read(b, w->of[so].nbits);
            }
            if (b.bitpos < 0) return -1;
            if (lp + (int)llen > regen || out + (long long)llen + mlen > dst_cap || offset > (unsigned)(out + llen)) return -1;
            for (unsigned k = 0; k < llen; ++k) dst[out++] = lit[lp++];
            for (unsigned k = 0; k < mlen; ++k) { dst[out] = dst[out - offset]; out++; }
        }
    }
    if (out + (regen - lp) > dst_cap) return -1;
    while (lp < regen) dst[out++] = lit[lp++];
    *produced = out - dst_pos;
    return 0;
}

// returns decompressed size or -1
__device__ int z_decompress(ZDecWork *w, const unsigned char *src, int n, unsigned char *dst, int dst_cap)
{
    if (n < 6) return -1;
    const unsigned magic = src[0] | (src[1] << 8) | (src[2] << 16) | ((unsigned)src[3] << 24);
    if (magic != Z_MAGIC) return -1;
    const int fhd = src[4], fcs_flag = fhd >> 6, single = (fhd >> 5) & 1, did = fhd & 3, checksum = (fhd >> 2) & 1;
    int pos = 5;
    if (!single) pos += 1;
    pos += did == 0 ? 0 : (did == 1 ? 1 : (did == 2 ? 2 : 4));
    pos += fcs_flag == 0 ? (single ? 1 : 0) : (fcs_flag == 1 ? 2 : (fcs_flag == 2 ? 4 : 8));
    w->huf.valid = 0; w->tables_valid = 0; w->rep[0] = 1; w->rep[1] = 4; w->rep[2] = 8;
    int out = 0;
    for (;;) {
        if (pos + 3 > n) return -1;
        const unsigned bh = src[pos] | (src[pos + 1] << 8) | (src[pos + 2] << 16);
        pos += 3;
        const int last = bh & 1, type = (bh >> 1) & 3, bsize = (int)(bh >> 3);
        if (type == 0) { if (pos + bsize > n || out + bsize > dst_cap) return -1; for (int i = 0; i < bsize; ++i) dst[out++] = src[pos + i]; pos += bsize; }
        else if (type == 1) { if (pos + 1 > n || out + bsize > dst_cap) return -1; for (int i = 0; i < bsize; ++i) dst[out++] = src[pos]; pos += 1; }
        else if (type == 2) {
            if (pos + bsize > n) return -1;
            int produced = 0;
            if (z_decode_block(w, src + pos, bsize, dst, out, dst_cap, &produced)) return -1;
            out += produced; pos += bsize;
        } else return -1;
        if (last) break;
    }
    (void)checksum;
    return out;
}

// grid = nframes, block = 32.  in: [frame][stride] zstd frames (the buffer must keep >= 8 readable bytes after each
// frame), in_len / out_cap per frame; out: [frame][out_stride].  out_len[frame] = bytes produced, or 0xffffffff.
__global__ void k_zstd_decompress(const unsigned char *__restrict__ in, size_t stride, const unsigned *__restrict__ in_len,
                                  unsigned char *out, size_t out_stride, const unsigned *__restrict__ out_cap, unsigned *out_len,
                                  ZDecWork *work)
{
    const int frame = blockIdx.x;
    if (threadIdx.x != 0) return;
    if (in_len[frame] == 0) { out_len[frame] = 0; return; }
    const int r = z_decompress(work + frame, in + (size_t)frame * stride, (int)in_len[frame], out + (size_t)frame * out_stride, (int)out_cap[frame]);
    out_len[frame] = r < 0 ? 0xffffffffu : (unsigned)r;
}

// ---- compressor -------------------------------------------------------------------------------------------
struct ZBitW { unsigned char *p; int cap; unsigned long long acc; int nacc; int pos; int err; };
__device__ __forceinline__ void zw_init(ZBitW &b, unsigned char *p, int cap) { b.p = p; b.cap = cap; b.acc = 0; b.nacc = 0; b.pos = 0; b.err = 0; }
__device__ __forceinline__ void zw_add(ZBitW &b, unsigned v, int nb)
{
    if (nb == 0) return;
    b.acc |= (unsigned long long)(v & ((nb >= 32) ? 0xffffffffu : ((1u << nb) - 1u))) << b.nacc;
    b.nacc += nb;
    while (b.nacc >= 8) { if (b.pos < b.cap) b.p[b.pos++] = (unsigned char)b.acc; else b.err = 1; b.acc >>= 8; b.nacc -= 8; }
}
// close with the end marker bit; returns size in bytes
__device__ __forceinline__ int zw_close(ZBitW &b)
{
    zw_add(b, 1, 1);
    if (b.nacc > 0) { if (b.pos < b.cap) b.p[b.pos++] = (unsigned char)b.acc; else b.err = 1; b.acc = 0; b.nacc = 0; }
    return b.err ? -1 : b.pos;
}

// Working set of one block of the compressor, in shared memory (one warp per block).
struct ZEncWork {
    unsigned char lit[Z_CHUNK + 16];     // the chunk, compacted in place to its literals
    unsigned short sll[Z_CHUNK / 5 + 8]; // sequences: literal run length, match length (offset: always repeat offset 1)
    unsigned short sml[Z_CHUNK / 5 + 8];
    ZFseEntry ll[64], ml[64], wt[64];
    unsigned hist[256];
    unsigned key[256];                   // Huffman build: weight << 8 | symbol
    unsigned short order[256];           // symbols by ascending weight
    unsigned nwt[512];                   // node weights (leaves in sorted order, then internal nodes)
    short parent[512];
    unsigned char depth[512];
    unsigned char nbits[256];
    unsigned short code[256];
    unsigned char weights[256];
    unsigned char tmp[Z_CHUNK + 1024];   // tree description and the Huffman streams
    unsigned char out[Z_CHUNK + 32];     // the block being built (header + body)
    int v[16];                           // scalars handed from one phase to the next
};

// tANS encode step through the decoding table: find the entry of symbol s whose state interval contains x
__device__ __forceinline__ unsigned z_fse_enc(ZBitW &b, const ZFseEntry *tab, int size, unsigned x, int s)
{
    for (int u = 0; u < size; ++u)
        if (tab[u].sym == s && x >= tab[u].base && x < (unsigned)tab[u].base + (1u << tab[u].nbits)) { zw_add(b, x - tab[u].base, tab[u].nbits); return (unsigned)u; }
    b.err = 1;
    return 0;
}
// a state of symbol s to start from (= the state the decoder ends in); the entry with the most update bits, so
// that an FSE stream without a symbol count (Huffman weights) terminates by running out of bits
__device__ __forceinline__ unsigned z_fse_first(const ZFseEntry *tab, int size, int s)
{
    int best = -1;
    for (int u = 0; u < size; ++u) if (tab[u].sym == s && (best < 0 || tab[u].nbits > tab[best].nbits)) best = u;
    return best < 0 ? 0u : (unsigned)best;
}

__device__ __forceinline__ int z_ll_code(unsigned v)
{
    if (v < 16) return (int)v;
    int c = 35;
    while (z_ll_base[c] > v) --c;
    return c;
}
__device__ __forceinline__ int z_ml_code(unsigned v)
{
    if (v < 35) return (int)v - 3;
    int c = 52;
    while (z_ml_base[c] > v) --c;
    return c;
}

// Huffman code lengths (<= Z_HUF_MAXLOG) for w->hist[0..255] into w->nbits; returns number of used symbols.  Called by
// the whole warp.  Symbols are rank-sorted by weight in parallel, lane 0 merges with the two-queue method.  Length
// limiting by flattening the histogram and rebuilding keeps the code complete (Kraft sum exactly 1), which the
// implicit last weight of the tree description relies on.
__device__ int z_huf_lengths(ZEncWork *w, int lane)
{
    int used = 0;
    for (int s = lane; s < 256; s += 32) { w->nbits[s] = 0; if (w->hist[s]) used++; }
    for (int o = 16; o > 0; o >>= 1) used += __shfl_xor_sync(0xffffffffu, used, o);
    if (used < 2) return used;
    for (int shift = 0; shift < 32; ++shift) {
        for (int s = lane; s < 256; s += 32) {
            const unsigned v = w->hist[s] >> shift;
            w->key[s] = w->hist[s] ? ((v ? v : 1u) << 8) | (unsigned)s : 0xffffffffu;
        }
        __syncwarp();
        for (int s = lane; s < 256; s += 32) {
            const unsigned k = w->key[s];
            if (k == 0xffffffffu) continue;
            int r = 0;
            for (int t = 0; t < 256; ++t) r += w->key[t] < k;
            w->order[r] = (unsigned short)s; w->nwt[r] = k >> 8;
        }
        __syncwarp();
        if (lane == 0) {
            // leaves 0..used-1 ascending, internal nodes used..2*used-2 are created in ascending weight order
            int lf = 0, in = used, n = used;
            for (int it = 0; it < used - 1; ++it) {
                int pick[2];
                for (int k = 0; k < 2; ++k) {
                    if (lf < used && (in >= n || w->nwt[lf] <= w->nwt[in])) pick[k] = lf++; else pick[k] = in++;
                }
                w->nwt[n] = w->nwt[pick[0]] + w->nwt[pick[1]];
                w->parent[pick[0]] = w->parent[pick[1]] = (short)n;
                n++;
            }
            w->depth[n - 1] = 0;
            int maxlen = 0;
            for (int i = n - 2; i >= 0; --i) {
                const int d = w->depth[w->parent[i]] + 1;
                w->depth[i] = (unsigned char)d;
                if (i < used) { w->nbits[w->order[i]] = (unsigned char)d; if (d > maxlen) maxlen = d; }
            }
            w->v[15] = maxlen;
        }
        __syncwarp();
        if (w->v[15] <= Z_HUF_MAXLOG) break;
        __syncwarp();
    }
    return used;
}

// writes one Huffman stream (symbols encoded last-to-first); returns bytes or -1
__device__ int z_huf_enc_stream(unsigned char *dst, int cap, const unsigned char *src, int n, const unsigned short *code, const unsigned char *nbits)
{
    ZBitW b; zw_init(b, dst, cap);
    for (int i = n - 1; i >= 0; --i) zw_add(b, code[src[i]], nbits[src[i]]);
    return zw_close(b);
}

// Huffman tree description with FSE-compressed weights; returns bytes or -1.  weights[0..nw) (last symbol implicit)
__device__ int z_huf_write_tree(unsigned char *dst, int cap, const unsigned char *weights, int nw, ZFseEntry *tab)
{
    if (nw < 2) return -1;
    short norm[13]; unsigned cnt[13];
    for (int i = 0; i < 13; ++i) cnt[i] = 0;
    int maxw = 0;
    for (int i = 0; i < nw; ++i) { cnt[weights[i]]++; if (weights[i] > maxw) maxw = weights[i]; }
    const int nsym = maxw + 1;
    int log = 6;
    while (log > 5 && (1 << (log - 1)) >= nw) log--;        // small tables for few weights
    // normalise: proportional, every present symbol >= 1, remainder to the largest
    const int size = 1 << log;
    int sum = 0, big = 0;
    for (int s = 0; s < nsym; ++s) {
        if (!cnt[s]) { norm[s] = 0; continue; }
        int v = (int)(((unsigned long long)cnt[s] * size) / nw);
        if (v < 1) v = 1;
        norm[s] = (short)v; sum += v;
        if (cnt[s] > cnt[big] || !cnt[big]) big = s;
    }
    while (sum != size) {
        if (sum < size) { norm[big] += (short)(size - sum); sum = size; }
        else {
            // take from the symbols with the largest normalised counts
            int m = 0;
            for (int s = 0; s < nsym; ++s) if (norm[s] > norm[m]) m = s;
            if (norm[m] <= 1) return -1;
            norm[m]--; sum--;
        }
    }
    // ---- table description (FSE_writeNCount) ----
    ZBitW b; zw_init(b, dst + 1, cap - 1);
    zw_add(b, (unsigned)(log - 5), 4);
    int remaining = size + 1, threshold = size, nbb = log + 1;
    bool prev0 = false;
    int s = 0;
    while (remaining > 1 && s < nsym) {
        if (prev0) {
            int start = s;
            while (s < nsym && norm[s] == 0) s++;
            int zeros = s - start;
            while (zeros >= 3) { zw_add(b, 3, 2); zeros -= 3; }
            zw_add(b, (unsigned)zeros, 2);
            prev0 = false;
            if (s >= nsym) break;
        }
        int count = norm[s++];
        const int maxv = (2 * threshold - 1) - remaining;
        remaining -= count < 0 ? -count : count;
        count++;
        if (count >= threshold) count += maxv;
        zw_add(b, (unsigned)count, nbb - (count < maxv ? 1 : 0));
        prev0 = count == 1;
        while (remaining < threshold && threshold > 1) { nbb--; threshold >>= 1; }
    }
    if (remaining != 1) return -1;
    if (b.nacc > 0) { zw_add(b, 0, 8 - b.nacc); }
    const int hl = b.pos;
    // ---- FSE stream of the weights, two interleaved states, encoded back to front ----
    z_fse_build(tab, log, norm, nsym);
    ZBitW f; zw_init(f, dst + 1 + hl, cap - 1 - hl);
    // decoder order: s1 emits w[0], s2 emits w[1], s1 w[2], ...; the final states carry the last two weights
    int i = nw - 1;
    unsigned st[2];
    // the weight with index i is produced by state (i & 1)
    st[i & 1] = z_fse_first(tab, size, weights[i]); if (tab[st[i & 1]].nbits == 0) return -1; i--;
    st[i & 1] = z_fse_first(tab, size, weights[i]); if (tab[st[i & 1]].nbits == 0) return -1; i--;
    for (; i >= 0; --i) st[i & 1] = z_fse_enc(f, tab, size, st[i & 1], weights[i]);
    zw_add(f, st[1], log);
    zw_add(f, st[0], log);
    const int fl = zw_close(f);
    if (fl < 0 || b.err) return -1;
    const int total = hl + fl;
    if (total >= 128) return -1;
    dst[0] = (unsigned char)total;
    return 1 + total;
}

// Literals section of the block into dst (called by the whole warp); returns bytes written or -1.
__device__ int z_write_literals(ZEncWork *w, unsigned char *dst, int cap, int nlit, int lane)
{
    const unsigned char *lit = w->lit;
    const int best_raw = nlit + (nlit < 32 ? 1 : (nlit < 4096 ? 2 : 3));
    int written = -1;
    if (nlit >= 64) {
        for (int i = lane; i < 256; i += 32) w->hist[i] = 0;
        __syncwarp();
        for (int i = lane; i < nlit; i += 32) atomicAdd(&w->hist[lit[i]], 1u);
        __syncwarp();
        const int used = z_huf_lengths(w, lane);
        unsigned char *t = w->tmp;
        if (lane == 0) {
            w->v[0] = 0;                                     // tree length, 0 = no Huffman coding
            int maxsym = 255;
            while (maxsym > 0 && !w->hist[maxsym]) maxsym--;
            if (used >= 2) {
                int maxlen = 0;
                unsigned long long bits = 0;
                for (int s = 0; s < 256; ++s) { if (w->nbits[s] > maxlen) maxlen = w->nbits[s]; bits += (unsigned long long)w->hist[s] * w->nbits[s]; }
                // worth trying only if the streams alone already beat raw
                if ((long long)(bits >> 3) + 16 < nlit) {
                    // canonical codes (HUF_buildCTable): longest codes first from 0, within a length ascending symbol order
                    unsigned nper[Z_HUF_MAXLOG + 2], val[Z_HUF_MAXLOG + 2];
                    for (int k = 0; k <= Z_HUF_MAXLOG + 1; ++k) nper[k] = 0;
                    for (int s = 0; s <= maxsym; ++s) nper[w->nbits[s]]++;
                    unsigned mn = 0;
                    for (int k = maxlen; k >= 1; --k) { val[k] = mn; mn += nper[k]; mn >>= 1; }
                    for (int s = 0; s <= maxsym; ++s) if (w->nbits[s]) w->code[s] = (unsigned short)val[w->nbits[s]]++;
                    for (int s = 0; s <= maxsym; ++s) w->weights[s] = w->nbits[s] ? (unsigned char)(maxlen + 1 - w->nbits[s]) : 0;
                    const int tl = z_huf_write_tree(t, 512, w->weights, maxsym, w->wt);
                    if (tl > 0) w->v[0] = tl;
                }
            }
        }
        __syncwarp();
        const int tl = w->v[0];
        if (tl > 0) {
            // streams 0..3 (or one) are coded by one lane each into their own quarter of tmp
            const int streams = nlit < 1024 ? 1 : 4;
            const int seg = (nlit + 3) / 4, scap = (Z_CHUNK + 1024 - 512) / 4;
            if (lane < streams) {
                const int a = streams == 1 ? 0 : lane * seg, len = streams == 1 ? nlit : (lane < 3 ? seg : nlit - 3 * seg);
                w->v[1 + lane] = z_huf_enc_stream(t + 512 + lane * scap, streams == 1 ? 4 * scap : scap, lit + a, len, w->code, w->nbits);
            }
            __syncwarp();
            int ok = 1, csize = tl + (streams == 4 ? 6 : 0);
            for (int k = 0; k < streams; ++k) { const int sl = w->v[1 + k]; if (sl < 0 || sl > 65535) ok = 0; else csize += sl; }
            if (ok) {
                int hdr; unsigned long long hv;
                if (streams == 1) { hdr = 3; hv = 2ull | (0ull << 2) | ((unsigned long long)nlit << 4) | ((unsigned long long)csize << 14); }
                else if (nlit < 1024 && csize < 1024) { hdr = 3; hv = 2ull | (1ull << 2) | ((unsigned long long)nlit << 4) | ((unsigned long long)csize << 14); }
                else if (nlit < 16384 && csize < 16384) { hdr = 4; hv = 2ull | (2ull << 2) | ((unsigned long long)nlit << 4) | ((unsigned long long)csize << 18); }
                else { hdr = 5; hv = 2ull | (3ull << 2) | ((unsigned long long)nlit << 4) | ((unsigned long long)csize << 22); }
                if (hdr + csize < best_raw && hdr + csize <= cap) {
                    int pos = 0;
                    if (lane == 0) for (int k = 0; k < hdr; ++k) dst[k] = (unsigned char)(hv >> (8 * k));
                    pos = hdr;
                    for (int k = lane; k < tl; k += 32) dst[pos + k] = t[k];
                    pos += tl;
                    if (streams == 4) {
                        if (lane < 3) { const int sl = w->v[1 + lane]; dst[pos + 2 * lane] = (unsigned char)sl; dst[pos + 2 * lane + 1] = (unsigned char)(sl >> 8); }
                        pos += 6;
                    }
                    for (int k = 0; k < streams; ++k) {
                        const int sl = w->v[1 + k];
                        const unsigned char *sp = t + 512 + k * scap;
                        for (int i = lane; i < sl; i += 32) dst[pos + i] = sp[i];
                        pos += sl;
                    }
                    written = pos;
                }
            }
        }
    }
    if (written < 0) {
        // raw literals
        if (best_raw > cap) return -1;
        int pos;
        if (nlit < 32) { if (lane == 0) dst[0] = (unsigned char)(nlit << 3); pos = 1; }
        else if (nlit < 4096) { if (lane == 0) { dst[0] = (unsigned char)((1 << 2) | ((nlit & 15) << 4)); dst[1] = (unsigned char)(nlit >> 4); } pos = 2; }
        else { if (lane == 0) { dst[0] = (unsigned char)((3 << 2) | ((nlit & 15) << 4)); dst[1] = (unsigned char)(nlit >> 4); dst[2] = (unsigned char)(nlit >> 12); } pos = 3; }
        for (int i = lane; i < nlit; i += 32) dst[pos + i] = lit[i];
        written = pos + nlit;
    }
    __syncwarp();
    return written;
}

// one compressed block from w->lit[0..n) (called by the whole warp); returns bytes written or -1 if the block
// should be stored raw
__device__ int z_compress_block(ZEncWork *w, int n, unsigned char *dst, int cap, int lane)
{
    // ---- does the chunk hold a run of >= 5 equal bytes at all?  (high-entropy chunks do not) ----
    int hit = 0;
    for (int i = lane; i + 4 < n; i += 32) {
        const unsigned char c = w->lit[i];
        if (w->lit[i + 1] == c && w->lit[i + 2] == c && w->lit[i + 3] == c && w->lit[i + 4] == c) { hit = 1; break; }
    }
    hit = __any_sync(0xffffffffu, hit);
    // ---- parse: runs of >= 5 equal bytes become (1 literal, match of run-1 at offset 1) ----
    if (lane == 0) {
        int nlit = n, nseq = 0;
        if (hit) {
            unsigned char *src = w->lit;
            int i = 0, pending = 0;
            nlit = 0;
            while (i < n) {
                const unsigned char c = src[i];
                int j = i + 1;
                while (j < n && src[j] == c) j++;
                const int run = j - i;
                if (run >= 5) {
                    src[nlit++] = c;
                    w->sll[nseq] = (unsigned short)(pending + 1); w->sml[nseq] = (unsigned short)(run - 1); nseq++;
                    pending = 0;
                } else {
                    for (int k = 0; k < run; ++k) src[nlit++] = c;
                    pending += run;
                }
                i = j;
            }
        }
        w->v[8] = nlit; w->v[9] = nseq;
    }
    __syncwarp();
    const int nlit = w->v[8], nseq = w->v[9];
    int pos = z_write_literals(w, dst, cap, nlit, lane);
    if (pos < 0) return -1;
    // ---- sequences ----
    if (lane == 0) {
        int r = -1;
        do {
            if (pos + 4 > cap) break;
            if (nseq == 0) { dst[pos++] = 0; r = pos; break; }
            if (nseq < 128) dst[pos++] = (unsigned char)nseq;
            else if (nseq < 0x7f00) { dst[pos++] = (unsigned char)((nseq >> 8) + 128); dst[pos++] = (unsigned char)nseq; }
            else { dst[pos++] = 255; dst[pos++] = (unsigned char)(nseq - 0x7f00); dst[pos++] = (unsigned char)((nseq - 0x7f00) >> 8); }
            dst[pos++] = (unsigned char)((0 << 6) | (1 << 4) | (0 << 2));     // LL predefined, OF RLE, ML predefined
            dst[pos++] = 0;                                                     // the single offset code: 0 = repeat offset 1
            z_fse_build(w->ll, 6, z_ll_default, 36);
            z_fse_build(w->ml, 6, z_ml_default, 53);
            ZBitW b; zw_init(b, dst + pos, cap - pos);
            unsigned sl, sm;
            {
                const unsigned ql = w->sll[nseq - 1], qm = w->sml[nseq - 1];
                const int lc = z_ll_code(ql), mc = z_ml_code(qm);
                sm = z_fse_first(w->ml, 64, mc); sl = z_fse_first(w->ll, 64, lc);
                zw_add(b, ql - z_ll_base[lc], z_ll_bits[lc]);
                zw_add(b, qm - z_ml_base[mc], z_ml_bits[mc]);
            }
            for (int k = nseq - 2; k >= 0; --k) {
                const unsigned ql = w->sll[k], qm = w->sml[k];
                const int lc = z_ll_code(ql), mc = z_ml_code(qm);
                // (offset state: RLE table, no bits)
                sm = z_fse_enc(b, w->ml, 64, sm, mc);
                sl = z_fse_enc(b, w->ll, 64, sl, lc);
                zw_add(b, ql - z_ll_base[lc], z_ll_bits[lc]);
                zw_add(b, qm - z_ml_base[mc], z_ml_bits[mc]);
            }
            zw_add(b, sm, 6);
            zw_add(b, sl, 6);
            const int sl_bytes = zw_close(b);
            if (sl_bytes >= 0) r = pos + sl_bytes;
        } while (0);
        w->v[10] = r;
    }
    __syncwarp();
    return w->v[10];
}

// Compressor, stage 1: grid = (maxchunks, nframes), one warp per block, working set in shared memory.  Chunk j of
// frame i (Z_CHUNK bytes of in[i]) becomes one zstd block (3-byte header + body) in blk[i][j]; blen[i][j] = its size.
// Blocks of a frame are independent: every block carries its own Huffman table, sequences use predefined tables
// and only the repeat offset 1, which no block ever changes.
#define Z_BLK_STRIDE (Z_CHUNK + 16)
__global__ void __launch_bounds__(ZC_THREADS)
k_zstd_compress_chunks(const unsigned char *__restrict__ in, size_t in_stride, const unsigned *__restrict__ in_len,
                       unsigned char *blk, unsigned *blen, int maxchunks)
{
    EBCC_DYN_SMEM(ZEncWork, w);
    const int frame = blockIdx.y, j = blockIdx.x, lane = threadIdx.x;
    const int n = (int)in_len[frame], a = j * Z_CHUNK;
    unsigned *bl = blen + (size_t)frame * maxchunks + j;
    if (a >= n) { if (lane == 0) *bl = 0; return; }
    const unsigned char *src = in + (size_t)frame * in_stride + a;
    unsigned char *dst = blk + ((size_t)frame * maxchunks + j) * Z_BLK_STRIDE;
    const int len = n - a < Z_CHUNK ? n - a : Z_CHUNK, last = a + len >= n;
    for (int i = lane; i < (len + 3) / 4; i += 32) ((unsigned *)w->lit)[i] = ((const unsigned *)src)[i];
    __syncwarp();
    int cs = z_compress_block(w, len, w->out, len - 1, lane);
    unsigned bh;
    if (cs > 0 && cs < len) {
        bh = (unsigned)last | (2u << 1) | ((unsigned)cs << 3);
        for (int k = lane; k < cs; k += 32) dst[3 + k] = w->out[k];
    } else {
        cs = len; bh = (unsigned)last | (0u << 1) | ((unsigned)len << 3);
        for (int k = lane; k < len; k += 32) dst[3 + k] = src[k];
    }
    if (lane == 0) {
        dst[0] = (unsigned char)bh; dst[1] = (unsigned char)(bh >> 8); dst[2] = (unsigned char)(bh >> 16);
        *bl = (unsigned)(3 + cs);
    }
}

// stage 2: grid = nframes, block = 256: frame header + the blocks at their prefix-summed offsets
__global__ void k_zstd_assemble(const unsigned char *__restrict__ blk, const unsigned *__restrict__ blen, int maxchunks,
                                const unsigned *__restrict__ in_len, unsigned char *out, size_t out_stride, unsigned *out_len)
{
    __shared__ unsigned off[512];
    const int frame = blockIdx.x;
    const int n = (int)in_len[frame];
    if (n == 0) { if (threadIdx.x == 0) out_len[frame] = 0; return; }
    const int nch = (n + Z_CHUNK - 1) / Z_CHUNK;
    unsigned char *dst = out + (size_t)frame * out_stride;
    const unsigned *bl = blen + (size_t)frame * maxchunks;
    for (int base = 0; base < nch; base += 511) {
        const int cnt = nch - base < 511 ? nch - base : 511;
        __syncthreads();
        if (threadIdx.x == 0) {
            unsigned o = base == 0 ? 9u : off[511];
            if (base == 0) {
                dst[0] = 0x28; dst[1] = 0xB5; dst[2] = 0x2F; dst[3] = 0xFD;
                dst[4] = (unsigned char)((2 << 6) | (1 << 5));       // 4-byte content size, single segment, no checksum
                dst[5] = (unsigned char)n; dst[6] = (unsigned char)(n >> 8); dst[7] = (unsigned char)(n >> 16); dst[8] = (unsigned char)(n >> 24);
            }
            for (int k = 0; k < cnt; ++k) { off[k] = o; o += bl[base + k]; }
            off[511] = o;
        }
        __syncthreads();
        for (int k = 0; k < cnt; ++k) {
            const unsigned char *src = blk + ((size_t)frame * maxchunks + base + k) * Z_BLK_STRIDE;
            const unsigned len = bl[base + k], o = off[k];
            for (unsigned t = threadIdx.x; t < len; t += blockDim.x) dst[o + t] = src[t];
        }
        __syncthreads();
        if (threadIdx.x == 0 && base + cnt >= nch) out_len[frame] = off[511];
    }
}
