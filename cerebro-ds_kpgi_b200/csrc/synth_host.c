/*
 * synth_host.c — host-side writer of synthetic Greenplum-5 heap / TOAST relation files.
 *
 * Not part of the decode path.  It produces the INPUT of the Direct-Access reader: 32 KB
 * pages holding MADlib packed-table tuples and TOAST chunk tuples, plus pglz streams.
 * The format it emits is the one the reference reader parses
 * (/root/reference/cerebro_gpdb/pg_page_reader.py:253-269 page header, :285-305 ItemId,
 *  :272-276 tuple header, :385-422 TOAST walk, :191-231 pglz stream) and is described in
 * SURVEY.md Appendix A.  The encoder below is our own (the reference has no encoder); any
 * stream the reference decoder accepts is a valid stream.
 *
 * Built with plain gcc into libcds_synth.so (see cerebro-ds_kpgi_b200/build.py).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define CDS_BLCKSZ 32768u
#define CDS_PAGE_HDR 24u
#define CDS_TUPLE_HDR 24u /* 23-byte HeapTupleHeader padded to MAXALIGN → t_hoff */
#define CDS_TOAST_MAX_CHUNK 8140u

typedef struct cds_synth_rel {
    uint8_t *buf;       /* caller-owned, cap_pages * 32768 bytes */
    uint64_t cap_pages;
    uint64_t n_pages;   /* pages in use; the open page is n_pages-1 */
    uint32_t natts;     /* goes into t_infomask2 */
    uint32_t xmin;
} cds_synth_rel;

static void put16(uint8_t *p, uint32_t v) { p[0] = (uint8_t)v; p[1] = (uint8_t)(v >> 8); }
static void put32(uint8_t *p, uint32_t v) { p[0] = (uint8_t)v; p[1] = (uint8_t)(v >> 8); p[2] = (uint8_t)(v >> 16); p[3] = (uint8_t)(v >> 24); }
static void put32be(uint8_t *p, uint32_t v) { p[3] = (uint8_t)v; p[2] = (uint8_t)(v >> 8); p[1] = (uint8_t)(v >> 16); p[0] = (uint8_t)(v >> 24); }
static uint32_t get16(const uint8_t *p) { return (uint32_t)p[0] | ((uint32_t)p[1] << 8); }

static int new_page(cds_synth_rel *r)
{
    if (r->n_pages >= r->cap_pages) return -1;
    uint8_t *pg = r->buf + r->n_pages * (uint64_t)CDS_BLCKSZ;
    memset(pg, 0, CDS_BLCKSZ);
    /* pd_lsn(8)=0, pd_tli=1, pd_flags=0, pd_lower, pd_upper, pd_special, pd_pagesize_version, pd_prune_xid */
    put16(pg + 8, 1);
    put16(pg + 10, 0);
    put16(pg + 12, CDS_PAGE_HDR);
    put16(pg + 14, CDS_BLCKSZ & 0xFFFF); /* 32768 does not fit in 15 bits but fits in u16 */
    put16(pg + 16, CDS_BLCKSZ & 0xFFFF);
    put16(pg + 18, (CDS_BLCKSZ | 4u) & 0xFFFF);
    put32(pg + 20, 0);
    r->n_pages++;
    return 0;
}

/* Reserve room for a tuple of `len` bytes on the open page (PostgreSQL PageAddItem rule:
 * tuples grow downward from pd_special, each start MAXALIGNed; ItemIds grow upward).
 * Returns a pointer to the tuple bytes (zeroed) or NULL when the relation is full. */
static uint8_t *reserve_tuple(cds_synth_rel *r, uint32_t len, uint32_t *item_index)
{
    uint32_t need = (len + 7u) & ~7u;
    if (need + 4u > CDS_BLCKSZ - CDS_PAGE_HDR) return NULL;
    for (;;) {
        if (r->n_pages == 0 && new_page(r) != 0) return NULL;
        uint8_t *pg = r->buf + (r->n_pages - 1) * (uint64_t)CDS_BLCKSZ;
        uint32_t lower = get16(pg + 12);
        uint32_t upper = get16(pg + 14);
        if (upper == 0) upper = CDS_BLCKSZ; /* never written as 0, defensive */
        if (upper - lower >= need + 4u) {
            upper -= need;
            uint32_t idx = (lower - CDS_PAGE_HDR) / 4u;
            /* ItemId: lp_off bits 0-14, lp_flags bits 15-16 (1 = LP_NORMAL), lp_len bits 17-31 */
            put32(pg + lower, (upper & 0x7FFFu) | (1u << 15) | ((len & 0x7FFFu) << 17));
            put16(pg + 12, lower + 4u);
            put16(pg + 14, upper);
            if (item_index) *item_index = idx;
            return pg + upper;
        }
        if (new_page(r) != 0) return NULL;
    }
}

static void write_tuple_header(cds_synth_rel *r, uint8_t *t, uint32_t item_index)
{
    uint64_t blk = r->n_pages - 1;
    put32(t + 0, r->xmin);            /* t_xmin */
    put32(t + 4, 0);                  /* t_xmax */
    put32(t + 8, 0);                  /* t_cid  */
    put16(t + 12, (uint32_t)(blk >> 16)); /* bi_hi */
    put16(t + 14, (uint32_t)(blk & 0xFFFF)); /* bi_lo */
    put16(t + 16, item_index + 1);    /* ip_posid */
    put16(t + 18, r->natts);          /* t_infomask2 */
    put16(t + 20, 0x0902);            /* t_infomask: HASVARWIDTH | XMIN_COMMITTED | XMAX_INVALID */
    t[22] = (uint8_t)CDS_TUPLE_HDR;   /* t_hoff */
    t[23] = 0;
}

/* Append one tuple whose user data (everything after t_hoff) is `data[0..len)`. */
int cds_synth_rel_append(cds_synth_rel *r, const uint8_t *data, uint32_t len)
{
    uint32_t idx = 0;
    uint8_t *t = reserve_tuple(r, CDS_TUPLE_HDR + len, &idx);
    if (!t) return -1;
    write_tuple_header(r, t, idx);
    memcpy(t + CDS_TUPLE_HDR, data, len);
    return 0;
}

/* Store `payload[0..n)` as the TOAST value `valueid`: chunks of at most `chunk_size`
 * (8140 in Greenplum 5) bytes, each a tuple (chunk_id u32, chunk_seq u32, bytea with a
 * big-endian 4-byte varlena header).  `first_seq` lets tests write out-of-order pieces. */
int cds_synth_toast_value(cds_synth_rel *r, uint32_t valueid, const uint8_t *payload, uint64_t n,
                          uint32_t chunk_size, uint32_t first_seq)
{
    if (chunk_size == 0 || chunk_size > CDS_TOAST_MAX_CHUNK) chunk_size = CDS_TOAST_MAX_CHUNK;
    uint32_t seq = first_seq;
    uint64_t off = 0;
    do {
        uint32_t c = (uint32_t)((n - off) < chunk_size ? (n - off) : chunk_size);
        uint32_t idx = 0;
        uint8_t *t = reserve_tuple(r, CDS_TUPLE_HDR + 8u + 4u + c, &idx);
        if (!t) return -1;
        write_tuple_header(r, t, idx);
        put32(t + CDS_TUPLE_HDR, valueid);
        put32(t + CDS_TUPLE_HDR + 4, seq);
        put32be(t + CDS_TUPLE_HDR + 8, 4u + c);
        memcpy(t + CDS_TUPLE_HDR + 12, payload + off, c);
        off += c;
        seq++;
    } while (off < n);
    return 0;
}

/* ---------------------------------------------------------------------------------------
 * pglz-compatible encoders (stream only: no 8-byte header; the caller prepends
 * vl_len_/rawsize as the storage form requires).
 * ------------------------------------------------------------------------------------- */

/* Literal-only stream: one zero control byte per 8 literals (SURVEY.md §8c "stand-in"). */
int64_t cds_synth_pglz_literal(const uint8_t *src, int64_t n, uint8_t *dst, int64_t cap)
{
    int64_t need = n + (n + 7) / 8;
    if (need > cap) return -1;
    int64_t o = 0;
    for (int64_t i = 0; i < n; i += 8) {
        int64_t k = n - i < 8 ? n - i : 8;
        dst[o++] = 0;
        memcpy(dst + o, src + i, (size_t)k);
        o += k;
    }
    return o;
}

#define HBITS 15
#define HSIZE (1 << HBITS)
#define WSIZE 4096
#define MAX_OFF 4095
#define MAX_LEN 273

static inline uint32_t h3(const uint8_t *p)
{
    uint32_t v = (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16);
    return (v * 2654435761u) >> (32 - HBITS);
}

/* Greedy hash-chain LZ producing the pglz item format (pg_page_reader.py:191-231 is the
 * decoder of record): control byte, then 8 items LSB first; literal = 1 byte; match =
 * [(off>>8)<<4 | (len-3)] [off&0xFF] for len 3..17, or low nibble 0xF + extension byte
 * (len-18) for len 18..273.  `effort` bounds the chain walk.  Returns the stream length or
 * -1 when `cap` is too small. */
int64_t cds_synth_pglz_encode(const uint8_t *src, int64_t n, uint8_t *dst, int64_t cap, int effort)
{
    int32_t *head = (int32_t *)malloc(sizeof(int32_t) * HSIZE);
    int32_t *prev = (int32_t *)malloc(sizeof(int32_t) * WSIZE);
    if (!head || !prev) { free(head); free(prev); return -1; }
    for (int i = 0; i < HSIZE; i++) head[i] = -1;
    if (effort < 1) effort = 1;

    int64_t o = 0, i = 0;
    int64_t ctrl_pos = -1;
    int nitems = 8;
    uint8_t ctrl = 0;
    int64_t rc = 0;

    while (i < n) {
        if (nitems == 8) {
            if (ctrl_pos >= 0) dst[ctrl_pos] = ctrl;
            if (o >= cap) { rc = -1; goto out; }
            ctrl_pos = o++;
            ctrl = 0;
            nitems = 0;
        }
        int best_len = 0, best_off = 0;
        if (i + 3 <= n) {
            uint32_t h = h3(src + i);
            int32_t cand = head[h];
            int tries = effort;
            int64_t maxl = n - i < MAX_LEN ? n - i : MAX_LEN;
            while (cand >= 0 && i - cand <= MAX_OFF && tries-- > 0) {
                const uint8_t *a = src + cand, *b = src + i;
                int l = 0;
                while (l < maxl && a[l] == b[l]) l++;
                if (l > best_len) { best_len = l; best_off = (int)(i - cand); if (l == maxl) break; }
                int32_t nx = prev[cand & (WSIZE - 1)];
                if (nx >= cand) break; /* stale slot */
                cand = nx;
            }
        }
        if (best_len >= 3) {
            int need = best_len >= 18 ? 3 : 2;
            if (o + need > cap) { rc = -1; goto out; }
            ctrl |= (uint8_t)(1u << nitems);
            if (best_len >= 18) {
                dst[o++] = (uint8_t)(((best_off >> 8) << 4) | 0x0F);
                dst[o++] = (uint8_t)(best_off & 0xFF);
                dst[o++] = (uint8_t)(best_len - 18);
            } else {
                dst[o++] = (uint8_t)(((best_off >> 8) << 4) | (best_len - 3));
                dst[o++] = (uint8_t)(best_off & 0xFF);
            }
            int64_t end = i + best_len;
            for (; i < end; i++) {
                if (i + 3 <= n) {
                    uint32_t h = h3(src + i);
                    prev[i & (WSIZE - 1)] = head[h];
                    head[h] = (int32_t)i;
                }
            }
        } else {
            if (o + 1 > cap) { rc = -1; goto out; }
            dst[o++] = src[i];
            if (i + 3 <= n) {
                uint32_t h = h3(src + i);
                prev[i & (WSIZE - 1)] = head[h];
                head[h] = (int32_t)i;
            }
            i++;
        }
        nitems++;
    }
    if (ctrl_pos >= 0) dst[ctrl_pos] = ctrl;
    rc = o;
out:
    free(head);
    free(prev);
    return rc;
}

/* Exact normalisation table of SURVEY.md §8c(3): ((x/255.0) - mean[c]) / std[c] evaluated in
 * double, rounded once to float (the commented-out formula of
 * preprocessing/imagenet/generate_h5_file.py:77-81).  Fills lut[3][256]. */
void cds_synth_norm_lut(float *lut)
{
    static const double mean[3] = {0.485, 0.456, 0.406};
    static const double sd[3] = {0.229, 0.224, 0.225};
    for (int c = 0; c < 3; c++)
        for (int x = 0; x < 256; x++)
            lut[c * 256 + x] = (float)((((double)x / 255.0) - mean[c]) / sd[c]);
}

/* u8 HWC pixels -> normalised f32 HWC through the table above. */
void cds_synth_apply_lut(const uint8_t *px, int64_t n_pixels, const float *lut, float *out)
{
    for (int64_t p = 0; p < n_pixels; p++) {
        out[3 * p + 0] = lut[0 * 256 + px[3 * p + 0]];
        out[3 * p + 1] = lut[1 * 256 + px[3 * p + 1]];
        out[3 * p + 2] = lut[2 * 256 + px[3 * p + 2]];
    }
}
