// page_scan.cuh — page / ItemId / tuple scan kernels and the detoast plan.
//
// Replaces, on the device (paths relative to /root/reference/cerebro_gpdb/):
//   deserialize_page_header    pg_page_reader.py:253-269
//   deserialize_item           pg_page_reader.py:285-305
//   desirialize_tupdata        pg_page_reader.py:328-351
//   generator_page(toast=True) pg_page_reader.py:385-422
//   read_toast_all_bytes       pg_page_reader.py:497-524
//   detoast bookkeeping        pg_page_reader.py:527-612 (order of results, size checks)
//
// All of this is metadata work: it touches a few sectors per 32 KB page (< 1 % of the bytes);
// the bulk traffic is in toast_gather.cuh.
#pragma once
#include <cooperative_groups.h>

#include "cds_common.cuh"

namespace cds {
namespace {   // internal linkage: with relocatable device code (-rdc) the optimiser then treats the file as a whole program

// Shared by the heap and the TOAST relation: item count per page from pd_lower, device-wide
// exclusive scan into base[0..n_pages], validation of the header.
//   toast != 0 additionally asserts pd_special == BLCKSZ (ppr:388-389).
__device__ __forceinline__ void d_page_index(const uint8_t *__restrict__ pages, uint32_t n_pages, uint32_t *__restrict__ base,
                                             DevState *st, int toast, uint32_t cap_items, ScanWs ws)
{
    const uint32_t stage = toast ? 1u : 0u;
    auto count_of = [&](uint32_t p) -> uint32_t {
        const uint8_t *pg = pages + (size_t)p * kPage;
        // pd_lower, pd_upper, pd_special live at bytes 12..17 of the header (ppr:629-638)
        const uint4 w = *reinterpret_cast<const uint4 *>(pg + 8);      // bytes 8..23
        const uint32_t pd_lower = w.y & 0xFFFFu, pd_special = w.z & 0xFFFFu;
        if (pd_lower < kPageHdr || ((pd_lower - kPageHdr) & 3u)) { report(st, stage, p, 0, 0, CDS_ERR_ITEMID_ALIGN); return 0u; }
        if (pd_lower > kPage) { report(st, stage, p, 0, 1, CDS_ERR_TUPLE); return 0u; }
        if (toast && pd_special != kPage) { report(st, stage, p, 0, 2, CDS_ERR_TOAST_SPECIAL); return 0u; }
        return (pd_lower - kPageHdr) >> 2;
    };
    bool last;
    const uint32_t total = device_exclusive_scan<uint32_t, 4>(count_of, base, n_pages, ws, true, &last);
    if (last && threadIdx.x == 0) {
        uint32_t n = total;
        if (total > cap_items) {
            // refuse everything: downstream kernels see zero items and skip their pages
            report(st, stage, 0, 0, 0, CDS_ERR_CAPACITY);
            atomicOr(&st->flags, toast ? kFlagChunkOverflow : kFlagRowOverflow);
            n = 0;
        }
        if (toast) st->n_chunks = n; else st->n_rows = n;
    }
}

__global__ void __launch_bounds__(kScanThreads)
k_page_index(const uint8_t *__restrict__ pages, uint32_t n_pages, uint32_t *__restrict__ base,
             DevState *st, int toast, uint32_t cap_items, ScanWs ws)
{
    d_page_index(pages, n_pages, base, st, toast, cap_items, ws);
}

// Heap (blockIdx.y == 0) and TOAST (blockIdx.y == 1) relation indexed by one launch (one-shot decode).
__global__ void __launch_bounds__(kScanThreads)
k_page_index2(const uint8_t *__restrict__ heap, uint32_t n_heap, uint32_t *__restrict__ heap_base, uint32_t cap_rows, ScanWs ws_h,
              const uint8_t *__restrict__ toast, uint32_t n_toast, uint32_t *__restrict__ toast_base, uint32_t cap_chunks, ScanWs ws_t,
              DevState *st)
{
    if (blockIdx.y == 0) d_page_index(heap, n_heap, heap_base, st, 0, cap_rows, ws_h);
    else d_page_index(toast, n_toast, toast_base, st, 1, cap_chunks, ws_t);
}

// Reset the per-value statistics of row r and enter its two va_valueid words into the hash table.
// value index v = 2*row + (0 independent_var | 1 dependent_var) = the row number in df_toast.
__device__ inline void index_row_values(const uint32_t *row, uint32_t r, unsigned long long *htab, uint32_t hmask,
                                        uint32_t *vfirst, uint32_t *vcount, unsigned long long *vbytes, uint32_t *vseq0)
{
#pragma unroll
    for (uint32_t a = 0; a < 2; a++) {
        const uint32_t v = 2 * r + a;
        vfirst[v] = kNone; vcount[v] = 0; vbytes[v] = 0; vseq0[v] = kNone;
    }
    if (row[RW_INLINE_LEN] == kNone) return;
    htab_insert(htab, hmask, row[RW_IND + 3], 2 * r);                        // always external (ppr:348)
    // an inline dependent_var still contributes the bytes at its va_valueid position to the
    // reference's chunk filter (ppr:516); mark it so a hit is reported the way .iloc[0] fails
    htab_insert(htab, hmask, row[RW_DEP + 3], (row[RW_DEP_EXTERNAL] ? 0u : 0x80000000u) | (2 * r + 1));
}

// kHeapSub blocks of 64 threads per heap page, one thread per ItemId: parse the MADlib packed row and
// write its 16-word descriptor in table order.  (A page holds at most 430 such tuples.)
constexpr uint32_t kHeapSub = 7;
// `base` may live in global or in shared memory.  The work is cut into groups of 64 threads: group `vb` of `nvb`, thread `vt`.
__device__ __forceinline__ void d_heap_rows(const uint8_t *__restrict__ heap, uint32_t n_pages, const uint32_t *base,
                                            uint32_t *__restrict__ rows, DevState *st, unsigned long long *htab, uint32_t hmask,
                                            uint32_t *vfirst, uint32_t *vcount, unsigned long long *vbytes, uint32_t *vseq0,
                                            uint32_t vb, uint32_t nvb, uint32_t vt)
{
    for (uint32_t w = vb; w < n_pages * kHeapSub; w += nvb) {
        const uint32_t p = w / kHeapSub, sub = w - p * kHeapSub;
        const uint8_t *pg = heap + (size_t)p * kPage;
        const uint32_t r0 = base[p], n = base[p + 1] - r0;
        for (uint32_t i = sub * 64u + vt; i < n; i += kHeapSub * 64u) {
            uint32_t *row = rows + (size_t)(r0 + i) * kRowWords;
            uint32_t iid = *reinterpret_cast<const uint32_t *>(pg + kPageHdr + 4u * i);
            uint32_t lp_off = iid & 0x7FFFu, lp_len = iid >> 17;            // ppr:287-300
            uint32_t end = min(lp_off + lp_len, kPage);                      // python slices clamp
            bool bad = end < lp_off + kItemHdr;
            uint32_t t_hoff = bad ? 0u : pg[lp_off + 22];                    // ppr:272-276
            bad = bad || (lp_off + t_hoff > end);
            uint32_t tdlen = bad ? 0u : end - (lp_off + t_hoff);
            bad = bad || tdlen < 44u;                                        // unpack of tupdata[24:44]
            if (bad) {
                report(st, 0, p, 1, i, CDS_ERR_TUPLE);
                for (uint32_t k = 0; k < kRowWords; k++) row[k] = 0;
                row[RW_INLINE_LEN] = kNone;                                  // marks the row unusable
                index_row_values(row, r0 + i, htab, hmask, vfirst, vcount, vbytes, vseq0);
                continue;
            }
            const uint8_t *td = pg + lp_off + t_hoff;
            row[RW_DIST_KEY] = ld_u32(td);
            row[RW_BUFFER_ID] = ld_u32(td + tdlen - 4);                      // ppr:330
#pragma unroll
            for (int k = 0; k < 5; k++) row[RW_IND + k] = ld_u32(td + 4 + 4 * k);    // ppr:337
#pragma unroll
            for (int k = 0; k < 5; k++) row[RW_DEP + k] = ld_u32(td + 24 + 4 * k);   // ppr:342
            // extended && !external && compressed  <=>  (first byte & 0xC0) == 0x40   (ppr:343)
            row[RW_DEP_EXTERNAL] = ((td[24] & 0xC0u) == 0x40u) ? 0u : 1u;
            unsigned long long off = (unsigned long long)(td + 24 - heap);
            row[RW_INLINE_OFF_LO] = (uint32_t)off;
            row[RW_INLINE_OFF_HI] = (uint32_t)(off >> 32);
            row[RW_INLINE_LEN] = tdlen - 24u;
            index_row_values(row, r0 + i, htab, hmask, vfirst, vcount, vbytes, vseq0);
        }
    }
}

__global__ void __launch_bounds__(64)
k_heap_rows(const uint8_t *__restrict__ heap, uint32_t n_pages, const uint32_t *__restrict__ base,
            uint32_t *__restrict__ rows, DevState *st, unsigned long long *htab, uint32_t hmask,
            uint32_t *vfirst, uint32_t *vcount, unsigned long long *vbytes, uint32_t *vseq0)
{
    if (st->flags & kFlagRowOverflow) return;
    d_heap_rows(heap, n_pages, base, rows, st, htab, hmask, vfirst, vcount, vbytes, vseq0, blockIdx.x, gridDim.x, threadIdx.x);
}

// One launch for the front of the one-call decode: plane 0 indexes the heap pages, plane 1 the TOAST pages, and
// plane 2 parses the heap rows at the same time.  The row parse needs the heap's item-count prefix; a heap of up to
// kFrontMaxHeapPages pages (a 1 GiB segment of one-image rows has 32) is small enough for every block of plane 2 to
// compute that prefix itself in shared memory, which makes the three planes independent.
constexpr uint32_t kFrontMaxHeapPages = 1020;                     // 4 pages per thread, and one more prefix entry than pages
__global__ void __launch_bounds__(kScanThreads)
k_front(const uint8_t *__restrict__ heap, uint32_t n_heap, uint32_t *__restrict__ heap_base, uint32_t cap_rows, ScanWs ws_h,
        const uint8_t *__restrict__ toast, uint32_t n_toast, uint32_t *__restrict__ toast_base, uint32_t cap_chunks, ScanWs ws_t,
        DevState *st, uint32_t *__restrict__ rows, unsigned long long *htab, uint32_t hmask,
        uint32_t *vfirst, uint32_t *vcount, unsigned long long *vbytes, uint32_t *vseq0)
{
    if (blockIdx.y == 0) { d_page_index(heap, n_heap, heap_base, st, 0, cap_rows, ws_h); return; }
    if (blockIdx.y == 1) { d_page_index(toast, n_toast, toast_base, st, 1, cap_chunks, ws_t); return; }
    if (blockIdx.x * 4u >= n_heap * kHeapSub) return;                        // no row work for this block
    __shared__ uint32_t sh_base[kFrontMaxHeapPages + 4];
    __shared__ uint32_t sh_warp[kScanThreads / 32];
    // item counts with the rules of d_page_index (plane 0 reports the errors), 4 pages per thread
    uint32_t c[4], tsum = 0;
#pragma unroll
    for (uint32_t k = 0; k < 4; k++) {
        const uint32_t p = threadIdx.x * 4u + k;
        c[k] = 0;
        if (p < n_heap) {
            const uint32_t pd_lower = *reinterpret_cast<const uint32_t *>(heap + (size_t)p * kPage + 12) & 0xFFFFu;
            if (pd_lower >= kPageHdr && !((pd_lower - kPageHdr) & 3u) && pd_lower <= kPage) c[k] = (pd_lower - kPageHdr) >> 2;
        }
        tsum += c[k];
    }
    const uint32_t winc = warp_inclusive_scan(tsum);
    if ((threadIdx.x & 31) == 31) sh_warp[threadIdx.x >> 5] = winc;
    __syncthreads();
    uint32_t off = 0, total = 0;
#pragma unroll
    for (uint32_t w = 0; w < kScanThreads / 32; w++) {
        if (w < (threadIdx.x >> 5)) off += sh_warp[w];
        total += sh_warp[w];
    }
    uint32_t run = off + winc - tsum;
#pragma unroll
    for (uint32_t k = 0; k < 4; k++) {
        const uint32_t p = threadIdx.x * 4u + k;
        if (p <= n_heap) sh_base[p] = run;
        run += c[k];
    }
    __syncthreads();
    if (total > cap_rows) return;                                            // plane 0 refuses the segment (CDS_ERR_CAPACITY)
    d_heap_rows(heap, n_heap, sh_base, rows, st, htab, hmask, vfirst, vcount, vbytes, vseq0,
                blockIdx.x * 4u + (threadIdx.x >> 6), gridDim.x * 4u, threadIdx.x & 63u);
}

// cds_set_rows path: rows arrive from the host; build the hash table from them.
__global__ void __launch_bounds__(256)
k_index_values(const uint32_t *__restrict__ rows, const DevState *st, unsigned long long *htab,
               uint32_t hmask, uint32_t *vfirst, uint32_t *vcount, unsigned long long *vbytes,
               uint32_t *vseq0)
{
    const uint32_t n_rows = st->n_rows;
    for (uint32_t r = blockIdx.x * blockDim.x + threadIdx.x; r < n_rows; r += gridDim.x * blockDim.x)
        index_row_values(rows + (size_t)r * kRowWords, r, htab, hmask, vfirst, vcount, vbytes, vseq0);
}

__global__ void k_set_n_rows(DevState *st, uint32_t n_rows) { st->n_rows = n_rows; }

// One thread per TOAST page: physical walk upward from pd_upper (ppr:391-414).
__global__ void __launch_bounds__(128)
k_toast_walk(const uint8_t *__restrict__ toast, uint32_t n_pages, const uint32_t *__restrict__ base,
             const unsigned long long *__restrict__ htab, uint32_t hmask,
             uint2 *__restrict__ dir_id, uint2 *__restrict__ dir_loc, uint32_t *__restrict__ dir_ref,
             uint32_t *__restrict__ chunk_flag,
             uint32_t *vfirst, uint32_t *vcount, unsigned long long *vbytes, uint32_t *vseq0, DevState *st, uint32_t page0)
{
    // (`toast`, `base`: the first page of this launch, which is page `page0` of the relation)
    uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_pages || (st->flags & kFlagChunkOverflow)) return;
    const uint8_t *pg = toast + (size_t)p * kPage;
    const uint32_t g0 = base[p], n = base[p + 1] - g0;
    uint32_t lp = ld_u16(pg + 14);                                           // pd_upper
    // The walk itself ignores the ItemIds, as the reference does; they only serve as prefetch hints so that
    // the dependent tuple-header reads below find their sectors on the way (or already in L2).
    for (uint32_t i = 0; i < n && i < 8u; i++) {
        const uint32_t hint = *reinterpret_cast<const uint32_t *>(pg + kPageHdr + 4u * i) & 0x7FFFu;
        if (hint + 40u <= kPage) asm volatile("prefetch.global.L2 [%0];" ::"l"(pg + hint + 16u));
    }
    bool dead = false;
    for (uint32_t i = 0; i < n; i++) {
        const uint32_t g = g0 + i;
        uint32_t id = 0, seq = 0, loc_off = 0, loc_len = 0, ref = kNone, hdr0 = 0;
        if (!dead) {
            lp = (lp + 7u) & ~7u;                                            // MAX_ALIGN (ppr:394)
            if (lp + kItemHdr > kPage) {                                     // the header unpack fails on a short slice (ppr:396)
                dead = true;
                report(st, 1, p, 1, i, CDS_ERR_TUPLE);
            }
        }
        if (!dead) {
            uint32_t t_hoff, hdr, chunkword;
            bool have = false;
            if (lp + 40u <= kPage) {
                // one round trip for the usual t_hoff == 24 tuple: bytes 16..39 hold t_hoff, chunk_id, chunk_seq, length word
                const uint2 a = *reinterpret_cast<const uint2 *>(pg + lp + 16), b = *reinterpret_cast<const uint2 *>(pg + lp + 24),
                            c = *reinterpret_cast<const uint2 *>(pg + lp + 32);
                t_hoff = (a.y >> 16) & 0xFFu;                                // byte 22 (ppr:272-276)
                if (t_hoff == 24u) { id = b.x; seq = b.y; chunkword = c.x; have = true; }
            } else {
                t_hoff = pg[lp + 22];
            }
            hdr = lp + t_hoff + 8u;                                          // ppr:399-403
            if (hdr + 4u > kPage) dead = true;
            else {
                if (!have) chunkword = ld_u32(pg + hdr);
                const uint32_t chunksize = __byte_perm(chunkword, 0, 0x0123) & 0x3FFFFFFFu;   // VARSIZE (ppr:130-131,404)
                if (chunksize < 4u || hdr + chunksize > kPage) dead = true;
                else {
                    if (!have) { id = ld_u32(pg + lp + t_hoff); seq = ld_u32(pg + lp + t_hoff + 4); }
                    loc_off = hdr + 4u;
                    loc_len = chunksize - 4u;
                    hdr0 = chunkword & 0xFFu;
                    lp += t_hoff + 8u + chunksize;                           // ppr:405-414
                }
            }
            if (dead) report(st, 1, p, 1, i, CDS_ERR_TUPLE);
        }
        if (!dead) {
            ref = htab_find(htab, hmask, id);                                // ppr:516
            if (ref != kNone) {
                if (ref & 0x80000000u) {
                    // only a non-external row carries this id: the reference dies in detoast (ppr:563-564)
                    report(st, 3, g, 0, 0, CDS_ERR_CATALOG_MISS);
                    ref = kNone;
                } else {
                    atomicMin(&vfirst[ref], g);
                    atomicAdd(&vcount[ref], 1u);
                    // low 36 bits: payload bytes of the value; high 28 bits: sum of its chunk_seq (mod 2^28)
                    atomicAdd(&vbytes[ref], (unsigned long long)loc_len + ((unsigned long long)(seq & 0x0FFFFFFFu) << 36));
                    if (seq == 0) atomicMin(&vseq0[ref], g);
                }
            }
        }
        dir_id[g] = make_uint2(id, seq);
        dir_loc[g] = make_uint2(page0 + p, loc_off | (loc_len << 16));
        dir_ref[g] = (ref == kNone) ? kNone : (ref | ((hdr0 & 0xC0u) ? 0x40000000u : 0u));
        chunk_flag[g] = 0;
    }
}

// ---- detoast plan ---------------------------------------------------------------------------------
// Five phases (rows -> ranks -> positions -> offsets -> copy descriptors).  They run either as one
// cooperative launch with grid-wide barriers in between (k_plan_fused) or as separate kernels.
struct PlanArgs {
    const uint32_t *rows;
    const uint8_t *heap, *toast;
    const uint2 *dir_id, *dir_loc;
    const uint32_t *dir_ref;
    uint32_t *vfirst, *vcount, *vseq0;
    unsigned long long *vbytes, *vscr;
    uint32_t *inline_flag, *chunk_flag, *row_first, *pos_of_row, *row_at_pos;
    unsigned long long *ind_len, *dep_len, *pos_ind, *pos_dep, *row_ind_off, *row_dep_off;
    CopyOp *ops;
    PglzJob *jobs;
    DevState *st;
    ScanWs ws[5];                 // inline rank, chunk rank, ind offsets, dep offsets, staging offsets
    uint32_t isb, dsb;            // sample sizes for the shape check (0: none)
    uint32_t ind_bf16;            // independent_var leaves as bfloat16 (stored float32): output offsets are halved
    uint32_t ind_u8norm;          // independent_var is uint8 pixels, leaves as float32 through the table: offsets x 4
    uint32_t lut_channels;
    uint32_t wide_threshold;      // count compressed values of at least this many decoded bytes (0: do not count)
    uint8_t *out_ind, *out_dep, *scratch;
    unsigned long long cap_ind, cap_dep, cap_scratch;
};

// Per row: decoded lengths, whether the row enters the result in the inline phase, and the
// first-seen chunk of its external values.
__device__ __forceinline__ void d_plan_rows(const PlanArgs &A)
{
    DevState *st = A.st;
    const uint32_t n_rows = st->n_rows;
    for (uint32_t r = blockIdx.x * blockDim.x + threadIdx.x; r < n_rows; r += gridDim.x * blockDim.x) {
        const uint32_t *row = A.rows + (size_t)r * kRowWords;
        unsigned long long len[2] = {0, 0};
        uint32_t first = kNone;
        const bool usable = row[RW_INLINE_LEN] != kNone;
        const bool is_inline = usable && row[RW_DEP_EXTERNAL] == 0;
        for (uint32_t a = 0; a < 2; a++) {
            const uint32_t v = 2 * r + a;
            A.vscr[v] = 0;
            if (!usable) continue;
            const uint32_t *ptr = row + (a ? RW_DEP : RW_IND);
            if (a == 1 && is_inline) {
                // inline pglz varlena: rawsize is the LE word after the BE length word (ppr:538)
                len[1] = ptr[1];
                continue;
            }
            if (A.vcount[v] == 0) continue;                                 // value has no chunks
            first = min(first, A.vfirst[v]);
            const uint32_t va_rawsize = ptr[1], va_extsize = ptr[2];
            if (va_extsize == va_rawsize - 4u) {
                len[a] = va_extsize;                                         // stored uncompressed
            } else {
                // compressed: rawsize is the first LE word of the stored payload (ppr:240, 574-594)
                uint32_t g = A.vseq0[v];
                if (g != kNone) {
                    uint2 loc = A.dir_loc[g];
                    if ((loc.y >> 16) >= 4u)
                        len[a] = ld_u32(A.toast + (size_t)loc.x * kPage + (loc.y & 0xFFFFu));
                }
                if (va_extsize >= 4u && (a == 1u || !(A.ind_bf16 | A.ind_u8norm)) && pglz_literal_only(va_extsize - 4u, len[a]))
                    continue;                                                // literal-only stream: unpacked by the gather itself
                A.vscr[v] = ((unsigned long long)va_extsize + 15ull) & ~15ull;
                if (A.wide_threshold && len[a] >= A.wide_threshold) atomicAdd(&st->n_wide, 1u);
            }
        }
        A.ind_len[r] = len[0];
        A.dep_len[r] = len[1];
        A.inline_flag[r] = is_inline ? 1u : 0u;
        A.row_first[r] = first;
        if (!is_inline && first != kNone) A.chunk_flag[first] = 1u;
    }
}

// which == 0: rank of inline rows (table order).  which == 1: rank of first-seen chunks.
__device__ __forceinline__ void d_plan_scan1(const PlanArgs &A, int which)
{
    bool last;
    if (which == 0) {
        uint32_t *f = A.inline_flag;
        const uint32_t tot = device_exclusive_scan<uint32_t, 8>([&](uint32_t i) { return f[i]; }, f, A.st->n_rows, A.ws[0], false, &last);
        if (last && threadIdx.x == 0) A.st->n_inline = tot;
    } else {
        uint32_t *f = A.chunk_flag;
        device_exclusive_scan<uint32_t, 8>([&](uint32_t i) { return f[i]; }, f, A.st->n_chunks, A.ws[1], true, &last);
    }
}

// Result position of every row: inline rows first in table order (ppr:532-552), then rows in
// the order their first external chunk is met (ppr:553-611; an existing entry keeps its place).
__device__ __forceinline__ void d_assign_pos(const PlanArgs &A)
{
    DevState *st = A.st;
    const uint32_t n_rows = st->n_rows, n_inline = st->n_inline, n_chunks = st->n_chunks;
    if (blockIdx.x == 0 && threadIdx.x == 0) st->n_pos = n_inline + A.chunk_flag[n_chunks];
    for (uint32_t r = blockIdx.x * blockDim.x + threadIdx.x; r < n_rows; r += gridDim.x * blockDim.x) {
        const uint32_t *row = A.rows + (size_t)r * kRowWords;
        const bool usable = row[RW_INLINE_LEN] != kNone;
        const bool is_inline = usable && row[RW_DEP_EXTERNAL] == 0;
        uint32_t pos = kNone;
        if (is_inline) pos = A.inline_flag[r];
        else if (A.row_first[r] != kNone) pos = n_inline + A.chunk_flag[A.row_first[r]];
        A.pos_of_row[r] = pos;
        if (pos == kNone) continue;     // no chunk of this row exists: the reference never emits it
        const unsigned long long il = A.ind_len[r], dl = A.dep_len[r];
        A.row_at_pos[pos] = r;
        A.pos_ind[pos] = il;
        A.pos_dep[pos] = dl;
        if (il == 0 || dl == 0) report(st, 4, pos, 0, 0, CDS_ERR_MISSING_VALUE);
        if (A.isb && A.dsb && il && dl) {
            if (il % A.isb || dl % A.dsb || il / A.isb != dl / A.dsb) report(st, 4, pos, 1, 0, CDS_ERR_SHAPE);
        }
    }
}

// Three independent 64-bit exclusive scans: output offsets of the two value streams in result order,
// and staging offsets of compressed external values.
__device__ __forceinline__ void d_plan_scan2(const PlanArgs &A, int which)
{
    bool last;
    DevState *st = A.st;
    if (which == 0) {
        unsigned long long *x = A.pos_ind;
        const unsigned long long t = device_exclusive_scan<unsigned long long, 8>([&](uint32_t i) { return x[i]; }, x, st->n_pos, A.ws[2], false, &last);
        if (last && threadIdx.x == 0) st->ind_bytes = t;
    } else if (which == 1) {
        unsigned long long *x = A.pos_dep;
        const unsigned long long t = device_exclusive_scan<unsigned long long, 8>([&](uint32_t i) { return x[i]; }, x, st->n_pos, A.ws[3], false, &last);
        if (last && threadIdx.x == 0) st->dep_bytes = t;
    } else {
        unsigned long long *x = A.vscr;
        const unsigned long long t = device_exclusive_scan<unsigned long long, 8>([&](uint32_t i) { return x[i]; }, x, 2u * st->n_rows, A.ws[4], false, &last);
        if (last && threadIdx.x == 0) st->scratch_bytes = t;
    }
}

// Copy descriptor of one chunk (sequence number idx, `size` payload bytes at `loc`) of value v = 2 * row + att whose row
// sits at result position pos.  The chunk checks are the caller's.  dst == 0: nothing to copy.
__device__ __forceinline__ CopyOp make_copy_op(const PlanArgs &A, uint32_t v, uint32_t pos, uint32_t idx, uint32_t size, uint2 loc,
                                               uint32_t va_rawsize, uint32_t va_extsize, unsigned long long rawlen)
{
    const uint32_t a = v & 1u;
    const uint32_t half = A.ind_bf16 ? 1u : 0u, quad = A.ind_u8norm ? 2u : 0u;
    CopyOp op;
    const bool compressed = va_extsize != va_rawsize - 4u;
    op.src_off = loc.y & 0xFFFFu;
    op.len = size;
    if (compressed && va_extsize >= 4u && (a == 1u || !(half | quad)) && pglz_literal_only(va_extsize - 4u, rawlen)) {
        // literal-only pglz stream: stream bytes [s_lo, s_hi) of the value live in this chunk
        const unsigned long long s_lo = idx ? (unsigned long long)idx * kChunkMax - 4ull : 0ull;
        const uint32_t skip = idx ? 0u : 4u;                                 // chunk 0 starts with the rawsize word
        const uint32_t n_stream = size - skip;
        const unsigned long long s_hi = s_lo + n_stream;
        const unsigned long long o_lo = s_lo - (s_lo + 8ull) / 9ull;
        const unsigned long long s_fin = 9ull * (((rawlen + 7ull) >> 3) - 1ull);   // control byte of the final group
        const uint32_t fin = (s_fin >= s_lo && s_fin < s_hi) ? (uint32_t)(rawlen & 7ull) : 0u;
        op.dst = (unsigned long long)((a ? A.out_dep + A.pos_dep[pos] : A.out_ind + A.pos_ind[pos]) + o_lo);
        op.src_off = ((loc.y & 0xFFFFu) + skip) | ((uint32_t)(s_lo % 9ull) << 16) | (fin << 20);
        op.len = n_stream | kOpLiteral;
    } else if (compressed) {
        op.dst = (unsigned long long)(A.scratch + A.vscr[v]) + (unsigned long long)idx * kChunkMax;
    } else if (a) {
        op.dst = (unsigned long long)(A.out_dep + A.pos_dep[pos]) + (unsigned long long)idx * kChunkMax;
    } else if (quad) {
        const unsigned long long voff = (unsigned long long)idx * kChunkMax;     // byte (= pixel component) inside the value
        op.dst = (unsigned long long)(A.out_ind + ((A.pos_ind[pos] + voff) << 2));
        op.src_off |= (uint32_t)(voff % A.lut_channels) << 16;
        op.len = size | kOpU8Norm;
    } else if (!half) {
        op.dst = (unsigned long long)(A.out_ind + A.pos_ind[pos]) + (unsigned long long)idx * kChunkMax;
    } else if (((size | va_extsize) & 3u) == 0u && (A.pos_ind[pos] & 3ull) == 0ull) {
        op.dst = (unsigned long long)(A.out_ind + ((A.pos_ind[pos] + (unsigned long long)idx * kChunkMax) >> 1));
        op.len = size | kOpToBf16;
    } else {
        report(A.st, 4, pos, 1, 1, CDS_ERR_SHAPE);                           // not a whole number of float32 elements
        op.dst = 0;
    }
    return op;
}

// pglz job of value v = 2 * row + att (row usable, at result position pos; the output fits): an inline datum, or a compressed
// external value whose payload the gather stages.  slen == kNone: the value needs no job (or was refused, and reported).
__device__ __forceinline__ PglzJob make_value_job(const PlanArgs &A, uint32_t v, uint32_t pos)
{
    DevState *st = A.st;
    const uint32_t half = A.ind_bf16 ? 1u : 0u, quad = A.ind_u8norm ? 2u : 0u;
    PglzJob job;
    job.src = 0; job.dst = 0; job.slen = kNone; job.rawsize = 0; job.err_key = 0;
    const uint32_t r = v >> 1, a = v & 1u;
    const uint32_t *row = A.rows + (size_t)r * kRowWords;
    const unsigned long long my_ind_off = A.pos_ind[pos], my_dep_off = A.pos_dep[pos];
    if (a == 1 && row[RW_DEP_EXTERNAL] == 0) {
        // inline datum: tupdata[24:] starts with the 8-byte PGLZ header (ppr:344, 537-543)
        const uint8_t *src = A.heap + (((unsigned long long)row[RW_INLINE_OFF_HI] << 32) | row[RW_INLINE_OFF_LO]);
        const uint32_t avail = row[RW_INLINE_LEN];
        const uint32_t varsize = ld_u32be(src) & 0x3FFFFFFFu;
        job.err_key = make_err_key(2, r, 0, 0, CDS_ERR_PGLZ_CORRUPT);
        if (avail < 8u || varsize < 8u || varsize > avail) {
            report_key(st, job.err_key);
        } else {
            job.src = (unsigned long long)(src + 8);
            job.slen = varsize - 8u;
            job.rawsize = (uint32_t)A.dep_len[r];
            job.dst = (unsigned long long)(A.out_dep + my_dep_off);
        }
    } else if (A.vcount[v] != 0) {
        const uint32_t *ptr = row + (a ? RW_DEP : RW_IND);
        const uint32_t va_rawsize = ptr[1], va_extsize = ptr[2];
        const unsigned long long len = a ? A.dep_len[r] : A.ind_len[r];
        const bool literal = va_extsize >= 4u && (a == 1u || !(half | quad)) && pglz_literal_only(va_extsize - 4u, len);
        if (va_extsize != va_rawsize - 4u && !literal) {
            job.err_key = make_err_key(3, A.vfirst[v], 3, 0, CDS_ERR_PGLZ_CORRUPT);
            if (va_extsize < 4u) {
                report_key(st, job.err_key);
            } else {
                // staged payload = LE rawsize + stream (ppr:574-594 builds header + payload)
                job.src = (unsigned long long)(A.scratch + A.vscr[v] + 4);
                job.slen = va_extsize - 4u;
                job.rawsize = (uint32_t)len;
                if (a) job.dst = (unsigned long long)(A.out_dep + my_dep_off);
                else if (quad) { job.dst = (unsigned long long)(A.out_ind + (my_ind_off << 2)); job.slen |= kJobU8Norm; }
                else if (!half) job.dst = (unsigned long long)(A.out_ind + my_ind_off);
                else if ((len & 3ull) == 0ull && (my_ind_off & 3ull) == 0ull) {
                    job.dst = (unsigned long long)(A.out_ind + (my_ind_off >> 1));
                    job.slen |= kJobToBf16;
                } else {
                    report(st, 4, pos, 1, 1, CDS_ERR_SHAPE);
                    job.slen = kNone;
                }
            }
        }
    }
    return job;
}

// One thread per chunk: the reference's per-chunk checks (ppr:570-596) and the copy descriptor.
// Also one pglz job per value (inline datums and compressed external values).
__device__ __forceinline__ void d_resolve(const PlanArgs &A)
{
    DevState *st = A.st;
    const uint32_t *rows = A.rows;
    const uint32_t n_chunks = st->n_chunks, n_rows = st->n_rows;
    const uint32_t half = A.ind_bf16 ? 1u : 0u;                             // bf16 output: byte offsets / 2
    const uint32_t quad = A.ind_u8norm ? 2u : 0u;                           // u8 -> f32 output: byte offsets x 4
    const bool fits = ((st->ind_bytes >> half) << quad) <= A.cap_ind && st->dep_bytes <= A.cap_dep && st->scratch_bytes <= A.cap_scratch;
    if (!fits && blockIdx.x == 0 && threadIdx.x == 0) report(st, 4, 0, 2, 0, CDS_ERR_CAPACITY);
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x, nth = gridDim.x * blockDim.x;

    for (uint32_t g = tid; g < n_chunks; g += nth) {
        CopyOp op;
        op.dst = 0; op.src_off = 0; op.len = 0;
        const uint32_t ref = A.dir_ref[g];
        if (ref != kNone) {
            const uint32_t v = ref & 0x3FFFFFFFu, r = v >> 1, a = v & 1u;
            const uint32_t *ptr = rows + (size_t)r * kRowWords + (a ? RW_DEP : RW_IND);
            const uint32_t va_rawsize = ptr[1], va_extsize = ptr[2];
            const uint32_t numchunks = va_extsize ? (va_extsize - 1u) / kChunkMax + 1u : 0u;   // ppr:155-156
            const uint32_t first = A.vfirst[v];
            const uint2 id = A.dir_id[g], loc = A.dir_loc[g];
            const uint32_t idx = id.y, size = loc.y >> 16;
            const unsigned long long vb = A.vbytes[v] & 0xFFFFFFFFFull;
            const uint32_t seqsum = (uint32_t)(A.vbytes[v] >> 36);
            bool ok = true;
            if (A.vcount[v] != numchunks) {                                 // ppr:570-572
                if (g == first) report(st, 3, first, 0, 0, CDS_ERR_CHUNK_COUNT);
                ok = false;
            }
            if (ref & 0x40000000u) { report(st, 3, first, 1, idx, CDS_ERR_CHUNK_EXTENDED); ok = false; }       // ppr:580
            if (idx == 0 && size < va_extsize && size != kChunkMax) { report(st, 3, first, 1, idx, CDS_ERR_MAX_CHUNK); ok = false; }   // ppr:584-586
            if (numchunks && idx < numchunks - 1u && size != kChunkMax) { report(st, 3, first, 1, idx, CDS_ERR_CHUNK_SIZE); ok = false; }   // ppr:587-588
            if (numchunks && idx == numchunks - 1u && (unsigned long long)idx * kChunkMax + size != va_extsize) { report(st, 3, first, 1, idx, CDS_ERR_CHUNK_SIZE); ok = false; }   // ppr:589-591
            if (idx >= numchunks) { report(st, 3, first, 1, idx, CDS_ERR_CHUNK_SIZE); ok = false; }   // outside the pointer's extent
            if (g == first && vb != va_extsize) { report(st, 3, first, 2, 0, CDS_ERR_FINAL_SIZE); }  // ppr:595-596
            // every chunk_seq 0..numchunks-1 exactly once?  (The reference sorts and joins whatever it finds; a repeated
            // sequence number would land two chunks on the same bytes here, so it is refused.)
            const uint32_t want_sum = (uint32_t)((((unsigned long long)numchunks * (numchunks ? numchunks - 1u : 0u)) >> 1) & 0x0FFFFFFFull);
            if (A.vcount[v] == numchunks && seqsum != want_sum) {
                if (g == first) report(st, 3, first, 1, 0x1FFFF, CDS_ERR_CHUNK_SIZE);
                ok = false;
            }
            const uint32_t pos = A.pos_of_row[r];
            if (ok && vb == va_extsize && fits && pos != kNone) {
                op = make_copy_op(A, v, pos, idx, size, loc, va_rawsize, va_extsize, a ? A.dep_len[r] : A.ind_len[r]);
            }
        }
        A.ops[g] = op;
    }

    // pglz jobs are compacted: expensive jobs grow from slot 0 upward, cheap ones from the last slot downward;
    // k_pglz_decode hands out the front list first.
    const uint32_t last_slot = 2u * n_rows - 1u;
    for (uint32_t v = tid; v < 2u * n_rows; v += nth) {
        PglzJob job;
        job.slen = kNone;
        const uint32_t r = v >> 1, a = v & 1u;
        const uint32_t *row = rows + (size_t)r * kRowWords;
        const uint32_t pos = A.pos_of_row[r];
        if (a == 0) {                                                        // exported to the host mirror
            A.row_ind_off[r] = pos == kNone ? ~0ull : A.pos_ind[pos];
            A.row_dep_off[r] = pos == kNone ? ~0ull : A.pos_dep[pos];
        }
        if (fits && pos != kNone && row[RW_INLINE_LEN] != kNone) job = make_value_job(A, v, pos);
        if (job.slen != kNone) {
            // front list = the expensive streams (really compressed, not tiny): handed out first, so that the cheap ones
            // (literal-only streams copy 256 bytes per step; small labels) fill the tail of the launch
            const bool expensive = (job.slen & ~kFlagBits) < job.rawsize && job.rawsize >= 4096u;
            if (expensive) A.jobs[atomicAdd(&st->n_jobs_front, 1u)] = job;
            else A.jobs[last_slot - atomicAdd(&st->n_jobs_back, 1u)] = job;
        }
    }
}

__global__ void __launch_bounds__(256) k_plan_rows(const PlanArgs A) { d_plan_rows(A); }
__global__ void __launch_bounds__(kScanThreads) k_plan_scan1(const PlanArgs A) { d_plan_scan1(A, (int)blockIdx.y); }
__global__ void __launch_bounds__(256) k_assign_pos(const PlanArgs A) { d_assign_pos(A); }
__global__ void __launch_bounds__(kScanThreads) k_plan_scan2(const PlanArgs A) { d_plan_scan2(A, (int)blockIdx.y); }
__global__ void __launch_bounds__(256) k_resolve(const PlanArgs A) { d_resolve(A); }

// All phases in one cooperative launch (every block resident, grid-wide barriers between phases).
__global__ void __launch_bounds__(256) k_plan_fused(const PlanArgs A, int do_resolve)
{
    cooperative_groups::grid_group grid = cooperative_groups::this_grid();
    d_plan_rows(A);
    grid.sync();
    d_plan_scan1(A, (int)(blockIdx.x & 1u));          // the two rank scans side by side
    grid.sync();
    d_assign_pos(A);
    grid.sync();
    d_plan_scan2(A, (int)(blockIdx.x % 3u));          // the three offset scans side by side
    if (do_resolve) {
        grid.sync();
        d_resolve(A);
    }
}

}  // namespace
}  // namespace cds
