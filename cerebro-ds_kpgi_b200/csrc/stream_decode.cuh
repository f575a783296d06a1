// stream_decode.cuh — kernels of the streamed cds_decode_host: the decode of a segment overlapped with its own
// host -> HBM copy (the reference reads and parses page by page, /root/reference/cerebro_gpdb/pg_page_reader.py:358-379).
//
// The classic decode needs the whole TOAST relation before it can plan: the result order is the order in which the
// physical walk first meets a row's values (ppr:553-611), and the sizes of compressed values are read from their
// first chunk.  The streamed decode plans from the heap alone, which is on the device first and is tiny, under two
// assumptions that hold for every table a loader writes: (1) external values appear in the TOAST relation in table
// order, so rows enter the result in table order behind the rows with an inline dependent_var; (2) a value's decoded
// size is its pointer's va_rawsize - 4.  With the plan known, every range of TOAST pages is indexed, walked, checked,
// gathered and its completed values decompressed as soon as the range has arrived, while the next range is still on
// the PCIe bus.  At the end the classic plan is computed over the complete chunk directory (tens of microseconds)
// and compared with the assumed one, row by row; any difference, and any error report at all, makes cds_finish
// decode the segment again the classic way from the pages that are all on the device by then — so results, error
// codes and their order are exactly those of the classic path, only slower in the cases that are wrong anyway.
#pragma once
#include "page_scan.cuh"

namespace cds {
namespace {

constexpr uint32_t kMaxStreamRanges = 64;

struct StreamState {             // zero when a decode starts
    uint32_t chunks_so_far;      // TOAST tuples of the ranges indexed so far
    uint32_t pad[15];
};

struct StreamArgs {
    uint32_t *ext_flag;          // [rows + 1] rows with an external dependent_var -> their rank in table order
    uint32_t *o_pos_of_row;      // [rows] the assumed plan, kept for the comparison at the end
    unsigned long long *o_ind_len, *o_dep_len;
    uint32_t *vjob;              // [2 * rows] the value's pglz job has been issued (zero when a decode starts)
    uint32_t *range_first;       // [ranges + 1] first job slot of each batch of jobs (batch 0: the inline datums)
    uint32_t *range_ticket;      // [ranges + 1] work counters of the decoder launches (zero when a decode starts)
    ScanWs ws[5];                // inline rank, external rank, independent_var / dependent_var / staging offsets
    StreamState *ss;
};

__device__ __forceinline__ bool stream_fits(const PlanArgs &A)
{
    const DevState *st = A.st;
    const uint32_t half = A.ind_bf16 ? 1u : 0u, quad = A.ind_u8norm ? 2u : 0u;
    return ((st->ind_bytes >> half) << quad) <= A.cap_ind && st->dep_bytes <= A.cap_dep && st->scratch_bytes <= A.cap_scratch;
}

// Per row, from the heap tuple alone: decoded sizes (va_rawsize - 4; the rawsize word of an inline datum), which rows have
// an inline dependent_var, staging bytes of compressed external values.
__global__ void __launch_bounds__(256) k_stream_rows(const PlanArgs A, const StreamArgs S)
{
    const uint32_t n_rows = A.st->n_rows;
    for (uint32_t r = blockIdx.x * blockDim.x + threadIdx.x; r < n_rows; r += gridDim.x * blockDim.x) {
        const uint32_t *row = A.rows + (size_t)r * kRowWords;
        const bool usable = row[RW_INLINE_LEN] != kNone;
        const bool is_inline = usable && row[RW_DEP_EXTERNAL] == 0;
        unsigned long long len[2] = {0, 0};
        for (uint32_t a = 0; a < 2; a++) {
            const uint32_t v = 2 * r + a;
            A.vscr[v] = 0;
            S.vjob[v] = 0;
            if (!usable) continue;
            const uint32_t *ptr = row + (a ? RW_DEP : RW_IND);
            if (a == 1 && is_inline) { len[1] = ptr[1]; continue; }          // ppr:538
            const uint32_t va_rawsize = ptr[1], va_extsize = ptr[2];
            len[a] = va_rawsize >= 4u ? va_rawsize - 4u : 0u;
            if (va_extsize != va_rawsize - 4u) {                             // compressed
                if (va_extsize >= 4u && (a == 1u || !(A.ind_bf16 | A.ind_u8norm)) && pglz_literal_only(va_extsize - 4u, len[a])) continue;
                A.vscr[v] = ((unsigned long long)va_extsize + 15ull) & ~15ull;
            }
        }
        A.ind_len[r] = len[0];
        A.dep_len[r] = len[1];
        S.o_ind_len[r] = len[0];
        S.o_dep_len[r] = len[1];
        A.inline_flag[r] = is_inline ? 1u : 0u;
        S.ext_flag[r] = (usable && !is_inline) ? 1u : 0u;
        A.row_first[r] = kNone;
    }
}

__global__ void __launch_bounds__(kScanThreads) k_stream_scan1(const PlanArgs A, const StreamArgs S)
{
    bool last;
    if (blockIdx.y == 0) {
        uint32_t *f = A.inline_flag;
        const uint32_t tot = device_exclusive_scan<uint32_t, 8>([&](uint32_t i) { return f[i]; }, f, A.st->n_rows, S.ws[0], false, &last);
        if (last && threadIdx.x == 0) A.st->n_inline = tot;
    } else {
        uint32_t *f = S.ext_flag;
        device_exclusive_scan<uint32_t, 8>([&](uint32_t i) { return f[i]; }, f, A.st->n_rows, S.ws[1], true, &last);
    }
}

// Result positions: rows with an inline dependent_var in table order (ppr:532-552), then the others, assumed to be met by
// the TOAST walk in table order too.
__global__ void __launch_bounds__(256) k_stream_assign(const PlanArgs A, const StreamArgs S)
{
    DevState *st = A.st;
    const uint32_t n_rows = st->n_rows, n_inline = st->n_inline;
    if (blockIdx.x == 0 && threadIdx.x == 0) st->n_pos = n_inline + S.ext_flag[n_rows];
    for (uint32_t r = blockIdx.x * blockDim.x + threadIdx.x; r < n_rows; r += gridDim.x * blockDim.x) {
        const uint32_t *row = A.rows + (size_t)r * kRowWords;
        const bool usable = row[RW_INLINE_LEN] != kNone;
        const bool is_inline = usable && row[RW_DEP_EXTERNAL] == 0;
        const uint32_t pos = is_inline ? A.inline_flag[r] : usable ? n_inline + S.ext_flag[r] : kNone;
        A.pos_of_row[r] = pos;
        S.o_pos_of_row[r] = pos;
        if (pos == kNone) continue;
        A.row_at_pos[pos] = r;
        A.pos_ind[pos] = A.ind_len[r];
        A.pos_dep[pos] = A.dep_len[r];
    }
}

__global__ void __launch_bounds__(kScanThreads) k_stream_scan2(const PlanArgs A, const StreamArgs S)
{
    bool last;
    DevState *st = A.st;
    if (blockIdx.y == 0) {
        unsigned long long *x = A.pos_ind;
        const unsigned long long t = device_exclusive_scan<unsigned long long, 8>([&](uint32_t i) { return x[i]; }, x, st->n_pos, S.ws[2], false, &last);
        if (last && threadIdx.x == 0) st->ind_bytes = t;
    } else if (blockIdx.y == 1) {
        unsigned long long *x = A.pos_dep;
        const unsigned long long t = device_exclusive_scan<unsigned long long, 8>([&](uint32_t i) { return x[i]; }, x, st->n_pos, S.ws[3], false, &last);
        if (last && threadIdx.x == 0) st->dep_bytes = t;
    } else {
        unsigned long long *x = A.vscr;
        const unsigned long long t = device_exclusive_scan<unsigned long long, 8>([&](uint32_t i) { return x[i]; }, x, 2u * st->n_rows, S.ws[4], false, &last);
        if (last && threadIdx.x == 0) st->scratch_bytes = t;
    }
}

// The next batch of jobs starts at the current end of the job list.
__global__ void k_stream_mark(const DevState *st, uint32_t *first) { *first = st->n_jobs_front; }

// Jobs of the inline datums (their source is the heap, which is complete).
__global__ void __launch_bounds__(256) k_stream_inline_jobs(const PlanArgs A, const StreamArgs S)
{
    DevState *st = A.st;
    if (!stream_fits(A)) return;
    const uint32_t n_rows = st->n_rows;
    for (uint32_t r = blockIdx.x * blockDim.x + threadIdx.x; r < n_rows; r += gridDim.x * blockDim.x) {
        const uint32_t *row = A.rows + (size_t)r * kRowWords;
        if (row[RW_INLINE_LEN] == kNone || row[RW_DEP_EXTERNAL] != 0) continue;
        const uint32_t pos = A.pos_of_row[r];
        if (pos == kNone) continue;
        const PglzJob job = make_value_job(A, 2u * r + 1u, pos);
        if (job.slen != kNone) A.jobs[atomicAdd(&st->n_jobs_front, 1u)] = job;
    }
}

// After the page index of one range (a scan that starts at zero): make its item bases global.
__global__ void __launch_bounds__(256) k_stream_fixup(uint32_t *base, uint32_t n_pages, DevState *st, StreamState *ss, uint32_t cap_chunks)
{
    const uint32_t carry = ss->chunks_so_far;
    const uint32_t local_total = base[n_pages];
    __syncthreads();                                                         // (one block: everybody has read before anybody writes)
    for (uint32_t i = threadIdx.x; i <= n_pages; i += blockDim.x) base[i] += carry;
    if (threadIdx.x == 0) {
        const unsigned long long total = (unsigned long long)carry + local_total;
        if (total > cap_chunks) {
            report(st, 1, 0, 0, 0, CDS_ERR_CAPACITY);
            atomicOr(&st->flags, kFlagChunkOverflow);
        }
        ss->chunks_so_far = (uint32_t)total;
        st->n_chunks = (st->flags & kFlagChunkOverflow) ? 0u : (uint32_t)total;
    }
}

// One thread per chunk of the range: the checks that need nothing but the chunk and its pointer (ppr:580-591), the
// copy descriptor, and — when this chunk completes its value — the value's pglz job.  The checks over whole values
// (chunk count, total size, repeated sequence numbers: ppr:570-572, 595-596) run at the end, over the complete directory.
__global__ void __launch_bounds__(256) k_stream_resolve(const PlanArgs A, const StreamArgs S, const uint32_t *base_lo, const uint32_t *base_hi)
{
    DevState *st = A.st;
    const bool fits = stream_fits(A);
    if (!fits && blockIdx.x == 0 && threadIdx.x == 0) report(st, 4, 0, 2, 0, CDS_ERR_CAPACITY);
    if (st->flags & kFlagChunkOverflow) return;
    const uint32_t g_lo = *base_lo, g_hi = *base_hi;
    for (uint32_t g = g_lo + blockIdx.x * blockDim.x + threadIdx.x; g < g_hi; g += gridDim.x * blockDim.x) {
        CopyOp op;
        op.dst = 0; op.src_off = 0; op.len = 0;
        const uint32_t ref = A.dir_ref[g];
        if (ref != kNone) {
            const uint32_t v = ref & 0x3FFFFFFFu, r = v >> 1, a = v & 1u;
            const uint32_t *ptr = A.rows + (size_t)r * kRowWords + (a ? RW_DEP : RW_IND);
            const uint32_t va_rawsize = ptr[1], va_extsize = ptr[2];
            const uint32_t numchunks = va_extsize ? (va_extsize - 1u) / kChunkMax + 1u : 0u;   // ppr:155-156
            const uint2 id = A.dir_id[g], loc = A.dir_loc[g];
            const uint32_t idx = id.y, size = loc.y >> 16;
            bool ok = !(ref & 0x40000000u) && idx < numchunks;
            if (ok && idx < numchunks - 1u && size != kChunkMax) ok = false;
            if (ok && idx == numchunks - 1u && (unsigned long long)idx * kChunkMax + size != va_extsize) ok = false;
            const uint32_t pos = A.pos_of_row[r];
            if (!ok) atomicOr(&st->flags, kFlagStreamRedo);                 // the classic pass names the error
            if (ok && fits && pos != kNone) {
                const unsigned long long rawlen = a ? A.dep_len[r] : A.ind_len[r];
                op = make_copy_op(A, v, pos, idx, size, loc, va_rawsize, va_extsize, rawlen);
                if (va_extsize != va_rawsize - 4u && A.vcount[v] == numchunks && atomicCAS(&S.vjob[v], 0u, 1u) == 0u) {
                    const PglzJob job = make_value_job(A, v, pos);           // (none for a literal-only value)
                    if (job.slen != kNone) A.jobs[atomicAdd(&st->n_jobs_front, 1u)] = job;
                }
            }
        }
        A.ops[g] = op;
    }
}

// Before the classic plan runs over the complete directory: it rebuilds the job list from slot 0.
__global__ void k_stream_prefinal(DevState *st) { st->n_jobs_front = 0; st->n_jobs_back = 0; }

// The assumed plan against the classic one, row by row.
__global__ void __launch_bounds__(256) k_stream_compare(const PlanArgs A, const StreamArgs S)
{
    DevState *st = A.st;
    const uint32_t n_rows = st->n_rows;
    bool differ = blockIdx.x == 0 && threadIdx.x == 0 && st->err_inv != 0ull;           // any report at all
    for (uint32_t r = blockIdx.x * blockDim.x + threadIdx.x; r < n_rows; r += gridDim.x * blockDim.x)
        differ = differ || A.pos_of_row[r] != S.o_pos_of_row[r] || A.ind_len[r] != S.o_ind_len[r] || A.dep_len[r] != S.o_dep_len[r];
    if (differ) atomicOr(&st->flags, kFlagStreamRedo);
}

}  // namespace
}  // namespace cds
