// pglz_wide.cuh — intra-datum parallel pglz decode for few, huge values (the paper's packed
// buffers, B = 3210 images = 483 MB per value, /root/reference/cerebro_gpdb/load_imagenet.py:27-31,118-135).
//
// Same algorithm as pglz_decompress_py (/root/reference/cerebro_gpdb/pg_page_reader.py:191-231) and
// pglz_decompress (/root/reference/cerebro_gpdb/pg_lzcompress.c:49-143), same accept/reject rule
// (ppr:229-230), but one value is decoded by the whole GPU instead of by one warp (pglz_decode.cuh).
//
// The serial chain of pglz has two parts, and both are broken exactly (no speculation):
//
//  A. Where do the control groups start?  A group is 1 control byte + 8 items of 1..3 bytes, so where a
//     group ends depends on its bytes, and where the next one starts depends on that.  The stream is cut
//     into 512-byte source segments.  The first group start at or after a segment boundary lies within
//     25 bytes of it, so there are only 25 possible entries into a segment.  k_wide_maps computes the
//     group length at EVERY source position, from registers only (a length depends on the control byte
//     and on one bit per following byte), then 25 lanes of a warp chase the 25 candidate chains through
//     the segment: entry -> exit offset into the next segment.  These 25-entry maps compose
//     associatively: k_wide_super folds 256 of them, one warp walks the folded maps of a value
//     (k_wide_chain: the chain must end exactly at srcend), k_wide_entries unfolds them into the true
//     entry of every segment.  k_wide_outs walks each segment's groups once: their output bytes and, per
//     32-byte sub-segment, where its first group starts and which output byte that group produces;
//     k_wide_scan / k_wide_check prefix-sum the segments (the total must be rawsize); k_wide_blkmap finds
//     the sub-segment each output block starts in.
//
//  B. What does a match copy?  Output is cut into blocks of 49152 bytes.  A block is rebuilt in shared
//     memory as a forest: every output byte points at the byte it copies (a literal points at itself;
//     byte k of a match points at start - off + (k mod off), i.e. always before its own item), offsets
//     are <= 4095 so pointers leave the block only into the last 4095 bytes of the previous block.
//     Pointer jumping (ptr[i] = ptr[ptr[ptr[i]]], in place) resolves every byte to a literal of the block
//     or to one of those 4095 previous bytes in O(log depth) rounds.  What a block copies from the previous
//     one is only known once that one is known, and so on down the value: a chain of nblk - 1 gathers of
//     4095 bytes, k_wide_tails<0,1,2> (folded over groups of 32 blocks, so the serial part is short).
//     Around it, one of two shapes:
//       * with a store of resolved blocks (2 bytes per output byte, WideArea::store): k_wide_blocks<2>
//         parses and resolves every block once and stores each byte as "value" or "index into the previous
//         tail"; the tail kernels turn the last 4095 entries of every stored block into values, in place;
//         k_wide_finish streams the store into the output (HBM-bound: 2 bytes read per byte written);
//       * without it (no room, or more blocks than it holds): k_wide_blocks<1> parses every block with the
//         previous tail unknown and publishes only its own last 4095 bytes as such a map; after the tail
//         kernels pass 3 (k_wide_blocks<3>) parses the block again with the previous tail known and writes
//         the output, 128 contiguous bytes per warp store.
//
// Work memory: 4 bytes per 32 bytes of source + 13 bytes per 512-byte segment + 12 bytes per output block (+ the
// optional store).  The
// segment maps, and in the two-parse shape the tail maps and the resolved tails, live inside the value's own
// (not yet written) output region and are overwritten by the last kernel; pass 3 therefore runs from the last
// block to the first behind a flag per block.
#pragma once
#include <cuda_bf16.h>

#include "cds_common.cuh"

namespace cds {
namespace {   // internal linkage: with relocatable device code (-rdc) the optimiser then treats the file as a whole program

constexpr uint32_t kWideSeg = 512;                 // source bytes per segment
constexpr uint32_t kWideChunkSegs = 8;             // segments per k_wide_maps step (one per warp)
constexpr uint32_t kWideChunk = kWideSeg * kWideChunkSegs;
constexpr uint32_t kWideSuper = 256;               // segments per folded map
constexpr uint32_t kWideSub = 32;                  // source bytes per sub-segment (one emitting thread each; > 25, so a group starts in each)
constexpr uint32_t kWideSubs = kWideSeg / kWideSub;
constexpr uint32_t kWideWin = 4095;                // largest pglz offset
constexpr uint32_t kWideBlock = 12u * 4096u;       // output bytes per block (at pglz ratios up to 0.66 one block is fed by at most 1024
                                                   // sub-segments: one emission round of the CTA; 14 * 4096 left a second, nearly empty round)
constexpr uint32_t kWideIdx = kWideBlock + 4096u;  // index space of a block: [previous tail | block], rounded up
constexpr uint32_t kWideThreads = 1024;
constexpr uint32_t kWidePer = kWideIdx / kWideThreads;   // entries per thread
static_assert(kWideIdx % kWideThreads == 0 && kWidePer <= 64 && kWideIdx < 65535, "block geometry");
constexpr uint32_t kWideEnd = 30, kWideBad = 31;   // map states besides the entry offsets 0..24
constexpr uint32_t kWideChase = 48;                // pass 1 follows a tail byte's chain this many hops before resolving the whole block
constexpr uint32_t kWideShort = 16;                // (cut) matches up to this long are written byte by byte by the emitting thread (the fill
                                                   // pass below costs ~10x more per byte; only long matches, whose bytes would serialise one thread, go there)
constexpr uint32_t kWideMinRaw = 1024;             // below this the maps would not fit the output region
static_assert(kWideShort == 16, "the emission writes short matches as two predicated groups of 8");
constexpr size_t kWideBlockSmem = (size_t)kWideIdx * 3 + kWideIdx / 8;

struct __align__(16) WideJob {
    unsigned long long src, dst, err_key;
    uint32_t slen, rawsize;
    uint32_t job;        // slot in the job array (status word)
    uint32_t seg0;       // first flat segment (multiple of kWideSuper)
    uint32_t blk0;       // first flat output block
    uint32_t nseg, nblk;
    uint32_t state;      // 0 well-formed so far, 1 corrupt, 2 not taken (capacity)
    uint32_t fmt;        // output: 0 the decoded bytes, 1 float32 -> bfloat16, 2 uint8 -> float32 through the table
};

// Output address of decoded byte x of a value (x a multiple of 4 for fmt 1).
__host__ __device__ __forceinline__ size_t wide_scale(uint32_t fmt, size_t x) { return fmt == 1u ? x >> 1 : fmt == 2u ? x << 2 : x; }


struct WideState {       // zero when a decode starts
    uint32_t n_jobs, n_segs, n_blks;
    uint32_t t_maps, t_l1, t_l3, t_scan;
    uint32_t n_taken;    // values this decoder accepted (the rest stay with the warp decoder)
    uint32_t t_l2;
    uint32_t pad[7];
};

struct WideArea {
    WideState *state;        // followed by the scan's status words and the block flags; `zero_bytes` from here are cleared per decode
    uint32_t *scan_status;
    uint32_t *blk_flag;      // per flat block: pass 3 has read the previous block's tail
    unsigned long long *scan_agg, *scan_incl;
    WideJob *jobs;
    uint8_t *seg_e;          // per flat segment: entry offset of its first group (0..24; 30: the stream ended before it)
    uint32_t *seg_out;       // per flat segment: output bytes of the groups that start in it
    unsigned long long *seg_base;   // exclusive prefix of seg_out over all flat segments (+ total)
    uint8_t *supermap;       // per flat super-segment: 32 entries
    uint8_t *sseg_e;         // per flat super-segment: entry offset
    uint32_t *sub_entry;     // per flat sub-segment: (output position of its first group relative to the segment's) << 5 | its offset
                             // (31: no group starts here)
    uint32_t *blk_sub0;      // per flat block: local sub-segment holding the group that produces its first byte
    uint32_t *blk_job;       // per flat block: slot of its value in `jobs`
    uint32_t cap_jobs, cap_segs, cap_blks;
    uint32_t zero_bytes;
    uint32_t threshold;      // values with rawsize >= threshold are candidates for this path
    uint32_t force;          // 1: take every candidate; 0: only when the cost model below says it pays
    int32_t *status;         // per job slot (may be null)
    DevState *st;            // first-error word (may be null)
    const float *lut;        // fmt 2: [channels][256]
    uint32_t lut_channels;
    // Optional store of resolved blocks (not part of the carved area; the caller owns it): kWideBlock uint16 entries per flat
    // output block, each "0x8000 | byte" or "index into the tail of the previous block".  With it a block is parsed once
    // (k_wide_blocks<2>) and written by a streaming kernel (k_wide_finish); without it, or when a decode has more blocks
    // than the store holds, the block is parsed twice (k_wide_blocks<1>, <3>).
    uint16_t *store;
    uint32_t store_blks;
};

__device__ __forceinline__ bool wide_stored(const WideArea &wa) { return wa.store != nullptr && wa.state->n_blks <= wa.store_blks; }

__host__ __device__ inline size_t wide_align(size_t x) { return (x + 255) & ~(size_t)255; }
__host__ __device__ inline size_t wide_scan_tiles(uint32_t cap_segs) { return (size_t)cap_segs / (kScanThreads * 8) + 2; }
__host__ __device__ inline size_t wide_area_bytes(uint32_t cap_jobs, uint32_t cap_segs, uint32_t cap_blks)
{
    const size_t tiles = wide_scan_tiles(cap_segs), nss = (size_t)cap_segs / kWideSuper + 1;
    return wide_align(256 + 4 * tiles + 4 * ((size_t)cap_blks + 1)) + wide_align(16 * tiles) + wide_align(sizeof(WideJob) * (size_t)cap_jobs) +
           wide_align((size_t)cap_segs) + wide_align(4 * (size_t)cap_segs) + wide_align(8 * ((size_t)cap_segs + 1)) +
           wide_align(32 * nss) + wide_align(nss) + wide_align(4 * (size_t)cap_segs * kWideSubs) + 2 * wide_align(4 * ((size_t)cap_blks + 1));
}

// Carve the work area out of one 256-byte aligned allocation (host side).
inline WideArea wide_area_carve(uint8_t *base, uint32_t cap_jobs, uint32_t cap_segs, uint32_t cap_blks)
{
    WideArea wa;
    cap_segs -= cap_segs % kWideSuper;
    const size_t tiles = wide_scan_tiles(cap_segs), nss = (size_t)cap_segs / kWideSuper + 1;
    uint8_t *p = base;
    wa.state = reinterpret_cast<WideState *>(p);
    wa.scan_status = reinterpret_cast<uint32_t *>(p + 256);
    wa.blk_flag = wa.scan_status + tiles;
    wa.zero_bytes = (uint32_t)wide_align(256 + 4 * tiles + 4 * ((size_t)cap_blks + 1));
    p += wa.zero_bytes;
    wa.scan_agg = reinterpret_cast<unsigned long long *>(p);
    wa.scan_incl = wa.scan_agg + tiles; p += wide_align(16 * tiles);
    wa.jobs = reinterpret_cast<WideJob *>(p); p += wide_align(sizeof(WideJob) * (size_t)cap_jobs);
    wa.seg_e = p; p += wide_align((size_t)cap_segs);
    wa.seg_out = reinterpret_cast<uint32_t *>(p); p += wide_align(4 * (size_t)cap_segs);
    wa.seg_base = reinterpret_cast<unsigned long long *>(p); p += wide_align(8 * ((size_t)cap_segs + 1));
    wa.supermap = p; p += wide_align(32 * nss);
    wa.sseg_e = p; p += wide_align(nss);
    wa.sub_entry = reinterpret_cast<uint32_t *>(p); p += wide_align(4 * (size_t)cap_segs * kWideSubs);
    wa.blk_sub0 = reinterpret_cast<uint32_t *>(p); p += wide_align(4 * ((size_t)cap_blks + 1));
    wa.blk_job = reinterpret_cast<uint32_t *>(p);
    wa.cap_jobs = cap_jobs; wa.cap_segs = cap_segs; wa.cap_blks = cap_blks;
    wa.threshold = 0; wa.force = 0; wa.status = nullptr; wa.st = nullptr; wa.lut = nullptr; wa.lut_channels = 1;
    wa.store = nullptr; wa.store_blks = 0;
    return wa;
}

__device__ inline void wide_fail(const WideArea &wa, uint32_t j)
{
    WideJob *wj = &wa.jobs[j];
    if (atomicExch(&wj->state, 1u) == 0u) {
        if (wa.status) wa.status[wj->job] = CDS_ERR_PGLZ_CORRUPT;
        if (wa.st && wj->err_key) report_key(wa.st, wj->err_key);
    }
}

// Last job whose first segment / block is <= key.
__device__ inline uint32_t wide_find_seg(const WideArea &wa, uint32_t n, uint32_t key)
{
    uint32_t lo = 0, hi = n;
    while (hi - lo > 1) {
        const uint32_t mid = (lo + hi) >> 1;
        if (wa.jobs[mid].seg0 <= key) lo = mid; else hi = mid;
    }
    return lo;
}
__device__ inline uint32_t wide_find_blk(const WideArea &wa, uint32_t n, uint32_t key)
{
    uint32_t lo = 0, hi = n;
    while (hi - lo > 1) {
        const uint32_t mid = (lo + hi) >> 1;
        if (wa.jobs[mid].blk0 <= key) lo = mid; else hi = mid;
    }
    return lo;
}

// ---- classification: which jobs take this path -----------------------------------------------------
// One block walks the job list in order (two-ended list of k_resolve when st_counts is given) and appends
// the large values to the wide list with their flat segment / block ranges; taken jobs are blanked in the
// job array so that the warp decoder skips them.  A value also needs slen <= 9/8 rawsize + 9 (anything
// longer cannot be well-formed and is left to the warp decoder to refuse).
//
// Whether the wide path pays is decided for the whole list from measured rates (profiles/README.md): one warp
// decodes about 0.06 GB/s of ordinary data and about 0.6 GB/s of run-dominated data (stream under 1/8 of the
// output: sparse Criteo features), and every value has a warp of its own up to ~4000 values; the wide decoder
// produces about 55 GB/s in total however few values there are.  So few huge values (the paper's 483 MB
// buffers) go wide, a thousand 7 MB Criteo buffers stay with their warps.
constexpr unsigned long long kWarpUnitsDense = 10, kWarpUnitsSparse = 1;    // cost per byte, unit = 1 / 0.6 GB/s
constexpr unsigned long long kWideBytesPerUnit = 92;                        // 55 GB/s in the same unit
constexpr unsigned long long kWarpsResident = 4144;                         // 148 SMs x 7 CTAs x 4 warps

__device__ __forceinline__ bool wide_candidate(const PglzJob &job, uint32_t threshold)
{
    return job.slen != kNone && job.rawsize >= threshold && job.rawsize >= kWideMinRaw &&
           (unsigned long long)(job.slen & ~kFlagBits) <= (unsigned long long)job.rawsize * 9ull / 8ull + 9ull;
}

// (A block of 1024 threads; every thread of the block must call it.)
__device__ __forceinline__ void d_wide_classify(PglzJob *jobs, const DevState *st_counts, uint32_t n_jobs_host, const WideArea &wa,
                                                const uint32_t *d_first, const uint32_t *d_end)
{
    __shared__ uint32_t sh_w[3][32];
    __shared__ uint32_t sh_base[3];
    const uint32_t first = d_first ? *d_first : 0u;                          // streamed decode: slots [*d_first, *d_end)
    const uint32_t front = (d_end ? *d_end : st_counts ? st_counts->n_jobs_front : n_jobs_host) - first;
    const uint32_t back = (st_counts && !d_first) ? st_counts->n_jobs_back : 0u;
    const uint32_t last_slot = st_counts ? 2u * st_counts->n_rows - 1u : 0u;
    const uint32_t n_jobs = front + back, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x < 3) sh_base[threadIdx.x] = 0;
    __shared__ unsigned long long sh_sum_raw, sh_sum_units, sh_max_units;
    if (threadIdx.x == 0) { sh_sum_raw = 0; sh_sum_units = 0; sh_max_units = 0; }
    __syncthreads();
    if (!wa.force) {
        unsigned long long sum_raw = 0, sum_units = 0, max_units = 0;
        for (uint32_t t = threadIdx.x; t < n_jobs; t += blockDim.x) {
            const PglzJob job = jobs[t < front ? first + t : last_slot - (t - front)];
            if (!wide_candidate(job, wa.threshold)) continue;
            const unsigned long long units = (unsigned long long)job.rawsize *
                                             ((unsigned long long)(job.slen & ~kFlagBits) * 8ull < job.rawsize ? kWarpUnitsSparse : kWarpUnitsDense);
            sum_raw += job.rawsize;
            sum_units += units;
            max_units = max(max_units, units);
        }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            sum_raw += __shfl_xor_sync(0xffffffffu, sum_raw, d);
            sum_units += __shfl_xor_sync(0xffffffffu, sum_units, d);
            max_units = max(max_units, __shfl_xor_sync(0xffffffffu, max_units, d));
        }
        if (lane == 0 && sum_raw) {
            atomicAdd(&sh_sum_raw, sum_raw);
            atomicAdd(&sh_sum_units, sum_units);
            atomicMax(&sh_max_units, max_units);
        }
        __syncthreads();
        const unsigned long long warp_units = max(sh_max_units, sh_sum_units / kWarpsResident);
        if (sh_sum_raw / kWideBytesPerUnit >= warp_units) {                   // the warps are at least as fast: nothing goes wide
            if (threadIdx.x == 0) { wa.state->n_jobs = 0; wa.state->n_segs = 0; wa.state->n_blks = 0; }
            return;
        }
    }
    for (uint32_t base = 0; base < n_jobs; base += blockDim.x) {
        const uint32_t t = base + threadIdx.x;
        const uint32_t j = t < front ? first + t : last_slot - (t - front);
        PglzJob job;
        job.slen = kNone;
        if (t < n_jobs) job = jobs[j];
        bool big = t < n_jobs && wide_candidate(job, wa.threshold);
        uint32_t nseg = 0, nblk = 0;
        const uint32_t slen = job.slen & ~kFlagBits;
        if (big) {
            nseg = slen ? (slen + kWideSeg - 1) / kWideSeg : 1u;
            nblk = (job.rawsize + kWideBlock - 1) / kWideBlock;
        }
        const uint32_t pseg = (nseg + kWideSuper - 1) / kWideSuper * kWideSuper;
        uint32_t v[3] = {big ? 1u : 0u, pseg, nblk}, inc[3];
#pragma unroll
        for (int q = 0; q < 3; q++) {
            inc[q] = warp_inclusive_scan(v[q]);
            if (lane == 31) sh_w[q][warp] = inc[q];
        }
        __syncthreads();
        uint32_t ex[3], tot[3];
#pragma unroll
        for (int q = 0; q < 3; q++) {
            uint32_t off = 0, total = 0;
            for (uint32_t w = 0; w < blockDim.x / 32; w++) {
                const uint32_t x = sh_w[q][w];
                if (w < warp) off += x;
                total += x;
            }
            ex[q] = sh_base[q] + off + inc[q] - v[q];
            tot[q] = total;
        }
        if (big) {
            const bool fits = ex[0] < wa.cap_jobs && (unsigned long long)ex[1] + pseg <= wa.cap_segs &&
                              (unsigned long long)ex[2] + nblk <= wa.cap_blks;
            if (ex[0] < wa.cap_jobs) {
                WideJob wj;
                wj.src = job.src; wj.dst = job.dst; wj.err_key = job.err_key;
                wj.slen = slen; wj.rawsize = job.rawsize; wj.job = j;
                wj.fmt = (job.slen & kJobToBf16) ? 1u : (job.slen & kJobU8Norm) ? 2u : 0u;
                wj.seg0 = ex[1]; wj.blk0 = ex[2]; wj.nseg = nseg; wj.nblk = nblk;
                wj.state = fits ? 0u : 2u;
                wa.jobs[ex[0]] = wj;
            }
            if (fits) { jobs[j].slen = kNone; atomicAdd(&wa.state->n_taken, 1u); }
        }
        __syncthreads();
        if (threadIdx.x < 3) sh_base[threadIdx.x] += tot[threadIdx.x];
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        wa.state->n_jobs = min(sh_base[0], wa.cap_jobs);
        wa.state->n_segs = min(sh_base[1], wa.cap_segs);                     // values that fit form a prefix of the list
        wa.state->n_blks = min(sh_base[2], wa.cap_blks);
    }
}

// The WideState header is cleared here (the rest of the zeroed area only matters to the kernels that follow).
// deferred != 0: those kernels were not enqueued with this decode; tell cds_finish when they are needed after all.
__global__ void __launch_bounds__(1024)
k_wide_classify(PglzJob *jobs, DevState *st_counts, uint32_t n_jobs_host, WideArea wa, int deferred, const uint32_t *d_first,
                const uint32_t *d_end)
{
    if (st_counts) {                                       // (the batch entry points clear the header with a memset)
        if (threadIdx.x < sizeof(WideState) / 4u) reinterpret_cast<uint32_t *>(wa.state)[threadIdx.x] = 0u;
        __syncthreads();
    }
    d_wide_classify(jobs, st_counts, n_jobs_host, wa, d_first, d_end);
    __syncthreads();
    if (threadIdx.x == 0 && st_counts) {
        if (wa.state->n_taken) atomicAdd(&st_counts->n_wide_taken, wa.state->n_taken);
        atomicMax(&st_counts->n_wide_blks, wa.state->n_blks);
        if (deferred && wa.state->n_jobs != 0u) atomicOr(&st_counts->flags, kFlagNeedWide);
    }
}

// Clears the scan status words and the block flags behind the WideState header (the device-launched pipeline:
// the header itself is cleared by the gate kernel before it classifies).
__global__ void __launch_bounds__(256)
k_wide_zero(WideArea wa)
{
    uint4 *p = reinterpret_cast<uint4 *>(reinterpret_cast<uint8_t *>(wa.state) + 256);
    const uint32_t n = (wa.zero_bytes - 256u) >> 4;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) p[i] = make_uint4(0u, 0u, 0u, 0u);
}

// ---- A: group starts ---------------------------------------------------------------------------------

// Length in bytes of the group that would start `avail` bytes before the end of the stream, given its control
// byte and the "length-extension" bits of the following bytes (bit i of `ext`: byte i after the control byte
// is 0x?F).  0: a tag runs past the source, or the control byte has no item; both end in "corrupt"
// (ppr:229-230).  Register-only: where a group ends depends on nothing but these bits.
// tags[c]: bit i = "byte i after control byte c starts a tag", as long as no earlier tag of the group is extended.
__device__ __forceinline__ uint32_t wide_tag_starts(uint32_t ctrl)
{
    uint32_t m = 0, q = 0;
#pragma unroll
    for (uint32_t k = 0; k < 8; k++) {
        if ((ctrl >> k) & 1u) { m |= 1u << q; q += 2u; } else q += 1u;
    }
    return m;
}

__device__ __forceinline__ uint32_t wide_group_len(uint32_t ctrl, uint32_t ext, uint32_t avail, const uint32_t *tags)
{
    if (avail >= 26u) {                                                      // the whole group lies inside the stream
        // 1 + 8 items + one more byte per tag + one more per extended tag.  Most groups have none of those: the tags'
        // first bytes are where the table says, and none of them has the extension bit.  Otherwise the first one that
        // has it is a real extended tag (nothing before it moved), and every tag after it starts one byte later.
        uint32_t len = 9u + (uint32_t)__popc(ctrl), tm = tags[ctrl], hit = ext & tm;
        while (hit) {
            const uint32_t p = (uint32_t)__ffs((int)hit) - 1u;
            len += 1u;
            tm = (tm & ~((2u << p) - 1u)) << 1;
            hit = ext & tm;
        }
        return len;
    }
    uint32_t q = 1;
#pragma unroll
    for (uint32_t k = 0; k < 8; k++) {
        if (q < avail) {                                                     // `sp < srcend` of the inner loop (ppr:203)
            if ((ctrl >> k) & 1u) {
                const uint32_t sz = 2u + ((ext >> (q - 1u)) & 1u);
                if (q + sz > avail) return 0;
                q += sz;
            } else {
                q += 1u;
            }
        }
    }
    return q == 1u ? 0u : q;
}

__device__ __forceinline__ uint32_t wide_ext_bits(uint32_t w)                // 4 bytes -> 4 bits "(byte & 0x0F) == 0x0F"
{
    const uint32_t t = w & (w >> 1) & (w >> 2) & (w >> 3) & 0x01010101u;
    return (t | (t >> 7) | (t >> 14) | (t >> 21)) & 0xFu;
}

// Per 4 KB of source: the group length at EVERY position (16 positions per thread, from registers), then one
// warp per 512-byte segment chases the 25 candidate entries to their exits.
__global__ void __launch_bounds__(256)
k_wide_maps(WideArea wa)
{
    __shared__ uint16_t sext[kWideChunk / 16 + 8];                           // extension bits, 16 positions per entry
    __shared__ uint8_t sctrl_extra[64];                                      // source bytes 4096..4159 of the chunk
    __shared__ __align__(16) uint8_t sdelta[kWideChunk + 32];
    __shared__ uint32_t sh_c;
    __shared__ uint32_t stags[256];
    const uint32_t n_jobs = wa.state->n_jobs, n_chunks = wa.state->n_segs / kWideChunkSegs;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5, t = threadIdx.x;
    if (n_jobs == 0) return;
    stags[t] = wide_tag_starts(t);                                           // (256 threads)
    for (;;) {
        __syncthreads();
        if (t == 0) sh_c = atomicAdd(&wa.state->t_maps, 1u);
        __syncthreads();
        const uint32_t c = sh_c;
        if (c >= n_chunks) break;
        const uint32_t seg = c * kWideChunkSegs, jn = wide_find_seg(wa, n_jobs, seg);
        const WideJob wj = wa.jobs[jn];
        const uint32_t lseg = seg - wj.seg0;
        if (wj.state != 0u || lseg >= wj.nseg) continue;
        const uint32_t coff = lseg * kWideSeg, slen = wj.slen;
        const uint8_t *src = reinterpret_cast<const uint8_t *>(wj.src);
        // my 16 source bytes (zeros past the end of the stream)
        uint32_t w[4];
        const uint32_t p0 = coff + 16u * t;
        if (p0 + 16u <= slen) {
#pragma unroll
            for (uint32_t j = 0; j < 4; j++) w[j] = ld_u32(src + p0 + 4u * j);
        } else {
#pragma unroll
            for (uint32_t j = 0; j < 4; j++) {
                w[j] = 0;
#pragma unroll
                for (uint32_t k = 0; k < 4; k++)
                    if (p0 + 4u * j + k < slen) w[j] |= (uint32_t)__ldg(src + p0 + 4u * j + k) << (8u * k);
            }
        }
        sext[t] = (uint16_t)(wide_ext_bits(w[0]) | (wide_ext_bits(w[1]) << 4) | (wide_ext_bits(w[2]) << 8) | (wide_ext_bits(w[3]) << 12));
        if (t < 64) {                                                        // the 64 bytes after the chunk
            const uint32_t a = coff + kWideChunk + t;
            const uint32_t bv = a < slen ? __ldg(src + a) : 0u;
            sctrl_extra[t] = (uint8_t)bv;
            const uint32_t bits = __ballot_sync(0xffffffffu, (bv & 0x0Fu) == 0x0Fu);
            if ((t & 31u) == 0u) { sext[256 + (t >> 4)] = (uint16_t)bits; sext[257 + (t >> 4)] = (uint16_t)(bits >> 16); }
        }
        if (t < 4) sext[260 + t] = 0;
        __syncthreads();
        {
            // bits for positions 16t .. 16t+63
            const unsigned long long e = (unsigned long long)sext[t] | ((unsigned long long)sext[t + 1] << 16) |
                                         ((unsigned long long)sext[t + 2] << 32) | ((unsigned long long)sext[t + 3] << 48);
            uint32_t packed[4];
#pragma unroll
            for (uint32_t j = 0; j < 16; j++) {
                const uint32_t a = p0 + j;
                uint32_t d = 0;
                if (a < slen) d = wide_group_len((w[j >> 2] >> (8u * (j & 3u))) & 0xFFu, (uint32_t)(e >> (j + 1u)), slen - a, stags);
                if ((j & 3u) == 0u) packed[j >> 2] = d; else packed[j >> 2] |= d << (8u * (j & 3u));
            }
            *reinterpret_cast<uint4 *>(sdelta + 16u * t) = make_uint4(packed[0], packed[1], packed[2], packed[3]);
        }
        if (t < 32) {                                                        // positions 4096..4127: entries of the next chunk's first segment
            const uint32_t a = coff + kWideChunk + t;
            const unsigned long long e = (unsigned long long)sext[256 + (t >> 4)] | ((unsigned long long)sext[257 + (t >> 4)] << 16) |
                                         ((unsigned long long)sext[258 + (t >> 4)] << 32);
            uint32_t d = 0;
            if (a < slen) d = wide_group_len(sctrl_extra[t], (uint32_t)(e >> ((t & 15u) + 1u)), slen - a, stags);
            sdelta[kWideChunk + t] = (uint8_t)d;
        }
        __syncthreads();
        if (lseg + warp < wj.nseg) {
            uint32_t p = warp * kWideSeg + lane, e;
            const uint32_t end = (warp + 1u) * kWideSeg;
            if (lane >= 25u) {
                e = kWideBad;
            } else {
                // follow the groups while they start inside the segment and inside the stream (a group that cannot be: far away)
                const uint32_t lim = min(end, slen - coff), kFar = 1u << 20;
                while (p < lim) {
                    const uint32_t d = sdelta[p];
                    p += d ? d : kFar;
                }
                if (p >= kFar) e = kWideBad;
                else if (coff + p >= slen) e = coff + p == slen ? kWideEnd : kWideBad;
                else e = p - end;
            }
            reinterpret_cast<uint8_t *>(wj.dst)[(size_t)(lseg + warp) * 32u + lane] = (uint8_t)e;
        }
    }
}

// Fold the maps of kWideSuper consecutive segments: one warp per super-segment, lane = entry candidate.
__global__ void __launch_bounds__(128)
k_wide_super(WideArea wa)
{
    const uint32_t n_jobs = wa.state->n_jobs, n_ss = wa.state->n_segs / kWideSuper;
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t wid = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = (gridDim.x * blockDim.x) >> 5;
    if (n_jobs == 0) return;
    for (uint32_t ss = wid; ss < n_ss; ss += nw) {
        const uint32_t seg = ss * kWideSuper, jn = wide_find_seg(wa, n_jobs, seg);
        const WideJob wj = wa.jobs[jn];
        const uint32_t lseg0 = seg - wj.seg0;
        if (wj.state != 0u || lseg0 >= wj.nseg) continue;
        const uint32_t n = min(kWideSuper, wj.nseg - lseg0);
        const uint8_t *maps = reinterpret_cast<const uint8_t *>(wj.dst) + (size_t)lseg0 * 32u;
        uint32_t e = lane < 25u ? lane : kWideBad;
        for (uint32_t s = 0; s < n; s += 8) {
            uint32_t r[8];
#pragma unroll
            for (uint32_t i = 0; i < 8; i++) r[i] = s + i < n ? maps[(size_t)(s + i) * 32u + lane] : 0u;
#pragma unroll
            for (uint32_t i = 0; i < 8; i++) {
                const uint32_t v = __shfl_sync(0xffffffffu, r[i], e < 25u ? e : 0u);
                if (s + i < n && e < 25u) e = v;
            }
        }
        wa.supermap[(size_t)ss * 32u + lane] = (uint8_t)e;
    }
}

// One warp per value: walk its folded maps from the true start (entry 0), record where every super-segment
// is entered, and decide the first half of ppr:229-230: the chain of groups must end exactly at srcend.
__global__ void __launch_bounds__(32)
k_wide_chain(WideArea wa)
{
    const uint32_t n_jobs = wa.state->n_jobs, lane = threadIdx.x;
    for (uint32_t jn = blockIdx.x; jn < n_jobs; jn += gridDim.x) {
        const WideJob wj = wa.jobs[jn];
        if (wj.state != 0u) continue;
        const uint32_t nss = (wj.nseg + kWideSuper - 1) / kWideSuper, ss0 = wj.seg0 / kWideSuper;
        uint32_t e = 0;
        for (uint32_t s = 0; s < nss; s += 8) {
            uint32_t r[8];
#pragma unroll
            for (uint32_t i = 0; i < 8; i++) r[i] = s + i < nss ? wa.supermap[(size_t)(ss0 + s + i) * 32u + lane] : 0u;
#pragma unroll
            for (uint32_t i = 0; i < 8; i++) {
                if (s + i < nss) {
                    if (lane == 0) wa.sseg_e[ss0 + s + i] = (uint8_t)e;
                    const uint32_t v = __shfl_sync(0xffffffffu, r[i], e < 25u ? e : 0u);
                    if (e < 25u) e = v;
                }
            }
        }
        if (lane == 0 && e != kWideEnd) wide_fail(wa, jn);
    }
}

// Unfold: the true entry offset of every segment.
__global__ void __launch_bounds__(128)
k_wide_entries(WideArea wa)
{
    const uint32_t n_jobs = wa.state->n_jobs, n_ss = wa.state->n_segs / kWideSuper;
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t wid = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = (gridDim.x * blockDim.x) >> 5;
    if (n_jobs == 0) return;
    for (uint32_t ss = wid; ss < n_ss; ss += nw) {
        const uint32_t seg = ss * kWideSuper, jn = wide_find_seg(wa, n_jobs, seg);
        const WideJob wj = wa.jobs[jn];
        const uint32_t lseg0 = seg - wj.seg0;
        if (wj.state != 0u || lseg0 >= wj.nseg) continue;
        const uint32_t n = min(kWideSuper, wj.nseg - lseg0);
        const uint8_t *maps = reinterpret_cast<const uint8_t *>(wj.dst) + (size_t)lseg0 * 32u;
        uint32_t e = wa.sseg_e[ss];
        for (uint32_t s = 0; s < n; s += 8) {
            uint32_t r[8];
#pragma unroll
            for (uint32_t i = 0; i < 8; i++) r[i] = s + i < n ? maps[(size_t)(s + i) * 32u + lane] : 0u;
#pragma unroll
            for (uint32_t i = 0; i < 8; i++) {
                if (s + i < n) {
                    if (lane == 0) wa.seg_e[seg + s + i] = (uint8_t)e;
                    const uint32_t v = __shfl_sync(0xffffffffu, r[i], e < 25u ? e : 0u);
                    if (e < 25u) e = v;
                }
            }
        }
    }
}

// The group at stream offset a of a well-formed stream: its end and its output bytes.
__device__ __forceinline__ uint32_t wide_group_walk(const uint8_t *__restrict__ src, uint32_t a, uint32_t slen, uint32_t *out_bytes)
{
    const uint32_t ctrl = __ldg(src + a);
    uint32_t q = a + 1u, out = 0;
#pragma unroll 1
    for (uint32_t it = 0; it < 8 && q < slen; it++) {
        if ((ctrl >> it) & 1u) {
            const uint32_t b0 = __ldg(src + q);
            const bool ext = (b0 & 0x0Fu) == 0x0Fu;
            out += (b0 & 0x0Fu) + 3u + (ext && q + 2u < slen ? (uint32_t)__ldg(src + q + 2u) : 0u);   // ppr:205,210-212
            q += ext ? 3u : 2u;
        } else {
            out += 1u;
            q += 1u;
        }
    }
    *out_bytes = out;
    return q;
}

// One thread per segment walks its chain of groups: output bytes of the groups that start in it (padding segments
// and values already refused count 0, so one flat prefix sum serves all values), and, for each 32-byte sub-segment,
// where its first group starts and which output byte (relative to the segment) that group produces first.
// Sub-segments are the units the block kernels hand to their threads.
__global__ void __launch_bounds__(256)
k_wide_outs(WideArea wa)
{
    const uint32_t n_jobs = wa.state->n_jobs, n_segs = wa.state->n_segs;
    if (n_jobs == 0) return;
    for (uint32_t seg = blockIdx.x * blockDim.x + threadIdx.x; seg < n_segs; seg += gridDim.x * blockDim.x) {
        const uint32_t jn = wide_find_seg(wa, n_jobs, seg);
        const WideJob *wjp = &wa.jobs[jn];
        const uint32_t lseg = seg - wjp->seg0, slen = wjp->slen;
        uint32_t total = 0;
        if (wjp->state == 0u && lseg < wjp->nseg) {
            const uint32_t e = wa.seg_e[seg];
            uint32_t *fine = wa.sub_entry + (size_t)seg * kWideSubs;
            const uint32_t end = min((lseg + 1u) * kWideSeg, slen);
            uint32_t sub = 0;                                                // next sub-segment of this segment without a record
            if (e < 25u) {
                const uint8_t *src = reinterpret_cast<const uint8_t *>(wjp->src);
                uint32_t a = lseg * kWideSeg + e;
                while (a < end) {
                    const uint32_t my_sub = (a - lseg * kWideSeg) / kWideSub;
                    if (my_sub >= sub) { fine[my_sub] = (total << 5) | (a - (lseg * kWideSeg + my_sub * kWideSub)); sub = my_sub + 1u; }
                    uint32_t out;
                    a = wide_group_walk(src, a, slen, &out);
                    total += out;
                }
            }
            // sub-segments of the stream's end in which no group starts
            const uint32_t nsub_here = (end - lseg * kWideSeg + kWideSub - 1u) / kWideSub;
            for (; sub < nsub_here; sub++) fine[sub] = 31u;
        }
        wa.seg_out[seg] = total;
    }
}

__global__ void __launch_bounds__(kScanThreads)
k_wide_scan(WideArea wa)
{
    bool last;
    const uint32_t *x = wa.seg_out;
    ScanWs ws;
    ws.ticket = &wa.state->t_scan; ws.status = wa.scan_status; ws.agg = wa.scan_agg; ws.incl = wa.scan_incl;
    device_exclusive_scan<unsigned long long, 8>([&](uint32_t i) { return (unsigned long long)x[i]; }, wa.seg_base, wa.state->n_segs, ws, true, &last);
}

// Second half of ppr:229-230: exactly rawsize bytes produced.
__global__ void __launch_bounds__(256)
k_wide_check(WideArea wa)
{
    const uint32_t n_jobs = wa.state->n_jobs;
    for (uint32_t jn = blockIdx.x * blockDim.x + threadIdx.x; jn < n_jobs; jn += gridDim.x * blockDim.x) {
        const WideJob *wjp = &wa.jobs[jn];
        if (wjp->state != 0u) continue;
        const uint32_t pseg = (wjp->nseg + kWideSuper - 1) / kWideSuper * kWideSuper;
        if (wa.seg_base[wjp->seg0 + pseg] - wa.seg_base[wjp->seg0] != (unsigned long long)wjp->rawsize) wide_fail(wa, jn);
    }
}

// One thread per output block: the sub-segment that holds the group producing the block's first byte = the last
// sub-segment whose first group starts at or before that byte (binary search; positions grow with the sub-segments,
// and sub-segments without a group exist only at the very end of a stream).
__global__ void __launch_bounds__(256)
k_wide_blkmap(WideArea wa)
{
    const uint32_t n_jobs = wa.state->n_jobs, n_blks = wa.state->n_blks;
    if (n_jobs == 0) return;
    for (uint32_t b = blockIdx.x * blockDim.x + threadIdx.x; b < n_blks; b += gridDim.x * blockDim.x) {
        const uint32_t jn = wide_find_blk(wa, n_jobs, b);
        wa.blk_job[b] = jn;
        const WideJob *wjp = &wa.jobs[jn];
        const uint32_t k = b - wjp->blk0;
        if (wjp->state != 0u || k >= wjp->nblk) continue;
        const unsigned long long target = (unsigned long long)k * kWideBlock;
        const unsigned long long base0 = wa.seg_base[wjp->seg0];
        const uint32_t *fine = wa.sub_entry + (size_t)wjp->seg0 * kWideSubs;
        const uint32_t nsub = wjp->slen ? (wjp->slen - 1u) / kWideSub + 1u : 0u;
        uint32_t lo = 0, hi = nsub;                                          // sub 0 starts at output 0 <= target
        while (hi - lo > 1u) {
            const uint32_t mid = (lo + hi) >> 1, ent = fine[mid];
            const bool le = (ent & 31u) != 31u &&
                            wa.seg_base[wjp->seg0 + mid / kWideSubs] - base0 + (ent >> 5) <= target;
            if (le) lo = mid; else hi = mid;
        }
        wa.blk_sub0[b] = lo;
    }
}

// ---- B: blocks -----------------------------------------------------------------------------------------


// Nearest long-match head at or below e (within 273 entries of any byte of a long match's body).
__device__ __forceinline__ uint32_t wide_head_of(const uint32_t *heads, uint32_t e)
{
    uint32_t w = e >> 5;
    uint32_t m = heads[w] & (0xFFFFFFFFu >> (31u - (e & 31u)));
    while (m == 0u && w > 0u) m = heads[--w];
    return m ? (w << 5) + 31u - (uint32_t)__clz((int)m) : kNone;
}

// Write bytes [0, n_out) of a block (byte(i) gives them) in the value's output format; `region` is the block's part of
// the output buffer, o0 the block's offset in the decoded value.
template <typename Byte>
__device__ __forceinline__ void wide_write_block(uint32_t fmt, uint8_t *region, uint32_t o0, uint32_t n_out, const float *lut,
                                                 uint32_t lut_channels, uint32_t tid, uint32_t nth, Byte byte)
{
    if (fmt == 1u) {
        // float32 -> bfloat16 (round to nearest even), one element per thread and step
        __nv_bfloat16 *g = reinterpret_cast<__nv_bfloat16 *>(region);
        for (uint32_t f = tid; f < (n_out >> 2); f += nth) {
            const uint32_t x = byte(4u * f) | (byte(4u * f + 1u) << 8) | (byte(4u * f + 2u) << 16) | (byte(4u * f + 3u) << 24);
            g[f] = __float2bfloat16_rn(__uint_as_float(x));
        }
    } else if (fmt == 2u) {
        // uint8 pixel component -> float32 table[channel][x]
        float *g = reinterpret_cast<float *>(region);
        for (uint32_t i = tid; i < n_out; i += nth) __stcs(g + i, __ldg(lut + ((o0 + i) % lut_channels) * 256u + byte(i)));
    } else {
        uint8_t *g = region;
        const uint32_t head = min(n_out, (uint32_t)((4u - (reinterpret_cast<uintptr_t>(g) & 3u)) & 3u));
        if (tid < head) g[tid] = (uint8_t)byte(tid);
        const uint32_t nw = (n_out - head) >> 2;
        for (uint32_t w = tid; w < nw; w += nth) {
            const uint32_t i0 = head + 4u * w;
            const uint32_t x = byte(i0) | (byte(i0 + 1u) << 8) | (byte(i0 + 2u) << 16) | (byte(i0 + 3u) << 24);
            __stcs(reinterpret_cast<uint32_t *>(g + i0), x);
        }
        const uint32_t t0 = head + 4u * nw;
        if (tid < n_out - t0) g[t0 + tid] = (uint8_t)byte(t0 + tid);
    }
}

template <int kMode>
__global__ void __launch_bounds__(kWideThreads, 1)
k_wide_blocks(WideArea wa)
{
    extern __shared__ __align__(16) uint8_t wide_smem[];
    uint16_t *ptr = reinterpret_cast<uint16_t *>(wide_smem);                 // [kWideIdx]
    uint8_t *val = wide_smem + 2 * (size_t)kWideIdx;                         // [kWideIdx]
    uint32_t *heads = reinterpret_cast<uint32_t *>(wide_smem + 3 * (size_t)kWideIdx);   // [kWideIdx / 32]
    __shared__ uint32_t sh_b, sh_j, sh_bad;
    __shared__ __align__(16) WideJob sh_wj[2];                               // mode 2: the value of this block and of the CTA's next one
    __shared__ uint32_t sh_pre[2][4];                                        // ... their slots in `jobs`, blk_sub0[b], blk_sub0[b + 1]
    const uint32_t n_jobs = wa.state->n_jobs, n_blks = wa.state->n_blks, tid = threadIdx.x;
    uint32_t *ticket = kMode == 1 ? &wa.state->t_l1 : &wa.state->t_l3;
    if (n_jobs == 0) return;
    if ((kMode == 2) != wide_stored(wa)) return;                             // one parse into the store, or two without it
    // Mode 2 takes the blocks blockIdx.x, + gridDim.x, ... (no block waits for another one there).  What a block needs to
    // start -- which value it belongs to, that value's record, where its source starts -- is three dependent loads; they are
    // issued for the CTA's next block while this one is built, one step per phase, so that nobody waits for them.
    uint32_t b_lin = blockIdx.x, cur = 0;
    if (kMode == 2 && tid < 16 && b_lin < n_blks) {
        const uint32_t j = wa.blk_job[b_lin];
        reinterpret_cast<uint32_t *>(&sh_wj[0])[tid] = reinterpret_cast<const uint32_t *>(&wa.jobs[j])[tid];
        if (tid == 0) sh_pre[0][0] = j;
        if (tid < 2) sh_pre[0][1 + tid] = wa.blk_sub0[b_lin + tid];
    }
    for (;;) {
        __syncthreads();
        uint32_t b, jn, pre_j = 0, pre_w = 0, pre_s = 0, pre_stage = 3;
        const uint32_t nb = b_lin + gridDim.x;
        if (kMode == 2) {
            if (b_lin >= n_blks) break;
            b = b_lin;
            jn = sh_pre[cur][0];
            if (tid == 0) sh_bad = 0;
            if (nb < n_blks) {
                if (tid < 32) pre_j = wa.blk_job[nb];                        // (every lane of warp 0: the same word)
                if (tid < 2) pre_s = wa.blk_sub0[nb + tid];
                pre_stage = 1;
            }
            __syncthreads();
        } else {
            if (tid == 0) {
                sh_b = atomicAdd(ticket, 1u);
                sh_bad = 0;
                if (sh_b < n_blks) sh_j = wide_find_blk(wa, n_jobs, kMode == 3 ? n_blks - 1u - sh_b : sh_b);
            }
            __syncthreads();
            if (sh_b >= n_blks) break;
            b = kMode == 3 ? n_blks - 1u - sh_b : sh_b;                      // pass 3 runs down the blocks (see the flags below)
            jn = sh_j;
        }
        const WideJob wj = kMode == 2 ? sh_wj[cur] : wa.jobs[jn];
        do {                                                                 // (`break`: done with this block)
        const uint32_t k = b - wj.blk0;
        if (wj.state != 0u || k >= wj.nblk) {
            // (a block that is skipped still answers its predecessor, which may have started before the value was refused)
            if (kMode == 3 && tid == 0) *reinterpret_cast<volatile uint32_t *>(&wa.blk_flag[b]) = 1u;
            break;
        }
        if (kMode == 1 && k + 1u == wj.nblk) break;                          // nobody needs the last block's tail
        const uint32_t o0 = k * kWideBlock, o1 = min(o0 + kWideBlock, wj.rawsize), n_out = o1 - o0;
        const uint8_t *src = reinterpret_cast<const uint8_t *>(wj.src);
        uint8_t *dst = reinterpret_cast<uint8_t *>(wj.dst);
        uint8_t *region = dst + wide_scale(wj.fmt, o0);                      // this block's part of the output buffer
        const uint32_t slen = wj.slen;

        for (uint32_t i = tid; i < kWideIdx / 32; i += kWideThreads) heads[i] = 0;
        // every entry of the block starts as "not written" (no pointer has this value: kWideIdx < 65535)
        for (uint32_t i = (kWideWin + 1u) / 8u + tid; i < kWideIdx / 8u; i += kWideThreads)
            reinterpret_cast<uint4 *>(ptr)[i] = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu);
        if (tid == 0) ptr[kWideWin] = (uint16_t)0xFFFFu;
        // the previous tail: terminal entries (their bytes are known in pass 3 only: k_wide_tails left them at the end
        // of the previous block's region)
        for (uint32_t e = tid; e < kWideWin; e += kWideThreads) {
            ptr[e] = (uint16_t)e;
            if (kMode == 3) val[e] = k > 0u ? *(region - kWideWin + e) : (uint8_t)0;
        }
        __syncthreads();
        if (kMode == 3 && tid == 0) {                                        // the previous block may now overwrite its region
            __threadfence();
            *reinterpret_cast<volatile uint32_t *>(&wa.blk_flag[b]) = 1u;
        }

        // ---- items: one thread per 32-byte source sub-segment that feeds this block.  Literals and every byte of a
        //      match of up to kWideShort bytes get their pointer here; of a longer match only the first byte does (its
        //      offset is parked in the two unused value bytes, its entry is marked in the bitmap of long-match heads), the
        //      body keeps the "not written" value and is filled in below.
        const uint32_t floor_idx = k == 0u ? kWideWin : 0u;                  // nothing exists before the output start
        const uint32_t sub0 = kMode == 2 ? sh_pre[cur][1] : wa.blk_sub0[b];
        const uint32_t sub1 = k + 1u < wj.nblk ? (kMode == 2 ? sh_pre[cur][2] : wa.blk_sub0[b + 1]) : (slen ? (slen - 1u) / kWideSub : 0u);
        const uint32_t *fine = wa.sub_entry + (size_t)wj.seg0 * kWideSubs;
        const unsigned long long value_base = wa.seg_base[wj.seg0];
        for (uint32_t sb = sub0 + tid; sb <= sub1; sb += kWideThreads) {
            const uint32_t ent = fine[sb];
            if ((ent & 31u) == 31u) continue;
            uint32_t a = sb * kWideSub + (ent & 31u);
            uint32_t o = (uint32_t)(wa.seg_base[wj.seg0 + sb / kWideSubs] - value_base) + (ent >> 5);
            const uint32_t send = min((sb + 1u) * kWideSub, slen);
            while (a < send && o < o1) {
                const uint32_t ctrl = __ldg(src + a);
                uint32_t q = a + 1u;
#pragma unroll 1
                for (uint32_t it = 0; it < 8 && q < slen && o < o1; it++) {
                    if ((ctrl >> it) & 1u) {
                        const uint32_t b0 = __ldg(src + q), b1 = q + 1u < slen ? __ldg(src + q + 1u) : 0u;
                        const bool ext = (b0 & 0x0Fu) == 0x0Fu;
                        uint32_t len = (b0 & 0x0Fu) + 3u;                        // ppr:205
                        const uint32_t off = ((b0 & 0xF0u) << 4) | b1;           // ppr:206-208
                        if (ext) len += q + 2u < slen ? __ldg(src + q + 2u) : 0u;   // ppr:210-212
                        q += ext ? 3u : 2u;
                        const uint32_t lo = max(o, o0), hi = min(o + len, o1);
                        if (lo < hi) {
                            // (a match cut by the block start is the same match started at the cut: it is periodic in off)
                            const uint32_t e0 = kWideWin + (lo - o0), n = hi - lo;
                            if (off == 0u || e0 - off < floor_idx) {             // offset 0, or before the start of the output
                                sh_bad = 1u;
                            } else if (n <= kWideShort) {
                                if (off >= n) {                                  // no overlap: every byte copies the byte `off` back
                                    const uint32_t p0 = e0 - off;
#pragma unroll
                                    for (uint32_t kk = 0; kk < 8u; kk++)         // straight-line, predicated: lanes sit in different items
                                        if (kk < n) ptr[e0 + kk] = (uint16_t)(p0 + kk);
                                    for (uint32_t kk = 8u; kk < n; kk++) ptr[e0 + kk] = (uint16_t)(p0 + kk);
                                } else {
                                    uint32_t r = 0;
                                    for (uint32_t kk = 0; kk < n; kk++) {        // byte kk copies start - off + kk mod off (ppr:216-219)
                                        ptr[e0 + kk] = (uint16_t)(e0 - off + r);
                                        if (++r == off) r = 0;
                                    }
                                }
                            } else {
                                ptr[e0] = (uint16_t)(e0 - off);
                                val[e0] = (uint8_t)off;
                                val[e0 + 1u] = (uint8_t)(off >> 8);
                                atomicOr(&heads[e0 >> 5], 1u << (e0 & 31u));
                            }
                        }
                        o += len;
                    } else {
                        if (o >= o0) {
                            const uint32_t e = kWideWin + (o - o0);
                            val[e] = __ldg(src + q);
                            ptr[e] = (uint16_t)e;
                        }
                        o += 1u;
                        q += 1u;
                    }
                }
                a = q;
            }
        }
        __syncthreads();
        if (kMode == 2 && pre_stage == 1) {                                  // the next block's value: its record
            if (tid < 16) pre_w = reinterpret_cast<const uint32_t *>(&wa.jobs[pre_j])[tid];
            pre_stage = 2;
        }
        if (sh_bad) {
            // A match with offset 0, or one that reaches before the start of the output (the group chain of such a stream
            // can still be well-formed).  Its bytes were not entered into the forest, so nothing below may run on it.
            if (tid == 0) wide_fail(wa, jn);
            break;
        }

        // ---- bodies of long matches: point at the byte they copy (always before their own item)
        // my entries i * kWideThreads + tid of the block: i in [i_lo, i_hi)
        const uint32_t i_lo = (kWideWin - tid + kWideThreads - 1u) / kWideThreads,
                       i_hi = (kWideWin + n_out - tid + kWideThreads - 1u) / kWideThreads;
        const unsigned long long act64 = (i_hi >= 64u ? ~0ull : (1ull << i_hi) - 1ull) & ~((1ull << i_lo) - 1ull);
        uint32_t act[2] = {(uint32_t)act64, (uint32_t)(act64 >> 32)};        // my entries that are not yet resolved
#pragma unroll 4
        for (uint32_t i = i_lo; i < i_hi; i++) {
            const uint32_t e = i * kWideThreads + tid;
            if (ptr[e] != (uint16_t)0xFFFFu) continue;                       // written by the emission
            const uint32_t h = wide_head_of(heads, e);
            if (h == kNone) { sh_bad = 1u; ptr[e] = (uint16_t)e; continue; }  // cannot happen on a stream k_wide_chain accepted
            const uint32_t off = (uint32_t)val[h] | ((uint32_t)val[h + 1u] << 8), kk = e - h;
            if (off == 0u || off > kWideWin || off > h) { sh_bad = 1u; ptr[e] = (uint16_t)e; continue; }   // (same: keeps every pointer backwards)
            uint32_t sidx;
            if (off > kk) sidx = e - off;
            else if (off <= 1u) sidx = h - 1u;                                // runs of one byte value
            else sidx = h - off + kk % off;                                   // overlapping copy (ppr:216-219)
            ptr[e] = (uint16_t)sidx;
        }
        __syncthreads();

        // Pass 1 only needs the roots of the block's last 4095 bytes: follow those chains directly, a bounded number
        // of hops; only when a chain is longer (run-dominated data) resolve the whole block as pass 3 does.
        bool resolve_all = true;
        uint32_t tail_root[4] = {0u, 0u, 0u, 0u};
        if (kMode == 1) {
            int unfinished = 0;
#pragma unroll
            for (uint32_t j = 0; j < 4; j++) {
                const uint32_t i = tid + kWideThreads * j;
                uint32_t r = i < kWideWin ? (uint32_t)ptr[kWideBlock + i] : 0u;      // entry kWideWin + (kWideBlock - kWideWin) + i
                uint32_t q = ptr[r];
                for (uint32_t hop = 0; hop < kWideChase && q != r; hop++) { r = q; q = ptr[r]; }
                if (q != r) unfinished = 1;
                tail_root[j] = r;
            }
            resolve_all = __syncthreads_or(unfinished) != 0;
        }
        // ---- pointer jumping (two hops a round) until every byte points at a literal of the block or into the
        //      previous tail; resolved entries leave the per-thread work mask
        // (pointers only ever point backwards, so the forest has no cycle and log2(kWideIdx) rounds suffice; the bound is a
        // safety net: a kernel must not spin on damaged input)
        for (uint32_t round = 0; resolve_all; round++) {
            if (round > 24u) { sh_bad = 1u; break; }
            int changed = 0;
#pragma unroll
            for (uint32_t half = 0; half < 2; half++) {
                uint32_t m = act[half];
                while (m) {
                    const uint32_t i = (uint32_t)__ffs((int)m) - 1u;
                    m &= m - 1u;
                    const uint32_t e = (32u * half + i) * kWideThreads + tid;
                    const uint32_t p = ptr[e];
                    const uint32_t q = ptr[p];
                    if (q == p) {
                        act[half] &= ~(1u << i);
                    } else {
                        ptr[e] = ptr[q];
                        changed = 1;
                    }
                }
            }
            if (!__syncthreads_or(changed)) break;
        }
        if (sh_bad) {                                                        // offset 0 or before the output start
            if (tid == 0) wide_fail(wa, jn);
            break;
        }

        if (kMode == 2) {
            // every byte of the block as "value" or "index into the previous tail", 4 entries per thread and step
            uint16_t *map = wa.store + (size_t)b * kWideBlock;
            for (uint32_t f = tid; f < (n_out + 3u) / 4u; f += kWideThreads) {
                uint32_t m[4];
#pragma unroll
                for (uint32_t j = 0; j < 4; j++) {
                    const uint32_t r = 4u * f + j < n_out ? (uint32_t)ptr[kWideWin + 4u * f + j] : 0u;
                    m[j] = r >= kWideWin ? 0x8000u | val[r] : r;
                }
                *reinterpret_cast<uint2 *>(map + 4u * f) = make_uint2(m[0] | (m[1] << 16), m[2] | (m[3] << 16));
            }
            break;
        }
        if (kMode == 1) {
            // the block's last 4095 bytes, each as "value" or "index into the previous tail"
            uint16_t *map = reinterpret_cast<uint16_t *>((reinterpret_cast<uintptr_t>(region) + 1u) & ~(uintptr_t)1u);
#pragma unroll
            for (uint32_t j = 0; j < 4; j++) {
                const uint32_t i = tid + kWideThreads * j;
                if (i < kWideWin) {
                    const uint32_t r = resolve_all ? (uint32_t)ptr[kWideBlock + i] : tail_root[j];
                    map[i] = (uint16_t)(r >= kWideWin ? 0x8000u | val[r] : r);
                }
            }
            break;
        }
        // Pass 3 overwrites the whole region, including the tail bytes the NEXT block's pass 3 reads: wait until it has
        // read them.  Blocks are handed out from the last to the first, so that block already runs (or is done).
        if (k + 1u < wj.nblk) {
            if (tid == 0) {
                while (*reinterpret_cast<volatile uint32_t *>(&wa.blk_flag[b + 1]) == 0u) { }
                __threadfence();
            }
            __syncthreads();
        }
        wide_write_block(wj.fmt, region, o0, n_out, wa.lut, wa.lut_channels, tid, kWideThreads,
                         [&](uint32_t i) { return (uint32_t)val[ptr[kWideWin + i]]; });
        } while (0);
        if (kMode == 2) {
            // the last steps of the next block's lookup (all of them here when this block ended early)
            if (pre_stage == 1) {
                if (tid < 16) pre_w = reinterpret_cast<const uint32_t *>(&wa.jobs[pre_j])[tid];
                pre_stage = 2;
            }
            if (pre_stage == 2) {
                if (tid < 16) reinterpret_cast<uint32_t *>(&sh_wj[cur ^ 1u])[tid] = pre_w;
                if (tid == 0) sh_pre[cur ^ 1u][0] = pre_j;
                if (tid < 2) sh_pre[cur ^ 1u][1 + tid] = pre_s;
            }
            b_lin = nb;
            cur ^= 1u;
        }
    }
}

// Tail chain.  Block k's pass-1 map says, for each of its last 4095 bytes, "value" or "index into the tail of block
// k-1"; applying the maps in order turns them into bytes.  That is a chain of nblk-1 dependent gathers per value,
// so it is run in three phases over groups of kTailGroup blocks:
//   phase 0  one CTA per group folds the group's maps into one (entries stay "value" or "index into the tail
//            that enters the group"), in parallel over groups;
//   phase 1  one CTA per value walks the folded maps: the tail entering every group;
//   phase 2  one CTA per group replays its blocks from that tail and writes every block's tail bytes where pass 3
//            of the next block reads them (the end of the block's own output region).
// The maps do not depend on the chain, so they are fetched kTailAhead steps ahead; the serial step is a
// shared-memory gather.  Folded maps and entering tails live in the group's first block's output region (regions
// are at least kWideBlock / 2 = 24576 bytes: float32 -> bfloat16 output halves them).
constexpr uint32_t kTailGroup = 32, kTailAhead = 6;
constexpr uint32_t kTailRing = 12;                                           // stored maps in flight per CTA (cp.async ring, 8 KB each)
constexpr size_t kWideTailsSmem = (size_t)kTailRing * 8192;
constexpr uint32_t kTailGmapOff = 8192, kTailGtailOff = 16384;               // inside the first block of a group
static_assert(kTailGtailOff + 4096u <= kWideBlock / 2u, "the tail scratch must fit the smallest block region");

__device__ __forceinline__ const uint16_t *wide_u16_at(const uint8_t *p)
{
    return reinterpret_cast<const uint16_t *>((reinterpret_cast<uintptr_t>(p) + 1u) & ~(uintptr_t)1u);
}

__device__ __forceinline__ void cp_async_16(void *smem, const void *gmem)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int kPending>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(kPending) : "memory"); }

template <int kPhase>
__global__ void __launch_bounds__(1024)
k_wide_tails(WideArea wa)
{
    extern __shared__ __align__(16) uint8_t tails_ring[];                    // phases 0, 2 with stored maps: kTailRing x 8192 bytes
    __shared__ uint16_t tail[2][4096];                                       // 0x8000 | byte, or an index (phase 0 only)
    __shared__ uint32_t work_u[1024], work_j[1024], n_work;
    const uint32_t n_jobs = wa.state->n_jobs, tid = threadIdx.x;
    if (n_jobs == 0) return;
    const uint32_t n_units = kPhase == 1 ? n_jobs : wa.state->n_blks;        // phase 1: values; phases 0, 2: group leaders
    const bool stored = wide_stored(wa);
    // This CTA's units are blockIdx.x, + gridDim.x, ...  Only one block in 32 leads a group, and finding a block's value is a
    // binary search: every thread looks one unit up, the units with work are listed, then the CTA goes through the list.
    for (uint32_t u0 = blockIdx.x; u0 < n_units; u0 += 1024u * gridDim.x) {
    __syncthreads();
    if (tid == 0) n_work = 0;
    __syncthreads();
    {
        const unsigned long long mine = (unsigned long long)u0 + (unsigned long long)tid * gridDim.x;
        if (mine < n_units) {
            const uint32_t u = (uint32_t)mine, j = kPhase == 1 ? u : wide_find_blk(wa, n_jobs, u);
            const WideJob *pj = &wa.jobs[j];
            bool work = pj->state == 0u && pj->nblk >= 2u;
            if (work && kPhase != 1) {
                const uint32_t k = u - pj->blk0;
                work = k < pj->nblk - 1u && k % kTailGroup == 0u;
            }
            if (work) {
                const uint32_t slot = atomicAdd(&n_work, 1u);
                work_u[slot] = u;
                work_j[slot] = j;
            }
        }
    }
    __syncthreads();
    for (uint32_t wi = 0; wi < n_work; wi++) {
        __syncthreads();                                                     // (the previous unit is done with `tail`)
        const uint32_t u = work_u[wi];
        const WideJob wj = wa.jobs[work_j[wi]];
        if (wj.state != 0u || wj.nblk < 2u) continue;
        const uint32_t nlinks = wj.nblk - 1u;                                // blocks 0 .. nblk-2 hand a tail on
        uint8_t *dst = reinterpret_cast<uint8_t *>(wj.dst);
        uint32_t first, nsteps;                                              // steps of this unit; step = block (phases 0, 2) or group (1)
        if (kPhase == 1) {
            first = 0;
            nsteps = (nlinks + kTailGroup - 1u) / kTailGroup;
        } else {
            const uint32_t k = u - wj.blk0;
            if (k >= nlinks || k % kTailGroup) continue;
            first = k;
            nsteps = min(kTailGroup, nlinks - k);
        }
        const uint32_t fmt = wj.fmt;
        auto region = [&](uint32_t blk) { return dst + wide_scale(fmt, (size_t)blk * kWideBlock); };   // block's part of the output buffer
        const size_t region_bytes = wide_scale(fmt, (size_t)kWideBlock);
        // a block's tail map: the first 8190 bytes of its output region, or the last 4095 entries of its stored map
        auto stored_map = [&](uint32_t blk) { return wa.store + (size_t)(wj.blk0 + blk) * kWideBlock + (kWideBlock - kWideWin); };
        auto map_of = [&](uint32_t step) {
            return kPhase == 1 ? wide_u16_at(region(step * kTailGroup) + kTailGmapOff)
                               : stored ? (const uint16_t *)stored_map(first + step) : wide_u16_at(region(first + step));
        };
        // state entering the unit
        for (uint32_t i = tid; i < 4096; i += blockDim.x) {
            uint32_t v;
            if (kPhase == 0) v = i;
            else if (kPhase == 1) v = 0x8000u;
            else v = 0x8000u | region(first)[kTailGtailOff + i];
            tail[0][i] = (uint16_t)v;
        }
        if (kPhase != 1 && stored) {
            // Stored maps: a block's last 4096 entries are 8192 aligned bytes; they arrive through a ring of asynchronous copies
            // kTailRing - 1 steps ahead (the store is far larger than L2: a step would otherwise wait a DRAM round trip).
            auto fetch = [&](uint32_t step) {
                if (step < nsteps && tid < 512u) {
                    const uint8_t *g = reinterpret_cast<const uint8_t *>(wa.store + (size_t)(wj.blk0 + first + step) * kWideBlock + (kWideBlock - 4096u));
                    cp_async_16(tails_ring + (size_t)(step % kTailRing) * 8192u + tid * 16u, g + tid * 16u);
                }
                cp_async_commit();
            };
            for (uint32_t a = 0; a + 1u < kTailRing; a++) fetch(a);
            uint32_t cur = 0;
            for (uint32_t k = 0; k < nsteps; k++) {
                cp_async_wait<(int)kTailRing - 2>();
                __syncthreads();                                             // step k's map is here; everybody is done with step k - 1
                fetch(k + kTailRing - 1u);                                   // (into the slot step k - 1 used)
                const uint16_t *mp = reinterpret_cast<const uint16_t *>(tails_ring + (size_t)(k % kTailRing) * 8192u) + 1;
#pragma unroll
                for (uint32_t j = 0; j < 4; j++) {
                    const uint32_t i = tid + 1024u * j;
                    if (i < kWideWin) {
                        const uint32_t mm = mp[i];
                        const uint32_t v = (mm & 0x8000u) ? mm : tail[cur][mm & 4095u];
                        tail[cur ^ 1u][i] = (uint16_t)v;
                        if (kPhase == 2) stored_map(first + k)[i] = (uint16_t)(0x8000u | (v & 0xFFu));
                    }
                }
                cur ^= 1u;
            }
            cp_async_wait<0>();
            __syncthreads();
            if (kPhase == 0) {                                               // the folded map of this group
                uint16_t *gm = const_cast<uint16_t *>(wide_u16_at(region(first) + kTailGmapOff));
                for (uint32_t i = tid; i < kWideWin; i += blockDim.x) gm[i] = tail[cur][i];
            }
            continue;
        }
        // thread t owns entries t, t + 1024, t + 2048, t + 3072 (the last one only below 4095)
        uint32_t m[kTailAhead][4];
#pragma unroll
        for (uint32_t a = 0; a < kTailAhead; a++) {
#pragma unroll
            for (uint32_t j = 0; j < 4; j++) {
                const uint32_t i = tid + 1024u * j;
                m[a][j] = a < nsteps && i < kWideWin ? map_of(a)[i] : 0x8000u;
            }
        }
        __syncthreads();
        uint32_t cur = 0;
        for (uint32_t k0 = 0; k0 < nsteps; k0 += kTailAhead) {
#pragma unroll
            for (uint32_t a = 0; a < kTailAhead; a++) {
                const uint32_t k = k0 + a;
                if (k < nsteps) {                                            // uniform
                    if (kPhase == 1) {                                       // the tail entering group k
                        uint8_t *gt = region(k * kTailGroup) + kTailGtailOff;
                        for (uint32_t i = tid; i < 4096; i += blockDim.x) gt[i] = (uint8_t)tail[cur][i];
                    }
                    uint8_t *out = region(first + k) + region_bytes - kWideWin;
#pragma unroll
                    for (uint32_t j = 0; j < 4; j++) {
                        const uint32_t i = tid + 1024u * j, mm = m[a][j];
                        if (i < kWideWin) {
                            const uint32_t v = (mm & 0x8000u) ? mm : tail[cur][mm & 4095u];
                            tail[cur ^ 1u][i] = (uint16_t)v;
                            if (kPhase == 2) {
                                if (stored) stored_map(first + k)[i] = (uint16_t)(0x8000u | (v & 0xFFu));   // (read above, by this thread)
                                else out[i] = (uint8_t)v;
                            }
                        }
                        const uint32_t kn = k + kTailAhead;
                        m[a][j] = kn < nsteps && i < kWideWin ? map_of(kn)[i] : 0x8000u;
                    }
                    __syncthreads();
                    cur ^= 1u;
                }
            }
        }
        if (kPhase == 0) {                                                   // the folded map of this group
            uint16_t *gm = const_cast<uint16_t *>(wide_u16_at(region(first) + kTailGmapOff));
            for (uint32_t i = tid; i < kWideWin; i += blockDim.x) gm[i] = tail[cur][i];
        }
    }
    }
}

// Stored blocks -> output (the streaming half of the one-parse path).  Only the previous block's tail (which
// k_wide_tails<2> has turned into bytes, in place) goes through shared memory; the block's own entries are read once,
// 8 per thread and step, and leave in the output format.
constexpr uint32_t kFinishThreads = 512;
__global__ void __launch_bounds__(kFinishThreads, 4)
k_wide_finish(WideArea wa)
{
    __shared__ __align__(16) uint8_t tail[4096];
    __shared__ uint32_t sh_j;
    const uint32_t n_jobs = wa.state->n_jobs, n_blks = wa.state->n_blks, tid = threadIdx.x;
    if (n_jobs == 0 || !wide_stored(wa)) return;
    for (uint32_t b = blockIdx.x; b < n_blks; b += gridDim.x) {
        __syncthreads();
        if (tid == 0) sh_j = wa.blk_job[b];
        __syncthreads();
        const WideJob wj = wa.jobs[sh_j];
        const uint32_t k = b - wj.blk0;
        if (wj.state != 0u || k >= wj.nblk) continue;
        const uint32_t o0 = k * kWideBlock, n_out = min(kWideBlock, wj.rawsize - o0);
        const uint16_t *map = wa.store + (size_t)b * kWideBlock;
        if (k > 0u) {
            const uint16_t *pt = map - kWideWin;                             // the last 4095 entries of the previous block
            for (uint32_t i = tid; i < kWideWin; i += kFinishThreads) tail[i] = (uint8_t)pt[i];
        }
        __syncthreads();
        auto byte_of = [&](uint32_t m) { return (m & 0x8000u) ? (m & 0xFFu) : (uint32_t)tail[m & 4095u]; };
        uint8_t *region = reinterpret_cast<uint8_t *>(wj.dst) + wide_scale(wj.fmt, o0);
        if (wj.fmt != 2u && (reinterpret_cast<uintptr_t>(region) & 3u) == 0u) {
            // 8 entries = 8 output bytes, or 2 float32 -> 2 bfloat16
            const uint4 *g = reinterpret_cast<const uint4 *>(map);
            const uint32_t n8 = n_out >> 3;
            for (uint32_t i = tid; i < n8; i += kFinishThreads) {
                const uint4 v = __ldcs(g + i);
                const uint32_t lo = byte_of(v.x & 0xFFFFu) | (byte_of(v.x >> 16) << 8) | (byte_of(v.y & 0xFFFFu) << 16) | (byte_of(v.y >> 16) << 24);
                const uint32_t hi = byte_of(v.z & 0xFFFFu) | (byte_of(v.z >> 16) << 8) | (byte_of(v.w & 0xFFFFu) << 16) | (byte_of(v.w >> 16) << 24);
                if (wj.fmt == 1u) {
                    const __nv_bfloat162 h = __floats2bfloat162_rn(__uint_as_float(lo), __uint_as_float(hi));
                    __stcs(reinterpret_cast<uint32_t *>(region) + i, *reinterpret_cast<const uint32_t *>(&h));
                } else {
                    __stcs(reinterpret_cast<uint32_t *>(region) + 2u * i, lo);
                    __stcs(reinterpret_cast<uint32_t *>(region) + 2u * i + 1u, hi);
                }
            }
            // the last 1..7 bytes of a value
            const uint32_t t0 = n8 << 3;
            if (wj.fmt == 1u) {
                if (tid == 0 && n_out - t0 >= 4u) {
                    const uint32_t x = byte_of(map[t0]) | (byte_of(map[t0 + 1u]) << 8) | (byte_of(map[t0 + 2u]) << 16) | (byte_of(map[t0 + 3u]) << 24);
                    reinterpret_cast<__nv_bfloat16 *>(region)[t0 >> 2] = __float2bfloat16_rn(__uint_as_float(x));
                }
            } else if (tid < n_out - t0) {
                region[t0 + tid] = (uint8_t)byte_of(map[t0 + tid]);
            }
        } else {
            wide_write_block(wj.fmt, region, o0, n_out, wa.lut, wa.lut_channels, tid, kFinishThreads,
                             [&](uint32_t i) { return byte_of(__ldg(map + i)); });
        }
    }
}

}  // namespace
}  // namespace cds
