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
//     25 bytes of it, so there are only 25 possible entries into a segment.  k_wide_maps evaluates
//     "group length / output bytes if a group started here" for EVERY source position (shared memory,
//     one thread per position), then 25 lanes of a warp chase the 25 candidate chains through the
//     segment: entry -> (exit offset into the next segment, output bytes).  These 25-entry maps compose
//     associatively: k_wide_super folds 256 of them, one warp walks the folded maps of a value
//     (k_wide_chain: also the stream's validity: it must end exactly at srcend having produced exactly
//     rawsize bytes), k_wide_entries unfolds them into the true entry and output position of every segment.
//
//  B. What does a match copy?  Output is cut into blocks of 57344 bytes.  A block is rebuilt in shared
//     memory as a forest: every output byte points at the byte it copies (a literal points at itself;
//     byte k of a match points at start - off + (k mod off), i.e. always before its own item), offsets
//     are <= 4095 so pointers leave the block only into the last 4095 bytes of the previous block.
//     Pointer jumping (ptr[i] = ptr[ptr[i]]) resolves every byte to a literal of the block or to one of
//     those 4095 previous bytes in O(log depth) rounds.  Pass 1 (k_wide_blocks<1>) does this with the
//     previous tail unknown and publishes, for the block's own last 4095 bytes, "value" or "index into
//     the previous tail"; k_wide_tails runs down each value applying these maps (4095 gathers per block
//     from shared memory, the only serial step left); pass 3 (k_wide_blocks<3>) repeats the block
//     with the previous tail known and writes the output, 128 contiguous bytes per warp store.
//
// Work memory: 8 bytes per source segment.  The segment maps, the tail maps and the resolved tails live
// inside the value's own (not yet written) output region and are overwritten by pass 3.
#pragma once
#include "cds_common.cuh"

namespace cds {

constexpr uint32_t kWideSeg = 512;                 // source bytes per segment
constexpr uint32_t kWideChunkSegs = 8;             // segments per k_wide_maps step (one per warp)
constexpr uint32_t kWideChunk = kWideSeg * kWideChunkSegs;
constexpr uint32_t kWideSuper = 256;               // segments per folded map
constexpr uint32_t kWideWin = 4095;                // largest pglz offset
constexpr uint32_t kWideBlock = 57344;             // output bytes per block
constexpr uint32_t kWideIdx = 61440;               // index space of a block: [previous tail | block], rounded up
constexpr uint32_t kWideThreads = 1024;
constexpr uint32_t kWidePer = kWideIdx / kWideThreads;   // 60 entries per thread
constexpr uint32_t kWideEnd = 30, kWideBad = 31;   // map states besides the entry offsets 0..24
constexpr uint32_t kWideMinRaw = 1024;             // below this the maps would not fit the output region
constexpr size_t kWideBlockSmem = (size_t)kWideIdx * 3 + kWideIdx / 8;

struct __align__(16) WideJob {
    unsigned long long src, dst, err_key;
    uint32_t slen, rawsize;
    uint32_t job;        // slot in the job array (status word)
    uint32_t seg0;       // first flat segment (multiple of kWideSuper)
    uint32_t blk0;       // first flat output block
    uint32_t nseg, nblk;
    uint32_t state;      // 0 well-formed so far, 1 corrupt, 2 not taken (capacity)
};

struct WideState {       // zero when a decode starts
    uint32_t n_jobs, n_segs, n_blks;
    uint32_t t_maps, t_l1, t_l3;
    uint32_t n_taken;    // values this decoder accepted (the rest stay with the warp decoder)
    uint32_t pad[9];
};

struct WideArea {
    WideState *state;
    WideJob *jobs;
    uint2 *seg_entry;        // per flat segment: (output position at its first group, entry offset)
    uint32_t *supermap;      // per flat super-segment: 32 words
    uint2 *sseg_entry;       // per flat super-segment: (output position, entry offset)
    uint32_t *blk_seg0;      // per flat block: local segment holding the group that produces its first byte
    uint32_t cap_jobs, cap_segs, cap_blks;
    uint32_t threshold;      // values with rawsize >= threshold take this path
    int32_t *status;         // per job slot (may be null)
    DevState *st;            // first-error word (may be null)
};

__host__ __device__ inline size_t wide_area_bytes(uint32_t cap_jobs, uint32_t cap_segs, uint32_t cap_blks)
{
    return 256 + sizeof(WideJob) * (size_t)cap_jobs + 8 * (size_t)cap_segs + (128 + 8) * (size_t)(cap_segs / kWideSuper + 1) +
           4 * (size_t)(cap_blks + 1) + 256;
}

// Carve the work area out of one 16-byte aligned allocation (host side).
inline WideArea wide_area_carve(uint8_t *base, uint32_t cap_jobs, uint32_t cap_segs, uint32_t cap_blks)
{
    WideArea wa;
    uint8_t *p = base;
    wa.state = reinterpret_cast<WideState *>(p); p += 256;
    wa.jobs = reinterpret_cast<WideJob *>(p); p += sizeof(WideJob) * (size_t)cap_jobs;
    wa.seg_entry = reinterpret_cast<uint2 *>(p); p += 8 * (size_t)cap_segs;
    wa.supermap = reinterpret_cast<uint32_t *>(p); p += 128 * (size_t)(cap_segs / kWideSuper + 1);
    wa.sseg_entry = reinterpret_cast<uint2 *>(p); p += 8 * (size_t)(cap_segs / kWideSuper + 1);
    wa.blk_seg0 = reinterpret_cast<uint32_t *>(p);
    wa.cap_jobs = cap_jobs; wa.cap_segs = cap_segs; wa.cap_blks = cap_blks;
    wa.threshold = 0; wa.status = nullptr; wa.st = nullptr;
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

__device__ inline uint32_t *wide_maps_of(const WideJob &wj)       // segment maps: start of the output region
{
    return reinterpret_cast<uint32_t *>((wj.dst + 3ull) & ~3ull);
}

// ---- classification: which jobs take this path -----------------------------------------------------
// One block walks the job list in order (two-ended list of k_resolve when st_counts is given) and appends
// the large values to the wide list with their flat segment / block ranges; taken jobs are blanked in the
// job array so that the warp decoder skips them.  A value also needs slen <= 9/8 rawsize + 9 (anything
// longer cannot be well-formed and is left to the warp decoder to refuse) and no output conversion.
__global__ void __launch_bounds__(1024)
k_wide_classify(PglzJob *jobs, const DevState *st_counts, uint32_t n_jobs_host, WideArea wa)
{
    __shared__ uint32_t sh_w[3][32];
    __shared__ uint32_t sh_base[3];
    const uint32_t front = st_counts ? st_counts->n_jobs_front : n_jobs_host;
    const uint32_t back = st_counts ? st_counts->n_jobs_back : 0u;
    const uint32_t last_slot = st_counts ? 2u * st_counts->n_rows - 1u : 0u;
    const uint32_t n_jobs = front + back, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x < 3) sh_base[threadIdx.x] = 0;
    __syncthreads();
    for (uint32_t base = 0; base < n_jobs; base += blockDim.x) {
        const uint32_t t = base + threadIdx.x;
        const uint32_t j = t < front ? t : last_slot - (t - front);
        PglzJob job;
        job.slen = kNone;
        if (t < n_jobs) job = jobs[j];
        bool big = t < n_jobs && job.slen != kNone && (job.slen & kFlagBits) == 0u && job.rawsize >= wa.threshold &&
                   job.rawsize >= kWideMinRaw && (unsigned long long)job.slen <= (unsigned long long)job.rawsize * 9ull / 8ull + 9ull;
        uint32_t nseg = 0, nblk = 0;
        if (big) {
            nseg = job.slen ? (job.slen + kWideSeg - 1) / kWideSeg : 1u;
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
                wj.slen = job.slen; wj.rawsize = job.rawsize; wj.job = j;
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
        wa.state->n_segs = sh_base[1];
        wa.state->n_blks = sh_base[2];
    }
}

// ---- A: group starts ---------------------------------------------------------------------------------

// The group that would start at stream offset a (a < slen), read from `b` (b[0] = the control byte).
// Returns its length in bytes (0: a tag runs past the source, or the control byte has no item: both
// end in "corrupt", ppr:229-230) and its output bytes.
__device__ __forceinline__ uint32_t wide_group(const uint8_t *b, uint32_t a, uint32_t slen, uint32_t *out_bytes)
{
    const uint32_t ctrl = b[0], avail = slen - a;                            // >= 1
    uint32_t q = 1, out = 0;
#pragma unroll
    for (uint32_t k = 0; k < 8; k++) {
        if (q < avail) {                                                     // `sp < srcend` of the inner loop (ppr:203)
            if ((ctrl >> k) & 1u) {
                const uint32_t b0 = b[q], ext = (b0 & 0x0Fu) == 0x0Fu ? 1u : 0u;
                if (q + 2u + ext > avail) { *out_bytes = 0; return 0; }
                out += (b0 & 0x0Fu) + 3u + (ext ? (uint32_t)b[q + 2u] : 0u);  // ppr:205,210-212
                q += 2u + ext;
            } else {
                out += 1u;
                q += 1u;
            }
        }
    }
    *out_bytes = out;
    return q == 1u ? 0u : q;
}

__global__ void __launch_bounds__(256)
k_wide_maps(WideArea wa)
{
    __shared__ uint8_t sb[kWideChunk + 64];
    __shared__ uint8_t sdelta[kWideChunk + 32];
    __shared__ uint16_t sgout[kWideChunk + 32];
    __shared__ uint32_t sh_c;
    const uint32_t n_jobs = wa.state->n_jobs, n_chunks = wa.state->n_segs / kWideChunkSegs;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (n_jobs == 0) return;
    for (;;) {
        __syncthreads();
        if (threadIdx.x == 0) sh_c = atomicAdd(&wa.state->t_maps, 1u);
        __syncthreads();
        const uint32_t c = sh_c;
        if (c >= n_chunks) break;
        const uint32_t seg = c * kWideChunkSegs, jn = wide_find_seg(wa, n_jobs, seg);
        const WideJob wj = wa.jobs[jn];
        const uint32_t lseg = seg - wj.seg0;
        if (wj.state != 0u || lseg >= wj.nseg) continue;
        const uint32_t coff = lseg * kWideSeg, slen = wj.slen;
        const uint8_t *src = reinterpret_cast<const uint8_t *>(wj.src);
        for (uint32_t i = threadIdx.x; i < kWideChunk + 64; i += blockDim.x) sb[i] = coff + i < slen ? __ldg(src + coff + i) : (uint8_t)0;
        __syncthreads();
        for (uint32_t p = threadIdx.x; p < kWideChunk + 25; p += blockDim.x) {
            uint32_t d = 0, o = 0;
            if (coff + p < slen) d = wide_group(sb + p, coff + p, slen, &o);
            sdelta[p] = (uint8_t)d;
            sgout[p] = (uint16_t)o;
        }
        __syncthreads();
        if (lseg + warp < wj.nseg) {
            uint32_t p = warp * kWideSeg + lane, out = 0, e;
            const uint32_t end = (warp + 1u) * kWideSeg;
            if (lane >= 25u) {
                e = kWideBad;
            } else {
                for (;;) {
                    const uint32_t a = coff + p;
                    if (a >= slen) { e = a == slen ? kWideEnd : kWideBad; break; }
                    if (p >= end) { e = p - end; break; }
                    const uint32_t d = sdelta[p];
                    if (d == 0u) { e = kWideBad; break; }
                    out += sgout[p];
                    p += d;
                }
            }
            wide_maps_of(wj)[(size_t)(lseg + warp) * 32u + lane] = e | (out << 5);
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
        const uint32_t *maps = wide_maps_of(wj) + (size_t)lseg0 * 32u;
        uint32_t e = lane < 25u ? lane : kWideBad, out = 0;
        for (uint32_t s = 0; s < n; s += 8) {
            uint32_t r[8];
#pragma unroll
            for (uint32_t i = 0; i < 8; i++) r[i] = s + i < n ? maps[(size_t)(s + i) * 32u + lane] : 0u;
#pragma unroll
            for (uint32_t i = 0; i < 8; i++) {
                const uint32_t v = __shfl_sync(0xffffffffu, r[i], e < 25u ? e : 0u);
                if (s + i < n && e < 25u) { out += v >> 5; e = v & 31u; }
            }
        }
        wa.supermap[(size_t)ss * 32u + lane] = e | (out << 5);
    }
}

// One warp per value: walk its folded maps from the true start (entry 0, output 0), record where every
// super-segment is entered, and decide ppr:229-230: the chain must end exactly at srcend with exactly
// rawsize bytes produced.
__global__ void __launch_bounds__(32)
k_wide_chain(WideArea wa)
{
    const uint32_t n_jobs = wa.state->n_jobs, lane = threadIdx.x;
    for (uint32_t jn = blockIdx.x; jn < n_jobs; jn += gridDim.x) {
        const WideJob wj = wa.jobs[jn];
        if (wj.state != 0u) continue;
        const uint32_t nss = (wj.nseg + kWideSuper - 1) / kWideSuper, ss0 = wj.seg0 / kWideSuper;
        uint32_t e = 0;
        unsigned long long ob = 0;
        for (uint32_t s = 0; s < nss; s += 8) {
            uint32_t r[8];
#pragma unroll
            for (uint32_t i = 0; i < 8; i++) r[i] = s + i < nss ? wa.supermap[(size_t)(ss0 + s + i) * 32u + lane] : 0u;
#pragma unroll
            for (uint32_t i = 0; i < 8; i++) {
                if (s + i < nss) {
                    if (lane == 0) wa.sseg_entry[ss0 + s + i] = make_uint2((uint32_t)ob, e);
                    const uint32_t v = __shfl_sync(0xffffffffu, r[i], e < 25u ? e : 0u);
                    if (e < 25u) { ob += v >> 5; e = v & 31u; }
                }
            }
        }
        if (lane == 0 && !(e == kWideEnd && ob == (unsigned long long)wj.rawsize)) wide_fail(wa, jn);
    }
}

// Unfold: true (output position, entry offset) of every segment, and the segment each output block starts in.
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
        const uint32_t *maps = wide_maps_of(wj) + (size_t)lseg0 * 32u;
        const uint2 ent = wa.sseg_entry[ss];
        uint32_t ob = ent.x, e = ent.y;
        for (uint32_t s = 0; s < n; s += 8) {
            uint32_t r[8];
#pragma unroll
            for (uint32_t i = 0; i < 8; i++) r[i] = s + i < n ? maps[(size_t)(s + i) * 32u + lane] : 0u;
#pragma unroll
            for (uint32_t i = 0; i < 8; i++) {
                if (s + i < n && e < 25u) {
                    const uint32_t v = __shfl_sync(0xffffffffu, r[i], e);
                    const uint32_t nb = ob + (v >> 5);
                    if (lane == 0) {
                        wa.seg_entry[seg + s + i] = make_uint2(ob, e);
                        const uint32_t k = (ob + kWideBlock - 1) / kWideBlock;       // first block boundary at or after ob
                        if ((unsigned long long)k * kWideBlock < nb && k < wj.nblk) wa.blk_seg0[wj.blk0 + k] = lseg0 + s + i;
                    }
                    ob = nb;
                    e = v & 31u;
                }
            }
        }
    }
}

// ---- B: blocks -----------------------------------------------------------------------------------------

__device__ __forceinline__ bool wide_is_head(const uint32_t *heads, uint32_t e) { return (heads[e >> 5] >> (e & 31u)) & 1u; }

// Nearest item start at or below e (there is one within 273 entries of any emitted byte).
__device__ __forceinline__ uint32_t wide_head_of(const uint32_t *heads, uint32_t e)
{
    uint32_t w = e >> 5;
    uint32_t m = heads[w] & (0xFFFFFFFFu >> (31u - (e & 31u)));
    while (m == 0u && w > 0u) m = heads[--w];
    return m ? (w << 5) + 31u - (uint32_t)__clz((int)m) : kNone;
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
    const uint32_t n_jobs = wa.state->n_jobs, n_blks = wa.state->n_blks, tid = threadIdx.x;
    uint32_t *ticket = kMode == 1 ? &wa.state->t_l1 : &wa.state->t_l3;
    if (n_jobs == 0) return;
    for (;;) {
        __syncthreads();
        if (tid == 0) {
            sh_b = atomicAdd(ticket, 1u);
            sh_bad = 0;
            if (sh_b < n_blks) sh_j = wide_find_blk(wa, n_jobs, sh_b);
        }
        __syncthreads();
        const uint32_t b = sh_b;
        if (b >= n_blks) break;
        const uint32_t jn = sh_j;
        const WideJob wj = wa.jobs[jn];
        const uint32_t k = b - wj.blk0;
        if (wj.state != 0u || k >= wj.nblk) continue;
        if (kMode == 1 && k + 1u == wj.nblk) continue;                       // nobody needs the last block's tail
        const uint32_t o0 = k * kWideBlock, o1 = min(o0 + kWideBlock, wj.rawsize), n_out = o1 - o0;
        const uint8_t *src = reinterpret_cast<const uint8_t *>(wj.src);
        uint8_t *dst = reinterpret_cast<uint8_t *>(wj.dst);
        const uint32_t slen = wj.slen;

        for (uint32_t i = tid; i < kWideIdx / 32; i += kWideThreads) heads[i] = 0;
        __syncthreads();

        // ---- item starts: one thread per source segment that feeds this block
        const uint32_t s0 = wa.blk_seg0[b], s1 = k + 1u < wj.nblk ? wa.blk_seg0[b + 1] : wj.nseg - 1u;
        if (tid <= s1 - s0) {
            const uint32_t s = s0 + tid;
            const uint2 ent = wa.seg_entry[wj.seg0 + s];
            uint32_t a = s * kWideSeg + ent.y, o = ent.x;
            const uint32_t send = min((s + 1u) * kWideSeg, slen);
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
                            if (off == 0u) sh_bad = 1u;
                            const uint32_t e = kWideWin + (lo - o0);
                            ptr[e] = (uint16_t)off;
                            atomicOr(&heads[e >> 5], 1u << (e & 31u));
                        }
                        o += len;
                    } else {
                        if (o >= o0) {
                            const uint32_t e = kWideWin + (o - o0);
                            val[e] = __ldg(src + q);
                            ptr[e] = 0;
                            atomicOr(&heads[e >> 5], 1u << (e & 31u));
                        }
                        o += 1u;
                        q += 1u;
                    }
                }
                a = q;
            }
        }
        __syncthreads();

        // ---- every other byte of a match: point at the byte it copies (always before its own item)
        const uint32_t floor_idx = k == 0u ? kWideWin : 0u;                  // nothing exists before the output start
#pragma unroll 4
        for (uint32_t i = 0; i < kWidePer; i++) {
            const uint32_t e = i * kWideThreads + tid;
            if (e < kWideWin || e >= kWideWin + n_out || wide_is_head(heads, e)) continue;
            const uint32_t h = wide_head_of(heads, e);
            if (h == kNone) { sh_bad = 1u; ptr[e] = (uint16_t)e; continue; }  // cannot happen on a stream k_wide_chain accepted
            const uint32_t off = ptr[h], kk = e - h;
            uint32_t sidx;
            if (off > kk) sidx = e - off;
            else if (off <= 1u) sidx = h - 1u;                                // runs of one byte value
            else sidx = h - off + kk % off;                                   // overlapping copy (ppr:216-219)
            if (sidx < floor_idx || off == 0u) { sh_bad = 1u; sidx = e; }
            ptr[e] = (uint16_t)sidx;
        }
        __syncthreads();
        // ---- item starts themselves, and the previous tail
#pragma unroll 4
        for (uint32_t i = 0; i < kWidePer; i++) {
            const uint32_t e = i * kWideThreads + tid;
            if (e < kWideWin) {
                ptr[e] = (uint16_t)e;
                if (kMode == 3) val[e] = k > 0u ? dst[o0 - kWideWin + e] : (uint8_t)0;
            } else if (e < kWideWin + n_out && wide_is_head(heads, e)) {
                const uint32_t off = ptr[e];
                uint32_t sidx = e - off;                                      // literal: off == 0
                if (off && sidx < floor_idx) { sh_bad = 1u; sidx = e; }
                ptr[e] = (uint16_t)sidx;
            }
        }
        __syncthreads();

        // ---- pointer jumping until every byte points at a literal of the block or into the previous tail
        unsigned long long done = 0;
#pragma unroll 4
        for (uint32_t i = 0; i < kWidePer; i++) {
            const uint32_t e = i * kWideThreads + tid;
            if (e < kWideWin || e >= kWideWin + n_out) done |= 1ull << i;
        }
        for (;;) {
            int changed = 0;
#pragma unroll 4
            for (uint32_t i = 0; i < kWidePer; i++) {
                if ((done >> i) & 1ull) continue;
                const uint32_t e = i * kWideThreads + tid;
                const uint32_t p = ptr[e];
                const uint32_t q = ptr[p];
                if (q == p) done |= 1ull << i;
                else { ptr[e] = (uint16_t)q; changed = 1; }
            }
            if (!__syncthreads_or(changed)) break;
        }
        if (sh_bad) {                                                        // offset 0 or before the output start
            if (tid == 0) wide_fail(wa, jn);
            continue;
        }

        if (kMode == 1) {
            // the block's last 4095 bytes, each as "value" or "index into the previous tail"
            uint16_t *map = reinterpret_cast<uint16_t *>((reinterpret_cast<uintptr_t>(dst + o0) + 1u) & ~(uintptr_t)1u);
            for (uint32_t i = tid; i < kWideWin; i += kWideThreads) {
                const uint32_t r = ptr[kWideWin + kWideBlock - kWideWin + i];
                map[i] = (uint16_t)(r >= kWideWin ? 0x8000u | val[r] : r);
            }
        } else {
            uint8_t *g = dst + o0;
            const uint32_t head = min(n_out, (uint32_t)((4u - (reinterpret_cast<uintptr_t>(g) & 3u)) & 3u));
            if (tid < head) g[tid] = val[ptr[kWideWin + tid]];
            const uint32_t nw = (n_out - head) >> 2;
            for (uint32_t w = tid; w < nw; w += kWideThreads) {
                const uint32_t e0 = kWideWin + head + 4u * w;
                const uint32_t x = (uint32_t)val[ptr[e0]] | ((uint32_t)val[ptr[e0 + 1u]] << 8) |
                                   ((uint32_t)val[ptr[e0 + 2u]] << 16) | ((uint32_t)val[ptr[e0 + 3u]] << 24);
                __stcs(reinterpret_cast<uint32_t *>(g + head + 4u * w), x);
            }
            const uint32_t t0 = head + 4u * nw;
            if (tid < n_out - t0) g[t0 + tid] = val[ptr[kWideWin + t0 + tid]];
        }
    }
}

// One CTA per value: run down its blocks turning the tail maps of pass 1 into tail bytes, written where
// pass 3 of the next block expects them (the end of each block's own output region).
__global__ void __launch_bounds__(1024)
k_wide_tails(WideArea wa)
{
    __shared__ uint8_t tail[2][4096];
    const uint32_t n_jobs = wa.state->n_jobs, tid = threadIdx.x;
    for (uint32_t jn = blockIdx.x; jn < n_jobs; jn += gridDim.x) {
        const WideJob wj = wa.jobs[jn];
        if (wj.state != 0u || wj.nblk < 2u) continue;
        uint8_t *dst = reinterpret_cast<uint8_t *>(wj.dst);
        __syncthreads();
        for (uint32_t i = tid; i < 4096; i += blockDim.x) tail[0][i] = 0;
        __syncthreads();
        uint32_t cur = 0;
        for (uint32_t k = 0; k + 1u < wj.nblk; k++) {
            const size_t o0 = (size_t)k * kWideBlock;
            const uint16_t *map = reinterpret_cast<const uint16_t *>((reinterpret_cast<uintptr_t>(dst + o0) + 1u) & ~(uintptr_t)1u);
            uint8_t *out = dst + o0 + kWideBlock - kWideWin;
            for (uint32_t i = tid; i < kWideWin; i += blockDim.x) {
                const uint32_t m = map[i];
                const uint8_t v = (m & 0x8000u) ? (uint8_t)m : tail[cur][m & 4095u];
                tail[cur ^ 1u][i] = v;
                out[i] = v;
            }
            __syncthreads();
            cur ^= 1u;
        }
    }
}

}  // namespace cds
