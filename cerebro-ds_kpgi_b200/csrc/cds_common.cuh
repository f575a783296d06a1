// cds_common.cuh — shared device-side types and helpers for libcds_da.so (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "cds_da.h"

namespace cds {
namespace {   // internal linkage: with relocatable device code (-rdc) the optimiser then treats the file as a whole program

constexpr uint32_t kPage = CDS_BLCKSZ;          // 32 KB Greenplum page           (ppr:34)
constexpr uint32_t kPageHdr = 24;               // PAGE_HEADER_LEN                (ppr:36)
constexpr uint32_t kItemHdr = 23;               // ITEM_HEADER_LEN                (ppr:40)
constexpr uint32_t kChunkMax = CDS_TOAST_MAX_CHUNK_SIZE;
constexpr uint32_t kRowWords = CDS_ROW_WORDS;
constexpr uint32_t kNone = 0xFFFFFFFFu;
constexpr unsigned long long kNoErr = ~0ull;
constexpr uint32_t kFlagRowOverflow = 1u, kFlagChunkOverflow = 2u;
constexpr uint32_t kFlagStreamRedo = 8u;  // the streamed decode assumed a plan that the complete directory does not bear out (or met an error): cds_finish decodes again
constexpr uint32_t kFlagNeedWide = 4u;   // the classification took values for the wide decoder, whose kernels were not enqueued (cds_finish runs them)

// row words (cds_export_rows layout)
enum RowWord : uint32_t {
    RW_DIST_KEY = 0, RW_BUFFER_ID = 1,
    RW_IND = 2,   // +0 header, +1 va_rawsize, +2 va_extsize, +3 va_valueid, +4 va_toastrelid
    RW_DEP = 7,
    RW_DEP_EXTERNAL = 12, RW_INLINE_OFF_LO = 13, RW_INLINE_OFF_HI = 14, RW_INLINE_LEN = 15
};

// Device-resident scalars of one decode; mirrored to pinned host memory by cds_finish.
// A decode starts from an all-zero state (one cudaMemsetAsync).
struct DevState {
    unsigned long long err_inv;   // ~(smallest error key seen), see make_err_key; 0 = no error
    unsigned long long ind_bytes;
    unsigned long long dep_bytes;
    unsigned long long scratch_bytes;
    uint32_t n_rows;
    uint32_t n_chunks;
    uint32_t n_inline;
    uint32_t n_pos;
    uint32_t flags;               // kFlagRowOverflow | kFlagChunkOverflow
    uint32_t pglz_ticket;         // work counter of k_pglz_decode
    uint32_t n_jobs_front;        // compacted pglz jobs: large values from slot 0 up ...
    uint32_t n_jobs_back;         // ... small values from the last slot down
    uint32_t n_wide;              // compressed values at or above the wide-decode threshold (pglz_wide.cuh)
    uint32_t n_wide_taken;        // values the wide decoder accepted, summed over the classifications of this decode
    uint32_t n_wide_blks;         // most output blocks any one of those classifications handed to the wide decoder
    uint32_t pad_[1];
};

// One TOAST chunk -> one copy.  dst == 0: chunk not referenced (or refused); nothing is copied.
constexpr uint32_t kJobToBf16 = 0x80000000u;  // PglzJob.slen flag, same meaning
constexpr uint32_t kOpToBf16 = 0x80000000u;   // CopyOp.len flag: payload is float32, destination is bfloat16
constexpr uint32_t kOpU8Norm = 0x40000000u;   // CopyOp.len flag: payload is uint8 pixels, destination float32 = lut[channel][x];
                                              // CopyOp.src_off bits 16.. hold the channel of the chunk's first byte
constexpr uint32_t kJobU8Norm = 0x40000000u;  // PglzJob.slen flag, same meaning
constexpr uint32_t kFlagBits = 0xC0000000u;
// CopyOp.len flag: the chunk holds part of a LITERAL-ONLY pglz stream (groups of one zero control byte + 8 literals;
// slen == rawsize + ceil(rawsize / 8), so no tag can exist in a well-formed stream): the gather drops the control
// bytes on the way (9 -> 8 bytes) and checks that they are zero, no staging copy and no pglz job.  len & 0xFFFF =
// stream bytes of this chunk; src_off bits 16..19 = stream offset of the first byte mod 9, bits 20..22 = items of
// the value's final group when its control byte lies in this chunk and the group is short (its unused control bits
// are never looked at by the reference, ppr:203), else 0.
constexpr uint32_t kOpLiteral = 0x20000000u;
__host__ __device__ __forceinline__ bool pglz_literal_only(unsigned long long slen, unsigned long long rawsize)
{
    return rawsize > 0ull && slen == rawsize + ((rawsize + 7ull) >> 3);
}
struct __align__(16) CopyOp {
    unsigned long long dst;   // absolute device address
    uint32_t src_off;         // payload offset inside its page
    uint32_t len;             // payload bytes (| kOpToBf16)
};

// One pglz datum.  slen == kNone: empty slot.
struct __align__(16) PglzJob {
    unsigned long long src;       // first stream byte (after the 8-byte header / 4-byte rawsize)
    unsigned long long dst;
    uint32_t slen;                // stream bytes (| kJobToBf16: decoded float32 leaves as bfloat16)
    uint32_t rawsize;             // bytes to produce
    unsigned long long err_key;   // key to publish when the stream is corrupt (0: none)
};

// Error keys order errors the way the reference would meet them:
// stage (heap scan < TOAST scan < inline rows < external values < plan), then where in that
// stage, then which check.  The smallest key wins (atomicMin).
__host__ __device__ inline unsigned long long make_err_key(uint32_t stage, unsigned long long primary,
                                                           uint32_t sub, uint32_t secondary, uint32_t code)
{
    if (secondary > 0x1FFFFu) secondary = 0x1FFFFu;
    return ((unsigned long long)(stage & 7u) << 61) | ((primary & 0x1FFFFFFFFull) << 28) |
           ((unsigned long long)(sub & 7u) << 25) | ((unsigned long long)secondary << 8) | (code & 0xFFu);
}

__device__ inline void report(DevState *st, uint32_t stage, unsigned long long primary, uint32_t sub,
                              uint32_t secondary, uint32_t code)
{
    atomicMax(&st->err_inv, ~make_err_key(stage, primary, sub, secondary, code));
}
__device__ inline void report_key(DevState *st, unsigned long long key) { atomicMax(&st->err_inv, ~key); }

__device__ inline uint32_t ld_u16(const uint8_t *p) { return (uint32_t)p[0] | ((uint32_t)p[1] << 8); }
__device__ inline uint32_t ld_u32(const uint8_t *p)
{
    if ((reinterpret_cast<uintptr_t>(p) & 3u) == 0) return *reinterpret_cast<const uint32_t *>(p);
    return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);
}
__device__ inline uint32_t ld_u32be(const uint8_t *p) { return __byte_perm(ld_u32(p), 0, 0x0123); }

// valueid -> value index hash table: 64-bit entries, key in the high word, min(value ref) in the low.
// value ref = (garbage << 31) | (2*row + att); garbage marks the bytes the reference parses as a
// va_valueid out of an inline datum (ppr:342,516).
constexpr unsigned long long kEmptySlot = 0ull;   // entries are stored +1 so that a zeroed table is empty
__device__ inline uint32_t hash_u32(uint32_t x)
{
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    return x;
}
__device__ inline void htab_insert(unsigned long long *tab, uint32_t mask, uint32_t key, uint32_t ref)
{
    unsigned long long entry = (((unsigned long long)key << 32) | ref) + 1ull;
    uint32_t slot = hash_u32(key) & mask;
    for (;;) {
        unsigned long long cur = tab[slot];
        if (cur == kEmptySlot) {
            unsigned long long prev = atomicCAS(&tab[slot], kEmptySlot, entry);
            if (prev == kEmptySlot) return;
            cur = prev;
        }
        if ((uint32_t)((cur - 1ull) >> 32) == key) { atomicMin(&tab[slot], entry); return; }
        slot = (slot + 1) & mask;
    }
}
__device__ inline uint32_t htab_find(const unsigned long long *tab, uint32_t mask, uint32_t key)
{
    uint32_t slot = hash_u32(key) & mask;
    for (;;) {
        unsigned long long cur = tab[slot];
        if (cur == kEmptySlot) return kNone;
        if ((uint32_t)((cur - 1ull) >> 32) == key) return (uint32_t)(cur - 1ull);
        slot = (slot + 1) & mask;
    }
}

// ---- device-wide exclusive scan: single pass, decoupled look-back ---------------------------------
// Tiles of kScanTile elements are handed out by a ticket counter (so a tile only ever waits for
// tiles whose blocks are already running); each tile publishes its aggregate, then looks back over
// its predecessors' aggregates / inclusive prefixes, 32 at a time.  The workspace (ticket + status
// words) must be zero when the kernel starts; one cudaMemsetAsync per decode clears all of them.
constexpr int kScanThreads = 256;
constexpr uint32_t kScanMinTile = kScanThreads * 4;   // smallest tile any scan uses (sizes the workspaces)

struct ScanWs {
    uint32_t *ticket;             // 1 word, zeroed
    uint32_t *status;             // per tile: 0 nothing, 1 aggregate published, 2 inclusive prefix published
    unsigned long long *agg;      // per tile
    unsigned long long *incl;     // per tile
};

template <typename T>
__device__ inline T warp_inclusive_scan(T v)
{
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        T o = __shfl_up_sync(0xffffffffu, v, d);
        if ((threadIdx.x & 31) >= d) v += o;
    }
    return v;
}

__device__ inline uint32_t ld_volatile_u32(const uint32_t *p) { return *reinterpret_cast<const volatile uint32_t *>(p); }
__device__ inline unsigned long long ld_volatile_u64(const unsigned long long *p)
{
    return *reinterpret_cast<const volatile unsigned long long *>(p);
}

// out[i] = sum_{j<i} load(j) for i in [0, n); out[n] = total when write_total_at_n.  Every block of
// the launch must call this (blockDim.x == kScanThreads).  Returns the grand total to the block that
// processed the last tile (others get 0) through *is_last.
template <typename T, int kScanItems, typename LoadFn>
__device__ inline T device_exclusive_scan(LoadFn load, T *out, uint32_t n, ScanWs ws, bool write_total_at_n, bool *is_last)
{
    constexpr uint32_t kScanTile = kScanThreads * kScanItems;
    __shared__ unsigned long long sh_warp[kScanThreads / 32];
    __shared__ unsigned long long sh_prefix;
    __shared__ uint32_t sh_tile;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t num_tiles = (n + kScanTile - 1) / kScanTile;
    *is_last = false;
    T grand = 0;
    if (n == 0) {
        // every participating block reports the (same) empty result: which blocks take part varies by caller
        if (threadIdx.x == 0 && write_total_at_n) out[0] = 0;
        *is_last = true;
        return 0;
    }
    for (;;) {
        if (threadIdx.x == 0) sh_tile = atomicAdd(ws.ticket, 1u);
        __syncthreads();
        const uint32_t tile = sh_tile;
        if (tile >= num_tiles) break;
        const uint32_t base = tile * kScanTile + threadIdx.x * kScanItems;
        T v[kScanItems];
        T tsum = 0;
#pragma unroll
        for (int k = 0; k < kScanItems; k++) {
            const uint32_t i = base + k;
            v[k] = i < n ? (T)load(i) : (T)0;
            tsum += v[k];
        }
        const T winc = warp_inclusive_scan(tsum);
        if (lane == 31) sh_warp[warp] = (unsigned long long)winc;
        __syncthreads();
        T warp_off = 0, total = 0;
#pragma unroll
        for (int w = 0; w < kScanThreads / 32; w++) {
            const T x = (T)sh_warp[w];
            if (w < (int)warp) warp_off += x;
            total += x;
        }
        if (warp == 0) {
            unsigned long long prefix = 0;
            if (tile == 0) {
                if (lane == 0) {
                    ws.incl[0] = (unsigned long long)total;
                    __threadfence();
                    *reinterpret_cast<volatile uint32_t *>(&ws.status[0]) = 2u;
                }
            } else {
                if (lane == 0) {
                    ws.agg[tile] = (unsigned long long)total;
                    __threadfence();
                    *reinterpret_cast<volatile uint32_t *>(&ws.status[tile]) = 1u;
                }
                int j = (int)tile - 1;
                for (;;) {
                    const int t = j - (int)lane;
                    uint32_t st = 2u;
                    unsigned long long val = 0;
                    if (t >= 0) {
                        do { st = ld_volatile_u32(&ws.status[t]); } while (st == 0u);
                        __threadfence();
                        val = st == 2u ? ld_volatile_u64(&ws.incl[t]) : ld_volatile_u64(&ws.agg[t]);
                    }
                    const uint32_t done = __ballot_sync(0xffffffffu, st == 2u);   // lanes past tile 0 count as done
                    const uint32_t first = (uint32_t)__ffs((int)done) - 1u;       // nearest tile with a full prefix
                    unsigned long long part = lane <= first ? val : 0ull;
#pragma unroll
                    for (int d = 16; d > 0; d >>= 1) part += __shfl_xor_sync(0xffffffffu, part, d);
                    prefix += part;
                    if (done) break;
                    j -= 32;
                }
                if (lane == 0) {
                    ws.incl[tile] = prefix + (unsigned long long)total;
                    __threadfence();
                    *reinterpret_cast<volatile uint32_t *>(&ws.status[tile]) = 2u;
                }
            }
            if (lane == 0) sh_prefix = prefix;
        }
        __syncthreads();
        T run = (T)sh_prefix + warp_off + (winc - tsum);
#pragma unroll
        for (int k = 0; k < kScanItems; k++) {
            const uint32_t i = base + k;
            if (i < n) out[i] = run;
            run += v[k];
        }
        if (tile == num_tiles - 1) {
            grand = (T)sh_prefix + total;
            *is_last = true;
            if (write_total_at_n && threadIdx.x == 0) out[n] = grand;
        }
        __syncthreads();
    }
    return grand;
}

}  // namespace
}  // namespace cds
