// cds_common.cuh — shared device-side types and helpers for libcds_da.so (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "cds_da.h"

namespace cds {

constexpr uint32_t kPage = CDS_BLCKSZ;          // 32 KB Greenplum page           (ppr:34)
constexpr uint32_t kPageHdr = 24;               // PAGE_HEADER_LEN                (ppr:36)
constexpr uint32_t kItemHdr = 23;               // ITEM_HEADER_LEN                (ppr:40)
constexpr uint32_t kChunkMax = CDS_TOAST_MAX_CHUNK_SIZE;
constexpr uint32_t kRowWords = CDS_ROW_WORDS;
constexpr uint32_t kNone = 0xFFFFFFFFu;
constexpr unsigned long long kNoErr = ~0ull;

// row words (cds_export_rows layout)
enum RowWord : uint32_t {
    RW_DIST_KEY = 0, RW_BUFFER_ID = 1,
    RW_IND = 2,   // +0 header, +1 va_rawsize, +2 va_extsize, +3 va_valueid, +4 va_toastrelid
    RW_DEP = 7,
    RW_DEP_EXTERNAL = 12, RW_INLINE_OFF_LO = 13, RW_INLINE_OFF_HI = 14, RW_INLINE_LEN = 15
};

// Device-resident scalars of one decode; mirrored to pinned host memory by cds_finish.
struct DevState {
    unsigned long long err_key;   // smallest error key seen (see make_err_key); kNoErr = none
    unsigned long long ind_bytes;
    unsigned long long dep_bytes;
    unsigned long long scratch_bytes;
    uint32_t n_rows;
    uint32_t n_chunks;
    uint32_t n_inline;
    uint32_t n_pos;
    uint32_t ind_sample_bytes;
    uint32_t dep_sample_bytes;
    uint32_t pad0, pad1;
};

// One TOAST chunk -> one copy.  dst == 0: chunk not referenced (or refused); nothing is copied.
struct __align__(16) CopyOp {
    unsigned long long dst;   // absolute device address
    uint32_t src_off;         // payload offset inside its page
    uint32_t len;             // payload bytes
};

// One pglz datum.  slen == kNone: empty slot.
struct __align__(16) PglzJob {
    unsigned long long src;       // first stream byte (after the 8-byte header / 4-byte rawsize)
    unsigned long long dst;
    uint32_t slen;                // stream bytes
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
    atomicMin(&st->err_key, make_err_key(stage, primary, sub, secondary, code));
}

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
constexpr unsigned long long kEmptySlot = ~0ull;
__device__ inline uint32_t hash_u32(uint32_t x)
{
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    return x;
}
__device__ inline void htab_insert(unsigned long long *tab, uint32_t mask, uint32_t key, uint32_t ref)
{
    unsigned long long entry = ((unsigned long long)key << 32) | ref;
    uint32_t slot = hash_u32(key) & mask;
    for (;;) {
        unsigned long long cur = tab[slot];
        if (cur == kEmptySlot) {
            unsigned long long prev = atomicCAS(&tab[slot], kEmptySlot, entry);
            if (prev == kEmptySlot) return;
            cur = prev;
        }
        if ((uint32_t)(cur >> 32) == key) { atomicMin(&tab[slot], entry); return; }
        slot = (slot + 1) & mask;
    }
}
__device__ inline uint32_t htab_find(const unsigned long long *tab, uint32_t mask, uint32_t key)
{
    uint32_t slot = hash_u32(key) & mask;
    for (;;) {
        unsigned long long cur = tab[slot];
        if (cur == kEmptySlot) return kNone;
        if ((uint32_t)(cur >> 32) == key) return (uint32_t)cur;
        slot = (slot + 1) & mask;
    }
}

// ---- block-wide exclusive scan over a long array by ONE block (plans are small) -------------
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

// out[i] = sum(in[0..i)), returns the grand total to every thread.  in/out may alias.
// Needs blockDim.x == 1024 and a __shared__ T[33] scratch.
template <typename T, typename TIn>
__device__ inline T block_exclusive_scan(const TIn *in, T *out, uint32_t n, T *sh)
{
    T carry = 0;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (uint32_t base = 0; base < n; base += 1024) {
        uint32_t i = base + threadIdx.x;
        T v = i < n ? (T)in[i] : (T)0;
        T inc = warp_inclusive_scan(v);
        if (lane == 31) sh[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            T w = sh[lane];
            T winc = warp_inclusive_scan(w);
            sh[lane] = winc - w;
            if (lane == 31) sh[32] = winc;
        }
        __syncthreads();
        if (i < n) out[i] = carry + sh[warp] + inc - v;
        carry += sh[32];
        __syncthreads();
    }
    return carry;
}

}  // namespace cds
