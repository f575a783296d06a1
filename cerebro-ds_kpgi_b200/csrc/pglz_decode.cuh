// pglz_decode.cuh — warp-per-datum pglz decoder.
//
// Replaces pglz_decompress_py (/root/reference/cerebro_gpdb/pg_page_reader.py:191-231), which is
// the same algorithm as pglz_decompress in /root/reference/cerebro_gpdb/pg_lzcompress.c:49-143.
//
// pglz is sequential inside a datum (a match may point at any of the previous 4095 output
// bytes), so parallelism comes from many datums: one warp owns one datum.
//   * The warp keeps the last 8 KB of output in a shared-memory ring; matches are served from it
//     (never from global memory), and finished bytes leave for HBM as aligned 16-byte stores.
//   * The source stream is read through a register window: every lane holds one aligned 32-bit
//     word of the current and of the next 128 bytes (coalesced 128-byte loads, the next window is
//     requested a whole window ahead of its use), so fetching a control group is two shuffles, not
//     a dependent global load.  Lane l then holds source byte sp+l (a group is <= 25 bytes);
//     `__ballot_sync` of "(byte & 0x0F) == 0x0F" tells every lane where 3-byte tags are, so the
//     item walk of the group is register-only and identical in all lanes.
//   * Up to four control groups (32 items) are laid out and decoded lane-parallel per step; short items are then
//     expanded one lane per item (dependency rounds when a match copies bytes of the same step), long ones item
//     after item with aligned 16-byte fills and copies.
//   * Streaks of all-literal groups (control byte 0: incompressible data) are found with one
//     ballot over the next 32 candidate control bytes and copied 256 bytes at a time, with the
//     loads of the following 288 source bytes already in flight.
//   * Datums are handed to warps through a ticket counter, large values first; when there are fewer datums than
//     warps the surplus warps of every CTA leave, so that the datums spread over all SMs.
//
// Corruption semantics follow the reference: the datum is corrupt unless decoding ends with
// dp == destend and sp == srcend (ppr:229-230).  Additionally reported as corrupt (the reference
// reads stale memory instead): tags that run past the source, offset 0, offsets before the
// start of the output.
#pragma once
#include <cuda_bf16.h>

#include "cds_common.cuh"

namespace cds {
namespace {   // internal linkage: with relocatable device code (-rdc) the optimiser then treats the file as a whole program

constexpr uint32_t kRing = 8192;              // bytes of output history per warp (power of two)
constexpr uint32_t kFlushAt = 2048;           // flush to HBM once this many bytes are pending
constexpr int kPglzWarps = 14;                // warps (datums) per CTA: two CTAs of 14 x 8 KB rings fill an SM's shared memory (28 datums per SM)
constexpr uint32_t kItemCopyMax = 20;         // matches up to this long are copied by one lane each
constexpr size_t kPglzSmemBytes = (size_t)kPglzWarps * kRing;

// Flush ring bytes for output positions [from, to) to global memory.  `shift` = dst & 15, so that
// ring index (pos + shift) has the alignment of the global address.
__device__ __forceinline__ void pglz_flush(const uint8_t *ring, uint8_t *dst, uint32_t shift, uint32_t from,
                                           uint32_t to, uint32_t lane)
{
    // head bytes up to the next 16-byte boundary of the global address
    uint32_t head = (16u - ((from + shift) & 15u)) & 15u;
    if (head > to - from) head = to - from;
    if (lane < head) dst[from + lane] = ring[(from + lane + shift) & (kRing - 1)];
    const uint32_t b0 = from + head;
    const uint32_t nvec = (to - b0) >> 4;
    for (uint32_t j = lane; j < nvec; j += 32) {
        const uint32_t pos = b0 + (j << 4);
        const uint4 v = *reinterpret_cast<const uint4 *>(ring + ((pos + shift) & (kRing - 1)));
        __stcs(reinterpret_cast<uint4 *>(dst + pos), v);
    }
    const uint32_t t0 = b0 + (nvec << 4);
    if (lane < to - t0) dst[t0 + lane] = ring[(t0 + lane + shift) & (kRing - 1)];
}

// bfloat16 variant: the ring holds float32 data, `dst` holds bfloat16; positions are float32 byte offsets
// (multiples of 4) and `shift` = (2 * dst) & 15, so that a 16-byte ring vector maps onto an aligned
// 8-byte store.
__device__ __forceinline__ void pglz_flush_bf16(const uint8_t *ring, uint8_t *dst, uint32_t shift, uint32_t from,
                                                uint32_t to, uint32_t lane)
{
    __nv_bfloat16 *dh = reinterpret_cast<__nv_bfloat16 *>(dst);
    uint32_t head = ((16u - ((from + shift) & 15u)) & 15u) >> 2;             // floats up to the next ring vector
    if (head > ((to - from) >> 2)) head = (to - from) >> 2;
    if (lane < head)
        dh[(from >> 2) + lane] = __float2bfloat16_rn(*reinterpret_cast<const float *>(ring + ((from + 4u * lane + shift) & (kRing - 1))));
    const uint32_t b0 = from + 4u * head;
    const uint32_t nvec = (to - b0) >> 4;
    for (uint32_t j = lane; j < nvec; j += 32) {
        const uint32_t pos = b0 + (j << 4);
        const float4 v = *reinterpret_cast<const float4 *>(ring + ((pos + shift) & (kRing - 1)));
        const __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
        uint2 q;
        q.x = *reinterpret_cast<const uint32_t *>(&lo);
        q.y = *reinterpret_cast<const uint32_t *>(&hi);
        __stcs(reinterpret_cast<uint2 *>(dst + (pos >> 1)), q);
    }
    const uint32_t t0 = b0 + (nvec << 4);
    if (lane < ((to - t0) >> 2))
        dh[(t0 >> 2) + lane] = __float2bfloat16_rn(*reinterpret_cast<const float *>(ring + ((t0 + 4u * lane + shift) & (kRing - 1))));
}

// uint8 -> float32 variant: the ring holds uint8 pixel components, `dst` holds float32 = lut[channel][x] with
// channel = position mod channels; `shift` = (dst >> 2) & 3 so that a 4-byte ring word maps onto an aligned 16-byte store.
__device__ __forceinline__ void pglz_flush_u8norm(const uint8_t *ring, float *dst, uint32_t shift, uint32_t from, uint32_t to,
                                                  uint32_t lane, const float *__restrict__ lut, uint32_t ch)
{
    uint32_t head = (4u - ((from + shift) & 3u)) & 3u;
    if (head > to - from) head = to - from;
    if (lane < head) dst[from + lane] = __ldg(lut + ((from + lane) % ch) * 256u + ring[(from + lane + shift) & (kRing - 1)]);
    const uint32_t b0 = from + head;
    const uint32_t nq = (to - b0) >> 2;
    for (uint32_t j = lane; j < nq; j += 32) {
        const uint32_t pos = b0 + (j << 2);
        const uint32_t w = *reinterpret_cast<const uint32_t *>(ring + ((pos + shift) & (kRing - 1)));
        uint32_t c = pos % ch;
        float4 o;
        o.x = __ldg(lut + c * 256u + (w & 0xFFu));          c = (c + 1u == ch) ? 0u : c + 1u;
        o.y = __ldg(lut + c * 256u + ((w >> 8) & 0xFFu));   c = (c + 1u == ch) ? 0u : c + 1u;
        o.z = __ldg(lut + c * 256u + ((w >> 16) & 0xFFu));  c = (c + 1u == ch) ? 0u : c + 1u;
        o.w = __ldg(lut + c * 256u + (w >> 24));
        __stcs(reinterpret_cast<float4 *>(dst + pos), o);
    }
    const uint32_t t0 = b0 + (nq << 2);
    if (lane < to - t0) dst[t0 + lane] = __ldg(lut + ((t0 + lane) % ch) * 256u + ring[(t0 + lane + shift) & (kRing - 1)]);
}

// Register window over the source stream.  The stream is addressed in the word-aligned frame
// y = offset + (src & 3); window k covers y in [128k, 128k + 128); lane l holds word 32k + l.
struct SrcWindow {
    const uint32_t *wp;   // src rounded down to 4 bytes
    uint32_t mis;         // src & 3
    uint32_t nwords;      // words holding at least one stream byte
    uint32_t k;           // window held in `cur`
    uint32_t cur, nxt;
};

__device__ __forceinline__ uint32_t win_word(const SrcWindow &w, uint32_t idx)
{
    return idx < w.nwords ? __ldg(w.wp + idx) : 0u;
}
__device__ __forceinline__ void win_seek(SrcWindow &w, uint32_t sp, uint32_t lane)
{
    w.k = (sp + w.mis) >> 7;
    w.cur = win_word(w, 32u * w.k + lane);
    w.nxt = win_word(w, 32u * (w.k + 1u) + lane);
}
// Source byte sp + lane (garbage past the end of the stream: callers mask with slen).
__device__ __forceinline__ uint32_t win_fetch(SrcWindow &w, uint32_t sp, uint32_t lane)
{
    const uint32_t y = sp + w.mis, kk = y >> 7;
    if (kk != w.k) {
        if (kk == w.k + 1u) {
            w.cur = w.nxt;
            w.k = kk;
            w.nxt = win_word(w, 32u * (kk + 1u) + lane);      // needed 128 source bytes from now
        } else {
            win_seek(w, sp, lane);
        }
    }
    const uint32_t rel = (y & 127u) + lane;                      // < 160
    const uint32_t a = __shfl_sync(0xffffffffu, w.cur, (rel >> 2) & 31u);
    const uint32_t b = __shfl_sync(0xffffffffu, w.nxt, (rel >> 2) & 31u);
    return ((rel < 128u ? a : b) >> ((rel & 3u) * 8u)) & 0xFFu;
}

// Source bytes sp + lane (returned) and sp + 32 + lane (`second`); same caveat past the end of the stream.
__device__ __forceinline__ uint32_t win_fetch2(SrcWindow &w, uint32_t sp, uint32_t lane, uint32_t &second)
{
    const uint32_t y = sp + w.mis, kk = y >> 7;
    if (kk != w.k) {
        if (kk == w.k + 1u) {
            w.cur = w.nxt;
            w.k = kk;
            w.nxt = win_word(w, 32u * (kk + 1u) + lane);      // needed 128 source bytes from now
        } else {
            win_seek(w, sp, lane);
        }
    }
    const uint32_t rel = (y & 127u) + lane, rel2 = rel + 32u;   // < 160, < 192: inside cur + nxt
    const uint32_t a = __shfl_sync(0xffffffffu, w.cur, (rel >> 2) & 31u);
    const uint32_t b = __shfl_sync(0xffffffffu, w.nxt, (rel >> 2) & 31u);
    const uint32_t c = __shfl_sync(0xffffffffu, w.cur, (rel2 >> 2) & 31u);
    const uint32_t d = __shfl_sync(0xffffffffu, w.nxt, (rel2 >> 2) & 31u);
    second = ((rel2 < 128u ? c : d) >> ((rel2 & 3u) * 8u)) & 0xFFu;
    return ((rel < 128u ? a : b) >> ((rel & 3u) * 8u)) & 0xFFu;
}

// The three aligned words that cover the 9 source bytes of the group at sp + 9*lane.
__device__ __forceinline__ void lit_load(const SrcWindow &w, uint32_t sp, uint32_t lane, uint32_t &w0, uint32_t &w1, uint32_t &w2)
{
    const uint32_t i0 = (sp + w.mis + 9u * lane) >> 2;
    w0 = win_word(w, i0);
    w1 = win_word(w, i0 + 1u);
    w2 = win_word(w, i0 + 2u);
}

// Expand up to 32 items, one per lane (lane i < nitems holds item i in stream order: a literal with
// its byte in `litbyte`, or a match (len, moff)).  Checks every item the way the reference would meet
// it; returns false when the datum is corrupt.  When the items produce more than `max_total` bytes
// nothing is written and *too_big is set.  Otherwise the bytes land in the ring and dp advances.
__device__ __forceinline__ bool pglz_expand_items(uint32_t nitems, bool is_match, uint32_t len, uint32_t moff,
                                                  uint32_t litbyte, uint32_t &dp, uint32_t rawsize, uint8_t *ring,
                                                  uint32_t shift, uint32_t lane, uint32_t max_total, bool *too_big)
{
    const bool is_item = lane < nitems;
    is_match = is_match && is_item;
    const uint32_t olen = is_item ? (is_match ? len : 1u) : 0u;
    uint32_t incl = olen;                                                    // inclusive scan of output lengths
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t o = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= (uint32_t)d) incl += o;
    }
    const uint32_t my_o = incl - olen;                                       // output offset of my item in this step
    const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
    *too_big = total > max_total;
    if (*too_big) return true;
    // Any item the reference would refuse ends in "corrupt" (ppr:213-215, 221-222 then 229-230):
    // a literal or a match that does not fit the output, offset 0, offset before the output start.
    const bool bad = is_item && ((dp + my_o + olen > rawsize) || (is_match && (moff == 0u || moff > dp + my_o)));
    if (__ballot_sync(0xffffffffu, bad)) return false;

    constexpr uint32_t M_ = kRing - 1;
    if (total <= 128u && __ballot_sync(0xffffffffu, is_match && len > kItemCopyMax) == 0u) {
        // ---- short items, none longer than kItemCopyMax: one lane per ITEM.  A literal lane stores its byte; a match lane
        //      copies its own bytes in order (so a copy that overlaps itself needs no modulo: it re-reads what it wrote).
        //      A match whose source lies inside this step's output waits until the items that produce those bytes are
        //      done: rounds over a "not done" ballot, the items it needs found through a bitmap of item starts.
        const uint32_t d0 = dp + my_o + shift;                               // ring position of my first byte
        if (is_item && !is_match) ring[d0 & M_] = (uint8_t)litbyte;
        const uint32_t ext = min(len, moff);                                 // source bytes that are not my own output
        const bool dep = is_match && my_o + ext > moff;                      // ... reach into this step's output
        uint32_t need = 0;                                                   // items (bit = lane) that produce them
        if (__ballot_sync(0xffffffffu, dep)) {
            const uint32_t wsel = my_o >> 5, bit = 1u << (my_o & 31u);
            const uint32_t s0 = __reduce_or_sync(0xffffffffu, (is_item && wsel == 0u) ? bit : 0u);
            const uint32_t s1 = __reduce_or_sync(0xffffffffu, (is_item && wsel == 1u) ? bit : 0u);
            const uint32_t s2 = __reduce_or_sync(0xffffffffu, (is_item && wsel == 2u) ? bit : 0u);
            const uint32_t s3 = __reduce_or_sync(0xffffffffu, (is_item && wsel == 3u) ? bit : 0u);
            const uint32_t p1 = __popc(s0), p2 = p1 + __popc(s1), p3 = p2 + __popc(s2);
            auto item_at = [&](uint32_t x) -> uint32_t {                     // index of the item that holds output offset x of the step
                const uint32_t w = x >> 5, m = 0xFFFFFFFFu >> (31u - (x & 31u));
                const uint32_t sw = w == 0u ? s0 : w == 1u ? s1 : w == 2u ? s2 : s3;
                const uint32_t below = w == 0u ? 0u : w == 1u ? p1 : w == 2u ? p2 : p3;
                return below + __popc(sw & m) - 1u;
            };
            if (dep) {
                const uint32_t lo = my_o > moff ? my_o - moff : 0u, hi = my_o + ext - moff - 1u;     // lo <= hi < my_o
                const uint32_t ja = item_at(lo), jb = item_at(hi);
                need = (0xFFFFFFFFu >> (31u - jb)) & (0xFFFFFFFFu << ja);
            }
        }
        bool done = !is_match;
        __syncwarp();
        for (;;) {
            const uint32_t undone = __ballot_sync(0xffffffffu, !done);
            if (undone == 0u) break;
            if (!done && (need & undone) == 0u) {
                const uint32_t sr = d0 - moff;
                if (((d0 | sr | len) & 3u) == 0u) {                          // whole aligned words
                    for (uint32_t k = 0; k < len; k += 4u)
                        *reinterpret_cast<uint32_t *>(ring + ((d0 + k) & M_)) = *reinterpret_cast<const uint32_t *>(ring + ((sr + k) & M_));
                } else if (moff >= len && len <= 8u) {
                    // up to 8 bytes that do not overlap their source: three aligned words re-aligned by funnel shifts, then
                    // straight-line predicated byte stores (no loop: the lanes of a warp sit in items of different lengths)
                    const uint32_t sa = sr & ~3u, sh = (sr & 3u) * 8u;
                    const uint32_t x0 = *reinterpret_cast<const uint32_t *>(ring + (sa & M_));
                    const uint32_t x1 = *reinterpret_cast<const uint32_t *>(ring + ((sa + 4u) & M_));
                    const uint32_t x2 = *reinterpret_cast<const uint32_t *>(ring + ((sa + 8u) & M_));
                    const uint32_t lo = __funnelshift_r(x0, x1, sh), hi = __funnelshift_r(x1, x2, sh);
#pragma unroll
                    for (uint32_t k = 0; k < 8u; k++)
                        if (k < len) ring[(d0 + k) & M_] = (uint8_t)((k < 4u ? lo : hi) >> (8u * (k & 3u)));
                } else {
                    for (uint32_t k = 0; k < len; k++) ring[(d0 + k) & M_] = ring[(sr + k) & M_];
                }
                done = true;
            }
            __syncwarp();
        }
    } else {
        // ---- long matches (or short items with a long one among them): item after item.  (Here total <= 2184 or <= 2048, so
        //      offsets fit 12 bits.)
        const uint32_t mbits = __ballot_sync(0xffffffffu, is_match);
        const uint32_t runbits = __ballot_sync(0xffffffffu, is_match && moff == 1u);
        const uint32_t packed = my_o | (len << 12);
        constexpr uint32_t M = kRing - 1;
        for (uint32_t k = 0; k < nitems; k++) {
            const uint32_t pk = __shfl_sync(0xffffffffu, packed, k);
            const uint32_t at = dp + (pk & 0xFFFu);
            if ((mbits >> k) & 1u) {
                uint32_t klen = pk >> 12;
                const uint32_t a0 = at + shift;                              // ring position of the first byte written
                if ((runbits >> k) & 1u) {
                    // consecutive runs repeat the same byte: one fill for all of them
                    const uint32_t more = k < 31u ? (uint32_t)__ffs((int)~(runbits >> (k + 1u))) - 1u : 0u;   // run items right after item k
                    if (more) {
                        const uint32_t last = k + more;                      // < nitems: runbits only has item lanes
                        const uint32_t pl = __shfl_sync(0xffffffffu, packed, last);
                        klen = (pl & 0xFFFu) + (pl >> 12) - (pk & 0xFFFu);
                        k = last;
                    }
                    // run of one byte value (zero runs of sparse features): aligned 16-byte stores, at most 18 of them.
                    // The last one may run up to 15 bytes past the item: those ring bytes belong to output that is not
                    // produced yet (they alias history 8 KB back, beyond the 4095-byte window and already flushed).
                    const uint32_t b = ring[(a0 - 1u) & M], v = b * 0x01010101u;
                    const uint32_t head = min(klen, (16u - (a0 & 15u)) & 15u);
                    if (lane < head) ring[(a0 + lane) & M] = (uint8_t)b;
                    const uint32_t nq = (klen - head + 15u) >> 4, q0 = a0 + head;
                    for (uint32_t i = lane; i < nq; i += 32)
                        *reinterpret_cast<uint4 *>(ring + ((q0 + 16u * i) & M)) = make_uint4(v, v, v, v);
                    __syncwarp();
                    continue;
                }
                const uint32_t kmoff = __shfl_sync(0xffffffffu, moff, k);
                if (kmoff >= klen && klen >= 16u) {
                    // source and destination do not overlap: aligned 16-byte stores (one per lane covers 273 bytes in one
                    // go), the source re-aligned with funnel shifts of five aligned words.  The last vector may run up to 15
                    // bytes past the item, like the run fill above.
                    const uint32_t s0 = a0 - kmoff;
                    const uint32_t head = (16u - (a0 & 15u)) & 15u;               // < 16 <= klen
                    if (lane < head) ring[(a0 + lane) & M] = ring[(s0 + lane) & M];
                    const uint32_t nv = (klen - head + 15u) >> 4, v0 = a0 + head, sv = s0 + head, r8 = (sv & 3u) * 8u;
                    for (uint32_t i = lane; i < nv; i += 32) {
                        const uint32_t sa = (sv + 16u * i) & ~3u;
                        const uint32_t x0 = *reinterpret_cast<const uint32_t *>(ring + (sa & M));
                        const uint32_t x1 = *reinterpret_cast<const uint32_t *>(ring + ((sa + 4u) & M));
                        const uint32_t x2 = *reinterpret_cast<const uint32_t *>(ring + ((sa + 8u) & M));
                        const uint32_t x3 = *reinterpret_cast<const uint32_t *>(ring + ((sa + 12u) & M));
                        const uint32_t x4 = *reinterpret_cast<const uint32_t *>(ring + ((sa + 16u) & M));
                        *reinterpret_cast<uint4 *>(ring + ((v0 + 16u * i) & M)) =
                            make_uint4(__funnelshift_r(x0, x1, r8), __funnelshift_r(x1, x2, r8), __funnelshift_r(x2, x3, r8), __funnelshift_r(x3, x4, r8));
                    }
                } else if (kmoff >= klen && klen >= 8u) {
                    // source and destination do not overlap: aligned 32-bit stores, the source re-aligned with a funnel
                    // shift of two aligned words (bytes read beyond the source are shifted out)
                    const uint32_t s0 = a0 - kmoff;
                    const uint32_t head = (4u - (a0 & 3u)) & 3u;                  // < 8 <= klen
                    if (lane < head) ring[(a0 + lane) & M] = ring[(s0 + lane) & M];
                    // (the last word may run up to 3 bytes past the item, as above)
                    const uint32_t nw = (klen - head + 3u) >> 2, w0 = a0 + head, sw = s0 + head, r8 = (sw & 3u) * 8u;
                    for (uint32_t i = lane; i < nw; i += 32) {
                        const uint32_t sa = (sw + 4u * i) & ~3u;
                        const uint32_t lo = *reinterpret_cast<const uint32_t *>(ring + (sa & M));
                        const uint32_t hi = *reinterpret_cast<const uint32_t *>(ring + ((sa + 4u) & M));
                        *reinterpret_cast<uint32_t *>(ring + ((w0 + 4u * i) & M)) = __funnelshift_r(lo, hi, r8);
                    }
                } else if (kmoff >= klen) {
                    if (lane < klen) ring[(a0 + lane) & M] = ring[(a0 - kmoff + lane) & M];
                } else {
                    for (uint32_t i = lane; i < klen; i += 32)
                        ring[(a0 + i) & M] = ring[(a0 - kmoff + (i % kmoff)) & M];
                }
            } else if (lane == k) {
                ring[(at + shift) & M] = (uint8_t)litbyte;
            }
            __syncwarp();
        }
    }
    dp += total;
    return true;
}

// Decode one datum with the calling warp.  Returns true when the stream is well-formed.
// kOut: 0 = the decoded bytes, 1 = float32 -> bfloat16, 2 = uint8 -> float32 through `lut`.  (A run-time value: only the
// flush differs between the formats, and three copies of this function are 17 000 SASS instructions — more than the
// instruction cache holds.)
__device__ bool pglz_warp_decode(const uint32_t kOut, const uint8_t *__restrict__ src, uint32_t slen, uint8_t *dst, uint32_t rawsize,
                                 uint8_t *ring, const float *__restrict__ lut, uint32_t lut_channels)
{
    const uint32_t lane = threadIdx.x & 31;
    // ring index = output position + shift, chosen so that ring vectors line up with aligned global stores
    const uint32_t shift = kOut == 1 ? (uint32_t)((2u * reinterpret_cast<uintptr_t>(dst)) & 15u)
                         : kOut == 2 ? (uint32_t)((reinterpret_cast<uintptr_t>(dst) >> 2) & 3u)
                                     : (uint32_t)(reinterpret_cast<uintptr_t>(dst) & 15u);
    auto flush = [&](uint32_t from, uint32_t to) {
        if (kOut == 1) pglz_flush_bf16(ring, dst, shift, from, to, lane);
        else if (kOut == 2) pglz_flush_u8norm(ring, reinterpret_cast<float *>(dst), shift, from, to, lane, lut, lut_channels);
        else pglz_flush(ring, dst, shift, from, to, lane);
    };
    uint32_t sp = 0, dp = 0, flushed = 0;
    uint32_t wide_pause = 0, wide_backoff = 4;                               // steps to stay on the one-group path
    bool corrupt = false;

    SrcWindow win;
    win.mis = (uint32_t)(reinterpret_cast<uintptr_t>(src) & 3u);
    win.wp = reinterpret_cast<const uint32_t *>(src - win.mis);
    win.nwords = (slen + win.mis + 3u) >> 2;
    win_seek(win, 0, lane);

    while (sp < slen && dp < rawsize && !corrupt) {
        // ---- lane l holds source bytes sp + l (and sp + 32 + l for the wide step) ---------------
        uint32_t byte1;
        const uint32_t fetched = win_fetch2(win, sp, lane, byte1);
        const bool inb = sp + lane < slen;
        const uint32_t byte = inb ? fetched : 0u;
        const uint32_t ctrl = __shfl_sync(0xffffffffu, byte, 0);
        const uint32_t zmask = __ballot_sync(0xffffffffu, inb && byte == 0u);

        // ---- streak of all-literal groups (control byte 0, 8 literals, 9 source bytes each) ----
        if ((zmask & 0x00040201u) == 0x00040201u && sp + 27u <= slen && rawsize - dp >= 24u) {
            uint32_t w0, w1, w2;
            lit_load(win, sp, lane, w0, w1, w2);
            for (;;) {
                uint32_t n0, n1, n2;
                lit_load(win, sp + 288u, lane, n0, n1, n2);                 // speculative: the streak goes on
                const uint32_t y0 = sp + win.mis + 9u * lane, sh = (y0 & 3u) * 8u;
                const bool full = sp + 9u * lane + 9u <= slen;
                const uint32_t m = __ballot_sync(0xffffffffu, full && ((w0 >> sh) & 0xFFu) == 0u);
                uint32_t streak = (m == 0xffffffffu) ? 32u : (uint32_t)(__ffs((int)~m) - 1);
                const uint32_t room = (rawsize - dp) >> 3;                   // whole groups that still fit
                if (streak > room) streak = room;
                if (lane < streak) {
                    const uint32_t lo = (uint32_t)(((((unsigned long long)w1) << 32) | w0) >> (sh + 8u));
                    const uint32_t hi = (uint32_t)(((((unsigned long long)w2) << 32) | w1) >> (sh + 8u));
                    const uint32_t o = dp + 8u * lane + shift;
                    if ((o & 3u) == 0u) {
                        *reinterpret_cast<uint32_t *>(ring + (o & (kRing - 1))) = lo;
                        *reinterpret_cast<uint32_t *>(ring + ((o + 4u) & (kRing - 1))) = hi;
                    } else {
#pragma unroll
                        for (int k = 0; k < 4; k++) {
                            ring[(o + k) & (kRing - 1)] = (uint8_t)(lo >> (8 * k));
                            ring[(o + 4 + k) & (kRing - 1)] = (uint8_t)(hi >> (8 * k));
                        }
                    }
                }
                sp += 9u * streak;
                dp += 8u * streak;
                __syncwarp();
                if (dp - flushed >= kFlushAt) {
                    const uint32_t to = dp - ((dp + shift) & 15u);
                    flush(flushed, to);
                    flushed = to;
                    __syncwarp();
                }
                if (streak < 32u) break;
                w0 = n0; w1 = n1; w2 = n2;
            }
            win_seek(win, sp, lane);
            continue;
        }

        bool stepped = false;
        if (slen - sp >= 64u && wide_pause == 0) {
            // ---- wide step: up to four whole groups (32 items) at once.  Lane 8g+k owns item k of group g.
            // Item positions depend on which earlier tags carry a length-extension byte; start from
            // "none do" and iterate to the fixed point (exact once the ext mask reproduces itself).
            const uint32_t W = byte | (byte1 << 8);                          // source bytes sp+lane, sp+32+lane
            const uint32_t e_lo = __ballot_sync(0xffffffffu, (byte & 0x0Fu) == 0x0Fu);
            const uint32_t e_hi = __ballot_sync(0xffffffffu, (byte1 & 0x0Fu) == 0x0Fu);
            const uint32_t g = lane >> 3, k = lane & 7u, below = (1u << k) - 1u;
            uint32_t M = 0, my_p = 0, my_bit = 0, ng = 0, consumed = 0;
            bool conv = false;
            {
                // First try, "no tag carries an extension byte" (always true of short matches): a group that starts at
                // position p is 9 + popc(byte p) long whatever else is in it, so every lane knows where the group after
                // "a group at my position" starts, and the four group starts are three hops through those values.
                const uint32_t nx_lo = lane + 9u + __popc(byte), nx_hi = lane + 41u + __popc(byte1);
                const uint32_t S1 = __shfl_sync(0xffffffffu, nx_lo, 0);                       // 9..17
                const uint32_t S2 = __shfl_sync(0xffffffffu, nx_lo, S1);                      // 18..34
                const uint32_t a3 = __shfl_sync(0xffffffffu, nx_lo, S2 & 31u), b3 = __shfl_sync(0xffffffffu, nx_hi, S2 & 31u);
                const uint32_t S3 = S2 < 32u ? a3 : b3;                                       // 27..51
                const uint32_t a4 = __shfl_sync(0xffffffffu, nx_lo, S3 & 31u), b4 = __shfl_sync(0xffffffffu, nx_hi, S3 & 31u);
                ng = S3 <= 39u ? 4u : 3u;                                                     // whole groups inside the 64 bytes
                consumed = ng == 4u ? (S3 < 32u ? a4 : b4) : S3;
                const uint32_t myS = g == 0u ? 0u : g == 1u ? S1 : g == 2u ? S2 : S3;
                const uint32_t myC = (__shfl_sync(0xffffffffu, W, myS & 31u) >> ((myS >> 5) * 8u)) & 0xFFu;
                my_bit = (myC >> k) & 1u;
                my_p = myS + 1u + k + __popc(myC & below);
                const uint32_t ebit = ((my_p < 32u ? e_lo : e_hi) >> (my_p & 31u)) & 1u;
                M = __ballot_sync(0xffffffffu, g < ng && my_bit && ebit);
                conv = M == 0u;
            }
            for (int it = 1; it < 4 && !conv; it++) {
                uint32_t S = 0, myS = 0, myC = 0;
                ng = 0;
#pragma unroll
                for (uint32_t gg = 0; gg < 4; gg++) {
                    if (ng == gg && S <= 39u) {                              // the whole group lies inside the 64 bytes
                        const uint32_t c = (__shfl_sync(0xffffffffu, W, S & 31u) >> ((S >> 5) * 8u)) & 0xFFu;
                        if (g == gg) { myS = S; myC = c; }
                        S += 9u + __popc(c) + __popc((M >> (8u * gg)) & 0xFFu);
                        ng = gg + 1u;
                    }
                }
                consumed = S;
                my_bit = (myC >> k) & 1u;
                my_p = myS + 1u + k + __popc(myC & below) + __popc((M >> (8u * g)) & below);
                const uint32_t ebit = ((my_p < 32u ? e_lo : e_hi) >> (my_p & 31u)) & 1u;
                const uint32_t M2 = __ballot_sync(0xffffffffu, g < ng && my_bit && ebit);
                conv = (M2 == M);
                M = M2;
            }
            if (conv) {
                const uint32_t w0 = __shfl_sync(0xffffffffu, W, my_p & 31u);
                const uint32_t w1 = __shfl_sync(0xffffffffu, W, (my_p + 1u) & 31u);
                const uint32_t w2 = __shfl_sync(0xffffffffu, W, (my_p + 2u) & 31u);
                const uint32_t b0 = (w0 >> ((my_p >> 5) * 8u)) & 0xFFu;
                const uint32_t b1 = (w1 >> (((my_p + 1u) >> 5) * 8u)) & 0xFFu;
                const uint32_t b2 = (w2 >> (((my_p + 2u) >> 5) * 8u)) & 0xFFu;
                uint32_t len = (b0 & 0x0Fu) + 3u;                            // ppr:205
                const uint32_t moff = ((b0 & 0xF0u) << 4) | b1;              // ppr:206-208
                if ((b0 & 0x0Fu) == 0x0Fu) len += b2;                        // ppr:210-212
                bool too_big;
                if (!pglz_expand_items(8u * ng, my_bit != 0u, len, moff, b0, dp, rawsize, ring, shift, lane, 2048u, &too_big)) {
                    corrupt = true;
                    break;
                }
                if (!too_big) { sp += consumed; stepped = true; wide_backoff = 4; }
            }
            if (!stepped) {                                                  // long matches / many extension bytes here:
                wide_pause = wide_backoff;                                   // stay narrow for a while, longer each time
                wide_backoff = min(wide_backoff * 2u, 1024u);
            }
        } else if (wide_pause) {
            wide_pause--;
        }

        if (!stepped) {
            // ---- one group.  Item walk: register-only and identical in all lanes; item k starts at source
            //      offset p_k inside the group and lane k keeps p_k (ppr:203-228 restated).
            const uint32_t extmask = __ballot_sync(0xffffffffu, (byte & 0x0Fu) == 0x0Fu);
            const uint32_t n_avail = min(32u, slen - sp);                    // a group is at most 25 bytes
            uint32_t pnext = 1, my_p = 0, nitems = 0;
            if (ctrl == 0xFFu && n_avail >= 25u && ((extmask >> 1) & 0x249249u) == 0x249249u) {
                // eight tags with a length-extension byte each (long matches: sparse data): item k starts at 1 + 3k
                my_p = lane < 8u ? 1u + 3u * lane : 0u;
                pnext = 25u;
                nitems = 8u;
            } else {
#pragma unroll
                for (uint32_t k = 0; k < 8; k++) {
                    if (pnext < n_avail) {                                   // `sp < srcend` of the inner loop
                        const uint32_t sz = ((ctrl >> k) & 1u) ? 2u + ((extmask >> pnext) & 1u) : 1u;
                        if (lane == k) my_p = pnext;
                        pnext += sz;
                        nitems = k + 1;
                    }
                }
            }
            if (pnext > n_avail) { corrupt = true; break; }                  // a tag runs past the source
            const uint32_t b0 = __shfl_sync(0xffffffffu, byte, my_p);
            const uint32_t b1 = __shfl_sync(0xffffffffu, byte, (my_p + 1u) & 31u);
            const uint32_t b2 = __shfl_sync(0xffffffffu, byte, (my_p + 2u) & 31u);
            uint32_t len = (b0 & 0x0Fu) + 3u;                                // ppr:205
            const uint32_t moff = ((b0 & 0xF0u) << 4) | b1;                  // ppr:206-208
            if ((b0 & 0x0Fu) == 0x0Fu) len += b2;                            // ppr:210-212
            bool too_big;
            if (!pglz_expand_items(nitems, ((ctrl >> lane) & 1u) != 0u, len, moff, b0, dp, rawsize, ring, shift, lane,
                                   0xFFFFFFFFu, &too_big)) {
                corrupt = true;
                break;
            }
            sp += pnext;
        }
        if (dp - flushed >= kFlushAt) {
            const uint32_t to = dp - ((dp + shift) & 15u);
            flush(flushed, to);
            flushed = to;
            __syncwarp();
        }
    }
    if (!corrupt) {
        __syncwarp();
        flush(flushed, dp);
    }
    return !corrupt && dp == rawsize && sp == slen;                          // ppr:229-230
}

// Jobs are taken through *ticket (zero at launch).  With `st_counts` the job array is the compacted
// two-ended list k_resolve builds (n_jobs_front slots from the start, n_jobs_back from slot 2*n_rows-1
// down): front jobs (large values) are handed out first, the small ones fill the tail.
__global__ void __launch_bounds__(32 * kPglzWarps, 2)
k_pglz_decode(const PglzJob *__restrict__ jobs, const DevState *st_counts, uint32_t n_jobs_host,
              int32_t *status, DevState *st, uint32_t *ticket, const float *__restrict__ lut, uint32_t lut_channels,
              const uint32_t *d_first = nullptr, const uint32_t *d_end = nullptr)
{
    extern __shared__ __align__(16) uint8_t smem_ring[];
    uint8_t *ring = smem_ring + (size_t)(threadIdx.x >> 5) * kRing;
    const uint32_t lane = threadIdx.x & 31;
    // d_first, d_end (streamed decode): the jobs are slots [*d_first, *d_end) of a plain list
    const uint32_t first = d_first ? *d_first : 0u;
    const uint32_t front = (d_end ? *d_end : st_counts ? st_counts->n_jobs_front : n_jobs_host) - first;
    const uint32_t back = (st_counts && !d_first) ? st_counts->n_jobs_back : 0u;
    const uint32_t last_slot = st_counts ? 2u * st_counts->n_rows - 1u : 0u;
    const uint32_t n_jobs = front + back;
    if (n_jobs == 0) return;
    // Fewer jobs than warps (Criteo: 2 048 on 4 144): the warps that come up first would take them all and fill a few SMs
    // with 14 busy warps each while the other SMs idle.  Only as many warps per CTA as the jobs need stay, so the jobs
    // spread over all CTAs; nothing has to wait for the warps that leave.  (A batch of a streamed decode keeps them: it
    // should fill CTA after CTA, because a CTA holds its 112 KB of shared memory until its last warp is done.)
    if (!d_first && (threadIdx.x >> 5) * gridDim.x >= n_jobs) return;
    // Jobs are taken by ticket.  When a launch has few expensive jobs (fewer than half its warps: Criteo, 1 024 values on
    // 4 144 warps) the warps that come up first would take them all and fill a few SMs, so only the first `quota` warps of
    // every CTA take part in the first round; a batch of a streamed decode does the opposite and fills CTA after CTA,
    // because a CTA keeps its 112 KB of shared memory until its last warp is done and the next batch needs room on the SMs.
    for (;;) {
        uint32_t t = 0;
        if (lane == 0) t = atomicAdd(ticket, 1u);
        t = __shfl_sync(0xffffffffu, t, 0);
        if (t >= n_jobs) break;
        const uint32_t j = t < front ? first + t : last_slot - (t - front);
        const PglzJob job = jobs[j];
        if (job.slen == kNone) continue;
        const uint8_t *jsrc = reinterpret_cast<const uint8_t *>(job.src);
        uint8_t *jdst = reinterpret_cast<uint8_t *>(job.dst);
        const uint32_t jlen = job.slen & ~kFlagBits;
        const bool ok = pglz_warp_decode((job.slen & kJobToBf16) ? 1u : (job.slen & kJobU8Norm) ? 2u : 0u, jsrc, jlen, jdst, job.rawsize, ring, lut,
                                         lut_channels);
        if (lane == 0) {
            if (status) status[j] = ok ? CDS_OK : CDS_ERR_PGLZ_CORRUPT;
            if (!ok && st && job.err_key) report_key(st, job.err_key);
        }
        __syncwarp();
    }
}

}  // namespace
}  // namespace cds
