// pglz_decode.cuh — warp-per-datum pglz decoder.
//
// Replaces pglz_decompress_py (/root/reference/cerebro_gpdb/pg_page_reader.py:191-231), which is
// the same algorithm as pglz_decompress in /root/reference/cerebro_gpdb/pg_lzcompress.c:49-143.
//
// pglz is sequential inside a datum (a match may point at any of the previous 4095 output
// bytes), so parallelism comes from many datums: one warp owns one datum.
//   * The warp keeps the last 8 KB of output in a shared-memory ring; matches are served from it
//     (never from global memory), and finished bytes leave for HBM as aligned 16-byte stores.
//   * A control group (1 control byte + up to 8 items, <= 25 source bytes) is fetched with one
//     coalesced 32-byte read: lane l holds source byte sp+l.  `__ballot_sync` of
//     "(byte & 0x0F) == 0x0F" tells every lane where 3-byte tags are, so the item walk of the
//     group is register-only and identical in all lanes.
//   * Literal runs are written by the lanes that already hold those bytes; a match of length L
//     is expanded by min(L,32) lanes at a time (overlapping matches read modulo the offset).
//   * Streaks of all-literal groups (control byte 0: incompressible data) are found with one
//     ballot over the next 32 candidate control bytes and copied 256 bytes at a time.
//
// Corruption semantics follow the reference: the datum is corrupt unless decoding ends with
// dp == destend and sp == srcend (ppr:229-230).  Additionally reported as corrupt (the reference
// reads stale memory instead): tags that run past the source, offset 0, offsets before the
// start of the output.
#pragma once
#include "cds_common.cuh"

namespace cds {

constexpr uint32_t kRing = 8192;              // bytes of output history per warp (power of two)
constexpr uint32_t kFlushAt = 2048;           // flush to HBM once this many bytes are pending
constexpr int kPglzWarps = 4;                 // warps (datums) per CTA
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
        *reinterpret_cast<uint4 *>(dst + pos) = v;
    }
    const uint32_t t0 = b0 + (nvec << 4);
    if (lane < to - t0) dst[t0 + lane] = ring[(t0 + lane + shift) & (kRing - 1)];
}

// Decode one datum with the calling warp.  Returns true when the stream is well-formed.
__device__ bool pglz_warp_decode(const uint8_t *__restrict__ src, uint32_t slen, uint8_t *dst, uint32_t rawsize,
                                 uint8_t *ring)
{
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t shift = (uint32_t)(reinterpret_cast<uintptr_t>(dst) & 15u);
    uint32_t sp = 0, dp = 0, flushed = 0;
    bool corrupt = false;

    while (sp < slen && dp < rawsize && !corrupt) {
        // ---- streak of all-literal groups: candidate control bytes at sp + 9*lane -------------
        {
            const uint32_t cpos = sp + 9u * lane;
            const bool zero_full = (cpos + 9u <= slen) && (src[cpos] == 0);
            const uint32_t m = __ballot_sync(0xffffffffu, zero_full);
            uint32_t streak = (m == 0xffffffffu) ? 32u : (uint32_t)(__ffs(~m) - 1);
            const uint32_t room = (rawsize - dp) >> 3;                       // whole groups that still fit
            if (streak > room) streak = room;
            if (streak >= 2) {
                if (lane < streak) {
                    const uint8_t *s = src + cpos + 1;
                    const uint32_t o = dp + 8u * lane + shift;
#pragma unroll
                    for (int k = 0; k < 8; k++) ring[(o + k) & (kRing - 1)] = s[k];
                }
                sp += 9u * streak;
                dp += 8u * streak;
                __syncwarp();
                if (dp - flushed >= kFlushAt) {
                    const uint32_t to = dp - ((dp + shift) & 15u);
                    pglz_flush(ring, dst, shift, flushed, to, lane);
                    flushed = to;
                    __syncwarp();
                }
                continue;
            }
        }
        // ---- one control group -------------------------------------------------------------
        const uint32_t byte = (sp + lane < slen) ? src[sp + lane] : 0u;
        const uint32_t ctrl = __shfl_sync(0xffffffffu, byte, 0);
        const uint32_t extmask = __ballot_sync(0xffffffffu, (byte & 0x0Fu) == 0x0Fu);
        uint32_t off = 1;                                                    // source offset inside the group
        uint32_t nout = 0;                                                   // bytes produced by this group
        bool dirty = false;                                                  // ring written since last sync
        bool stop = false;                                                   // inner `break` of the reference
        for (uint32_t k = 0; k < 8 && !stop && !corrupt; ) {
            if (sp + off >= slen) break;                                     // `sp < srcend` of the inner loop
            if ((ctrl >> k) & 1u) {
                if (sp + off + 2u > slen) { corrupt = true; break; }
                const uint32_t b0 = __shfl_sync(0xffffffffu, byte, off);
                const uint32_t b1 = __shfl_sync(0xffffffffu, byte, off + 1u);
                uint32_t len = (b0 & 0x0Fu) + 3u;                            // ppr:205
                const uint32_t moff = ((b0 & 0xF0u) << 4) | b1;              // ppr:206-208
                const bool ext = (extmask >> off) & 1u;
                off += 2u;
                if (ext) {                                                   // len == 18 (ppr:210-212)
                    if (sp + off >= slen) { corrupt = true; break; }
                    len += __shfl_sync(0xffffffffu, byte, off);
                    off += 1u;
                }
                const uint32_t at = dp + nout;
                if (at + len > rawsize) { corrupt = true; break; }           // ppr:213-215 always ends corrupt
                if (moff == 0 || moff > at) { corrupt = true; break; }
                if (dirty) { __syncwarp(); dirty = false; }
                for (uint32_t i = lane; i < len; i += 32) {
                    const uint32_t back = (moff >= len) ? i : (i % moff);    // overlapping copy (ppr:216-219)
                    ring[(at + i + shift) & (kRing - 1)] = ring[(at - moff + back + shift) & (kRing - 1)];
                }
                dirty = true;
                nout += len;
                k++;
            } else {
                // run of literals: consecutive zero control bits, bounded by source and room
                uint32_t run = (uint32_t)__ffs((int)((ctrl >> k) | 0x100u >> k)) - 1u;   // zeros before next set bit
                if (run > 8u - k) run = 8u - k;
                if (run > slen - (sp + off)) run = slen - (sp + off);
                const uint32_t at = dp + nout;
                if (at >= rawsize) { stop = true; break; }                   // ppr:221-222 -> corrupt at the end
                const uint32_t roomb = rawsize - at;
                bool hit_end = false;
                if (run > roomb) { run = roomb; hit_end = true; }
                if (lane >= off && lane < off + run) ring[(at + (lane - off) + shift) & (kRing - 1)] = (uint8_t)byte;
                dirty = true;
                nout += run;
                off += run;
                k += run;
                if (hit_end) { stop = true; }                                // next literal would break with sp < srcend
            }
        }
        if (stop && sp + off < slen) corrupt = true;                         // literal refused: sp != srcend
        sp += off;
        dp += nout;
        __syncwarp();
        if (dp - flushed >= kFlushAt) {
            const uint32_t to = dp - ((dp + shift) & 15u);
            pglz_flush(ring, dst, shift, flushed, to, lane);
            flushed = to;
            __syncwarp();
        }
    }
    if (!corrupt) {
        __syncwarp();
        pglz_flush(ring, dst, shift, flushed, dp, lane);
    }
    return !corrupt && dp == rawsize && sp == slen;                          // ppr:229-230
}

__global__ void __launch_bounds__(32 * kPglzWarps)
k_pglz_decode(const PglzJob *__restrict__ jobs, const uint32_t *n_jobs_dev, uint32_t n_jobs_mul, uint32_t n_jobs_host,
              int32_t *status, DevState *st)
{
    extern __shared__ __align__(16) uint8_t smem_ring[];
    uint8_t *ring = smem_ring + (size_t)(threadIdx.x >> 5) * kRing;
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t n_jobs = n_jobs_dev ? (*n_jobs_dev) * n_jobs_mul : n_jobs_host;
    const uint32_t wid = blockIdx.x * kPglzWarps + (threadIdx.x >> 5), nw = gridDim.x * kPglzWarps;
    for (uint32_t j = wid; j < n_jobs; j += nw) {
        const PglzJob job = jobs[j];
        if (job.slen == kNone) continue;
        const bool ok = pglz_warp_decode(reinterpret_cast<const uint8_t *>(job.src), job.slen,
                                         reinterpret_cast<uint8_t *>(job.dst), job.rawsize, ring);
        if (lane == 0) {
            if (status) status[j] = ok ? CDS_OK : CDS_ERR_PGLZ_CORRUPT;
            if (!ok && st && job.err_key) report_key(st, job.err_key);
        }
        __syncwarp();
    }
}

}  // namespace cds
