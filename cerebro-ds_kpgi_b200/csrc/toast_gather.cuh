// toast_gather.cuh — the bulk kernel: TOAST chunk payloads -> contiguous values.
//
// Replaces the chunk reassembly of detoast (/root/reference/cerebro_gpdb/pg_page_reader.py:577-594:
// sort by chunk_seq, strip the 4-byte chunk header, b''.join) fused with the typed
// reinterpretation (pg_page_reader.py:165-182, which is a byte copy), writing straight into the
// caller's output tensor (uncompressed values) or into the pglz staging area (compressed values).
//
// Shape of the work: every 32 KB page is read exactly once and every payload byte written
// exactly once -> HBM-bound.  Payloads start at page offset = 4 (mod 8) and are 8140 = 12 (mod 16)
// bytes long, so source and destination are never mutually 16-byte aligned.  Design:
//   * persistent CTAs, one per SM; a producer warp streams whole pages into a shared-memory ring
//     with TMA bulk copies (cp.async.bulk ... mbarrier::complete_tx), kStages pages in flight;
//   * the producer warp also prefetches the page's copy descriptors into shared memory;
//   * consumer warps realign in shared memory (two aligned LDS.128 + word select) and emit
//     aligned, coalesced, streaming 16-byte global stores.
#pragma once
#include <cuda_bf16.h>

#include "cds_common.cuh"
#include "tma_helpers.cuh"

namespace cds {
namespace {   // internal linkage: with relocatable device code (-rdc) the optimiser then treats the file as a whole program

constexpr int kGatherStages = 6;
constexpr int kGatherConsumerWarps = 16;
constexpr int kGatherThreads = 32 * (1 + kGatherConsumerWarps);
constexpr int kGatherMaxStageOps = 48;
constexpr uint32_t kMaxLutChannels = 4;
constexpr size_t kGatherSmemBytes =
    (size_t)kGatherStages * kPage + 16 /* realign over-read pad */ +
    (size_t)kGatherStages * kGatherMaxStageOps * sizeof(CopyOp) +
    (size_t)kGatherStages * 2 * sizeof(uint32_t) + (size_t)kGatherStages * 2 * sizeof(unsigned long long) + 64 +
    kMaxLutChannels * 256 * sizeof(float);

__device__ __forceinline__ void st_stream_v4(void *gdst, uint4 v)
{
    asm volatile("st.global.cs.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(gdst), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// Copy `len` bytes from shared `src` to global `dst` with the consumer threads (tid of nth).
__device__ __forceinline__ void copy_payload(const uint8_t *src, uint8_t *dst, uint32_t len, uint32_t tid, uint32_t nth)
{
    const uint32_t sa = smem_u32(src);
    const uintptr_t da = reinterpret_cast<uintptr_t>(dst);
    if (((sa | (uint32_t)da) & 3u) != 0) {
        for (uint32_t i = tid; i < len; i += nth) dst[i] = src[i];          // odd alignment: bytes
        return;
    }
    uint32_t head = (16u - (uint32_t)(da & 15u)) & 15u;                      // bytes up to a 16-byte line
    if (head > (len & ~3u)) head = len & ~3u;
    const uint32_t body = (len - head) >> 4;                                 // whole 16-byte stores
    const uint32_t done = head + (body << 4);
    if (tid < (head >> 2))
        *reinterpret_cast<uint32_t *>(dst + 4u * tid) = *reinterpret_cast<const uint32_t *>(src + 4u * tid);
    const uint8_t *sb = src + head;
    const uint32_t r = smem_u32(sb) & 15u;                                   // 0, 4, 8 or 12
    const uint4 *s16 = reinterpret_cast<const uint4 *>(sb - r);
    uint4 *d16 = reinterpret_cast<uint4 *>(dst + head);
    if (r == 0) {
        for (uint32_t j = tid; j < body; j += nth) st_stream_v4(d16 + j, s16[j]);
    } else {
        for (uint32_t j = tid; j < body; j += nth) {
            const uint4 a = s16[j], b = s16[j + 1];
            uint4 o;
            if (r == 4)       o = make_uint4(a.y, a.z, a.w, b.x);
            else if (r == 8)  o = make_uint4(a.z, a.w, b.x, b.y);
            else              o = make_uint4(a.w, b.x, b.y, b.z);
            st_stream_v4(d16 + j, o);
        }
    }
    const uint32_t tail_words = (len - done) >> 2;
    if (tid < tail_words)
        *reinterpret_cast<uint32_t *>(dst + done + 4u * tid) = *reinterpret_cast<const uint32_t *>(src + done + 4u * tid);
    const uint32_t rem = done + (tail_words << 2);
    if (tid < len - rem) dst[rem + tid] = src[rem + tid];
}

__device__ __forceinline__ uint32_t pack_bf16x2(float a, float b)
{
    const __nv_bfloat162 h = __floats2bfloat162_rn(a, b);                    // round to nearest even
    return *reinterpret_cast<const uint32_t *>(&h);
}

// Same job as copy_payload, but the payload is float32 data and the destination holds bfloat16: `len`
// source bytes (a multiple of 4, 4-byte aligned in shared memory) become len/2 bytes at `dst`.
__device__ __forceinline__ void copy_payload_bf16(const uint8_t *src, uint8_t *dst, uint32_t len, uint32_t tid, uint32_t nth)
{
    const uint32_t n = len >> 2;                                             // floats
    const uintptr_t da = reinterpret_cast<uintptr_t>(dst);
    uint32_t head = ((8u - (uint32_t)(da & 7u)) & 7u) >> 1;                   // floats until dst is 8-byte aligned
    if (head > n) head = n;
    const float *sf = reinterpret_cast<const float *>(src);
    __nv_bfloat16 *dh = reinterpret_cast<__nv_bfloat16 *>(dst);
    if (tid < head) dh[tid] = __float2bfloat16_rn(sf[tid]);
    const uint32_t body = (n - head) >> 2;                                   // groups of 4 floats -> one 8-byte store
    const uint8_t *sb = src + 4u * head;
    const uint32_t r = smem_u32(sb) & 15u;                                   // 0, 4, 8 or 12
    const uint4 *s16 = reinterpret_cast<const uint4 *>(sb - r);
    uint2 *d8 = reinterpret_cast<uint2 *>(dst + 2u * head);
    for (uint32_t j = tid; j < body; j += nth) {
        const uint4 a = s16[j];
        uint4 o = a;
        if (r != 0) {
            const uint4 b = s16[j + 1];
            if (r == 4)       o = make_uint4(a.y, a.z, a.w, b.x);
            else if (r == 8)  o = make_uint4(a.z, a.w, b.x, b.y);
            else              o = make_uint4(a.w, b.x, b.y, b.z);
        }
        uint2 q;
        q.x = pack_bf16x2(__uint_as_float(o.x), __uint_as_float(o.y));
        q.y = pack_bf16x2(__uint_as_float(o.z), __uint_as_float(o.w));
        asm volatile("st.global.cs.v2.u32 [%0], {%1, %2};" ::"l"(d8 + j), "r"(q.x), "r"(q.y) : "memory");
    }
    const uint32_t done = head + (body << 2);
    if (tid < n - done) dh[done + tid] = __float2bfloat16_rn(sf[done + tid]);
}

// uint8 pixels -> normalised float32 through a [channels][256] table in shared memory (the reference's
// (x/255 - mean)/std, evaluated once per table entry in double on the host: SURVEY.md §8c(3)).  `phase` is the
// channel of the first payload byte.  len source bytes become 4*len bytes at dst (16-byte aligned here).
__device__ __forceinline__ void copy_payload_u8norm(const uint8_t *src, float *dst, uint32_t len, uint32_t phase, uint32_t ch,
                                                    const float *lut, uint32_t tid, uint32_t nth)
{
    const uint32_t quads = len >> 2;
    const bool aligned = ((smem_u32(src) & 3u) == 0) && ((reinterpret_cast<uintptr_t>(dst) & 15u) == 0);
    for (uint32_t j = tid; j < quads; j += nth) {
        uint32_t w;
        if (aligned) w = *reinterpret_cast<const uint32_t *>(src + 4u * j);
        else w = (uint32_t)src[4u * j] | ((uint32_t)src[4u * j + 1] << 8) | ((uint32_t)src[4u * j + 2] << 16) | ((uint32_t)src[4u * j + 3] << 24);
        uint32_t c = (phase + 4u * j) % ch;
        float4 o;
        o.x = lut[c * 256u + (w & 0xFFu)];          c = (c + 1u == ch) ? 0u : c + 1u;
        o.y = lut[c * 256u + ((w >> 8) & 0xFFu)];   c = (c + 1u == ch) ? 0u : c + 1u;
        o.z = lut[c * 256u + ((w >> 16) & 0xFFu)];  c = (c + 1u == ch) ? 0u : c + 1u;
        o.w = lut[c * 256u + (w >> 24)];
        if (aligned) st_stream_v4(dst + 4u * j, make_uint4(__float_as_uint(o.x), __float_as_uint(o.y), __float_as_uint(o.z), __float_as_uint(o.w)));
        else { dst[4u * j] = o.x; dst[4u * j + 1] = o.y; dst[4u * j + 2] = o.z; dst[4u * j + 3] = o.w; }
    }
    const uint32_t done = quads << 2;
    if (tid < len - done) dst[done + tid] = lut[((phase + done + tid) % ch) * 256u + src[done + tid]];
}

// Part of a literal-only pglz stream (see kOpLiteral): `n` stream bytes at shared `src`, the first one at stream
// offset = phase (mod 9).  Stream byte s is a control byte when s % 9 == 0 (must be zero: a set bit would announce a
// tag, which a stream of this length cannot hold, ppr:203-230) and literal number (s / 9) * 8 + s % 9 - 1 otherwise.
// Writes the literals to `dst` (the address of the first literal in this part); returns false when a control byte is
// not zero.  `fin`: 1..7 = only that many low bits of the LAST control byte of this part count.
__device__ __forceinline__ bool copy_payload_literal(const uint8_t *src, uint8_t *dst, uint32_t n, uint32_t phase, uint32_t fin,
                                                     uint32_t tid, uint32_t nth)
{
    bool good = true;
    const uint32_t c0 = phase ? 9u - phase : 0u;                             // first control byte inside the part
    const uint32_t n_ctrl = n > c0 ? (n - c0 + 8u) / 9u : 0u;
    for (uint32_t i = tid; i < n_ctrl; i += nth) {
        uint32_t b = src[c0 + 9u * i];
        if (fin && i + 1u == n_ctrl) b &= (1u << fin) - 1u;
        good = good && b == 0u;
    }
    const uint32_t n_out = n - n_ctrl;
    const uint32_t q = phase ? phase - 1u : 0u;                              // literal index of the first output byte inside its group
    // output byte j of this part = literal O = q + j counted from the group's first literal -> src[(O / 8) * 9 + O % 8 + 1 - phase]
    const uintptr_t da = reinterpret_cast<uintptr_t>(dst);
    if (((da - q) & 7u) != 0u) {                                             // groups do not line up with 8-byte words of the output
        for (uint32_t j = tid; j < n_out; j += nth) {
            const uint32_t O = q + j;
            dst[j] = src[(O >> 3) * 9u + (O & 7u) + 1u - phase];
        }
        return good;
    }
    uint32_t head = (16u - (uint32_t)(da & 15u)) & 15u;
    if (head > n_out) head = n_out;
    if (tid < head) {
        const uint32_t O = q + tid;
        dst[tid] = src[(O >> 3) * 9u + (O & 7u) + 1u - phase];
    }
    const uint32_t nvec = (n_out - head) >> 4, m0 = (q + head) >> 3;         // q + head is a multiple of 8 here
    uint4 *d16 = reinterpret_cast<uint4 *>(dst + head);
    for (uint32_t v = tid; v < nvec; v += nth) {
        // two groups: literals at src[a .. a+8) and src[a+9 .. a+17)
        const uint32_t a = smem_u32(src) + (m0 + 2u * v) * 9u + 1u - phase, b = a + 9u;
        const uint32_t a4 = a & ~3u, b4 = b & ~3u, sa = (a & 3u) * 8u, sb = (b & 3u) * 8u;
        uint32_t x0, x1, x2, y0, y1, y2;
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(x0) : "r"(a4));
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(x1) : "r"(a4 + 4u));
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(x2) : "r"(a4 + 8u));
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(y0) : "r"(b4));
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(y1) : "r"(b4 + 4u));
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(y2) : "r"(b4 + 8u));
        st_stream_v4(d16 + v, make_uint4(__funnelshift_r(x0, x1, sa), __funnelshift_r(x1, x2, sa),
                                         __funnelshift_r(y0, y1, sb), __funnelshift_r(y1, y2, sb)));
    }
    const uint32_t done = head + (nvec << 4);
    if (tid < n_out - done) {
        const uint32_t O = q + done + tid;
        dst[done + tid] = src[(O >> 3) * 9u + (O & 7u) + 1u - phase];
    }
    return good;
}

__global__ void __launch_bounds__(kGatherThreads, 1)
k_toast_gather(const uint8_t *__restrict__ pages, uint32_t n_pages, const uint32_t *__restrict__ base,
               const CopyOp *__restrict__ ops, DevState *st, const float *__restrict__ lut, uint32_t lut_channels,
               const uint32_t *__restrict__ dir_ref, const uint32_t *__restrict__ vfirst)
{
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t *stage_buf = smem;
    CopyOp *stage_ops = reinterpret_cast<CopyOp *>(smem + (size_t)kGatherStages * kPage + 16);
    uint32_t *stage_g0 = reinterpret_cast<uint32_t *>(stage_ops + kGatherStages * kGatherMaxStageOps);
    uint32_t *stage_n = stage_g0 + kGatherStages;
    unsigned long long *full = reinterpret_cast<unsigned long long *>(stage_n + kGatherStages);
    unsigned long long *empty = full + kGatherStages;
    float *lut_s = reinterpret_cast<float *>(empty + kGatherStages);        // [channels][256], only in the u8 -> f32 mode
    if (lut) for (uint32_t i = threadIdx.x; i < 256u * lut_channels; i += blockDim.x) lut_s[i] = lut[i];

    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int s = 0; s < kGatherStages; s++) {
            mbar_init(&full[s], 2);                       // expect_tx arrival + descriptor arrival
            mbar_init(&empty[s], kGatherConsumerWarps);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (blockIdx.x >= n_pages || (st->flags & kFlagChunkOverflow)) return;   // no directory: nothing to copy
    const uint32_t my_pages = (n_pages - blockIdx.x + gridDim.x - 1) / gridDim.x;

    if (warp == 0) {
        // ---- producer: page p of this CTA's strided sequence goes to ring slot it % kStages
        for (uint32_t it = 0; it < my_pages; it++) {
            const uint32_t s = it % kGatherStages, round = it / kGatherStages;
            const uint32_t p = blockIdx.x + it * gridDim.x;
            if (round > 0) mbar_wait(&empty[s], (round - 1) & 1u);
            if (lane == 0) {
                mbar_arrive_expect_tx(&full[s], kPage);
                tma_load_bulk(stage_buf + (size_t)s * kPage, pages + (size_t)p * kPage, kPage, &full[s]);
            }
            const uint32_t g0 = base[p], n = base[p + 1] - g0;
            for (uint32_t j = lane; j < n && j < (uint32_t)kGatherMaxStageOps; j += 32)
                stage_ops[s * kGatherMaxStageOps + j] = ops[g0 + j];
            if (lane == 0) { stage_g0[s] = g0; stage_n[s] = n; }
            __syncwarp();
            if (lane == 0) mbar_arrive(&full[s]);
        }
    } else {
        // ---- consumers
        const uint32_t ctid = threadIdx.x - 32, cnth = 32 * kGatherConsumerWarps;
        for (uint32_t it = 0; it < my_pages; it++) {
            const uint32_t s = it % kGatherStages, round = it / kGatherStages;
            mbar_wait(&full[s], round & 1u);
            const uint32_t n = stage_n[s], g0 = stage_g0[s];
            const uint8_t *pg = stage_buf + (size_t)s * kPage;
            for (uint32_t j = 0; j < n; j++) {
                const CopyOp op = j < (uint32_t)kGatherMaxStageOps ? stage_ops[s * kGatherMaxStageOps + j] : ops[g0 + j];
                if (op.dst == 0 || op.len == 0) continue;
                if (op.len & kOpLiteral) {
                    if (!copy_payload_literal(pg + (op.src_off & 0xFFFFu), reinterpret_cast<uint8_t *>(op.dst), op.len & 0xFFFFu,
                                              (op.src_off >> 16) & 15u, (op.src_off >> 20) & 7u, ctid, cnth)) {
                        // "compressed data is corrupt" for this value (ppr:229-230), the key the pglz decoder would publish
                        const uint32_t v = dir_ref[g0 + j] & 0x3FFFFFFFu;
                        report(st, 3, vfirst[v], 3, 0, CDS_ERR_PGLZ_CORRUPT);
                    }
                }
                else if (op.len & kOpToBf16) copy_payload_bf16(pg + op.src_off, reinterpret_cast<uint8_t *>(op.dst), op.len & ~kFlagBits, ctid, cnth);
                else if (op.len & kOpU8Norm) copy_payload_u8norm(pg + (op.src_off & 0xFFFFu), reinterpret_cast<float *>(op.dst), op.len & ~kFlagBits,
                                                                op.src_off >> 16, lut_channels, lut_s, ctid, cnth);
                else copy_payload(pg + op.src_off, reinterpret_cast<uint8_t *>(op.dst), op.len, ctid, cnth);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[s]);
        }
    }
}

}  // namespace
}  // namespace cds
