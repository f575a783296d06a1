// cast_pack.cuh — consumer-side casts of the decoded partition.
//
//   one-hot int16 row -> int64 class index   np.argmax(dependent_var[idx])
//                                             /root/reference/cerebro_gpdb/run_pytorchddp_da.py:55
//   NHWC f32 -> NCHW f32                      np.rollaxis(image, 2, 0) + DataLoader collate
//                                             /root/reference/cerebro_gpdb/run_pytorchddp_da.py:56-58
//   f32 -> bf16 (RNE)                         optional training-buffer dtype (not in the reference)
#pragma once
#include <cuda_bf16.h>

#include "cds_common.cuh"
#include "tma_helpers.cuh"

namespace cds {
namespace {   // internal linkage: with relocatable device code (-rdc) the optimiser then treats the file as a whole program

// One warp per row; first maximum wins, as np.argmax.  Rows whose bytes are 16-byte aligned are read
// eight classes per lane and load.
__global__ void __launch_bounds__(256)
k_onehot_argmax_i16(const int16_t *__restrict__ onehot, unsigned long long n_rows, uint32_t classes,
                    long long *__restrict__ labels)
{
    const uint32_t lane = threadIdx.x & 31;
    const unsigned long long w0 = (unsigned long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const unsigned long long nw = (unsigned long long)gridDim.x * (blockDim.x >> 5);
    const bool vec = ((classes & 7u) == 0u) && ((reinterpret_cast<uintptr_t>(onehot) & 15u) == 0u);
    for (unsigned long long r = w0; r < n_rows; r += nw) {
        const int16_t *row = onehot + r * classes;
        int best = -32769;
        uint32_t best_i = 0xFFFFFFFFu;
        if (vec) {
            const uint4 *row4 = reinterpret_cast<const uint4 *>(row);
            for (uint32_t c8 = lane; c8 < (classes >> 3); c8 += 32) {
                const uint4 q = __ldcs(row4 + c8);
                const uint32_t wds[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
                for (int k = 0; k < 8; k++) {
                    const int v = (int)(int16_t)(wds[k >> 1] >> ((k & 1) * 16));
                    if (v > best) { best = v; best_i = c8 * 8u + k; }
                }
            }
        } else {
            for (uint32_t c = lane; c < classes; c += 32) {
                const int v = row[c];
                if (v > best) { best = v; best_i = c; }
            }
        }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            const int ov = __shfl_xor_sync(0xffffffffu, best, d);
            const uint32_t oi = __shfl_xor_sync(0xffffffffu, best_i, d);
            if (ov > best || (ov == best && oi < best_i)) { best = ov; best_i = oi; }
        }
        if (lane == 0) labels[r] = (long long)best_i;
    }
}

// dst[n][c][p] = src[n][p][c].  Thread per destination element: coalesced stores; the strided
// loads of one warp stay inside 32*channels*4 contiguous bytes and are served by L1.
__global__ void __launch_bounds__(256)
k_nhwc_to_nchw_f32(const float *__restrict__ src, unsigned long long n, uint32_t hw, uint32_t ch,
                   float *__restrict__ dst)
{
    const unsigned long long per = (unsigned long long)hw * ch, total = n * per;
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const unsigned long long img = i / per;
        const uint32_t rem = (uint32_t)(i - img * per);
        const uint32_t c = rem / hw, p = rem - c * hw;
        dst[i] = src[img * per + (unsigned long long)p * ch + c];
    }
}

__global__ void __launch_bounds__(256)
k_f32_to_bf16(const float *__restrict__ src, unsigned long long n, __nv_bfloat16 *__restrict__ dst)
{
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    const unsigned long long tid = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned long long n4 = n >> 2;
    const bool aligned = ((reinterpret_cast<uintptr_t>(src) & 15u) == 0) && ((reinterpret_cast<uintptr_t>(dst) & 7u) == 0);
    if (aligned) {
        for (unsigned long long i = tid; i < n4; i += stride) {
            const float4 v = reinterpret_cast<const float4 *>(src)[i];
            __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
            uint2 o;
            o.x = *reinterpret_cast<uint32_t *>(&lo);
            o.y = *reinterpret_cast<uint32_t *>(&hi);
            reinterpret_cast<uint2 *>(dst)[i] = o;
        }
        for (unsigned long long i = (n4 << 2) + tid; i < n; i += stride) dst[i] = __float2bfloat16_rn(src[i]);
    } else {
        for (unsigned long long i = tid; i < n; i += stride) dst[i] = __float2bfloat16_rn(src[i]);
    }
}

// Batch assembly for the torch consumer: DADataset.__getitem__ + DataLoader collate
// (/root/reference/cerebro_gpdb/run_pytorchddp_da.py:51-62, run_pytorchddp.py:397-425) as one kernel:
//   i = indices[b] % n_samples      (train mode pads the epoch by wrapping, run_pytorchddp_da.py:52-53)
//   out[b][c][p] = images[i][p][c]  (np.rollaxis(image, 2, 0)), optionally rounded to bf16
//   out_labels[b] = labels[i]       (np.argmax(dependent_var[idx]), precomputed once per partition)
// channels == 1 degenerates to a row gather (Criteo features).
template <typename TOut>
__global__ void __launch_bounds__(256)
k_batch_assemble(const float *__restrict__ images, unsigned long long n_samples, uint32_t hw, uint32_t ch,
                 const long long *__restrict__ labels, const long long *__restrict__ indices, uint32_t batch,
                 TOut *__restrict__ out, long long *__restrict__ out_labels)
{
    const unsigned long long per = (unsigned long long)hw * ch, total = (unsigned long long)batch * per;
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    const unsigned long long tid = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    for (unsigned long long i = tid; i < total; i += stride) {
        const uint32_t b = (uint32_t)(i / per);
        const uint32_t rem = (uint32_t)(i - (unsigned long long)b * per);
        const uint32_t c = rem / hw, p = rem - c * hw;
        long long s = indices[b] % (long long)n_samples;
        if (s < 0) s += (long long)n_samples;
        const float v = images[(unsigned long long)s * per + (unsigned long long)p * ch + c];
        if constexpr (sizeof(TOut) == 2) out[i] = __float2bfloat16_rn(v); else out[i] = v;
    }
    if (labels && out_labels) {
        for (unsigned long long b = tid; b < batch; b += stride) {
            long long s = indices[b] % (long long)n_samples;
            if (s < 0) s += (long long)n_samples;
            out_labels[b] = labels[s];
        }
    }
}

// Staged version of the two NHWC -> NCHW kernels above for images whose samples are 16-byte multiples:
// persistent CTAs, a producer warp streams tiles of `tp` whole pixels (tp * ch * 4 bytes, a multiple of 16)
// into a shared-memory ring with TMA bulk copies, consumer warps read one channel of consecutive pixels
// (stride `ch` words: conflict-free for odd ch) and write coalesced channel planes.  Every source byte is
// read once and every destination byte written once.
//   indices == NULL: sample b of the source goes to slot b (whole-partition transpose);
//   otherwise sample = indices[b] mod n_samples (batch assembly).
constexpr int kAsmStages = 4;
constexpr int kAsmConsumerWarps = 12;
constexpr int kAsmThreads = 32 * (1 + kAsmConsumerWarps);
constexpr uint32_t kAsmMaxTileBytes = 48 * 1024;

template <typename TOut>
__global__ void __launch_bounds__(kAsmThreads, 1)
k_assemble_tma(const float *__restrict__ images, unsigned long long n_samples, uint32_t hw, uint32_t ch, uint32_t tp,
               const long long *__restrict__ labels, const long long *__restrict__ indices, uint32_t batch,
               TOut *__restrict__ out, long long *__restrict__ out_labels)
{
    extern __shared__ __align__(128) uint8_t smem_raw[];
    const uint32_t tile_bytes = tp * ch * 4u;
    unsigned long long *full = reinterpret_cast<unsigned long long *>(smem_raw + (size_t)kAsmStages * tile_bytes);
    unsigned long long *empty = full + kAsmStages;
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int s = 0; s < kAsmStages; s++) { mbar_init(&full[s], 1); mbar_init(&empty[s], kAsmConsumerWarps); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const uint32_t tiles_per = hw / tp;
    const unsigned long long total = (unsigned long long)batch * tiles_per;
    if (blockIdx.x >= total) return;
    const uint32_t mine = (uint32_t)((total - blockIdx.x + gridDim.x - 1) / gridDim.x);
    const unsigned long long per = (unsigned long long)hw * ch;

    auto sample_of = [&](uint32_t b) -> unsigned long long {
        if (!indices) return b;
        long long s = indices[b] % (long long)n_samples;
        return (unsigned long long)(s < 0 ? s + (long long)n_samples : s);
    };

    if (warp == 0) {
        for (uint32_t it = 0; it < mine; it++) {
            const uint32_t s = it % kAsmStages, round = it / kAsmStages;
            const unsigned long long w = blockIdx.x + (unsigned long long)it * gridDim.x;
            const uint32_t b = (uint32_t)(w / tiles_per), t = (uint32_t)(w - (unsigned long long)b * tiles_per);
            if (round > 0) mbar_wait(&empty[s], (round - 1) & 1u);
            if (lane == 0) {
                const float *src = images + sample_of(b) * per + (unsigned long long)t * tp * ch;
                mbar_arrive_expect_tx(&full[s], tile_bytes);
                tma_load_bulk(smem_raw + (size_t)s * tile_bytes, src, tile_bytes, &full[s]);
                if (t == 0 && labels && out_labels) out_labels[b] = labels[sample_of(b)];
            }
        }
    } else {
        const uint32_t ctid = threadIdx.x - 32, cnth = 32 * kAsmConsumerWarps;
        for (uint32_t it = 0; it < mine; it++) {
            const uint32_t s = it % kAsmStages, round = it / kAsmStages;
            const unsigned long long w = blockIdx.x + (unsigned long long)it * gridDim.x;
            const uint32_t b = (uint32_t)(w / tiles_per), t = (uint32_t)(w - (unsigned long long)b * tiles_per);
            mbar_wait(&full[s], round & 1u);
            const float *tile = reinterpret_cast<const float *>(smem_raw + (size_t)s * tile_bytes);
            TOut *dst = out + (unsigned long long)b * per + (unsigned long long)t * tp;
            if constexpr (sizeof(TOut) == 4) {
                for (uint32_t c = 0; c < ch; c++)
                    for (uint32_t p = ctid; p < tp; p += cnth)
                        __stcs(reinterpret_cast<float *>(dst) + (unsigned long long)c * hw + p, tile[p * ch + c]);
            } else {
                // two pixels per lane -> one 32-bit store of two bfloat16 (tp and hw are even here)
                for (uint32_t c = 0; c < ch; c++)
                    for (uint32_t p2 = ctid; p2 < (tp >> 1); p2 += cnth) {
                        const __nv_bfloat162 h = __floats2bfloat162_rn(tile[(2u * p2) * ch + c], tile[(2u * p2 + 1u) * ch + c]);
                        __stcs(reinterpret_cast<unsigned int *>(dst + (unsigned long long)c * hw) + p2, *reinterpret_cast<const unsigned int *>(&h));
                    }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[s]);
        }
    }
}

}  // namespace
}  // namespace cds
