// util_kernels.cuh — two small bandwidth kernels outside the decode itself:
//   * k_checksum64: an order-sensitive 64-bit checksum of a device buffer (bench.py's correctness gate compares it
//     with the same sum computed by the segment generator over the arrays it packed);
//   * k_peer_copy: the model-hopper's state transfer, a plain vectorised copy whose destination
//     may be a peer GPU's buffer mapped into this process (NVLink stores; no TMA, nothing multicast).
#pragma once
#include "cds_common.cuh"

namespace cds {
namespace {   // internal linkage: with relocatable device code (-rdc) the optimiser then treats the file as a whole program

constexpr unsigned long long kSumMul = 0x9E3779B97F4A7C15ull;

// sum over 64-bit little-endian words w_i (the tail zero-padded) of w_i * (2 i + 1) * kSumMul, modulo 2^64.
__global__ void __launch_bounds__(256)
k_checksum64(const uint8_t *__restrict__ buf, unsigned long long n_bytes, unsigned long long *out)
{
    const unsigned long long n_words = n_bytes >> 3;
    unsigned long long acc = 0;
    const bool aligned = (reinterpret_cast<uintptr_t>(buf) & 7u) == 0;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_words; i += (unsigned long long)gridDim.x * blockDim.x) {
        unsigned long long w;
        if (aligned) w = __ldg(reinterpret_cast<const unsigned long long *>(buf) + i);
        else {
            w = 0;
            for (int k = 0; k < 8; k++) w |= (unsigned long long)buf[8 * i + k] << (8 * k);
        }
        acc += w * ((2ull * i + 1ull) * kSumMul);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0 && (n_bytes & 7ull)) {
        unsigned long long w = 0;
        for (unsigned k = 0; k < (unsigned)(n_bytes & 7ull); k++) w |= (unsigned long long)buf[8 * n_words + k] << (8 * k);
        acc += w * ((2ull * n_words + 1ull) * kSumMul);
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, d);
    __shared__ unsigned long long sh[8];
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long t = 0;
        for (int w = 0; w < 8; w++) t += sh[w];
        atomicAdd(out, t);
    }
}

// dst[0, n) = src[0, n).  Both 16-byte aligned in the body; head/tail bytes are copied one by one.
__global__ void __launch_bounds__(512)
k_peer_copy(const uint8_t *__restrict__ src, uint8_t *__restrict__ dst, unsigned long long n_bytes)
{
    const unsigned long long tid = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x, nth = (unsigned long long)gridDim.x * blockDim.x;
    if (((reinterpret_cast<uintptr_t>(src) ^ reinterpret_cast<uintptr_t>(dst)) & 15u) != 0) {
        for (unsigned long long i = tid; i < n_bytes; i += nth) dst[i] = src[i];
        return;
    }
    unsigned long long head = (16u - (reinterpret_cast<uintptr_t>(src) & 15u)) & 15u;
    if (head > n_bytes) head = n_bytes;
    if (tid < head) dst[tid] = src[tid];
    const unsigned long long nvec = (n_bytes - head) >> 4;
    const uint4 *s = reinterpret_cast<const uint4 *>(src + head);
    uint4 *d = reinterpret_cast<uint4 *>(dst + head);
    unsigned long long i = tid;
    for (; i + 3 * nth < nvec; i += 4 * nth) {                               // four independent 16-byte loads in flight per thread
        const uint4 a = __ldg(s + i), b = __ldg(s + i + nth), c = __ldg(s + i + 2 * nth), e = __ldg(s + i + 3 * nth);
        d[i] = a; d[i + nth] = b; d[i + 2 * nth] = c; d[i + 3 * nth] = e;
    }
    for (; i < nvec; i += nth) d[i] = __ldg(s + i);
    const unsigned long long done = head + (nvec << 4);
    if (tid < n_bytes - done) dst[done + tid] = src[done + tid];
}

}  // namespace
}  // namespace cds
