// orbx_device.cuh -- small device helpers shared by the sm_100a kernels.
#ifndef ORBX_DEVICE_CUH
#define ORBX_DEVICE_CUH

#include <cuda_runtime.h>
#include <stdint.h>

#include "orbx_internal.h"

#define ORBX_FULL 0xffffffffu

__device__ __forceinline__ int orbx_lane() { return threadIdx.x & 31; }
__device__ __forceinline__ int orbx_warp() { return threadIdx.x >> 5; }
__device__ __forceinline__ uint32_t orbx_lanemask_lt() {
    uint32_t m;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
    return m;
}

// Base pointer and pitch of level l of one frame (level 0 lives in the caller's image buffer).
__device__ __forceinline__ const uint8_t* orbx_level_ptr(const OrbxPlan& P, const OrbxSrc& src,
                                                         const uint8_t* pyr, int frame, int l,
                                                         int* pitch) {
    if (l == 0) {
        *pitch = src.pitch0;
        return src.img0 + (size_t)frame * src.stride0;
    }
    *pitch = P.lvl[l].pitch;
    return pyr + (size_t)frame * P.frame_stride + P.lvl[l].off;
}

__device__ __forceinline__ uint32_t warp_incl_scan(uint32_t v) {
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t n = __shfl_up_sync(ORBX_FULL, v, d);
        if (orbx_lane() >= d) v += n;
    }
    return v;
}

__device__ __forceinline__ int warp_sum(int v) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(ORBX_FULL, v, d);
    return v;
}

// In-place exclusive scan of a[0..n) in shared memory by the whole block; returns the total.
// tmp: >= 34 words of shared memory.  Every thread of the block must call it.
__device__ __forceinline__ uint32_t block_excl_scan(uint32_t* a, int n, uint32_t* tmp) {
    const int T = blockDim.x, t = threadIdx.x;
    const int per = (n + T - 1) / T;
    const int b = min(t * per, n), e = min(b + per, n);
    uint32_t s = 0;
    for (int i = b; i < e; ++i) s += a[i];
    uint32_t incl = warp_incl_scan(s);
    if (orbx_lane() == 31) tmp[orbx_warp()] = incl;
    __syncthreads();
    if (orbx_warp() == 0) {
        const int nw = (T + 31) >> 5;
        uint32_t v = orbx_lane() < nw ? tmp[orbx_lane()] : 0;
        uint32_t iv = warp_incl_scan(v);
        tmp[orbx_lane()] = iv - v;
        if (orbx_lane() == 31) tmp[33] = iv;
    }
    __syncthreads();
    uint32_t base = incl - s + tmp[orbx_warp()];
    for (int i = b; i < e; ++i) {
        uint32_t v = a[i];
        a[i] = base;
        base += v;
    }
    uint32_t total = tmp[33];
    __syncthreads();
    return total;
}

#endif  // ORBX_DEVICE_CUH
