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

// E7 (ORBextractor.cc:1124-1144): prefix over the levels' survivors of ONE frame, lapping-area test
// on the scaled x, stable TWO-ENDED compaction: a keypoint with lap0 <= pt.x <= lap1 goes to the
// back (stereoIndex--), others to the front (monoIndex++), both in level-major order.  Writes one
// record {key, output row | level << 24} per survivor for k_describe and the frame header
// {n, monoIndex, status}.  Called by every thread of ONE CTA (the last quadtree CTA of the frame to
// finish); surv / survcnt were written by other CTAs, hence the L2 loads.  sm: >= surv_total + 66 words.
__device__ __forceinline__ void orbx_assemble_frame(const OrbxPlan& P, const uint32_t* surv,
                                                    const uint32_t* survcnt, const int* __restrict__ lap,
                                                    uint32_t* __restrict__ dest, int32_t* __restrict__ header,
                                                    int frame, uint32_t* sm) {
    uint32_t* flags = sm;                    // [surv_total + 2]
    uint32_t* tmp = sm + P.surv_total + 2;   // [64]
    const int tid = threadIdx.x;
    const uint32_t* sc = survcnt + (size_t)frame * ORBX_MAX_LEVELS;
    const uint32_t* sv = surv + (size_t)frame * P.surv_total;
    const float lap0 = (float)lap[2 * frame], lap1 = (float)lap[2 * frame + 1];
    int off[ORBX_MAX_LEVELS + 1];
    off[0] = 0;
#pragma unroll
    for (int l = 0; l < ORBX_MAX_LEVELS; ++l)
        off[l + 1] = off[l] + (l < P.nlevels ? min((int)__ldcg(sc + l), P.lvl[l].surv_cap) : 0);
    const int n = off[ORBX_MAX_LEVELS];
    for (int idx = tid; idx < n; idx += blockDim.x) {
        int l = 0, base = 0;   // level of survivor idx and the first index of that level (selects: off[] stays in registers)
#pragma unroll
        for (int k = 1; k < ORBX_MAX_LEVELS; ++k) {
            const bool ge = idx >= off[k] && k < P.nlevels;
            l += ge ? 1 : 0;
            base = ge ? off[k] : base;
        }
        const uint32_t key = __ldcg(sv + P.lvl[l].surv_off + (idx - base));
        float x = (float)(ORBX_KEY_X(key) + ORBX_FAST_BORDER);
        if (l != 0) x = __fmul_rn(x, P.lvl[l].scale);
        flags[idx] = (x >= lap0 && x <= lap1) ? 1u : 0u;
    }
    __syncthreads();
    const uint32_t nstereo = block_excl_scan(flags, n, tmp);
    if (tid == 0) flags[n] = nstereo;
    __syncthreads();
    uint2* d = reinterpret_cast<uint2*>(dest) + (size_t)frame * P.surv_total;
    for (int idx = tid; idx < n; idx += blockDim.x) {
        const uint32_t sb = flags[idx];
        const bool stereo = flags[idx + 1] != sb;
        int l = 0, base = 0;   // level of survivor idx and the first index of that level (selects: off[] stays in registers)
#pragma unroll
        for (int k = 1; k < ORBX_MAX_LEVELS; ++k) {
            const bool ge = idx >= off[k] && k < P.nlevels;
            l += ge ? 1 : 0;
            base = ge ? off[k] : base;
        }
        const uint32_t key = __ldcg(sv + P.lvl[l].surv_off + (idx - base));
        const uint32_t row = stereo ? (uint32_t)(n - 1) - sb : (uint32_t)idx - sb;
        d[idx] = make_uint2(key, row | ((uint32_t)l << 24));
    }
    if (tid == 0) {
        int32_t* h = header + 4 * (size_t)frame;
        h[0] = n;
        h[1] = n - (int)nstereo;
        h[2] = n > P.out_cap ? ORBX_E_CAPACITY : ORBX_OK;
        h[3] = 0;
    }
}

#endif  // ORBX_DEVICE_CUH
