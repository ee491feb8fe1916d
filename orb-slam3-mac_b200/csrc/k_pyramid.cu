// k_pyramid.cu -- E1: ORBextractor::ComputePyramid (src/ORBextractor.cc:1152-1177).
//
// Level l is cv::resize(level l-1, INTER_LINEAR) of the previous ROUNDED 8-bit level.  The
// arithmetic is OpenCV's 8-bit fixed-point bilinear (HResizeLinear / VResizeLinear, 11-bit
// coefficients): H = S[sx]*a0 + S[sx+1]*a1; dst = (((b0*(H0>>4))>>16) + ((b1*(H1>>4))>>16) + 2) >> 2.
// The per-column / per-row (offset, coefficient) tables are built once per image shape on the
// host (orbx_api.cu, build_plan) and live in HBM/L2.  The reflect-101 border the reference
// also writes (copyMakeBorder) is never read by the path and is not materialised on device;
// orbx_get_pyramid_level(pad=19) reproduces it on demand.
//
// Mapping: one thread produces 4 adjacent output pixels of one row (one 32-bit store); a warp
// covers 128 contiguous output bytes.  Source bytes come through the read-only path; each
// source byte feeds ~1.2 outputs horizontally and vertically, which L1 absorbs.
#include "orbx_device.cuh"

__global__ void __launch_bounds__(256)
k_resize_level(const __grid_constant__ OrbxPlan P, const OrbxSrc isrc, int level,
               uint8_t* __restrict__ pyr) {
    const OrbxLevel& D = P.lvl[level];
    const OrbxLevel& S = P.lvl[level - 1];
    const int y = blockIdx.y;
    const int x0 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (x0 >= D.w) return;
    int spitch;
    const uint8_t* src = orbx_level_ptr(P, isrc, pyr, blockIdx.z, level - 1, &spitch);
    uint8_t* dst = pyr + (size_t)blockIdx.z * P.frame_stride + D.off;

    const int sy0 = D.yofs[y];
    const int sy1 = min(sy0 + 1, S.h - 1);
    const short2 bb = reinterpret_cast<const short2*>(D.ya)[y];
    const uint8_t* r0 = src + (size_t)sy0 * spitch;
    const uint8_t* r1 = src + (size_t)sy1 * spitch;

    uint32_t out = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int x = min(x0 + k, D.w - 1);
        const int sx = D.xofs[x];
        const int sx1 = min(sx + 1, S.w - 1);
        const short2 aa = reinterpret_cast<const short2*>(D.xa)[x];
        const int h0 = __ldg(r0 + sx) * aa.x + __ldg(r0 + sx1) * aa.y;
        const int h1 = __ldg(r1 + sx) * aa.x + __ldg(r1 + sx1) * aa.y;
        const int v = (((bb.x * (h0 >> 4)) >> 16) + ((bb.y * (h1 >> 4)) >> 16) + 2) >> 2;
        out |= (uint32_t)(v & 0xff) << (8 * k);
    }
    // pitch is a multiple of 16 and x0 of 4: the 32-bit store is aligned; bytes past w land in
    // the row's padding.
    *reinterpret_cast<uint32_t*>(dst + (size_t)y * D.pitch + x0) = out;
}

void orbx_launch_pyramid(const OrbxPlan& P, const OrbxSrc& src, uint8_t* pyr, int nframes, cudaStream_t st, int* launches) {
    for (int l = 1; l < P.nlevels; ++l) {
        const OrbxLevel& D = P.lvl[l];
        dim3 block(128);
        dim3 grid((D.w + 4 * 128 - 1) / (4 * 128), D.h, nframes);
        k_resize_level<<<grid, block, 0, st>>>(P, src, l, pyr);
        ++*launches;
    }
}
