// k_input.cu -- N4: the step in front of pyramid level 0.
//
//   cv::remap(im, imRect, M1, M2, cv::INTER_LINEAR)      Examples/Stereo/stereo_euroc.cc:170-171
//   cvtColor(mImGray, mImGray, CV_{RGB,BGR,RGBA,BGRA}2GRAY)   src/Tracking.cc:354-379, 425-438, 479-492
//
// in the reference's order (rectify every channel first, then luma), fused into one pass that
// writes the 8-bit level 0 the pyramid kernels read.  Arithmetic (oracle/cvprims.h, pinned against
// cv2 4.13): remap = OpenCV's fixed-point bilinear -- map coordinates are converted ONCE on the host
// to (integer part, 5-bit fractions) exactly like cv::remap does per call (cvRound(map * 32)),
// pixel = (a00*v00 + a01*v01 + a10*v10 + a11*v11 + 512) >> 10 with source pixels outside the image
// reading 0 (BORDER_CONSTANT); luma = (R*cr + G*cg + B*cb + half) >> bits with the 15-bit
// coefficients of OpenCV 4 or the 14-bit ones of OpenCV 2/3.
//
// Thread -> 4 adjacent output pixels of one row (one 32-bit store).  HBM/L2 bound: a frame reads its
// raw bytes once (the 8-byte map entries are shared by every frame of a batch and stay in L2).
#include "orbx_device.cuh"
#include "orbx_internal.h"

struct InArgs {
    const uint8_t* raw;       // device raw frames
    long long raw_fstride;    // bytes between frames
    int raw_pitch;            // bytes between rows
    int w, h;                 // image size (source == destination)
    int channels;             // 1, 3, 4
    int rgb_order;            // first channel is R
    int cr, cg, cb, bits;     // luma coefficients
    const uint2* map;         // w*h entries {ix & 0xffff | iy << 16, fx | fy << 8} or nullptr
    uint8_t* dst;
    long long dst_fstride;
    int dst_pitch;
};

__device__ __forceinline__ int in_luma(const InArgs& A, int c0, int c1, int c2) {
    const int r = A.rgb_order ? c0 : c2, b = A.rgb_order ? c2 : c0;
    return (r * A.cr + c1 * A.cg + b * A.cb + (1 << (A.bits - 1))) >> A.bits;
}

// one rectified pixel, channel values in v[0..channels)
__device__ __forceinline__ void in_remap_px(const InArgs& A, const uint8_t* __restrict__ S, uint2 m, int v[3]) {
    const int ix = (int)(short)(m.x & 0xffffu), iy = (int)(short)(m.x >> 16);
    const int fx = (int)(m.y & 0xffu), fy = (int)(m.y >> 8);
    const int a00 = (32 - fx) * (32 - fy), a01 = fx * (32 - fy), a10 = (32 - fx) * fy, a11 = fx * fy;
    const bool x0 = ix >= 0 && ix < A.w, x1 = ix + 1 >= 0 && ix + 1 < A.w;
    const bool y0 = iy >= 0 && iy < A.h, y1 = iy + 1 >= 0 && iy + 1 < A.h;
    const int nc = A.channels == 1 ? 1 : 3;   // alpha never reaches the luma
    const uint8_t* p00 = S + (long long)iy * A.raw_pitch + (long long)ix * A.channels;
    const uint8_t* p10 = p00 + A.raw_pitch;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        if (c < nc) {
            const int v00 = (x0 && y0) ? p00[c] : 0, v01 = (x1 && y0) ? p00[A.channels + c] : 0;
            const int v10 = (x0 && y1) ? p10[c] : 0, v11 = (x1 && y1) ? p10[A.channels + c] : 0;
            v[c] = (a00 * v00 + a01 * v01 + a10 * v10 + a11 * v11 + 512) >> 10;
        }
    }
}

__global__ void __launch_bounds__(128)
k_input(const __grid_constant__ InArgs A) {
    const int x0 = (blockIdx.x * 128 + threadIdx.x) * 4;
    const int y = blockIdx.y;
    if (x0 >= A.w) return;
    const uint8_t* S = A.raw + (long long)blockIdx.z * A.raw_fstride;
    uint32_t out = 0;
    if (A.map) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int x = min(x0 + k, A.w - 1);
            int v[3] = {0, 0, 0};
            in_remap_px(A, S, __ldg(A.map + (size_t)y * A.w + x), v);
            const int g = A.channels == 1 ? v[0] : in_luma(A, v[0], v[1], v[2]);
            out |= (uint32_t)g << (8 * k);
        }
    } else {
        const uint8_t* row = S + (long long)y * A.raw_pitch;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int x = min(x0 + k, A.w - 1);
            const uint8_t* p = row + (long long)x * A.channels;
            const int g = A.channels == 1 ? p[0] : in_luma(A, p[0], p[1], p[2]);
            out |= (uint32_t)g << (8 * k);
        }
    }
    *reinterpret_cast<uint32_t*>(A.dst + (long long)blockIdx.z * A.dst_fstride + (long long)y * A.dst_pitch + x0) = out;
}

cudaError_t orbx_launch_input(const uint8_t* raw, long long raw_fstride, int raw_pitch, int w, int h, int channels,
                              int rgb_order, int gray_bits, const uint2* map, uint8_t* dst, long long dst_fstride,
                              int dst_pitch, int nframes, cudaStream_t st, int* launches) {
    InArgs A;
    A.raw = raw; A.raw_fstride = raw_fstride; A.raw_pitch = raw_pitch;
    A.w = w; A.h = h; A.channels = channels; A.rgb_order = rgb_order;
    A.bits = gray_bits;
    A.cr = gray_bits == 15 ? 9798 : 4899;
    A.cg = gray_bits == 15 ? 19235 : 9617;
    A.cb = gray_bits == 15 ? 3735 : 1868;
    A.map = map;
    A.dst = dst; A.dst_fstride = dst_fstride; A.dst_pitch = dst_pitch;
    dim3 grid((w + 511) / 512, h, nframes);
    k_input<<<grid, 128, 0, st>>>(A);
    ++*launches;
    return cudaGetLastError();
}
