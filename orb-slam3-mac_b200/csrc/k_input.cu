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
// The device map holds, per output pixel, everything that does not depend on the frame: the
// top-left tap CLAMPED into the image, whether the right / bottom taps are one step away or collapse
// onto it, and the four 11-bit weights with the weights of taps outside the image zeroed -- so the
// kernel loads four always-valid bytes and needs no bounds logic:
//   {cx | cy << 16,  a00 | a01 << 16,  a10 | a11 << 16,  dx | dy << 1}          (16 bytes)
//
// Thread -> 4 adjacent output pixels of one row (one 32-bit store).  HBM/L2 bound: a frame reads its
// raw bytes once (the map entries are shared by every frame of a batch and stay in L2).
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
    const uint4* map;         // w*h entries (see above) or nullptr
    uint8_t* dst;
    long long dst_fstride;
    int dst_pitch;
    int vec_ok;               // raw rows are 4-byte aligned: colour pixels are read as whole words
};

__device__ __forceinline__ int in_luma(const InArgs& A, int c0, int c1, int c2) {
    const int r = A.rgb_order ? c0 : c2, b = A.rgb_order ? c2 : c0;
    return (r * A.cr + c1 * A.cg + b * A.cb + (1 << (A.bits - 1))) >> A.bits;
}

// one rectified pixel, channel values in v[0..channels)
__device__ __forceinline__ void in_remap_px(const InArgs& A, const uint8_t* __restrict__ S, uint4 m, int v[3]) {
    const int cx = (int)(m.x & 0xffffu), cy = (int)(m.x >> 16);
    const uint8_t* p00 = S + (long long)cy * A.raw_pitch + cx * A.channels;
    const int dx = (m.w & 1u) ? A.channels : 0;
    const uint8_t* p10 = p00 + ((m.w & 2u) ? A.raw_pitch : 0);
    const int nc = A.channels == 1 ? 1 : 3;   // alpha never reaches the luma
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        if (c < nc) {
            const uint32_t top = (uint32_t)p00[c] | ((uint32_t)p00[dx + c] << 16);
            const uint32_t bot = (uint32_t)p10[c] | ((uint32_t)p10[dx + c] << 16);
            // two 16x16 dot products (weights <= 1024, pixels <= 255: no overflow)
            const uint32_t acc = (m.y & 0xffffu) * (top & 0xffffu) + (m.y >> 16) * (top >> 16) +
                                 (m.z & 0xffffu) * (bot & 0xffffu) + (m.z >> 16) * (bot >> 16);
            v[c] = (int)((acc + 512u) >> 10);
        }
    }
}

// 4 output pixels of frame `S` (no rectification): colour -> luma, or a plain copy
__device__ __forceinline__ uint32_t in_convert4(const InArgs& A, const uint8_t* __restrict__ S, int x0, int y) {
    uint32_t out = 0;
    if (A.vec_ok && A.channels == 3 && x0 + 3 < A.w) {
        // 4 BGR pixels = 12 bytes = three aligned words
        const uint32_t* q = reinterpret_cast<const uint32_t*>(S + (long long)y * A.raw_pitch + x0 * 3);
        const uint32_t w0 = __ldg(q), w1 = __ldg(q + 1), w2 = __ldg(q + 2);
        const int g0 = in_luma(A, w0 & 0xff, (w0 >> 8) & 0xff, (w0 >> 16) & 0xff);
        const int g1 = in_luma(A, w0 >> 24, w1 & 0xff, (w1 >> 8) & 0xff);
        const int g2 = in_luma(A, (w1 >> 16) & 0xff, w1 >> 24, w2 & 0xff);
        const int g3 = in_luma(A, (w2 >> 8) & 0xff, (w2 >> 16) & 0xff, w2 >> 24);
        out = (uint32_t)g0 | ((uint32_t)g1 << 8) | ((uint32_t)g2 << 16) | ((uint32_t)g3 << 24);
    } else if (A.vec_ok && A.channels == 4 && x0 + 3 < A.w) {
        const uint4 p = __ldg(reinterpret_cast<const uint4*>(S + (long long)y * A.raw_pitch + x0 * 4));
        out = (uint32_t)in_luma(A, p.x & 0xff, (p.x >> 8) & 0xff, (p.x >> 16) & 0xff) |
              ((uint32_t)in_luma(A, p.y & 0xff, (p.y >> 8) & 0xff, (p.y >> 16) & 0xff) << 8) |
              ((uint32_t)in_luma(A, p.z & 0xff, (p.z >> 8) & 0xff, (p.z >> 16) & 0xff) << 16) |
              ((uint32_t)in_luma(A, p.w & 0xff, (p.w >> 8) & 0xff, (p.w >> 16) & 0xff) << 24);
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
    return out;
}

// Thread -> 4 adjacent output pixels (flattened over the image so that no lane idles on the row
// tail) of IN_FPT consecutive frames: the four map entries are loaded once and reused for every
// frame of the group -- read per frame, the map (16 B per pixel from L2) would outweigh the image.
#define IN_FPT 8

__global__ void __launch_bounds__(256)
k_input(const __grid_constant__ InArgs A, int nframes) {
    const int wq = (A.w + 3) >> 2;
    const int item = blockIdx.x * 256 + threadIdx.x;
    if (item >= wq * A.h) return;
    const int y = item / wq, x0 = (item - y * wq) * 4;
    const int f0 = blockIdx.y * IN_FPT, f1 = min(f0 + IN_FPT, nframes);
    const uint8_t* S = A.raw + (long long)f0 * A.raw_fstride;
    uint8_t* D = A.dst + (long long)f0 * A.dst_fstride + (long long)y * A.dst_pitch + x0;
    if (A.map) {
        uint4 m[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) m[k] = __ldg(A.map + (size_t)y * A.w + min(x0 + k, A.w - 1));
        for (int f = f0; f < f1; ++f, S += A.raw_fstride, D += A.dst_fstride) {
            uint32_t out = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                int v[3] = {0, 0, 0};
                in_remap_px(A, S, m[k], v);
                const int g = A.channels == 1 ? v[0] : in_luma(A, v[0], v[1], v[2]);
                out |= (uint32_t)g << (8 * k);
            }
            *reinterpret_cast<uint32_t*>(D) = out;
        }
    } else {
        for (int f = f0; f < f1; ++f, S += A.raw_fstride, D += A.dst_fstride)
            *reinterpret_cast<uint32_t*>(D) = in_convert4(A, S, x0, y);
    }
}

cudaError_t orbx_launch_input(const uint8_t* raw, long long raw_fstride, int raw_pitch, int w, int h, int channels,
                              int rgb_order, int gray_bits, const uint4* map, uint8_t* dst, long long dst_fstride,
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
    A.vec_ok = ((uintptr_t)raw % 4 == 0) && (raw_pitch % 4 == 0) && (raw_fstride % 4 == 0);
    dim3 grid((((w + 3) / 4) * h + 255) / 256, (nframes + IN_FPT - 1) / IN_FPT);
    k_input<<<grid, 256, 0, st>>>(A, nframes);
    ++*launches;
    return cudaGetLastError();
}
