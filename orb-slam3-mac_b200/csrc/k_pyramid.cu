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
// Mapping (k_resize_rows): one thread owns 4 adjacent output columns (one 32-bit store per row) and
// walks 16 output rows.  Per thread, once: the two source-word bases of its pixel pairs, four
// PRMT selectors that pull (S[sx], S[sx+1]) out of a word pair, four u16x2 coefficient words.
// Per SOURCE row: 4 aligned word loads, 4 PRMT, 4 IDP.2A (H = S[sx]*a0 + S[sx+1]*a1), 4 shifts;
// the horizontal results of the lower source row are carried to the next output row (each source
// row is interpolated ~1.26 times instead of 2).  Per OUTPUT row: one 8-byte table load, 8
// IMAD.HI ((b*(H>>4))>>16 = umulhi(b<<16, H>>4), accumulating), 5 pack instructions, one store.
// A warp covers 128 contiguous output bytes of one 16-row band; warps are flattened over
// (band, column chunk) so that lanes are only idle at a level's right edge.
// k_resize_level (one thread per 4 pixels of ONE row, byte loads) is kept for scale factors
// whose source step exceeds 3 columns.
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

#define RZ_ROWS 16

// Horizontal pass of one source row for the thread's 4 columns, pre-shifted by 4 (VResizeLinear
// works on H >> 4).
__device__ __forceinline__ void rz_hrow(const uint32_t* __restrict__ row, int wa0, int wa1, int wb0,
                                        int wb1, const uint32_t (&sel)[4], const uint32_t (&cf)[4],
                                        uint32_t (&H)[4]) {
    const uint32_t A0 = __ldg(row + wa0), A1 = __ldg(row + wa1);
    const uint32_t B0 = __ldg(row + wb0), B1 = __ldg(row + wb1);
    H[0] = __dp2a_lo(cf[0], __byte_perm(A0, A1, sel[0]), 0u) >> 4;
    H[1] = __dp2a_lo(cf[1], __byte_perm(A0, A1, sel[1]), 0u) >> 4;
    H[2] = __dp2a_lo(cf[2], __byte_perm(B0, B1, sel[2]), 0u) >> 4;
    H[3] = __dp2a_lo(cf[3], __byte_perm(B0, B1, sel[3]), 0u) >> 4;
}

__global__ void __launch_bounds__(256)
k_resize_rows(const __grid_constant__ OrbxPlan P, const OrbxSrc isrc, int level, int warps_x,
              int nbands, uint8_t* __restrict__ pyr) {
    const OrbxLevel& D = P.lvl[level];
    const OrbxLevel& S = P.lvl[level - 1];
    const int gw = blockIdx.x * 8 + (threadIdx.x >> 5);   // warp -> (band, column chunk)
    const int band = gw / warps_x, wx = gw - band * warps_x;
    if (band >= nbands) return;
    const int x0 = (wx * 32 + (threadIdx.x & 31)) * 4;
    if (x0 >= D.w) return;
    int spitch;
    const uint8_t* src = orbx_level_ptr(P, isrc, pyr, blockIdx.y, level - 1, &spitch);
    uint8_t* dst = pyr + (size_t)blockIdx.y * P.frame_stride + D.off;

    // per-thread column setup
    int sx[4];
    uint32_t cf[4], sel[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int x = min(x0 + k, D.w - 1);
        sx[k] = D.xofs[x];
        cf[k] = reinterpret_cast<const uint32_t*>(D.xa)[x];  // a0 | a1 << 16
    }
    const int wlast = (S.w - 1) >> 2;  // last source word holding valid bytes
    const int wa0 = sx[0] >> 2, wa1 = min(wa0 + 1, wlast);
    const int wb0 = sx[2] >> 2, wb1 = min(wb0 + 1, wlast);
    {
        // byte offsets of S[sx] inside its word pair; S[sx+1] is the next byte (offset <= 6:
        // x_step_max <= 3 is checked by the launcher).  A clamped upper word is only ever read
        // for a byte whose coefficient is 0 (sx = S.w - 1).
        const int o0 = sx[0] - 4 * wa0, o1 = sx[1] - 4 * wa0, o2 = sx[2] - 4 * wb0, o3 = sx[3] - 4 * wb0;
        sel[0] = (uint32_t)o0 | ((uint32_t)(o0 + 1) << 4);
        sel[1] = (uint32_t)o1 | ((uint32_t)(o1 + 1) << 4);
        sel[2] = (uint32_t)o2 | ((uint32_t)(o2 + 1) << 4);
        sel[3] = (uint32_t)o3 | ((uint32_t)(o3 + 1) << 4);
    }
    const int sp4 = spitch >> 2;
    const uint32_t* src4 = reinterpret_cast<const uint32_t*>(src);

    const int y0 = band * RZ_ROWS, y1 = min(y0 + RZ_ROWS, D.h);
    uint32_t HA[4], HB[4];
    int ra = -1, rb = -1;  // source rows held in HA / HB
    uint8_t* drow = dst + (size_t)y0 * D.pitch + x0;
    for (int y = y0; y < y1; ++y, drow += D.pitch) {
        const uint2 yt = __ldg(D.ytab + y);
        const int s0 = (int)yt.x, s1 = min(s0 + 1, S.h - 1);
        if (s0 != ra) {
            if (s0 == rb) {
#pragma unroll
                for (int k = 0; k < 4; ++k) HA[k] = HB[k];
            } else {
                rz_hrow(src4 + (size_t)s0 * sp4, wa0, wa1, wb0, wb1, sel, cf, HA);
            }
            ra = s0;
        }
        if (s1 == ra) {
#pragma unroll
            for (int k = 0; k < 4; ++k) HB[k] = HA[k];
            rb = ra;
        } else if (s1 != rb) {
            rz_hrow(src4 + (size_t)s1 * sp4, wa0, wa1, wb0, wb1, sel, cf, HB);
            rb = s1;
        }
        const uint32_t b0 = yt.y << 16, b1 = yt.y & 0xffff0000u;
        uint32_t t[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) t[k] = __umulhi(b0, HA[k]) + (__umulhi(b1, HB[k]) + 2u);
        const uint32_t t01 = __byte_perm(t[0], t[1], 0x5410) >> 2;
        const uint32_t t23 = __byte_perm(t[2], t[3], 0x5410) >> 2;
        // pitch is a multiple of 16 and x0 of 4: the 32-bit store is aligned; bytes past w land
        // in the row's padding.
        *reinterpret_cast<uint32_t*>(drow) = __byte_perm(t01, t23, 0x6420);
    }
}

void orbx_launch_pyramid(const OrbxPlan& P, const OrbxSrc& src, uint8_t* pyr, int nframes, cudaStream_t st, int* launches) {
    for (int l = 1; l < P.nlevels; ++l) {
        const OrbxLevel& D = P.lvl[l];
        if (D.x_step_max <= 3) {
            const int warps_x = (D.w + 127) / 128, nbands = (D.h + RZ_ROWS - 1) / RZ_ROWS;
            dim3 grid((warps_x * nbands + 7) / 8, nframes);
            k_resize_rows<<<grid, 256, 0, st>>>(P, src, l, warps_x, nbands, pyr);
        } else {
            dim3 block(128);
            dim3 grid((D.w + 4 * 128 - 1) / (4 * 128), D.h, nframes);
            k_resize_level<<<grid, block, 0, st>>>(P, src, l, pyr);
        }
        ++*launches;
    }
}
