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
#include <cstdlib>

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

#ifndef RZ_ROWS
#define RZ_ROWS 16
#endif

// Everything k_resize_rows needs about one (source level -> destination level) step, as plain
// kernel-parameter scalars (constant-bank operands; no indexed loads from the plan).
struct RzArgs {
    const uint8_t* src;      // source level of frame 0
    uint8_t* dst;            // destination level of frame 0
    long long src_fstride;   // bytes between frames
    long long dst_fstride;
    const uint16_t* xofs;    // [dw]
    const uint32_t* xa;      // [dw] a0 | a1 << 16
    const uint2* ytab;       // [dh] {s0, b0 | b1 << 16}, bottom source row = s0 + 1
    int spitch, dpitch;
    int sw, sh, dw, dh;
    int warps_x, nbands, rows;   // rows = output rows per warp (band height)
};

// Horizontal pass of one source row for the thread's 4 columns from its 4 loaded words,
// pre-shifted by 4 (VResizeLinear works on H >> 4).
__device__ __forceinline__ void rz_hrow(const uint32_t (&W)[4], const uint32_t (&sel)[4],
                                        const uint32_t (&cf)[4], uint32_t (&H)[4]) {
    H[0] = __dp2a_lo(cf[0], __byte_perm(W[0], W[1], sel[0]), 0u) >> 4;
    H[1] = __dp2a_lo(cf[1], __byte_perm(W[0], W[1], sel[1]), 0u) >> 4;
    H[2] = __dp2a_lo(cf[2], __byte_perm(W[2], W[3], sel[2]), 0u) >> 4;
    H[3] = __dp2a_lo(cf[3], __byte_perm(W[2], W[3], sel[3]), 0u) >> 4;
}

// One output row: dst = ((b0*T>>16) + (b1*B>>16) + 2) >> 2 for the 4 columns, one 32-bit store.
__device__ __forceinline__ void rz_emit(uint8_t* drow, uint32_t bw, const uint32_t (&T)[4],
                                        const uint32_t (&B)[4]) {
    const uint32_t b0 = bw << 16, b1 = bw & 0xffff0000u;
    uint32_t t[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) t[k] = __umulhi(b0, T[k]) + (__umulhi(b1, B[k]) + 2u);
    const uint32_t t01 = __byte_perm(t[0], t[1], 0x5410) >> 2;
    const uint32_t t23 = __byte_perm(t[2], t[3], 0x5410) >> 2;
    // pitch is a multiple of 16 and x0 of 4: the 32-bit store is aligned; bytes past w land in
    // the row's padding.
    *reinterpret_cast<uint32_t*>(drow) = __byte_perm(t01, t23, 0x6420);
}

__global__ void __launch_bounds__(256)
k_resize_rows(const __grid_constant__ RzArgs A) {
    const int gw = blockIdx.x * 8 + (threadIdx.x >> 5);   // warp -> (band, column chunk)
    const int band = gw / A.warps_x, wx = gw - band * A.warps_x;
    if (band >= A.nbands) return;
    const int x0 = (wx * 32 + (threadIdx.x & 31)) * 4;
    if (x0 >= A.dw) return;

    // per-thread column setup
    int sx[4];
    uint32_t cf[4], sel[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int x = min(x0 + k, A.dw - 1);
        sx[k] = A.xofs[x];
        cf[k] = A.xa[x];  // a0 | a1 << 16
    }
    // Word pair (w, w+1) of each pixel pair, kept inside the row's valid words; S[sx] sits at byte
    // offset o <= 7 of the pair and S[sx+1] at o+1 (x_step_max <= 3 is checked by the launcher).
    // o+1 == 8 only happens for sx == sw-1, whose second coefficient is 0.
    const int wlast = (A.sw - 1) >> 2;
    const int wa = min(sx[0] >> 2, wlast - 1), wb = min(sx[2] >> 2, wlast - 1);
    {
        const int o0 = sx[0] - 4 * wa, o1 = sx[1] - 4 * wa, o2 = sx[2] - 4 * wb, o3 = sx[3] - 4 * wb;
        sel[0] = (uint32_t)o0 | ((uint32_t)min(o0 + 1, 7) << 4);
        sel[1] = (uint32_t)o1 | ((uint32_t)min(o1 + 1, 7) << 4);
        sel[2] = (uint32_t)o2 | ((uint32_t)min(o2 + 1, 7) << 4);
        sel[3] = (uint32_t)o3 | ((uint32_t)min(o3 + 1, 7) << 4);
    }
    const int y0 = band * A.rows, y1 = min(y0 + A.rows, A.dh);
    uint8_t* drow = A.dst + (size_t)blockIdx.y * A.dst_fstride + (size_t)y0 * A.dpitch + x0;
    const uint2* yp = A.ytab + y0;
    uint2 yt = __ldg(yp);
    int y = y0;
    // Source rows are visited once each, in increasing order, starting at an even row: even rows
    // go through (W0 -> E), odd rows through (W1 -> O); two rows of loads are always in flight.
    int s = (int)yt.x & ~1;
    const uint8_t* sbase = A.src + (size_t)blockIdx.y * A.src_fstride;
    const uint8_t* pa = sbase + 4 * wa;
    const uint8_t* pb = sbase + 4 * wb;
    uint32_t ro = (uint32_t)s * (uint32_t)A.spitch;             // byte offset of the row to load
    const uint32_t ro_max = (uint32_t)(A.sh - 1) * (uint32_t)A.spitch;

    uint32_t E[4], O[4], W0[4], W1[4];
#define RZ_LOAD(W)                                                          \
    {                                                                       \
        const uint32_t* qa = reinterpret_cast<const uint32_t*>(pa + ro);    \
        const uint32_t* qb = reinterpret_cast<const uint32_t*>(pb + ro);    \
        W[0] = __ldg(qa); W[1] = __ldg(qa + 1);                             \
        W[2] = __ldg(qb); W[3] = __ldg(qb + 1);                             \
        ro = min(ro + (uint32_t)A.spitch, ro_max);                          \
    }
    // After source row s is interpolated into CUR, the output row whose bottom source row is s
    // (ytab: s0 + 1 == s; strictly increasing s0, checked by the launcher) is complete.
#define RZ_STEP(CUR, PREV, W)                                               \
    {                                                                       \
        rz_hrow(W, sel, cf, CUR);                                           \
        RZ_LOAD(W)                                                          \
        if ((int)yt.x + 1 == s) {                                           \
            rz_emit(drow, yt.y, PREV, CUR);                                 \
            if (++y >= y1) break;                                           \
            drow += A.dpitch;                                               \
            yt = __ldg(++yp);                                               \
        }                                                                   \
        ++s;                                                                \
    }
    RZ_LOAD(W0)
    RZ_LOAD(W1)
#pragma unroll 1
    while (s < A.sh) {   // the break above is the normal exit; this bound only guards bad tables
        RZ_STEP(E, O, W0)
        RZ_STEP(O, E, W1)
    }
#undef RZ_STEP
#undef RZ_LOAD
}

// ---- cp.async-staged variant ------------------------------------------------------------------
// Same arithmetic and walk as k_resize_rows, with PX = 4 or 8 adjacent output columns per thread,
// but the source rows a warp needs (its 32*PX output columns = <= 256 / 384 source bytes) are
// streamed into a per-warp shared-memory ring by 16-byte cp.async.cg copies, RZ_DEPTH-1 rows
// ahead of their use, so that DRAM latency is covered by ~1 KB of loads in flight per warp
// instead of two dependent row loads.  PX = 8 halves the per-row bookkeeping per pixel (ring
// wait, copy issue, row-table lookup, stores) and is used when the level is wide enough to fill
// the lanes.  Needs 16-byte aligned source rows (always true for pyramid levels; level 0 when
// the caller's pitch allows it).
#ifndef RZ_DEPTH
#define RZ_DEPTH 4
#endif

__device__ __forceinline__ void rz_cp16(uint32_t saddr, const void* g) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(saddr), "l"(g) : "memory");
}
__device__ __forceinline__ void rz_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void rz_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <int PX>
struct RzCfg {
    static constexpr int SLOT_BYTES = PX == 4 ? 256 : 384;   // staged segment of one source row
    static constexpr int PAIRS = PX / 2;
};

// One warp: 32*PX output columns starting at xbase, output rows of `band`.
template <int PX>
__device__ __forceinline__ void rz_walk(const RzArgs& A, int band, int xbase, uint8_t* ring) {
    constexpr int SLOT = RzCfg<PX>::SLOT_BYTES, NP = RzCfg<PX>::PAIRS;
    const int lane = threadIdx.x & 31;
    const int x0 = xbase + lane * PX;
    const bool live = x0 < A.dw;              // dead lanes still help with the copies
    const int xc = live ? x0 : xbase + ((A.dw - 1 - xbase) / PX) * PX;

    int sx[PX];
    uint32_t cf[PX], sel[PX];
#pragma unroll
    for (int k = 0; k < PX; ++k) {
        const int x = min(xc + k, A.dw - 1);
        sx[k] = A.xofs[x];
        cf[k] = A.xa[x];
    }
    // Word pair (w, w+1) of each pixel pair, kept inside the row's valid words; S[sx] sits at byte
    // offset o <= 7 of the pair and S[sx+1] at o+1 (o+1 == 8 only for sx == sw-1: coefficient 0).
    const int wlast = (A.sw - 1) >> 2;
    int wp[NP];
#pragma unroll
    for (int q = 0; q < NP; ++q) {
        wp[q] = min(sx[2 * q] >> 2, wlast - 1);
        const int o0 = sx[2 * q] - 4 * wp[q], o1 = sx[2 * q + 1] - 4 * wp[q];
        sel[2 * q] = (uint32_t)o0 | ((uint32_t)min(o0 + 1, 7) << 4);
        sel[2 * q + 1] = (uint32_t)o1 | ((uint32_t)min(o1 + 1, 7) << 4);
    }
    // staged segment of a source row: 16-byte chunks [c0, c0 + nchunk) covering every lane's words
    const int w_first = __shfl_sync(ORBX_FULL, wp[0], 0) & ~3;
    const int w_end = __reduce_max_sync(ORBX_FULL, wp[NP - 1] + 2);   // one past the last word needed
    const int nchunk = min((w_end - w_first + 3) >> 2, SLOT / 16);

    const int y0 = band * A.rows, y1 = min(y0 + A.rows, A.dh);
    uint8_t* drow = A.dst + (size_t)blockIdx.y * A.dst_fstride + (size_t)y0 * A.dpitch + xc;
    const uint2* yp = A.ytab + y0;
    uint2 yt = __ldg(yp);
    int y = y0;
    int s = (int)yt.x & ~1;
    const uint8_t* gsrc = A.src + (size_t)blockIdx.y * A.src_fstride + 4 * (size_t)w_first + 16 * lane;
    uint32_t ro = (uint32_t)s * (uint32_t)A.spitch;
    const uint32_t ro_max = (uint32_t)(A.sh - 1) * (uint32_t)A.spitch;
    const uint32_t ring_sa = (uint32_t)__cvta_generic_to_shared(ring);
    uint32_t cp_sa = ring_sa + 16 * lane;                // this lane's 16-byte chunk of slot 0
    uint32_t wsa[NP];
#pragma unroll
    for (int q = 0; q < NP; ++q) wsa[q] = ring_sa + 4 * (wp[q] - w_first);
    uint32_t spitch_r = (uint32_t)A.spitch;
    // opaque to the compiler: otherwise it rematerialises the shared-window address arithmetic
    // (S2R + LEA chain) inside every step of the loop
    asm volatile("" : "+r"(cp_sa), "+r"(spitch_r));
#pragma unroll
    for (int q = 0; q < NP; ++q) asm volatile("" : "+r"(wsa[q]));
    const bool copier = lane < nchunk;
    uint32_t ioff = (uint32_t)(s % RZ_DEPTH) * SLOT;     // slot being staged
    uint32_t soff = ioff;                                // slot being consumed

#define RZ_ISSUE()                                                                          \
    {                                                                                       \
        if (copier) rz_cp16(cp_sa + ioff, gsrc + ro);                                       \
        rz_commit();                                                                        \
        ro = min(ro + spitch_r, ro_max);                                                    \
        ioff += SLOT;                                                                       \
        if (ioff == RZ_DEPTH * SLOT) ioff = 0;                                              \
    }
#pragma unroll
    for (int i = 0; i < RZ_DEPTH - 1; ++i) RZ_ISSUE()

    uint32_t E[PX], O[PX], W[PX];
#define RZ_LDS(dst, addr) asm volatile("ld.shared.u32 %0, [%1];" : "=r"(dst) : "r"(addr))
#define RZ_STEP(CUR, PREV)                                                                  \
    {                                                                                       \
        rz_wait<RZ_DEPTH - 2>();                                                            \
        __syncwarp();                                                                       \
        _Pragma("unroll") for (int q = 0; q < NP; ++q) {                                    \
            RZ_LDS(W[2 * q], wsa[q] + soff);                                                \
            RZ_LDS(W[2 * q + 1], wsa[q] + soff + 4);                                        \
        }                                                                                   \
        soff += SLOT;                                                                       \
        if (soff == RZ_DEPTH * SLOT) soff = 0;                                              \
        RZ_ISSUE()                                                                          \
        _Pragma("unroll") for (int q = 0; q < NP; ++q) {                                    \
            CUR[2 * q] = __dp2a_lo(cf[2 * q], __byte_perm(W[2 * q], W[2 * q + 1], sel[2 * q]), 0u) >> 4;             \
            CUR[2 * q + 1] = __dp2a_lo(cf[2 * q + 1], __byte_perm(W[2 * q], W[2 * q + 1], sel[2 * q + 1]), 0u) >> 4; \
        }                                                                                   \
        if ((int)yt.x + 1 == s) {                                                           \
            if (live) {                                                                     \
                const uint32_t b0 = yt.y << 16, b1 = yt.y & 0xffff0000u;                    \
                uint32_t o[PX / 4];                                                         \
                _Pragma("unroll") for (int g = 0; g < PX / 4; ++g) {                        \
                    uint32_t t[4];                                                          \
                    _Pragma("unroll") for (int k = 0; k < 4; ++k)                           \
                        t[k] = __umulhi(b0, PREV[4 * g + k]) + (__umulhi(b1, CUR[4 * g + k]) + 2u); \
                    o[g] = __byte_perm(__byte_perm(t[0], t[1], 0x5410) >> 2,                \
                                       __byte_perm(t[2], t[3], 0x5410) >> 2, 0x6420);       \
                }                                                                           \
                if (PX == 4) *reinterpret_cast<uint32_t*>(drow) = o[0];                     \
                else *reinterpret_cast<uint2*>(drow) = make_uint2(o[0], o[PX / 4 - 1]);     \
            }                                                                               \
            if (++y >= y1) break;                                                           \
            drow += A.dpitch;                                                               \
            yt = __ldg(++yp);                                                               \
        }                                                                                   \
        ++s;                                                                                \
    }
#pragma unroll 1
    while (s < A.sh) {
        RZ_STEP(E, O)
        RZ_STEP(O, E)
    }
#undef RZ_LDS
#undef RZ_STEP
#undef RZ_ISSUE
    rz_wait<0>();
}

// warps_x = n8 + n4 warps per band: the first n8 take 256 columns each (8 per thread), the rest
// 128 columns each (4 per thread), so that a level's width never leaves most of a warp idle.
#ifndef RZ_MINB
#define RZ_MINB 4   // 64 registers, no spill (A/B: 5 -> 48 regs + 88 B of spill 0.386 ms, 6 -> 40 regs 0.493 ms, against 0.314)
#endif
__global__ void __launch_bounds__(256, RZ_MINB)
k_resize_rows_async(const __grid_constant__ RzArgs A, int n8) {
    __shared__ __align__(16) uint8_t ring_all[8 * RZ_DEPTH * RzCfg<8>::SLOT_BYTES];
    const int gw = blockIdx.x * 8 + (threadIdx.x >> 5);   // warp -> (band, column chunk)
    const int band = gw / A.warps_x, wx = gw - band * A.warps_x;
    if (band >= A.nbands) return;
    uint8_t* ring = ring_all + (threadIdx.x >> 5) * (RZ_DEPTH * RzCfg<8>::SLOT_BYTES);
    if (wx < n8) rz_walk<8>(A, band, wx * 256, ring);
    else rz_walk<4>(A, band, n8 * 256 + (wx - n8) * 128, ring);
}

// ---- the coarse levels of a SMALL batch in one launch (single-frame latency path) ----------------------------------
// For one frame the per-level launches of levels 4..7 are four dependent kernels of ~5 us each that are almost all
// launch + first-load latency (211 k pixels together).  k_resize_tail runs them as phases of ONE launch: a thread-block
// cluster per frame (RZ_TAIL_CLUSTER CTAs; the cluster's warps share a level's (band, column chunk) units), a
// cluster barrier between levels.  A level is written to HBM/L2 as before and read back by the next phase through
// cp.async.cg (L2, never a stale L1 line), after __threadfence() + barrier.cluster release/acquire.
// MEASURED AND LEFT OFF (bit-exact, numbers in profiles/README.md): the single-frame p50 gets WORSE
// (C1 102 -> 109 us from level 4, 129 us from level 2) -- eight SMs per frame instead of every SM per level, and the
// FAST / quadtree branches of levels 4..7 all wait for the end of the launch instead of starting level by level.
#ifndef RZ_TAIL_WARPS
#define RZ_TAIL_WARPS 16
#endif
#ifndef RZ_TAIL_CLUSTER
#define RZ_TAIL_CLUSTER 8
#endif
#define RZ_TAIL_MAX 6
struct RzTail {
    RzArgs a[RZ_TAIL_MAX];
    int n8[RZ_TAIL_MAX];
    int n;
};

__global__ void __launch_bounds__(RZ_TAIL_WARPS * 32)
k_resize_tail(const __grid_constant__ RzTail T) {
    __shared__ __align__(16) uint8_t ring_all[RZ_TAIL_WARPS * RZ_DEPTH * RzCfg<8>::SLOT_BYTES];
    const int warp = threadIdx.x >> 5;
    uint8_t* ring = ring_all + warp * (RZ_DEPTH * RzCfg<8>::SLOT_BYTES);
    const int stride = gridDim.x * RZ_TAIL_WARPS;   // gridDim.x = the cluster's CTAs, blockIdx.y = frame
    for (int p = 0; p < T.n; ++p) {
        const RzArgs& A = T.a[p];
        const int n8 = T.n8[p];
        const int total = A.warps_x * A.nbands;
        for (int gw = blockIdx.x * RZ_TAIL_WARPS + warp; gw < total; gw += stride) {
            const int band = gw / A.warps_x, wx = gw - band * A.warps_x;
            if (wx < n8) rz_walk<8>(A, band, wx * 256, ring);
            else rz_walk<4>(A, band, n8 * 256 + (wx - n8) * 128, ring);
            __syncwarp();   // every lane is done with the ring before the next unit stages into it
        }
        if (p + 1 < T.n) {
            __threadfence();
            asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
        }
    }
}

#ifndef RZ_MIN_WARPS
#define RZ_MIN_WARPS (148 * 16)
#endif
// Levels [from, nlevels) in one k_resize_tail launch when the batch is small and every one of them qualifies for the
// cp.async kernel; returns the first level it did NOT take (nlevels when it ran, `from` when it did not).
static bool rz_args(const OrbxPlan& P, const OrbxSrc& src, uint8_t* pyr, int nframes, int l, RzArgs* out, int* n8_out, bool* async_out);
// OFF by default (measured slower, see the comment above): ORBX_TAIL_FROM=<first level, >= 2> switches it on for
// launches of at most sms / cluster frames, ORBX_TAIL_BATCH=1 for every batch, ORBX_TAIL_CLUSTER=1|2|4|8 sets the
// cluster size (developer A/B, tools/tail_ab.sh).
static int rz_env(const char* name, int dflt) {
    const char* e = getenv(name);
    return e ? atoi(e) : dflt;
}
static int rz_tail_from() { static const int v = rz_env("ORBX_TAIL_FROM", 0); return v; }
static int rz_tail_batch() { static const int v = rz_env("ORBX_TAIL_BATCH", 0); return v; }
static int rz_tail_cluster() {
    static const int v = [] { const int c = rz_env("ORBX_TAIL_CLUSTER", RZ_TAIL_CLUSTER); return (c == 1 || c == 2 || c == 4 || c == 8) ? c : RZ_TAIL_CLUSTER; }();
    return v;
}

void orbx_launch_pyramid(const OrbxPlan& P, const OrbxSrc& src, uint8_t* pyr, int nframes, cudaStream_t st, int* launches,
                         cudaEvent_t* level_done) {
    // small batch: the coarse levels as phases of one cluster launch
    int tail = P.nlevels;
    RzTail T;
    T.n = 0;
    {
        const int from = rz_tail_from();
        int dev = 0, sms = 148;
        if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) {
            cudaGetLastError();
            sms = 148;
        }
        if (from >= 2 && from < P.nlevels && P.nlevels - from <= RZ_TAIL_MAX && (rz_tail_batch() || nframes * rz_tail_cluster() <= sms)) {
            bool ok = true;
            for (int l = from; l < P.nlevels && ok; ++l) {
                bool async_ok = false;
                ok = rz_args(P, src, pyr, nframes, l, &T.a[l - from], &T.n8[l - from], &async_ok) && async_ok;
            }
            if (ok) {
                tail = from;
                T.n = P.nlevels - from;
            }
        }
    }
    for (int l = 1; l < tail; ++l) {
        const OrbxLevel& D = P.lvl[l];
        RzArgs A;
        int n8 = 0;
        bool async_ok = false;
        if (rz_args(P, src, pyr, nframes, l, &A, &n8, &async_ok)) {
            dim3 grid((A.warps_x * A.nbands + 7) / 8, nframes);
            if (async_ok)
                k_resize_rows_async<<<grid, 256, 0, st>>>(A, n8);
            else
                k_resize_rows<<<grid, 256, 0, st>>>(A);
        } else {
            dim3 block(128);
            dim3 grid((D.w + 4 * 128 - 1) / (4 * 128), D.h, nframes);
            k_resize_level<<<grid, block, 0, st>>>(P, src, l, pyr);
        }
        ++*launches;
        if (level_done) cudaEventRecord(level_done[l], st);   // level l is complete (forked latency graph)
    }
    if (T.n) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(rz_tail_cluster(), nframes);
        cfg.blockDim = dim3(RZ_TAIL_WARPS * 32);
        cfg.stream = st;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = rz_tail_cluster();
        at[0].val.clusterDim.y = 1;
        at[0].val.clusterDim.z = 1;
        cfg.attrs = at;
        cfg.numAttrs = 1;
        cudaLaunchKernelEx(&cfg, k_resize_tail, T);
        ++*launches;
        if (level_done)
            for (int l = tail; l < P.nlevels; ++l) cudaEventRecord(level_done[l], st);
    }
}

// The (source level -> level l) step as k_resize_rows / k_resize_rows_async arguments; false when the level needs
// k_resize_level (source step above 3 columns, non-monotone rows, tiny source).
static bool rz_args(const OrbxPlan& P, const OrbxSrc& src, uint8_t* pyr, int nframes, int l, RzArgs* out, int* n8_out,
                    bool* async_out) {
    const OrbxLevel& D = P.lvl[l];
    const OrbxLevel& S = P.lvl[l - 1];
    if (!(D.x_step_max <= 3 && D.y_monotone && S.w >= 8 && S.h >= 2)) return false;
    RzArgs A;
    A.src = l == 1 ? src.img0 : pyr + S.off;
    A.src_fstride = l == 1 ? src.stride0 : P.frame_stride;
    A.spitch = l == 1 ? src.pitch0 : S.pitch;
    A.dst = pyr + D.off;
    A.dst_fstride = P.frame_stride;
    A.dpitch = D.pitch;
    A.xofs = D.xofs;
    A.xa = reinterpret_cast<const uint32_t*>(D.xa);
    A.ytab = D.ytab;
    A.sw = S.w; A.sh = S.h; A.dw = D.w; A.dh = D.h;
    // output rows per warp: RZ_ROWS when the launch fills the GPU anyway; small batches take
    // shorter bands so that a level of one frame still spreads over the SMs (latency path)
    A.rows = RZ_ROWS;
    while (A.rows > 2 && (long long)nframes * ((D.w + 255) / 256) * ((D.h + A.rows - 1) / A.rows) < RZ_MIN_WARPS)
        A.rows >>= 1;
    A.nbands = (D.h + A.rows - 1) / A.rows;
    const bool aligned = ((uintptr_t)A.src % 16 == 0) && (A.spitch % 16 == 0) && (A.src_fstride % 16 == 0);
    const double ratio = (double)S.w / D.w;
    const bool async_ok = aligned && D.x_step_max <= 2 && ratio <= 1.7;
    // 8 columns per thread need the 384-byte slot: 256 * ratio + 32 bytes
    const int n8 = (async_ok && ratio <= 1.35) ? D.w / 256 : 0;
    A.warps_x = async_ok ? n8 + (D.w - 256 * n8 + 127) / 128 : (D.w + 127) / 128;
    *out = A;
    *n8_out = n8;
    *async_out = async_ok;
    return true;
}
