// k_describe.cu -- E7 layout + E4 orientation + E5 blur + E6 rotated BRIEF, fused.
//
//   k_assemble   one CTA per frame: prefix over the levels' survivors, lapping-area test on the
//                scaled x, stable TWO-ENDED compaction (ORBextractor.cc:1124-1144): a keypoint
//                with lap0 <= pt.x <= lap1 goes to the back (stereoIndex--), others to the front
//                (monoIndex++), both in level-major order.  Writes every survivor's output row
//                and the frame header {n, monoIndex, status}.
//   k_describe   one warp per keypoint.  The 43x43 neighbourhood of the UN-blurred level is
//                staged in shared memory once (aligned 32-bit loads; reflect-101 of the level at
//                image borders, which is what GaussianBlur sees on the reference's clone(),
//                ORBextractor.cc:1114-1115) and serves three consumers:
//                  - IC_Angle (:75-102): one lane per row of the radius-15 disc, integer
//                    moments, shuffle reduction, cv::fastAtan2's polynomial evaluated with
//                    explicitly non-contracted float ops;
//                  - the horizontal pass of the 7x7 sigma-2 Gaussian in OpenCV 4's 8.8 fixed
//                    point ({18,34,48,56,48,34,18}/256, exact in 16 bits) over 43 rows x 37 cols;
//                  - the vertical pass, evaluated ONLY at the 512 rotated sample positions
//                    ((sum + 32768) >> 16), i.e. the blur is never written to HBM.
//                Lane i produces descriptor byte i (16 samples, 8 comparisons).
#include "orbx_device.cuh"

#define DESC_WARPS 8
#define DESC_WP 13                       // words per staged row (52 B; odd -> conflict-free rows)
#define DESC_HP 19                       // words per H row (38 u16)
#define DESC_S_WORDS (43 * DESC_WP)      // 559
#define DESC_H_WORDS (43 * DESC_HP)      // 817
#define DESC_WARP_WORDS (DESC_S_WORDS + DESC_H_WORDS + 4)

__device__ const int8_t g_pattern[1024] = {
#include "orb_pattern.inc"
};

__device__ __forceinline__ int reflect101(int p, int len) {
    if (p < 0) p = -p;
    if (p >= len) p = 2 * (len - 1) - p;
    return p;
}

// cv::fastAtan2 (degrees), float32, no FMA contraction.
__device__ __forceinline__ float fast_atan2_deg(float y, float x) {
    const float scale = (float)(180.0 / 3.14159265358979323846);
    const float p1 = 0.9997878412794807f * scale;
    const float p3 = -0.3258083974640975f * scale;
    const float p5 = 0.1555786518463281f * scale;
    const float p7 = -0.04432655554792128f * scale;
    const float eps = 2.2204460492503131e-16f;  // (float)DBL_EPSILON
    const float ax = fabsf(x), ay = fabsf(y);
    float a, c, c2;
    if (ax >= ay) {
        c = __fdiv_rn(ay, __fadd_rn(ax, eps));
        c2 = __fmul_rn(c, c);
        a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
    } else {
        c = __fdiv_rn(ax, __fadd_rn(ay, eps));
        c2 = __fmul_rn(c, c);
        a = __fsub_rn(90.f, __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c));
    }
    if (x < 0) a = __fsub_rn(180.f, a);
    if (y < 0) a = __fsub_rn(360.f, a);
    return a;
}

__global__ void __launch_bounds__(256)
k_assemble(const __grid_constant__ OrbxPlan P, const uint32_t* __restrict__ surv,
           const uint32_t* __restrict__ survcnt, const int* __restrict__ lap,
           uint32_t* __restrict__ dest, int32_t* __restrict__ header) {
    extern __shared__ __align__(16) uint32_t sm[];
    uint32_t* flags = sm;               // [surv_total + 2]
    uint32_t* tmp = sm + P.surv_total + 2;  // [64]
    const int frame = blockIdx.x, tid = threadIdx.x;
    const uint32_t* sc = survcnt + (size_t)frame * ORBX_MAX_LEVELS;
    const uint32_t* sv = surv + (size_t)frame * P.surv_total;
    const float lap0 = (float)lap[2 * frame], lap1 = (float)lap[2 * frame + 1];

    int off[ORBX_MAX_LEVELS + 1];
    off[0] = 0;
    for (int l = 0; l < P.nlevels; ++l) off[l + 1] = off[l] + min((int)sc[l], P.lvl[l].surv_cap);
    const int n = off[P.nlevels];

    for (int idx = tid; idx < n; idx += blockDim.x) {
        int l = 0;
        while (idx >= off[l + 1]) ++l;
        const uint32_t key = sv[P.lvl[l].surv_off + (idx - off[l])];
        float x = (float)(ORBX_KEY_X(key) + ORBX_FAST_BORDER);
        if (l != 0) x = __fmul_rn(x, P.lvl[l].scale);
        flags[idx] = (x >= lap0 && x <= lap1) ? 1u : 0u;
    }
    __syncthreads();
    const uint32_t nstereo = block_excl_scan(flags, n, tmp);
    if (tid == 0) flags[n] = nstereo;
    __syncthreads();
    uint32_t* d = dest + (size_t)frame * P.surv_total;
    for (int idx = tid; idx < n; idx += blockDim.x) {
        const uint32_t sb = flags[idx];
        const bool stereo = flags[idx + 1] != sb;
        d[idx] = stereo ? (uint32_t)(n - 1) - sb : (uint32_t)idx - sb;
    }
    if (tid == 0) {
        int32_t* h = header + 4 * (size_t)frame;
        h[0] = n;
        h[1] = n - (int)nstereo;
        h[2] = n > P.out_cap ? ORBX_E_CAPACITY : ORBX_OK;
        h[3] = 0;
    }
}

__global__ void __launch_bounds__(DESC_WARPS * 32)
k_describe(const __grid_constant__ OrbxPlan P, const OrbxSrc isrc, const uint8_t* __restrict__ pyr,
           const uint32_t* __restrict__ surv, const uint32_t* __restrict__ survcnt,
           const uint32_t* __restrict__ dest, orbx_keypoint* __restrict__ kps,
           uint8_t* __restrict__ desc) {
    __shared__ __align__(16) uint32_t sm[DESC_WARPS * DESC_WARP_WORDS];
    const int warp = orbx_warp(), lane = orbx_lane();
    const int frame = blockIdx.y;
    const int idx = blockIdx.x * DESC_WARPS + warp;

    // which (level, i) is keypoint idx of this frame (level-major order)?
    const uint32_t* sc = survcnt + (size_t)frame * ORBX_MAX_LEVELS;
    int l = 0, base = 0;
    for (;; ++l) {
        if (l == P.nlevels) return;  // idx >= n
        const int nl = min((int)sc[l], P.lvl[l].surv_cap);
        if (idx < base + nl) break;
        base += nl;
    }
    const OrbxLevel& L = P.lvl[l];
    const uint32_t key = surv[(size_t)frame * P.surv_total + L.surv_off + (idx - base)];
    const uint32_t out_row = dest[(size_t)frame * P.surv_total + idx];
    if ((int)out_row >= P.out_cap) return;
    const int cx = ORBX_KEY_X(key) + ORBX_FAST_BORDER, cy = ORBX_KEY_Y(key) + ORBX_FAST_BORDER;

    uint32_t* S = sm + warp * DESC_WARP_WORDS;       // staged source rows, DESC_WP words each
    uint16_t* H = reinterpret_cast<uint16_t*>(S + DESC_S_WORDS);  // 43 x 38 u16
    int lpitch;
    const uint8_t* src = orbx_level_ptr(P, isrc, pyr, frame, l, &lpitch);

    // ---- stage the 43x43 window --------------------------------------------------------------
    const int ox = cx - 21, oy = cy - 21;
    int xo;  // byte offset of window column 0 inside a staged row
    if (ox >= 0 && oy >= 0 && ox + 43 <= L.w && oy + 43 <= L.h) {
        const int ax = ox & ~3;
        xo = ox - ax;
        const int nw = (xo + 43 + 3) >> 2;  // <= 12
        for (int i = lane; i < 43 * 12; i += 32) {
            const int r = i / 12, cw = i - r * 12;
            if (cw < nw)
                S[r * DESC_WP + cw] = __ldg(reinterpret_cast<const uint32_t*>(src + (size_t)(oy + r) * lpitch + ax) + cw);
        }
    } else {
        xo = 0;
        uint8_t* Sb = reinterpret_cast<uint8_t*>(S);
        for (int i = lane; i < 43 * 43; i += 32) {
            const int r = i / 43, c = i - r * 43;
            Sb[r * (DESC_WP * 4) + c] = __ldg(src + (size_t)reflect101(oy + r, L.h) * lpitch + reflect101(ox + c, L.w));
        }
    }
    __syncwarp();
    const uint8_t* Sb = reinterpret_cast<const uint8_t*>(S) + xo;  // Sb[r * 52 + c], c in [0,43)

    // ---- E4: intensity centroid over the radius-15 disc --------------------------------------
    int m10 = 0, m01 = 0;
    if (lane < 31) {
        const int v = lane - 15;
        const int av = v < 0 ? -v : v;
        const int d = (int)((0x3689ABCDDEEEFFFFull >> (4 * av)) & 15ull);  // umax[|v|], :452-467
        const uint8_t* row = Sb + (21 + v) * (DESC_WP * 4) + 21;
        int rs = 0;
        for (int u = -d; u <= d; ++u) {
            const int val = row[u];
            rs += val;
            m10 += u * val;
        }
        m01 = v * rs;
    }
    m10 = warp_sum(m10);
    m01 = warp_sum(m01);
    const float angle = fast_atan2_deg((float)m01, (float)m10);

    // ---- E5 (horizontal): H[r][c] = sum_k g[k] * S[r][c + k], c in [0,37) --------------------
    for (int t = lane; t < 86; t += 32) {
        const int r = t >> 1, c0 = (t & 1) * 19;
        const int nout = (t & 1) ? 18 : 19;
        const uint8_t* row = Sb + r * (DESC_WP * 4) + c0;
        uint16_t* hrow = H + r * (DESC_HP * 2) + c0;
        int w0 = row[0], w1 = row[1], w2 = row[2], w3 = row[3], w4 = row[4], w5 = row[5];
        for (int c = 0; c < nout; ++c) {
            const int w6 = row[c + 6];
            hrow[c] = (uint16_t)(18 * (w0 + w6) + 34 * (w1 + w5) + 48 * (w2 + w4) + 56 * w3);
            w0 = w1; w1 = w2; w2 = w3; w3 = w4; w4 = w5; w5 = w6;
        }
    }
    __syncwarp();

    // ---- E6: steered BRIEF; lane i -> descriptor byte i (points 16i .. 16i+15) ----------------
    const float factorPI = (float)(3.1415926535897932384626433832795 / 180.f);
    const float ang = __fmul_rn(angle, factorPI);
    // correctly rounded float cos/sin via double (glibc's cosf/sinf are within 0.55 ulp;
    // see DESIGN.md on the residual cvRound-boundary mismatches)
    const float a = (float)cos((double)ang), b = (float)sin((double)ang);
    const int8_t* pat = g_pattern + lane * 32;
    uint32_t byte = 0;
#pragma unroll 2
    for (int k = 0; k < 8; ++k) {
        int tv[2];
#pragma unroll
        for (int e = 0; e < 2; ++e) {
            const float px = (float)pat[4 * k + 2 * e], py = (float)pat[4 * k + 2 * e + 1];
            const int dy = __float2int_rn(__fadd_rn(__fmul_rn(px, b), __fmul_rn(py, a)));
            const int dx = __float2int_rn(__fsub_rn(__fmul_rn(px, a), __fmul_rn(py, b)));
            const uint16_t* h = H + (18 + dy) * (DESC_HP * 2) + (18 + dx);
            const uint32_t acc = 18u * (h[0] + h[6 * DESC_HP * 2]) + 34u * (h[DESC_HP * 2] + h[5 * DESC_HP * 2]) +
                                 48u * (h[2 * DESC_HP * 2] + h[4 * DESC_HP * 2]) + 56u * h[3 * DESC_HP * 2];
            tv[e] = (int)((acc + 32768u) >> 16);
        }
        byte |= (tv[0] < tv[1] ? 1u : 0u) << k;
    }
    desc[((size_t)frame * P.out_cap + out_row) * 32 + lane] = (uint8_t)byte;

    if (lane == 0) {
        orbx_keypoint kp;
        float x = (float)cx, y = (float)cy;
        if (l != 0) {
            x = __fmul_rn(x, L.scale);
            y = __fmul_rn(y, L.scale);
        }
        kp.x = x;
        kp.y = y;
        kp.size = L.size;
        kp.angle = angle;
        kp.response = (float)ORBX_KEY_R(key);
        kp.octave = l;
        kp.class_id = -1;
        kps[(size_t)frame * P.out_cap + out_row] = kp;
    }
}

cudaError_t orbx_launch_assemble(const OrbxPlan& P, const uint32_t* surv, const uint32_t* survcnt,
                                 const int* lap, uint32_t* dest, int32_t* header, int nframes,
                                 cudaStream_t st, int* launches) {
    const size_t sm_a = (size_t)(P.surv_total + 2 + 64) * 4;
    k_assemble<<<nframes, 256, sm_a, st>>>(P, surv, survcnt, lap, dest, header);
    ++*launches;
    return cudaGetLastError();
}

cudaError_t orbx_launch_describe(const OrbxPlan& P, const OrbxSrc& src, const uint8_t* pyr,
                                 const uint32_t* surv, const uint32_t* survcnt, const uint32_t* dest,
                                 orbx_keypoint* kps, uint8_t* desc, int nframes, cudaStream_t st,
                                 int* launches) {
    dim3 grid((P.surv_total + DESC_WARPS - 1) / DESC_WARPS, nframes);
    k_describe<<<grid, DESC_WARPS * 32, 0, st>>>(P, src, pyr, surv, survcnt, dest, kps, desc);
    ++*launches;
    return cudaGetLastError();
}
