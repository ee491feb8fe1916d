// k_describe.cu -- E7 layout + E4 orientation + E5 blur + E6 rotated BRIEF, fused.
//
//   k_assemble   one CTA per frame: prefix over the levels' survivors, lapping-area test on the
//                scaled x, stable TWO-ENDED compaction (ORBextractor.cc:1124-1144): a keypoint
//                with lap0 <= pt.x <= lap1 goes to the back (stereoIndex--), others to the front
//                (monoIndex++), both in level-major order.  Writes every survivor's output row
//                and the frame header {n, monoIndex, status}.
//   k_describe   one warp per keypoint.  The 43x43 neighbourhood of the UN-blurred level is
//                staged in shared memory once (4-byte cp.async; reflect-101 of the level at
//                image borders, which is what GaussianBlur sees on the reference's clone(),
//                ORBextractor.cc:1114-1115) and serves three consumers:
//                  - IC_Angle (:75-102): one lane per row of the radius-15 disc, integer
//                    moments, shuffle reduction, cv::fastAtan2's polynomial evaluated with
//                    explicitly non-contracted float ops;
//                  - the horizontal pass of the 7x7 sigma-2 Gaussian in OpenCV 4's 8.8 fixed
//                    point ({18,34,48,56,48,34,18}/256, exact in 16 bits): a lane produces 4
//                    adjacent outputs from 3 staged words with 6 funnel shifts + 8 IDP.4A;
//                  - the vertical pass ((sum + 32768) >> 16) over the 37x37 region the rotated
//                    pattern can reach: H rows are re-paired vertically (PRMT) so that one
//                    IDP.2A applies two taps; 8 IDP.2A give an even and an odd output row.
//                The blurred patch lives only in shared memory (never written to HBM); lane i
//                samples its 16 pattern points from it (cvRound by the 1.5*2^23 magic add) and
//                produces descriptor byte i.
#include "orbx_device.cuh"

#ifndef DESC_WARPS
#define DESC_WARPS 4
#endif
#define DESC_SP 13                       // words per staged row (52 B; odd -> conflict-free rows)
#define DESC_HP 22                       // words per H row (40 u16 + pad; even: 8-byte rows)
#define DESC_BP 10                       // words per blurred row (40 B)
#define DESC_S_WORDS 560                 // 43 * DESC_SP rounded up; reused for 38 * DESC_BP
#define DESC_H_WORDS (44 * DESC_HP)      // 43 rows + one the last odd output row reads
#define DESC_WARP_WORDS (DESC_S_WORDS + DESC_H_WORDS)
#define DESC_SMEM_BYTES ((DESC_WARPS * DESC_WARP_WORDS + 8 * 32 * 4) * 4)

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

__device__ __forceinline__ void cp_async4(uint32_t saddr, const void* g) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(saddr), "l"(g) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

__global__ void __launch_bounds__(DESC_WARPS * 32)
k_describe(const __grid_constant__ OrbxPlan P, const OrbxSrc isrc, const uint8_t* __restrict__ pyr,
           const uint32_t* __restrict__ surv, const uint32_t* __restrict__ survcnt,
           const uint32_t* __restrict__ dest, orbx_keypoint* __restrict__ kps,
           uint8_t* __restrict__ desc) {
    extern __shared__ __align__(16) uint32_t sm[];
    float4* pat4 = reinterpret_cast<float4*>(sm);  // [8][32]: comparison k of lane i
    const int warp = orbx_warp(), lane = orbx_lane();
    {
        // pattern points 16*lane + 2k, 16*lane + 2k + 1 = one aligned word of the int8 table
        for (int e = threadIdx.x; e < 256; e += DESC_WARPS * 32) {
            const int k = e >> 5, i = e & 31;
            const uint32_t w = reinterpret_cast<const uint32_t*>(g_pattern)[8 * i + k];
            pat4[e] = make_float4((float)(int8_t)(w & 0xff), (float)(int8_t)((w >> 8) & 0xff),
                                  (float)(int8_t)((w >> 16) & 0xff), (float)(int8_t)(w >> 24));
        }
    }
    __syncthreads();
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

    uint32_t* S = sm + 8 * 32 * 4 + warp * DESC_WARP_WORDS;  // staged source rows, DESC_SP words each
    uint32_t* Hs = S + DESC_S_WORDS;                         // 44 x 20 packed u16 pairs
    int lpitch;
    const uint8_t* src = orbx_level_ptr(P, isrc, pyr, frame, l, &lpitch);

    // ---- stage the 43x43 window --------------------------------------------------------------
    const int ox = cx - 21, oy = cy - 21;
    int xo;  // byte offset of window column 0 inside a staged row
    if (ox >= 0 && oy >= 0 && ox + 43 <= L.w && oy + 43 <= L.h) {
        const int ax = ox & ~3;
        xo = ox - ax;
        const int nw = (xo + 43 + 3) >> 2;  // <= 12
        const int ro = lane >> 4, cw = lane & 15;
        if (cw < nw) {
            const uint8_t* gp = src + (size_t)(oy + ro) * lpitch + ax + 4 * cw;
            uint32_t sa = (uint32_t)__cvta_generic_to_shared(S + ro * DESC_SP + cw);
#pragma unroll 4
            for (int r = ro; r < 43; r += 2) {
                cp_async4(sa, gp);
                gp += 2 * (size_t)lpitch;
                sa += 2 * DESC_SP * 4;
            }
        }
        cp_async_wait_all();
    } else {
        xo = 0;
        uint8_t* Sb = reinterpret_cast<uint8_t*>(S);
        for (int i = lane; i < 43 * 43; i += 32) {
            const int r = i / 43, c = i - r * 43;
            Sb[r * (DESC_SP * 4) + c] = __ldg(src + (size_t)reflect101(oy + r, L.h) * lpitch + reflect101(ox + c, L.w));
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
        const uint8_t* row = Sb + (21 + v) * (DESC_SP * 4) + 21;
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

    // ---- E5 (horizontal): H[r][b] = sum_k g[k] * S[r][b + k] for staged byte columns b < 40 ----
    if (lane < 30) {
        const uint32_t GA = 0x38302212u, GB = 0x00122230u;  // {18,34,48,56}, {48,34,18,0}
        const int ro = lane / 10, q = lane - ro * 10;
        const uint32_t* sp = S + ro * DESC_SP + q;
        uint2* hp = reinterpret_cast<uint2*>(Hs + ro * DESC_HP) + q;
#pragma unroll 5
        for (int r = ro; r < 43; r += 3) {
            const uint32_t w0 = sp[0], w1 = sp[1], w2 = sp[2];
            const uint32_t o0 = __dp4a(w0, GA, __dp4a(w1, GB, 0u));
            const uint32_t o1 = __dp4a(__funnelshift_r(w0, w1, 8), GA, __dp4a(__funnelshift_r(w1, w2, 8), GB, 0u));
            const uint32_t o2 = __dp4a(__funnelshift_r(w0, w1, 16), GA, __dp4a(__funnelshift_r(w1, w2, 16), GB, 0u));
            const uint32_t o3 = __dp4a(__funnelshift_r(w0, w1, 24), GA, __dp4a(__funnelshift_r(w1, w2, 24), GB, 0u));
            *hp = make_uint2(__byte_perm(o0, o1, 0x5410), __byte_perm(o2, o3, 0x5410));
            sp += 3 * DESC_SP;
            hp += 3 * DESC_HP / 2;
        }
    }
    __syncwarp();

    // ---- E5 (vertical): B[r][b] = (sum_k g[k] * H[r + k][b] + 32768) >> 16, r < 38 -------------
    uint32_t* Bw = S;  // the source window is dead: reuse it for the blurred patch
    if (lane < 30) {
        const int band = lane / 10, g = lane - band * 10;
        const int m0 = 6 * band;  // pair rows m0 .. m0+6 (band 1/2 recompute one shared pair row)
        const uint2* hp = reinterpret_cast<const uint2*>(Hs + 2 * m0 * DESC_HP) + g;
        uint32_t* bp = Bw + 2 * m0 * DESC_BP + g;
        uint32_t Q[4][4];  // Q[j][c] = {H[2(m+j)][c], H[2(m+j)+1][c]} as u16x2
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            const uint2 A = hp[0], Bv = hp[DESC_HP / 2];
            Q[j][0] = __byte_perm(A.x, Bv.x, 0x5410);
            Q[j][1] = __byte_perm(A.x, Bv.x, 0x7632);
            Q[j][2] = __byte_perm(A.y, Bv.y, 0x5410);
            Q[j][3] = __byte_perm(A.y, Bv.y, 0x7632);
            hp += DESC_HP;
        }
#pragma unroll
        for (int s = 0; s < 7; ++s) {
            const uint2 A = hp[0], Bv = hp[DESC_HP / 2];
            Q[3][0] = __byte_perm(A.x, Bv.x, 0x5410);
            Q[3][1] = __byte_perm(A.x, Bv.x, 0x7632);
            Q[3][2] = __byte_perm(A.y, Bv.y, 0x5410);
            Q[3][3] = __byte_perm(A.y, Bv.y, 0x7632);
            hp += DESC_HP;
            uint32_t e[4], o[4];
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                e[c] = __dp2a_lo(Q[3][c], 0x0012u, __dp2a_lo(Q[2][c], 0x2230u, __dp2a_lo(Q[1][c], 0x3830u, __dp2a_lo(Q[0][c], 0x2212u, 32768u))));
                o[c] = __dp2a_lo(Q[3][c], 0x1222u, __dp2a_lo(Q[2][c], 0x3038u, __dp2a_lo(Q[1][c], 0x3022u, __dp2a_lo(Q[0][c], 0x1200u, 32768u))));
            }
            bp[0] = __byte_perm(__byte_perm(e[0], e[1], 0x0062), __byte_perm(e[2], e[3], 0x0062), 0x5410);
            bp[DESC_BP] = __byte_perm(__byte_perm(o[0], o[1], 0x0062), __byte_perm(o[2], o[3], 0x0062), 0x5410);
            bp += 2 * DESC_BP;
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                Q[0][c] = Q[1][c];
                Q[1][c] = Q[2][c];
                Q[2][c] = Q[3][c];
            }
        }
    }
    __syncwarp();

    // ---- E6: steered BRIEF; lane i -> descriptor byte i (points 16i .. 16i+15) ----------------
    const float factorPI = (float)(3.1415926535897932384626433832795 / 180.f);
    const float ang = __fmul_rn(angle, factorPI);
    // correctly rounded float cos/sin via double (glibc's cosf/sinf are within 0.55 ulp;
    // see DESIGN.md on the residual cvRound-boundary mismatches)
    double sd, cd;
    sincos((double)ang, &sd, &cd);
    const float a = (float)cd, b = (float)sd;
    // B[18 + dy][xo + 18 + dx]; cvRound(v) = bits(v + 1.5*2^23) - 0x4B400000 (round half even)
    const float MAGIC = 12582912.f;
    const uint8_t* Bb = reinterpret_cast<const uint8_t*>(Bw);
    const uint32_t C0 = (uint32_t)(18 * DESC_BP * 4 + 18 + xo) - 0x4B400000u * (uint32_t)(DESC_BP * 4 + 1);
    uint32_t byte = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const float4 p = pat4[k * 32 + lane];
        const uint32_t y0 = __float_as_uint(__fadd_rn(__fadd_rn(__fmul_rn(p.x, b), __fmul_rn(p.y, a)), MAGIC));
        const uint32_t x0 = __float_as_uint(__fadd_rn(__fsub_rn(__fmul_rn(p.x, a), __fmul_rn(p.y, b)), MAGIC));
        const uint32_t y1 = __float_as_uint(__fadd_rn(__fadd_rn(__fmul_rn(p.z, b), __fmul_rn(p.w, a)), MAGIC));
        const uint32_t x1 = __float_as_uint(__fadd_rn(__fsub_rn(__fmul_rn(p.z, a), __fmul_rn(p.w, b)), MAGIC));
        const int t0 = Bb[y0 * (uint32_t)(DESC_BP * 4) + x0 + C0];
        const int t1 = Bb[y1 * (uint32_t)(DESC_BP * 4) + x1 + C0];
        byte |= (t0 < t1 ? 1u : 0u) << k;
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
    static bool configured = false;  // per process; the attribute is per function and device-wide
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(k_describe, cudaFuncAttributeMaxDynamicSharedMemorySize, DESC_SMEM_BYTES);
        if (e != cudaSuccess) return e;
        configured = true;
    }
    k_describe<<<grid, DESC_WARPS * 32, DESC_SMEM_BYTES, st>>>(P, src, pyr, surv, survcnt, dest, kps, desc);
    ++*launches;
    return cudaGetLastError();
}
