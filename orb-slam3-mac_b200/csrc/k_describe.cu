// k_describe.cu -- E7 layout + E4 orientation + E5 blur + E6 rotated BRIEF, fused.
//
//   (E7, the two-ended output layout, is orbx_assemble_frame in orbx_device.cuh: it runs in the
//   tail of k_octree, in the last CTA of a frame to finish.)
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
#include <cstring>
#include <mutex>

#include "orbx_device.cuh"

#ifndef DESC_WARPS
#define DESC_WARPS 4
#endif
#define DESC_SP 13                       // words per staged row (52 B; odd -> conflict-free rows)
#define DESC_BP 10                       // words per blurred row (40 B)
#define DESC_S_WORDS (44 * DESC_SP)      // 43 rows + one the last (unused) odd output row reads
#define DESC_B_WORDS (38 * DESC_BP)
#define DESC_WARP_WORDS (DESC_S_WORDS + DESC_B_WORDS)
#define DESC_SMEM_BYTES (DESC_WARPS * DESC_WARP_WORDS * 4)

__device__ const int8_t g_pattern[1024] = {
#include "orb_pattern.inc"
};

// The pattern as floats, lane-major per comparison: g_pat4[k * 32 + i] = points 16i + 2k and
// 16i + 2k + 1 (x0, y0, x1, y1) -- what lane i needs for bit k of descriptor byte i.  Filled once
// per device by k_init_pattern; read through L1 (one coalesced 512-byte request per comparison).
__device__ float4 g_pat4[256];

__global__ void k_init_pattern() {
    const int e = threadIdx.x, k = e >> 5, i = e & 31;
    const uint32_t w = reinterpret_cast<const uint32_t*>(g_pattern)[8 * i + k];
    g_pat4[e] = make_float4((float)(int8_t)(w & 0xff), (float)(int8_t)((w >> 8) & 0xff),
                            (float)(int8_t)((w >> 16) & 0xff), (float)(int8_t)(w >> 24));
}

// PTX shifts clamp the amount: anything >= 32 yields 0 (C++ << / >> would be undefined)
__device__ __forceinline__ uint32_t desc_shl(uint32_t x, uint32_t n) {
    uint32_t r;
    asm("shl.b32 %0, %1, %2;" : "=r"(r) : "r"(x), "r"(n));
    return r;
}
__device__ __forceinline__ uint32_t desc_shr(uint32_t x, uint32_t n) {
    uint32_t r;
    asm("shr.u32 %0, %1, %2;" : "=r"(r) : "r"(x), "r"(n));
    return r;
}

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

// cos / sin of a float angle in [0, 2 pi] to double accuracy, rounded once to float -- what the reference's
// (float)cos(angle) needs -- without the general sincos(double): the quadrant comes from one multiply, the reduced
// argument r = x - q * pi/2 from a two-term pi/2 (q <= 4, so q * PIO2_HI is exact), and sin r / cos r on
// |r| <= pi/4 from the classical degree-13 / degree-14 minimax polynomials (fdlibm's __kernel_sin / __kernel_cos
// coefficients, error < 2^-57).  ~30 FP64 instructions instead of ~75.  tests/test_gpu_extractor.py checks it against
// sincos() on EVERY float in [0, 6.2832] (orbx_debug_sincos_check): identical float results.  NOT the default: measured
// slower than sincos() inside k_describe (see DESC_POLY_SINCOS below); kept as the evidence of the attempt.
__device__ __forceinline__ void desc_sincos(float ang, float* cs, float* sn) {
    const double x = (double)ang;
    const double PIO2_HI = 1.57079632679489655800e+00, PIO2_LO = 6.12323399573676603587e-17;
    const double S1 = -1.66666666666666324348e-01, S2 = 8.33333333332248946124e-03, S3 = -1.98412698298579493134e-04,
                 S4 = 2.75573137070700676789e-06, S5 = -2.50507602534068634195e-08, S6 = 1.58969099521155010221e-10;
    const double C1 = 4.16666666666666019037e-02, C2 = -1.38888888888741095749e-03, C3 = 2.48015872894767294178e-05,
                 C4 = -2.75573143513906633035e-07, C5 = 2.08757232129817482790e-09, C6 = -1.13596475577881948265e-11;
    const int q = __double2int_rn(x * 6.36619772367581382433e-01);   // x * 2 / pi
    const double qd = (double)q;
    const double r = fma(-qd, PIO2_LO, fma(-qd, PIO2_HI, x));
    const double z = r * r;
    const double ps = fma(fma(fma(fma(fma(S6, z, S5), z, S4), z, S3), z, S2), z, S1);
    const double s = fma(r * z, ps, r);
    const double pc = fma(fma(fma(fma(fma(C6, z, C5), z, C4), z, C3), z, C2), z, C1);
    const double c = fma(z * z, pc, fma(-0.5, z, 1.0));
    // quadrant: (cos, sin)(x) = (c, s), (-s, c), (-c, -s), (s, -c)
    const double cq = (q & 1) ? s : c, sq = (q & 1) ? c : s;
    *cs = (float)(((q + 1) & 2) ? -cq : cq);
    *sn = (float)((q & 2) ? -sq : sq);
}

__device__ __forceinline__ void cp_async4(uint32_t saddr, const void* g) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(saddr), "l"(g) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

#ifndef DESC_MINB
#define DESC_MINB 12   // 40 registers (40 B of spill): 12 CTAs = 48 warps per SM; same-box A/B 0.618 (MINB 1, 63 regs) -> 0.600 ms
#endif
__global__ void __launch_bounds__(DESC_WARPS * 32, DESC_MINB)
k_describe(const __grid_constant__ OrbxPlan P, const OrbxSrc isrc, const uint8_t* __restrict__ pyr,
           const int32_t* __restrict__ header, const uint32_t* __restrict__ dest,
           orbx_keypoint* __restrict__ kps, uint8_t* __restrict__ desc, const OrbxSinkArgs sink) {
    extern __shared__ __align__(16) uint32_t sm[];
    const int warp = orbx_warp(), lane = orbx_lane();
    const int frame = blockIdx.y;
    const int idx = blockIdx.x * DESC_WARPS + warp;

    // fused result gather: the frame header also goes to the sink (rank 0's slab, peer-mapped over NVLink)
    if (sink.hdr && blockIdx.x == 0 && threadIdx.x < 4)
        sink.hdr[4 * (size_t)(sink.first_frame + frame) + threadIdx.x] = header[4 * frame + threadIdx.x];
    // the keypoint's record, written by orbx_assemble_frame: {key, output row | level << 24}
    if (idx >= min(header[4 * frame], P.surv_total)) return;
    const uint2 rec = __ldg(reinterpret_cast<const uint2*>(dest) + (size_t)frame * P.surv_total + idx);
    const uint32_t key = rec.x;
    const uint32_t out_row = rec.y & 0xffffffu;
    const int l = (int)(rec.y >> 24);
    if ((int)out_row >= P.out_cap) return;
    const OrbxLevel& L = P.lvl[l];
    const int cx = ORBX_KEY_X(key) + ORBX_FAST_BORDER, cy = ORBX_KEY_Y(key) + ORBX_FAST_BORDER;

    uint32_t* S = sm + warp * DESC_WARP_WORDS;  // staged source rows, DESC_SP words each
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

    // ---- E4: intensity centroid over the radius-15 disc --------------------------------------
    // lane = row v of the disc; the row's bytes [21-d, 21+d] are read as words (rows are 13 words
    // apart: conflict-free) and reduced with IDP.4A against 1 and against the biased column
    // weights u + 32 (>= 0 in every byte, so the packed weights advance by 0x04040404 per word).
    int m10 = 0, m01 = 0;
    if (lane < 31) {
        const int v = lane - 15;
        const int av = v < 0 ? -v : v;
        const int d = (int)((0x3689ABCDDEEEFFFFull >> (4 * av)) & 15ull);  // umax[|v|], :452-467
        const int lo8 = 8 * (xo + 21 - d), hi8 = 8 * (xo + 21 + d);   // staged byte columns [lo, hi], in bits
        const uint32_t* row = S + (21 + v) * DESC_SP;
        // weight of staged byte b is b - xo - 21 + 32; the disc spans words 1..9 of a staged row
        uint32_t uw = (uint32_t)(11 - xo) * 0x01010101u + 0x03020100u + 0x04040404u;
        uint32_t rs = 0, mu = 0;
#pragma unroll
        for (int j = 1; j <= 9; ++j, uw += 0x04040404u) {
            // keep the bytes of word j that lie in [lo, hi]: PTX shifts clamp (>= 32 gives 0)
            const uint32_t w = row[j] & desc_shl(0xffffffffu, (uint32_t)max(lo8 - 32 * j, 0)) &
                               desc_shr(0xffffffffu, (uint32_t)max(32 * j + 24 - hi8, 0));
            rs = __dp4a(w, 0x01010101u, rs);
            mu = __dp4a(w, uw, mu);
        }
        m10 = (int)mu - 32 * (int)rs;
        m01 = v * (int)rs;
    }
    m10 = warp_sum(m10);
    m01 = warp_sum(m01);
    const float angle = fast_atan2_deg((float)m01, (float)m10);

    // ---- E5: 7x7 sigma-2 Gaussian, horizontal and vertical pass fused in registers -----------
    // lane = (band, column group): 4 adjacent staged byte columns b = 4g .. 4g+3, 14 output rows.
    // Walking down the source rows, H[r][b] = sum_k g[k] * S[r][b + k] comes from 3 staged words
    // (6 funnel shifts + 8 IDP.4A); two consecutive rows are paired per column as u16x2 (PRMT) so
    // that one IDP.2A applies two vertical taps: 8 IDP.2A give an even and an odd output row of
    // B[r][b] = (sum_k g[k] * H[r + k][b] + 32768) >> 16.  Only B goes to shared memory.
    uint32_t* Bw = S + DESC_S_WORDS;
    if (lane < 30) {
        const uint32_t GA = 0x38302212u, GB = 0x00122230u;  // {18,34,48,56}, {48,34,18,0}
        const int band = lane / 10, g = lane - band * 10;
        const uint32_t* sp = S + (12 * band) * DESC_SP + g;   // source rows 12*band ..
        uint32_t* bp = Bw + (12 * band) * DESC_BP + g;        // output rows 12*band ..
        uint32_t Q[4][4];  // Q[j][c] = {H[2(m+j)][c], H[2(m+j)+1][c]} as u16x2
#define DESC_HROW(h)                                                                              \
        {                                                                                         \
            const uint32_t w0 = sp[0], w1 = sp[1], w2 = sp[2];                                    \
            sp += DESC_SP;                                                                        \
            h[0] = __dp4a(w0, GA, __dp4a(w1, GB, 0u));                                            \
            h[1] = __dp4a(__funnelshift_r(w0, w1, 8), GA, __dp4a(__funnelshift_r(w1, w2, 8), GB, 0u));   \
            h[2] = __dp4a(__funnelshift_r(w0, w1, 16), GA, __dp4a(__funnelshift_r(w1, w2, 16), GB, 0u)); \
            h[3] = __dp4a(__funnelshift_r(w0, w1, 24), GA, __dp4a(__funnelshift_r(w1, w2, 24), GB, 0u)); \
        }
#define DESC_PAIR(q)                                                                              \
        {                                                                                         \
            uint32_t he[4], ho[4];                                                                \
            DESC_HROW(he)                                                                         \
            DESC_HROW(ho)                                                                         \
            _Pragma("unroll") for (int c = 0; c < 4; ++c) q[c] = __byte_perm(he[c], ho[c], 0x5410); \
        }
        DESC_PAIR(Q[0])
        DESC_PAIR(Q[1])
        DESC_PAIR(Q[2])
#pragma unroll
        for (int st = 0; st < 7; ++st) {
            DESC_PAIR(Q[3])
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
#undef DESC_PAIR
#undef DESC_HROW
    }
    __syncwarp();

    // ---- E6: steered BRIEF; lane i -> descriptor byte i (points 16i .. 16i+15) ----------------
    const float factorPI = (float)(3.1415926535897932384626433832795 / 180.f);
    const float ang = __fmul_rn(angle, factorPI);
    // correctly rounded float cos/sin via double (glibc's cosf/sinf are within 0.55 ulp;
    // see DESIGN.md on the residual cvRound-boundary mismatches)
    float a, b;
#ifdef DESC_POLY_SINCOS   // A/B (r2): bit-identical on every angle, but SLOWER -- describe 0.662 vs 0.617 ms per 384 frames:
    desc_sincos(ang, &a, &b);   // 7 % fewer instructions, yet a 20-deep dependent FP64 FMA chain per warp
#else
    {
        double sd, cd;
        sincos((double)ang, &sd, &cd);
        a = (float)cd;
        b = (float)sd;
    }
#endif
    // B[18 + dy][xo + 18 + dx]; cvRound(v) = bits(v + 1.5*2^23) - 0x4B400000 (round half even)
    const float MAGIC = 12582912.f;
    const uint8_t* Bb = reinterpret_cast<const uint8_t*>(Bw);
    const uint32_t C0 = (uint32_t)(18 * DESC_BP * 4 + 18 + xo) - 0x4B400000u * (uint32_t)(DESC_BP * 4 + 1);
    uint32_t byte = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const float4 p = __ldg(&g_pat4[k * 32 + lane]);
        const uint32_t y0 = __float_as_uint(__fadd_rn(__fadd_rn(__fmul_rn(p.x, b), __fmul_rn(p.y, a)), MAGIC));
        const uint32_t x0 = __float_as_uint(__fadd_rn(__fsub_rn(__fmul_rn(p.x, a), __fmul_rn(p.y, b)), MAGIC));
        const uint32_t y1 = __float_as_uint(__fadd_rn(__fadd_rn(__fmul_rn(p.z, b), __fmul_rn(p.w, a)), MAGIC));
        const uint32_t x1 = __float_as_uint(__fadd_rn(__fsub_rn(__fmul_rn(p.z, a), __fmul_rn(p.w, b)), MAGIC));
        const int t0 = Bb[y0 * (uint32_t)(DESC_BP * 4) + x0 + C0];
        const int t1 = Bb[y1 * (uint32_t)(DESC_BP * 4) + x1 + C0];
        byte |= (t0 < t1 ? 1u : 0u) << k;
    }
    desc[((size_t)frame * P.out_cap + out_row) * 32 + lane] = (uint8_t)byte;
    // fused result gather: the same row straight into the gathering GPU's slab (one 32-byte write per
    // keypoint over NVLink when the sink is peer memory) -- no collective follows the kernel
    if (sink.desc) sink.desc[((size_t)(sink.first_frame + frame) * P.out_cap + out_row) * 32 + lane] = (uint8_t)byte;

    // the 28-byte keypoint record as ONE coalesced store: lane i < 7 writes word i (one 28-byte request
    // instead of seven 4-byte ones -- it also travels over NVLink when a sink is set)
    float x = (float)cx, y = (float)cy;
    if (l != 0) {
        x = __fmul_rn(x, L.scale);
        y = __fmul_rn(y, L.scale);
    }
    uint32_t word = __float_as_uint(x);                       // pt.x
    word = lane == 1 ? __float_as_uint(y) : word;             // pt.y
    word = lane == 2 ? __float_as_uint(L.size) : word;        // size
    word = lane == 3 ? __float_as_uint(angle) : word;         // angle
    word = lane == 4 ? __float_as_uint((float)ORBX_KEY_R(key)) : word;   // response
    word = lane == 5 ? (uint32_t)l : word;                    // octave
    word = lane == 6 ? 0xffffffffu : word;                    // class_id = -1
    if (lane < 7) {
        static_assert(sizeof(orbx_keypoint) == 28, "cv::KeyPoint layout");
        reinterpret_cast<uint32_t*>(kps + (size_t)frame * P.out_cap + out_row)[lane] = word;
        if (sink.kps)
            reinterpret_cast<uint32_t*>(sink.kps + (size_t)(sink.first_frame + frame) * P.out_cap + out_row)[lane] = word;
    }
}

// ---- fused gather: per-batch completion word and back-pressure ----------------------------------
// A consumer on the gathering GPU cannot use the producers' streams; it needs to know when batch e of producer r is
// complete in its slab, and a producer must not overwrite a region the consumer is still reading.  Both are two words
// per region in the sink: flag (written by the producer after the batch, read by the consumer) and ack (the reverse).
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
// wait until *word - need >= 0 (wrap-safe); returns false after timeout_ns
__device__ __forceinline__ bool spin_until(const uint32_t* word, uint32_t need, unsigned long long timeout_ns) {
    const unsigned long long t0 = global_ns();
    while ((int32_t)(ld_acquire_sys(word) - need) < 0) {
        __nanosleep(200);
        if (global_ns() - t0 > timeout_ns) return false;
    }
    return true;
}
__global__ void k_sink_wait_ack(const uint32_t* ack, const uint32_t* epoch_dev, uint32_t* err) {
    // everything this slot has produced so far (epoch_dev batches) must have been released by the consumer
    if (!spin_until(ack, *epoch_dev, 2000000000ull)) *err = 1;
}
__global__ void k_sink_signal(uint32_t* flag, uint32_t* epoch_dev) {
    const uint32_t e = *epoch_dev + 1;
    *epoch_dev = e;
    __threadfence_system();   // the batch's rows (written by the kernel before this one on the stream) before the flag
    st_release_sys(flag, e);
}
__global__ void k_sink_wait_flag(const uint32_t* flag, uint32_t epoch, unsigned long long timeout_ns, uint32_t* err) {
    if (!spin_until(flag, epoch, timeout_ns)) *err = 1;
}
__global__ void k_sink_release(uint32_t* ack, uint32_t epoch) {
    __threadfence_system();
    st_release_sys(ack, epoch);
}
// With CUDA's lazy module loading the FIRST launch of a kernel loads its code, and that load serialises behind kernels
// already running on the device -- e.g. behind a producer parked in k_sink_wait_ack (measured: the consumer's first
// k_sink_release waited 1.85 s, until the parked kernel timed out).  So all four are loaded when a sink is attached,
// before anything can park.
cudaError_t orbx_sink_preload() {
    cudaFuncAttributes a;
    cudaError_t e;
    if ((e = cudaFuncGetAttributes(&a, k_sink_wait_ack)) != cudaSuccess) return e;
    if ((e = cudaFuncGetAttributes(&a, k_sink_signal)) != cudaSuccess) return e;
    if ((e = cudaFuncGetAttributes(&a, k_sink_wait_flag)) != cudaSuccess) return e;
    return cudaFuncGetAttributes(&a, k_sink_release);
}

cudaError_t orbx_launch_sink_wait_flag(const uint32_t* flag, uint32_t epoch, unsigned long long timeout_ns, uint32_t* err,
                                       cudaStream_t st) {
    k_sink_wait_flag<<<1, 1, 0, st>>>(flag, epoch, timeout_ns, err);
    return cudaGetLastError();
}
cudaError_t orbx_launch_sink_release(uint32_t* ack, uint32_t epoch, cudaStream_t st) {
    k_sink_release<<<1, 1, 0, st>>>(ack, epoch);
    return cudaGetLastError();
}

// Exhaustive check of desc_sincos against sincos(): every float bit pattern in [first, last]; counts the values whose
// float cos or sin differs and keeps the first one.
__global__ void k_sincos_check(uint32_t first, uint32_t last, unsigned long long* bad, uint32_t* first_bad) {
    const uint64_t n = (uint64_t)last - first + 1;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        const float x = __uint_as_float(first + (uint32_t)i);
        float c, s;
        desc_sincos(x, &c, &s);
        double sd, cd;
        sincos((double)x, &sd, &cd);
        if (__float_as_uint(c) != __float_as_uint((float)cd) || __float_as_uint(s) != __float_as_uint((float)sd)) {
            atomicAdd(bad, 1ull);
            atomicMin(first_bad, first + (uint32_t)i);
        }
    }
}
cudaError_t orbx_run_sincos_check(float lo, float hi, unsigned long long* n_bad, float* first_bad) {
    unsigned long long* d_bad;
    uint32_t* d_first;
    cudaError_t e;
    if ((e = cudaMalloc((void**)&d_bad, 16)) != cudaSuccess) return e;
    d_first = (uint32_t*)(d_bad + 1);
    const unsigned long long zero = 0;
    const uint32_t ones = 0xffffffffu;
    cudaMemcpy(d_bad, &zero, 8, cudaMemcpyHostToDevice);
    cudaMemcpy(d_first, &ones, 4, cudaMemcpyHostToDevice);
    uint32_t a, b;
    memcpy(&a, &lo, 4);
    memcpy(&b, &hi, 4);
    k_sincos_check<<<148 * 16, 256>>>(a, b, d_bad, d_first);
    e = cudaDeviceSynchronize();
    uint32_t fb = 0;
    cudaMemcpy(n_bad, d_bad, 8, cudaMemcpyDeviceToHost);
    cudaMemcpy(&fb, d_first, 4, cudaMemcpyDeviceToHost);
    memcpy(first_bad, &fb, 4);
    cudaFree(d_bad);
    return e;
}

cudaError_t orbx_launch_describe(const OrbxPlan& P, const OrbxSrc& src, const uint8_t* pyr,
                                 const int32_t* header, const uint32_t* dest,
                                 orbx_keypoint* kps, uint8_t* desc, int nframes, cudaStream_t st,
                                 int* launches, const OrbxSinkArgs& sink) {
    dim3 grid((P.surv_total + DESC_WARPS - 1) / DESC_WARPS, nframes);
    // once per device: shared-memory opt-in and the float pattern table
    static std::mutex mu;
    static bool ready[64] = {false};
    {
        int dev = 0;
        cudaGetDevice(&dev);
        std::lock_guard<std::mutex> lk(mu);
        if (dev >= 0 && dev < 64 && !ready[dev]) {
            cudaError_t e = cudaFuncSetAttribute(k_describe, cudaFuncAttributeMaxDynamicSharedMemorySize, DESC_SMEM_BYTES);
            if (e != cudaSuccess) return e;
            k_init_pattern<<<1, 256, 0, st>>>();
            if ((e = cudaStreamSynchronize(st)) != cudaSuccess) return e;
            ready[dev] = true;
        }
    }
    if (sink.hdr && sink.ack) {
        k_sink_wait_ack<<<1, 1, 0, st>>>(sink.ack, sink.epoch_dev, sink.err_dev);
        ++*launches;
    }
    k_describe<<<grid, DESC_WARPS * 32, DESC_SMEM_BYTES, st>>>(P, src, pyr, header, dest, kps, desc, sink);
    ++*launches;
    if (sink.hdr && sink.flag) {
        k_sink_signal<<<1, 1, 0, st>>>(sink.flag, sink.epoch_dev);
        ++*launches;
    }
    return cudaGetLastError();
}
