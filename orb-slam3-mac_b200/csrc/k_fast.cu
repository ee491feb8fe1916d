// k_fast.cu -- E2: per-cell FAST-9/16 with the iniThFAST -> minThFAST fallback
// (ORBextractor::ComputeKeyPointsOctTree, src/ORBextractor.cc:763-854; cv::FAST semantics as
// pinned in oracle/cvprims.h against cv2 4.13).
//
// One warp owns one 30-px cell of one level of one frame (all levels in a single launch):
//   1. the (wCell+6) x (hCell+6) tile is staged in shared memory with aligned 32-bit loads;
//   2. every detection pixel takes a 4-point rejection test; survivors are compacted through a
//      64-entry shared-memory queue (ballot/popc) so that the exact score is always evaluated
//      by 32 busy lanes;
//   3. the exact score = max over the 16 arcs of 9 of min(ring - c) / min(c - ring), minus 1
//      (cornerScore<16>), is computed for bright and dark arcs AT ONCE on packed s16x2 values
//      (one IMAD builds {ring-c+256, c-ring+256}; __vimin3_s16x2 / __vimax3_s16x2 do the rest);
//   4. 3x3 non-max suppression inside the cell only (scores outside the cell's detection region
//      are 0 -- NMS never crosses a cell), strict '>' on all 8 neighbours;
//   5. survivors are emitted in raster order (ballot/popc compaction) into the cell's
//      fixed-capacity slot; if none survived, steps 2-5 repeat with minThFAST.
// Order of candidates = cells row-major, raster inside a cell, which is what the quadtree
// needs to reproduce the reference's first-maximum-wins choice.
#include "orbx_device.cuh"

#define FAST_WARPS 8

struct FastWarpSmem {
    uint8_t* tile;   // tile_pitch * tile_h
    uint8_t* score;  // score_pitch * (dh + 2), 1-px zero frame around the detection region
    uint16_t* queue; // 64 entries
};

__device__ __forceinline__ int fast_exact_score(const uint8_t* t, int tp) {
    // t points at the centre pixel inside the tile (pitch tp).  Returns raw - 1.
    const int c = t[0];
    const uint32_t K = (uint32_t)(256 - c) + ((uint32_t)(c + 256) << 16);
    uint32_t x[16];
    x[0] = t[3 * tp] * 0xFFFF0001u + K;
    x[1] = t[3 * tp + 1] * 0xFFFF0001u + K;
    x[2] = t[2 * tp + 2] * 0xFFFF0001u + K;
    x[3] = t[tp + 3] * 0xFFFF0001u + K;
    x[4] = t[3] * 0xFFFF0001u + K;
    x[5] = t[-tp + 3] * 0xFFFF0001u + K;
    x[6] = t[-2 * tp + 2] * 0xFFFF0001u + K;
    x[7] = t[-3 * tp + 1] * 0xFFFF0001u + K;
    x[8] = t[-3 * tp] * 0xFFFF0001u + K;
    x[9] = t[-3 * tp - 1] * 0xFFFF0001u + K;
    x[10] = t[-2 * tp - 2] * 0xFFFF0001u + K;
    x[11] = t[-tp - 3] * 0xFFFF0001u + K;
    x[12] = t[-3] * 0xFFFF0001u + K;
    x[13] = t[tp - 3] * 0xFFFF0001u + K;
    x[14] = t[2 * tp - 2] * 0xFFFF0001u + K;
    x[15] = t[3 * tp - 1] * 0xFFFF0001u + K;
    uint32_t m3[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) m3[i] = __vimin3_s16x2(x[i], x[(i + 1) & 15], x[(i + 2) & 15]);
    uint32_t best = 0;  // halves are >= 1
#pragma unroll
    for (int i = 0; i < 16; i += 2) {
        uint32_t a = __vimin3_s16x2(m3[i], m3[(i + 3) & 15], m3[(i + 6) & 15]);
        uint32_t b = __vimin3_s16x2(m3[i + 1], m3[(i + 4) & 15], m3[(i + 7) & 15]);
        best = __vimax3_s16x2(best, a, b);
    }
    const int bright = (int)(best & 0xffffu) - 256;
    const int dark = (int)(best >> 16) - 256;
    return max(bright, dark) - 1;
}

// Computes the score map for threshold `th` (s = raw-1 stored where raw > th, else 0 kept).
__device__ __forceinline__ void fast_scores(const FastWarpSmem& s, int tp, int sp, int dw, int dh,
                                            uint32_t dw_inv, int th) {
    const int lane = orbx_lane();
    const int npix = dw * dh;
    int qhead = 0, qn = 0;
    for (int base = 0; base < npix; base += 32) {
        const int p = base + lane;
        bool pass = false;
        if (p < npix) {
            const int y = (int)(((uint32_t)p * dw_inv) >> 20);
            const int x = p - y * dw;
            const uint8_t* t = s.tile + (y + 3) * tp + (x + 3);
            const int c = t[0];
            const int r0 = t[3 * tp], r8 = t[-3 * tp], r4 = t[3], r12 = t[-3];
            const int hi = c + th, lo = c - th;
            const bool bright = (max(r0, r8) > hi) && (max(r4, r12) > hi);
            const bool dark = (min(r0, r8) < lo) && (min(r4, r12) < lo);
            pass = bright || dark;
        }
        const uint32_t m = __ballot_sync(ORBX_FULL, pass);
        if (pass) s.queue[(qhead + qn + __popc(m & orbx_lanemask_lt())) & 63] = (uint16_t)p;
        qn += __popc(m);
        __syncwarp();
        if (qn >= 32 || (base + 32 >= npix && qn > 0)) {
            // drain up to 32 queued pixels with all lanes busy
            const int take = min(qn, 32);
            if (lane < take) {
                const int q = s.queue[(qhead + lane) & 63];
                const int y = (int)(((uint32_t)q * dw_inv) >> 20);
                const int x = q - y * dw;
                const int sc = fast_exact_score(s.tile + (y + 3) * tp + (x + 3), tp);
                if (sc >= th) s.score[(y + 1) * sp + (x + 1)] = (uint8_t)sc;
            }
            qhead = (qhead + take) & 63;
            qn -= take;
            __syncwarp();
        }
    }
    // the loop's last iteration drains at most 32; a second drain covers 32 < qn < 64
    if (qn > 0) {
        if (lane < qn) {
            const int q = s.queue[(qhead + lane) & 63];
            const int y = (int)(((uint32_t)q * dw_inv) >> 20);
            const int x = q - y * dw;
            const int sc = fast_exact_score(s.tile + (y + 3) * tp + (x + 3), tp);
            if (sc >= th) s.score[(y + 1) * sp + (x + 1)] = (uint8_t)sc;
        }
    }
    __syncwarp();
}

// NMS + ordered emission.  Returns the number of keypoints written.
__device__ __forceinline__ int fast_nms_emit(const FastWarpSmem& s, int sp, int dw, int dh,
                                             uint32_t dw_inv, int xadd, int yadd,
                                             uint32_t* __restrict__ slot) {
    const int lane = orbx_lane();
    const int npix = dw * dh;
    int count = 0;
    for (int base = 0; base < npix; base += 32) {
        const int p = base + lane;
        bool keep = false;
        int x = 0, y = 0, sc = 0;
        if (p < npix) {
            y = (int)(((uint32_t)p * dw_inv) >> 20);
            x = p - y * dw;
            const uint8_t* c = s.score + (y + 1) * sp + (x + 1);
            sc = c[0];
            if (sc > 0) {
                keep = sc > c[-1] && sc > c[1] && sc > c[-sp - 1] && sc > c[-sp] && sc > c[-sp + 1] &&
                       sc > c[sp - 1] && sc > c[sp] && sc > c[sp + 1];
            }
        }
        const uint32_t m = __ballot_sync(ORBX_FULL, keep);
        if (keep) slot[count + __popc(m & orbx_lanemask_lt())] = ORBX_MAKE_KEY(x + xadd, y + yadd, sc);
        count += __popc(m);
    }
    return count;
}

__global__ void __launch_bounds__(FAST_WARPS * 32)
k_fast_cells(const __grid_constant__ OrbxPlan P, const OrbxSrc isrc, const uint8_t* __restrict__ pyr,
             uint32_t* __restrict__ cellcnt, uint32_t* __restrict__ cellkeys, int tile_pitch,
             int score_pitch, int warp_smem_bytes) {
    extern __shared__ __align__(16) uint8_t smem[];
    const int warp = orbx_warp(), lane = orbx_lane();
    const int g = blockIdx.x * FAST_WARPS + warp;  // global cell index inside the frame
    if (g >= P.cells_total) return;
    const int frame = blockIdx.y;

    int l = 0;
    while (l + 1 < P.nlevels && g >= P.lvl[l + 1].cell_off) ++l;
    const OrbxLevel& L = P.lvl[l];
    const int c = g - L.cell_off;
    const int ci = c / L.n_cols, cj = c - ci * L.n_cols;

    uint32_t* cnt_out = cellcnt + (size_t)frame * P.cells_total + g;
    uint32_t* slot = cellkeys + (size_t)frame * P.keys_total + L.key_off + (size_t)c * L.cell_cap;

    // ORBextractor.cc:787-804
    const int iniY = ORBX_FAST_BORDER + ci * L.h_cell;
    const int iniX = ORBX_FAST_BORDER + cj * L.w_cell;
    const int maxY = min(iniY + L.h_cell + 6, L.max_by);
    const int maxX = min(iniX + L.w_cell + 6, L.max_bx);
    const int tw = maxX - iniX, th_ = maxY - iniY;
    if (iniY >= L.max_by - 3 || iniX >= L.max_bx - 6 || tw < 7 || th_ < 7) {
        if (lane == 0) *cnt_out = 0;
        return;
    }
    const int dw = tw - 6, dh = th_ - 6;  // detection region = tile pixels [3, t-4]

    FastWarpSmem s;
    uint8_t* base = smem + (size_t)warp * warp_smem_bytes;
    s.queue = reinterpret_cast<uint16_t*>(base);
    s.tile = base + 128;
    s.score = s.tile + tile_pitch * P.max_tile_h;

    // 1. stage the tile: aligned words covering [iniX & ~3, maxX)
    int lpitch;
    const uint8_t* src = orbx_level_ptr(P, isrc, pyr, frame, l, &lpitch);
    const int ax = iniX & ~3;
    const int xo = iniX - ax;  // tile x -> smem byte x + xo
    const int nw = (maxX - ax + 3) >> 2;
    const int nwords = nw * th_;
    const uint32_t nw_inv = (1u << 20) / (uint32_t)nw + 1;
    for (int i = lane; i < nwords; i += 32) {
        const int r = (int)(((uint32_t)i * nw_inv) >> 20);
        const int cw = i - r * nw;
        const uint32_t v = __ldg(reinterpret_cast<const uint32_t*>(src + (size_t)(iniY + r) * lpitch + ax) + cw);
        reinterpret_cast<uint32_t*>(s.tile + r * tile_pitch)[cw] = v;
    }
    // zero the score map (frame included)
    const int swords = (score_pitch * (dh + 2)) >> 2;
    for (int i = lane; i < swords; i += 32) reinterpret_cast<uint32_t*>(s.score)[i] = 0;
    __syncwarp();

    FastWarpSmem t = s;
    t.tile = s.tile + xo;
    const uint32_t dw_inv = (1u << 20) / (uint32_t)dw + 1;
    const int xadd = 3 + cj * L.w_cell, yadd = 3 + ci * L.h_cell;

    fast_scores(t, tile_pitch, score_pitch, dw, dh, dw_inv, P.ini_th);
    int n = fast_nms_emit(t, score_pitch, dw, dh, dw_inv, xadd, yadd, slot);
    if (n == 0) {
        // ORBextractor.cc:824-828: rerun the cell with minThFAST.  Scores already stored are
        // threshold independent, so the map is only extended.
        fast_scores(t, tile_pitch, score_pitch, dw, dh, dw_inv, P.min_th);
        n = fast_nms_emit(t, score_pitch, dw, dh, dw_inv, xadd, yadd, slot);
    }
    if (lane == 0) *cnt_out = (uint32_t)n;
}

void orbx_fast_smem_layout(const OrbxPlan& P, int* tile_pitch, int* score_pitch, int* warp_bytes) {
    // tile rows hold up to max_tile_w + 3 (alignment slack) bytes, rounded to words
    *tile_pitch = ((P.max_tile_w + 3 + 3) / 4) * 4;
    *score_pitch = ((P.max_tile_w - 6 + 2 + 3) / 4) * 4;
    int bytes = 128 + *tile_pitch * P.max_tile_h + *score_pitch * (P.max_tile_h - 6 + 2);
    *warp_bytes = (bytes + 15) & ~15;
}

// Called once per (device, plan): opt in to the dynamic shared memory the plan needs.
cudaError_t orbx_configure_fast(const OrbxPlan& P) {
    int tp, sp, wb;
    orbx_fast_smem_layout(P, &tp, &sp, &wb);
    return cudaFuncSetAttribute(k_fast_cells, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                wb * FAST_WARPS);
}

cudaError_t orbx_launch_fast(const OrbxPlan& P, const OrbxSrc& src, const uint8_t* pyr, uint32_t* cellcnt,
                             uint32_t* cellkeys, int nframes, cudaStream_t st, int* launches) {
    int tp, sp, wb;
    orbx_fast_smem_layout(P, &tp, &sp, &wb);
    const int smem = wb * FAST_WARPS;
    dim3 grid((P.cells_total + FAST_WARPS - 1) / FAST_WARPS, nframes);
    k_fast_cells<<<grid, FAST_WARPS * 32, smem, st>>>(P, src, pyr, cellcnt, cellkeys, tp, sp, wb);
    ++*launches;
    return cudaGetLastError();
}
