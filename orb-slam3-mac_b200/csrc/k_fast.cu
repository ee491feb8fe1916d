// k_fast.cu -- E2: per-cell FAST-9/16 with the iniThFAST -> minThFAST fallback
// (ORBextractor::ComputeKeyPointsOctTree, src/ORBextractor.cc:763-854; cv::FAST semantics as
// pinned in oracle/cvprims.h against cv2 4.13).
//
// One warp owns one 30-px cell of one level of one frame (all levels in a single launch); the
// cell's geometry comes from a per-shape record table (OrbxPlan::cell_tab, one 16-byte load):
//   1. the (wCell+6) x (hCell+6) tile is staged in shared memory with 16-byte cp.async chunks
//      (aligned words when the caller's level-0 buffer is not 16-byte aligned), a fixed
//      lane -> (row offset, chunk) mapping (no per-chunk division);
//   2. REJECTION TEST, 8 pixels per lane per step: 64-bit shared loads of the centre row and of
//      the rows 3 above / below, funnel shifts for the +-3 column neighbours, VABSDIFF4.U8 for
//      |ring - c| on four pixels at once, and a carry-free per-byte "> th" flag
//      (((d >> 1) & 0x7f) + K, bit 7).  A 9-arc contains one pixel of every opposite ring pair,
//      so (|r0-c|>th or |r8-c|>th) and (|r4-c|>th or |r12-c|>th) is necessary; ~7 % of the
//      pixels of a textured frame pass (the sign-aware scalar test passes 5.3 %);
//   3. passing pixels go to a shared-memory ring in raster order (warp scan of the per-lane
//      counts, each lane writes its own pixels);
//      whenever 32 are queued the exact score = max over the 16 arcs of 9 of min(ring - c) /
//      min(c - ring), minus 1 (cornerScore<16>) is evaluated by 32 busy lanes, for bright and
//      dark arcs AT ONCE on packed s16x2 values (one IMAD builds {ring-c+256, c-ring+256};
//      VIMNMX3.S16x2 does the rest); corners go to the score map and to a raster-ordered list;
//   4. 3x3 non-max suppression inside the cell only (scores outside the cell's detection region
//      are 0 -- NMS never crosses a cell), strict '>' on all 8 neighbours, run over the corner
//      list (a dense pass is kept for cells with more corners than the list holds);
//   5. survivors are emitted in raster order (ballot/popc compaction) into the cell's
//      fixed-capacity slot; if none survived, steps 2-5 repeat with minThFAST.
// Order of candidates = cells row-major, raster inside a cell, which is what the quadtree
// needs to reproduce the reference's first-maximum-wins choice.
#include "orbx_device.cuh"

#ifndef FAST_WARPS
#define FAST_WARPS 4
#endif
#ifndef FAST_DIAG
#define FAST_DIAG 0   // A/B: neutral on B200 (0.736 vs 0.738 ms): fewer exact scores, 52 more instructions per step
#endif
#ifndef FAST_BITMAP
#define FAST_BITMAP 1    // 1: rejection writes a per-row flag bitmap, ONE ordered compaction per 32 detection rows;
                         // 0: the round-1 scheme (per-step warp scan + per-lane push loop into a ring), kept for A/B
#endif
#if FAST_BITMAP
#ifndef FAST_Q
#define FAST_Q 512       // queued pixels (u16) of one group of 32 detection rows; a group with more goes in chunks
#endif
#define FAST_BM_PITCH 12 // bitmap bytes per detection row: 8 pixels per byte, <= 9 byte pairs (dw <= 59), word aligned
#define FAST_BM_ROWS 64  // dh <= 59
#define FAST_BM_BYTES (FAST_BM_PITCH * FAST_BM_ROWS)
#else
#define FAST_Q 512       // ring of queued pixels (u16): <= 31 left over + 256 pushed per step
#define FAST_BM_BYTES 0
#endif
#define FAST_CLIST 256   // corners kept for the sparse NMS pass
#define FAST_HDR_BYTES (2 * (FAST_Q + FAST_CLIST) + FAST_BM_BYTES)

struct FastWarpSmem {
    uint8_t* tile;    // tile_pitch * tile_h, rows 8-byte aligned
    uint8_t* score;   // score_pitch * (dh + 2), 1-px zero frame around the detection region
    uint16_t* queue;  // FAST_Q entries: (tile row << 7) | tile byte column
    uint16_t* clist;  // FAST_CLIST entries, same encoding
    uint8_t* bitmap;  // FAST_BITMAP: flag bits of the detection rows, FAST_BM_PITCH bytes per row
    int xo;           // tile byte column of tile x = 0 (alignment slack of the staged words)
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

// Exact score of up to 32 queued pixels (lane < take); corners (raw > th) go to the score map
// and, in queue (= raster) order, to the corner list.
__device__ __forceinline__ void fast_drain(const FastWarpSmem& s, int tp, int sp, int th, int qhead,
                                           int take, int& ncorn) {
    const int lane = orbx_lane();
    bool corner = false;
    int e = 0;
    if (lane < take) {
#if FAST_BITMAP
        e = s.queue[qhead + lane];
#else
        e = s.queue[(qhead + lane) & (FAST_Q - 1)];
#endif
        const int r = e >> 7, bc = e & 127;
        const int sc = fast_exact_score(s.tile + r * tp + bc, tp);
        if (sc >= th) {
            corner = true;
            s.score[(r - 2) * sp + (bc - s.xo - 2)] = (uint8_t)sc;
        }
    }
    const uint32_t m = __ballot_sync(ORBX_FULL, corner);
    const int at = ncorn + __popc(m & orbx_lanemask_lt());
    if (corner && at < FAST_CLIST) s.clist[at] = (uint16_t)e;
    ncorn += __popc(m);
}

// Per-byte flag (bit 7) of d > th for four packed absolute differences; K4 = (128 - ((th+1)>>1))
// replicated.  Conservative by at most one grey level for even th (d >= th passes).
__device__ __forceinline__ uint32_t fast_gt4(uint32_t d, uint32_t K4) {
    return ((d >> 1) & 0x7f7f7f7fu) + K4;
}
#ifndef FAST_GT_EXACT
#define FAST_GT_EXACT 1
#endif
// Bit 7 of every byte: d1 > th OR d2 > th, for two words of packed absolute differences.
// FAST_GT_EXACT: exact, K = max(127 - th, 0) replicated: the low seven bits plus K carry into bit 7 iff they exceed
// th, and a difference of 128 or more has bit 7 set by itself (th >= 127: every d >= 128 is flagged, a superset);
// two three-input LOP3 merge the four terms.  Else the round-1 form on halved differences (conservative by one grey
// level for even th), K = 128 - ((th + 1) >> 1) replicated.
__device__ __forceinline__ uint32_t fast_gt_or2(uint32_t d1, uint32_t d2, uint32_t K) {
#if FAST_GT_EXACT
    return (((d1 & 0x7f7f7f7fu) + K) | d1) | (((d2 & 0x7f7f7f7fu) + K) | d2);
#else
    return fast_gt4(d1, K) | fast_gt4(d2, K);
#endif
}

// q = x / d for 0 <= x < 64, 1 <= d <= 32 without an integer division: M = 2^16/d rounded up
// (MUFU.RCP), exact because x * d * (M - 2^16/d) < 2^16.
__constant__ uint32_t c_fast_inv[33] = {
    0,     65537, 32769, 21846, 16385, 13108, 10923, 9363, 8193, 7282, 6554, 5958, 5462, 5042, 4682, 4370, 4097,
    3856,  3641,  3450,  3277,  3121,  2979,  2850,  2731, 2622, 2521, 2428, 2341, 2260, 2185, 2115, 2049};
__device__ __forceinline__ uint32_t fast_small_inv(int d) { return c_fast_inv[d]; }
__device__ __forceinline__ int fast_small_div(int x, uint32_t inv) { return (int)(((uint32_t)x * inv) >> 16); }

// The rejection test of eight adjacent detection pixels (one lane's 8-byte pair of a tile row): a FAST-9 corner needs
// |p - c| > th for the pixel three above or three below AND for the one three left or three right (a 9-arc of the
// 16-ring holds two ring-adjacent ones of those four).  Returns the eight flags as a byte.
__device__ __forceinline__ uint32_t fast_reject8(const uint8_t* row, int tp, uint32_t K4, uint32_t vlo, uint32_t vhi) {
    uint32_t flo, fhi;
    const uint2 c = *reinterpret_cast<const uint2*>(row);
    const uint2 up = *reinterpret_cast<const uint2*>(row - 3 * tp);
    const uint2 dn = *reinterpret_cast<const uint2*>(row + 3 * tp);
    const uint32_t l1 = *reinterpret_cast<const uint32_t*>(row - 4);
    const uint32_t r1 = *reinterpret_cast<const uint32_t*>(row + 8);
    // pixels x-3 and x+3 of the eight centres
    const uint32_t wl = __funnelshift_r(l1, c.x, 8), wh_ = __funnelshift_r(c.x, c.y, 8);
    const uint32_t el = __funnelshift_r(c.x, c.y, 24), eh = __funnelshift_r(c.y, r1, 24);
    const uint32_t v_lo = fast_gt_or2(__vabsdiffu4(up.x, c.x), __vabsdiffu4(dn.x, c.x), K4);
    const uint32_t h_lo = fast_gt_or2(__vabsdiffu4(wl, c.x), __vabsdiffu4(el, c.x), K4);
    const uint32_t v_hi = fast_gt_or2(__vabsdiffu4(up.y, c.y), __vabsdiffu4(dn.y, c.y), K4);
    const uint32_t h_hi = fast_gt_or2(__vabsdiffu4(wh_, c.y), __vabsdiffu4(eh, c.y), K4);
    flo = v_lo & h_lo & vlo;
    fhi = v_hi & h_hi & vhi;
#if FAST_DIAG
    // the two diagonal ring pairs (+-2, +-2) as well: a 9-arc also holds one pixel of each of
    // them, and on textured frames this cuts the pixels that reach the exact score from ~10 %
    // to ~3.4 % (evaluated only where the axis pairs left something)
    if (flo | fhi) {
        const uint2 u2 = *reinterpret_cast<const uint2*>(row - 2 * tp);
        const uint2 d2 = *reinterpret_cast<const uint2*>(row + 2 * tp);
        const uint32_t u2l = *reinterpret_cast<const uint32_t*>(row - 2 * tp - 4);
        const uint32_t u2r = *reinterpret_cast<const uint32_t*>(row - 2 * tp + 8);
        const uint32_t d2l = *reinterpret_cast<const uint32_t*>(row + 2 * tp - 4);
        const uint32_t d2r = *reinterpret_cast<const uint32_t*>(row + 2 * tp + 8);
        // (x-2, y-2) / (x+2, y+2) and (x+2, y-2) / (x-2, y+2) of the eight centres
        const uint32_t nw_lo = __funnelshift_r(u2l, u2.x, 16), nw_hi = __funnelshift_r(u2.x, u2.y, 16);
        const uint32_t ne_lo = __funnelshift_r(u2.x, u2.y, 16), ne_hi = __funnelshift_r(u2.y, u2r, 16);
        const uint32_t sw_lo = __funnelshift_r(d2l, d2.x, 16), sw_hi = __funnelshift_r(d2.x, d2.y, 16);
        const uint32_t se_lo = __funnelshift_r(d2.x, d2.y, 16), se_hi = __funnelshift_r(d2.y, d2r, 16);
        const uint32_t a_lo = fast_gt_or2(__vabsdiffu4(nw_lo, c.x), __vabsdiffu4(se_lo, c.x), K4);
        const uint32_t b_lo = fast_gt_or2(__vabsdiffu4(ne_lo, c.x), __vabsdiffu4(sw_lo, c.x), K4);
        const uint32_t a_hi = fast_gt_or2(__vabsdiffu4(nw_hi, c.y), __vabsdiffu4(se_hi, c.y), K4);
        const uint32_t b_hi = fast_gt_or2(__vabsdiffu4(ne_hi, c.y), __vabsdiffu4(sw_hi, c.y), K4);
        flo &= a_lo & b_lo;
        fhi &= a_hi & b_hi;
    }
#endif
    // flags sit in bit 7 of each byte: one multiply per word gathers them into a nibble.
    return ((flo * 0x00204081u) >> 28) | (((fhi * 0x00204081u) >> 24) & 0xf0u);
}

// Computes the score map for threshold `th` (s = raw-1 stored where raw > th, else 0 kept) and
// the raster-ordered corner list.  Returns the number of corners (may exceed FAST_CLIST).
__device__ __forceinline__ int fast_scores(const FastWarpSmem& s, int tp, int sp, int dw, int dh, int th) {
    const int lane = orbx_lane();
    // detection bytes of a tile row: [b0, b1]; 8-byte pairs [plo, phi]
    const int b0 = s.xo + 3, b1 = s.xo + 2 + dw;
    const int plo = b0 >> 3, npair = (b1 >> 3) - plo + 1;
    const uint32_t inv = fast_small_inv(npair);
    const int rpi = fast_small_div(32, inv);   // rows per step
    const int ro = fast_small_div(lane, inv), pc = plo + (lane - ro * npair);
    const bool active = ro < rpi;
    // valid-byte mask (bit 7 of each detection byte) of this lane's pair
    uint32_t vlo, vhi;
    {
        const int lo = max(b0 - pc * 8, 0), hi = min(b1 - pc * 8, 7);  // valid bytes [lo, hi]
        unsigned long long m = 0x8080808080808080ull;
        m = (m >> (8 * lo)) << (8 * lo);
        m = (m << (8 * (7 - hi))) >> (8 * (7 - hi));
        vlo = (uint32_t)m;
        vhi = (uint32_t)(m >> 32);
    }
#if FAST_GT_EXACT
    const uint32_t K4 = (uint32_t)max(127 - th, 0) * 0x01010101u;
#else
    const uint32_t K4 = (uint32_t)(128 - ((th + 1) >> 1)) * 0x01010101u;
#endif
    int ncorn = 0;
#if !FAST_BITMAP
    int qhead = 0, qn = 0;
#endif
    const uint8_t* row = s.tile + (3 + ro) * tp + pc * 8;
    const int rstep = rpi * tp;
    const int rend = 3 + dh;
#if FAST_BITMAP
    // one byte per lane and row into the row bitmap: bit j of byte (pc - plo) of row r - 3 = tile byte column
    // pc * 8 + j.  No warp collective inside: every lane walks its own rows.
    if (active) {
        uint8_t* bm = s.bitmap + ro * FAST_BM_PITCH + (pc - plo);
        const int bstep = rpi * FAST_BM_PITCH;
        for (int r = 3 + ro; r < rend; r += rpi, row += rstep, bm += bstep) *bm = (uint8_t)fast_reject8(row, tp, K4, vlo, vhi);
    }
#else
    for (int r = 3 + ro; r < rend + ro; r += rpi, row += rstep) {   // r - ro is warp-uniform
        const uint32_t nib0 = (active && r < rend) ? fast_reject8(row, tp, K4, vlo, vhi) : 0u;
#endif
#if FAST_BITMAP
    __syncwarp();
    // ORDERED COMPACTION, once per group of 32 detection rows: lane = row, its flags are <= 3 words (the bytes
    // past the row's last pair are stale and masked); raster order = (row, word, bit) = lane-major, so one warp
    // scan of the per-row counts gives every flagged pixel its position in the queue.  Then the exact score runs
    // on full 32-lane batches.  A group with more flagged pixels than the queue holds goes in chunks.
    const int nwords = (npair + 3) >> 2;
    for (int g0 = 0; g0 < dh; g0 += 32) {
        const int rr = g0 + lane;
        uint32_t w[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            w[k] = 0;
            if (rr < dh && k < nwords) {
                const int nb = min(npair - 4 * k, 4);   // valid bytes of this word (>= 1)
                w[k] = *reinterpret_cast<const uint32_t*>(s.bitmap + rr * FAST_BM_PITCH + 4 * k) &
                       (0xffffffffu >> (32 - 8 * nb));
            }
        }
        const uint32_t cnt = __popc(w[0]) + __popc(w[1]) + __popc(w[2]);
        const uint32_t incl = warp_incl_scan(cnt);
        const int total = (int)__shfl_sync(ORBX_FULL, incl, 31);
        const int ebase = ((rr + 3) << 7) | (plo * 8);
        for (int done = 0; done < total; done += FAST_Q) {
            int at = (int)(incl - cnt) - done;
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                uint32_t m = w[k];
                while (m) {
                    const int bit = __ffs(m) - 1;
                    m &= m - 1;
                    if ((unsigned)at < (unsigned)FAST_Q) s.queue[at] = (uint16_t)(ebase + 32 * k + bit);
                    ++at;
                }
            }
            __syncwarp();
            const int nq = min(total - done, FAST_Q);
            for (int q0 = 0; q0 < nq; q0 += 32) fast_drain(s, tp, sp, th, q0, min(nq - q0, 32), ncorn);
            __syncwarp();
        }
    }
#else
        // Flagged pixels go straight to the pixel ring in raster order (lane order = (row, pair)
        // order): warp scan of the per-lane counts, then each lane writes its own pixels.
        if (__any_sync(ORBX_FULL, nib0 != 0)) {
            uint32_t nib = nib0;
            const uint32_t cnt = __popc(nib);
            const uint32_t incl = warp_incl_scan(cnt);
            int at = qhead + qn + (int)(incl - cnt);
            const int ebase = (r << 7) | (pc * 8);
            while (nib) {
                const int k = __ffs(nib) - 1;
                nib &= nib - 1;
                s.queue[at & (FAST_Q - 1)] = (uint16_t)(ebase + k);
                ++at;
            }
            qn += (int)__shfl_sync(ORBX_FULL, incl, 31);
            __syncwarp();
            while (qn >= 32) {
                fast_drain(s, tp, sp, th, qhead, 32, ncorn);
                qhead = (qhead + 32) & (FAST_Q - 1);
                qn -= 32;
            }
        }
    }
    while (qn > 0) {
        const int take = min(qn, 32);
        fast_drain(s, tp, sp, th, qhead, take, ncorn);
        qhead = (qhead + take) & (FAST_Q - 1);
        qn -= take;
    }
#endif
    __syncwarp();
    return ncorn;
}

// NMS + ordered emission over the corner list.  Returns the number of keypoints written.
__device__ __forceinline__ int fast_nms_sparse(const FastWarpSmem& s, int sp, int ncorn, int xadd,
                                               int yadd, uint32_t* __restrict__ slot) {
    const int lane = orbx_lane();
    int count = 0;
    for (int base = 0; base < ncorn; base += 32) {
        bool keep = false;
        int x = 0, y = 0, sc = 0;
        if (base + lane < ncorn) {
            const int e = s.clist[base + lane];
            y = (e >> 7) - 3;
            x = (e & 127) - s.xo - 3;
            const uint8_t* c = s.score + (y + 1) * sp + (x + 1);
            sc = c[0];
            keep = sc > c[-1] && sc > c[1] && sc > c[-sp - 1] && sc > c[-sp] && sc > c[-sp + 1] &&
                   sc > c[sp - 1] && sc > c[sp] && sc > c[sp + 1];
        }
        const uint32_t m = __ballot_sync(ORBX_FULL, keep);
        if (keep) slot[count + __popc(m & orbx_lanemask_lt())] = ORBX_MAKE_KEY(x + xadd, y + yadd, sc);
        count += __popc(m);
    }
    return count;
}

// Dense NMS + ordered emission (cells with more than FAST_CLIST corners).
__device__ __forceinline__ int fast_nms_emit(const FastWarpSmem& s, int sp, int dw, int dh,
                                             uint32_t dw_inv, int xadd, int yadd,
                                             uint32_t* __restrict__ slot) {
    const int lane = orbx_lane();
    const int npix = dw * dh;
    dw_inv = (1u << 20) / (uint32_t)dw + 1;
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

#ifndef FAST_MINB
#define FAST_MINB 8
#endif
__global__ void __launch_bounds__(FAST_WARPS * 32, FAST_MINB)
k_fast_cells(const __grid_constant__ OrbxPlan P, const OrbxSrc isrc, const uint8_t* __restrict__ pyr,
             uint32_t* __restrict__ cellcnt, uint32_t* __restrict__ cellkeys, int tile_pitch,
             int score_pitch, int warp_smem_bytes, int score_off, int src16, int cell_base, int cell_end) {
    extern __shared__ __align__(16) uint8_t smem[];
    const int warp = orbx_warp(), lane = orbx_lane();
    const int g = cell_base + blockIdx.x * FAST_WARPS + warp;  // global cell index inside the frame
    if (g >= cell_end) return;
    const int frame = blockIdx.y;

    // the cell's record (built once per shape on the host, ORBextractor.cc:787-804)
    const uint4 rec = __ldg(P.cell_tab + g);
    const int iniX = (int)(rec.x & 0xffffu), iniY = (int)(rec.x >> 16);
    const int tw = (int)(rec.y & 0xffu), th_ = (int)((rec.y >> 8) & 0xffu), l = (int)(rec.y >> 16);
    uint32_t* cnt_out = cellcnt + (size_t)frame * P.cells_total + g;
    if (tw == 0) {
        if (lane == 0) *cnt_out = 0;
        return;
    }
    uint32_t* slot = cellkeys + (size_t)frame * P.keys_total + rec.z;
    const int maxX = iniX + tw;
    const int dw = tw - 6, dh = th_ - 6;  // detection region = tile pixels [3, t-4]

    FastWarpSmem s;
    uint8_t* base = smem + (size_t)warp * warp_smem_bytes;
    s.queue = reinterpret_cast<uint16_t*>(base);
    s.clist = s.queue + FAST_Q;
    s.bitmap = base + 2 * (FAST_Q + FAST_CLIST);
    s.tile = base + FAST_HDR_BYTES;
    s.score = s.tile + score_off;

    // 1. stage the tile with cp.async: 16-byte chunks covering [iniX & ~15, maxX) when the
    //    level's rows are 16-byte aligned (every pyramid level; level 0 unless the caller's
    //    device buffer is not), else aligned words; lane -> (row offset, chunk)
    int lpitch;
    const uint8_t* src = orbx_level_ptr(P, isrc, pyr, frame, l, &lpitch);
    if (l != 0 || src16) {
        const int ax = iniX & ~15;
        s.xo = iniX - ax;  // tile x -> smem byte x + xo
        const int nch = (maxX - ax + 15) >> 4;   // <= tile_pitch / 16 <= 7
        const uint32_t inv = fast_small_inv(nch);
        const int rpi = fast_small_div(32, inv);
        const int ro = fast_small_div(lane, inv), cw = lane - ro * nch;
        if (ro < rpi) {
            const uint8_t* gp = src + (size_t)(iniY + ro) * lpitch + ax + 16 * cw;
            uint32_t sa = (uint32_t)__cvta_generic_to_shared(s.tile + ro * tile_pitch + 16 * cw);
            const size_t gstep = (size_t)rpi * lpitch;
            const uint32_t tstep = (uint32_t)(rpi * tile_pitch);
#pragma unroll 2
            for (int r = ro; r < th_; r += rpi) {
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(gp) : "memory");
                gp += gstep;
                sa += tstep;
            }
        }
    } else {
        const int ax = iniX & ~3;
        s.xo = iniX - ax;
        const int nw = (maxX - ax + 3) >> 2;   // <= 21
        const uint32_t inv = fast_small_inv(nw);
        const int rpi = fast_small_div(32, inv);
        const int ro = fast_small_div(lane, inv), cw = lane - ro * nw;
        if (ro < rpi) {
            const uint8_t* gp = src + (size_t)(iniY + ro) * lpitch + ax + 4 * cw;
            uint32_t sa = (uint32_t)__cvta_generic_to_shared(s.tile + ro * tile_pitch + 4 * cw);
            const size_t gstep = (size_t)rpi * lpitch;
            const uint32_t tstep = (uint32_t)(rpi * tile_pitch);
#pragma unroll 4
            for (int r = ro; r < th_; r += rpi) {
                asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sa), "l"(gp) : "memory");
                gp += gstep;
                sa += tstep;
            }
        }
    }
    // zero the score map (frame included; 16-byte stores, the region is padded to 16)
    {
        const int sq = (score_pitch * (dh + 2) + 15) >> 4;
        for (int i = lane; i < sq; i += 32) reinterpret_cast<uint4*>(s.score)[i] = make_uint4(0, 0, 0, 0);
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
    __syncwarp();

    const int xadd = (int)(rec.w & 0xffffu), yadd = (int)(rec.w >> 16);
    const uint32_t dw_inv = 0;  // computed inside the (rare) dense NMS path

    int nc = fast_scores(s, tile_pitch, score_pitch, dw, dh, P.ini_th);
    int n = nc <= FAST_CLIST ? fast_nms_sparse(s, score_pitch, nc, xadd, yadd, slot)
                             : fast_nms_emit(s, score_pitch, dw, dh, dw_inv, xadd, yadd, slot);
    if (n == 0) {
        // ORBextractor.cc:824-828: rerun the cell with minThFAST.  Scores already stored are
        // threshold independent, so the map is only extended; the corner list is rebuilt.
        nc = fast_scores(s, tile_pitch, score_pitch, dw, dh, P.min_th);
        n = nc <= FAST_CLIST ? fast_nms_sparse(s, score_pitch, nc, xadd, yadd, slot)
                             : fast_nms_emit(s, score_pitch, dw, dh, dw_inv, xadd, yadd, slot);
    }
    if (lane == 0) *cnt_out = (uint32_t)n;
}

#ifndef FAST_PITCH_ODD
#define FAST_PITCH_ODD 1
#endif
void orbx_fast_smem_layout(const OrbxPlan& P, int* tile_pitch, int* score_pitch, int* warp_bytes, int* score_off) {
    // tile rows hold the 16-byte chunks covering up to 15 (alignment slack) + max_tile_w bytes;
    // an odd number of chunks per row keeps rows r and r + 2 on different banks.  The 8-byte pair /
    // +-1 word reach of the rejection test may run into the next row (one spare row is allocated;
    // only bytes outside the detection region, which are masked, can come from there).
    int tp = ((P.max_tile_w + 15 + 15) / 16) * 16;
    if (FAST_PITCH_ODD && (tp / 16) % 2 == 0) tp += 16;
    *tile_pitch = tp;
    *score_pitch = ((P.max_tile_w - 6 + 2 + 3) / 4) * 4;
    *score_off = (tp * (P.max_tile_h + 1) + 15) & ~15;
    int bytes = FAST_HDR_BYTES + *score_off + ((*score_pitch * (P.max_tile_h - 6 + 2) + 15) & ~15);
    *warp_bytes = (bytes + 15) & ~15;
}

// Called per (device, plan).  The attribute belongs to the KERNEL on the device, not to the handle: a second handle
// with a smaller plan must not lower it under a live handle with a larger one (mono builds an N- and a 5N-feature
// extractor, src/Tracking.cc:206-212).  So it is always raised to the opt-in maximum; the plan's need is only checked.
cudaError_t orbx_configure_fast(const OrbxPlan& P) {
    int tp, sp, wb, so;
    orbx_fast_smem_layout(P, &tp, &sp, &wb, &so);
    if (wb * FAST_WARPS > ORBX_MAX_DYN_SMEM) return cudaErrorInvalidConfiguration;
    return cudaFuncSetAttribute(k_fast_cells, cudaFuncAttributeMaxDynamicSharedMemorySize, ORBX_MAX_DYN_SMEM);
}

cudaError_t orbx_launch_fast(const OrbxPlan& P, const OrbxSrc& src, const uint8_t* pyr, uint32_t* cellcnt,
                             uint32_t* cellkeys, int nframes, cudaStream_t st, int* launches, int level) {
    int tp, sp, wb, so;
    orbx_fast_smem_layout(P, &tp, &sp, &wb, &so);
    const int smem = wb * FAST_WARPS;
    const int src16 = ((((uintptr_t)src.img0) | (uintptr_t)src.stride0 | (uintptr_t)src.pitch0) & 15) == 0;
    // level < 0: every level in one launch; else only that level's cells (the forked latency graph)
    const int c0 = level < 0 ? 0 : P.lvl[level].cell_off;
    const int c1 = level < 0 ? P.cells_total : c0 + P.lvl[level].n_cols * P.lvl[level].n_rows;
    dim3 grid((c1 - c0 + FAST_WARPS - 1) / FAST_WARPS, nframes);
    k_fast_cells<<<grid, FAST_WARPS * 32, smem, st>>>(P, src, pyr, cellcnt, cellkeys, tp, sp, wb, so, src16, c0, c1);
    ++*launches;
    return cudaGetLastError();
}
