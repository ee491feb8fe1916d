// k_match.cu -- M1/M5/M6: batched 256-bit Hamming distances (ORBmatcher::DescriptorDistance,
// src/ORBmatcher.cc:2353-2369; identical to DBoW2 FORB::distance) as __popc over 8 words.
// Nothing here is a dense contraction: descriptors are 32-byte rows read as two uint4, the
// working sets (tens of KB to ~1 MB) live in L1/L2, and the kernels are popcount/LSU bound.
//
//   k_hamming_csr   one thread per (query, candidate) pair of a CSR candidate list; the row of a
//                   pair is found by binary search in row_ptr (L1-resident); distances are written
//                   in candidate order so the host can replay the reference's order-dependent
//                   accept loops bit-exactly.
//   k_knn2          cv::BFMatcher(NORM_HAMMING).knnMatch(k=2): one warp per query, lanes stride
//                   the train rows keeping their two best (distance, index) pairs; shuffle merge;
//                   ties resolve to the LOWEST train index (packed key = dist << 20 | idx).
//   k_stereo_rows   Frame::ComputeStereoMatches' Hamming search (src/Frame.cc:840-885): one warp
//                   per left keypoint, gates evaluated per right keypoint, warp-min reduction.
#include "orbx_device.cuh"
#include "../../include/orbm.h"
#include "orbm_internal.h"

__device__ __forceinline__ int hamming256(const uint4 a0, const uint4 a1, const uint4* __restrict__ b) {
    const uint4 b0 = __ldg(b), b1 = __ldg(b + 1);
    return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
           __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

__global__ void __launch_bounds__(256)
k_hamming_csr(const uint4* __restrict__ q, int nq, const uint4* __restrict__ t,
              const int32_t* __restrict__ row_ptr, const int32_t* __restrict__ cand, int nnz,
              uint16_t* __restrict__ dist) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nnz) return;
    // largest r with row_ptr[r] <= k
    int lo = 0, hi = nq;
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (__ldg(row_ptr + mid) <= k) lo = mid; else hi = mid;
    }
    const uint4 a0 = __ldg(q + 2 * lo), a1 = __ldg(q + 2 * lo + 1);
    dist[k] = (uint16_t)hamming256(a0, a1, t + 2 * (size_t)__ldg(cand + k));
}

__global__ void __launch_bounds__(256)
k_knn2(const uint4* __restrict__ q, int nq, const uint4* __restrict__ t, int nt,
       int32_t* __restrict__ idx_out, int32_t* __restrict__ dist_out) {
    const int qi = blockIdx.x * (blockDim.x >> 5) + orbx_warp();
    if (qi >= nq) return;
    const int lane = orbx_lane();
    const uint4 a0 = __ldg(q + 2 * qi), a1 = __ldg(q + 2 * qi + 1);
    // packed key: distance (<= 256) above a 20-bit train index; smaller key = better, and equal
    // distances order by index, i.e. lowest train index first
    uint32_t b1 = 0xffffffffu, b2 = 0xffffffffu;
    for (int j = lane; j < nt; j += 32) {
        const uint32_t key = ((uint32_t)hamming256(a0, a1, t + 2 * (size_t)j) << 20) | (uint32_t)j;
        if (key < b1) { b2 = b1; b1 = key; }
        else if (key < b2) b2 = key;
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        const uint32_t o1 = __shfl_xor_sync(ORBX_FULL, b1, d), o2 = __shfl_xor_sync(ORBX_FULL, b2, d);
        // merge two sorted pairs, keep the two smallest
        const uint32_t lo = min(b1, o1), hi = max(b1, o1);
        b2 = min(hi, min(b2, o2));
        b1 = lo;
    }
    if (lane == 0) {
        idx_out[2 * qi] = b1 == 0xffffffffu ? -1 : (int32_t)(b1 & 0xfffffu);
        dist_out[2 * qi] = b1 == 0xffffffffu ? -1 : (int32_t)(b1 >> 20);
        idx_out[2 * qi + 1] = b2 == 0xffffffffu ? -1 : (int32_t)(b2 & 0xfffffu);
        dist_out[2 * qi + 1] = b2 == 0xffffffffu ? -1 : (int32_t)(b2 >> 20);
    }
}

__global__ void __launch_bounds__(256)
k_stereo_rows(const orbx_keypoint* __restrict__ kl, const uint4* __restrict__ dl, int nl,
              const orbx_keypoint* __restrict__ kr, const uint4* __restrict__ dr, int nr,
              const float* __restrict__ scale, float min_d, float max_d, int32_t* __restrict__ best_idx,
              int32_t* __restrict__ best_dist) {
    const int il = blockIdx.x * (blockDim.x >> 5) + orbx_warp();
    if (il >= nl) return;
    const int lane = orbx_lane();
    const orbx_keypoint L = kl[il];
    const int row = (int)L.y;                  // vRowIndices[vL], Frame.cc:846
    const float min_u = L.x - max_d, max_u = L.x - min_d;  // :851-852
    uint32_t best = ((uint32_t)ORBM_TH_HIGH << 20) | 0xfffffu;  // bestDist = TH_HIGH, no index
    if (!(max_u < 0)) {                        // :854
        const uint4 a0 = __ldg(dl + 2 * il), a1 = __ldg(dl + 2 * il + 1);
        for (int j = lane; j < nr; j += 32) {
            const orbx_keypoint R = kr[j];
            const float r = 2.0f * scale[R.octave];   // :823
            const int maxr = (int)ceilf(R.y + r), minr = (int)floorf(R.y - r);
            if (row < minr || row > maxr) continue;
            if (R.octave < L.octave - 1 || R.octave > L.octave + 1) continue;  // :869
            if (!(R.x >= min_u && R.x <= max_u)) continue;                       // :874
            const uint32_t key = ((uint32_t)hamming256(a0, a1, dr + 2 * (size_t)j) << 20) | (uint32_t)j;
            best = min(best, key);
        }
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) best = min(best, __shfl_xor_sync(ORBX_FULL, best, d));
    if (lane == 0) {
        const int dist = (int)(best >> 20);
        const bool found = dist < ORBM_TH_HIGH;   // strict '<' against bestDist = TH_HIGH, :880
        best_idx[il] = found ? (int32_t)(best & 0xfffffu) : -1;
        best_dist[il] = found ? dist : ORBM_TH_HIGH;
    }
}

// ---- N1: sub-pixel stereo refinement, Frame::ComputeStereoMatches src/Frame.cc:888-963 -----------
// One warp per left keypoint whose Hamming match is better than thOrbDist.  The 11x11 left patch
// and the 11x21 right strip (11 window positions, incR = -5..5) of level kpL.octave are staged in
// shared memory straight from the two extractors' device-resident pyramids; the reference reads a
// 19-px BORDER_REFLECT_101 border there (mvImagePyramid is a padded buffer), here the reflection
// is applied while staging.  SAD of the centre-subtracted int16 patches (cv::norm L1), first
// minimum wins, parabola fit in non-contracted float32.
struct OrbmPyr {
    const uint8_t* data[ORBX_MAX_LEVELS];
    int pitch[ORBX_MAX_LEVELS], w[ORBX_MAX_LEVELS], h[ORBX_MAX_LEVELS];
};

__device__ __forceinline__ int orbm_reflect101(int p, int len) {
    if (p < 0) p = -p;
    if (p >= len) p = 2 * (len - 1) - p;
    return p;
}

__global__ void __launch_bounds__(256)
k_stereo_refine(const orbx_keypoint* __restrict__ kl, int nl, const orbx_keypoint* __restrict__ kr,
                const int32_t* __restrict__ best_idx, const int32_t* __restrict__ best_dist,
                const __grid_constant__ OrbmPyr PL, const __grid_constant__ OrbmPyr PR,
                const float* __restrict__ scale, const float* __restrict__ inv_scale, int th_orb,
                float min_d, float max_d, float bf, float* __restrict__ u_right,
                float* __restrict__ depth, int32_t* __restrict__ sad) {
    __shared__ int16_t sm[8][11 * 11 + 11 * 21];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int il = blockIdx.x * 8 + warp;
    if (il >= nl) return;
    if (lane == 0) {
        u_right[il] = -1.0f;
        depth[il] = -1.0f;
        sad[il] = -1;
    }
    const int ir = best_idx[il];
    if (ir < 0 || best_dist[il] >= th_orb) return;   // :889
    const orbx_keypoint L = kl[il];
    const int oct = L.octave;
    const float uR0 = kr[ir].x;
    const float sf = inv_scale[oct];
    const float scaleduL = roundf(__fmul_rn(L.x, sf));       // :893-895 (round half away from zero)
    const float scaledvL = roundf(__fmul_rn(L.y, sf));
    const float scaleduR0 = roundf(__fmul_rn(uR0, sf));
    const int w = 5, Lw = 5;
    const float iniu = scaleduR0 + Lw - w, endu = scaleduR0 + Lw + w + 1;   // :911-912
    if (iniu < 0 || endu >= (float)PR.w[oct]) return;                        // :913
    const int xl = (int)scaleduL, yl = (int)scaledvL, xr = (int)scaleduR0;
    int16_t* IL = sm[warp];
    int16_t* IR = IL + 121;
    const uint8_t* pl = PL.data[oct];
    const uint8_t* pr = PR.data[oct];
    const int cl = pl[(size_t)orbm_reflect101(yl, PL.h[oct]) * PL.pitch[oct] + orbm_reflect101(xl, PL.w[oct])];
    for (int i = lane; i < 121; i += 32) {
        const int r = i / 11, c = i - r * 11;
        const int y = orbm_reflect101(yl - w + r, PL.h[oct]), x = orbm_reflect101(xl - w + c, PL.w[oct]);
        IL[i] = (int16_t)((int)pl[(size_t)y * PL.pitch[oct] + x] - cl);      // IL - IL.at<short>(w,w)
    }
    for (int i = lane; i < 231; i += 32) {
        const int r = i / 21, c = i - r * 21;
        const int y = orbm_reflect101(yl - w + r, PR.h[oct]), x = orbm_reflect101(xr - Lw - w + c, PR.w[oct]);
        IR[i] = (int16_t)pr[(size_t)y * PR.pitch[oct] + x];
    }
    __syncwarp();
    // lane k < 11 owns window incR = k - 5: centre pixel IR[5][k + 5]
    int dist = 0x7fffffff;
    if (lane < 11) {
        const int cr = IR[5 * 21 + lane + 5];
        int acc = 0;
        for (int r = 0; r < 11; ++r)
            for (int c = 0; c < 11; ++c) {
                const int d = (int)IL[r * 11 + c] - ((int)IR[r * 21 + lane + c] - cr);
                acc += d < 0 ? -d : d;
            }
        dist = acc;
    }
    // first minimum (strict '<', :927-931): smallest (dist, incR)
    int bd = dist, bk = lane;
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        const int od = __shfl_xor_sync(ORBX_FULL, bd, d), ok = __shfl_xor_sync(ORBX_FULL, bk, d);
        if (od < bd || (od == bd && ok < bk)) { bd = od; bk = ok; }
    }
    const int incR = bk - Lw;
    if (incR == -Lw || incR == Lw) return;                                   // :936
    const float dist1 = (float)__shfl_sync(ORBX_FULL, dist, bk - 1);
    const float dist2 = (float)bd;
    const float dist3 = (float)__shfl_sync(ORBX_FULL, dist, bk + 1);
    if (lane != 0) return;
    const float deltaR = __fdiv_rn(__fsub_rn(dist1, dist3),
                                   __fmul_rn(2.0f, __fsub_rn(__fadd_rn(dist1, dist3), __fmul_rn(2.0f, dist2))));  // :944
    if (deltaR < -1 || deltaR > 1) return;                                   // :946 (NaN passes, as in C++)
    float bestuR = __fmul_rn(scale[oct], __fadd_rn(__fadd_rn(scaleduR0, (float)incR), deltaR));   // :950
    float disparity = __fsub_rn(L.x, bestuR);
    if (disparity >= min_d && disparity < max_d) {                           // :954
        if (disparity <= 0) {
            disparity = 0.01f;
            bestuR = (float)((double)L.x - 0.01);                            // uL-0.01: double arithmetic
        }
        depth[il] = __fdiv_rn(bf, disparity);
        u_right[il] = bestuR;
        sad[il] = bd;
    }
}

cudaError_t orbm_launch_stereo_refine(const orbx_keypoint* kl, int nl, const orbx_keypoint* kr,
                                      const int32_t* best_idx, const int32_t* best_dist,
                                      const orbx_pyramid_view* vl, const orbx_pyramid_view* vr,
                                      const float* scale, const float* inv_scale, int th_orb, float min_d,
                                      float max_d, float bf, float* u_right, float* depth, int32_t* sad,
                                      cudaStream_t st) {
    if (nl <= 0) return cudaSuccess;
    OrbmPyr A, B;
    for (int l = 0; l < ORBX_MAX_LEVELS; ++l) {
        A.data[l] = vl->data[l]; A.pitch[l] = vl->pitch[l]; A.w[l] = vl->width[l]; A.h[l] = vl->height[l];
        B.data[l] = vr->data[l]; B.pitch[l] = vr->pitch[l]; B.w[l] = vr->width[l]; B.h[l] = vr->height[l];
    }
    k_stereo_refine<<<(nl + 7) / 8, 256, 0, st>>>(kl, nl, kr, best_idx, best_dist, A, B, scale, inv_scale, th_orb,
                                                  min_d, max_d, bf, u_right, depth, sad);
    return cudaGetLastError();
}

cudaError_t orbm_launch_hamming_csr(const uint8_t* q, int nq, const uint8_t* t, const int32_t* row_ptr,
                                    const int32_t* cand, int nnz, uint16_t* dist, cudaStream_t st) {
    if (nnz <= 0) return cudaSuccess;
    k_hamming_csr<<<(nnz + 255) / 256, 256, 0, st>>>((const uint4*)q, nq, (const uint4*)t, row_ptr, cand, nnz, dist);
    return cudaGetLastError();
}

cudaError_t orbm_launch_knn2(const uint8_t* q, int nq, const uint8_t* t, int nt, int32_t* idx,
                             int32_t* dist, cudaStream_t st) {
    if (nq <= 0) return cudaSuccess;
    k_knn2<<<(nq + 7) / 8, 256, 0, st>>>((const uint4*)q, nq, (const uint4*)t, nt, idx, dist);
    return cudaGetLastError();
}

cudaError_t orbm_launch_stereo_rows(const orbx_keypoint* kl, const uint8_t* dl, int nl,
                                    const orbx_keypoint* kr, const uint8_t* dr, int nr,
                                    const float* scale, float min_d, float max_d, int32_t* best_idx,
                                    int32_t* best_dist, cudaStream_t st) {
    if (nl <= 0) return cudaSuccess;
    k_stereo_rows<<<(nl + 7) / 8, 256, 0, st>>>(kl, (const uint4*)dl, nl, kr, (const uint4*)dr, nr, scale, min_d,
                                                max_d, best_idx, best_dist);
    return cudaGetLastError();
}

// ---- N2 / M7: Frame::AssignFeaturesToGrid + Frame::GetFeaturesInArea on the device ---------------
// (src/Frame.cc:377-408, 645-726).  The 64 x 48 bucket grid of a frame is built by one CTA
// (counting sort by cell, cell index = ix * 48 + iy, insertion order = keypoint index inside a
// cell), then one warp per query walks the window's cells in the reference's order -- columns
// outer, rows inner; the cells of one column are adjacent in the item array, so a column of the
// window is ONE contiguous item range -- filters by octave range and |dx| < r && |dy| < r, and
// compacts the survivors in order with ballots.  A count pass sizes the CSR rows; the fill pass
// also evaluates the Hamming distance of every (query, candidate) pair, so the order-dependent
// accept loops of M2-M4 only have to read (candidate, distance) lists.
#define ORBM_GRID_CELLS (ORBM_GRID_COLS * ORBM_GRID_ROWS)

struct OrbmGrid {
    float min_x, min_y, w_inv, h_inv;
};

// Frames of up to ORBM_GRID_SMEM_ITEMS keypoints (every config of BASELINE.json) keep the cell of each
// keypoint in registers between the two passes and the item array in shared memory until it is sorted:
// the kernel is one CTA's dependent chain, and every global round trip in it (cellof[], the insertion
// sort on items[]) is a microsecond of a 12 us kernel.
#define ORBM_GRID_SMEM_ITEMS 4096
#define ORBM_GRID_PER_THREAD (ORBM_GRID_SMEM_ITEMS / 1024)

__device__ __forceinline__ void grid_sort_cells(int32_t* items, const uint32_t* cnt, int tid, int nthreads) {
    // restore insertion order (ascending keypoint index) inside every cell: a handful of items
    for (int c = tid; c < ORBM_GRID_CELLS; c += nthreads) {
        const int s = (int)cnt[c], e = (int)cnt[c + 1];
        for (int a = s + 1; a < e; ++a) {
            const int32_t v = items[a];
            int b = a - 1;
            while (b >= s && items[b] > v) {
                items[b + 1] = items[b];
                --b;
            }
            items[b + 1] = v;
        }
    }
}

__global__ void __launch_bounds__(1024)
k_grid_build(const orbx_keypoint* __restrict__ keys, int n, OrbmGrid G, int32_t* __restrict__ start,
             int32_t* __restrict__ items, int32_t* __restrict__ cellof) {
    __shared__ uint32_t cnt[ORBM_GRID_CELLS + 1];
    __shared__ uint32_t cur[ORBM_GRID_CELLS];
    __shared__ uint32_t tmp[64];
    __shared__ int32_t sitems[ORBM_GRID_SMEM_ITEMS];
    const int tid = threadIdx.x;
    const bool small = n <= ORBM_GRID_SMEM_ITEMS;
    int mycell[ORBM_GRID_PER_THREAD];
    for (int c = tid; c <= ORBM_GRID_CELLS; c += blockDim.x) cnt[c] = 0;
    __syncthreads();
    // PosInGrid (:716-726): round, not floor; keypoints outside the grid are not inserted
    auto cell_of = [&](int i) {
        const int px = (int)roundf((keys[i].x - G.min_x) * G.w_inv);
        const int py = (int)roundf((keys[i].y - G.min_y) * G.h_inv);
        return (px >= 0 && px < ORBM_GRID_COLS && py >= 0 && py < ORBM_GRID_ROWS) ? px * ORBM_GRID_ROWS + py : -1;
    };
    if (small) {
#pragma unroll
        for (int j = 0; j < ORBM_GRID_PER_THREAD; ++j) {
            const int i = tid + j * 1024;
            mycell[j] = i < n ? cell_of(i) : -1;
            if (mycell[j] >= 0) atomicAdd(&cnt[mycell[j]], 1u);
        }
    } else {
        for (int i = tid; i < n; i += blockDim.x) {
            const int cell = cell_of(i);
            if (cell >= 0) atomicAdd(&cnt[cell], 1u);
            cellof[i] = cell;
        }
    }
    __syncthreads();
    const uint32_t total = block_excl_scan(cnt, ORBM_GRID_CELLS, tmp);
    if (tid == 0) cnt[ORBM_GRID_CELLS] = total;
    __syncthreads();
    for (int c = tid; c <= ORBM_GRID_CELLS; c += blockDim.x) {
        start[c] = (int32_t)cnt[c];
        if (c < ORBM_GRID_CELLS) cur[c] = cnt[c];
    }
    __syncthreads();
    if (small) {
#pragma unroll
        for (int j = 0; j < ORBM_GRID_PER_THREAD; ++j)
            if (mycell[j] >= 0) sitems[atomicAdd(&cur[mycell[j]], 1u)] = tid + j * 1024;
        __syncthreads();
        grid_sort_cells(sitems, cnt, tid, blockDim.x);
        __syncthreads();
        for (int k = tid; k < (int)total; k += blockDim.x) items[k] = sitems[k];
        return;
    }
    for (int i = tid; i < n; i += blockDim.x) {
        const int cell = cellof[i];
        if (cell >= 0) items[atomicAdd(&cur[cell], 1u)] = i;
    }
    __syncthreads();
    grid_sort_cells(items, cnt, tid, blockDim.x);
}

// One window query by one warp: GetFeaturesInArea (src/Frame.cc:645-714) over the 64 x 48 grid.  FILL: candidates (in
// the reference's order) and their Hamming distances go to cand / dist from position `base`; else they are counted.
template <bool FILL, bool PAIR = false>
__device__ __forceinline__ int area_walk(const orbx_keypoint* __restrict__ keys, const int32_t* __restrict__ start,
                                         const int32_t* __restrict__ items, const OrbmGrid& G, const orbm_area_query& q,
                                         int base, const uint4* __restrict__ qdesc, const uint4* __restrict__ tdesc,
                                         int32_t* __restrict__ cand, uint16_t* __restrict__ dist, int cap) {
    const int lane = orbx_lane();
    const uint32_t lt = orbx_lanemask_lt();
    int total = 0;
    const float r = q.r;
    // GetFeaturesInArea :652-676
    const int c0 = max(0, (int)floorf((q.x - G.min_x - r) * G.w_inv));
    const int c1 = min(ORBM_GRID_COLS - 1, (int)ceilf((q.x - G.min_x + r) * G.w_inv));
    const int r0 = max(0, (int)floorf((q.y - G.min_y - r) * G.h_inv));
    const int r1 = min(ORBM_GRID_ROWS - 1, (int)ceilf((q.y - G.min_y + r) * G.h_inv));
    if (q.valid && c0 < ORBM_GRID_COLS && c1 >= 0 && r0 < ORBM_GRID_ROWS && r1 >= 0) {
        const bool check_levels = (q.min_level > 0) || (q.max_level >= 0);
        uint4 a0 = make_uint4(0, 0, 0, 0), a1 = a0;
        if (FILL) {
            a0 = __ldg(qdesc + 2 * (size_t)q.desc_row);
            a1 = __ldg(qdesc + 2 * (size_t)q.desc_row + 1);
        }
        // A column of the window is ONE contiguous range of the cell-sorted item array.  Lane i fetches the range of
        // column cb + i, a warp scan concatenates the ranges, and the items are then walked 32 at a time ACROSS the
        // columns (item j's column by a 5-step search of the scan): a window of two or three short columns is one
        // round of dependent loads (range -> item -> keypoint) instead of one per column.  Order = column, row, cell
        // order = the reference's (:678-707).
        for (int cb = c0; cb <= c1 && r0 <= r1; cb += 32) {
            const int ix = cb + lane;
            int s = 0, len = 0;
            if (ix <= c1) {
                s = __ldg(start + ix * ORBM_GRID_ROWS + r0);
                len = __ldg(start + ix * ORBM_GRID_ROWS + r1 + 1) - s;
            }
            const int incl = (int)warp_incl_scan((uint32_t)len);
            const int T = __shfl_sync(ORBX_FULL, incl, 31);
            for (int j0 = 0; j0 < T; j0 += 32) {
                const int j = j0 + lane;
                int lo = 0, hi = 31;   // smallest lane whose inclusive count exceeds j
#pragma unroll
                for (int step = 0; step < 5; ++step) {
                    const int mid = (lo + hi) >> 1;
                    const int v = __shfl_sync(ORBX_FULL, incl, mid);
                    if (v > j) hi = mid; else lo = mid + 1;
                }
                const int s_col = __shfl_sync(ORBX_FULL, s, lo), incl_col = __shfl_sync(ORBX_FULL, incl, lo);
                const int len_col = __shfl_sync(ORBX_FULL, len, lo);
                bool ok = false;
                int idx = 0;
                if (j < T) {
                    idx = __ldg(items + s_col + (j - (incl_col - len_col)));
                    const orbx_keypoint kp = keys[idx];
                    ok = true;
                    if (check_levels) {
                        if (kp.octave < q.min_level) ok = false;
                        if (q.max_level >= 0 && kp.octave > q.max_level) ok = false;
                    }
                    const float dx = kp.x - q.x, dy = kp.y - q.y;
                    if (!(fabsf(dx) < r && fabsf(dy) < r)) ok = false;
                }
                const uint32_t m = __ballot_sync(ORBX_FULL, ok);
                const int pos = base + total + __popc(m & lt);
                if (FILL && ok && pos < cap) {   // cap: the lists were sized before the counts were known
                    const uint32_t hd = (uint32_t)hamming256(a0, a1, tdesc + 2 * (size_t)idx);
                    if (PAIR) {   // {candidate, distance} as one 8-byte record (`cand` is then a uint2 array)
                        reinterpret_cast<uint2*>(cand)[pos] = make_uint2((uint32_t)idx, hd);
                    } else {
                        cand[pos] = idx;
                        dist[pos] = (uint16_t)hd;
                    }
                }
                total += __popc(m);
            }
        }
    }
    return total;
}

template <bool FILL>
__global__ void __launch_bounds__(256)
k_area(const orbx_keypoint* __restrict__ keys, const int32_t* __restrict__ start,
       const int32_t* __restrict__ items, OrbmGrid G, const orbm_area_query* __restrict__ Q, int nq,
       int32_t* __restrict__ cnt, const int32_t* __restrict__ row_ptr, const uint4* __restrict__ qdesc,
       const uint4* __restrict__ tdesc, int32_t* __restrict__ cand, uint16_t* __restrict__ dist, int cap) {
    const int qi = blockIdx.x * (blockDim.x >> 5) + orbx_warp();
    if (qi >= nq) return;
    const orbm_area_query q = Q[qi];
    const int total = area_walk<FILL>(keys, start, items, G, q, FILL ? row_ptr[qi] : 0, qdesc, tdesc, cand, dist, cap);
    if (!FILL && orbx_lane() == 0) cnt[qi] = total;
}

// Count and fill in ONE launch (the two tracking searches): a warp counts its query's candidates, reserves their span
// of cand / dist with one atomicAdd on a running counter and fills it.  Rows then sit in the lists in the order the
// warps got there -- row_info[q] = {first position, count} instead of a CSR row pointer -- but a row's own candidates
// keep the reference's order, which is all the accept loops depend on.  The counter is never reset: `counter0` is its
// value before this launch (the host reads it back with the results), positions are relative to it.
__global__ void __launch_bounds__(256)
k_area_onepass(const orbx_keypoint* __restrict__ keys, const int32_t* __restrict__ start,
               const int32_t* __restrict__ items, OrbmGrid G, const orbm_area_query* __restrict__ Q, int nq,
               uint2* __restrict__ row_info, uint32_t* __restrict__ counter, uint32_t counter0,
               const uint4* __restrict__ qdesc, const uint4* __restrict__ tdesc, uint2* __restrict__ pairs, int cap) {
    // (row_info, the counter and the {candidate, distance} records live in ONE buffer: a single download brings them)
    const int qi = blockIdx.x * (blockDim.x >> 5) + orbx_warp();
    if (qi >= nq) return;
    const orbm_area_query q = Q[qi];
    int32_t* cand = reinterpret_cast<int32_t*>(pairs);
    uint16_t* dist = nullptr;
    const int total = area_walk<false>(keys, start, items, G, q, 0, qdesc, tdesc, cand, dist, cap);
    uint32_t base = 0;
    if (total) {   // (warp-uniform)
        if (orbx_lane() == 0) base = atomicAdd(counter, (uint32_t)total) - counter0;
        base = __shfl_sync(ORBX_FULL, base, 0);
        // a span that would pass 2^31 is not written (cap < 2^31 drops it anyway; the host sees the total and reruns)
        if (base < 0x7fffffffu - (uint32_t)total)
            area_walk<true, true>(keys, start, items, G, q, (int)base, qdesc, tdesc, cand, dist, cap);
    }
    if (orbx_lane() == 0) row_info[qi] = make_uint2(base, (uint32_t)total);
}

// row_ptr[0..n] = exclusive scan of cnt[0..n) (one CTA)
__global__ void __launch_bounds__(1024) k_rowptr_scan(const int32_t* __restrict__ cnt, int n, int32_t* __restrict__ row_ptr) {
    __shared__ uint32_t part[1024];
    __shared__ uint32_t tmp[64];
    const int tid = threadIdx.x;
    const int per = (n + 1023) / 1024;
    const int b = min(tid * per, n), e = min(b + per, n);
    uint32_t s = 0;
    for (int i = b; i < e; ++i) s += (uint32_t)cnt[i];
    part[tid] = s;
    __syncthreads();
    const uint32_t total = block_excl_scan(part, 1024, tmp);
    uint32_t run = part[tid];
    for (int i = b; i < e; ++i) {
        row_ptr[i] = (int32_t)run;
        run += (uint32_t)cnt[i];
    }
    if (tid == 0) row_ptr[n] = (int32_t)total;
}

cudaError_t orbm_launch_grid_build(const orbx_keypoint* keys, int n, float min_x, float min_y, float w_inv,
                                   float h_inv, int32_t* start, int32_t* items, int32_t* cellof, cudaStream_t st) {
    OrbmGrid G{min_x, min_y, w_inv, h_inv};
    k_grid_build<<<1, 1024, 0, st>>>(keys, n, G, start, items, cellof);
    return cudaGetLastError();
}

cudaError_t orbm_launch_area_count(const orbx_keypoint* keys, const int32_t* start, const int32_t* items,
                                   float min_x, float min_y, float w_inv, float h_inv, const orbm_area_query* Q,
                                   int nq, int32_t* cnt, int32_t* row_ptr, cudaStream_t st) {
    OrbmGrid G{min_x, min_y, w_inv, h_inv};
    k_area<false><<<(nq + 7) / 8, 256, 0, st>>>(keys, start, items, G, Q, nq, cnt, nullptr, nullptr, nullptr,
                                                 nullptr, nullptr, 0);
    k_rowptr_scan<<<1, 1024, 0, st>>>(cnt, nq, row_ptr);
    return cudaGetLastError();
}

cudaError_t orbm_launch_area_fill(const orbx_keypoint* keys, const int32_t* start, const int32_t* items,
                                  float min_x, float min_y, float w_inv, float h_inv, const orbm_area_query* Q,
                                  int nq, const int32_t* row_ptr, const uint8_t* qdesc, const uint8_t* tdesc,
                                  int32_t* cand, uint16_t* dist, int cap, cudaStream_t st) {
    OrbmGrid G{min_x, min_y, w_inv, h_inv};
    k_area<true><<<(nq + 7) / 8, 256, 0, st>>>(keys, start, items, G, Q, nq, nullptr, row_ptr,
                                                reinterpret_cast<const uint4*>(qdesc),
                                                reinterpret_cast<const uint4*>(tdesc), cand, dist, cap);
    return cudaGetLastError();
}

cudaError_t orbm_launch_area_onepass(const orbx_keypoint* keys, const int32_t* start, const int32_t* items,
                                     float min_x, float min_y, float w_inv, float h_inv, const orbm_area_query* Q,
                                     int nq, uint2* row_info, uint32_t* counter, uint32_t counter0, const uint8_t* qdesc,
                                     const uint8_t* tdesc, uint2* pairs, int cap, cudaStream_t st) {
    OrbmGrid G{min_x, min_y, w_inv, h_inv};
    k_area_onepass<<<(nq + 7) / 8, 256, 0, st>>>(keys, start, items, G, Q, nq, row_info, counter, counter0,
                                                  reinterpret_cast<const uint4*>(qdesc),
                                                  reinterpret_cast<const uint4*>(tdesc), pairs, cap);
    return cudaGetLastError();
}

// ---- measurement helper: the device's POPC issue rate (the bound SURVEY 8(d) names for M1-M6) ----
// Every thread runs 8 independent xor -> popc -> add chains; the result is written so that nothing
// is optimised away.  Returns popc instructions per launch through *n_popc.
__global__ void __launch_bounds__(256) k_popc_peak(uint32_t* __restrict__ out, int iters) {
    uint32_t x[8], acc[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        x[k] = threadIdx.x * 2654435761u + k * 40503u + blockIdx.x;
        acc[k] = 0;
    }
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            acc[k] += __popc(x[k] ^ acc[k]);
        }
    }
    uint32_t s = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) s += acc[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

cudaError_t orbm_launch_popc_peak(uint32_t* out, int blocks, int iters, cudaStream_t st) {
    k_popc_peak<<<blocks, 256, 0, st>>>(out, iters);
    return cudaGetLastError();
}

// ---- N3: MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:330-404), batched --------------
// A map point is a group of N observed descriptors.  The median of row i is the k-th smallest
// (k = (int)(0.5 * (N - 1)), the reference's index into the sorted row, which includes d(i,i) = 0) of
// d(i, .).  The host cuts every group into work items of at most DIST_ITEM_ROWS rows (N <= 256) or 32
// rows (larger groups), one warp per item, so that a 150-observation point does not leave one warp
// running long after the two-observation points have finished.  N <= 256: the warp takes one row at a
// time, lane j holds d(i, j), d(i, j + 32), .. in registers (each distance computed once), and the
// k-th smallest comes from a 9-step binary search on the value (distances are integers in [0, 256])
// with one vote or REDUX per step.  Larger groups: lane r takes row row0 + r and recomputes it per
// search step from L1 (no N x N matrix is ever stored).  Every item folds (median << 20 | row) into
// key[group] with atomicMin: the first row with the least median wins, as the reference's strict '<'
// scan does.  key[] is preset to 0xffffffff, which also reads as the -1 of an empty group.
#ifndef BOW_TOPK
#define BOW_TOPK 4
#endif
#define DIST_CHUNKS 8
#ifndef DIST_ITEM_ROWS
#define DIST_ITEM_ROWS 8
#endif
// NCH is a template parameter so that d[] stays in registers and the usual map point (a few to a few
// tens of observations: NCH = 1) pays one compare + one vote per search step instead of eight compares
// and a REDUX.
template <int NCH>
__device__ __forceinline__ uint32_t distinctive_rows(const uint4* __restrict__ D, int N, int k, int lane, int r0,
                                                     int r1) {
    uint32_t best = 0xffffffffu;
    for (int i = r0; i < r1; ++i) {
        const uint4 a0 = __ldg(D + 2 * i), a1 = __ldg(D + 2 * i + 1);
        int d[NCH];
#pragma unroll
        for (int c = 0; c < NCH; ++c) {
            const int j = c * 32 + lane;
            d[c] = j < N ? hamming256(a0, a1, D + 2 * j) : 0x7fff;
        }
        int lo = 0, hi = 256;
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            int cnt;
            if (NCH == 1) {
                cnt = __popc(__ballot_sync(ORBX_FULL, d[0] <= mid));
            } else {
                cnt = 0;
#pragma unroll
                for (int c = 0; c < NCH; ++c) cnt += d[c] <= mid ? 1 : 0;
                cnt = __reduce_add_sync(ORBX_FULL, cnt);
            }
            if (cnt >= k + 1) hi = mid; else lo = mid + 1;
        }
        best = min(best, ((uint32_t)lo << 20) | (uint32_t)i);
    }
    return best;
}

__global__ void __launch_bounds__(128)
k_distinctive(const uint4* __restrict__ desc, const int32_t* __restrict__ group_ptr, const int2* __restrict__ items,
              int nitems, uint32_t* __restrict__ key_out) {
    const int it = blockIdx.x * (blockDim.x >> 5) + orbx_warp();
    if (it >= nitems) return;
    const int lane = orbx_lane();
    const int2 item = __ldg(items + it);   // x = group, y = first row of the item
    const int g = item.x, r0 = item.y;
    const int base = __ldg(group_ptr + g), N = __ldg(group_ptr + g + 1) - base;
    const uint4* D = desc + 2 * (size_t)base;
    const int k = (int)(0.5 * (double)(N - 1));
    uint32_t key;
    if (N <= 32 * DIST_CHUNKS) {
        const int r1 = min(N, r0 + DIST_ITEM_ROWS);
        if (N <= 32) key = distinctive_rows<1>(D, N, k, lane, r0, r1);
        else if (N <= 64) key = distinctive_rows<2>(D, N, k, lane, r0, r1);
        else if (N <= 128) key = distinctive_rows<4>(D, N, k, lane, r0, r1);
        else key = distinctive_rows<DIST_CHUNKS>(D, N, k, lane, r0, r1);
    } else {
        key = 0xffffffffu;
        const int i = r0 + lane;
        if (i < N) {
            const uint4 a0 = __ldg(D + 2 * i), a1 = __ldg(D + 2 * i + 1);
            // smallest v with #{j : d(i,j) <= v} >= k + 1
            int lo = 0, hi = 256;
            while (lo < hi) {
                const int mid = (lo + hi) >> 1;
                int cnt = 0;
                for (int j = 0; j < N; ++j) cnt += hamming256(a0, a1, D + 2 * j) <= mid ? 1 : 0;
                if (cnt >= k + 1) hi = mid; else lo = mid + 1;
            }
            key = ((uint32_t)lo << 20) | (uint32_t)i;
        }
        key = __reduce_min_sync(ORBX_FULL, key);
    }
    if (lane == 0) atomicMin(key_out + g, key);
}

int orbm_distinctive_item_rows() { return DIST_ITEM_ROWS; }

cudaError_t orbm_launch_distinctive(const uint8_t* desc, const int32_t* group_ptr, const int32_t* items, int nitems,
                                    uint32_t* key, cudaStream_t st) {
    if (nitems <= 0) return cudaSuccess;
    k_distinctive<<<(nitems + 3) / 4, 128, 0, st>>>(reinterpret_cast<const uint4*>(desc), group_ptr,
                                                    reinterpret_cast<const int2*>(items), nitems, key);
    return cudaGetLastError();
}

// ---- SearchByBoW (both overloads), SearchForTriangulation: per-row distances + the K best candidates ----
// Row r = one feature of side A inside a vocabulary node both sides share; its candidates are side B's
// features of that node, read in place from B's node_items (no expanded candidate list is uploaded).
// rows[r] = {feature of A, first candidate in items[], candidate count, first slot in dist[]}.  One
// warp per row: every distance is written to dist[] (the host's full-scan fallback), and the K
// smallest keys (distance << 20 | rank) of the candidates that pass the static gate and max_dist go to
// topk[r * K ..], ascending, 0xffffffff = none.  rank = position in the row (SearchByBoW: strict '<'
// scans, the first of equal distances wins) or len - 1 - position (SearchForTriangulation: its
// 'dist > bestDist -> skip' scan lets the last of equal distances win).  Ascending keys are the order in
// which the reference's scan would prefer the candidates, so the host replays a row from these few
// keys unless they are all refused (matched by earlier rows / failing the epipolar test).
__global__ void __launch_bounds__(128)
k_bow_topk(const uint4* __restrict__ A, const uint4* __restrict__ B, const uint32_t* __restrict__ items,
           const int4* __restrict__ rows, int nrows, const uint8_t* __restrict__ gate_b, int K, int max_dist,
           int last_wins, uint16_t* __restrict__ dist, uint32_t* __restrict__ topk) {
    const int r = blockIdx.x * (blockDim.x >> 5) + orbx_warp();
    if (r >= nrows) return;
    const int lane = orbx_lane();
    const int4 row = __ldg(rows + r);
    const int len = row.z;
    const uint32_t* it = items + row.y;
    uint16_t* drow = dist + row.w;
    const uint4 a0 = __ldg(A + 2 * (size_t)row.x), a1 = __ldg(A + 2 * (size_t)row.x + 1);
    uint32_t mine = 0xffffffffu;   // key of position `lane` (rows of at most 32 candidates: the usual case)
    for (int pos = lane; pos < len; pos += 32) {
        const uint32_t f = __ldg(it + pos);
        const int d = hamming256(a0, a1, B + 2 * (size_t)f);
        drow[pos] = (uint16_t)d;
        if (pos < 32 && d <= max_dist && (!gate_b || __ldg(gate_b + f)))
            mine = ((uint32_t)d << 20) | (uint32_t)(last_wins ? len - 1 - pos : pos);
    }
    uint32_t prev = 0;
    for (int k = 0; k < K; ++k) {
        uint32_t m = mine;
        if (len > 32) {   // long row: this lane's least key above the last one taken, from its own dist[] writes
            m = 0xffffffffu;
            for (int pos = lane; pos < len; pos += 32) {
                const int d = drow[pos];
                const uint32_t key = ((uint32_t)d << 20) | (uint32_t)(last_wins ? len - 1 - pos : pos);
                if ((k == 0 || key > prev) && key < m && d <= max_dist && (!gate_b || __ldg(gate_b + __ldg(it + pos)))) m = key;
            }
        }
        const uint32_t sel = __reduce_min_sync(ORBX_FULL, m);
        if (lane == 0) topk[(size_t)r * K + k] = sel;
        if (mine == sel) mine = 0xffffffffu;
        prev = sel;
        if (sel == 0xffffffffu) {
            if (lane == 0)
                for (int j = k + 1; j < K; ++j) topk[(size_t)r * K + j] = 0xffffffffu;
            break;
        }
    }
}

int orbm_bow_topk_k() { return BOW_TOPK; }

cudaError_t orbm_launch_bow_topk(const uint8_t* A, const uint8_t* B, const uint32_t* items, const int32_t* rows, int nrows,
                                 const uint8_t* gate_b, int K, int max_dist, int last_wins, uint16_t* dist,
                                 uint32_t* topk, cudaStream_t st) {
    if (nrows <= 0) return cudaSuccess;
    k_bow_topk<<<(nrows + 3) / 4, 128, 0, st>>>(reinterpret_cast<const uint4*>(A), reinterpret_cast<const uint4*>(B), items,
                                                reinterpret_cast<const int4*>(rows), nrows, gate_b, K, max_dist,
                                                last_wins, dist, topk);
    return cudaGetLastError();
}

// ---- window searches: the K best candidates of every CSR row -----------------------------------------
// After k_area<1> has written a row's candidates (GetFeaturesInArea order) and distances, one warp per
// row returns the K smallest (distance << 20 | position) keys, ascending, and the candidate each one
// names (0xffffffff / -1 = none): the order in which a strict '<' best / second-best scan prefers the
// candidates, so the host's accept loop reads K entries per row instead of the whole row.
__global__ void __launch_bounds__(128)
k_csr_topk(const int32_t* __restrict__ row_ptr, const int32_t* __restrict__ cand, const uint16_t* __restrict__ dist,
           int nrows, int K, int cap, uint32_t* __restrict__ keys, int32_t* __restrict__ key_cand) {
    const int r = blockIdx.x * (blockDim.x >> 5) + orbx_warp();
    if (r >= nrows || __ldg(row_ptr + nrows) > cap) return;   // lists cut at cap: the host reruns with larger ones
    const int lane = orbx_lane();
    const int b = __ldg(row_ptr + r), len = __ldg(row_ptr + r + 1) - b;
    uint32_t mine = 0xffffffffu;
    if (lane < len) mine = ((uint32_t)__ldg(dist + b + lane) << 20) | (uint32_t)lane;
    uint32_t prev = 0;
    for (int k = 0; k < K; ++k) {
        uint32_t m = mine;
        if (len > 32) {
            m = 0xffffffffu;
            for (int pos = lane; pos < len; pos += 32) {
                const uint32_t key = ((uint32_t)__ldg(dist + b + pos) << 20) | (uint32_t)pos;
                if ((k == 0 || key > prev) && key < m) m = key;
            }
        }
        const uint32_t sel = __reduce_min_sync(ORBX_FULL, m);
        if (mine == sel) mine = 0xffffffffu;
        prev = sel;
        if (lane == 0) {
            keys[(size_t)r * K + k] = sel;
            key_cand[(size_t)r * K + k] = sel == 0xffffffffu ? -1 : __ldg(cand + b + (int)(sel & 0xfffffu));
        }
        if (sel == 0xffffffffu) {
            if (lane == 0)
                for (int j = k + 1; j < K; ++j) {
                    keys[(size_t)r * K + j] = 0xffffffffu;
                    key_cand[(size_t)r * K + j] = -1;
                }
            break;
        }
    }
}

cudaError_t orbm_launch_csr_topk(const int32_t* row_ptr, const int32_t* cand, const uint16_t* dist, int nrows, int K,
                                 int cap, uint32_t* keys, int32_t* key_cand, cudaStream_t st) {
    if (nrows <= 0) return cudaSuccess;
    k_csr_topk<<<(nrows + 3) / 4, 128, 0, st>>>(row_ptr, cand, dist, nrows, K, cap, keys, key_cand);
    return cudaGetLastError();
}

// ---- upload scatter -------------------------------------------------------------------------------
// A matcher call's inputs arrive as ONE copy of the context's pinned arena into its device mirror; this
// kernel hands the segments to the buffers the search kernels read (orbm_api.cu, ctx_flush): one DMA
// and one launch instead of a copy-engine operation per array.
__global__ void __launch_bounds__(256)
k_scatter(const __grid_constant__ OrbmScatter S) {
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, nth = (size_t)gridDim.x * blockDim.x;
    for (int k = 0; k < S.n; ++k) {
        const uint4* a = reinterpret_cast<const uint4*>(S.seg[k].src);
        uint4* d = reinterpret_cast<uint4*>(S.seg[k].dst);
        const size_t n16 = S.seg[k].bytes >> 4;
        for (size_t i = tid; i < n16; i += nth) d[i] = a[i];
        const uint32_t tail = S.seg[k].bytes & 15u;
        if (tid < tail)
            reinterpret_cast<uint8_t*>(S.seg[k].dst)[16 * n16 + tid] = reinterpret_cast<const uint8_t*>(S.seg[k].src)[16 * n16 + tid];
    }
}

cudaError_t orbm_launch_scatter(const OrbmScatter& S, cudaStream_t st) {
    if (S.n <= 0) return cudaSuccess;
    size_t most = 0;
    for (int k = 0; k < S.n; ++k) most = most > S.seg[k].bytes ? most : S.seg[k].bytes;
    const int blocks = (int)((most / 16 + 255) / 256);
    k_scatter<<<blocks < 1 ? 1 : (blocks > 296 ? 296 : blocks), 256, 0, st>>>(S);
    return cudaGetLastError();
}

// ---- N4 (second half): DBoW2 TemplatedVocabulary<FORB>::transform, the tree descent ---------------
// (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1230-1275, called per feature by transform(features,
// v, fv, levelsup) :1139-1217 from Frame::ComputeBoW, src/Frame.cc:729-736).  One warp per feature:
// at every level the lanes take the node's children (k = 10 for ORBvoc: one round), FORB::distance =
// the same 256-bit popcount, and the first child with the least distance wins (packed key
// distance << 16 | position in Node::children, warp min = the reference's strict '<' scan).  The
// vocabulary (child lists + 32-byte node descriptors, ~36 MB for the 10^6-word ORBvoc) stays resident in
// HBM/L2.  word_out / weight_out[i] = id and weight of the word reached, nid[i] = the node passed at level L - levelsup (0 when that level is
// <= 0 or the branch ends above it).
__global__ void __launch_bounds__(128)
k_bow_descend(const uint4* __restrict__ feat, int n, const int32_t* __restrict__ child_ptr,
              const uint32_t* __restrict__ child_ids, const uint4* __restrict__ node_desc,
              const uint32_t* __restrict__ word_id, const double* __restrict__ weight, int nid_level,
              uint32_t* __restrict__ word_out, double* __restrict__ weight_out, int32_t* __restrict__ nid_out) {
    const int i = blockIdx.x * (blockDim.x >> 5) + orbx_warp();
    if (i >= n) return;
    const int lane = orbx_lane();
    const uint4 a0 = __ldg(feat + 2 * i), a1 = __ldg(feat + 2 * i + 1);
    uint32_t node = 0, nid = 0;
    int level = 0;
    for (;;) {
        const int cp = __ldg(child_ptr + node), ce = __ldg(child_ptr + node + 1);
        if (ce <= cp) break;   // leaf: a word
        ++level;
        uint32_t key = 0xffffffffu;
        for (int r = lane; r < ce - cp; r += 32) {
            const uint32_t id = __ldg(child_ids + cp + r);
            key = min(key, ((uint32_t)hamming256(a0, a1, node_desc + 2 * (size_t)id) << 16) | (uint32_t)r);
        }
        key = __reduce_min_sync(ORBX_FULL, key);
        node = __ldg(child_ids + cp + (int)(key & 0xffffu));
        if (level == nid_level) nid = node;
    }
    if (lane == 0) {   // the word's id and weight travel with the result: the host never indexes the 10^6-entry tables
        word_out[i] = __ldg(word_id + node);
        weight_out[i] = __ldg(weight + node);
        nid_out[i] = (int32_t)nid;
    }
}

cudaError_t orbm_launch_bow_descend(const uint8_t* feat, int n, const int32_t* child_ptr, const uint32_t* child_ids,
                                    const uint8_t* node_desc, const uint32_t* word_id, const double* weight,
                                    int nid_level, uint32_t* word_out, double* weight_out, int32_t* nid,
                                    cudaStream_t st) {
    if (n <= 0) return cudaSuccess;
    k_bow_descend<<<(n + 3) / 4, 128, 0, st>>>(reinterpret_cast<const uint4*>(feat), n, child_ptr, child_ids,
                                                reinterpret_cast<const uint4*>(node_desc), word_id, weight, nid_level,
                                                word_out, weight_out, nid);
    return cudaGetLastError();
}

// ---- N2: Frame::UndistortKeyPoints + Frame::ComputeImageBounds on the device (src/Frame.cc:738-800) --------
// cv::undistortPoints(mat, mat, K, distCoef, Mat(), K) with its default criteria: all arithmetic in double, five
// fixed-point iterations of the inverse of (k1, k2, p1, p2, k3), then P = K (modules/calib3d undistort.dispatch.cpp
// cvUndistortPointsInternal; restated and pinned bit-for-bit against cv2.undistortPoints in oracle/cvshim).  Every
// operation is an explicit IEEE round-to-nearest intrinsic: a contracted multiply-add would round differently.
// Thread i < n: keypoint i (pt replaced in the output record, the other fields copied); threads n .. n+3: the image
// corners (0,0), (w,0), (0,h), (w,h) into corners_xy (ComputeImageBounds undistorts them the same way).
struct OrbmCam {
    double fx, fy, cx, cy, k1, k2, p1, p2, k3;
};

__device__ __forceinline__ void undistort_point(const OrbmCam& C, double u, double v, float* xo, float* yo) {
    const double ifx = __ddiv_rn(1.0, C.fx), ify = __ddiv_rn(1.0, C.fy);
    double x = __dmul_rn(__dsub_rn(u, C.cx), ifx), y = __dmul_rn(__dsub_rn(v, C.cy), ify);
    const double x0 = x, y0 = y;
    for (int j = 0; j < 5; ++j) {
        const double r2 = __dadd_rn(__dmul_rn(x, x), __dmul_rn(y, y));
        // icdist = (1 + ((k[7]*r2 + k[6])*r2 + k[5])*r2) / (1 + ((k[4]*r2 + k[1])*r2 + k[0])*r2), k[5..7] = 0
        const double num = 1.0;   // the rational model's numerator coefficients are zero
        const double den = __dadd_rn(1.0, __dmul_rn(__dadd_rn(__dmul_rn(__dadd_rn(__dmul_rn(C.k3, r2), C.k2), r2), C.k1), r2));
        const double icdist = __ddiv_rn(num, den);
        if (icdist < 0) {
            x = x0;
            y = y0;
            break;
        }
        // deltaX = 2*p1*x*y + p2*(r2 + 2*x*x) [+ 0*r2 + 0*r2*r2], evaluated left to right as the C expression
        const double dX = __dadd_rn(__dmul_rn(__dmul_rn(__dmul_rn(2.0, C.p1), x), y),
                                    __dmul_rn(C.p2, __dadd_rn(r2, __dmul_rn(__dmul_rn(2.0, x), x))));
        const double dY = __dadd_rn(__dmul_rn(C.p1, __dadd_rn(r2, __dmul_rn(__dmul_rn(2.0, y), y))),
                                    __dmul_rn(__dmul_rn(__dmul_rn(2.0, C.p2), x), y));
        x = __dmul_rn(__dsub_rn(x0, dX), icdist);
        y = __dmul_rn(__dsub_rn(y0, dY), icdist);
    }
    // P = K: xx = fx*x + 0*y + cx, yy = 0*x + fy*y + cy, ww = 1 / (0*x + 0*y + 1)
    const double xx = __dadd_rn(__dadd_rn(__dmul_rn(C.fx, x), __dmul_rn(0.0, y)), C.cx);
    const double yy = __dadd_rn(__dadd_rn(__dmul_rn(0.0, x), __dmul_rn(C.fy, y)), C.cy);
    const double ww = __ddiv_rn(1.0, __dadd_rn(__dadd_rn(__dmul_rn(0.0, x), __dmul_rn(0.0, y)), 1.0));
    *xo = (float)__dmul_rn(xx, ww);
    *yo = (float)__dmul_rn(yy, ww);
}

__global__ void __launch_bounds__(128)
k_undistort(const orbx_keypoint* __restrict__ keys, int n, OrbmCam C, float width, float height,
            orbx_keypoint* __restrict__ keys_un, float* __restrict__ corners_xy) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        orbx_keypoint k = keys[i];
        undistort_point(C, (double)k.x, (double)k.y, &k.x, &k.y);
        keys_un[i] = k;
    } else if (i < n + 4) {
        const int c = i - n;
        undistort_point(C, (c & 1) ? (double)width : 0.0, (c & 2) ? (double)height : 0.0, corners_xy + 2 * c,
                        corners_xy + 2 * c + 1);
    }
}

cudaError_t orbm_launch_undistort(const orbx_keypoint* keys, int n, const float* cam9, float width, float height,
                                  orbx_keypoint* keys_un, float* corners_xy, cudaStream_t st) {
    OrbmCam C{cam9[0], cam9[1], cam9[2], cam9[3], cam9[4], cam9[5], cam9[6], cam9[7], cam9[8]};
    k_undistort<<<(n + 4 + 127) / 128, 128, 0, st>>>(keys, n, C, width, height, keys_un, corners_xy);
    return cudaGetLastError();
}

// ---- orbm_area_candidates_static: the window search with its scratch in a static device array ------
// No counterpart in the reference; added for callers that cannot
// call cudaMalloc (CUDA graph capture, serving loops, harnesses that forbid it).  Same kernels as
// orbm_area_candidates (k_grid_build -> k_area<count> -> k_rowptr_scan -> k_area<fill>); what that entry
// point keeps in cudaMalloc'ed context buffers -- the cell of every keypoint, the 64 x 48 grid (start
// offsets + item array) and the per-query candidate counts -- lives in the file-scope array below.
// Every GPU that loads the library holds these 274,688 bytes until the process exits.
#define ORBM_STATIC_START_INTS 3136   /* 64 * 48 + 1 = 3073 offsets, padded to a multiple of 64 ints */
#define ORBM_STATIC_BYTES ((2 * ORBM_STATIC_MAX_KEYS + ORBM_STATIC_START_INTS + ORBM_STATIC_MAX_QUERIES) * 4)
static __device__ __align__(256) uint8_t g_orbm_static_scratch[ORBM_STATIC_BYTES];

#include <mutex>
int orbx_set_error(int code, const char* fmt, ...);   // orbx_api.cu

namespace {
const int kStaticMaxDevices = 64;
std::mutex g_static_mu;
uint8_t* g_static_ptr[kStaticMaxDevices];   // device address of g_orbm_static_scratch, per device (0 = not looked up)
}  // namespace

extern "C" int orbm_area_candidates_static(int device, void* stream, const orbx_keypoint* d_keys,
                                           const uint8_t* d_desc, int n, float min_x, float min_y,
                                           float grid_w_inv, float grid_h_inv, const orbm_area_query* d_q, int nq,
                                           const uint8_t* d_q_desc, int32_t* d_row_ptr, int32_t* d_cand,
                                           uint16_t* d_dist, int cap) {
    if (device < 0 || device >= kStaticMaxDevices || n < 0 || nq < 0 || cap < 0)
        return orbx_set_error(ORBX_E_BAD_ARG, "orbm_area_candidates_static: bad argument");
    if (n > ORBM_STATIC_MAX_KEYS)
        return orbx_set_error(ORBX_E_TOO_LARGE, "orbm_area_candidates_static: %d keypoints > ORBM_STATIC_MAX_KEYS (%d)",
                              n, ORBM_STATIC_MAX_KEYS);
    if (nq > ORBM_STATIC_MAX_QUERIES)
        return orbx_set_error(ORBX_E_TOO_LARGE, "orbm_area_candidates_static: %d queries > ORBM_STATIC_MAX_QUERIES (%d)",
                              nq, ORBM_STATIC_MAX_QUERIES);
    if (!d_row_ptr || (n && (!d_keys || !d_desc)) || (nq && (!d_q || !d_q_desc)) || (cap && (!d_cand || !d_dist)))
        return orbx_set_error(ORBX_E_BAD_ARG, "orbm_area_candidates_static: null argument");
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess)
        return orbx_set_error(ORBX_E_CUDA, "cudaSetDevice(%d) failed: %s", device, cudaGetErrorString(e));
    uint8_t* base;
    {
        // The array's name is a host-side stand-in in host code, not a device address: look the address up once
        // per device, with that device current, and hand THAT to the kernels.
        std::lock_guard<std::mutex> lk(g_static_mu);
        if (!g_static_ptr[device]) {
            void* p = nullptr;
            e = cudaGetSymbolAddress(&p, g_orbm_static_scratch);
            if (e != cudaSuccess || !p)
                return orbx_set_error(ORBX_E_CUDA, "cudaGetSymbolAddress(g_orbm_static_scratch) failed on device %d: %s",
                                      device, cudaGetErrorString(e));
            g_static_ptr[device] = (uint8_t*)p;
        }
        base = g_static_ptr[device];
    }
    int32_t* cellof = (int32_t*)base;
    int32_t* start = cellof + ORBM_STATIC_MAX_KEYS;
    int32_t* items = start + ORBM_STATIC_START_INTS;
    int32_t* cnt = items + ORBM_STATIC_MAX_KEYS;
    cudaStream_t st = (cudaStream_t)stream;
    const OrbmGrid G{min_x, min_y, grid_w_inv, grid_h_inv};
    // Nothing in the scratch has to start at zero: k_grid_build writes all 3073 offsets and the items of
    // every keypoint inside the grid (its counters are shared memory, cleared by the kernel), k_area<count>
    // writes cnt[0..nq), and only those ranges are read back.  The previous call's contents beyond them stay.
    k_grid_build<<<1, 1024, 0, st>>>(d_keys, n, G, start, items, cellof);
    if (nq) {
        k_area<false><<<(nq + 7) / 8, 256, 0, st>>>(d_keys, start, items, G, d_q, nq, cnt, nullptr, nullptr, nullptr,
                                                     nullptr, nullptr, 0);
    }
    k_rowptr_scan<<<1, 1024, 0, st>>>(cnt, nq, d_row_ptr);
    if (nq) {
        k_area<true><<<(nq + 7) / 8, 256, 0, st>>>(d_keys, start, items, G, d_q, nq, nullptr, d_row_ptr,
                                                    reinterpret_cast<const uint4*>(d_q_desc),
                                                    reinterpret_cast<const uint4*>(d_desc), d_cand, d_dist, cap);
    }
    e = cudaGetLastError();
    if (e != cudaSuccess)
        return orbx_set_error(ORBX_E_CUDA, "orbm_area_candidates_static: launch failed: %s", cudaGetErrorString(e));
    return ORBX_OK;
}
