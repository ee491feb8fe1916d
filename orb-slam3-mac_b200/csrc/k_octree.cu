// k_octree.cu -- E3: ORBextractor::DistributeOctTree (src/ORBextractor.cc:537-761) and
// ExtractorNode::DivideNode (:479-535) as a deterministic GPU quadtree.
//
// One CTA owns one (frame, level).  The reference's std::list algorithm is restated as array
// passes (the formulation proven bit-identical to the verbatim reference by oracle/orb_oracle.cpp):
//
//   * the node list is an array in list order; a pass splits a set of nodes given in
//     PROCESSING order; children are created in order n1..n4 per parent, and because the
//     reference push_front()s them the new list is reverse(creation order) ++ untouched nodes;
//   * full pass (:597-661): processing order = list order of all multi-key nodes;
//   * careful pass (:671-735): processing order = multi-key children of the previous pass sorted
//     by (key count, creation order) DESCENDING -- creation order is this project's canonical
//     replacement for the reference's heap-address tie-break at :682 -- and the
//     pass stops at the first node after which size >= N (prefix scan + atomicMin);
//   * keys of a node stay contiguous: a split is a STABLE 4-way partition of the node's own key
//     range (ballot/popc ranks), ping-ponged between two buffers (shared memory when the level's
//     keys fit, HBM otherwise), so the reference's intra-node key order -- which decides
//     response ties in the final first-maximum-wins scan (:740-758) -- is preserved.
//
// Small nodes are partitioned one per warp, nodes with >= OCT_BIG keys by the whole CTA.
#include "orbx_device.cuh"

#ifndef OCT_THREADS
#define OCT_THREADS 256
#endif
#ifndef OCT_SMEM_KB
#define OCT_SMEM_KB 36
#endif
#define OCT_WARPS (OCT_THREADS / 32)
#ifndef OCT_BIG
#define OCT_BIG 1024
#endif
#define OCT_MAX_INI 8
#ifndef OCT_MINB
#define OCT_MINB 4
#endif

struct OctCtx {
    // node list, double buffered (index = list position)
    // (pairs are picked with a select, never indexed: a dynamically indexed pointer array would
    // put the whole context into local memory)
    short4 *nb0, *nb1;
    uint32_t *nst0, *nst1;  // key range start | buffer bit << 31
    uint32_t *ncn0, *ncn1;  // key count
    __device__ __forceinline__ short4* nb(int i) const { return i ? nb1 : nb0; }
    __device__ __forceinline__ uint32_t* nst(int i) const { return i ? nst1 : nst0; }
    __device__ __forceinline__ uint32_t* ncn(int i) const { return i ? ncn1 : ncn0; }
    __device__ __forceinline__ uint32_t* expand(int i) const { return i ? ex1 : ex0; }
    __device__ __forceinline__ uint32_t* buf(uint32_t i) const { return i ? buf1 : buf0; }
    uint32_t* cc;      // [4 * NC] child counts by processing index
    uint32_t* proc;    // [NC] list positions in processing order
    uint32_t *ex0, *ex1;  // [NC] each: multi-key children in creation order (full passes: creation
                          // indices, ping-ponged; careful passes: list positions in expand[0])
    uint32_t* done;    // [NC] key locations of the single-key nodes retired by the full passes
    uint32_t* aux;     // [NC + 1] packed scans
    uint32_t* flag;    // [NC + 1] processed flags / kept ranks
    unsigned long long* sortk;  // [NCP2]
    uint32_t* wc;      // [OCT_WARPS * 8] per-warp bin counts for CTA-wide partitions
    uint32_t* tmp;     // [64] scan scratch + scalars
    uint32_t *buf0, *buf1;  // key buffers
};

__device__ __forceinline__ int oct_quadrant(uint32_t key, int mx, int my) {
    return (ORBX_KEY_X(key) < mx ? 0 : 1) + (ORBX_KEY_Y(key) < my ? 0 : 2);
}

// Stable 4-way partition of one node by ONE warp.  Returns the four child counts (all lanes).
__device__ __forceinline__ void oct_warp_partition(const uint32_t* __restrict__ src,
                                                   uint32_t* __restrict__ dst, uint32_t start,
                                                   uint32_t cnt, int mx, int my, uint32_t c[4]) {
    const int lane = orbx_lane();
    const uint32_t lt = orbx_lanemask_lt();
    if (cnt <= 32) {
        // one key per lane: counts and stable ranks from the same four ballots
        uint32_t key = 0;
        int q = 4;
        if ((uint32_t)lane < cnt) {
            key = src[start + lane];
            q = oct_quadrant(key, mx, my);
        }
        uint32_t m[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            m[k] = __ballot_sync(ORBX_FULL, q == k);
            c[k] = __popc(m[k]);
        }
        if (q < 4) {
            const uint32_t base = start + (q > 0 ? c[0] : 0) + (q > 1 ? c[1] : 0) + (q > 2 ? c[2] : 0);
            const uint32_t mine = q == 0 ? m[0] : q == 1 ? m[1] : q == 2 ? m[2] : m[3];
            dst[base + __popc(mine & lt)] = key;
        }
        return;
    }
    c[0] = c[1] = c[2] = c[3] = 0;
    for (uint32_t b = 0; b < cnt; b += 32) {
        const uint32_t i = b + lane;
        const int q = i < cnt ? oct_quadrant(src[start + i], mx, my) : 4;
#pragma unroll
        for (int k = 0; k < 4; ++k) c[k] += __popc(__ballot_sync(ORBX_FULL, q == k));
    }
    uint32_t o[4] = {start, start + c[0], start + c[0] + c[1], start + c[0] + c[1] + c[2]};
    for (uint32_t b = 0; b < cnt; b += 32) {
        const uint32_t i = b + lane;
        uint32_t key = 0;
        int q = 4;
        if (i < cnt) {
            key = src[start + i];
            q = oct_quadrant(key, mx, my);
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const uint32_t m = __ballot_sync(ORBX_FULL, q == k);
            if (q == k) dst[o[k] + __popc(m & lt)] = key;
            o[k] += __popc(m);
        }
    }
}

// Stable partition of src[start, start+cnt) into `nbins` (<= 8) bins by the whole CTA.
// MODE 0: quadrant(mx, my); MODE 1: root index (int)(x / hx).  Writes the bin counts to
// out_counts[0..nbins) (shared memory).  All threads must call; contains __syncthreads.
template <int MODE>
__device__ __forceinline__ void oct_block_partition(const OctCtx& C, const uint32_t* __restrict__ src,
                                                    uint32_t* __restrict__ dst, uint32_t start,
                                                    uint32_t cnt, int nbins, int mx, int my, float hx,
                                                    uint32_t* out_counts) {
    const int lane = orbx_lane(), warp = orbx_warp();
    const uint32_t lt = orbx_lanemask_lt();
    const uint32_t per = ((cnt + OCT_WARPS * 32 - 1) / (OCT_WARPS * 32)) * 32;  // keys per warp
    const uint32_t b0 = min(cnt, warp * per), b1 = min(cnt, b0 + per);
    uint32_t c[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) c[k] = 0;
    for (uint32_t b = b0; b < b1; b += 32) {
        const uint32_t i = b + lane;
        int q = 8;
        if (i < b1) {
            const uint32_t key = src[start + i];
            q = MODE == 0 ? oct_quadrant(key, mx, my) : (int)((float)ORBX_KEY_X(key) / hx);
        }
#pragma unroll
        for (int k = 0; k < 8; ++k)
            if (k < nbins) c[k] += __popc(__ballot_sync(ORBX_FULL, q == k));
    }
    if (lane == 0) {
#pragma unroll
        for (int k = 0; k < 8; ++k) C.wc[warp * 8 + k] = c[k];
    }
    __syncthreads();
    // bin totals and this warp's base inside every bin
    uint32_t o[8];
    uint32_t run = start;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        uint32_t before = 0, tot = 0;
        if (k < nbins) {
            for (int w = 0; w < OCT_WARPS; ++w) {
                const uint32_t v = C.wc[w * 8 + k];
                if (w < warp) before += v;
                tot += v;
            }
        }
        o[k] = run + before;
        run += tot;
        if (threadIdx.x == 0 && k < nbins) out_counts[k] = tot;
    }
    for (uint32_t b = b0; b < b1; b += 32) {
        const uint32_t i = b + lane;
        uint32_t key = 0;
        int q = 8;
        if (i < b1) {
            key = src[start + i];
            q = MODE == 0 ? oct_quadrant(key, mx, my) : (int)((float)ORBX_KEY_X(key) / hx);
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            if (k < nbins) {
                const uint32_t m = __ballot_sync(ORBX_FULL, q == k);
                if (q == k) dst[o[k] + __popc(m & lt)] = key;
                o[k] += __popc(m);
            }
        }
    }
    __syncthreads();
}

// shared scalars inside C.tmp
#define OT_EPROC 40
#define OT_NBIG 41
#define OT_BIGNEXT 43

// Exclusive scan of a[0..n) by the whole CTA, one barrier when n <= OCT_THREADS.  Returns the total.
// The caller must put a barrier between this call and the next use of C.tmp[0..OCT_WARPS) or any
// read of a[] by a thread other than the one that owns the element (thread i owns a[i]).
__device__ __forceinline__ uint32_t oct_scan(uint32_t* a, int n, uint32_t* tmp) {
    if (n > OCT_THREADS) return block_excl_scan(a, n, tmp);
    const int t = threadIdx.x;
    const uint32_t v = t < n ? a[t] : 0u;
    const uint32_t incl = warp_incl_scan(v);
    if (orbx_lane() == 31) tmp[orbx_warp()] = incl;
    __syncthreads();
    uint32_t base = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < OCT_WARPS; ++w) {
        const uint32_t x = tmp[w];
        base += w < orbx_warp() ? x : 0u;
        tot += x;
    }
    if (t < n) a[t] = base + incl - v;
    return tot;
}

__device__ __forceinline__ uint32_t oct_child_code(const uint32_t c[4]) {
    uint32_t v = 0;
#pragma unroll
    for (int q = 0; q < 4; ++q) v += (c[q] > 0 ? 1u : 0u) + (c[q] > 1 ? 0x10000u : 0u);
    return v;
}

#ifndef OCT_SMALL
#define OCT_SMALL 8
#endif

// Stable 4-way partition of a node with <= OCT_SMALL keys by ONE thread: counts and running
// offsets are four byte lanes of one word.  Returns the packed counts.
__device__ __forceinline__ uint32_t oct_thread_partition(const uint32_t* __restrict__ src,
                                                         uint32_t* __restrict__ dst, uint32_t start,
                                                         uint32_t cnt, int mx, int my) {
    uint32_t cpack = 0;
    for (uint32_t t = 0; t < cnt; ++t) cpack += 1u << (8 * oct_quadrant(src[start + t], mx, my));
    uint32_t opack = (cpack << 8) + (cpack << 16) + (cpack << 24);   // exclusive prefix per byte lane
    for (uint32_t t = 0; t < cnt; ++t) {
        const uint32_t key = src[start + t];
        const int sh = 8 * oct_quadrant(key, mx, my);
        dst[start + ((opack >> sh) & 0xffu)] = key;
        opack += 1u << sh;
    }
    return cpack;
}

// Step (a) of a pass: split the E nodes order[j] (reversed: order[E-1-j]) of array set `cur` by a
// stable 4-way partition of their key ranges into the other key buffer; cc[4j..4j+3] = child
// counts, aux[j] = #children | #multi-key children << 16.  Nodes with >= OCT_BIG keys take the
// whole CTA, nodes with <= OCT_SMALL keys one thread, the rest one warp.  No trailing barrier.
__device__ __forceinline__ void oct_split_nodes(const OctCtx& C, int cur, int E, bool maybe_big,
                                                const uint32_t* order, bool reversed) {
    const int tid = threadIdx.x, lane = orbx_lane(), warp = orbx_warp();
    if (maybe_big) {
        for (int j = 0; j < E; ++j) {   // CTA-uniform
            const int i = order[reversed ? E - 1 - j : j];
            const uint32_t cnt = C.ncn(cur)[i];
            if (cnt < OCT_BIG) continue;
            const short4 bb = C.nb(cur)[i];
            const uint32_t st = C.nst(cur)[i];
            const int mx = bb.x + ((bb.y - bb.x + 1) >> 1), my = bb.z + ((bb.w - bb.z + 1) >> 1);
            oct_block_partition<0>(C, C.buf(st >> 31), C.buf((st >> 31) ^ 1), st & 0x7fffffffu, cnt, 4, mx,
                                   my, 0.f, &C.cc[4 * j]);
            if (tid == 0) C.aux[j] = oct_child_code(&C.cc[4 * j]);
        }
    }
    for (int j = warp; j < E; j += OCT_WARPS) {
        const int i = order[reversed ? E - 1 - j : j];
        const uint32_t cnt = C.ncn(cur)[i];
        if (cnt >= OCT_BIG || cnt <= OCT_SMALL) continue;
        const short4 bb = C.nb(cur)[i];
        const uint32_t st = C.nst(cur)[i];
        const int mx = bb.x + ((bb.y - bb.x + 1) >> 1), my = bb.z + ((bb.w - bb.z + 1) >> 1);
        uint32_t c[4];
        oct_warp_partition(C.buf(st >> 31), C.buf((st >> 31) ^ 1), st & 0x7fffffffu, cnt, mx, my, c);
        if (lane == 0) {
            C.cc[4 * j] = c[0]; C.cc[4 * j + 1] = c[1]; C.cc[4 * j + 2] = c[2]; C.cc[4 * j + 3] = c[3];
            C.aux[j] = oct_child_code(c);
        }
    }
    for (int j = tid; j < E; j += OCT_THREADS) {
        const int i = order[reversed ? E - 1 - j : j];
        const uint32_t cnt = C.ncn(cur)[i];
        if (cnt > OCT_SMALL) continue;
        const short4 bb = C.nb(cur)[i];
        const uint32_t st = C.nst(cur)[i];
        const int mx = bb.x + ((bb.y - bb.x + 1) >> 1), my = bb.z + ((bb.w - bb.z + 1) >> 1);
        const uint32_t cp = oct_thread_partition(C.buf(st >> 31), C.buf((st >> 31) ^ 1), st & 0x7fffffffu, cnt, mx, my);
        uint32_t c[4] = {cp & 0xffu, (cp >> 8) & 0xffu, (cp >> 16) & 0xffu, cp >> 24};
        C.cc[4 * j] = c[0]; C.cc[4 * j + 1] = c[1]; C.cc[4 * j + 2] = c[2]; C.cc[4 * j + 3] = c[3];
        C.aux[j] = oct_child_code(c);
    }
}

// Full pass (:597-661) in CREATION-ORDER form.  While only full passes run, the list is always
//     reverse(children of the last pass, creation order) ++ reverse(done[0..Dbefore))
// because the untouched nodes of a full pass are exactly the single-key ones, which the reference
// leaves in place behind the push_front()ed children.  So a pass needs no list positions at all:
// the children go to nb/nst/ncn[nxt] by creation index, the multi-key ones are listed in
// expand[nxt], the single-key ones are appended to done[].  Processing order = list order =
// expand[cur] backwards.  Three barriers per pass.
// In: E multi-key nodes of the previous pass.  Out: M children, *nm of them multi-key;
// done[D .. D + M - *nm) appended.
__device__ __forceinline__ int oct_full_pass(const OctCtx& C, int cur, int E, int D, bool maybe_big,
                                             int* nm) {
    const int nxt = cur ^ 1;
    const int tid = threadIdx.x, lane = orbx_lane(), warp = orbx_warp();
    const uint32_t* mul = C.expand(cur);
    oct_split_nodes(C, cur, E, maybe_big, mul, true);
    __syncthreads();
    if (tid == 0) C.tmp[OT_BIGNEXT] = 0;
    const uint32_t tot = oct_scan(C.aux, E, C.tmp);
    const int M = (int)(tot & 0xffffu);
    *nm = (int)(tot >> 16);
    if (E > OCT_THREADS) __syncthreads();
    for (int j = tid; j < E; j += OCT_THREADS) {
        const int i = mul[E - 1 - j];
        const short4 bb = C.nb(cur)[i];
        const uint32_t st = C.nst(cur)[i];
        const int mx = bb.x + ((bb.y - bb.x + 1) >> 1), my = bb.z + ((bb.w - bb.z + 1) >> 1);
        uint32_t run = (st & 0x7fffffffu);
        const uint32_t bit = (st & 0x80000000u) ^ 0x80000000u;
        int cidx = (int)(C.aux[j] & 0xffffu);
        int midx = (int)(C.aux[j] >> 16);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const uint32_t c = C.cc[4 * j + q];
            if (c > 0) {
                short4 nbv;
                nbv.x = (q & 1) ? mx : bb.x;
                nbv.y = (q & 1) ? bb.y : mx;
                nbv.z = (q & 2) ? my : bb.z;
                nbv.w = (q & 2) ? bb.w : my;
                C.nb(nxt)[cidx] = nbv;
                C.nst(nxt)[cidx] = run | bit;
                C.ncn(nxt)[cidx] = c;
                if (c > 1) {
                    C.expand(nxt)[midx++] = cidx;
                    if (c >= OCT_BIG) C.tmp[OT_BIGNEXT] = 1;
                } else {
                    C.done[D + cidx - midx] = run | bit;
                }
                ++cidx;
            }
            run += c;
        }
    }
    __syncthreads();
    return M;
}

// One pass over the nodes proc[0..E) of list `cur` (S nodes).  Builds list cur^1.
// careful == true applies the reference's early stop "size >= N" (:730-731).
// Returns the new list size; *n_multi = number of multi-key children created (nToExpand).
__device__ __forceinline__ int oct_pass(const OctCtx& C, int cur, int S, int E, bool careful, int N,
                                        int* n_multi, bool maybe_big) {
    const int nxt = cur ^ 1;
    const int tid = threadIdx.x, lane = orbx_lane(), warp = orbx_warp();

    // ---- (a) split every node of the pass: stable 4-way partition into the other buffer;
    //      aux[j] = #children | #multi-key children << 16
    if (tid == 0) C.tmp[OT_EPROC] = (uint32_t)E;
    oct_split_nodes(C, cur, E, maybe_big, C.proc, false);
    __syncthreads();

    // ---- (b) creation indices: exclusive scan; early stop "size >= N" of the careful pass
    const uint32_t tot = oct_scan(C.aux, E, C.tmp);
    if (tid == 0) C.aux[E] = tot;
    if (E > OCT_THREADS) __syncthreads();
    if (careful) {
        for (int j = tid; j < E; j += OCT_THREADS) {
            uint32_t nch = 0;
#pragma unroll
            for (int q = 0; q < 4; ++q) nch += C.cc[4 * j + q] > 0 ? 1u : 0u;
            const int size_after = S + (int)((C.aux[j] & 0xffffu) + nch) - (j + 1);
            if (size_after >= N) atomicMin(&C.tmp[OT_EPROC], (uint32_t)(j + 1));
        }
    }
    for (int i = tid; i < S; i += OCT_THREADS) C.flag[i] = 1;
    __syncthreads();
    const int Eproc = (int)C.tmp[OT_EPROC];
    const int M = (int)(C.aux[Eproc] & 0xffffu);
    *n_multi = (int)(C.aux[Eproc] >> 16);

    // ---- (c) rank of the untouched nodes
    for (int j = tid; j < Eproc; j += OCT_THREADS) C.flag[C.proc[j]] = 0;
    __syncthreads();
    const uint32_t kept = oct_scan(C.flag, S, C.tmp);
    if (tid == 0) C.flag[S] = kept;
    __syncthreads();

    // ---- (d) new list = reverse(creation order) ++ untouched nodes
    for (int j = tid; j < Eproc; j += OCT_THREADS) {
        const int i = C.proc[j];
        const short4 bb = C.nb(cur)[i];
        const uint32_t st = C.nst(cur)[i];
        const int mx = bb.x + ((bb.y - bb.x + 1) >> 1), my = bb.z + ((bb.w - bb.z + 1) >> 1);
        uint32_t run = (st & 0x7fffffffu);
        const uint32_t bit = (st & 0x80000000u) ^ 0x80000000u;
        int cidx = (int)(C.aux[j] & 0xffffu);
        int midx = (int)(C.aux[j] >> 16);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const uint32_t c = C.cc[4 * j + q];
            if (c > 0) {
                const int pos = M - 1 - cidx;
                short4 nbv;
                nbv.x = (q & 1) ? mx : bb.x;
                nbv.y = (q & 1) ? bb.y : mx;
                nbv.z = (q & 2) ? my : bb.z;
                nbv.w = (q & 2) ? bb.w : my;
                C.nb(nxt)[pos] = nbv;
                C.nst(nxt)[pos] = run | bit;
                C.ncn(nxt)[pos] = c;
                if (c > 1) C.expand(0)[midx++] = pos;
                ++cidx;
            }
            run += c;
        }
    }
    for (int i = tid; i < S; i += OCT_THREADS) {
        const uint32_t r = C.flag[i];
        if (C.flag[i + 1] != r) {
            const int pos = M + (int)r;
            C.nb(nxt)[pos] = C.nb(cur)[i];
            C.nst(nxt)[pos] = C.nst(cur)[i];
            C.ncn(nxt)[pos] = C.ncn(cur)[i];
        }
    }
    __syncthreads();
    return M + (int)kept;
}

// Descending bitonic sort of C.sortk[0..P2) (P2 power of two).
__device__ __forceinline__ void oct_bitonic_desc(unsigned long long* a, int P2) {
    for (int k = 2; k <= P2; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = threadIdx.x; i < P2; i += OCT_THREADS) {
                const int p = i ^ j;
                if (p > i) {
                    const unsigned long long x = a[i], y = a[p];
                    const bool desc = (i & k) == 0;
                    if (desc ? (x < y) : (x > y)) {
                        a[i] = y;
                        a[p] = x;
                    }
                }
            }
            __syncthreads();
        }
    }
}

// proc[0..E) = the expandable nodes expand[0][0..E) sorted by (key count, creation order), largest
// first (:680-682).  Small E: every thread ranks one element by counting the larger keys
// (unique keys -> a permutation), one barrier instead of a bitonic network.
__device__ __forceinline__ void oct_sort_expandable(const OctCtx& C, int cur, int E, int NCP2) {
    const int tid = threadIdx.x;
    if (E <= 2 * OCT_THREADS) {
        for (int e = tid; e < E; e += OCT_THREADS)
            C.sortk[e] = ((unsigned long long)C.ncn(cur)[C.expand(0)[e]] << 32) | (unsigned long long)(uint32_t)(e + 1);
        __syncthreads();
        for (int e = tid; e < E; e += OCT_THREADS) {
            const unsigned long long mine = C.sortk[e];
            int r = 0;
            for (int k = 0; k < E; ++k) r += C.sortk[k] > mine ? 1 : 0;
            C.proc[r] = C.expand(0)[e];
        }
        __syncthreads();
        return;
    }
    for (int e = tid; e < NCP2; e += OCT_THREADS)
        C.sortk[e] = e < E ? (((unsigned long long)C.ncn(cur)[C.expand(0)[e]] << 32) |
                              (unsigned long long)(uint32_t)(e + 1))
                           : 0ull;
    __syncthreads();
    int P2 = 1;
    while (P2 < E) P2 <<= 1;
    oct_bitonic_desc(C.sortk, P2);
    for (int j = tid; j < E; j += OCT_THREADS)
        C.proc[j] = C.expand(0)[(uint32_t)(C.sortk[j] & 0xffffffffull) - 1u];
    __syncthreads();
}

// Shared memory carve-up (words).  NC = node capacity, NCP2 = next power of two >= NC,
// CW = max(4 * NC, cells_max + 1) (cc[] doubles as the cell-offset table during the gather).
__host__ __device__ inline size_t oct_fixed_words(int NC, int NCP2, int CW) {
    return (size_t)2 * 2 * NC      // nb[2] (short4 = 2 words)
           + 2 * NC + 2 * NC       // nst[2], ncn[2]
           + CW                    // cc / cell offsets
           + NC                    // proc
           + 3 * NC                // expand[2], done
           + (NC + 2) + (NC + 2)   // aux, flag
           + 2 * (size_t)NCP2      // sortk (u64)
           + OCT_WARPS * 8 + 64;   // wc, tmp
}

__global__ void __launch_bounds__(OCT_THREADS, OCT_MINB)
k_octree(const __grid_constant__ OrbxPlan P, const uint32_t* __restrict__ cellcnt,
         const uint32_t* __restrict__ cellkeys, uint32_t* __restrict__ keysA,
         uint32_t* __restrict__ keysB, uint32_t* surv, uint32_t* survcnt,
         const int* __restrict__ lap, uint32_t* __restrict__ dest, int32_t* __restrict__ header,
         uint32_t* frame_done /* [nframes], zero between launches */,
         int NC, int NCP2, int CW, int key_words /* shared-memory words per key buffer */,
         int level_base /* first level of this launch (per-level launches of the latency graph) */) {
    extern __shared__ __align__(16) uint32_t sm[];
#ifdef OCT_FRAME_MAJOR
    const int level = level_base + blockIdx.x, frame = blockIdx.y;
#else
    // level-major launch order: the long level-0 CTAs of every frame start first, the short top levels fill
    // the tail (longest-processing-time-first packing of the 8 very different per-level durations)
    const int level = level_base + blockIdx.y, frame = blockIdx.x;
#endif
    const OrbxLevel& L = P.lvl[level];
    const int tid = threadIdx.x, lane = orbx_lane(), warp = orbx_warp();

    OctCtx C;
    uint32_t* p = sm;
    C.sortk = reinterpret_cast<unsigned long long*>(p); p += 2 * (size_t)NCP2;
    C.nb0 = reinterpret_cast<short4*>(p); p += 2 * NC;
    C.nb1 = reinterpret_cast<short4*>(p); p += 2 * NC;
    C.nst0 = p; p += NC;
    C.nst1 = p; p += NC;
    C.ncn0 = p; p += NC;
    C.ncn1 = p; p += NC;
    C.cc = p; p += CW;
    C.proc = p; p += NC;
    C.ex0 = p; p += NC;
    C.ex1 = p; p += NC;
    C.done = p; p += NC;
    C.aux = p; p += NC + 2;
    C.flag = p; p += NC + 2;
    C.wc = p; p += OCT_WARPS * 8;
    C.tmp = p; p += 64;
    uint32_t* smkeys = p;

    // ---- gather the level's candidates in the reference's order (cells row-major) ----------
    const int ncells = L.n_cols * L.n_rows;
    const uint32_t* ccnt = cellcnt + (size_t)frame * P.cells_total + L.cell_off;
    const uint32_t* ckeys = cellkeys + (size_t)frame * P.keys_total + L.key_off;
    uint32_t* cofs = C.cc;
    for (int c = tid; c < ncells; c += OCT_THREADS) cofs[c] = ccnt[c];
    __syncthreads();
    const uint32_t nk = block_excl_scan(cofs, ncells, C.tmp);
    if (tid == 0) cofs[ncells] = nk;
    __syncthreads();
    if ((int)nk <= key_words) {
        C.buf0 = smkeys;
        C.buf1 = smkeys + key_words;
    } else {
        C.buf0 = keysA + (size_t)frame * P.keys_total + L.key_off;
        C.buf1 = keysB + (size_t)frame * P.keys_total + L.key_off;
    }
    // thread per cell: a cell holds a handful of keys, and all cells' loads are in flight at once
    for (int c = tid; c < ncells; c += OCT_THREADS) {
        const uint32_t o = cofs[c], n = cofs[c + 1] - o;
        const uint32_t* s = ckeys + (size_t)c * L.cell_cap;
        for (uint32_t k = 0; k < n; ++k) C.buf0[o + k] = s[k];
    }
    __syncthreads();

    // ---- root nodes (:541-581): keys -> vpIniNodes[kp.pt.x / hX], empty roots erased --------
    const int nIni = L.n_ini;
    const int N = L.n_target;
    const int Wd = L.max_bx - ORBX_FAST_BORDER, Hd = L.max_by - ORBX_FAST_BORDER;
    (void)Wd;
    uint32_t* rootcnt = C.tmp + 48;  // 8 words
    oct_block_partition<1>(C, C.buf0, C.buf1, 0, nk, nIni, 0, 0, L.hx, rootcnt);
    // roots in creation order = reversed list order (the list is rev(creation) ++ rev(done))
    if (tid == 0) {
        uint32_t start = nk;
        int s = 0, m = 0, d = 0;
        for (int i = nIni - 1; i >= 0; --i) {
            const uint32_t c = rootcnt[i];
            start -= c;
            if (c == 0) continue;
            short4 b;
            b.x = (short)(int)(L.hx * (float)i);
            b.y = (short)(int)(L.hx * (float)(i + 1));
            b.z = 0;
            b.w = (short)Hd;
            C.nb(0)[s] = b;
            C.nst0[s] = start | 0x80000000u;  // keys now live in buf[1]
            C.ncn(0)[s] = c;
            if (c > 1) C.expand(0)[m++] = (uint32_t)s;
            else C.done[d++] = start | 0x80000000u;
            ++s;
        }
        C.tmp[42] = (uint32_t)s;
        C.tmp[44] = (uint32_t)m;
    }
    __syncthreads();
    int M = (int)C.tmp[42];   // children of the last pass (creation-order arrays [cur])
    int E = (int)C.tmp[44];   // ... of which multi-key
    int Dbefore = 0;          // done[] entries older than that pass
    int S = M;
    int cur = 0;
    bool maybe_big = nk >= OCT_BIG;

    // ---- full passes (:583-661), list kept implicit -------------------------------------------
    bool careful = false;
    while (true) {
        const int D = Dbefore + M - E;
        int nm = 0;
        const int Mn = oct_full_pass(C, cur, E, D, maybe_big, &nm);
        const int Sn = S - E + Mn;
        const bool stalled = Sn == S;
        maybe_big = maybe_big && C.tmp[OT_BIGNEXT] != 0;
        cur ^= 1;
        S = Sn; M = Mn; E = nm; Dbefore = D;
        if (S >= N || stalled) break;
        if (S + 3 * E > N) {
            careful = true;
            break;
        }
    }
    // ---- materialise the list: rev(children) ++ rev(done) into the free array set ---------------
    {
        const int nxt = cur ^ 1;
        for (int p2 = tid; p2 < S; p2 += OCT_THREADS) {
            if (p2 < M) {
                const int c = M - 1 - p2;
                C.nb(nxt)[p2] = C.nb(cur)[c];
                C.nst(nxt)[p2] = C.nst(cur)[c];
                C.ncn(nxt)[p2] = C.ncn(cur)[c];
            } else {
                C.nst(nxt)[p2] = C.done[Dbefore - 1 - (p2 - M)];
                C.ncn(nxt)[p2] = 1;
            }
        }
        for (int e = tid; e < E; e += OCT_THREADS) C.expand(0)[e] = (uint32_t)(M - 1) - C.expand(cur)[e];
        __syncthreads();
        cur = nxt;
    }

    // ---- careful passes (:663-736) --------------------------------------------------------------
    if (careful) {
        int Ecur = E;
        bool finish = false;
        while (!finish) {
            const int prev2 = S;
            oct_sort_expandable(C, cur, Ecur, NCP2);
            int nm = 0;
            S = oct_pass(C, cur, S, Ecur, true, N, &nm, maybe_big);
            cur ^= 1;
            Ecur = nm;
            if (S >= N || S == prev2) finish = true;
        }
    }

    // ---- best response per node, first maximum wins (:738-758) --------------------------------
    uint32_t* out = surv + (size_t)frame * P.surv_total + L.surv_off;
    const int Sout = min(S, L.surv_cap);
    // thread per node (a final node holds a few keys); nodes with many keys take a warp each
    for (int i = tid; i < Sout; i += OCT_THREADS) {
        const uint32_t cnt = C.ncn(cur)[i];
        if (cnt > 64) continue;
        const uint32_t st = C.nst(cur)[i];
        const uint32_t* src = C.buf(st >> 31) + (st & 0x7fffffffu);
        uint32_t bestk = src[0];
        for (uint32_t k = 1; k < cnt; ++k) {
            const uint32_t v = src[k];
            if ((v >> 24) > (bestk >> 24)) bestk = v;   // strict '>': the first maximum wins
        }
        out[i] = bestk;
    }
    for (int i = warp; i < Sout; i += OCT_WARPS) {
        const uint32_t cnt = C.ncn(cur)[i];
        if (cnt <= 64) continue;
        const uint32_t st = C.nst(cur)[i];
        const uint32_t* src = C.buf(st >> 31) + (st & 0x7fffffffu);
        unsigned long long best = 0;
        for (uint32_t k = lane; k < cnt; k += 32) {
            const unsigned long long v = ((unsigned long long)(src[k] >> 24) << 32) |
                                         (unsigned long long)(0xffffffffu - k);
            best = v > best ? v : best;
        }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            const unsigned long long o = __shfl_xor_sync(ORBX_FULL, best, d);
            best = o > best ? o : best;
        }
        if (lane == 0) out[i] = src[0xffffffffu - (uint32_t)(best & 0xffffffffull)];
    }
    if (tid == 0) survcnt[(size_t)frame * ORBX_MAX_LEVELS + level] = (uint32_t)S;

    // ---- E7 in the tail: the last CTA of the frame to get here lays out the frame's output --------
    // (release: every thread publishes its survivors, then one ticket per CTA; acquire: the winner
    // fences again and reads the other levels' results through L2)
    __threadfence();
    __syncthreads();
    if (tid == 0) C.tmp[45] = atomicAdd(&frame_done[frame], 1u) == (uint32_t)P.nlevels - 1 ? 1u : 0u;
    __syncthreads();
    if (C.tmp[45] == 0) return;
    __threadfence();
    if (tid == 0) frame_done[frame] = 0;   // ready for the next launch on this slot
    __syncthreads();                        // C.tmp is about to be reused as scan scratch
    orbx_assemble_frame(P, surv, survcnt, lap, dest, header, frame, sm);
}

struct OctLaunchCfg {
    int NC, NCP2, CW, key_words;
    size_t smem;
};

OctLaunchCfg orbx_octree_cfg(const OrbxPlan& P) {
    OctLaunchCfg c;
    int nc = 8, cells = 1;
    long long kmax = 0;
    for (int l = 0; l < P.nlevels; ++l) {
        nc = nc > P.lvl[l].node_cap ? nc : P.lvl[l].node_cap;
        int cl = P.lvl[l].n_cols * P.lvl[l].n_rows;
        cells = cells > cl ? cells : cl;
        kmax = kmax > P.lvl[l].key_cap ? kmax : P.lvl[l].key_cap;
    }
    c.NC = nc;
    c.NCP2 = 1;
    while (c.NCP2 < nc) c.NCP2 <<= 1;
    c.CW = 4 * nc > cells + 2 ? 4 * nc : cells + 2;
    const size_t fixed = oct_fixed_words(c.NC, c.NCP2, c.CW) * 4;
    // key buffers: sized so that 5 CTAs fit an SM (larger levels spill to the HBM ping-pong buffers), capped by the worst case
    size_t budget = OCT_SMEM_KB * 1024;
    long long kw = budget > fixed ? (long long)((budget - fixed) / 8) : 0;
    if (kw > kmax) kw = kmax;
    kw &= ~3ll;
    c.key_words = (int)kw;
    c.smem = fixed + (size_t)kw * 8;
    const size_t assemble = (size_t)(P.surv_total + 2 + 64) * 4;   // orbx_assemble_frame in the kernel's tail
    if (c.smem < assemble) c.smem = assemble;
    return c;
}

cudaError_t orbx_configure_octree(const OrbxPlan& P) {
    OctLaunchCfg c = orbx_octree_cfg(P);
    if (c.smem > ORBX_MAX_DYN_SMEM) return cudaErrorInvalidConfiguration;
    // always the opt-in maximum: the attribute is per kernel and device, another live handle may need more than this plan
    return cudaFuncSetAttribute(k_octree, cudaFuncAttributeMaxDynamicSharedMemorySize, ORBX_MAX_DYN_SMEM);
}

cudaError_t orbx_launch_octree(const OrbxPlan& P, const uint32_t* cellcnt, const uint32_t* cellkeys,
                               uint32_t* keysA, uint32_t* keysB, uint32_t* surv, uint32_t* survcnt,
                               const int* lap, uint32_t* dest, int32_t* header, uint32_t* frame_done,
                               int nframes, cudaStream_t st, int* launches, int level) {
    OctLaunchCfg c = orbx_octree_cfg(P);
    // level < 0: one CTA per (level, frame); else only that level (the forked latency graph).  The
    // frame ticket counts P.nlevels CTAs either way: the last level to finish lays out the frame.
#ifdef OCT_FRAME_MAJOR
    dim3 grid(level < 0 ? P.nlevels : 1, nframes);
#else
    dim3 grid(nframes, level < 0 ? P.nlevels : 1);
#endif
    k_octree<<<grid, OCT_THREADS, c.smem, st>>>(P, cellcnt, cellkeys, keysA, keysB, surv, survcnt, lap, dest, header,
                                               frame_done, c.NC,
                                               c.NCP2, c.CW, c.key_words, level < 0 ? 0 : level);
    ++*launches;
    return cudaGetLastError();
}
