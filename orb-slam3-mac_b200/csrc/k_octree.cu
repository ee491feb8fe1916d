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
#include <cstdlib>

#include "orbx_device.cuh"

#ifndef OCT_SMEM_KB
#define OCT_SMEM_KB 36       // 256-thread CTAs: 5 fit an SM
#endif
#ifndef OCT_SMEM_KB_128
#define OCT_SMEM_KB_128 26   // 128-thread CTAs: 8 per SM
#endif
#ifndef OCT_SMEM_KB_512
#define OCT_SMEM_KB_512 36   // 512-thread CTAs: 2 per SM by their threads and registers
#endif
#ifndef OCT_BIG
#define OCT_BIG 1024
#endif
#define OCT_MAX_INI 8

// Shared memory carve-up (words).  NC = node capacity, NCP2 = next power of two >= NC,
// CW = max(4 * NC, cells_max + 1) (cc[] doubles as the cell-offset table during the gather).
__host__ __device__ inline size_t oct_fixed_words(int NC, int NCP2, int CW, int warps) {
    return (size_t)2 * 2 * NC      // nb[2] (short4 = 2 words)
           + 2 * NC + 2 * NC       // nst[2], ncn[2]
           + CW                    // cc / cell offsets
           + NC                    // proc
           + 3 * NC                // expand[2], done
           + (NC + 2) + (NC + 2)   // aux, flag
           + 2 * (size_t)NCP2      // sortk (u64)
           + warps * 8 + 64;   // wc, tmp
}

// The kernel is compiled for three CTA sizes: a (frame, level) with a few hundred keys is a chain of short dependent
// passes, which wants many small CTAs per SM (latency hiding across CTAs); one with thousands of keys (C5's fine
// levels, the 5 x nFeatures initialisation extractor) wants the passes themselves wide.  orbx_octree_threads() picks
// per plan.  Same-box A/B (ms per launch batch, C1 / C5): 128 x 8 CTAs 0.163 / 0.513, 256 x 4 0.177 / 0.300,
// 512 x 2 0.277 / 0.186.
#ifndef OCT_MINB_128
#define OCT_MINB_128 8
#endif
#define OCT_THREADS 128
#define OCT_MINB OCT_MINB_128
#define OCT_NS oct128
#include "k_octree_body.inc"
#undef OCT_THREADS
#undef OCT_MINB
#undef OCT_NS
#define OCT_THREADS 256
#define OCT_MINB 4
#define OCT_NS oct256
#include "k_octree_body.inc"
#undef OCT_THREADS
#undef OCT_MINB
#undef OCT_NS
#define OCT_THREADS 512
#define OCT_MINB 2
#define OCT_NS oct512
#include "k_octree_body.inc"
#undef OCT_THREADS
#undef OCT_MINB
#undef OCT_NS

struct OctLaunchCfg {
    int NC, NCP2, CW, key_words, threads;
    size_t smem;
};
#ifndef OCT_QUOTA_128
#define OCT_QUOTA_128 240   // largest per-level quota up to which 128-thread CTAs win (C1: 217; C2: 261 is a tie)
#endif
#ifndef OCT_QUOTA_256
#define OCT_QUOTA_256 650   // ... 256-thread CTAs (C3: 326, C4: 434); above: 512 (C5: 869, 5000-feature mono init: 1086)
#endif

// CTA size of a launch of `ctas` CTAs.  A launch that leaves SMs to spare (the single-frame latency path: 8 CTAs, or
// one per level branch of the forked graph) takes the widest CTAs -- only the length of the dependent pass chain counts
// (p50 of one C1 frame: 147 / 110 / 103 us with 128 / 256 / 512 threads; C4 243 / 173 / 145; C5 555 / 358 / 271).  A
// batch goes by the plan's largest per-level feature quota (the finest level's), which sets how many nodes and keys
// the heaviest CTAs walk.  ORBX_OCT_THREADS=128|256|512 overrides (developer A/B).
static int orbx_octree_threads(const OrbxPlan& P, int ctas) {
    static const int forced = [] {
        const char* e = getenv("ORBX_OCT_THREADS");
        const int v = e ? atoi(e) : 0;
        return (v == 128 || v == 256 || v == 512) ? v : 0;
    }();
    if (forced) return forced;
    int dev = 0, sms = 148;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) {
        cudaGetLastError();
        sms = 148;
    }
    if (ctas <= 2 * sms) return 512;
    int quota = 0;
    for (int l = 0; l < P.nlevels; ++l) quota = quota > P.lvl[l].n_target ? quota : P.lvl[l].n_target;
    // 128-thread CTAs only pay when there are at least two full waves of them (8 per SM): C1, ms per launch with
    // 128 / 256 threads: 48 frames 0.093 / 0.054, 96 frames 0.103 / 0.063, 192 frames 0.120 / 0.101, 384 frames 0.163 / 0.177
    if (quota <= OCT_QUOTA_128 && ctas >= 2 * OCT_MINB_128 * sms) return 128;
    return quota <= OCT_QUOTA_256 ? 256 : 512;
}

OctLaunchCfg orbx_octree_cfg(const OrbxPlan& P, int threads) {
    OctLaunchCfg c;
    c.threads = threads;
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
    const size_t fixed = oct_fixed_words(c.NC, c.NCP2, c.CW, c.threads / 32) * 4;
    // key buffers: sized so that the CTAs the launch bounds ask for fit an SM (larger levels spill to the HBM ping-pong
    // buffers), capped by the worst case
    size_t budget = (c.threads == 128 ? OCT_SMEM_KB_128 : c.threads == 256 ? OCT_SMEM_KB : OCT_SMEM_KB_512) * 1024;
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
    for (int t : {128, 256, 512})
        if (orbx_octree_cfg(P, t).smem > ORBX_MAX_DYN_SMEM) return cudaErrorInvalidConfiguration;
    // always the opt-in maximum: the attribute is per kernel and device, another live handle may need more than this plan
    cudaError_t e = cudaFuncSetAttribute(oct128::k_octree, cudaFuncAttributeMaxDynamicSharedMemorySize, ORBX_MAX_DYN_SMEM);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(oct256::k_octree, cudaFuncAttributeMaxDynamicSharedMemorySize, ORBX_MAX_DYN_SMEM);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(oct512::k_octree, cudaFuncAttributeMaxDynamicSharedMemorySize, ORBX_MAX_DYN_SMEM);
    return e;
}

cudaError_t orbx_launch_octree(const OrbxPlan& P, const uint32_t* cellcnt, const uint32_t* cellkeys,
                               uint32_t* keysA, uint32_t* keysB, uint32_t* surv, uint32_t* survcnt,
                               const int* lap, uint32_t* dest, int32_t* header, uint32_t* frame_done,
                               int nframes, cudaStream_t st, int* launches, int level) {
    // level < 0: one CTA per (level, frame); else only that level (the forked latency graph).  The
    // frame ticket counts P.nlevels CTAs either way: the last level to finish lays out the frame.
    const int nl = level < 0 ? P.nlevels : 1;
    const OctLaunchCfg c = orbx_octree_cfg(P, orbx_octree_threads(P, nframes * nl));
    dim3 grid(nframes, nl);
    auto kern = c.threads == 128 ? oct128::k_octree : (c.threads == 256 ? oct256::k_octree : oct512::k_octree);
    kern<<<grid, c.threads, c.smem, st>>>(P, cellcnt, cellkeys, keysA, keysB, surv, survcnt, lap, dest, header, frame_done,
                                          c.NC, c.NCP2, c.CW, c.key_words, level < 0 ? 0 : level);
    ++*launches;
    return cudaGetLastError();
}
