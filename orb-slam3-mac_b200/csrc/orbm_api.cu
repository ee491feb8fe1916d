// orbm_api.cu -- host runtime behind include/orbm.h: the Hamming searches of ORBmatcher / Frame.
//
// Layer 1 ships descriptors to the device, runs one batched distance kernel (k_match.cu) and
// returns distances / arg-mins.  Layer 2 restates the reference's search methods on POD
// snapshots: candidate generation on the 64x48 frame grid (Frame::GetFeaturesInArea order), ONE
// batched orbm_hamming_csr over every (query, candidate) pair, then the method's own sequential
// accept/reject loop replayed over the precomputed distances in the original order, so results
// equal the reference's even where acceptance depends on earlier acceptances
// (src/ORBmatcher.cc:93-95, 749-750, 2040-2042).  No distance is ever computed on the host.
#include <cuda_runtime.h>

#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <map>
#include <string>
#include <memory>
#include <unordered_map>
#include <vector>

#include "../../include/orbm.h"
#include "orbm_internal.h"

cudaError_t orbm_launch_hamming_csr(const uint8_t* q, int nq, const uint8_t* t, const int32_t* row_ptr,
                                    const int32_t* cand, int nnz, uint16_t* dist, cudaStream_t st);
cudaError_t orbm_launch_grid_build(const orbx_keypoint* keys, int n, float min_x, float min_y, float w_inv,
                                   float h_inv, int32_t* start, int32_t* items, int32_t* cellof, cudaStream_t st);
cudaError_t orbm_launch_area_count(const orbx_keypoint* keys, const int32_t* start, const int32_t* items,
                                   float min_x, float min_y, float w_inv, float h_inv, const orbm_area_query* Q,
                                   int nq, int32_t* cnt, int32_t* row_ptr, cudaStream_t st);
cudaError_t orbm_launch_area_onepass(const orbx_keypoint* keys, const int32_t* start, const int32_t* items,
                                     float min_x, float min_y, float w_inv, float h_inv, const orbm_area_query* Q,
                                     int nq, uint2* row_info, uint32_t* counter, uint32_t counter0, const uint8_t* qdesc,
                                     const uint8_t* tdesc, uint2* pairs, int cap, cudaStream_t st);
cudaError_t orbm_launch_area_fill(const orbx_keypoint* keys, const int32_t* start, const int32_t* items,
                                  float min_x, float min_y, float w_inv, float h_inv, const orbm_area_query* Q,
                                  int nq, const int32_t* row_ptr, const uint8_t* qdesc, const uint8_t* tdesc,
                                  int32_t* cand, uint16_t* dist, int cap, cudaStream_t st);
cudaError_t orbm_launch_csr_topk(const int32_t* row_ptr, const int32_t* cand, const uint16_t* dist, int nrows, int K,
                                 int cap, uint32_t* keys, int32_t* key_cand, cudaStream_t st);
cudaError_t orbm_launch_scatter(const OrbmScatter& S, cudaStream_t st);
cudaError_t orbm_launch_popc_peak(uint32_t* out, int blocks, int iters, cudaStream_t st);
cudaError_t orbm_launch_knn2(const uint8_t* q, int nq, const uint8_t* t, int nt, int32_t* idx,
                             int32_t* dist, cudaStream_t st);
cudaError_t orbm_launch_stereo_refine(const orbx_keypoint* kl, int nl, const orbx_keypoint* kr,
                                      const int32_t* best_idx, const int32_t* best_dist,
                                      const orbx_pyramid_view* vl, const orbx_pyramid_view* vr,
                                      const float* scale, const float* inv_scale, int th_orb, float min_d,
                                      float max_d, float bf, float* u_right, float* depth, int32_t* sad,
                                      cudaStream_t st);
cudaError_t orbm_launch_stereo_rows(const orbx_keypoint* kl, const uint8_t* dl, int nl,
                                    const orbx_keypoint* kr, const uint8_t* dr, int nr,
                                    const float* scale, float min_d, float max_d, int32_t* best_idx,
                                    int32_t* best_dist, cudaStream_t st);

// error string shared with orbx_api.cu (orbx_last_error_string reads it)
int orbx_set_error(int code, const char* fmt, ...);

// A failed CUDA call ends the API call early.  Copies that the call had already staged (ctx_h2d / ctx_d2h: scatter
// segments, pending deliveries into the CALLER's arrays -- some of them stack locals of the failed call) must not
// survive into the next call's ctx_sync: orbm_abort_staged() drains the stream and drops them.
struct orbm_context;
void orbm_abort_staged(orbm_context* c);
static thread_local orbm_context* g_staging_ctx = nullptr;   // the context whose call is staging copies on this thread
#define M_TRY(expr)                                                                              \
    do {                                                                                         \
        cudaError_t _e = (expr);                                                                 \
        if (_e != cudaSuccess) {                                                                 \
            if (g_staging_ctx) orbm_abort_staged(g_staging_ctx);                                 \
            return orbx_set_error(ORBX_E_CUDA, "%s failed: %s (%s:%d)", #expr,                  \
                                  cudaGetErrorString(_e), __FILE__, __LINE__);                   \
        }                                                                                        \
    } while (0)

namespace {

struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    int ensure(size_t n) {
        if (n <= cap) return ORBX_OK;
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
        n = std::max(n + n / 4, (size_t)4096);
        M_TRY(cudaMalloc(&p, n));
        cap = n;
        return ORBX_OK;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};

}  // namespace

// Developer probe (make EXTRA=-DORBM_TRACE): host-side phase times of a call, printed to stderr.
#ifdef ORBM_TRACE
#include <chrono>
struct PhaseTimer {   // marks are kept in memory and printed when the call returns: a write per mark disturbs what it measures
    const char* what;
    const char* name[16];
    double us[16];
    int n = 0;
    std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
    explicit PhaseTimer(const char* w) : what(w) {}
    void mark(const char* phase) {
        auto t1 = std::chrono::steady_clock::now();
        if (n < 16) {
            name[n] = phase;
            us[n++] = std::chrono::duration<double, std::micro>(t1 - t0).count();
        }
        t0 = t1;
    }
    ~PhaseTimer() {
        for (int i = 0; i < n; ++i) fprintf(stderr, "[orbm %s] %s %.1f us\n", what, name[i], us[i]);
    }
};
#define ORBM_PHASE_BEGIN(w) PhaseTimer _pt(w)
#define ORBM_PHASE(p) _pt.mark(p)
#else
#define ORBM_PHASE_BEGIN(w) do {} while (0)
#define ORBM_PHASE(p) do {} while (0)
#endif

#ifndef INIT_TOPK
#define INIT_TOPK 4   // keys per row of SearchForInitialization (k_csr_topk)
#endif

struct orbm_context {
    int device = 0;
    cudaStream_t st = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;   // around the kernels of the last call (orbm_last_device_ms)
    bool timed = false;
    DevBuf q, t, rowptr, cand, dist, aux1, aux2, kl, kr, sc, isc, ur, dp, sd, gstart, gitems, gcell, aq, acnt;
    std::vector<int32_t> h_rowptr, h_cand;   // candidate lists of the last layer-2 search
    std::vector<uint16_t> hdist;     // distances of the last layer-2 search
    std::vector<int32_t> h_cand2;    // (scratch of the CSR re-pack in device_candidates)
    std::vector<uint16_t> hdist2;
    std::vector<uint32_t> h_topk;    // k_bow_topk's / k_csr_topk's keys of the last search
    std::vector<int32_t> h_topc;     // k_csr_topk: the candidate each key names
    int nnz_hint = 0;                // pairs of the last window search: sizes the lists of the next one before its counts are known
    bool have_full = true;           // h_cand / hdist hold the last search's full lists (top-K mode fetches them lazily)
    std::vector<float> h_rad;        // window radius per query of the last SearchByProjection (the accept loop reads it)
    std::vector<int32_t> h_items;    // work items of the last orbm_distinctive_descriptors call
    long long last_pairs = 0;        // number of (query, candidate) pairs of the last search
    // pinned staging arena (ctx_h2d / ctx_d2h / ctx_sync below)
    uint8_t* pin = nullptr;
    size_t pin_cap = 0, pin_used = 0, pin_want = 0;
    struct PendingCopy { void* dst; const void* src; size_t bytes; };
    std::vector<PendingCopy> pending;
    DevBuf darena;                   // device mirror of the arena: uploads arrive in one copy (ctx_flush)
    OrbmScatter up = {};                // staged uploads not yet sent (src = offset in the arena until ctx_flush)
    size_t up_first = 0, up_end = 0;    // span of the arena the next ctx_flush copies ([first, end), empty when end == 0)
    // one-pass window search (device_candidates_onepass): {first position, count} per query, the running counter
    std::vector<uint2> h_rowinfo;
    std::vector<uint8_t> h_one;        // (fallback landing area of the one-buffer download)
    bool hold_arena = false;           // a ctx_d2h_view is being read in place: ctx_sync keeps the arena as it is
    DevBuf counter;                     // [counter | rows | records] of the one-launch search (device_candidates_onepass)
    uint32_t counter_host = 0;          // the counter's value after the last search that used it
    bool counter_known = false;         // false: zero it before the next use (first use, or a call that failed mid-way)
    std::unordered_map<const orbm_frame_view*, struct orbm_frame*> frames;   // device-resident frames of this context
    std::vector<struct orbm_bank*> banks;         // descriptor banks of this context still alive (orbm_destroy frees them)
    std::vector<struct orbm_frame*> frame_pool;   // destroyed frames whose device buffers the next frame reuses (no cudaMalloc per frame)
};

cudaError_t orbm_launch_undistort(const orbx_keypoint* keys, int n, const float* cam9, float width, float height,
                                  orbx_keypoint* keys_un, float* corners_xy, cudaStream_t st);

// Descriptor rows that stay on the device across calls: the descriptors of the map points (one row per MapPoint,
// rewritten when MapPoint::ComputeDistinctiveDescriptors changes it) or of a resident frame (orbm_frame_bank).  The
// *_bank searches name their queries' descriptors by row instead of sending 32 bytes per query with every call.
struct orbm_bank {
    orbm_context* ctx = nullptr;
    DevBuf buf;                   // rows owned by the bank (orbm_bank_create)
    const uint8_t* d = nullptr;   // the rows: buf.p, or a resident frame's descriptors
    int rows = 0;
};

// A frame whose (undistorted) keypoints, descriptors and 64 x 48 grid stay on the device across the two or three
// searches it goes through (TrackWithMotionModel -> SearchLocalPoints -> relocalisation).  `view` is what the
// searches take: its address is the frame's identity (device_candidates looks it up in c->frames).
struct orbm_frame {
    orbm_context* ctx = nullptr;
    orbm_frame_view view = {};
    DevBuf d_keys, d_desc, d_start, d_items, d_cell;
    std::vector<orbx_keypoint> h_keys;   // mvKeysUn (the accept loops read octaves on the host)
    std::vector<uint8_t> h_desc;
    std::vector<float> h_scale, h_uright;
    orbm_bank bank;   // view of d_desc (orbm_frame_bank)
};

// Host <-> device copies of a call go through one pinned arena per context: a caller's (pageable) array
// is copied into the arena and sent with a truly asynchronous cudaMemcpyAsync, results land in the
// arena and are handed to the caller's arrays by ctx_sync().  A pageable cudaMemcpyAsync is staged by
// the driver synchronously, copy by copy (4-6 copies per search: most of a small search's 40-80 us);
// from pinned memory the copies only queue.  Payloads that do not fit the arena take the direct path
// and the arena grows at the next ctx_sync (nothing is in flight then).
static const size_t ORBM_PIN_INIT = 2u << 20, ORBM_PIN_MAX = 32u << 20;
// larger single copies go direct: the driver pipelines its own staging with the DMA, one memcpy here would not
static const size_t ORBM_PIN_DIRECT = 1u << 20;

// Sends the staged uploads: one copy of their span of the arena into the device mirror, one k_scatter
// launch that hands the segments to their buffers.  Called before the first kernel of a search (every
// kernel sequence starts at the ev0 record), before a download is queued and at the sync.
static cudaError_t ctx_flush(orbm_context* c) {
    if (c->up.n == 0 && c->up_end == 0) return cudaSuccess;
#ifdef ORBM_NO_SCATTER   // A/B switch: a copy per segment
    const bool mirror = false;
#else
    const bool mirror = c->darena.ensure(c->pin_cap) == ORBX_OK;
#endif
    if (!mirror) {   // no mirror: a copy per segment
        cudaGetLastError();
        for (int k = 0; k < c->up.n; ++k) {
            cudaError_t e = cudaMemcpyAsync(c->up.seg[k].dst, c->pin + (uintptr_t)c->up.seg[k].src, c->up.seg[k].bytes,
                                            cudaMemcpyHostToDevice, c->st);
            if (e != cudaSuccess) return e;
        }
        c->up.n = 0;
        c->up_end = 0;
        return cudaSuccess;
    }
    cudaError_t e = cudaMemcpyAsync((uint8_t*)c->darena.p + c->up_first, c->pin + c->up_first, c->up_end - c->up_first,
                                    cudaMemcpyHostToDevice, c->st);
    c->up_end = 0;
    if (e != cudaSuccess) return e;
    if (c->up.n == 0) return cudaSuccess;   // everything staged was read in place (ctx_stage_alias)
    for (int k = 0; k < c->up.n; ++k) c->up.seg[k].src = (const uint8_t*)c->darena.p + (uintptr_t)c->up.seg[k].src;
    e = orbm_launch_scatter(c->up, c->st);
    c->up.n = 0;
    return e;
}

void orbm_abort_staged(orbm_context* c) {
    cudaStreamSynchronize(c->st);   // whatever was queued may still read the arena
    cudaGetLastError();
    c->pending.clear();
    c->up.n = 0;
    c->up_end = 0;
    c->pin_used = 0;
    c->counter_known = false;
    c->hold_arena = false;
    g_staging_ctx = nullptr;
}

static cudaError_t ctx_h2d(orbm_context* c, void* dst, const void* src, size_t bytes) {
    if (bytes == 0) return cudaSuccess;
    g_staging_ctx = c;
    const size_t off = (c->pin_used + 63) & ~(size_t)63;
    if (bytes > ORBM_PIN_DIRECT) return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, c->st);
    if (off + bytes > c->pin_cap) {
        c->pin_want = std::max(c->pin_want, off + bytes);
        return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, c->st);
    }
    memcpy(c->pin + off, src, bytes);
    c->pin_used = off + bytes;
    if (c->up.n == ORBM_SCATTER_MAX) {
        cudaError_t e = ctx_flush(c);
        if (e != cudaSuccess) return e;
    }
    if (c->up_end == 0) c->up_first = off;
    c->up_end = off + bytes;
    c->up.seg[c->up.n++] = OrbmScatter::Seg{dst, (const void*)(uintptr_t)off, (uint32_t)bytes, 0};
    return cudaSuccess;
}

// Room for `bytes` of an upload to `dst` inside the arena, for a caller that builds the data in place (the window
// queries of a search: no intermediate array, no memcpy).  nullptr when the arena cannot take it: the caller then
// builds its own array and sends it with ctx_h2d.
static void* ctx_stage(orbm_context* c, void* dst, size_t bytes) {
    const size_t off = (c->pin_used + 63) & ~(size_t)63;
    if (bytes == 0 || bytes > ORBM_PIN_DIRECT || c->up.n == ORBM_SCATTER_MAX) return nullptr;
    if (off + bytes > c->pin_cap) {
        c->pin_want = std::max(c->pin_want, off + bytes);
        return nullptr;
    }
    g_staging_ctx = c;
    c->pin_used = off + bytes;
    if (c->up_end == 0) c->up_first = off;
    c->up_end = off + bytes;
    c->up.seg[c->up.n++] = OrbmScatter::Seg{dst, (const void*)(uintptr_t)off, (uint32_t)bytes, 0};
    return c->pin + off;
}

// The same, for data the kernels read IN PLACE in the arena's device mirror: no scatter, the flush's one copy brings
// it over.  *dev = where it will be on the device.  nullptr when the arena (or its mirror) cannot take it.
static void* ctx_stage_alias(orbm_context* c, size_t bytes, const void** dev) {
#ifdef ORBM_NO_SCATTER
    return nullptr;
#endif
    const size_t off = (c->pin_used + 255) & ~(size_t)255;
    if (bytes == 0 || bytes > ORBM_PIN_DIRECT || off + bytes > c->pin_cap) {
        if (bytes && bytes <= ORBM_PIN_DIRECT) c->pin_want = std::max(c->pin_want, off + bytes);
        return nullptr;
    }
    if (c->darena.ensure(c->pin_cap) != ORBX_OK) {
        cudaGetLastError();
        return nullptr;
    }
    g_staging_ctx = c;
    c->pin_used = off + bytes;
    if (c->up_end == 0) c->up_first = off;
    c->up_end = off + bytes;
    *dev = (const uint8_t*)c->darena.p + off;
    return c->pin + off;
}

// Drops what a call has staged but not sent (an argument error found while the queries were being built).
static void ctx_unstage(orbm_context* c) {
    c->up.n = 0;
    c->up_end = 0;
    c->pin_used = 0;
    g_staging_ctx = nullptr;
}

static cudaError_t ctx_d2h(orbm_context* c, void* dst, const void* src, size_t bytes) {
    if (bytes == 0) return cudaSuccess;
    g_staging_ctx = c;
    cudaError_t ef = ctx_flush(c);
    if (ef != cudaSuccess) return ef;
    const size_t off = (c->pin_used + 63) & ~(size_t)63;
    if (bytes > ORBM_PIN_DIRECT) return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, c->st);
    if (off + bytes > c->pin_cap) {
        c->pin_want = std::max(c->pin_want, off + bytes);
        return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, c->st);
    }
    c->pin_used = off + bytes;
    c->pending.push_back(orbm_context::PendingCopy{dst, c->pin + off, bytes});
    return cudaMemcpyAsync(c->pin + off, src, bytes, cudaMemcpyDeviceToHost, c->st);
}

// A download the caller reads IN PLACE in the arena after ctx_sync (no second copy into an array of its own): *view
// stays valid until the call stages again.  Falls back to nullptr when the arena cannot take it.
static cudaError_t ctx_d2h_view(orbm_context* c, const void* src, size_t bytes, const uint8_t** view) {
    *view = nullptr;
    if (bytes == 0) return cudaSuccess;
    g_staging_ctx = c;
    cudaError_t ef = ctx_flush(c);
    if (ef != cudaSuccess) return ef;
    const size_t off = (c->pin_used + 63) & ~(size_t)63;
    if (bytes > ORBM_PIN_DIRECT || off + bytes > c->pin_cap) {
        c->pin_want = std::max(c->pin_want, off + bytes);
        return cudaSuccess;
    }
    c->pin_used = off + bytes;
    *view = c->pin + off;   // (the caller holds the arena: ArenaHold)
    return cudaMemcpyAsync(c->pin + off, src, bytes, cudaMemcpyDeviceToHost, c->st);
}

// Replaces the arena by a larger one when the last call wanted more than it holds (nothing is in flight).
static void ctx_grow(orbm_context* c) {
    if (c->pin_cap && c->pin_want > c->pin_cap && c->pin_cap < ORBM_PIN_MAX) {
        size_t want = std::min(std::max(c->pin_want * 2, ORBM_PIN_INIT), ORBM_PIN_MAX);
        uint8_t* np = nullptr;
        if (cudaHostAlloc((void**)&np, want, cudaHostAllocDefault) == cudaSuccess) {
            if (c->pin) cudaFreeHost(c->pin);
            c->darena.release();
            c->pin = np;
            c->pin_cap = want;
        } else {
            cudaGetLastError();   // keep the arena we have
        }
    }
    c->pin_want = 0;
}

// Waits for the context's stream, delivers the staged results and recycles the arena.
static cudaError_t ctx_sync(orbm_context* c) {
    cudaError_t e = ctx_flush(c);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->st);
    else cudaStreamSynchronize(c->st);
    if (e == cudaSuccess)
        for (const orbm_context::PendingCopy& pc : c->pending) memcpy(pc.dst, pc.src, pc.bytes);
    c->pending.clear();
    c->pin_used = 0;
    c->up.n = 0;
    c->up_end = 0;
    g_staging_ctx = nullptr;
    if (c->hold_arena) return e;   // a view of the arena is being read: the reader calls ctx_grow when it is done
    if (e == cudaSuccess) ctx_grow(c);
    else c->pin_want = 0;
    return e;
}

extern "C" {

int orbm_create(int device, orbm_context** out) {
    if (!out) return orbx_set_error(ORBX_E_BAD_ARG, "null argument");
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return orbx_set_error(ORBX_E_CUDA, "no CUDA device (%s); this library has no CPU fallback", cudaGetErrorString(e));
    if (device < 0 || device >= ndev) return orbx_set_error(ORBX_E_BAD_ARG, "device %d out of range", device);
    M_TRY(cudaSetDevice(device));
    orbm_context* c = new orbm_context();
    c->device = device;
    if (cudaStreamCreateWithFlags(&c->st, cudaStreamNonBlocking) != cudaSuccess) {
        delete c;
        return orbx_set_error(ORBX_E_CUDA, "cudaStreamCreate failed");
    }
    if (cudaEventCreate(&c->ev0) != cudaSuccess || cudaEventCreate(&c->ev1) != cudaSuccess) {
        cudaStreamDestroy(c->st);
        delete c;
        return orbx_set_error(ORBX_E_CUDA, "cudaEventCreate failed");
    }
    // the arena is an optimisation: without it every copy takes the direct (pageable) path
#ifndef ORBM_NO_PIN   // A/B switch: every copy direct
    if (cudaHostAlloc((void**)&c->pin, ORBM_PIN_INIT, cudaHostAllocDefault) == cudaSuccess) c->pin_cap = ORBM_PIN_INIT;
    else cudaGetLastError();
#endif
    *out = c;
    return ORBX_OK;
}

int orbm_destroy(orbm_context* c) {
    if (!c) return ORBX_OK;
    cudaSetDevice(c->device);
    if (c->st) ctx_sync(c);
    while (!c->frames.empty()) orbm_frame_destroy(c->frames.begin()->second);   // frames die with their context
    while (!c->banks.empty()) orbm_bank_destroy(c->banks.back());               // ... and so do its descriptor banks
    for (orbm_frame* f : c->frame_pool) {
        for (DevBuf* b : {&f->d_keys, &f->d_desc, &f->d_start, &f->d_items, &f->d_cell}) b->release();
        delete f;
    }
    c->frame_pool.clear();
    for (DevBuf* b : {&c->q, &c->t, &c->rowptr, &c->cand, &c->dist, &c->aux1, &c->aux2, &c->kl, &c->kr, &c->sc, &c->isc,
                      &c->ur, &c->dp, &c->sd, &c->gstart, &c->gitems, &c->gcell, &c->aq, &c->acnt, &c->counter, &c->darena})
        b->release();
    if (c->ev0) cudaEventDestroy(c->ev0);
    if (c->ev1) cudaEventDestroy(c->ev1);
    if (c->st) cudaStreamDestroy(c->st);
    if (c->pin) cudaFreeHost(c->pin);
    delete c;
    return ORBX_OK;
}

int orbm_hamming_csr(orbm_context* c, const uint8_t* q, int nq, const uint8_t* t, int nt,
                     const int32_t* row_ptr, const int32_t* cand, int nnz, uint16_t* dist_out) {
    if (!c || nq < 0 || nt < 0 || nnz < 0) return orbx_set_error(ORBX_E_BAD_ARG, "bad argument");
    if (nnz == 0) return ORBX_OK;
    if (!q || !t || !row_ptr || !cand || !dist_out) return orbx_set_error(ORBX_E_BAD_ARG, "null argument");
    M_TRY(cudaSetDevice(c->device));
    int r;
    if ((r = c->q.ensure((size_t)nq * 32)) || (r = c->t.ensure((size_t)nt * 32)) ||
        (r = c->rowptr.ensure((size_t)(nq + 1) * 4)) || (r = c->cand.ensure((size_t)nnz * 4)) ||
        (r = c->dist.ensure((size_t)nnz * 2)))
        return r;
    M_TRY(ctx_h2d(c, c->q.p, q, (size_t)nq * 32));
    M_TRY(ctx_h2d(c, c->t.p, t, (size_t)nt * 32));
    M_TRY(ctx_h2d(c, c->rowptr.p, row_ptr, (size_t)(nq + 1) * 4));
    M_TRY(ctx_h2d(c, c->cand.p, cand, (size_t)nnz * 4));
    c->last_pairs = nnz;
    M_TRY(ctx_flush(c));
    M_TRY(cudaEventRecord(c->ev0, c->st));
    M_TRY(orbm_launch_hamming_csr((const uint8_t*)c->q.p, nq, (const uint8_t*)c->t.p, (const int32_t*)c->rowptr.p,
                                  (const int32_t*)c->cand.p, nnz, (uint16_t*)c->dist.p, c->st));
    M_TRY(cudaEventRecord(c->ev1, c->st));
    c->timed = true;
    M_TRY(ctx_d2h(c, dist_out, c->dist.p, (size_t)nnz * 2));
    M_TRY(ctx_sync(c));
    return ORBX_OK;
}

int orbm_knn2(orbm_context* c, const uint8_t* q, int nq, const uint8_t* t, int nt, int32_t* idx_out,
              int32_t* dist_out) {
    if (!c || nq < 0 || nt < 0) return orbx_set_error(ORBX_E_BAD_ARG, "bad argument");
    if (nq == 0) return ORBX_OK;
    if (!q || (nt && !t) || !idx_out || !dist_out) return orbx_set_error(ORBX_E_BAD_ARG, "null argument");
    if (nt >= (1 << 20)) return orbx_set_error(ORBX_E_TOO_LARGE, "more than 2^20 train descriptors");
    M_TRY(cudaSetDevice(c->device));
    int r;
    if ((r = c->q.ensure((size_t)nq * 32)) || (r = c->t.ensure((size_t)std::max(nt, 1) * 32)) ||
        (r = c->aux1.ensure((size_t)nq * 8)) || (r = c->aux2.ensure((size_t)nq * 8)))
        return r;
    M_TRY(ctx_h2d(c, c->q.p, q, (size_t)nq * 32));
    if (nt) M_TRY(ctx_h2d(c, c->t.p, t, (size_t)nt * 32));
    c->last_pairs = (long long)nq * nt;
    M_TRY(ctx_flush(c));
    M_TRY(cudaEventRecord(c->ev0, c->st));
    M_TRY(orbm_launch_knn2((const uint8_t*)c->q.p, nq, (const uint8_t*)c->t.p, nt, (int32_t*)c->aux1.p,
                           (int32_t*)c->aux2.p, c->st));
    M_TRY(cudaEventRecord(c->ev1, c->st));
    c->timed = true;
    M_TRY(ctx_d2h(c, idx_out, c->aux1.p, (size_t)nq * 8));
    M_TRY(ctx_d2h(c, dist_out, c->aux2.p, (size_t)nq * 8));
    M_TRY(ctx_sync(c));
    return ORBX_OK;
}

int orbm_stereo_rows(orbm_context* c, const orbx_keypoint* kps_l, const uint8_t* desc_l, int n_l,
                     const orbx_keypoint* kps_r, const uint8_t* desc_r, int n_r,
                     const float* scale_factors, int nlevels, float min_d, float max_d,
                     int32_t* best_idx, int32_t* best_dist) {
    if (!c || n_l < 0 || n_r < 0 || nlevels < 1) return orbx_set_error(ORBX_E_BAD_ARG, "bad argument");
    if (n_l == 0) return ORBX_OK;
    if (!kps_l || !desc_l || (n_r && (!kps_r || !desc_r)) || !scale_factors || !best_idx || !best_dist)
        return orbx_set_error(ORBX_E_BAD_ARG, "null argument");
    if (n_r >= (1 << 20)) return orbx_set_error(ORBX_E_TOO_LARGE, "more than 2^20 right keypoints");
    for (int i = 0; i < n_r; ++i)
        if (kps_r[i].octave < 0 || kps_r[i].octave >= nlevels) return orbx_set_error(ORBX_E_BAD_ARG, "octave out of range");
    M_TRY(cudaSetDevice(c->device));
    int r;
    if ((r = c->q.ensure((size_t)n_l * 32)) || (r = c->t.ensure((size_t)std::max(n_r, 1) * 32)) ||
        (r = c->kl.ensure((size_t)n_l * sizeof(orbx_keypoint))) ||
        (r = c->kr.ensure((size_t)std::max(n_r, 1) * sizeof(orbx_keypoint))) ||
        (r = c->sc.ensure((size_t)nlevels * 4)) || (r = c->aux1.ensure((size_t)n_l * 4)) ||
        (r = c->aux2.ensure((size_t)n_l * 4)))
        return r;
    M_TRY(ctx_h2d(c, c->q.p, desc_l, (size_t)n_l * 32));
    M_TRY(ctx_h2d(c, c->kl.p, kps_l, (size_t)n_l * sizeof(orbx_keypoint)));
    if (n_r) {
        M_TRY(ctx_h2d(c, c->t.p, desc_r, (size_t)n_r * 32));
        M_TRY(ctx_h2d(c, c->kr.p, kps_r, (size_t)n_r * sizeof(orbx_keypoint)));
    }
    M_TRY(ctx_h2d(c, c->sc.p, scale_factors, (size_t)nlevels * 4));
    c->last_pairs = (long long)n_l * n_r;
    M_TRY(ctx_flush(c));
    M_TRY(cudaEventRecord(c->ev0, c->st));
    M_TRY(orbm_launch_stereo_rows((const orbx_keypoint*)c->kl.p, (const uint8_t*)c->q.p, n_l,
                                  (const orbx_keypoint*)c->kr.p, (const uint8_t*)c->t.p, n_r, (const float*)c->sc.p,
                                  min_d, max_d, (int32_t*)c->aux1.p, (int32_t*)c->aux2.p, c->st));
    M_TRY(cudaEventRecord(c->ev1, c->st));
    c->timed = true;
    M_TRY(ctx_d2h(c, best_idx, c->aux1.p, (size_t)n_l * 4));
    M_TRY(ctx_d2h(c, best_dist, c->aux2.p, (size_t)n_l * 4));
    M_TRY(ctx_sync(c));
    return ORBX_OK;
}

// Frame::ComputeStereoMatches, complete (src/Frame.cc:802-980): row-band Hamming argmin
// (k_stereo_rows), SAD sub-pixel refinement on the two device-resident pyramids
// (k_stereo_refine), then the median-based outlier cut of :966-979 on the host (a k-th element
// over <= N integers).
int orbm_compute_stereo_matches(orbm_context* c, const orbx_keypoint* kps_l, const uint8_t* desc_l, int n_l,
                                const orbx_keypoint* kps_r, const uint8_t* desc_r, int n_r,
                                const float* scale_factors, const float* inv_scale_factors, int nlevels,
                                const orbx_pyramid_view* pyr_l, const orbx_pyramid_view* pyr_r, float mb,
                                float mbf, float* u_right, float* depth) {
    if (!c || n_l < 0 || n_r < 0 || nlevels < 1 || nlevels > ORBX_MAX_LEVELS)
        return orbx_set_error(ORBX_E_BAD_ARG, "bad argument");
    if (n_l == 0) return ORBX_OK;
    if (!kps_l || !desc_l || (n_r && (!kps_r || !desc_r)) || !scale_factors || !inv_scale_factors ||
        !pyr_l || !pyr_r || !u_right || !depth)
        return orbx_set_error(ORBX_E_BAD_ARG, "null argument");
    if (pyr_l->device != c->device || pyr_r->device != c->device)
        return orbx_set_error(ORBX_E_BAD_ARG, "pyramids live on another device than the matcher context");
    if (pyr_l->nlevels < nlevels || pyr_r->nlevels < nlevels) return orbx_set_error(ORBX_E_BAD_ARG, "pyramid has fewer levels");
    if (n_r >= (1 << 20)) return orbx_set_error(ORBX_E_TOO_LARGE, "more than 2^20 right keypoints");
    for (int i = 0; i < n_r; ++i)
        if (kps_r[i].octave < 0 || kps_r[i].octave >= nlevels) return orbx_set_error(ORBX_E_BAD_ARG, "octave out of range");
    for (int i = 0; i < n_l; ++i)
        if (kps_l[i].octave < 0 || kps_l[i].octave >= nlevels) return orbx_set_error(ORBX_E_BAD_ARG, "octave out of range");
    M_TRY(cudaSetDevice(c->device));
    int r;
    if ((r = c->q.ensure((size_t)n_l * 32)) || (r = c->t.ensure((size_t)std::max(n_r, 1) * 32)) ||
        (r = c->kl.ensure((size_t)n_l * sizeof(orbx_keypoint))) ||
        (r = c->kr.ensure((size_t)std::max(n_r, 1) * sizeof(orbx_keypoint))) ||
        (r = c->sc.ensure((size_t)nlevels * 4)) || (r = c->isc.ensure((size_t)nlevels * 4)) ||
        (r = c->aux1.ensure((size_t)n_l * 4)) || (r = c->aux2.ensure((size_t)n_l * 4)) ||
        (r = c->ur.ensure((size_t)n_l * 4)) || (r = c->dp.ensure((size_t)n_l * 4)) ||
        (r = c->sd.ensure((size_t)n_l * 4)))
        return r;
    M_TRY(ctx_h2d(c, c->q.p, desc_l, (size_t)n_l * 32));
    M_TRY(ctx_h2d(c, c->kl.p, kps_l, (size_t)n_l * sizeof(orbx_keypoint)));
    if (n_r) {
        M_TRY(ctx_h2d(c, c->t.p, desc_r, (size_t)n_r * 32));
        M_TRY(ctx_h2d(c, c->kr.p, kps_r, (size_t)n_r * sizeof(orbx_keypoint)));
    }
    M_TRY(ctx_h2d(c, c->sc.p, scale_factors, (size_t)nlevels * 4));
    M_TRY(ctx_h2d(c, c->isc.p, inv_scale_factors, (size_t)nlevels * 4));
    // :831-833
    const float minZ = mb, minD = 0, maxD = mbf / minZ;
    const int thOrbDist = (ORBM_TH_HIGH + ORBM_TH_LOW) / 2;   // :807
    c->last_pairs = (long long)n_l * n_r;
    M_TRY(ctx_flush(c));
    M_TRY(cudaEventRecord(c->ev0, c->st));
    M_TRY(orbm_launch_stereo_rows((const orbx_keypoint*)c->kl.p, (const uint8_t*)c->q.p, n_l,
                                  (const orbx_keypoint*)c->kr.p, (const uint8_t*)c->t.p, n_r, (const float*)c->sc.p,
                                  minD, maxD, (int32_t*)c->aux1.p, (int32_t*)c->aux2.p, c->st));
    M_TRY(orbm_launch_stereo_refine((const orbx_keypoint*)c->kl.p, n_l, (const orbx_keypoint*)c->kr.p,
                                    (const int32_t*)c->aux1.p, (const int32_t*)c->aux2.p, pyr_l, pyr_r,
                                    (const float*)c->sc.p, (const float*)c->isc.p, thOrbDist, minD, maxD, mbf,
                                    (float*)c->ur.p, (float*)c->dp.p, (int32_t*)c->sd.p, c->st));
    M_TRY(cudaEventRecord(c->ev1, c->st));
    c->timed = true;
    std::vector<int32_t> sad(n_l);
    M_TRY(ctx_d2h(c, u_right, c->ur.p, (size_t)n_l * 4));
    M_TRY(ctx_d2h(c, depth, c->dp.p, (size_t)n_l * 4));
    M_TRY(ctx_d2h(c, sad.data(), c->sd.p, (size_t)n_l * 4));
    M_TRY(ctx_sync(c));
    // :966-979: median of the accepted SAD distances, reject >= 1.5 * 1.4 * median
    std::vector<int32_t> acc;
    acc.reserve(n_l);
    for (int i = 0; i < n_l; ++i)
        if (sad[i] >= 0) acc.push_back(sad[i]);
    if (!acc.empty()) {   // (the reference indexes an empty vector here)
        std::nth_element(acc.begin(), acc.begin() + acc.size() / 2, acc.end());
        const float median = (float)acc[acc.size() / 2];
        const float thDist = 1.5f * 1.4f * median;
        for (int i = 0; i < n_l; ++i)
            if (sad[i] >= 0 && !((float)sad[i] < thDist)) {
                u_right[i] = -1;
                depth[i] = -1;
            }
    }
    return ORBX_OK;
}

}  // extern "C"

// =================================================================================================
// layer 2: the reference's searches on POD snapshots
// =================================================================================================
namespace {

// ORBmatcher::ComputeThreeMaxima (src/ORBmatcher.cc:2307-2348)
void three_maxima(const std::vector<int>* histo, int L, int& ind1, int& ind2, int& ind3) {
    int max1 = 0, max2 = 0, max3 = 0;
    for (int i = 0; i < L; i++) {
        const int s = (int)histo[i].size();
        if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
        else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
        else if (s > max3) { max3 = s; ind3 = i; }
    }
    if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
    else if (max3 < 0.1f * (float)max1) { ind3 = -1; }
}

int check_view(const orbm_frame_view* F) {
    if (!F || F->n < 0 || (F->n && (!F->keys || !F->desc)) || !F->scale_factors || F->nlevels < 1)
        return orbx_set_error(ORBX_E_BAD_ARG, "bad frame view");
    return ORBX_OK;
}

// Candidate generation + distances of a search on the device (N2): grid build, window queries in
// the reference's order, Hamming distance per (query, candidate).  Leaves the CSR lists in
// c->h_rowptr / c->h_cand and the distances in c->hdist.
// topk > 0: the rows' full lists stay on the device (window_full_lists() fetches them on demand) and the
// host gets k_csr_topk's topk best (distance, position) keys + candidates per row instead.
// The queries of a window search, built in place in the pinned arena when they fit (else in `own`).
struct QueryBuf {
    orbm_area_query* q = nullptr;            // where the caller writes them
    const orbm_area_query* dev = nullptr;    // staged: where the kernels will read them
    bool staged = false;
    std::vector<orbm_area_query> own;
};
int device_candidates_onepass(orbm_context* c, const orbm_frame_view* F, const QueryBuf& qb, int nq, const uint8_t* qdesc,
                              int n_qdesc, const uint8_t* d_qdesc);

// q_staged: the queries are already in the arena (stage_queries).  d_qdesc: the queries' descriptors are rows of a
// device-resident bank (qdesc is then null and nothing is sent for them).
// Without top-K the lists come from the one-launch search (device_candidates_onepass) and are put into CSR order
// here, on the host (a copy of nnz entries: a few microseconds against two launches and a scatter); with top-K
// (SearchForInitialization) the count -> scan -> fill -> top-K chain below runs, whose lists stay on the device.
int device_candidates(orbm_context* c, const orbm_frame_view* F, const orbm_area_query* Qp, int nq, bool q_staged,
                      const uint8_t* qdesc, int n_qdesc, const uint8_t* d_qdesc, int topk = 0) {
#ifndef ORBM_NO_ONEPASS   // (A/B switch: the three-launch path for every search)
    if (topk == 0) {
        QueryBuf qb;
        qb.q = const_cast<orbm_area_query*>(Qp);
        qb.staged = q_staged;
        qb.dev = q_staged ? (const orbm_area_query*)c->aq.p : nullptr;
        int rc = device_candidates_onepass(c, F, qb, nq, qdesc, n_qdesc, d_qdesc);
        if (rc) return rc;
        c->h_topk.clear();
        c->h_topc.clear();
        c->h_rowptr.resize((size_t)nq + 1);
        c->h_rowptr[0] = 0;
        const uint2* row = c->h_rowinfo.data();
        for (int i = 0; i < nq; ++i) c->h_rowptr[i + 1] = c->h_rowptr[i] + (int32_t)row[i].y;
        const size_t nnz = (size_t)c->h_rowptr[nq];
        c->h_cand2.resize(nnz);
        c->hdist2.resize(nnz);
        for (int i = 0; i < nq; ++i) {
            const uint32_t s0 = row[i].x, n = row[i].y;
            int32_t* dc = c->h_cand2.data() + c->h_rowptr[i];
            uint16_t* dd = c->hdist2.data() + c->h_rowptr[i];
            for (uint32_t k = 0; k < n; ++k) {
                dc[k] = c->h_cand[s0 + k];
                dd[k] = c->hdist[s0 + k];
            }
        }
        c->h_cand.swap(c->h_cand2);
        c->hdist.swap(c->hdist2);
        return ORBX_OK;
    }
#endif
    ORBM_PHASE_BEGIN("device_candidates");
    c->h_rowptr.assign(nq + 1, 0);
    c->h_cand.clear();
    c->hdist.clear();
    c->have_full = true;
    c->h_topk.assign((size_t)nq * topk, 0xffffffffu);
    c->h_topc.assign((size_t)nq * topk, -1);
    c->last_pairs = 0;
    if (nq == 0 || F->n == 0) {
        if (q_staged) ctx_unstage(c);
        return ORBX_OK;
    }
    M_TRY(cudaSetDevice(c->device));
    // a device-resident frame (orbm_frame_upload / orbm_frame_from_device): keypoints, descriptors and the grid are
    // already there -- no upload of the frame, no grid build
    const orbm_frame* RF = nullptr;
    {
        auto it = c->frames.find(F);
        if (it != c->frames.end()) RF = it->second;
    }
    int r;
    if ((r = c->q.ensure((size_t)std::max(n_qdesc, 1) * 32)) || (r = c->aq.ensure((size_t)nq * sizeof(orbm_area_query))) ||
        (r = c->acnt.ensure((size_t)nq * 4)) || (r = c->rowptr.ensure((size_t)(nq + 1) * 4)))
        return r;
    if (!RF && ((r = c->kl.ensure((size_t)F->n * sizeof(orbx_keypoint))) || (r = c->t.ensure((size_t)F->n * 32)) ||
                (r = c->gstart.ensure((size_t)(ORBM_GRID_COLS * ORBM_GRID_ROWS + 1) * 4)) ||
                (r = c->gitems.ensure((size_t)F->n * 4)) || (r = c->gcell.ensure((size_t)F->n * 4))))
        return r;
    if (!RF) {
        M_TRY(ctx_h2d(c, c->kl.p, F->keys, (size_t)F->n * sizeof(orbx_keypoint)));
        M_TRY(ctx_h2d(c, c->t.p, F->desc, (size_t)F->n * 32));
    }
    if (qdesc) M_TRY(ctx_h2d(c, c->q.p, qdesc, (size_t)n_qdesc * 32));
    if (!q_staged) M_TRY(ctx_h2d(c, c->aq.p, Qp, (size_t)nq * sizeof(orbm_area_query)));
    ORBM_PHASE("uploads");
    const orbx_keypoint* dk = RF ? (const orbx_keypoint*)RF->d_keys.p : (const orbx_keypoint*)c->kl.p;
    const uint8_t* dt = RF ? (const uint8_t*)RF->d_desc.p : (const uint8_t*)c->t.p;
    const int32_t* gstart = RF ? (const int32_t*)RF->d_start.p : (const int32_t*)c->gstart.p;
    const int32_t* gitems = RF ? (const int32_t*)RF->d_items.p : (const int32_t*)c->gitems.p;
    M_TRY(ctx_flush(c));
    ORBM_PHASE("flush (one copy + scatter queued)");
    M_TRY(cudaEventRecord(c->ev0, c->st));
    if (!RF)
        M_TRY(orbm_launch_grid_build(dk, F->n, F->min_x, F->min_y, F->grid_w_inv, F->grid_h_inv, (int32_t*)c->gstart.p,
                                     (int32_t*)c->gitems.p, (int32_t*)c->gcell.p, c->st));
    M_TRY(orbm_launch_area_count(dk, gstart, gitems, F->min_x, F->min_y,
                                 F->grid_w_inv, F->grid_h_inv, (const orbm_area_query*)c->aq.p, nq,
                                 (int32_t*)c->acnt.p, (int32_t*)c->rowptr.p, c->st));
    // One blocking wait per search: the candidate / distance lists are sized from the last search's pair
    // count (twice that, at least 4096) before this one's counts are known, k_area<1> writes inside that
    // capacity, and the counts come back together with the results.  A search that outgrows the guess
    // (first call, larger windows) reruns the fill with lists of the right size.
    {
        const size_t guess = std::max<size_t>(4096, 2 * (size_t)c->nnz_hint);
        if ((r = c->cand.ensure(guess * 4)) || (r = c->dist.ensure(guess * 2))) return r;
    }
    if (topk > 0 && ((r = c->aux1.ensure((size_t)nq * topk * 4)) || (r = c->aux2.ensure((size_t)nq * topk * 4)))) return r;
    int nnz = 0;
    for (int pass = 0; pass < 2; ++pass) {
        const int cap = (int)std::min<size_t>({c->cand.cap / 4, c->dist.cap / 2, (size_t)INT_MAX});
        M_TRY(orbm_launch_area_fill(dk, gstart, gitems, F->min_x, F->min_y,
                                    F->grid_w_inv, F->grid_h_inv, (const orbm_area_query*)c->aq.p, nq,
                                    (const int32_t*)c->rowptr.p,
                                    d_qdesc ? d_qdesc : (qdesc ? (const uint8_t*)c->q.p : dt),
                                    dt, (int32_t*)c->cand.p, (uint16_t*)c->dist.p, cap, c->st));
        if (topk > 0)
            M_TRY(orbm_launch_csr_topk((const int32_t*)c->rowptr.p, (const int32_t*)c->cand.p, (const uint16_t*)c->dist.p,
                                       nq, topk, cap, (uint32_t*)c->aux1.p, (int32_t*)c->aux2.p, c->st));
        M_TRY(cudaEventRecord(c->ev1, c->st));
        c->timed = true;
        if (pass == 0) M_TRY(ctx_d2h(c, c->h_rowptr.data(), c->rowptr.p, (size_t)(nq + 1) * 4));
        // without top-K the lists themselves are wanted: fetch what the last search would have needed (+ slack)
        // now, the rest (if any) once the count is known
        size_t spec = 0;
        if (topk > 0) {
            M_TRY(ctx_d2h(c, c->h_topk.data(), c->aux1.p, (size_t)nq * topk * 4));
            M_TRY(ctx_d2h(c, c->h_topc.data(), c->aux2.p, (size_t)nq * topk * 4));
        } else {
            spec = pass == 0 ? std::min<size_t>({(size_t)cap, (size_t)c->nnz_hint + (size_t)c->nnz_hint / 8 + 1024, (size_t)65536})
                             : (size_t)nnz;   // at most 384 KB fetched on a guess
            c->h_cand.resize(spec);
            c->hdist.resize(spec);
            M_TRY(ctx_d2h(c, c->h_cand.data(), c->cand.p, spec * 4));
            M_TRY(ctx_d2h(c, c->hdist.data(), c->dist.p, spec * 2));
        }
        ORBM_PHASE("launches + downloads queued");
        M_TRY(ctx_sync(c));
        ORBM_PHASE("wait + deliver");
        nnz = c->h_rowptr[nq];
        c->last_pairs = nnz;
        c->nnz_hint = nnz;
        if (nnz > cap) {   // the guess was too small: lists of the right size, fill again
            if ((r = c->cand.ensure((size_t)nnz * 4)) || (r = c->dist.ensure((size_t)nnz * 2))) return r;
            continue;
        }
        if (topk > 0) {
            c->have_full = nnz == 0;
        } else {
            if ((size_t)nnz > spec) {   // the tail the speculative fetch did not cover
                c->h_cand.resize(nnz);
                c->hdist.resize(nnz);
                M_TRY(ctx_d2h(c, c->h_cand.data() + spec, (const int32_t*)c->cand.p + spec, ((size_t)nnz - spec) * 4));
                M_TRY(ctx_d2h(c, c->hdist.data() + spec, (const uint16_t*)c->dist.p + spec, ((size_t)nnz - spec) * 2));
                M_TRY(ctx_sync(c));
            }
            c->h_cand.resize(nnz);
            c->hdist.resize(nnz);
        }
        break;
    }
    ORBM_PHASE("tail");
    return ORBX_OK;
}

int device_candidates(orbm_context* c, const orbm_frame_view* F, const std::vector<orbm_area_query>& Q,
                      const uint8_t* qdesc, int n_qdesc, int topk = 0) {
    return device_candidates(c, F, Q.data(), (int)Q.size(), false, qdesc, n_qdesc, nullptr, topk);
}

// alias: the kernels read the queries in the arena's device mirror (no scatter launch); else they go to c->aq.
int stage_queries(orbm_context* c, int n, QueryBuf* b, bool alias = false) {
    int r;
    M_TRY(cudaSetDevice(c->device));
    if ((r = c->aq.ensure((size_t)std::max(n, 1) * sizeof(orbm_area_query)))) return r;
    const void* dev = nullptr;
    if (alias) b->q = (orbm_area_query*)ctx_stage_alias(c, (size_t)n * sizeof(orbm_area_query), &dev);
    if (b->q) {
        b->dev = (const orbm_area_query*)dev;
    } else {
        b->q = (orbm_area_query*)ctx_stage(c, c->aq.p, (size_t)n * sizeof(orbm_area_query));
        b->dev = (const orbm_area_query*)c->aq.p;
    }
    b->staged = b->q != nullptr;
    if (!b->staged) {
        b->own.resize(n);
        b->q = b->own.data();
        b->dev = nullptr;
    }
    return ORBX_OK;
}

// While a call reads staged data in place -- queries / descriptors in the arena's device mirror, results in the arena
// itself -- neither may be replaced: ctx_sync leaves the arena alone, and it grows (if the call wanted more than it
// holds) when the guard goes out of scope.
struct ArenaHold {
    orbm_context* c;
    explicit ArenaHold(orbm_context* c_) : c(c_) { c->hold_arena = true; }
    ~ArenaHold() {
        c->hold_arena = false;
        if (!g_staging_ctx) ctx_grow(c);   // (not between a failed call's staging and its abort)
    }
};

// Candidates + distances of the two tracking searches in ONE kernel (k_area_onepass): no count launch, no row scan,
// and -- with the queries (and host descriptors) read in place in the arena's mirror -- no scatter launch either:
// one H2D copy, one kernel, the downloads.  Leaves {first position, count} per query in c->h_rowinfo and the lists
// in c->h_cand / c->hdist (rows in the order the warps reserved them, a row's candidates in the reference's order).
int device_candidates_onepass(orbm_context* c, const orbm_frame_view* F, const QueryBuf& qb, int nq, const uint8_t* qdesc,
                              int n_qdesc, const uint8_t* d_qdesc) {
    ORBM_PHASE_BEGIN("device_candidates_onepass");
    c->h_rowinfo.assign(nq, make_uint2(0, 0));
    c->h_cand.clear();
    c->hdist.clear();
    c->have_full = true;
    c->last_pairs = 0;
    if (nq == 0 || F->n == 0) {
        if (qb.staged) ctx_unstage(c);
        return ORBX_OK;
    }
    M_TRY(cudaSetDevice(c->device));
    ArenaHold hold(c);
    const orbm_frame* RF = nullptr;
    {
        auto it = c->frames.find(F);
        if (it != c->frames.end()) RF = it->second;
    }
    int r;
    // ONE device buffer: [counter (64 B) | {first, count} per query | {candidate, distance} records]: one download
    // brings the counter, the rows and the records the last search would have needed (+ slack)
    const size_t rows_off = 64, pairs_off = rows_off + (((size_t)nq * 8 + 63) & ~(size_t)63);
    auto ensure_one = [&](size_t records) -> int {
        const size_t before = c->counter.cap;
        int rr = c->counter.ensure(pairs_off + records * 8);
        if (rr) return rr;
        if (c->counter.cap != before) c->counter_known = false;   // a new buffer: its counter starts at an unknown value
        return ORBX_OK;
    };
    if (!RF && ((r = c->kl.ensure((size_t)F->n * sizeof(orbx_keypoint))) || (r = c->t.ensure((size_t)F->n * 32)) ||
                (r = c->gstart.ensure((size_t)(ORBM_GRID_COLS * ORBM_GRID_ROWS + 1) * 4)) ||
                (r = c->gitems.ensure((size_t)F->n * 4)) || (r = c->gcell.ensure((size_t)F->n * 4))))
        return r;
    if (!RF) {
        M_TRY(ctx_h2d(c, c->kl.p, F->keys, (size_t)F->n * sizeof(orbx_keypoint)));
        M_TRY(ctx_h2d(c, c->t.p, F->desc, (size_t)F->n * 32));
    }
    const orbx_keypoint* dk = RF ? (const orbx_keypoint*)RF->d_keys.p : (const orbx_keypoint*)c->kl.p;
    const uint8_t* dt = RF ? (const uint8_t*)RF->d_desc.p : (const uint8_t*)c->t.p;
    const int32_t* gstart = RF ? (const int32_t*)RF->d_start.p : (const int32_t*)c->gstart.p;
    const int32_t* gitems = RF ? (const int32_t*)RF->d_items.p : (const int32_t*)c->gitems.p;
    const uint8_t* dq = d_qdesc ? d_qdesc : dt;
    if (!d_qdesc && qdesc) {   // host descriptors: into the mirror, read in place
        const void* dev = nullptr;
        void* h = ctx_stage_alias(c, (size_t)n_qdesc * 32, &dev);
        if (h) {
            memcpy(h, qdesc, (size_t)n_qdesc * 32);
            dq = (const uint8_t*)dev;
        } else {
            if ((r = c->q.ensure((size_t)std::max(n_qdesc, 1) * 32))) return r;
            M_TRY(ctx_h2d(c, c->q.p, qdesc, (size_t)n_qdesc * 32));
            dq = (const uint8_t*)c->q.p;
        }
    }
    const orbm_area_query* dQ = qb.dev;
    if (!qb.staged) {
        if ((r = c->aq.ensure((size_t)nq * sizeof(orbm_area_query)))) return r;
        M_TRY(ctx_h2d(c, c->aq.p, qb.q, (size_t)nq * sizeof(orbm_area_query)));
        dQ = (const orbm_area_query*)c->aq.p;
    }
    // records sized from the last search's pair count, as in the three-launch path
    if ((r = ensure_one(std::max<size_t>(4096, 2 * (size_t)c->nnz_hint)))) return r;
    ORBM_PHASE("uploads");
    M_TRY(ctx_flush(c));
    ORBM_PHASE("flush");
    M_TRY(cudaEventRecord(c->ev0, c->st));
    if (!RF)
        M_TRY(orbm_launch_grid_build(dk, F->n, F->min_x, F->min_y, F->grid_w_inv, F->grid_h_inv, (int32_t*)c->gstart.p,
                                     (int32_t*)c->gitems.p, (int32_t*)c->gcell.p, c->st));
    uint32_t nnz = 0;
    for (int pass = 0; pass < 2; ++pass) {
        if (!c->counter_known) {
            M_TRY(cudaMemsetAsync(c->counter.p, 0, 64, c->st));
            c->counter_host = 0;
            c->counter_known = true;
        }
        uint8_t* base = (uint8_t*)c->counter.p;
        const int cap = (int)std::min<size_t>((c->counter.cap - pairs_off) / 8, (size_t)0x3fffffff);
        M_TRY(orbm_launch_area_onepass(dk, gstart, gitems, F->min_x, F->min_y, F->grid_w_inv, F->grid_h_inv, dQ, nq,
                                       (uint2*)(base + rows_off), (uint32_t*)base, c->counter_host, dq, dt,
                                       (uint2*)(base + pairs_off), cap, c->st));
        M_TRY(cudaEventRecord(c->ev1, c->st));
        c->timed = true;
        const size_t spec = pass == 0 ? std::min<size_t>({(size_t)cap, (size_t)c->nnz_hint + (size_t)c->nnz_hint / 8 + 1024, (size_t)65536})
                                      : (size_t)nnz;
        const uint8_t* view = nullptr;
        M_TRY(ctx_d2h_view(c, base, pairs_off + spec * 8, &view));
        if (!view) {   // the arena cannot take it (first calls, very large searches): an array of this call's own
            c->h_one.resize(pairs_off + spec * 8);
            M_TRY(ctx_d2h(c, c->h_one.data(), base, pairs_off + spec * 8));
        }
        ORBM_PHASE("launch + download queued");
        M_TRY(ctx_sync(c));
        ORBM_PHASE("wait");
        const bool in_arena = view != nullptr;
        if (!view) view = c->h_one.data();
        uint32_t after;
        memcpy(&after, view, 4);
        nnz = after - c->counter_host;
        c->counter_host = after;
        c->last_pairs = nnz;
        c->nnz_hint = (int)std::min<uint32_t>(nnz, 0x3fffffffu);
        if (nnz > (uint32_t)cap) {   // the guess was too small: room for all of them, once more
            if (nnz > 0x3fffffffu) return orbx_set_error(ORBX_E_CAPACITY, "%u candidate pairs", nnz);
            if ((r = ensure_one(nnz))) return r;
            continue;
        }
        memcpy(c->h_rowinfo.data(), view + rows_off, (size_t)nq * 8);
        c->h_cand.resize(nnz);
        c->hdist.resize(nnz);
        const size_t have = std::min<size_t>(nnz, spec);
        const uint2* rec = reinterpret_cast<const uint2*>(view + pairs_off);
        for (size_t k = 0; k < have; ++k) {
            c->h_cand[k] = (int32_t)rec[k].x;
            c->hdist[k] = (uint16_t)rec[k].y;
        }
        (void)in_arena;
        if ((size_t)nnz > spec) {   // the tail the speculative fetch did not cover
            c->h_one.resize(((size_t)nnz - spec) * 8);
            M_TRY(ctx_d2h(c, c->h_one.data(), base + pairs_off + spec * 8, ((size_t)nnz - spec) * 8));
            M_TRY(ctx_sync(c));
            const uint2* rec2 = reinterpret_cast<const uint2*>(c->h_one.data());
            for (size_t k = spec; k < nnz; ++k) {
                c->h_cand[k] = (int32_t)rec2[k - spec].x;
                c->hdist[k] = (uint16_t)rec2[k - spec].y;
            }
        }
        break;
    }
    ORBM_PHASE("tail");
    return ORBX_OK;
}

// The full candidate / distance lists of the last top-K search (c->h_cand, c->hdist), fetched on the first
// row that needs them.
int window_full_lists(orbm_context* c) {
    if (c->have_full) return ORBX_OK;
    const int nnz = c->h_rowptr.back();
    c->h_cand.resize(nnz);
    c->hdist.resize(nnz);
    M_TRY(ctx_d2h(c, c->h_cand.data(), c->cand.p, (size_t)nnz * 4));
    M_TRY(ctx_d2h(c, c->hdist.data(), c->dist.p, (size_t)nnz * 2));
    M_TRY(ctx_sync(c));
    c->have_full = true;
    return ORBX_OK;
}

}  // namespace

extern "C" {

int orbm_last_device_ms(orbm_context* c, float* ms, long long* pairs) {
    if (!c || !ms) return orbx_set_error(ORBX_E_BAD_ARG, "null argument");
    *ms = 0.f;
    if (pairs) *pairs = c->last_pairs;
    if (!c->timed) return ORBX_OK;
    M_TRY(cudaSetDevice(c->device));
    M_TRY(cudaEventSynchronize(c->ev1));
    M_TRY(cudaEventElapsedTime(ms, c->ev0, c->ev1));
    return ORBX_OK;
}

int orbm_measure_popc_peak(orbm_context* c, double* gpopc_per_s) {
    if (!c || !gpopc_per_s) return orbx_set_error(ORBX_E_BAD_ARG, "null argument");
    M_TRY(cudaSetDevice(c->device));
    cudaDeviceProp prop;
    M_TRY(cudaGetDeviceProperties(&prop, c->device));
    const int blocks = prop.multiProcessorCount * 8, iters = 4096;
    int r;
    if ((r = c->aux1.ensure((size_t)blocks * 256 * 4))) return r;
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        M_TRY(cudaEventRecord(c->ev0, c->st));
        M_TRY(orbm_launch_popc_peak((uint32_t*)c->aux1.p, blocks, iters, c->st));
        M_TRY(cudaEventRecord(c->ev1, c->st));
        M_TRY(cudaEventSynchronize(c->ev1));
        float ms = 0.f;
        M_TRY(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
        if (rep > 0 && ms < best) best = ms;
    }
    *gpopc_per_s = (double)blocks * 256 * 8 * iters / (best * 1e-3) * 1e-9;
    c->timed = false;
    return ORBX_OK;
}

int orbm_area_candidates(orbm_context* c, const orbm_frame_view* F, const orbm_area_query* q, int nq,
                         const uint8_t* q_desc, int n_q_desc, int32_t* row_ptr, int32_t* cand, uint16_t* dist,
                         int cap) {
    int rc = check_view(F);
    if (rc) return rc;
    if (!c || nq < 0 || (nq && !q) || !row_ptr || cap < 0 || (q_desc && n_q_desc <= 0))
        return orbx_set_error(ORBX_E_BAD_ARG, "bad argument");
    for (int i = 0; i < nq; ++i)
        if (q_desc && q[i].valid && (q[i].desc_row < 0 || q[i].desc_row >= n_q_desc))
            return orbx_set_error(ORBX_E_BAD_ARG, "query %d: desc_row out of range", i);
    std::vector<orbm_area_query> Q(q, q + nq);
    if (!q_desc)
        for (auto& e : Q) e.desc_row = 0;
    if ((rc = device_candidates(c, F, Q, q_desc, n_q_desc))) return rc;
    for (int i = 0; i <= nq; ++i) row_ptr[i] = c->h_rowptr[i];
    const int nnz = c->h_rowptr[nq];
    if (nnz > cap) return orbx_set_error(ORBX_E_CAPACITY, "%d candidates exceed the capacity %d", nnz, cap);
    for (int k = 0; k < nnz; ++k) {
        if (cand) cand[k] = c->h_cand[k];
        if (dist && q_desc) dist[k] = c->hdist[k];
    }
    return nnz;
}

// ---- device-resident frames -------------------------------------------------------
static orbm_frame* frame_new(orbm_context* c) {
    orbm_frame* f;
    if (!c->frame_pool.empty()) {   // a frame per camera frame: its buffers come from the one that just retired
        f = c->frame_pool.back();
        c->frame_pool.pop_back();
        f->view = orbm_frame_view{};
        f->h_uright.clear();
    } else {
        f = new orbm_frame();
    }
    f->ctx = c;
    return f;
}
static void frame_retire(orbm_context* c, orbm_frame* f) {
    if (c->frame_pool.size() < 4) {
        c->frame_pool.push_back(f);
        return;
    }
    for (DevBuf* b : {&f->d_keys, &f->d_desc, &f->d_start, &f->d_items, &f->d_cell}) b->release();
    delete f;
}
struct FrameGuard {   // a frame under construction goes back to the pool on an early return
    orbm_context* c;
    orbm_frame* f;
    ~FrameGuard() { if (f) frame_retire(c, f); }
    orbm_frame* release() { orbm_frame* r = f; f = nullptr; return r; }
    orbm_frame* operator->() { return f; }
    orbm_frame* get() { return f; }
};

static int frame_finish(orbm_context* c, orbm_frame* f, int n) {
    // grid of the resident keypoints, built once (Frame::AssignFeaturesToGrid, src/Frame.cc:377-408)
    int r;
    if ((r = f->d_start.ensure((size_t)(ORBM_GRID_COLS * ORBM_GRID_ROWS + 1) * 4)) ||
        (r = f->d_items.ensure((size_t)std::max(n, 1) * 4)) || (r = f->d_cell.ensure((size_t)std::max(n, 1) * 4)))
        return r;
    M_TRY(orbm_launch_grid_build((const orbx_keypoint*)f->d_keys.p, n, f->view.min_x, f->view.min_y, f->view.grid_w_inv,
                                 f->view.grid_h_inv, (int32_t*)f->d_start.p, (int32_t*)f->d_items.p, (int32_t*)f->d_cell.p,
                                 c->st));
    M_TRY(cudaStreamSynchronize(c->st));
    c->frames[&f->view] = f;
    f->bank.ctx = c;
    f->bank.d = (const uint8_t*)f->d_desc.p;
    f->bank.rows = n;
    return ORBX_OK;
}

int orbm_frame_upload(orbm_context* c, const orbm_frame_view* view, orbm_frame** out) {
    if (!c || !out) return orbx_set_error(ORBX_E_BAD_ARG, "null argument");
    *out = nullptr;
    int rc = check_view(view);
    if (rc) return rc;
    M_TRY(cudaSetDevice(c->device));
    FrameGuard f{c, frame_new(c)};
    const int n = view->n;
    f->h_keys.assign(view->keys, view->keys + n);
    f->h_desc.assign(view->desc, view->desc + (size_t)n * 32);
    f->h_scale.assign(view->scale_factors, view->scale_factors + view->nlevels);
    if (view->u_right) f->h_uright.assign(view->u_right, view->u_right + n);
    f->view = *view;
    f->view.keys = f->h_keys.data();
    f->view.desc = f->h_desc.data();
    f->view.scale_factors = f->h_scale.data();
    f->view.u_right = view->u_right ? f->h_uright.data() : nullptr;
    if ((rc = f->d_keys.ensure((size_t)std::max(n, 1) * sizeof(orbx_keypoint))) || (rc = f->d_desc.ensure((size_t)std::max(n, 1) * 32)))
        return rc;
    if (n) {
        M_TRY(cudaMemcpyAsync(f->d_keys.p, f->h_keys.data(), (size_t)n * sizeof(orbx_keypoint), cudaMemcpyHostToDevice, c->st));
        M_TRY(cudaMemcpyAsync(f->d_desc.p, f->h_desc.data(), (size_t)n * 32, cudaMemcpyHostToDevice, c->st));
    }
    if ((rc = frame_finish(c, f.get(), n))) return rc;
    *out = f.release();
    return ORBX_OK;
}

int orbm_frame_from_device(orbm_context* c, const orbx_keypoint* d_keys, const uint8_t* d_desc, int n,
                           const orbm_camera* cam, int width, int height, const float* scale_factors, int nlevels,
                           orbm_frame** out) {
    if (!c || !out || n < 0 || (n && (!d_keys || !d_desc)) || !scale_factors || nlevels < 1 || width < 1 || height < 1)
        return orbx_set_error(ORBX_E_BAD_ARG, "orbm_frame_from_device: bad argument");
    *out = nullptr;
    M_TRY(cudaSetDevice(c->device));
    FrameGuard f{c, frame_new(c)};
    int rc;
    if ((rc = f->d_keys.ensure((size_t)std::max(n, 1) * sizeof(orbx_keypoint))) || (rc = f->d_desc.ensure((size_t)std::max(n, 1) * 32)) ||
        (rc = f->d_cell.ensure(64)))
        return rc;
    f->h_keys.resize(n);
    f->h_desc.resize((size_t)n * 32);
    f->h_scale.assign(scale_factors, scale_factors + nlevels);
    float corners[8] = {0.f, 0.f, (float)width, 0.f, 0.f, (float)height, (float)width, (float)height};
    const bool distorted = cam && cam->k1 != 0.0f;   // src/Frame.cc:740: mDistCoef.at<float>(0) == 0.0 -> mvKeysUn = mvKeys
    if (distorted) {
        const float cam9[9] = {cam->fx, cam->fy, cam->cx, cam->cy, cam->k1, cam->k2, cam->p1, cam->p2, cam->k3};
        M_TRY(orbm_launch_undistort(d_keys, n, cam9, (float)width, (float)height, (orbx_keypoint*)f->d_keys.p,
                                    (float*)f->d_cell.p, c->st));
        M_TRY(cudaMemcpyAsync(corners, f->d_cell.p, sizeof(corners), cudaMemcpyDeviceToHost, c->st));
        M_TRY(cudaStreamSynchronize(c->st));   // the bounds (and so the grid pitch) come from the undistorted corners
    } else if (n) {
        M_TRY(cudaMemcpyAsync(f->d_keys.p, d_keys, (size_t)n * sizeof(orbx_keypoint), cudaMemcpyDeviceToDevice, c->st));
    }
    if (n) {
        M_TRY(cudaMemcpyAsync(f->d_desc.p, d_desc, (size_t)n * 32, cudaMemcpyDeviceToDevice, c->st));
        M_TRY(cudaMemcpyAsync(f->h_keys.data(), f->d_keys.p, (size_t)n * sizeof(orbx_keypoint), cudaMemcpyDeviceToHost, c->st));
        M_TRY(cudaMemcpyAsync(f->h_desc.data(), f->d_desc.p, (size_t)n * 32, cudaMemcpyDeviceToHost, c->st));
    }
    // Frame::ComputeImageBounds (src/Frame.cc:773-800) from the undistorted corners, then the grid pitch (:119-120)
    orbm_frame_view& v = f->view;
    v.keys = f->h_keys.data();
    v.desc = f->h_desc.data();
    v.n = n;
    v.min_x = std::min(corners[0], corners[4]);
    v.max_x = std::max(corners[2], corners[6]);
    v.min_y = std::min(corners[1], corners[3]);
    v.max_y = std::max(corners[5], corners[7]);
    v.grid_w_inv = (float)ORBM_GRID_COLS / (v.max_x - v.min_x);
    v.grid_h_inv = (float)ORBM_GRID_ROWS / (v.max_y - v.min_y);
    v.scale_factors = f->h_scale.data();
    v.nlevels = nlevels;
    v.u_right = nullptr;
    if ((rc = frame_finish(c, f.get(), n))) return rc;
    *out = f.release();
    return ORBX_OK;
}

const orbm_frame_view* orbm_frame_get_view(const orbm_frame* f) { return f ? &f->view : nullptr; }

int orbm_frame_set_u_right(orbm_frame* f, const float* u_right) {
    if (!f) return orbx_set_error(ORBX_E_BAD_ARG, "null argument");
    if (u_right) {
        f->h_uright.assign(u_right, u_right + f->view.n);
        f->view.u_right = f->h_uright.data();
    } else {
        f->view.u_right = nullptr;
    }
    return ORBX_OK;
}

int orbm_frame_destroy(orbm_frame* f) {
    if (!f) return ORBX_OK;
    cudaSetDevice(f->ctx->device);
    ctx_sync(f->ctx);
    f->ctx->frames.erase(&f->view);
    frame_retire(f->ctx, f);
    return ORBX_OK;
}

}  // extern "C"

namespace {

int bank_row(const orbm_bank* bank, const int32_t* rows, int i, int* row) {
    *row = rows ? rows[i] : i;
    if (bank && (*row < 0 || *row >= bank->rows)) {
        ctx_unstage(bank->ctx);
        return orbx_set_error(ORBX_E_BAD_ARG, "query %d: descriptor row %d outside the bank's %d rows", i, *row, bank->rows);
    }
    return ORBX_OK;
}

// mp_desc (host rows, one per map point) or bank + rows (device-resident rows named per map point; rows == null: row i)
int search_by_projection_points_impl(orbm_context* c, const orbm_frame_view* F, const orbm_map_point* mps,
                                     const uint8_t* mp_desc, const orbm_bank* bank, const int32_t* rows, int m, float th,
                                     int far_points, float th_far, float nn_ratio, int32_t* kp_obs, int32_t* kp_match) {
    int rc = check_view(F);
    if (rc) return rc;
    if (!c || m < 0 || (m && (!mps || (!mp_desc && !bank))) || !kp_obs || !kp_match) return orbx_set_error(ORBX_E_BAD_ARG, "null argument");
    if (bank && bank->ctx != c) return orbx_set_error(ORBX_E_BAD_ARG, "the bank belongs to another context");
    ORBM_PHASE_BEGIN("search_by_projection_points");
    for (int i = 0; i < F->n; ++i) kp_match[i] = -1;
    const bool bFactor = th != 1.0;
    // pass 1: window queries (state independent), written straight into the pinned arena; candidates + distances come
    // from the device
    QueryBuf qb;
    if ((rc = stage_queries(c, m, &qb, true))) return rc;
    orbm_area_query* Q = qb.q;
    std::vector<float>& rad = c->h_rad;
    rad.resize(m);
    for (int i = 0; i < m; ++i) {
        const orbm_map_point& p = mps[i];
        int row;
        if ((rc = bank_row(bank, rows, i, &row))) return rc;
        const bool live = p.in_view && !(far_points && p.depth > th_far) && !p.bad;
        if (live && (p.level < 0 || p.level >= F->nlevels)) {
            ctx_unstage(c);
            return orbx_set_error(ORBX_E_BAD_ARG, "map point %d: level out of range", i);
        }
        float r = ((double)p.view_cos > 0.998) ? 2.5f : 4.0f;  // RadiusByViewingCos, :220-226
        if (bFactor) r *= th;
        rad[i] = live ? r * F->scale_factors[p.level] : 0.f;
        Q[i] = live ? orbm_area_query{p.proj_x, p.proj_y, rad[i], p.level - 1, p.level, 1, row}
                    : orbm_area_query{0.f, 0.f, 0.f, 0, 0, 0, row};
    }
    ORBM_PHASE("queries");
    if ((rc = device_candidates_onepass(c, F, qb, m, bank ? nullptr : mp_desc, bank ? 0 : m, bank ? bank->d : nullptr))) return rc;
    ORBM_PHASE("device_candidates");
    const uint2* row = c->h_rowinfo.data();   // {first position, count} of each map point's candidates
    const std::vector<int32_t>& cand = c->h_cand;
    // pass 2: the reference's accept loop, original order (:54-147)
    int nmatches = 0;
    // Rows with candidates (filtered-out points and empty windows have none) are listed without a branch per
    // row, and the scan below uses selects instead of early-outs: whether a keypoint is taken, has a right
    // coordinate or beats the running best is a coin flip to the branch predictor (same-box A/B of the accept
    // loop for 10 k map points: 35 -> 15 us at th = 1, 125 -> 92 us at th = 3; a select-based verdict on top
    // of it measured slower, 18 / 104 us, and was dropped).
    std::vector<int32_t>& live_rows = c->h_items;
    live_rows.resize((size_t)m + 1);
    int n_live = 0;
    for (int i = 0; i < m; ++i) {
        live_rows[n_live] = i;
        n_live += row[i].y != 0;
    }
    const float* ur = F->u_right;
    const uint16_t* hd = c->hdist.data();
    for (int t = 0; t < n_live; ++t) {
        const int i = live_rows[t];
        const orbm_map_point& p = mps[i];
        const float radi = rad[i];
        int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
        for (int k = (int)row[i].x, ke = (int)(row[i].x + row[i].y); k < ke; ++k) {
            const int idx = cand[k];
            // :90-101: a keypoint whose map point has observations, or whose right coordinate is further
            // than the radius from the projection, is skipped = takes a distance nothing accepts
            bool skip = kp_obs[idx] > 0;
            if (ur) {
                const float u = ur[idx];
                skip |= (u > 0) & (fabsf(p.proj_xr - u) > radi);
            }
            const int dist = skip ? 0x10000 : (int)hd[k];
            const int lvl = F->keys[idx].octave;
            const bool lt1 = dist < bestDist, lt2 = dist < bestDist2;   // bestDist <= bestDist2 throughout
            bestLevel2 = lt1 ? bestLevel : (lt2 ? lvl : bestLevel2);
            bestDist2 = lt1 ? bestDist : (lt2 ? dist : bestDist2);
            bestLevel = lt1 ? lvl : bestLevel;
            bestIdx = lt1 ? idx : bestIdx;
            bestDist = lt1 ? dist : bestDist;
        }
        if (bestDist <= ORBM_TH_HIGH) {
            if (bestLevel == bestLevel2 && bestDist > nn_ratio * bestDist2) continue;
            if (bestLevel != bestLevel2 || bestDist <= nn_ratio * bestDist2) {
                kp_match[bestIdx] = i;
                kp_obs[bestIdx] = p.observations;
                nmatches++;
            }
        }
    }
    ORBM_PHASE("replay");
    return nmatches;
}

int search_by_projection_frames_impl(orbm_context* c, const orbm_frame_view* cur, const orbm_last_obs* last,
                                     const uint8_t* last_desc, const orbm_bank* bank, const int32_t* rows, int n_last,
                                     float th, int forward, int backward, float bf, int check_orientation,
                                     int32_t* kp_obs, int32_t* kp_match) {
    int rc = check_view(cur);
    if (rc) return rc;
    if (!c || n_last < 0 || (n_last && (!last || (!last_desc && !bank))) || !kp_obs || !kp_match) return orbx_set_error(ORBX_E_BAD_ARG, "null argument");
    if (bank && bank->ctx != c) return orbx_set_error(ORBX_E_BAD_ARG, "the bank belongs to another context");
    for (int i = 0; i < cur->n; ++i) kp_match[i] = -1;
    std::vector<int> rotHist[ORBM_HISTO_LENGTH];
    const float factor = 1.0f / ORBM_HISTO_LENGTH;
    QueryBuf qb;
    if ((rc = stage_queries(c, n_last, &qb, true))) return rc;
    orbm_area_query* Q = qb.q;
    std::vector<float>& rad = c->h_rad;
    rad.resize(n_last);
    for (int i = 0; i < n_last; ++i) {
        const orbm_last_obs& o = last[i];
        int row;
        if ((rc = bank_row(bank, rows, i, &row))) return rc;
        orbm_area_query& q = Q[i];
        q = orbm_area_query{0.f, 0.f, 0.f, 0, 0, 0, row};
        rad[i] = 0.f;
        if (!o.valid) continue;
        if (o.inv_z < 0) continue;
        if (o.u < cur->min_x || o.u > cur->max_x) continue;
        if (o.v < cur->min_y || o.v > cur->max_y) continue;
        if (o.octave < 0 || o.octave >= cur->nlevels) {
            ctx_unstage(c);
            return orbx_set_error(ORBX_E_BAD_ARG, "observation %d: octave out of range", i);
        }
        rad[i] = th * cur->scale_factors[o.octave];
        if (forward) q = orbm_area_query{o.u, o.v, rad[i], o.octave, -1, 1, row};
        else if (backward) q = orbm_area_query{o.u, o.v, rad[i], 0, o.octave, 1, row};
        else q = orbm_area_query{o.u, o.v, rad[i], o.octave - 1, o.octave + 1, 1, row};
    }
    if ((rc = device_candidates_onepass(c, cur, qb, n_last, bank ? nullptr : last_desc, bank ? 0 : n_last,
                                        bank ? bank->d : nullptr)))
        return rc;
    const uint2* row = c->h_rowinfo.data();   // {first position, count} of each observation's candidates
    const std::vector<int32_t>& cand = c->h_cand;
    int nmatches = 0;
    for (int i = 0; i < n_last; ++i) {
        if (row[i].y == 0) continue;
        const orbm_last_obs& o = last[i];
        int bestDist = 256, bestIdx2 = -1;
        // selects instead of early-outs, as in SearchByProjection(Frame, map points) above (:2040-2056)
        const float ur = o.u - bf * o.inv_z, radi = rad[i];
        for (int k = (int)row[i].x, ke = (int)(row[i].x + row[i].y); k < ke; ++k) {
            const int i2 = cand[k];
            bool skip = kp_obs[i2] > 0;
            if (cur->u_right) {
                const float u2 = cur->u_right[i2];
                skip |= (u2 > 0) & (fabsf(ur - u2) > radi);
            }
            const int dist = skip ? 0x10000 : (int)c->hdist[k];
            const bool lt = dist < bestDist;
            bestIdx2 = lt ? i2 : bestIdx2;
            bestDist = lt ? dist : bestDist;
        }
        if (bestDist <= ORBM_TH_HIGH) {
            kp_match[bestIdx2] = i;
            kp_obs[bestIdx2] = o.observations;
            nmatches++;
            if (check_orientation) {
                float rot = o.angle - cur->keys[bestIdx2].angle;
                if (rot < 0.0) rot += 360.0f;
                int bin = (int)roundf(rot * factor);
                if (bin == ORBM_HISTO_LENGTH) bin = 0;
                if (bin < 0 || bin >= ORBM_HISTO_LENGTH) return orbx_set_error(ORBX_E_BAD_ARG, "angle outside [0,360)");
                rotHist[bin].push_back(bestIdx2);
            }
        }
    }
    if (check_orientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, ORBM_HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < ORBM_HISTO_LENGTH; i++) {
            if (i != ind1 && i != ind2 && i != ind3) {
                for (size_t j = 0; j < rotHist[i].size(); j++) {
                    kp_match[rotHist[i][j]] = -1;
                    kp_obs[rotHist[i][j]] = -1;
                    nmatches--;
                }
            }
        }
    }
    return nmatches;
}

}  // namespace

extern "C" {

int orbm_search_by_projection_points(orbm_context* c, const orbm_frame_view* F, const orbm_map_point* mps,
                                     const uint8_t* mp_desc, int m, float th, int far_points, float th_far,
                                     float nn_ratio, int32_t* kp_obs, int32_t* kp_match) {
    if (m && !mp_desc) return orbx_set_error(ORBX_E_BAD_ARG, "null argument");
    return search_by_projection_points_impl(c, F, mps, mp_desc, nullptr, nullptr, m, th, far_points, th_far, nn_ratio,
                                            kp_obs, kp_match);
}

int orbm_search_by_projection_points_bank(orbm_context* c, const orbm_frame_view* F, const orbm_map_point* mps,
                                          const orbm_bank* bank, const int32_t* rows, int m, float th, int far_points,
                                          float th_far, float nn_ratio, int32_t* kp_obs, int32_t* kp_match) {
    if (!bank) return orbx_set_error(ORBX_E_BAD_ARG, "null argument");
    return search_by_projection_points_impl(c, F, mps, nullptr, bank, rows, m, th, far_points, th_far, nn_ratio, kp_obs,
                                            kp_match);
}

int orbm_search_by_projection_frames(orbm_context* c, const orbm_frame_view* cur, const orbm_last_obs* last,
                                     const uint8_t* last_desc, int n_last, float th, int forward, int backward,
                                     float bf, int check_orientation, int32_t* kp_obs, int32_t* kp_match) {
    if (n_last && !last_desc) return orbx_set_error(ORBX_E_BAD_ARG, "null argument");
    return search_by_projection_frames_impl(c, cur, last, last_desc, nullptr, nullptr, n_last, th, forward, backward, bf,
                                            check_orientation, kp_obs, kp_match);
}

int orbm_search_by_projection_frames_bank(orbm_context* c, const orbm_frame_view* cur, const orbm_last_obs* last,
                                          const orbm_bank* bank, const int32_t* rows, int n_last, float th, int forward,
                                          int backward, float bf, int check_orientation, int32_t* kp_obs,
                                          int32_t* kp_match) {
    if (!bank) return orbx_set_error(ORBX_E_BAD_ARG, "null argument");
    return search_by_projection_frames_impl(c, cur, last, nullptr, bank, rows, n_last, th, forward, backward, bf,
                                            check_orientation, kp_obs, kp_match);
}

// ---- descriptor banks ---------------------------------------------------------------------------------
int orbm_bank_create(orbm_context* c, int capacity_rows, orbm_bank** out) {
    if (!c || !out || capacity_rows < 1) return orbx_set_error(ORBX_E_BAD_ARG, "bad argument");
    *out = nullptr;
    M_TRY(cudaSetDevice(c->device));
    orbm_bank* b = new orbm_bank();
    b->ctx = c;
    int rc = b->buf.ensure((size_t)capacity_rows * 32);
    if (rc) {
        delete b;
        return rc;
    }
    b->d = (const uint8_t*)b->buf.p;
    b->rows = capacity_rows;
    c->banks.push_back(b);
    *out = b;
    return ORBX_OK;
}

int orbm_bank_write(orbm_bank* b, int first_row, int n, const uint8_t* desc) {
    if (!b || !b->buf.p) return orbx_set_error(ORBX_E_BAD_ARG, "not a writable bank");
    if (first_row < 0 || n < 0 || first_row > b->rows - n || (n && !desc)) return orbx_set_error(ORBX_E_BAD_ARG, "rows outside the bank");
    if (n == 0) return ORBX_OK;
    M_TRY(cudaSetDevice(b->ctx->device));
    // on the context's stream, so that it is ordered with the searches; returns when the rows are in place
    M_TRY(cudaMemcpyAsync((uint8_t*)b->buf.p + (size_t)first_row * 32, desc, (size_t)n * 32, cudaMemcpyHostToDevice, b->ctx->st));
    M_TRY(cudaStreamSynchronize(b->ctx->st));
    return ORBX_OK;
}

int orbm_bank_rows(const orbm_bank* b) { return b ? b->rows : orbx_set_error(ORBX_E_BAD_ARG, "null argument"); }

int orbm_bank_destroy(orbm_bank* b) {
    if (!b) return ORBX_OK;
    if (!b->buf.p) return orbx_set_error(ORBX_E_BAD_ARG, "a frame's bank goes with its frame (orbm_frame_destroy)");
    cudaSetDevice(b->ctx->device);
    cudaStreamSynchronize(b->ctx->st);
    std::vector<orbm_bank*>& v = b->ctx->banks;
    v.erase(std::remove(v.begin(), v.end(), b), v.end());
    b->buf.release();
    delete b;
    return ORBX_OK;
}

const orbm_bank* orbm_frame_bank(const orbm_frame* f) { return f ? &f->bank : nullptr; }

}  // extern "C"

namespace {

// the right camera's half as a frame view over the same grid geometry (PosInGrid uses the frame's
// bounds for both cameras, src/Frame.cc:377-408)
orbm_frame_view right_view(const orbm_frame_view* F, const orbm_frame_right* R) {
    orbm_frame_view v = *F;
    v.keys = R->keys;
    v.desc = R->desc;
    v.n = R->n;
    v.u_right = nullptr;
    return v;
}

int check_right(const orbm_frame_right* R) {
    if (!R || R->n < 0 || (R->n && (!R->keys || !R->desc))) return orbx_set_error(ORBX_E_BAD_ARG, "bad right-camera view");
    return ORBX_OK;
}

struct Csr {
    std::vector<int32_t> row_ptr, cand;
    std::vector<uint16_t> dist;
};

void take_csr(orbm_context* c, Csr& o) {
    o.row_ptr = c->h_rowptr;
    o.cand = c->h_cand;
    o.dist = c->hdist;
}

}  // namespace

extern "C" {

int orbm_search_by_projection_points_stereo(orbm_context* c, const orbm_frame_view* F, const orbm_frame_right* R,
                                            const orbm_map_point* mps, const orbm_map_point_right* mps_r,
                                            const uint8_t* mp_desc, int m, float th, int far_points, float th_far,
                                            float nn_ratio, int32_t* kp_obs, int32_t* kp_match) {
    int rc = check_view(F);
    if (rc || (rc = check_right(R))) return rc;
    if (!c || m < 0 || (m && (!mps || !mps_r || !mp_desc)) || !kp_obs || !kp_match)
        return orbx_set_error(ORBX_E_BAD_ARG, "null argument");
    const int Nleft = F->n;
    for (int i = 0; i < Nleft + R->n; ++i) kp_match[i] = -1;
    const bool bFactor = th != 1.0;
    // pass 1: the left and the right window of every map point (state independent)
    std::vector<orbm_area_query> QL(m), QR(m);
    for (int i = 0; i < m; ++i) {
        const orbm_map_point& p = mps[i];
        const orbm_map_point_right& pr = mps_r[i];
        QL[i] = QR[i] = orbm_area_query{0.f, 0.f, 0.f, 0, 0, 0, i};
        if (!p.in_view && !pr.in_view) continue;            // :57-58
        if (far_points && p.depth > th_far) continue;
        if (p.bad) continue;
        if (p.in_view) {
            if (p.level < 0 || p.level >= F->nlevels) return orbx_set_error(ORBX_E_BAD_ARG, "map point %d: level out of range", i);
            float r = ((double)p.view_cos > 0.998) ? 2.5f : 4.0f;
            if (bFactor) r *= th;
            QL[i] = orbm_area_query{p.proj_x, p.proj_y, r * F->scale_factors[p.level], p.level - 1, p.level, 1, i};
        }
        if (pr.in_view && pr.level != -1) {                  // :149-151
            if (pr.level < 0 || pr.level >= F->nlevels) return orbx_set_error(ORBX_E_BAD_ARG, "map point %d: right level out of range", i);
            const float r = ((double)pr.view_cos > 0.998) ? 2.5f : 4.0f;   // not scaled by th (:152)
            QR[i] = orbm_area_query{pr.proj_x, pr.proj_y, r * F->scale_factors[pr.level], pr.level - 1, pr.level, 1, i};
        }
    }
    Csr L, Rr;
    if ((rc = device_candidates(c, F, QL, mp_desc, m))) return rc;
    take_csr(c, L);
    const orbm_frame_view FR = right_view(F, R);
    if ((rc = device_candidates(c, &FR, QR, mp_desc, m))) return rc;
    take_csr(c, Rr);
    c->last_pairs = (long long)L.cand.size() + (long long)Rr.cand.size();
    // pass 2: the reference's accept loop in the original order
    int nmatches = 0;
    for (int i = 0; i < m; ++i) {
        const orbm_map_point& p = mps[i];
        if (L.row_ptr[i + 1] != L.row_ptr[i]) {              // mbTrackInView && !vIndices.empty()
            int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
            for (int k = L.row_ptr[i]; k < L.row_ptr[i + 1]; ++k) {
                const int idx = L.cand[k];
                if (kp_obs[idx] > 0) continue;
                const int dist = L.dist[k];
                if (dist < bestDist) {
                    bestDist2 = bestDist;
                    bestDist = dist;
                    bestLevel2 = bestLevel;
                    bestLevel = F->keys[idx].octave;
                    bestIdx = idx;
                } else if (dist < bestDist2) {
                    bestLevel2 = F->keys[idx].octave;
                    bestDist2 = dist;
                }
            }
            if (bestDist <= ORBM_TH_HIGH) {
                if (bestLevel == bestLevel2 && bestDist > nn_ratio * bestDist2) continue;   // skips the right branch too (:125-126)
                if (bestLevel != bestLevel2 || bestDist <= nn_ratio * bestDist2) {
                    kp_match[bestIdx] = i;
                    kp_obs[bestIdx] = p.observations;
                    const int l2r = R->left_to_right ? R->left_to_right[bestIdx] : -1;
                    if (l2r != -1) {                             // :131-135
                        kp_match[l2r + Nleft] = i;
                        kp_obs[l2r + Nleft] = p.observations;
                        nmatches++;
                    }
                    nmatches++;
                }
            }
        }
        if (Rr.row_ptr[i + 1] == Rr.row_ptr[i]) continue;     // not in the right view, or vIndices.empty()
        int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
        for (int k = Rr.row_ptr[i]; k < Rr.row_ptr[i + 1]; ++k) {
            const int idx = Rr.cand[k];
            if (kp_obs[idx + Nleft] > 0) continue;
            const int dist = Rr.dist[k];
            if (dist < bestDist) {
                bestDist2 = bestDist;
                bestDist = dist;
                bestLevel2 = bestLevel;
                bestLevel = R->keys[idx].octave;
                bestIdx = idx;
            } else if (dist < bestDist2) {
                bestLevel2 = R->keys[idx].octave;
                bestDist2 = dist;
            }
        }
        if (bestDist <= ORBM_TH_HIGH) {
            if (bestLevel == bestLevel2 && bestDist > nn_ratio * bestDist2) continue;
            const int r2l = R->right_to_left ? R->right_to_left[bestIdx] : -1;
            if (r2l != -1) {                                     // :201-205
                kp_match[r2l] = i;
                kp_obs[r2l] = p.observations;
                nmatches++;
            }
            kp_match[bestIdx + Nleft] = i;
            kp_obs[bestIdx + Nleft] = p.observations;
            nmatches++;
        }
    }
    return nmatches;
}

int orbm_search_by_projection_frames_stereo(orbm_context* c, const orbm_frame_view* cur, const orbm_frame_right* R,
                                            const orbm_last_obs* last, const float* uv_right, const uint8_t* last_desc,
                                            int n_last, float th, int forward, int backward, int check_orientation,
                                            int32_t* kp_obs, int32_t* kp_match) {
    int rc = check_view(cur);
    if (rc || (rc = check_right(R))) return rc;
    if (!c || n_last < 0 || (n_last && (!last || !last_desc || !uv_right)) || !kp_obs || !kp_match)
        return orbx_set_error(ORBX_E_BAD_ARG, "null argument");
    const int Nleft = cur->n;
    for (int i = 0; i < Nleft + R->n; ++i) kp_match[i] = -1;
    std::vector<int> rotHist[ORBM_HISTO_LENGTH];
    const float factor = 1.0f / ORBM_HISTO_LENGTH;
    std::vector<orbm_area_query> QL(n_last), QR(n_last);
    std::vector<char> live(n_last, 0);
    for (int i = 0; i < n_last; ++i) {
        const orbm_last_obs& o = last[i];
        QL[i] = QR[i] = orbm_area_query{0.f, 0.f, 0.f, 0, 0, 0, i};
        if (!o.valid) continue;
        if (o.inv_z < 0) continue;
        if (o.u < cur->min_x || o.u > cur->max_x) continue;
        if (o.v < cur->min_y || o.v > cur->max_y) continue;
        if (o.octave < 0 || o.octave >= cur->nlevels) return orbx_set_error(ORBX_E_BAD_ARG, "observation %d: octave out of range", i);
        const float radius = th * cur->scale_factors[o.octave];
        const int mn = forward ? o.octave : backward ? 0 : o.octave - 1;
        const int mx = forward ? -1 : backward ? o.octave : o.octave + 1;
        QL[i] = orbm_area_query{o.u, o.v, radius, mn, mx, 1, i};
        // the right window is searched whenever the left one is reached, wherever it lies (:2089-2107)
        QR[i] = orbm_area_query{uv_right[2 * i], uv_right[2 * i + 1], radius, mn, mx, 1, i};
        live[i] = 1;
    }
    Csr L, Rr;
    if ((rc = device_candidates(c, cur, QL, last_desc, n_last))) return rc;
    take_csr(c, L);
    const orbm_frame_view FR = right_view(cur, R);
    if ((rc = device_candidates(c, &FR, QR, last_desc, n_last))) return rc;
    take_csr(c, Rr);
    c->last_pairs = (long long)L.cand.size() + (long long)Rr.cand.size();
    int nmatches = 0;
    auto bin_of = [&](float rot, int* bin) {
        if (rot < 0.0) rot += 360.0f;
        int b = (int)roundf(rot * factor);
        if (b == ORBM_HISTO_LENGTH) b = 0;
        *bin = b;
        return b >= 0 && b < ORBM_HISTO_LENGTH;
    };
    for (int i = 0; i < n_last; ++i) {
        if (!live[i]) continue;
        const orbm_last_obs& o = last[i];
        // left camera: `if(vIndices2.empty()) continue;` skips the right camera as well (:2028-2029)
        if (L.row_ptr[i + 1] == L.row_ptr[i]) continue;
        {
            int bestDist = 256, bestIdx2 = -1;
            for (int k = L.row_ptr[i]; k < L.row_ptr[i + 1]; ++k) {
                const int i2 = L.cand[k];
                if (kp_obs[i2] > 0) continue;
                const int dist = L.dist[k];
                if (dist < bestDist) { bestDist = dist; bestIdx2 = i2; }
            }
            if (bestDist <= ORBM_TH_HIGH) {
                kp_match[bestIdx2] = i;
                kp_obs[bestIdx2] = o.observations;
                nmatches++;
                if (check_orientation) {
                    int bin;
                    if (!bin_of(o.angle - cur->keys[bestIdx2].angle, &bin)) return orbx_set_error(ORBX_E_BAD_ARG, "angle outside [0,360)");
                    rotHist[bin].push_back(bestIdx2);
                }
            }
        }
        {
            int bestDist = 256, bestIdx2 = -1;
            for (int k = Rr.row_ptr[i]; k < Rr.row_ptr[i + 1]; ++k) {
                const int i2 = Rr.cand[k];
                if (kp_obs[i2 + Nleft] > 0) continue;
                const int dist = Rr.dist[k];
                if (dist < bestDist) { bestDist = dist; bestIdx2 = i2; }
            }
            if (bestDist <= ORBM_TH_HIGH) {
                kp_match[bestIdx2 + Nleft] = i;
                kp_obs[bestIdx2 + Nleft] = o.observations;
                nmatches++;
                if (check_orientation) {
                    int bin;
                    if (!bin_of(o.angle - R->keys[bestIdx2].angle, &bin)) return orbx_set_error(ORBX_E_BAD_ARG, "angle outside [0,360)");
                    rotHist[bin].push_back(bestIdx2 + Nleft);
                }
            }
        }
    }
    if (check_orientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, ORBM_HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < ORBM_HISTO_LENGTH; i++) {
            if (i != ind1 && i != ind2 && i != ind3) {
                for (size_t j = 0; j < rotHist[i].size(); j++) {
                    kp_match[rotHist[i][j]] = -1;
                    kp_obs[rotHist[i][j]] = -1;
                    nmatches--;
                }
            }
        }
    }
    return nmatches;
}

// SearchByBoW(KeyFrame*, Frame&) with every (row, candidate) pair expanded and scanned on the host: the
// fisheye-rig form (f_nleft != -1, separate left / right bests); orbm_search_by_bow (orbm_n3.inc) answers
// the usual form from k_bow_topk.
static int search_by_bow_full(orbm_context* c, const orbm_bow_view* KF, const uint8_t* kf_mp_ok, const orbm_bow_view* F,
                              int f_nleft, float nn_ratio, int check_orientation, int32_t* matches) {
    if (!c || !KF || !F || !matches || KF->n < 0 || F->n < 0 || KF->nnodes < 0 || F->nnodes < 0 ||
        (KF->n && (!KF->keys || !KF->desc || !kf_mp_ok)) || (F->n && (!F->keys || !F->desc)) ||
        (KF->nnodes && (!KF->node_ids || !KF->node_ptr || !KF->node_items)) ||
        (F->nnodes && (!F->node_ids || !F->node_ptr || !F->node_items)))
        return orbx_set_error(ORBX_E_BAD_ARG, "bad argument");
    ORBM_PHASE_BEGIN("search_by_bow");
    for (int i = 0; i < F->n; ++i) matches[i] = -1;
    // pass 1: walk the two sorted feature vectors like the reference (:291-449; lower_bound on a
    // sorted key array) and list every (keyframe feature, frame feature) pair that shares a node
    struct NodePair { int a, b; };
    std::vector<NodePair> shared;
    {
        int a = 0, b = 0;
        while (a < KF->nnodes && b < F->nnodes) {
            if (KF->node_ids[a] == F->node_ids[b]) {
                shared.push_back(NodePair{a, b});
                ++a;
                ++b;
            } else if (KF->node_ids[a] < F->node_ids[b]) {
                a = (int)(std::lower_bound(KF->node_ids, KF->node_ids + KF->nnodes, F->node_ids[b]) - KF->node_ids);
            } else {
                b = (int)(std::lower_bound(F->node_ids, F->node_ids + F->nnodes, KF->node_ids[a]) - F->node_ids);
            }
        }
    }
    std::vector<int32_t> row_ptr, cand, row_q;   // one row per (node, valid keyframe feature)
    row_ptr.push_back(0);
    for (const NodePair& np : shared) {
        for (int ia = KF->node_ptr[np.a]; ia < KF->node_ptr[np.a + 1]; ++ia) {
            const uint32_t realIdxKF = KF->node_items[ia];
            if (realIdxKF >= (uint32_t)KF->n) return orbx_set_error(ORBX_E_BAD_ARG, "keyframe feature index out of range");
            if (!kf_mp_ok[realIdxKF]) continue;
            for (int ib = F->node_ptr[np.b]; ib < F->node_ptr[np.b + 1]; ++ib) {
                if (F->node_items[ib] >= (uint32_t)F->n) return orbx_set_error(ORBX_E_BAD_ARG, "frame feature index out of range");
                cand.push_back((int32_t)F->node_items[ib]);
            }
            row_q.push_back((int32_t)realIdxKF);
            row_ptr.push_back((int32_t)cand.size());
        }
    }
    // the CSR kernel takes one query descriptor per row: gather the rows' keyframe descriptors
    const int nrows = (int)row_q.size();
    ORBM_PHASE("csr");
    std::vector<uint8_t> qdesc((size_t)nrows * 32);
    for (int r = 0; r < nrows; ++r) memcpy(&qdesc[(size_t)r * 32], KF->desc + (size_t)row_q[r] * 32, 32);
    c->hdist.assign(cand.size(), 0);
    ORBM_PHASE("gather");
    int rc;
    if (!cand.empty() && (rc = orbm_hamming_csr(c, qdesc.data(), nrows, F->desc, F->n, row_ptr.data(), cand.data(),
                                                (int)cand.size(), c->hdist.data())))
        return rc;
    c->last_pairs = (long long)cand.size();
    ORBM_PHASE("device");
    // pass 2: the reference's accept loop over the precomputed distances
    std::vector<int> rotHist[ORBM_HISTO_LENGTH];
    const float factor = 1.0f / ORBM_HISTO_LENGTH;
    int nmatches = 0;
    auto push_rot = [&](int idxKF, int idxF) {
        float rot = KF->keys[idxKF].angle - F->keys[idxF].angle;
        if (rot < 0.0) rot += 360.0f;
        int bin = (int)roundf(rot * factor);
        if (bin == ORBM_HISTO_LENGTH) bin = 0;
        if (bin < 0 || bin >= ORBM_HISTO_LENGTH) return false;
        rotHist[bin].push_back(idxF);
        return true;
    };
    for (int r = 0; r < nrows; ++r) {
        const int realIdxKF = row_q[r];
        int bestDist1 = 256, bestIdxF = -1, bestDist2 = 256;
        int bestDist1R = 256, bestIdxFR = -1, bestDist2R = 256;
        for (int k = row_ptr[r]; k < row_ptr[r + 1]; ++k) {
            const int realIdxF = cand[k];
            if (matches[realIdxF] >= 0) continue;
            const int dist = c->hdist[k];
            if (f_nleft == -1) {
                if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdxF = realIdxF; }
                else if (dist < bestDist2) bestDist2 = dist;
            } else {
                if (realIdxF < f_nleft && dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdxF = realIdxF; }
                else if (realIdxF < f_nleft && dist < bestDist2) bestDist2 = dist;
                if (realIdxF >= f_nleft && dist < bestDist1R) { bestDist2R = bestDist1R; bestDist1R = dist; bestIdxFR = realIdxF; }
                else if (realIdxF >= f_nleft && dist < bestDist2R) bestDist2R = dist;
            }
        }
        if (bestDist1 <= ORBM_TH_LOW) {
            if ((float)bestDist1 < nn_ratio * (float)bestDist2) {
                matches[bestIdxF] = realIdxKF;
                if (check_orientation && !push_rot(realIdxKF, bestIdxF)) return orbx_set_error(ORBX_E_BAD_ARG, "angle outside [0,360)");
                nmatches++;
            }
            if (bestDist1R <= ORBM_TH_LOW) {   // the ratio test is disabled in the reference ("|| true", :396)
                matches[bestIdxFR] = realIdxKF;
                if (check_orientation && !push_rot(realIdxKF, bestIdxFR)) return orbx_set_error(ORBX_E_BAD_ARG, "angle outside [0,360)");
                nmatches++;
            }
        }
    }
    if (check_orientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, ORBM_HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < ORBM_HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (size_t j = 0; j < rotHist[i].size(); j++) {
                matches[rotHist[i][j]] = -1;
                nmatches--;
            }
        }
    }
    ORBM_PHASE("replay");
    return nmatches;
}

int orbm_search_for_initialization(orbm_context* c, const orbm_frame_view* F1, const orbm_frame_view* F2,
                                   float* prev_matched, int32_t* matches12, int window_size, float nn_ratio,
                                   int check_orientation) {
    int rc = check_view(F1);
    if (rc || (rc = check_view(F2))) return rc;
    if (!c || !prev_matched || !matches12) return orbx_set_error(ORBX_E_BAD_ARG, "null argument");
    int nmatches = 0;
    for (int i = 0; i < F1->n; ++i) matches12[i] = -1;
    std::vector<int> rotHist[ORBM_HISTO_LENGTH];
    const float factor = 1.0f / ORBM_HISTO_LENGTH;
    std::vector<int> vMatchedDistance(F2->n, INT_MAX), vnMatches21(F2->n, -1);
    std::vector<orbm_area_query> Q(F1->n);
    for (int i1 = 0; i1 < F1->n; ++i1) {
        const int level1 = F1->keys[i1].octave;
        Q[i1] = orbm_area_query{prev_matched[2 * i1], prev_matched[2 * i1 + 1], (float)window_size, level1, level1,
                                level1 > 0 ? 0 : 1, i1};
    }
    // 100-pixel windows hold tens to hundreds of candidates per row: the accept loop reads the device's
    // INIT_TOPK best keys per row and scans a row in full only when they cannot settle it
    const int K = INIT_TOPK;
    if ((rc = device_candidates(c, F2, Q, F1->desc, F1->n, K))) return rc;
    const std::vector<int32_t>& row_ptr = c->h_rowptr;
    for (int i1 = 0; i1 < F1->n; ++i1) {
        const int len = row_ptr[i1 + 1] - row_ptr[i1];
        if (len == 0) continue;
        int bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx2 = -1;
        // :741-762: candidates already matched at a distance <= this one are skipped; of the others the
        // first two in (distance, position) order are the scan's best and second best
        const uint32_t* keys = &c->h_topk[(size_t)i1 * K];
        const int32_t* kc = &c->h_topc[(size_t)i1 * K];
        int found = 0, seen = 0, lastd = 0;
        for (int k = 0; k < K && keys[k] != 0xffffffffu; ++k) {
            ++seen;
            const int dist = lastd = (int)(keys[k] >> 20);
            const int i2 = kc[k];
            if (vMatchedDistance[i2] <= dist) continue;
            if (!found) {
                bestDist = dist;
                bestIdx2 = i2;
                found = 1;
            } else {
                bestDist2 = dist;
                found = 2;
                break;
            }
        }
        if (found < 2 && seen == K && len > K) {   // keys exhausted on a longer row
            bool settled;
            if (found == 0) settled = lastd > ORBM_TH_LOW;              // nothing acceptable is left
            else if (bestDist > ORBM_TH_LOW) settled = true;
            else if ((settled = bestDist < (float)lastd * nn_ratio)) bestDist2 = lastd;   // second best >= lastd
            if (!settled) {
                if ((rc = window_full_lists(c))) return rc;
                bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx2 = -1;
                for (int k = row_ptr[i1]; k < row_ptr[i1 + 1]; ++k) {
                    const int i2 = c->h_cand[k];
                    const int dist = c->hdist[k];
                    if (vMatchedDistance[i2] <= dist) continue;
                    if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestIdx2 = i2; }
                    else if (dist < bestDist2) bestDist2 = dist;
                }
            }
        }
        if (bestDist <= ORBM_TH_LOW) {
            if (bestDist < (float)bestDist2 * nn_ratio) {
                if (vnMatches21[bestIdx2] >= 0) {
                    matches12[vnMatches21[bestIdx2]] = -1;
                    nmatches--;
                }
                matches12[i1] = bestIdx2;
                vnMatches21[bestIdx2] = i1;
                vMatchedDistance[bestIdx2] = bestDist;
                nmatches++;
                if (check_orientation) {
                    float rot = F1->keys[i1].angle - F2->keys[bestIdx2].angle;
                    if (rot < 0.0) rot += 360.0f;
                    int bin = (int)roundf(rot * factor);
                    if (bin == ORBM_HISTO_LENGTH) bin = 0;
                    if (bin < 0 || bin >= ORBM_HISTO_LENGTH) return orbx_set_error(ORBX_E_BAD_ARG, "angle outside [0,360)");
                    rotHist[bin].push_back(i1);
                }
            }
        }
    }
    if (check_orientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, ORBM_HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < ORBM_HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (size_t j = 0; j < rotHist[i].size(); j++) {
                const int idx1 = rotHist[i][j];
                if (matches12[idx1] >= 0) {
                    matches12[idx1] = -1;
                    nmatches--;
                }
            }
        }
    }
    for (int i1 = 0; i1 < F1->n; ++i1)
        if (matches12[i1] >= 0) {
            prev_matched[2 * i1] = F2->keys[matches12[i1]].x;
            prev_matched[2 * i1 + 1] = F2->keys[matches12[i1]].y;
        }
    return nmatches;
}

}  // extern "C"

#include "orbm_n3.inc"
