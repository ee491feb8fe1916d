// orbx_api.cu -- host runtime behind the C ABI of include/orbx.h.
//
// Mirrors the host side of ORB_SLAM3::ORBextractor (reference include/ORBextractor.h:43-109):
// the constructor's tables (src/ORBextractor.cc:408-468), per-shape planning of everything the
// reference recomputes per frame (level sizes :1156-1157, FAST cell grid :771-785, quadtree
// roots :541-556) and the device pipeline  H2D -> pyramid -> FAST -> quadtree -> assemble ->
// describe -> D2H  on one CUDA stream per pipeline slot.  No CPU implementation of any stage
// exists here: without a CUDA device orbx_create fails.
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "orbx_internal.h"

// kernel launchers (k_*.cu)
void orbx_launch_pyramid(const OrbxPlan& P, const OrbxSrc& src, uint8_t* pyr, int nframes,
                         cudaStream_t st, int* launches, cudaEvent_t* level_done = nullptr);
cudaError_t orbx_launch_input(const uint8_t* raw, long long raw_fstride, int raw_pitch, int w, int h, int channels,
                              int rgb_order, int gray_bits, const uint4* map, uint8_t* dst, long long dst_fstride,
                              int dst_pitch, int nframes, cudaStream_t st, int* launches);
cudaError_t orbx_configure_fast(const OrbxPlan& P);
cudaError_t orbx_launch_fast(const OrbxPlan& P, const OrbxSrc& src, const uint8_t* pyr,
                             uint32_t* cellcnt, uint32_t* cellkeys, int nframes, cudaStream_t st,
                             int* launches, int level = -1);
cudaError_t orbx_configure_octree(const OrbxPlan& P);
cudaError_t orbx_launch_octree(const OrbxPlan& P, const uint32_t* cellcnt, const uint32_t* cellkeys,
                               uint32_t* keysA, uint32_t* keysB, uint32_t* surv, uint32_t* survcnt,
                               const int* lap, uint32_t* dest, int32_t* header, uint32_t* frame_done,
                               int nframes, cudaStream_t st, int* launches, int level = -1);
cudaError_t orbx_launch_describe(const OrbxPlan& P, const OrbxSrc& src, const uint8_t* pyr,
                                 const int32_t* header, const uint32_t* dest,
                                 orbx_keypoint* kps, uint8_t* desc, int nframes, cudaStream_t st,
                                 int* launches, const OrbxSinkArgs& sink);

cudaError_t orbx_launch_sink_wait_flag(const uint32_t* flag, uint32_t epoch, unsigned long long timeout_ns, uint32_t* err,
                                       cudaStream_t st);
cudaError_t orbx_launch_sink_release(uint32_t* ack, uint32_t epoch, cudaStream_t st);
cudaError_t orbx_sink_preload();
cudaError_t orbx_run_sincos_check(float lo, float hi, unsigned long long* n_bad, float* first_bad);

namespace {

thread_local std::string g_err = "";

int vfail(int code, const char* fmt, va_list ap) {
    char buf[512];
    vsnprintf(buf, sizeof buf, fmt, ap);
    g_err = buf;
    return code;
}

int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vfail(code, fmt, ap);
    va_end(ap);
    return code;
}

#define CUDA_TRY(expr)                                                                        \
    do {                                                                                      \
        cudaError_t _e = (expr);                                                              \
        if (_e != cudaSuccess)                                                                \
            return fail(ORBX_E_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), \
                        __FILE__, __LINE__);                                                  \
    } while (0)

inline int cv_round(float v) { return (int)lrintf(v); }  // cvRound: round half to even
inline int cv_floor(double v) { int i = (int)v; return i - (i > v); }
inline int round_up(int v, int a) { return (v + a - 1) / a * a; }

}  // namespace

// used by orbm_api.cu so that both halves of the library share orbx_last_error_string()
int orbx_set_error(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vfail(code, fmt, ap);
    va_end(ap);
    return code;
}

namespace {

struct SlotCaps {  // bytes currently allocated per device buffer
    size_t img_in = 0, pyr = 0, cellcnt = 0, cellkeys = 0, keysA = 0, keysB = 0, surv = 0, survcnt = 0,
           dest = 0, header = 0, kps = 0, desc = 0, lap = 0, frame_done = 0;
};

struct Slot {
    SlotCaps caps;
    cudaStream_t st = nullptr;
    uint8_t* img_in = nullptr;   // device staging of the H2D copy (level 0)
    uint8_t* raw_in = nullptr;   // raw (colour / unrectified) frames when the input stage is on
    size_t raw_cap = 0;
    uint8_t* pyr = nullptr;
    uint32_t *cellcnt = nullptr, *cellkeys = nullptr, *keysA = nullptr, *keysB = nullptr;
    uint32_t *surv = nullptr, *survcnt = nullptr, *dest = nullptr;
    int32_t* header = nullptr;
    orbx_keypoint* kps = nullptr;
    uint8_t* desc = nullptr;
    int* lap_d = nullptr;
    uint32_t* frame_done = nullptr;   // per-frame CTA tickets of k_octree (self-resetting)
    int32_t* h_header = nullptr;  // pinned
    int* h_lap = nullptr;         // pinned
    int lap_valid = 0;            // frames whose lapping area on the device equals h_lap
    // in-flight bookkeeping
    bool in_flight = false;
    int nframes = 0, cap = 0, launches = 0;
    int *u_n = nullptr, *u_mono = nullptr, *u_status = nullptr;
    cudaEvent_t ev[6] = {};      // stage boundaries (ORBX_FLAG_STAGE_TIMING)
    cudaEvent_t ev_in[2] = {};   // around the input stage (N4)
    bool in_timed = false;
    OrbxSrc src{};               // level-0 source of the last run (for the debug taps)
    int run_w = 0, run_h = 0;
    // CUDA graph of the kernel sequence for small batches (single-frame latency path)
    cudaGraphExec_t gexec = nullptr;
    unsigned long long g_epoch = 0, warm_epoch = 0;
    int g_nframes = 0, g_launches = 0;
    OrbxSrc g_src{};
    // forked form of that graph: one branch per pyramid level (created on first capture)
    OrbxSinkArgs sink{};         // fused result gather (orbx_set_result_sink); hdr == nullptr: off
    uint32_t* sink_words = nullptr;   // device: [0] batches sent to the sink (epoch), [1] back-pressure timeout flag
    uint32_t sink_epoch = 0;          // host count of the same
    cudaStream_t aux[ORBX_MAX_LEVELS] = {};
    cudaEvent_t ev_fork = nullptr, ev_lvl[ORBX_MAX_LEVELS] = {}, ev_join[ORBX_MAX_LEVELS] = {};
};

}  // namespace

struct orbx_handle {
    orbx_config cfg;
    std::vector<float> sf, isf, s2, is2;
    std::vector<int> nper;
    OrbxPlan plan;       // for plan.w x plan.h (0 = none yet)
    OrbxPlan cap_plan;   // for max_width x max_height: sizes every buffer
    uint8_t* tables = nullptr;  // device resize tables of `plan`
    size_t tables_bytes = 0;
    int pitch0_cap = 0;
    unsigned long long epoch = 1;   // bumped whenever the plan or a device buffer changes
    // input stage (N4)
    int in_format = ORBX_IN_GRAY8, gray_bits = 15;
    uint4* d_map = nullptr;         // fixed-point rectification map, map_w x map_h entries (k_input.cu)
    int map_w = 0, map_h = 0;
    bool input_stage() const { return in_format != ORBX_IN_GRAY8 || d_map != nullptr; }
    int channels() const { return in_format == ORBX_IN_GRAY8 ? 1 : (in_format <= ORBX_IN_RGB8 ? 3 : 4); }
    std::vector<Slot> slots;
};

namespace {

// ---- ORBextractor::ORBextractor, src/ORBextractor.cc:408-444 ---------------------------------
void build_tables(orbx_handle* h) {
    const int nl = h->cfg.nlevels;
    const double scaleFactor = (double)h->cfg.scale_factor;  // `double scaleFactor` member
    h->sf.assign(nl, 1.f);
    h->s2.assign(nl, 1.f);
    h->isf.assign(nl, 1.f);
    h->is2.assign(nl, 1.f);
    for (int i = 1; i < nl; ++i) {
        h->sf[i] = (float)(h->sf[i - 1] * scaleFactor);
        h->s2[i] = h->sf[i] * h->sf[i];
    }
    for (int i = 0; i < nl; ++i) {
        h->isf[i] = 1.0f / h->sf[i];
        h->is2[i] = 1.0f / h->s2[i];
    }
    h->nper.assign(nl, 0);
    const float factor = (float)(1.0f / scaleFactor);
    float nDesired = (float)(h->cfg.nfeatures * (1 - factor) /
                             (1 - (float)pow((double)factor, (double)nl)));
    int sum = 0;
    for (int l = 0; l < nl - 1; ++l) {
        h->nper[l] = cv_round(nDesired);
        sum += h->nper[l];
        nDesired *= factor;
    }
    h->nper[nl - 1] = std::max(h->cfg.nfeatures - sum, 0);
}

// OpenCV's INTER_LINEAR tables for one axis (imgproc/resize.cpp): source offset and the two
// 11-bit coefficients per destination index.  Filled into host vectors, uploaded by the caller.
void linear_axis(int ssize, int dsize, std::vector<uint16_t>& ofs, std::vector<short>& coef) {
    ofs.resize(dsize);
    coef.resize((size_t)2 * dsize);
    const double inv_scale = (double)dsize / ssize;
    const double scale = 1. / inv_scale;
    for (int d = 0; d < dsize; ++d) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = cv_floor(f);
        f -= s;
        if (s < 0) { f = 0; s = 0; }
        if (s >= ssize - 1) { f = 0; s = ssize - 1; }
        ofs[d] = (uint16_t)s;
        coef[2 * d] = (short)cv_round((1.f - f) * 2048.f);
        coef[2 * d + 1] = (short)cv_round(f * 2048.f);
    }
}

// Everything that depends on the image shape.  Returns an orbx_status.
int build_plan(const orbx_handle* h, int w, int hgt, OrbxPlan* out, std::vector<uint8_t>* tab_host) {
    OrbxPlan& P = *out;
    memset(&P, 0, sizeof P);
    const int nl = h->cfg.nlevels;
    P.nlevels = nl;
    P.w = w;
    P.h = hgt;
    P.ini_th = std::min(std::max(h->cfg.ini_th_fast, 0), 255);
    P.min_th = std::min(std::max(h->cfg.min_th_fast, 0), 255);
    if (w > ORBX_MAX_DIM || hgt > ORBX_MAX_DIM) return fail(ORBX_E_UNSUPPORTED, "image larger than %d", ORBX_MAX_DIM);
    long long off = 0, key_off = 0;
    int cell_off = 0, surv_off = 0;
    size_t tab_off = 0;
    if (tab_host) tab_host->clear();
    auto put = [&](const void* p, size_t n) -> size_t {
        size_t at = tab_off;
        if (tab_host) {
            tab_host->resize(at + ((n + 15) & ~(size_t)15));
            memcpy(tab_host->data() + at, p, n);
        }
        tab_off += (n + 15) & ~(size_t)15;
        return at;
    };
    for (int l = 0; l < nl; ++l) {
        OrbxLevel& L = P.lvl[l];
        // ORBextractor.cc:1156-1157
        L.w = cv_round((float)w * h->isf[l]);
        L.h = cv_round((float)hgt * h->isf[l]);
        L.scale = h->sf[l];
        L.size = (float)(int)(31 * h->sf[l]);  // PATCH_SIZE * mvScaleFactor[level], :862
        if (l == 0) {
            L.pitch = 0;  // level 0 lives in the source buffer (OrbxSrc)
            L.off = 0;
        } else {
            L.pitch = round_up(L.w, 64);
            L.off = off;
            off += (long long)L.pitch * L.h;
            off = (off + 255) & ~255ll;
        }
        // ORBextractor.cc:771-785
        L.max_bx = L.w - ORBX_FAST_BORDER;
        L.max_by = L.h - ORBX_FAST_BORDER;
        const float width = (float)(L.max_bx - ORBX_FAST_BORDER);
        const float height = (float)(L.max_by - ORBX_FAST_BORDER);
        L.n_cols = (int)(width / 30.f);
        L.n_rows = (int)(height / 30.f);
        if (L.n_cols < 1 || L.n_rows < 1)
            return fail(ORBX_E_UNSUPPORTED, "level %d (%dx%d) has no 30-px FAST cell; the reference divides by zero here", l, L.w, L.h);
        L.n_cols_inv = L.n_cols == 1 ? 0u : (uint32_t)(((1ull << 32) + L.n_cols - 1) / L.n_cols);
        L.w_cell = (int)ceilf(width / L.n_cols);
        L.h_cell = (int)ceilf(height / L.n_rows);
        L.cell_cap = ((L.w_cell + 1) / 2) * ((L.h_cell + 1) / 2);
        L.cell_off = cell_off;
        cell_off += L.n_cols * L.n_rows;
        L.key_off = key_off;
        L.key_cap = L.n_cols * L.n_rows * L.cell_cap;
        key_off += L.key_cap;
        P.max_tile_w = std::max(P.max_tile_w, L.w_cell + 6);
        P.max_tile_h = std::max(P.max_tile_h, L.h_cell + 6);
        // ORBextractor.cc:541-543
        L.n_target = h->nper[l];
        const int Wd = L.max_bx - ORBX_FAST_BORDER, Hd = L.max_by - ORBX_FAST_BORDER;
        L.n_ini = (int)roundf((float)Wd / Hd);
        if (L.n_ini < 1 || L.n_ini > 64)
            return fail(ORBX_E_UNSUPPORTED, "level %d: nIni=%d (aspect ratio outside [0.5, 64.5)); the reference divides by zero for nIni=0", l, L.n_ini);
        L.hx = (float)Wd / L.n_ini;
        L.node_cap = std::max(L.n_target + 3, 4 * L.n_ini) + 5;
        if (L.node_cap > 8192) return fail(ORBX_E_UNSUPPORTED, "more than 8184 features on one level");
        L.surv_off = surv_off;
        L.surv_cap = L.node_cap;
        surv_off += L.surv_cap;
        // resize tables (level l from level l-1)
        if (l > 0) {
            std::vector<uint16_t> xo, yo;
            std::vector<short> xa, ya;
            linear_axis(P.lvl[l - 1].w, L.w, xo, xa);
            linear_axis(P.lvl[l - 1].h, L.h, yo, ya);
            // stored as offsets; patched to device pointers by the caller
            L.xofs = (const uint16_t*)put(xo.data(), xo.size() * 2);
            L.xa = (const short*)put(xa.data(), xa.size() * 2);
            L.yofs = (const uint16_t*)put(yo.data(), yo.size() * 2);
            L.ya = (const short*)put(ya.data(), ya.size() * 2);
            std::vector<uint32_t> yt((size_t)2 * L.h);
            // Row table of the walking kernel: {s0, b0 | b1 << 16}.  A row clamped to the last
            // source row (s0 == s1 == sh-1, b1 == 0) is re-expressed as top = sh-2 with weight 0,
            // bottom = sh-1 with weight b0: the same value, and "bottom row = s0 + 1" always holds.
            const int sh_last = P.lvl[l - 1].h - 1;
            L.y_monotone = 1;
            for (int y = 0; y < L.h; ++y) {
                int s0 = yo[y];
                uint32_t b0 = (uint16_t)ya[2 * y], b1 = (uint16_t)ya[2 * y + 1];
                if (s0 >= sh_last && sh_last >= 1) {
                    s0 = sh_last - 1;
                    b1 = b0;
                    b0 = 0;
                }
                yt[2 * y] = (uint32_t)s0;
                yt[2 * y + 1] = b0 | (b1 << 16);
                if (y > 0 && yt[2 * y] <= yt[2 * (y - 1)]) L.y_monotone = 0;
            }
            L.ytab = (const uint2*)put(yt.data(), yt.size() * 4);
            L.x_step_max = 0;
            for (int x = 0; x + 1 < L.w; ++x) L.x_step_max = std::max(L.x_step_max, (int)xo[x + 1] - (int)xo[x]);
        }
    }
    P.frame_stride = std::max(off, 256ll);
    P.cells_total = cell_off;
    P.keys_total = key_off;
    P.surv_total = surv_off;
    P.out_cap = surv_off;
    if (P.max_tile_w > 80 || P.max_tile_h > 80) return fail(ORBX_E_UNSUPPORTED, "FAST cell larger than 74 px");
    // per-cell records of the FAST kernel (ORBextractor.cc:787-804)
    {
        std::vector<uint32_t> ct((size_t)4 * P.cells_total, 0u);
        for (int l = 0; l < nl; ++l) {
            const OrbxLevel& L = P.lvl[l];
            for (int ci = 0; ci < L.n_rows; ++ci)
                for (int cj = 0; cj < L.n_cols; ++cj) {
                    const int c = ci * L.n_cols + cj;
                    uint32_t* r = &ct[(size_t)4 * (L.cell_off + c)];
                    const int iniY = ORBX_FAST_BORDER + ci * L.h_cell, iniX = ORBX_FAST_BORDER + cj * L.w_cell;
                    const int maxY = std::min(iniY + L.h_cell + 6, L.max_by), maxX = std::min(iniX + L.w_cell + 6, L.max_bx);
                    int tw = maxX - iniX, th = maxY - iniY;
                    if (iniY >= L.max_by - 3 || iniX >= L.max_bx - 6 || tw < 7 || th < 7) tw = th = 0;
                    r[0] = (uint32_t)iniX | ((uint32_t)iniY << 16);
                    r[1] = (uint32_t)tw | ((uint32_t)th << 8) | ((uint32_t)l << 16);
                    r[2] = (uint32_t)(L.key_off + (long long)c * L.cell_cap);
                    r[3] = (uint32_t)(3 + cj * L.w_cell) | ((uint32_t)(3 + ci * L.h_cell) << 16);
                }
        }
        if (P.keys_total >= (1ll << 32)) return fail(ORBX_E_UNSUPPORTED, "key slab too large");
        P.cell_tab = (const uint4*)put(ct.data(), ct.size() * 4);
    }
    return ORBX_OK;
}

void patch_tables(OrbxPlan& P, uint8_t* dev_base) {
    P.cell_tab = (const uint4*)(dev_base + (size_t)P.cell_tab);
    for (int l = 1; l < P.nlevels; ++l) {
        OrbxLevel& L = P.lvl[l];
        L.xofs = (const uint16_t*)(dev_base + (size_t)L.xofs);
        L.xa = (const short*)(dev_base + (size_t)L.xa);
        L.yofs = (const uint16_t*)(dev_base + (size_t)L.yofs);
        L.ya = (const short*)(dev_base + (size_t)L.ya);
        L.ytab = (const uint2*)(dev_base + (size_t)L.ytab);
    }
}

}  // namespace

// =================================================================================================
// buffers
// =================================================================================================
namespace {

struct Sizes {
    size_t img_in, pyr, cellcnt, keys, surv, survcnt, header, kps, desc, lap;
};

Sizes sizes_for(const OrbxPlan& P, int B) {
    Sizes z;
    z.img_in = (size_t)B * round_up(P.w, 16) * P.h + 64;
    z.pyr = (size_t)B * P.frame_stride + 64;
    z.cellcnt = (size_t)B * P.cells_total * 4;
    z.keys = (size_t)B * P.keys_total * 4;
    z.surv = (size_t)B * P.surv_total * 4;
    z.survcnt = (size_t)B * ORBX_MAX_LEVELS * 4;
    z.header = (size_t)B * 16;
    z.kps = (size_t)B * P.out_cap * sizeof(orbx_keypoint);
    z.desc = (size_t)B * P.out_cap * 32;
    z.lap = (size_t)B * 8;
    return z;
}

template <typename T>
int grow(T** p, size_t* have, size_t need) {
    if (need <= *have && *p) return ORBX_OK;
    if (*p) cudaFree(*p);
    *p = nullptr;
    *have = 0;
    need += need / 8;  // headroom for the next, slightly different shape
    CUDA_TRY(cudaMalloc((void**)p, need));
    *have = need;
    return ORBX_OK;
}

int ensure_buffers(orbx_handle* h, const OrbxPlan& P) {
    const Sizes z = sizes_for(P, h->cfg.max_batch);
    for (size_t i = 0; i < h->slots.size(); ++i) {
        Slot& s = h->slots[i];
        SlotCaps& c = s.caps;
        int r;
        if ((r = grow(&s.img_in, &c.img_in, z.img_in))) return r;
        if ((r = grow(&s.pyr, &c.pyr, z.pyr))) return r;
        if ((r = grow(&s.cellcnt, &c.cellcnt, z.cellcnt))) return r;
        if ((r = grow(&s.cellkeys, &c.cellkeys, z.keys))) return r;
        if ((r = grow(&s.keysA, &c.keysA, z.keys))) return r;
        if ((r = grow(&s.keysB, &c.keysB, z.keys))) return r;
        if ((r = grow(&s.surv, &c.surv, z.surv))) return r;
        if ((r = grow(&s.survcnt, &c.survcnt, z.survcnt))) return r;
        if ((r = grow(&s.dest, &c.dest, 2 * z.surv))) return r;  // uint2 records
        if ((r = grow(&s.header, &c.header, z.header))) return r;
        if ((r = grow(&s.kps, &c.kps, z.kps))) return r;
        if ((r = grow(&s.desc, &c.desc, z.desc))) return r;
        {
            int* before = s.lap_d;
            if ((r = grow(&s.lap_d, &c.lap, z.lap))) return r;
            if (s.lap_d != before) s.lap_valid = 0;
        }
        if (c.frame_done < z.lap) {
            if ((r = grow(&s.frame_done, &c.frame_done, z.lap))) return r;
            CUDA_TRY(cudaMemset(s.frame_done, 0, c.frame_done));
        }
    }
    return ORBX_OK;
}

int sync_all(orbx_handle* h) {
    for (Slot& s : h->slots) CUDA_TRY(cudaStreamSynchronize(s.st));
    return ORBX_OK;
}

// Switch the handle to image shape (w, hgt): plan, tables upload, kernel attributes, buffers.
int set_shape(orbx_handle* h, int w, int hgt) {
    if (h->plan.w == w && h->plan.h == hgt) return ORBX_OK;
    OrbxPlan P;
    std::vector<uint8_t> tab;
    int r = build_plan(h, w, hgt, &P, &tab);
    if (r) return r;
    // a new shape reallocates the buffers of EVERY slot: device pointers handed out for a slot that is still in flight
    // (orbx_extract_device results, orbx_device_pyramid views, a sink) would dangle -- refuse instead
    for (size_t i = 0; i < h->slots.size(); ++i)
        if (h->slots[i].in_flight)
            return fail(ORBX_E_BUSY, "image shape changes to %dx%d while slot %d has an un-waited submit", w, hgt, (int)i);
    if ((r = sync_all(h))) return r;
    size_t need = std::max(tab.size(), (size_t)256);
    if ((r = grow(&h->tables, &h->tables_bytes, need))) return r;
    if (!tab.empty()) CUDA_TRY(cudaMemcpy(h->tables, tab.data(), tab.size(), cudaMemcpyHostToDevice));
    patch_tables(P, h->tables);
    CUDA_TRY(orbx_configure_fast(P));
    cudaError_t e = orbx_configure_octree(P);
    if (e != cudaSuccess)
        return fail(ORBX_E_UNSUPPORTED, "quadtree working set exceeds shared memory (nfeatures too large)");
    if ((r = ensure_buffers(h, P))) return r;
    h->plan = P;
    ++h->epoch;   // captured graphs hold the old plan and buffer addresses
    return ORBX_OK;
}

}  // namespace

// =================================================================================================
// pipeline
// =================================================================================================
namespace {

int launch_stages(orbx_handle* h, Slot& s, const OrbxSrc& src, int nframes, bool timing) {
    const OrbxPlan& P = h->plan;
    s.launches = 0;
    if (timing) CUDA_TRY(cudaEventRecord(s.ev[0], s.st));
    orbx_launch_pyramid(P, src, s.pyr, nframes, s.st, &s.launches);
    CUDA_TRY(cudaGetLastError());
    if (timing) CUDA_TRY(cudaEventRecord(s.ev[1], s.st));
    CUDA_TRY(orbx_launch_fast(P, src, s.pyr, s.cellcnt, s.cellkeys, nframes, s.st, &s.launches));
    if (timing) CUDA_TRY(cudaEventRecord(s.ev[2], s.st));
    // quadtree + (in its tail) the two-ended output layout of every frame
    CUDA_TRY(orbx_launch_octree(P, s.cellcnt, s.cellkeys, s.keysA, s.keysB, s.surv, s.survcnt, s.lap_d, s.dest,
                                s.header, s.frame_done, nframes, s.st, &s.launches));
    if (timing) CUDA_TRY(cudaEventRecord(s.ev[3], s.st));
    if (timing) CUDA_TRY(cudaEventRecord(s.ev[4], s.st));
    CUDA_TRY(orbx_launch_describe(P, src, s.pyr, s.header, s.dest, s.kps, s.desc, nframes,
                                  s.st, &s.launches, s.sink));
    if (timing) CUDA_TRY(cudaEventRecord(s.ev[5], s.st));
    return ORBX_OK;
}

// The same stages as a fork-join, for capture into the latency graph: FAST and the quadtree run per
// level, each level on its own branch as soon as its image exists (level 0 at once, level l after
// pyramid launch l), while the pyramid chain continues on the main branch; describe joins them.  The
// critical path of one frame drops from pyramid + FAST + quadtree to about max(pyramid chain + the
// last level's FAST/quadtree, level 0's FAST + quadtree).  Results are identical: the kernels and
// their per-level work are the same, only launch granularity and ordering between levels change.
int launch_stages_forked(orbx_handle* h, Slot& s, const OrbxSrc& src, int nframes) {
    const OrbxPlan& P = h->plan;
    const int nl = P.nlevels;
    if (!s.ev_fork) CUDA_TRY(cudaEventCreateWithFlags(&s.ev_fork, cudaEventDisableTiming));
    for (int l = 0; l < nl; ++l) {
        if (!s.aux[l]) CUDA_TRY(cudaStreamCreateWithFlags(&s.aux[l], cudaStreamNonBlocking));
        if (!s.ev_lvl[l]) CUDA_TRY(cudaEventCreateWithFlags(&s.ev_lvl[l], cudaEventDisableTiming));
        if (!s.ev_join[l]) CUDA_TRY(cudaEventCreateWithFlags(&s.ev_join[l], cudaEventDisableTiming));
    }
    s.launches = 0;
    CUDA_TRY(cudaEventRecord(s.ev_fork, s.st));
    orbx_launch_pyramid(P, src, s.pyr, nframes, s.st, &s.launches, s.ev_lvl);
    CUDA_TRY(cudaGetLastError());
    for (int l = 0; l < nl; ++l) {
        CUDA_TRY(cudaStreamWaitEvent(s.aux[l], l == 0 ? s.ev_fork : s.ev_lvl[l], 0));
        CUDA_TRY(orbx_launch_fast(P, src, s.pyr, s.cellcnt, s.cellkeys, nframes, s.aux[l], &s.launches, l));
        CUDA_TRY(orbx_launch_octree(P, s.cellcnt, s.cellkeys, s.keysA, s.keysB, s.surv, s.survcnt, s.lap_d, s.dest,
                                    s.header, s.frame_done, nframes, s.aux[l], &s.launches, l));
        CUDA_TRY(cudaEventRecord(s.ev_join[l], s.aux[l]));
    }
    for (int l = 0; l < nl; ++l) CUDA_TRY(cudaStreamWaitEvent(s.st, s.ev_join[l], 0));
    CUDA_TRY(orbx_launch_describe(P, src, s.pyr, s.header, s.dest, s.kps, s.desc, nframes, s.st, &s.launches, s.sink));
    return ORBX_OK;
}

// Enqueue the whole device pipeline for `nframes` frames whose level 0 is `src`.
// Small batches (the single-frame latency path) replay a CUDA graph of the kernel sequence: the
// second run of a (plan, buffers, nframes, source) combination on a slot captures it, later runs
// launch the instantiated graph -- one launch instead of eleven, no host gaps between the levels.
int enqueue_kernels(orbx_handle* h, Slot& s, const OrbxSrc& src, int nframes) {
    const OrbxPlan& P = h->plan;
    const bool timing = (h->cfg.flags & ORBX_FLAG_STAGE_TIMING) != 0;
    // (only for the handle's own staging buffer: a caller-owned device source may change every call)
    const bool graph_ok = !timing && !(h->cfg.flags & ORBX_FLAG_NO_GRAPH) && nframes <= ORBX_GRAPH_MAX_FRAMES &&
                          src.img0 == s.img_in;
    bool done = false;
    if (graph_ok) {
        const bool same = s.gexec && s.g_epoch == h->epoch && s.g_nframes == nframes && s.g_src.img0 == src.img0 &&
                          s.g_src.pitch0 == src.pitch0 && s.g_src.stride0 == src.stride0;
        if (!same && s.warm_epoch == h->epoch) {
            if (s.gexec) {
                cudaGraphExecDestroy(s.gexec);
                s.gexec = nullptr;
            }
            cudaGraph_t g = nullptr;
            if (cudaStreamBeginCapture(s.st, cudaStreamCaptureModeThreadLocal) == cudaSuccess) {
                const int r = (h->cfg.flags & ORBX_FLAG_NO_FORK) ? launch_stages(h, s, src, nframes, false)
                                                                 : launch_stages_forked(h, s, src, nframes);
                const cudaError_t e = cudaStreamEndCapture(s.st, &g);
                if (r == ORBX_OK && e == cudaSuccess && g &&
                    cudaGraphInstantiate(&s.gexec, g, 0) == cudaSuccess) {
                    s.g_epoch = h->epoch;
                    s.g_nframes = nframes;
                    s.g_src = src;
                    s.g_launches = s.launches;
                } else {
                    s.gexec = nullptr;
                    cudaGetLastError();
                }
                if (g) cudaGraphDestroy(g);
            } else {
                cudaGetLastError();
            }
        }
        if (s.gexec && s.g_epoch == h->epoch && s.g_nframes == nframes && s.g_src.img0 == src.img0 &&
            s.g_src.pitch0 == src.pitch0 && s.g_src.stride0 == src.stride0) {
            CUDA_TRY(cudaGraphLaunch(s.gexec, s.st));
            s.launches = s.g_launches;
            done = true;
        }
    }
    if (!done) {
        int r = launch_stages(h, s, src, nframes, timing);
        if (r) return r;
        s.warm_epoch = h->epoch;   // lazy one-time initialisation has run: the next call may capture
    }
    s.src = src;
    s.run_w = P.w;
    s.run_h = P.h;
    s.nframes = nframes;
    if (s.sink.hdr) ++s.sink_epoch;   // k_sink_signal publishes the same count on the device side
    return ORBX_OK;
}

// N4: raw frames (device) -> rectified gray level 0 in s.img_in.  Returns the level-0 source.
int run_input_stage(orbx_handle* h, Slot& s, const uint8_t* d_raw, long long raw_fstride, int raw_pitch, int w,
                    int hgt, int nframes, OrbxSrc* src, int* launches) {
    if (h->d_map && (h->map_w != w || h->map_h != hgt))
        return fail(ORBX_E_BAD_ARG, "rectification maps are %dx%d, the image is %dx%d", h->map_w, h->map_h, w, hgt);
    const int pitch0 = round_up(w, 16);
    const long long fstride = (long long)pitch0 * hgt;
    const bool timing = (h->cfg.flags & ORBX_FLAG_STAGE_TIMING) != 0;
    if (timing) CUDA_TRY(cudaEventRecord(s.ev_in[0], s.st));
    CUDA_TRY(orbx_launch_input(d_raw, raw_fstride, raw_pitch, w, hgt, h->channels(),
                               h->in_format == ORBX_IN_RGB8 || h->in_format == ORBX_IN_RGBA8, h->gray_bits, h->d_map,
                               s.img_in, fstride, pitch0, nframes, s.st, launches));
    if (timing) CUDA_TRY(cudaEventRecord(s.ev_in[1], s.st));
    s.in_timed = timing;
    src->img0 = s.img_in;
    src->pitch0 = pitch0;
    src->stride0 = fstride;
    src->_pad = 0;
    return ORBX_OK;
}

int check_common(orbx_handle* h, int slot, int nframes, int w, int hgt) {
    if (!h) return fail(ORBX_E_BAD_ARG, "null handle");
    if (slot < 0 || slot >= (int)h->slots.size()) return fail(ORBX_E_BAD_ARG, "slot %d out of range", slot);
    if (nframes < 1) return fail(ORBX_E_BAD_ARG, "nframes must be >= 1");
    if (nframes > h->cfg.max_batch) return fail(ORBX_E_TOO_LARGE, "nframes %d > max_batch %d", nframes, h->cfg.max_batch);
    if (w <= 0 || hgt <= 0) return fail(ORBX_E_EMPTY_IMAGE, "empty image");
    if (w > h->cfg.max_width || hgt > h->cfg.max_height)
        return fail(ORBX_E_TOO_LARGE, "%dx%d exceeds the handle's %dx%d", w, hgt, h->cfg.max_width, h->cfg.max_height);
    if (h->slots[slot].in_flight) return fail(ORBX_E_BUSY, "slot %d has an un-waited submit", slot);
    return ORBX_OK;
}

int upload_lap(Slot& s, const int* lap, int nframes) {
    // the lapping areas rarely change between calls: skip the copy when the device already holds them
    bool same = s.lap_valid >= nframes;
    for (int i = 0; i < nframes; ++i) {
        const int a = lap ? lap[2 * i] : 0, b = lap ? lap[2 * i + 1] : 0;
        same = same && s.h_lap[2 * i] == a && s.h_lap[2 * i + 1] == b;
        s.h_lap[2 * i] = a;
        s.h_lap[2 * i + 1] = b;
    }
    if (same) return ORBX_OK;
    s.lap_valid = nframes;
    CUDA_TRY(cudaMemcpyAsync(s.lap_d, s.h_lap, (size_t)nframes * 8, cudaMemcpyHostToDevice, s.st));
    return ORBX_OK;
}

}  // namespace

extern "C" {

int orbx_abi_version(void) { return ORBX_ABI_VERSION; }
const char* orbx_last_error_string(void) { return g_err.c_str(); }

int orbx_create(const orbx_config* cfg, orbx_handle** out) {
    if (!cfg || !out) return fail(ORBX_E_BAD_ARG, "null argument");
    *out = nullptr;
    if (cfg->nlevels < 1 || cfg->nlevels > ORBX_MAX_LEVELS) return fail(ORBX_E_BAD_ARG, "nlevels must be 1..%d", ORBX_MAX_LEVELS);
    if (cfg->nfeatures < 1) return fail(ORBX_E_BAD_ARG, "nfeatures must be >= 1");
    if (!(cfg->scale_factor > 1.0f)) return fail(ORBX_E_BAD_ARG, "scale_factor must be > 1");
    if (cfg->max_width < 1 || cfg->max_height < 1 || cfg->max_batch < 1) return fail(ORBX_E_BAD_ARG, "max_width/max_height/max_batch must be >= 1");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail(ORBX_E_CUDA, "no CUDA device (%s); this library has no CPU fallback", cudaGetErrorString(e));
    if (cfg->device < 0 || cfg->device >= ndev) return fail(ORBX_E_BAD_ARG, "device %d out of range", cfg->device);
    CUDA_TRY(cudaSetDevice(cfg->device));
    orbx_handle* h = new orbx_handle();
    h->cfg = *cfg;
    if (h->cfg.n_slots <= 0) h->cfg.n_slots = 2;
    memset(&h->plan, 0, sizeof h->plan);
    build_tables(h);
    h->slots.resize(h->cfg.n_slots);
    int r = ORBX_OK;
    for (Slot& s : h->slots) {
        if (cudaStreamCreateWithFlags(&s.st, cudaStreamNonBlocking) != cudaSuccess ||
            cudaMallocHost((void**)&s.h_header, (size_t)h->cfg.max_batch * 16) != cudaSuccess ||
            cudaMallocHost((void**)&s.h_lap, (size_t)h->cfg.max_batch * 8) != cudaSuccess) {
            r = fail(ORBX_E_CUDA, "stream / pinned allocation failed: %s", cudaGetErrorString(cudaGetLastError()));
            break;
        }
        for (int k = 0; k < 6; ++k) cudaEventCreate(&s.ev[k]);
        for (int k = 0; k < 2; ++k) cudaEventCreate(&s.ev_in[k]);
    }
    if (r == ORBX_OK) r = set_shape(h, cfg->max_width, cfg->max_height);
    if (r != ORBX_OK) {
        std::string keep = g_err;
        orbx_destroy(h);
        g_err = keep;
        return r;
    }
    *out = h;
    return ORBX_OK;
}

int orbx_destroy(orbx_handle* h) {
    if (!h) return ORBX_OK;
    cudaSetDevice(h->cfg.device);
    for (Slot& s : h->slots) {
        if (s.st) cudaStreamSynchronize(s.st);
        cudaFree(s.raw_in);
        cudaFree(s.img_in); cudaFree(s.pyr); cudaFree(s.cellcnt); cudaFree(s.cellkeys);
        cudaFree(s.keysA); cudaFree(s.keysB); cudaFree(s.surv); cudaFree(s.survcnt);
        cudaFree(s.dest); cudaFree(s.header); cudaFree(s.kps); cudaFree(s.desc); cudaFree(s.lap_d);
        cudaFree(s.frame_done);
        if (s.sink_words) cudaFree(s.sink_words);
        if (s.h_header) cudaFreeHost(s.h_header);
        if (s.h_lap) cudaFreeHost(s.h_lap);
        for (int k = 0; k < 6; ++k) if (s.ev[k]) cudaEventDestroy(s.ev[k]);
        for (int k = 0; k < 2; ++k) if (s.ev_in[k]) cudaEventDestroy(s.ev_in[k]);
        if (s.gexec) cudaGraphExecDestroy(s.gexec);
        for (int l = 0; l < ORBX_MAX_LEVELS; ++l) {
            if (s.aux[l]) cudaStreamDestroy(s.aux[l]);
            if (s.ev_lvl[l]) cudaEventDestroy(s.ev_lvl[l]);
            if (s.ev_join[l]) cudaEventDestroy(s.ev_join[l]);
        }
        if (s.ev_fork) cudaEventDestroy(s.ev_fork);
        if (s.st) cudaStreamDestroy(s.st);
    }
    cudaFree(h->tables);
    cudaFree(h->d_map);
    delete h;
    return ORBX_OK;
}

int orbx_get_tables(const orbx_handle* h, float* scale, float* inv_scale, float* sigma2,
                    float* inv_sigma2, int32_t* n_per_level) {
    if (!h) return fail(ORBX_E_BAD_ARG, "null handle");
    for (int i = 0; i < h->cfg.nlevels; ++i) {
        if (scale) scale[i] = h->sf[i];
        if (inv_scale) inv_scale[i] = h->isf[i];
        if (sigma2) sigma2[i] = h->s2[i];
        if (inv_sigma2) inv_sigma2[i] = h->is2[i];
        if (n_per_level) n_per_level[i] = h->nper[i];
    }
    return ORBX_OK;
}

int orbx_max_keypoints(const orbx_handle* h) { return h ? h->plan.out_cap : ORBX_E_BAD_ARG; }

void* orbx_host_alloc(size_t bytes) {
    void* p = nullptr;
    if (cudaMallocHost(&p, bytes ? bytes : 1) != cudaSuccess) {
        fail(ORBX_E_CUDA, "cudaMallocHost(%zu) failed", bytes);
        return nullptr;
    }
    return p;
}
void orbx_host_free(void* p) { if (p) cudaFreeHost(p); }

void* orbx_slot_stream(orbx_handle* h, int slot) {
    if (!h || slot < 0 || slot >= (int)h->slots.size()) return nullptr;
    return (void*)h->slots[slot].st;
}

int orbx_last_launch_count(const orbx_handle* h, int slot) {
    if (!h || slot < 0 || slot >= (int)h->slots.size()) return ORBX_E_BAD_ARG;
    return h->slots[slot].launches;
}

int orbx_last_stage_ms(orbx_handle* h, int slot, float* ms, int n) {
    if (!h || slot < 0 || slot >= (int)h->slots.size() || !ms) return fail(ORBX_E_BAD_ARG, "bad argument");
    if (!(h->cfg.flags & ORBX_FLAG_STAGE_TIMING)) return fail(ORBX_E_BAD_ARG, "handle created without ORBX_FLAG_STAGE_TIMING");
    Slot& s = h->slots[slot];
    if (s.in_flight) return fail(ORBX_E_BUSY, "slot %d not waited", slot);
    for (int k = 0; k < 5 && k < n; ++k) CUDA_TRY(cudaEventElapsedTime(&ms[k], s.ev[k], s.ev[k + 1]));
    if (n > 5) CUDA_TRY(cudaEventElapsedTime(&ms[5], s.ev[0], s.ev[5]));
    if (n > 6) {
        ms[6] = 0.f;
        if (s.in_timed && h->input_stage()) CUDA_TRY(cudaEventElapsedTime(&ms[6], s.ev_in[0], s.ev_in[1]));
    }
    return n < 7 ? n : 7;
}

int orbx_level_size(const orbx_handle* h, int w, int hgt, int level, int* lw, int* lh) {
    if (!h || level < 0 || level >= h->cfg.nlevels) return fail(ORBX_E_BAD_ARG, "bad level");
    if (lw) *lw = cv_round((float)w * h->isf[level]);
    if (lh) *lh = cv_round((float)hgt * h->isf[level]);
    return ORBX_OK;
}

int orbx_set_input_format(orbx_handle* h, int format, int gray_bits) {
    if (!h) return fail(ORBX_E_BAD_ARG, "null handle");
    if (format < ORBX_IN_GRAY8 || format > ORBX_IN_RGBA8) return fail(ORBX_E_BAD_TYPE, "unknown input format %d", format);
    if (gray_bits != 14 && gray_bits != 15) return fail(ORBX_E_BAD_ARG, "gray_bits must be 14 or 15");
    for (Slot& s : h->slots)
        if (s.in_flight) return fail(ORBX_E_BUSY, "a slot has an un-waited submit");
    h->in_format = format;
    h->gray_bits = gray_bits;
    return ORBX_OK;
}

int orbx_set_rectify_maps(orbx_handle* h, const float* map_x, const float* map_y, int w, int hgt,
                          size_t map_stride_floats) {
    if (!h) return fail(ORBX_E_BAD_ARG, "null handle");
    for (Slot& s : h->slots)
        if (s.in_flight) return fail(ORBX_E_BUSY, "a slot has an un-waited submit");
    CUDA_TRY(cudaSetDevice(h->cfg.device));
    int r;
    if ((r = sync_all(h))) return r;
    if (h->d_map) cudaFree(h->d_map);
    h->d_map = nullptr;
    h->map_w = h->map_h = 0;
    if (!map_x && !map_y) return ORBX_OK;
    if (!map_x || !map_y || w <= 0 || hgt <= 0 || map_stride_floats < (size_t)w)
        return fail(ORBX_E_BAD_ARG, "bad rectification maps");
    if (w > h->cfg.max_width || hgt > h->cfg.max_height)
        return fail(ORBX_E_TOO_LARGE, "maps %dx%d exceed the handle's %dx%d", w, hgt, h->cfg.max_width, h->cfg.max_height);
    // cv::remap's per-call conversion of CV_32FC1 maps, done once: sx = cvRound(map * INTER_TAB_SIZE)
    // (float product, round half to even), integer part saturated to short, 5-bit fractions
    // then everything frame independent is folded into the entry: taps clamped into the image, the
    // weights of taps outside it zeroed (BORDER_CONSTANT 0)
    std::vector<uint4> m((size_t)w * hgt);
    for (int y = 0; y < hgt; ++y)
        for (int x = 0; x < w; ++x) {
            const int sx = (int)lrintf(map_x[(size_t)y * map_stride_floats + x] * 32.0f);
            const int sy = (int)lrintf(map_y[(size_t)y * map_stride_floats + x] * 32.0f);
            const int ix = std::min(32767, std::max(-32768, sx >> 5)), iy = std::min(32767, std::max(-32768, sy >> 5));
            const int fx = sx & 31, fy = sy & 31;
            const bool x0 = ix >= 0 && ix < w, x1 = ix + 1 >= 0 && ix + 1 < w;
            const bool y0 = iy >= 0 && iy < hgt, y1 = iy + 1 >= 0 && iy + 1 < hgt;
            const uint32_t a00 = (x0 && y0) ? (32 - fx) * (32 - fy) : 0, a01 = (x1 && y0) ? fx * (32 - fy) : 0;
            const uint32_t a10 = (x0 && y1) ? (32 - fx) * fy : 0, a11 = (x1 && y1) ? fx * fy : 0;
            const int cx0 = std::min(w - 1, std::max(0, ix)), cx1 = std::min(w - 1, std::max(0, ix + 1));
            const int cy0 = std::min(hgt - 1, std::max(0, iy)), cy1 = std::min(hgt - 1, std::max(0, iy + 1));
            m[(size_t)y * w + x] = make_uint4((uint32_t)cx0 | ((uint32_t)cy0 << 16), a00 | (a01 << 16), a10 | (a11 << 16),
                                              (uint32_t)(cx1 - cx0) | ((uint32_t)(cy1 - cy0) << 1));
        }
    CUDA_TRY(cudaMalloc(&h->d_map, m.size() * sizeof(uint4)));
    CUDA_TRY(cudaMemcpy(h->d_map, m.data(), m.size() * sizeof(uint4), cudaMemcpyHostToDevice));
    h->map_w = w;
    h->map_h = hgt;
    return ORBX_OK;
}

int orbx_extract_device(orbx_handle* h, int slot, const uint8_t* d_imgs, int nframes, int w, int hgt,
                        size_t pitch_bytes, size_t frame_stride_bytes, const int* lap,
                        const int32_t** d_header, const orbx_keypoint** d_kps, const uint8_t** d_desc) {
    int r = check_common(h, slot, nframes, w, hgt);
    if (r) return r;
    if (!d_imgs) return fail(ORBX_E_EMPTY_IMAGE, "null image");
    if (!h->input_stage() &&
        (((uintptr_t)d_imgs & 3) || (pitch_bytes & 3) || (frame_stride_bytes & 3) || pitch_bytes < (size_t)round_up(w, 4)))
        return fail(ORBX_E_BAD_ARG, "device images need 4-byte aligned pointer/pitch/stride and pitch >= round_up(w,4)");
    if (h->input_stage() && pitch_bytes < (size_t)w * h->channels())
        return fail(ORBX_E_BAD_ARG, "pitch < width * channels");
    CUDA_TRY(cudaSetDevice(h->cfg.device));
    if ((r = set_shape(h, w, hgt))) return r;
    Slot& s = h->slots[slot];
    if ((r = upload_lap(s, lap, nframes))) return r;
    OrbxSrc src;
    src.img0 = d_imgs;
    src.pitch0 = (int)pitch_bytes;
    src.stride0 = (long long)frame_stride_bytes;
    src._pad = 0;
    int in_launches = 0;
    if (h->input_stage() &&
        (r = run_input_stage(h, s, d_imgs, (long long)frame_stride_bytes, (int)pitch_bytes, w, hgt, nframes, &src,
                             &in_launches)))
        return r;
    if ((r = enqueue_kernels(h, s, src, nframes))) return r;
    s.launches += in_launches;
    s.in_flight = true;
    s.u_n = s.u_mono = s.u_status = nullptr;
    s.cap = 0;
    if (d_header) *d_header = s.header;
    if (d_kps) *d_kps = s.kps;
    if (d_desc) *d_desc = s.desc;
    return ORBX_OK;
}

int orbx_submit_batch(orbx_handle* h, int slot, const uint8_t* const* imgs, int nframes, int w, int hgt,
                      size_t stride_bytes, const int* lap, orbx_keypoint* kps, uint8_t* desc, int cap,
                      int* n_out, int* mono_index, int* status) {
    int r = check_common(h, slot, nframes, w, hgt);
    if (r) return r;
    if (!imgs || !kps || !desc || !n_out) return fail(ORBX_E_BAD_ARG, "null argument");
    if (cap < 1) return fail(ORBX_E_BAD_ARG, "cap must be >= 1");
    if (stride_bytes < (size_t)w) return fail(ORBX_E_BAD_ARG, "stride < width");
    for (int i = 0; i < nframes; ++i)
        if (!imgs[i]) return fail(ORBX_E_EMPTY_IMAGE, "frame %d is null", i);
    CUDA_TRY(cudaSetDevice(h->cfg.device));
    if ((r = set_shape(h, w, hgt))) return r;
    Slot& s = h->slots[slot];
    const OrbxPlan& P = h->plan;
    OrbxSrc src;
    int in_launches = 0;
    if ((r = upload_lap(s, lap, nframes))) return r;
    if (h->input_stage()) {
        // raw frames -> s.raw_in (rows raw_pitch apart), then remap / luma into the level-0 staging
        const int ch = h->channels();
        if (stride_bytes < (size_t)w * ch) return fail(ORBX_E_BAD_ARG, "stride < width * channels");
        const int raw_pitch = round_up(w * ch, 16);
        const size_t raw_fstride = (size_t)raw_pitch * hgt;
        const size_t need = raw_fstride * h->cfg.max_batch + 64;
        if (s.raw_cap < need) {
            CUDA_TRY(cudaStreamSynchronize(s.st));
            if (s.raw_in) cudaFree(s.raw_in);
            s.raw_in = nullptr;
            s.raw_cap = 0;
            CUDA_TRY(cudaMalloc(&s.raw_in, need));
            s.raw_cap = need;
        }
        bool contiguous = stride_bytes == (size_t)raw_pitch;
        for (int i = 1; i < nframes && contiguous; ++i) contiguous = imgs[i] == imgs[0] + (size_t)i * raw_fstride;
        if (contiguous) {
            CUDA_TRY(cudaMemcpyAsync(s.raw_in, imgs[0], raw_fstride * nframes, cudaMemcpyHostToDevice, s.st));
        } else {
            for (int i = 0; i < nframes; ++i)
                CUDA_TRY(cudaMemcpy2DAsync(s.raw_in + (size_t)i * raw_fstride, raw_pitch, imgs[i], stride_bytes,
                                           (size_t)w * ch, hgt, cudaMemcpyHostToDevice, s.st));
        }
        if ((r = run_input_stage(h, s, s.raw_in, (long long)raw_fstride, raw_pitch, w, hgt, nframes, &src, &in_launches)))
            return r;
    } else {
        // H2D: one copy when the frames are contiguous and tightly packed at the device pitch
        const int pitch0 = round_up(w, 16);
        const size_t fstride = (size_t)pitch0 * hgt;
        bool contiguous = stride_bytes == (size_t)pitch0;
        for (int i = 1; i < nframes && contiguous; ++i) contiguous = imgs[i] == imgs[0] + (size_t)i * fstride;
        if (contiguous) {
            CUDA_TRY(cudaMemcpyAsync(s.img_in, imgs[0], fstride * nframes, cudaMemcpyHostToDevice, s.st));
        } else {
            for (int i = 0; i < nframes; ++i)
                CUDA_TRY(cudaMemcpy2DAsync(s.img_in + (size_t)i * fstride, pitch0, imgs[i], stride_bytes, w, hgt,
                                           cudaMemcpyHostToDevice, s.st));
        }
        src.img0 = s.img_in;
        src.pitch0 = pitch0;
        src.stride0 = (long long)fstride;
        src._pad = 0;
    }
    if ((r = enqueue_kernels(h, s, src, nframes))) return r;
    s.launches += in_launches;
    // D2H: header, keypoints, descriptors (device rows are out_cap apart, caller rows cap apart)
    const int rows = std::min(cap, P.out_cap);
    CUDA_TRY(cudaMemcpyAsync(s.h_header, s.header, (size_t)nframes * 16, cudaMemcpyDeviceToHost, s.st));
    CUDA_TRY(cudaMemcpy2DAsync(kps, (size_t)cap * sizeof(orbx_keypoint), s.kps, (size_t)P.out_cap * sizeof(orbx_keypoint),
                               (size_t)rows * sizeof(orbx_keypoint), nframes, cudaMemcpyDeviceToHost, s.st));
    CUDA_TRY(cudaMemcpy2DAsync(desc, (size_t)cap * 32, s.desc, (size_t)P.out_cap * 32, (size_t)rows * 32, nframes,
                               cudaMemcpyDeviceToHost, s.st));
    s.in_flight = true;
    s.cap = cap;
    s.u_n = n_out;
    s.u_mono = mono_index;
    s.u_status = status;
    return ORBX_OK;
}

int orbx_wait(orbx_handle* h, int slot) {
    if (!h || slot < 0 || slot >= (int)h->slots.size()) return fail(ORBX_E_BAD_ARG, "bad slot");
    Slot& s = h->slots[slot];
    if (!s.in_flight) return ORBX_OK;
    s.in_flight = false;
    CUDA_TRY(cudaStreamSynchronize(s.st));
    int worst = ORBX_OK;
    if (s.u_n) {
        for (int i = 0; i < s.nframes; ++i) {
            const int n = s.h_header[4 * i], mono = s.h_header[4 * i + 1];
            int st = s.h_header[4 * i + 2];
            if (n > s.cap) st = ORBX_E_CAPACITY;
            s.u_n[i] = n;
            if (s.u_mono) s.u_mono[i] = mono;
            if (s.u_status) s.u_status[i] = st;
            if (st != ORBX_OK && worst == ORBX_OK)
                worst = fail(st, "frame %d: %d keypoints exceed the output capacity %d", i, n, s.cap);
        }
    }
    return worst;
}

int orbx_extract_batch(orbx_handle* h, const uint8_t* const* imgs, int nframes, int w, int hgt,
                       size_t stride_bytes, const int* lap, orbx_keypoint* kps, uint8_t* desc, int cap,
                       int* n_out, int* mono_index, int* status) {
    int r = orbx_submit_batch(h, 0, imgs, nframes, w, hgt, stride_bytes, lap, kps, desc, cap, n_out,
                              mono_index, status);
    if (r) return r;
    return orbx_wait(h, 0);
}

int orbx_extract(orbx_handle* h, const uint8_t* img, int w, int hgt, size_t stride_bytes, int lap_x0,
                 int lap_x1, orbx_keypoint* kps, uint8_t* desc, int cap, int* n_out, int* mono_index) {
    if (!img || w <= 0 || hgt <= 0) return fail(ORBX_E_EMPTY_IMAGE, "empty image");
    const int lap[2] = {lap_x0, lap_x1};
    const uint8_t* imgs[1] = {img};
    return orbx_extract_batch(h, imgs, 1, w, hgt, stride_bytes, lap, kps, desc, cap, n_out, mono_index, nullptr);
}

}  // extern "C"

// =================================================================================================
// pyramid access and stage taps (results of the last run on a slot)
// =================================================================================================
namespace {

inline int reflect101_host(int p, int len) {
    if (len == 1) return 0;
    while (p < 0 || p >= len) p = p < 0 ? -p : 2 * (len - 1) - p;
    return p;
}

int tap_ready(orbx_handle* h, int slot, int frame, int level) {
    if (!h || slot < 0 || slot >= (int)h->slots.size()) return fail(ORBX_E_BAD_ARG, "bad slot");
    Slot& s = h->slots[slot];
    if (s.in_flight) return fail(ORBX_E_BUSY, "slot %d not waited", slot);
    if (frame < 0 || frame >= s.nframes) return fail(ORBX_E_BAD_ARG, "frame %d not in the last batch", frame);
    if (level < 0 || level >= h->cfg.nlevels) return fail(ORBX_E_BAD_ARG, "bad level");
    if (s.run_w != h->plan.w || s.run_h != h->plan.h) return fail(ORBX_E_BAD_ARG, "shape changed since the last run");
    CUDA_TRY(cudaSetDevice(h->cfg.device));
    return ORBX_OK;
}

}  // namespace

extern "C" {

int orbx_get_pyramid_level(orbx_handle* h, int slot, int frame, int level, int pad, uint8_t* dst,
                           size_t dst_stride) {
    int r = tap_ready(h, slot, frame, level);
    if (r) return r;
    if (!dst || pad < 0) return fail(ORBX_E_BAD_ARG, "bad argument");
    Slot& s = h->slots[slot];
    const OrbxLevel& L = h->plan.lvl[level];
    const uint8_t* src;
    size_t spitch;
    if (level == 0) {
        src = s.src.img0 + (size_t)frame * s.src.stride0;
        spitch = (size_t)s.src.pitch0;
    } else {
        src = s.pyr + (size_t)frame * h->plan.frame_stride + L.off;
        spitch = (size_t)L.pitch;
    }
    if (dst_stride < (size_t)(L.w + 2 * pad)) return fail(ORBX_E_BAD_ARG, "dst_stride too small");
    CUDA_TRY(cudaMemcpy2D(dst + (size_t)pad * dst_stride + pad, dst_stride, src, spitch, L.w, L.h,
                          cudaMemcpyDeviceToHost));
    if (pad > 0) {
        // the reference's copyMakeBorder(..., BORDER_REFLECT_101) around the level
        // (ORBextractor.cc:1167-1173); a pure host-side re-indexing of the copied pixels
        for (int y = 0; y < L.h; ++y) {
            uint8_t* row = dst + (size_t)(y + pad) * dst_stride;
            for (int x = 0; x < pad; ++x) {
                row[x] = row[pad + reflect101_host(x - pad, L.w)];
                row[pad + L.w + x] = row[pad + reflect101_host(L.w + x, L.w)];
            }
        }
        for (int y = 0; y < pad; ++y) {
            memcpy(dst + (size_t)y * dst_stride, dst + (size_t)(pad + reflect101_host(y - pad, L.h)) * dst_stride, L.w + 2 * pad);
            memcpy(dst + (size_t)(pad + L.h + y) * dst_stride,
                   dst + (size_t)(pad + reflect101_host(L.h + y, L.h)) * dst_stride, L.w + 2 * pad);
        }
    }
    return ORBX_OK;
}

int orbx_device_pyramid(orbx_handle* h, int slot, int frame, orbx_pyramid_view* out) {
    int r = tap_ready(h, slot, frame, 0);
    if (r) return r;
    if (!out) return fail(ORBX_E_BAD_ARG, "null argument");
    Slot& s = h->slots[slot];
    memset(out, 0, sizeof *out);
    out->nlevels = h->cfg.nlevels;
    out->device = h->cfg.device;
    for (int l = 0; l < h->cfg.nlevels; ++l) {
        const OrbxLevel& L = h->plan.lvl[l];
        out->data[l] = l == 0 ? s.src.img0 + (size_t)frame * s.src.stride0
                              : s.pyr + (size_t)frame * h->plan.frame_stride + L.off;
        out->pitch[l] = l == 0 ? s.src.pitch0 : L.pitch;
        out->width[l] = L.w;
        out->height[l] = L.h;
    }
    return ORBX_OK;
}

int orbx_debug_sincos_check(int device, float lo, float hi, unsigned long long* n_bad, float* first_bad) {
    if (!n_bad || !first_bad || !(lo >= 0.f) || !(hi >= lo)) return fail(ORBX_E_BAD_ARG, "bad argument");
    CUDA_TRY(cudaSetDevice(device));
    CUDA_TRY(orbx_run_sincos_check(lo, hi, n_bad, first_bad));
    return ORBX_OK;
}

int orbx_debug_candidates(orbx_handle* h, int slot, int frame, int level, int32_t* xyr, int cap) {
    int r = tap_ready(h, slot, frame, level);
    if (r) return r;
    Slot& s = h->slots[slot];
    const OrbxPlan& P = h->plan;
    const OrbxLevel& L = P.lvl[level];
    const int ncells = L.n_cols * L.n_rows;
    std::vector<uint32_t> cnt(ncells), keys((size_t)L.key_cap);
    CUDA_TRY(cudaMemcpy(cnt.data(), s.cellcnt + (size_t)frame * P.cells_total + L.cell_off, (size_t)ncells * 4,
                        cudaMemcpyDeviceToHost));
    CUDA_TRY(cudaMemcpy(keys.data(), s.cellkeys + (size_t)frame * P.keys_total + L.key_off, (size_t)L.key_cap * 4,
                        cudaMemcpyDeviceToHost));
    int n = 0;
    for (int c = 0; c < ncells; ++c) {
        if ((int)cnt[c] > L.cell_cap) return fail(ORBX_E_CUDA, "cell %d count %u exceeds its slot", c, cnt[c]);
        for (uint32_t k = 0; k < cnt[c]; ++k, ++n) {
            if (xyr && n < cap) {
                const uint32_t key = keys[(size_t)c * L.cell_cap + k];
                xyr[3 * n] = ORBX_KEY_X(key);
                xyr[3 * n + 1] = ORBX_KEY_Y(key);
                xyr[3 * n + 2] = ORBX_KEY_R(key);
            }
        }
    }
    return n;
}

int orbx_debug_survivors(orbx_handle* h, int slot, int frame, int level, int32_t* xyr, int cap) {
    int r = tap_ready(h, slot, frame, level);
    if (r) return r;
    Slot& s = h->slots[slot];
    const OrbxPlan& P = h->plan;
    const OrbxLevel& L = P.lvl[level];
    uint32_t n = 0;
    CUDA_TRY(cudaMemcpy(&n, s.survcnt + (size_t)frame * ORBX_MAX_LEVELS + level, 4, cudaMemcpyDeviceToHost));
    if ((int)n > L.surv_cap) return fail(ORBX_E_CUDA, "level %d: %u nodes exceed the bound %d", level, n, L.surv_cap);
    std::vector<uint32_t> keys(n);
    if (n) CUDA_TRY(cudaMemcpy(keys.data(), s.surv + (size_t)frame * P.surv_total + L.surv_off, (size_t)n * 4,
                               cudaMemcpyDeviceToHost));
    for (uint32_t i = 0; i < n && xyr && (int)i < cap; ++i) {
        xyr[3 * i] = ORBX_KEY_X(keys[i]);
        xyr[3 * i + 1] = ORBX_KEY_Y(keys[i]);
        xyr[3 * i + 2] = ORBX_KEY_R(keys[i]);
    }
    return (int)n;
}

}  // extern "C"

// =================================================================================================
// fused multi-GPU result gather: result sinks
// =================================================================================================
struct orbx_sink {
    int device = 0;        // device the pointers are valid on (the creator's, or the opener's)
    int owner_device = -1; // device that holds the memory when it was created in this process
    int frames = 0, cap = 0;
    bool opened = false;   // mapped from another process through CUDA IPC
    uint8_t* base = nullptr;
    size_t off_sync = 0, off_kps = 0, off_desc = 0, bytes = 0;
    cudaStream_t st = nullptr;     // for the blocking forms of orbx_sink_wait / orbx_sink_release
    uint32_t* err = nullptr;       // device word: a wait timed out
    uint8_t* pin = nullptr;        // pinned staging of orbx_sink_read_frame (a pageable D2H copy serialises behind
                                   // running kernels, i.e. behind a producer parked in its back-pressure wait)
    uint32_t* pin_err = nullptr;   // pinned copy of `err`
};

namespace {
void sink_layout(orbx_sink* k) {
    // [frames][4] headers | [frames][2] {flag, ack} completion words | keypoints | descriptors
    k->off_sync = ((size_t)k->frames * 16 + 255) & ~(size_t)255;
    k->off_kps = (k->off_sync + (size_t)k->frames * 8 + 255) & ~(size_t)255;
    k->off_desc = (k->off_kps + (size_t)k->frames * k->cap * sizeof(orbx_keypoint) + 255) & ~(size_t)255;
    k->bytes = k->off_desc + (size_t)k->frames * k->cap * 32;
}
}  // namespace

extern "C" {

int orbx_sink_create(int device, int frames, int cap, orbx_sink** out) {
    if (!out || frames < 1 || cap < 1) return fail(ORBX_E_BAD_ARG, "bad argument");
    *out = nullptr;
    CUDA_TRY(cudaSetDevice(device));
    orbx_sink* k = new orbx_sink();
    k->device = k->owner_device = device;
    k->frames = frames;
    k->cap = cap;
    sink_layout(k);
    // plain cudaMalloc (not a pool / VMM allocation): the block can be exported through CUDA IPC
    cudaError_t e = cudaMalloc((void**)&k->base, k->bytes);
    if (e == cudaSuccess) e = cudaMemset(k->base, 0, k->bytes);
    if (e != cudaSuccess) {
        const size_t bytes = k->bytes;
        if (k->base) cudaFree(k->base);
        delete k;
        return fail(ORBX_E_CUDA, "sink allocation of %zu bytes failed: %s", bytes, cudaGetErrorString(e));
    }
    *out = k;
    return ORBX_OK;
}

int orbx_sink_export(orbx_sink* k, uint8_t handle[64]) {
    if (!k || !handle || k->opened) return fail(ORBX_E_BAD_ARG, "bad sink");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handle size");
    CUDA_TRY(cudaSetDevice(k->device));
    cudaIpcMemHandle_t hnd;
    CUDA_TRY(cudaIpcGetMemHandle(&hnd, k->base));
    memcpy(handle, &hnd, 64);
    return ORBX_OK;
}

int orbx_sink_open(int device, const uint8_t handle[64], int frames, int cap, orbx_sink** out) {
    if (!out || !handle || frames < 1 || cap < 1) return fail(ORBX_E_BAD_ARG, "bad argument");
    *out = nullptr;
    CUDA_TRY(cudaSetDevice(device));
    cudaIpcMemHandle_t hnd;
    memcpy(&hnd, handle, 64);
    void* p = nullptr;
    CUDA_TRY(cudaIpcOpenMemHandle(&p, hnd, cudaIpcMemLazyEnablePeerAccess));
    orbx_sink* k = new orbx_sink();
    k->device = device;
    k->frames = frames;
    k->cap = cap;
    k->opened = true;
    k->base = (uint8_t*)p;
    sink_layout(k);
    *out = k;
    return ORBX_OK;
}

int orbx_sink_destroy(orbx_sink* k) {
    if (!k) return ORBX_OK;
    cudaSetDevice(k->device);
    if (k->st) cudaStreamDestroy(k->st);
    if (k->err) cudaFree(k->err);
    if (k->pin) cudaFreeHost(k->pin);
    if (k->pin_err) cudaFreeHost(k->pin_err);
    if (k->base) {
        if (k->opened) cudaIpcCloseMemHandle(k->base);
        else cudaFree(k->base);
    }
    delete k;
    return ORBX_OK;
}

static int sink_sync_ptr(orbx_sink* k, int first_frame, uint32_t** sync);

int orbx_sink_pointers(orbx_sink* k, void** header, void** kps, void** desc) {
    if (!k) return fail(ORBX_E_BAD_ARG, "null sink");
    if (header) *header = k->base;
    if (kps) *kps = k->base + k->off_kps;
    if (desc) *desc = k->base + k->off_desc;
    return ORBX_OK;
}

int orbx_sink_read_frame(orbx_sink* k, int frame, int32_t header[4], orbx_keypoint* kps, uint8_t* desc, int rows) {
    if (!k || frame < 0 || frame >= k->frames || rows < 0 || rows > k->cap) return fail(ORBX_E_BAD_ARG, "bad argument");
    // on the sink's own (non-blocking) stream, NOT a device-wide synchronisation: a producer may legitimately be
    // parked in its back-pressure wait.  Completeness of the frame is the caller's business (orbx_sink_wait, or the
    // producers' streams / a barrier as in round 1).
    uint32_t* unused = nullptr;
    int r = sink_sync_ptr(k, 0, &unused);
    if (r) return r;
    uint8_t* ph = k->pin;
    uint8_t* pk = k->pin + 16;
    uint8_t* pd = pk + (size_t)k->cap * sizeof(orbx_keypoint);
    if (header) CUDA_TRY(cudaMemcpyAsync(ph, k->base + (size_t)frame * 16, 16, cudaMemcpyDeviceToHost, k->st));
    if (kps && rows)
        CUDA_TRY(cudaMemcpyAsync(pk, k->base + k->off_kps + (size_t)frame * k->cap * sizeof(orbx_keypoint),
                                 (size_t)rows * sizeof(orbx_keypoint), cudaMemcpyDeviceToHost, k->st));
    if (desc && rows)
        CUDA_TRY(cudaMemcpyAsync(pd, k->base + k->off_desc + (size_t)frame * k->cap * 32, (size_t)rows * 32,
                                 cudaMemcpyDeviceToHost, k->st));
    CUDA_TRY(cudaStreamSynchronize(k->st));
    if (header) memcpy(header, ph, 16);
    if (kps && rows) memcpy(kps, pk, (size_t)rows * sizeof(orbx_keypoint));
    if (desc && rows) memcpy(desc, pd, (size_t)rows * 32);
    return ORBX_OK;
}

int orbx_set_result_sink(orbx_handle* h, int slot, orbx_sink* k, int first_frame) {
    return orbx_set_result_sink_ex(h, slot, k, first_frame, 0);
}

int orbx_set_result_sink_ex(orbx_handle* h, int slot, orbx_sink* k, int first_frame, int flags) {
    if (!h || slot < 0 || slot >= (int)h->slots.size()) return fail(ORBX_E_BAD_ARG, "bad slot");
    Slot& s = h->slots[slot];
    if (s.in_flight) return fail(ORBX_E_BUSY, "slot %d has an un-waited submit", slot);
    CUDA_TRY(cudaSetDevice(h->cfg.device));
    CUDA_TRY(cudaStreamSynchronize(s.st));
    if (s.gexec) {   // the captured graph holds the old sink arguments
        cudaGraphExecDestroy(s.gexec);
        s.gexec = nullptr;
    }
    if (!k) {
        s.sink = OrbxSinkArgs{};
        return ORBX_OK;
    }
    const int cap = orbx_max_keypoints(h);
    if (k->cap != cap) return fail(ORBX_E_BAD_ARG, "sink rows per frame (%d) differ from the handle's capacity (%d)", k->cap, cap);
    if (first_frame < 0 || first_frame + h->cfg.max_batch > k->frames)
        return fail(ORBX_E_BAD_ARG, "frames [%d, %d) do not fit the sink's %d frames", first_frame,
                    first_frame + h->cfg.max_batch, k->frames);
    if (!k->opened && k->owner_device != h->cfg.device) {
        // same process, other GPU: map the owner's memory into this device (NVLink peer access)
        int can = 0;
        CUDA_TRY(cudaDeviceCanAccessPeer(&can, h->cfg.device, k->owner_device));
        if (!can) return fail(ORBX_E_UNSUPPORTED, "device %d cannot access device %d", h->cfg.device, k->owner_device);
        cudaError_t e = cudaDeviceEnablePeerAccess(k->owner_device, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled)
            return fail(ORBX_E_CUDA, "cudaDeviceEnablePeerAccess failed: %s", cudaGetErrorString(e));
        cudaGetLastError();
    } else if (k->opened && k->device != h->cfg.device) {
        return fail(ORBX_E_BAD_ARG, "the sink was opened on device %d, the handle runs on %d", k->device, h->cfg.device);
    }
    CUDA_TRY(orbx_sink_preload());
    if (!s.sink_words) CUDA_TRY(cudaMalloc((void**)&s.sink_words, 8));
    CUDA_TRY(cudaMemset(s.sink_words, 0, 8));
    s.sink_epoch = 0;
    s.sink.hdr = (int32_t*)k->base;
    s.sink.kps = (orbx_keypoint*)(k->base + k->off_kps);
    s.sink.desc = k->base + k->off_desc;
    s.sink.first_frame = first_frame;
    // the region's completion words live in the sink: word 2 * first_frame = flag, + 1 = ack.  A region that is
    // (re)attached starts a new count: the producer clears both (the consumer counts from 1 again).
    uint32_t* sync = (uint32_t*)(k->base + k->off_sync) + 2 * (size_t)first_frame;
    CUDA_TRY(cudaMemset(sync, 0, 8));
    s.sink.flag = sync;
    s.sink.ack = (flags & ORBX_SINK_BACKPRESSURE) ? sync + 1 : nullptr;
    s.sink.epoch_dev = s.sink_words;
    s.sink.err_dev = s.sink_words + 1;
    return ORBX_OK;
}

int orbx_sink_epoch(orbx_handle* h, int slot, uint32_t* epoch, int* timed_out) {
    if (!h || slot < 0 || slot >= (int)h->slots.size() || !epoch) return fail(ORBX_E_BAD_ARG, "bad argument");
    Slot& s = h->slots[slot];
    *epoch = s.sink_epoch;
    if (timed_out) {
        *timed_out = 0;
        if (s.sink_words && !s.in_flight) {
            uint32_t w[2] = {0, 0};
            CUDA_TRY(cudaSetDevice(h->cfg.device));
            CUDA_TRY(cudaMemcpy(w, s.sink_words, 8, cudaMemcpyDeviceToHost));
            *timed_out = (int)w[1];
        }
    }
    return ORBX_OK;
}

static int sink_sync_ptr(orbx_sink* k, int first_frame, uint32_t** sync) {
    if (!k || first_frame < 0 || first_frame >= k->frames) return fail(ORBX_E_BAD_ARG, "bad sink region");
    CUDA_TRY(cudaSetDevice(k->device));
    if (!k->st) CUDA_TRY(cudaStreamCreateWithFlags(&k->st, cudaStreamNonBlocking));
    if (!k->err) {
        CUDA_TRY(cudaMalloc((void**)&k->err, 4));
        CUDA_TRY(cudaMemsetAsync(k->err, 0, 4, k->st));
        CUDA_TRY(cudaHostAlloc((void**)&k->pin_err, 4, cudaHostAllocDefault));
        CUDA_TRY(cudaHostAlloc((void**)&k->pin, 16 + (size_t)k->cap * (sizeof(orbx_keypoint) + 32), cudaHostAllocDefault));
        CUDA_TRY(cudaStreamSynchronize(k->st));
    }
    CUDA_TRY(orbx_sink_preload());
    *sync = (uint32_t*)(k->base + k->off_sync) + 2 * (size_t)first_frame;
    return ORBX_OK;
}

int orbx_sink_wait(orbx_sink* k, int first_frame, uint32_t epoch, void* stream, int timeout_ms) {
    uint32_t* sync = nullptr;
    int r = sink_sync_ptr(k, first_frame, &sync);
    if (r) return r;
    const unsigned long long ns = (unsigned long long)(timeout_ms > 0 ? timeout_ms : 2000) * 1000000ull;
    cudaStream_t st = stream ? (cudaStream_t)stream : k->st;
    CUDA_TRY(orbx_launch_sink_wait_flag(sync, epoch, ns, k->err, st));
    if (stream) return ORBX_OK;   // ordered on the caller's stream: work enqueued after it sees the batch
    CUDA_TRY(cudaMemcpyAsync(k->pin_err, k->err, 4, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    if (*k->pin_err) {
        CUDA_TRY(cudaMemsetAsync(k->err, 0, 4, st));
        CUDA_TRY(cudaStreamSynchronize(st));
        return fail(ORBX_E_BUSY, "sink region %d: batch %u did not arrive within %d ms", first_frame, epoch, timeout_ms);
    }
    return ORBX_OK;
}

int orbx_sink_release(orbx_sink* k, int first_frame, uint32_t epoch, void* stream) {
    uint32_t* sync = nullptr;
    int r = sink_sync_ptr(k, first_frame, &sync);
    if (r) return r;
    cudaStream_t st = stream ? (cudaStream_t)stream : k->st;
    CUDA_TRY(orbx_launch_sink_release(sync + 1, epoch, st));
    if (!stream) CUDA_TRY(cudaStreamSynchronize(st));
    return ORBX_OK;
}

}  // extern "C"
