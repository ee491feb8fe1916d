// oracle/orb_oracle.cpp -- TEST INFRASTRUCTURE ONLY (CPU oracle, "port").
//
// Our own CPU restatement of the reference's ORB extraction path.  It exists so that parity
// tests have an oracle that builds from this repository alone (the verbatim reference build in
// oracle/_ref needs /root/reference).  PARITY PINNING: tests/test_oracle_*.py require this
// file to agree bit-for-bit with oracle/_ref/liborbref.so (the reference's own ORBextractor.cc)
// on every synthetic fixture, and cvprims.h to agree with cv2 4.13.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load this library.
//
// Reference functions followed (file:line relative to /root/reference):
//   tables            ORBextractor::ORBextractor          src/ORBextractor.cc:408-468
//   pyramid           ORBextractor::ComputePyramid        src/ORBextractor.cc:1152-1177
//   fast_cells        ORBextractor::ComputeKeyPointsOctTree src/ORBextractor.cc:763-854
//   distribute        ORBextractor::DistributeOctTree     src/ORBextractor.cc:537-761
//                     ExtractorNode::DivideNode           src/ORBextractor.cc:479-535
//   ic_angle          IC_Angle                            src/ORBextractor.cc:75-102
//   descriptor        computeOrbDescriptor                src/ORBextractor.cc:106-145
//   extract           ORBextractor::operator()            src/ORBextractor.cc:1068-1150
//
// The quadtree is written as array passes (not std::list) with the tie-break of the sort at
// src/ORBextractor.cc:682 made explicit: equal key counts are ordered by node CREATION ORDER
// (what the pointer comparison yields under a monotonic allocator).  This is
// also the formulation the CUDA kernel implements.
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>
#include <algorithm>

#include "cvprims.h"

namespace {

const int8_t kPattern[1024] = {
#include "../orb-slam3-mac_b200/csrc/orb_pattern.inc"
};

const int PATCH_SIZE = 31, HALF_PATCH_SIZE = 15, EDGE_THRESHOLD = 19;

struct Kp {  // same layout as cv::KeyPoint (28 bytes)
    float x, y, size, angle, response;
    int32_t octave, class_id;
};

struct Cand {
    int x, y, r;  // relative to (16,16); r = FAST response
};

struct Level {
    int w, h;
    std::vector<uint8_t> px;  // w*h, unpadded
};

struct Node {
    int x0, x1, y0, y1;       // UL.x, UR.x, UL.y, BL.y
    std::vector<int> keys;    // indices into the candidate array, reference order preserved
};

struct Oracle {
    int nfeatures, nlevels, iniTh, minTh;
    double scaleFactor;
    std::vector<float> sf, isf, s2, is2;
    std::vector<int> nper, umax;
    std::vector<Level> pyr;
    std::vector<std::vector<Cand> > cand;
    std::vector<std::vector<Kp> > surv;

    Oracle(int nf, float s, int nl, int a, int b)
        : nfeatures(nf), nlevels(nl), iniTh(a), minTh(b), scaleFactor(s) {
        // src/ORBextractor.cc:413-429 (scaleFactor is a double member, ORBextractor.h:96)
        sf.resize(nl); isf.resize(nl); s2.resize(nl); is2.resize(nl);
        sf[0] = 1.0f; s2[0] = 1.0f;
        for (int i = 1; i < nl; ++i) {
            sf[i] = (float)(sf[i - 1] * scaleFactor);
            s2[i] = sf[i] * sf[i];
        }
        for (int i = 0; i < nl; ++i) { isf[i] = 1.0f / sf[i]; is2[i] = 1.0f / s2[i]; }
        // :431-444
        nper.resize(nl);
        float factor = (float)(1.0f / scaleFactor);
        float nDesired = (float)(nfeatures * (1 - factor) /
                                 (1 - (float)pow((double)factor, (double)nlevels)));
        int sum = 0;
        for (int l = 0; l < nl - 1; ++l) {
            nper[l] = cvprims::cv_round(nDesired);
            sum += nper[l];
            nDesired *= factor;
        }
        nper[nl - 1] = std::max(nfeatures - sum, 0);
        // :452-467
        umax.assign(HALF_PATCH_SIZE + 1, 0);
        int v, v0, vmax = cvprims::cv_floor(HALF_PATCH_SIZE * sqrt(2.f) / 2 + 1);
        int vmin = cvprims::cv_ceil(HALF_PATCH_SIZE * sqrt(2.f) / 2);
        const double hp2 = HALF_PATCH_SIZE * HALF_PATCH_SIZE;
        for (v = 0; v <= vmax; ++v) umax[v] = cvprims::cv_round(sqrt(hp2 - v * v));
        for (v = HALF_PATCH_SIZE, v0 = 0; v >= vmin; --v) {
            while (umax[v0] == umax[v0 + 1]) ++v0;
            umax[v] = v0;
            ++v0;
        }
        pyr.resize(nl);
        cand.resize(nl);
        surv.resize(nl);
    }

    // ---- E1 ------------------------------------------------------------------------------
    void pyramid(const uint8_t* img, int w, int h, size_t stride) {
        for (int l = 0; l < nlevels; ++l) {
            float scale = isf[l];
            Level& L = pyr[l];
            L.w = cvprims::cv_round((float)w * scale);
            L.h = cvprims::cv_round((float)h * scale);
            L.px.resize((size_t)L.w * L.h);
            if (l == 0) {
                for (int y = 0; y < h; ++y) memcpy(&L.px[(size_t)y * w], img + (size_t)y * stride, w);
            } else {
                const Level& P = pyr[l - 1];
                cvprims::resize_linear_u8(P.px.data(), P.h, P.w, P.w, L.px.data(), L.h, L.w, L.w);
            }
        }
    }

    // ---- E2 ------------------------------------------------------------------------------
    void fast_cells(int l) {
        const Level& L = pyr[l];
        std::vector<Cand>& out = cand[l];
        out.clear();
        const float W = 30;
        const int minBX = EDGE_THRESHOLD - 3, minBY = minBX;
        const int maxBX = L.w - EDGE_THRESHOLD + 3, maxBY = L.h - EDGE_THRESHOLD + 3;
        const float width = (float)(maxBX - minBX), height = (float)(maxBY - minBY);
        const int nCols = (int)(width / W), nRows = (int)(height / W);
        if (nCols <= 0 || nRows <= 0) return;  // reference would divide by zero
        const int wCell = (int)ceil(width / nCols), hCell = (int)ceil(height / nRows);
        std::vector<cvprims::FastKp> kps;
        for (int i = 0; i < nRows; ++i) {
            const float iniY = (float)(minBY + i * hCell);
            float maxY = iniY + hCell + 6;
            if (iniY >= maxBY - 3) continue;
            if (maxY > maxBY) maxY = (float)maxBY;
            for (int j = 0; j < nCols; ++j) {
                const float iniX = (float)(minBX + j * wCell);
                float maxX = iniX + wCell + 6;
                if (iniX >= maxBX - 6) continue;
                if (maxX > maxBX) maxX = (float)maxBX;
                const uint8_t* sub = &L.px[(size_t)(int)iniY * L.w + (int)iniX];
                const int sr = (int)maxY - (int)iniY, sc = (int)maxX - (int)iniX;
                cvprims::fast9_nms(sub, sr, sc, L.w, iniTh, kps);
                if (kps.empty()) cvprims::fast9_nms(sub, sr, sc, L.w, minTh, kps);
                for (const auto& k : kps) out.push_back(Cand{k.x + j * wCell, k.y + i * hCell, k.score});
            }
        }
    }

    // ---- E3 ------------------------------------------------------------------------------
    // Returns indices into `c` in the reference's output order.
    static void divide(const Node& n, const std::vector<Cand>& c, Node ch[4]) {
        const int halfX = (int)ceil((float)(n.x1 - n.x0) / 2);
        const int halfY = (int)ceil((float)(n.y1 - n.y0) / 2);
        const int mx = n.x0 + halfX, my = n.y0 + halfY;
        ch[0] = Node{n.x0, mx, n.y0, my, {}};
        ch[1] = Node{mx, n.x1, n.y0, my, {}};
        ch[2] = Node{n.x0, mx, my, n.y1, {}};
        ch[3] = Node{mx, n.x1, my, n.y1, {}};
        for (int k : n.keys) {
            const Cand& kp = c[k];
            int q = ((float)kp.x < (float)mx ? 0 : 1) + ((float)kp.y < (float)my ? 0 : 2);
            ch[q].keys.push_back(k);
        }
    }

    static int distribute(const std::vector<Cand>& c, int minX, int maxX, int minY, int maxY, int N,
                          std::vector<int>& result) {
        result.clear();
        const int nIni = (int)round((float)(maxX - minX) / (maxY - minY));
        if (nIni <= 0) return -1;  // reference divides by zero here (tall portrait levels)
        const float hX = (float)(maxX - minX) / nIni;
        // `nodes` is the std::list in list order (front first).
        std::vector<Node> nodes(nIni);
        for (int i = 0; i < nIni; ++i)
            nodes[i] = Node{(int)(hX * (float)i), (int)(hX * (float)(i + 1)), 0, maxY - minY, {}};
        for (int k = 0; k < (int)c.size(); ++k) {
            int idx = (int)((float)c[k].x / hX);
            nodes[idx].keys.push_back(k);
        }
        {
            std::vector<Node> keep;
            for (auto& n : nodes)
                if (!n.keys.empty()) keep.push_back(std::move(n));
            nodes.swap(keep);
        }
        bool finish = false;
        // (count, creation index within the pass) of the multi-key children made by the last pass,
        // plus where each of them currently sits in `nodes`.
        while (!finish) {
            const int prevSize = (int)nodes.size();
            // ---- full pass (:597-661): split every multi-key node in list order; children are
            // push_front'ed, so the new list is reverse(creation order) ++ untouched nodes.
            std::vector<Node> created, kept;
            for (auto& n : nodes) {
                if (n.keys.size() == 1) { kept.push_back(std::move(n)); continue; }
                Node ch[4];
                divide(n, c, ch);
                for (int q = 0; q < 4; ++q)
                    if (!ch[q].keys.empty()) created.push_back(std::move(ch[q]));
            }
            int nToExpand = 0;
            for (auto& n : created) nToExpand += n.keys.size() > 1;
            // creation order of multi-key children = vSizeAndPointerToNode order
            nodes.clear();
            for (int i = (int)created.size() - 1; i >= 0; --i) nodes.push_back(std::move(created[i]));
            const int nCreated = (int)nodes.size();
            for (auto& n : kept) nodes.push_back(std::move(n));
            // `expand` = positions in `nodes` of the multi-key children, in CREATION order
            std::vector<int> expand;
            for (int i = nCreated - 1; i >= 0; --i)
                if (nodes[i].keys.size() > 1) expand.push_back(i);

            if ((int)nodes.size() >= N || (int)nodes.size() == prevSize) {
                finish = true;
            } else if ((int)nodes.size() + nToExpand * 3 > N) {
                // ---- careful phase (:671-735)
                while (!finish) {
                    const int prev2 = (int)nodes.size();
                    // sort ascending by (count, creation order), walk from the back (:682-683)
                    std::vector<int> order(expand.size());
                    for (size_t i = 0; i < order.size(); ++i) order[i] = (int)i;
                    std::stable_sort(order.begin(), order.end(), [&](int a, int b) {
                        return nodes[expand[a]].keys.size() < nodes[expand[b]].keys.size();
                    });
                    std::vector<Node> made;           // creation order in this pass
                    std::vector<char> erased(nodes.size(), 0);
                    int size = (int)nodes.size();
                    for (int j = (int)order.size() - 1; j >= 0; --j) {
                        const int pos = expand[order[j]];
                        Node ch[4];
                        divide(nodes[pos], c, ch);
                        for (int q = 0; q < 4; ++q)
                            if (!ch[q].keys.empty()) { made.push_back(std::move(ch[q])); ++size; }
                        erased[pos] = 1;
                        --size;
                        if (size >= N) break;
                    }
                    std::vector<Node> next;
                    for (int i = (int)made.size() - 1; i >= 0; --i) next.push_back(std::move(made[i]));
                    const int nMade = (int)next.size();
                    for (size_t i = 0; i < nodes.size(); ++i)
                        if (!erased[i]) next.push_back(std::move(nodes[i]));
                    nodes.swap(next);
                    expand.clear();
                    for (int i = nMade - 1; i >= 0; --i)
                        if (nodes[i].keys.size() > 1) expand.push_back(i);
                    if ((int)nodes.size() >= N || (int)nodes.size() == prev2) finish = true;
                }
            }
        }
        // :738-758 best response per node, first maximum wins
        for (const auto& n : nodes) {
            int best = n.keys[0];
            for (size_t k = 1; k < n.keys.size(); ++k)
                if ((float)c[n.keys[k]].r > (float)c[best].r) best = n.keys[k];
            result.push_back(best);
        }
        return 0;
    }

    // ---- E4 ------------------------------------------------------------------------------
    float ic_angle(const Level& L, int cx, int cy) const {
        int m01 = 0, m10 = 0;
        const uint8_t* center = &L.px[(size_t)cy * L.w + cx];
        for (int u = -HALF_PATCH_SIZE; u <= HALF_PATCH_SIZE; ++u) m10 += u * center[u];
        const int step = L.w;
        for (int v = 1; v <= HALF_PATCH_SIZE; ++v) {
            int v_sum = 0, d = umax[v];
            for (int u = -d; u <= d; ++u) {
                int p = center[u + v * step], m = center[u - v * step];
                v_sum += (p - m);
                m10 += u * (p + m);
            }
            m01 += v * v_sum;
        }
        return cvprims::fast_atan2_deg((float)m01, (float)m10);
    }

    int keypoints_octree() {
        for (int l = 0; l < nlevels; ++l) {
            const Level& L = pyr[l];
            fast_cells(l);
            const int minB = EDGE_THRESHOLD - 3;
            const int maxBX = L.w - EDGE_THRESHOLD + 3, maxBY = L.h - EDGE_THRESHOLD + 3;
            std::vector<int> res;
            surv[l].clear();
            if (distribute(cand[l], minB, maxBX, minB, maxBY, nper[l], res) != 0) return -1;
            const int scaledPatch = (int)(PATCH_SIZE * sf[l]);
            for (int k : res) {
                const Cand& cd = cand[l][k];
                Kp kp;
                kp.x = (float)(cd.x + minB);
                kp.y = (float)(cd.y + minB);
                kp.size = (float)scaledPatch;
                kp.angle = -1;
                kp.response = (float)cd.r;
                kp.octave = l;
                kp.class_id = -1;
                surv[l].push_back(kp);
            }
        }
        for (int l = 0; l < nlevels; ++l)
            for (auto& kp : surv[l])
                kp.angle = ic_angle(pyr[l], cvprims::cv_round(kp.x), cvprims::cv_round(kp.y));
        return 0;
    }

    // ---- E6 ------------------------------------------------------------------------------
    static void descriptor(const Kp& kpt, const uint8_t* img, int step, uint8_t* desc) {
        const float factorPI = (float)(3.1415926535897932384626433832795 / 180.f);
        float angle = (float)kpt.angle * factorPI;
        float a = cosf(angle), b = sinf(angle);
        const uint8_t* center = img + (size_t)cvprims::cv_round(kpt.y) * step + cvprims::cv_round(kpt.x);
        const int8_t* p = kPattern;
        auto val = [&](int idx) -> int {
            float px = (float)p[2 * idx], py = (float)p[2 * idx + 1];
            int dy = cvprims::cv_round(px * b + py * a);
            int dx = cvprims::cv_round(px * a - py * b);
            return center[dy * step + dx];
        };
        for (int i = 0; i < 32; ++i, p += 32) {
            int v = 0;
            for (int k = 0; k < 8; ++k) v |= (val(2 * k) < val(2 * k + 1)) << k;
            desc[i] = (uint8_t)v;
        }
    }

    // ---- E7 ------------------------------------------------------------------------------
    int extract(const uint8_t* img, int w, int h, size_t stride, int lap0, int lap1,
                std::vector<Kp>& keys, std::vector<uint8_t>& descs) {
        keys.clear();
        descs.clear();
        if (!img || w <= 0 || h <= 0) return -1;
        pyramid(img, w, h, stride);
        if (keypoints_octree() != 0) return -3;
        int n = 0;
        for (int l = 0; l < nlevels; ++l) n += (int)surv[l].size();
        keys.resize(n);
        descs.resize((size_t)n * 32);
        int mono = 0, stereo = n - 1;
        std::vector<uint8_t> blur;
        for (int l = 0; l < nlevels; ++l) {
            if (surv[l].empty()) continue;
            const Level& L = pyr[l];
            blur.resize(L.px.size());
            cvprims::gaussian7_sigma2_u8(L.px.data(), L.h, L.w, L.w, blur.data(), L.w);
            const float scale = sf[l];
            for (Kp kp : surv[l]) {
                uint8_t d[32];
                descriptor(kp, blur.data(), L.w, d);
                if (l != 0) { kp.x *= scale; kp.y *= scale; }
                int at;
                if (kp.x >= (float)lap0 && kp.x <= (float)lap1) at = stereo--;
                else at = mono++;
                keys[at] = kp;
                memcpy(&descs[(size_t)at * 32], d, 32);
            }
        }
        return mono;
    }
};

}  // namespace

// ---------------------------------------------------------------------------------------
// C surface (same shape as oracle/ref_capi.cpp, prefix orbo_), plus raw primitive taps used
// to pin cvprims.h against cv2.
// ---------------------------------------------------------------------------------------
extern "C" {

void* orbo_create(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST) {
    return new Oracle(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST);
}
void orbo_destroy(void* h) { delete (Oracle*)h; }
void orbo_set_malloc_mode(int) {}  // the port has no allocator dependence

int orbo_tables(void* h, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2,
                int* n_per_level) {
    Oracle* o = (Oracle*)h;
    for (int i = 0; i < o->nlevels; ++i) {
        scale[i] = o->sf[i];
        inv_scale[i] = o->isf[i];
        sigma2[i] = o->s2[i];
        inv_sigma2[i] = o->is2[i];
        n_per_level[i] = o->nper[i];
    }
    return o->nlevels;
}
int orbo_umax(void* h, int* out16) {
    Oracle* o = (Oracle*)h;
    for (int i = 0; i < 16; ++i) out16[i] = o->umax[i];
    return 16;
}
int orbo_pattern(void*, int* out1024) {
    for (int i = 0; i < 1024; ++i) out1024[i] = kPattern[i];
    return 512;
}

int orbo_extract(void* h, const uint8_t* img, int w, int h_, size_t stride, int lap0, int lap1,
                 void* kps, uint8_t* desc, int cap, int* n_out) {
    Oracle* o = (Oracle*)h;
    std::vector<Kp> keys;
    std::vector<uint8_t> d;
    int mono = o->extract(img, w, h_, stride, lap0, lap1, keys, d);
    int n = (int)keys.size();
    if (n > cap) { *n_out = -2; return mono; }
    *n_out = n;
    if (n) {
        memcpy(kps, keys.data(), (size_t)n * sizeof(Kp));
        memcpy(desc, d.data(), (size_t)n * 32);
    }
    return mono;
}

int orbo_level_size(void* h, int level, int* w, int* h_) {
    Oracle* o = (Oracle*)h;
    *w = o->pyr[level].w;
    *h_ = o->pyr[level].h;
    return 0;
}

int orbo_get_level(void* h, int level, int pad, uint8_t* dst, size_t dst_stride) {
    Oracle* o = (Oracle*)h;
    const Level& L = o->pyr[level];
    if (pad == 0) {
        for (int y = 0; y < L.h; ++y) memcpy(dst + (size_t)y * dst_stride, &L.px[(size_t)y * L.w], L.w);
    } else {
        cvprims::copy_make_border_reflect101(L.px.data(), L.h, L.w, L.w, dst, dst_stride, pad, pad,
                                             pad, pad);
    }
    return 0;
}

int orbo_stages(void* h, const uint8_t* img, int w, int h_, size_t stride) {
    Oracle* o = (Oracle*)h;
    o->pyramid(img, w, h_, stride);
    return o->keypoints_octree();
}
int orbo_stage_candidates(void* h, int level, int* xyr, int cap) {
    Oracle* o = (Oracle*)h;
    int n = (int)o->cand[level].size();
    if (xyr && n <= cap)
        for (int i = 0; i < n; ++i) {
            xyr[3 * i] = o->cand[level][i].x;
            xyr[3 * i + 1] = o->cand[level][i].y;
            xyr[3 * i + 2] = o->cand[level][i].r;
        }
    return n;
}
int orbo_stage_survivors(void* h, int level, void* kps, int cap) {
    Oracle* o = (Oracle*)h;
    int n = (int)o->surv[level].size();
    if (kps && n <= cap && n) memcpy(kps, o->surv[level].data(), (size_t)n * sizeof(Kp));
    return n;
}

// Standalone quadtree on an arbitrary candidate list (x, y, response triplets, coordinates
// relative to the detection window of size W x H).  out_idx receives indices into the list.
int orbo_distribute(const int* xyr, int n, int W, int H, int N, int* out_idx, int cap) {
    std::vector<Cand> c(n);
    for (int i = 0; i < n; ++i) c[i] = Cand{xyr[3 * i], xyr[3 * i + 1], xyr[3 * i + 2]};
    std::vector<int> res;
    if (Oracle::distribute(c, 0, W, 0, H, N, res) != 0) return -1;
    if ((int)res.size() > cap) return -2;
    for (size_t i = 0; i < res.size(); ++i) out_idx[i] = res[i];
    return (int)res.size();
}

// ---- primitive taps (tests/test_oracle_prims.py compares these with cv2 4.13) ----------------
void orbo_prim_resize(const uint8_t* src, int sh, int sw, size_t sstep, uint8_t* dst, int dh,
                      int dw, size_t dstep) {
    cvprims::resize_linear_u8(src, sh, sw, sstep, dst, dh, dw, dstep);
}
void orbo_prim_blur(const uint8_t* src, int h, int w, size_t sstep, uint8_t* dst, size_t dstep) {
    cvprims::gaussian7_sigma2_u8(src, h, w, sstep, dst, dstep);
}
void orbo_prim_border(const uint8_t* src, int h, int w, size_t sstep, uint8_t* dst, size_t dstep,
                      int pad) {
    cvprims::copy_make_border_reflect101(src, h, w, sstep, dst, dstep, pad, pad, pad, pad);
}
int orbo_prim_fast(const uint8_t* img, int h, int w, size_t step, int th, int* xyr, int cap) {
    std::vector<cvprims::FastKp> v;
    cvprims::fast9_nms(img, h, w, step, th, v);
    if ((int)v.size() > cap) return -2;
    for (size_t i = 0; i < v.size(); ++i) {
        xyr[3 * i] = v[i].x;
        xyr[3 * i + 1] = v[i].y;
        xyr[3 * i + 2] = v[i].score;
    }
    return (int)v.size();
}
void orbo_prim_atan2(const float* y, const float* x, float* out, int n) {
    for (int i = 0; i < n; ++i) out[i] = cvprims::fast_atan2_deg(y[i], x[i]);
}
int orbo_prim_round(float v) { return cvprims::cv_round(v); }
void orbo_prim_cvt_gray(const uint8_t* src, int h, int w, size_t sstep, int channels, int rgb_order, int bits,
                        uint8_t* dst, size_t dstep) {
    cvprims::cvt_gray_u8(src, h, w, sstep, channels, rgb_order, bits, dst, dstep);
}
void orbo_prim_remap(const uint8_t* src, int sh, int sw, size_t sstep, int channels, const float* mapx,
                     const float* mapy, uint8_t* dst, int dh, int dw, size_t dstep) {
    cvprims::remap_linear_u8(src, sh, sw, sstep, channels, mapx, mapy, (size_t)dw, dst, dh, dw, dstep);
}

}  // extern "C"
