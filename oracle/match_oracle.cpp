// oracle/match_oracle.cpp -- TEST INFRASTRUCTURE ONLY (CPU oracle for the ORBmatcher path).
//
// PINNED TO THE REFERENCE: every function here is checked, bit for bit, against the reference's own
// src/ORBmatcher.cc / src/Frame.cc / src/MapPoint.cc / Thirdparty/DBoW2 compiled VERBATIM into
// oracle/_ref/liborbmref.so (oracle/ref_match_capi.cpp drives it with the same POD snapshots) by
// tests/test_oracle_match.py, tests/test_oracle_match_ref.py and the goldens in tests/golden/ (generated
// from the verbatim build).  This restatement stays because it builds from this repo alone, is the timed
// CPU arm of bench_match.py, and holds the one piece the reference has no separate function for
// (orbo_stereo_rows, the Hamming half of Frame::ComputeStereoMatches; the whole function is pinned).
// Additionally: the distance against numpy popcount, the 2-NN search against cv2.BFMatcher(NORM_HAMMING)
// .knnMatch (the un-vendored OpenCV call at src/Frame.cc:1146), the candidate order against a brute-force scan.
//
// Every distance here is computed inline with the reference's SWAR popcount, sequentially, per
// candidate -- the opposite decomposition of the product (which batches all distances on the GPU
// first and replays), so agreement is meaningful.
//
// Reference functions followed (file:line relative to /root/reference):
//   descriptor_distance   ORBmatcher::DescriptorDistance          src/ORBmatcher.cc:2353-2369
//   three_maxima          ORBmatcher::ComputeThreeMaxima          src/ORBmatcher.cc:2307-2348
//   Frame grid            AssignFeaturesToGrid/PosInGrid/GetFeaturesInArea  src/Frame.cc:377-408,716-726,645-714
//   sbp_points            SearchByProjection(Frame&, vector<MapPoint*>&,..) src/ORBmatcher.cc:48-147
//   sbp_frames            SearchByProjection(Frame&, const Frame&,..)      src/ORBmatcher.cc:1965-2076,2158-2181
//   search_init           SearchForInitialization                 src/ORBmatcher.cc:710-825
//   stereo_rows           Frame::ComputeStereoMatches (Hamming part)      src/Frame.cc:802-885
//   knn2                  BFMatcher(NORM_HAMMING).knnMatch(k=2)   src/Frame.cc:1146 (OpenCV semantics)
//   N3 (LocalMapping / LoopClosing / reloc): SearchByBoW(KF,KF) :827-967, SearchByProjection(pKF,Scw,..) :477-708,
//   reloc SearchByProjection :2183-2305, Fuse :1403-1737, SearchBySim3 :1739-1963, SearchForTriangulation :969-1401,
//   MapPoint::ComputeDistinctiveDescriptors src/MapPoint.cc:330-404
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <map>
#include <vector>
#include <algorithm>

namespace {

const int TH_HIGH = 100, TH_LOW = 50, HISTO_LENGTH = 30;
const int GRID_COLS = 64, GRID_ROWS = 48;

struct Kp {  // cv::KeyPoint layout
    float x, y, size, angle, response;
    int32_t octave, class_id;
};

struct FrameView {  // same layout as orbm_frame_view (include/orbm.h)
    const Kp* keys;
    const uint8_t* desc;
    int32_t n;
    float min_x, min_y, max_x, max_y;
    float grid_w_inv, grid_h_inv;
    const float* scale_factors;
    int32_t nlevels;
    const float* u_right;
};

struct MapPointPod {  // orbm_map_point
    float proj_x, proj_y, proj_xr, view_cos, depth;
    int32_t level, observations;
    uint8_t in_view, bad, pad[2];
};

struct FrameRight {  // orbm_frame_right
    const Kp* keys;
    const uint8_t* desc;
    int32_t n;
    const int32_t* left_to_right;
    const int32_t* right_to_left;
};

struct MapPointRightPod {  // orbm_map_point_right
    float proj_x, proj_y, view_cos;
    int32_t level;
    uint8_t in_view, pad[3];
};

struct BowView {  // orbm_bow_view
    const Kp* keys;
    const uint8_t* desc;
    int32_t n, nnodes;
    const uint32_t* node_ids;
    const int32_t* node_ptr;
    const uint32_t* node_items;
};

struct LastObs {  // orbm_last_obs
    float u, v, inv_z, angle;
    int32_t octave, observations;
    uint8_t valid, pad[3];
};

int descriptor_distance(const uint8_t* a, const uint8_t* b) {
    const int32_t* pa = (const int32_t*)a;
    const int32_t* pb = (const int32_t*)b;
    int dist = 0;
    for (int i = 0; i < 8; i++, pa++, pb++) {
        unsigned int v = *pa ^ *pb;
        v = v - ((v >> 1) & 0x55555555);
        v = (v & 0x33333333) + ((v >> 2) & 0x33333333);
        dist += (((v + (v >> 4)) & 0xF0F0F0F) * 0x1010101) >> 24;
    }
    return dist;
}

void three_maxima(std::vector<int>* histo, const int L, int& ind1, int& ind2, int& ind3) {
    int max1 = 0, max2 = 0, max3 = 0;
    for (int i = 0; i < L; i++) {
        const int s = histo[i].size();
        if (s > max1) {
            max3 = max2; max2 = max1; max1 = s;
            ind3 = ind2; ind2 = ind1; ind1 = i;
        } else if (s > max2) {
            max3 = max2; max2 = s;
            ind3 = ind2; ind2 = i;
        } else if (s > max3) {
            max3 = s; ind3 = i;
        }
    }
    if (max2 < 0.1f * (float)max1) {
        ind2 = -1; ind3 = -1;
    } else if (max3 < 0.1f * (float)max1) {
        ind3 = -1;
    }
}

struct GridO {
    std::vector<size_t> cell[GRID_COLS][GRID_ROWS];
    explicit GridO(const FrameView& F) {
        for (int i = 0; i < F.n; i++) {
            const Kp& kp = F.keys[i];
            int posX = round((kp.x - F.min_x) * F.grid_w_inv);
            int posY = round((kp.y - F.min_y) * F.grid_h_inv);
            if (posX < 0 || posX >= GRID_COLS || posY < 0 || posY >= GRID_ROWS) continue;
            cell[posX][posY].push_back(i);
        }
    }
};

std::vector<size_t> get_features_in_area(const FrameView& F, const GridO& G, const float& x, const float& y,
                                         const float& r, const int minLevel = -1, const int maxLevel = -1) {
    std::vector<size_t> vIndices;
    float factorX = r, factorY = r;
    const int nMinCellX = std::max(0, (int)floor((x - F.min_x - factorX) * F.grid_w_inv));
    if (nMinCellX >= GRID_COLS) return vIndices;
    const int nMaxCellX = std::min((int)GRID_COLS - 1, (int)ceil((x - F.min_x + factorX) * F.grid_w_inv));
    if (nMaxCellX < 0) return vIndices;
    const int nMinCellY = std::max(0, (int)floor((y - F.min_y - factorY) * F.grid_h_inv));
    if (nMinCellY >= GRID_ROWS) return vIndices;
    const int nMaxCellY = std::min((int)GRID_ROWS - 1, (int)ceil((y - F.min_y + factorY) * F.grid_h_inv));
    if (nMaxCellY < 0) return vIndices;
    const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
    for (int ix = nMinCellX; ix <= nMaxCellX; ix++) {
        for (int iy = nMinCellY; iy <= nMaxCellY; iy++) {
            const std::vector<size_t>& vCell = G.cell[ix][iy];
            for (size_t j = 0, jend = vCell.size(); j < jend; j++) {
                const Kp& kpUn = F.keys[vCell[j]];
                if (bCheckLevels) {
                    if (kpUn.octave < minLevel) continue;
                    if (maxLevel >= 0)
                        if (kpUn.octave > maxLevel) continue;
                }
                const float distx = kpUn.x - x;
                const float disty = kpUn.y - y;
                if (fabs(distx) < factorX && fabs(disty) < factorY) vIndices.push_back(vCell[j]);
            }
        }
    }
    return vIndices;
}

float radius_by_viewing_cos(const float& viewCos) {
    if (viewCos > 0.998) return 2.5;
    else return 4.0;
}

}  // namespace

extern "C" {

int orbo_descriptor_distance(const uint8_t* a, const uint8_t* b) { return descriptor_distance(a, b); }

int orbo_features_in_area(const FrameView* F, float x, float y, float r, int minLevel, int maxLevel,
                          int32_t* out, int cap) {
    GridO G(*F);
    std::vector<size_t> v = get_features_in_area(*F, G, x, y, r, minLevel, maxLevel);
    for (size_t i = 0; i < v.size() && (int)i < cap; ++i) out[i] = (int32_t)v[i];
    return (int)v.size();
}

// src/ORBmatcher.cc:48-147 (F.Nleft == -1)
int orbo_search_by_projection_points(const FrameView* Fp, const MapPointPod* mps, const uint8_t* mp_desc, int m,
                                     float th, int bFarPoints, float thFarPoints, float mfNNratio,
                                     int32_t* kp_obs, int32_t* kp_match) {
    const FrameView& F = *Fp;
    GridO G(F);
    for (int i = 0; i < F.n; ++i) kp_match[i] = -1;
    int nmatches = 0;
    const bool bFactor = th != 1.0;
    for (int iMP = 0; iMP < m; iMP++) {
        const MapPointPod& pMP = mps[iMP];
        if (!pMP.in_view) continue;
        if (bFarPoints && pMP.depth > thFarPoints) continue;
        if (pMP.bad) continue;
        const int& nPredictedLevel = pMP.level;
        float r = radius_by_viewing_cos(pMP.view_cos);
        if (bFactor) r *= th;
        const std::vector<size_t> vIndices = get_features_in_area(
            F, G, pMP.proj_x, pMP.proj_y, r * F.scale_factors[nPredictedLevel], nPredictedLevel - 1, nPredictedLevel);
        if (vIndices.empty()) continue;
        const uint8_t* MPdescriptor = mp_desc + (size_t)iMP * 32;
        int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
        for (std::vector<size_t>::const_iterator vit = vIndices.begin(), vend = vIndices.end(); vit != vend; vit++) {
            const size_t idx = *vit;
            if (kp_obs[idx] > 0) continue;
            if (F.u_right && F.u_right[idx] > 0) {
                const float er = fabs(pMP.proj_xr - F.u_right[idx]);
                if (er > r * F.scale_factors[nPredictedLevel]) continue;
            }
            const int dist = descriptor_distance(MPdescriptor, F.desc + idx * 32);
            if (dist < bestDist) {
                bestDist2 = bestDist;
                bestDist = dist;
                bestLevel2 = bestLevel;
                bestLevel = F.keys[idx].octave;
                bestIdx = idx;
            } else if (dist < bestDist2) {
                bestLevel2 = F.keys[idx].octave;
                bestDist2 = dist;
            }
        }
        if (bestDist <= TH_HIGH) {
            if (bestLevel == bestLevel2 && bestDist > mfNNratio * bestDist2) continue;
            if (bestLevel != bestLevel2 || bestDist <= mfNNratio * bestDist2) {
                kp_match[bestIdx] = iMP;
                kp_obs[bestIdx] = pMP.observations;
                nmatches++;
            }
        }
    }
    return nmatches;
}

// src/ORBmatcher.cc:1965-2076 + 2158-2181 (CurrentFrame.Nleft == -1)
int orbo_search_by_projection_frames(const FrameView* Cp, const LastObs* last, const uint8_t* last_desc, int n_last,
                                     float th, int bForward, int bBackward, float mbf, int mbCheckOrientation,
                                     int32_t* kp_obs, int32_t* kp_match) {
    const FrameView& CurrentFrame = *Cp;
    GridO G(CurrentFrame);
    for (int i = 0; i < CurrentFrame.n; ++i) kp_match[i] = -1;
    int nmatches = 0;
    std::vector<int> rotHist[HISTO_LENGTH];
    for (int i = 0; i < HISTO_LENGTH; i++) rotHist[i].reserve(500);
    const float factor = 1.0f / HISTO_LENGTH;
    for (int i = 0; i < n_last; i++) {
        const LastObs& o = last[i];
        if (!o.valid) continue;
        const float invzc = o.inv_z;
        if (invzc < 0) continue;
        if (o.u < CurrentFrame.min_x || o.u > CurrentFrame.max_x) continue;
        if (o.v < CurrentFrame.min_y || o.v > CurrentFrame.max_y) continue;
        int nLastOctave = o.octave;
        float radius = th * CurrentFrame.scale_factors[nLastOctave];
        std::vector<size_t> vIndices2;
        if (bForward) vIndices2 = get_features_in_area(CurrentFrame, G, o.u, o.v, radius, nLastOctave);
        else if (bBackward) vIndices2 = get_features_in_area(CurrentFrame, G, o.u, o.v, radius, 0, nLastOctave);
        else vIndices2 = get_features_in_area(CurrentFrame, G, o.u, o.v, radius, nLastOctave - 1, nLastOctave + 1);
        if (vIndices2.empty()) continue;
        const uint8_t* dMP = last_desc + (size_t)i * 32;
        int bestDist = 256;
        int bestIdx2 = -1;
        for (std::vector<size_t>::const_iterator vit = vIndices2.begin(), vend = vIndices2.end(); vit != vend; vit++) {
            const size_t i2 = *vit;
            if (kp_obs[i2] > 0) continue;
            if (CurrentFrame.u_right && CurrentFrame.u_right[i2] > 0) {
                const float ur = o.u - mbf * invzc;
                const float er = fabs(ur - CurrentFrame.u_right[i2]);
                if (er > radius) continue;
            }
            const int dist = descriptor_distance(dMP, CurrentFrame.desc + i2 * 32);
            if (dist < bestDist) {
                bestDist = dist;
                bestIdx2 = i2;
            }
        }
        if (bestDist <= TH_HIGH) {
            kp_match[bestIdx2] = i;
            kp_obs[bestIdx2] = o.observations;
            nmatches++;
            if (mbCheckOrientation) {
                float rot = o.angle - CurrentFrame.keys[bestIdx2].angle;
                if (rot < 0.0) rot += 360.0f;
                int bin = round(rot * factor);
                if (bin == HISTO_LENGTH) bin = 0;
                rotHist[bin].push_back(bestIdx2);
            }
        }
    }
    if (mbCheckOrientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i != ind1 && i != ind2 && i != ind3) {
                for (size_t j = 0, jend = rotHist[i].size(); j < jend; j++) {
                    kp_match[rotHist[i][j]] = -1;
                    kp_obs[rotHist[i][j]] = -1;
                    nmatches--;
                }
            }
        }
    }
    return nmatches;
}

// src/ORBmatcher.cc:710-825
int orbo_search_for_initialization(const FrameView* F1p, const FrameView* F2p, float* vbPrevMatched,
                                   int32_t* vnMatches12, int windowSize, float mfNNratio, int mbCheckOrientation) {
    const FrameView &F1 = *F1p, &F2 = *F2p;
    GridO G2(F2);
    int nmatches = 0;
    for (int i = 0; i < F1.n; ++i) vnMatches12[i] = -1;
    std::vector<int> rotHist[HISTO_LENGTH];
    for (int i = 0; i < HISTO_LENGTH; i++) rotHist[i].reserve(500);
    const float factor = 1.0f / HISTO_LENGTH;
    std::vector<int> vMatchedDistance(F2.n, INT_MAX);
    std::vector<int> vnMatches21(F2.n, -1);
    for (size_t i1 = 0, iend1 = F1.n; i1 < iend1; i1++) {
        Kp kp1 = F1.keys[i1];
        int level1 = kp1.octave;
        if (level1 > 0) continue;
        std::vector<size_t> vIndices2 = get_features_in_area(F2, G2, vbPrevMatched[2 * i1], vbPrevMatched[2 * i1 + 1],
                                                             windowSize, level1, level1);
        if (vIndices2.empty()) continue;
        const uint8_t* d1 = F1.desc + i1 * 32;
        int bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx2 = -1;
        for (std::vector<size_t>::iterator vit = vIndices2.begin(); vit != vIndices2.end(); vit++) {
            size_t i2 = *vit;
            int dist = descriptor_distance(d1, F2.desc + i2 * 32);
            if (vMatchedDistance[i2] <= dist) continue;
            if (dist < bestDist) {
                bestDist2 = bestDist;
                bestDist = dist;
                bestIdx2 = i2;
            } else if (dist < bestDist2) {
                bestDist2 = dist;
            }
        }
        if (bestDist <= TH_LOW) {
            if (bestDist < (float)bestDist2 * mfNNratio) {
                if (vnMatches21[bestIdx2] >= 0) {
                    vnMatches12[vnMatches21[bestIdx2]] = -1;
                    nmatches--;
                }
                vnMatches12[i1] = bestIdx2;
                vnMatches21[bestIdx2] = i1;
                vMatchedDistance[bestIdx2] = bestDist;
                nmatches++;
                if (mbCheckOrientation) {
                    float rot = F1.keys[i1].angle - F2.keys[bestIdx2].angle;
                    if (rot < 0.0) rot += 360.0f;
                    int bin = round(rot * factor);
                    if (bin == HISTO_LENGTH) bin = 0;
                    rotHist[bin].push_back(i1);
                }
            }
        }
    }
    if (mbCheckOrientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (size_t j = 0, jend = rotHist[i].size(); j < jend; j++) {
                int idx1 = rotHist[i][j];
                if (vnMatches12[idx1] >= 0) {
                    vnMatches12[idx1] = -1;
                    nmatches--;
                }
            }
        }
    }
    for (size_t i1 = 0, iend1 = F1.n; i1 < iend1; i1++)
        if (vnMatches12[i1] >= 0) {
            vbPrevMatched[2 * i1] = F2.keys[vnMatches12[i1]].x;
            vbPrevMatched[2 * i1 + 1] = F2.keys[vnMatches12[i1]].y;
        }
    return nmatches;
}

// src/Frame.cc:802-885: row table + per-left-keypoint Hamming argmin.  n_rows = mvImagePyramid[0].rows.
int orbo_stereo_rows(const Kp* mvKeys, const uint8_t* mDescriptors, int N, const Kp* mvKeysRight,
                     const uint8_t* mDescriptorsRight, int Nr, const float* mvScaleFactors, int n_rows,
                     float minD, float maxD, int32_t* best_idx, int32_t* best_dist) {
    std::vector<std::vector<size_t> > vRowIndices(n_rows, std::vector<size_t>());
    for (int iR = 0; iR < Nr; iR++) {
        const Kp& kp = mvKeysRight[iR];
        const float& kpY = kp.y;
        const float r = 2.0f * mvScaleFactors[mvKeysRight[iR].octave];
        const int maxr = ceil(kpY + r);
        const int minr = floor(kpY - r);
        for (int yi = minr; yi <= maxr; yi++)
            if (yi >= 0 && yi < n_rows) vRowIndices[yi].push_back(iR);  // (the reference indexes unchecked)
    }
    for (int iL = 0; iL < N; iL++) {
        best_idx[iL] = -1;
        best_dist[iL] = TH_HIGH;
        const Kp& kpL = mvKeys[iL];
        const int& levelL = kpL.octave;
        const float& vL = kpL.y;
        const float& uL = kpL.x;
        const std::vector<size_t>& vCandidates = vRowIndices[(size_t)vL];
        if (vCandidates.empty()) continue;
        const float minU = uL - maxD;
        const float maxU = uL - minD;
        if (maxU < 0) continue;
        int bestDist = TH_HIGH;
        size_t bestIdxR = 0;
        bool found = false;
        const uint8_t* dL = mDescriptors + (size_t)iL * 32;
        for (size_t iC = 0; iC < vCandidates.size(); iC++) {
            const size_t iR = vCandidates[iC];
            const Kp& kpR = mvKeysRight[iR];
            if (kpR.octave < levelL - 1 || kpR.octave > levelL + 1) continue;
            const float& uR = kpR.x;
            if (uR >= minU && uR <= maxU) {
                const int dist = descriptor_distance(dL, mDescriptorsRight + iR * 32);
                if (dist < bestDist) {
                    bestDist = dist;
                    bestIdxR = iR;
                    found = true;
                }
            }
        }
        if (found) {
            best_idx[iL] = (int32_t)bestIdxR;
            best_dist[iL] = bestDist;
        }
    }
    return 0;
}

// src/Frame.cc:802-980, complete: the Hamming search above, then the SAD sub-pixel refinement on
// the two padded pyramids (mvImagePyramid[l] is a view into a buffer with a 19-px
// BORDER_REFLECT_101 border: pyr_l[l] / pyr_r[l] point at pixel (0,0) of level l inside such a
// buffer with row stride stride_[l]), the parabola fit, the disparity gate and the median cut.
// cv::Mat::convertTo(CV_16S), "IL - IL.at<short>(w,w)" and cv::norm(NORM_L1) are restated as
// integer arithmetic (exact: |values| <= 510, 121 terms).
int orbo_compute_stereo_matches(const Kp* mvKeys, const uint8_t* mDescriptors, int N, const Kp* mvKeysRight,
                                const uint8_t* mDescriptorsRight, int Nr, const float* mvScaleFactors,
                                const float* mvInvScaleFactors, const uint8_t* const* pyr_l,
                                const int* stride_l, const uint8_t* const* pyr_r, const int* stride_r,
                                const int* cols_r, int n_rows, float mb, float mbf, float* mvuRight,
                                float* mvDepth) {
    for (int i = 0; i < N; ++i) mvuRight[i] = mvDepth[i] = -1.0f;
    const int thOrbDist = (TH_HIGH + TH_LOW) / 2;
    const float minZ = mb;
    const float minD = 0;
    const float maxD = mbf / minZ;
    std::vector<int32_t> best_idx(N), best_dist(N);
    orbo_stereo_rows(mvKeys, mDescriptors, N, mvKeysRight, mDescriptorsRight, Nr, mvScaleFactors, n_rows, minD,
                     maxD, best_idx.data(), best_dist.data());
    std::vector<std::pair<int, int> > vDistIdx;
    vDistIdx.reserve(N);
    for (int iL = 0; iL < N; iL++) {
        if (best_idx[iL] < 0) continue;
        const Kp& kpL = mvKeys[iL];
        const float& uL = kpL.x;
        const int bestIdxR = best_idx[iL];
        if (best_dist[iL] < thOrbDist) {
            const float uR0 = mvKeysRight[bestIdxR].x;
            const float scaleFactor = mvInvScaleFactors[kpL.octave];
            const float scaleduL = round(kpL.x * scaleFactor);
            const float scaledvL = round(kpL.y * scaleFactor);
            const float scaleduR0 = round(uR0 * scaleFactor);
            const int w = 5;
            const uint8_t* PL = pyr_l[kpL.octave];
            const int sl = stride_l[kpL.octave];
            const uint8_t* PR = pyr_r[kpL.octave];
            const int sr = stride_r[kpL.octave];
            short IL[11][11];
            {
                const int y0 = (int)(scaledvL - w), x0 = (int)(scaleduL - w);
                for (int r = 0; r < 11; ++r)
                    for (int c = 0; c < 11; ++c) IL[r][c] = (short)PL[(y0 + r) * sl + (x0 + c)];
                const short ctr = IL[w][w];
                for (int r = 0; r < 11; ++r)
                    for (int c = 0; c < 11; ++c) IL[r][c] = (short)(IL[r][c] - ctr);
            }
            int bestDist = INT_MAX;
            int bestincR = 0;
            const int L = 5;
            std::vector<float> vDists;
            vDists.resize(2 * L + 1);
            const float iniu = scaleduR0 + L - w;
            const float endu = scaleduR0 + L + w + 1;
            if (iniu < 0 || endu >= cols_r[kpL.octave]) continue;
            for (int incR = -L; incR <= +L; incR++) {
                short IR[11][11];
                const int y0 = (int)(scaledvL - w), x0 = (int)(scaleduR0 + incR - w);
                for (int r = 0; r < 11; ++r)
                    for (int c = 0; c < 11; ++c) IR[r][c] = (short)PR[(y0 + r) * sr + (x0 + c)];
                const short ctr = IR[w][w];
                int l1 = 0;
                for (int r = 0; r < 11; ++r)
                    for (int c = 0; c < 11; ++c) l1 += abs((int)IL[r][c] - (int)(short)(IR[r][c] - ctr));
                float dist = (float)l1;
                if (dist < bestDist) {
                    bestDist = dist;
                    bestincR = incR;
                }
                vDists[L + incR] = dist;
            }
            if (bestincR == -L || bestincR == L) continue;
            const float dist1 = vDists[L + bestincR - 1];
            const float dist2 = vDists[L + bestincR];
            const float dist3 = vDists[L + bestincR + 1];
            const float deltaR = (dist1 - dist3) / (2.0f * (dist1 + dist3 - 2.0f * dist2));
            if (deltaR < -1 || deltaR > 1) continue;
            float bestuR = mvScaleFactors[kpL.octave] * ((float)scaleduR0 + (float)bestincR + deltaR);
            float disparity = (uL - bestuR);
            if (disparity >= minD && disparity < maxD) {
                if (disparity <= 0) {
                    disparity = 0.01;
                    bestuR = uL - 0.01;
                }
                mvDepth[iL] = mbf / disparity;
                mvuRight[iL] = bestuR;
                vDistIdx.push_back(std::pair<int, int>(bestDist, iL));
            }
        }
    }
    if (vDistIdx.empty()) return 0;  // (the reference indexes an empty vector here)
    sort(vDistIdx.begin(), vDistIdx.end());
    const float median = vDistIdx[vDistIdx.size() / 2].first;
    const float thDist = 1.5f * 1.4f * median;
    for (int i = vDistIdx.size() - 1; i >= 0; i--) {
        if (vDistIdx[i].first < thDist)
            break;
        else {
            mvuRight[vDistIdx[i].second] = -1;
            mvDepth[vDistIdx[i].second] = -1;
        }
    }
    return (int)vDistIdx.size();
}

// cv::BFMatcher(NORM_HAMMING).knnMatch(k = 2) semantics (verified against cv2 in tests): the two
// smallest distances, lowest train index first on ties; missing entries are -1.
int orbo_knn2(const uint8_t* q, int nq, const uint8_t* t, int nt, int32_t* idx_out, int32_t* dist_out) {
    for (int i = 0; i < nq; ++i) {
        int b1 = INT_MAX, b2 = INT_MAX, i1 = -1, i2 = -1;
        for (int j = 0; j < nt; ++j) {
            const int d = descriptor_distance(q + (size_t)i * 32, t + (size_t)j * 32);
            if (d < b1) { b2 = b1; i2 = i1; b1 = d; i1 = j; }
            else if (d < b2) { b2 = d; i2 = j; }
        }
        idx_out[2 * i] = i1;
        dist_out[2 * i] = i1 < 0 ? -1 : b1;
        idx_out[2 * i + 1] = i2;
        dist_out[2 * i + 1] = i2 < 0 ? -1 : b2;
    }
    return 0;
}

// The right camera's half of a fisheye stereo frame seen through the same grid geometry:
// AssignFeaturesToGrid (src/Frame.cc:377-408) puts right keypoints into mGridRight with PosInGrid
// of the frame's own bounds, GetFeaturesInArea(..., bRight = true) reads mGridRight / mvKeysRight.
static FrameView right_half(const FrameView& F, const FrameRight& R) {
    FrameView v = F;
    v.keys = R.keys;
    v.desc = R.desc;
    v.n = R.n;
    v.u_right = nullptr;
    return v;
}

// src/ORBmatcher.cc:48-218 with F.Nleft != -1 (fisheye stereo): mvpMapPoints / descriptors are
// indexed [0, Nleft) for the left and Nleft + j for the right camera.
int orbo_search_by_projection_points_stereo(const FrameView* Fp, const FrameRight* Rp, const MapPointPod* mps,
                                            const MapPointRightPod* mps_r, const uint8_t* mp_desc, int m, float th,
                                            int bFarPoints, float thFarPoints, float mfNNratio, int32_t* kp_obs,
                                            int32_t* kp_match) {
    const FrameView& F = *Fp;
    const FrameRight& R = *Rp;
    const FrameView FR = right_half(F, R);
    GridO G(F), GR(FR);
    const int Nleft = F.n;
    for (int i = 0; i < Nleft + R.n; ++i) kp_match[i] = -1;
    int nmatches = 0, left = 0, right = 0;
    const bool bFactor = th != 1.0;
    for (int iMP = 0; iMP < m; iMP++) {
        const MapPointPod& pMP = mps[iMP];
        const MapPointRightPod& pR = mps_r[iMP];
        if (!pMP.in_view && !pR.in_view) continue;
        if (bFarPoints && pMP.depth > thFarPoints) continue;
        if (pMP.bad) continue;
        if (pMP.in_view) {
            const int& nPredictedLevel = pMP.level;
            float r = radius_by_viewing_cos(pMP.view_cos);
            if (bFactor) r *= th;
            const std::vector<size_t> vIndices = get_features_in_area(
                F, G, pMP.proj_x, pMP.proj_y, r * F.scale_factors[nPredictedLevel], nPredictedLevel - 1, nPredictedLevel);
            if (!vIndices.empty()) {
                const uint8_t* MPdescriptor = mp_desc + (size_t)iMP * 32;
                int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
                for (std::vector<size_t>::const_iterator vit = vIndices.begin(), vend = vIndices.end(); vit != vend; vit++) {
                    const size_t idx = *vit;
                    if (kp_obs[idx] > 0) continue;
                    // (F.Nleft == -1 && F.mvuRight[idx] > 0) is false here
                    const int dist = descriptor_distance(MPdescriptor, F.desc + idx * 32);
                    if (dist < bestDist) {
                        bestDist2 = bestDist;
                        bestDist = dist;
                        bestLevel2 = bestLevel;
                        bestLevel = F.keys[idx].octave;   // idx < Nleft: mvKeys[idx]
                        bestIdx = idx;
                    } else if (dist < bestDist2) {
                        bestLevel2 = F.keys[idx].octave;
                        bestDist2 = dist;
                    }
                }
                if (bestDist <= TH_HIGH) {
                    if (bestLevel == bestLevel2 && bestDist > mfNNratio * bestDist2) continue;
                    if (bestLevel != bestLevel2 || bestDist <= mfNNratio * bestDist2) {
                        kp_match[bestIdx] = iMP;
                        kp_obs[bestIdx] = pMP.observations;
                        if (R.left_to_right && R.left_to_right[bestIdx] != -1) {
                            kp_match[R.left_to_right[bestIdx] + Nleft] = iMP;
                            kp_obs[R.left_to_right[bestIdx] + Nleft] = pMP.observations;
                            nmatches++;
                            right++;
                        }
                        nmatches++;
                        left++;
                    }
                }
            }
        }
        if (pR.in_view) {
            const int& nPredictedLevel = pR.level;
            if (nPredictedLevel != -1) {
                float r = radius_by_viewing_cos(pR.view_cos);
                const std::vector<size_t> vIndices = get_features_in_area(
                    FR, GR, pR.proj_x, pR.proj_y, r * F.scale_factors[nPredictedLevel], nPredictedLevel - 1, nPredictedLevel);
                if (vIndices.empty()) continue;
                const uint8_t* MPdescriptor = mp_desc + (size_t)iMP * 32;
                int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
                for (std::vector<size_t>::const_iterator vit = vIndices.begin(), vend = vIndices.end(); vit != vend; vit++) {
                    const size_t idx = *vit;
                    if (kp_obs[idx + Nleft] > 0) continue;
                    const int dist = descriptor_distance(MPdescriptor, R.desc + idx * 32);
                    if (dist < bestDist) {
                        bestDist2 = bestDist;
                        bestDist = dist;
                        bestLevel2 = bestLevel;
                        bestLevel = R.keys[idx].octave;
                        bestIdx = idx;
                    } else if (dist < bestDist2) {
                        bestLevel2 = R.keys[idx].octave;
                        bestDist2 = dist;
                    }
                }
                if (bestDist <= TH_HIGH) {
                    if (bestLevel == bestLevel2 && bestDist > mfNNratio * bestDist2) continue;
                    if (R.right_to_left && R.right_to_left[bestIdx] != -1) {
                        kp_match[R.right_to_left[bestIdx]] = iMP;
                        kp_obs[R.right_to_left[bestIdx]] = pMP.observations;
                        nmatches++;
                        left++;
                    }
                    kp_match[bestIdx + Nleft] = iMP;
                    kp_obs[bestIdx + Nleft] = pMP.observations;
                    nmatches++;
                    right++;
                }
            }
        }
    }
    (void)left; (void)right;
    return nmatches;
}

// src/ORBmatcher.cc:1965-2181 with CurrentFrame.Nleft != -1.  uv_right[2i..2i+1] = project(mTrl * x3Dc).
int orbo_search_by_projection_frames_stereo(const FrameView* Cp, const FrameRight* Rp, const LastObs* last,
                                            const float* uv_right, const uint8_t* last_desc, int n_last, float th,
                                            int bForward, int bBackward, int mbCheckOrientation, int32_t* kp_obs,
                                            int32_t* kp_match) {
    const FrameView& CurrentFrame = *Cp;
    const FrameRight& R = *Rp;
    const FrameView FR = right_half(CurrentFrame, R);
    GridO G(CurrentFrame), GR(FR);
    const int Nleft = CurrentFrame.n;
    for (int i = 0; i < Nleft + R.n; ++i) kp_match[i] = -1;
    int nmatches = 0;
    std::vector<int> rotHist[HISTO_LENGTH];
    for (int i = 0; i < HISTO_LENGTH; i++) rotHist[i].reserve(500);
    const float factor = 1.0f / HISTO_LENGTH;
    for (int i = 0; i < n_last; i++) {
        const LastObs& o = last[i];
        if (!o.valid) continue;
        const float invzc = o.inv_z;
        if (invzc < 0) continue;
        if (o.u < CurrentFrame.min_x || o.u > CurrentFrame.max_x) continue;
        if (o.v < CurrentFrame.min_y || o.v > CurrentFrame.max_y) continue;
        int nLastOctave = o.octave;
        float radius = th * CurrentFrame.scale_factors[nLastOctave];
        std::vector<size_t> vIndices2;
        if (bForward) vIndices2 = get_features_in_area(CurrentFrame, G, o.u, o.v, radius, nLastOctave);
        else if (bBackward) vIndices2 = get_features_in_area(CurrentFrame, G, o.u, o.v, radius, 0, nLastOctave);
        else vIndices2 = get_features_in_area(CurrentFrame, G, o.u, o.v, radius, nLastOctave - 1, nLastOctave + 1);
        if (vIndices2.empty()) continue;
        const uint8_t* dMP = last_desc + (size_t)i * 32;
        {
            int bestDist = 256, bestIdx2 = -1;
            for (std::vector<size_t>::const_iterator vit = vIndices2.begin(), vend = vIndices2.end(); vit != vend; vit++) {
                const size_t i2 = *vit;
                if (kp_obs[i2] > 0) continue;
                const int dist = descriptor_distance(dMP, CurrentFrame.desc + i2 * 32);
                if (dist < bestDist) {
                    bestDist = dist;
                    bestIdx2 = i2;
                }
            }
            if (bestDist <= TH_HIGH) {
                kp_match[bestIdx2] = i;
                kp_obs[bestIdx2] = o.observations;
                nmatches++;
                if (mbCheckOrientation) {
                    float rot = o.angle - CurrentFrame.keys[bestIdx2].angle;
                    if (rot < 0.0) rot += 360.0f;
                    int bin = round(rot * factor);
                    if (bin == HISTO_LENGTH) bin = 0;
                    rotHist[bin].push_back(bestIdx2);
                }
            }
        }
        {
            const float ux = uv_right[2 * i], uy = uv_right[2 * i + 1];
            std::vector<size_t> vR;
            if (bForward) vR = get_features_in_area(FR, GR, ux, uy, radius, nLastOctave, -1);
            else if (bBackward) vR = get_features_in_area(FR, GR, ux, uy, radius, 0, nLastOctave);
            else vR = get_features_in_area(FR, GR, ux, uy, radius, nLastOctave - 1, nLastOctave + 1);
            int bestDist = 256, bestIdx2 = -1;
            for (std::vector<size_t>::const_iterator vit = vR.begin(), vend = vR.end(); vit != vend; vit++) {
                const size_t i2 = *vit;
                if (kp_obs[i2 + Nleft] > 0) continue;
                const int dist = descriptor_distance(dMP, R.desc + i2 * 32);
                if (dist < bestDist) {
                    bestDist = dist;
                    bestIdx2 = i2;
                }
            }
            if (bestDist <= TH_HIGH) {
                kp_match[bestIdx2 + Nleft] = i;
                kp_obs[bestIdx2 + Nleft] = o.observations;
                nmatches++;
                if (mbCheckOrientation) {
                    float rot = o.angle - R.keys[bestIdx2].angle;
                    if (rot < 0.0) rot += 360.0f;
                    int bin = round(rot * factor);
                    if (bin == HISTO_LENGTH) bin = 0;
                    rotHist[bin].push_back(bestIdx2 + Nleft);
                }
            }
        }
    }
    if (mbCheckOrientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i != ind1 && i != ind2 && i != ind3) {
                for (size_t j = 0, jend = rotHist[i].size(); j < jend; j++) {
                    kp_match[rotHist[i][j]] = -1;
                    kp_obs[rotHist[i][j]] = -1;
                    nmatches--;
                }
            }
        }
    }
    return nmatches;
}

// src/ORBmatcher.cc:273-475.  DBoW2::FeatureVector = std::map<NodeId, std::vector<unsigned int>>; the
// snapshot carries its (key, vector) pairs in key order, rebuilt here into the same std::map so that
// the iterator walk and the lower_bound calls are the reference's own.
int orbo_search_by_bow(const BowView* KFp, const uint8_t* kf_mp_ok, const BowView* Fp, int Nleft, float mfNNratio,
                       int mbCheckOrientation, int32_t* matches) {
    const BowView& KF = *KFp;
    const BowView& F = *Fp;
    typedef std::map<unsigned int, std::vector<unsigned int> > FeatureVector;
    FeatureVector vFeatVecKF, vFeatVecF;
    for (int a = 0; a < KF.nnodes; ++a)
        vFeatVecKF[KF.node_ids[a]] = std::vector<unsigned int>(KF.node_items + KF.node_ptr[a], KF.node_items + KF.node_ptr[a + 1]);
    for (int b = 0; b < F.nnodes; ++b)
        vFeatVecF[F.node_ids[b]] = std::vector<unsigned int>(F.node_items + F.node_ptr[b], F.node_items + F.node_ptr[b + 1]);
    for (int i = 0; i < F.n; ++i) matches[i] = -1;
    int nmatches = 0;
    std::vector<int> rotHist[HISTO_LENGTH];
    for (int i = 0; i < HISTO_LENGTH; i++) rotHist[i].reserve(500);
    const float factor = 1.0f / HISTO_LENGTH;
    FeatureVector::const_iterator KFit = vFeatVecKF.begin(), Fit = vFeatVecF.begin();
    FeatureVector::const_iterator KFend = vFeatVecKF.end(), Fend = vFeatVecF.end();
    while (KFit != KFend && Fit != Fend) {
        if (KFit->first == Fit->first) {
            const std::vector<unsigned int> vIndicesKF = KFit->second;
            const std::vector<unsigned int> vIndicesF = Fit->second;
            for (size_t iKF = 0; iKF < vIndicesKF.size(); iKF++) {
                const unsigned int realIdxKF = vIndicesKF[iKF];
                if (!kf_mp_ok[realIdxKF]) continue;   // !pMP || pMP->isBad()
                const uint8_t* dKF = KF.desc + (size_t)realIdxKF * 32;
                int bestDist1 = 256, bestIdxF = -1, bestDist2 = 256;
                int bestDist1R = 256, bestIdxFR = -1, bestDist2R = 256;
                for (size_t iF = 0; iF < vIndicesF.size(); iF++) {
                    const unsigned int realIdxF = vIndicesF[iF];
                    if (matches[realIdxF] >= 0) continue;
                    const int dist = descriptor_distance(dKF, F.desc + (size_t)realIdxF * 32);
                    if (Nleft == -1) {
                        if (dist < bestDist1) {
                            bestDist2 = bestDist1;
                            bestDist1 = dist;
                            bestIdxF = realIdxF;
                        } else if (dist < bestDist2) {
                            bestDist2 = dist;
                        }
                    } else {
                        if ((int)realIdxF < Nleft && dist < bestDist1) {
                            bestDist2 = bestDist1;
                            bestDist1 = dist;
                            bestIdxF = realIdxF;
                        } else if ((int)realIdxF < Nleft && dist < bestDist2) {
                            bestDist2 = dist;
                        }
                        if ((int)realIdxF >= Nleft && dist < bestDist1R) {
                            bestDist2R = bestDist1R;
                            bestDist1R = dist;
                            bestIdxFR = realIdxF;
                        } else if ((int)realIdxF >= Nleft && dist < bestDist2R) {
                            bestDist2R = dist;
                        }
                    }
                }
                if (bestDist1 <= TH_LOW) {
                    if (static_cast<float>(bestDist1) < mfNNratio * static_cast<float>(bestDist2)) {
                        matches[bestIdxF] = realIdxKF;
                        if (mbCheckOrientation) {
                            float rot = KF.keys[realIdxKF].angle - F.keys[bestIdxF].angle;
                            if (rot < 0.0) rot += 360.0f;
                            int bin = round(rot * factor);
                            if (bin == HISTO_LENGTH) bin = 0;
                            rotHist[bin].push_back(bestIdxF);
                        }
                        nmatches++;
                    }
                    if (bestDist1R <= TH_LOW) {
                        if (static_cast<float>(bestDist1R) < mfNNratio * static_cast<float>(bestDist2R) || true) {
                            matches[bestIdxFR] = realIdxKF;
                            if (mbCheckOrientation) {
                                float rot = KF.keys[realIdxKF].angle - F.keys[bestIdxFR].angle;
                                if (rot < 0.0) rot += 360.0f;
                                int bin = round(rot * factor);
                                if (bin == HISTO_LENGTH) bin = 0;
                                rotHist[bin].push_back(bestIdxFR);
                            }
                            nmatches++;
                        }
                    }
                }
            }
            KFit++;
            Fit++;
        } else if (KFit->first < Fit->first) {
            KFit = vFeatVecKF.lower_bound(Fit->first);
        } else {
            Fit = vFeatVecF.lower_bound(KFit->first);
        }
    }
    if (mbCheckOrientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (size_t j = 0, jend = rotHist[i].size(); j < jend; j++) {
                matches[rotHist[i][j]] = -1;
                nmatches--;
            }
        }
    }
    return nmatches;
}

// ---- the LocalMapping / LoopClosing / relocalisation searches ---------------------
struct ProjPoint {  // orbm_proj_point
    float u, v, ur, angle;
    int32_t level;
    uint8_t valid, pad[3];
};

typedef std::map<unsigned int, std::vector<unsigned int> > FeatVec;
static FeatVec feat_vec_of(const BowView& V) {
    FeatVec fv;
    for (int a = 0; a < V.nnodes; ++a)
        fv[V.node_ids[a]] = std::vector<unsigned int>(V.node_items + V.node_ptr[a], V.node_items + V.node_ptr[a + 1]);
    return fv;
}

// src/ORBmatcher.cc:827-967.  n_un = mvKeysUn.size() when NLeft != -1, else -1.
int orbo_search_by_bow_kf(const BowView* K1, const uint8_t* ok1, int n_un1, const BowView* K2, const uint8_t* ok2,
                          int n_un2, float mfNNratio, int mbCheckOrientation, int32_t* vpMatches12) {
    const FeatVec fv1 = feat_vec_of(*K1), fv2 = feat_vec_of(*K2);
    for (int i = 0; i < K1->n; ++i) vpMatches12[i] = -1;
    std::vector<bool> vbMatched2(K2->n, false);
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    int nmatches = 0;
    FeatVec::const_iterator f1it = fv1.begin(), f2it = fv2.begin();
    while (f1it != fv1.end() && f2it != fv2.end()) {
        if (f1it->first == f2it->first) {
            for (size_t i1 = 0; i1 < f1it->second.size(); i1++) {
                const size_t idx1 = f1it->second[i1];
                if (n_un1 != -1 && idx1 >= (size_t)n_un1) continue;
                if (!ok1[idx1]) continue;
                const uint8_t* d1 = K1->desc + idx1 * 32;
                int bestDist1 = 256, bestIdx2 = -1, bestDist2 = 256;
                for (size_t i2 = 0; i2 < f2it->second.size(); i2++) {
                    const size_t idx2 = f2it->second[i2];
                    if (n_un2 != -1 && idx2 >= (size_t)n_un2) continue;
                    if (vbMatched2[idx2] || !ok2[idx2]) continue;
                    const int dist = descriptor_distance(d1, K2->desc + idx2 * 32);
                    if (dist < bestDist1) {
                        bestDist2 = bestDist1;
                        bestDist1 = dist;
                        bestIdx2 = (int)idx2;
                    } else if (dist < bestDist2) {
                        bestDist2 = dist;
                    }
                }
                if (bestDist1 < TH_LOW) {
                    if (static_cast<float>(bestDist1) < mfNNratio * static_cast<float>(bestDist2)) {
                        vpMatches12[idx1] = bestIdx2;
                        vbMatched2[bestIdx2] = true;
                        if (mbCheckOrientation) {
                            float rot = K1->keys[idx1].angle - K2->keys[bestIdx2].angle;
                            if (rot < 0.0) rot += 360.0f;
                            int bin = round(rot * factor);
                            if (bin == HISTO_LENGTH) bin = 0;
                            rotHist[bin].push_back((int)idx1);
                        }
                        nmatches++;
                    }
                }
            }
            f1it++;
            f2it++;
        } else if (f1it->first < f2it->first) {
            f1it = fv1.lower_bound(f2it->first);
        } else {
            f2it = fv2.lower_bound(f1it->first);
        }
    }
    if (mbCheckOrientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (size_t j = 0; j < rotHist[i].size(); j++) {
                vpMatches12[rotHist[i][j]] = -1;
                nmatches--;
            }
        }
    }
    return nmatches;
}

// src/ORBmatcher.cc:477-590 and :592-708 (the loop body after the geometric gates).
int orbo_search_by_projection_sim3(const FrameView* KFp, const ProjPoint* pts, const uint8_t* mp_desc, int m, int th,
                                   float ratioHamming, const uint8_t* kp_taken, int32_t* kp_match) {
    const FrameView& KF = *KFp;
    const GridO G(KF);
    std::vector<bool> vpMatched(KF.n);
    for (int i = 0; i < KF.n; ++i) {
        vpMatched[i] = kp_taken && kp_taken[i];
        kp_match[i] = -1;
    }
    int nmatches = 0;
    for (int iMP = 0; iMP < m; iMP++) {
        if (!pts[iMP].valid) continue;
        const int nPredictedLevel = pts[iMP].level;
        const float radius = th * KF.scale_factors[nPredictedLevel];
        const std::vector<size_t> vIndices = get_features_in_area(KF, G, pts[iMP].u, pts[iMP].v, radius);
        if (vIndices.empty()) continue;
        const uint8_t* dMP = mp_desc + (size_t)iMP * 32;
        int bestDist = 256, bestIdx = -1;
        for (std::vector<size_t>::const_iterator vit = vIndices.begin(); vit != vIndices.end(); vit++) {
            const size_t idx = *vit;
            if (vpMatched[idx]) continue;
            const int kpLevel = KF.keys[idx].octave;
            if (kpLevel < nPredictedLevel - 1 || kpLevel > nPredictedLevel) continue;
            const int dist = descriptor_distance(dMP, KF.desc + idx * 32);
            if (dist < bestDist) {
                bestDist = dist;
                bestIdx = (int)idx;
            }
        }
        if (bestDist <= TH_LOW * ratioHamming) {
            vpMatched[bestIdx] = true;
            kp_match[bestIdx] = iMP;
            nmatches++;
        }
    }
    return nmatches;
}

// src/ORBmatcher.cc:2183-2305 (the loop body after the gates :2210-2233).
int orbo_search_by_projection_reloc(const FrameView* Fp, const ProjPoint* pts, const uint8_t* mp_desc, int m, float th,
                                    int ORBdist, int mbCheckOrientation, const uint8_t* kp_taken, int32_t* kp_match) {
    const FrameView& CF = *Fp;
    const GridO G(CF);
    std::vector<bool> has_mp(CF.n);
    for (int i = 0; i < CF.n; ++i) {
        has_mp[i] = kp_taken && kp_taken[i];
        kp_match[i] = -1;
    }
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    int nmatches = 0;
    for (int i = 0; i < m; i++) {
        if (!pts[i].valid) continue;
        const int nPredictedLevel = pts[i].level;
        const float radius = th * CF.scale_factors[nPredictedLevel];
        const std::vector<size_t> vIndices2 =
            get_features_in_area(CF, G, pts[i].u, pts[i].v, radius, nPredictedLevel - 1, nPredictedLevel + 1);
        if (vIndices2.empty()) continue;
        const uint8_t* dMP = mp_desc + (size_t)i * 32;
        int bestDist = 256, bestIdx2 = -1;
        for (size_t k = 0; k < vIndices2.size(); k++) {
            const size_t i2 = vIndices2[k];
            if (has_mp[i2]) continue;
            const int dist = descriptor_distance(dMP, CF.desc + i2 * 32);
            if (dist < bestDist) {
                bestDist = dist;
                bestIdx2 = (int)i2;
            }
        }
        if (bestDist <= ORBdist) {
            has_mp[bestIdx2] = true;
            kp_match[bestIdx2] = i;
            nmatches++;
            if (mbCheckOrientation) {
                float rot = pts[i].angle - CF.keys[bestIdx2].angle;
                if (rot < 0.0) rot += 360.0f;
                int bin = round(rot * factor);
                if (bin == HISTO_LENGTH) bin = 0;
                rotHist[bin].push_back(bestIdx2);
            }
        }
    }
    if (mbCheckOrientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i != ind1 && i != ind2 && i != ind3) {
                for (size_t j = 0; j < rotHist[i].size(); j++) {
                    kp_match[rotHist[i][j]] = -1;
                    nmatches--;
                }
            }
        }
    }
    return nmatches;
}

// The search of Fuse: src/ORBmatcher.cc:1519-1585 (inv_sigma2 != NULL) and :1685-1717 (NULL).
int orbo_fuse(const FrameView* KFp, const ProjPoint* pts, const uint8_t* mp_desc, int m, float th,
              const float* mvInvLevelSigma2, int idx_offset, int32_t* best_idx, int32_t* best_dist) {
    const FrameView& KF = *KFp;
    const GridO G(KF);
    int nFused = 0;
    for (int i = 0; i < m; i++) {
        best_idx[i] = -1;
        best_dist[i] = 256;
        if (!pts[i].valid) continue;
        const int nPredictedLevel = pts[i].level;
        const float radius = th * KF.scale_factors[nPredictedLevel];
        const float u = pts[i].u, v = pts[i].v, ur = pts[i].ur;
        const std::vector<size_t> vIndices = get_features_in_area(KF, G, u, v, radius);
        if (vIndices.empty()) continue;
        const uint8_t* dMP = mp_desc + (size_t)i * 32;
        int bestDist = 256, bestIdx = -1;
        for (std::vector<size_t>::const_iterator vit = vIndices.begin(); vit != vIndices.end(); vit++) {
            size_t idx = *vit;
            const Kp& kp = KF.keys[idx];
            const int kpLevel = kp.octave;
            if (kpLevel < nPredictedLevel - 1 || kpLevel > nPredictedLevel) continue;
            if (mvInvLevelSigma2) {
                if (KF.u_right && KF.u_right[idx] >= 0) {
                    const float kpx = kp.x, kpy = kp.y, kpr = KF.u_right[idx];
                    const float ex = u - kpx, ey = v - kpy, er = ur - kpr;
                    const float e2 = ex * ex + ey * ey + er * er;
                    if (e2 * mvInvLevelSigma2[kpLevel] > 7.8) continue;
                } else {
                    const float kpx = kp.x, kpy = kp.y;
                    const float ex = u - kpx, ey = v - kpy;
                    const float e2 = ex * ex + ey * ey;
                    if (e2 * mvInvLevelSigma2[kpLevel] > 5.99) continue;
                }
            }
            const int dist = descriptor_distance(dMP, KF.desc + idx * 32);
            if (dist < bestDist) {
                bestDist = dist;
                bestIdx = (int)idx + idx_offset;
            }
        }
        if (bestDist <= TH_LOW) {
            best_idx[i] = bestIdx;
            best_dist[i] = bestDist;
            nFused++;
        }
    }
    return nFused;
}

// src/ORBmatcher.cc:1739-1963 (the two projection passes after their gates, then the agreement test).
static void sim3_pass(const FrameView& KF, const ProjPoint* pts, const uint8_t* mp_desc, int n, float th,
                      std::vector<int>& vnMatch) {
    const GridO G(KF);
    vnMatch.assign(n, -1);
    for (int i1 = 0; i1 < n; i1++) {
        if (!pts[i1].valid) continue;
        const int nPredictedLevel = pts[i1].level;
        const float radius = th * KF.scale_factors[nPredictedLevel];
        const std::vector<size_t> vIndices = get_features_in_area(KF, G, pts[i1].u, pts[i1].v, radius);
        if (vIndices.empty()) continue;
        const uint8_t* dMP = mp_desc + (size_t)i1 * 32;
        int bestDist = INT_MAX, bestIdx = -1;
        for (size_t k = 0; k < vIndices.size(); k++) {
            const size_t idx = vIndices[k];
            const Kp& kp = KF.keys[idx];
            if (kp.octave < nPredictedLevel - 1 || kp.octave > nPredictedLevel) continue;
            const int dist = descriptor_distance(dMP, KF.desc + idx * 32);
            if (dist < bestDist) {
                bestDist = dist;
                bestIdx = (int)idx;
            }
        }
        if (bestDist <= TH_HIGH) vnMatch[i1] = bestIdx;
    }
}

int orbo_search_by_sim3(const FrameView* KF1, const FrameView* KF2, const ProjPoint* pts1, const uint8_t* mp_desc1,
                        const ProjPoint* pts2, const uint8_t* mp_desc2, float th, int32_t* match12) {
    const int N1 = KF1->n, N2 = KF2->n;
    std::vector<int> vnMatch1, vnMatch2;
    sim3_pass(*KF2, pts1, mp_desc1, N1, th, vnMatch1);
    sim3_pass(*KF1, pts2, mp_desc2, N2, th, vnMatch2);
    int nFound = 0;
    for (int i1 = 0; i1 < N1; i1++) {
        match12[i1] = -1;
        const int idx2 = vnMatch1[i1];
        if (idx2 >= 0) {
            const int idx1 = vnMatch2[idx2];
            if (idx1 == i1) {
                match12[i1] = idx2;
                nFound++;
            }
        }
    }
    return nFound;
}

// src/ORBmatcher.cc:969-1210 and :1212-1401.  pair_ok: the camera model's epipolarConstrain /
// matchAndtriangulate as a table lookup supplied by the test (row-major KF1.n x KF2.n bytes), NULL with
// F12 != NULL: Pinhole::epipolarConstrain (src/CameraModels/Pinhole.cpp:129-143), both NULL: bCoarse.
int orbo_search_for_triangulation(const BowView* K1, const uint8_t* has_mp1, const uint8_t* stereo1, const BowView* K2,
                                  const uint8_t* has_mp2, const uint8_t* stereo2, int bOnlyStereo, const float* ep,
                                  const float* mvScaleFactors2, const uint8_t* pair_ok, const float* F12,
                                  const float* mvLevelSigma2_2, int mbCheckOrientation, int32_t* vMatches12) {
    const FeatVec fv1 = feat_vec_of(*K1), fv2 = feat_vec_of(*K2);
    int nmatches = 0;
    std::vector<bool> vbMatched2(K2->n, false);
    for (int i = 0; i < K1->n; ++i) vMatches12[i] = -1;
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    FeatVec::const_iterator f1it = fv1.begin(), f2it = fv2.begin();
    while (f1it != fv1.end() && f2it != fv2.end()) {
        if (f1it->first == f2it->first) {
            for (size_t i1 = 0; i1 < f1it->second.size(); i1++) {
                const size_t idx1 = f1it->second[i1];
                if (has_mp1[idx1]) continue;
                const bool bStereo1 = stereo1 && stereo1[idx1];
                if (bOnlyStereo)
                    if (!bStereo1) continue;
                const Kp& kp1 = K1->keys[idx1];
                const uint8_t* d1 = K1->desc + idx1 * 32;
                int bestDist = TH_LOW, bestIdx2 = -1;
                for (size_t i2 = 0; i2 < f2it->second.size(); i2++) {
                    const size_t idx2 = f2it->second[i2];
                    if (vbMatched2[idx2] || has_mp2[idx2]) continue;
                    const bool bStereo2 = stereo2 && stereo2[idx2];
                    if (bOnlyStereo)
                        if (!bStereo2) continue;
                    const int dist = descriptor_distance(d1, K2->desc + idx2 * 32);
                    if (dist > TH_LOW || dist > bestDist) continue;
                    const Kp& kp2 = K2->keys[idx2];
                    if (!bStereo1 && !bStereo2 && ep) {
                        const float distex = ep[0] - kp2.x, distey = ep[1] - kp2.y;
                        if (distex * distex + distey * distey < 100 * mvScaleFactors2[kp2.octave]) continue;
                    }
                    bool ok = true;
                    if (pair_ok) {
                        ok = pair_ok[idx1 * (size_t)K2->n + idx2] != 0;
                    } else if (F12) {
                        const float a = kp1.x * F12[0] + kp1.y * F12[3] + F12[6];
                        const float b = kp1.x * F12[1] + kp1.y * F12[4] + F12[7];
                        const float c = kp1.x * F12[2] + kp1.y * F12[5] + F12[8];
                        const float num = a * kp2.x + b * kp2.y + c;
                        const float den = a * a + b * b;
                        if (den == 0) {
                            ok = false;
                        } else {
                            const float dsqr = num * num / den;
                            ok = dsqr < 3.84 * mvLevelSigma2_2[kp2.octave];
                        }
                    }
                    if (ok) {
                        bestIdx2 = (int)idx2;
                        bestDist = dist;
                    }
                }
                if (bestIdx2 >= 0) {
                    const Kp& kp2 = K2->keys[bestIdx2];
                    vMatches12[idx1] = bestIdx2;
                    nmatches++;
                    if (mbCheckOrientation) {
                        float rot = kp1.angle - kp2.angle;
                        if (rot < 0.0) rot += 360.0f;
                        int bin = round(rot * factor);
                        if (bin == HISTO_LENGTH) bin = 0;
                        rotHist[bin].push_back((int)idx1);
                    }
                }
            }
            f1it++;
            f2it++;
        } else if (f1it->first < f2it->first) {
            f1it = fv1.lower_bound(f2it->first);
        } else {
            f2it = fv2.lower_bound(f1it->first);
        }
    }
    if (mbCheckOrientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (size_t j = 0; j < rotHist[i].size(); j++) {
                vMatches12[rotHist[i][j]] = -1;
                nmatches--;
            }
        }
    }
    return nmatches;
}

// src/MapPoint.cc:330-404 for a batch of map points (group g = rows group_ptr[g] .. group_ptr[g+1]).
int orbo_distinctive_descriptors(const uint8_t* desc, const int32_t* group_ptr, int ngroups, int32_t* best) {
    for (int g = 0; g < ngroups; ++g) {
        const uint8_t* vDescriptors = desc + (size_t)group_ptr[g] * 32;
        const size_t N = (size_t)(group_ptr[g + 1] - group_ptr[g]);
        if (N == 0) {
            best[g] = -1;
            continue;
        }
        std::vector<float> Distances(N * N);
        for (size_t i = 0; i < N; i++) {
            Distances[i * N + i] = 0;
            for (size_t j = i + 1; j < N; j++) {
                int distij = descriptor_distance(vDescriptors + i * 32, vDescriptors + j * 32);
                Distances[i * N + j] = distij;
                Distances[j * N + i] = distij;
            }
        }
        int BestMedian = INT_MAX, BestIdx = 0;
        for (size_t i = 0; i < N; i++) {
            std::vector<int> vDists(Distances.begin() + i * N, Distances.begin() + i * N + N);
            std::sort(vDists.begin(), vDists.end());
            int median = vDists[0.5 * (N - 1)];
            if (median < BestMedian) {
                BestMedian = median;
                BestIdx = (int)i;
            }
        }
        best[g] = BestIdx;
    }
    return 0;
}

// DBoW2 TemplatedVocabulary<FORB>::transform(features, v, fv, levelsup) with its per-feature descent
// (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1139-1217, 1230-1275; BowVector.cpp addWeight / addIfNotExist /
// normalize; FeatureVector.cpp addFeature) on the POD vocabulary of orbm_vocab_create.
int orbo_bow_transform(int n_nodes, const int32_t* child_ptr, const uint32_t* child_ids, const uint8_t* node_desc,
                       const double* weight, const uint32_t* word_id_of, int m_L, int m_weighting, int norm_kind,
                       const uint8_t* features, int n, int levelsup, uint32_t* bow_ids, double* bow_vals,
                       int32_t* n_words, uint32_t* fv_node_ids, int32_t* fv_ptr, uint32_t* fv_items,
                       int32_t* n_fv_nodes) {
    std::map<unsigned int, double> v;                       // BowVector
    std::map<unsigned int, std::vector<unsigned int> > fv;  // FeatureVector
    const bool must = norm_kind != 0;
    for (int i_feature = 0; i_feature < n; ++i_feature) {
        const uint8_t* feature = features + (size_t)i_feature * 32;
        unsigned int nid = 0;
        const int nid_level = m_L - levelsup;
        unsigned int final_id = 0;
        int current_level = 0;
        do {
            ++current_level;
            const uint32_t* nodes = child_ids + child_ptr[final_id];
            const int nchild = child_ptr[final_id + 1] - child_ptr[final_id];
            final_id = nodes[0];
            double best_d = descriptor_distance(feature, node_desc + (size_t)final_id * 32);
            for (int c = 1; c < nchild; ++c) {
                const unsigned int id = nodes[c];
                double d = descriptor_distance(feature, node_desc + (size_t)id * 32);
                if (d < best_d) {
                    best_d = d;
                    final_id = id;
                }
            }
            if (current_level == nid_level) nid = final_id;
        } while (child_ptr[final_id + 1] > child_ptr[final_id]);
        const unsigned int id = word_id_of[final_id];
        const double w = weight[final_id];
        if (w > 0) {
            std::map<unsigned int, double>::iterator vit = v.lower_bound(id);
            if (vit != v.end() && !(id < vit->first)) {
                if (m_weighting == 0 || m_weighting == 1) vit->second += w;   // addWeight; addIfNotExist keeps it
            } else {
                v.insert(vit, std::make_pair(id, w));
            }
            fv[nid].push_back(i_feature);
        }
    }
    if ((m_weighting == 0 || m_weighting == 1) && !v.empty() && !must) {
        const double nd = v.size();
        for (std::map<unsigned int, double>::iterator vit = v.begin(); vit != v.end(); vit++) vit->second /= nd;
    }
    if (must) {
        double norm = 0.0;
        if (norm_kind == 1) {
            for (std::map<unsigned int, double>::iterator it = v.begin(); it != v.end(); ++it) norm += fabs(it->second);
        } else {
            for (std::map<unsigned int, double>::iterator it = v.begin(); it != v.end(); ++it) norm += it->second * it->second;
            norm = sqrt(norm);
        }
        if (norm > 0.0)
            for (std::map<unsigned int, double>::iterator it = v.begin(); it != v.end(); ++it) it->second /= norm;
    }
    int k = 0;
    for (std::map<unsigned int, double>::iterator it = v.begin(); it != v.end(); ++it, ++k) {
        bow_ids[k] = it->first;
        bow_vals[k] = it->second;
    }
    *n_words = k;
    int a = 0, items = 0;
    fv_ptr[0] = 0;
    for (std::map<unsigned int, std::vector<unsigned int> >::iterator it = fv.begin(); it != fv.end(); ++it) {
        fv_node_ids[a] = it->first;
        for (size_t j = 0; j < it->second.size(); ++j) fv_items[items++] = it->second[j];
        fv_ptr[++a] = items;
    }
    *n_fv_nodes = a;
    return 0;
}

}  // extern "C"
