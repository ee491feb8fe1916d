/* include/orbm.h -- C ABI of the B200-native descriptor-distance search (liborbx_b200.so).
 *
 * Drop-in boundary for the Hamming searches of ORB-SLAM3 (reference: zhoujoey/ORB-SLAM3-Mac; cites
 * are file:line relative to the reference tree).  Every ORBmatcher::Search* method walks pointer
 * graphs (Frame&, MapPoint*, KeyFrame*; include/ORBmatcher.h:46-88), so the seam is cut INSIDE the
 * methods: the caller snapshots the POD the method reads, all 256-bit distances are computed in
 * one batched CUDA launch, and the method's own order-dependent accept/reject logic is replayed
 * over those distances in the original order.
 *
 *   layer 1 (device primitives)
 *     orbm_hamming_csr   ORBmatcher::DescriptorDistance          src/ORBmatcher.cc:2353-2369
 *                        over every (query, candidate) pair of a CSR candidate list
 *     orbm_knn2          cv::BFMatcher(NORM_HAMMING).knnMatch(k=2) src/Frame.cc:43,1146
 *     orbm_stereo_rows   Frame::ComputeStereoMatches, Hamming part src/Frame.cc:802-885
 *   layer 2 (whole searches on POD snapshots; grid, candidate lists and distances on the GPU)
 *     orbm_search_by_projection_points   ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>&..)
 *                                        src/ORBmatcher.cc:48-218 (monocular/rectified branch)
 *     orbm_search_by_projection_frames   ORBmatcher::SearchByProjection(Frame&, const Frame&..)
 *                                        src/ORBmatcher.cc:1965-2181 (Nleft == -1 branch)
 *     orbm_search_for_initialization     ORBmatcher::SearchForInitialization src/ORBmatcher.cc:710-825
 *     orbm_search_by_bow / _bow_kf       SearchByBoW x2                     src/ORBmatcher.cc:273-475, 827-967
 *     orbm_search_by_projection_sim3     SearchByProjection(pKF, Scw, ..) x2  :477-590, 592-708
 *     orbm_search_by_projection_reloc    SearchByProjection(F, pKF, sAlreadyFound, ..)  :2183-2305
 *     orbm_fuse                          Fuse x2 (the search part)          :1403-1613, 1615-1737
 *     orbm_search_by_sim3                SearchBySim3                       :1739-1963
 *     orbm_search_for_triangulation      SearchForTriangulation x2          :969-1210, 1212-1401
 *     orbm_distinctive_descriptors       MapPoint::ComputeDistinctiveDescriptors  src/MapPoint.cc:330-404
 *     orbm_vocab_create / orbm_bow_transform   DBoW2 transform (Frame::ComputeBoW)  src/Frame.cc:729-736
 *     candidate generation               Frame::AssignFeaturesToGrid / PosInGrid / GetFeaturesInArea
 *                                        src/Frame.cc:377-408, 716-726, 645-714
 *
 * Plain C, POD only.  Contexts are independent and may be used concurrently from different
 * threads (Tracking, LocalMapping and LoopClosing all construct ORBmatcher objects); one context
 * is not re-entrant.  No CPU fallback: without a CUDA device orbm_create fails.
 */
#ifndef ORBM_H
#define ORBM_H

#include <stddef.h>
#include <stdint.h>

#include "orbx.h"

#ifdef __cplusplus
extern "C" {
#endif

#define ORBM_TH_HIGH 100     /* ORBmatcher::TH_HIGH, src/ORBmatcher.cc:40 */
#define ORBM_TH_LOW 50       /* ORBmatcher::TH_LOW,  :41 */
#define ORBM_HISTO_LENGTH 30 /* ORBmatcher::HISTO_LENGTH, :42 */
#define ORBM_GRID_COLS 64    /* FRAME_GRID_COLS, include/Frame.h:39 */
#define ORBM_GRID_ROWS 48    /* FRAME_GRID_ROWS, include/Frame.h:38 */

typedef struct orbm_context orbm_context;

int orbm_create(int device, orbm_context** out);
int orbm_destroy(orbm_context* c);

/* ---- layer 1 ---------------------------------------------------------------------------------- */
/* dist_out[k] = Hamming(q[row(k)], t[cand[k]]) for every k in [0, nnz); row_ptr has nq+1 entries,
 * row(k) is the r with row_ptr[r] <= k < row_ptr[r+1].  Order of candidates is preserved, which is
 * what lets the caller replay the reference's first-strictly-smaller-wins loops.  q, t: rows of 32
 * bytes (cv::Mat N x 32 CV_8U, 4-byte aligned as DescriptorDistance requires). */
int orbm_hamming_csr(orbm_context* c, const uint8_t* q, int nq, const uint8_t* t, int nt,
                     const int32_t* row_ptr, const int32_t* cand, int nnz, uint16_t* dist_out);

/* Two nearest train rows per query row, smallest distance first, LOWEST train index first on ties
 * (verified against cv2 4.13).  idx_out / dist_out: nq x 2; with nt < 2 the missing
 * entries are idx -1, dist -1 (the reference then fails its size()>=2 test, src/Frame.cc:1153). */
int orbm_knn2(orbm_context* c, const uint8_t* q, int nq, const uint8_t* t, int nt, int32_t* idx_out,
              int32_t* dist_out);

/* Rectified stereo (src/Frame.cc:812-885): for each left keypoint the right keypoint with the
 * smallest distance < TH_HIGH among those whose row band [floor(y-r), ceil(y+r)], r = 2*scale[octave],
 * contains row (int)yL, whose octave is within +-1 and whose x lies in [xL - max_d, xL - min_d];
 * ties -> lowest right index.  best_idx[i] = -1 / best_dist[i] = TH_HIGH when there is none. */
int orbm_stereo_rows(orbm_context* c, const orbx_keypoint* kps_l, const uint8_t* desc_l, int n_l,
                     const orbx_keypoint* kps_r, const uint8_t* desc_r, int n_r,
                     const float* scale_factors, int nlevels, float min_d, float max_d,
                     int32_t* best_idx, int32_t* best_dist);

/* Frame::ComputeStereoMatches, complete (src/Frame.cc:802-980): the row-band
 * Hamming search above with minD = 0, maxD = mbf / mb (:831-833), then for matches better than
 * thOrbDist = (TH_HIGH + TH_LOW) / 2 the 11x11 SAD search over +-5 px on level kpL.octave of BOTH
 * extractors' pyramids (:888-934; the reference reads mvImagePyramid, a buffer with a 19-px
 * BORDER_REFLECT_101 border -- here the two DEVICE-RESIDENT pyramids from orbx_device_pyramid, the
 * reflection applied on the fly), the parabola fit (:936-947), the disparity gate (:950-963) and the
 * median-based SAD outlier cut (:966-979).  u_right / depth: mvuRight / mvDepth (n_l floats, -1 =
 * no stereo match).  The pyramid views must come from handles on the context's device. */
int orbm_compute_stereo_matches(orbm_context* c, const orbx_keypoint* kps_l, const uint8_t* desc_l, int n_l,
                                const orbx_keypoint* kps_r, const uint8_t* desc_r, int n_r,
                                const float* scale_factors, const float* inv_scale_factors, int nlevels,
                                const orbx_pyramid_view* pyr_l, const orbx_pyramid_view* pyr_r, float mb,
                                float mbf, float* u_right, float* depth);

/* ---- layer 2 ---------------------------------------------------------------------------------- */
/* The parts of ORB_SLAM3::Frame the searches read (include/Frame.h). */
typedef struct orbm_frame_view {
    const orbx_keypoint* keys; /* mvKeysUn (N records) */
    const uint8_t* desc;       /* mDescriptors, N x 32 */
    int32_t n;                 /* N */
    float min_x, min_y, max_x, max_y; /* mnMinX, mnMinY, mnMaxX, mnMaxY */
    float grid_w_inv, grid_h_inv;     /* mfGridElementWidthInv / HeightInv */
    const float* scale_factors;       /* mvScaleFactors */
    int32_t nlevels;
    const float* u_right;      /* mvuRight (N) or NULL when not stereo */
} orbm_frame_view;

/* One local map point as SearchByProjection(Frame&, vector<MapPoint*>&) sees it after
 * Frame::isInFrustum filled the tracking fields (src/Frame.cc:483-572). */
typedef struct orbm_map_point {
    float proj_x, proj_y;   /* mTrackProjX, mTrackProjY */
    float proj_xr;          /* mTrackProjXR */
    float view_cos;         /* mTrackViewCos */
    float depth;            /* mTrackDepth */
    int32_t level;          /* mnTrackScaleLevel */
    int32_t observations;   /* Observations() */
    uint8_t in_view;        /* mbTrackInView */
    uint8_t bad;            /* isBad() */
    uint8_t _pad[2];
} orbm_map_point;

/* SearchByProjection(F, vpMapPoints, th, bFarPoints, thFarPoints), src/ORBmatcher.cc:48-147.
 * mp_desc: M x 32 (MapPoint::GetDescriptor()).  kp_obs (in/out, N ints): Observations() of the map
 * point currently attached to keypoint i, or -1 if none (F.mvpMapPoints[i] == NULL).
 * kp_match (out, N ints): index of the map point this call attached to keypoint i, else -1.
 * Returns nmatches (>= 0) or a negative orbx_status. */
int orbm_search_by_projection_points(orbm_context* c, const orbm_frame_view* F,
                                     const orbm_map_point* mps, const uint8_t* mp_desc, int m,
                                     float th, int far_points, float th_far, float nn_ratio,
                                     int32_t* kp_obs, int32_t* kp_match);

/* One last-frame observation for SearchByProjection(CurrentFrame, LastFrame, th, bMono),
 * src/ORBmatcher.cc:1965-2076: the caller projects the map point with the current pose
 * (camera model stays on the host) and passes the result. */
typedef struct orbm_last_obs {
    float u, v;           /* CurrentFrame.mpCamera->project(x3Dc) */
    float inv_z;          /* 1 / x3Dc.z */
    float angle;          /* kpLF.angle */
    int32_t octave;       /* nLastOctave */
    int32_t observations; /* pMP->Observations() */
    uint8_t valid;        /* LastFrame.mvpMapPoints[i] != NULL && !mvbOutlier[i] */
    uint8_t _pad[3];
} orbm_last_obs;

/* forward / backward: bForward / bBackward (:1985-1986); bf = CurrentFrame.mbf.
 * kp_obs / kp_match as above (kp_match[i] = index into `last`).  Applies the rotation-histogram
 * consistency check when check_orientation != 0 (:2158-2178).  Returns nmatches. */
int orbm_search_by_projection_frames(orbm_context* c, const orbm_frame_view* cur,
                                     const orbm_last_obs* last, const uint8_t* last_desc, int n_last,
                                     float th, int forward, int backward, float bf,
                                     int check_orientation, int32_t* kp_obs, int32_t* kp_match);

/* ---- fisheye stereo frames (Frame::Nleft != -1; KannalaBrandt8 rigs, config C3) ------------------
 * The frame holds the left keypoints in `F` (keys = mvKeys, n = Nleft, u_right = NULL) and the right
 * camera's half here; descriptor rows, mvpMapPoints and therefore kp_obs / kp_match are indexed
 * [0, Nleft) for the left and Nleft + j for right keypoint j, as in the reference. */
typedef struct orbm_frame_right {
    const orbx_keypoint* keys;     /* mvKeysRight (n records) */
    const uint8_t* desc;           /* mDescriptors rows Nleft.. (n x 32) */
    int32_t n;                     /* Nright */
    const int32_t* left_to_right;  /* mvLeftToRightMatch (Nleft ints) or NULL = all -1 */
    const int32_t* right_to_left;  /* mvRightToLeftMatch (Nright ints) or NULL = all -1 */
} orbm_frame_right;

/* The right-camera tracking fields of a map point (Frame::isInFrustum, src/Frame.cc:548-569). */
typedef struct orbm_map_point_right {
    float proj_x, proj_y;  /* mTrackProjXR, mTrackProjYR */
    float view_cos;        /* mTrackViewCosR */
    int32_t level;         /* mnTrackScaleLevelR (-1: none) */
    uint8_t in_view;       /* mbTrackInViewR */
    uint8_t _pad[3];
} orbm_map_point_right;

/* SearchByProjection(F, vpMapPoints, th, bFarPoints, thFarPoints) for a fisheye stereo frame:
 * left branch src/ORBmatcher.cc:66-147 (no mvuRight gate; a match is mirrored onto
 * mvLeftToRightMatch), right branch :149-215 (window in mGridRight, radius NOT scaled by th,
 * mirrored onto mvRightToLeftMatch).  kp_obs / kp_match: Nleft + Nright ints. */
int orbm_search_by_projection_points_stereo(orbm_context* c, const orbm_frame_view* F, const orbm_frame_right* R,
                                            const orbm_map_point* mps, const orbm_map_point_right* mps_r,
                                            const uint8_t* mp_desc, int m, float th, int far_points,
                                            float th_far, float nn_ratio, int32_t* kp_obs, int32_t* kp_match);

/* SearchByProjection(CurrentFrame, LastFrame, th, bMono) for a fisheye stereo current frame
 * (src/ORBmatcher.cc:1965-2181): uv_right = 2 floats per observation, the projection of
 * mTrl * x3Dc (:2090-2092).  kp_obs / kp_match: Nleft + Nright ints. */
int orbm_search_by_projection_frames_stereo(orbm_context* c, const orbm_frame_view* cur, const orbm_frame_right* R,
                                            const orbm_last_obs* last, const float* uv_right,
                                            const uint8_t* last_desc, int n_last, float th, int forward,
                                            int backward, int check_orientation, int32_t* kp_obs,
                                            int32_t* kp_match);

/* ---- SearchByBoW(KeyFrame*, Frame&, vpMapPointMatches), src/ORBmatcher.cc:273-475 --
 * Called on the tracking thread (src/Tracking.cc:1760, 2665).  One side of the search: descriptors,
 * keypoints (only .angle is read) and the DBoW2::FeatureVector (node id -> feature indices) as CSR.
 * Fisheye rigs pass the concatenation left ++ right, as mDescriptors is laid out. */
typedef struct orbm_bow_view {
    const orbx_keypoint* keys; /* n records */
    const uint8_t* desc;       /* n x 32 */
    int32_t n;
    int32_t nnodes;
    const uint32_t* node_ids;   /* FeatureVector keys, strictly ascending (nnodes) */
    const int32_t* node_ptr;    /* nnodes + 1 offsets into node_items */
    const uint32_t* node_items; /* feature indices of each node in the FeatureVector's order */
} orbm_bow_view;

/* kf_mp_ok[i] != 0: keyframe feature i has a map point that is not bad (:302-308).  f_nleft: F.Nleft
 * (-1 for mono / rectified stereo).  matches (out, F.n ints): index of the keyframe feature whose map
 * point was attached to frame feature j, else -1.  Returns nmatches.  The distances of all
 * (keyframe feature, frame feature) pairs that share a vocabulary node are computed on the device in
 * one launch; the accept loop, which depends on earlier acceptances (:323-324), runs on the host. */
int orbm_search_by_bow(orbm_context* c, const orbm_bow_view* KF, const uint8_t* kf_mp_ok, const orbm_bow_view* F,
                       int f_nleft, float nn_ratio, int check_orientation, int32_t* matches);

/* ---- SearchByBoW(KeyFrame*, KeyFrame*, vpMatches12), src/ORBmatcher.cc:827-967 (LoopClosing) -----------
 * mp_ok1 / mp_ok2 [i] != 0: feature i of that keyframe has a map point that is not bad (:868-872,
 * :892-898).  n_un1 / n_un2: pKF->mvKeysUn.size() when pKF->NLeft != -1 (features with a larger index
 * are skipped, :863-865, :887-889), else -1.  matches12 (out, KF1->n ints): index of the KF2 feature
 * whose map point is assigned to KF1 feature i, else -1.  Returns nmatches. */
int orbm_search_by_bow_kf(orbm_context* c, const orbm_bow_view* KF1, const uint8_t* mp_ok1, int n_un1,
                          const orbm_bow_view* KF2, const uint8_t* mp_ok2, int n_un2, float nn_ratio,
                          int check_orientation, int32_t* matches12);

/* ---- searches that project map points into a keyframe / frame (LoopClosing, LocalMapping, reloc) ---
 * One map point as those loops see it after their geometric gates (Sim3 / pose transform, positive
 * depth, camera projection, IsInImage, distance-invariance range, viewing angle, PredictScale): the
 * camera models and the map stay on the host, the caller passes the result. */
typedef struct orbm_proj_point {
    float u, v;      /* projection into the target image */
    float ur;        /* Fuse: u - bf * invz (src/ORBmatcher.cc:1490); unused elsewhere */
    float angle;     /* reloc SearchByProjection: pKF->mvKeysUn[i].angle (:2267); unused elsewhere */
    int32_t level;   /* nPredictedLevel */
    uint8_t valid;   /* 0: a gate before GetFeaturesInArea rejected the point (it is skipped) */
    uint8_t _pad[3];
} orbm_proj_point;

/* SearchByProjection(pKF, Scw, vpPoints, vpMatched, th, ratioHamming), src/ORBmatcher.cc:477-590, and
 * its overload with vpPointsKFs / vpMatchedKF, :592-708 (same search; the caller maps kp_match back to
 * vpPointsKFs).  Window = KeyFrame::GetFeaturesInArea(u, v, th * scale[level]) (src/KeyFrame.cc:770-814),
 * keypoint octave in [level-1, level], skip keypoints with vpMatched[idx] != NULL, smallest distance
 * (first wins), accepted when bestDist <= TH_LOW * ratioHamming.
 * kp_taken (in, KF->n bytes): vpMatched[idx] != NULL on entry.  kp_match (out, KF->n ints): index of the
 * point this call assigned to keypoint idx, else -1.  Returns nmatches. */
int orbm_search_by_projection_sim3(orbm_context* c, const orbm_frame_view* KF, const orbm_proj_point* pts,
                                   const uint8_t* mp_desc, int m, int th, float ratio_hamming,
                                   const uint8_t* kp_taken, int32_t* kp_match);

/* SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist), src/ORBmatcher.cc:2183-2305
 * (relocalisation).  Point i = map point of keyframe feature i (valid = exists, not bad, not already
 * found, inside the image bounds, inside its distance range).  Window with levels [level-1, level+1],
 * skip keypoints with CurrentFrame.mvpMapPoints[i2] != NULL (kp_taken), accept bestDist <= orb_dist,
 * rotation-histogram check on pts[i].angle - F.keys[i2].angle.  kp_match as above. */
int orbm_search_by_projection_reloc(orbm_context* c, const orbm_frame_view* F, const orbm_proj_point* pts,
                                    const uint8_t* mp_desc, int m, float th, int orb_dist,
                                    int check_orientation, const uint8_t* kp_taken, int32_t* kp_match);

/* The search of ORBmatcher::Fuse, both overloads: Fuse(pKF, vpMapPoints, th, bRight)
 * src/ORBmatcher.cc:1403-1613 (inv_level_sigma2 != NULL: the chi-square reprojection gates :1543-1570,
 * stereo form where KF->u_right[idx] >= 0) and Fuse(pKF, Scw, vpPoints, th, vpReplacePoint) :1615-1737
 * (inv_level_sigma2 == NULL).  For bRight the caller passes the right camera's view (keys =
 * mvKeysRight, desc = rows NLeft.., u_right as the reference indexes it) and idx_offset = NLeft.
 * best_idx[i] = keypoint (+ idx_offset) with the smallest distance among the gated candidates when that
 * distance is <= TH_LOW, else -1; best_dist[i] likewise (256 when none).  The map-graph updates
 * (Replace / AddObservation / AddMapPoint, :1588-1604) stay with the caller, in point order: the search
 * itself does not depend on them.  Returns nFused (the number of points with best_idx >= 0). */
int orbm_fuse(orbm_context* c, const orbm_frame_view* KF, const orbm_proj_point* pts, const uint8_t* mp_desc,
              int m, float th, const float* inv_level_sigma2, int idx_offset, int32_t* best_idx,
              int32_t* best_dist);

/* SearchBySim3(pKF1, pKF2, vpMatches12, s12, R12, t12, th), src/ORBmatcher.cc:1739-1963.
 * pts1[i] = map point of KF1 feature i projected into KF2 (valid = exists, not already matched, not
 * bad, passed the gates :1790-1823), mp_desc1 = those points' descriptors (n1 x 32); pts2 / mp_desc2 the
 * reverse direction (:1871-1935).  Each direction keeps the best keypoint (octave in [level-1, level])
 * when its distance is <= TH_HIGH; a pair is accepted when both directions agree (:1941-1957).
 * match12 (out, n1 ints): index of the agreeing KF2 feature, else -1.  Returns nFound. */
int orbm_search_by_sim3(orbm_context* c, const orbm_frame_view* KF1, const orbm_frame_view* KF2,
                        const orbm_proj_point* pts1, const uint8_t* mp_desc1, const orbm_proj_point* pts2,
                        const uint8_t* mp_desc2, float th, int32_t* match12);

/* The camera model's pair test of SearchForTriangulation: GeometricCamera::epipolarConstrain
 * (src/ORBmatcher.cc:1112) or matchAndtriangulate (:1316) for KF1 feature idx1 and KF2 feature idx2;
 * non-zero = the pair is geometrically consistent.  It must be a pure predicate of (idx1, idx2): the
 * library calls it for a row's candidates in order of preference (nearest first) until one passes, which
 * is a subset of the pairs the reference's scan would test, in another order. */
typedef int (*orbm_pair_test)(void* user, int32_t idx1, int32_t idx2);

/* SearchForTriangulation, src/ORBmatcher.cc:969-1210 (epipolarConstrain, bOnlyStereo, bCoarse) and
 * :1212-1401 (matchAndtriangulate).  has_mp1 / has_mp2 [i] != 0: pKF->GetMapPoint(i) != NULL.
 * stereo1 / stereo2 (may be NULL = all 0): bStereo of each feature (:1046, :1073).  epipole != NULL
 * enables the monocular epipole gate (:1096-1104; (ep.x, ep.y), tested against 100 * KF2 scale factor
 * of kp2.octave); pass NULL for rigs with a second camera and for the second overload.
 * Pair test, in this order: `test` (the caller's camera model) if not NULL; else F12 (row-major 3x3)
 * with level_sigma2_2 = pKF2->mvLevelSigma2: the Pinhole::epipolarConstrain arithmetic
 * (src/CameraModels/Pinhole.cpp:129-143); else every pair passes (bCoarse).
 * matches12 (out, KF1->n ints): vMatches12.  Returns nmatches. */
int orbm_search_for_triangulation(orbm_context* c, const orbm_bow_view* KF1, const uint8_t* has_mp1,
                                  const uint8_t* stereo1, const orbm_bow_view* KF2, const uint8_t* has_mp2,
                                  const uint8_t* stereo2, int only_stereo, const float* epipole,
                                  const float* scale_factors2, int nlevels2, orbm_pair_test test, void* user,
                                  const float* F12, const float* level_sigma2_2, int check_orientation,
                                  int32_t* matches12);

/* MapPoint::ComputeDistinctiveDescriptors, src/MapPoint.cc:330-404, for a batch of map points: group g
 * owns the observed descriptors desc[group_ptr[g] .. group_ptr[g+1]) (rows of 32 bytes, in the order of
 * the reference's vDescriptors).  best[g] = row (relative to the group) with the least median distance
 * to the others -- median = sorted row of the N x N distance matrix at index (int)(0.5 * (N - 1)),
 * first minimum wins -- or -1 for an empty group.  All distances, medians and the argmin run on the
 * device (one warp per map point). */
int orbm_distinctive_descriptors(orbm_context* c, const uint8_t* desc, const int32_t* group_ptr, int ngroups,
                                 int32_t* best);

/* ---- DBoW2 transform: Frame::ComputeBoW src/Frame.cc:729-736, KeyFrame::ComputeBoW ----
 * DBoW2::TemplatedVocabulary<FORB::TDescriptor, FORB> as POD (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h):
 * node i = m_nodes[i] (node 0 = root); its children are child_ids[child_ptr[i] .. child_ptr[i+1]) in the order of
 * Node::children; a node without children is a word with weight[i] (Node::weight) and word_id[i].  node_desc:
 * n_nodes x 32 (Node::descriptor; the root's row is unused).  L = m_L.  weighting: 0 TF_IDF, 1 TF, 2 IDF,
 * 3 BINARY (WeightingType); norm: what ScoringObject::mustNormalize returns -- 0 none, 1 L1, 2 L2 (ORBvoc.txt:
 * k = 10, L = 6, TF_IDF, L1_NORM -> weighting 0, norm 1).  The tree stays resident on the context's device. */
typedef struct orbm_vocabulary orbm_vocabulary;
int orbm_vocab_create(orbm_context* c, int n_nodes, const int32_t* child_ptr, const uint32_t* child_ids,
                      const uint8_t* node_desc, const double* weight, const uint32_t* word_id, int L, int weighting,
                      int norm, orbm_vocabulary** out);
int orbm_vocab_destroy(orbm_vocabulary* v);

/* transform(features, v, fv, levelsup), TemplatedVocabulary.h:1139-1217: every feature descends the tree on the
 * device (first child with the least FORB::distance per level, :1246-1266); the BowVector (word id -> summed
 * weight, then normalised) and the FeatureVector (node at level L - levelsup -> feature indices) are assembled in
 * feature order as the reference does, so the double sums are identical.  Outputs (capacity n each, fv_ptr n + 1):
 * bow_ids ascending with bow_vals, *n_words; fv_node_ids ascending with the CSR fv_ptr / fv_items, *n_fv_nodes --
 * the layout orbm_bow_view takes. */
int orbm_bow_transform(orbm_context* c, const orbm_vocabulary* v, const uint8_t* desc, int n, int levelsup,
                       uint32_t* bow_ids, double* bow_vals, int32_t* n_words, uint32_t* fv_node_ids,
                       int32_t* fv_ptr, uint32_t* fv_items, int32_t* n_fv_nodes);

/* SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize), src/ORBmatcher.cc:710-825.
 * prev_matched: F1.n (x, y) pairs, in/out.  matches12: F1.n ints, out.  Returns nmatches. */
int orbm_search_for_initialization(orbm_context* c, const orbm_frame_view* F1,
                                   const orbm_frame_view* F2, float* prev_matched, int32_t* matches12,
                                   int window_size, float nn_ratio, int check_orientation);

/* One window query of Frame::GetFeaturesInArea(x, y, r, minLevel, maxLevel), src/Frame.cc:645. */
typedef struct orbm_area_query {
    float x, y, r;
    int32_t min_level, max_level; /* as the reference's arguments (max_level < 0: no upper bound) */
    int32_t valid;                /* 0: the caller's gates rejected this query, it gets an empty row */
    int32_t desc_row;             /* row of the query descriptor matrix this query compares with */
} orbm_area_query;

/* Device-side AssignFeaturesToGrid + GetFeaturesInArea (src/Frame.cc:377-408, 645-726) for a batch
 * of queries: builds the frame's 64 x 48 grid on the GPU and returns the candidate lists as CSR
 * (row_ptr: nq + 1 ints; cand: keypoint indices in the reference's order) plus, when q_desc is not
 * NULL, the Hamming distance of every (query, candidate) pair (dist: one uint16 per candidate).
 * cap = capacity of cand / dist.  Returns nnz, or ORBX_E_CAPACITY with row_ptr filled when
 * cap < nnz.  The three searches below use exactly this path. */
int orbm_area_candidates(orbm_context* c, const orbm_frame_view* F, const orbm_area_query* q, int nq,
                         const uint8_t* q_desc, int n_q_desc, int32_t* row_ptr, int32_t* cand,
                         uint16_t* dist, int cap);

/* ---- device-resident frames ---------------------------------------------------------
 * A frame goes through two or three searches (TrackWithMotionModel's SearchByProjection, SearchLocalPoints,
 * relocalisation): an orbm_frame keeps its undistorted keypoints (mvKeysUn), descriptors and the 64 x 48 grid of
 * Frame::AssignFeaturesToGrid (src/Frame.cc:377-408) on the device, so a search uploads only its own queries.
 * Use: pass orbm_frame_get_view(f) wherever a search takes an orbm_frame_view* -- the library recognises the
 * address and skips the frame upload and the grid build.  Frames belong to their context (not to be shared between
 * contexts or used after orbm_destroy, which destroys the ones still alive). */
typedef struct orbm_frame orbm_frame;

/* Camera intrinsics and OpenCV distortion coefficients (Frame::mK, Frame::mDistCoef: k1 k2 p1 p2 [k3]). */
typedef struct orbm_camera {
    float fx, fy, cx, cy;
    float k1, k2, p1, p2, k3;
} orbm_camera;

/* From a host snapshot (the arrays are copied: they need not outlive the call). */
int orbm_frame_upload(orbm_context* c, const orbm_frame_view* view, orbm_frame** out);

/* From the extractor's DEVICE results (orbx_extract_device: d_keys = mvKeys, d_desc), without a round trip through
 * the host: Frame::UndistortKeyPoints (src/Frame.cc:738-771: cv::undistortPoints with K, the distortion and P = K;
 * skipped when cam is NULL or k1 == 0, as :740 does) and Frame::ComputeImageBounds (:773-800: the undistorted image
 * corners) run on the device, then the grid is built there.  The view's keys are mvKeysUn (host copy kept by the
 * frame: the accept loops read octaves), its bounds / grid pitch as the reference computes them, u_right NULL
 * (orbm_frame_set_u_right attaches mvuRight after the stereo matcher ran). */
int orbm_frame_from_device(orbm_context* c, const orbx_keypoint* d_keys, const uint8_t* d_desc, int n,
                           const orbm_camera* cam, int width, int height, const float* scale_factors, int nlevels,
                           orbm_frame** out);
const orbm_frame_view* orbm_frame_get_view(const orbm_frame* f);
int orbm_frame_set_u_right(orbm_frame* f, const float* u_right);
int orbm_frame_destroy(orbm_frame* f);

/* ---- device-resident descriptor rows ("banks") -------------------------------------------------------------
 * A MapPoint keeps its descriptor (MapPoint::mDescriptor, read through pMP->GetDescriptor() at
 * src/ORBmatcher.cc:80 and :2031) from one frame to the next, and tracking projects the same few thousand map points
 * into every frame (SearchLocalPoints -> :48; TrackWithMotionModel -> :1965, the last frame's map points).  A bank
 * keeps such 32-byte rows on the device: the *_bank searches below take, per query, the ROW of its descriptor instead
 * of the descriptor itself, so a call uploads its window queries only (32 bytes less per query, and no copy of them
 * on the host).  Results are identical to the host-descriptor entry points.
 *   orbm_bank_create / _write / _destroy : rows owned by the bank; _write (row = e.g. the map point's slot; call it
 *       where MapPoint::ComputeDistinctiveDescriptors stores a new mDescriptor, src/MapPoint.cc:402; the constructor
 *       from a frame copies the keypoint's row, :109) returns when
 *       the rows are in place and is ordered with the context's searches.
 *   orbm_frame_bank : a resident frame's descriptors as a bank (row = keypoint index; a map point created from a
 *       keypoint starts with that keypoint's descriptor); lives and dies with the frame.
 * `rows` = the bank row per query, NULL for row i.  A bank belongs to the context that made it (orbm_destroy frees
 * the banks that are still alive, as it does the frames). */
typedef struct orbm_bank orbm_bank;
int orbm_bank_create(orbm_context* c, int capacity_rows, orbm_bank** out);
int orbm_bank_write(orbm_bank* b, int first_row, int n, const uint8_t* desc);
int orbm_bank_rows(const orbm_bank* b);
int orbm_bank_destroy(orbm_bank* b);
const orbm_bank* orbm_frame_bank(const orbm_frame* f);

/* orbm_search_by_projection_points (src/ORBmatcher.cc:48-218) with the map points' descriptors in a bank. */
int orbm_search_by_projection_points_bank(orbm_context* c, const orbm_frame_view* F, const orbm_map_point* mps,
                                          const orbm_bank* bank, const int32_t* rows, int m, float th, int far_points,
                                          float th_far, float nn_ratio, int32_t* kp_obs, int32_t* kp_match);
/* orbm_search_by_projection_frames (src/ORBmatcher.cc:1965-2181) with the descriptors of the last frame's map
 * points in a bank (rows = the bank row of each observation's map point). */
int orbm_search_by_projection_frames_bank(orbm_context* c, const orbm_frame_view* cur, const orbm_last_obs* last,
                                          const orbm_bank* bank, const int32_t* rows, int n_last, float th, int forward,
                                          int backward, float bf, int check_orientation, int32_t* kp_obs,
                                          int32_t* kp_match);

/* ---- window search without any allocation ----------------------------------------------------------
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot
 * call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * The same computation as orbm_area_candidates -- grid build, per-query candidate lists in
 * GetFeaturesInArea order as CSR, the Hamming distance of every (query, candidate) pair -- but
 *   - every input and output is a DEVICE pointer the caller owns (keys: n records, desc: n x 32, q: nq
 *     queries, q_desc: the rows desc_row names, row_ptr: nq + 1 ints, cand / dist: cap entries);
 *   - the work is enqueued on `stream` (a cudaStream_t, NULL = the default stream) and the call returns
 *     without waiting; nnz = row_ptr[nq], pairs at positions >= cap are counted but not written;
 *   - the scratch (cell of every keypoint, the 64 x 48 grid, per-query counts) is ONE file-scope
 *     `static __device__` array of 274,688 bytes that every GPU loading the library holds until exit, so
 *     sizes are bounded: n <= ORBM_STATIC_MAX_KEYS, nq <= ORBM_STATIC_MAX_QUERIES, else ORBX_E_TOO_LARGE;
 *   - it calls no cudaMalloc / cudaFree / allocating library routine.  The first call on a device looks the
 *     array's device address up (cudaGetSymbolAddress): make that call outside a graph capture.
 * All calls on a device share the array and it keeps the previous call's contents (nothing in it has to
 * start at zero; each call overwrites the ranges it reads).  Calls that overlap on the GPU -- two streams, or
 * two captured graphs launched concurrently -- RACE on it: serialise them on one stream or with events. */
#define ORBM_STATIC_MAX_KEYS 16384
#define ORBM_STATIC_MAX_QUERIES 32768
int orbm_area_candidates_static(int device, void* stream, const orbx_keypoint* d_keys, const uint8_t* d_desc,
                                int n, float min_x, float min_y, float grid_w_inv, float grid_h_inv,
                                const orbm_area_query* d_q, int nq, const uint8_t* d_q_desc, int32_t* d_row_ptr,
                                int32_t* d_cand, uint16_t* d_dist, int cap);

/* Measurement hook: device time between the first and the last kernel of the previous call on this
 * context (CUDA events on its stream; for the searches this spans the candidate kernels and the
 * row_ptr read-back between them) and the number of descriptor pairs it compared. */
int orbm_last_device_ms(orbm_context* c, float* ms, long long* pairs);

/* Measurement hook: the device's POPC issue rate in 1e9 popc/s (8 independent xor-popc-add chains
 * per thread, all SMs), the peak the Hamming kernels are held against (8 popc per pair). */
int orbm_measure_popc_peak(orbm_context* c, double* gpopc_per_s);

#ifdef __cplusplus
}
#endif
#endif /* ORBM_H */
