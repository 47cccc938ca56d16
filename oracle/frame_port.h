/* frame_port.h -- TEST INFRASTRUCTURE (oracle).  CPU restatement of the grid-gated searches of SURVEY.md section 8(f2):
 *   Frame::AssignFeaturesToGrid / PosInGrid / GetFeaturesInArea   src/Frame.cc:230-245, 327-391
 *   ORBmatcher::SearchForInitialization                           src/ORBmatcher.cc:405-520
 *   ORBmatcher::SearchByProjection(Frame&, const Frame&, th, mono) src/ORBmatcher.cc:1328-1470
 *   ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>&, th) src/ORBmatcher.cc:44-133 (local-map search)
 *   ORBmatcher::SearchByBoW (both overloads)                      src/ORBmatcher.cc:159-288, 522-655
 *   ORBmatcher::SearchForTriangulation + CheckDistEpipolarLine    src/ORBmatcher.cc:657-820, 140-157
 *   Frame::isInFrustum + MapPoint::PredictScale (Tracking::SearchLocalPoints)  src/Frame.cc:269-325, src/MapPoint.cc:402-417
 *   ORBmatcher::SearchByProjection(Frame&, KeyFrame*, sAlreadyFound, th, ORBdist)  src/ORBmatcher.cc:1472-1597
 *   the Sim3 routines Fuse(KeyFrame*, Scw, ...), SearchByProjection(KeyFrame*, Scw, ...), SearchBySim3
 *                                                                 src/ORBmatcher.cc:969-1097, 290-403, 1102-1326
 *   the search of ORBmatcher::Fuse(KeyFrame*, vector<MapPoint*>, th) + KeyFrame::GetFeaturesInArea / IsInImage
 *                                                                 src/ORBmatcher.cc:825-957, src/KeyFrame.cc:569-613
 *   ORBmatcher::ComputeThreeMaxima                                src/ORBmatcher.cc:1601-1642
 * PINNED: the reference's own src/Frame.cc, src/ORBmatcher.cc, src/KeyFrame.cc and src/MapPoint.cc compile unchanged
 * against oracle/cvshim2 into oracle/_ref/libframeref.so (oracle/ref_frame_wrap.cc, recipe in oracle/Makefile), and
 * tests/test_ref_frame.py checks every function here bit-for-bit against them (real two-view scenes, randomised frames
 * with ties / validity / occupancy, and a knife-edge pose for the bForward flag).  The one library arithmetic on this
 * path -- the cv::Mat expressions "Rcw*x3Dw+tcw" and "-Rcw.t()*tcw" = cv::gemm's 3x3 float special case -- is pinned to
 * cv2 4.13.0 golden vectors (tests/golden/gemm_golden.npz); how cv::MatExpr lowers "-R.t()*t" (a transposed COPY times
 * -1, not GEMM_1_T) is restated from OpenCV's matrix_expressions.cpp, which cv2 cannot execute.
 * Only tests/ may use this file; the product never does. */
#ifndef ORACLE_FRAME_PORT_H
#define ORACLE_FRAME_PORT_H
#include <stdint.h>
#include "orb_port.h"
#ifdef __cplusplus
extern "C" {
#endif

/* same layout as orbgpu_projection_args (include/orbgpu.h) */
typedef struct frameport_projection_args {
    int32_t n_last;
    const float* world_pos;        /* n_last x 3: pMP->GetWorldPos() */
    const uint8_t* mp_desc;        /* n_last x 32: pMP->GetDescriptor() */
    const uint8_t* mp_valid;       /* LastFrame.mvpMapPoints[i] && !LastFrame.mvbOutlier[i] */
    const uint8_t* mp_observed;    /* pMP->Observations() > 0 */
    const int32_t* last_octave;    /* LastFrame.mvKeys[i].octave */
    const float* last_angle;       /* LastFrame.mvKeysUn[i].angle */
    const float* Tcw_last;         /* 4 x 4 row-major */
    int32_t n_cur;
    const orbport_keypoint* cur_keys_un;
    const uint8_t* cur_desc;
    const float* cur_u_right;      /* CurrentFrame.mvuRight */
    const uint8_t* cur_occupied;   /* CurrentFrame.mvpMapPoints[i2] && ->Observations() > 0 on entry */
    const float* Tcw_cur;
    float fx, fy, cx, cy, bf, b;
    float bounds[4];               /* mnMinX, mnMaxX, mnMinY, mnMaxY */
    const float* scale_factors;
    int32_t nlevels;
    float th;
    int32_t mono;
    int32_t check_orientation;
} frameport_projection_args;

/* ordered index list of GetFeaturesInArea on a grid built from keys_un; returns the count */
int frameport_features_in_area(const orbport_keypoint* keys_un, int n, const float* bounds4, float x, float y, float r,
                               int min_level, int max_level, int* out, int cap);
/* returns nmatches; matches12[n1]; prev_matched (n1 x 2) updated in place */
int frameport_search_for_initialization(const orbport_keypoint* k1, const uint8_t* d1, int n1, const orbport_keypoint* k2,
                                        const uint8_t* d2, int n2, const float* bounds4, float* prev_matched, int window_size,
                                        float nnratio, int check_orientation, int* matches12);
/* returns nmatches; cur_match[n_cur]: -1 untouched, -2 reset to NULL by the rotation check, >= 0 index into the last frame */
int frameport_search_by_projection(const frameport_projection_args* a, int* cur_match);

/* same layout as orbgpu_local_points_args (include/orbgpu.h): ORBmatcher::SearchByProjection(Frame&, const
 * vector<MapPoint*>&, th) (src/ORBmatcher.cc:44-125) on the fields Frame::isInFrustum left in the map points */
typedef struct frameport_local_points_args {
    int32_t n_points;
    const uint8_t* in_view;        /* pMP->mbTrackInView && !pMP->isBad() */
    const float* proj_x;
    const float* proj_y;
    const float* proj_xr;
    const int32_t* scale_level;
    const float* view_cos;
    const uint8_t* mp_desc;
    const uint8_t* mp_observed;    /* pMP->Observations() > 0 */
    int32_t n_cur;
    const orbport_keypoint* cur_keys_un;
    const uint8_t* cur_desc;
    const float* cur_u_right;
    const uint8_t* cur_occupied;
    float bounds[4];
    const float* scale_factors;
    int32_t nlevels;
    float th;
    float nnratio;
} frameport_local_points_args;
/* returns nmatches; cur_match[n_cur]: -1 untouched, i >= 0 = F.mvpMapPoints[idx] set to vpMapPoints[i] */
int frameport_search_local_points(const frameport_local_points_args* a, int* cur_match);

/* ORBmatcher::SearchByBoW, both overloads (src/ORBmatcher.cc:159-288 KeyFrame/Frame, :522-655 KeyFrame/KeyFrame).
 * Side 1 = the keyframe whose features are the queries (pKF / pKF1), side 2 = the searched one (F / pKF2).  A DBoW2
 * FeatureVector is passed flattened: ascending node ids, node_start[n_nodes + 1], feature indices in the map's order.
 * valid = map point present and not bad (all 1 for a Frame).  strict_th_low: 0 = "bestDist1 <= TH_LOW" (:241),
 * 1 = "bestDist1 < TH_LOW" (:600).  match12[n1] / match21[n2] = the accepted pairs after the rotation check
 * (-1 elsewhere): overload 1 fills vpMapPointMatches[idx2] = map point of match21[idx2], overload 2 fills
 * vpMatches12[idx1] = map point of match12[idx1].  Same layout as orbgpu_bow_side (include/orbgpu.h). */
typedef struct frameport_bow_side {
    const uint8_t* desc; int32_t n; const uint8_t* valid; const float* angle;
    const uint32_t* node_ids; const int32_t* node_start; const uint32_t* indices; int32_t n_nodes;
} frameport_bow_side;
int frameport_search_by_bow(const frameport_bow_side* s1, const frameport_bow_side* s2, float nnratio, int check_orientation,
                            int strict_th_low, int* match12, int* match21);

/* same layout as orbgpu_triangulation_side / orbgpu_triangulation_args (include/orbgpu.h):
 * ORBmatcher::SearchForTriangulation (src/ORBmatcher.cc:657-820) with CheckDistEpipolarLine (:140-157) */
typedef struct frameport_triangulation_side {
    const orbport_keypoint* keys_un; const uint8_t* desc; int32_t n; const uint8_t* has_map_point; const float* u_right;
    const uint32_t* node_ids; const int32_t* node_start; const uint32_t* indices; int32_t n_nodes;
} frameport_triangulation_side;
typedef struct frameport_triangulation_args {
    frameport_triangulation_side kf1, kf2;
    float F12[9]; float Cw1[3]; float T2w[16]; float fx, fy, cx, cy;
    const float* scale_factors2; const float* level_sigma2; int32_t nlevels; int32_t only_stereo; int32_t check_orientation;
} frameport_triangulation_args;
/* returns nmatches; matches12[kf1.n] = vMatches12 after the rotation check */
int frameport_search_for_triangulation(const frameport_triangulation_args* a, int* matches12);

/* same layout as orbgpu_local_map_args / orbgpu_track_fields (include/orbgpu.h): Tracking::SearchLocalPoints
 * (src/Tracking.cc:1150-1200) = Frame::isInFrustum per candidate point, then the local-map SearchByProjection */
typedef struct frameport_local_map_args {
    int32_t n_points;
    const uint8_t* candidate; const float* world_pos; const float* normal; const float* max_distance; const float* min_distance;
    const uint8_t* mp_desc; const uint8_t* mp_observed;
    const float* Tcw; float fx, fy, cx, cy, bf; float viewing_cos_limit; float log_scale_factor;
    int32_t n_cur; const orbport_keypoint* cur_keys_un; const uint8_t* cur_desc; const float* cur_u_right; const uint8_t* cur_occupied;
    float bounds[4]; const float* scale_factors; int32_t nlevels; float th; float nnratio;
} frameport_local_map_args;
typedef struct frameport_track_fields {
    uint8_t* in_view; float* proj_x; float* proj_y; float* proj_xr; int32_t* scale_level; float* view_cos;
} frameport_track_fields;
/* returns nmatches; *n_in_view = nToMatch */
int frameport_search_local_map(const frameport_local_map_args* a, const frameport_track_fields* out, int* n_in_view, int* cur_match);
/* glibc logf on an array (what MapPoint::PredictScale calls), for the device logf test */
void frameport_logf_array(const float* x, int n, float* out);
/* the device's logf algorithm restated on the CPU vs the host's logf over float bit patterns first + k*step, k < count */
long long frameport_logf_port_mismatches(uint32_t first, uint32_t count, uint32_t step);

/* same layout as orbgpu_fuse_args (include/orbgpu.h): per map point, the keypoint of pKF the loop body of ORBmatcher::Fuse
 * (src/ORBmatcher.cc:845-957) would fuse it with, or -1 */
typedef struct frameport_fuse_args {
    int32_t n_points; const uint8_t* present; const float* world_pos; const float* normal; const float* max_distance;
    const float* min_distance; const uint8_t* mp_desc; const float* Tcw; float Ow[3]; float fx, fy, cx, cy, bf; float log_scale_factor;
    int32_t n_kf; const orbport_keypoint* keys_un; const uint8_t* desc; const float* u_right; float bounds[4];
    const float* scale_factors; const float* inv_level_sigma2; int32_t nlevels; float th;
} frameport_fuse_args;
void frameport_fuse_search(const frameport_fuse_args* a, int* best_idx);

/* same layout as orbgpu_reloc_projection_args (include/orbgpu.h): ORBmatcher::SearchByProjection(CurrentFrame, pKF,
 * sAlreadyFound, th, ORBdist) (src/ORBmatcher.cc:1472-1597) */
typedef struct frameport_reloc_projection_args {
    int32_t n_kf; const uint8_t* mp_valid; const float* world_pos; const float* max_distance; const float* min_distance;
    const uint8_t* mp_desc; const float* kf_angle;
    int32_t n_cur; const orbport_keypoint* cur_keys_un; const uint8_t* cur_desc; const uint8_t* cur_occupied; const float* Tcw_cur;
    float fx, fy, cx, cy; float bounds[4]; const float* scale_factors; float log_scale_factor; int32_t nlevels; float th;
    int32_t orb_dist; int32_t check_orientation;
} frameport_reloc_projection_args;
/* returns nmatches; cur_match[n_cur]: -1 untouched, -2 reset by the rotation check, >= 0 keyframe feature index */
int frameport_search_by_projection_keyframe(const frameport_reloc_projection_args* a, int* cur_match);

/* MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:242-303) on groups of descriptors given in the order the
 * reference collects them; best_idx[g] = BestIdx, -1 for an empty group */
void frameport_distinctive_descriptors(const uint8_t* desc, const int32_t* group_start, int n_groups, int* best_idx);

/* the Sim3 routines on the frameport_fuse_args snapshot (see include/orbgpu.h for what each field means there) */
void frameport_fuse_search_sim3(const frameport_fuse_args* a, const float* Scw16, int* best_idx);
int frameport_search_by_projection_sim3(const frameport_fuse_args* a, const float* Scw16, const uint8_t* matched_in, int* matched);
int frameport_search_by_sim3(const frameport_fuse_args* a12, const frameport_fuse_args* a21, float s12, const float* R12,
                             const float* t12, int* match12);

/* The DBoW2 vocabulary (oracle/vocab_port.cc): transform(features, BowVector, FeatureVector, levelsup) and the L1 score on
 * flat arrays; the arguments mean what they mean in include/orbgpu.h (orbgpu_vocabulary_*). */
typedef struct frameport_vocab frameport_vocab;
frameport_vocab* frameport_vocab_create(int k, int L, int scoring, int weighting, int n_nodes, const int32_t* parent,
                                        const uint8_t* is_leaf, const uint8_t* desc, const double* weight);
void frameport_vocab_free(frameport_vocab* v);
void frameport_vocab_transform(const frameport_vocab* v, const uint8_t* desc, int n, int levelsup, uint32_t* words,
                               double* values, uint32_t* node_ids, int32_t* start, uint32_t* idx, int* counts);
void frameport_vocab_score(const uint32_t* qw, const double* qv, int nq, const uint32_t* kw, const double* kv,
                           const int32_t* kstart, int nk, double* out);

/* KeyFrameDatabase::DetectLoopCandidates / DetectRelocalizationCandidates (src/KeyFrameDatabase.cc:76-197, :199-309) on flat
 * arrays.  Keyframes are numbered in the order they were added to the database. */
typedef struct {
    int n_kf;
    const uint32_t* kw; const double* kv; const int32_t* kstart;   /* the keyframes' BowVectors (CSR, ascending words) */
    const uint8_t* erased;        /* [n_kf] 1 = KeyFrameDatabase::erase was called for it (it stays in the covisibility graph) */
    const uint32_t* qw; const double* qv; int nq;                  /* the query's BowVector */
    const uint8_t* connected;     /* [n_kf] loop detection: member of pKF->GetConnectedKeyFrames() */
    const int32_t* neigh_start; const int32_t* neigh;              /* mvpOrderedConnectedKeyFrames of every keyframe (CSR) */
    const float* prev_score;      /* [n_kf] mLoopScore / mRelocScore as an earlier query left it */
    float min_score;              /* loop detection */
    int reloc;                    /* 0 DetectLoopCandidates, 1 DetectRelocalizationCandidates */
} frameport_db_args;
/* marked[i] = the keyframe was put into lKFsSharingWords; words[i] = mnLoopWords / mnRelocWords (where marked);
 * score[i] = mLoopScore / mRelocScore after the call; returns the number of candidates written (keyframe numbers, in order) */
int frameport_db_detect(const frameport_db_args* a, int32_t* marked, int32_t* words, float* score, int32_t* candidates);

/* MapPoint::UpdateNormalAndDepth (src/MapPoint.cc:330-371), batched over map points.  Point g owns the camera centres
 * centers[3 * obs_start[g] .. 3 * obs_start[g + 1]) of its observing keyframes, in the order the reference walks them (its
 * std::map<KeyFrame*, size_t>); ref_center / ref_scale / last_scale: the reference keyframe's camera centre,
 * mvScaleFactors[octave of the point's keypoint there] and mvScaleFactors[mnScaleLevels - 1].  Points without observations
 * keep normal / max_dist / min_dist as they came in. */
void frameport_update_normal_and_depth(const float* pos, const int32_t* obs_start, const float* centers, const float* ref_center,
                                       const float* ref_scale, const float* last_scale, int n_points, float* normal,
                                       float* max_dist, float* min_dist);

#ifdef __cplusplus
}
#endif
#endif
