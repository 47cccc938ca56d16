// oracle/ref_frame_wrap.cc -- TEST INFRASTRUCTURE, not product code.
//
// C entry points around the UNMODIFIED reference translation units src/Frame.cc, src/ORBmatcher.cc, src/KeyFrame.cc,
// src/MapPoint.cc, src/ORBextractor.cc and Thirdparty/DBoW2/DBoW2/{FeatureVector,BowVector}.cpp, which oracle/Makefile
// compiles from where they lie under /root/reference against oracle/cvshim2 into oracle/_ref/libframeref.so.
// This is what pins the SURVEY section-8(f) rows: the restatements in oracle/frame_port.cc / orb_port.cc and the CUDA
// path are compared with what the reference's own code returns on the same inputs.
//
// The wrapper only builds the objects the reference functions expect (Frame / KeyFrame / MapPoint with their members
// filled from plain arrays -- the same arrays the C ABI takes, include/orbgpu.h) and reads the results back:
//   frameref_stereo_matches               Frame::ComputeStereoMatches                     src/Frame.cc:466-640
//   frameref_features_in_area             Frame::AssignFeaturesToGrid + GetFeaturesInArea src/Frame.cc:230-245, 327-380
//   frameref_search_for_initialization    ORBmatcher::SearchForInitialization             src/ORBmatcher.cc:405-520
//   frameref_search_by_projection         ORBmatcher::SearchByProjection(Frame&, const Frame&, th, mono)  :1328-1470
//   frameref_search_local_points          ORBmatcher::SearchByProjection(Frame&, const vector<MapPoint*>&, th)  :44-125
//   frameref_search_local_map             Frame::isInFrustum (src/Frame.cc:269-325) per point + that search, as Tracking::SearchLocalPoints
//   frameref_fuse_search                  ORBmatcher::Fuse(KeyFrame*, vector<MapPoint*>, th) :825-957, one map point per call
//   frameref_search_by_projection_keyframe  ORBmatcher::SearchByProjection(Frame&, KeyFrame*, sAlreadyFound, th, ORBdist) :1472-1597
//   frameref_fuse_search_sim3 / frameref_search_by_projection_sim3 / frameref_search_by_sim3   the Sim3 routines :969-1097, :290-403, :1102-1326
//   frameref_search_by_bow                ORBmatcher::SearchByBoW, both overloads         :159-288, :522-655
//   frameref_distinctive_descriptors      MapPoint::ComputeDistinctiveDescriptors         src/MapPoint.cc:242-303
//   frameref_search_for_triangulation     ORBmatcher::SearchForTriangulation + CheckDistEpipolarLine  :657-820, :140-157
//   frameref_undistort_keypoints          Frame::UndistortKeyPoints + ComputeImageBounds  src/Frame.cc:404-463
//   frameref_descriptor_distance / frameref_three_maxima   ORBmatcher::DescriptorDistance :1647-1663, ComputeThreeMaxima :1601-1642
//   frameref_gemm3 / frameref_neg_rt_t    the shim's cv::gemm and "-R.t()*t" evaluation (golden-vector check)
// Signatures mirror oracle/frame_port.h / orb_port.h so the tests run the same cases through both.
// Symbols of the reference that none of these paths reaches (Map, KeyFrameDatabase, Converter) are defined here as
// aborting stand-ins.  "private"/"protected" are opened for the reference headers in THIS file only (class layout is
// unchanged) so the harness can fill members such as MapPoint::mDescriptor without running the whole SLAM system.
#include <opencv2/core/core.hpp>
#include <algorithm>
#include <cstdio>
#include <map>
#include <set>
#include <mutex>
#include <thread>
#include <vector>

#define private public
#define protected public
#include "Frame.h"
#include "KeyFrame.h"
#include "KeyFrameDatabase.h"
#include "Map.h"
#include "MapPoint.h"
#include "ORBmatcher.h"
#include "ORBextractor.h"
#undef private
#undef protected

#include "frame_port.h"

using namespace ORB_SLAM2;

namespace {
[[noreturn]] void never(const char* what)
{
    std::fprintf(stderr, "oracle/ref_frame_wrap: %s is outside the tested paths\n", what);
    std::abort();
}
}  // namespace

// ---- reference symbols outside the tested paths ------------------------------------------------------------------------
namespace ORB_SLAM2 {
Map::Map() : mnMaxKFid(0), mnBigChangeIdx(0) {}
void Map::EraseMapPoint(MapPoint*) {}     // reached from MapPoint::Replace under Fuse; the harness owns the points
void Map::EraseKeyFrame(KeyFrame*) { never("Map::EraseKeyFrame"); }
}  // namespace ORB_SLAM2
#include "Converter.h"
// stand-in for src/Converter.cc:27-35 (that file needs Eigen and g2o): one single-row header per descriptor
std::vector<cv::Mat> ORB_SLAM2::Converter::toDescriptorVector(const cv::Mat& D)
{
    std::vector<cv::Mat> rows((size_t)D.rows);
    for (int r = 0; r < D.rows; ++r) rows[(size_t)r] = D.row(r);
    return rows;
}

namespace {

typedef orbport_keypoint KP;   // 28-byte cv::KeyPoint layout

std::vector<cv::KeyPoint> to_keys(const KP* k, int n)
{
    std::vector<cv::KeyPoint> v((size_t)n);
    static_assert(sizeof(cv::KeyPoint) == sizeof(KP), "cv::KeyPoint is a 28-byte POD");
    if (n) std::memcpy((void*)v.data(), k, sizeof(KP) * (size_t)n);
    return v;
}

cv::Mat desc_mat(const uint8_t* d, int n)
{
    cv::Mat m(n > 0 ? n : 1, 32, CV_8U);
    if (n > 0) std::memcpy(m.data, d, (size_t)n * 32);
    return n > 0 ? m : m.rowRange(0, 0);
}

cv::Mat pose_mat(const float* T16)
{
    cv::Mat m(4, 4, CV_32F);
    std::memcpy(m.data, T16, 16 * sizeof(float));
    return m;
}

// the static members Frame's constructors compute once (src/Frame.cc:96-115)
void set_frame_statics(const float* bounds4, float fx, float fy, float cx, float cy)
{
    Frame::mnMinX = bounds4[0]; Frame::mnMaxX = bounds4[1]; Frame::mnMinY = bounds4[2]; Frame::mnMaxY = bounds4[3];
    Frame::mfGridElementWidthInv = static_cast<float>(FRAME_GRID_COLS) / static_cast<float>(Frame::mnMaxX - Frame::mnMinX);
    Frame::mfGridElementHeightInv = static_cast<float>(FRAME_GRID_ROWS) / static_cast<float>(Frame::mnMaxY - Frame::mnMinY);
    Frame::fx = fx; Frame::fy = fy; Frame::cx = cx; Frame::cy = cy;
    Frame::invfx = 1.0f / fx; Frame::invfy = 1.0f / fy;
    Frame::mbInitialComputations = false;
}

// a Frame as its constructor leaves it, from plain arrays
void fill_frame(Frame& F, const KP* keys_un, const uint8_t* desc, int n)
{
    F.N = n;
    F.mvKeysUn = to_keys(keys_un, n);
    F.mvKeys = F.mvKeysUn;
    if (desc) F.mDescriptors = desc_mat(desc, n);
    F.mvuRight.assign((size_t)n, -1.f);
    F.mvDepth.assign((size_t)n, -1.f);
    F.mvpMapPoints.assign((size_t)n, static_cast<MapPoint*>(NULL));
    F.mvbOutlier.assign((size_t)n, false);
    F.AssignFeaturesToGrid();
}

void set_scale_tables(Frame& F, const float* scale_factors, int nlevels)
{
    F.mnScaleLevels = nlevels;
    F.mvScaleFactors.assign(scale_factors, scale_factors + nlevels);
    F.mvInvScaleFactors.resize((size_t)nlevels);
    F.mvLevelSigma2.resize((size_t)nlevels);
    F.mvInvLevelSigma2.resize((size_t)nlevels);
    for (int i = 0; i < nlevels; ++i) {
        F.mvInvScaleFactors[i] = 1.0f / scale_factors[i];
        F.mvLevelSigma2[i] = scale_factors[i] * scale_factors[i];
        F.mvInvLevelSigma2[i] = 1.0f / F.mvLevelSigma2[i];
    }
    F.mfScaleFactor = nlevels > 1 ? scale_factors[1] : 1.2f;
    F.mfLogScaleFactor = std::log(F.mfScaleFactor);
}

struct World {           // owns the MapPoints / KeyFrames of one call
    Map map;
    Frame anchor;        // reference frame handed to KeyFrame's constructor for the map points
    KeyFrame* anchor_kf;
    std::vector<MapPoint*> points;
    std::vector<KeyFrame*> keyframes;
    World() : anchor_kf(0)
    {
        anchor.N = 0;
        anchor.mTcw = cv::Mat::eye(4, 4, CV_32F);
        anchor.mnId = 0;
        anchor.mbf = 1.f; anchor.mb = 1.f; anchor.mThDepth = 1.f; anchor.mTimeStamp = 0;
        anchor.mnScaleLevels = 1; anchor.mfScaleFactor = 1.2f; anchor.mfLogScaleFactor = std::log(1.2f);
        anchor_kf = new KeyFrame(anchor, &map, static_cast<KeyFrameDatabase*>(NULL));
        keyframes.push_back(anchor_kf);
    }
    MapPoint* point(const float* xyz, const uint8_t* desc, int observations, bool bad)
    {
        cv::Mat pos(3, 1, CV_32F);
        const float zero[3] = {0, 0, 0};
        std::memcpy(pos.data, xyz ? xyz : zero, 3 * sizeof(float));
        MapPoint* p = new MapPoint(pos, anchor_kf, &map);      // src/MapPoint.cc:32-44
        if (desc) {
            cv::Mat d(1, 32, CV_8U);
            std::memcpy(d.data, desc, 32);
            p->mDescriptor = d;                                   // what ComputeDistinctiveDescriptors would leave
        }
        p->nObs = observations;
        p->mbBad = bad;
        points.push_back(p);
        return p;
    }
    ~World()
    {
        for (size_t i = 0; i < points.size(); ++i) delete points[i];
        for (size_t i = 0; i < keyframes.size(); ++i) delete keyframes[i];
    }
};

void fill_feature_vector(DBoW2::FeatureVector& fv, const frameport_bow_side* s)
{
    fv.clear();
    for (int n = 0; n < s->n_nodes; ++n)
        for (int k = s->node_start[n]; k < s->node_start[n + 1]; ++k) fv.addFeature(s->node_ids[n], s->indices[k]);
}

}  // namespace

extern "C" {

// d = alpha * op(A) * x + beta * c through the shim's cv::gemm (A 3x3, x and c 3x1, float); c may be NULL
void frameref_gemm3(const float* A, const float* x, const float* c, double alpha, double beta, int flags, float* d)
{
    cv::Mat mA(3, 3, CV_32F, (void*)A), mx(3, 1, CV_32F, (void*)x), mc, out;
    if (c) mc = cv::Mat(3, 1, CV_32F, (void*)c);
    cv::cv_gemm32f(mA, mx, alpha, c ? &mc : 0, beta, out, flags);
    for (int i = 0; i < 3; ++i) d[i] = out.at<float>(i);
}

// "-Rcw.t()*tcw" and "Rlw*twc+tlw" exactly as src/ORBmatcher.cc:1338-1346 writes them, through the shim's MatExpr
void frameref_neg_rt_t(const float* Tcw16, const float* Tlw16, float* twc3, float* tlc3)
{
    const cv::Mat Tc = pose_mat(Tcw16), Tl = pose_mat(Tlw16);
    const cv::Mat Rcw = Tc.rowRange(0, 3).colRange(0, 3);
    const cv::Mat tcw = Tc.rowRange(0, 3).col(3);
    const cv::Mat twc = -Rcw.t() * tcw;
    const cv::Mat Rlw = Tl.rowRange(0, 3).colRange(0, 3);
    const cv::Mat tlw = Tl.rowRange(0, 3).col(3);
    const cv::Mat tlc = Rlw * twc + tlw;
    for (int i = 0; i < 3; ++i) { twc3[i] = twc.at<float>(i); tlc3[i] = tlc.at<float>(i); }
}

// "M / s" for a CV_32F row vector, as the Sim3 routines of src/ORBmatcher.cc form it (:301-302): the shim's MatExpr
// lowering (convertTo with scale 1/s), checked against cv2 golden vectors
void frameref_mat_div(const float* v, int n, float s, float* out)
{
    cv::Mat M(1, n, CV_32F);
    std::memcpy(M.data, v, sizeof(float) * (size_t)n);
    cv::Mat R = M / s;
    std::memcpy(out, R.data, sizeof(float) * (size_t)n);
}

// "acc + v / s" for CV_32F vectors, as MapPoint::UpdateNormalAndDepth forms it (src/MapPoint.cc:355): the shim's MatExpr
// lowering (MatOp_AddEx -> cv::scaleAdd), checked against cv2 golden vectors
void frameref_mat_add_div(const float* acc, const float* v, int n, double s, float* out)
{
    cv::Mat A(n, 1, CV_32F), V(n, 1, CV_32F);
    std::memcpy(A.data, acc, sizeof(float) * (size_t)n);
    std::memcpy(V.data, v, sizeof(float) * (size_t)n);
    cv::Mat R = A + V / s;
    std::memcpy(out, R.data, sizeof(float) * (size_t)n);
}

int frameref_features_in_area(const KP* keys_un, int n, const float* bounds4, float x, float y, float r, int min_level,
                              int max_level, int* out, int cap)
{
    set_frame_statics(bounds4, 1.f, 1.f, 0.f, 0.f);
    Frame F;
    fill_frame(F, keys_un, 0, n);
    const std::vector<size_t> v = F.GetFeaturesInArea(x, y, r, min_level, max_level);
    for (size_t i = 0; i < v.size() && (int)i < cap; ++i) out[i] = (int)v[i];
    return (int)v.size();
}

int frameref_search_for_initialization(const KP* k1, const uint8_t* d1, int n1, const KP* k2, const uint8_t* d2, int n2,
                                       const float* bounds4, float* prev_matched, int window_size, float nnratio,
                                       int check_orientation, int* matches12)
{
    set_frame_statics(bounds4, 1.f, 1.f, 0.f, 0.f);
    Frame F1, F2;
    fill_frame(F1, k1, d1, n1);
    fill_frame(F2, k2, d2, n2);
    std::vector<cv::Point2f> prev((size_t)n1);
    for (int i = 0; i < n1; ++i) prev[i] = cv::Point2f(prev_matched[2 * i], prev_matched[2 * i + 1]);
    std::vector<int> m12;
    ORBmatcher matcher(nnratio, check_orientation != 0);
    const int nm = matcher.SearchForInitialization(F1, F2, prev, m12, window_size);
    for (int i = 0; i < n1; ++i) {
        matches12[i] = m12[i];
        prev_matched[2 * i] = prev[i].x; prev_matched[2 * i + 1] = prev[i].y;
    }
    return nm;
}

// cur_match: index of the last-frame map point CurrentFrame.mvpMapPoints[i2] ends up with, -1 where it is NULL or still
// the entry's own (occupied) map point.  (The restatement's extra "-2 = reset by the rotation check" is not observable.)
int frameref_search_by_projection(const frameport_projection_args* a, int* cur_match)
{
    set_frame_statics(a->bounds, a->fx, a->fy, a->cx, a->cy);
    World world;
    Frame Last, Cur;
    set_scale_tables(Last, a->scale_factors, a->nlevels);
    set_scale_tables(Cur, a->scale_factors, a->nlevels);

    // LastFrame: N, mvpMapPoints, mvbOutlier, mvKeys[i].octave, mvKeysUn[i].angle, mTcw
    Last.N = a->n_last;
    Last.mvKeys.assign((size_t)a->n_last, cv::KeyPoint());
    for (int i = 0; i < a->n_last; ++i) { Last.mvKeys[i].octave = a->last_octave[i]; Last.mvKeys[i].angle = a->last_angle[i]; }
    Last.mvKeysUn = Last.mvKeys;
    Last.mvpMapPoints.assign((size_t)a->n_last, static_cast<MapPoint*>(NULL));
    Last.mvbOutlier.assign((size_t)a->n_last, false);
    std::map<MapPoint*, int> index_of;
    for (int i = 0; i < a->n_last; ++i) {
        // mp_valid = map point present and not an outlier: alternate the two ways of being invalid
        MapPoint* p = world.point(a->world_pos + 3 * i, a->mp_desc + 32 * i, a->mp_observed[i] ? 1 : 0, false);
        index_of[p] = i;
        if (a->mp_valid[i]) Last.mvpMapPoints[i] = p;
        else if (i & 1) { Last.mvpMapPoints[i] = p; Last.mvbOutlier[i] = true; }
    }
    Last.SetPose(pose_mat(a->Tcw_last));

    fill_frame(Cur, a->cur_keys_un, a->cur_desc, a->n_cur);
    Cur.mvuRight.assign(a->cur_u_right, a->cur_u_right + a->n_cur);
    Cur.mbf = a->bf; Cur.mb = a->b;
    Last.mbf = a->bf; Last.mb = a->b;
    const float origin[3] = {0, 0, 0};
    for (int i = 0; i < a->n_cur; ++i)
        if (a->cur_occupied[i]) Cur.mvpMapPoints[i] = world.point(origin, 0, 1, false);
    Cur.SetPose(pose_mat(a->Tcw_cur));

    ORBmatcher matcher(0.9f, a->check_orientation != 0);
    const int nm = matcher.SearchByProjection(Cur, Last, a->th, a->mono != 0);
    for (int i = 0; i < a->n_cur; ++i) {
        MapPoint* p = Cur.mvpMapPoints[i];
        std::map<MapPoint*, int>::const_iterator it = p ? index_of.find(p) : index_of.end();
        cur_match[i] = it == index_of.end() ? -1 : it->second;
    }
    return nm;
}

// The local-map search.  Map points carry what Frame::isInFrustum would have left in them; in_view = 0 alternates between
// the two ways the reference skips a point (mbTrackInView false, or a bad point still flagged in view).
int frameref_search_local_points(const frameport_local_points_args* a, int* cur_match)
{
    set_frame_statics(a->bounds, 1.f, 1.f, 0.f, 0.f);
    World world;
    Frame F;
    set_scale_tables(F, a->scale_factors, a->nlevels);
    fill_frame(F, a->cur_keys_un, a->cur_desc, a->n_cur);
    F.mvuRight.assign(a->cur_u_right, a->cur_u_right + a->n_cur);
    const float origin[3] = {0, 0, 0};
    for (int i = 0; i < a->n_cur; ++i)
        if (a->cur_occupied[i]) F.mvpMapPoints[i] = world.point(origin, 0, 1, false);
        else if (i % 7 == 3) F.mvpMapPoints[i] = world.point(origin, 0, 0, false);       // present but unobserved: not skipped (:89-91)
    std::vector<MapPoint*> vp((size_t)a->n_points);
    std::map<MapPoint*, int> index_of;
    for (int i = 0; i < a->n_points; ++i) {
        const bool bad = !a->in_view[i] && (i & 1);
        MapPoint* p = world.point(origin, a->mp_desc + 32 * i, a->mp_observed[i] ? 2 : 0, bad);
        p->mbTrackInView = a->in_view[i] || bad;
        p->mTrackProjX = a->proj_x[i]; p->mTrackProjY = a->proj_y[i]; p->mTrackProjXR = a->proj_xr[i];
        p->mnTrackScaleLevel = a->scale_level[i];
        p->mTrackViewCos = a->view_cos[i];
        vp[i] = p; index_of[p] = i;
    }
    ORBmatcher matcher(a->nnratio, true);
    const int nm = matcher.SearchByProjection(F, vp, a->th);
    for (int i = 0; i < a->n_cur; ++i) {
        MapPoint* p = F.mvpMapPoints[i];
        std::map<MapPoint*, int>::const_iterator it = p ? index_of.find(p) : index_of.end();
        cur_match[i] = it == index_of.end() ? -1 : it->second;
    }
    return nm;
}

// The second half of Tracking::SearchLocalPoints (src/Tracking.cc:1171-1199) on a Frame posed by SetPose and MapPoints
// carrying world position, normal and the scale-invariance distances.
int frameref_search_local_map(const frameport_local_map_args* a, const frameport_track_fields* out, int* n_in_view, int* cur_match)
{
    set_frame_statics(a->bounds, a->fx, a->fy, a->cx, a->cy);
    World world;
    Frame F;
    set_scale_tables(F, a->scale_factors, a->nlevels);
    F.mfLogScaleFactor = a->log_scale_factor;
    fill_frame(F, a->cur_keys_un, a->cur_desc, a->n_cur);
    if (a->n_cur) F.mvuRight.assign(a->cur_u_right, a->cur_u_right + a->n_cur);
    F.mbf = a->bf; F.mb = 1.f; F.mnId = 77;
    F.SetPose(pose_mat(a->Tcw));                               // UpdatePoseMatrices: mRcw, mtcw, mOw
    const float origin[3] = {0, 0, 0};
    for (int i = 0; i < a->n_cur; ++i)
        if (a->cur_occupied[i]) F.mvpMapPoints[i] = world.point(origin, 0, 1, false);
    std::vector<MapPoint*> vp((size_t)a->n_points);
    std::map<MapPoint*, int> index_of;
    for (int i = 0; i < a->n_points; ++i) {
        // candidate = 0: alternately a bad point and one already seen in this frame (src/Tracking.cc:1177-1180)
        const bool bad = !a->candidate[i] && (i & 1);
        MapPoint* p = world.point(a->world_pos + 3 * i, a->mp_desc + 32 * i, a->mp_observed[i] ? 2 : 0, bad);
        cv::Mat nrm(3, 1, CV_32F);
        std::memcpy(nrm.data, a->normal + 3 * i, 3 * sizeof(float));
        p->mNormalVector = nrm;
        p->mfMaxDistance = a->max_distance[i]; p->mfMinDistance = a->min_distance[i];
        p->mnLastFrameSeen = (!a->candidate[i] && !bad) ? F.mnId : 0;
        p->mbTrackInView = false;
        vp[i] = p; index_of[p] = i;
    }
    int nToMatch = 0;
    for (int i = 0; i < a->n_points; ++i) {
        MapPoint* pMP = vp[i];
        out->in_view[i] = 0; out->proj_x[i] = out->proj_y[i] = out->proj_xr[i] = out->view_cos[i] = 0.f; out->scale_level[i] = 0;
        if (pMP->mnLastFrameSeen == F.mnId) continue;
        if (pMP->isBad()) continue;
        if (F.isInFrustum(pMP, a->viewing_cos_limit)) {
            nToMatch++;
            out->in_view[i] = pMP->mbTrackInView ? 1 : 0;
            out->proj_x[i] = pMP->mTrackProjX; out->proj_y[i] = pMP->mTrackProjY; out->proj_xr[i] = pMP->mTrackProjXR;
            out->scale_level[i] = pMP->mnTrackScaleLevel; out->view_cos[i] = pMP->mTrackViewCos;
        }
    }
    *n_in_view = nToMatch;
    for (int i = 0; i < a->n_cur; ++i) cur_match[i] = -1;
    if (nToMatch == 0 || a->n_cur == 0) return 0;
    ORBmatcher matcher(a->nnratio, true);
    const int nm = matcher.SearchByProjection(F, vp, a->th);
    for (int i = 0; i < a->n_cur; ++i) {
        MapPoint* p = F.mvpMapPoints[i];
        std::map<MapPoint*, int>::const_iterator it = p ? index_of.find(p) : index_of.end();
        cur_match[i] = it == index_of.end() ? -1 : it->second;
    }
    return nm;
}

// ORBmatcher::Fuse on a KeyFrame built by the reference's constructor.  Fuse's map surgery makes its per-point decisions
// observable only through the map, so the harness runs the reference's Fuse once per map point (a one-element vector) on
// a keyframe without map points, reads which keypoint the point was attached to, and detaches it again.
void frameref_fuse_search(const frameport_fuse_args* a, int* best_idx, float* ow_out)
{
    set_frame_statics(a->bounds, a->fx, a->fy, a->cx, a->cy);
    World world;
    Frame F;
    set_scale_tables(F, a->scale_factors, a->nlevels);
    F.mvInvLevelSigma2.assign(a->inv_level_sigma2, a->inv_level_sigma2 + a->nlevels);
    F.mfLogScaleFactor = a->log_scale_factor;
    fill_frame(F, a->keys_un, a->desc, a->n_kf);
    if (a->n_kf) F.mvuRight.assign(a->u_right, a->u_right + a->n_kf);
    F.mbf = a->bf; F.mb = 1.f; F.mThDepth = 1.f; F.mTimeStamp = 0; F.mnId = 5;
    F.mTcw = pose_mat(a->Tcw);
    KeyFrame* pKF = new KeyFrame(F, &world.map, static_cast<KeyFrameDatabase*>(NULL));
    world.keyframes.push_back(pKF);
    const cv::Mat Ow = pKF->GetCameraCenter();
    for (int i = 0; i < 3; ++i) ow_out[i] = Ow.at<float>(i);
    ORBmatcher matcher(0.6f, true);
    for (int i = 0; i < a->n_points; ++i) {
        best_idx[i] = -1;
        if (!a->present[i]) continue;
        MapPoint* p = world.point(a->world_pos + 3 * i, a->mp_desc + 32 * i, 3, false);
        cv::Mat nrm(3, 1, CV_32F);
        std::memcpy(nrm.data, a->normal + 3 * i, 3 * sizeof(float));
        p->mNormalVector = nrm;
        p->mfMaxDistance = a->max_distance[i]; p->mfMinDistance = a->min_distance[i];
        std::vector<MapPoint*> one(1, p);
        const int nf = matcher.Fuse(pKF, one, a->th);
        const int idx = p->GetIndexInKeyFrame(pKF);
        if ((nf == 1) != (idx >= 0)) never("Fuse: return value and observation disagree");
        if (idx >= 0) { best_idx[i] = idx; pKF->EraseMapPointMatch((size_t)idx); p->mObservations.erase(pKF); p->nObs = 3; }
    }
}

// The relocalisation search.  mp_valid = 0 cycles through the three ways the reference skips a keyframe feature: no map
// point, a bad one, one already in sAlreadyFound.
int frameref_search_by_projection_keyframe(const frameport_reloc_projection_args* a, int* cur_match)
{
    set_frame_statics(a->bounds, a->fx, a->fy, a->cx, a->cy);
    World world;
    Frame Cur, Src;
    set_scale_tables(Cur, a->scale_factors, a->nlevels);
    Cur.mfLogScaleFactor = a->log_scale_factor;
    fill_frame(Cur, a->cur_keys_un, a->cur_desc, a->n_cur);
    const float origin[3] = {0, 0, 0};
    for (int i = 0; i < a->n_cur; ++i)
        if (a->cur_occupied[i]) Cur.mvpMapPoints[i] = world.point(origin, 0, (i & 1), false);     // observed or not: any map point blocks
    Cur.SetPose(pose_mat(a->Tcw_cur));
    // the keyframe: only its map point matches and keypoint angles are read
    std::vector<KP> kk((size_t)a->n_kf);
    for (int i = 0; i < a->n_kf; ++i) { std::memset(&kk[i], 0, sizeof(KP)); kk[i].x = 10.f + (float)(i % 600); kk[i].y = 10.f + (float)(i % 400); kk[i].angle = a->kf_angle[i]; }
    set_scale_tables(Src, a->scale_factors, a->nlevels);
    fill_frame(Src, kk.data(), 0, a->n_kf);
    Src.mDescriptors = cv::Mat(a->n_kf > 0 ? a->n_kf : 1, 32, CV_8U);
    Src.mTcw = cv::Mat::eye(4, 4, CV_32F);
    Src.mbf = 1.f; Src.mb = 1.f; Src.mThDepth = 1.f; Src.mTimeStamp = 0; Src.mnId = 9;
    std::set<MapPoint*> already;
    std::map<MapPoint*, int> index_of;
    for (int i = 0; i < a->n_kf; ++i) {
        const int how = a->mp_valid[i] ? -1 : i % 3;
        if (how == 0) continue;                                // no map point
        MapPoint* p = world.point(a->world_pos + 3 * i, a->mp_desc + 32 * i, 2, how == 1);
        p->mfMaxDistance = a->max_distance[i]; p->mfMinDistance = a->min_distance[i];
        Src.mvpMapPoints[i] = p;
        index_of[p] = i;
        if (how == 2) already.insert(p);
    }
    KeyFrame* pKF = new KeyFrame(Src, &world.map, static_cast<KeyFrameDatabase*>(NULL));
    world.keyframes.push_back(pKF);
    ORBmatcher matcher(0.9f, a->check_orientation != 0);
    const int nm = matcher.SearchByProjection(Cur, pKF, already, a->th, a->orb_dist);
    for (int i = 0; i < a->n_cur; ++i) {
        MapPoint* p = Cur.mvpMapPoints[i];
        std::map<MapPoint*, int>::const_iterator it = p ? index_of.find(p) : index_of.end();
        cur_match[i] = it == index_of.end() ? -1 : it->second;
    }
    return nm;
}

namespace {
// a KeyFrame carrying the keypoints / descriptors / tables of a frameport_fuse_args snapshot, posed at T16
KeyFrame* keyframe_of(World& world, const frameport_fuse_args* a, const float* T16, unsigned long id, Frame& F)
{
    set_scale_tables(F, a->scale_factors, a->nlevels);
    if (a->inv_level_sigma2) F.mvInvLevelSigma2.assign(a->inv_level_sigma2, a->inv_level_sigma2 + a->nlevels);
    F.mfLogScaleFactor = a->log_scale_factor;
    fill_frame(F, a->keys_un, a->desc, a->n_kf);
    if (a->n_kf && a->u_right) F.mvuRight.assign(a->u_right, a->u_right + a->n_kf);
    F.mbf = a->bf; F.mb = 1.f; F.mThDepth = 1.f; F.mTimeStamp = 0; F.mnId = id;
    F.mTcw = pose_mat(T16);
    KeyFrame* pKF = new KeyFrame(F, &world.map, static_cast<KeyFrameDatabase*>(NULL));
    world.keyframes.push_back(pKF);
    return pKF;
}

MapPoint* point_of(World& world, const frameport_fuse_args* a, int i, bool bad)
{
    MapPoint* p = world.point(a->world_pos + 3 * i, a->mp_desc + 32 * i, 3, bad);
    if (a->normal) {
        cv::Mat nrm(3, 1, CV_32F);
        std::memcpy(nrm.data, a->normal + 3 * i, 3 * sizeof(float));
        p->mNormalVector = nrm;
    }
    p->mfMaxDistance = a->max_distance[i]; p->mfMinDistance = a->min_distance[i];
    return p;
}

cv::Mat mat44(const float* S16)
{
    cv::Mat m(4, 4, CV_32F);
    std::memcpy(m.data, S16, 16 * sizeof(float));
    return m;
}
}  // namespace

// Fuse(pKF, Scw, vpPoints, th, vpReplacePoint).  present = 0: alternately a bad point and a point the keyframe already
// holds (spAlreadyFound).  The keypoint a point was fused with is read back from the map: its own observation of pKF, or
// the observation of the point it was told to replace.
int frameref_fuse_search_sim3(const frameport_fuse_args* a, const float* Scw, int* best_idx)
{
    set_frame_statics(a->bounds, a->fx, a->fy, a->cx, a->cy);
    World world;
    Frame F;
    const float eye[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
    KeyFrame* pKF = keyframe_of(world, a, eye, 5, F);
    std::vector<MapPoint*> vp((size_t)a->n_points), repl((size_t)a->n_points, static_cast<MapPoint*>(NULL));
    int slot = 0;
    for (int i = 0; i < a->n_points; ++i) {
        const bool bad = !a->present[i] && (i & 1);
        vp[i] = point_of(world, a, i, bad);
        if (!a->present[i] && !bad && a->n_kf > 0) {          // already in the keyframe
            while (slot < a->n_kf && pKF->GetMapPoint(slot)) ++slot;
            if (slot < a->n_kf) { pKF->AddMapPoint(vp[i], slot); vp[i]->mObservations[pKF] = slot; }
            else vp[i]->mbBad = true;
        }
    }
    ORBmatcher matcher(0.8f, true);
    const int nf = matcher.Fuse(pKF, mat44(Scw), vp, a->th, repl);
    int seen = 0;
    for (int i = 0; i < a->n_points; ++i) {
        best_idx[i] = -1;
        if (!a->present[i]) continue;
        best_idx[i] = repl[i] ? repl[i]->GetIndexInKeyFrame(pKF) : vp[i]->GetIndexInKeyFrame(pKF);
        seen += best_idx[i] >= 0;
    }
    if (seen != nf) never("Fuse(Scw): return value and the map disagree");
    return nf;
}

// SearchByProjection(pKF, Scw, vpPoints, vpMatched, th).  matched_in marks keypoints that already hold a match; present = 0
// points are bad or (every third one) sit in vpMatched themselves.
int frameref_search_by_projection_sim3(const frameport_fuse_args* a, const float* Scw, const uint8_t* matched_in, int* matched)
{
    set_frame_statics(a->bounds, a->fx, a->fy, a->cx, a->cy);
    World world;
    Frame F;
    const float eye[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
    KeyFrame* pKF = keyframe_of(world, a, eye, 5, F);
    const float origin[3] = {0, 0, 0};
    std::vector<MapPoint*> vpMatched((size_t)a->n_kf, static_cast<MapPoint*>(NULL));
    std::vector<int> occupied;
    for (int j = 0; j < a->n_kf; ++j)
        if (matched_in[j]) { vpMatched[j] = world.point(origin, 0, 1, false); occupied.push_back(j); }
    std::vector<MapPoint*> vp((size_t)a->n_points);
    std::map<MapPoint*, int> index_of;
    size_t next = 0;
    for (int i = 0; i < a->n_points; ++i) {
        const bool in_matched = !a->present[i] && i % 3 == 0 && next < occupied.size();
        vp[i] = point_of(world, a, i, !a->present[i] && !in_matched);
        if (in_matched) vpMatched[occupied[next++]] = vp[i];
        index_of[vp[i]] = i;
    }
    ORBmatcher matcher(0.75f, true);
    const int nm = matcher.SearchByProjection(pKF, mat44(Scw), vp, vpMatched, (int)a->th);
    for (int j = 0; j < a->n_kf; ++j) {
        matched[j] = -1;
        if (matched_in[j] || !vpMatched[j]) continue;
        matched[j] = index_of[vpMatched[j]];
    }
    return nm;
}

// SearchBySim3(pKF1, pKF2, vpMatches12, s12, R12, t12, th).  a12: the map points of pKF1 (its pose in a12->Tcw) searched in
// pKF2, whose keypoints a12 carries; a21 the other way round.  present = 0 cycles through: no map point, a bad one, one
// already matched on entry (vpMatches12 / an observation of pKF2 by a pre-matched point).
int frameref_search_by_sim3(const frameport_fuse_args* a12, const frameport_fuse_args* a21, float s12, const float* R12,
                            const float* t12, int* match12)
{
    set_frame_statics(a12->bounds, a12->fx, a12->fy, a12->cx, a12->cy);
    World world;
    Frame F1, F2;
    const int N1 = a12->n_points, N2 = a21->n_points;
    KeyFrame* pKF1 = keyframe_of(world, a21, a12->Tcw, 5, F1);      // pKF1 carries the keypoints a21 searches in
    KeyFrame* pKF2 = keyframe_of(world, a12, a21->Tcw, 6, F2);
    const float origin[3] = {0, 0, 0};
    std::vector<MapPoint*> vpMatches12((size_t)N1, static_cast<MapPoint*>(NULL));
    std::map<MapPoint*, int> index2;
    std::vector<uint8_t> pre2((size_t)N2, 0);
    for (int i = 0; i < N2; ++i) {
        const int how = a21->present[i] ? -1 : i % 3;
        if (how == 0) continue;
        MapPoint* p = point_of(world, a21, i, how == 1);
        pKF2->AddMapPoint(p, i);
        index2[p] = i;
        if (how == 2) pre2[i] = 1;                             // will be reached through a pre-matched point of pKF1, or stays bad
    }
    std::vector<int> want2;                                    // keypoints of pKF2 that must read "already matched"
    for (int i = 0; i < N2; ++i) if (pre2[i]) want2.push_back(i);
    size_t next = 0;
    for (int i = 0; i < N1; ++i) {
        const int how = a12->present[i] ? -1 : i % 3;
        if (how == 0) continue;
        MapPoint* p = point_of(world, a12, i, how == 1);
        pKF1->AddMapPoint(p, i);
        if (how == 2) {                                        // pre-matched: blocks i1 and, when it observes pKF2, that keypoint
            MapPoint* m = world.point(origin, 0, 1, false);
            if (next < want2.size()) m->mObservations[pKF2] = (size_t)want2[next++];
            vpMatches12[i] = m;
        }
    }
    for (; next < want2.size(); ++next) pKF2->GetMapPoint(want2[next])->mbBad = true;      // leftovers: skipped as bad instead
    cv::Mat R(3, 3, CV_32F), t(3, 1, CV_32F);
    std::memcpy(R.data, R12, 9 * sizeof(float));
    std::memcpy(t.data, t12, 3 * sizeof(float));
    std::vector<MapPoint*> before = vpMatches12;
    ORBmatcher matcher(0.75f, true);
    const int nf = matcher.SearchBySim3(pKF1, pKF2, vpMatches12, s12, R, t, a12->th);
    for (int i = 0; i < N1; ++i) {
        match12[i] = -1;
        if (before[i] || !vpMatches12[i]) continue;
        match12[i] = index2[vpMatches12[i]];
    }
    return nf;
}

// strict_th_low = 0: SearchByBoW(KeyFrame* pKF, Frame& F, vpMapPointMatches)   (side 1 = pKF, side 2 = F)
// strict_th_low = 1: SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, vpMatches12)
int frameref_search_by_bow(const frameport_bow_side* s1, const frameport_bow_side* s2, float nnratio, int check_orientation,
                           int strict_th_low, int* match12, int* match21)
{
    const float bounds[4] = {0.f, 640.f, 0.f, 480.f};
    set_frame_statics(bounds, 1.f, 1.f, 0.f, 0.f);
    World world;
    const float origin[3] = {0, 0, 0};
    const frameport_bow_side* sides[2] = {s1, s2};
    Frame F[2];
    KeyFrame* KF[2] = {0, 0};
    std::map<MapPoint*, int> index_of[2];
    for (int s = 0; s < 2; ++s) {
        const frameport_bow_side* sd = sides[s];
        std::vector<KP> keys((size_t)sd->n);
        for (int i = 0; i < sd->n; ++i) {
            std::memset(&keys[i], 0, sizeof(KP));
            keys[i].x = 10.f + (float)(i % 600); keys[i].y = 10.f + (float)(i % 400);
            keys[i].angle = sd->angle[i];
        }
        fill_frame(F[s], keys.data(), sd->desc, sd->n);
        F[s].mTcw = cv::Mat::eye(4, 4, CV_32F);
        F[s].mbf = 1.f; F[s].mb = 1.f; F[s].mThDepth = 1.f; F[s].mTimeStamp = 0; F[s].mnId = (unsigned long)(s + 1);
        F[s].mnScaleLevels = 1; F[s].mfScaleFactor = 1.2f; F[s].mfLogScaleFactor = std::log(1.2f);
        fill_feature_vector(F[s].mFeatVec, sd);
        const bool as_keyframe = s == 0 || strict_th_low;
        if (!as_keyframe) continue;
        for (int i = 0; i < sd->n; ++i) {
            // valid = map point present and not bad: alternate the two ways of being invalid
            const bool valid = !sd->valid || sd->valid[i];
            if (valid || (i & 1)) {
                MapPoint* p = world.point(origin, sd->desc + 32 * (size_t)i, 1, !valid);
                F[s].mvpMapPoints[i] = p;
                index_of[s][p] = i;
            }
        }
        KF[s] = new KeyFrame(F[s], &world.map, static_cast<KeyFrameDatabase*>(NULL));   // src/KeyFrame.cc:34-60
        world.keyframes.push_back(KF[s]);
    }
    for (int i = 0; i < s1->n; ++i) match12[i] = -1;
    for (int i = 0; i < s2->n; ++i) match21[i] = -1;
    ORBmatcher matcher(nnratio, check_orientation != 0);
    int nm;
    if (!strict_th_low) {
        std::vector<MapPoint*> matches;
        nm = matcher.SearchByBoW(KF[0], F[1], matches);
        for (int i2 = 0; i2 < s2->n; ++i2)
            if (matches[i2]) { const int i1 = index_of[0][matches[i2]]; match21[i2] = i1; match12[i1] = i2; }
    } else {
        std::vector<MapPoint*> matches;
        nm = matcher.SearchByBoW(KF[0], KF[1], matches);
        for (int i1 = 0; i1 < s1->n; ++i1)
            if (matches[i1]) { const int i2 = index_of[1][matches[i1]]; match12[i1] = i2; match21[i2] = i1; }
    }
    return nm;
}

// SearchByBoW(pKF1, pKF2) three times on the SAME two KeyFrame objects, as a running system calls it: (a) before pKF2's
// FeatureVector exists (ComputeBoW not run yet: nothing can match), (b) after it has been filled in, (c) after every third
// map point of pKF1 has been erased from the keyframe.  match12_abc = three arrays of s1->n entries; returns the three
// counts packed as a + 1000 * (b + 1000 * c).  (For a matcher that keeps keyframes resident: hit, refresh, changed flags.)
long frameref_search_by_bow_thrice(const frameport_bow_side* s1, const frameport_bow_side* s2, float nnratio, int check_orientation,
                                   int* match12_abc)
{
    const float bounds[4] = {0.f, 640.f, 0.f, 480.f};
    set_frame_statics(bounds, 1.f, 1.f, 0.f, 0.f);
    World world;
    const float origin[3] = {0, 0, 0};
    const frameport_bow_side* sides[2] = {s1, s2};
    Frame F[2];
    KeyFrame* KF[2] = {0, 0};
    std::map<MapPoint*, int> index_of2;
    for (int s = 0; s < 2; ++s) {
        const frameport_bow_side* sd = sides[s];
        std::vector<KP> keys((size_t)sd->n);
        for (int i = 0; i < sd->n; ++i) {
            std::memset(&keys[i], 0, sizeof(KP));
            keys[i].x = 10.f + (float)(i % 600); keys[i].y = 10.f + (float)(i % 400);
            keys[i].angle = sd->angle[i];
        }
        fill_frame(F[s], keys.data(), sd->desc, sd->n);
        F[s].mTcw = cv::Mat::eye(4, 4, CV_32F);
        F[s].mbf = 1.f; F[s].mb = 1.f; F[s].mThDepth = 1.f; F[s].mTimeStamp = 0; F[s].mnId = (unsigned long)(s + 1);
        F[s].mnScaleLevels = 1; F[s].mfScaleFactor = 1.2f; F[s].mfLogScaleFactor = std::log(1.2f);
        if (s == 0) fill_feature_vector(F[s].mFeatVec, sd);
        for (int i = 0; i < sd->n; ++i)
            if (!sd->valid || sd->valid[i]) {
                MapPoint* p = world.point(origin, sd->desc + 32 * (size_t)i, 1, false);
                F[s].mvpMapPoints[i] = p;
                if (s == 1) index_of2[p] = i;
            }
        KF[s] = new KeyFrame(F[s], &world.map, static_cast<KeyFrameDatabase*>(NULL));
        world.keyframes.push_back(KF[s]);
    }
    ORBmatcher matcher(nnratio, check_orientation != 0);
    long nm[3];
    for (int pass = 0; pass < 3; ++pass) {
        if (pass == 1) fill_feature_vector(KF[1]->mFeatVec, s2);
        if (pass == 2)
            for (int i = 0; i < s1->n; i += 3) KF[0]->EraseMapPointMatch((size_t)i);
        std::vector<MapPoint*> matches;
        nm[pass] = matcher.SearchByBoW(KF[0], KF[1], matches);
        int* out = match12_abc + (size_t)pass * s1->n;
        for (int i1 = 0; i1 < s1->n; ++i1) out[i1] = matches[i1] ? index_of2[matches[i1]] : -1;
    }
    return nm[0] + 1000 * (nm[1] + 1000 * nm[2]);
}

// SearchForTriangulation between two KeyFrames built by the reference's constructor from plain arrays.  T1w is the pose
// of pKF1; the camera centre the reference derives from it (KeyFrame::SetPose, GetCameraCenter) is returned in cw1_out so
// the restatement can be given exactly the value the reference used (a->Cw1 is ignored here).
int frameref_search_for_triangulation(const frameport_triangulation_args* a, const float* T1w, int* matches12, float* cw1_out)
{
    const float bounds[4] = {0.f, 640.f, 0.f, 480.f};
    set_frame_statics(bounds, a->fx, a->fy, a->cx, a->cy);
    World world;
    const float origin[3] = {0, 0, 0};
    const frameport_triangulation_side* sides[2] = {&a->kf1, &a->kf2};
    Frame F[2];
    KeyFrame* KF[2] = {0, 0};
    for (int s = 0; s < 2; ++s) {
        const frameport_triangulation_side* sd = sides[s];
        set_scale_tables(F[s], a->scale_factors2, a->nlevels);
        F[s].mvLevelSigma2.assign(a->level_sigma2, a->level_sigma2 + a->nlevels);
        fill_frame(F[s], sd->keys_un, sd->desc, sd->n);
        F[s].mvuRight.assign(sd->u_right, sd->u_right + sd->n);
        F[s].mTcw = pose_mat(s == 0 ? T1w : a->T2w);
        F[s].mbf = 1.f; F[s].mb = 1.f; F[s].mThDepth = 1.f; F[s].mTimeStamp = 0; F[s].mnId = (unsigned long)(s + 1);
        frameport_bow_side fv; std::memset(&fv, 0, sizeof(fv));
        fv.node_ids = sd->node_ids; fv.node_start = sd->node_start; fv.indices = sd->indices; fv.n_nodes = sd->n_nodes;
        fill_feature_vector(F[s].mFeatVec, &fv);
        for (int i = 0; i < sd->n; ++i)
            if (sd->has_map_point[i]) F[s].mvpMapPoints[i] = world.point(origin, sd->desc + 32 * (size_t)i, 1, false);
        KF[s] = new KeyFrame(F[s], &world.map, static_cast<KeyFrameDatabase*>(NULL));   // src/KeyFrame.cc:34-60
        world.keyframes.push_back(KF[s]);
    }
    const cv::Mat Cw = KF[0]->GetCameraCenter();
    for (int i = 0; i < 3; ++i) cw1_out[i] = Cw.at<float>(i);
    cv::Mat F12(3, 3, CV_32F);
    std::memcpy(F12.data, a->F12, 9 * sizeof(float));
    std::vector<std::pair<size_t, size_t> > pairs;
    ORBmatcher matcher(0.6f, a->check_orientation != 0);
    const int nm = matcher.SearchForTriangulation(KF[0], KF[1], F12, pairs, a->only_stereo != 0);
    for (int i = 0; i < a->kf1.n; ++i) matches12[i] = -1;
    size_t last = 0;
    for (size_t k = 0; k < pairs.size(); ++k) {
        if (k && pairs[k].first <= last) never("vMatchedPairs not ascending");
        last = pairs[k].first;
        matches12[pairs[k].first] = (int)pairs[k].second;
    }
    return nm;
}

// MapPoint::ComputeDistinctiveDescriptors on groups of descriptors.  The reference walks std::map<KeyFrame*, size_t>, i.e.
// the observing keyframes in ADDRESS order: a pool of keyframes sorted by address carries the group's descriptors in the
// given order (one one-row mDescriptors each); best_idx[g] = the row whose descriptor the map point ends up with (the
// first such row when several are equal -- equal rows have equal medians, and the reference keeps the first).
void frameref_distinctive_descriptors(const uint8_t* desc, const int32_t* group_start, int n_groups, int* best_idx)
{
    const float bounds[4] = {0.f, 640.f, 0.f, 480.f};
    set_frame_statics(bounds, 1.f, 1.f, 0.f, 0.f);
    World world;
    int nmax = 0;
    for (int g = 0; g < n_groups; ++g) nmax = std::max(nmax, group_start[g + 1] - group_start[g]);
    std::vector<KeyFrame*> pool;
    for (int j = 0; j < nmax; ++j) {
        Frame F;
        F.N = 0; F.mTcw = cv::Mat::eye(4, 4, CV_32F); F.mnId = (unsigned long)(j + 1);
        F.mbf = 1.f; F.mb = 1.f; F.mThDepth = 1.f; F.mTimeStamp = 0;
        F.mnScaleLevels = 1; F.mfScaleFactor = 1.2f; F.mfLogScaleFactor = std::log(1.2f);
        KeyFrame* kf = new KeyFrame(F, &world.map, static_cast<KeyFrameDatabase*>(NULL));
        world.keyframes.push_back(kf);
        pool.push_back(kf);
    }
    std::sort(pool.begin(), pool.end());
    const float origin[3] = {0, 0, 0};
    for (int g = 0; g < n_groups; ++g) {
        const int N = group_start[g + 1] - group_start[g];
        const uint8_t* d = desc + 32 * (size_t)group_start[g];
        best_idx[g] = -1;
        MapPoint* p = world.point(origin, 0, N, false);
        p->mDescriptor = cv::Mat();
        p->mObservations.clear();
        for (int j = 0; j < N; ++j) {
            cv::Mat row(1, 32, CV_8U);
            std::memcpy(row.data, d + 32 * (size_t)j, 32);
            const_cast<cv::Mat&>(pool[j]->mDescriptors) = row;
            p->mObservations[pool[j]] = 0;
        }
        p->ComputeDistinctiveDescriptors();
        const cv::Mat out = p->GetDescriptor();
        if (N == 0) { if (!out.empty()) never("ComputeDistinctiveDescriptors touched an unobserved point"); continue; }
        for (int j = 0; j < N && best_idx[g] < 0; ++j)
            if (std::memcmp(out.data, d + 32 * (size_t)j, 32) == 0) best_idx[g] = j;
    }
}

// Frame::ComputeStereoMatches on caller-provided keypoints, descriptors and UNBLURRED pyramid levels (dense rows).
// nlevels = number of entries of lvlL / lvlR / w / h / scale / inv_scale.
void frameref_stereo_matches(const KP* kL, const uint8_t* dL, int nL, const KP* kR, const uint8_t* dR, int nR,
                             const uint8_t* const* lvlL, const uint8_t* const* lvlR, const int* w, const int* h,
                             const float* scale, const float* inv_scale, int nlevels, float bf, float b, float* u_right,
                             float* depth)
{
    ORBextractor exL(1000, scale[nlevels > 1 ? 1 : 0], nlevels, 20, 7), exR(1000, scale[nlevels > 1 ? 1 : 0], nlevels, 20, 7);
    exL.mvImagePyramid.resize((size_t)nlevels);
    exR.mvImagePyramid.resize((size_t)nlevels);
    for (int l = 0; l < nlevels; ++l) {
        exL.mvImagePyramid[l] = cv::Mat(h[l], w[l], CV_8U, (void*)lvlL[l]);
        exR.mvImagePyramid[l] = cv::Mat(h[l], w[l], CV_8U, (void*)lvlR[l]);
    }
    Frame F;
    F.mpORBextractorLeft = &exL; F.mpORBextractorRight = &exR;
    F.N = nL;
    F.mvKeys = to_keys(kL, nL);
    F.mvKeysRight = to_keys(kR, nR);
    F.mDescriptors = desc_mat(dL, nL);
    F.mDescriptorsRight = desc_mat(dR, nR);
    F.mvScaleFactors.assign(scale, scale + nlevels);
    F.mvInvScaleFactors.assign(inv_scale, inv_scale + nlevels);
    F.mbf = bf; F.mb = b;
    F.ComputeStereoMatches();
    for (int i = 0; i < nL; ++i) { u_right[i] = F.mvuRight[i]; depth[i] = F.mvDepth[i]; }
}

// ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:1647-1663) on two 32-byte descriptors
int frameref_descriptor_distance(const uint8_t* a, const uint8_t* b)
{
    cv::Mat ma(1, 32, CV_8U, (void*)a), mb(1, 32, CV_8U, (void*)b);
    return ORBmatcher::DescriptorDistance(ma, mb);
}

// ORBmatcher::ComputeThreeMaxima (src/ORBmatcher.cc:1601-1642) on a histogram given by its bin sizes
void frameref_three_maxima(const int* bin_sizes, int L, int* ind3)
{
    std::vector<std::vector<int> > histo((size_t)L);
    for (int i = 0; i < L; ++i) histo[i].assign((size_t)bin_sizes[i], 0);
    ORBmatcher matcher(0.6f, true);
    ind3[0] = ind3[1] = ind3[2] = -1;
    matcher.ComputeThreeMaxima(histo.data(), L, ind3[0], ind3[1], ind3[2]);
}

// Frame::UndistortKeyPoints (and, with bounds4 != NULL, Frame::ComputeImageBounds on a w x h image)
void frameref_undistort_keypoints(const KP* keys, int n, const float* K4, const float* dist, int ndist, int w, int h,
                                  KP* keys_un, float* bounds4)
{
    Frame F;
    F.N = n;
    F.mvKeys = to_keys(keys, n);
    F.mK = cv::Mat::eye(3, 3, CV_32F);
    F.mK.at<float>(0, 0) = K4[0]; F.mK.at<float>(1, 1) = K4[1]; F.mK.at<float>(0, 2) = K4[2]; F.mK.at<float>(1, 2) = K4[3];
    F.mDistCoef = cv::Mat(ndist, 1, CV_32F);
    for (int i = 0; i < ndist; ++i) F.mDistCoef.at<float>(i) = dist[i];
    F.UndistortKeyPoints();
    if (n) std::memcpy(keys_un, F.mvKeysUn.data(), sizeof(KP) * (size_t)n);
    if (bounds4) {
        cv::Mat im(h, w, CV_8U);
        F.ComputeImageBounds(im);
        bounds4[0] = Frame::mnMinX; bounds4[1] = Frame::mnMaxX; bounds4[2] = Frame::mnMinY; bounds4[3] = Frame::mnMaxY;
    }
}

}  // extern "C"

// The reference's own KeyFrameDatabase (src/KeyFrameDatabase.cc, compiled unchanged): add / erase / DetectLoopCandidates /
// DetectRelocalizationCandidates on KeyFrames whose BowVectors, covisibility lists and earlier scores come from the caller.
// `voc` is a vocabulary read by oracle/ref_vocab_wrap.cc (it sizes the inverted file and scores).
extern "C" int frameref_db_detect(void* voc, const frameport_db_args* a, int32_t* marked, int32_t* words, float* score,
                                  int32_t* candidates)
{
    const float bounds[4] = {0.f, 640.f, 0.f, 480.f};
    set_frame_statics(bounds, 1.f, 1.f, 0.f, 0.f);
    const ORBVocabulary* V = static_cast<const ORBVocabulary*>(voc);
    World world;
    KeyFrameDatabase db(*V);
    const int n = a->n_kf;
    std::vector<KeyFrame*> kfs;
    for (int k = 0; k < n; ++k) {
        Frame F;
        F.N = 0; F.mTcw = cv::Mat::eye(4, 4, CV_32F); F.mnId = (unsigned long)(k + 1);
        F.mbf = 1.f; F.mb = 1.f; F.mThDepth = 1.f; F.mTimeStamp = 0;
        F.mnScaleLevels = 1; F.mfScaleFactor = 1.2f; F.mfLogScaleFactor = std::log(1.2f);
        for (int i = a->kstart[k]; i < a->kstart[k + 1]; ++i) F.mBowVec.insert(F.mBowVec.end(), std::make_pair(a->kw[i], a->kv[i]));
        KeyFrame* kf = new KeyFrame(F, &world.map, &db);
        world.keyframes.push_back(kf);
        kfs.push_back(kf);
        kf->mLoopScore = a->prev_score[k]; kf->mRelocScore = a->prev_score[k];
        db.add(kf);
    }
    for (int k = 0; k < n; ++k) {
        for (int j = a->neigh_start[k]; j < a->neigh_start[k + 1]; ++j) kfs[k]->mvpOrderedConnectedKeyFrames.push_back(kfs[a->neigh[j]]);
        if (a->erased[k]) db.erase(kfs[k]);
    }
    // the query: a keyframe outside the database (LoopClosing::DetectLoop adds it afterwards) / the current frame
    Frame Q;
    Q.N = 0; Q.mTcw = cv::Mat::eye(4, 4, CV_32F); Q.mnId = (unsigned long)(n + 100);
    Q.mbf = 1.f; Q.mb = 1.f; Q.mThDepth = 1.f; Q.mTimeStamp = 0;
    Q.mnScaleLevels = 1; Q.mfScaleFactor = 1.2f; Q.mfLogScaleFactor = std::log(1.2f);
    for (int i = 0; i < a->nq; ++i) Q.mBowVec.insert(Q.mBowVec.end(), std::make_pair(a->qw[i], a->qv[i]));
    std::vector<KeyFrame*> got;
    unsigned long qid;
    if (a->reloc) {
        got = db.DetectRelocalizationCandidates(&Q);
        qid = Q.mnId;
    } else {
        KeyFrame* q = new KeyFrame(Q, &world.map, &db);
        world.keyframes.push_back(q);
        for (int k = 0; k < n; ++k)
            if (a->connected[k]) q->mConnectedKeyFrameWeights[kfs[k]] = 20;
        got = db.DetectLoopCandidates(q, a->min_score);
        qid = q->mnId;
    }
    for (int k = 0; k < n; ++k) {
        marked[k] = a->reloc ? (kfs[k]->mnRelocQuery == qid) : (kfs[k]->mnLoopQuery == qid);
        words[k] = a->reloc ? kfs[k]->mnRelocWords : kfs[k]->mnLoopWords;
        score[k] = a->reloc ? kfs[k]->mRelocScore : kfs[k]->mLoopScore;
    }
    for (size_t i = 0; i < got.size(); ++i)
        candidates[i] = (int32_t)(std::find(kfs.begin(), kfs.end(), got[i]) - kfs.begin());
    return (int)got.size();
}

// MapPoint::UpdateNormalAndDepth (src/MapPoint.cc:330-371) executed by the reference on a pool of keyframes sorted by
// address (the order of its std::map<KeyFrame*, size_t>): keyframe j of a point's walk gets that observation's camera
// centre, keyframe `ref_slot[g]` of the walk is the point's reference keyframe.
extern "C" void frameref_update_normal_and_depth(const float* pos, const int32_t* obs_start, const float* centers,
                                                 const int32_t* ref_slot, const int32_t* ref_level, const float* scale_factors,
                                                 int n_levels, int n_points, float* normal, float* max_dist, float* min_dist)
{
    const float bounds[4] = {0.f, 640.f, 0.f, 480.f};
    set_frame_statics(bounds, 1.f, 1.f, 0.f, 0.f);
    World world;
    int nmax = 0;
    for (int g = 0; g < n_points; ++g) nmax = std::max(nmax, obs_start[g + 1] - obs_start[g]);
    std::vector<KeyFrame*> pool;
    for (int j = 0; j < nmax; ++j) {
        Frame F;
        F.N = 0; F.mTcw = cv::Mat::eye(4, 4, CV_32F); F.mnId = (unsigned long)(j + 1);
        F.mbf = 1.f; F.mb = 1.f; F.mThDepth = 1.f; F.mTimeStamp = 0;
        F.mnScaleLevels = n_levels; F.mfScaleFactor = scale_factors[1]; F.mfLogScaleFactor = std::log(scale_factors[1]);
        F.mvScaleFactors.assign(scale_factors, scale_factors + n_levels);
        F.mvKeysUn.resize(1);
        KeyFrame* kf = new KeyFrame(F, &world.map, static_cast<KeyFrameDatabase*>(NULL));
        world.keyframes.push_back(kf);
        pool.push_back(kf);
    }
    std::sort(pool.begin(), pool.end());
    for (int g = 0; g < n_points; ++g) {
        const int N = obs_start[g + 1] - obs_start[g];
        MapPoint* p = world.point(pos + 3 * (size_t)g, 0, N, false);
        p->mObservations.clear();
        p->mNormalVector = cv::Mat(3, 1, CV_32F);
        std::memcpy(p->mNormalVector.data, normal + 3 * (size_t)g, 12);
        p->mfMaxDistance = max_dist[g]; p->mfMinDistance = min_dist[g];
        for (int j = 0; j < N; ++j) {
            cv::Mat ow(3, 1, CV_32F);
            std::memcpy(ow.data, centers + 3 * (size_t)(obs_start[g] + j), 12);
            pool[j]->Ow = ow;
            p->mObservations[pool[j]] = 0;
        }
        if (N > 0) {
            p->mpRefKF = pool[ref_slot[g]];
            const_cast<std::vector<cv::KeyPoint>&>(p->mpRefKF->mvKeysUn)[0].octave = ref_level[g];
        }
        p->UpdateNormalAndDepth();
        std::memcpy(normal + 3 * (size_t)g, p->mNormalVector.data, 12);
        max_dist[g] = p->mfMaxDistance; min_dist[g] = p->mfMinDistance;
    }
}
