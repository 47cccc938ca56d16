// ORBmatcher_gpu.cc -- drop-in replacement of the reference's src/ORBmatcher.cc: every member of ORB_SLAM2::ORBmatcher with
// the reference's own signature (include/ORBmatcher.h:37-102), compiled against the reference's own headers (Frame.h,
// KeyFrame.h, MapPoint.h).  Each Search* / Fuse body takes a plain-array snapshot of what the reference body reads through
// the Frame / KeyFrame / MapPoint getters, makes ONE call into the C ABI (include/orbgpu.h) and applies the result to the
// objects exactly where the reference body writes (mvpMapPoints, vpMatched, Replace / AddObservation / AddMapPoint ...).
// The parts that walk or edit the map graph stay host code, in the reference's order.
//
// In an ORB-SLAM2 tree: replace src/ORBmatcher.cc by this file, add `friend class ORBmatcher;` to class MapPoint
// (include/MapPoint.h; the searches need the raw mfMaxDistance / mfMinDistance, the public getters return them scaled by
// 1.2f / 0.8f, src/MapPoint.cc:373-383) and link liborbgpu.so.  The test build (oracle/Makefile,
// oracle/_ref/libframegpu.so) defines ORBGPU_OPEN_MAPPOINT instead of editing the header.
#ifdef ORBGPU_OPEN_MAPPOINT
#define protected public
#include <MapPoint.h>
#undef protected
#endif
#include <ORBmatcher.h>      // the REFERENCE's include/ORBmatcher.h (angle brackets: not the batched host/ORBmatcher.h beside this file)

#include <stdint.h>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <set>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <vector>

#include "orbgpu.h"

using namespace std;

namespace ORB_SLAM2 {

const int ORBmatcher::TH_HIGH = 100;      // src/ORBmatcher.cc:37
const int ORBmatcher::TH_LOW = 50;        // :38
const int ORBmatcher::HISTO_LENGTH = 30;  // :39

static_assert(sizeof(cv::KeyPoint) == sizeof(orbgpu_keypoint), "cv::KeyPoint must be the 28-byte POD of the C ABI");

#ifndef ORBGPU_MATCHER_DEVICE
#define ORBGPU_MATCHER_DEVICE 0           // CUDA ordinal the matcher's calls run on
#endif
#ifndef ORBGPU_MATCHER_RESIDENT_KEYFRAMES
#define ORBGPU_MATCHER_RESIDENT_KEYFRAMES 4096   // keyframes kept in HBM by the BoW-gated searches (~70 KB each), least recently used out first
#endif

namespace {

// The reference's matcher cannot fail; there is no CPU path to fall back to, so a failed call is fatal and loud.
void check(int rc, const char* what)
{
    if (rc == ORBGPU_OK) return;
    const std::string msg = std::string("ORBmatcher (") + what + "): " + orbgpu_strerror(rc) + ": " + orbgpu_last_cuda_error();
    std::fprintf(stderr, "%s\n", msg.c_str());
    throw std::runtime_error(msg);
}

const orbgpu_keypoint* keys_of(const std::vector<cv::KeyPoint>& v) { return reinterpret_cast<const orbgpu_keypoint*>(v.data()); }

// dense N x 32 view of a descriptor matrix (copied only when the rows are strided)
const uint8_t* dense_desc(const cv::Mat& D, std::vector<uint8_t>& tmp)
{
    if (D.rows == 0) { tmp.assign(32, 0); return tmp.data(); }
    if (D.isContinuous()) return D.ptr(0);
    tmp.resize((size_t)D.rows * 32);
    for (int r = 0; r < D.rows; ++r) std::memcpy(&tmp[(size_t)32 * r], D.ptr(r), 32);
    return tmp.data();
}

void pose16(const cv::Mat& T, float out[16])
{
    for (int r = 0; r < 4; ++r)
        for (int c = 0; c < 4; ++c) out[4 * r + c] = T.at<float>(r, c);
}

// what the searches read from a list of map points (entries may be NULL)
struct PointSnap {
    std::vector<float> pos, normal, max_d, min_d;
    std::vector<uint8_t> desc, observed;
    explicit PointSnap(size_t n) : pos(3 * n + 3, 0.f), normal(3 * n + 3, 0.f), max_d(n + 1, 0.f), min_d(n + 1, 0.f), desc(32 * n + 32, 0), observed(n + 1, 0) {}
    void set(size_t i, MapPoint* p)
    {
        const cv::Mat P = p->GetWorldPos(), Nv = p->GetNormal(), D = p->GetDescriptor();
        for (int k = 0; k < 3; ++k) { pos[3 * i + k] = P.at<float>(k); normal[3 * i + k] = Nv.at<float>(k); }
        max_d[i] = p->mfMaxDistance; min_d[i] = p->mfMinDistance;
        if (D.rows > 0) std::memcpy(&desc[32 * i], D.ptr(0), 32);
        observed[i] = p->Observations() > 0;
    }
};

// DBoW2::FeatureVector (std::map<node id, vector<feature index>>) flattened in the map's order
struct FlatFeatVec {
    std::vector<uint32_t> ids, idx;
    std::vector<int32_t> start;
    explicit FlatFeatVec(const DBoW2::FeatureVector& fv)
    {
        for (DBoW2::FeatureVector::const_iterator it = fv.begin(); it != fv.end(); ++it) {
            ids.push_back(it->first);
            start.push_back((int32_t)idx.size());
            idx.insert(idx.end(), it->second.begin(), it->second.end());
        }
        start.push_back((int32_t)idx.size());
        if (ids.empty()) ids.push_back(0);
        if (idx.empty()) idx.push_back(0);
    }
    int n_nodes() const { return (int)start.size() - 1; }
};

// the keyframe half of orbgpu_fuse_args (pose, intrinsics, features, grid bounds, scale tables)
struct KeyFrameSnap {
    std::vector<uint8_t> tmp;
    float T[16];
    void fill(orbgpu_fuse_args& a, KeyFrame* pKF, float th)
    {
        pose16(pKF->GetPose(), T);
        a.Tcw = T;
        const cv::Mat Ow = pKF->GetCameraCenter();
        for (int k = 0; k < 3; ++k) a.Ow[k] = Ow.at<float>(k);
        a.fx = pKF->fx; a.fy = pKF->fy; a.cx = pKF->cx; a.cy = pKF->cy; a.bf = pKF->mbf;
        a.log_scale_factor = pKF->mfLogScaleFactor;
        a.n_kf = pKF->N;
        a.keys_un = keys_of(pKF->mvKeysUn);
        a.desc = dense_desc(pKF->mDescriptors, tmp);
        a.u_right = pKF->mvuRight.data();
        a.bounds[0] = Frame::mnMinX; a.bounds[1] = Frame::mnMaxX; a.bounds[2] = Frame::mnMinY; a.bounds[3] = Frame::mnMaxY;
        a.scale_factors = pKF->mvScaleFactors.data();
        a.inv_level_sigma2 = pKF->mvInvLevelSigma2.data();
        a.nlevels = pKF->mnScaleLevels;
        a.th = th;
    }
};

// Keyframes resident in HBM.  What the BoW-gated searches read of a KeyFrame -- mvKeysUn, mDescriptors, mvuRight, mFeatVec,
// the level tables and intrinsics -- is written once (KeyFrame constructor, ComputeBoW: src/KeyFrame.cc:32-58, :66-76), so it
// is uploaded the first time a search names the keyframe (orbgpu_kf_add) and referred to by id afterwards; a call then
// carries only the map-point flags and poses.  An entry is refreshed when the keyframe's FeatureVector has been computed
// since (a keyframe searched before LocalMapping::ProcessNewKeyFrame ran ComputeBoW), and the least recently used entries
// go when the table is full (culled keyframes are never searched again, so they age out).  Shared by the matcher objects of
// all threads (Tracking, LocalMapping, LoopClosing), as the keyframes are.
class ResidentKeyFrames {
public:
    orbgpu_kfstore* store()
    {
        std::lock_guard<std::mutex> g(mu_);
        open();
        return store_;
    }
    int32_t id_of(KeyFrame* pKF)
    {
        std::lock_guard<std::mutex> g(mu_);
        open();
        const size_t fv = pKF->mFeatVec.size();
        std::unordered_map<const KeyFrame*, Entry>::iterator it = map_.find(pKF);
        if (it != map_.end()) {
            if (it->second.mnId == pKF->mnId && it->second.fv_nodes == fv) { it->second.used = ++clock_; return it->second.id; }
            orbgpu_kf_erase(store_, it->second.id);
            map_.erase(it);
        }
        if (map_.size() >= (size_t)ORBGPU_MATCHER_RESIDENT_KEYFRAMES) {
            std::unordered_map<const KeyFrame*, Entry>::iterator lru = map_.begin();
            for (it = map_.begin(); it != map_.end(); ++it) if (it->second.used < lru->second.used) lru = it;
            orbgpu_kf_erase(store_, lru->second.id);
            map_.erase(lru);
        }
        const FlatFeatVec f(pKF->mFeatVec);
        std::vector<uint8_t> tmp;
        orbgpu_kf_data k;
        std::memset(&k, 0, sizeof(k));
        k.n = pKF->N; k.keys_un = keys_of(pKF->mvKeysUn); k.desc = dense_desc(pKF->mDescriptors, tmp);
        k.u_right = pKF->mvuRight.empty() ? 0 : pKF->mvuRight.data();
        k.node_ids = f.ids.data(); k.node_start = f.start.data(); k.indices = f.idx.data(); k.n_nodes = f.n_nodes();
        k.scale_factors = pKF->mvScaleFactors.data(); k.level_sigma2 = pKF->mvLevelSigma2.data(); k.nlevels = pKF->mnScaleLevels;
        k.fx = pKF->fx; k.fy = pKF->fy; k.cx = pKF->cx; k.cy = pKF->cy;
        Entry e;
        check(orbgpu_kf_add(store_, &k, &e.id), "keyframe upload");
        e.mnId = pKF->mnId; e.fv_nodes = fv; e.used = ++clock_;
        map_[pKF] = e;
        return e.id;
    }

private:
    struct Entry { int32_t id; long unsigned int mnId; size_t fv_nodes; uint64_t used; };
    void open() { if (!store_) check(orbgpu_kfstore_create(&store_, ORBGPU_MATCHER_DEVICE), "keyframe store"); }
    std::mutex mu_;
    orbgpu_kfstore* store_ = 0;
    std::unordered_map<const KeyFrame*, Entry> map_;
    uint64_t clock_ = 0;
};
ResidentKeyFrames g_keyframes;      // released with the process

void points_into(orbgpu_fuse_args& a, const PointSnap& s, const std::vector<uint8_t>& present)
{
    a.n_points = (int32_t)present.size();
    a.present = present.empty() ? s.observed.data() : present.data();
    a.world_pos = s.pos.data(); a.normal = s.normal.data();
    a.max_distance = s.max_d.data(); a.min_distance = s.min_d.data();
    a.mp_desc = s.desc.data();
}

}  // namespace

ORBmatcher::ORBmatcher(float nnratio, bool checkOri) : mfNNratio(nnratio), mbCheckOrientation(checkOri) {}

// src/ORBmatcher.cc:1647-1663
int ORBmatcher::DescriptorDistance(const cv::Mat& a, const cv::Mat& b) { return orbgpu_hamming256(a.ptr(0), b.ptr(0)); }

// ---- Tracking::SearchLocalPoints: src/ORBmatcher.cc:44-125 ---------------------------------------------------------------
int ORBmatcher::SearchByProjection(Frame& F, const vector<MapPoint*>& vpMapPoints, const float th)
{
    const size_t n = vpMapPoints.size();
    if (n == 0 || F.N == 0) return 0;
    std::vector<uint8_t> inView(n), observed(n), desc(32 * n), occupied((size_t)F.N), tmp;
    std::vector<float> px(n), py(n), pxr(n), vcos(n);
    std::vector<int32_t> lvl(n);
    for (size_t i = 0; i < n; ++i) {
        MapPoint* p = vpMapPoints[i];
        inView[i] = p->mbTrackInView && !p->isBad();
        px[i] = p->mTrackProjX; py[i] = p->mTrackProjY; pxr[i] = p->mTrackProjXR;
        lvl[i] = p->mnTrackScaleLevel; vcos[i] = p->mTrackViewCos;
        observed[i] = p->Observations() > 0;
        const cv::Mat D = p->GetDescriptor();
        if (D.rows > 0) std::memcpy(&desc[32 * i], D.ptr(0), 32);
        if (!inView[i]) lvl[i] = 0;                              // members of points outside the frustum are never read
    }
    for (int j = 0; j < F.N; ++j) occupied[(size_t)j] = F.mvpMapPoints[j] && F.mvpMapPoints[j]->Observations() > 0;
    orbgpu_local_points_args a;
    std::memset(&a, 0, sizeof(a));
    a.n_points = (int32_t)n; a.in_view = inView.data(); a.proj_x = px.data(); a.proj_y = py.data(); a.proj_xr = pxr.data();
    a.scale_level = lvl.data(); a.view_cos = vcos.data(); a.mp_desc = desc.data(); a.mp_observed = observed.data();
    a.n_cur = F.N; a.cur_keys_un = keys_of(F.mvKeysUn); a.cur_desc = dense_desc(F.mDescriptors, tmp);
    a.cur_u_right = F.mvuRight.data(); a.cur_occupied = occupied.data();
    a.bounds[0] = F.mnMinX; a.bounds[1] = F.mnMaxX; a.bounds[2] = F.mnMinY; a.bounds[3] = F.mnMaxY;
    a.scale_factors = F.mvScaleFactors.data(); a.nlevels = F.mnScaleLevels; a.th = th; a.nnratio = mfNNratio;
    std::vector<int32_t> curMatch((size_t)F.N, -1);
    int nmatches = 0;
    check(orbgpu_search_local_points(ORBGPU_MATCHER_DEVICE, &a, curMatch.data(), &nmatches), "SearchByProjection(F, points)");
    for (int j = 0; j < F.N; ++j)
        if (curMatch[(size_t)j] >= 0) F.mvpMapPoints[j] = vpMapPoints[(size_t)curMatch[(size_t)j]];
    return nmatches;
}

float ORBmatcher::RadiusByViewingCos(const float& viewCos) { return viewCos > 0.998 ? 2.5 : 4.0; }      // :127-133 (evaluated on the device)

// :140-157 (evaluated on the device inside SearchForTriangulation; kept for callers of the protected member)
bool ORBmatcher::CheckDistEpipolarLine(const cv::KeyPoint& kp1, const cv::KeyPoint& kp2, const cv::Mat& F12, const KeyFrame* pKF2)
{
    const float a = kp1.pt.x * F12.at<float>(0, 0) + kp1.pt.y * F12.at<float>(1, 0) + F12.at<float>(2, 0);
    const float b = kp1.pt.x * F12.at<float>(0, 1) + kp1.pt.y * F12.at<float>(1, 1) + F12.at<float>(2, 1);
    const float c = kp1.pt.x * F12.at<float>(0, 2) + kp1.pt.y * F12.at<float>(1, 2) + F12.at<float>(2, 2);
    const float num = a * kp2.pt.x + b * kp2.pt.y + c;
    const float den = a * a + b * b;
    if (den == 0) return false;
    const float dsqr = num * num / den;
    return dsqr < 3.84 * pKF2->mvLevelSigma2[kp2.octave];
}

// ---- SearchByBoW, both overloads: :159-288 and :522-655 --------------------------------------------------------------------
static void bow_side_of(orbgpu_bow_side& s, const cv::Mat& D, int n, const std::vector<uint8_t>* valid, const std::vector<float>& angle,
                        const FlatFeatVec& fv, std::vector<uint8_t>& tmp)
{
    s.desc = dense_desc(D, tmp); s.n = n; s.valid = valid ? valid->data() : 0; s.angle = angle.data();
    s.node_ids = fv.ids.data(); s.node_start = fv.start.data(); s.indices = fv.idx.data(); s.n_nodes = fv.n_nodes();
}

int ORBmatcher::SearchByBoW(KeyFrame* pKF, Frame& F, vector<MapPoint*>& vpMapPointMatches)
{
    const vector<MapPoint*> vpMapPointsKF = pKF->GetMapPointMatches();
    vpMapPointMatches = vector<MapPoint*>(F.N, static_cast<MapPoint*>(NULL));
    const int n1 = pKF->N, n2 = F.N;
    if (n1 == 0 || n2 == 0) return 0;
    std::vector<uint8_t> valid1((size_t)n1), valid2((size_t)n2, 1), t2;              // a Frame's features are all eligible
    std::vector<float> ang2((size_t)n2 + 1);
    for (int i = 0; i < n1; ++i) {
        MapPoint* p = vpMapPointsKF[(size_t)i];
        valid1[(size_t)i] = p && !p->isBad();
    }
    for (int i = 0; i < n2; ++i) ang2[(size_t)i] = F.mvKeys[(size_t)i].angle;        // the frame side reads mvKeys (:245)
    const FlatFeatVec f2(F.mFeatVec);
    orbgpu_bow_side s2;
    bow_side_of(s2, F.mDescriptors, n2, &valid2, ang2, f2, t2);
    std::vector<int32_t> m12((size_t)n1, -1), m21((size_t)n2, -1);
    int nmatches = 0;
    // the keyframe is resident (uploaded the first time it is searched); the frame travels with the call
    check(orbgpu_kf_search_by_bow(g_keyframes.store(), g_keyframes.id_of(pKF), valid1.data(), -1, 0, &s2, mfNNratio, mbCheckOrientation, 0,
                                  m12.data(), m21.data(), &nmatches), "SearchByBoW(KF, F)");
    for (int i2 = 0; i2 < n2; ++i2)
        if (m21[(size_t)i2] >= 0) vpMapPointMatches[(size_t)i2] = vpMapPointsKF[(size_t)m21[(size_t)i2]];
    return nmatches;
}

int ORBmatcher::SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, vector<MapPoint*>& vpMatches12)
{
    const vector<MapPoint*> vpMapPoints1 = pKF1->GetMapPointMatches(), vpMapPoints2 = pKF2->GetMapPointMatches();
    vpMatches12 = vector<MapPoint*>(vpMapPoints1.size(), static_cast<MapPoint*>(NULL));
    const int n1 = (int)vpMapPoints1.size(), n2 = (int)vpMapPoints2.size();
    if (n1 == 0 || n2 == 0) return 0;
    std::vector<uint8_t> valid1((size_t)n1), valid2((size_t)n2);
    for (int i = 0; i < n1; ++i) {
        MapPoint* p = vpMapPoints1[(size_t)i];
        valid1[(size_t)i] = p && !p->isBad();
    }
    for (int i = 0; i < n2; ++i) {
        MapPoint* p = vpMapPoints2[(size_t)i];
        valid2[(size_t)i] = p && !p->isBad();
    }
    std::vector<int32_t> m12((size_t)n1, -1), m21((size_t)n2, -1);
    int nmatches = 0;
    const int32_t id1 = g_keyframes.id_of(pKF1), id2 = g_keyframes.id_of(pKF2);      // both resident: only the flags travel
    check(orbgpu_kf_search_by_bow(g_keyframes.store(), id1, valid1.data(), id2, valid2.data(), 0, mfNNratio, mbCheckOrientation, 1,
                                  m12.data(), m21.data(), &nmatches), "SearchByBoW(KF, KF)");
    for (int i1 = 0; i1 < n1; ++i1)
        if (m12[(size_t)i1] >= 0) vpMatches12[(size_t)i1] = vpMapPoints2[(size_t)m12[(size_t)i1]];
    return nmatches;
}

// ---- LoopClosing: SearchByProjection with a Sim3, :290-403 ---------------------------------------------------------------------
int ORBmatcher::SearchByProjection(KeyFrame* pKF, cv::Mat Scw, const vector<MapPoint*>& vpPoints, vector<MapPoint*>& vpMatched, int th)
{
    set<MapPoint*> spAlreadyFound(vpMatched.begin(), vpMatched.end());
    spAlreadyFound.erase(static_cast<MapPoint*>(NULL));
    const size_t n = vpPoints.size();
    if (n == 0 || pKF->N == 0) return 0;
    PointSnap snap(n);
    std::vector<uint8_t> present(n), matchedIn((size_t)pKF->N);
    for (size_t i = 0; i < n; ++i) {
        MapPoint* p = vpPoints[i];
        present[i] = !p->isBad() && !spAlreadyFound.count(p);
        snap.set(i, p);
    }
    for (int j = 0; j < pKF->N; ++j) matchedIn[(size_t)j] = vpMatched[(size_t)j] != NULL;
    orbgpu_fuse_args a;
    std::memset(&a, 0, sizeof(a));
    KeyFrameSnap ks;
    ks.fill(a, pKF, (float)th);
    points_into(a, snap, present);
    float S[16];
    pose16(Scw, S);
    std::vector<int32_t> matched((size_t)pKF->N, -1);
    int nmatches = 0;
    check(orbgpu_search_by_projection_sim3(ORBGPU_MATCHER_DEVICE, &a, S, matchedIn.data(), matched.data(), &nmatches), "SearchByProjection(KF, Scw)");
    for (int j = 0; j < pKF->N; ++j)
        if (matched[(size_t)j] >= 0) vpMatched[(size_t)j] = vpPoints[(size_t)matched[(size_t)j]];
    return nmatches;
}

// ---- Tracking::MonocularInitialization: :405-520 -------------------------------------------------------------------------------
int ORBmatcher::SearchForInitialization(Frame& F1, Frame& F2, vector<cv::Point2f>& vbPrevMatched, vector<int>& vnMatches12, int windowSize)
{
    vnMatches12 = vector<int>(F1.mvKeysUn.size(), -1);
    const int n1 = (int)F1.mvKeysUn.size(), n2 = F2.N;
    if (n1 == 0 || n2 == 0) return 0;
    static_assert(sizeof(cv::Point2f) == 8, "cv::Point2f is two floats");
    std::vector<uint8_t> t1, t2;
    const float bounds[4] = {F2.mnMinX, F2.mnMaxX, F2.mnMinY, F2.mnMaxY};
    std::vector<int32_t> m12((size_t)n1, -1);
    int nmatches = 0;
    check(orbgpu_search_for_initialization(ORBGPU_MATCHER_DEVICE, keys_of(F1.mvKeysUn), dense_desc(F1.mDescriptors, t1), n1, keys_of(F2.mvKeysUn),
                                           dense_desc(F2.mDescriptors, t2), n2, bounds, reinterpret_cast<float*>(vbPrevMatched.data()), windowSize,
                                           mfNNratio, mbCheckOrientation, m12.data(), &nmatches),
          "SearchForInitialization");
    for (int i = 0; i < n1; ++i) vnMatches12[(size_t)i] = m12[(size_t)i];
    return nmatches;
}

// ---- LocalMapping::CreateNewMapPoints: :657-820 ------------------------------------------------------------------------------------
int ORBmatcher::SearchForTriangulation(KeyFrame* pKF1, KeyFrame* pKF2, cv::Mat F12, vector<pair<size_t, size_t> >& vMatchedPairs,
                                       const bool bOnlyStereo)
{
    vMatchedPairs.clear();
    const int n1 = pKF1->N, n2 = pKF2->N;
    if (n1 == 0 || n2 == 0) return 0;
    orbgpu_kf_triangulation_job a;
    std::memset(&a, 0, sizeof(a));
    std::vector<uint8_t> has1((size_t)n1), has2((size_t)n2);
    for (int i = 0; i < n1; ++i) has1[(size_t)i] = pKF1->GetMapPoint((size_t)i) != NULL;
    for (int i = 0; i < n2; ++i) has2[(size_t)i] = pKF2->GetMapPoint((size_t)i) != NULL;
    a.id1 = g_keyframes.id_of(pKF1); a.id2 = g_keyframes.id_of(pKF2);                // keypoints, descriptors, mvuRight, mFeatVec: resident
    a.has_map_point1 = has1.data(); a.has_map_point2 = has2.data();
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) a.F12[3 * r + c] = F12.at<float>(r, c);
    const cv::Mat Cw = pKF1->GetCameraCenter();
    for (int k = 0; k < 3; ++k) a.Cw1[k] = Cw.at<float>(k);
    pose16(pKF2->GetPose(), a.T2w);
    a.only_stereo = bOnlyStereo; a.check_orientation = mbCheckOrientation;
    std::vector<int32_t> m12((size_t)n1, -1);
    int nmatches = 0;
    check(orbgpu_kf_search_for_triangulation(g_keyframes.store(), &a, m12.data(), &nmatches), "SearchForTriangulation");
    vMatchedPairs.reserve((size_t)nmatches);
    for (int i = 0; i < n1; ++i)
        if (m12[(size_t)i] >= 0) vMatchedPairs.push_back(make_pair((size_t)i, (size_t)m12[(size_t)i]));
    return nmatches;
}

// ---- LocalMapping::SearchInNeighbors: :825-957 -------------------------------------------------------------------------------------
int ORBmatcher::Fuse(KeyFrame* pKF, const vector<MapPoint*>& vpMapPoints, const float th)
{
    const size_t n = vpMapPoints.size();
    if (n == 0 || pKF->N == 0) return 0;
    PointSnap snap(n);
    std::vector<uint8_t> present(n);
    for (size_t i = 0; i < n; ++i) {
        present[i] = vpMapPoints[i] != NULL;
        if (present[i]) snap.set(i, vpMapPoints[i]);
    }
    orbgpu_fuse_args a;
    std::memset(&a, 0, sizeof(a));
    KeyFrameSnap ks;
    ks.fill(a, pKF, th);
    points_into(a, snap, present);
    std::vector<int32_t> best(n, -1);
    check(orbgpu_fuse_search(ORBGPU_MATCHER_DEVICE, &a, best.data()), "Fuse");
    // the reference's loop with the search taken out: entry tests and map surgery in index order
    int nFused = 0;
    for (size_t i = 0; i < n; ++i) {
        MapPoint* pMP = vpMapPoints[i];
        if (!pMP || pMP->isBad() || pMP->IsInKeyFrame(pKF) || best[i] < 0) continue;
        const int bestIdx = best[i];
        MapPoint* pMPinKF = pKF->GetMapPoint((size_t)bestIdx);
        if (pMPinKF) {
            if (!pMPinKF->isBad()) {
                if (pMPinKF->Observations() > pMP->Observations()) pMP->Replace(pMPinKF);
                else pMPinKF->Replace(pMP);
            }
        } else {
            pMP->AddObservation(pKF, (size_t)bestIdx);
            pKF->AddMapPoint(pMP, (size_t)bestIdx);
        }
        ++nFused;
    }
    return nFused;
}

// ---- LoopClosing::SearchAndFuse: :969-1097 -----------------------------------------------------------------------------------------
int ORBmatcher::Fuse(KeyFrame* pKF, cv::Mat Scw, const vector<MapPoint*>& vpPoints, float th, vector<MapPoint*>& vpReplacePoint)
{
    const set<MapPoint*> spAlreadyFound = pKF->GetMapPoints();
    const size_t n = vpPoints.size();
    if (n == 0 || pKF->N == 0) return 0;
    PointSnap snap(n);
    std::vector<uint8_t> present(n);
    for (size_t i = 0; i < n; ++i) {
        MapPoint* p = vpPoints[i];
        present[i] = !p->isBad() && !spAlreadyFound.count(p);
        snap.set(i, p);
    }
    orbgpu_fuse_args a;
    std::memset(&a, 0, sizeof(a));
    KeyFrameSnap ks;
    ks.fill(a, pKF, th);
    points_into(a, snap, present);
    float S[16];
    pose16(Scw, S);
    std::vector<int32_t> best(n, -1);
    check(orbgpu_fuse_search_sim3(ORBGPU_MATCHER_DEVICE, &a, S, best.data()), "Fuse(Scw)");
    int nFused = 0;
    for (size_t i = 0; i < n; ++i) {
        if (!present[i] || best[i] < 0) continue;
        MapPoint* pMP = vpPoints[i];
        MapPoint* pMPinKF = pKF->GetMapPoint((size_t)best[i]);
        if (pMPinKF) {
            if (!pMPinKF->isBad()) vpReplacePoint[i] = pMPinKF;
        } else {
            pMP->AddObservation(pKF, (size_t)best[i]);
            pKF->AddMapPoint(pMP, (size_t)best[i]);
        }
        ++nFused;
    }
    return nFused;
}

// ---- LoopClosing::ComputeSim3: :1102-1326 ------------------------------------------------------------------------------------------
int ORBmatcher::SearchBySim3(KeyFrame* pKF1, KeyFrame* pKF2, vector<MapPoint*>& vpMatches12, const float& s12, const cv::Mat& R12,
                             const cv::Mat& t12, const float th)
{
    const vector<MapPoint*> vpMapPoints1 = pKF1->GetMapPointMatches(), vpMapPoints2 = pKF2->GetMapPointMatches();
    const int N1 = (int)vpMapPoints1.size(), N2 = (int)vpMapPoints2.size();
    if (N1 == 0 || N2 == 0) return 0;
    std::vector<uint8_t> done1((size_t)N1, 0), done2((size_t)N2, 0);
    for (int i = 0; i < N1; ++i) {
        MapPoint* pMP = vpMatches12[(size_t)i];
        if (!pMP) continue;
        done1[(size_t)i] = 1;
        const int idx2 = pMP->GetIndexInKeyFrame(pKF2);
        if (idx2 >= 0 && idx2 < N2) done2[(size_t)idx2] = 1;
    }
    PointSnap snap1((size_t)N1), snap2((size_t)N2);
    std::vector<uint8_t> present1((size_t)N1), present2((size_t)N2);
    for (int i = 0; i < N1; ++i) {
        MapPoint* p = vpMapPoints1[(size_t)i];
        present1[(size_t)i] = p && !done1[(size_t)i] && !p->isBad();
        if (p) snap1.set((size_t)i, p);
    }
    for (int i = 0; i < N2; ++i) {
        MapPoint* p = vpMapPoints2[(size_t)i];
        present2[(size_t)i] = p && !done2[(size_t)i] && !p->isBad();
        if (p) snap2.set((size_t)i, p);
    }
    // a12: the points of pKF1 (camera 1 from world) searched among pKF2's features; a21 the other direction
    orbgpu_fuse_args a12, a21;
    std::memset(&a12, 0, sizeof(a12)); std::memset(&a21, 0, sizeof(a21));
    KeyFrameSnap k2, k1;
    float T1[16], T2[16];
    k2.fill(a12, pKF2, th); k1.fill(a21, pKF1, th);
    pose16(pKF1->GetPose(), T1); pose16(pKF2->GetPose(), T2);
    a12.Tcw = T1; a21.Tcw = T2;
    a12.fx = a21.fx = pKF1->fx; a12.fy = a21.fy = pKF1->fy; a12.cx = a21.cx = pKF1->cx; a12.cy = a21.cy = pKF1->cy;   // :1105-1108: pKF1's for both
    points_into(a12, snap1, present1); points_into(a21, snap2, present2);
    float R[9], t[3];
    for (int r = 0; r < 3; ++r) { t[r] = t12.at<float>(r); for (int c = 0; c < 3; ++c) R[3 * r + c] = R12.at<float>(r, c); }
    std::vector<int32_t> m12((size_t)N1, -1);
    int nFound = 0;
    check(orbgpu_search_by_sim3(ORBGPU_MATCHER_DEVICE, &a12, &a21, s12, R, t, m12.data(), &nFound), "SearchBySim3");
    for (int i1 = 0; i1 < N1; ++i1)
        if (m12[(size_t)i1] >= 0) vpMatches12[(size_t)i1] = vpMapPoints2[(size_t)m12[(size_t)i1]];
    return nFound;
}

// ---- Tracking::TrackWithMotionModel: :1328-1470 ------------------------------------------------------------------------------------
int ORBmatcher::SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, const float th, const bool bMono)
{
    const int nl = LastFrame.N, nc = CurrentFrame.N;
    if (nl == 0 || nc == 0) return 0;
    const size_t n = (size_t)nl;
    std::vector<float> pos(3 * n, 0.f), ang(n);
    std::vector<uint8_t> desc(32 * n, 0), valid(n), observed(n, 0), occupied((size_t)nc), tmp;
    std::vector<int32_t> oct(n);
    for (size_t i = 0; i < n; ++i) {
        MapPoint* p = LastFrame.mvpMapPoints[i];
        valid[i] = p && !LastFrame.mvbOutlier[i];
        oct[i] = LastFrame.mvKeys[i].octave;
        ang[i] = LastFrame.mvKeysUn[i].angle;
        if (!p) continue;
        const cv::Mat P = p->GetWorldPos(), D = p->GetDescriptor();
        for (int k = 0; k < 3; ++k) pos[3 * i + k] = P.at<float>(k);
        if (D.rows > 0) std::memcpy(&desc[32 * i], D.ptr(0), 32);
        observed[i] = p->Observations() > 0;
    }
    for (int j = 0; j < nc; ++j) occupied[(size_t)j] = CurrentFrame.mvpMapPoints[j] && CurrentFrame.mvpMapPoints[j]->Observations() > 0;
    float Tl[16], Tc[16];
    pose16(LastFrame.mTcw, Tl); pose16(CurrentFrame.mTcw, Tc);
    orbgpu_projection_args a;
    std::memset(&a, 0, sizeof(a));
    a.n_last = nl; a.world_pos = pos.data(); a.mp_desc = desc.data(); a.mp_valid = valid.data(); a.mp_observed = observed.data();
    a.last_octave = oct.data(); a.last_angle = ang.data(); a.Tcw_last = Tl;
    a.n_cur = nc; a.cur_keys_un = keys_of(CurrentFrame.mvKeysUn); a.cur_desc = dense_desc(CurrentFrame.mDescriptors, tmp);
    a.cur_u_right = CurrentFrame.mvuRight.data(); a.cur_occupied = occupied.data(); a.Tcw_cur = Tc;
    a.fx = CurrentFrame.fx; a.fy = CurrentFrame.fy; a.cx = CurrentFrame.cx; a.cy = CurrentFrame.cy; a.bf = CurrentFrame.mbf; a.b = CurrentFrame.mb;
    a.bounds[0] = CurrentFrame.mnMinX; a.bounds[1] = CurrentFrame.mnMaxX; a.bounds[2] = CurrentFrame.mnMinY; a.bounds[3] = CurrentFrame.mnMaxY;
    a.scale_factors = CurrentFrame.mvScaleFactors.data(); a.nlevels = CurrentFrame.mnScaleLevels;
    a.th = th; a.mono = bMono; a.check_orientation = mbCheckOrientation;
    std::vector<int32_t> curMatch((size_t)nc, -1);
    int nmatches = 0;
    check(orbgpu_search_by_projection(ORBGPU_MATCHER_DEVICE, &a, curMatch.data(), &nmatches), "SearchByProjection(Cur, Last)");
    for (int j = 0; j < nc; ++j) {
        const int m = curMatch[(size_t)j];
        if (m >= 0) CurrentFrame.mvpMapPoints[j] = LastFrame.mvpMapPoints[(size_t)m];
        else if (m == -2) CurrentFrame.mvpMapPoints[j] = static_cast<MapPoint*>(NULL);        // set, then removed by the rotation check
    }
    return nmatches;
}

// ---- Tracking::Relocalization: :1472-1597 -------------------------------------------------------------------------------------------
int ORBmatcher::SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, const set<MapPoint*>& sAlreadyFound, const float th, const int ORBdist)
{
    const vector<MapPoint*> vpMPs = pKF->GetMapPointMatches();
    const size_t n = vpMPs.size();
    const int nc = CurrentFrame.N;
    if (n == 0 || nc == 0) return 0;
    PointSnap snap(n);
    std::vector<uint8_t> valid(n), occupied((size_t)nc), tmp;
    std::vector<float> ang(n);
    for (size_t i = 0; i < n; ++i) {
        MapPoint* p = vpMPs[i];
        valid[i] = p && !p->isBad() && !sAlreadyFound.count(p);
        ang[i] = pKF->mvKeysUn[i].angle;
        if (p) snap.set(i, p);
    }
    for (int j = 0; j < nc; ++j) occupied[(size_t)j] = CurrentFrame.mvpMapPoints[j] != NULL;
    float Tc[16];
    pose16(CurrentFrame.mTcw, Tc);
    orbgpu_reloc_projection_args a;
    std::memset(&a, 0, sizeof(a));
    a.n_kf = (int32_t)n; a.mp_valid = valid.data(); a.world_pos = snap.pos.data(); a.max_distance = snap.max_d.data();
    a.min_distance = snap.min_d.data(); a.mp_desc = snap.desc.data(); a.kf_angle = ang.data();
    a.n_cur = nc; a.cur_keys_un = keys_of(CurrentFrame.mvKeysUn); a.cur_desc = dense_desc(CurrentFrame.mDescriptors, tmp);
    a.cur_occupied = occupied.data(); a.Tcw_cur = Tc;
    a.fx = CurrentFrame.fx; a.fy = CurrentFrame.fy; a.cx = CurrentFrame.cx; a.cy = CurrentFrame.cy;
    a.bounds[0] = CurrentFrame.mnMinX; a.bounds[1] = CurrentFrame.mnMaxX; a.bounds[2] = CurrentFrame.mnMinY; a.bounds[3] = CurrentFrame.mnMaxY;
    a.scale_factors = CurrentFrame.mvScaleFactors.data(); a.log_scale_factor = CurrentFrame.mfLogScaleFactor; a.nlevels = CurrentFrame.mnScaleLevels;
    a.th = th; a.orb_dist = ORBdist; a.check_orientation = mbCheckOrientation;
    std::vector<int32_t> curMatch((size_t)nc, -1);
    int nmatches = 0;
    check(orbgpu_search_by_projection_keyframe(ORBGPU_MATCHER_DEVICE, &a, curMatch.data(), &nmatches), "SearchByProjection(Cur, KF)");
    for (int j = 0; j < nc; ++j) {
        const int m = curMatch[(size_t)j];
        if (m >= 0) CurrentFrame.mvpMapPoints[j] = vpMPs[(size_t)m];
        else if (m == -2) CurrentFrame.mvpMapPoints[j] = static_cast<MapPoint*>(NULL);
    }
    return nmatches;
}

// :1601-1642 (the rotation histograms of the searches above are evaluated inside the library; kept for callers of the member).
// Three fullest bins; the first bin reaching a size wins ties; the 2nd / 3rd are dropped below a tenth of the fullest.
void ORBmatcher::ComputeThreeMaxima(vector<int>* histo, const int L, int& ind1, int& ind2, int& ind3)
{
    int top[3] = {0, 0, 0};
    int* ind[3] = {&ind1, &ind2, &ind3};
    for (int i = 0; i < L; ++i) {
        const int s = (int)histo[i].size();
        int rank = 0;
        while (rank < 3 && s <= top[rank]) ++rank;
        if (rank == 3) continue;
        for (int k = 2; k > rank; --k) { top[k] = top[k - 1]; *ind[k] = *ind[k - 1]; }
        top[rank] = s; *ind[rank] = i;
    }
    if ((float)top[1] < 0.1f * (float)top[0]) { ind2 = -1; ind3 = -1; }
    else if ((float)top[2] < 0.1f * (float)top[0]) { ind3 = -1; }
}

}  // namespace ORB_SLAM2
