/* orbgpu.h -- C ABI of the B200-native ORB front end (drop-in boundary).
 *
 * The reference has no FFI layer on this path: the boundary is the C++ surface of two classes.
 * Each entry point below names the reference interface it stands behind (paths relative to the
 * reference tree).  The C++ drop-in classes in my-slam_b200/host/ (ORB_SLAM2::ORBextractor,
 * ORB_SLAM2::ORBmatcher, same signatures as include/ORBextractor.h:45-111 and
 * include/ORBmatcher.h:37-102) are thin wrappers over these calls; INTEGRATION.md shows the binding.
 *
 * Conventions: plain C types only; the caller owns every host buffer passed in or out; the library
 * owns device and pinned memory; every call returns 0 (ORBGPU_OK) or a negative error code, never
 * throws and never exits; one call in flight per extractor handle (the reference extractor is not
 * re-entrant either: it mutates mvImagePyramid, src/ORBextractor.cc:1117), distinct handles may be
 * used concurrently from different threads (src/Frame.cc:78-81).  There is no CPU fallback: without a
 * CUDA device every compute call fails with ORBGPU_ERR_CUDA.
 */
#ifndef ORBGPU_H
#define ORBGPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORBGPU_OK                0
#define ORBGPU_ERR_ARG          -1   /* null pointer, bad size, bad level ...                       */
#define ORBGPU_ERR_CAPACITY     -2   /* caller's output capacity < orbgpu_max_keypoints()           */
#define ORBGPU_ERR_CUDA         -3   /* CUDA runtime/driver error (orbgpu_last_cuda_error())        */
#define ORBGPU_ERR_NOMEM        -4
#define ORBGPU_ERR_UNSUPPORTED  -5   /* image shape / parameters outside the supported envelope     */
#define ORBGPU_ERR_STATE        -6   /* ticket unknown, queue full, nothing extracted yet ...       */

#define ORBGPU_MAX_LEVELS 16

typedef struct orbgpu_extractor orbgpu_extractor;   /* opaque: streams, device + pinned buffers     */
typedef struct orbgpu_matcher orbgpu_matcher;       /* opaque: resident descriptor sets            */

/* == cv::KeyPoint (28-byte POD) as filled by the reference: include/ORBextractor.h:39,60;
 * field values per src/ORBextractor.cc:842-849 (octave, size), :479 (angle), :1097-1103 (pt*=scale) */
typedef struct {
    float x, y, size, angle, response;
    int32_t octave, class_id;
} orbgpu_keypoint;

/* ---- extractor ----------------------------------------------------------------------------- */

/* ORBextractor::ORBextractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST)
 * include/ORBextractor.h:51-52, src/ORBextractor.cc:412-472.  Builds the scale / features-per-level /
 * umax / pattern tables on the host with the reference's float-double mix and uploads them. */
int orbgpu_create(orbgpu_extractor** out, int device, int nfeatures, float scale_factor, int nlevels,
                  int ini_th_fast, int min_th_fast);
void orbgpu_destroy(orbgpu_extractor* e);

/* Getters GetScaleFactors / GetInverseScaleFactors / GetScaleSigmaSquares / GetInverseScaleSigmaSquares
 * (include/ORBextractor.h:63-83): scale4 = [scale | 1/scale | sigma^2 | 1/sigma^2], nlevels floats each;
 * features_per_level[nlevels]; umax[16].  Any pointer may be NULL.  Returns nlevels. */
int orbgpu_get_tables(orbgpu_extractor* e, float* scale4, int* features_per_level, int* umax);

/* Upper bound of keypoints per frame for this shape (sum over levels of max(N_l, 4*nIni_l) + 3;
 * the reference can return N_l + 2 per level, src/ORBextractor.cc:730-733).  Negative = error code. */
int orbgpu_max_keypoints(orbgpu_extractor* e, int w, int h);

/* ORBextractor::operator()(image, mask, keypoints, descriptors), include/ORBextractor.h:59-61,
 * src/ORBextractor.cc:1045-1107.  Synchronous, one frame: img is CV_8UC1 with row stride `stride`
 * bytes (host memory); kps[capacity], desc[capacity*32].  The mask of the reference is ignored there
 * and absent here.  w==0 || h==0 || img==NULL -> returns ORBGPU_OK with *n_out = -1 and outputs
 * untouched (the reference's silent return on an empty image, :1048-1049). */
int orbgpu_extract(orbgpu_extractor* e, const uint8_t* img, int w, int h, size_t stride,
                   orbgpu_keypoint* kps, uint8_t* desc, int capacity, int* n_out);

/* Same operator over n same-shape frames (frame i at frames + i*frame_stride, rows row_stride apart),
 * host buffers; H2D copy, kernels and D2H copy of successive chunks overlap on internal streams.
 * Output for frame i: kps + i*capacity, desc + i*capacity*32, n_out[i]; records at index n_out[i] and beyond are
 * undefined (pinned output buffers receive whole rows of the device arrays).  Pinned (cudaHostAlloc /
 * cudaHostRegister) caller buffers are used in place; pageable ones are staged through the
 * library's pinned ring. */
int orbgpu_extract_batch(orbgpu_extractor* e, const uint8_t* frames, int n, int w, int h,
                         size_t row_stride, size_t frame_stride,
                         orbgpu_keypoint* kps, uint8_t* desc, int capacity, int* n_out);

/* Asynchronous form of orbgpu_extract_batch for a continuous stream of batches: enqueues the batch on the
 * internal streams and returns (it only blocks while all stream contexts are still busy with earlier chunks:
 * back-pressure), so the upload of batch k+1 overlaps the kernels of batch k and the pipeline never drains.
 * Every buffer must stay valid and untouched until orbgpu_wait returns.  orbgpu_wait blocks until everything
 * enqueued since the previous wait has completed, delivers staged (pageable) outputs and returns the first
 * error; orbgpu_last_batch_ms then covers the whole window.  The synchronous calls, the frame queue and
 * shape changes return ORBGPU_ERR_STATE while a window is open. */
int orbgpu_extract_batch_async(orbgpu_extractor* e, const uint8_t* frames, int n, int w, int h,
                               size_t row_stride, size_t frame_stride,
                               orbgpu_keypoint* kps, uint8_t* desc, int capacity, int* n_out);
int orbgpu_wait(orbgpu_extractor* e);

/* orbgpu_extract_batch over several GPUs of one process: handle r (created on its own device) extracts the r-th contiguous
 * block of the n frames on its own host thread, results land in the caller's arrays in frame order.  Frames are
 * independent, so nothing is exchanged between the GPUs (the one-extractor-per-thread pattern of src/Frame.cc:78-81 with
 * one GPU per thread).  Returns the first error of any handle. */
int orbgpu_extract_batch_multi(orbgpu_extractor** handles, int num_handles, const uint8_t* frames, int n, int w, int h,
                               size_t row_stride, size_t frame_stride, orbgpu_keypoint* kps, uint8_t* desc, int capacity,
                               int* n_out);

/* Same, with every buffer already resident in device memory (frames, kps, desc, n_out are device
 * pointers of this extractor's device).  Returns after the work has completed. */
int orbgpu_extract_batch_device(orbgpu_extractor* e, const uint8_t* d_frames, int n, int w, int h,
                                size_t row_stride, size_t frame_stride,
                                orbgpu_keypoint* d_kps, uint8_t* d_desc, int capacity, int* d_n_out);

/* Page-locked host memory for the caller's frame / result buffers (the batch entry points use pinned caller buffers in place;
 * this spares a caller that does not link the CUDA runtime a cudaHostAlloc of its own).  write_combined != 0 asks for
 * write-combined memory: meant for buffers the CPU only WRITES (camera frames on their way to the GPU) -- the device reads
 * them without snooping the CPU caches, CPU reads of such memory are very slow, so never use it for results.  The memory
 * is portable across the devices of the process.  The reference keeps frames in cv::Mat heap memory (src/Tracking.cc:239). */
int orbgpu_host_alloc(void** out, size_t bytes, int write_combined);
void orbgpu_host_free(void* p);

/* Input pixel format of the frames handed to every extraction entry point of this handle: 1 (gray, default),
 * 3 or 4 interleaved 8-bit channels; rgb_order != 0 means R,G,B[,A], else B,G,R[,A].  Colour frames are converted
 * on the device exactly as Tracking::GrabImageMonocular does on the host before it constructs the Frame
 * (cv::cvtColor CV_RGB2GRAY / BGR2GRAY / RGBA2GRAY / BGRA2GRAY, src/Tracking.cc:244-257; OpenCV 4.13 fixed point);
 * strides are in bytes.  orbgpu_pyramid_level(0) then returns the gray image (mImGray). */
int orbgpu_set_input_format(orbgpu_extractor* e, int channels, int rgb_order);

/* The pinned-memory frame queue for Frame construction (north_star): submit copies the image into
 * a pinned ring slot and enqueues upload + kernels + download without blocking; collect blocks on
 * that frame only.  Frame::Frame (src/Frame.cc:174-228) or Tracking::GrabImage* one frame ahead
 * submits; Frame::ExtractORB (src/Frame.cc:247-253) collects.  Tickets are collected in any order;
 * at most orbgpu_queue_depth() tickets may be outstanding (else ORBGPU_ERR_STATE). */
int orbgpu_submit(orbgpu_extractor* e, const uint8_t* img, int w, int h, size_t stride, uint64_t* ticket);
int orbgpu_collect(orbgpu_extractor* e, uint64_t ticket, orbgpu_keypoint* kps, uint8_t* desc,
                   int capacity, int* n_out);
int orbgpu_queue_depth(orbgpu_extractor* e);

/* std::vector<cv::Mat> mvImagePyramid (include/ORBextractor.h:85; read by src/Frame.cc:473,563,575,580
 * in stereo mode): level pixels of the most recent orbgpu_extract / orbgpu_collect'ed frame, copied to
 * dst (dst may be NULL to query the size).  The 19-px REFLECT_101 frame of the reference's buffers is
 * rebuilt on the host by the C++ wrapper. */
int orbgpu_pyramid_level(orbgpu_extractor* e, int level, uint8_t* dst, size_t dst_stride, int* w, int* h);

/* Frame::ComputeStereoMatches (src/Frame.cc:466-640) between the most recent frame of `left` and of `right`
 * (two extractor handles on one device, each after its own orbgpu_extract / orbgpu_collect, as the stereo Frame
 * constructor runs them, src/Frame.cc:78-81).  Pyramids, keypoints and descriptors are used where they lie in HBM:
 * nothing is uploaded and mvImagePyramid need not be downloaded.  bf = mbf, b = mb.  Outputs for the n_left left
 * keypoints: u_right = mvuRight, depth = mvDepth (-1 where unmatched), after the median-SAD outlier rejection. */
int orbgpu_stereo_match(orbgpu_extractor* left, orbgpu_extractor* right, float bf, float b,
                        float* u_right, float* depth, int capacity, int* n_left);

/* Frame::UndistortKeyPoints (src/Frame.cc:404-435): the keypoints of the most recent frame of `e`, read where they
 * lie in HBM, with pt replaced by cv::undistortPoints(pt, K, D, noArray(), K) (OpenCV 4.13: 5 iterations in double);
 * the other fields are copied.  K4 = fx, fy, cx, cy of mK; dist = k1, k2, p1, p2[, k3] of mDistCoef (ndist = 4 or 5,
 * up to 12 accepted).  dist[0] == 0 returns the keypoints unchanged, as the reference does. */
int orbgpu_undistort_keypoints(orbgpu_extractor* e, const float* K4, const float* dist, int ndist,
                               orbgpu_keypoint* kps_un, int capacity, int* n_out);

/* orbgpu_extract followed by orbgpu_undistort_keypoints behind ONE synchronisation: what the monocular Frame constructor
 * does with each image (ExtractORB, then UndistortKeyPoints: src/Frame.cc:118-126, :404-435).  kps = mvKeys, kps_un =
 * mvKeysUn (both `capacity` records), desc = mDescriptors.  The frame stays the handle's "most recent frame" for
 * orbgpu_stereo_match / orbgpu_pyramid_level. */
int orbgpu_extract_undistorted(orbgpu_extractor* e, const uint8_t* img, int w, int h, size_t stride, const float* K4,
                               const float* dist, int ndist, orbgpu_keypoint* kps, orbgpu_keypoint* kps_un, uint8_t* desc,
                               int capacity, int* n_out);

/* cv::undistortPoints(xy, K, D, noArray(), K) on n arbitrary points (host buffers, x0 y0 x1 y1 ...): the four image
 * corners of Frame::ComputeImageBounds (src/Frame.cc:437-463), or any point list. */
int orbgpu_undistort_points(int device, const float* xy, int n, const float* K4, const float* dist, int ndist, float* out_xy);

/* ---- grid-gated searches between two frames ------------------------------------------
 * Plain-array snapshots of what the reference reads through Frame / MapPoint getters.  bounds4 = mnMinX, mnMaxX,
 * mnMinY, mnMaxY of the searched frame (its 64 x 48 grid, src/Frame.cc:230-245, is rebuilt on the device). */

/* ORBmatcher::SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize) (src/ORBmatcher.cc:405-520).
 * keys*_un = mvKeysUn, desc* = mDescriptors; prev_matched (n1 x 2, in/out) = vbPrevMatched; matches12[n1] = vnMatches12;
 * *nmatches = the return value.  nnratio / check_orientation = the matcher's mfNNratio / mbCheckOrientation. */
int orbgpu_search_for_initialization(int device, const orbgpu_keypoint* keys1_un, const uint8_t* desc1, int n1,
                                     const orbgpu_keypoint* keys2_un, const uint8_t* desc2, int n2, const float* bounds4,
                                     float* prev_matched, int window_size, float nnratio, int check_orientation,
                                     int32_t* matches12, int* nmatches);

/* ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono) (src/ORBmatcher.cc:1328-1470). */
typedef struct orbgpu_projection_args {
    int32_t n_last;
    const float* world_pos;        /* n_last x 3: pMP->GetWorldPos() */
    const uint8_t* mp_desc;        /* n_last x 32: pMP->GetDescriptor() */
    const uint8_t* mp_valid;       /* LastFrame.mvpMapPoints[i] && !LastFrame.mvbOutlier[i] */
    const uint8_t* mp_observed;    /* pMP->Observations() > 0 */
    const int32_t* last_octave;    /* LastFrame.mvKeys[i].octave */
    const float* last_angle;       /* LastFrame.mvKeysUn[i].angle */
    const float* Tcw_last;         /* LastFrame.mTcw, 4 x 4 row-major */
    int32_t n_cur;
    const orbgpu_keypoint* cur_keys_un;   /* CurrentFrame.mvKeysUn */
    const uint8_t* cur_desc;       /* CurrentFrame.mDescriptors */
    const float* cur_u_right;      /* CurrentFrame.mvuRight */
    const uint8_t* cur_occupied;   /* CurrentFrame.mvpMapPoints[i2] && ->Observations() > 0, on entry */
    const float* Tcw_cur;          /* CurrentFrame.mTcw */
    float fx, fy, cx, cy, bf, b;   /* CurrentFrame.fx .. mbf, mb */
    float bounds[4];               /* mnMinX, mnMaxX, mnMinY, mnMaxY */
    const float* scale_factors;    /* CurrentFrame.mvScaleFactors[nlevels] */
    int32_t nlevels;
    float th;
    int32_t mono;                  /* bMono */
    int32_t check_orientation;     /* mbCheckOrientation */
} orbgpu_projection_args;
/* cur_match[n_cur]: what the call did to CurrentFrame.mvpMapPoints[i2]: -1 untouched, i >= 0 set to the map point of
 * last-frame keypoint i, -2 set and then reset to NULL by the rotation-consistency check.  *nmatches = return value. */
int orbgpu_search_by_projection(int device, const orbgpu_projection_args* args, int32_t* cur_match, int* nmatches);

/* ORBmatcher::SearchByProjection(Frame& F, const vector<MapPoint*>& vpMapPoints, th) (src/ORBmatcher.cc:44-125): the
 * local-map search of Tracking::SearchLocalPoints (src/Tracking.cc:1150-1200), run on every tracked frame.  The map
 * points arrive as the plain arrays Frame::isInFrustum (src/Frame.cc:269-325, stays host) left in them. */
typedef struct orbgpu_local_points_args {
    int32_t n_points;
    const uint8_t* in_view;        /* pMP->mbTrackInView && !pMP->isBad() */
    const float* proj_x;           /* pMP->mTrackProjX */
    const float* proj_y;           /* pMP->mTrackProjY */
    const float* proj_xr;          /* pMP->mTrackProjXR */
    const int32_t* scale_level;    /* pMP->mnTrackScaleLevel, 0 <= level < nlevels where in_view */
    const float* view_cos;         /* pMP->mTrackViewCos */
    const uint8_t* mp_desc;        /* n_points x 32: pMP->GetDescriptor() */
    const uint8_t* mp_observed;    /* pMP->Observations() > 0 */
    int32_t n_cur;
    const orbgpu_keypoint* cur_keys_un;   /* F.mvKeysUn */
    const uint8_t* cur_desc;       /* F.mDescriptors */
    const float* cur_u_right;      /* F.mvuRight */
    const uint8_t* cur_occupied;   /* F.mvpMapPoints[idx] && ->Observations() > 0, on entry */
    float bounds[4];               /* mnMinX, mnMaxX, mnMinY, mnMaxY */
    const float* scale_factors;    /* F.mvScaleFactors[nlevels] */
    int32_t nlevels;
    float th;
    float nnratio;                 /* mfNNratio */
} orbgpu_local_points_args;
/* cur_match[n_cur]: -1 untouched, i >= 0: F.mvpMapPoints[idx] was set to vpMapPoints[i].  *nmatches = return value. */
int orbgpu_search_local_points(int device, const orbgpu_local_points_args* args, int32_t* cur_match, int* nmatches);

/* Tracking::SearchLocalPoints (src/Tracking.cc:1150-1200) in one call: Frame::isInFrustum (src/Frame.cc:269-325, with
 * MapPoint::PredictScale, src/MapPoint.cc:402-417) for every candidate local map point, then
 * ORBmatcher::SearchByProjection(F, vpLocalMapPoints, th) (src/ORBmatcher.cc:44-125) on what it left in the points. */
typedef struct orbgpu_local_map_args {
    int32_t n_points;
    const uint8_t* candidate;      /* pMP->mnLastFrameSeen != F.mnId && !pMP->isBad() (src/Tracking.cc:1177-1180) */
    const float* world_pos;        /* n_points x 3: pMP->GetWorldPos() */
    const float* normal;           /* n_points x 3: pMP->GetNormal() */
    const float* max_distance;     /* pMP->mfMaxDistance ( = GetMaxDistanceInvariance() / 1.2f ) */
    const float* min_distance;     /* pMP->mfMinDistance */
    const uint8_t* mp_desc;        /* n_points x 32: pMP->GetDescriptor() */
    const uint8_t* mp_observed;    /* pMP->Observations() > 0 */
    const float* Tcw;              /* F.mTcw, 4 x 4 row-major */
    float fx, fy, cx, cy, bf;      /* F.fx .. F.mbf */
    float viewing_cos_limit;       /* 0.5 at src/Tracking.cc:1182 */
    float log_scale_factor;        /* F.mfLogScaleFactor */
    int32_t n_cur;
    const orbgpu_keypoint* cur_keys_un;   /* F.mvKeysUn */
    const uint8_t* cur_desc;       /* F.mDescriptors */
    const float* cur_u_right;      /* F.mvuRight */
    const uint8_t* cur_occupied;   /* F.mvpMapPoints[idx] && ->Observations() > 0, on entry */
    float bounds[4];               /* mnMinX, mnMaxX, mnMinY, mnMaxY */
    const float* scale_factors;    /* F.mvScaleFactors[nlevels] */
    int32_t nlevels;
    float th;
    float nnratio;                 /* mfNNratio (0.8 at src/Tracking.cc:1191) */
} orbgpu_local_map_args;
/* what isInFrustum leaves in each map point (arrays of n_points); where in_view = 0 the reference leaves the members
 * untouched and the other five entries are 0 */
typedef struct orbgpu_track_fields {
    uint8_t* in_view;              /* mbTrackInView = the return value of isInFrustum (caller: IncreaseVisible()) */
    float* proj_x;                 /* mTrackProjX */
    float* proj_y;                 /* mTrackProjY */
    float* proj_xr;                /* mTrackProjXR */
    int32_t* scale_level;          /* mnTrackScaleLevel */
    float* view_cos;               /* mTrackViewCos */
} orbgpu_track_fields;
/* *n_in_view = nToMatch; the search runs when it is > 0 (src/Tracking.cc:1189); cur_match / *nmatches as for
 * orbgpu_search_local_points.  n_cur == 0 makes this a batched isInFrustum. */
int orbgpu_search_local_map(int device, const orbgpu_local_map_args* args, const orbgpu_track_fields* out, int* n_in_view,
                            int32_t* cur_match, int* nmatches);

/* The search inside ORBmatcher::Fuse(KeyFrame* pKF, const vector<MapPoint*>& vpMapPoints, th) (src/ORBmatcher.cc:825-957,
 * LocalMapping::SearchInNeighbors): projection into pKF, image / distance / viewing-angle tests, PredictScale,
 * KeyFrame::GetFeaturesInArea (src/KeyFrame.cc:569-608) and the chi-square gated best-descriptor scan (:892-940) for every
 * map point at once.  best_idx[i] = the keypoint of pKF map point i fuses with (bestDist <= TH_LOW), else -1.  The map
 * surgery that follows (:943-961: Replace / AddObservation / AddMapPoint) and the entry tests that depend on it
 * (isBad(), IsInKeyFrame(pKF), :850-851) stay host and run in index order over best_idx. */
typedef struct orbgpu_fuse_args {
    int32_t n_points;
    const uint8_t* present;        /* vpMapPoints[i] != NULL */
    const float* world_pos;        /* n_points x 3: GetWorldPos() */
    const float* normal;           /* n_points x 3: GetNormal() */
    const float* max_distance;     /* mfMaxDistance */
    const float* min_distance;     /* mfMinDistance */
    const uint8_t* mp_desc;        /* n_points x 32: GetDescriptor() */
    const float* Tcw;              /* pKF pose, 4 x 4 row-major (GetRotation / GetTranslation) */
    float Ow[3];                   /* pKF->GetCameraCenter() */
    float fx, fy, cx, cy, bf;      /* pKF->fx .. mbf */
    float log_scale_factor;        /* pKF->mfLogScaleFactor */
    int32_t n_kf;                  /* pKF->N */
    const orbgpu_keypoint* keys_un;/* pKF->mvKeysUn */
    const uint8_t* desc;           /* pKF->mDescriptors */
    const float* u_right;          /* pKF->mvuRight */
    float bounds[4];               /* Frame::mnMinX, mnMaxX, mnMinY, mnMaxY the keyframe's grid was built with; the keyframe
                                      itself keeps them truncated to int (include/KeyFrame.h:185-188), as the window does */
    const float* scale_factors;    /* pKF->mvScaleFactors[nlevels] */
    const float* inv_level_sigma2; /* pKF->mvInvLevelSigma2[nlevels] */
    int32_t nlevels;
    float th;
} orbgpu_fuse_args;
int orbgpu_fuse_search(int device, const orbgpu_fuse_args* args, int32_t* best_idx);
/* The same for n_jobs (keyframe, point set) problems in one call: LocalMapping::SearchInNeighbors runs Fuse once per target
 * keyframe (src/LocalMapping.cc:485-492; the call back into the current keyframe, :516, is one more job), each an independent search.  All jobs are enqueued back
 * to back and waited for once; job p writes n_points entries of best_idx right after those of job p - 1. */
int orbgpu_fuse_search_batch(int device, const orbgpu_fuse_args* args, int n_jobs, int32_t* best_idx);

/* ORBmatcher::SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, const set<MapPoint*>& sAlreadyFound, th, ORBdist)
 * (src/ORBmatcher.cc:1472-1597), the guided search of Tracking::Relocalization (src/Tracking.cc:1459, :1473). */
typedef struct orbgpu_reloc_projection_args {
    int32_t n_kf;                  /* pKF->GetMapPointMatches().size() */
    const uint8_t* mp_valid;       /* pMP && !pMP->isBad() && !sAlreadyFound.count(pMP) */
    const float* world_pos;        /* n_kf x 3: pMP->GetWorldPos() */
    const float* max_distance;     /* pMP->mfMaxDistance */
    const float* min_distance;     /* pMP->mfMinDistance */
    const uint8_t* mp_desc;        /* n_kf x 32: pMP->GetDescriptor() */
    const float* kf_angle;         /* pKF->mvKeysUn[i].angle */
    int32_t n_cur;
    const orbgpu_keypoint* cur_keys_un;   /* CurrentFrame.mvKeysUn */
    const uint8_t* cur_desc;       /* CurrentFrame.mDescriptors */
    const uint8_t* cur_occupied;   /* CurrentFrame.mvpMapPoints[i2] != NULL, on entry */
    const float* Tcw_cur;          /* CurrentFrame.mTcw, 4 x 4 row-major */
    float fx, fy, cx, cy;          /* CurrentFrame.fx .. cy */
    float bounds[4];               /* mnMinX, mnMaxX, mnMinY, mnMaxY */
    const float* scale_factors;    /* CurrentFrame.mvScaleFactors[nlevels] */
    float log_scale_factor;        /* CurrentFrame.mfLogScaleFactor */
    int32_t nlevels;
    float th;
    int32_t orb_dist;              /* ORBdist */
    int32_t check_orientation;     /* mbCheckOrientation */
} orbgpu_reloc_projection_args;
/* cur_match[n_cur]: -1 untouched, i >= 0 set to the map point of keyframe feature i, -2 set and then reset to NULL by the
 * rotation-consistency check.  *nmatches = return value. */
int orbgpu_search_by_projection_keyframe(int device, const orbgpu_reloc_projection_args* args, int32_t* cur_match, int* nmatches);

/* The three Sim3 routines of loop closing, on the same engine and the same orbgpu_fuse_args snapshot (Tcw / Ow / bf /
 * u_right / inv_level_sigma2 are not read where a Sim3 replaces the pose; th may hold the reference's int th):
 *  - Fuse(KeyFrame* pKF, cv::Mat Scw, vpPoints, th, vpReplacePoint) (src/ORBmatcher.cc:969-1097): present[i] = !isBad() &&
 *    !spAlreadyFound.count(pMP); best_idx[i] = bestIdx where bestDist <= TH_LOW; the branch on pKF->GetMapPoint(bestIdx)
 *    (:1082-1094) stays host, in index order.
 *  - SearchByProjection(KeyFrame* pKF, cv::Mat Scw, vpPoints, vpMatched, th) (:290-403): matched_in[idx] = vpMatched[idx] != NULL
 *    on entry, present[i] = !isBad() && !spAlreadyFound.count(pMP); matched[n_kf] = the point index vpMatched[idx] was set to,
 *    else -1; *nmatches = return value.
 *  - SearchBySim3(pKF1, pKF2, vpMatches12, s12, R12, t12, th) (:1102-1326): a12 = the map points of pKF1 (present[i] = pMP &&
 *    !vbAlreadyMatched1[i] && !isBad(); Tcw = pose of pKF1) searched among pKF2's keypoints, a21 the other direction;
 *    match12[N1] = idx2 of the pairs both directions agree on (vpMatches12[i1] = vpMapPoints2[idx2]); *nfound = return value. */
int orbgpu_fuse_search_sim3(int device, const orbgpu_fuse_args* args, const float* Scw16, int32_t* best_idx);
int orbgpu_search_by_projection_sim3(int device, const orbgpu_fuse_args* args, const float* Scw16, const uint8_t* matched_in,
                                     int32_t* matched, int* nmatches);
int orbgpu_search_by_sim3(int device, const orbgpu_fuse_args* a12, const orbgpu_fuse_args* a21, float s12, const float* R12,
                          const float* t12, int32_t* match12, int* nfound);

/* ORBmatcher::SearchByBoW, both overloads (src/ORBmatcher.cc:159-288 KeyFrame/Frame, :522-655 KeyFrame/KeyFrame).
 * Side 1 = the keyframe whose features are the queries (pKF / pKF1), side 2 = the searched one (F / pKF2).  A DBoW2
 * FeatureVector is passed flattened: ascending node ids, node_start[n_nodes + 1], feature indices in the map's order
 * (DBoW2 itself stays on the reference path).  valid = map point present and not bad (all 1 for a Frame); angle =
 * keypoint angles.  strict_th_low: 0 = "bestDist1 <= TH_LOW" (:241), 1 = "bestDist1 < TH_LOW" (:600).
 * match12[n1] / match21[n2] = the accepted pairs after the rotation check, -1 elsewhere: overload 1 sets
 * vpMapPointMatches[idx2] to the map point of keyframe feature match21[idx2]; overload 2 sets vpMatches12[idx1] to the
 * map point of feature match12[idx1] of pKF2.  *nmatches = the return value. */
typedef struct orbgpu_bow_side {
    const uint8_t* desc; int32_t n; const uint8_t* valid; const float* angle;
    const uint32_t* node_ids; const int32_t* node_start; const uint32_t* indices; int32_t n_nodes;
} orbgpu_bow_side;
int orbgpu_search_by_bow(int device, const orbgpu_bow_side* side1, const orbgpu_bow_side* side2, float nnratio,
                         int check_orientation, int strict_th_low, int32_t* match12, int32_t* match21, int* nmatches);

/* ORBmatcher::SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs, bOnlyStereo) + CheckDistEpipolarLine
 * (src/ORBmatcher.cc:657-820, :140-157): the BoW-gated, epipolar-checked search LocalMapping::CreateNewMapPoints runs
 * against every neighbouring keyframe.  FeatureVectors flattened as for orbgpu_search_by_bow; F12 is computed by the
 * caller (LocalMapping::ComputeF12 stays host). */
typedef struct orbgpu_triangulation_side {
    const orbgpu_keypoint* keys_un;  /* pKF->mvKeysUn */
    const uint8_t* desc;             /* pKF->mDescriptors */
    int32_t n;                       /* pKF->N */
    const uint8_t* has_map_point;    /* pKF->GetMapPoint(idx) != NULL */
    const float* u_right;            /* pKF->mvuRight */
    const uint32_t* node_ids; const int32_t* node_start; const uint32_t* indices; int32_t n_nodes;   /* pKF->mFeatVec */
} orbgpu_triangulation_side;
typedef struct orbgpu_triangulation_args {
    orbgpu_triangulation_side kf1, kf2;
    float F12[9];                    /* row-major 3 x 3 */
    float Cw1[3];                    /* pKF1->GetCameraCenter() */
    float T2w[16];                   /* pose of pKF2, 4 x 4 row-major (GetRotation / GetTranslation) */
    float fx, fy, cx, cy;            /* pKF2->fx .. cy */
    const float* scale_factors2;     /* pKF2->mvScaleFactors[nlevels] */
    const float* level_sigma2;       /* pKF2->mvLevelSigma2[nlevels] */
    int32_t nlevels;
    int32_t only_stereo;             /* bOnlyStereo */
    int32_t check_orientation;       /* mbCheckOrientation */
} orbgpu_triangulation_args;
/* matches12[kf1.n] = vMatches12 after the rotation check (vMatchedPairs = the pairs (i, matches12[i]) with matches12[i]
 * >= 0, ascending i); *nmatches = the return value. */
int orbgpu_search_for_triangulation(int device, const orbgpu_triangulation_args* args, int32_t* matches12, int* nmatches);
/* The same for n_pairs keyframe pairs in one call: LocalMapping::CreateNewMapPoints runs the search once per neighbouring
 * keyframe (src/LocalMapping.cc:239-270), each an independent problem.  Uploads, kernels and downloads of all pairs are
 * enqueued back to back and waited for once.  Pair p writes kf1.n entries of matches12 right after those of pair p - 1
 * and nmatches[p]. */
int orbgpu_search_for_triangulation_batch(int device, const orbgpu_triangulation_args* args, int n_pairs, int32_t* matches12,
                                          int* nmatches);

/* ---- resident keyframe store ------------------------------------------------------------------------------------
 * The reference passes KeyFrame objects; a KeyFrame's mvKeysUn, mDescriptors, mvuRight, mFeatVec and level tables are
 * written once (src/KeyFrame.cc:32-58, ComputeBoW :66-76) and read by every later search.  orbgpu_kf_add uploads them once;
 * the searches below name keyframes by id and send only what changes between calls (map-point flags, poses, F12), so
 * SearchByBoW / SearchForTriangulation stop re-staging 100-250 KB per call.  Ids are small integers, reused after
 * orbgpu_kf_erase.  Add / erase / search may be called from different threads (Tracking, LocalMapping, LoopClosing);
 * erasing a keyframe while a search that names it is running is the caller's race. */
typedef struct orbgpu_kfstore orbgpu_kfstore;
typedef struct orbgpu_kf_data {
    int32_t n;                        /* pKF->N */
    const orbgpu_keypoint* keys_un;   /* pKF->mvKeysUn */
    const uint8_t* desc;              /* pKF->mDescriptors, n x 32 */
    const float* u_right;             /* pKF->mvuRight; NULL = monocular (all -1) */
    const uint32_t* node_ids; const int32_t* node_start; const uint32_t* indices; int32_t n_nodes;   /* pKF->mFeatVec, flattened
                                         as for orbgpu_search_by_bow (call after ComputeBoW) */
    const float* scale_factors;       /* pKF->mvScaleFactors[nlevels]; read by the triangulation search only (NULL = zeros) */
    const float* level_sigma2;        /* pKF->mvLevelSigma2[nlevels]; likewise */
    int32_t nlevels;
    float fx, fy, cx, cy;             /* pKF->fx .. cy */
} orbgpu_kf_data;
int orbgpu_kfstore_create(orbgpu_kfstore** out, int device);
void orbgpu_kfstore_destroy(orbgpu_kfstore* s);
int orbgpu_kfstore_size(orbgpu_kfstore* s);                                  /* live keyframes */
int orbgpu_kf_add(orbgpu_kfstore* s, const orbgpu_kf_data* kf, int32_t* id); /* returns when the keyframe is resident */
int orbgpu_kf_erase(orbgpu_kfstore* s, int32_t id);                          /* KeyFrame::SetBadFlag / map reset */

/* ORBmatcher::SearchByBoW with side 1 = stored keyframe id1 (valid1[n1] as in orbgpu_bow_side).  Side 2 is either a
 * stored keyframe (id2 >= 0, valid2[n2], frame2 = NULL: the KeyFrame/KeyFrame overload, src/ORBmatcher.cc:522-655) or a
 * Frame given by its arrays (id2 = -1, valid2 ignored, frame2 != NULL: the KeyFrame/Frame overload, :159-288).
 * Outputs as orbgpu_search_by_bow. */
int orbgpu_kf_search_by_bow(orbgpu_kfstore* s, int32_t id1, const uint8_t* valid1, int32_t id2, const uint8_t* valid2,
                            const orbgpu_bow_side* frame2, float nnratio, int check_orientation, int strict_th_low,
                            int32_t* match12, int32_t* match21, int* nmatches);

/* ORBmatcher::SearchForTriangulation between two stored keyframes (src/ORBmatcher.cc:657-820); the camera and level
 * tables of pKF2 come from the store.  Outputs as orbgpu_search_for_triangulation(_batch). */
typedef struct orbgpu_kf_triangulation_job {
    int32_t id1, id2;
    const uint8_t* has_map_point1;   /* pKF1->GetMapPoint(idx) != NULL, n1 flags */
    const uint8_t* has_map_point2;
    float F12[9];                    /* row-major 3 x 3 (LocalMapping::ComputeF12 stays host) */
    float Cw1[3];                    /* pKF1->GetCameraCenter() */
    float T2w[16];                   /* pose of pKF2, 4 x 4 row-major */
    int32_t only_stereo;             /* bOnlyStereo */
    int32_t check_orientation;       /* mbCheckOrientation */
} orbgpu_kf_triangulation_job;
int orbgpu_kf_search_for_triangulation(orbgpu_kfstore* s, const orbgpu_kf_triangulation_job* job, int32_t* matches12, int* nmatches);
int orbgpu_kf_search_for_triangulation_batch(orbgpu_kfstore* s, const orbgpu_kf_triangulation_job* jobs, int n_jobs,
                                             int32_t* matches12, int* nmatches);

/* Per-stage views of the most recent orbgpu_extract call, for parity tests of the intermediate
 * results (SURVEY.md section 4 test plan): blurred level (GaussianBlur, src/ORBextractor.cc:1088),
 * FAST candidates in list order (:791-831) and octree survivors in list order (:541-765); x, y are
 * relative to the (16,16) cell-grid origin as at :824-825.  Return the count or a negative code. */
int orbgpu_debug_level_blurred(orbgpu_extractor* e, int level, uint8_t* dst, size_t dst_stride);
int orbgpu_debug_level_candidates(orbgpu_extractor* e, int level, int* x, int* y, int* score, int cap);
int orbgpu_debug_level_selected(orbgpu_extractor* e, int level, int* x, int* y, int* score, int cap);

/* ORBextractor::DistributeOctTree (src/ORBextractor.cc:541-765) run by the octree kernel on an
 * arbitrary candidate list (coordinates relative to the cell-grid origin, 0 <= x < W <= 4095,
 * 0 <= y < H <= 4095, 0 <= resp <= 255); out_idx receives indices into the input list in result
 * order.  Returns the count or a negative code. */
int orbgpu_octree(orbgpu_extractor* e, const int* xs, const int* ys, const int* resp, int n,
                  int W, int H, int N, int* out_idx, int cap);

/* Test hooks for the float steps that decide descriptor bits: cv::fastAtan2 (src/ORBextractor.cc:105)
 * and cos/sin of the keypoint angle (:115) exactly as the kernels evaluate them.  Host arrays.
 * sincos_checksum folds (x, sin x, cos x) of the float bit patterns first_bits + i*step, i < count,
 * into an order-independent 64-bit sum (oracle/orb_prims.c computes the same sum with glibc's sincosf). */
int orbgpu_debug_atan2(orbgpu_extractor* e, const float* y, const float* x, int n, float* out_deg);
int orbgpu_debug_sincosf(orbgpu_extractor* e, const float* rad, int n, float* out_sin, float* out_cos);
int orbgpu_debug_sincos_checksum(orbgpu_extractor* e, uint32_t first_bits, uint32_t count, uint32_t step, uint64_t* checksum);

/* Test hook for the float step that decides MapPoint::PredictScale (src/MapPoint.cc:410): glibc's logf exactly as
 * k_frustum evaluates it.  Host arrays. */
int orbgpu_debug_logf(int device, const float* x, int n, float* out);

/* Profiling: with profiling on, CUDA events bracket every stage of each chunk on its stream.
 * orbgpu_get_stage_ms returns accumulated milliseconds per stage since the last reset:
 * [0] upload, [1] pyramid, [2] FAST cells, [3] octree, [4] blur, [5] orientation+descriptors,
 * [6] download; launches = kernels launched by this handle since creation. */
#define ORBGPU_NUM_STAGES 7
int orbgpu_set_profiling(orbgpu_extractor* e, int on);
int orbgpu_get_stage_ms(orbgpu_extractor* e, double* ms, int n, int reset);
uint64_t orbgpu_launch_count(orbgpu_extractor* e);
/* Device time (ms) of the most recent orbgpu_extract / _batch / _batch_device call: CUDA events on the
 * library's own streams, from before the first upload/kernel to after the last kernel/download. */
double orbgpu_last_batch_ms(orbgpu_extractor* e);
/* How a batch call is cut up: chunks of at most max_chunk_frames frames are dealt round-robin to
 * `streams` internal CUDA streams (1..orbgpu_queue_depth()), so upload, kernels and download of
 * successive chunks overlap.  streams = 1 serialises the chunks (per-stage event times are then those of
 * each kernel running alone).  Small chunks keep a chunk's pyramid + blurred levels inside L2. */
int orbgpu_set_batching(orbgpu_extractor* e, int streams, int max_chunk_frames);
/* Largest chunk (frames) the device buffers of one stream context hold for this shape: min(128, 1.5 GB / bytes per frame). */
int orbgpu_chunk_capacity(orbgpu_extractor* e, int w, int h);

/* Host-buffer entry points (orbgpu_extract, _batch, _batch_async, orbgpu_submit) replay the kernels of a chunk from a CUDA
 * graph captured on first use per (stream context, frame count): one graph launch per chunk instead of 11 kernel launches
 * and 4 event calls from the host thread.  on = 0 switches back to direct launches (default on; profiling mode always
 * launches directly so that the stage events bracket one kernel each).  Results are identical either way. */
int orbgpu_set_graphs(orbgpu_extractor* e, int on);

/* ---- matcher ------------------------------------------------------------------------------- */

/* ORBmatcher::DescriptorDistance(a, b), include/ORBmatcher.h:44, src/ORBmatcher.cc:1647-1663:
 * Hamming distance of two 256-bit descriptors (host, hardware popcount). */
int orbgpu_hamming256(const void* a, const void* b);

/* The best / second-best scan shared by the Search* routines (canonical form
 * src/ORBmatcher.cc:201-226): every row of A (nA x 32 B) against every row of B in index order,
 * strict '<' updates, so ties keep the lowest index; best/second start at 256, best_idx at -1.
 * Host buffers; device = CUDA ordinal. */
int orbgpu_match_bruteforce(int device, const uint8_t* A, int nA, const uint8_t* B, int nB,
                            int32_t* best_idx, uint16_t* best, uint16_t* second);

/* orbgpu_match_bruteforce without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same arguments, same kernel and bit-identical results as orbgpu_match_bruteforce, but the device
 * staging area is a fixed file-scope __device__ array of the library (288 KB, held by every GPU that loads the
 * library until exit) instead of scratch grown with cudaMalloc / cudaHostAlloc: the call makes no cudaMalloc,
 * cudaFree, cudaHostAlloc, stream / event creation or other allocating call; copies are plain cudaMemcpy on the
 * legacy default stream.  Limit: nA and nB at most ORBGPU_NOALLOC_MAX_DESCRIPTORS each, beyond it the call returns
 * ORBGPU_ERR_CAPACITY (text in orbgpu_last_cuda_error()) and touches nothing.  0 <= device < 64.  All calls on
 * one device share the array, which keeps the previous call's contents (every byte a call reads is written by that
 * call first); calls on the same device that overlap in time (two host threads) race on it -- serialise them. */
#define ORBGPU_NOALLOC_MAX_DESCRIPTORS 4096
int orbgpu_match_bruteforce_noalloc(int device, const uint8_t* A, int nA, const uint8_t* B, int nB,
                                    int32_t* best_idx, uint16_t* best, uint16_t* second);

/* All-pairs keyframe matching (BASELINE config 5).  A matcher holds K keyframes x n descriptors
 * resident in HBM; pair p = (a[p], b[p]) yields counts[p] = number of rows of keyframe a whose best
 * match in keyframe b passes best <= th_low && float(best) < ratio*float(second)
 * (accept rule of src/ORBmatcher.cc:228-232 with TH_LOW, mfNNratio). */
int orbgpu_matcher_create(orbgpu_matcher** out, int device);
void orbgpu_matcher_destroy(orbgpu_matcher* m);
int orbgpu_matcher_set_keyframes(orbgpu_matcher* m, const uint8_t* desc, int num_keyframes, int n_per_keyframe);
int orbgpu_matcher_set_keyframes_device(orbgpu_matcher* m, const uint8_t* d_desc, int num_keyframes, int n_per_keyframe);
int orbgpu_matcher_count_pairs(orbgpu_matcher* m, const int32_t* kf_a, const int32_t* kf_b, int num_pairs,
                               int th_low, float ratio, int32_t* counts);
/* device-resident variant: d_pairs = num_pairs x (a, b) int32 pairs, d_counts int32[num_pairs] */
int orbgpu_matcher_count_pairs_device(orbgpu_matcher* m, const int32_t* d_pairs, int num_pairs,
                                      int th_low, float ratio, int32_t* d_counts);
/* All-pairs keyframe matching sharded over GPUs (BASELINE config 5: "pair blocks sharded over 8 B200 with NCCL result
 * gather"; the one-extractor-per-thread pattern of src/Frame.cc:78-81 carried over to one matcher per GPU).  Every rank's
 * matcher holds ALL keyframes; the upper-triangular pair list (a < b, row-major, K (K - 1) / 2 pairs) is cut into blocks of
 * min(4096, pairs / (16 world)) pairs dealt round-robin to the ranks; each rank counts its pairs and the counts are
 * exchanged with ONE ncclAllGather enqueued on the kernel's own stream (no host synchronisation between the last tile
 * and the gather) and put in pair order by one scatter kernel.  d_counts = device int32[K (K - 1) / 2] on every rank;
 * ms3 (may be NULL) = kernel, gather + scatter, total milliseconds (CUDA events on that stream).  NCCL is bound at run
 * time (dlopen libnccl.so.2); without it the comm calls return ORBGPU_ERR_UNSUPPORTED.
 *  - one process per GPU: rank 0 calls orbgpu_comm_unique_id (128 bytes), hands the id to the other ranks by any means,
 *    every rank calls orbgpu_comm_init_rank and then orbgpu_matcher_all_pairs_sharded (collective: all ranks call it);
 *    comm == NULL runs the single-GPU form (no NCCL);
 *  - one process, n GPUs: orbgpu_comm_init_all(comms, devices, n) and orbgpu_matcher_all_pairs_multi from ONE host
 *    thread (matcher r on devices[r]; the gather is one ncclGroup). */
typedef struct orbgpu_comm orbgpu_comm;
int orbgpu_comm_unique_id(void* id128);
int orbgpu_comm_init_rank(orbgpu_comm** out, int device, int rank, int world, const void* id128);
int orbgpu_comm_init_all(orbgpu_comm** out, const int* devices, int n);
void orbgpu_comm_destroy(orbgpu_comm* c);
int orbgpu_matcher_all_pairs_sharded(orbgpu_matcher* m, orbgpu_comm* c, int th_low, float ratio, int32_t* d_counts, double* ms3);
int orbgpu_matcher_all_pairs_multi(orbgpu_matcher** ms, orbgpu_comm** cs, int n, int th_low, float ratio, int32_t** d_counts,
                                   double* ms3);
/* milliseconds of the most recent count_pairs kernel (CUDA events on its stream) */
double orbgpu_matcher_last_kernel_ms(orbgpu_matcher* m);

/* MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:242-303), batched over map points: LocalMapping calls it for
 * every map point a new keyframe touches (src/LocalMapping.cc ProcessNewKeyFrame / CreateNewMapPoints / SearchInNeighbors),
 * each call an N x N table of DescriptorDistance (:281) and a median per row.  Group g = the descriptors
 * desc[group_start[g] .. group_start[g + 1]) of one map point's observations, in the order the reference collects them
 * (its std::map<KeyFrame*, size_t> iteration, bad keyframes left out).  best_idx[g] = BestIdx: the row with the smallest
 * median distance (element int(0.5 * (N - 1)) of the sorted row, the row's own 0 included), the first one on ties; -1 for
 * an empty group (the reference returns without touching mDescriptor). */
int orbgpu_distinctive_descriptors(int device, const uint8_t* desc, const int32_t* group_start, int n_groups, int32_t* best_idx);

/* MapPoint::UpdateNormalAndDepth (src/MapPoint.cc:330-371), batched over map points: LocalMapping and Tracking call it next to
 * ComputeDistinctiveDescriptors for every map point a new keyframe touches.  Point g: world position pos[3 g ..], the camera
 * centres (KeyFrame::GetCameraCenter) of its observing keyframes centers[3 obs_start[g] .. 3 obs_start[g + 1]) in the order the
 * reference walks its std::map<KeyFrame*, size_t>, its reference keyframe's centre ref_center[3 g ..], ref_scale[g] =
 * pRefKF->mvScaleFactors[octave of the point's keypoint there], last_scale[g] = pRefKF->mvScaleFactors[mnScaleLevels - 1].
 * Out: normal[3 g ..] = mNormalVector, max_dist[g] = mfMaxDistance, min_dist[g] = mfMinDistance; a point without
 * observations keeps the values the three arrays held on entry (the reference returns early, :345-346).  OpenCV 4.13 arithmetic:
 * cv::norm in double, "normal + normali / norm" as cv::scaleAdd (a fused multiply-add with the scale narrowed to float). */
int orbgpu_update_normals_and_depths(int device, const float* pos, const int32_t* obs_start, const float* centers,
                                     const float* ref_center, const float* ref_scale, const float* last_scale, int n_points,
                                     float* normal, float* max_dist, float* min_dist);

/* ---- bag of words: the DBoW2 vocabulary tree ----------------------------------------------------
 * Frame::ComputeBoW / KeyFrame::ComputeBoW (src/Frame.cc:395-402, src/KeyFrame.cc:59-68) call
 * ORBVocabulary::transform(descriptors, BowVector, FeatureVector, 4)
 * (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1127-1194): every descriptor walks the k-ary tree from the root, at each
 * node to the child with the smallest FORB::distance (FORB.cpp:81-101), the first child on ties (:1237-1250); the leaf's
 * word id and weight go into the BowVector (BowVector.cpp:34-60), the ancestor `levelsup` levels above the leaves and the
 * feature index into the FeatureVector (FeatureVector.cpp:29-43); the BowVector is then normalised (BowVector.cpp:64-86).
 * The tree arrives as flat arrays in node-id order, node 0 = the root (its parent / descriptor / weight are ignored):
 *   parent[i]   id of the parent (children keep the order in which they appear, as m_nodes[pid].children.push_back does,
 *               TemplatedVocabulary.h:1389-1390)
 *   is_leaf[i]  the file's leaf flag: flagged nodes get consecutive word ids in node order (:1407-1414)
 *   desc[i*32]  node descriptor, weight[i] node weight (double, as parsed)
 * scoring / weighting are DBoW2's enums (BowVector.h:36-53): weighting 0 TF_IDF, 1 TF, 2 IDF, 3 BINARY; scoring
 * 0 L1_NORM ... 5 DOT_PRODUCT decides the norm applied by transform (L1, except L2 for L2_NORM and none for DOT_PRODUCT). */
typedef struct orbgpu_vocabulary orbgpu_vocabulary;
int orbgpu_vocabulary_create(orbgpu_vocabulary** out, int device, int k, int L, int scoring, int weighting, int n_nodes,
                             const int32_t* parent, const uint8_t* is_leaf, const uint8_t* desc, const double* weight);
/* TemplatedVocabulary::loadFromTextFile (TemplatedVocabulary.h:1338-1425), the ORBvoc.txt format: "k L scoring weighting",
 * then one line per node "parent is_leaf d0 .. d31 weight".  Same header check (:1358-1362 -> ORBGPU_ERR_ARG).  Empty lines
 * are skipped: the reference's `while(!f.eof())` turns a trailing newline into one more child of the root whose descriptor
 * is uninitialised memory (undefined there, not reproduced here). */
int orbgpu_vocabulary_load_text(orbgpu_vocabulary** out, int device, const char* path);
void orbgpu_vocabulary_destroy(orbgpu_vocabulary* v);
/* out6 = k, L, scoring, weighting, number of nodes (root included), number of words */
int orbgpu_vocabulary_info(const orbgpu_vocabulary* v, int* out6);
/* transform(features, BowVector, FeatureVector, levelsup) for n descriptors (n x 32 bytes, host).  Outputs, caller-sized:
 *   words[n], values[n]     the BowVector in ascending word order (std::map iteration order)
 *   node_ids[n], start[n+1], idx[n]   the FeatureVector in ascending node order, node j owning idx[start[j] .. start[j+1]),
 *                           feature indices ascending: the layout orbgpu_bow_side takes
 *   counts3 = BowVector size, FeatureVector size, number of feature indices (features with weight <= 0 are "stopped")
 * A descriptor whose walk ends on a leaf above the FeatureVector level (an unbalanced tree) leaves `nid` uninitialised in
 * the reference (:1150-1155); here it gets the leaf's own node id.  n <= 16384, else ORBGPU_ERR_CAPACITY. */
int orbgpu_vocabulary_transform(orbgpu_vocabulary* v, const uint8_t* desc, int n, int levelsup, uint32_t* words,
                                double* values, uint32_t* node_ids, int32_t* start, uint32_t* idx, int* counts3);
/* The same for a batch of frames in one call (a frame's ComputeBoW is independent of every other's): frame f owns the
 * descriptors [frame_start[f], frame_start[f + 1]) of desc (frame_start[0] = 0, at most 16384 per frame); its outputs go to
 * the same offsets of words / values / node_ids / idx, its start list (count + 1 entries) to start + frame_start[f] + f, so
 * start needs frame_start[n_frames] + n_frames entries; counts3[3 f .. 3 f + 2] as above.  One descriptor walk over all
 * frames, then one CTA per frame. */
int orbgpu_vocabulary_transform_batch(orbgpu_vocabulary* v, const uint8_t* desc, const int32_t* frame_start, int n_frames,
                                      int levelsup, uint32_t* words, double* values, uint32_t* node_ids, int32_t* start,
                                      uint32_t* idx, int* counts3);
/* Vocabulary::score(query, keyframe) for nk keyframes at once (L1Scoring::score, ScoringObject.cpp:25-72): the call
 * KeyFrameDatabase::DetectLoopCandidates / DetectRelocalizationCandidates make per candidate keyframe
 * (src/KeyFrameDatabase.cc:133, :249) and LoopClosing::DetectLoop per covisible keyframe (src/LoopClosing.cc:135).
 * BowVectors as ascending (word, value) arrays; keyframe k owns kw/kv[kstart[k] .. kstart[k+1]).  The terms
 * |vi - wi| - |vi| - |wi| of the common words are added in ascending word order, in double.  Only L1_NORM vocabularies
 * (the one ORB-SLAM2 ships); others return ORBGPU_ERR_UNSUPPORTED. */
int orbgpu_vocabulary_score(orbgpu_vocabulary* v, const uint32_t* qw, const double* qv, int nq, const uint32_t* kw,
                            const double* kv, const int32_t* kstart, int nk, double* scores);

/* KeyFrameDatabase (include/KeyFrameDatabase.h:44-75, src/KeyFrameDatabase.cc) with the keyframes' BowVectors resident on the
 * device.  add (:40-46) appends a BowVector (ascending (word, value) arrays) and returns its slot, slots count up from 0 in
 * insertion order -- the order in which the reference's inverted-file lists hold the keyframes; erase (:48-67) empties a
 * slot (its number is not reused); clear (:69-73) forgets everything.
 * orbgpu_bowdb_query is the data-parallel part of DetectLoopCandidates / DetectRelocalizationCandidates for one query
 * BowVector, over every slot s (arrays of orbgpu_bowdb_size() entries):
 *   common_words[s]  words the keyframe shares with the query = mnLoopWords / mnRelocWords after the walk of :86-105 / :207-222
 *                    (0: the keyframe is not in lKFsSharingWords; erased slots: 0)
 *   first_common[s]  position in the query of the first shared word (nq if none): the walk meets the keyframes in ascending
 *                    (first_common, slot) order, which is the order of lKFsSharingWords
 *   scores[s]        Vocabulary::score(query, keyframe), L1 (ScoringObject.cpp:25-72), for every slot; `float si = score`
 *                    (:133, :249) narrows it on the host
 * The connected-keyframe filter, the 0.8 * maxCommonWords and minScore thresholds and the covisibility accumulation
 * (:107-197, :224-309) walk the keyframe graph and stay with the caller (my-slam_b200/orbgpu.py KeyFrameDatabase). */
typedef struct orbgpu_bowdb orbgpu_bowdb;
int orbgpu_bowdb_create(orbgpu_bowdb** out, int device);
void orbgpu_bowdb_destroy(orbgpu_bowdb* db);
int orbgpu_bowdb_add(orbgpu_bowdb* db, const uint32_t* words, const double* values, int n, int* slot);
int orbgpu_bowdb_erase(orbgpu_bowdb* db, int slot);
int orbgpu_bowdb_clear(orbgpu_bowdb* db);
int orbgpu_bowdb_size(const orbgpu_bowdb* db);
int orbgpu_bowdb_query(orbgpu_bowdb* db, const uint32_t* qw, const double* qv, int nq, int32_t* common_words,
                       int32_t* first_common, double* scores);

/* Measured instruction-rate ceilings behind the roofline of the matching kernels (SURVEY.md section 8(d) asks for a
 * popc micro-benchmark beside the HBM figure; the reference has no counterpart): which = 0 thread-level POPC per second,
 * which = 1 thread-level LOP3 per second, best of three launches of a register-only kernel (8 CTAs of 256 threads per SM,
 * eight independent chains per thread), CUDA events on the launching stream.  k_match_* spends 5 POPC per 256-bit distance
 * (src/ORBmatcher.cc:1647-1663 spends 8 SWAR popcounts). */
int orbgpu_debug_pipe_peak(int device, int which, double* ops_per_s);

/* ---- misc ---------------------------------------------------------------------------------- */
const char* orbgpu_strerror(int code);
const char* orbgpu_last_cuda_error(void);   /* thread-local text of the last CUDA failure */
int orbgpu_device_count(void);              /* 0 when no CUDA device / driver */
const char* orbgpu_version(void);

#ifdef __cplusplus
}
#endif
#endif /* ORBGPU_H */
