// ORBextractor.cc -- host side of the drop-in extractor: argument checks, the reference's side effects
// on its outputs (src/ORBextractor.cc:1045-1107) and buffer marshalling around the orbgpu_* C ABI.
// No pixel arithmetic happens here.
#include "ORBextractor.h"

#include <cassert>
#include <cstring>
#include <mutex>
#include <stdexcept>
#include <string>

#include "../../include/orbgpu.h"

namespace ORB_SLAM2 {

static_assert(sizeof(cv::KeyPoint) == sizeof(orbgpu_keypoint), "cv::KeyPoint must be the 28-byte POD of the C ABI");

const int EDGE_THRESHOLD = 19;    // src/ORBextractor.cc:76: width of the frame around every pyramid level

// live extractors of this process: the second one (a stereo pair) switches the pyramid download on for all of them
static std::mutex g_live_mu;
static std::vector<ORBextractor*> g_live;

ORBextractor::ORBextractor(int _nfeatures, float _scaleFactor, int _nlevels, int _iniThFAST, int _minThFAST)
    : ORBextractor(_nfeatures, _scaleFactor, _nlevels, _iniThFAST, _minThFAST, 0) {}

ORBextractor::ORBextractor(int _nfeatures, float _scaleFactor, int _nlevels, int _iniThFAST, int _minThFAST, int cudaDevice)
    : nfeatures(_nfeatures), scaleFactor(_scaleFactor), nlevels(_nlevels), iniThFAST(_iniThFAST), minThFAST(_minThFAST),
      mpHandle(0), mbDownloadPyramid(false), mbDownloadExplicit(false), mLastError("ok"), mLastW(0), mLastH(0), mChannels(1)
{
    const int rc = orbgpu_create(&mpHandle, cudaDevice, _nfeatures, _scaleFactor, _nlevels, _iniThFAST, _minThFAST);
    if (rc != ORBGPU_OK)      // the reference has no failure path here; there is no CPU path to fall back to
        throw std::runtime_error(std::string("ORBextractor: orbgpu_create failed: ") + orbgpu_strerror(rc) + " (" +
                                 orbgpu_last_cuda_error() + ")");
    // the tables are built by the library with the reference's float/double mix (src/ORBextractor.cc:417-471)
    std::vector<float> s4((size_t)4 * nlevels);
    mnFeaturesPerLevel.resize(nlevels);
    umax.resize(16);
    orbgpu_get_tables(mpHandle, s4.data(), mnFeaturesPerLevel.data(), umax.data());
    mvScaleFactor.assign(s4.begin(), s4.begin() + nlevels);
    mvInvScaleFactor.assign(s4.begin() + nlevels, s4.begin() + 2 * nlevels);
    mvLevelSigma2.assign(s4.begin() + 2 * nlevels, s4.begin() + 3 * nlevels);
    mvInvLevelSigma2.assign(s4.begin() + 3 * nlevels, s4.begin() + 4 * nlevels);
    mvImagePyramid.resize(nlevels);
    std::lock_guard<std::mutex> lk(g_live_mu);
    g_live.push_back(this);
    if (g_live.size() >= 2)
        for (size_t i = 0; i < g_live.size(); ++i)
            if (!g_live[i]->mbDownloadExplicit) g_live[i]->mbDownloadPyramid = true;
}

ORBextractor::~ORBextractor()
{
    {
        std::lock_guard<std::mutex> lk(g_live_mu);
        for (size_t i = 0; i < g_live.size(); ++i)
            if (g_live[i] == this) { g_live.erase(g_live.begin() + (long)i); break; }
    }
    orbgpu_destroy(mpHandle);
}

// keypoints.clear() + fill, descriptors.create(N,32,CV_8U) or release() when N == 0 (:1066-1074)
void ORBextractor::Finish(int n, std::vector<cv::KeyPoint>& keypoints, cv::OutputArray descriptors)
{
    keypoints.clear();
    if (n <= 0) { descriptors.release(); return; }
    keypoints.resize((size_t)n);
    std::memcpy((void*)keypoints.data(), mKpScratch.data(), sizeof(orbgpu_keypoint) * (size_t)n);
    descriptors.create(n, 32, CV_8U);
    cv::Mat d = descriptors.getMat();
    for (int i = 0; i < n; ++i) std::memcpy(d.ptr(i), mDescScratch.data() + (size_t)32 * i, 32);
}

void ORBextractor::operator()(cv::InputArray _image, cv::InputArray /*_mask*/, std::vector<cv::KeyPoint>& _keypoints,
                              cv::OutputArray _descriptors)
{
    if (_image.empty()) return;                            // :1048-1049: outputs untouched
    cv::Mat image = _image.getMat();
    assert(image.type() == (mChannels == 1 ? CV_8UC1 : mChannels == 3 ? CV_8UC3 : CV_8UC4));   // :1052 (CV_8UC1 in the reference)
    const int w = image.cols, h = image.rows;
    const int cap = orbgpu_max_keypoints(mpHandle, w, h);
    if (cap < 0) { mLastError = orbgpu_strerror(cap); throw std::runtime_error(std::string("ORBextractor: ") + mLastError); }
    mKpScratch.resize(sizeof(orbgpu_keypoint) * (size_t)(cap > 0 ? cap : 1));
    mDescScratch.resize((size_t)32 * (cap > 0 ? cap : 1));
    int n = 0;
    const int rc = orbgpu_extract(mpHandle, image.ptr(0), w, h, (size_t)image.step, (orbgpu_keypoint*)mKpScratch.data(),
                                  mDescScratch.data(), cap, &n);
    if (rc != ORBGPU_OK) { mLastError = orbgpu_strerror(rc); throw std::runtime_error(std::string("ORBextractor: ") + mLastError + " (" + orbgpu_last_cuda_error() + ")"); }
    mLastW = w; mLastH = h;
    Finish(n, _keypoints, _descriptors);
    if (mbDownloadPyramid) FillPyramid();
}

bool ORBextractor::SetColorInput(int channels, bool rgbOrder)
{
    const int rc = orbgpu_set_input_format(mpHandle, channels, rgbOrder ? 1 : 0);
    if (rc != ORBGPU_OK) { mLastError = orbgpu_strerror(rc); return false; }
    mChannels = channels;
    return true;
}

cv::Mat ORBextractor::GrayImage()
{
    int w = 0, h = 0;
    if (orbgpu_pyramid_level(mpHandle, 0, 0, 0, &w, &h) != ORBGPU_OK) return cv::Mat();
    cv::Mat g(h, w, CV_8UC1);
    orbgpu_pyramid_level(mpHandle, 0, g.ptr(0), (size_t)g.step, &w, &h);
    return g;
}

uint64_t ORBextractor::Submit(cv::InputArray _image)
{
    if (_image.empty()) return 0;
    cv::Mat image = _image.getMat();
    assert(image.channels() == mChannels);
    uint64_t ticket = 0;
    const int rc = orbgpu_submit(mpHandle, image.ptr(0), image.cols, image.rows, (size_t)image.step, &ticket);
    if (rc != ORBGPU_OK) { mLastError = orbgpu_strerror(rc); return 0; }
    mLastW = image.cols; mLastH = image.rows;
    return ticket;
}

bool ORBextractor::Collect(uint64_t ticket, std::vector<cv::KeyPoint>& keypoints, cv::OutputArray descriptors)
{
    const int cap = orbgpu_max_keypoints(mpHandle, mLastW, mLastH);
    if (cap < 0) { mLastError = orbgpu_strerror(cap); return false; }
    mKpScratch.resize(sizeof(orbgpu_keypoint) * (size_t)(cap > 0 ? cap : 1));
    mDescScratch.resize((size_t)32 * (cap > 0 ? cap : 1));
    int n = 0;
    const int rc = orbgpu_collect(mpHandle, ticket, (orbgpu_keypoint*)mKpScratch.data(), mDescScratch.data(), cap, &n);
    if (rc != ORBGPU_OK) { mLastError = orbgpu_strerror(rc); return false; }
    Finish(n, keypoints, descriptors);
    if (mbDownloadPyramid) FillPyramid();
    return true;
}

int ORBextractor::QueueDepth() const { return orbgpu_queue_depth(mpHandle); }

bool ORBextractor::ExtractBatch(const std::vector<cv::Mat>& frames, std::vector<std::vector<cv::KeyPoint> >& keypoints,
                                std::vector<cv::Mat>& descriptors)
{
    const int n = (int)frames.size();
    keypoints.assign((size_t)n, std::vector<cv::KeyPoint>());
    descriptors.assign((size_t)n, cv::Mat());
    if (n == 0) return true;
    const int w = frames[0].cols, h = frames[0].rows;
    for (int i = 0; i < n; ++i)
        if (frames[i].cols != w || frames[i].rows != h || frames[i].empty() || frames[i].channels() != mChannels) {
            mLastError = "ExtractBatch: frames must share one shape and the handle's channel count";
            return false;
        }
    const int cap = orbgpu_max_keypoints(mpHandle, w, h);
    if (cap < 0) { mLastError = orbgpu_strerror(cap); return false; }
    const size_t rb = (size_t)w * mChannels;                    // bytes per dense row (colour input: SetColorInput)
    std::vector<unsigned char> dense((size_t)n * rb * h);
    for (int i = 0; i < n; ++i)
        for (int y = 0; y < h; ++y) std::memcpy(&dense[((size_t)i * h + y) * rb], frames[i].ptr(y), rb);
    std::vector<orbgpu_keypoint> kps((size_t)n * cap);
    std::vector<unsigned char> desc((size_t)n * cap * 32);
    std::vector<int> nout((size_t)n);
    const int rc = orbgpu_extract_batch(mpHandle, dense.data(), n, w, h, rb, rb * h, kps.data(), desc.data(), cap, nout.data());
    if (rc != ORBGPU_OK) { mLastError = orbgpu_strerror(rc); return false; }
    for (int i = 0; i < n; ++i) {
        const int k = nout[i];
        if (k <= 0) continue;
        keypoints[i].resize((size_t)k);
        std::memcpy((void*)keypoints[i].data(), &kps[(size_t)i * cap], sizeof(orbgpu_keypoint) * (size_t)k);
        descriptors[i].create(k, 32, CV_8U);
        for (int r = 0; r < k; ++r) std::memcpy(descriptors[i].ptr(r), &desc[((size_t)i * cap + r) * 32], 32);
    }
    mLastW = w; mLastH = h;
    return true;
}

bool ORBextractor::ComputeStereoMatches(ORBextractor& left, ORBextractor& right, float bf, float b,
                                        std::vector<float>& mvuRight, std::vector<float>& mvDepth)
{
    const int cap = orbgpu_max_keypoints(left.mpHandle, left.mLastW, left.mLastH);
    if (cap < 0) { left.mLastError = orbgpu_strerror(cap); return false; }
    mvuRight.assign((size_t)(cap > 0 ? cap : 1), -1.0f);
    mvDepth.assign((size_t)(cap > 0 ? cap : 1), -1.0f);
    int n = 0;
    const int rc = orbgpu_stereo_match(left.mpHandle, right.mpHandle, bf, b, mvuRight.data(), mvDepth.data(), cap, &n);
    if (rc != ORBGPU_OK) { left.mLastError = orbgpu_strerror(rc); mvuRight.clear(); mvDepth.clear(); return false; }
    mvuRight.resize((size_t)n);
    mvDepth.resize((size_t)n);
    return true;
}

static inline int reflect101(int i, int n)
{
    if (n == 1) return 0;
    while (i < 0 || i >= n) i = i < 0 ? -i : 2 * (n - 1) - i;
    return i;
}

// mvImagePyramid[l] = ROI (19,19,w,h) of a (w+38) x (h+38) buffer whose frame is the REFLECT_101
// continuation of the level, as the reference builds it (src/ORBextractor.cc:1113-1135).
void ORBextractor::FillPyramid()
{
    for (int l = 0; l < nlevels; ++l) {
        int w = 0, h = 0;
        if (orbgpu_pyramid_level(mpHandle, l, 0, 0, &w, &h) != ORBGPU_OK) { mvImagePyramid[l] = cv::Mat(); continue; }
        const int E = EDGE_THRESHOLD;
        cv::Mat temp(h + 2 * E, w + 2 * E, CV_8UC1);
        cv::Mat roi = temp(cv::Rect(E, E, w, h));
        orbgpu_pyramid_level(mpHandle, l, roi.ptr(0), (size_t)roi.step, &w, &h);
        for (int y = 0; y < h; ++y) {                       // left / right frame of the level's own rows
            unsigned char* row = temp.ptr(y + E);
            for (int x = 0; x < E; ++x) {
                row[x] = row[E + reflect101(x - E, w)];
                row[E + w + x] = row[E + reflect101(w + x, w)];
            }
        }
        for (int y = 0; y < E; ++y) {                       // top / bottom frame rows
            std::memcpy(temp.ptr(y), temp.ptr(E + reflect101(y - E, h)), (size_t)w + 2 * E);
            std::memcpy(temp.ptr(E + h + y), temp.ptr(E + reflect101(h + y, h)), (size_t)w + 2 * E);
        }
        mvImagePyramid[l] = roi;
    }
}

}  // namespace ORB_SLAM2
