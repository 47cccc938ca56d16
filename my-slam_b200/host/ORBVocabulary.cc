// ORBVocabulary.cc -- thin C++ layer over orbgpu_vocabulary_* / orbgpu_bowdb_* (include/orbgpu.h); see ORBVocabulary.h.
#include "ORBVocabulary.h"

#include <cstring>
#include <stdexcept>

namespace ORB_SLAM2 {

namespace {
void check(int rc, const char* what)
{
    if (rc == ORBGPU_OK) return;
    std::string msg = std::string(what) + ": " + orbgpu_strerror(rc);
    if (rc == ORBGPU_ERR_CUDA) msg += std::string(" (") + orbgpu_last_cuda_error() + ")";
    throw std::runtime_error(msg);
}

void flatten(const DBoW2::BowVector& v, std::vector<uint32_t>& w, std::vector<double>& x)
{
    w.reserve(w.size() + v.size()); x.reserve(x.size() + v.size());
    for (DBoW2::BowVector::const_iterator it = v.begin(); it != v.end(); ++it) { w.push_back(it->first); x.push_back(it->second); }
}
}  // namespace

ORBVocabulary::ORBVocabulary(int device) : mDevice(device), mHandle(NULL) {}
ORBVocabulary::~ORBVocabulary() { orbgpu_vocabulary_destroy(mHandle); }

bool ORBVocabulary::loadFromTextFile(const std::string& filename)
{
    orbgpu_vocabulary* h = NULL;
    const int rc = orbgpu_vocabulary_load_text(&h, mDevice, filename.c_str());
    if (rc == ORBGPU_ERR_ARG) return false;              // unreadable / "not a correct text file": the reference returns false
    check(rc, "orbgpu_vocabulary_load_text");
    orbgpu_vocabulary_destroy(mHandle);
    mHandle = h;
    return true;
}

unsigned int ORBVocabulary::size() const
{
    if (!mHandle) return 0;
    int info[6];
    check(orbgpu_vocabulary_info(mHandle, info), "orbgpu_vocabulary_info");
    return (unsigned int)info[5];
}

bool ORBVocabulary::empty() const { return size() == 0; }

void ORBVocabulary::transform_rows(const unsigned char* rows, int n, DBoW2::BowVector& v, DBoW2::FeatureVector& fv, int levelsup) const
{
    v.clear();
    fv.clear();
    if (!mHandle || n == 0) return;
    std::vector<uint32_t> words((size_t)n), nodes((size_t)n), idx((size_t)n);
    std::vector<double> values((size_t)n);
    std::vector<int32_t> start((size_t)n + 1);
    int counts[3];
    check(orbgpu_vocabulary_transform(mHandle, rows, n, levelsup, words.data(), values.data(), nodes.data(), start.data(), idx.data(), counts),
          "orbgpu_vocabulary_transform");
    for (int i = 0; i < counts[0]; ++i) v.insert(v.end(), std::make_pair(words[(size_t)i], values[(size_t)i]));       // ascending already
    for (int j = 0; j < counts[1]; ++j)
        fv.insert(fv.end(), std::make_pair(nodes[(size_t)j], std::vector<unsigned int>(idx.begin() + start[(size_t)j], idx.begin() + start[(size_t)j + 1])));
}

void ORBVocabulary::transform(const std::vector<cv::Mat>& features, DBoW2::BowVector& v, DBoW2::FeatureVector& fv, int levelsup) const
{
    std::vector<unsigned char> rows(features.size() * 32);
    for (size_t i = 0; i < features.size(); ++i) std::memcpy(&rows[i * 32], features[i].ptr(), 32);
    transform_rows(rows.data(), (int)features.size(), v, fv, levelsup);
}

void ORBVocabulary::transform(const cv::Mat& descriptors, DBoW2::BowVector& v, DBoW2::FeatureVector& fv, int levelsup) const
{
    if (descriptors.rows > 0 && descriptors.isContinuous()) { transform_rows(descriptors.ptr(), descriptors.rows, v, fv, levelsup); return; }
    std::vector<unsigned char> rows((size_t)descriptors.rows * 32);
    for (int i = 0; i < descriptors.rows; ++i) std::memcpy(&rows[(size_t)i * 32], descriptors.ptr(i), 32);
    transform_rows(rows.data(), descriptors.rows, v, fv, levelsup);
}

std::vector<double> ORBVocabulary::score(const DBoW2::BowVector& query, const std::vector<const DBoW2::BowVector*>& keyframes) const
{
    std::vector<double> out(keyframes.size());
    if (keyframes.empty()) return out;
    if (!mHandle) throw std::runtime_error("ORBVocabulary::score: no vocabulary loaded");
    std::vector<uint32_t> qw, kw; std::vector<double> qv, kv;
    flatten(query, qw, qv);
    std::vector<int32_t> ks(1, 0);
    for (size_t k = 0; k < keyframes.size(); ++k) { flatten(*keyframes[k], kw, kv); ks.push_back((int32_t)kw.size()); }
    check(orbgpu_vocabulary_score(mHandle, qw.data(), qv.data(), (int)qw.size(), kw.data(), kv.data(), ks.data(), (int)keyframes.size(), out.data()),
          "orbgpu_vocabulary_score");
    return out;
}

double ORBVocabulary::score(const DBoW2::BowVector& v1, const DBoW2::BowVector& v2) const
{
    return score(v1, std::vector<const DBoW2::BowVector*>(1, &v2))[0];
}

BowDatabase::BowDatabase(int device) : mHandle(NULL) { check(orbgpu_bowdb_create(&mHandle, device), "orbgpu_bowdb_create"); }
BowDatabase::~BowDatabase() { orbgpu_bowdb_destroy(mHandle); }

int BowDatabase::add(const DBoW2::BowVector& v)
{
    std::vector<uint32_t> w; std::vector<double> x;
    flatten(v, w, x);
    int slot = -1;
    check(orbgpu_bowdb_add(mHandle, w.data(), x.data(), (int)w.size(), &slot), "orbgpu_bowdb_add");
    return slot;
}

void BowDatabase::erase(int slot) { check(orbgpu_bowdb_erase(mHandle, slot), "orbgpu_bowdb_erase"); }
void BowDatabase::clear() { check(orbgpu_bowdb_clear(mHandle), "orbgpu_bowdb_clear"); }
int BowDatabase::size() const { return orbgpu_bowdb_size(mHandle); }

void BowDatabase::query(const DBoW2::BowVector& q, std::vector<int>& commonWords, std::vector<int>& firstCommon, std::vector<double>& scores) const
{
    std::vector<uint32_t> w; std::vector<double> x;
    flatten(q, w, x);
    const size_t n = (size_t)size();
    commonWords.assign(n, 0); firstCommon.assign(n, 0); scores.assign(n, 0.0);
    if (n == 0) return;
    check(orbgpu_bowdb_query(mHandle, w.data(), x.data(), (int)w.size(), commonWords.data(), firstCommon.data(), scores.data()), "orbgpu_bowdb_query");
}

}  // namespace ORB_SLAM2
