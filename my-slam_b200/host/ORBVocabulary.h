// ORBVocabulary.h -- drop-in for the reference's include/ORBVocabulary.h:30-31
//   typedef DBoW2::TemplatedVocabulary<DBoW2::FORB::TDescriptor, DBoW2::FORB> ORBVocabulary;
// restricted to what ORB-SLAM2 calls on it: loadFromTextFile (src/System.cc:66), transform(features, BowVector,
// FeatureVector, levelsup) (src/Frame.cc:400, src/KeyFrame.cc:66), score (src/KeyFrameDatabase.cc:133, :249,
// src/LoopClosing.cc:135), size (src/KeyFrameDatabase.cc:36).  The tree lives on the device (orbgpu_vocabulary_*, include/orbgpu.h);
// every call throws std::runtime_error when the CUDA library cannot run -- there is no CPU fallback.
// BowDatabase is the device half of KeyFrameDatabase (include/KeyFrameDatabase.h:44-75): see orbgpu_bowdb_* in include/orbgpu.h.
#ifndef ORBGPU_HOST_ORBVOCABULARY_H
#define ORBGPU_HOST_ORBVOCABULARY_H

#include <map>
#include <string>
#include <vector>

#include "cv_compat.h"
#include "../../include/orbgpu.h"

#ifdef ORBGPU_USE_DBOW2                      // inside the reference tree: its own BowVector / FeatureVector
#include "Thirdparty/DBoW2/DBoW2/BowVector.h"
#include "Thirdparty/DBoW2/DBoW2/FeatureVector.h"
#else
namespace DBoW2 {                            // Thirdparty/DBoW2/DBoW2/BowVector.h:20-26, :56-58; FeatureVector.h:21-23
typedef unsigned int WordId;
typedef double WordValue;
typedef unsigned int NodeId;
class BowVector : public std::map<WordId, WordValue> {};
class FeatureVector : public std::map<NodeId, std::vector<unsigned int> > {};
}  // namespace DBoW2
#endif

namespace ORB_SLAM2 {

class ORBVocabulary {
public:
    explicit ORBVocabulary(int device = 0);
    virtual ~ORBVocabulary();
    bool loadFromTextFile(const std::string& filename);
    virtual unsigned int size() const;       // number of words
    virtual bool empty() const;
    virtual void transform(const std::vector<cv::Mat>& features, DBoW2::BowVector& v, DBoW2::FeatureVector& fv, int levelsup) const;
    // the same on the N x 32 descriptor matrix itself (Frame::mDescriptors), without Converter::toDescriptorVector
    void transform(const cv::Mat& descriptors, DBoW2::BowVector& v, DBoW2::FeatureVector& fv, int levelsup) const;
    double score(const DBoW2::BowVector& v1, const DBoW2::BowVector& v2) const;
    // score(query, each of the keyframes) in one device call
    std::vector<double> score(const DBoW2::BowVector& query, const std::vector<const DBoW2::BowVector*>& keyframes) const;
    orbgpu_vocabulary* Handle() const { return mHandle; }

private:
    ORBVocabulary(const ORBVocabulary&);
    ORBVocabulary& operator=(const ORBVocabulary&);
    void transform_rows(const unsigned char* rows, int n, DBoW2::BowVector& v, DBoW2::FeatureVector& fv, int levelsup) const;
    int mDevice;
    orbgpu_vocabulary* mHandle;
};

class BowDatabase {
public:
    explicit BowDatabase(int device = 0);
    ~BowDatabase();
    int add(const DBoW2::BowVector& v);      // KeyFrameDatabase::add: returns the keyframe's slot
    void erase(int slot);                    // KeyFrameDatabase::erase
    void clear();                            // KeyFrameDatabase::clear
    int size() const;
    // for every slot: words shared with the query (mnLoopWords / mnRelocWords), query position of the first shared word
    // (orders lKFsSharingWords), Vocabulary::score(query, keyframe)
    void query(const DBoW2::BowVector& q, std::vector<int>& commonWords, std::vector<int>& firstCommon, std::vector<double>& scores) const;

private:
    BowDatabase(const BowDatabase&);
    BowDatabase& operator=(const BowDatabase&);
    orbgpu_bowdb* mHandle;
};

}  // namespace ORB_SLAM2
#endif
