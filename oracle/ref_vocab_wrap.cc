// oracle/ref_vocab_wrap.cc -- TEST INFRASTRUCTURE (never linked into the product).
// Drives the reference's own bag-of-words code, compiled unchanged by oracle/Makefile into oracle/_ref/libframeref.so:
//   Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h  loadFromTextFile :1338-1425, transform(features, BowVector,
//                                                 FeatureVector, levelsup) :1127-1194, transform(feature, ...) :1218-1261
//   Thirdparty/DBoW2/DBoW2/FORB.cpp               distance :81-101, fromString :121-137
//   Thirdparty/DBoW2/DBoW2/BowVector.cpp          addWeight / addIfNotExist / normalize :34-86
//   Thirdparty/DBoW2/DBoW2/FeatureVector.cpp      addFeature :29-43
//   Thirdparty/DBoW2/DBoW2/ScoringObject.cpp      L1Scoring::score :25-72 (and the others through Vocabulary::score)
//   src/Frame.cc:395-402                          Frame::ComputeBoW (descriptor Mat -> rows -> transform(..., 4))
// The vocabulary is read by the reference's own loader from a text file the test writes (the ORBvoc.txt format).
#include <opencv2/core/core.hpp>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#define private public
#define protected public
#include "Frame.h"
#include "ORBVocabulary.h"
#include "Converter.h"
#undef private
#undef protected

using namespace ORB_SLAM2;

extern "C" {

void* frameref_vocab_load(const char* path)
{
    ORBVocabulary* v = new ORBVocabulary();
    if (!v->loadFromTextFile(path)) { delete v; return NULL; }
    return v;
}

void frameref_vocab_free(void* h) { delete static_cast<ORBVocabulary*>(h); }

// k, L, scoring, weighting, number of nodes (root included), number of words
void frameref_vocab_info(void* h, int* out6)
{
    const ORBVocabulary* v = static_cast<ORBVocabulary*>(h);
    out6[0] = v->m_k; out6[1] = v->m_L; out6[2] = (int)v->m_scoring; out6[3] = (int)v->m_weighting;
    out6[4] = (int)v->m_nodes.size(); out6[5] = (int)v->m_words.size();
}

// Frame::ComputeBoW on a frame that holds only the descriptors.  Outputs sized by the caller: words/values [n],
// node_ids [n], start [n + 1], idx [n].  counts = {BowVector size, FeatureVector size, total feature indices}.
void frameref_compute_bow(void* h, const uint8_t* desc, int n, uint32_t* words, double* values, uint32_t* node_ids,
                          int32_t* start, uint32_t* idx, int* counts)
{
    Frame F;
    F.mpORBvocabulary = static_cast<ORBVocabulary*>(h);
    F.N = n;
    F.mDescriptors = cv::Mat(n, 32, CV_8U);
    if (n > 0) std::memcpy(F.mDescriptors.data, desc, (size_t)n * 32);
    F.ComputeBoW();
    int nw = 0;
    for (DBoW2::BowVector::const_iterator it = F.mBowVec.begin(); it != F.mBowVec.end(); ++it, ++nw) {
        words[nw] = it->first; values[nw] = it->second;
    }
    int nn = 0, ni = 0;
    start[0] = 0;
    for (DBoW2::FeatureVector::const_iterator it = F.mFeatVec.begin(); it != F.mFeatVec.end(); ++it, ++nn) {
        node_ids[nn] = it->first;
        for (size_t j = 0; j < it->second.size(); ++j) idx[ni++] = it->second[j];
        start[nn + 1] = ni;
    }
    counts[0] = nw; counts[1] = nn; counts[2] = ni;
}

// the general transform with a caller-chosen levelsup (LoopClosing and the relocaliser use 4 through ComputeBoW)
void frameref_vocab_transform(void* h, const uint8_t* desc, int n, int levelsup, uint32_t* words, double* values,
                              uint32_t* node_ids, int32_t* start, uint32_t* idx, int* counts)
{
    const ORBVocabulary* v = static_cast<ORBVocabulary*>(h);
    cv::Mat D(n, 32, CV_8U);
    if (n > 0) std::memcpy(D.data, desc, (size_t)n * 32);
    std::vector<cv::Mat> rows = Converter::toDescriptorVector(D);
    DBoW2::BowVector bv; DBoW2::FeatureVector fv;
    v->transform(rows, bv, fv, levelsup);
    int nw = 0;
    for (DBoW2::BowVector::const_iterator it = bv.begin(); it != bv.end(); ++it, ++nw) { words[nw] = it->first; values[nw] = it->second; }
    int nn = 0, ni = 0;
    start[0] = 0;
    for (DBoW2::FeatureVector::const_iterator it = fv.begin(); it != fv.end(); ++it, ++nn) {
        node_ids[nn] = it->first;
        for (size_t j = 0; j < it->second.size(); ++j) idx[ni++] = it->second[j];
        start[nn + 1] = ni;
    }
    counts[0] = nw; counts[1] = nn; counts[2] = ni;
}

// Vocabulary::score(query, keyframe k) for a batch of keyframes given as one CSR of (word, value) pairs: the call
// KeyFrameDatabase::DetectLoopCandidates / DetectRelocalizationCandidates make per candidate (src/KeyFrameDatabase.cc:133,
// :249) and LoopClosing::DetectLoop per covisible keyframe (src/LoopClosing.cc:135)
void frameref_vocab_score(void* h, const uint32_t* qw, const double* qv, int nq, const uint32_t* kw, const double* kv,
                          const int32_t* kstart, int nk, double* out)
{
    const ORBVocabulary* v = static_cast<ORBVocabulary*>(h);
    DBoW2::BowVector q;
    for (int i = 0; i < nq; ++i) q.insert(q.end(), DBoW2::BowVector::value_type(qw[i], qv[i]));
    for (int k = 0; k < nk; ++k) {
        DBoW2::BowVector w;
        for (int i = kstart[k]; i < kstart[k + 1]; ++i) w.insert(w.end(), DBoW2::BowVector::value_type(kw[i], kv[i]));
        out[k] = v->score(q, w);
    }
}

}  // extern "C"
