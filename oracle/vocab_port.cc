// oracle/vocab_port.cc -- TEST INFRASTRUCTURE (never linked into the product; only tests/, smoke() and bench.py's CPU
// baseline use it).  CPU restatement of the reference's bag-of-words path on flat arrays:
//   ORBVocabulary::transform(features, BowVector, FeatureVector, levelsup)   Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1127-1194
//   the walk of one descriptor                                              :1218-1261 (strict '<': first child on ties)
//   FORB::distance                                                          Thirdparty/DBoW2/DBoW2/FORB.cpp:81-101
//   BowVector::addWeight / addIfNotExist / normalize                        Thirdparty/DBoW2/DBoW2/BowVector.cpp:34-86
//   FeatureVector::addFeature                                               Thirdparty/DBoW2/DBoW2/FeatureVector.cpp:29-43
//   L1Scoring::score                                                        Thirdparty/DBoW2/DBoW2/ScoringObject.cpp:25-72
//   KeyFrameDatabase::DetectLoopCandidates / DetectRelocalizationCandidates src/KeyFrameDatabase.cc:76-197, :199-309
// Pinned by tests/test_vocab.py against the reference's own code (oracle/_ref/libframeref.so: the same vocabulary read by
// TemplatedVocabulary::loadFromTextFile, the same descriptors through Frame::ComputeBoW and Vocabulary::score).
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <map>
#include <vector>

#include "frame_port.h"

struct frameport_vocab {
    int k, L, scoring, weighting;
    std::vector<std::vector<int> > children;      // in order of appearance (m_nodes[pid].children.push_back, :1389-1390)
    std::vector<uint8_t> desc;
    std::vector<double> weight;
    std::vector<uint32_t> word_id;
    int n_words;
};

static int distance256(const uint8_t* a, const uint8_t* b)
{
    int d = 0;
    for (int i = 0; i < 32; i += 8) {
        uint64_t x, y;
        std::memcpy(&x, a + i, 8); std::memcpy(&y, b + i, 8);
        d += __builtin_popcountll(x ^ y);
    }
    return d;
}

extern "C" {

frameport_vocab* frameport_vocab_create(int k, int L, int scoring, int weighting, int n_nodes, const int32_t* parent,
                                        const uint8_t* is_leaf, const uint8_t* desc, const double* weight)
{
    frameport_vocab* v = new frameport_vocab();
    v->k = k; v->L = L; v->scoring = scoring; v->weighting = weighting;
    v->children.resize((size_t)n_nodes);
    v->desc.assign(desc, desc + (size_t)n_nodes * 32);
    v->weight.assign(weight, weight + n_nodes);
    v->word_id.assign((size_t)n_nodes, 0);
    v->n_words = 0;
    for (int i = 1; i < n_nodes; ++i) {
        v->children[(size_t)parent[i]].push_back(i);
        if (is_leaf[i]) v->word_id[(size_t)i] = (uint32_t)v->n_words++;          // :1407-1414
    }
    return v;
}

void frameport_vocab_free(frameport_vocab* v) { delete v; }

void frameport_vocab_transform(const frameport_vocab* v, const uint8_t* desc, int n, int levelsup, uint32_t* words,
                               double* values, uint32_t* node_ids, int32_t* start, uint32_t* idx, int* counts)
{
    counts[0] = counts[1] = counts[2] = 0;
    start[0] = 0;
    if (v->n_words == 0) return;                                                  // empty() (:1135-1138)
    std::map<uint32_t, double> bow;
    std::map<uint32_t, std::vector<uint32_t> > fv;
    const int nid_level = v->L - levelsup;
    const bool accumulate = v->weighting == 0 || v->weighting == 1;               // TF_IDF, TF
    for (int f = 0; f < n; ++f) {
        const uint8_t* q = desc + (size_t)f * 32;
        int final_id = 0, level = 0;
        bool have_nid = nid_level <= 0;
        uint32_t nid = 0;
        while (!v->children[(size_t)final_id].empty()) {                          // do { } while(!isLeaf) from the root
            ++level;
            const std::vector<int>& ch = v->children[(size_t)final_id];
            int best = ch[0], best_d = distance256(q, &v->desc[(size_t)ch[0] * 32]);
            for (size_t c = 1; c < ch.size(); ++c) {
                const int d = distance256(q, &v->desc[(size_t)ch[c] * 32]);
                if (d < best_d) { best_d = d; best = ch[c]; }
            }
            final_id = best;
            if (level == nid_level) { nid = (uint32_t)final_id; have_nid = true; }
        }
        if (!have_nid) nid = (uint32_t)final_id;      // the reference leaves nid uninitialised here (see include/orbgpu.h)
        const double w = v->weight[(size_t)final_id];
        if (w > 0) {
            const uint32_t id = v->word_id[(size_t)final_id];
            std::map<uint32_t, double>::iterator it = bow.find(id);
            if (it == bow.end()) bow[id] = w;
            else if (accumulate) it->second += w;
            fv[nid].push_back((uint32_t)f);
        }
    }
    const int norm_mode = v->scoring == 5 ? 0 : (v->scoring == 1 ? 2 : 1);        // ScoringObject.h:74-89
    if (norm_mode == 0) {
        if (accumulate && !bow.empty()) {
            const double nd = (double)bow.size();
            for (std::map<uint32_t, double>::iterator it = bow.begin(); it != bow.end(); ++it) it->second /= nd;
        }
    } else {
        double norm = 0.0;
        if (norm_mode == 1) {
            for (std::map<uint32_t, double>::iterator it = bow.begin(); it != bow.end(); ++it) norm += std::fabs(it->second);
        } else {
            for (std::map<uint32_t, double>::iterator it = bow.begin(); it != bow.end(); ++it) norm += it->second * it->second;
            norm = std::sqrt(norm);
        }
        if (norm > 0.0)
            for (std::map<uint32_t, double>::iterator it = bow.begin(); it != bow.end(); ++it) it->second /= norm;
    }
    int nw = 0;
    for (std::map<uint32_t, double>::iterator it = bow.begin(); it != bow.end(); ++it, ++nw) { words[nw] = it->first; values[nw] = it->second; }
    int nn = 0, ni = 0;
    for (std::map<uint32_t, std::vector<uint32_t> >::iterator it = fv.begin(); it != fv.end(); ++it, ++nn) {
        node_ids[nn] = it->first;
        for (size_t j = 0; j < it->second.size(); ++j) idx[ni++] = it->second[j];
        start[nn + 1] = ni;
    }
    counts[0] = nw; counts[1] = nn; counts[2] = ni;
}

void frameport_vocab_score(const uint32_t* qw, const double* qv, int nq, const uint32_t* kw, const double* kv,
                           const int32_t* kstart, int nk, double* out)
{
    for (int k = 0; k < nk; ++k) {
        int i = 0, j = kstart[k];
        const int je = kstart[k + 1];
        double score = 0;
        while (i < nq && j < je) {
            if (qw[i] == kw[j]) {
                const double vi = qv[i], wi = kv[j];
                score += std::fabs(vi - wi) - std::fabs(vi) - std::fabs(wi);
                ++i; ++j;
            } else if (qw[i] < kw[j]) ++i;            // lower_bound jumps land where single steps do
            else ++j;
        }
        out[k] = -score / 2.0;
    }
}

}  // extern "C"

// KeyFrameDatabase::DetectLoopCandidates (src/KeyFrameDatabase.cc:76-197) / DetectRelocalizationCandidates (:199-309)
#include <list>
#include <set>
#include <utility>

extern "C" int frameport_db_detect(const frameport_db_args* a, int32_t* marked, int32_t* words, float* score, int32_t* candidates)
{
    const int n = a->n_kf;
    // the inverted file as add() leaves it (:40-46), erase() having removed its keyframes (:48-67).  It persists between
    // queries in the reference, so repeated calls on the same arrays (tests/rows_bench.py timing the query) reuse it.
    static std::vector<std::vector<int> > inverted;
    static const void* built_for = NULL;
    static int built_n = -1, built_total = -1;
    static uint64_t built_sample = 0;
    uint64_t sample = 1469598103934665603ull;                      // a strided sample of the words guards a reused address
    for (int i = 0; i < a->kstart[n]; i += (i < 256 ? 1 : 257)) sample = (sample ^ a->kw[i]) * 1099511628211ull;
    static std::vector<uint8_t> built_erased;
    const bool same = built_for == a->kw && built_n == n && built_total == a->kstart[n] && built_sample == sample && built_erased.size() == (size_t)n &&
                      std::memcmp(built_erased.data(), a->erased, (size_t)n) == 0;
    if (!same) {
        uint32_t max_word = 0;
        for (int i = 0; i < a->kstart[n]; ++i) max_word = std::max(max_word, a->kw[i]);
        inverted.assign((size_t)max_word + 1, std::vector<int>());
        for (int k = 0; k < n; ++k)
            if (!a->erased[k])
                for (int i = a->kstart[k]; i < a->kstart[k + 1]; ++i) inverted[a->kw[i]].push_back(k);
        built_for = a->kw; built_sample = sample; built_n = n; built_total = a->kstart[n]; built_erased.assign(a->erased, a->erased + n);
    }
    static const std::vector<int> no_keyframes;
    for (int k = 0; k < n; ++k) { marked[k] = 0; words[k] = 0; score[k] = a->prev_score[k]; }
    std::list<int> sharing;
    for (int i = 0; i < a->nq; ++i) {                                              // :86-105 / :207-222
        const std::vector<int>& l = a->qw[i] < inverted.size() ? inverted[a->qw[i]] : no_keyframes;
        for (size_t j = 0; j < l.size(); ++j) {
            const int k = l[j];
            if (!marked[k]) {
                words[k] = 0;
                if (a->reloc || !a->connected[k]) { marked[k] = 1; sharing.push_back(k); }
            }
            words[k]++;
        }
    }
    if (sharing.empty()) return 0;
    int max_common = 0;
    for (std::list<int>::iterator it = sharing.begin(); it != sharing.end(); ++it) max_common = std::max(max_common, words[*it]);
    const int min_common = max_common * 0.8f;                                       // :120 / :235
    std::list<std::pair<float, int> > scored;
    for (std::list<int>::iterator it = sharing.begin(); it != sharing.end(); ++it) {
        const int k = *it;
        if (words[k] > min_common) {
            double s;
            const int32_t span[2] = {0, a->kstart[k + 1] - a->kstart[k]};
            frameport_vocab_score(a->qw, a->qv, a->nq, a->kw + a->kstart[k], a->kv + a->kstart[k], span, 1, &s);
            const float si = (float)s;
            score[k] = si;
            if (a->reloc || si >= a->min_score) scored.push_back(std::make_pair(si, k));
        }
    }
    if (scored.empty()) return 0;
    std::list<std::pair<float, int> > acc;
    float best_acc = a->reloc ? 0.f : a->min_score;
    for (std::list<std::pair<float, int> >::iterator it = scored.begin(); it != scored.end(); ++it) {
        const int k = it->second;
        float best = it->first, sum = it->first;
        int best_k = k;
        const int nb = std::min(a->neigh_start[k + 1] - a->neigh_start[k], 10);    // GetBestCovisibilityKeyFrames(10)
        for (int j = 0; j < nb; ++j) {
            const int k2 = a->neigh[a->neigh_start[k] + j];
            if (!marked[k2]) continue;
            if (!a->reloc && !(words[k2] > min_common)) continue;                   // :159 (the relocaliser has no such test, :273-274)
            sum += score[k2];
            if (score[k2] > best) { best_k = k2; best = score[k2]; }
        }
        acc.push_back(std::make_pair(sum, best_k));
        if (sum > best_acc) best_acc = sum;
    }
    const float keep = 0.75f * best_acc;
    std::set<int> added;
    int nc = 0;
    for (std::list<std::pair<float, int> >::iterator it = acc.begin(); it != acc.end(); ++it)
        if (it->first > keep && !added.count(it->second)) { candidates[nc++] = it->second; added.insert(it->second); }
    return nc;
}
