// orb_host_cli.cc -- small driver of the C++ drop-in classes, used by tests/test_gpu_host_cpp.py.
//
//   orb_host_cli extract <raw.gray> <w> <h> <nfeatures> <scale> <nlevels> <iniTh> <minTh> <out.bin> [mode]
//       mode: call (operator()), queue (FrameQueue push/pop, 3 frames in flight), batch (ExtractBatch x4),
//             pyramid (operator() with SetPyramidDownload; appends the bordered levels)
//       out.bin: int32 N, N x 28 B keypoints, N x 32 B descriptors [, per level: int32 w+38, h+38, bytes]
//   orb_host_cli match <A.bin> <nA> <B.bin> <nB> <out.bin>      -> int32 best idx / dist / second per row,
//       then int32 accepted count of SearchBruteForce and one DescriptorDistance(A0, B0)
//   orb_host_cli maxima  (reads histogram sizes from stdin)     -> prints ind1 ind2 ind3
//   orb_host_cli stereo <left.gray> <right.gray> <w> <h> <nfeatures> <bf> <b> <out.bin>
//       two extractors on two threads as the stereo Frame constructor runs them (src/Frame.cc:78-81), then
//       ORBextractor::ComputeStereoMatches -> int32 N, N x float mvuRight, N x float mvDepth
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <exception>
#include <string>
#include <thread>
#include <vector>

#include "FrameQueue.h"
#include "ORBextractor.h"
#include "ORBmatcher.h"
#include "ORBVocabulary.h"

using namespace ORB_SLAM2;

static bool read_file(const char* path, std::vector<unsigned char>& buf, size_t want)
{
    FILE* f = fopen(path, "rb");
    if (!f) return false;
    buf.resize(want);
    const size_t got = fread(buf.data(), 1, want, f);
    fclose(f);
    return got == want;
}

static void put(FILE* f, const std::vector<cv::KeyPoint>& kps, const cv::Mat& desc)
{
    const int n = (int)kps.size();
    fwrite(&n, 4, 1, f);
    if (n) fwrite(kps.data(), sizeof(cv::KeyPoint), (size_t)n, f);
    for (int i = 0; i < n; ++i) fwrite(desc.ptr(i), 1, 32, f);
}

static int cmd_extract(int argc, char** argv)
{
    if (argc < 11) return 2;
    const int w = atoi(argv[3]), h = atoi(argv[4]);
    std::vector<unsigned char> raw;
    if (!read_file(argv[2], raw, (size_t)w * h)) { fprintf(stderr, "cannot read %s\n", argv[2]); return 2; }
    const std::string mode = argc > 11 ? argv[11] : "call";
    ORBextractor ex(atoi(argv[5]), (float)atof(argv[6]), atoi(argv[7]), atoi(argv[8]), atoi(argv[9]));
    cv::Mat image(h, w, CV_8UC1, raw.data(), (size_t)w);
    FILE* out = fopen(argv[10], "wb");
    if (!out) return 2;
    std::vector<cv::KeyPoint> kps;
    cv::Mat desc;
    if (mode == "call" || mode == "pyramid") {
        if (mode == "pyramid") ex.SetPyramidDownload(true);
        ex(image, cv::Mat(), kps, desc);
        put(out, kps, desc);
        if (mode == "pyramid")
            for (int l = 0; l < ex.GetLevels(); ++l) {
                const cv::Mat& m = ex.mvImagePyramid[l];
                const int bw = m.cols + 38, bh = m.rows + 38;
                fwrite(&bw, 4, 1, out); fwrite(&bh, 4, 1, out);
                for (int y = -19; y < m.rows + 19; ++y) fwrite(m.data + (ptrdiff_t)y * (ptrdiff_t)m.step - 19, 1, (size_t)bw, out);
            }
    } else if (mode == "queue") {
        FrameQueue q(&ex);
        int pushed = 0;
        while (pushed < 3 && !q.Full() && q.Push(image, 0.1 * pushed)) ++pushed;
        double ts;
        int popped = 0;
        while (q.Pop(ts, kps, desc)) { put(out, kps, desc); ++popped; }
        if (popped != pushed || pushed < 2) { fprintf(stderr, "queue: pushed %d popped %d\n", pushed, popped); return 1; }
    } else if (mode == "batch") {
        std::vector<cv::Mat> frames(4, image);
        std::vector<std::vector<cv::KeyPoint> > vk;
        std::vector<cv::Mat> vd;
        if (!ex.ExtractBatch(frames, vk, vd)) { fprintf(stderr, "batch: %s\n", ex.LastError()); return 1; }
        for (size_t i = 0; i < vk.size(); ++i) put(out, vk[i], vd[i]);
    } else return 2;
    fclose(out);
    // getters the Frame constructors copy (src/Frame.cc:182-188)
    printf("levels %d scale %.9g", ex.GetLevels(), ex.GetScaleFactor());
    const std::vector<float> s = ex.GetScaleFactors(), is = ex.GetInverseScaleFactors(), g = ex.GetScaleSigmaSquares(),
                             ig = ex.GetInverseScaleSigmaSquares();
    for (size_t i = 0; i < s.size(); ++i) printf(" %.9g %.9g %.9g %.9g", s[i], is[i], g[i], ig[i]);
    printf("\n");
    return 0;
}

static int cmd_match(int argc, char** argv)
{
    if (argc < 7) return 2;
    const int nA = atoi(argv[3]), nB = atoi(argv[5]);
    std::vector<unsigned char> a, b;
    if (!read_file(argv[2], a, (size_t)nA * 32) || !read_file(argv[4], b, (size_t)nB * 32)) return 2;
    cv::Mat A(nA, 32, CV_8U, a.data(), 32), B(nB, 32, CV_8U, b.data(), 32);
    ORBmatcher m(0.6f, true);
    std::vector<int> idx, d1, d2, m12;
    if (!m.BestTwo(A, B, idx, d1, d2)) return 1;
    const int acc = m.SearchBruteForce(A, B, m12);
    const int d00 = ORBmatcher::DescriptorDistance(A.row(0), B.row(0));
    FILE* out = fopen(argv[6], "wb");
    if (!out) return 2;
    fwrite(idx.data(), 4, idx.size(), out); fwrite(d1.data(), 4, d1.size(), out); fwrite(d2.data(), 4, d2.size(), out);
    fwrite(m12.data(), 4, m12.size(), out);
    fwrite(&acc, 4, 1, out); fwrite(&d00, 4, 1, out);
    fclose(out);
    return 0;
}

static int cmd_stereo(int argc, char** argv)
{
    if (argc < 10) return 2;
    const int w = atoi(argv[4]), h = atoi(argv[5]), nf = atoi(argv[6]);
    std::vector<unsigned char> rawL, rawR;
    if (!read_file(argv[2], rawL, (size_t)w * h) || !read_file(argv[3], rawR, (size_t)w * h)) return 2;
    ORBextractor exL(nf, 1.2f, 8, 20, 7), exR(nf, 1.2f, 8, 20, 7);
    cv::Mat imL(h, w, CV_8UC1, rawL.data(), (size_t)w), imR(h, w, CV_8UC1, rawR.data(), (size_t)w);
    std::vector<cv::KeyPoint> kL, kR;
    cv::Mat dL, dR;
    std::thread tl([&] { exL(imL, cv::Mat(), kL, dL); });
    std::thread tr([&] { exR(imR, cv::Mat(), kR, dR); });
    tl.join(); tr.join();
    std::vector<float> uR, depth;
    if (!ORBextractor::ComputeStereoMatches(exL, exR, (float)atof(argv[7]), (float)atof(argv[8]), uR, depth)) {
        fprintf(stderr, "ComputeStereoMatches: %s\n", exL.LastError());
        return 3;
    }
    FILE* out = fopen(argv[9], "wb");
    if (!out) return 2;
    const int n = (int)uR.size();
    fwrite(&n, 4, 1, out);
    fwrite(uR.data(), 4, (size_t)n, out); fwrite(depth.data(), 4, (size_t)n, out);
    fclose(out);
    // an unchanged stereo Frame.cc reads mvImagePyramid of both extractors (src/Frame.cc:473, :563-580): with two live
    // extractors the download is on by default, without any call to SetPyramidDownload
    printf("%d %d %d %d %d %d\n", (int)kL.size(), (int)kR.size(), exL.mvImagePyramid[0].cols, exL.mvImagePyramid[0].rows,
           exR.mvImagePyramid[7].cols, exR.mvImagePyramid[7].rows);
    return 0;
}

static int cmd_maxima()
{
    std::vector<std::vector<int> > histo;
    int s;
    while (scanf("%d", &s) == 1) histo.push_back(std::vector<int>((size_t)s, 0));
    ORBmatcher m;
    int i1 = -1, i2 = -1, i3 = -1;
    m.ComputeThreeMaxima(histo.data(), (int)histo.size(), i1, i2, i3);
    printf("%d %d %d\n", i1, i2, i3);
    return 0;
}

// bow <voc.txt> <desc.bin> <n> <levelsup> <out.bin>: ORBVocabulary::loadFromTextFile + transform on the descriptor rows, then the
// BowVector scored against itself and against a BowDatabase holding it.  out.bin: int32 nw, nn, ni | words u32[nw] | values f64[nw]
// | nodes u32[nn] | per node: int32 count + u32 indices | f64 self score | int32 db common words | f64 db score
static int cmd_bow(int argc, char** argv)
{
    if (argc < 7) return 2;
    const int n = atoi(argv[4]), levelsup = atoi(argv[5]);
    std::vector<unsigned char> raw;
    if (!read_file(argv[3], raw, (size_t)n * 32)) { fprintf(stderr, "cannot read %s\n", argv[3]); return 2; }
    ORBVocabulary voc;
    if (!voc.loadFromTextFile(argv[2])) { fprintf(stderr, "cannot load %s\n", argv[2]); return 4; }
    std::vector<cv::Mat> rows;
    cv::Mat D(n, 32, CV_8U);
    if (n) memcpy(D.data, raw.data(), (size_t)n * 32);
    for (int i = 0; i < n; ++i) rows.push_back(D.row(i));                       // Converter::toDescriptorVector
    DBoW2::BowVector bv, bv2; DBoW2::FeatureVector fv, fv2;
    voc.transform(rows, bv, fv, levelsup);
    voc.transform(D, bv2, fv2, levelsup);
    if (bv != bv2 || fv != fv2) { fprintf(stderr, "the two transform overloads disagree\n"); return 5; }
    BowDatabase db;
    const int slot = db.add(bv);
    std::vector<int> common, first; std::vector<double> scores;
    db.query(bv, common, first, scores);
    FILE* out = fopen(argv[6], "wb");
    if (!out) return 2;
    int total = 0;
    for (DBoW2::FeatureVector::const_iterator it = fv.begin(); it != fv.end(); ++it) total += (int)it->second.size();
    const int hdr[3] = {(int)bv.size(), (int)fv.size(), total};
    fwrite(hdr, 4, 3, out);
    for (DBoW2::BowVector::const_iterator it = bv.begin(); it != bv.end(); ++it) fwrite(&it->first, 4, 1, out);
    for (DBoW2::BowVector::const_iterator it = bv.begin(); it != bv.end(); ++it) fwrite(&it->second, 8, 1, out);
    for (DBoW2::FeatureVector::const_iterator it = fv.begin(); it != fv.end(); ++it) fwrite(&it->first, 4, 1, out);
    for (DBoW2::FeatureVector::const_iterator it = fv.begin(); it != fv.end(); ++it) {
        const int c = (int)it->second.size();
        fwrite(&c, 4, 1, out);
        fwrite(it->second.data(), 4, (size_t)c, out);
    }
    const double self = voc.score(bv, bv);
    fwrite(&self, 8, 1, out);
    fwrite(&common[(size_t)slot], 4, 1, out);
    fwrite(&scores[(size_t)slot], 8, 1, out);
    fclose(out);
    printf("%u words in the vocabulary, %d in the frame\n", voc.size(), hdr[0]);
    return 0;
}

int main(int argc, char** argv)
{
    try {
        if (argc >= 2 && !strcmp(argv[1], "extract")) return cmd_extract(argc, argv);
        if (argc >= 2 && !strcmp(argv[1], "match")) return cmd_match(argc, argv);
        if (argc >= 2 && !strcmp(argv[1], "maxima")) return cmd_maxima();
        if (argc >= 2 && !strcmp(argv[1], "stereo")) return cmd_stereo(argc, argv);
        if (argc >= 2 && !strcmp(argv[1], "bow")) return cmd_bow(argc, argv);
    } catch (const std::exception& e) {
        fprintf(stderr, "orb_host_cli: %s\n", e.what());     // e.g. no CUDA device: there is no CPU path
        return 3;
    }
    fprintf(stderr, "usage: orb_host_cli extract|match|maxima|stereo|bow ...\n");
    return 2;
}
