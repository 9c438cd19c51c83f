// bow_ref_cli.cpp — TEST INFRASTRUCTURE ONLY.  Driver around the reference's OWN, unmodified
// Thirdparty/DBoW2 (TemplatedVocabulary.h, FORB.cpp, BowVector.cpp, FeatureVector.cpp, ScoringObject.cpp,
// DUtils/Random.cpp), compiled from /root/reference by oracle/Makefile into oracle/_ref/ygz_ref_bow.
//   ygz_ref_bow create <k> <L> <seed> <train.bin> <voc.bin>
//        train.bin = int32 n_images | per image: int32 n | n x 32 B descriptors (several images, or every idf is log(1/1) = 0).
//        TemplatedVocabulary::create (:~560-700, k-means++ seeded through
//        DUtils::Random::SeedRandOnce(seed)) with TF_IDF / L1_NORM, then saveToBinaryFile (:1528-1550).
//   ygz_ref_bow transform <voc.bin> <levelsup> <feats.bin> <out.bin> [threads] [repeat]
//        feats.bin = int32 n_frames | per frame: int32 n | n x 32 B.  loadFromBinaryFile (:1478-1523) then
//        transform(features, BowVector, FeatureVector, levelsup) (:1149-1218), exactly as Frame::ComputeBoW calls it
//        (src/Frame.cc:495-500).  out.bin per frame: int32 n | n x {int32 word, int32 node, double weight} (the
//        per-feature transform, :1242-1282) | int32 nb | nb x {int32 word, double value} | int32 nn |
//        nn x {int32 node, int32 cnt, cnt x int32 feature}.  Prints the wall time of the transform loop.
//   ygz_ref_bow dump <voc.bin> <flat.bin>
//        the loaded tree as flat arrays: int32 n_nodes, k, L | n_nodes x {int32 parent, int32 n_children, int32 word_id,
//        double weight, 32 B descriptor} | per node its children ids in vector order.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include "DBoW2/FORB.h"
#include "DBoW2/TemplatedVocabulary.h"

typedef DBoW2::TemplatedVocabulary<DBoW2::FORB::TDescriptor, DBoW2::FORB> ORBVocabulary;   // include/ORBVocabulary.h

struct VocAccess : public ORBVocabulary {
    VocAccess(int k, int L) : ORBVocabulary(k, L, DBoW2::TF_IDF, DBoW2::L1_NORM) {}
    const std::vector<Node> &nodes() const { return m_nodes; }
    int k() const { return m_k; }
    int L() const { return m_L; }
    void one(const cv::Mat &f, DBoW2::WordId &wid, DBoW2::WordValue &w, DBoW2::NodeId *nid, int levelsup) const {
        transform(f, wid, w, nid, levelsup);   // the protected per-feature descent (:1242-1282)
    }
};

static std::vector<cv::Mat> readDescs(FILE *f, int n) {
    std::vector<cv::Mat> v(n);
    for (int i = 0; i < n; i++) {
        v[i].create(1, 32, CV_8U);
        if (fread(v[i].data, 1, 32, f) != 32) { fprintf(stderr, "short read\n"); exit(2); }
    }
    return v;
}
static int readInt(FILE *f) { int32_t v; if (fread(&v, 4, 1, f) != 1) { fprintf(stderr, "short read\n"); exit(2); } return v; }

int main(int argc, char **argv) {
    if (argc >= 7 && !strcmp(argv[1], "create")) {
        const int k = atoi(argv[2]), L = atoi(argv[3]), seed = atoi(argv[4]);
        FILE *f = fopen(argv[5], "rb");
        if (!f) { perror("train"); return 2; }
        const int nImages = readInt(f);
        std::vector<std::vector<cv::Mat>> training(nImages);
        for (int i = 0; i < nImages; i++) { const int n = readInt(f); training[i] = readDescs(f, n); }
        fclose(f);
        DUtils::Random::SeedRandOnce(seed);
        VocAccess voc(k, L);
        voc.create(training);
        voc.saveToBinaryFile(argv[6]);
        printf("created vocabulary: k=%d L=%d nodes=%zu words=%u\n", k, L, voc.nodes().size(), voc.size());
        return 0;
    }
    if (argc >= 4 && !strcmp(argv[1], "dump")) {
        VocAccess voc(10, 6);
        if (!voc.loadFromBinaryFile(argv[2])) return 2;
        FILE *f = fopen(argv[3], "wb");
        const int32_t hdr[3] = {(int32_t)voc.nodes().size(), voc.k(), voc.L()};
        fwrite(hdr, 4, 3, f);
        for (const auto &nd : voc.nodes()) {
            const int32_t a[3] = {(int32_t)nd.parent, (int32_t)nd.children.size(), nd.isLeaf() ? (int32_t)nd.word_id : -1};
            const double w = nd.weight;
            uint8_t d[32] = {0};
            if (!nd.descriptor.empty()) memcpy(d, nd.descriptor.data, 32);
            fwrite(a, 4, 3, f); fwrite(&w, 8, 1, f); fwrite(d, 1, 32, f);
        }
        for (const auto &nd : voc.nodes())
            for (auto c : nd.children) { const int32_t ci = (int32_t)c; fwrite(&ci, 4, 1, f); }
        fclose(f);
        return 0;
    }
    if (argc >= 6 && !strcmp(argv[1], "transform")) {
        VocAccess voc(10, 6);
        if (!voc.loadFromBinaryFile(argv[2])) return 2;
        const int levelsup = atoi(argv[3]);
        const int threads = argc > 6 ? atoi(argv[6]) : 1, repeat = argc > 7 ? atoi(argv[7]) : 1;
        FILE *f = fopen(argv[4], "rb");
        if (!f) { perror("feats"); return 2; }
        const int nFrames = readInt(f);
        std::vector<std::vector<cv::Mat>> frames(nFrames);
        for (int i = 0; i < nFrames; i++) { const int n = readInt(f); frames[i] = readDescs(f, n); }
        fclose(f);
        std::vector<DBoW2::BowVector> bow(nFrames);
        std::vector<DBoW2::FeatureVector> fv(nFrames);
        const auto t0 = std::chrono::steady_clock::now();
        for (int r = 0; r < repeat; r++) {
            std::vector<std::thread> th;
            for (int t = 0; t < threads; t++)
                th.emplace_back([&, t]() {
                    for (int i = t; i < nFrames; i += threads) voc.transform(frames[i], bow[i], fv[i], levelsup);   // Frame.cc:498
                });
            for (auto &x : th) x.join();
        }
        const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        long long nFeat = 0;
        for (auto &fr : frames) nFeat += (long long)fr.size();
        printf("transform: %d frames, %lld features, %d threads, %d repeats: %.6f s (%.1f features/s)\n", nFrames, nFeat, threads,
               repeat, sec, nFeat * (double)repeat / sec);
        FILE *o = fopen(argv[5], "wb");
        for (int i = 0; i < nFrames; i++) {
            const int32_t n = (int32_t)frames[i].size();
            fwrite(&n, 4, 1, o);
            for (int j = 0; j < n; j++) {
                DBoW2::WordId wid; DBoW2::WordValue w; DBoW2::NodeId nid;
                voc.one(frames[i][j], wid, w, &nid, levelsup);
                const int32_t a[2] = {(int32_t)wid, (int32_t)nid};
                const double wd = w;
                fwrite(a, 4, 2, o); fwrite(&wd, 8, 1, o);
            }
            const int32_t nb = (int32_t)bow[i].size();
            fwrite(&nb, 4, 1, o);
            for (const auto &e : bow[i]) { const int32_t wi = (int32_t)e.first; const double v = e.second; fwrite(&wi, 4, 1, o); fwrite(&v, 8, 1, o); }
            const int32_t nn = (int32_t)fv[i].size();
            fwrite(&nn, 4, 1, o);
            for (const auto &e : fv[i]) {
                const int32_t a[2] = {(int32_t)e.first, (int32_t)e.second.size()};
                fwrite(a, 4, 2, o);
                for (unsigned idx : e.second) { const int32_t ii = (int32_t)idx; fwrite(&ii, 4, 1, o); }
            }
        }
        fclose(o);
        return 0;
    }
    fprintf(stderr, "usage: ygz_ref_bow create|dump|transform ... (see the header of bow_ref_cli.cpp)\n");
    return 2;
}
