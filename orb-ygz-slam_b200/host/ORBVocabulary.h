// ORBVocabulary.h — drop-in for the part of ygz::ORBVocabulary (include/ORBVocabulary.h: DBoW2::TemplatedVocabulary<
// cv::Mat, FORB>) that the front end uses: loadFromTextFile / loadFromBinaryFile (src/System.cc:183-190) and
// transform(features, BowVector&, FeatureVector&, levelsup) (src/Frame.cc:495-500, src/KeyFrame.cc:295-302).  The tree
// lives on the device; the work is done by libygzorb_cuda.so (ygzorb_vocabulary_*, ygzorb_bow_transform).
// BowVector / FeatureVector are the reference's container types (Thirdparty/DBoW2/DBoW2/BowVector.h:20-48,
// FeatureVector.h:23-24): inside the reference tree include its own headers first and define YGZ_HAVE_DBOW2_TYPES.
#ifndef YGZ_HOST_ORBVOCABULARY_H_
#define YGZ_HOST_ORBVOCABULARY_H_

#include <cstring>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

#include "cvcompat.h"
#include "../../include/ygzorb.h"

#ifndef YGZ_HAVE_DBOW2_TYPES
namespace DBoW2 {
typedef unsigned int WordId;
typedef double WordValue;
typedef unsigned int NodeId;
class BowVector : public std::map<WordId, WordValue> {};
class FeatureVector : public std::map<NodeId, std::vector<unsigned int>> {};
}  // namespace DBoW2
#endif

namespace ygz {

class ORBVocabulary {
public:
    explicit ORBVocabulary(int device = 0) : mDevice(device), mVoc(nullptr), mMatcher(nullptr) {}
    ~ORBVocabulary() { ygzorb_vocabulary_destroy(mVoc); ygzorb_matcher_destroy(mMatcher); }
    ORBVocabulary(const ORBVocabulary &) = delete;
    ORBVocabulary &operator=(const ORBVocabulary &) = delete;

    bool loadFromTextFile(const std::string &filename) { return load(filename, true); }      // TemplatedVocabulary.h:243
    bool loadFromBinaryFile(const std::string &filename) { return load(filename, false); }   // TemplatedVocabulary.h:255
    bool empty() const { return !mVoc || info(3) == 0; }                                     // m_words.empty()
    unsigned int size() const { return mVoc ? (unsigned int)info(3) : 0; }                   // number of words
    const ygzorb_vocabulary *handle() const { return mVoc; }

    // TemplatedVocabulary.h:1149-1218
    void transform(const std::vector<cv::Mat> &features, DBoW2::BowVector &v, DBoW2::FeatureVector &fv, int levelsup) const {
        v.clear();
        fv.clear();
        if (empty() || features.empty()) return;
        const int n = (int)features.size();
        if (n > 8192) throw std::runtime_error("ORBVocabulary::transform: more than 8192 features");
        std::vector<uint8_t> desc((size_t)n * 32);
        for (int i = 0; i < n; i++) std::memcpy(&desc[(size_t)i * 32], features[i].ptr(0), 32);
        const int32_t nFeat = n;
        int32_t nb = 0, nn = 0;
        std::vector<int32_t> bw(n), fs(n + 1);
        std::vector<double> bv(n);
        std::vector<uint32_t> fnode(n), ffeat(n);
        if (ygzorb_bow_transform(mMatcher, mVoc, desc.data(), &nFeat, 1, n, levelsup, 0, nullptr, nullptr, nullptr, &nb, bw.data(),
                                 bv.data(), &nn, fnode.data(), fs.data(), ffeat.data()) != 0)
            throw std::runtime_error(std::string("ygzorb_bow_transform: ") + ygzorb_last_error());
        for (int j = 0; j < nb; j++) v.insert(v.end(), std::make_pair((DBoW2::WordId)bw[j], bv[j]));     // ascending: O(1) each
        for (int j = 0; j < nn; j++) {
            auto it = fv.insert(fv.end(), std::make_pair((DBoW2::NodeId)fnode[j], std::vector<unsigned int>()));
            it->second.assign(ffeat.begin() + fs[j], ffeat.begin() + fs[j + 1]);
        }
    }

private:
    bool load(const std::string &filename, bool text) {
        ygzorb_vocabulary *voc = nullptr;
        const int rc = text ? ygzorb_vocabulary_load_text(filename.c_str(), mDevice, &voc)
                            : ygzorb_vocabulary_load_binary(filename.c_str(), mDevice, &voc);
        if (rc != 0) return false;
        ygzorb_vocabulary_destroy(mVoc);
        mVoc = voc;
        if (!mMatcher && ygzorb_matcher_create(mDevice, &mMatcher) != 0) return false;
        return true;
    }
    int info(int i) const { int32_t a[6]; ygzorb_vocabulary_info(mVoc, a); return a[i]; }
    int mDevice;
    ygzorb_vocabulary *mVoc;
    ygzorb_matcher *mMatcher;
};

}  // namespace ygz
#endif
