// bow_oracle.cpp — TEST INFRASTRUCTURE ONLY.  CPU restatement of the reference's bag-of-words assignment
// (SURVEY.md §8(f) row f1): DBoW2::TemplatedVocabulary<cv::Mat, FORB>::transform as Frame::ComputeBoW calls it
// (src/Frame.cc:495-500, src/KeyFrame.cc:295-302), and the vocabulary file formats it is loaded from.
// Linked into oracle/libygz_oracle.so; used by tests/ and bench CPU legs only — never by the product.
//
// PARITY PINNED: the reference's own Thirdparty/DBoW2 compiles here unmodified (oracle/_ref/ygz_ref_bow, see
// oracle/Makefile); tests/test_bow_oracle.py checks this restatement against it (tree arrays, per-feature
// word / node / weight, BowVector doubles bit for bit, FeatureVector) and against tests/golden/bow_*.npz.
// Citations are file:line in /root/reference/Thirdparty/DBoW2/DBoW2.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#define YGZO_API extern "C" __attribute__((visibility("default")))

namespace {

struct Voc {
    int k = 0, L = 0, scoring = 0, weighting = 0;
    std::vector<int32_t> parent, wordId;      // wordId: -1 for inner nodes
    std::vector<double> weight;
    std::vector<uint8_t> desc;                // 32 B per node (root: zeros)
    std::vector<std::vector<int32_t>> children;
    // flattened (CSR) children, built on demand
    std::vector<int32_t> childStart, childList;
    int nWords = 0;
    void flatten() {
        childStart.assign(parent.size() + 1, 0);
        childList.clear();
        for (size_t i = 0; i < parent.size(); i++) {
            childStart[i] = (int32_t)childList.size();
            childList.insert(childList.end(), children[i].begin(), children[i].end());
        }
        childStart[parent.size()] = (int32_t)childList.size();
    }
};

// FORB::distance — FORB.cpp:82-102 (bit-count hack per 32-bit word == popcount)
inline int forbDistance(const uint8_t *a, const uint8_t *b) {
    int d = 0;
    for (int i = 0; i < 8; i++) { uint32_t x, y; std::memcpy(&x, a + 4 * i, 4); std::memcpy(&y, b + 4 * i, 4); d += __builtin_popcount(x ^ y); }
    return d;
}

// TemplatedVocabulary::transform(feature, word_id, weight, nid, levelsup) — TemplatedVocabulary.h:1242-1282.
// First minimum wins among the children (strict <, :1267); a node is a leaf when it has no children (isLeaf, :118).
void descend(const Voc &v, const uint8_t *f, int levelsup, int32_t *word, int32_t *node, double *w) {
    const int nidLevel = v.L - levelsup;
    int32_t nid = 0;
    if (nidLevel <= 0) nid = 0;
    int32_t finalId = 0;
    int level = 0;
    do {
        ++level;
        const std::vector<int32_t> &nodes = v.children[finalId];
        finalId = nodes[0];
        double best = forbDistance(f, &v.desc[(size_t)finalId * 32]);
        for (size_t c = 1; c < nodes.size(); c++) {
            const double d = forbDistance(f, &v.desc[(size_t)nodes[c] * 32]);
            if (d < best) { best = d; finalId = nodes[c]; }
        }
        if (level == nidLevel) nid = finalId;
    } while (!v.children[finalId].empty());
    *word = v.wordId[finalId]; *w = v.weight[finalId]; *node = nid;
}

}  // namespace

// TemplatedVocabulary::loadFromBinaryFile — TemplatedVocabulary.h:1478-1523, including its read loop quirk: the
// `while (!f.eof())` loop runs once more after the last record, so the last node is appended a second time (same
// parent, descriptor, weight, leaf flag) as node nb_nodes; being an exact duplicate listed after the original it can
// never win the strict `<` of the descent, but it exists (and owns the last word id when it is a leaf).
YGZO_API void *ygzo_voc_load_binary(const char *path) {
    FILE *f = fopen(path, "rb");
    if (!f) return nullptr;
    uint32_t nbNodes = 0, sizeNode = 0;
    int32_t k = 0, L = 0, scoring = 0, weighting = 0;
    if (fread(&nbNodes, 4, 1, f) != 1 || fread(&sizeNode, 4, 1, f) != 1 || fread(&k, 4, 1, f) != 1 || fread(&L, 4, 1, f) != 1 ||
        fread(&scoring, 4, 1, f) != 1 || fread(&weighting, 4, 1, f) != 1 || sizeNode < 41) { fclose(f); return nullptr; }
    Voc *v = new Voc();
    v->k = k; v->L = L; v->scoring = scoring; v->weighting = weighting;
    const size_t n = (size_t)nbNodes + 1;
    v->parent.assign(n, 0); v->wordId.assign(n, -1); v->weight.assign(n, 0.0); v->desc.assign(n * 32, 0); v->children.resize(n);
    std::vector<char> buf(sizeNode, 0);
    size_t nid = 1;
    bool eof = false;
    while (!eof && nid < n) {
        const size_t got = fread(buf.data(), 1, sizeNode, f);   // a short / empty read leaves the previous record in buf
        if (got < sizeNode) eof = true;
        int32_t par; std::memcpy(&par, buf.data(), 4);
        v->parent[nid] = par;
        v->children[par].push_back((int32_t)nid);
        std::memcpy(&v->desc[nid * 32], buf.data() + 4, 32);
        float w; std::memcpy(&w, buf.data() + 36, 4);
        v->weight[nid] = w;
        if (buf[40]) v->wordId[nid] = v->nWords++;
        nid++;
    }
    fclose(f);
    // nodes the file did not reach stay childless with word -1 (cannot be reached: nobody lists them as a child)
    v->flatten();
    return v;
}

// A vocabulary from flat arrays (what ygzorb_vocabulary_create takes): children of node i are
// childList[childStart[i] .. childStart[i+1]) in vector order.
YGZO_API void *ygzo_voc_from_arrays(int nNodes, int k, int L, const int32_t *childStart, const int32_t *childList,
                                    const uint8_t *desc, const int32_t *wordId, const double *weight) {
    Voc *v = new Voc();
    v->k = k; v->L = L; v->scoring = 0; v->weighting = 0;
    v->parent.assign(nNodes, 0); v->wordId.assign(wordId, wordId + nNodes); v->weight.assign(weight, weight + nNodes);
    v->desc.assign(desc, desc + (size_t)nNodes * 32); v->children.resize(nNodes);
    for (int i = 0; i < nNodes; i++)
        for (int c = childStart[i]; c < childStart[i + 1]; c++) { v->children[i].push_back(childList[c]); v->parent[childList[c]] = i; }
    for (int i = 0; i < nNodes; i++) if (wordId[i] >= 0) v->nWords = std::max(v->nWords, wordId[i] + 1);
    v->flatten();
    return v;
}

YGZO_API void ygzo_voc_free(void *h) { delete (Voc *)h; }
YGZO_API void ygzo_voc_info(const void *h, int32_t *out5) {
    const Voc *v = (const Voc *)h;
    out5[0] = (int32_t)v->parent.size(); out5[1] = v->k; out5[2] = v->L; out5[3] = v->nWords; out5[4] = (int32_t)v->childList.size();
}
YGZO_API void ygzo_voc_arrays(const void *h, int32_t *parent, int32_t *childStart, int32_t *childList, uint8_t *desc,
                              int32_t *wordId, double *weight) {
    const Voc *v = (const Voc *)h;
    const size_t n = v->parent.size();
    if (parent) std::memcpy(parent, v->parent.data(), n * 4);
    if (childStart) std::memcpy(childStart, v->childStart.data(), (n + 1) * 4);
    if (childList) std::memcpy(childList, v->childList.data(), v->childList.size() * 4);
    if (desc) std::memcpy(desc, v->desc.data(), n * 32);
    if (wordId) std::memcpy(wordId, v->wordId.data(), n * 4);
    if (weight) std::memcpy(weight, v->weight.data(), n * 8);
}

// TemplatedVocabulary::transform(features, BowVector&, FeatureVector&, levelsup) — TemplatedVocabulary.h:1149-1218 for
// the TF_IDF weighting / L1 norm the ORB vocabulary is built with: per feature in order, if its word's weight is > 0,
// BowVector::addWeight (BowVector.cpp:33-46: value += w on an existing word) and FeatureVector::addFeature
// (FeatureVector.cpp:33-46: push_back in feature order); then BowVector::normalize(L1) (BowVector.cpp:70-98:
// norm = sum of |value| in ascending word order; values /= norm when norm > 0).
// Outputs: per feature word / node / weight; the BowVector as (word, value) ascending; the FeatureVector as CSR
// (ascending node ids, features in insertion order).  Returns the BowVector size; *nNodesOut = FeatureVector size.
YGZO_API int ygzo_bow_transform(const void *h, const uint8_t *desc, int n, int levelsup, int32_t *word, int32_t *node,
                                double *weight, int32_t *bowWord, double *bowValue, int32_t *nNodesOut, int32_t *fvNode,
                                int32_t *fvStart, int32_t *fvFeat) {
    const Voc *v = (const Voc *)h;
    std::map<uint32_t, double> bow;
    std::map<uint32_t, std::vector<uint32_t>> fv;
    if (v->parent.empty() || v->children[0].empty()) { if (nNodesOut) *nNodesOut = 0; return 0; }   // empty(): :1161-1164
    for (int i = 0; i < n; i++) {
        int32_t wid, nid; double w;
        descend(*v, desc + (size_t)i * 32, levelsup, &wid, &nid, &w);
        if (word) word[i] = wid;
        if (node) node[i] = nid;
        if (weight) weight[i] = w;
        if (w > 0) {
            auto it = bow.lower_bound((uint32_t)wid);
            if (it != bow.end() && !(bow.key_comp()((uint32_t)wid, it->first))) it->second += w;
            else bow.insert(it, std::make_pair((uint32_t)wid, w));
            fv[(uint32_t)nid].push_back((uint32_t)i);
        }
    }
    double norm = 0.0;
    for (auto &e : bow) norm += std::fabs(e.second);
    if (norm > 0.0) for (auto &e : bow) e.second /= norm;
    int nb = 0;
    for (auto &e : bow) { if (bowWord) bowWord[nb] = (int32_t)e.first; if (bowValue) bowValue[nb] = e.second; nb++; }
    int nn = 0, nf = 0;
    for (auto &e : fv) {
        if (fvNode) fvNode[nn] = (int32_t)e.first;
        if (fvStart) fvStart[nn] = nf;
        for (uint32_t idx : e.second) { if (fvFeat) fvFeat[nf] = (int32_t)idx; nf++; }
        nn++;
    }
    if (fvStart) fvStart[nn] = nf;
    if (nNodesOut) *nNodesOut = nn;
    return nb;
}
