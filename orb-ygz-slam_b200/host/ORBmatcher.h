// ORBmatcher.h — drop-in for the Hamming paths of the reference's include/ORBmatcher.h: same class
// name, constructor, constants and method signatures for DescriptorDistance, SearchByProjection(Frame&,
// vector<MapPoint*>&, th, checkLevel), SearchByProjection(Frame&, const Frame&, th, bMono, checkLevel) and the stereo matcher Frame::ComputeStereoMatches (as a free
// function over the Frame members it uses).  The searches are templates over the reference's Frame /
// MapPoint types so this header compiles both inside the reference tree and against host/FrameLite.h.
// The work is done by libygzorb_cuda.so; there is no CPU implementation here except the 8-word
// popcount of the scalar static DescriptorDistance (a single pair is not worth a kernel launch).
#ifndef YGZ_HOST_ORBMATCHER_H_
#define YGZ_HOST_ORBMATCHER_H_

#include <algorithm>
#include <cstring>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "ORBextractor.h"
#include "cvcompat.h"
#include "../../include/ygzorb.h"

namespace ygz {

class ORBmatcher {
public:
    ORBmatcher(float nnratio = 0.6, bool checkOri = true, int device = 0)          // ORBmatcher.h:43
        : mfNNratio(nnratio), mbCheckOrientation(checkOri), mHandle(nullptr) {
        if (ygzorb_matcher_create(device, &mHandle) != 0)
            throw std::runtime_error(std::string("ygzorb_matcher_create: ") + ygzorb_last_error());
    }
    ~ORBmatcher() { ygzorb_matcher_destroy(mHandle); }
    ORBmatcher(const ORBmatcher &) = delete;
    ORBmatcher &operator=(const ORBmatcher &) = delete;

    static const int TH_LOW = 50, TH_HIGH = 100, HISTO_LENGTH = 30;                // src/ORBmatcher.cc:36-38

    // src/ORBmatcher.cc:1507-1523
    static int DescriptorDistance(const cv::Mat &a, const cv::Mat &b) {
        const uint32_t *pa = a.ptr<uint32_t>(), *pb = b.ptr<uint32_t>();
        int dist = 0;
        for (int i = 0; i < 8; i++) dist += __builtin_popcount(pa[i] ^ pb[i]);
        return dist;
    }

    // src/ORBmatcher.cc:44-126.  Writes F.mvpMapPoints like the reference and returns nmatches.
    template <class FrameT, class MapPointT>
    int SearchByProjection(FrameT &F, const std::vector<MapPointT *> &vpMapPoints, const float th = 3, bool checkLevel = true) {
        const int n = (int)F.mvKeys.size(), m = (int)vpMapPoints.size();
        if (n == 0 || m == 0) return 0;
        std::vector<float> px(m), py(m), pxr(m), vc(m);
        std::vector<int32_t> lvl(m);
        std::vector<uint8_t> inView(m), hasObs(m), mdesc((size_t)m * 32), claimed(n), desc((size_t)n * 32);
        for (int i = 0; i < m; i++) {
            MapPointT *p = vpMapPoints[i];
            inView[i] = p->mbTrackInView && !p->isBad();
            hasObs[i] = p->Observations() > 0;
            px[i] = p->mTrackProjX; py[i] = p->mTrackProjY; pxr[i] = p->mTrackProjXR; vc[i] = p->mTrackViewCos;
            lvl[i] = p->mnTrackScaleLevel;
            if (inView[i]) { const cv::Mat d = p->GetDescriptor(); std::memcpy(&mdesc[(size_t)i * 32], d.ptr(0), 32); }
        }
        bool anyRight = false;
        for (int i = 0; i < n; i++) {
            claimed[i] = F.mvpMapPoints[i] && F.mvpMapPoints[i]->Observations() > 0;
            std::memcpy(&desc[(size_t)i * 32], F.mDescriptors.ptr(i), 32);
            anyRight |= F.mvuRight[i] > 0;
        }
        const int32_t nKp = n, nMp = m;
        ygzorb_frame_features fv{1, n, &nKp, (const ygzorb_keypoint *)F.mvKeys.data(), desc.data(),
                                 anyRight ? F.mvuRight.data() : nullptr, claimed.data(), F.mnMinX, F.mnMinY, F.mnMaxX, F.mnMaxY,
                                 (int)F.mvScaleFactors.size(), F.mvScaleFactors.data()};
        ygzorb_mappoint_proj pv{m, &nMp, px.data(), py.data(), pxr.data(), vc.data(), lvl.data(), inView.data(), hasObs.data(),
                                mdesc.data()};
        std::vector<int32_t> kpToMp(n);
        int32_t nmatches = 0;
        if (ygzorb_search_by_projection(mHandle, &fv, &pv, th, checkLevel ? 1 : 0, mfNNratio, 0, kpToMp.data(), nullptr, &nmatches) != 0)
            throw std::runtime_error(std::string("ygzorb_search_by_projection: ") + ygzorb_last_error());
        for (int i = 0; i < n; i++)
            if (kpToMp[i] >= 0) F.mvpMapPoints[i] = vpMapPoints[kpToMp[i]];      // F.mvpMapPoints[bestIdx] = pMP (:120)
        return nmatches;
    }

    // src/ORBmatcher.cc:1218-1350 (the motion-model tracker).  FrameT needs mTcw with rotationMatrix() / translation()
    // (Sophus SE3f in the reference) whose results index as (r, c) / [r], plus the members the reference reads.
    template <class FrameT>
    int SearchByProjection(FrameT &CurrentFrame, const FrameT &LastFrame, const float th, const bool bMono, bool checkLevel = true) {
        const int n = (int)CurrentFrame.mvKeys.size(), m = LastFrame.N;
        if (n == 0 || m <= 0) return 0;
        float pose[30];
        const auto Rcw = CurrentFrame.mTcw.rotationMatrix();
        const auto tcw = CurrentFrame.mTcw.translation();
        const auto Rlw = LastFrame.mTcw.rotationMatrix();
        const auto tlw = LastFrame.mTcw.translation();
        for (int r = 0; r < 3; r++) {
            for (int c = 0; c < 3; c++) { pose[3 * r + c] = Rcw(r, c); pose[12 + 3 * r + c] = Rlw(r, c); }
            pose[9 + r] = tcw[r]; pose[21 + r] = tlw[r];
        }
        pose[24] = CurrentFrame.fx; pose[25] = CurrentFrame.fy; pose[26] = CurrentFrame.cx; pose[27] = CurrentFrame.cy;
        pose[28] = CurrentFrame.mbf; pose[29] = CurrentFrame.mb;
        std::vector<uint8_t> valid(m), hasObs(m), pdesc((size_t)m * 32), claimed(n), desc((size_t)n * 32);
        std::vector<float> wpos((size_t)m * 3), angle(m);
        std::vector<int32_t> octave(m);
        for (int i = 0; i < m; i++) {
            auto *p = LastFrame.mvpMapPoints[i];
            valid[i] = p && !LastFrame.mvbOutlier[i];
            octave[i] = LastFrame.mvKeys[i].octave; angle[i] = LastFrame.mvKeys[i].angle;
            if (!valid[i]) continue;
            hasObs[i] = p->Observations() > 0;
            const auto x3Dw = p->GetWorldPos();
            for (int r = 0; r < 3; r++) wpos[(size_t)i * 3 + r] = x3Dw[r];
            const cv::Mat d = p->GetDescriptor();
            std::memcpy(&pdesc[(size_t)i * 32], d.ptr(0), 32);
        }
        bool anyRight = false;
        for (int i = 0; i < n; i++) {
            claimed[i] = CurrentFrame.mvpMapPoints[i] && CurrentFrame.mvpMapPoints[i]->Observations() > 0;
            std::memcpy(&desc[(size_t)i * 32], CurrentFrame.mDescriptors.ptr(i), 32);
            anyRight |= CurrentFrame.mvuRight[i] > 0;
        }
        const int32_t nKp = n, nPt = m;
        ygzorb_frame_features fv{1, n, &nKp, (const ygzorb_keypoint *)CurrentFrame.mvKeys.data(), desc.data(),
                                 anyRight ? CurrentFrame.mvuRight.data() : nullptr, claimed.data(), CurrentFrame.mnMinX,
                                 CurrentFrame.mnMinY, CurrentFrame.mnMaxX, CurrentFrame.mnMaxY,
                                 (int)CurrentFrame.mvScaleFactors.size(), CurrentFrame.mvScaleFactors.data()};
        ygzorb_lastframe_points lv{m, &nPt, valid.data(), hasObs.data(), wpos.data(), pdesc.data(), octave.data(), angle.data(), pose};
        std::vector<int32_t> kpToPt(n), best(m);
        int32_t nmatches = 0;
        if (ygzorb_search_by_projection_last(mHandle, &fv, &lv, th, bMono ? 1 : 0, checkLevel ? 1 : 0, mbCheckOrientation ? 1 : 0, 0,
                                             kpToPt.data(), best.data(), &nmatches) != 0)
            throw std::runtime_error(std::string("ygzorb_search_by_projection_last: ") + ygzorb_last_error());
        // CurrentFrame.mvpMapPoints[bestIdx2] = pMP (:1312), then the rotation vote NULLs its losers (:1339-1342)
        for (int i = 0; i < m; i++)
            if (best[i] >= 0) CurrentFrame.mvpMapPoints[best[i]] = nullptr;
        for (int i = 0; i < n; i++)
            if (kpToPt[i] >= 0) CurrentFrame.mvpMapPoints[i] = LastFrame.mvpMapPoints[kpToPt[i]];
        return nmatches;
    }

    // Frame::ComputeStereoMatches() (src/Frame.cc:509-682) over the Frame members it uses; the patch search reads
    // the device-resident pyramids of F.mpORBextractorLeft / Right (their last extraction).
    template <class FrameT> void ComputeStereoMatches(FrameT &F) {
        const int nL = (int)F.mvKeys.size(), nR = (int)F.mvKeysRight.size();
        F.mvuRight.assign(nL, -1.0f);
        F.mvDepth.assign(nL, -1.0f);
        if (nL == 0) return;
        std::vector<uint8_t> dl((size_t)nL * 32), dr((size_t)nR * 32);
        for (int i = 0; i < nL; i++) std::memcpy(&dl[(size_t)i * 32], F.mDescriptors.ptr(i), 32);
        for (int i = 0; i < nR; i++) std::memcpy(&dr[(size_t)i * 32], F.mDescriptorsRight.ptr(i), 32);
        if (ygzorb_stereo_match(mHandle, F.mpORBextractorLeft->handle(), F.mpORBextractorRight->handle(), 0,
                                (const ygzorb_keypoint *)F.mvKeys.data(), dl.data(), nL,
                                (const ygzorb_keypoint *)F.mvKeysRight.data(), dr.data(), nR, F.mbf, F.mb, 0, F.mvuRight.data(),
                                F.mvDepth.data()) != 0)
            throw std::runtime_error(std::string("ygzorb_stereo_match: ") + ygzorb_last_error());
    }

    // src/ORBmatcher.cc:597-746.  KeyFrameT = the reference's KeyFrame (mFeatVec is a DBoW2::FeatureVector, i.e. a
    // std::map<unsigned, std::vector<unsigned>>); Mat3 = Matrix3f (anything with operator()(r, c)); the epipole
    // (:604-610) is computed here with the caller's own vector types.
    template <class KeyFrameT, class Mat3>
    int SearchForTriangulation(KeyFrameT *pKF1, KeyFrameT *pKF2, Mat3 &F12, std::vector<std::pair<size_t, size_t>> &vMatchedPairs,
                               const bool bOnlyStereo) {
        auto Cw = pKF1->GetCameraCenter();
        auto R2w = pKF2->GetRotation();
        auto t2w = pKF2->GetTranslation();
        auto C2 = R2w * Cw + t2w;
        const float invz = 1.0f / C2[2];
        const float ex = pKF2->fx * C2[0] * invz + pKF2->cx;
        const float ey = pKF2->fy * C2[1] * invz + pKF2->cy;
        struct Flat {
            std::vector<uint8_t> desc, hasMP; std::vector<uint32_t> ids, feats; std::vector<int32_t> start;
            ygzorb_keyframe_features v;
        };
        auto flatten = [](KeyFrameT *kf, Flat &f) {
            const int n = kf->N;
            f.desc.resize((size_t)n * 32); f.hasMP.resize(n);
            for (int i = 0; i < n; i++) {
                std::memcpy(&f.desc[(size_t)i * 32], kf->mDescriptors.ptr(i), 32);
                f.hasMP[i] = kf->GetMapPoint(i) != nullptr;
            }
            f.start.push_back(0);
            for (const auto &node : kf->mFeatVec) {           // std::map iterates in ascending NodeId
                f.ids.push_back(node.first);
                for (unsigned idx : node.second) f.feats.push_back(idx);
                f.start.push_back((int32_t)f.feats.size());
            }
            f.v = ygzorb_keyframe_features{n, (int)f.ids.size(), (int)kf->mvScaleFactors.size(),
                                           (const ygzorb_keypoint *)kf->mvKeys.data(), f.desc.data(), kf->mvuRight.data(),
                                           f.hasMP.data(), f.ids.data(), f.start.data(), f.feats.data(),
                                           kf->mvScaleFactors.data(), kf->mvLevelSigma2.data()};
        };
        Flat a, b;
        flatten(pKF1, a); flatten(pKF2, b);
        float f12[9];
        for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) f12[3 * r + c] = F12(r, c);
        std::vector<int32_t> m12(std::max(pKF1->N, 1));
        int32_t nmatches = 0;
        if (ygzorb_search_for_triangulation(mHandle, &a.v, &b.v, f12, ex, ey, bOnlyStereo ? 1 : 0, mbCheckOrientation ? 1 : 0, 0,
                                            m12.data(), &nmatches) != 0)
            throw std::runtime_error(std::string("ygzorb_search_for_triangulation: ") + ygzorb_last_error());
        vMatchedPairs.clear();
        vMatchedPairs.reserve(nmatches);
        for (int i = 0; i < pKF1->N; i++)
            if (m12[i] >= 0) vMatchedPairs.push_back(std::make_pair((size_t)i, (size_t)m12[i]));   // :736-743
        return nmatches;
    }

    // src/ORBmatcher.cc:155-263 (Tracking::TrackReferenceKeyFrame, Relocalization).  mFeatVec of both sides is a
    // DBoW2::FeatureVector (std::map<NodeId, std::vector<unsigned>>) as ORBVocabulary::transform fills it.
    template <class KeyFrameT, class FrameT, class MapPointT>
    int SearchByBoW(KeyFrameT *pKF, FrameT &F, std::vector<MapPointT *> &vpMapPointMatches) {
        const std::vector<MapPointT *> vpMapPointsKF = pKF->GetMapPointMatches();
        vpMapPointMatches.assign(F.N, nullptr);
        BowSide a, b;
        flattenBow(pKF->mvKeys, pKF->mDescriptors, pKF->mFeatVec, a);
        flattenBow(F.mvKeys, F.mDescriptors, F.mFeatVec, b);
        for (size_t i = 0; i < a.hasMP.size(); i++) a.hasMP[i] = vpMapPointsKF[i] && !vpMapPointsKF[i]->isBad();
        std::vector<int32_t> m(std::max<size_t>(b.hasMP.size(), 1));
        int32_t nmatches = 0;
        if (ygzorb_search_by_bow(mHandle, &a.v, &b.v, 1, mfNNratio, mbCheckOrientation ? 1 : 0, 0, m.data(), &nmatches) != 0)
            throw std::runtime_error(std::string("ygzorb_search_by_bow: ") + ygzorb_last_error());
        for (int i = 0; i < F.N; i++)
            if (m[i] >= 0) vpMapPointMatches[i] = vpMapPointsKF[m[i]];   // :209
        return nmatches;
    }

    // src/ORBmatcher.cc:480-596 (LoopClosing::ComputeSim3)
    template <class KeyFrameT, class MapPointT>
    int SearchByBoW(KeyFrameT *pKF1, KeyFrameT *pKF2, std::vector<MapPointT *> &vpMatches12) {
        const std::vector<MapPointT *> vpMapPoints1 = pKF1->GetMapPointMatches(), vpMapPoints2 = pKF2->GetMapPointMatches();
        vpMatches12.assign(vpMapPoints1.size(), nullptr);
        BowSide a, b;
        flattenBow(pKF1->mvKeys, pKF1->mDescriptors, pKF1->mFeatVec, a);
        flattenBow(pKF2->mvKeys, pKF2->mDescriptors, pKF2->mFeatVec, b);
        for (size_t i = 0; i < a.hasMP.size(); i++) a.hasMP[i] = vpMapPoints1[i] && !vpMapPoints1[i]->isBad();
        for (size_t i = 0; i < b.hasMP.size(); i++) b.hasMP[i] = vpMapPoints2[i] && !vpMapPoints2[i]->isBad();
        std::vector<int32_t> m(std::max<size_t>(a.hasMP.size(), 1));
        int32_t nmatches = 0;
        if (ygzorb_search_by_bow(mHandle, &a.v, &b.v, 0, mfNNratio, mbCheckOrientation ? 1 : 0, 0, m.data(), &nmatches) != 0)
            throw std::runtime_error(std::string("ygzorb_search_by_bow: ") + ygzorb_last_error());
        for (size_t i = 0; i < vpMapPoints1.size(); i++)
            if (m[i] >= 0) vpMatches12[i] = vpMapPoints2[m[i]];           // :546
        return nmatches;
    }

    ygzorb_matcher *handle() const { return mHandle; }

protected:
    struct BowSide {
        std::vector<uint8_t> desc, hasMP; std::vector<uint32_t> ids, feats; std::vector<int32_t> start;
        ygzorb_keyframe_features v;
    };
    template <class KeysT, class DescT, class FeatVecT>
    static void flattenBow(const KeysT &keys, const DescT &descriptors, const FeatVecT &featVec, BowSide &f) {
        const int n = (int)keys.size();
        f.desc.resize((size_t)std::max(n, 1) * 32); f.hasMP.assign(n, 1);
        for (int i = 0; i < n; i++) std::memcpy(&f.desc[(size_t)i * 32], descriptors.ptr(i), 32);
        f.start.push_back(0);
        for (const auto &node : featVec) {                   // std::map iterates in ascending NodeId
            f.ids.push_back(node.first);
            for (unsigned idx : node.second) f.feats.push_back(idx);
            f.start.push_back((int32_t)f.feats.size());
        }
        f.v = ygzorb_keyframe_features{n, (int)f.ids.size(), 0, (const ygzorb_keypoint *)keys.data(), f.desc.data(), nullptr,
                                       f.hasMP.data(), f.ids.data(), f.start.data(), f.feats.data(), nullptr, nullptr};
    }
    float mfNNratio;
    bool mbCheckOrientation;
    ygzorb_matcher *mHandle;
};

}  // namespace ygz
#endif
