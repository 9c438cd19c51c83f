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
#include <cmath>
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

    // src/ORBmatcher.cc:375-478 (Tracking::MonocularInitialization).  Point2T = cv::Point2f (anything with x, y).
    template <class FrameT, class Point2T>
    int SearchForInitialization(FrameT &F1, FrameT &F2, std::vector<Point2T> &vbPrevMatched, std::vector<int> &vnMatches12,
                                int windowSize = 10) {
        const int n1 = (int)F1.mvKeys.size(), n2 = (int)F2.mvKeys.size();
        vnMatches12.assign(n1, -1);
        if (n1 == 0 || n2 == 0) return 0;
        std::vector<uint8_t> d1((size_t)n1 * 32), d2((size_t)n2 * 32);
        for (int i = 0; i < n1; i++) std::memcpy(&d1[(size_t)i * 32], F1.mDescriptors.ptr(i), 32);
        for (int i = 0; i < n2; i++) std::memcpy(&d2[(size_t)i * 32], F2.mDescriptors.ptr(i), 32);
        std::vector<float> prev((size_t)n1 * 2);
        for (int i = 0; i < n1; i++) { prev[2 * i] = vbPrevMatched[i].x; prev[2 * i + 1] = vbPrevMatched[i].y; }
        const int32_t c1 = n1, c2 = n2;
        ygzorb_frame_features v1{1, n1, &c1, (const ygzorb_keypoint *)F1.mvKeys.data(), d1.data(), nullptr, nullptr, F2.mnMinX,
                                 F2.mnMinY, F2.mnMaxX, F2.mnMaxY, 0, nullptr};
        ygzorb_frame_features v2{1, n2, &c2, (const ygzorb_keypoint *)F2.mvKeys.data(), d2.data(), nullptr, nullptr, F2.mnMinX,
                                 F2.mnMinY, F2.mnMaxX, F2.mnMaxY, 0, nullptr};
        std::vector<int32_t> m12(n1);
        int32_t nmatches = 0;
        if (ygzorb_search_for_initialization(mHandle, &v1, &v2, prev.data(), windowSize, mfNNratio, mbCheckOrientation ? 1 : 0, 0,
                                             m12.data(), &nmatches) != 0)
            throw std::runtime_error(std::string("ygzorb_search_for_initialization: ") + ygzorb_last_error());
        for (int i = 0; i < n1; i++) {
            vnMatches12[i] = m12[i];
            if (m12[i] >= 0) { vbPrevMatched[i].x = prev[2 * i]; vbPrevMatched[i].y = prev[2 * i + 1]; }
        }
        return nmatches;
    }

    // src/ORBmatcher.cc:1352-1469 (Tracking::Relocalization).  SetT = std::set<MapPoint*>.  MapPointT needs GetMaxDistance() /
    // GetMinDistance() returning mfMaxDistance / mfMinDistance (two one-line getters to add to MapPoint).
    template <class FrameT, class KeyFrameT, class SetT>
    int SearchByProjection(FrameT &CurrentFrame, KeyFrameT *pKF, const SetT &sAlreadyFound, const float th, const int ORBdist) {
        const auto vpMPs = pKF->GetMapPointMatches();
        const int n = (int)CurrentFrame.mvKeys.size(), m = (int)vpMPs.size();
        if (n == 0 || m == 0) return 0;
        PointArrays P(m);
        const auto Rcw = CurrentFrame.mTcw.rotationMatrix();
        const auto tcw = CurrentFrame.mTcw.translation();
        for (int r = 0; r < 3; r++) { for (int c = 0; c < 3; c++) P.pose[3 * r + c] = Rcw(r, c); P.pose[9 + r] = tcw[r]; }
        P.pose[24] = CurrentFrame.fx; P.pose[25] = CurrentFrame.fy; P.pose[26] = CurrentFrame.cx; P.pose[27] = CurrentFrame.cy;
        for (int i = 0; i < m; i++) {
            auto *pMP = vpMPs[i];
            P.angle[i] = pKF->mvKeys[i].angle;
            if (pMP && !pMP->isBad() && !sAlreadyFound.count(pMP)) P.set(i, pMP, false);
        }
        std::vector<uint8_t> claimed(n), desc((size_t)n * 32);
        for (int i = 0; i < n; i++) {
            claimed[i] = CurrentFrame.mvpMapPoints[i] != nullptr;
            std::memcpy(&desc[(size_t)i * 32], CurrentFrame.mDescriptors.ptr(i), 32);
        }
        const int32_t nKp = n, nPt = m;
        ygzorb_frame_features fv{1, n, &nKp, (const ygzorb_keypoint *)CurrentFrame.mvKeys.data(), desc.data(), nullptr, claimed.data(),
                                 CurrentFrame.mnMinX, CurrentFrame.mnMinY, CurrentFrame.mnMaxX, CurrentFrame.mnMaxY,
                                 (int)CurrentFrame.mvScaleFactors.size(), CurrentFrame.mvScaleFactors.data()};
        ygzorb_keyframe_points pv = P.view(&nPt, CurrentFrame.mfLogScaleFactor, CurrentFrame.mnScaleLevels);
        std::vector<int32_t> kpToPt(n);
        int32_t nmatches = 0;
        if (ygzorb_search_by_projection_reloc(mHandle, &fv, &pv, th, ORBdist, mbCheckOrientation ? 1 : 0, 0, kpToPt.data(), nullptr,
                                              &nmatches) != 0)
            throw std::runtime_error(std::string("ygzorb_search_by_projection_reloc: ") + ygzorb_last_error());
        for (int i = 0; i < n; i++)
            if (kpToPt[i] >= 0) CurrentFrame.mvpMapPoints[i] = vpMPs[kpToPt[i]];   // :1431; rotation losers stay NULL (:1459)
        return nmatches;
    }

    // src/ORBmatcher.cc:265-373 (LoopClosing::ComputeSim3).  MatT = cv::Mat: the decomposition of Scw (:275-278) is written
    // with the reference's own cv::Mat expressions, so this overload instantiates against OpenCV only.
    template <class KeyFrameT, class MatT, class MapPointT>
    int SearchByProjection(KeyFrameT *pKF, MatT Scw, const std::vector<MapPointT *> &vpPoints, std::vector<MapPointT *> &vpMatched, int th) {
        const int n = (int)pKF->mvKeys.size(), m = (int)vpPoints.size();
        if (n == 0 || m == 0) return 0;
        PointArrays P(m);
        decomposeSim3(Scw, P.pose);
        P.pose[24] = pKF->fx; P.pose[25] = pKF->fy; P.pose[26] = pKF->cx; P.pose[27] = pKF->cy;
        std::vector<MapPointT *> found(vpMatched.begin(), vpMatched.end());
        std::sort(found.begin(), found.end());
        for (int i = 0; i < m; i++) {
            MapPointT *pMP = vpPoints[i];
            if (!pMP->isBad() && !std::binary_search(found.begin(), found.end(), pMP)) P.set(i, pMP, true);
        }
        std::vector<uint8_t> claimed(n), desc((size_t)n * 32);
        for (int i = 0; i < n; i++) { claimed[i] = vpMatched[i] != nullptr; std::memcpy(&desc[(size_t)i * 32], pKF->mDescriptors.ptr(i), 32); }
        const int32_t nKp = n, nPt = m;
        ygzorb_frame_features fv{1, n, &nKp, (const ygzorb_keypoint *)pKF->mvKeys.data(), desc.data(), nullptr, claimed.data(),
                                 (float)pKF->mnMinX, (float)pKF->mnMinY, (float)pKF->mnMaxX, (float)pKF->mnMaxY,
                                 (int)pKF->mvScaleFactors.size(), pKF->mvScaleFactors.data()};
        ygzorb_keyframe_points pv = P.view(&nPt, pKF->mfLogScaleFactor, pKF->mnScaleLevels);
        std::vector<int32_t> kpToPt(n);
        int32_t nmatches = 0;
        if (ygzorb_search_by_projection_sim3(mHandle, &fv, &pv, th, 0, kpToPt.data(), nullptr, &nmatches) != 0)
            throw std::runtime_error(std::string("ygzorb_search_by_projection_sim3: ") + ygzorb_last_error());
        for (int i = 0; i < n; i++)
            if (kpToPt[i] >= 0) vpMatched[i] = vpPoints[kpToPt[i]];               // :364
        return nmatches;
    }

    // src/ORBmatcher.cc:748-886 (LocalMapping::SearchInNeighbors).  The window search of every MapPoint runs on the device;
    // the map updates (:862-877) are applied here in the reference's order with the reference's own calls.
    template <class KeyFrameT, class MapPointT>
    int Fuse(KeyFrameT *pKF, const std::vector<MapPointT *> &vpMapPoints, const float th = 3.0) {
        const int n = (int)pKF->mvKeys.size(), m = (int)vpMapPoints.size();
        if (n == 0 || m == 0) return 0;
        PointArrays P(m);
        const auto Rcw = pKF->GetRotation();
        const auto tcw = pKF->GetTranslation();
        const auto Ow = pKF->GetCameraCenter();
        for (int r = 0; r < 3; r++) { for (int c = 0; c < 3; c++) P.pose[3 * r + c] = Rcw(r, c); P.pose[9 + r] = tcw[r]; P.pose[12 + r] = Ow[r]; }
        P.pose[24] = pKF->fx; P.pose[25] = pKF->fy; P.pose[26] = pKF->cx; P.pose[27] = pKF->cy; P.pose[28] = pKF->mbf;
        for (int i = 0; i < m; i++) {
            MapPointT *pMP = vpMapPoints[i];
            if (pMP && !pMP->isBad() && !pMP->IsInKeyFrame(pKF)) P.set(i, pMP, true);
        }
        std::vector<int32_t> best(m);
        int32_t nFusedDev = 0;
        fuseSearch(pKF, P, th, 0, best, nFusedDev);
        int nFused = 0;
        for (int i = 0; i < m; i++) {
            MapPointT *pMP = vpMapPoints[i];
            // a duplicate entry may have been fused by an earlier iteration: the reference re-evaluates :767-771 per entry
            if (best[i] < 0 || !pMP || pMP->isBad() || pMP->IsInKeyFrame(pKF)) continue;
            MapPointT *pMPinKF = pKF->GetMapPoint(best[i]);
            if (pMPinKF) {
                if (!pMPinKF->isBad()) {
                    if (pMPinKF->Observations() > pMP->Observations()) pMP->Replace(pMPinKF);
                    else pMPinKF->Replace(pMP);
                }
            } else {
                pMP->AddObservation(pKF, best[i]);
                pKF->AddMapPoint(pMP, best[i]);
            }
            nFused++;
        }
        return nFused;
    }

    // src/ORBmatcher.cc:888-1004 (LoopClosing::SearchAndFuse).  MatT = cv::Mat, see the Sim3 SearchByProjection above.
    template <class KeyFrameT, class MatT, class MapPointT>
    int Fuse(KeyFrameT *pKF, MatT Scw, const std::vector<MapPointT *> &vpPoints, float th, std::vector<MapPointT *> &vpReplacePoint) {
        const int n = (int)pKF->mvKeys.size(), m = (int)vpPoints.size();
        if (n == 0 || m == 0) return 0;
        PointArrays P(m);
        decomposeSim3(Scw, P.pose);
        P.pose[24] = pKF->fx; P.pose[25] = pKF->fy; P.pose[26] = pKF->cx; P.pose[27] = pKF->cy;
        const auto spAlreadyFound = pKF->GetMapPoints();
        for (int i = 0; i < m; i++) {
            MapPointT *pMP = vpPoints[i];
            if (!pMP->isBad() && !spAlreadyFound.count(pMP)) P.set(i, pMP, true);
        }
        std::vector<int32_t> best(m);
        int32_t nFusedDev = 0;
        fuseSearch(pKF, P, th, 1, best, nFusedDev);
        int nFused = 0;
        for (int i = 0; i < m; i++) {
            if (best[i] < 0) continue;
            MapPointT *pMP = vpPoints[i];
            MapPointT *pMPinKF = pKF->GetMapPoint(best[i]);
            if (pMPinKF) {
                if (!pMPinKF->isBad()) vpReplacePoint[i] = pMPinKF;
            } else {
                pMP->AddObservation(pKF, best[i]);
                pKF->AddMapPoint(pMP, best[i]);
            }
            nFused++;
        }
        return nFused;
    }

    // src/ORBmatcher.cc:1006-1216 (LoopClosing::ComputeSim3).  MatT = cv::Mat; sR12, sR21, t21 (:1023-1025) and the
    // conversions of the keyframe poses are evaluated with the reference's expressions.
    template <class KeyFrameT, class MapPointT, class MatT>
    int SearchBySim3(KeyFrameT *pKF1, KeyFrameT *pKF2, std::vector<MapPointT *> &vpMatches12, const float &s12, const MatT &R12,
                     const MatT &t12, const float th) {
        const auto vpMapPoints1 = pKF1->GetMapPointMatches(), vpMapPoints2 = pKF2->GetMapPointMatches();
        const int N1 = (int)vpMapPoints1.size(), N2 = (int)vpMapPoints2.size();
        if (N1 == 0 || N2 == 0) return 0;
        MatT sR12 = s12 * R12;
        MatT sR21 = (1.0 / s12) * R12.t();
        MatT t21 = -sR21 * t12;
        std::vector<bool> vbAlreadyMatched1(N1, false), vbAlreadyMatched2(N2, false);
        for (int i = 0; i < N1; i++) {
            MapPointT *pMP = vpMatches12[i];
            if (pMP) {
                vbAlreadyMatched1[i] = true;
                int idx2 = pMP->GetIndexInKeyFrame(pKF2);
                if (idx2 >= 0 && idx2 < N2) vbAlreadyMatched2[idx2] = true;
            }
        }
        PointArrays P1(N1), P2(N2);
        auto side = [&](KeyFrameT *from, const MatT &sR, const MatT &t, PointArrays &P) {
            const auto Rw = from->GetRotation();
            const auto tw = from->GetTranslation();
            for (int r = 0; r < 3; r++) {
                for (int c = 0; c < 3; c++) { P.pose[3 * r + c] = Rw(r, c); P.pose[12 + 3 * r + c] = sR.template at<float>(r, c); }
                P.pose[9 + r] = tw[r]; P.pose[21 + r] = t.template at<float>(r);
            }
            P.pose[24] = pKF1->fx; P.pose[25] = pKF1->fy; P.pose[26] = pKF1->cx; P.pose[27] = pKF1->cy;   // :1008-1011
        };
        side(pKF1, sR21, t21, P1); side(pKF2, sR12, t12, P2);
        for (int i = 0; i < N1; i++) if (vpMapPoints1[i] && !vbAlreadyMatched1[i] && !vpMapPoints1[i]->isBad()) P1.set(i, vpMapPoints1[i], false);
        for (int i = 0; i < N2; i++) if (vpMapPoints2[i] && !vbAlreadyMatched2[i] && !vpMapPoints2[i]->isBad()) P2.set(i, vpMapPoints2[i], false);
        std::vector<uint8_t> d1((size_t)N1 * 32), d2((size_t)N2 * 32);
        for (int i = 0; i < N1; i++) std::memcpy(&d1[(size_t)i * 32], pKF1->mDescriptors.ptr(i), 32);
        for (int i = 0; i < N2; i++) std::memcpy(&d2[(size_t)i * 32], pKF2->mDescriptors.ptr(i), 32);
        const int32_t c1 = N1, c2 = N2;
        ygzorb_frame_features v1{1, N1, &c1, (const ygzorb_keypoint *)pKF1->mvKeys.data(), d1.data(), nullptr, nullptr, (float)pKF1->mnMinX,
                                 (float)pKF1->mnMinY, (float)pKF1->mnMaxX, (float)pKF1->mnMaxY, (int)pKF1->mvScaleFactors.size(),
                                 pKF1->mvScaleFactors.data()};
        ygzorb_frame_features v2{1, N2, &c2, (const ygzorb_keypoint *)pKF2->mvKeys.data(), d2.data(), nullptr, nullptr, (float)pKF2->mnMinX,
                                 (float)pKF2->mnMinY, (float)pKF2->mnMaxX, (float)pKF2->mnMaxY, (int)pKF2->mvScaleFactors.size(),
                                 pKF2->mvScaleFactors.data()};
        ygzorb_keyframe_points p1 = P1.view(&c1, pKF2->mfLogScaleFactor, pKF2->mnScaleLevels);
        ygzorb_keyframe_points p2 = P2.view(&c2, pKF1->mfLogScaleFactor, pKF1->mnScaleLevels);
        std::vector<int32_t> m12(N1);
        int32_t nFound = 0;
        if (ygzorb_search_by_sim3(mHandle, &v1, &v2, &p1, &p2, th, 0, m12.data(), &nFound) != 0)
            throw std::runtime_error(std::string("ygzorb_search_by_sim3: ") + ygzorb_last_error());
        for (int i = 0; i < N1; i++)
            if (m12[i] >= 0) vpMatches12[i] = vpMapPoints2[m12[i]];               // :1190
        return nFound;
    }

    ygzorb_matcher *handle() const { return mHandle; }

protected:
    // flat per-point arrays of ygzorb_keyframe_points
    struct PointArrays {
        std::vector<uint8_t> valid, desc;
        std::vector<float> wpos, normal, angle, maxD, minD;
        float pose[30];
        explicit PointArrays(int m) : valid(m, 0), desc((size_t)m * 32), wpos((size_t)m * 3), normal((size_t)m * 3), angle(m, 0.f),
                                      maxD(m, 1.f), minD(m, 1.f) { std::memset(pose, 0, sizeof pose); }
        template <class MapPointT> void set(int i, MapPointT *pMP, bool withNormal) {
            valid[i] = 1;
            const auto X = pMP->GetWorldPos();
            for (int r = 0; r < 3; r++) wpos[(size_t)i * 3 + r] = X[r];
            if (withNormal) { const auto Pn = pMP->GetNormal(); for (int r = 0; r < 3; r++) normal[(size_t)i * 3 + r] = Pn[r]; }
            maxD[i] = pMP->GetMaxDistance(); minD[i] = pMP->GetMinDistance();
            const cv::Mat d = pMP->GetDescriptor();
            std::memcpy(&desc[(size_t)i * 32], d.ptr(0), 32);
        }
        ygzorb_keyframe_points view(const int32_t *nPts, float logScaleFactor, int nScaleLevels) const {
            return ygzorb_keyframe_points{(int)valid.size(), nPts, valid.data(), wpos.data(), normal.data(), desc.data(), angle.data(),
                                          maxD.data(), minD.data(), pose, logScaleFactor, nScaleLevels};
        }
    };
    // Scw -> Rcw, tcw, Ow with the reference's cv::Mat expressions (src/ORBmatcher.cc:275-278, :897-901)
    template <class MatT> static void decomposeSim3(const MatT &Scw, float *pose) {
        MatT sRcw = Scw.rowRange(0, 3).colRange(0, 3);
        const float scw = sqrt(sRcw.row(0).dot(sRcw.row(0)));
        MatT Rcw = sRcw / scw;
        MatT tcw = Scw.rowRange(0, 3).col(3) / scw;
        MatT Ow = -Rcw.t() * tcw;
        for (int r = 0; r < 3; r++) {
            for (int c = 0; c < 3; c++) pose[3 * r + c] = Rcw.template at<float>(r, c);
            pose[9 + r] = tcw.template at<float>(r); pose[12 + r] = Ow.template at<float>(r);
        }
    }
    template <class KeyFrameT> void fuseSearch(KeyFrameT *pKF, const PointArrays &P, float th, int sim3, std::vector<int32_t> &best,
                                               int32_t &nFused) {
        const int n = (int)pKF->mvKeys.size();
        std::vector<uint8_t> desc((size_t)n * 32);
        for (int i = 0; i < n; i++) std::memcpy(&desc[(size_t)i * 32], pKF->mDescriptors.ptr(i), 32);
        const int32_t nKp = n, nPt = (int32_t)P.valid.size();
        ygzorb_frame_features fv{1, n, &nKp, (const ygzorb_keypoint *)pKF->mvKeys.data(), desc.data(), pKF->mvuRight.data(), nullptr,
                                 (float)pKF->mnMinX, (float)pKF->mnMinY, (float)pKF->mnMaxX, (float)pKF->mnMaxY,
                                 (int)pKF->mvScaleFactors.size(), pKF->mvScaleFactors.data()};
        ygzorb_keyframe_points pv = P.view(&nPt, pKF->mfLogScaleFactor, pKF->mnScaleLevels);
        if (ygzorb_fuse(mHandle, &fv, &pv, th, sim3, pKF->mvInvLevelSigma2.data(), 0, best.data(), &nFused) != 0)
            throw std::runtime_error(std::string("ygzorb_fuse: ") + ygzorb_last_error());
    }
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
