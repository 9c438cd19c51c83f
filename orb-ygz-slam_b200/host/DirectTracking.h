// DirectTracking.h — drop-ins for SURVEY.md 8(f) row f4 over the reference's object types: MapPoint::ComputeDistinctiveDescriptors
// (src/MapPoint.cc:211-271; batched), ORBmatcher::FindDirectProjection (src/ORBmatcher.cc:1573-1602) and SparseImgAlign::run
// (src/SparseImageAlign.cc:20-55).  Templates over the Frame / KeyFrame / MapPoint types, so the header compiles against the
// reference's own classes and against host/FrameLite.h.  What an integration adds to Frame and KeyFrame is one member,
// `ygzorb_pyramid *mpDevPyramid`: the device-side copy of mvImagePyramid (taken with ygzorb_pyramid_from_extractor right after
// extraction — the pyramid never leaves the device — and shared by the KeyFrame made from the Frame, like the cv::Mats are,
// src/KeyFrame.cc:259-261).  Poses: the templates read a pose `T` through `T.unit_quaternion().coeffs().data()` (x, y, z, w)
// and `T.translation().data()` — Sophus::SE3f's own accessors — and compose TCR with the pose type's own operator* /
// inverse(), so the Sophus arithmetic stays the reference's.  No CPU implementation here: the work is in libygzorb_cuda.so.
#ifndef YGZ_HOST_DIRECTTRACKING_H_
#define YGZ_HOST_DIRECTTRACKING_H_

#include <cstring>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

#include "cvcompat.h"
#include "../../include/ygzorb.h"

namespace ygz {

template <class PoseT> inline void PoseToFloats(const PoseT &T, float *out7) {
    std::memcpy(out7, T.unit_quaternion().coeffs().data(), 4 * sizeof(float));
    std::memcpy(out7 + 4, T.translation().data(), 3 * sizeof(float));
}

// MapPoint::ComputeDistinctiveDescriptors for every point of `points` in one call (LocalMapping::ProcessNewKeyFrame,
// CreateNewMapPoints and SearchInNeighbors call it point by point in loops, src/LocalMapping.cc).  Reads mObservations and the
// keyframes' mDescriptors, writes mDescriptor; bad points and points without a live observation are left alone (:219-235).
template <class MapPointT> void ComputeDistinctiveDescriptors(ygzorb_matcher *handle, const std::vector<MapPointT *> &points) {
    std::vector<uint8_t> rows;
    std::vector<int32_t> start(1, 0);
    for (MapPointT *mp : points) {
        if (mp && !mp->isBad())
            for (const auto &o : mp->mObservations)
                if (!o.first->isBad()) { const uint8_t *d = o.first->mDescriptors.ptr(o.second); rows.insert(rows.end(), d, d + 32); }
        start.push_back((int32_t)(rows.size() / 32));
    }
    const int n = (int)points.size();
    if (n == 0) return;
    std::vector<int32_t> idx(n), med(n);
    std::vector<uint8_t> best((size_t)n * 32);
    if (ygzorb_distinctive_descriptors(handle, rows.data(), start.data(), n, 0, idx.data(), med.data(), best.data()) != 0)
        throw std::runtime_error(std::string("ygzorb_distinctive_descriptors: ") + ygzorb_last_error());
    for (int i = 0; i < n; i++)
        if (idx[i] >= 0) {
            points[i]->mDescriptor.create(1, 32, cv::CV_8U);
            std::memcpy(points[i]->mDescriptor.ptr(0), &best[(size_t)i * 32], 32);
        }
}

// One request of Tracking::SearchLocalPointsDirect (src/Tracking.cc:2284-2305): align MapPoint mp, observed in keyframe ref,
// in the current frame.  px_curr in: (mTrackProjX, mTrackProjY); out: the aligned pixel.
template <class KeyFrameT, class MapPointT> struct DirectRequest {
    KeyFrameT *ref; MapPointT *mp;
    float px_curr[2]; int search_level; bool success;
};

// ORBmatcher::FindDirectProjection over a batch of requests (the member function of the same name = a batch of one).
template <class KeyFrameT, class FrameT, class MapPointT>
void FindDirectProjectionBatch(ygzorb_matcher *handle, FrameT *curr, std::vector<DirectRequest<KeyFrameT, MapPointT>> &reqs) {
    const int n = (int)reqs.size();
    if (n == 0) return;
    std::map<KeyFrameT *, int> kfIndex;
    std::vector<ygzorb_direct_keyframe> kfs;
    std::vector<int32_t> kfi(n), oct(n), level(n);
    std::vector<float> pxRef(2 * (size_t)n), wpos(3 * (size_t)n), pxCurr(2 * (size_t)n);
    std::vector<uint8_t> ok(n);
    for (int i = 0; i < n; i++) {
        KeyFrameT *ref = reqs[i].ref;
        auto it = kfIndex.find(ref);
        if (it == kfIndex.end()) {
            ygzorb_direct_keyframe k;
            k.pyramid = ref->mpDevPyramid;
            const auto poseRef = ref->GetPose();
            PoseToFloats(poseRef, k.pose_ref);
            PoseToFloats(curr->mTcw * poseRef.inverse(), k.t_cur_ref);          // TCR (:1581-1582)
            k.fx = ref->fx; k.fy = ref->fy; k.cx = ref->cx; k.cy = ref->cy;
            it = kfIndex.emplace(ref, (int)kfs.size()).first;
            kfs.push_back(k);
        }
        kfi[i] = it->second;
        const size_t index = reqs[i].mp->mObservations.at(ref);                 // mp->GetObservations()[ref] (:1577)
        const cv::KeyPoint &kp = ref->mvKeys[index];
        pxRef[2 * i] = kp.pt.x; pxRef[2 * i + 1] = kp.pt.y; oct[i] = kp.octave;
        const float *w = reqs[i].mp->GetWorldPos();
        wpos[3 * i] = w[0]; wpos[3 * i + 1] = w[1]; wpos[3 * i + 2] = w[2];
        pxCurr[2 * i] = reqs[i].px_curr[0]; pxCurr[2 * i + 1] = reqs[i].px_curr[1];
    }
    const float intr[4] = {curr->fx, curr->fy, curr->cx, curr->cy};
    if (ygzorb_find_direct_projection(handle, kfs.data(), (int)kfs.size(), curr->mpDevPyramid, intr, curr->mvScaleFactors.data(),
                                      curr->mvInvScaleFactors.data(), reqs[0].ref->mvInvLevelSigma2[1], n, kfi.data(), pxRef.data(),
                                      oct.data(), wpos.data(), 0, pxCurr.data(), level.data(), ok.data()) != 0)
        throw std::runtime_error(std::string("ygzorb_find_direct_projection: ") + ygzorb_last_error());
    for (int i = 0; i < n; i++) {
        reqs[i].px_curr[0] = pxCurr[2 * i]; reqs[i].px_curr[1] = pxCurr[2 * i + 1];
        reqs[i].search_level = level[i]; reqs[i].success = ok[i] != 0;
    }
}

// SparseImgAlign (include/SparseImageAlign.h:19-118): same constructor arguments and run() signature; Gauss-Newton, unit weights.
class SparseImgAlign {
public:
    SparseImgAlign(ygzorb_matcher *handle, int max_level, int min_level, int n_iter = 10)
        : max_level_(max_level), min_level_(min_level), n_iter_(n_iter), sa_(nullptr), iterations_(0), chi2_(0) {
        if (n_iter != 10) throw std::runtime_error("SparseImgAlign: run() uses 10 iterations per level (src/SparseImageAlign.cc:39)");
        if (ygzorb_sparse_align_create(handle, &sa_) != 0)
            throw std::runtime_error(std::string("ygzorb_sparse_align_create: ") + ygzorb_last_error());
    }
    ~SparseImgAlign() { ygzorb_sparse_align_destroy(sa_); }
    SparseImgAlign(const SparseImgAlign &) = delete;
    SparseImgAlign &operator=(const SparseImgAlign &) = delete;

    // size_t run(Frame *ref_frame, Frame *cur_frame, SE3f &TCR) (src/SparseImageAlign.cc:20-55)
    template <class FrameT, class PoseT> size_t run(FrameT *ref_frame, FrameT *cur_frame, PoseT &TCR) {
        if (ref_frame->mvKeys.empty()) return 0;                                  // :24-27
        const int n = ref_frame->N;
        std::vector<float> kp(2 * (size_t)n), wpos(3 * (size_t)n, 0.f);
        std::vector<uint8_t> valid(n);
        for (int i = 0; i < n; i++) {
            auto *mp = ref_frame->mvpMapPoints[i];
            valid[i] = mp != nullptr && !mp->isBad() && !ref_frame->mvbOutlier[i];  // :68-70
            kp[2 * i] = ref_frame->mvKeys[i].pt.x; kp[2 * i + 1] = ref_frame->mvKeys[i].pt.y;
            if (valid[i]) { const float *w = mp->GetWorldPos(); wpos[3 * i] = w[0]; wpos[3 * i + 1] = w[1]; wpos[3 * i + 2] = w[2]; }
        }
        float refPose[7], T[7];
        PoseToFloats(ref_frame->mTcw, refPose);
        PoseToFloats(cur_frame->mTcw * ref_frame->mTcw.inverse(), T);             // T_cur_from_ref (:36)
        const float intr[4] = {cur_frame->fx, cur_frame->fy, cur_frame->cx, cur_frame->cy};
        if (ygzorb_sparse_align_begin(sa_, ref_frame->mpDevPyramid, cur_frame->mpDevPyramid, n, kp.data(), valid.data(), wpos.data(), refPose,
                                      ref_frame->fx, intr, ref_frame->mvInvScaleFactors.data()) != 0)
            throw std::runtime_error(std::string("ygzorb_sparse_align_begin: ") + ygzorb_last_error());
        int64_t tracked = 0;
        float stats[2];
        if (ygzorb_sparse_align_run(sa_, max_level_, min_level_, T, &tracked, stats) != 0)
            throw std::runtime_error(std::string("ygzorb_sparse_align_run: ") + ygzorb_last_error());
        iterations_ = (int)stats[0]; chi2_ = stats[1];
        TCR = PoseT::FromFloats(T);      // Sophus: SE3f(Eigen::Quaternionf(T[3], T[0], T[1], T[2]), Vector3f(T[4], T[5], T[6]))
        return (size_t)tracked;
    }
    int iterations() const { return iterations_; }
    float getChi2() const { return chi2_; }

private:
    int max_level_, min_level_, n_iter_;
    ygzorb_sparse_align *sa_;
    int iterations_;
    float chi2_;
};

}  // namespace ygz
#endif
