// host_latency.cc — single-frame latency THROUGH the drop-in class ygz::ORBextractor::operator() (the call Frame::ExtractORB
// makes, src/Frame.cc:332-348), i.e. including the pageable-image upload, the keypoint / descriptor download and the copy of the
// pyramid into mvImagePyramid that the reference's callers rely on.  BASELINE.json configs[0] (640x480, 1000 features) and
// configs[2] (1241x376 stereo pair, 2000 features per image, Left / Right extractors on two host threads like
// src/Frame.cc:728-731, + ComputeStereoMatches).
// usage: host_latency <frames.u8> <w> <h> <nfeatures> <reps> [stereo]     (frames.u8: one image, or left|right for stereo)
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <thread>
#include <vector>

#include "FrameLite.h"
#include "ORBextractor.h"
#include "ORBmatcher.h"

using namespace ygz;

int main(int argc, char **argv) {
    if (argc < 6) { fprintf(stderr, "usage: host_latency frames.u8 w h nfeatures reps [stereo]\n"); return 2; }
    const int w = atoi(argv[2]), h = atoi(argv[3]), nf = atoi(argv[4]), reps = atoi(argv[5]);
    const bool stereo = argc > 6;
    FILE *f = fopen(argv[1], "rb");
    if (!f) { perror("frames"); return 2; }
    cv::Mat left(h, w, cv::CV_8UC1), right(h, w, cv::CV_8UC1);
    if (fread(left.data, 1, (size_t)w * h, f) != (size_t)w * h) { fprintf(stderr, "short read\n"); return 2; }
    if (stereo && fread(right.data, 1, (size_t)w * h, f) != (size_t)w * h) { fprintf(stderr, "short read\n"); return 2; }
    fclose(f);
    ORBextractor exL(nf, 1.2f, 8, 20, 7), exR(nf, 1.2f, 8, 20, 7);
    ORBmatcher matcher;
    std::vector<double> us;
    int nL = 0, nR = 0, nStereo = 0;
    for (int r = 0; r < reps + 5; r++) {
        Frame F;
        F.mpORBextractorLeft = &exL; F.mpORBextractorRight = &exR;
        const auto t0 = std::chrono::steady_clock::now();
        if (!stereo) {
            exL(left, cv::_InputArray(), F.mvKeys, F.mDescriptors);
        } else {
            std::thread tl([&]() { exL(left, cv::_InputArray(), F.mvKeys, F.mDescriptors); });
            std::thread tr([&]() { exR(right, cv::_InputArray(), F.mvKeysRight, F.mDescriptorsRight); });
            tl.join(); tr.join();
            F.N = (int)F.mvKeys.size();
            F.mvScaleFactors = exL.GetScaleFactors(); F.mvInvScaleFactors = exL.GetInverseScaleFactors();
            F.mbf = 386.1448f; F.mb = F.mbf / 718.856f;
            matcher.ComputeStereoMatches(F);
        }
        const double dt = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t0).count();
        if (r >= 5) us.push_back(dt);
        nL = (int)F.mvKeys.size(); nR = (int)F.mvKeysRight.size();
        nStereo = 0;
        for (float u : F.mvuRight) nStereo += u >= 0;
    }
    std::sort(us.begin(), us.end());
    double sum = 0;
    for (double v : us) sum += v;
    printf("{\"through\": \"ygz::ORBextractor::operator() (host class, pageable image in, keypoints + descriptors + mvImagePyramid out)\", "
           "\"w\": %d, \"h\": %d, \"nfeatures\": %d, \"stereo\": %s, \"reps\": %d, \"latency_us_median\": %.1f, \"latency_us_mean\": %.1f, "
           "\"latency_us_p90\": %.1f, \"keypoints_left\": %d, \"keypoints_right\": %d, \"stereo_matches\": %d}\n",
           w, h, nf, stereo ? "true" : "false", reps, us[us.size() / 2], sum / us.size(), us[(size_t)(us.size() * 0.9)], nL, nR, nStereo);
    return 0;
}
