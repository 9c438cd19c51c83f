// ref_prelude.h — TEST INFRASTRUCTURE ONLY.  Force-included (-include) when the reference's own
// src/ORBextractor.cc is compiled, unmodified and in place, into oracle/_ref/ (see oracle/Makefile).
// The reference's include/Common.h drags in Eigen, sophus, g2o, pangolin and glog, none of which
// exist in this image and none of which src/ORBextractor.cc uses; defining its include guard makes
// it a no-op, and what the file really needs (std, using-directives, OpenCV surface, LOG) comes
// from here.  OpenCV is routed to oracle/minicv (bit-exact against cv2 4.13, see its header).
#ifndef YGZ_REF_PRELUDE_H_
#define YGZ_REF_PRELUDE_H_
#define YGZ_COMMON_H_  // include/Common.h:1-2 guard -> skipped

#include <algorithm>
#include <iostream>
#include <list>
#include <map>
#include <memory>
#include <set>
#include <string>
#include <vector>
using namespace std;  // include/Common.h:19

#include "../minicv/minicv.hpp"
using cv::Mat;  // include/Common.h:62

// glog stand-in (include/Common.h:70): messages are dropped.
struct YgzNullLog {
    template <class T> YgzNullLog &operator<<(const T &) { return *this; }
    YgzNullLog &operator<<(std::ostream &(*)(std::ostream &)) { return *this; }
};
#define LOG(severity) YgzNullLog()
#endif
