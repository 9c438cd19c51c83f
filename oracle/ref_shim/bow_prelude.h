// bow_prelude.h — TEST INFRASTRUCTURE ONLY.  Force-included (-include) when the reference's own
// Thirdparty/DBoW2 sources are compiled, unmodified and in place, into oracle/_ref/ygz_ref_bow
// (see oracle/Makefile).  DBoW2 needs <opencv2/core/core.hpp> only for cv::Mat (routed to
// oracle/minicv) and for cv::FileStorage in its YAML save/load members, which this oracle never
// calls: they get a do-nothing stand-in so that the header parses.  <sstream> is included here
// because TemplatedVocabulary.h relies on OpenCV dragging it in.
#ifndef YGZ_BOW_PRELUDE_H_
#define YGZ_BOW_PRELUDE_H_
#include <sstream>
#include <string>
#include <vector>

#include "../minicv/minicv.hpp"
using cv::CV_8U;
using cv::CV_32F;
namespace cv {
struct FileNode {
    FileNode operator[](const char *) const { return FileNode(); }
    FileNode operator[](const std::string &) const { return FileNode(); }
    FileNode operator[](int) const { return FileNode(); }
    size_t size() const { return 0; }
    operator int() const { return 0; }
    operator double() const { return 0; }
    operator std::string() const { return std::string(); }
    int type() const { return 0; }
    enum { SEQ = 5 };
};
struct FileStorage {
    enum { READ = 0, WRITE = 1 };
    FileStorage() {}
    FileStorage(const std::string &, int) {}
    bool isOpened() const { return false; }
    void release() {}
    FileNode operator[](const char *) const { return FileNode(); }
    FileNode operator[](const std::string &) const { return FileNode(); }
};
template <class T> inline FileStorage &operator<<(FileStorage &fs, const T &) { return fs; }
}  // namespace cv
#endif
