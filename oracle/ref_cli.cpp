// ref_cli.cpp — TEST INFRASTRUCTURE ONLY.  Driver around the reference's own ygz::ORBextractor
// (src/ORBextractor.cc, compiled unmodified from /root/reference by oracle/Makefile into
// oracle/_ref/ygz_ref_orb).  It is the checker for the CUDA path (tests/) and the timed CPU
// reference arm of bench.py; nothing in the product links or executes it.
//
// Canonicalisation of the one place where the reference is allocator-dependent: DistributeOctTree
// sorts pair<int, ExtractorNode*> (src/ORBextractor.cc:653-657), i.e. breaks size ties by heap
// address.  This binary replaces global operator new with a per-thread monotonic bump arena, so
// address order == creation order — a legal execution of the reference, and the rule the CUDA
// kernel and oracle/orb_oracle.cpp implement explicitly.
//
// usage: ygz_ref_orb --in frames.u8 --w W --h H --n N [--nfeatures 1000 --scale 1.2 --levels 8
//                    --ini 20 --min 7] [--out out.bin] [--pyr] [--frame-api] [--repeat R] [--threads T]
//                    [--dso [--exist exist.bin] [--grid G0]]
// out.bin, per frame: int32 n | n x 28 B cv::KeyPoint | n x 32 B descriptors |
//                     (--pyr) per level: int32 w, int32 h, w*h bytes | (--dso) int32 mnGridSize after the frame.
// --dso: Frame::ExtractORB's direct-tracking branch (src/Frame.cc:335-337): ONE extractor walks the frames in order (its
// mnGridSize carries over, src/ORBextractor.cc:1295-1376) and runs operator()(Frame*, mvKeys, desc, DSO_KEYPOINT) on a Frame
// that already holds the keypoints of exist.bin (per frame: int32 n | n x 28 B cv::KeyPoint); --grid seeds mnGridSize.
#include <sys/mman.h>

#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <thread>
#include <vector>

#define protected public   // mnGridSize (include/ORBextractor.h:191) is read back for the DSO keypoint mode; layout is unaffected
#include "ORBextractor.h"
#undef protected
#include "Frame.h"

// ------------------------------------------------------------------ monotonic bump arena
namespace {
constexpr size_t kArenaBytes = size_t(6) << 30;  // virtual; committed lazily
struct Arena { char *base, *top, *end; };
thread_local Arena t_arena = {nullptr, nullptr, nullptr};
struct ArenaRange { char *base, *end; };
ArenaRange g_ranges[512];
std::atomic<int> g_nranges{0};
std::mutex g_mu;

void arenaInit() {
    void *p = mmap(nullptr, kArenaBytes, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
    if (p == MAP_FAILED) { std::fputs("ygz_ref_orb: mmap failed\n", stderr); std::abort(); }
    t_arena.base = t_arena.top = (char *)p;
    t_arena.end = t_arena.base + kArenaBytes;
    std::lock_guard<std::mutex> lk(g_mu);
    int i = g_nranges.load();
    if (i >= 512) std::abort();
    g_ranges[i] = {t_arena.base, t_arena.end};
    g_nranges.store(i + 1);
}
inline bool inArena(void *p) {
    const int n = g_nranges.load(std::memory_order_acquire);
    for (int i = 0; i < n; i++)
        if ((char *)p >= g_ranges[i].base && (char *)p < g_ranges[i].end) return true;
    return false;
}
}  // namespace

void *operator new(size_t n) {
    if (!t_arena.base) arenaInit();
    n = (n + 15) & ~size_t(15);
    char *p = t_arena.top;
    if (p + n > t_arena.end) { std::fputs("ygz_ref_orb: arena exhausted\n", stderr); std::abort(); }
    t_arena.top = p + n;
    return p;
}
void *operator new[](size_t n) { return operator new(n); }
void operator delete(void *p) noexcept { if (p && !inArena(p)) std::free(p); }
void operator delete[](void *p) noexcept { operator delete(p); }
void operator delete(void *p, size_t) noexcept { operator delete(p); }
void operator delete[](void *p, size_t) noexcept { operator delete(p); }

// ------------------------------------------------------------------ driver
struct Opts {
    std::string in, out;
    int w = 0, h = 0, n = 1, nfeatures = 1000, levels = 8, ini = 20, mn = 7, repeat = 1, threads = 1;
    float scale = 1.2f;
    bool pyr = false, frameApi = false, dso = false;
    std::string exist;
    int grid = -1;
};

struct FrameOut {
    int n = 0;
    unsigned char *kps = nullptr;   // malloc'd, n*28
    unsigned char *desc = nullptr;  // malloc'd, n*32
    std::vector<int> lw, lh;
    std::vector<unsigned char *> lv;
    int gridAfter = 0;
};

static void storeOut(FrameOut *fo, const std::vector<cv::KeyPoint> &kps, const cv::Mat &desc) {
    fo->n = (int)kps.size();
    fo->kps = (unsigned char *)std::malloc((size_t)fo->n * 28 + 1);
    fo->desc = (unsigned char *)std::malloc((size_t)fo->n * 32 + 1);
    if (fo->n) std::memcpy(fo->kps, kps.data(), (size_t)fo->n * 28);
    for (int i = 0; i < fo->n; i++) std::memcpy(fo->desc + (size_t)i * 32, desc.ptr(i), 32);
}

// --dso (DSO_KEYPOINT) or --frame-api --exist (ORBSLAM_KEYPOINT on Frames that already hold keypoints,
// src/ORBextractor.cc:1086-1099): all frames through one extractor, in order
static void runDso(const Opts &o, const unsigned char *frames, std::vector<FrameOut> &outs) {
    ygz::ORBextractor ex(o.nfeatures, o.scale, o.levels, o.ini, o.mn);
    if (o.grid > 0) ex.mnGridSize = o.grid;
    FILE *fe = o.exist.empty() ? nullptr : std::fopen(o.exist.c_str(), "rb");
    const size_t fsz = (size_t)o.w * o.h;
    for (int k = 0; k < o.n; k++) {
        cv::Mat im(o.h, o.w, cv::CV_8UC1, (void *)(frames + fsz * k), (size_t)o.w);
        ygz::Frame f;
        ex.ComputePyramid(im);
        f.mvImagePyramid.resize(o.levels);
        for (int l = 0; l < o.levels; l++) f.mvImagePyramid[l] = ex.mvImagePyramid[l].clone();
        if (fe) {
            int ne = 0;
            if (std::fread(&ne, 4, 1, fe) != 1) { std::fputs("short exist file\n", stderr); std::exit(1); }
            f.mvKeys.resize(ne);
            if (ne && std::fread(f.mvKeys.data(), 28, (size_t)ne, fe) != (size_t)ne) { std::fputs("short exist file\n", stderr); std::exit(1); }
        }
        f.N = (int)f.mvKeys.size();
        cv::Mat desc;
        ex(&f, f.mvKeys, desc, o.dso ? ygz::ORBextractor::DSO_KEYPOINT : ygz::ORBextractor::ORBSLAM_KEYPOINT, true);
        storeOut(&outs[k], f.mvKeys, desc);
        outs[k].gridAfter = ex.mnGridSize;
    }
    if (fe) std::fclose(fe);
}

static void runOne(const Opts &o, const unsigned char *img, FrameOut *fo) {
    // Everything allocated between mark and release dies inside this scope.
    char *mark = t_arena.base ? t_arena.top : nullptr;
    {
        ygz::ORBextractor ex(o.nfeatures, o.scale, o.levels, o.ini, o.mn);
        cv::Mat im(o.h, o.w, cv::CV_8UC1, (void *)img, (size_t)o.w);
        std::vector<cv::KeyPoint> kps;
        cv::Mat desc;
        if (!o.frameApi) {
            ex(im, cv::Mat(), kps, desc);  // src/ORBextractor.cc:970-1028
        } else {
            // Frame.cc:773-814 + 332-348: Frame ctor computes and clones the pyramid, then the
            // Frame* overload (src/ORBextractor.cc:1031-1127) extracts into frame->mvKeys.
            ygz::Frame f;
            ex.ComputePyramid(im);
            f.mvImagePyramid.resize(o.levels);
            for (int l = 0; l < o.levels; l++) f.mvImagePyramid[l] = ex.mvImagePyramid[l].clone();
            ex(&f, f.mvKeys, desc, ygz::ORBextractor::ORBSLAM_KEYPOINT, true);
            kps = f.mvKeys;
        }
        if (fo) {
            fo->n = (int)kps.size();
            fo->kps = (unsigned char *)std::malloc((size_t)fo->n * 28 + 1);
            fo->desc = (unsigned char *)std::malloc((size_t)fo->n * 32 + 1);
            if (fo->n) std::memcpy(fo->kps, kps.data(), (size_t)fo->n * 28);
            for (int i = 0; i < fo->n; i++) std::memcpy(fo->desc + (size_t)i * 32, desc.ptr(i), 32);
            if (o.pyr) {
                for (int l = 0; l < o.levels; l++) {
                    const cv::Mat &m = ex.mvImagePyramid[l];
                    unsigned char *b = (unsigned char *)std::malloc((size_t)m.rows * m.cols + 1);
                    for (int r = 0; r < m.rows; r++) std::memcpy(b + (size_t)r * m.cols, m.ptr(r), (size_t)m.cols);
                    fo->lw.push_back(m.cols); fo->lh.push_back(m.rows); fo->lv.push_back(b);
                }
            }
        }
    }
    if (mark) t_arena.top = mark;
}

int main(int argc, char **argv) {
    Opts o;
    for (int i = 1; i < argc; i++) {
        std::string a = argv[i];
        auto next = [&]() -> const char * { if (i + 1 >= argc) { std::fprintf(stderr, "missing value for %s\n", a.c_str()); std::exit(2); } return argv[++i]; };
        if (a == "--in") o.in = next();
        else if (a == "--out") o.out = next();
        else if (a == "--w") o.w = std::atoi(next());
        else if (a == "--h") o.h = std::atoi(next());
        else if (a == "--n") o.n = std::atoi(next());
        else if (a == "--nfeatures") o.nfeatures = std::atoi(next());
        else if (a == "--scale") o.scale = (float)std::atof(next());
        else if (a == "--levels") o.levels = std::atoi(next());
        else if (a == "--ini") o.ini = std::atoi(next());
        else if (a == "--min") o.mn = std::atoi(next());
        else if (a == "--repeat") o.repeat = std::atoi(next());
        else if (a == "--threads") o.threads = std::atoi(next());
        else if (a == "--pyr") o.pyr = true;
        else if (a == "--frame-api") o.frameApi = true;
        else if (a == "--dso") o.dso = true;
        else if (a == "--exist") o.exist = next();
        else if (a == "--grid") o.grid = std::atoi(next());
        else { std::fprintf(stderr, "unknown option %s\n", a.c_str()); return 2; }
    }
    if (o.in.empty() || o.w <= 0 || o.h <= 0 || o.n <= 0) { std::fputs("need --in --w --h --n\n", stderr); return 2; }
    const size_t fsz = (size_t)o.w * o.h;
    unsigned char *frames = (unsigned char *)std::malloc(fsz * o.n);
    FILE *fi = std::fopen(o.in.c_str(), "rb");
    if (!fi || std::fread(frames, 1, fsz * o.n, fi) != fsz * o.n) { std::fputs("cannot read input frames\n", stderr); return 1; }
    std::fclose(fi);

    std::vector<FrameOut> outs((size_t)o.n);
    // touch the main arena so mark/release is active from the first frame
    delete new int(0);

    const auto t0 = std::chrono::steady_clock::now();
    const int total = o.n * o.repeat;
    if (o.dso || (o.frameApi && !o.exist.empty())) {
        runDso(o, frames, outs);
    } else if (o.threads <= 1) {
        for (int k = 0; k < total; k++) {
            const int f = k % o.n;
            runOne(o, frames + fsz * f, k < o.n ? &outs[f] : nullptr);
        }
    } else {
        std::atomic<int> next{0};
        std::vector<std::thread> th;
        for (int t = 0; t < o.threads; t++)
            th.emplace_back([&]() {
                delete new int(0);
                for (;;) {
                    const int k = next.fetch_add(1);
                    if (k >= total) break;
                    const int f = k % o.n;
                    runOne(o, frames + fsz * f, k < o.n ? &outs[f] : nullptr);
                }
            });
        for (auto &t : th) t.join();
    }
    const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();

    if (!o.out.empty()) {
        FILE *fo = std::fopen(o.out.c_str(), "wb");
        if (!fo) { std::fputs("cannot open output\n", stderr); return 1; }
        for (int f = 0; f < o.n; f++) {
            const FrameOut &r = outs[f];
            std::fwrite(&r.n, 4, 1, fo);
            std::fwrite(r.kps, 28, (size_t)r.n, fo);
            std::fwrite(r.desc, 32, (size_t)r.n, fo);
            for (size_t l = 0; l < r.lv.size(); l++) {
                std::fwrite(&r.lw[l], 4, 1, fo);
                std::fwrite(&r.lh[l], 4, 1, fo);
                std::fwrite(r.lv[l], 1, (size_t)r.lw[l] * r.lh[l], fo);
            }
            if (o.dso) std::fwrite(&r.gridAfter, 4, 1, fo);
        }
        std::fclose(fo);
    }
    std::printf("{\"frames\": %d, \"seconds\": %.6f, \"frames_per_s\": %.3f, \"threads\": %d, \"w\": %d, \"h\": %d, \"nfeatures\": %d}\n",
                total, sec, total / sec, o.threads, o.w, o.h, o.nfeatures);
    return 0;
}
