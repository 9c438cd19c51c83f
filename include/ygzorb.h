/* ygzorb.h — C ABI of libygzorb_cuda.so: the B200 (sm_100a) implementation of ORB-YGZ-SLAM's ORB
 * front-end hot path.  Plain C types only; no exceptions, no ownership transfer.  Every function
 * returns 0 on success or a negative YGZORB_E_* code; ygzorb_last_error() gives the text of the
 * calling thread's last failure.  The reference has no FFI layer: its boundary is the C++ class API
 * (SURVEY.md §8(b)); each entry point cites the reference member it stands behind (file:line in
 * Ewenwan/ORB-YGZ-SLAM).  The C++ classes with the reference's own signatures that call this ABI
 * live in orb-ygz-slam_b200/host/.
 *
 * There is no CPU fallback in this library: without a CUDA device every compute entry point fails
 * with YGZORB_E_CUDA.
 */
#ifndef YGZORB_H_
#define YGZORB_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define YGZORB_OK 0
#define YGZORB_E_ARG (-1)      /* bad argument (null pointer, size out of range, ...) */
#define YGZORB_E_CUDA (-2)     /* CUDA runtime failure, text in ygzorb_last_error() */
#define YGZORB_E_CAPACITY (-3) /* a caller-provided output buffer is too small */
#define YGZORB_E_STATE (-4)    /* call sequence error (e.g. reading a pyramid before any extract) */

/* Buffer-placement flags of the batch entry points. */
#define YGZORB_IN_DEVICE 1u    /* input frames pointer is device memory (default: host) */
#define YGZORB_OUT_DEVICE 2u   /* counts / keypoints / descriptors pointers are device memory */
#define YGZORB_KEEP_STAGES 4u  /* keep blurred levels + FAST candidates readable (parity tests) */
#define YGZORB_NO_SYNC 8u      /* return after enqueueing; caller calls ygzorb_extractor_sync() */

const char *ygzorb_last_error(void);
const char *ygzorb_version(void);
/* Number of CUDA devices visible, or a negative error code. */
int ygzorb_device_count(void);

/* ---- ORBextractor ------------------------------------------------------------------------- */

/* The five constructor arguments of ygz::ORBextractor (include/ORBextractor.h:53-57,
 * src/ORBextractor.cc:412-470; read from the settings YAML at src/Tracking.cc:249-261). */
typedef struct ygzorb_params {
    int nfeatures;
    float scaleFactor;
    int nlevels;
    int iniThFAST;
    int minThFAST;
} ygzorb_params;

/* cv::KeyPoint, byte for byte (28 B): pt.x, pt.y, size, angle, response, octave, class_id. */
typedef struct ygzorb_keypoint {
    float x, y, size, angle, response;
    int32_t octave, class_id;
} ygzorb_keypoint;

typedef struct ygzorb_extractor ygzorb_extractor;

/* One extractor = one CUDA stream + workspace for up to max_batch frames of at most
 * max_w x max_h pixels on `device` (so the Left and Right extractors of a stereo rig,
 * src/Frame.cc:728-731, run concurrently).  Not re-entrant, like the reference object
 * (it owns mvImagePyramid, include/ORBextractor.h:112). */
int ygzorb_extractor_create(const ygzorb_params *params, int max_w, int max_h, int max_batch, int device,
                            ygzorb_extractor **out);
void ygzorb_extractor_destroy(ygzorb_extractor *ex);

/* Scale tables computed by the constructor (src/ORBextractor.cc:416-445) and umax (:451-466);
 * getters of include/ORBextractor.h:87-109.  Arrays of nlevels (umax16: 16 ints); any may be NULL. */
int ygzorb_extractor_tables(const ygzorb_extractor *ex, float *scale, float *inv_scale, float *sigma2,
                            float *inv_sigma2, int *n_per_level, int *umax16);
/* Level sizes for a w x h input (src/ORBextractor.cc:1131-1133). */
int ygzorb_extractor_level_sizes(const ygzorb_extractor *ex, int w, int h, int *lw, int *lh);
/* Upper bound of keypoints one frame can yield (sum over levels of the quadtree's list bound);
 * size per-frame output records with this. */
int ygzorb_extractor_max_keypoints(const ygzorb_extractor *ex, int w, int h);

/* ORBextractor::operator()(InputArray image, InputArray mask, vector<KeyPoint>&, OutputArray)
 * (src/ORBextractor.cc:970-1028; the Frame* overload :1031-1127 with a fresh Frame is the same
 * computation) on n_frames 8-bit single-channel frames of w x h, row stride `stride` bytes and
 * `frame_stride` bytes between frames.  Outputs, per frame f: counts[f] keypoints, records at
 * keypoints[f*cap .. f*cap+counts[f]) and descriptor rows (32 B, CV_8U) at desc[(f*cap+i)*32].
 * cap >= ygzorb_extractor_max_keypoints().  The mask is ignored, as in the reference
 * (include/ORBextractor.h:69).  An empty image (w or h 0) yields counts 0 (:972-973). */
int ygzorb_extract_batch(ygzorb_extractor *ex, const uint8_t *frames, int n_frames, int w, int h, int stride,
                         size_t frame_stride, unsigned flags, int cap, int32_t *counts,
                         ygzorb_keypoint *keypoints, uint8_t *desc);

/* ORBextractor::ComputePyramid(cv::Mat) alone (src/ORBextractor.cc:1129-1150) — what the Frame
 * constructor runs (src/Frame.cc:773-814).  Host or device input per flags. */
int ygzorb_compute_pyramid_batch(ygzorb_extractor *ex, const uint8_t *frames, int n_frames, int w, int h,
                                 int stride, size_t frame_stride, unsigned flags);

/* Copy level `level` of frame `frame` of the last batch into a continuous (step == cols) host
 * buffer of lw*lh bytes — mvImagePyramid[level] as Frame clones it (src/Frame.cc:810-813). */
int ygzorb_get_pyramid_level(ygzorb_extractor *ex, int frame, int level, uint8_t *dst);
/* All levels of one frame, packed back to back (continuous each), into dst (sum of lw*lh bytes). */
int ygzorb_get_pyramid(ygzorb_extractor *ex, int frame, uint8_t *dst);

/* Parity-test taps, valid after a call with YGZORB_KEEP_STAGES: the 7x7 sigma-2 blurred level
 * (src/ORBextractor.cc:1009-1010) and the per-cell FAST candidates of a level before the quadtree
 * (vToDistributeKeys, :747-781; x,y relative to (16,16) like the reference; UNORDERED — sort by
 * order_key to get the reference's cell-major/raster order).  Returns the candidate count. */
int ygzorb_get_blurred_level(ygzorb_extractor *ex, int frame, int level, uint8_t *dst);
int ygzorb_get_candidates(ygzorb_extractor *ex, int frame, int level, int cap, int16_t *x, int16_t *y,
                          uint8_t *response, uint32_t *order_key);

int ygzorb_extractor_sync(ygzorb_extractor *ex);
/* The extractor's cudaStream_t (as void*), for callers that time with CUDA events. */
void *ygzorb_extractor_stream(ygzorb_extractor *ex);
/* Kernel launches enqueued by this extractor since creation (bench.py's gpu_launches). */
long long ygzorb_extractor_launch_count(const ygzorb_extractor *ex);
/* Per-stage device time (ms) of the last batch when profiling is on: 0 upload, 1 pyramid, 2 FAST,
 * 3 quadtree, 4 blur, 5 orientation+descriptor, 6 download.  n <= 8. */
int ygzorb_extractor_set_profiling(ygzorb_extractor *ex, int on);
int ygzorb_extractor_stage_ms(ygzorb_extractor *ex, float *ms, int n);

/* Stand-alone stages on one host image (parity tests of a single kernel against the oracle /
 * cv2): cv::resize INTER_LINEAR 8UC1 (src/ORBextractor.cc:1139) and cv::GaussianBlur 7x7 s=2
 * REFLECT_101 (:1010).  Continuous buffers. */
int ygzorb_resize_linear_u8(const uint8_t *src, int sw, int sh, uint8_t *dst, int dw, int dh, int device);
int ygzorb_gaussian_blur7_u8(const uint8_t *src, int w, int h, uint8_t *dst, int device);

#ifdef __cplusplus
}
#endif
#endif /* YGZORB_H_ */
