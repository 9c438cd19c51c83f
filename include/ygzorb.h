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
#define YGZORB_KEEP_STAGES 4u  /* one chunk per max_batch frames (no upload pipelining), so that the stage taps
                                  (pyramid, blurred levels, FAST candidates) cover every frame of the call */
#define YGZORB_OUT_PINNED 32u   /* counts / keypoints / descriptors are PAGE-LOCKED host memory (ygzorb_host_alloc, cudaHostAlloc,
                                 * cudaHostRegister): calls of up to 8 frames let the last kernel write them directly (no device slot,
                                 * no three device-to-host copies: ~30 us less per live-tracking frame).  Ignored for larger calls and
                                 * together with YGZORB_OUT_DEVICE.  Passing pageable memory with this flag is an error the library
                                 * cannot detect cheaply: the kernel would fault. */
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

/* The undistortion in front of the pyramid, Frame::ComputeImagePyramid (src/Frame.cc:775-805): cv::remap(mImGray, dst, map1,
 * map2, cv::INTER_LINEAR) with the fixed-point maps cv::initUndistortRectifyMap(..., CV_16SC2, map1, map2) gives (map1: w*h
 * pairs of int16 (x, y); map2: w*h uint16 interpolation-table indices; HOST pointers, copied).  Once set, every frame given to
 * ygzorb_extract_batch / ygzorb_compute_pyramid_batch / ygzorb_extract_dso_batch (which must be w x h) is remapped on the
 * device first — bit-exact with OpenCV's 8-bit remapBilinear, border value 0 — and level 0 of the pyramid is the undistorted
 * image (ygzorb_get_pyramid_level(.., 0, ..) returns it, what Frame keeps as mImGray).  map1 == NULL removes the maps. */
int ygzorb_extractor_set_undistort_maps(ygzorb_extractor *ex, const int16_t *map1, const uint16_t *map2, int w, int h);

/* ORBextractor::operator()(Frame *frame, vector<KeyPoint> &_keypoints, OutputArray _descriptors, DSO_KEYPOINT, leftEye = true)
 * (src/ORBextractor.cc:1031-1127 with :1052-1058) as Frame::ExtractORB calls it on frames that come out of direct tracking
 * (src/Frame.cc:335-337): ComputeKeyPointsDSOSingleLevel (:1275-1386) — cells of mnGridSize pixels on level 0, FAST-10
 * (Thirdparty/fast; the scalar detector without a margin for cells narrower than 22 px, the SSE2 one with its 3-px margin
 * otherwise) at threshold 20 then 5, pixels of existing keypoints excluded, Shi-Tomasi score (:1152-1187), the best three per
 * cell, the grid shrunk by 5 until nfeatures corners are found — then IC_Angle for the new keypoints and the existing ones
 * (:1383-1385) and descriptors of all of them on the blurred pyramid (:1079-1113).
 * keypoints is in-out: on entry the first n_exist[f] records of frame f (at keypoints + f*cap) are frame->mvKeys (level-0
 * coordinates, octave = their level); on return records [0, counts[f]) are the existing keypoints with refreshed angles
 * followed by the new ones (size 7, octave 0, response 0), desc holds one row per record.  keypoints / desc / counts follow
 * YGZORB_OUT_DEVICE (the existing keypoints are then read from device memory too).
 * grid_size (HOST, in-out, one per frame): mnGridSize before / after the frame (< 0: not yet initialised, :1295-1296).  Frames
 * of one batch are independent extractors; a single camera stream feeds grid_size[0] of one call into the next.
 * n_exist: HOST.  n_frames <= max_batch.  Canonical where the reference is undefined: equal Shi-Tomasi scores keep raster
 * order (std::sort is unstable beyond 16 corners), a frame without any corner stops instead of looping forever (:1299-1301),
 * an existing keypoint closer than 16 px to the border of its level gets a zero descriptor and keeps its angle. */
int ygzorb_extract_dso_batch(ygzorb_extractor *ex, const uint8_t *frames, int n_frames, int w, int h, int stride,
                             size_t frame_stride, unsigned flags, int32_t *grid_size, const int32_t *n_exist, int cap,
                             int32_t *counts, ygzorb_keypoint *keypoints, uint8_t *desc);

/* The "existing keys" rows of ORBextractor::operator()(Frame *frame, ..., ORBSLAM_KEYPOINT, leftEye = true) on a Frame that
 * already holds keypoints (src/ORBextractor.cc:1086-1099): rBRIEF descriptors of n keypoints (HOST records in level-0
 * coordinates, octave = their level, angle as they carry it — not recomputed, :1091) sampled on the blurred level `octave` of
 * frame `frame` of the LAST ygzorb_extract_batch call, at cvRound(pt * mvInvScaleFactor[octave]) (:1095-1096, :111).  desc: HOST,
 * n x 32 B.  Canonical where the reference is undefined: a centre closer than 16 px to the border of its level (the reference
 * would sample outside the image) gets a zero descriptor.  The host class puts these rows in front of the new OctTree keys. */
int ygzorb_describe_existing(ygzorb_extractor *ex, int frame, int n, const ygzorb_keypoint *keypoints, uint8_t *desc);

/* ORBextractor::ComputePyramid(cv::Mat) alone (src/ORBextractor.cc:1129-1150) — what the Frame
 * constructor runs (src/Frame.cc:773-814).  Host or device input per flags. */
int ygzorb_compute_pyramid_batch(ygzorb_extractor *ex, const uint8_t *frames, int n_frames, int w, int h,
                                 int stride, size_t frame_stride, unsigned flags);

/* Copy level `level` of frame `frame` of the last batch into a continuous (step == cols) host
 * buffer of lw*lh bytes — mvImagePyramid[level] as Frame clones it (src/Frame.cc:810-813). */
int ygzorb_get_pyramid_level(ygzorb_extractor *ex, int frame, int level, uint8_t *dst);
/* All levels of one frame, packed back to back (continuous each), into dst (sum of lw*lh bytes). */
int ygzorb_get_pyramid(ygzorb_extractor *ex, int frame, uint8_t *dst);

/* Live tracking (one frame per call): the host buffer that receives mvImagePyramid.  Once set (NULL removes it), every
 * ygzorb_extract_batch / ygzorb_compute_pyramid_batch call with n_frames == 1 copies all levels of the frame, continuous and back
 * to back like ygzorb_get_pyramid, to host_dst on the download stream right after the pyramid kernels, i.e. while FAST, the
 * quadtree, the blur and the descriptors run; the data is complete when the call returns (or after ygzorb_extractor_sync with
 * YGZORB_NO_SYNC).  host_dst must hold the sum of lw*lh bytes and should be page-locked (ygzorb_host_alloc).  Replaces the
 * separate ygzorb_get_pyramid round trip in Frame::ComputeImagePyramid's clone (src/Frame.cc:807-813). */
int ygzorb_extractor_set_pyramid_sink(ygzorb_extractor *ex, uint8_t *host_dst);

/* Page-locked host memory (cudaHostAlloc / cudaFreeHost) for the buffers a caller hands to the host-pointer entry points:
 * copies from and to it run at the full rate of the link and asynchronously.  No reference counterpart; the host classes keep
 * their staging (the pyramid slab behind mvImagePyramid, src/Frame.cc:810-813 copies out of it) in such memory. */
int ygzorb_host_alloc(size_t bytes, void **out);
void ygzorb_host_free(void *p);

/* Parity-test taps, valid after a call with YGZORB_KEEP_STAGES: the 7x7 sigma-2 blurred level
 * (src/ORBextractor.cc:1009-1010) and the per-cell FAST candidates of a level before the quadtree
 * (vToDistributeKeys, :747-781; x,y relative to (16,16) like the reference; UNORDERED — sort by
 * order_key to get the reference's cell-major/raster order).  Returns the candidate count. */
int ygzorb_get_blurred_level(ygzorb_extractor *ex, int frame, int level, uint8_t *dst);
int ygzorb_get_candidates(ygzorb_extractor *ex, int frame, int level, int cap, int16_t *x, int16_t *y,
                          uint8_t *response, uint32_t *order_key);

/* Wait for everything the extractor enqueued (kernels and the host-output downloads). */
int ygzorb_extractor_sync(ygzorb_extractor *ex);
/* Host-input calls are pipelined in chunks of `frames_per_chunk` frames (default 256): the upload of
 * chunk k+1 and the download of chunk k-1 run on their own streams beside the kernels of chunk k, and
 * with YGZORB_NO_SYNC also beside the tail of the previous call.  0 = one chunk per max_batch.  The stage
 * taps then hold the LAST chunk only (pass YGZORB_KEEP_STAGES for parity tests). */
int ygzorb_extractor_set_pipeline(ygzorb_extractor *ex, int frames_per_chunk);
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

/* ---- ORBmatcher Hamming paths and Frame::ComputeStereoMatches ------------------------------- */

#define YGZORB_TH_HIGH 100   /* ORBmatcher::TH_HIGH, src/ORBmatcher.cc:36 */
#define YGZORB_TH_LOW 50     /* ORBmatcher::TH_LOW,  src/ORBmatcher.cc:37 */

/* One matcher = one CUDA stream + scratch on `device`.  ORBmatcher objects are stack-local per call
 * site and used concurrently from the Tracking / LocalMapping / LoopClosing threads (SURVEY.md
 * §8(b)): create one handle per calling thread; calls on different handles run concurrently.
 * Pointer placement per call: YGZORB_IN_DEVICE = every input array pointer is device memory,
 * YGZORB_OUT_DEVICE = every output array pointer is device memory; default host. */
typedef struct ygzorb_matcher ygzorb_matcher;
int ygzorb_matcher_create(int device, ygzorb_matcher **out);
void ygzorb_matcher_destroy(ygzorb_matcher *m);
int ygzorb_matcher_sync(ygzorb_matcher *m);
void *ygzorb_matcher_stream(ygzorb_matcher *m);
/* Run this matcher's work on another cudaStream_t of the same device (e.g. ygzorb_extractor_stream(), so
 * that a search consumes an extraction's device-resident outputs in stream order); NULL = its own stream. */
int ygzorb_matcher_set_stream(ygzorb_matcher *m, void *stream);
long long ygzorb_matcher_launch_count(const ygzorb_matcher *m);

/* ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:1507-1523) on n row pairs a[i], b[i] (32 B rows). */
int ygzorb_descriptor_distance(ygzorb_matcher *m, const uint8_t *a, const uint8_t *b, int n, unsigned flags,
                               int32_t *dist);

/* Brute-force Hamming kNN (k = 2) + ratio test (BASELINE.json config 5): for every query row the
 * best / second-best rule of ORBmatcher::SearchByBoW (src/ORBmatcher.cc:193-217) over all nt train
 * rows in index order.  best_idx/d1/d2 = bestIdxF/bestDist1/bestDist2 (-1/256/256 if nt == 0);
 * match[q] = best_idx if d1 <= th and (float)d1 < ratio*(float)d2 (:207-209), else -1.  Any output
 * pointer may be NULL. */
int ygzorb_hamming_knn2(ygzorb_matcher *m, const uint8_t *query, int nq, const uint8_t *train, int nt, float ratio,
                        int th, unsigned flags, int32_t *best_idx, int32_t *d1, int32_t *d2, int32_t *match);
/* Selects the POPC/LOP3 balance of the kNN kernel (0: 8 POPC per pair ... 3: 4 POPC + carry-save adders);
 * all variants return identical results.  Default 2. */
int ygzorb_matcher_set_knn_variant(ygzorb_matcher *m, int variant);
/* Measured integer-pipe peak of the device in 32-bit POPC per second (register-only popc/xor/add
 * microkernel) — the denominator of the Hamming roofline. */
int ygzorb_measure_popc_peak(ygzorb_matcher *m, double *popc_per_s);

/* ygzorb_hamming_knn2 for callers that cannot allocate.  NO COUNTERPART IN THE REFERENCE: added
 * for code under CUDA graph capture, tight serving loops and harnesses that forbid cudaMalloc.  Same computation
 * and results as ygzorb_hamming_knn2 (the same kernels), but: no matcher handle, every pointer is DEVICE memory on `device`
 * (query nq x 32 B, train nt x 32 B; outputs nq x int32, any may be NULL), the work is enqueued on `stream` (a cudaStream_t
 * as void*, NULL = the default stream) and NOT synchronised, and the kernels' scratch (12 B per query and train split) is a
 * file-scope static __device__ array of YGZORB_NOALLOC_SCRATCH_BYTES that every GPU loading the library holds until exit.
 * The call makes no cudaMalloc / cudaFree / allocating library call; the first call on a device resolves the array's device
 * address (cudaGetSymbolAddress, once per device).  Limit: nq x splits x 12 B <= YGZORB_NOALLOC_SCRATCH_BYTES, which every
 * nq <= YGZORB_NOALLOC_MAX_QUERIES satisfies; beyond it the call fails with YGZORB_E_CAPACITY and a message (use
 * ygzorb_hamming_knn2).  All calls on a device share the array: it keeps the last call's contents (nothing in it has to
 * start at zero: every partial result the merge kernel reads is written by the scan kernel of the same call), and two calls
 * that overlap on the GPU (different streams, or concurrent host threads) RACE on it — serialise them. */
#define YGZORB_NOALLOC_SCRATCH_BYTES (8 * 1024 * 1024)
#define YGZORB_NOALLOC_MAX_QUERIES 600000
int ygzorb_hamming_knn2_noalloc(int device, void *stream, const uint8_t *query, int nq, const uint8_t *train, int nt,
                                float ratio, int th, int32_t *best_idx, int32_t *d1, int32_t *d2, int32_t *match);

/* The Frame arrays the searches read (include/Frame.h:248-332) for n_frames frames at fixed strides
 * (frame f: n_kp[f] records starting at f*kp_stride). */
typedef struct ygzorb_frame_features {
    int n_frames, kp_stride;
    const int32_t *n_kp;          /* Frame::N */
    const ygzorb_keypoint *kps;   /* mvKeys / mvKeysUn */
    const uint8_t *desc;          /* mDescriptors, 32 B rows */
    const float *u_right;         /* mvuRight, or NULL (monocular: all -1) */
    const uint8_t *claimed;       /* 1 where mvpMapPoints[i] is set with Observations() > 0 on entry, or NULL */
    float min_x, min_y, max_x, max_y; /* mnMinX, mnMinY, mnMaxX, mnMaxY (src/Frame.cc:502-507) */
    int nlevels;
    const float *scale_factors;   /* mvScaleFactors[nlevels], HOST pointer always */
} ygzorb_frame_features;

/* MapPoints after Frame::isInFrustum (src/Frame.cc:363-422; MapPoint.h:95-102), frame f: n_mp[f]
 * entries starting at f*mp_stride, in vpMapPoints order. */
typedef struct ygzorb_mappoint_proj {
    int mp_stride;
    const int32_t *n_mp;
    const float *proj_x, *proj_y, *proj_xr, *view_cos;   /* mTrackProjX/Y/XR, mTrackViewCos */
    const int32_t *level;                                /* mnTrackScaleLevel */
    const uint8_t *in_view;                              /* mbTrackInView && !isBad() */
    const uint8_t *has_obs;                              /* Observations() > 0 */
    const uint8_t *desc;                                 /* GetDescriptor(), 32 B rows */
} ygzorb_mappoint_proj;

/* ORBmatcher::SearchByProjection(Frame&, const vector<MapPoint*>&, th, checkLevel)
 * (src/ORBmatcher.cc:44-126), batched over frames, reproducing the serial claim order.
 * kp_to_mp[f*kp_stride + i] = index of the MapPoint left in F.mvpMapPoints[i] by the call, or -1;
 * mp_best_idx[f*mp_stride + j] = keypoint MapPoint j was assigned to, or -1 (may be NULL);
 * n_matches[f] = the function's return value. */
int ygzorb_search_by_projection(ygzorb_matcher *m, const ygzorb_frame_features *frames, const ygzorb_mappoint_proj *mps,
                                float th, int check_level, float nnratio, unsigned flags, int32_t *kp_to_mp,
                                int32_t *mp_best_idx, int32_t *n_matches);

/* LastFrame side of ORBmatcher::SearchByProjection(Frame &CurrentFrame, const Frame &LastFrame, th, bMono,
 * checkLevel) (src/ORBmatcher.cc:1218-1350): frame pair f has n_pts[f] entries starting at f*pt_stride, entry i =
 * LastFrame keypoint i. */
typedef struct ygzorb_lastframe_points {
    int pt_stride;
    const int32_t *n_pts;       /* LastFrame.N */
    const uint8_t *valid;       /* LastFrame.mvpMapPoints[i] != NULL && !LastFrame.mvbOutlier[i] (:1244-1246) */
    const uint8_t *has_obs;     /* that MapPoint's Observations() > 0 */
    const float *world_pos;     /* GetWorldPos(), 3 floats per entry */
    const uint8_t *desc;        /* GetDescriptor(), 32 B rows */
    const int32_t *octave;      /* LastFrame.mvKeys[i].octave */
    const float *angle;         /* LastFrame.mvKeys[i].angle */
    /* per frame pair 30 floats: Rcw (row-major 9), tcw (3) of CurrentFrame.mTcw; Rlw (9), tlw (3) of
     * LastFrame.mTcw; CurrentFrame fx, fy, cx, cy, mbf, mb */
    const float *pose;
} ygzorb_lastframe_points;

/* The motion-model matcher (src/ORBmatcher.cc:1218-1350 + ComputeThreeMaxima :1471-1502), batched over frame
 * pairs and reproducing the serial claim / overwrite order and the rotation-histogram rejection.
 * `cur` is CurrentFrame (claimed = entries of mvpMapPoints that hold a MapPoint with observations on entry;
 * Tracking clears them first, Tracking.cc:1017).  Float evaluation order of the projection: 3-term dot products
 * as a0 + (a1 + a2), uncontracted; invzc = (float)(1.0 / (double)z); a non-finite projection is out of view.
 * kp_to_pt[f*kp_stride + i2] = LastFrame index whose MapPoint is left in CurrentFrame.mvpMapPoints[i2], or -1;
 * pt_best_idx[f*pt_stride + i] = keypoint entry i matched before the rotation check, or -1 (may be NULL);
 * n_matches[f] = the function's return value. */
int ygzorb_search_by_projection_last(ygzorb_matcher *m, const ygzorb_frame_features *cur,
                                     const ygzorb_lastframe_points *last, float th, int mono, int check_level,
                                     int check_orientation, unsigned flags, int32_t *kp_to_pt, int32_t *pt_best_idx,
                                     int32_t *n_matches);

/* The MapPoints of a KeyFrame (or a candidate list) for the two projection searches that predict the scale level from
 * the distance: entry i of frame pair f at f*pt_stride + i. */
typedef struct ygzorb_keyframe_points {
    int pt_stride;
    const int32_t *n_pts;
    const uint8_t *valid;        /* pMP != NULL && !pMP->isBad() && not in the already-found set (:1369-1372, :291-292) */
    const float *world_pos;      /* GetWorldPos() */
    const float *normal;         /* GetNormal() (Sim3 variant only; may be NULL otherwise) */
    const uint8_t *desc;         /* GetDescriptor() */
    const float *angle;          /* pKF->mvKeys[i].angle (relocalisation variant with check_orientation; else may be NULL) */
    const float *max_distance;   /* mfMaxDistance (GetMaxDistanceInvariance() / 1.2f: the raw member, PredictScale reads it) */
    const float *min_distance;   /* mfMinDistance */
    /* per frame pair 30 floats: [0..8] Rcw row-major, [9..11] tcw, [12..14] Ow (Sim3 variant: -Rcw^T tcw as the caller's
     * cv::Mat arithmetic gives it, :278; the relocalisation variant derives it itself), [24..27] fx, fy, cx, cy */
    const float *pose;
    float log_scale_factor;      /* mfLogScaleFactor */
    int n_scale_levels;          /* mnScaleLevels */
} ygzorb_keyframe_points;

/* ORBmatcher::SearchByProjection(Frame &CurrentFrame, KeyFrame *pKF, const set<MapPoint*> &sAlreadyFound, th, ORBdist)
 * (src/ORBmatcher.cc:1352-1469; Tracking::Relocalization).  cur->claimed = CurrentFrame.mvpMapPoints[i] != NULL on entry;
 * a keypoint taken by an earlier point of this call is skipped by later ones, as there.  Eigen evaluation order as in
 * ygzorb_search_by_projection_last; PredictScale uses the correctly rounded float logarithm.  Outputs as in
 * ygzorb_search_by_projection_last. */
int ygzorb_search_by_projection_reloc(ygzorb_matcher *m, const ygzorb_frame_features *cur, const ygzorb_keyframe_points *pts,
                                      float th, int orb_dist, int check_orientation, unsigned flags, int32_t *kp_to_pt,
                                      int32_t *pt_best_idx, int32_t *n_matches);

/* ORBmatcher::SearchByProjection(KeyFrame *pKF, cv::Mat Scw, const vector<MapPoint*> &vpPoints, vector<MapPoint*>
 * &vpMatched, int th) (src/ORBmatcher.cc:265-373; LoopClosing::ComputeSim3).  kf = the KeyFrame's keypoints / descriptors
 * with claimed = vpMatched[idx] != NULL on entry and min/max = mnMinX.. (IsInImage); the decomposition of Scw (:275-278)
 * stays with the caller and arrives in pose.  cv::Mat arithmetic as OpenCV 4.13 evaluates it: the 3x3 by 3x1 gemm as
 * ((a0 + a1) + a2) + c in float, cv::norm / Mat::dot in double.  kp_to_pt[idx] = index into vpPoints newly stored in
 * vpMatched[idx], or -1. */
int ygzorb_search_by_projection_sim3(ygzorb_matcher *m, const ygzorb_frame_features *kf, const ygzorb_keyframe_points *pts,
                                     int th, unsigned flags, int32_t *kp_to_pt, int32_t *pt_best_idx, int32_t *n_matches);

/* ORBmatcher::Fuse, the search half of both overloads (the MapPoint graph updates — Replace / AddObservation /
 * AddMapPoint, :862-877 and :989-998 — stay with the caller, who walks pt_best_idx in order):
 *   sim3 = 0: Fuse(KeyFrame *pKF, const vector<MapPoint*> &vpMapPoints, th) (src/ORBmatcher.cc:748-886).  Eigen
 *     arithmetic; pose[12..14] = pKF->GetCameraCenter(), pose[28] = mbf; kf->u_right = mvuRight and inv_level_sigma2 =
 *     mvInvLevelSigma2 feed the chi-square gate (:835-858).  valid = pMP && !isBad() && !IsInKeyFrame(pKF).
 *   sim3 = 1: Fuse(KeyFrame *pKF, cv::Mat Scw, vpPoints, th, vpReplacePoint) (:888-1004).  cv::Mat arithmetic, pose as in
 *     ygzorb_search_by_projection_sim3; valid = !isBad() && not in pKF->GetMapPoints().
 * pt_best_idx[f*pt_stride + i] = bestIdx when bestDist <= TH_LOW, else -1; n_fused[f] = number of such points (nFused). */
int ygzorb_fuse(ygzorb_matcher *m, const ygzorb_frame_features *kf, const ygzorb_keyframe_points *pts, float th, int sim3,
                const float *inv_level_sigma2, unsigned flags, int32_t *pt_best_idx, int32_t *n_fused);

/* ORBmatcher::SearchBySim3(pKF1, pKF2, vpMatches12, s12, R12, t12, th) (src/ORBmatcher.cc:1006-1216).  pts1 = KF1's
 * MapPoints, entry i1 per keypoint (pt_stride == kf1->kp_stride), valid = pMP && !vbAlreadyMatched1[i1] && !isBad();
 * its pose = R1w (0..8), t1w (9..11), sR21 (12..20), t21 (21..23), fx, fy, cx, cy (24..27; the reference uses KF1's
 * intrinsics for both directions), and the image bounds of kf2 gate the projection.  pts2 likewise with R2w, t2w, sR12,
 * t12.  sR12, sR21, t21 (:1023-1025) are the caller's cv::Mat results.  matches12[f*kp_stride1 + i1] = idx2 when both
 * directions agree (:1182-1196), else -1; n_found[f] = nFound. */
int ygzorb_search_by_sim3(ygzorb_matcher *m, const ygzorb_frame_features *kf1, const ygzorb_frame_features *kf2,
                          const ygzorb_keyframe_points *pts1, const ygzorb_keyframe_points *pts2, float th, unsigned flags,
                          int32_t *matches12, int32_t *n_found);

/* ORBmatcher::SearchForInitialization(Frame &F1, Frame &F2, vector<cv::Point2f> &vbPrevMatched, vector<int> &vnMatches12,
 * int windowSize) (src/ORBmatcher.cc:375-478; Tracking::MonocularInitialization), batched over frame pairs with the serial
 * take-over order (:412, :426-433) reproduced.  f1 / f2: keypoints and descriptors (claimed, u_right unused; the grid
 * bounds of f2 are used).  prev_matched: vbPrevMatched as (x, y) floats, [n_frames * f1->kp_stride * 2], read and updated
 * in place (:472-475; device memory when YGZORB_IN_DEVICE, else host).  matches12[f*kp_stride1 + i1] = vnMatches12. */
int ygzorb_search_for_initialization(ygzorb_matcher *m, const ygzorb_frame_features *f1, const ygzorb_frame_features *f2,
                                     float *prev_matched, int window_size, float nnratio, int check_orientation,
                                     unsigned flags, int32_t *matches12, int32_t *n_matches);

/* KeyFrame arrays read by SearchForTriangulation (include/KeyFrame.h:236-271); the DBoW2::FeatureVector
 * (Thirdparty/DBoW2/DBoW2/FeatureVector.h:23-24, std::map<NodeId, vector<unsigned>>) flattened to CSR:
 * node_ids ascending, features of node k = node_feats[node_start[k] .. node_start[k+1]). */
typedef struct ygzorb_keyframe_features {
    int n, n_nodes, nlevels;
    const ygzorb_keypoint *kps;    /* mvKeysUn */
    const uint8_t *desc;
    const float *u_right;          /* mvuRight (< 0: monocular keypoint) */
    const uint8_t *has_mappoint;   /* GetMapPoint(i) != NULL */
    const uint32_t *node_ids;
    const int32_t *node_start;
    const uint32_t *node_feats;
    const float *scale_factors;    /* HOST pointers always */
    const float *level_sigma2;
} ygzorb_keyframe_features;

/* ORBmatcher::SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs, bOnlyStereo)
 * (src/ORBmatcher.cc:597-746).  f12 row-major (f12[3*r+c] = F12(r,c)); (ex, ey) = the epipole of :604-610.
 * matches12[i] = matched keypoint of KF2 or -1 (vMatches12; the pair list is its non-negative entries
 * in index order); *n_matches = return value. */
int ygzorb_search_for_triangulation(ygzorb_matcher *m, const ygzorb_keyframe_features *kf1,
                                    const ygzorb_keyframe_features *kf2, const float *f12, float ex, float ey,
                                    int only_stereo, int check_orientation, unsigned flags, int32_t *matches12,
                                    int32_t *n_matches);

/* ORBmatcher::SearchByBoW, both overloads, over FeatureVectors flattened like above (the output of
 * ygzorb_bow_transform plugs in directly):
 *   second_is_frame = 1: SearchByBoW(KeyFrame *pKF, Frame &F, vector<MapPoint*> &vpMapPointMatches)
 *     (src/ORBmatcher.cc:155-263; Tracking::TrackReferenceKeyFrame, Relocalization).  kf1->has_mappoint = the
 *     keyframe's MapPoint exists and !isBad(); matches[iF] (length second->n) = keyframe keypoint whose MapPoint frame
 *     feature iF received, or -1.  Accepts bestDist1 <= TH_LOW and bestDist1 < nnratio * bestDist2.
 *   second_is_frame = 0: SearchByBoW(KeyFrame *pKF1, KeyFrame *pKF2, vector<MapPoint*> &vpMatches12) (:480-596;
 *     LoopClosing).  has_mappoint on both sides; matches[idx1] (length kf1->n) = matched keypoint of KF2, or -1.
 *     Accepts bestDist1 < TH_LOW (strict, as written there) and the ratio.
 * u_right, scale_factors and level_sigma2 of the views are not read.  *n_matches = return value. */
int ygzorb_search_by_bow(ygzorb_matcher *m, const ygzorb_keyframe_features *kf1, const ygzorb_keyframe_features *second,
                         int second_is_frame, float nnratio, int check_orientation, unsigned flags, int32_t *matches,
                         int32_t *n_matches);

/* Frame::ComputeStereoMatches() (src/Frame.cc:509-682).  The patch search reads the pyramids of the
 * Left and Right extractors (mpORBextractorLeft/Right->mvImagePyramid, :604-623): frame `frame` of the
 * last batch of `left` / `right`.  kps/desc: mvKeys / mDescriptors and mvKeysRight / mDescriptorsRight.
 * Outputs mvuRight[n_left], mvDepth[n_left] (-1 = no match).  mbf, mb as in Frame (bf, bf/fx). */
int ygzorb_stereo_match(ygzorb_matcher *m, ygzorb_extractor *left, ygzorb_extractor *right, int frame,
                        const ygzorb_keypoint *kps_left, const uint8_t *desc_left, int n_left,
                        const ygzorb_keypoint *kps_right, const uint8_t *desc_right, int n_right, float mbf, float mb,
                        unsigned flags, float *u_right, float *depth);

/* ---- Bag of words: DBoW2::TemplatedVocabulary<cv::Mat, FORB> on the device --------------------
 * A vocabulary is the tree of Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:98-121 (struct Node) as flat arrays: node 0 is
 * the root; children of node i are child_list[child_start[i] .. child_start[i+1]) in the order of Node::children; a node
 * without children is a leaf (isLeaf, :118) and carries word_id >= 0 and its weight; desc = 32 B per node. */
typedef struct ygzorb_vocabulary ygzorb_vocabulary;
int ygzorb_vocabulary_create(int n_nodes, int k, int L, const int32_t *child_start, const int32_t *child_list,
                             const uint8_t *desc, const int32_t *word_id, const double *weight, int device,
                             ygzorb_vocabulary **out);
/* The reference's two file formats, read like TemplatedVocabulary::loadFromBinaryFile (:1478-1523; including the extra
 * pass of its read loop, which appends the last node a second time) and ::loadFromTextFile (:1362-1448; ORBvoc.txt —
 * blank lines are skipped, where the reference reads an uninitialised parent id from its trailing empty line).
 * src/System.cc:183-190 picks one by file extension. */
int ygzorb_vocabulary_load_binary(const char *path, int device, ygzorb_vocabulary **out);
int ygzorb_vocabulary_load_text(const char *path, int device, ygzorb_vocabulary **out);
void ygzorb_vocabulary_destroy(ygzorb_vocabulary *voc);
/* info[0..5] = n_nodes, k, L, n_words, n_children (length of child_list), scoring*16 + weighting */
int ygzorb_vocabulary_info(const ygzorb_vocabulary *voc, int32_t *info6);
/* Host copies of the flat arrays (any pointer may be NULL). */
int ygzorb_vocabulary_arrays(const ygzorb_vocabulary *voc, int32_t *child_start, int32_t *child_list, uint8_t *desc,
                             int32_t *word_id, double *weight);

/* TemplatedVocabulary::transform(features, BowVector&, FeatureVector&, levelsup) (:1149-1218, per-feature descent
 * :1242-1282) for TF_IDF weighting and L1 scoring (the ORB vocabulary), batched over frames: frame f has n_feat[f]
 * descriptors starting at desc + f*stride*32, as Frame::ComputeBoW passes them (src/Frame.cc:495-500).  stride <= 8192.
 * Per feature (may be NULL): word[f*stride+i], node[f*stride+i] (the ancestor `levelsup` levels above the leaf level,
 * 0 = root when L - levelsup <= 0), weight[f*stride+i].
 * BowVector of frame f: bow_count[f] pairs (bow_word, bow_value)[f*stride + j] in ascending WordId — the std::map's
 * iteration order — with the reference's doubles bit for bit (sequential += in feature order, sequential L1 norm).
 * FeatureVector of frame f as CSR: fv_count[f] nodes, ids fv_node[f*stride + j] ascending, features of node j =
 * fv_feat[f*stride + fv_start[f*(stride+1)+j] .. fv_start[f*(stride+1)+j+1]) in insertion order — the layout
 * ygzorb_keyframe_features takes (node_ids, node_start, node_feats). */
int ygzorb_bow_transform(ygzorb_matcher *m, const ygzorb_vocabulary *voc, const uint8_t *desc, const int32_t *n_feat,
                         int n_frames, int stride, int levelsup, unsigned flags, int32_t *word, int32_t *node, double *weight,
                         int32_t *bow_count, int32_t *bow_word, double *bow_value, int32_t *fv_count, uint32_t *fv_node,
                         int32_t *fv_start, uint32_t *fv_feat);

/* ---- descriptor post-processing and the pyramid consumers of the direct tracker ---- */

/* MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:211-271), batched over n_points MapPoints (the reference calls it
 * point by point from LocalMapping::ProcessNewKeyFrame / CreateNewMapPoints / SearchInNeighbors and Tracking).  The
 * observation descriptors of point p are rows obs_start[p] .. obs_start[p+1]-1 of obs_desc (32 B each): mObservations in
 * map order, bad keyframes skipped (:226-232).  best_idx[p] = the row (relative to obs_start[p]) with the least median
 * Hamming distance to the others — element (int)(0.5*(N-1)) of the sorted row, first row wins ties (:254-264) — or -1 when
 * the point has no observation (the reference returns without touching mDescriptor); best_median[p] = that median
 * (INT_MAX for -1); best_desc + 32*p = the chosen row (mDescriptor; untouched for -1).  obs_start has n_points + 1 entries. */
int ygzorb_distinctive_descriptors(ygzorb_matcher *m, const uint8_t *obs_desc, const int32_t *obs_start, int n_points,
                                   unsigned flags, int32_t *best_idx, int32_t *best_median, uint8_t *best_desc);

/* A device-resident copy of one frame's image pyramid (Frame::mvImagePyramid / KeyFrame::mvImagePyramid, src/Frame.cc:810-813,
 * src/KeyFrame.cc:259-261: continuous 8U images, step == cols) for the direct-tracking entry points below.
 *   ygzorb_pyramid_create: from n_levels images (host pointers, or device pointers with YGZORB_IN_DEVICE), any pitch.
 *   ygzorb_pyramid_from_extractor: snapshot of frame `frame` of the extractor's last batch, device to device — the pyramid
 *     never visits the host (what a keyframe keeps when the tracker promotes the current frame).
 * n_levels <= 16.  The object is immutable after creation and may be read by several matchers concurrently. */
typedef struct ygzorb_pyramid ygzorb_pyramid;
int ygzorb_pyramid_create(int device, const uint8_t *const *levels, const int32_t *w, const int32_t *h, const int32_t *pitch,
                          int n_levels, unsigned flags, ygzorb_pyramid **out);
int ygzorb_pyramid_from_extractor(ygzorb_extractor *ex, int frame, ygzorb_pyramid **out);
void ygzorb_pyramid_destroy(ygzorb_pyramid *p);
/* *n_levels, and w[l], h[l] for every level (w / h may be NULL) */
int ygzorb_pyramid_info(const ygzorb_pyramid *p, int32_t *n_levels, int32_t *w, int32_t *h);
/* continuous host copy of level `level` (w*h bytes) */
int ygzorb_pyramid_get_level(const ygzorb_pyramid *p, int level, uint8_t *out);

/* What ORBmatcher::FindDirectProjection reads of a reference KeyFrame.  Poses are Sophus::SE3f as stored: the unit
 * quaternion's coefficients x, y, z, w, then the translation.  t_cur_ref = curr->mTcw * ref->GetPose().inverse()
 * (src/ORBmatcher.cc:1581-1582), composed by the caller once per keyframe. */
typedef struct ygzorb_direct_keyframe {
    const ygzorb_pyramid *pyramid;   /* ref->mvImagePyramid */
    float pose_ref[7];               /* ref->GetPose() */
    float t_cur_ref[7];              /* TCR */
    float fx, fy, cx, cy;            /* ref intrinsics (KeyFrame::Pixel2Camera, include/KeyFrame.h:146-152) */
} ygzorb_direct_keyframe;

/* ORBmatcher::FindDirectProjection(KeyFrame *ref, Frame *curr, MapPoint *mp, Vector2f &px_curr, int &search_level)
 * (src/ORBmatcher.cc:1573-1602), batched over n (MapPoint, reference keyframe) pairs as Tracking::SearchLocalPointsDirect
 * issues them (src/Tracking.cc:2284-2305, 2366-2385): GetWarpAffineMatrix (:1525-1547), GetBestSearchLevel
 * (include/ORBmatcher.h:226-238), WarpAffine of a 10x10 patch from ref level `octave` (:1549-1571), ygz::Align2D with 10
 * iterations on level search_level of `cur` (src/Align.cc:8-105).
 * Pair i: keyframe kfs[kf_index[i]]; px_ref[2i..] = ref->mvKeys[idx].pt, octave[i] = its octave; world_pos[3i..] =
 * mp->GetWorldPos(); px_curr[2i..] in = (mp->mTrackProjX, mp->mTrackProjY), out = the aligned pixel (level-0 coordinates);
 * search_level[i]; success[i] = the bool the reference returns.  cur_intrinsics = curr fx, fy, cx, cy; scale_factors /
 * inv_scale_factors = mvScaleFactors / mvInvScaleFactors (n_levels of `cur`); inv_level_sigma2_1 = ref->mvInvLevelSigma2[1].
 * Float results follow the reference's evaluation order (Eigen 3.3 fixed-size expressions, uncontracted) and equal the CPU
 * restatement bit for bit. */
int ygzorb_find_direct_projection(ygzorb_matcher *m, const ygzorb_direct_keyframe *kfs, int n_kfs, const ygzorb_pyramid *cur,
                                  const float *cur_intrinsics, const float *scale_factors, const float *inv_scale_factors,
                                  float inv_level_sigma2_1, int n, const int32_t *kf_index, const float *px_ref,
                                  const int32_t *octave, const float *world_pos, unsigned flags, float *px_curr,
                                  int32_t *search_level, uint8_t *success);

/* SparseImgAlign (include/SparseImageAlign.h, src/SparseImageAlign.cc): sparse direct alignment of the current frame against
 * the reference frame's features, 4x4 patches, inverse compositional, as Tracking::TrackWithSparseAlignment runs it
 * (src/Tracking.cc:284, :2171: max_level = nLevels - 1, min_level = 1, Gauss-Newton, 10 iterations per level, unit weights).
 * A handle owns the class's caches on the device (ref_patch_cache_, jacobian_cache_, visible_fts_) and runs on the matcher's
 * stream; destroy it before the matcher it was created on.  Poses are Sophus::SE3f storage (quaternion x, y, z, w, translation).
 *   ygzorb_sparse_align_begin: what run() sets up (:20-37) — the reference frame's N features: kp_xy[2i..] = mvKeys[i].pt,
 *     valid[i] = mvpMapPoints[i] && !isBad() && !mvbOutlier[i] (:68-70), world_pos[3i..] = GetWorldPos(), ref_pose =
 *     ref_frame->mTcw, focal = ref fx, inv_scale_factors = mvInvScaleFactors; clears visible_fts_.
 *   ygzorb_sparse_align_run: the whole coarse-to-fine optimisation in ONE kernel launch: t_cur_ref in = cur->mTcw *
 *     ref->mTcw.inverse(), out = TCR; returns through *n_tracked the value run() returns (n_meas_ / patch_area_);
 *     stats (may be NULL): [0] Gauss-Newton iterations done, [1] final chi2_.
 *   ygzorb_sparse_align_precompute / _residuals: the two hot loops alone, for a caller that keeps the reference's own
 *     NLLSSolver<6, SE3f> driver: precomputeReferencePatches() at `level` (:57-128; also zeroes jacobian_cache_ like :42) and
 *     computeResiduals(model, true) at that level (:130-225): H (6x6 row-major), Jres (6), chi2 SUM (the caller divides by
 *     n_meas like :224), n_meas.
 * Per feature the arithmetic is the reference's in its order; across features the sums are reduced by a fixed tree in
 * double, so H / Jres / chi2 equal the reference's sequential float sums to rounding (rel. ~1e-6), not bit for bit. */
typedef struct ygzorb_sparse_align ygzorb_sparse_align;
int ygzorb_sparse_align_create(ygzorb_matcher *m, ygzorb_sparse_align **out);
void ygzorb_sparse_align_destroy(ygzorb_sparse_align *sa);
int ygzorb_sparse_align_begin(ygzorb_sparse_align *sa, const ygzorb_pyramid *ref, const ygzorb_pyramid *cur, int n,
                              const float *kp_xy, const uint8_t *valid, const float *world_pos, const float *ref_pose,
                              float focal, const float *cur_intrinsics, const float *inv_scale_factors);
int ygzorb_sparse_align_run(ygzorb_sparse_align *sa, int max_level, int min_level, float *t_cur_ref, int64_t *n_tracked,
                            float *stats);
int ygzorb_sparse_align_precompute(ygzorb_sparse_align *sa, int level);
int ygzorb_sparse_align_residuals(ygzorb_sparse_align *sa, int level, const float *t_cur_ref, float *h36, float *jres6,
                                  float *chi2_sum, int64_t *n_meas);
/* host copies of the caches after precompute (any may be NULL): visible[n], ref_patch[n*16], jacobian[n*16*6] */
int ygzorb_sparse_align_caches(ygzorb_sparse_align *sa, uint8_t *visible, float *ref_patch, float *jacobian);

/* ---- multi-GPU: the two workloads that shard (BASELINE.json configs[3], [4]) ------------------------------- */

/* Frames of a batched extraction and query rows of a brute-force kNN are independent: each rank takes the contiguous range
 * ygzorb_shard_range gives it (sizes differ by at most one) and runs the single-GPU path; the only communication is the gather
 * of the results on rank 0 over NCCL.  No reference counterpart (the reference is a single-process CPU system); the interface is
 * SURVEY.md 8(b)'s.  NCCL is bound at run time (dlopen): these entry points fail with YGZORB_E_STATE where libnccl is absent,
 * everything else works without it. */
typedef struct ygzorb_comm ygzorb_comm;
int ygzorb_shard_range(long long n, int rank, int world, long long *begin, long long *end);
/* One process per GPU: rank 0 calls ygzorb_comm_unique_id (ncclGetUniqueId, 128 bytes), the host code hands the bytes to the
 * other ranks by any means (bench.py: torch.distributed broadcast), every rank calls ygzorb_comm_create_rank (ncclCommInitRank).
 * world == 1 gives a handle without NCCL (the sharded calls then reduce to the single-GPU path + the packing). */
int ygzorb_comm_unique_id(uint8_t *id128);
int ygzorb_comm_create_rank(const uint8_t *id128, int rank, int world, int device, ygzorb_comm **out);
/* One process driving n_gpus devices (ncclCommInitAll over devices 0..n_gpus-1): fills out[0..n_gpus); use one host thread per
 * handle for the sharded calls (they synchronise with the device between chunks). */
int ygzorb_comm_create(int n_gpus, ygzorb_comm **out);
void ygzorb_comm_destroy(ygzorb_comm *comm);
int ygzorb_comm_info(const ygzorb_comm *comm, int *rank, int *world, int *device);
long long ygzorb_comm_launch_count(const ygzorb_comm *comm);

#define YGZORB_BROADCAST_TRAIN 16u   /* ygzorb_hamming_knn2_sharded: replicate the train rows from rank 0 first (ncclBroadcast) */

/* ygzorb_hamming_knn2 (the best / second-best + ratio rule of src/ORBmatcher.cc:193-217 over all pairs), query-sharded.
 * query_shard: this rank's rows [begin, end) of the nq_total queries; train: nt rows on every rank (with
 * YGZORB_BROADCAST_TRAIN only rank 0's content matters).  best_idx / d1 / d2 / match: full-size arrays (nq_total) on every rank;
 * a rank fills its own slice, rank 0 ends up with all of them (one grouped ncclSend / ncclRecv on the matcher's stream).
 * All pointers are device memory.  YGZORB_NO_SYNC: return after enqueueing (events on the matcher's stream then include
 * the gather). */
int ygzorb_hamming_knn2_sharded(ygzorb_comm *comm, ygzorb_matcher *m, const uint8_t *query_shard, long long nq_total, uint8_t *train,
                                int nt, float ratio, int th, unsigned flags, int32_t *best_idx, int32_t *d1, int32_t *d2,
                                int32_t *match);

/* ygzorb_extract_batch, frame-sharded.  frames_shard: this rank's frames [begin, end) of the n_total (host or device per
 * YGZORB_IN_DEVICE).  The shard is extracted in chunks of chunk_frames (<= the extractor's max_batch); each chunk's records are
 * packed back to back on the device and gathered on rank 0 — counts first, then the packed keypoint records and descriptor rows
 * of exactly the keypoints found — on the communicator's own stream while the next chunk extracts.  Outputs on rank 0 (device
 * memory; ignored elsewhere): counts[n_total]; keypoints / desc with room for n_total * cap records: the records of rank r's
 * frames lie back to back from record r_begin * cap on; frame_offset[f] = record index of frame f's first keypoint.
 * The extractor's stream waits for the gather, so events recorded on it after the call include it. */
int ygzorb_extract_batch_sharded(ygzorb_comm *comm, ygzorb_extractor *ex, const uint8_t *frames_shard, long long n_total, int w, int h,
                                 int stride, size_t frame_stride, unsigned flags, int cap, int chunk_frames, int32_t *counts,
                                 int64_t *frame_offset, ygzorb_keypoint *keypoints, uint8_t *desc);

#ifdef __cplusplus
}
#endif
#endif /* YGZORB_H_ */
