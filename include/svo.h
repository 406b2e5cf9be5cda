/*
 * svo.h -- flat C ABI of the B200-native stereo-VO front end.
 *
 * This is the drop-in boundary: everything the reference's `include/vo.h`
 * (namespace visual_odometry) exposes for the per-frame hot path is reachable
 * through these entry points; `include/vo.h` in THIS repo is a thin C++
 * wrapper with the reference's class / function signatures over this ABI.
 *
 * Conventions
 *   - plain pointers + sizes, no C++ / torch types, no exceptions.
 *   - every function returns an int: >= 0 on success (meaning documented per
 *     call), < 0 is an `svo_status` error.  There is NO CPU fallback: without
 *     a CUDA device every compute call returns SVO_ERR_NO_DEVICE.
 *   - all pointers are HOST pointers unless the name ends in `_dev`.
 *   - images are 8-bit single channel (or interleaved 3-channel when svo_params.channels == 3; the stage-level
 *     entry points other than svo_lk_track are single channel), row stride in bytes given explicitly.
 *   - points are interleaved float32 (x, y); 3-D points interleaved (x, y, z).
 *
 * Reference citations (file:line) are relative to Alex7Li/stereo_visual_odometry.
 */
#ifndef SVO_H_
#define SVO_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum svo_status {
    SVO_OK = 0,
    SVO_ERR_INVALID = -1,   /* bad argument                                  */
    SVO_ERR_NO_DEVICE = -2, /* no CUDA device / driver                       */
    SVO_ERR_CUDA = -3,      /* a CUDA runtime call failed (svo_last_error)   */
    SVO_ERR_CAPACITY = -4,  /* caller-provided output buffer too small       */
    SVO_ERR_STATE = -5      /* call order violated (e.g. no projection set)  */
} svo_status;

/* Compile-time constants of the reference, include/vo.h:53-127, 251-252 and
 * the literals at src/vo.cpp:183-184, 295.  svo_default_params() fills these. */
typedef struct svo_params {
    int width, height;            /* image size (all frames of a context)        */
    int n_seq;                    /* independent sequences batched in this ctx   */
    int buckets_along_height;     /* 92   vo.h BUCKETS_ALONG_HEIGHT              */
    int buckets_along_width;      /* 160  vo.h BUCKETS_ALONG_WIDTH               */
    int bucket_start_row;         /* 4    vo.h BUCKET_START_ROW                  */
    int features_per_bucket;      /* 1    vo.h FEATURES_PER_BUCKET (ctx path: 1) */
    int features_threshold;       /* 15   vo.h FEATURES_THRESHOLD                */
    int pre_matching_threshold;   /* 100  vo.h PRE_MATCHING_FEATURE_THRESHOLD    */
    int age_threshold;            /* 20   vo.h AGE_THRESHOLD                     */
    int fast_threshold;           /* 20   vo.h FAST_THRESHOLD (retry uses /4)    */
    int win_size;                 /* 10   vo.h:251 winSize (square)              */
    int max_level;                /* 3    vo.h:252 maxLevel                      */
    int lk_max_iters;             /* 30   vo.cpp:183 TermCriteria COUNT          */
    float lk_epsilon;             /* 1e-4 vo.cpp:184 TermCriteria EPS            */
    float lk_min_eig;             /* 1e-3 vo.h OPTICAL_FLOW_MIN_EIG_THRESHOLD    */
    float circular_threshold;     /* .15  vo.h CIRCULAR_MATCHING_SUCCESS_THRESHOLD */
    float ransac_reproj_error;    /* 8    vo.h RANSAC_REPROJECTION_ERROR         */
    int ransac_iterations;        /* 100  vo.h RANSAC_ITERATIONS (reference's sequential cap) */
    int ransac_hypotheses;        /* hypotheses scored in parallel per frame (default 100 = the cap above) */
    float ransac_confidence;      /* .98  vo.cpp:295 (kept for the record)       */
    double max_translation_norm;  /* .1   vo.h MAX_TRANSLATION_NORM              */
    double max_rotation_norm;     /* .5   vo.h MAX_ROTATION_NORM                 */
    uint64_t ransac_seed;         /* reserved (ABI layout): the sample stream is OpenCV's fixed RNG((uint64)-1) replayed */
    int device;                   /* CUDA device ordinal                         */
    int channels;                 /* 1 = 8UC1 (default); 3 = interleaved 8UC3 exactly as the reference CLI feeds
                                     cv::imread's BGR images (src/main.cpp:37-46): FAST scans the first `width`
                                     bytes of every row, pyramids and LK run on all three channels        */
    int lookahead;                /* K > 1: svo_submit_frames may hand over up to K frames per call; the context then
                                     keeps 2 K + 3 pyramid ring slots instead of 3 (single sequence, 8UC1)        */
} svo_params;

/* Per-frame result of the pipeline (one per sequence). */
typedef struct svo_frame_result {
    int success;          /* stereo_callback().first, vo.cpp:41-137            */
    int reason;           /* svo_fail_reason                                    */
    int n_bucketed;       /* features entering circular matching                */
    int n_matched;        /* survivors of circular matching + bounds            */
    int n_inliers;        /* RANSAC inliers (0 if RANSAC did not run)           */
    int retry;            /* 1 if the low-threshold FAST retry fired (vo.cpp:327-332) */
    int pnp_ok;           /* solvePnPRansac-equivalent success                  */
    int lk_iters;         /* total LK iterations over all features/legs/levels  */
    double R[9];          /* rotation state after this frame (row major)        */
    double t[3];          /* translation state after this frame                 */
    double transform[16]; /* stereo_callback().second (row-major 4x4)           */
} svo_frame_result;

typedef enum svo_fail_reason {
    SVO_REASON_OK = 0,
    SVO_REASON_FIRST_FRAME = 1,   /* vo.cpp:47-56   */
    SVO_REASON_FEW_FEATURES = 2,  /* vo.cpp:82-84   */
    SVO_REASON_PNP_FAILED = 3,    /* vo.cpp:106-113 */
    SVO_REASON_TOO_BIG = 4        /* vo.cpp:123-132 */
} svo_fail_reason;

typedef struct svo_ctx svo_ctx;

/* ---- library -------------------------------------------------------------- */
const char* svo_version(void);
const char* svo_last_error(void);               /* thread-local message of the last failure */
int svo_device_count(void);                     /* >=0, or SVO_ERR_NO_DEVICE                */
void svo_default_params(svo_params* p, int width, int height);

/* ---- pipeline context: replaces class VisualOdometry (vo.h:231-380) -------- */
svo_ctx* svo_create(const svo_params* p);       /* NULL on failure (svo_last_error)         */
void svo_destroy(svo_ctx* ctx);
/* VisualOdometry::initalize_projection_matricies, vo.h:307-309 / vo.cpp:8-26.
 * Pl, Pr: 3x4 row-major float32.  Applies to every sequence of the context. */
int svo_set_projection(svo_ctx* ctx, const float Pl[12], const float Pr[12]);
/* VisualOdometry::stereo_callback, vo.h:333-334 / vo.cpp:41-137, for sequence 0
 * of the context.  HOST images; H2D copy, all kernels and the D2H of the result
 * happen inside the call; synchronous.  Returns 1 (success) / 0 (soft failure:
 * T_out = last successful transform) / <0 error.  T_out: row-major 4x4. */
int svo_process_frame(svo_ctx* ctx, const uint8_t* left, const uint8_t* right,
                      int stride, double T_out[16]);
/* Same for all n_seq sequences at once: images are n_seq consecutive planes
 * (plane pitch = seq_stride bytes).  results[n_seq].  Returns #successes. */
int svo_process_frame_batch(svo_ctx* ctx, const uint8_t* left, const uint8_t* right,
                            int stride, long long seq_stride, svo_frame_result* results);
/* Device-resident variant: images already in HBM (same layout), result struct
 * array on the device; nothing is copied and nothing synchronises (stream order).
 * results_dev may be NULL. */
int svo_process_frame_batch_dev(svo_ctx* ctx, const uint8_t* left_dev, const uint8_t* right_dev,
                                int stride, long long seq_stride, svo_frame_result* results_dev);
/* Streaming form of svo_process_frame[_batch] for a camera that delivers frames faster than one
 * frame's latency: svo_submit_frame enqueues the H2D copy (left/right should be pinned host memory and
 * stay valid until the frame is collected), the ingest kernels and the tracking chain and returns at
 * once; svo_collect blocks until the OLDEST uncollected frame is done and returns its result exactly as
 * svo_process_frame_batch would (frames are processed strictly in submission order, results are
 * identical to the synchronous calls).  The ingest of frame f + 1 (its own stream) overlaps the
 * tracking chain of frame f.  At most 32 frames may be pending (SVO_ERR_STATE beyond that). */
int svo_submit_frame(svo_ctx* ctx, const uint8_t* left, const uint8_t* right, int stride);
int svo_submit_frame_batch(svo_ctx* ctx, const uint8_t* left, const uint8_t* right, int stride,
                           long long seq_stride);
/* Look-ahead form for ONE trajectory whose next K frames are already there (a recording, a camera buffer): the
 * state-independent stages -- the pyramids of both images and FAST on the left one, src/vo.cpp:200-201, 325 -- run
 * for all K frames in ONE launch set (the batched kernels, frame index in place of the sequence index), then the K
 * frames go through the tracking chain one after the other exactly like K svo_submit_frame calls; results are
 * identical and are collected with svo_collect in order.  Needs svo_params.lookahead >= K, n_seq == 1, 8UC1.
 * lefts / rights: K host pointers (pinned memory recommended), all frames with the same stride. */
int svo_submit_frames(svo_ctx* ctx, int n_frames, const uint8_t* const* lefts, const uint8_t* const* rights, int stride);
int svo_collect(svo_ctx* ctx, svo_frame_result* results /* [n_seq], may be NULL */);
int svo_pending(svo_ctx* ctx);                  /* submitted, not yet collected */
/* Makes the context's primary stream (svo_stream) wait for everything enqueued so far on the context's
 * other streams, so that an event recorded on svo_stream afterwards covers all submitted frames. */
int svo_stream_join(svo_ctx* ctx);
/* Pipeline introspection (synchronises): out = { speculative tracking on, SMs of the chain / ingest / KLT
 * partitions (0 = whole GPU, no partitioning), features of sequence 0 the last frame had to track outside the
 * speculative cache, persistent KLT blocks per sequence (chain side / KLT partition), FAST-winner candidates of the
 * last speculative pass: tracked | skipped << 16 }. */
int svo_pipeline_info(svo_ctx* ctx, int out[8]);
int svo_synchronize(svo_ctx* ctx);
/* Last frame's result for sequence `seq` (host copy; synchronises). */
int svo_get_result(svo_ctx* ctx, int seq, svo_frame_result* out);
/* Stage taps for parity tests: copies of the last frame's intermediate arrays
 * of sequence `seq`.  Each pointer may be NULL.  cap = capacity in features.
 *   which = 0: feature set entering circular matching (x,y / age / strength)
 *   which = 1: LK outputs pL1,pR1,pR0,pC (4 x n x 2 float) + status (4 x n u8)
 *   which = 2: matched sets pL0,pL1,pR0,pR1 (4 x n x 2) and X3 (n x 3)
 *   which = 3: feature set after the frame (x,y / age / strength)              */
int svo_get_tap(svo_ctx* ctx, int seq, int which, int cap, float* f_out, int* i_out,
                uint8_t* u8_out);
/* The images the NEXT frame will be matched against (level 0 of the cached t0 pyramids of sequence `seq`,
 * lastLeftPyramid / lastRightPyramid of vo.h:257-258, including the "kept after an empty frame" case of
 * vo.cpp:179-181), copied to host buffers of `stride` bytes per row (>= width * channels).  Synchronises.
 * Lets a wrapper serve VisualOdometry::circularMatching (vo.h:374-379) without keeping host copies per frame. */
int svo_get_frame_images(svo_ctx* ctx, int seq, uint8_t* left_out, uint8_t* right_out, int stride);
/* The CUDA stream the context launches on (cudaStream_t as void*). */
void* svo_stream(svo_ctx* ctx);
/* Number of kernel launches issued by the context so far. */
long long svo_launch_count(svo_ctx* ctx);
/* Accumulated CUDA-event time (ms) and launch count of one stage since the last
 * svo_reset_timers(); only measured when svo_enable_timers(ctx,1) was called.
 * stage: 0 ingest(FAST+pyramid) 1 bucket 2 klt 3 select+triangulate 4 pnp hypotheses 5 pnp refit */
int svo_enable_timers(svo_ctx* ctx, int on);
int svo_get_stage_time(svo_ctx* ctx, int stage, double* ms_out, long long* launches_out);
int svo_reset_timers(svo_ctx* ctx);

/* ---- stage-level entry points (host pointers; synchronous) ------------------
 * featureDetectionFast, vo.h:393-395 / feature_set.cpp:55-68 == cv::FAST(img,
 * kps, threshold, nonmaxSuppression=true), TYPE_9_16.  Row-major keypoint order.
 * xy_out[2*cap], resp_out[cap].  Returns the number of keypoints found (may
 * exceed cap, in which case only cap are written and SVO_ERR_CAPACITY is NOT
 * raised: the count tells the caller). */
int svo_fast_detect(const uint8_t* img, int width, int height, int stride, int threshold,
                    float* xy_out, float* resp_out, int cap);
/* FeatureSet::filterByBucketLocationInternal, vo.h:167-174 / feature_set.cpp:95-147
 * (+ Bucket::add_feature, feature_set.cpp:20-53).  In-place on the three arrays;
 * returns the new count. */
int svo_bucket_filter(float* xy, int* ages, int* strengths, int n, int image_height,
                      int image_width, int buckets_along_height, int buckets_along_width,
                      int bucket_start_row, int features_per_bucket, int age_threshold,
                      int fast_threshold);
/* cv::buildOpticalFlowPyramid image levels (vo.cpp:50-53, 200-201): level l+1 =
 * pyrDown(level l).  out receives levels 1..max_level packed back to back
 * (level l is ceil(w/2^l) x ceil(h/2^l), dense rows).  Returns #levels built
 * (incl. level 0; stops early like OpenCV when a level would be <= win). */
int svo_build_pyramid(const uint8_t* img, int width, int height, int stride, int max_level,
                      int win_size, uint8_t* out, long long out_cap);
/* One cv::calcOpticalFlowPyrLK call as at vo.cpp:203-215 (winSize/maxLevel/
 * criteria/minEig from p; flags = 0).  pts_in/pts_out: n x 2; status_out: n. */
int svo_lk_track(const svo_params* p, const uint8_t* prev_img, const uint8_t* next_img,
                 int stride, const float* pts_in, int n, float* pts_out, uint8_t* status_out);
/* VisualOdometry::circularMatching, vo.h:374-379 / vo.cpp:169-240, stateless
 * form: the four images are given explicitly.  pL0 (n x 2) in; the five arrays
 * are compacted in place / written; keep_out[n] (optional) is the is_ok mask
 * over the inputs.  Returns the survivor count. */
int svo_circular_match(const svo_params* p, const uint8_t* L0, const uint8_t* R0,
                       const uint8_t* L1, const uint8_t* R1, int stride, float* pL0, int n,
                       float* pR0, float* pL1, float* pR1, uint8_t* keep_out);
/* VisualOdometry::matchingFeatures, vo.h:354-362 / vo.cpp:315-366, stateless form:
 * FAST on L0 (+ the low-threshold retry) appended to the feature set, bucketing,
 * circular matching, in-bounds filter, compaction.  fs_* (capacity fs_cap records,
 * *n_fs valid on entry) is the FeatureSet in/out (ages are NOT incremented, that is
 * stereo_callback's job); the four point arrays need fs_cap capacity.  Returns the
 * matched count (== *n_fs on exit). */
int svo_matching_features(const svo_params* p, const uint8_t* L0, const uint8_t* R0,
                          const uint8_t* L1, const uint8_t* R1, int stride, float* fs_xy,
                          int* fs_age, int* fs_str, int* n_fs, int fs_cap, float* pL0,
                          float* pR0, float* pL1, float* pR1);
/* findClosePoints, vo.h:430-432 / vo.cpp:265-280. ok_out[n] in {0,1}. */
int svo_find_close_points(const float* p1, const float* p2, int n, float threshold,
                          uint8_t* ok_out);
/* deletePointsWithFailureStatus / deleteFeaturesWithFailureStatus, vo.h:406-417 /
 * vo.cpp:144-168: stable stream compaction of `n` records of `elem_bytes` each. */
int svo_compact(void* data, int elem_bytes, int n, const uint8_t* keep);
/* cv::triangulatePoints + convertPointsFromHomogeneous as at vo.cpp:89-94. */
int svo_triangulate(const float Pl[12], const float Pr[12], const float* pL, const float* pR,
                    int n, float* X3_out);
/* cameraToWorld, vo.h:452-456 / vo.cpp:282-313 (solvePnPRansac, extrinsic guess,
 * 8 px, refit on inliers).  R (3x3 row-major) and t (3) are in/out exactly like
 * the reference (R updated on success only; t restored on failure).
 * inliers_out[n] ascending indices.  Returns #inliers; *success_out = bool. */
int svo_camera_to_world(const svo_params* p, const float K[9], const float* cam_pts,
                        const float* world_pts, int n, double R[9], double t[3],
                        int* inliers_out, int* success_out);
/* getInverseTransform, vo.h:469-470 / vo.cpp:246-258 (host, FP64). */
int svo_get_inverse_transform(const double R[9], const double t[3], double T_out[16]);

/* ---- allocation-free entry point ---------------------------------------------
 * No counterpart in the reference: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  Same result as svo_fast_detect, but every pointer
 * is a DEVICE pointer and the call allocates nothing: the response map and the row
 * counts of the ordered compaction live in one static device array of
 * SVO_NOALLOC_SCRATCH_BYTES that every GPU loading the library holds until exit.
 *   - img_dev: width x height bytes, `pitch` bytes per row (a 16-byte aligned base and
 *     pitch let the kernel use TMA tile loads; any pitch >= width works);
 *   - xy_out_dev[2*cap], resp_out_dev[cap], n_out_dev[1] (the keypoint count, which may
 *     exceed cap exactly like svo_fast_detect's return value);
 *   - stream: a cudaStream_t (NULL = the default stream).  The launches are enqueued
 *     and the call returns; it does not synchronise.
 * Limit: ((width+127)&~127) * height, rounded up to 256, + 4 * (height+1) bytes must fit
 * in SVO_NOALLOC_SCRATCH_BYTES (e.g. 2048 x 2040); larger images are rejected with
 * SVO_ERR_CAPACITY and a message naming both numbers.
 * Every call on one device shares the array: calls that overlap on the GPU (different
 * streams, or graphs replayed concurrently) race; serialise them.  The array keeps the
 * last call's contents, and every byte a call reads it has written first.  The first
 * call on each device looks the array's device address up (cudaGetSymbolAddress): make
 * that call outside stream capture. */
#define SVO_NOALLOC_SCRATCH_BYTES (4 << 20)
int svo_fast_detect_noalloc(const uint8_t* img_dev, int width, int height, int pitch, int threshold,
                            float* xy_out_dev, float* resp_out_dev, int cap, int* n_out_dev,
                            void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SVO_H_ */
