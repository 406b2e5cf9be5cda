/*
 * vo.h -- drop-in replacement for the reference's include/vo.h (namespace visual_odometry).
 *
 * Same classes, member and free-function signatures, constants and failure behaviour as
 * Alex7Li/stereo_visual_odometry include/vo.h; every numerical stage runs as sm_100a CUDA kernels
 * through the C ABI in svo.h (libsvo_b200.so).  Header-only: link with -lsvo_b200.
 * There is no CPU fallback: without a CUDA device the calls throw std::runtime_error.
 *
 * Reference line numbers (vo.h:NNN, vo.cpp:NNN, feature_set.cpp:NNN) cite the reference repo.
 * If OpenCV C++ is installed the real cv:: types are used, otherwise svo_cv_shim.h supplies the
 * few types that appear in the signatures.
 */
#ifndef SVO_DROPIN_VO_H_
#define SVO_DROPIN_VO_H_

#define CFE_ES_WriteToSysLog printf   /* vo.h:17 */
#define float32 float                 /* vo.h:18 */
#define float64 double                /* vo.h:19 */

#include <math.h>
#include <stdio.h>

#include <algorithm>
#include <ctime>
#include <fstream>
#include <iostream>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#if defined(__has_include)
#if __has_include(<opencv2/core.hpp>) && !defined(SVO_FORCE_CV_SHIM)
#include <opencv2/core.hpp>
#define SVO_HAVE_OPENCV 1
#endif
#endif
#ifndef SVO_HAVE_OPENCV
#include "svo_cv_shim.h"
#endif

#include "svo.h"

namespace visual_odometry {

/* vo.h:53-127 */
const int BUCKET_START_ROW = 4;
const int BUCKETS_ALONG_HEIGHT = 92;
const int BUCKETS_ALONG_WIDTH = 160;
const int FEATURES_PER_BUCKET = 1;
const int FEATURES_THRESHOLD = 15;
const int PRE_MATCHING_FEATURE_THRESHOLD = 100;
const int AGE_THRESHOLD = 20;
const int FAST_THRESHOLD = 20;
const float32 RANSAC_REPROJECTION_ERROR = 8;
const int RANSAC_ITERATIONS = 100;
const float64 OPTICAL_FLOW_MIN_EIG_THRESHOLD = 0.001;
const float64 CIRCULAR_MATCHING_SUCCESS_THRESHOLD = .15;
const float64 MAX_TRANSLATION_NORM = .1;
const float64 MAX_ROTATION_NORM = .5;

namespace detail {
inline int check(int rc) {
    if (rc < 0) throw std::runtime_error(std::string("svo: ") + svo_last_error());
    return rc;
}
/* 8-bit view of a cv::Mat: 8UC1 (the contract) or 8UC3 exactly as the reference's CLI passes cv::imread's BGR
 * images through (main.cpp:37-46): FAST then scans the first `cols` bytes of every row, the pyramids and
 * LK use all three channels -- svo_params.channels = 3 */
inline const uint8_t* gray(const cv::Mat& m, int& stride) {
    if (m.empty() || (m.type() != CV_8UC1 && m.type() != CV_8UC3))
        throw std::runtime_error("svo: images must be CV_8UC1 (or CV_8UC3, CLI-faithful mode)");
    stride = (int)m.step;
    return m.data;
}
inline svo_params params_for(const cv::Mat& img) {
    svo_params p;
    svo_default_params(&p, img.cols, img.rows);
    p.channels = img.channels();
    return p;
}
}  // namespace detail

/* ---- vo.h:132-189 ---------------------------------------------------------------------- */
class FeatureSet {
   public:
    std::vector<cv::Point2f> points;
    std::vector<int> ages;
    std::vector<int> strengths;

    int size() { return (int)points.size(); }

    /* vo.h:167 / feature_set.cpp:148-152 */
    void filterByBucketLocation(const cv::Mat& image) {
        filterByBucketLocationInternal(image, BUCKETS_ALONG_HEIGHT, BUCKETS_ALONG_WIDTH, BUCKET_START_ROW,
                                       FEATURES_PER_BUCKET);
    }
    /* vo.h:170-174 / feature_set.cpp:95-147 */
    void filterByBucketLocationInternal(const cv::Mat& image, const int buckets_along_height,
                                        const int buckets_along_width, const int bucket_start_row,
                                        const int features_per_bucket) {
        const int n = size();
        std::vector<float> xy(2 * (size_t)n + 2);
        for (int i = 0; i < n; ++i) {
            xy[2 * i] = points[i].x;
            xy[2 * i + 1] = points[i].y;
        }
        ages.resize(n + 1);
        strengths.resize(n + 1);
        const int m = detail::check(svo_bucket_filter(xy.data(), ages.data(), strengths.data(), n, image.rows, image.cols,
                                                      buckets_along_height, buckets_along_width, bucket_start_row,
                                                      features_per_bucket, AGE_THRESHOLD, FAST_THRESHOLD));
        points.resize(m);
        ages.resize(m);
        strengths.resize(m);
        for (int i = 0; i < m; ++i) points[i] = cv::Point2f(xy[2 * i], xy[2 * i + 1]);
    }
    /* vo.h:183-184 / feature_set.cpp:75-89 */
    void appendFeaturesFromImage(const cv::Mat& image, const int fast_threshold);
    void clear() {
        points.clear();
        strengths.clear();
        ages.clear();
    }
};

/* ---- vo.h:196-229 / feature_set.cpp:4-53.  Plain container bookkeeping used by the reference's
 * tests; the per-frame bucketing itself runs on the GPU (filterByBucketLocationInternal above). */
class Bucket {
   public:
    int max_size;
    FeatureSet features;
    Bucket(int size) : max_size(size) {}
    ~Bucket() {}
    int compute_score(const int age, const int strength) { return age + (strength - FAST_THRESHOLD) / 20; }
    int size() { return (int)features.points.size(); }
    void add_feature(const cv::Point2f point, const int age, const int strength) {
        if (!max_size) return;
        if (age >= AGE_THRESHOLD) return;
        if (size() < max_size) {
            features.points.push_back(point);
            features.ages.push_back(age);
            features.strengths.push_back(strength);
            return;
        }
        int lo = 0;
        for (int i = 1; i < size(); ++i)
            if (compute_score(features.ages[i], features.strengths[i]) <
                compute_score(features.ages[lo], features.strengths[lo]))
                lo = i;
        if (compute_score(age, strength) > compute_score(features.ages[lo], features.strengths[lo])) {
            features.points[lo] = point;
            features.ages[lo] = age;
            features.strengths[lo] = strength;
        }
    }
};

/* ---- free functions, vo.h:393-470 ------------------------------------------------------- */
/* vo.h:393-395 / feature_set.cpp:55-68 */
inline std::vector<cv::Point2f> featureDetectionFast(const cv::Mat image, const int fast_threshold,
                                                     std::vector<float32>& response_strengths) {
    int stride;
    const uint8_t* px = detail::gray(image, stride);
    int cap = 4096;
    std::vector<float> xy, resp;
    int n;
    for (;;) {
        xy.resize(2 * (size_t)cap);
        resp.resize(cap);
        n = detail::check(svo_fast_detect(px, image.cols, image.rows, stride, fast_threshold, xy.data(), resp.data(), cap));
        if (n <= cap) break;
        cap = n;
    }
    std::vector<cv::Point2f> pts(n);
    response_strengths.reserve(response_strengths.size() + n);
    for (int i = 0; i < n; ++i) {
        pts[i] = cv::Point2f(xy[2 * i], xy[2 * i + 1]);
        response_strengths.push_back(resp[i]);
    }
    return pts;
}

inline void FeatureSet::appendFeaturesFromImage(const cv::Mat& image, const int fast_threshold) {
    std::vector<float> response_strength;
    std::vector<cv::Point2f> points_new = featureDetectionFast(image, fast_threshold, response_strength);
    points.insert(points.end(), points_new.begin(), points_new.end());
    ages.insert(ages.end(), points_new.size(), 0);
    for (float r : response_strength) strengths.push_back((int)r);
    filterByBucketLocation(image);
}

namespace detail {
inline std::vector<uint8_t> mask_bytes(const std::vector<bool>& ok) {
    std::vector<uint8_t> k(ok.size() + 1);
    for (size_t i = 0; i < ok.size(); ++i) k[i] = ok[i] ? 1 : 0;
    return k;
}
}  // namespace detail

/* vo.h:406-408 / vo.cpp:144-153 */
inline void deletePointsWithFailureStatus(std::vector<cv::Point2f>& point_vector, const std::vector<bool>& isok) {
    if (isok.empty()) return;
    std::vector<uint8_t> k = detail::mask_bytes(isok);
    static_assert(sizeof(cv::Point2f) == 8, "Point2f must be two packed floats");
    const int m = detail::check(svo_compact(point_vector.data(), 8, (int)isok.size(), k.data()));
    point_vector.resize(m + (point_vector.size() - isok.size()));
}
/* vo.h:415-417 / vo.cpp:154-168 */
inline void deleteFeaturesWithFailureStatus(FeatureSet& currentFeatures, const std::vector<bool>& isok) {
    if (isok.empty()) return;
    std::vector<uint8_t> k = detail::mask_bytes(isok);
    const int n = (int)isok.size();
    detail::check(svo_compact(currentFeatures.points.data(), 8, n, k.data()));
    detail::check(svo_compact(currentFeatures.ages.data(), 4, n, k.data()));
    const int m = detail::check(svo_compact(currentFeatures.strengths.data(), 4, n, k.data()));
    currentFeatures.points.resize(m);
    currentFeatures.ages.resize(m);
    currentFeatures.strengths.resize(m);
}
/* vo.h:430-432 / vo.cpp:265-280 */
inline std::vector<bool> findClosePoints(const std::vector<cv::Point2f>& points_1,
                                         const std::vector<cv::Point2f>& points_2, float32 threshold) {
    const int n = (int)points_1.size();
    std::vector<uint8_t> ok(n + 1);
    if (n) detail::check(svo_find_close_points(&points_1[0].x, &points_2[0].x, n, threshold, ok.data()));
    std::vector<bool> out(n);
    for (int i = 0; i < n; ++i) out[i] = ok[i] != 0;
    return out;
}
/* vo.h:452-456 / vo.cpp:282-313.  worldPoints: N x 1 CV_32FC3 (or N x 3 CV_32F). */
inline std::pair<cv::Mat, bool> cameraToWorld(const cv::Mat_<float32>& cameraProjection,
                                              const std::vector<cv::Point2f>& cameraPoints,
                                              const cv::Mat_<float32>& worldPoints, cv::Mat& rotation,
                                              cv::Mat& translation) {
    const int n = (int)cameraPoints.size();
    float K[9];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) K[i * 3 + j] = cameraProjection(i, j);
    std::vector<float> world(3 * (size_t)n);
    const cv::Mat& wm = worldPoints;
    for (int i = 0; i < n; ++i) std::memcpy(&world[3 * i], wm.data + (size_t)i * (wm.rows == n ? wm.step : 12), 12);
    double R[9], t[3];
    for (int i = 0; i < 3; ++i) {
        for (int j = 0; j < 3; ++j) R[i * 3 + j] = rotation.at<double>(i, j);
        t[i] = translation.at<double>(i);
    }
    svo_params p;
    svo_default_params(&p, 64, 64);
    std::vector<int> inl(n + 1);
    int success = 0;
    const int m = detail::check(svo_camera_to_world(&p, K, &cameraPoints[0].x, world.data(), n, R, t, inl.data(), &success));
    for (int i = 0; i < 3; ++i) {
        for (int j = 0; j < 3; ++j) rotation.at<double>(i, j) = R[i * 3 + j];
        translation.at<double>(i) = t[i];
    }
    cv::Mat inliers;
    if (success && m > 0) {
        inliers = cv::Mat(m, 1, CV_32SC1);
        for (int i = 0; i < m; ++i) inliers.at<int>(i, 0) = inl[i];
    }
    return std::make_pair(inliers, success != 0);
}
/* vo.h:469-470 / vo.cpp:246-258 */
inline cv::Mat getInverseTransform(const cv::Mat& rotation, const cv::Mat& translation_stereo) {
    double R[9], t[3];
    for (int i = 0; i < 3; ++i) {
        for (int j = 0; j < 3; ++j) R[i * 3 + j] = rotation.at<double>(i, j);
        t[i] = translation_stereo.at<double>(i);
    }
    cv::Mat T(4, 4, CV_64FC1);
    detail::check(svo_get_inverse_transform(R, t, T.ptr<double>(0)));
    return T;
}

/* ---- vo.h:231-380 ---------------------------------------------------------------------- */
class VisualOdometry {
   private:
    int frame_id = 0;
    cv::Mat pendingLeft_, pendingRight_;     /* last submitted frame (streaming calls) */
    svo_ctx* ctx_ = nullptr;                 /* device-resident state (pyramids, currentVOFeatures, R|t, last_transform) */
    cv::Mat imageRightT0_, imageLeftT0_;     /* t0 images for the public circularMatching(): fetched from the device on demand */
    bool t0_fetched_ = false;
    int lookahead_ = 0;
    std::vector<cv::Mat> keepAlive_;
    int cols_ = 0, rows_ = 0, type_ = 0;     /* geometry the context was created for */
    cv::Mat_<float64> last_transform = cv::Mat_<float64>::eye(4, 4);
    std::ofstream vo_log;

    void ensure_ctx(const cv::Mat& img) {
        if (ctx_) {   /* the reference would hit an OpenCV size / type assertion; never read past a smaller buffer */
            if (img.cols != cols_ || img.rows != rows_ || img.type() != type_)
                throw std::runtime_error("svo: image size or type differs from the first frame of this VisualOdometry");
            return;
        }
        cols_ = img.cols;
        rows_ = img.rows;
        type_ = img.type();
        svo_params p = detail::params_for(img);
        p.lookahead = lookahead_;
        ctx_ = svo_create(&p);
        if (!ctx_) throw std::runtime_error(std::string("svo_create: ") + svo_last_error());
        if (!leftCameraProjection_.empty())
            detail::check(svo_set_projection(ctx_, leftCameraProjection_.ptr<float>(0), rightCameraProjection_.ptr<float>(0)));
    }

   public:
    cv::Mat_<float32> leftCameraProjection_, rightCameraProjection_;

    VisualOdometry() {}
    ~VisualOdometry() { svo_destroy(ctx_); }
    VisualOdometry(const VisualOdometry&) = delete;
    VisualOdometry& operator=(const VisualOdometry&) = delete;

    void setLogFile(std::string logpath) {   /* vo.h:279-290 */
        vo_log.open(logpath);
        if (vo_log.good()) {
            CFE_ES_WriteToSysLog("Pose Estimator: Opened VO Logs in executable directory");
            auto now = std::time(0);
            vo_log << "Pose Estimator Test: " << std::ctime(&now) << std::endl;
        } else {
            CFE_ES_WriteToSysLog("Pose Estimator: Unable to open log files");
        }
    }

    /* vo.h:307-309 / vo.cpp:8-26 (sic) */
    void initalize_projection_matricies(const cv::Mat_<float32> leftCameraProjection,
                                        const cv::Mat_<float32> rightCameraProjection) {
        leftCameraProjection.copyTo(leftCameraProjection_);
        rightCameraProjection.copyTo(rightCameraProjection_);
        if (ctx_)
            detail::check(svo_set_projection(ctx_, leftCameraProjection_.ptr<float>(0), rightCameraProjection_.ptr<float>(0)));
    }

    /* vo.h:333-334 / vo.cpp:41-137 */
    std::pair<bool, cv::Mat_<float64>> stereo_callback(const cv::Mat& image_left, const cv::Mat& image_right) {
        int sl, sr;
        const uint8_t* L = detail::gray(image_left, sl);
        const uint8_t* R = detail::gray(image_right, sr);
        if (image_left.rows != image_right.rows || image_left.cols != image_right.cols ||
            image_left.type() != image_right.type())
            throw std::runtime_error("svo: left / right image sizes or types differ");
        cv::Mat rtmp;
        if (sl != sr) {   /* the C ABI takes one stride: repack the right image */
            rtmp = cv::Mat(image_right.rows, image_right.cols, image_left.type());
            cv::Mat ltmp(image_left.rows, image_left.cols, image_left.type());
            const size_t row_bytes = (size_t)image_left.cols * image_left.channels();
            for (int r = 0; r < image_left.rows; ++r) {
                std::memcpy(ltmp.data + (size_t)r * ltmp.step, L + (size_t)r * sl, row_bytes);
                std::memcpy(rtmp.data + (size_t)r * rtmp.step, R + (size_t)r * sr, row_bytes);
            }
            return stereo_callback(ltmp, rtmp);
        }
        ensure_ctx(image_left);
        if (leftCameraProjection_.empty()) {   /* the reference would run with empty matrices; be explicit */
            float z[12] = {0};
            detail::check(svo_set_projection(ctx_, z, z));
        }
        cv::Mat_<float64> T(4, 4);
        const int ok = detail::check(svo_process_frame(ctx_, L, R, sl, T.ptr<double>(0)));
        t0_fetched_ = false;   /* vo.cpp:74-75: the t0 images live in the context's pyramid ring; no host copies per frame */
        ++frame_id;
        if (ok) last_transform = T;
        return std::make_pair(ok != 0, T);
    }

    /* Streaming form of stereo_callback (not in the reference): submit_frame enqueues a frame and returns at
     * once, collect_frame returns the result of the oldest submitted frame -- exactly what stereo_callback would
     * have returned, in order.  With a few frames in flight the ingest and the speculative tracking of the next
     * frame overlap the pose chain of the current one (svo_submit_frame / svo_collect in svo.h).  The images must
     * stay alive (ideally in pinned memory) until the frame is collected; 8UC1 with equal strides. */
    void submit_frame(const cv::Mat& image_left, const cv::Mat& image_right) {
        int sl, sr;
        const uint8_t* L = detail::gray(image_left, sl);
        const uint8_t* R = detail::gray(image_right, sr);
        if (image_left.rows != image_right.rows || image_left.cols != image_right.cols || sl != sr)
            throw std::runtime_error("svo: submit_frame needs equally sized images with equal strides");
        ensure_ctx(image_left);
        if (leftCameraProjection_.empty()) {
            float z[12] = {0};
            detail::check(svo_set_projection(ctx_, z, z));
        }
        detail::check(svo_submit_frame(ctx_, L, R, sl));
        pendingLeft_ = image_left;   /* shallow references: keep the buffers alive */
        pendingRight_ = image_right;
    }
    /* Look-ahead form (svo_submit_frames): up to `lookahead` frames that are already there; their pyramids and FAST
     * corners are computed in one launch set, the results come back one by one through collect_frame().  Set the
     * look-ahead depth before the first frame with set_lookahead(K), K <= 16. */
    void set_lookahead(int k) { lookahead_ = k; }
    void submit_frames(const std::vector<cv::Mat>& lefts, const std::vector<cv::Mat>& rights) {
        if (lefts.empty() || lefts.size() != rights.size()) throw std::runtime_error("svo: submit_frames needs equally many left and right frames");
        std::vector<const uint8_t*> lp, rp;
        int s0 = 0;
        for (size_t i = 0; i < lefts.size(); ++i) {
            int sl, sr;
            lp.push_back(detail::gray(lefts[i], sl));
            rp.push_back(detail::gray(rights[i], sr));
            if (i == 0) s0 = sl;
            if (sl != s0 || sr != s0 || lefts[i].rows != lefts[0].rows || lefts[i].cols != lefts[0].cols ||
                rights[i].rows != lefts[0].rows || rights[i].cols != lefts[0].cols)
                throw std::runtime_error("svo: submit_frames needs equally sized frames with equal strides");
        }
        ensure_ctx(lefts[0]);
        if (leftCameraProjection_.empty()) {
            float z[12] = {0};
            detail::check(svo_set_projection(ctx_, z, z));
        }
        detail::check(svo_submit_frames(ctx_, (int)lp.size(), lp.data(), rp.data(), s0));
        pendingLeft_ = lefts.back();
        pendingRight_ = rights.back();
        keepAlive_.insert(keepAlive_.end(), lefts.begin(), lefts.end());   /* shallow references until collected */
        keepAlive_.insert(keepAlive_.end(), rights.begin(), rights.end());
    }
    std::pair<bool, cv::Mat_<float64>> collect_frame() {
        svo_frame_result r;
        detail::check(svo_collect(ctx_, &r));
        cv::Mat_<float64> T(4, 4);
        std::memcpy(T.ptr<double>(0), r.transform, sizeof(double) * 16);
        ++frame_id;
        if (r.success) last_transform = T;
        t0_fetched_ = false;
        if (svo_pending(ctx_) == 0) {   /* nothing in flight any more: drop the references to the caller's buffers */
            pendingLeft_ = cv::Mat();
            pendingRight_ = cv::Mat();
            keepAlive_.clear();
        }
        return std::make_pair(r.success != 0, T);
    }

    /* vo.h:354-362 / vo.cpp:315-366 */
    void matchingFeatures(const cv::Mat& imageLeftT0, const cv::Mat& imageRightT0, const cv::Mat& imageLeftT1,
                          const cv::Mat& imageRightT1, FeatureSet& currentVOFeatures,
                          std::vector<cv::Point2f>& pointsLeftT0, std::vector<cv::Point2f>& pointsRightT0,
                          std::vector<cv::Point2f>& pointsLeftT1, std::vector<cv::Point2f>& pointsRightT1) {
        int s0, s1, s2, s3;
        const uint8_t *a = detail::gray(imageLeftT0, s0), *b = detail::gray(imageRightT0, s1),
                      *c = detail::gray(imageLeftT1, s2), *d = detail::gray(imageRightT1, s3);
        if (s0 != s1 || s0 != s2 || s0 != s3) throw std::runtime_error("svo: the four images must share one row stride");
        svo_params p = detail::params_for(imageLeftT0);
        const int n0 = currentVOFeatures.size();
        const int cap = p.buckets_along_height * p.buckets_along_width + n0;
        std::vector<float> xy(2 * (size_t)cap), o0(2 * (size_t)cap), o1(2 * (size_t)cap), o2(2 * (size_t)cap), o3(2 * (size_t)cap);
        std::vector<int> age(cap), str(cap);
        for (int i = 0; i < n0; ++i) {
            xy[2 * i] = currentVOFeatures.points[i].x;
            xy[2 * i + 1] = currentVOFeatures.points[i].y;
            age[i] = currentVOFeatures.ages[i];
            str[i] = currentVOFeatures.strengths[i];
        }
        int n = n0;
        const int m = detail::check(svo_matching_features(&p, a, b, c, d, s0, xy.data(), age.data(), str.data(), &n, cap,
                                                          o0.data(), o1.data(), o2.data(), o3.data()));
        auto fill = [m](std::vector<cv::Point2f>& v, const std::vector<float>& src) {
            v.resize(m);
            for (int i = 0; i < m; ++i) v[i] = cv::Point2f(src[2 * i], src[2 * i + 1]);
        };
        fill(pointsLeftT0, o0);
        fill(pointsRightT0, o1);
        fill(pointsLeftT1, o2);
        fill(pointsRightT1, o3);
        fill(currentVOFeatures.points, xy);
        currentVOFeatures.ages.assign(age.begin(), age.begin() + m);
        currentVOFeatures.strengths.assign(str.begin(), str.begin() + m);
    }

    /* vo.h:374-379 / vo.cpp:169-240: tracks against the images of the previous stereo_callback */
    void circularMatching(const cv::Mat& imgLeftT1, const cv::Mat& imgRightT1, std::vector<cv::Point2f>& pointsLeftT0,
                          std::vector<cv::Point2f>& pointsRightT0, std::vector<cv::Point2f>& pointsLeftT1,
                          std::vector<cv::Point2f>& pointsRightT1, FeatureSet& current_features) {
        if (pointsLeftT0.size() == 0) return;   /* vo.cpp:179-181 */
        if (!ctx_ || frame_id == 0) throw std::runtime_error("svo: circularMatching needs a prior stereo_callback");
        if (svo_pending(ctx_) != 0) throw std::runtime_error("svo: collect the submitted frames before circularMatching");
        if (!t0_fetched_) {   /* the images of the cached t0 pyramids (vo.h:257-258), once per frame at most */
            imageLeftT0_ = cv::Mat(rows_, cols_, type_);
            imageRightT0_ = cv::Mat(rows_, cols_, type_);
            detail::check(svo_get_frame_images(ctx_, 0, imageLeftT0_.data, imageRightT0_.data, (int)imageLeftT0_.step));
            t0_fetched_ = true;
        }
        int s0, s1, s2, s3;
        const uint8_t *a = detail::gray(imageLeftT0_, s0), *b = detail::gray(imageRightT0_, s1),
                      *c = detail::gray(imgLeftT1, s2), *d = detail::gray(imgRightT1, s3);
        if (s0 != s1 || s0 != s2 || s0 != s3) throw std::runtime_error("svo: the four images must share one row stride");
        svo_params p = detail::params_for(imgLeftT1);
        const int n = (int)pointsLeftT0.size();
        std::vector<float> r0(2 * (size_t)n), l1(2 * (size_t)n), r1(2 * (size_t)n);
        std::vector<uint8_t> keep(n + 1);
        const int m = detail::check(svo_circular_match(&p, a, b, c, d, s0, &pointsLeftT0[0].x, n, r0.data(), l1.data(),
                                                       r1.data(), keep.data()));
        if (current_features.size() == n) {
            std::vector<bool> ok(n);
            for (int i = 0; i < n; ++i) ok[i] = keep[i] != 0;
            deleteFeaturesWithFailureStatus(current_features, ok);
        }
        auto fill = [m](std::vector<cv::Point2f>& v, const std::vector<float>& src) {
            v.resize(m);
            for (int i = 0; i < m; ++i) v[i] = cv::Point2f(src[2 * i], src[2 * i + 1]);
        };
        pointsLeftT0.resize(m);   /* compacted in place by the ABI call */
        fill(pointsRightT0, r0);
        fill(pointsLeftT1, l1);
        fill(pointsRightT1, r1);
    }

    /* per-frame counters of the last stereo_callback (what the reference printf's) */
    svo_frame_result lastResult() {
        svo_frame_result r;
        std::memset(&r, 0, sizeof(r));
        if (ctx_) detail::check(svo_get_result(ctx_, 0, &r));
        return r;
    }
};

/* ---- upstream (hjamal3) names used by the README / north_star, as thin aliases ------------ */
class StereoVO : public VisualOdometry {
   public:
    std::pair<bool, cv::Mat_<float64>> run(const cv::Mat& l, const cv::Mat& r) { return stereo_callback(l, r); }
};
inline void bucketingFeatures(const cv::Mat& image, FeatureSet& fs) { fs.filterByBucketLocation(image); }
inline std::pair<cv::Mat, bool> trackingFrame2Frame(const cv::Mat_<float32>& K, const std::vector<cv::Point2f>& cam,
                                                    const cv::Mat_<float32>& world, cv::Mat& R, cv::Mat& t) {
    return cameraToWorld(K, cam, world, R, t);
}
inline void removeInvalidPoints(std::vector<cv::Point2f>& pts, const std::vector<bool>& ok) {
    deletePointsWithFailureStatus(pts, ok);
}
inline std::vector<bool> checkValidMatch(const std::vector<cv::Point2f>& a, const std::vector<cv::Point2f>& b, float32 thr) {
    return findClosePoints(a, b, thr);
}
/* frame_pose = frame_pose * inverse([R|t]) (main.cpp:396): the 4x4 update stays on the host */
inline void integrateOdometryStereo(cv::Mat& frame_pose, const cv::Mat& rotation, const cv::Mat& translation) {
    frame_pose = frame_pose * getInverseTransform(rotation, translation);
}

}  // namespace visual_odometry

#endif /* SVO_DROPIN_VO_H_ */
