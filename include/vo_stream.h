/*
 * vo_stream.h -- streaming wrapper around the drop-in VisualOdometry class: what the reference's ROS node
 * (src/stereo_vo.cpp:6-68 of Alex7Li/stereo_visual_odometry) does around stereo_callback, without ROS:
 *   - to_mono8()            cv_bridge::toCvCopy(img, MONO8)              (stereo_vo.cpp:6-14)
 *   - load_calibration()    fx / fy / cx / cy / bf from the YAML          (stereo_vo.cpp:33-47; both key styles:
 *                           `fx:` as the node reads them, `Camera.fx:` as calibration/ spells them)
 *   - ApproximateTimePairer message_filters ApproximateTime policy for two topics, queue size 10
 *                           (stereo_vo.cpp:53-58): every left image is paired with the right image closest in time
 *   - StereoVONode          two image callbacks -> pairing -> VisualOdometry::stereo_callback -> pose integration
 *                           frame_pose = frame_pose * T (src/main.cpp:396)
 *   - displayPoints / displayTracking   the debug overlays of src/vo.cpp:28-39 / include/vo.h:292-296 (host only;
 *                           the reference draws with cv::circle, radius 2, value 255)
 * Host code over include/vo.h; the per-frame work runs on the GPU through the C ABI like every other call.
 */
#ifndef SVO_DROPIN_VO_STREAM_H_
#define SVO_DROPIN_VO_STREAM_H_

#include <cmath>
#include <cstdlib>
#include <deque>
#include <fstream>
#include <functional>
#include <string>
#include <utility>
#include <vector>

#include "vo.h"

namespace visual_odometry {

/* cv_bridge::toCvCopy(..., MONO8): 8UC1 passes through (deep copy), 8UC3 is BGR -> gray exactly as
 * cv::cvtColor(BGR2GRAY) (15-bit fixed point, round to nearest), 16UC1 keeps the high byte. */
inline cv::Mat to_mono8(const cv::Mat& img) {
    if (img.empty()) return cv::Mat();
    cv::Mat g(img.rows, img.cols, CV_8UC1);
    for (int y = 0; y < img.rows; ++y) {
        const uint8_t* p = img.ptr<uint8_t>(y);
        uint8_t* d = g.ptr<uint8_t>(y);
        if (img.type() == CV_8UC1) {
            std::memcpy(d, p, (size_t)img.cols);
        } else if (img.type() == CV_8UC3) {
            for (int x = 0; x < img.cols; ++x)
                d[x] = (uint8_t)((p[3 * x] * 3735 + p[3 * x + 1] * 19235 + p[3 * x + 2] * 9798 + (1 << 14)) >> 15);
        } else if (img.type() == CV_16UC1) {
            const uint16_t* q = reinterpret_cast<const uint16_t*>(p);
            for (int x = 0; x < img.cols; ++x) d[x] = (uint8_t)(q[x] >> 8);
        } else {
            return cv::Mat();   /* cv_bridge would throw; the node then drops the frame (stereo_vo.cpp:10-12) */
        }
    }
    return g;
}

/* `key: value` lines; keys with or without the `Camera.` prefix; P_right[0][3] = bf (stereo_vo.cpp:46-47) */
inline bool load_calibration(const std::string& path, cv::Mat_<float>& P_left, cv::Mat_<float>& P_right) {
    std::ifstream f(path);
    if (!f.good()) return false;
    double v[5] = {0, 0, 0, 0, 0};
    const char* names[5] = {"fx", "fy", "cx", "cy", "bf"};
    int seen = 0;
    std::string line;
    while (std::getline(f, line)) {
        const size_t hash = line.find('#');
        if (hash != std::string::npos) line.erase(hash);
        const size_t colon = line.find(':');
        if (colon == std::string::npos) continue;
        std::string key = line.substr(0, colon);
        key.erase(0, key.find_first_not_of(" \t"));
        key.erase(key.find_last_not_of(" \t") + 1);
        if (key.rfind("Camera.", 0) == 0) key.erase(0, 7);
        const std::string val = line.substr(colon + 1);
        char* end = nullptr;
        const double x = std::strtod(val.c_str(), &end);
        if (end == val.c_str()) continue;
        for (int k = 0; k < 5; ++k)
            if (key == names[k]) {
                v[k] = x;
                seen |= 1 << k;
            }
    }
    if (seen != 31) return false;
    P_left = cv::Mat_<float>(3, 4);
    P_right = cv::Mat_<float>(3, 4);
    const float l[12] = {(float)v[0], 0.f, (float)v[2], 0.f, 0.f, (float)v[1], (float)v[3], 0.f, 0.f, 0.f, 1.f, 0.f};
    for (int i = 0; i < 12; ++i) P_left(i / 4, i % 4) = P_right(i / 4, i % 4) = l[i];
    P_right(0, 3) = (float)v[4];
    return true;
}

/* Two-topic ApproximateTime pairing (message_filters::sync_policies::ApproximateTime, queue size 10 per topic):
 * a pair is emitted once both queues are non-empty; the older head is matched with the message of the other queue
 * that is closest in time, messages older than the match are dropped.  Emission is in time order and every message is
 * used at most once, which is what the policy guarantees for two topics. */
class ApproximateTimePairer {
   public:
    typedef std::function<void(const cv::Mat&, const cv::Mat&, double, double)> Callback;
    explicit ApproximateTimePairer(Callback cb, size_t queue_size = 10, double max_interval = 1e300)
        : cb_(std::move(cb)), qsize_(queue_size), max_dt_(max_interval) {}
    void addLeft(double stamp, const cv::Mat& img) { push(left_, stamp, img); }
    void addRight(double stamp, const cv::Mat& img) { push(right_, stamp, img); }
    size_t pairs() const { return n_pairs_; }
    size_t dropped() const { return n_dropped_; }

   private:
    struct Msg { double t; cv::Mat img; };
    void push(std::deque<Msg>& q, double stamp, const cv::Mat& img) {
        q.push_back(Msg{stamp, img});
        if (q.size() > qsize_) {
            q.pop_front();
            ++n_dropped_;
        }
        match();
    }
    void match() {
        for (;;) {
            if (left_.empty() || right_.empty()) return;
            const bool pivot_left = left_.front().t > right_.front().t;   /* the pivot is the NEWER of the two heads */
            std::deque<Msg>& P = pivot_left ? left_ : right_;
            std::deque<Msg>& O = pivot_left ? right_ : left_;
            const double tp = P.front().t;
            size_t k = 0;   /* last message of the other topic that is not newer than the pivot */
            while (k + 1 < O.size() && O[k + 1].t <= tp) ++k;
            size_t pick = k;
            if (O[k].t != tp) {
                if (k + 1 < O.size()) {
                    if (std::fabs(O[k + 1].t - tp) < std::fabs(O[k].t - tp)) pick = k + 1;
                } else if (!closed_) {
                    return;   /* a closer message of the other topic may still arrive */
                }
            }
            n_dropped_ += pick;
            O.erase(O.begin(), O.begin() + (long)pick);
            const Msg mo = O.front(), mp = P.front();
            O.pop_front();
            P.pop_front();
            if (std::fabs(mo.t - mp.t) > max_dt_) {
                n_dropped_ += 2;
                continue;
            }
            ++n_pairs_;
            if (pivot_left) cb_(mp.img, mo.img, mp.t, mo.t);
            else cb_(mo.img, mp.img, mo.t, mp.t);
        }
    }

   public:
    /* no more messages will arrive: match what is queued */
    void flush() {
        closed_ = true;
        match();
        closed_ = false;
    }

   private:
    Callback cb_;
    size_t qsize_;
    double max_dt_;
    std::deque<Msg> left_, right_;
    size_t n_pairs_ = 0, n_dropped_ = 0;
    bool closed_ = false;
};

/* src/vo.cpp:28-39: cv::circle(image, point, 2, 255) for every point, on an 8UC1 image (in place).  Radius 2 with
 * thickness 1 is OpenCV's 8-pixel ring (checked against cv2.circle in tests/test_gpu_cpp.py). */
inline void displayPoints(cv::Mat& image, const std::vector<cv::Point2f>& points) {
    static const int ring[8][2] = {{0, -2}, {-1, -1}, {1, -1}, {-2, 0}, {2, 0}, {-1, 1}, {1, 1}, {0, 2}};   /* == cv::circle(r = 2) */
    for (const cv::Point2f& p : points) {
        const int cx = (int)p.x, cy = (int)p.y;   /* cv::Point(points[i].x, points[i].y): truncation */
        for (const auto& o : ring) {
            const int x = cx + o[0], y = cy + o[1];
            if (x >= 0 && x < image.cols && y >= 0 && y < image.rows) image.at<uint8_t>(y, x) = 255;
        }
    }
}
/* include/vo.h:294-296 of the reference declares displayTracking without defining it; here: image_2 with the
 * displacement of every feature drawn as a line from its t0 to its t1 position and the t1 end marked. */
inline cv::Mat displayTracking(const cv::Mat& image_1, const cv::Mat& image_2, const std::vector<cv::Point2f>& pointsLeftT0,
                               const std::vector<cv::Point2f>& pointsLeftT1) {
    (void)image_1;
    cv::Mat out = to_mono8(image_2);
    const size_t n = pointsLeftT0.size() < pointsLeftT1.size() ? pointsLeftT0.size() : pointsLeftT1.size();
    for (size_t i = 0; i < n; ++i) {
        int x0 = (int)pointsLeftT0[i].x, y0 = (int)pointsLeftT0[i].y;
        const int x1 = (int)pointsLeftT1[i].x, y1 = (int)pointsLeftT1[i].y;
        const int dx = std::abs(x1 - x0), dy = -std::abs(y1 - y0), sx = x0 < x1 ? 1 : -1, sy = y0 < y1 ? 1 : -1;
        int err = dx + dy;
        for (int guard = 0; guard < 4096; ++guard) {   /* Bresenham */
            if (x0 >= 0 && x0 < out.cols && y0 >= 0 && y0 < out.rows) out.at<uint8_t>(y0, x0) = 255;
            if (x0 == x1 && y0 == y1) break;
            const int e2 = 2 * err;
            if (e2 >= dy) { err += dy; x0 += sx; }
            if (e2 <= dx) { err += dx; y0 += sy; }
        }
    }
    displayPoints(out, pointsLeftT1);
    return out;
}

/* The node of src/stereo_vo.cpp:17-68 without ROS: feed it the two image topics, it pairs them, converts to MONO8,
 * runs stereo_callback and integrates the pose.  `on_pose` (optional) is called after every processed pair. */
class StereoVONode {
   public:
    typedef std::function<void(bool, const cv::Mat_<double>& /*transform*/, const cv::Mat_<double>& /*frame_pose*/)> PoseCallback;
    StereoVONode(const cv::Mat_<float>& P_left, const cv::Mat_<float>& P_right, PoseCallback on_pose = PoseCallback())
        : pairer_([this](const cv::Mat& l, const cv::Mat& r, double, double) { process(l, r); }), on_pose_(std::move(on_pose)) {
        vo_.initalize_projection_matricies(P_left, P_right);
        frame_pose_ = cv::Mat_<double>::eye(4, 4);
    }
    explicit StereoVONode(const std::string& calib_yaml, PoseCallback on_pose = PoseCallback())
        : pairer_([this](const cv::Mat& l, const cv::Mat& r, double, double) { process(l, r); }), on_pose_(std::move(on_pose)) {
        cv::Mat_<float> Pl, Pr;
        if (!load_calibration(calib_yaml, Pl, Pr)) throw std::runtime_error("cv failed to load yaml");   /* stereo_vo.cpp:35-38 */
        vo_.initalize_projection_matricies(Pl, Pr);
        frame_pose_ = cv::Mat_<double>::eye(4, 4);
    }
    void leftImage(double stamp, const cv::Mat& img) { pairer_.addLeft(stamp, img); }     /* left/image_rect  */
    void rightImage(double stamp, const cv::Mat& img) { pairer_.addRight(stamp, img); }   /* right/image_rect */
    void flush() { pairer_.flush(); }
    const cv::Mat_<double>& pose() const { return frame_pose_; }
    size_t framesProcessed() const { return n_frames_; }
    VisualOdometry& vo() { return vo_; }

   private:
    void process(const cv::Mat& left, const cv::Mat& right) {
        const cv::Mat l = to_mono8(left), r = to_mono8(right);
        if (l.empty() || r.empty()) return;   /* conversion failed: the node passes an empty Mat on; nothing to do */
        const std::pair<bool, cv::Mat_<double>> res = vo_.stereo_callback(l, r);
        cv::Mat_<double> next(4, 4);
        for (int i = 0; i < 4; ++i)
            for (int j = 0; j < 4; ++j) {
                double s = 0;
                for (int k = 0; k < 4; ++k) s += frame_pose_(i, k) * res.second(k, j);
                next(i, j) = s;
            }
        frame_pose_ = next;   /* main.cpp:396, also on failure (the transform is then the last good one) */
        ++n_frames_;
        if (on_pose_) on_pose_(res.first, res.second, frame_pose_);
    }
    VisualOdometry vo_;
    ApproximateTimePairer pairer_;
    PoseCallback on_pose_;
    cv::Mat_<double> frame_pose_;
    size_t n_frames_ = 0;
};

}  // namespace visual_odometry
#endif /* SVO_DROPIN_VO_STREAM_H_ */
