/*
 * svo_cv_shim.h -- the handful of OpenCV core types that the reference's include/vo.h uses in
 * its signatures (cv::Mat, cv::Mat_<T>, cv::Point2f, cv::Size), for builds where OpenCV C++ is
 * not installed (this image has no OpenCV headers/libs).  When <opencv2/core.hpp> is available
 * include/vo.h uses the real thing instead and this file is not compiled.
 *
 * Only what the hot-path API and the reference's own tests/driver (src/main.cpp) touch is here:
 * construction from external data, rows/cols/type/step/data, at<T>, Mat_<T>(r,c) with (i,j)
 * access, eye/zeros, clone/copyTo, 4x4 / 3x3 double products, size().height, empty().
 */
#ifndef SVO_CV_SHIM_H_
#define SVO_CV_SHIM_H_

#include <cstdint>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <vector>

#define CV_8U 0
#define CV_16U 2
#define CV_32S 4
#define CV_32F 5
#define CV_64F 6
#define CV_MAKETYPE(depth, cn) ((depth) + (((cn)-1) << 3))
#define CV_8UC1 CV_MAKETYPE(CV_8U, 1)
#define CV_8UC3 CV_MAKETYPE(CV_8U, 3)
#define CV_16UC1 CV_MAKETYPE(CV_16U, 1)
#define CV_32FC1 CV_MAKETYPE(CV_32F, 1)
#define CV_32FC3 CV_MAKETYPE(CV_32F, 3)
#define CV_64FC1 CV_MAKETYPE(CV_64F, 1)
#define CV_32SC1 CV_MAKETYPE(CV_32S, 1)

namespace cv {

struct Point2f {
    float x, y;
    Point2f() : x(0), y(0) {}
    Point2f(float x_, float y_) : x(x_), y(y_) {}
};
struct Size {
    int width, height;
    Size() : width(0), height(0) {}
    Size(int w, int h) : width(w), height(h) {}
};

class Mat {
   public:
    int rows = 0, cols = 0;
    size_t step = 0;
    uint8_t* data = nullptr;

    Mat() {}
    Mat(int r, int c, int type) { create(r, c, type); }
    /* external data, not owned (like cv::Mat(rows, cols, type, void*)) */
    Mat(int r, int c, int type, void* ext, size_t step_ = 0) : rows(r), cols(c), type_(type) {
        step = step_ ? step_ : (size_t)c * elemSize();
        data = static_cast<uint8_t*>(ext);
    }
    void create(int r, int c, int type) {
        rows = r;
        cols = c;
        type_ = type;
        step = (size_t)c * elemSize();
        buf_ = std::shared_ptr<uint8_t>(new uint8_t[(size_t)r * step + 1](), std::default_delete<uint8_t[]>());
        data = buf_.get();
    }
    int type() const { return type_; }
    int depth() const { return type_ & 7; }
    int channels() const { return (type_ >> 3) + 1; }
    size_t elemSize() const {
        static const int sz[8] = {1, 1, 2, 2, 4, 4, 8, 2};
        return (size_t)sz[depth()] * channels();
    }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    Size size() const { return Size(cols, rows); }
    bool isContinuous() const { return step == (size_t)cols * elemSize(); }
    template <class T>
    T& at(int r, int c) { return *reinterpret_cast<T*>(data + (size_t)r * step + (size_t)c * sizeof(T)); }
    template <class T>
    const T& at(int r, int c) const { return *reinterpret_cast<const T*>(data + (size_t)r * step + (size_t)c * sizeof(T)); }
    template <class T>
    T& at(int i) { return rows == 1 ? at<T>(0, i) : at<T>(i, 0); }
    template <class T>
    const T& at(int i) const { return rows == 1 ? at<T>(0, i) : at<T>(i, 0); }
    template <class T>
    T* ptr(int r = 0) { return reinterpret_cast<T*>(data + (size_t)r * step); }
    template <class T>
    const T* ptr(int r = 0) const { return reinterpret_cast<const T*>(data + (size_t)r * step); }
    Mat clone() const {
        Mat m;
        copyTo(m);
        return m;
    }
    void copyTo(Mat& dst) const {
        if (empty()) {
            dst = Mat();
            return;
        }
        Mat m(rows, cols, type_);
        for (int r = 0; r < rows; ++r) std::memcpy(m.data + (size_t)r * m.step, data + (size_t)r * step, (size_t)cols * elemSize());
        dst = m;
    }
    static Mat eye(int r, int c, int type) {
        Mat m(r, c, type);
        for (int i = 0; i < r && i < c; ++i) {
            if (m.depth() == CV_64F) m.at<double>(i, i) = 1.0;
            else if (m.depth() == CV_32F) m.at<float>(i, i) = 1.f;
            else throw std::runtime_error("cv shim: eye() supports CV_32F / CV_64F");
        }
        return m;
    }
    static Mat zeros(int r, int c, int type) { return Mat(r, c, type); }
    /* FP64 matrix product (the CLI integrates frame_pose = frame_pose * transform, main.cpp:396) */
    Mat operator*(const Mat& b) const {
        if (depth() != CV_64F || b.depth() != CV_64F || cols != b.rows) throw std::runtime_error("cv shim: operator* needs CV_64F");
        Mat m(rows, b.cols, CV_64F);
        for (int i = 0; i < rows; ++i)
            for (int j = 0; j < b.cols; ++j) {
                double s = 0;
                for (int k = 0; k < cols; ++k) s += at<double>(i, k) * b.at<double>(k, j);
                m.at<double>(i, j) = s;
            }
        return m;
    }

   protected:
    int type_ = 0;
    std::shared_ptr<uint8_t> buf_;
};

template <class T>
struct DepthOf;
template <> struct DepthOf<unsigned char> { enum { value = CV_8U }; };
template <> struct DepthOf<int> { enum { value = CV_32S }; };
template <> struct DepthOf<float> { enum { value = CV_32F }; };
template <> struct DepthOf<double> { enum { value = CV_64F }; };

template <class T>
class Mat_ : public Mat {
   public:
    Mat_() { type_ = DepthOf<T>::value; }
    Mat_(int r, int c) : Mat(r, c, DepthOf<T>::value) {}
    Mat_(int r, int c, int /*type, ignored like cv::Mat_*/) : Mat(r, c, DepthOf<T>::value) {}
    Mat_(Size s) : Mat(s.height, s.width, DepthOf<T>::value) {}
    Mat_(const Mat& m) {
        if (!m.empty() && m.type() != (int)DepthOf<T>::value) throw std::runtime_error("cv shim: Mat_ type mismatch");
        static_cast<Mat&>(*this) = m;
        type_ = DepthOf<T>::value;
    }
    T& operator()(int r, int c) { return this->template at<T>(r, c); }
    const T& operator()(int r, int c) const { return this->template at<T>(r, c); }
    T* operator[](int r) { return this->template ptr<T>(r); }
    const T* operator[](int r) const { return this->template ptr<T>(r); }
    static Mat_<T> eye(int r, int c) { return Mat_<T>(Mat::eye(r, c, DepthOf<T>::value)); }
    static Mat_<T> zeros(int r, int c) { return Mat_<T>(r, c); }
};

}  // namespace cv

#endif /* SVO_CV_SHIM_H_ */
