// TEST INFRASTRUCTURE — NOT PRODUCT CODE, NOT OpenCV.
// The smallest cv:: surface that lets the reference's feature/orb_extractor.cc, orb_extractor_node.cc, orb_params.cc and
// match/stereo.cc compile UNMODIFIED from /root/reference (the OpenCV C++ SDK is not installed here and cannot be). The
// image primitives route to oracle/cvprim.cc, whose arithmetic is pinned bit-for-bit against cv2 4.13 by
// tests/test_oracle_vs_cv2.py. Only 8-bit single-channel and 32-bit float single-channel matrices exist.
#pragma once

#include <algorithm>
#include <array>
#include <cassert>
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <vector>

#include "cvprim.h"

typedef unsigned char uchar;

#define CV_8U 0
#define CV_32F 5
#define CV_8UC1 CV_8U
#define CV_32FC1 CV_32F
#define CV_SSE3 0
#define CV_MAJOR_VERSION 4

inline int cvRound(double v) { return cvprim::round_half_even(v); }
inline int cvRound(float v) { return cvprim::round_half_even(v); }
inline int cvRound(int v) { return v; }
inline int cvFloor(double v) { return cvprim::floor_int(v); }
inline int cvFloor(float v) { return cvprim::floor_int(v); }
inline int cvCeil(double v) { return cvprim::ceil_int(v); }
inline int cvCeil(float v) { return cvprim::ceil_int(v); }

namespace cv {

enum { INTER_LINEAR = 1, BORDER_REFLECT_101 = 4, LINE_AA = 16, NORM_L1 = 2 };

template<typename T>
struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T x_, T y_) : x(x_), y(y_) {}
    template<typename U>
    Point_& operator*=(U s) {
        x = static_cast<T>(x * s);
        y = static_cast<T>(y * s);
        return *this;
    }
};
typedef Point_<int> Point2i;
typedef Point_<float> Point2f;
typedef Point2i Point;

struct Size {
    int width, height;
    Size() : width(0), height(0) {}
    Size(int w, int h) : width(w), height(h) {}
    Size(double w, double h) : width(static_cast<int>(w)), height(static_cast<int>(h)) {}
};

struct Scalar {
    double v[4];
    Scalar(double a = 0) : v{a, 0, 0, 0} {}
};

struct KeyPoint {
    Point2f pt;
    float size = 0, angle = -1, response = 0;
    int octave = 0, class_id = -1;
    KeyPoint() {}
    KeyPoint(float x, float y, float s, float a = -1, float r = 0, int o = 0, int c = -1)
        : pt(x, y), size(s), angle(a), response(r), octave(o), class_id(c) {}
};
static_assert(sizeof(KeyPoint) == 28, "cv::KeyPoint layout");

struct Range {
    int start, end;
};

// cv::MatExpr of Mat::zeros / Mat::ones: assigning it to a Mat that already has the right size and type fills that Mat's
// buffer IN PLACE (MatOp_Initializer::assign -> Mat::create keeps the allocation). feature/orb_extractor.cc:637 relies
// on this: `descriptors = cv::Mat::zeros(...)` must keep writing into the rows of the caller's output matrix.
struct MatInit {
    int rows, cols, type;
    double value;
};

class Mat {
public:
    int rows = 0, cols = 0;
    size_t step = 0;  // bytes per row
    uchar* data = nullptr;

    Mat() {}
    Mat(int r, int c, int type) { create(r, c, type); }
    Mat(int r, int c, int type, const Scalar& s) {
        create(r, c, type);
        fill(s.v[0]);
    }
    // header over foreign memory (not owned)
    Mat(int r, int c, int type, void* ext, size_t step_) : rows(r), cols(c), step(step_), data(static_cast<uchar*>(ext)), type_(type) {}

    void create(int r, int c, int type) {
        if (rows == r && cols == c && type_ == type && data && owner_) return;
        rows = r;
        cols = c;
        type_ = type;
        step = static_cast<size_t>(c) * elem();
        owner_ = std::shared_ptr<uchar>(new uchar[std::max<size_t>(step * r, 1)], std::default_delete<uchar[]>());
        data = owner_.get();
    }
    void release() { *this = Mat(); }
    // reshape(cn): same rows, the row's elements regrouped into cn channels (camera/perspective.h:36-43 turns an N x 2 float
    // matrix into N two-channel points and back)
    Mat reshape(int cn) const {
        Mat m(*this);
        m.cols = cols * channels_ / cn;
        m.channels_ = cn;
        return m;
    }
    // reshape(cn, rows): a continuous matrix (here: one row, or rows packed back to back) seen with another row count
    Mat reshape(int cn, int new_rows) const {
        if (new_rows == 0) return reshape(cn);
        if (rows > 1 && step != static_cast<size_t>(cols) * elem() * channels_) throw std::invalid_argument("cv::Mat::reshape: not continuous");
        const size_t total = static_cast<size_t>(rows) * cols * channels_;
        if (total % (static_cast<size_t>(new_rows) * cn) != 0) throw std::invalid_argument("cv::Mat::reshape: bad row count");
        Mat m(*this);
        m.rows = new_rows;
        m.cols = static_cast<int>(total / (static_cast<size_t>(new_rows) * cn));
        m.channels_ = cn;
        m.step = static_cast<size_t>(m.cols) * elem() * cn;
        return m;
    }
    int channels() const { return channels_; }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    bool isContinuous() const { return rows <= 1 || step == static_cast<size_t>(cols) * elem(); }
    int type() const { return type_; }
    size_t elem() const { return type_ == CV_32F ? 4 : 1; }
    size_t step1() const { return step / elem(); }

    template<typename T>
    T& at(int y, int x) {
        return *reinterpret_cast<T*>(data + static_cast<size_t>(y) * step + static_cast<size_t>(x) * sizeof(T));
    }
    template<typename T>
    const T& at(int y, int x) const {
        return *reinterpret_cast<const T*>(data + static_cast<size_t>(y) * step + static_cast<size_t>(x) * sizeof(T));
    }
    uchar* ptr(int y = 0) { return data + static_cast<size_t>(y) * step; }
    const uchar* ptr(int y = 0) const { return data + static_cast<size_t>(y) * step; }
    template<typename T>
    T* ptr(int y = 0) { return reinterpret_cast<T*>(data + static_cast<size_t>(y) * step); }
    template<typename T>
    const T* ptr(int y = 0) const { return reinterpret_cast<const T*>(data + static_cast<size_t>(y) * step); }

    Mat rowRange(int a, int b) const {
        if (a < 0 || b > rows || a > b) throw std::out_of_range("cv::Mat::rowRange");
        Mat m(*this);
        m.data = data + static_cast<size_t>(a) * step;
        m.rows = b - a;
        return m;
    }
    Mat colRange(int a, int b) const {
        if (a < 0 || b > cols || a > b) throw std::out_of_range("cv::Mat::colRange");
        Mat m(*this);
        m.data = data + static_cast<size_t>(a) * elem();
        m.cols = b - a;
        return m;
    }
    Mat row(int y) const { return rowRange(y, y + 1); }
    Mat clone() const {
        Mat m(rows, cols, type_);
        for (int y = 0; y < rows; ++y) std::memcpy(m.ptr(y), ptr(y), static_cast<size_t>(cols) * elem());
        return m;
    }
    void convertTo(Mat& dst, int type) const {
        assert(type_ == CV_8U && type == CV_32F);
        Mat out(rows, cols, type);
        for (int y = 0; y < rows; ++y)
            for (int x = 0; x < cols; ++x) out.at<float>(y, x) = static_cast<float>(at<uchar>(y, x));
        dst = out;
    }
    Mat& operator-=(const Mat& o) {
        assert(type_ == CV_32F && o.type_ == CV_32F && rows == o.rows && cols == o.cols);
        for (int y = 0; y < rows; ++y)
            for (int x = 0; x < cols; ++x) at<float>(y, x) -= o.at<float>(y, x);
        return *this;
    }
    static MatInit zeros(int r, int c, int type) { return MatInit{r, c, type, 0.0}; }
    static MatInit ones(int r, int c, int type) { return MatInit{r, c, type, 1.0}; }
    Mat(const MatInit& e) {
        create(e.rows, e.cols, e.type);
        fill(e.value);
    }
    Mat& operator=(const MatInit& e) {
        // same geometry: keep the buffer (even a non-owned row range of another matrix), exactly like cv::Mat::create
        if (!(rows == e.rows && cols == e.cols && type_ == e.type && data)) create(e.rows, e.cols, e.type);
        fill(e.value);
        return *this;
    }

private:
    void fill(double v) {
        for (int y = 0; y < rows; ++y)
            for (int x = 0; x < cols; ++x) {
                if (type_ == CV_32F) at<float>(y, x) = static_cast<float>(v);
                else at<uchar>(y, x) = static_cast<uchar>(v);
            }
    }
    int type_ = CV_8U;
    int channels_ = 1;
    std::shared_ptr<uchar> owner_;
};

// cv::Mat_<float>(r, c) << a, b, ... (camera/perspective.cc:21-22): values are converted to float one by one
template<typename T>
class Mat_ : public Mat {
public:
    Mat_(int r, int c) : Mat(r, c, CV_32F) {}
};
template<typename T>
class MatCommaInit_ {
public:
    MatCommaInit_(Mat_<T>& m, T first) : m_(m), n_(0) { put(first); }
    template<typename U>
    MatCommaInit_& operator,(U v) {
        put(static_cast<T>(v));
        return *this;
    }
    operator Mat() const { return m_; }

private:
    void put(T v) {
        m_.template at<T>(n_ / m_.cols, n_ % m_.cols) = v;
        ++n_;
    }
    Mat_<T> m_;
    int n_;
};
template<typename T, typename U>
MatCommaInit_<T> operator<<(Mat_<T> m, U first) {
    return MatCommaInit_<T>(m, static_cast<T>(first));
}

struct TermCriteria {
    enum { COUNT = 1, MAX_ITER = COUNT, EPS = 2 };
    int type, maxCount;
    double epsilon;
    TermCriteria(int t, int c, double e) : type(t), maxCount(c), epsilon(e) {}
};

inline Mat operator*(float s, const Mat& m) {
    Mat out = m.clone();
    for (int y = 0; y < out.rows; ++y)
        for (int x = 0; x < out.cols; ++x) out.at<float>(y, x) = s * m.at<float>(y, x);
    return out;
}
inline Mat operator*(double s, const Mat& m) { return static_cast<float>(s) * m; }
inline Mat operator*(float s, const MatInit& e) { return s * Mat(e); }
inline Mat operator*(double s, const MatInit& e) { return static_cast<float>(s) * Mat(e); }

// cv::norm(a, b, NORM_L1) on CV_32F: absolute differences accumulated in double
inline double norm(const Mat& a, const Mat& b, int /*type*/) {
    double s = 0.0;
    for (int y = 0; y < a.rows; ++y)
        for (int x = 0; x < a.cols; ++x) s += std::fabs(a.at<float>(y, x) - b.at<float>(y, x));
    return s;
}

class _InputArray {
public:
    _InputArray() {}
    _InputArray(const Mat& m) : m_(const_cast<Mat*>(&m)) {}
    bool empty() const { return !m_ || m_->empty(); }
    Mat getMat() const { return m_ ? *m_ : Mat(); }

protected:
    Mat* m_ = nullptr;
};
class _OutputArray : public _InputArray {
public:
    _OutputArray() {}
    _OutputArray(Mat& m) { m_ = &m; }
    void create(int r, int c, int type) const { m_->create(r, c, type); }
    void release() const {
        if (m_) m_->release();
    }
};
typedef const _InputArray& InputArray;
typedef const _OutputArray& OutputArray;
inline const _InputArray& noArray() {
    static _InputArray none;
    return none;
}

inline void resize(const Mat& src, Mat& dst, Size size, double, double, int) {
    Mat out(size.height, size.width, CV_8U);
    cvprim::resize_linear_u8(src.data, src.cols, src.rows, src.step, out.data, out.cols, out.rows, out.step);
    dst = out;
}
inline void GaussianBlur(const Mat& src, Mat& dst, Size, double, double, int) {
    Mat out(src.rows, src.cols, CV_8U);
    cvprim::gaussian7_u8(src.data, src.cols, src.rows, src.step, out.data, out.step);
    dst = out;
}
inline void FAST(const Mat& img, std::vector<KeyPoint>& kps, int thr, bool /*nonmax*/) {
    std::vector<cvprim::FastPoint> pts;
    cvprim::fast9_nms(img.data, img.cols, img.rows, img.step, thr, pts);
    kps.clear();
    for (const auto& p : pts) kps.emplace_back(static_cast<float>(p.x), static_cast<float>(p.y), 7.f, -1.f, static_cast<float>(p.score));
}
inline float fastAtan2(float y, float x) { return cvprim::fast_atan2(y, x); }
inline void rectangle(Mat& img, Point2i a, Point2i b, const Scalar&, int, int) {
    // filled, value 0 (the only use): the LINE_AA fringe is non-zero and only `== 0` is ever tested, see cvprim.h
    cvprim::fill_rect_zero(img.data, img.cols, img.rows, img.step, a.x, a.y, b.x, b.y);
}

}  // namespace cv
