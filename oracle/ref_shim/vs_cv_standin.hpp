// TEST INFRASTRUCTURE ONLY — stand-in for the slice of the OpenCV C++ API that the reference's
// feature_tracker.cpp / camera headers use (OpenCV's C++ SDK is not installed; the cv2 Python wheel is).
// Types (Mat, Point_, Size, Scalar, Ptr, TermCriteria) are minimal re-declarations of the public API; the
// six compute entry points (CLAHE::apply, calcOpticalFlowPyrLK, goodFeaturesToTrack, findFundamentalMat,
// circle, imread) forward through vs_ref_cb to the REAL cv2 4.13.0 functions, so no OpenCV arithmetic is
// restated here.  The only API semantics restated are:
//   * Point2f -> Point conversion = saturate_cast<int>(float) = cvRound (round half to even);
//   * Mat copy = shared buffer, clone() = deep copy;
//   * buildOpticalFlowPyramid(img, pyr, win, maxLevel) is represented by {img, maxLevel}: calcOpticalFlowPyrLK
//     on image inputs builds the identical pyramid internally and, like OpenCV with pyramid
//     inputs, uses min(maxLevel, levels of either pyramid).
// Nothing in the product includes this file.
#ifndef VS_CV_STANDIN_HPP
#define VS_CV_STANDIN_HPP

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <memory>
#include <string>
#include <type_traits>
#include <vector>

#include "vs_ref_callbacks.h"

typedef unsigned char uchar;

#define CV_8U 0
#define CV_8UC1 0
#define CV_32F 5
#define CV_64F 6

static inline int cvRound(double v) { return (int)std::lrint(v); }   // default rounding mode: half to even
static inline int cvRound(float v) { return (int)std::lrintf(v); }
static inline int cvRound(int v) { return v; }

namespace cv {

// saturate_cast<int>(float/double) rounds (cvRound); every other combination used here is a plain conversion
template <class U, class T>
inline U saturate_cast(T v) {
    if constexpr (std::is_integral<U>::value && std::is_floating_point<T>::value)
        return (U)cvRound(v);
    else
        return (U)v;
}

template <class T>
struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T x_, T y_) : x(x_), y(y_) {}
    template <class U>
    operator Point_<U>() const { return Point_<U>(saturate_cast<U>(x), saturate_cast<U>(y)); }
};
typedef Point_<int> Point;
typedef Point_<float> Point2f;
typedef Point_<double> Point2d;

template <class T>
struct Point3_ {
    T x, y, z;
    Point3_() : x(0), y(0), z(0) {}
    Point3_(T x_, T y_, T z_) : x(x_), y(y_), z(z_) {}
};
typedef Point3_<float> Point3f;

struct Size {
    int width, height;
    Size() : width(0), height(0) {}
    Size(int w, int h) : width(w), height(h) {}
    bool operator==(const Size& o) const { return width == o.width && height == o.height; }
    bool operator!=(const Size& o) const { return !(*this == o); }
};

struct Scalar {
    double val[4];
    Scalar() : val{0, 0, 0, 0} {}
    Scalar(double v0) : val{v0, 0, 0, 0} {}
    Scalar(double v0, double v1, double v2 = 0, double v3 = 0) : val{v0, v1, v2, v3} {}
};

template <class T>
using Ptr = std::shared_ptr<T>;

struct _OutputArray {};
typedef const _OutputArray& OutputArray;
inline OutputArray noArray() {
    static _OutputArray a;
    return a;
}

class Mat {
public:
    int rows = 0, cols = 0;
    uchar* data = nullptr;
    size_t step = 0;
    int pyr_levels = -1;  // stand-in bookkeeping: maxLevel given to buildOpticalFlowPyramid for this image

    Mat() {}
    Mat(int r, int c, int type, const Scalar& s) { create(r, c, type); fill(s); }
    // wraps caller memory without copying (vio_engine.cpp:180)
    Mat(int r, int c, int type, void* ext, size_t ext_step = 0) : rows(r), cols(c), data((uchar*)ext), type_(type) {
        step = ext_step ? ext_step : size_t(c) * elem();
    }
    void create(int r, int c, int type) {
        rows = r;
        cols = c;
        type_ = type;
        step = size_t(c) * elem();
        buf_ = std::make_shared<std::vector<uchar>>(step * size_t(r));
        data = buf_->data();
    }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    int type() const { return type_; }
    Size size() const { return Size(cols, rows); }
    Mat clone() const {
        Mat m;
        if (empty()) return m;
        m.create(rows, cols, type_);
        for (int y = 0; y < rows; y++) std::memcpy(m.data + y * m.step, data + y * step, size_t(cols) * elem());
        return m;
    }
    Mat rowRange(int a, int b) const {
        Mat m = *this;
        m.data = data + size_t(a) * step;
        m.rows = b - a;
        return m;
    }
    template <class T>
    T& at(Point p) { return *(T*)(data + size_t(p.y) * step + size_t(p.x) * sizeof(T)); }
    template <class T>
    T& at(int r, int c) { return *(T*)(data + size_t(r) * step + size_t(c) * sizeof(T)); }
    template <class T>
    T* ptr(int r = 0) { return (T*)(data + size_t(r) * step); }
    template <class T>
    const T* ptr(int r = 0) const { return (const T*)(data + size_t(r) * step); }
    static Mat eye(int r, int c, int type) {
        Mat m(r, c, type, Scalar(0));
        for (int i = 0; i < std::min(r, c); i++) {
            if (type == CV_32F) m.at<float>(i, i) = 1.f;
            else if (type == CV_64F) m.at<double>(i, i) = 1.0;
            else m.at<uchar>(i, i) = 1;
        }
        return m;
    }

private:
    size_t elem() const { return type_ == CV_64F ? 8 : type_ == CV_32F ? 4 : 1; }
    void fill(const Scalar& s) {
        for (int y = 0; y < rows; y++)
            for (int x = 0; x < cols; x++) {
                if (type_ == CV_64F) at<double>(y, x) = s.val[0];
                else if (type_ == CV_32F) at<float>(y, x) = (float)s.val[0];
                else at<uchar>(y, x) = (uchar)s.val[0];
            }
    }
    int type_ = CV_8UC1;
    std::shared_ptr<std::vector<uchar>> buf_;
};

struct TermCriteria {
    enum Type { COUNT = 1, MAX_ITER = COUNT, EPS = 2 };
    int type, maxCount;
    double epsilon;
    TermCriteria() : type(0), maxCount(0), epsilon(0) {}
    TermCriteria(int t, int c, double e) : type(t), maxCount(c), epsilon(e) {}
};

enum { IMREAD_GRAYSCALE = 0 };
enum { FM_7POINT = 1, FM_8POINT = 2, FM_LMEDS = 4, FM_RANSAC = 8 };

static inline void vs_need(const void* fn) {
    if (!fn) {
        std::cerr << "oracle/_ref: OpenCV callback not set (vsref_set_callbacks)" << std::endl;
        std::abort();
    }
}

class CLAHE {
public:
    CLAHE(double clip, Size tiles) : clip_(clip), tiles_(tiles) {}
    virtual ~CLAHE() {}
    virtual void apply(const Mat& src, Mat& dst) {
        vs_need((const void*)vs_ref_cb.clahe_apply);
        Mat out;
        out.create(src.rows, src.cols, CV_8UC1);
        vs_ref_cb.clahe_apply(clip_, tiles_.width, tiles_.height, src.data, src.rows, src.cols, (long)src.step, out.data);
        dst = out;
    }

private:
    double clip_;
    Size tiles_;
};
inline Ptr<CLAHE> createCLAHE(double clipLimit = 40.0, Size tileGridSize = Size(8, 8)) {
    return std::make_shared<CLAHE>(clipLimit, tileGridSize);
}

inline int buildOpticalFlowPyramid(const Mat& img, std::vector<Mat>& pyramid, Size /*winSize*/, int maxLevel,
                                   bool /*withDerivatives*/ = true) {
    pyramid.clear();
    pyramid.push_back(img.clone());       // OpenCV copies the image into its padded level 0: the pyramid owns its pixels
    pyramid.back().pyr_levels = maxLevel;
    return maxLevel;
}

inline void calcOpticalFlowPyrLK(const std::vector<Mat>& prevPyr, const std::vector<Mat>& nextPyr,
                                 const std::vector<Point2f>& prevPts, std::vector<Point2f>& nextPts,
                                 std::vector<uchar>& status, std::vector<float>& err, Size winSize = Size(21, 21),
                                 int maxLevel = 3, TermCriteria criteria = TermCriteria(3, 30, 0.01)) {
    vs_need((const void*)vs_ref_cb.lk);
    const Mat& a = prevPyr.at(0);
    const Mat& b = nextPyr.at(0);
    int lv = maxLevel;
    if (a.pyr_levels >= 0) lv = std::min(lv, a.pyr_levels);
    if (b.pyr_levels >= 0) lv = std::min(lv, b.pyr_levels);
    const int n = (int)prevPts.size();
    nextPts.assign(n, Point2f());
    status.assign(n, 0);
    err.assign(n, 0.f);
    static_assert(sizeof(Point2f) == 8, "Point2f must be two packed floats");
    vs_ref_cb.lk(a.data, (long)a.step, b.data, (long)b.step, a.rows, a.cols, (const float*)prevPts.data(), n,
                 (float*)nextPts.data(), status.data(), err.data(), winSize.width, winSize.height, lv, criteria.type,
                 criteria.maxCount, criteria.epsilon);
}

inline void goodFeaturesToTrack(const Mat& image, std::vector<Point2f>& corners, int maxCorners, double qualityLevel,
                                double minDistance, const Mat& mask = Mat()) {
    vs_need((const void*)vs_ref_cb.gftt);
    std::vector<float> out(size_t(std::max(maxCorners, 0)) * 2);
    int n = vs_ref_cb.gftt(image.data, (long)image.step, mask.empty() ? nullptr : mask.data, (long)mask.step, image.rows,
                           image.cols, maxCorners, qualityLevel, minDistance, out.data());
    corners.clear();
    for (int i = 0; i < n; i++) corners.push_back(Point2f(out[2 * i], out[2 * i + 1]));
}

inline Mat findFundamentalMat(const std::vector<Point2f>& p1, const std::vector<Point2f>& p2, int method,
                              double ransacReprojThreshold, double confidence, std::vector<uchar>& mask) {
    vs_need((const void*)vs_ref_cb.find_f);
    const int n = (int)p1.size();
    double f9[9];
    std::vector<uchar> st(n, 0);
    int has_status = 0;
    int ok = vs_ref_cb.find_f((const float*)p1.data(), (const float*)p2.data(), n, method, ransacReprojThreshold,
                              confidence, f9, st.data(), &has_status);
    // OpenCV leaves the mask untouched when it bails out early; an all-zero mask of the right length keeps the
    // caller's filterByStatus in bounds (the case does not occur with >= 30 distinct points)
    mask = st;
    Mat F;
    if (ok) {
        F.create(3, 3, CV_64F);
        std::memcpy(F.data, f9, sizeof(f9));
    }
    return F;
}

inline void circle(Mat& img, Point center, int radius, const Scalar& color, int thickness = 1, int /*lineType*/ = 8,
                   int /*shift*/ = 0) {
    vs_need((const void*)vs_ref_cb.circle);
    vs_ref_cb.circle(img.data, (long)img.step, img.rows, img.cols, center.x, center.y, radius, (int)color.val[0], thickness);
}

inline Mat imread(const std::string& path, int /*flags*/ = IMREAD_GRAYSCALE) {
    vs_need((const void*)vs_ref_cb.imread_gray);
    int r = 0, c = 0;
    vs_ref_cb.imread_gray(path.c_str(), nullptr, &r, &c);
    Mat m;
    if (r > 0 && c > 0) {
        m.create(r, c, CV_8UC1);
        vs_ref_cb.imread_gray(path.c_str(), m.data, &r, &c);
    }
    return m;
}

// cv::FileStorage / cv::FileNode, read side only (CameraFactory::generateCameraFromYamlFile and the cameras'
// Parameters::readFromYamlFile): every lookup is answered by the real cv2.FileStorage through vs_ref_cb.yaml_get
class FileNode {
public:
    FileNode() {}
    FileNode(const std::string& path, const std::string& k1, const std::string& k2) : path_(path), k1_(k1), k2_(k2) { fetch(); }
    bool isNone() const { return kind_ <= 0; }
    bool empty() const { return kind_ <= 0; }
    FileNode operator[](const std::string& key) const { return k1_.empty() ? FileNode(path_, key, "") : FileNode(path_, k1_, key); }
    FileNode operator[](const char* key) const { return (*this)[std::string(key)]; }
    explicit operator int() const { return kind_ == 1 ? (int)num_ : kind_ == 2 ? cvRound(num_) : 0; }
    explicit operator double() const { return (kind_ == 1 || kind_ == 2) ? num_ : 0.0; }
    explicit operator float() const { return (float)(double)(*this); }
    const std::string& str() const { return str_; }
    int kind() const { return kind_; }

private:
    void fetch() {
        vs_need((const void*)vs_ref_cb.yaml_get);
        char buf[512];
        buf[0] = 0;
        kind_ = vs_ref_cb.yaml_get(path_.c_str(), k1_.c_str(), k2_.c_str(), &num_, buf, (int)sizeof(buf));
        if (kind_ == 3) str_ = buf;
    }
    std::string path_, k1_, k2_, str_;
    double num_ = 0.0;
    int kind_ = 0;
};
inline void operator>>(const FileNode& n, std::string& out) { out = n.kind() == 3 ? n.str() : std::string(); }

class FileStorage {
public:
    enum Mode { READ = 0, WRITE = 1 };
    FileStorage(const std::string& path, int mode) : path_(path) {
        vs_need((const void*)vs_ref_cb.yaml_get);
        double d;
        char b[8];
        opened_ = mode == READ && vs_ref_cb.yaml_get(path.c_str(), "", "", &d, b, (int)sizeof(b)) >= 0;
    }
    bool isOpened() const { return opened_; }
    FileNode operator[](const char* key) const { return FileNode(path_, key, ""); }
    FileNode operator[](const std::string& key) const { return FileNode(path_, key, ""); }
    void release() {}

private:
    std::string path_;
    bool opened_ = false;
};

}  // namespace cv

#endif
