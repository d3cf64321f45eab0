// Stand-in for the OpenCV surface that the reference's include/legacy/mantis/MonteCarlo.cpp touches, written for this
// repository (NOT OpenCV code).  TEST INFRASTRUCTURE ONLY: it lets that file compile unchanged, from where it lies
// under /root/reference, into oracle/_ref/libmontecarlo_ref.so, so that the oracle's restatement of the 2016
// generation's particle weight (oracle/legacy.c: mo_camera_error) is checked against the reference's own
// project2d / colorError / computeCameraError.  What those three functions rely on is modelled faithfully:
//   cv::Mat           8-bit 3-channel images and 32-bit float matrices, rows / cols, at<float>(i, j),
//                     at<Vec3b>(Point): UNCHECKED addressing data + step * y + 3 * x like a release build of OpenCV;
//                     images made by mc_image() carry zeroed guard rows, so a read one row below the image (a
//                     projected y in (rows - 0.5, rows) rounds to rows) returns black -- the oracle's convention
//   Point2f -> Point  saturate_cast<int>(float) = cvRound (lrint: round half to even)
//   Vec3i -> Vec3b    saturate_cast<uchar>
// Everything else (imshow, drawMarker, undistortImage, RNG, inv) only has to compile and link: the wrapper never
// calls the functions that use it.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <memory>
#include <sstream>
#include <string>
#include <vector>

#define CV_8UC3 16
#define CV_32F 5
#define CV_PI 3.1415926535897932384626433832795

namespace cv {

inline int cvRoundD(double v) { return (int)lrint(v); }

template <typename T> struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T a, T b) : x(a), y(b) {}
    template <typename U> Point_(const Point_<U>& o);
};
template <> template <> inline Point_<int>::Point_(const Point_<float>& o) : x(cvRoundD(o.x)), y(cvRoundD(o.y)) {}
typedef Point_<int> Point;
typedef Point_<float> Point2f;

template <typename T> struct Vec3_ {
    T v[3];
    Vec3_() { v[0] = v[1] = v[2] = 0; }
    Vec3_(T a, T b, T c) { v[0] = a; v[1] = b; v[2] = c; }
    template <typename U> Vec3_(const Vec3_<U>& o);
    T& operator[](int i) { return v[i]; }
    const T& operator[](int i) const { return v[i]; }
};
typedef Vec3_<unsigned char> Vec3b;
typedef Vec3_<int> Vec3i;
template <> template <> inline Vec3_<unsigned char>::Vec3_(const Vec3_<int>& o)
{
    for (int k = 0; k < 3; ++k) v[k] = (unsigned char)std::min(255, std::max(0, o.v[k]));
}

struct Scalar {
    double v[4];
    Scalar(double a = 0, double b = 0, double c = 0, double d = 0) { v[0] = a; v[1] = b; v[2] = c; v[3] = d; }
};

class Mat {
public:
    int rows = 0, cols = 0, type_ = 0;
    size_t step = 0;
    unsigned char* data = nullptr;
    std::shared_ptr<std::vector<unsigned char>> buf;

    Mat() {}
    Mat(int r, int c, int type) { create(r, c, type, 0); }
    template <typename T> Mat(const std::vector<T>& v, bool = false)   // (cv::Mat(req...D, CV_32F): only has to compile)
    {
        create((int)v.size(), 1, CV_32F, 0);
        for (size_t i = 0; i < v.size(); ++i) ((float*)data)[i] = (float)v[i];
    }
    void create(int r, int c, int type, int guard_rows)
    {
        rows = r; cols = c; type_ = type;
        const size_t elem = type == CV_8UC3 ? 3 : 4;
        step = elem * (size_t)c;
        buf = std::make_shared<std::vector<unsigned char>>(step * (size_t)(r + guard_rows) + 16, 0);
        data = buf->data();
    }
    template <typename T> T& at(int i, int j) { return ((T*)(data + step * (size_t)i))[j]; }
    template <typename T> const T& at(int i, int j) const { return ((const T*)(data + step * (size_t)i))[j]; }
    template <typename T> T& at(Point p) { return ((T*)(data + step * (size_t)p.y))[p.x]; }   // unchecked, as in OpenCV
    Mat clone() const
    {
        Mat m;
        m.rows = rows; m.cols = cols; m.type_ = type_; m.step = step;
        m.buf = std::make_shared<std::vector<unsigned char>>(*buf);
        m.data = m.buf->data();
        return m;
    }
    Mat inv() const { return clone(); }   // (parseRequest's K_inv: never used by the functions under test)
};
inline std::ostream& operator<<(std::ostream& o, const Mat&) { return o; }

class RNG {   // (the wrapper never draws: generateRandomTransform / generateGaussianTransform are not called)
public:
    uint64_t state;
    RNG(uint64_t s = 0xffffffff) : state(s) {}
    double uniform(double a, double) { return a; }
    float uniform(float a, float) { return a; }
    double gaussian(double) { return 0; }
};

enum { MARKER_SQUARE = 3 };
inline void imshow(const std::string&, const Mat&) {}
inline int waitKey(int = 0) { return -1; }
inline void drawMarker(Mat&, Point2f, const Scalar&, int = 0) {}
namespace fisheye {
inline void undistortImage(const Mat& src, Mat& dst, const Mat&, const Mat&, const Mat&) { dst = src.clone(); }
}

}  // namespace cv

// a BGR image for the wrapper: copies `bgr` (tight rows) and leaves two zeroed guard rows below it
inline cv::Mat mc_image(const unsigned char* bgr, int w, int h, int stride)
{
    cv::Mat m;
    m.create(h, w, CV_8UC3, 2);
    for (int y = 0; y < h; ++y) memcpy(m.data + m.step * (size_t)y, bgr + (size_t)stride * y, 3 * (size_t)w);
    return m;
}
