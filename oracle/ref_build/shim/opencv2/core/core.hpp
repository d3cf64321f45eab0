// Minimal stand-in for <opencv2/core/core.hpp>, written for this repository (NOT OpenCV code).
//
// TEST INFRASTRUCTURE ONLY.  It lets the reference's own RobustPlanarPose/RPP.cpp compile
// unchanged, from where it lies under /root/reference, into oracle/_ref/librpp_ref.so, so
// that the oracle's restatement of RPP can be validated against the real reference code.
// Only the surface RPP.cpp touches is provided: a reference-counted CV_64F 2-D cv::Mat with
// at<double>, t(), inv(), clone(), create(), zeros/eye/ones, + - * / operators; cv::SVD of
// small matrices (one-sided Jacobi, singular values descending, like OpenCV);
// cv::determinant; cv::Vec3d; cv::Point2d/Point3d.
#pragma once
#include <cassert>
#include <cmath>
#include <cstring>
#include <memory>
#include <vector>

#define CV_64F 6

namespace cv {

template <typename T> struct Point_ { T x, y; Point_() : x(0), y(0) {} Point_(T a, T b) : x(a), y(b) {} };
template <typename T> struct Point3_ { T x, y, z; Point3_() : x(0), y(0), z(0) {} Point3_(T a, T b, T c) : x(a), y(b), z(c) {} };
typedef Point_<double> Point2d;
typedef Point3_<double> Point3d;

struct Vec3d {
    double v[3];
    Vec3d() { v[0] = v[1] = v[2] = 0; }
    Vec3d(double a, double b, double c) { v[0] = a; v[1] = b; v[2] = c; }
    double& operator[](int i) { return v[i]; }
    const double& operator[](int i) const { return v[i]; }
    double dot(const Vec3d& o) const { return v[0] * o.v[0] + v[1] * o.v[1] + v[2] * o.v[2]; }
    Vec3d cross(const Vec3d& o) const
    {
        return Vec3d(v[1] * o.v[2] - v[2] * o.v[1], v[2] * o.v[0] - v[0] * o.v[2], v[0] * o.v[1] - v[1] * o.v[0]);
    }
    Vec3d operator-() const { return Vec3d(-v[0], -v[1], -v[2]); }
};

class Mat {
public:
    int rows, cols;
    double* data;  // non-null iff allocated (RPP.cpp tests `.data`)
    std::shared_ptr<std::vector<double>> buf;

    Mat() : rows(0), cols(0), data(nullptr) {}
    Mat(int r, int c, int /*type*/) { alloc(r, c); }
    void create(int r, int c, int /*type*/)
    {
        if (data && r == rows && c == cols) return;
        alloc(r, c);
    }
    void alloc(int r, int c)
    {
        rows = r; cols = c;
        buf = std::make_shared<std::vector<double>>((size_t)r * c, 0.0);
        data = buf->data();
    }
    template <typename T> T& at(int i, int j) { return data[(size_t)i * cols + j]; }
    template <typename T> const T& at(int i, int j) const { return data[(size_t)i * cols + j]; }
    template <typename T> T& at(int i) { return data[i]; }
    template <typename T> const T& at(int i) const { return data[i]; }
    Mat clone() const
    {
        Mat m;
        if (data) { m.alloc(rows, cols); std::memcpy(m.data, data, sizeof(double) * rows * cols); }
        return m;
    }
    Mat t() const
    {
        Mat m(cols, rows, CV_64F);
        for (int i = 0; i < rows; ++i)
            for (int j = 0; j < cols; ++j) m.data[(size_t)j * rows + i] = data[(size_t)i * cols + j];
        return m;
    }
    Mat inv() const;  // square, via Gauss-Jordan with partial pivoting
    static Mat zeros(int r, int c, int t) { return Mat(r, c, t); }
    static Mat ones(int r, int c, int t)
    {
        Mat m(r, c, t);
        for (int i = 0; i < r * c; ++i) m.data[i] = 1.0;
        return m;
    }
    static Mat eye(int r, int c, int t)
    {
        Mat m(r, c, t);
        for (int i = 0; i < (r < c ? r : c); ++i) m.data[(size_t)i * c + i] = 1.0;
        return m;
    }
    Mat& operator+=(const Mat& o)
    {
        assert(rows == o.rows && cols == o.cols);
        for (int i = 0; i < rows * cols; ++i) data[i] += o.data[i];
        return *this;
    }
};

inline Mat operator+(const Mat& a, const Mat& b)
{
    assert(a.rows == b.rows && a.cols == b.cols);
    Mat m(a.rows, a.cols, CV_64F);
    for (int i = 0; i < a.rows * a.cols; ++i) m.data[i] = a.data[i] + b.data[i];
    return m;
}
inline Mat operator-(const Mat& a, const Mat& b)
{
    assert(a.rows == b.rows && a.cols == b.cols);
    Mat m(a.rows, a.cols, CV_64F);
    for (int i = 0; i < a.rows * a.cols; ++i) m.data[i] = a.data[i] - b.data[i];
    return m;
}
inline Mat operator-(const Mat& a)
{
    Mat m(a.rows, a.cols, CV_64F);
    for (int i = 0; i < a.rows * a.cols; ++i) m.data[i] = -a.data[i];
    return m;
}
inline Mat operator*(const Mat& a, const Mat& b)
{
    assert(a.cols == b.rows);
    Mat m(a.rows, b.cols, CV_64F);
    for (int i = 0; i < a.rows; ++i)
        for (int j = 0; j < b.cols; ++j) {
            double s = 0;
            for (int k = 0; k < a.cols; ++k) s += a.data[(size_t)i * a.cols + k] * b.data[(size_t)k * b.cols + j];
            m.data[(size_t)i * b.cols + j] = s;
        }
    return m;
}
inline Mat operator*(const Mat& a, double s)
{
    Mat m(a.rows, a.cols, CV_64F);
    for (int i = 0; i < a.rows * a.cols; ++i) m.data[i] = a.data[i] * s;
    return m;
}
inline Mat operator*(double s, const Mat& a) { return a * s; }
inline Mat operator/(const Mat& a, double s)
{
    Mat m(a.rows, a.cols, CV_64F);
    for (int i = 0; i < a.rows * a.cols; ++i) m.data[i] = a.data[i] / s;
    return m;
}

inline Mat Mat::inv() const
{
    assert(rows == cols);
    int n = rows;
    std::vector<double> a(data, data + (size_t)n * n);
    Mat r = Mat::eye(n, n, CV_64F);
    for (int c = 0; c < n; ++c) {
        int p = c;
        for (int i = c + 1; i < n; ++i)
            if (std::fabs(a[(size_t)i * n + c]) > std::fabs(a[(size_t)p * n + c])) p = i;
        if (a[(size_t)p * n + c] == 0.0) return Mat::zeros(n, n, CV_64F);
        if (p != c)
            for (int j = 0; j < n; ++j) {
                std::swap(a[(size_t)p * n + j], a[(size_t)c * n + j]);
                std::swap(r.data[(size_t)p * n + j], r.data[(size_t)c * n + j]);
            }
        double d = a[(size_t)c * n + c];
        for (int j = 0; j < n; ++j) { a[(size_t)c * n + j] /= d; r.data[(size_t)c * n + j] /= d; }
        for (int i = 0; i < n; ++i) {
            if (i == c) continue;
            double f = a[(size_t)i * n + c];
            if (f == 0.0) continue;
            for (int j = 0; j < n; ++j) {
                a[(size_t)i * n + j] -= f * a[(size_t)c * n + j];
                r.data[(size_t)i * n + j] -= f * r.data[(size_t)c * n + j];
            }
        }
    }
    return r;
}

inline double determinant(const Mat& m)
{
    assert(m.rows == 3 && m.cols == 3);
    const double* a = m.data;
    return a[0] * (a[4] * a[8] - a[5] * a[7]) - a[1] * (a[3] * a[8] - a[5] * a[6]) + a[2] * (a[3] * a[7] - a[4] * a[6]);
}

// A = u * diag(w) * vt, singular values descending.  One-sided (Hestenes) Jacobi on A^T.
class SVD {
public:
    Mat u, w, vt;
    explicit SVD(const Mat& A)
    {
        int m = A.rows, n = A.cols;
        assert(m >= n);
        // work on columns of A: rotate pairs until orthogonal
        std::vector<double> a(A.data, A.data + (size_t)m * n), v((size_t)n * n, 0.0);
        for (int i = 0; i < n; ++i) v[(size_t)i * n + i] = 1.0;
        for (int sweep = 0; sweep < 60; ++sweep) {
            bool changed = false;
            for (int p = 0; p < n - 1; ++p)
                for (int q = p + 1; q < n; ++q) {
                    double alpha = 0, beta = 0, gamma = 0;
                    for (int k = 0; k < m; ++k) {
                        double ap = a[(size_t)k * n + p], aq = a[(size_t)k * n + q];
                        alpha += ap * ap; beta += aq * aq; gamma += ap * aq;
                    }
                    if (std::fabs(gamma) <= 1e-300 || std::fabs(gamma) <= 2.220446049250313e-16 * std::sqrt(alpha * beta)) continue;
                    changed = true;
                    double zeta = (beta - alpha) / (2.0 * gamma);
                    double t = (zeta >= 0 ? 1.0 : -1.0) / (std::fabs(zeta) + std::sqrt(1.0 + zeta * zeta));
                    double c = 1.0 / std::sqrt(1.0 + t * t), s = c * t;
                    for (int k = 0; k < m; ++k) {
                        double ap = a[(size_t)k * n + p], aq = a[(size_t)k * n + q];
                        a[(size_t)k * n + p] = c * ap - s * aq;
                        a[(size_t)k * n + q] = s * ap + c * aq;
                    }
                    for (int k = 0; k < n; ++k) {
                        double vp = v[(size_t)k * n + p], vq = v[(size_t)k * n + q];
                        v[(size_t)k * n + p] = c * vp - s * vq;
                        v[(size_t)k * n + q] = s * vp + c * vq;
                    }
                }
            if (!changed) break;
        }
        std::vector<double> sv(n);
        std::vector<int> ord(n);
        for (int j = 0; j < n; ++j) {
            double s = 0;
            for (int k = 0; k < m; ++k) s += a[(size_t)k * n + j] * a[(size_t)k * n + j];
            sv[j] = std::sqrt(s);
            ord[j] = j;
        }
        for (int i = 0; i < n; ++i)
            for (int j = i + 1; j < n; ++j)
                if (sv[ord[j]] > sv[ord[i]]) std::swap(ord[i], ord[j]);
        u = Mat(m, n, CV_64F); w = Mat(n, 1, CV_64F); vt = Mat(n, n, CV_64F);
        std::vector<int> null_col(n, 0);
        for (int jj = 0; jj < n; ++jj) {
            int j = ord[jj];
            w.data[jj] = sv[j];
            for (int k = 0; k < n; ++k) vt.data[(size_t)jj * n + k] = v[(size_t)k * n + j];
            if (sv[j] > 1e-12 * sv[ord[0]] && sv[j] > 1e-300)
                for (int k = 0; k < m; ++k) u.data[(size_t)k * n + jj] = a[(size_t)k * n + j] / sv[j];
            else
                w.data[jj] = 0.0 * (null_col[jj] = 1);
        }
        // complete u for zero singular values (3x3 only): cross product of the other two columns
        if (m == 3 && n == 3) {
            for (int jj = 0; jj < 3; ++jj)
                if (null_col[jj]) {
                    int a1 = (jj + 1) % 3, a2 = (jj + 2) % 3;
                    double x1[3], x2[3];
                    for (int k = 0; k < 3; ++k) { x1[k] = u.data[k * 3 + a1]; x2[k] = u.data[k * 3 + a2]; }
                    u.data[0 * 3 + jj] = x1[1] * x2[2] - x1[2] * x2[1];
                    u.data[1 * 3 + jj] = x1[2] * x2[0] - x1[0] * x2[2];
                    u.data[2 * 3 + jj] = x1[0] * x2[1] - x1[1] * x2[0];
                }
        }
    }
};

}  // namespace cv
