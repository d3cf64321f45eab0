// Stand-in for the ROS / tf / cv_bridge / boost surface of the reference's include/legacy/mantis/MonteCarlo.cpp,
// written for this repository (NOT ROS code).  TEST INFRASTRUCTURE ONLY (see mc_cv.hpp).  The tf algebra that
// project2d relies on -- tf::Transform * tf::Vector3 = (row_i . v) + origin_i, rows summed x, y, z in that order, all
// in double -- follows tf's LinearMath; the rest only has to compile and link.
#pragma once
#include <array>
#include <cmath>
#include <memory>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

#include "mc_cv.hpp"

#define ROS_DEBUG_STREAM(x) do { } while (0)
#define ROS_DEBUG_STREAM_ONCE(x) do { } while (0)
#define ROS_WARN_STREAM(x) do { } while (0)

namespace boost {
template <class T, std::size_t N> using array = std::array<T, N>;
}

namespace ros {
struct Time { Time(double = 0) {} };
namespace param {
template <class T, class U> bool param(const std::string&, T& var, const U& def) { var = (T)def; return false; }
}
}  // namespace ros

namespace tf {
typedef double tfScalar;
class Vector3 {
public:
    tfScalar m[4];
    Vector3() { m[0] = m[1] = m[2] = m[3] = 0; }
    Vector3(tfScalar x, tfScalar y, tfScalar z) { m[0] = x; m[1] = y; m[2] = z; m[3] = 0; }
    tfScalar x() const { return m[0]; }
    tfScalar y() const { return m[1]; }
    tfScalar z() const { return m[2]; }
    tfScalar& operator[](int i) { return m[i]; }
    const tfScalar& operator[](int i) const { return m[i]; }
    tfScalar dot(const Vector3& v) const { return m[0] * v.m[0] + m[1] * v.m[1] + m[2] * v.m[2]; }
};
class Quaternion {
public:
    tfScalar x_, y_, z_, w_;
    Quaternion() : x_(0), y_(0), z_(0), w_(1) {}
    Quaternion(tfScalar x, tfScalar y, tfScalar z, tfScalar w) : x_(x), y_(y), z_(z), w_(w) {}
};
class Matrix3x3 {
public:
    Vector3 r[3];
    Matrix3x3() { r[0] = Vector3(1, 0, 0); r[1] = Vector3(0, 1, 0); r[2] = Vector3(0, 0, 1); }
    Matrix3x3(const Quaternion& q)
    {
        const tfScalar d = q.x_ * q.x_ + q.y_ * q.y_ + q.z_ * q.z_ + q.w_ * q.w_, s = 2.0 / d;
        const tfScalar xs = q.x_ * s, ys = q.y_ * s, zs = q.z_ * s;
        const tfScalar wx = q.w_ * xs, wy = q.w_ * ys, wz = q.w_ * zs, xx = q.x_ * xs, xy = q.x_ * ys, xz = q.x_ * zs,
                       yy = q.y_ * ys, yz = q.y_ * zs, zz = q.z_ * zs;
        r[0] = Vector3(1.0 - (yy + zz), xy - wz, xz + wy);
        r[1] = Vector3(xy + wz, 1.0 - (xx + zz), yz - wx);
        r[2] = Vector3(xz - wy, yz + wx, 1.0 - (xx + yy));
    }
    const Vector3& operator[](int i) const { return r[i]; }
    void getRPY(tfScalar& roll, tfScalar& pitch, tfScalar& yaw) const   // (only generateGaussianTransform, never called)
    {
        pitch = -asin(r[2].x());
        roll = atan2(r[2].y(), r[2].z());
        yaw = atan2(r[1].x(), r[0].x());
    }
    Matrix3x3 transpose() const
    {
        Matrix3x3 t;
        for (int i = 0; i < 3; ++i) t.r[i] = Vector3(r[0][i], r[1][i], r[2][i]);
        return t;
    }
};
class Transform {
public:
    Matrix3x3 basis;
    Vector3 origin;
    Transform() {}
    Transform(const Quaternion& q, const Vector3& t = Vector3(0, 0, 0)) : basis(q), origin(t) {}
    void setBasis(const Matrix3x3& b) { basis = b; }
    void setOrigin(const Vector3& o) { origin = o; }
    Vector3 operator*(const Vector3& x) const
    {
        return Vector3(basis[0].dot(x) + origin.x(), basis[1].dot(x) + origin.y(), basis[2].dot(x) + origin.z());
    }
    Transform operator*(const Transform& t) const
    {
        Transform o;
        const Matrix3x3 tt = t.basis.transpose();
        for (int i = 0; i < 3; ++i) o.basis.r[i] = Vector3(basis[i].dot(tt[0]), basis[i].dot(tt[1]), basis[i].dot(tt[2]));
        o.origin = (*this) * t.origin;
        return o;
    }
    Transform inverse() const
    {
        Transform o;
        o.basis = basis.transpose();
        const Vector3 n(-origin.x(), -origin.y(), -origin.z());
        o.origin = Vector3(o.basis[0].dot(n), o.basis[1].dot(n), o.basis[2].dot(n));
        return o;
    }
};
class StampedTransform : public Transform {};
class TransformException : public std::runtime_error {
public:
    TransformException(const std::string& s) : std::runtime_error(s) {}
};
class TransformListener {
public:
    void lookupTransform(const std::string&, const std::string&, const ros::Time&, StampedTransform&) const {}
};
}  // namespace tf

namespace std_msgs { struct Header { std::string frame_id; }; }
namespace sensor_msgs {
struct Image { std_msgs::Header header; std::string encoding; };
struct CameraInfo { boost::array<double, 9> K; std::vector<double> D; };
}
namespace geometry_msgs {
struct Vector3 { double x = 0, y = 0, z = 0; };
struct Quaternion { double x = 0, y = 0, z = 0, w = 1; };
}
namespace mantis {
struct mantisServiceRequest {
    std::vector<sensor_msgs::Image> image;
    std::vector<sensor_msgs::CameraInfo> camera_info;
    geometry_msgs::Vector3 delta_pos;
    geometry_msgs::Quaternion delta_quat;
};
}
namespace cv_bridge {
struct CvImage { cv::Mat image; };
inline std::shared_ptr<CvImage> toCvCopy(const sensor_msgs::Image&, const std::string&) { return std::make_shared<CvImage>(); }
}
