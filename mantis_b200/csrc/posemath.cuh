// posemath.cuh -- fp64 device math of the pose stages: tf (Bullet LinearMath) algebra as used by
// the reference (Mantis3Types.h:68-136, HypothesisGeneration.h:91-120, PoseAdjustment.h:13-23,
// PoseClusterer.h:198-205), the Kannala-Brandt fisheye model of cv::fisheye::distortPoints /
// undistortPoints (Mantis3Types.h:125-136, QuadDetection.h:289-298), and small 3x3 helpers.
// Compiled with -fmad=false: operation order decides rounding, and the scoring decisions
// downstream (cvRound pixel addressing, greedy `<` selection) are sensitive to it.
#pragma once
#include <math.h>
#include "pipeline.cuh"

namespace pm {

struct M3 { double m[9]; };
struct V3 { double v[3]; };

__host__ __device__ inline M3 mul(const M3& a, const M3& b)
{
    M3 r;
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            double s = 0;
            for (int k = 0; k < 3; ++k) s += a.m[i * 3 + k] * b.m[k * 3 + j];
            r.m[i * 3 + j] = s;
        }
    return r;
}
__host__ __device__ inline M3 transpose(const M3& a)
{
    M3 r;
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) r.m[i * 3 + j] = a.m[j * 3 + i];
    return r;
}
__host__ __device__ inline V3 mulv(const M3& a, const V3& x)
{
    V3 r;
    for (int i = 0; i < 3; ++i) r.v[i] = a.m[i * 3] * x.v[0] + a.m[i * 3 + 1] * x.v[1] + a.m[i * 3 + 2] * x.v[2];
    return r;
}
__host__ __device__ inline M3 eye()
{
    M3 r;
    for (int i = 0; i < 9; ++i) r.m[i] = (i % 4 == 0) ? 1.0 : 0.0;
    return r;
}
__host__ __device__ inline double det(const M3& a)
{
    const double* m = a.m;
    return m[0] * (m[4] * m[8] - m[5] * m[7]) - m[1] * (m[3] * m[8] - m[5] * m[6]) + m[2] * (m[3] * m[7] - m[4] * m[6]);
}
__host__ __device__ inline M3 inverse(const M3& a)
{
    const double* m = a.m;
    double d = det(a);
    M3 r;
    r.m[0] = (m[4] * m[8] - m[5] * m[7]) / d;
    r.m[1] = (m[2] * m[7] - m[1] * m[8]) / d;
    r.m[2] = (m[1] * m[5] - m[2] * m[4]) / d;
    r.m[3] = (m[5] * m[6] - m[3] * m[8]) / d;
    r.m[4] = (m[0] * m[8] - m[2] * m[6]) / d;
    r.m[5] = (m[2] * m[3] - m[0] * m[5]) / d;
    r.m[6] = (m[3] * m[7] - m[4] * m[6]) / d;
    r.m[7] = (m[1] * m[6] - m[0] * m[7]) / d;
    r.m[8] = (m[0] * m[4] - m[1] * m[3]) / d;
    return r;
}

// ---- tf::Transform ----
__host__ __device__ inline void tf_mul(const Tf& a, const Tf& b, Tf& r)
{
    Tf o;
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j)
            o.R[i * 3 + j] = a.R[i * 3] * b.R[j] + a.R[i * 3 + 1] * b.R[3 + j] + a.R[i * 3 + 2] * b.R[6 + j];
    for (int i = 0; i < 3; ++i)
        o.t[i] = (a.R[i * 3] * b.t[0] + a.R[i * 3 + 1] * b.t[1] + a.R[i * 3 + 2] * b.t[2]) + a.t[i];
    r = o;
}
__host__ __device__ inline void tf_inv(const Tf& a, Tf& r)
{
    Tf o;
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) o.R[i * 3 + j] = a.R[j * 3 + i];
    double nt[3] = {-a.t[0], -a.t[1], -a.t[2]};
    for (int i = 0; i < 3; ++i) o.t[i] = o.R[i * 3] * nt[0] + o.R[i * 3 + 1] * nt[1] + o.R[i * 3 + 2] * nt[2];
    r = o;
}
// tf::Matrix3x3::getRotation
__host__ __device__ inline void mat_to_quat(const double* m, double* q)
{
    double trace = m[0] + m[4] + m[8];
    double temp[4];
    if (trace > 0.0) {
        double s = sqrt(trace + 1.0);
        temp[3] = s * 0.5;
        s = 0.5 / s;
        temp[0] = (m[7] - m[5]) * s;
        temp[1] = (m[2] - m[6]) * s;
        temp[2] = (m[3] - m[1]) * s;
    } else {
        int i = m[0] < m[4] ? (m[4] < m[8] ? 2 : 1) : (m[0] < m[8] ? 2 : 0);
        int j = (i + 1) % 3, k = (i + 2) % 3;
        double s = sqrt(m[i * 3 + i] - m[j * 3 + j] - m[k * 3 + k] + 1.0);
        temp[i] = s * 0.5;
        s = 0.5 / s;
        temp[3] = (m[k * 3 + j] - m[j * 3 + k]) * s;
        temp[j] = (m[j * 3 + i] + m[i * 3 + j]) * s;
        temp[k] = (m[k * 3 + i] + m[i * 3 + k]) * s;
    }
    for (int i = 0; i < 4; ++i) q[i] = temp[i];
}
// tf::Matrix3x3::setRotation
__host__ __device__ inline void quat_to_mat(const double* q, double* m)
{
    double d = q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3];
    double s = 2.0 / d;
    double xs = q[0] * s, ys = q[1] * s, zs = q[2] * s;
    double wx = q[3] * xs, wy = q[3] * ys, wz = q[3] * zs;
    double xx = q[0] * xs, xy = q[0] * ys, xz = q[0] * zs;
    double yy = q[1] * ys, yz = q[1] * zs, zz = q[2] * zs;
    m[0] = 1.0 - (yy + zz); m[1] = xy - wz;         m[2] = xz + wy;
    m[3] = xy + wz;         m[4] = 1.0 - (xx + zz); m[5] = yz - wx;
    m[6] = xz - wy;         m[7] = yz + wx;         m[8] = 1.0 - (xx + yy);
}
// tf::Matrix3x3::setRPY
__host__ __device__ inline void set_rpy(double roll, double pitch, double yaw, double* m)
{
    double ci = cos(roll), cj = cos(pitch), ch = cos(yaw);
    double si = sin(roll), sj = sin(pitch), sh = sin(yaw);
    double cc = ci * ch, cs = ci * sh, sc = si * ch, ss = si * sh;
    m[0] = cj * ch; m[1] = sj * sc - cs; m[2] = sj * cc + ss;
    m[3] = cj * sh; m[4] = sj * ss + cc; m[5] = sj * cs - sc;
    m[6] = -sj;     m[7] = cj * si;      m[8] = cj * ci;
}
// tf::Matrix3x3::getRPY, solution 1
__host__ __device__ inline void get_rpy(const double* m, double* rpy)
{
    const double PI = 3.14159265358979323846;
    if (fabs(m[6]) >= 1) {
        rpy[2] = 0;
        rpy[1] = (m[6] < 0) ? PI / 2.0 : -PI / 2.0;
        rpy[0] = atan2(m[7], m[8]);
    } else {
        double pitch = -asin(m[6]);
        rpy[1] = pitch;
        rpy[0] = atan2(m[7] / cos(pitch), m[8] / cos(pitch));
        rpy[2] = atan2(m[3] / cos(pitch), m[0] / cos(pitch));
    }
}
__host__ __device__ inline void hyp_set_c2w(Hyp& h, const Tf& c2w)
{
    h.c2w = c2w;
    tf_inv(h.c2w, h.w2c);
    mat_to_quat(h.w2c.R, h.q);
}
__host__ __device__ inline void hyp_set_w2c(Hyp& h, const Tf& w2c)
{
    h.w2c = w2c;
    tf_inv(h.w2c, h.c2w);
    mat_to_quat(h.w2c.R, h.q);
}

// ---- fisheye (Kannala-Brandt) ----
struct Camera {
    double fx, fy, cx, cy;  // from the CV_32F K of the reference, widened
    double D[4];
};

__host__ __device__ inline bool fisheye_undistort(const Camera& cam, double u, double v, double& xn, double& yn)
{
    const double PI = 3.14159265358979323846, eps = 1e-8;
    double pwx = (u - cam.cx) / cam.fx, pwy = (v - cam.cy) / cam.fy;
    double theta_d = sqrt(pwx * pwx + pwy * pwy);
    theta_d = fmin(fmax(-PI / 2., theta_d), PI / 2.);
    bool converged = false;
    double theta = theta_d, scale = 0.0;
    if (fabs(theta_d) > eps) {
        for (int j = 0; j < 10; ++j) {
            double t2 = theta * theta, t4 = t2 * t2, t6 = t4 * t2, t8 = t6 * t2;
            double k0 = cam.D[0] * t2, k1 = cam.D[1] * t4, k2 = cam.D[2] * t6, k3 = cam.D[3] * t8;
            double fix = (theta * (1 + k0 + k1 + k2 + k3) - theta_d) / (1 + 3 * k0 + 5 * k1 + 7 * k2 + 9 * k3);
            theta = theta - fix;
            if (fabs(fix) < eps) { converged = true; break; }
        }
        scale = tan(theta) / theta_d;
    } else {
        converged = true;
    }
    bool flipped = (theta_d < 0 && theta > 0) || (theta_d > 0 && theta < 0);
    if (converged && !flipped) {
        xn = pwx * scale;
        yn = pwy * scale;
        return true;
    }
    xn = yn = -1000000.0;
    return false;
}

__host__ __device__ inline void fisheye_distort(const Camera& cam, double x, double y, double& u, double& v)
{
    double r2 = x * x + y * y;
    double r = sqrt(r2);
    double theta = atan(r);
    double t2 = theta * theta, t3 = t2 * theta, t4 = t2 * t2, t5 = t4 * theta, t6 = t3 * t3, t7 = t6 * theta,
           t8 = t4 * t4, t9 = t8 * theta;
    double theta_d = theta + cam.D[0] * t3 + cam.D[1] * t5 + cam.D[2] * t7 + cam.D[3] * t9;
    double inv_r = r > 1e-8 ? 1.0 / r : 1;
    double cdist = r > 1e-8 ? theta_d * inv_r : 1;
    double xd = x * cdist, yd = y * cdist;
    u = xd * cam.fx + cam.cx;
    v = yd * cam.fy + cam.cy;
}

}  // namespace pm
