// mantis_host.h -- host-side pieces shared by the two drop-in executables (mantis3, mantis_server):
// what the reference's nodes do around the hot path, without ROS.
//   getParameters() / parseCoordinatesFromString()   include/mantis3/Mantis3Params.h:125-194
//   publishPose() gating and covariance              include/mantis3/PosePub.h:12-61
//   mantisService request / response                 srv/mantisService.srv:1-12
// Everything numerical happens behind include/mantis_cuda.h.
#pragma once
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <fstream>
#include <thread>
#include <sstream>
#include <string>
#include <vector>

#include "../../include/mantis_cuda.h"

namespace mantis_host {

// defines of Mantis3Params.h that act outside the library
constexpr double MINIMUM_YAW_DIFFERENCE = 4000;          // :67
constexpr double PIXEL_ERROR_VARIANCE_COEFF = 1.0 / 600; // :68

struct Vec3 { double x, y, z; };

// "x, y, z; x, y, z; ..." -> points.  Same tokenisation as the reference: split on ';', drop
// newlines and spaces, three comma-separated atof() fields (missing fields read as 0).
inline std::vector<Vec3> parseCoordinatesFromString(const std::string& str)
{
    std::vector<Vec3> fin;
    std::stringstream text(str);
    std::string row;
    while (std::getline(text, row, ';')) {
        row.erase(std::remove(row.begin(), row.end(), '\n'), row.end());
        row.erase(std::remove(row.begin(), row.end(), ' '), row.end());
        std::stringstream rs(row);
        std::string f;
        double v[3] = {0, 0, 0};
        for (int k = 0; k < 3; ++k) {
            f.clear();
            std::getline(rs, f, ',');
            v[k] = atof(f.c_str());
        }
        fin.push_back({v[0], v[1], v[2]});
    }
    return fin;
}

struct Map {
    std::vector<Vec3> white, red, green;
};

// params/map.yaml: three keys (whiteMap, redMap, greenMap) each holding one quoted multi-line string.
inline bool loadMapYaml(const std::string& path, Map& m)
{
    std::ifstream in(path);
    if (!in) return false;
    std::stringstream ss;
    ss << in.rdbuf();
    const std::string all = ss.str();
    auto value_of = [&](const char* key) -> std::string {
        size_t k = all.find(std::string(key) + ":");
        if (k == std::string::npos) return "";
        size_t a = all.find('\'', k);
        if (a == std::string::npos) return "";
        size_t b = all.find('\'', a + 1);
        if (b == std::string::npos) return "";
        return all.substr(a + 1, b - a - 1);
    };
    m.white = parseCoordinatesFromString(value_of("whiteMap"));
    m.red = parseCoordinatesFromString(value_of("redMap"));
    m.green = parseCoordinatesFromString(value_of("greenMap"));
    // a trailing ';' in the yaml string yields one empty row = (0, 0, 0), exactly as in the reference
    return !m.white.empty();
}

inline int applyMap(mantis_ctx* ctx, const Map& m)
{
    std::vector<double> xyz;
    for (const std::vector<Vec3>* v : {&m.white, &m.red, &m.green})
        for (const Vec3& p : *v) { xyz.push_back(p.x); xyz.push_back(p.y); xyz.push_back(p.z); }
    return mantis_set_map(ctx, xyz.data(), (int)m.white.size(), (int)m.red.size(), (int)m.green.size());
}

// sensor_msgs/Image (the fields the reference reads) with owned storage
struct Image {
    int width = 0, height = 0, step = 0;
    std::string encoding = "bgr8", frame_id;
    uint64_t stamp_ns = 0;
    std::vector<uint8_t> data;
};
// sensor_msgs/CameraInfo: K row-major doubles, D fisheye coefficients
struct CameraInfo {
    double K[9] = {0, 0, 0, 0, 0, 0, 0, 0, 1};
    std::vector<double> D;
};

// Binary PPM (P6, maxval 255) or raw BGR ("file.bgr:WxH").  PPM is RGB: swapped to BGR on load.
inline bool loadImage(const std::string& spec, Image& img)
{
    size_t colon = spec.rfind(':');
    if (colon != std::string::npos && spec.find(".bgr") != std::string::npos) {
        int w = 0, h = 0;
        if (sscanf(spec.c_str() + colon + 1, "%dx%d", &w, &h) != 2 || w <= 0 || h <= 0) return false;
        std::ifstream in(spec.substr(0, colon), std::ios::binary);
        if (!in) return false;
        img.width = w; img.height = h; img.step = 3 * w;
        img.data.resize((size_t)3 * w * h);
        in.read((char*)img.data.data(), (std::streamsize)img.data.size());
        return (size_t)in.gcount() == img.data.size();
    }
    std::ifstream in(spec, std::ios::binary);
    if (!in) return false;
    std::string magic;
    in >> magic;
    if (magic != "P6") return false;
    int vals[3], got = 0;
    while (got < 3) {
        int c = in.peek();
        if (c == '#') { std::string skip; std::getline(in, skip); continue; }
        if (isspace(c)) { in.get(); continue; }
        if (!(in >> vals[got])) return false;
        ++got;
    }
    in.get();  // single whitespace after maxval
    if (vals[2] != 255) return false;
    img.width = vals[0]; img.height = vals[1]; img.step = 3 * vals[0];
    img.data.resize((size_t)img.step * img.height);
    in.read((char*)img.data.data(), (std::streamsize)img.data.size());
    if ((size_t)in.gcount() != img.data.size()) return false;
    for (size_t i = 0; i + 2 < img.data.size(); i += 3) std::swap(img.data[i], img.data[i + 2]);
    return true;
}

inline mantis_frame toFrame(const Image& img, const CameraInfo& ci, const double* cam_to_base7 = nullptr)
{
    mantis_frame f;
    memset(&f, 0, sizeof(f));
    f.bgr = img.data.data();
    f.width = img.width; f.height = img.height; f.stride_bytes = img.step;
    for (int i = 0; i < 9; ++i) f.K[i] = (float)ci.K[i];                  // get3x3FromVector, QuadDetection.h:189-201
    for (int i = 0; i < 4; ++i) f.D[i] = i < (int)ci.D.size() ? ci.D[i] : 0.0;
    if (cam_to_base7) for (int i = 0; i < 7; ++i) f.cam_to_base[i] = cam_to_base7[i];
    f.stamp_ns = img.stamp_ns;
    return f;
}

// geometry_msgs/PoseWithCovarianceStamped as publishPose() fills it
struct PoseEstimate {
    bool published = false;   // false: gated out (min_yaw_diff <= 4000) or the frame had no result
    double position[3] = {0, 0, 0}, orientation_xyzw[4] = {0, 0, 0, 1};
    double variance = 0;      // all six diagonal entries of the covariance
    uint64_t stamp_ns = 0;
};

inline PoseEstimate publishPose(const mantis_result& r, uint64_t stamp_ns)
{
    PoseEstimate p;
    p.stamp_ns = stamp_ns;
    if (r.status != MANTIS_FRAME_OK || !(r.min_yaw_diff > MINIMUM_YAW_DIFFERENCE)) return p;
    p.published = true;
    for (int k = 0; k < 3; ++k) p.position[k] = r.pos[k];
    for (int k = 0; k < 4; ++k) p.orientation_xyzw[k] = r.quat_xyzw[k];
    p.variance = r.error * PIXEL_ERROR_VARIANCE_COEFF;
    return p;
}

// pose (x y z qx qy qz qw) algebra for the rig: world<-base = (world<-cam) * (cam<-base)
inline void quatMul(const double* a, const double* b, double* o)
{
    o[0] = a[3] * b[0] + a[0] * b[3] + a[1] * b[2] - a[2] * b[1];
    o[1] = a[3] * b[1] - a[0] * b[2] + a[1] * b[3] + a[2] * b[0];
    o[2] = a[3] * b[2] + a[0] * b[1] - a[1] * b[0] + a[2] * b[3];
    o[3] = a[3] * b[3] - a[0] * b[0] - a[1] * b[1] - a[2] * b[2];
}
inline void quatRotate(const double* q, const double* v, double* o)
{
    const double p[4] = {v[0], v[1], v[2], 0}, qc[4] = {-q[0], -q[1], -q[2], q[3]};
    double t[4], r[4];
    quatMul(q, p, t);
    quatMul(t, qc, r);
    o[0] = r[0]; o[1] = r[1]; o[2] = r[2];
}
// out = a * b, each (x y z qx qy qz qw)
inline void poseCompose(const double* a, const double* b, double* out)
{
    double rt[3];
    quatRotate(a + 3, b, rt);
    for (int k = 0; k < 3; ++k) out[k] = a[k] + rt[k];
    quatMul(a + 3, b + 3, out + 3);
}
inline void poseInverse(const double* a, double* out)
{
    const double qc[4] = {-a[3], -a[4], -a[5], a[6]};
    double t[3];
    quatRotate(qc, a, t);
    for (int k = 0; k < 3; ++k) out[k] = -t[k];
    for (int k = 0; k < 4; ++k) out[3 + k] = qc[k];
}

inline bool parseDoubles(const char* s, double* out, int n)
{
    for (int i = 0; i < n; ++i) {
        char* end = nullptr;
        out[i] = strtod(s, &end);
        if (end == s) return false;
        s = end;
        while (*s == ',' || *s == ' ') ++s;
    }
    return true;
}


// "0-7", "0,2,5", "3" -> device indices
inline std::vector<int> parseDevices(const std::string& spec)
{
    std::vector<int> d;
    std::stringstream ss(spec);
    std::string part;
    while (std::getline(ss, part, ',')) {
        const size_t dash = part.find('-');
        if (dash != std::string::npos && dash > 0) {
            const int a = atoi(part.substr(0, dash).c_str()), b = atoi(part.substr(dash + 1).c_str());
            for (int k = a; k <= b; ++k) d.push_back(k);
        } else if (!part.empty()) {
            d.push_back(atoi(part.c_str()));
        }
    }
    return d;
}

// Several GPUs of one box behind one call: one mantis_ctx and one host thread per device, the
// batch cut into contiguous shards -- device g of G takes units [g n / G, (g + 1) n / G) -- and every shard's
// records written straight into the caller's array at the shard's offset.  Frames are independent, so there is no
// exchange between devices: no collective, no peer copies.  A unit is a frame, or a whole rig (rigs are never
// split).  The reference serves one frame at a time on one thread (src/mantis_server.cpp:23-38).
class DevicePool {
public:
    DevicePool() = default;
    DevicePool(const DevicePool&) = delete;
    DevicePool& operator=(const DevicePool&) = delete;
    ~DevicePool() { close(); }

    // 0 on success, else the mantis_status of the first context that could not be created
    int open(const std::vector<int>& devices, const mantis_config* cfg, const Map* map = nullptr)
    {
        close();
        for (int dev : devices) {
            mantis_ctx* c = nullptr;
            int rc = mantis_create(cfg, dev, &c);
            if (rc == 0 && map) rc = applyMap(c, *map);
            if (rc) { if (c) mantis_destroy(c); close(); return rc; }
            ctx_.push_back(c);
            dev_.push_back(dev);
        }
        return ctx_.empty() ? (int)MANTIS_ERR_INVALID : 0;
    }
    void close()
    {
        for (mantis_ctx* c : ctx_) mantis_destroy(c);
        ctx_.clear();
        dev_.clear();
    }
    int size() const { return (int)ctx_.size(); }
    mantis_ctx* ctx(int g) const { return ctx_[g]; }
    const std::string& error() const { return err_; }

    // frames: n_units * cams_per_unit records (cams_per_unit == 0: plain frames, mantis_localize_batch);
    // out (and trace, optional): n_units records.  Host or device frames as `mem` says -- device frames must
    // live on the device their shard goes to, so multi-device callers normally pass host frames.
    int localize(const mantis_frame* frames, int n_units, int cams_per_unit, mantis_result* out, mantis_frame_trace* trace,
                 int mem)
    {
        const int G = size();
        if (G == 0) return MANTIS_ERR_INVALID;
        const int C = cams_per_unit > 0 ? cams_per_unit : 1;
        std::vector<int> rc((size_t)G, 0);
        auto shard = [&](int g) {
            const int u0 = (int)((long long)n_units * g / G), u1 = (int)((long long)n_units * (g + 1) / G);
            if (u1 <= u0) return;
            const mantis_frame* f = frames + (size_t)u0 * C;
            mantis_frame_trace* t = trace ? trace + u0 : nullptr;
            rc[g] = cams_per_unit > 0 ? mantis_localize_rig_batch(ctx_[g], f, u1 - u0, C, out + u0, t, mem)
                                      : mantis_localize_batch(ctx_[g], f, u1 - u0, out + u0, t, mem);
        };
        std::vector<std::thread> th;
        for (int g = 1; g < G; ++g) th.emplace_back(shard, g);
        shard(0);
        for (std::thread& t : th) t.join();
        for (int g = 0; g < G; ++g)
            if (rc[g]) {
                err_ = "device " + std::to_string(dev_[g]) + ": " + mantis_strerror(rc[g]) + " [" + mantis_last_error(ctx_[g]) + "]";
                return rc[g];
            }
        return MANTIS_OK;
    }

private:
    std::vector<mantis_ctx*> ctx_;
    std::vector<int> dev_;
    std::string err_;
};

}  // namespace mantis_host
