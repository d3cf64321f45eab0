// mantis_server -- drop-in for the reference's `mantis_server` (src/mantis_server.cpp): the service
// handler runMantis(Request&, Response&) (:23-36) over the request/response of
// srv/mantisService.srv:1-12, without ROS.  The reference's handler parses the request and never
// fills the response (:27-29); this one localizes every image of the request in ONE batched
// mantis_localize_batch() call and fills pose / weight / num_particles.
//
// Requests arrive as text on stdin (one service call per block):
//     request
//     image <file.ppm | file.bgr:WxH>          sensor_msgs/Image  (BGR8)
//     K k0 ... k8                              sensor_msgs/CameraInfo.K of that image
//     D d0 d1 d2 d3                            sensor_msgs/CameraInfo.D
//     cam_to_base x y z qx qy qz qw            optional: pose of this camera in base_link
//     ... more image / K / D / cam_to_base groups (a multi-camera rig) ...
//     delta_pos x y z                          parsed and unused, as in MonteCarlo.cpp:274-276
//     delta_quat x y z w
//     end
// One response line per request on stdout.
#include <iostream>

#include "mantis_host.h"

using namespace mantis_host;

struct Request {   // mantisService.srv:1-7
    std::vector<Image> image;
    std::vector<CameraInfo> camera_info;
    std::vector<std::vector<double>> cam_to_base;  // per image; empty = camera frame is the base frame
    double delta_pos[3] = {0, 0, 0}, delta_quat[4] = {0, 0, 0, 1};
};
struct Response {  // mantisService.srv:9-12
    double pose[7] = {0, 0, 0, 0, 0, 0, 1};  // position xyz + orientation xyzw (of base_link in the world)
    double weight = 0;
    int32_t num_particles = 0;
    bool valid = false;
    int best_camera = -1;
};

static mantis_ctx* g_ctx;

static bool g_per_camera = false;

// The service handler.  Returns false on an API failure (the ROS service call would fail).
static bool runMantis(const Request& req, Response& res)
{
    const int n = (int)req.image.size();
    if (n == 0 || req.camera_info.size() != req.image.size()) return false;
    std::vector<mantis_frame> frames;
    for (int i = 0; i < n; ++i)
        frames.push_back(toFrame(req.image[i], req.camera_info[i], req.cam_to_base[i].size() == 7 ? req.cam_to_base[i].data() : nullptr));
    std::vector<mantis_result> out((size_t)n);
    bool same = true, rig = n > 1;
    for (const mantis_frame& f : frames) same = same && f.width == frames[0].width && f.height == frames[0].height && f.stride_bytes == frames[0].stride_bytes;
    for (int i = 0; i < n; ++i) rig = rig && req.cam_to_base[i].size() == 7 && req.cam_to_base[i][6] * req.cam_to_base[i][6] + req.cam_to_base[i][3] * req.cam_to_base[i][3] + req.cam_to_base[i][4] * req.cam_to_base[i][4] + req.cam_to_base[i][5] * req.cam_to_base[i][5] > 0;
    res = Response();
    if (rig && same && n <= MANTIS_MAX_RIG_CAMERAS && !g_per_camera) {
        // Every camera comes with its base->camera transform: the request is one rig, localized JOINTLY -- hypotheses
        // of all cameras pooled in the base frame, every score summed over the cameras (what the reference's server
        // states as its intent: MonteCarlo.cpp:233-241, Mantis2Error.h:55-99).  The pose is the base's.
        mantis_result r;
        if (mantis_localize_rig_batch(g_ctx, frames.data(), 1, n, &r, nullptr, MANTIS_MEM_HOST)) return false;
        res.num_particles = r.n_particles;
        PoseEstimate p = publishPose(r, req.image[0].stamp_ns);
        if (p.published) {
            for (int k = 0; k < 3; ++k) res.pose[k] = p.position[k];
            for (int k = 0; k < 4; ++k) res.pose[3 + k] = p.orientation_xyzw[k];
            res.weight = r.error;
            res.valid = true;
            res.best_camera = -1;   // joint
        }
        return true;
    }
    if (same) {
        if (mantis_localize_batch(g_ctx, frames.data(), n, out.data(), nullptr, MANTIS_MEM_HOST)) return false;
    } else {  // one call per resolution group; here simply one per frame
        for (int i = 0; i < n; ++i)
            if (mantis_localize_batch(g_ctx, &frames[i], 1, &out[i], nullptr, MANTIS_MEM_HOST)) return false;
    }
    // Rig fusion: every camera that passes publishPose()'s gate votes with its pose moved into the
    // base frame (world<-base = world<-cam * (base<-cam)^-1); the lowest colour error wins.
    // (The reference computes `hyp.getW2C() * c2b` and then publishes the camera pose, PosePub.h:28-44.)
    for (int i = 0; i < n; ++i) {
        res.num_particles += out[i].n_particles;
        PoseEstimate p = publishPose(out[i], req.image[i].stamp_ns);
        if (!p.published) continue;
        if (res.valid && !(out[i].error < res.weight)) continue;
        double w2c[7] = {p.position[0], p.position[1], p.position[2], p.orientation_xyzw[0], p.orientation_xyzw[1],
                         p.orientation_xyzw[2], p.orientation_xyzw[3]};
        if (req.cam_to_base[i].size() == 7 && req.cam_to_base[i][6] != 0.0) {
            double b2c[7];
            poseInverse(req.cam_to_base[i].data(), b2c);
            poseCompose(w2c, b2c, res.pose);
        } else {
            for (int k = 0; k < 7; ++k) res.pose[k] = w2c[k];
        }
        res.weight = out[i].error;
        res.valid = true;
        res.best_camera = i;
    }
    return true;
}

int main(int argc, char** argv)
{
    mantis_config cfg;
    mantis_default_config(&cfg);
    int device = 0;
    std::string map_path;
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        if (a == "--device" && i + 1 < argc) device = atoi(argv[++i]);
        else if (a == "--map" && i + 1 < argc) map_path = argv[++i];
        else if (a == "--per-camera") g_per_camera = true;   // rigs: localize every camera alone, best one wins (no joint refinement)
        else { fprintf(stderr, "usage: mantis_server [--device N] [--map map.yaml] [--per-camera] < requests\n"); return 2; }
    }
    int rc = mantis_create(&cfg, device, &g_ctx);
    if (rc) { fprintf(stderr, "mantis_create: %s\n", mantis_strerror(rc)); return 1; }
    if (!map_path.empty()) {
        Map m;
        if (!loadMapYaml(map_path, m) || applyMap(g_ctx, m)) { fprintf(stderr, "cannot load map %s\n", map_path.c_str()); return 1; }
    }

    std::string line;
    Request req;
    bool open = false;
    int served = 0;
    while (std::getline(std::cin, line)) {
        std::stringstream ls(line);
        std::string key;
        if (!(ls >> key) || key[0] == '#') continue;
        if (key == "request") { req = Request(); open = true; continue; }
        if (!open) continue;
        std::string rest;
        std::getline(ls, rest);
        if (key == "image") {
            size_t a = rest.find_first_not_of(" \t");
            Image im;
            if (a == std::string::npos || !loadImage(rest.substr(a), im)) { fprintf(stderr, "cannot read image%s\n", rest.c_str()); return 1; }
            im.stamp_ns = (uint64_t)served;
            req.image.push_back(im);
            req.camera_info.push_back(CameraInfo());
            req.camera_info.back().D.assign(4, 0.0);
            req.cam_to_base.push_back({});
        } else if (key == "K" && !req.camera_info.empty()) {
            parseDoubles(rest.c_str(), req.camera_info.back().K, 9);
        } else if (key == "D" && !req.camera_info.empty()) {
            parseDoubles(rest.c_str(), req.camera_info.back().D.data(), 4);
        } else if (key == "cam_to_base" && !req.cam_to_base.empty()) {
            req.cam_to_base.back().assign(7, 0.0);
            parseDoubles(rest.c_str(), req.cam_to_base.back().data(), 7);
        } else if (key == "delta_pos") {
            parseDoubles(rest.c_str(), req.delta_pos, 3);
        } else if (key == "delta_quat") {
            parseDoubles(rest.c_str(), req.delta_quat, 4);
        } else if (key == "end") {
            Response res;
            const bool ok = runMantis(req, res);
            printf("response %d ok=%d valid=%d best_camera=%d pose=(%.9f %.9f %.9f %.9f %.9f %.9f %.9f) weight=%.9g num_particles=%d\n",
                   served, (int)ok, (int)res.valid, res.best_camera, res.pose[0], res.pose[1], res.pose[2], res.pose[3], res.pose[4],
                   res.pose[5], res.pose[6], res.weight, res.num_particles);
            fflush(stdout);
            if (!ok) fprintf(stderr, "runMantis: %s\n", mantis_last_error(g_ctx));
            ++served;
            open = false;
        }
    }
    mantis_destroy(g_ctx);
    return 0;
}
