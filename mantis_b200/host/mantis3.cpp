// mantis3 -- drop-in for the reference's `mantis3` node (src/mantis3.cpp) without ROS: the topic
// callback quadDetection(image, camera_info) (:68-135) becomes one mantis_localize_batch() call per
// frame (or one call for all frames with --batch), publishPose() (PosePub.h:12-61) becomes a line on
// stdout.  main() mirrors src/mantis3.cpp:138-158: parameters -> context -> frames.
//
//   mantis3 --K fx,0,cx,0,fy,cy,0,0,1 [--D k1,k2,k3,k4] [--map params/map.yaml] [--device N | --devices 0-7] [--batch]
//           [--canny_hysteresis T] [--polygon_epsilon E] [--grid_spacing S]
//           [--neighborhood_search_radius_multiplier M]   image.ppm | frame.bgr:WxH ...
#include "mantis_host.h"

using namespace mantis_host;

static void usage()
{
    fprintf(stderr,
            "usage: mantis3 --K k0,...,k8 [--D d0,d1,d2,d3] [--map map.yaml] [--device N | --devices 0-7] [--batch]\n"
            "               [--canny_hysteresis T] [--polygon_epsilon E] [--grid_spacing S]\n"
            "               [--neighborhood_search_radius_multiplier M] image.ppm|frame.bgr:WxH ...\n");
}

static void printPose(const char* name, const mantis_result& r, const PoseEstimate& p)
{
    printf("%s status=%d quads=%d hyps=%d scored=%d error=%.9g min_yaw_diff=%.9g ", name, r.status, r.n_quads, r.n_hyps,
           r.n_particles, r.error, r.min_yaw_diff);
    if (p.published)
        printf("pose_estimate position=(%.9f %.9f %.9f) orientation_xyzw=(%.9f %.9f %.9f %.9f) variance=%.9g\n",
               p.position[0], p.position[1], p.position[2], p.orientation_xyzw[0], p.orientation_xyzw[1],
               p.orientation_xyzw[2], p.orientation_xyzw[3], p.variance);
    else
        printf("not_published candidate_position=(%.9f %.9f %.9f)\n", r.pos[0], r.pos[1], r.pos[2]);
}

int main(int argc, char** argv)
{
    mantis_config cfg;
    mantis_default_config(&cfg);  // the rosparam defaults of getParameters(), Mantis3Params.h:154-194
    CameraInfo ci;
    ci.D.assign(4, 0.0);
    std::string map_path;
    std::vector<std::string> inputs;
    int device = 0;
    std::vector<int> devices;   // --devices 0-7: the batch is sharded over several GPUs (implies --batch)
    bool have_K = false, batch = false;
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        auto next = [&]() -> const char* { return i + 1 < argc ? argv[++i] : ""; };
        if (a == "--K") have_K = parseDoubles(next(), ci.K, 9);
        else if (a == "--D") { if (!parseDoubles(next(), ci.D.data(), 4)) { usage(); return 2; } }
        else if (a == "--map") map_path = next();
        else if (a == "--check-map") {  // host-only: parse a map file the way getParameters() does and report
            Map m;
            if (!loadMapYaml(next(), m)) { fprintf(stderr, "cannot read map\n"); return 1; }
            printf("white %zu red %zu green %zu first=(%.6f %.6f %.6f) last_green=(%.6f %.6f %.6f)\n", m.white.size(),
                   m.red.size(), m.green.size(), m.white[0].x, m.white[0].y, m.white[0].z, m.green.back().x,
                   m.green.back().y, m.green.back().z);
            return 0;
        }
        else if (a == "--device") device = atoi(next());
        else if (a == "--devices") { devices = parseDevices(next()); batch = true; }
        else if (a == "--batch") batch = true;
        else if (a == "--canny_hysteresis") cfg.canny_hysteresis = atoi(next());
        else if (a == "--polygon_epsilon") cfg.polygon_epsilon = atoi(next());
        else if (a == "--grid_spacing") cfg.grid_spacing = atof(next());
        else if (a == "--neighborhood_search_radius_multiplier") cfg.search_radius_multiplier = atof(next());
        else if (a == "-h" || a == "--help") { usage(); return 0; }
        else inputs.push_back(a);
    }
    if (!have_K || inputs.empty()) { usage(); return 2; }

    if (devices.empty()) devices.push_back(device);
    Map m;
    const bool have_map = !map_path.empty();   // otherwise the library's built-in copy of params/map.yaml is used
    if (have_map) {
        if (!loadMapYaml(map_path, m)) { fprintf(stderr, "cannot read map %s\n", map_path.c_str()); return 1; }
        fprintf(stderr, "size of white map %zu\nsize of red map %zu\nsize of greenMap %zu\n", m.white.size(), m.red.size(),
                m.green.size());
    }
    DevicePool pool;
    int rc = pool.open(devices, &cfg, have_map ? &m : nullptr);
    if (rc != MANTIS_OK) { fprintf(stderr, "mantis_create: %s\n", mantis_strerror(rc)); return 1; }
    mantis_ctx* ctx = pool.ctx(0);

    std::vector<Image> images(inputs.size());
    for (size_t i = 0; i < inputs.size(); ++i) {
        if (!loadImage(inputs[i], images[i])) { fprintf(stderr, "cannot read image %s\n", inputs[i].c_str()); return 1; }
        images[i].stamp_ns = (uint64_t)i;
    }
    std::vector<mantis_frame> frames;
    for (const Image& im : images) frames.push_back(toFrame(im, ci));
    std::vector<mantis_result> res(frames.size());

    bool same = true;
    for (const mantis_frame& f : frames) same = same && f.width == frames[0].width && f.height == frames[0].height;
    if (batch && same) {
        rc = pool.localize(frames.data(), (int)frames.size(), 0, res.data(), nullptr, MANTIS_MEM_HOST);
        if (rc) { fprintf(stderr, "mantis_localize_batch: %s\n", pool.error().c_str()); return 1; }
    } else {
        for (size_t i = 0; i < frames.size(); ++i) {  // one callback per image, as the node receives them
            rc = mantis_localize_batch(ctx, &frames[i], 1, &res[i], nullptr, MANTIS_MEM_HOST);
            if (rc) { fprintf(stderr, "mantis_localize_batch: %s [%s]\n", mantis_strerror(rc), mantis_last_error(ctx)); return 1; }
        }
    }
    for (size_t i = 0; i < frames.size(); ++i) printPose(inputs[i].c_str(), res[i], publishPose(res[i], images[i].stamp_ns));
    return 0;
}
