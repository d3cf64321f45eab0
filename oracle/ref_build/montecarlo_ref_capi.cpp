// C wrapper around the reference's 2016-generation particle weight, compiled together with the reference's own,
// unmodified include/legacy/mantis/MonteCarlo.cpp (from /root/reference) into oracle/_ref/libmontecarlo_ref.so.
// TEST INFRASTRUCTURE ONLY: it pins oracle/legacy.c's mo_camera_error (and through it csrc/legacy.cu's
// mantis_particle_weights_batch) to the reference's project2d / colorError / computeCameraError / computeWeight
// (MonteCarlo.cpp:166-241).  OpenCV, ROS and tf come from ref_build/shim_mc (stand-ins written for this repo).
#include "MonteCarlo.cpp"

extern "C" {

// MonteCarlo::computeCameraError (MonteCarlo.cpp:183-226) on a BGR image with the camera matrix K (9 floats, what
// get3x3FromVector makes of CameraInfo.K) and the world -> camera transform w2c = R (9, row-major) + t (3).
double montecarlo_ref_camera_error(const unsigned char* bgr, int w, int h, int stride, const float* K, const double* w2c12,
                                   const double* map_xyz, int n_white, int n_red, int n_green)
{
    MonteCarlo mc;
    MantisRequest req;
    const double* p = map_xyz;
    for (int i = 0; i < n_white; ++i, p += 3) req.white_map.push_back(tf::Vector3(p[0], p[1], p[2]));
    for (int i = 0; i < n_red; ++i, p += 3) req.red_map.push_back(tf::Vector3(p[0], p[1], p[2]));
    for (int i = 0; i < n_green; ++i, p += 3) req.green_map.push_back(tf::Vector3(p[0], p[1], p[2]));
    req.c1.img = mc_image(bgr, w, h, stride);
    req.c1.K = cv::Mat(3, 3, CV_32F);
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) req.c1.K.at<float>(i, j) = K[3 * i + j];
    tf::Transform t;
    for (int i = 0; i < 3; ++i) t.basis.r[i] = tf::Vector3(w2c12[3 * i], w2c12[3 * i + 1], w2c12[3 * i + 2]);
    t.origin = tf::Vector3(w2c12[9], w2c12[10], w2c12[11]);
    return mc.computeCameraError(req, req.c1, t);
}

// MonteCarlo::computeWeight (MonteCarlo.cpp:233-241): the particle (base pose in the world, R + t) and the camera's
// base -> camera transform b2c (R + t); the camera error is evaluated at particle^-1 * b2c^-1.
double montecarlo_ref_weight(const unsigned char* bgr, int w, int h, int stride, const float* K, const double* particle12,
                             const double* b2c12, const double* map_xyz, int n_white, int n_red, int n_green)
{
    MonteCarlo mc;
    MantisRequest req;
    const double* p = map_xyz;
    for (int i = 0; i < n_white; ++i, p += 3) req.white_map.push_back(tf::Vector3(p[0], p[1], p[2]));
    for (int i = 0; i < n_red; ++i, p += 3) req.red_map.push_back(tf::Vector3(p[0], p[1], p[2]));
    for (int i = 0; i < n_green; ++i, p += 3) req.green_map.push_back(tf::Vector3(p[0], p[1], p[2]));
    req.c1.img = mc_image(bgr, w, h, stride);
    req.c1.K = cv::Mat(3, 3, CV_32F);
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) req.c1.K.at<float>(i, j) = K[3 * i + j];
    auto load = [](const double* a) {
        tf::Transform t;
        for (int i = 0; i < 3; ++i) t.basis.r[i] = tf::Vector3(a[3 * i], a[3 * i + 1], a[3 * i + 2]);
        t.origin = tf::Vector3(a[9], a[10], a[11]);
        return t;
    };
    tf::Transform b2c = load(b2c12);
    req.c1.b2c_inv = b2c.inverse();
    Particle part;
    part.trans = load(particle12);
    part.weight = 0;
    return mc.computeWeight(req, part);
}

}  // extern "C"
