// rpp_batch.cu -- stage-level entry points of the pose / scoring stages:
//   mantis_rpp_batch    RPP::Rpp  (include/mantis3/RobustPlanarPose/RPP.h:54-55, RPP.cpp:13-64) per problem
//   mantis_score_batch  cleanImageByEdge + evaluateHypothesisWithImageWHITE (HypothesisEvaluation.h:107-158, 319-386)
// Planar 4-point RPP problems and all scoring go through the kernels of the localization path (pose.cu,
// search.cu); this file adds the general-shape RPP kernel (3 .. 16 points, any z) and the host plumbing.
#include "ctx.h"
#include "posemath.cuh"
#include "rpp.cuh"

namespace {

constexpr int GNT = 64;

constexpr int CAP = 16;  // most points per problem (array capacity of the one instantiation; the count is a run-time value)

// One thread per problem (fp64, latency-bound: a stage-level checker, not a hot path).
__global__ void __launch_bounds__(GNT) k_rpp_generic(const double* model_xyz, int model_shared, const double* image_xy, int n, int m,
                                                     mantis_rpp_result* out)
{
    const int i = blockIdx.x * GNT + threadIdx.x;
    if (i >= n) return;
    const double* mp = model_xyz + (model_shared ? 0 : (size_t)i * m * 3);
    const double* ip = image_xy + (size_t)i * m * 2;
    pm::V3 model[CAP], ray[CAP];
    for (int k = 0; k < m; ++k) {
        model[k].v[0] = mp[3 * k]; model[k].v[1] = mp[3 * k + 1]; model[k].v[2] = mp[3 * k + 2];
        ray[k].v[0] = ip[2 * k]; ray[k].v[1] = ip[2 * k + 1]; ray[k].v[2] = 1.0;
    }
    mantis_rpp_result r;
    for (int k = 0; k < 9; ++k) r.R[k] = 0;
    for (int k = 0; k < 3; ++k) r.t[k] = 0;
    r.obj_err = r.img_err = 0;
    pm::M3 R;
    pm::V3 t;
    double oe = 0, ie = 0;
    int its[3] = {0, 0, 0};
    const bool ok = rpp::solve<CAP>(model, ray, R, t, oe, ie, its, m);
    r.iters_first = its[0]; r.iters_best = its[1]; r.n_candidates = its[2];
    r.status = ok ? MANTIS_RPP_OK : MANTIS_RPP_FAILED;
    if (ok) {
        for (int k = 0; k < 9; ++k) r.R[k] = R.m[k];
        for (int k = 0; k < 3; ++k) r.t[k] = t.v[k];
        r.obj_err = oe; r.img_err = ie;
    }
    out[i] = r;
}

}  // namespace

int launch_rpp_generic(mantis_ctx* c, const double* d_model_xyz, int model_shared, const double* d_image_xy, int n, int m,
                       mantis_rpp_result* d_out)
{
    if (m < 3 || m > CAP) return MANTIS_ERR_INVALID;
    k_rpp_generic<<<(n + GNT - 1) / GNT, GNT, 0, c->stream>>>(d_model_xyz, model_shared, d_image_xy, n, m, d_out);
    c->launches++;
    MCHECK(c, cudaGetLastError());
    return MANTIS_OK;
}

extern "C" int mantis_rpp_batch(mantis_ctx* c, const double* model_xyz, int model_shared, const double* image_xy, int n, int m,
                                mantis_rpp_result* out, int mem)
{
    if (!c || !model_xyz || !image_xy || !out || n < 0 || m < 3 || m > 16 || (mem != MANTIS_MEM_HOST && mem != MANTIS_MEM_DEVICE))
        return MANTIS_ERR_INVALID;
    if (n == 0) return MANTIS_OK;
    MCHECK(c, cudaSetDevice(c->device));
    const size_t mb = sizeof(double) * 3 * (size_t)m * (model_shared ? 1 : (size_t)n), ib = sizeof(double) * 2 * (size_t)m * n,
                 ob = sizeof(mantis_rpp_result) * (size_t)n;
    const double* d_model = model_xyz;
    const double* d_img = image_xy;
    mantis_rpp_result* d_out = out;
    bool planar = false;
    int rc;
    if (mem == MANTIS_MEM_HOST) {
        planar = m == 4;
        for (size_t k = 0; planar && k < (model_shared ? (size_t)1 : (size_t)n) * m; ++k) planar = model_xyz[3 * k + 2] == 0.0;
        if ((rc = ctx_reserve(c, c->leg_a, mb)) || (rc = ctx_reserve(c, c->leg_b, ib)) || (rc = ctx_reserve(c, c->leg_c, ob))) return rc;
        MCHECK(c, cudaMemcpyAsync(c->leg_a.p, model_xyz, mb, cudaMemcpyHostToDevice, c->stream));
        MCHECK(c, cudaMemcpyAsync(c->leg_b.p, image_xy, ib, cudaMemcpyHostToDevice, c->stream));
        d_model = (const double*)c->leg_a.p; d_img = (const double*)c->leg_b.p; d_out = (mantis_rpp_result*)c->leg_c.p;
    }
    // (device callers: the general kernel, which needs no look at the model on the host)
    rc = planar ? launch_rpp_planar4(c, d_model, model_shared, d_img, n, d_out)
                : launch_rpp_generic(c, d_model, model_shared, d_img, n, m, d_out);
    if (rc) return rc;
    if (mem == MANTIS_MEM_HOST) MCHECK(c, cudaMemcpyAsync(out, d_out, ob, cudaMemcpyDeviceToHost, c->stream));
    MCHECK(c, cudaStreamSynchronize(c->stream));
    return MANTIS_OK;
}

extern "C" int mantis_score_batch(mantis_ctx* c, const uint8_t* bgr, int n, int width, int height, int stride_bytes,
                                  size_t frame_pitch_bytes, const float* K, const double* D, const double* w2c, int n_hyps,
                                  double* scores, int mem)
{
    int rc = check_frames(c, bgr, n, width, height, stride_bytes, frame_pitch_bytes, mem);
    if (rc) return rc;
    if (!K || !D || !w2c || !scores || n_hyps <= 0) return MANTIS_ERR_INVALID;
    if (n == 0) return MANTIS_OK;
    MCHECK(c, cudaSetDevice(c->device));
    const uint8_t* d_in;
    if ((rc = stage_input(c, bgr, n, width, height, stride_bytes, frame_pitch_bytes, mem, &d_in))) return rc;
    const int stride = stride_bytes;            // (stage_input keeps the caller's layout)
    const size_t pitch = frame_pitch_bytes;
    const size_t plane = (size_t)n * height * words_per_row(width) * sizeof(uint32_t);
    const size_t wb = sizeof(double) * 12 * (size_t)n * n_hyps, sb = sizeof(double) * (size_t)n * n_hyps;
    if ((rc = ctx_reserve(c, c->canny_bits, plane)) || (rc = ctx_reserve(c, c->mask_bits, plane)) ||
        (rc = ctx_reserve(c, c->cams, sizeof(pm::Camera) * (size_t)n)))
        return rc;
    // K, D are host arrays whatever `mem` says (like mantis_frame's); poses and scores follow `mem`
    std::vector<pm::Camera> cams((size_t)n);
    for (int i = 0; i < n; ++i) {
        const float* k = K + 9 * (size_t)i;
        if (k[1] != 0.0f || !(k[0] > 0.0f) || !(k[4] > 0.0f)) return MANTIS_ERR_INVALID;
        cams[i].fx = k[0]; cams[i].fy = k[4]; cams[i].cx = k[2]; cams[i].cy = k[5];
        for (int j = 0; j < 4; ++j) cams[i].D[j] = D[4 * (size_t)i + j];
    }
    MCHECK(c, cudaMemcpyAsync(c->cams.p, cams.data(), sizeof(pm::Camera) * (size_t)n, cudaMemcpyHostToDevice, c->stream));
    const double* d_w2c = w2c;
    double* d_scores = scores;
    if (mem == MANTIS_MEM_HOST) {
        if ((rc = ctx_reserve(c, c->leg_a, wb)) || (rc = ctx_reserve(c, c->leg_b, sb))) return rc;
        MCHECK(c, cudaMemcpyAsync(c->leg_a.p, w2c, wb, cudaMemcpyHostToDevice, c->stream));
        d_w2c = (const double*)c->leg_a.p; d_scores = (double*)c->leg_b.p;
    }
    EdgeOutputs eo;
    eo.canny_bits = (uint32_t*)c->canny_bits.p;
    eo.mask_bits = (uint32_t*)c->mask_bits.p;
    if ((rc = run_edge_stage(c, d_in, n, width, height, stride, pitch, eo))) return rc;
    if ((rc = launch_score(c, d_in, pitch, stride, eo.mask_bits, n, width, height, (const pm::Camera*)c->cams.p, d_w2c, n_hyps,
                           d_scores)))
        return rc;
    if (mem == MANTIS_MEM_HOST) MCHECK(c, cudaMemcpyAsync(scores, d_scores, sb, cudaMemcpyDeviceToHost, c->stream));
    MCHECK(c, cudaStreamSynchronize(c->stream));   // (cams is a local: the upload must be over before it goes away)
    return MANTIS_OK;
}
