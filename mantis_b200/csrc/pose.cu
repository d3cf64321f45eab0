// pose.cu -- K7: per-quad planar pose -> hypotheses -> rotation clustering, one CTA per frame.
//
//   undistortAndNormalizeQuadTestPoints   include/mantis3/QuadDetection.h:289-298
//   generateHypotheses / generateCentralHypotheses   HypothesisGeneration.h:23-109
//   PoseClusterer::clusterByAngle / BFcluster / keepLargestCluster / convert2Hypotheses
//                                          PoseClusterer.cpp:33-116, PoseClusterer.h:59-115
// One thread per quad runs RPP (rpp.cuh) for up to two corner orderings; passing quads emit four
// yaw-rotated hypotheses in quad order.  The greedy Euler-space clustering is sequential over
// seeds (as in the reference) and parallel over members.
#include "ctx.h"
#include "rpp.cuh"

namespace {

constexpr int PNT = 128;

struct PoseArgs {
    const FrameQuads* quads;
    const pm::Camera* cams;  // per frame
    Hyp* quad_hyp;           // [n][MANTIS_MAX_QUADS] RPP result of each passing quad
    Hyp* hyps;               // [n][MANTIS_MAX_HYPS]
    short* sel;              // [n][MANTIS_MAX_HYPS] members of the kept cluster
    FramePose* info;         // [n]
    double half_spacing, max_quad_error, cluster_radius;
    Tf rotz;                 // tf::Transform(tf::Quaternion(0, 0, 1/sqrt2, 1/sqrt2))
};

// Exclusive prefix sum of the per-frame quad counts: item i of the flattened work list belongs to
// the frame f with offs[f] <= i < offs[f+1].
__global__ void k_pose_index(const FrameQuads* quads, int n, int* offs)
{
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        int acc = 0;
        for (int f = 0; f < n; ++f) { offs[f] = acc; acc += quads[f].n_quads; }
        offs[n] = acc;
    }
}

// One thread per (frame, quad): undistort the test points, RPP for up to two corner orderings
// (generateCentralHypotheses, HypothesisGeneration.h:57-89).  Work items are flattened across the
// frames of the batch so that warps are fully populated.
__global__ void __launch_bounds__(PNT) k_pose_solve(PoseArgs a, const int* offs, int n_frames)
{
    const int item = blockIdx.x * PNT + threadIdx.x;
    if (item >= offs[n_frames]) return;
    int lo = 0, hi = n_frames;  // last f with offs[f] <= item
    while (hi - lo > 1) {
        int mid = (lo + hi) >> 1;
        if (offs[mid] <= item) lo = mid; else hi = mid;
    }
    const int frame = lo, q = item - offs[lo];
    const FrameQuads* fq = a.quads + frame;
    const pm::Camera cam = a.cams[frame];
    FramePose* info = a.info + frame;

    pm::V3 ray[4];
    for (int k = 0; k < 4; ++k) {
        pm::fisheye_undistort(cam, fq->test_px[q][2 * k], fq->test_px[q][2 * k + 1], ray[k].v[0], ray[k].v[1]);
        ray[k].v[2] = 1.0;
    }
    const double s = a.half_spacing;
    // gridSquarePossibilities, Mantis3Params.h:102-123
    const double poss[2][4][2] = {{{s, s}, {-s, s}, {-s, -s}, {s, -s}}, {{s, -s}, {-s, -s}, {-s, s}, {s, s}}};
    double error = 1.7976931348623157e308;
    bool failed = false;
    Hyp h0;
    for (int e = 0; e < 2; ++e) {
        pm::V3 model[4];
        for (int k = 0; k < 4; ++k) { model[k].v[0] = poss[e][k][0]; model[k].v[1] = poss[e][k][1]; model[k].v[2] = 0; }
        pm::M3 R;
        pm::V3 t;
        double oe, ie;
        if (!rpp::solve<4>(model, ray, R, t, oe, ie)) { failed = true; break; }
        error = ie;
        Tf c2w;
        for (int k = 0; k < 9; ++k) c2w.R[k] = R.m[k];
        for (int k = 0; k < 3; ++k) c2w.t[k] = t.v[k];
        pm::hyp_set_c2w(h0, c2w);
        if (h0.w2c.t[2] >= 0) break;
    }
    if (failed) error = 1.7976931348623157e308;
    const bool pass = !failed && !(error > a.max_quad_error);
    info->quad_img_err[q] = error;
    info->quad_pass[q] = pass;
    if (pass) {
        h0.error = 0;
        a.quad_hyp[(size_t)frame * MANTIS_MAX_QUADS + q] = h0;
    }
}

// Per frame: hypotheses of the passing quads in quad order (4 yaw copies each), Euler angles,
// greedy clustering, largest cluster.
__global__ void __launch_bounds__(PNT) k_pose_cluster(PoseArgs a)
{
    __shared__ float s_ang[MANTIS_MAX_HYPS][3];
    __shared__ unsigned char s_taken[MANTIS_MAX_HYPS];
    __shared__ unsigned char s_member[MANTIS_MAX_HYPS];
    __shared__ unsigned char s_best[MANTIS_MAX_HYPS];
    __shared__ short s_slot[MANTIS_MAX_QUADS];
    __shared__ float s_center[3];
    __shared__ int s_cnt, s_nh;

    const int frame = blockIdx.x, tid = threadIdx.x;
    const FrameQuads* fq = a.quads + frame;
    Hyp* hyps = a.hyps + (size_t)frame * MANTIS_MAX_HYPS;
    short* sel = a.sel + (size_t)frame * MANTIS_MAX_HYPS;
    FramePose* info = a.info + frame;
    const int nq = fq->n_quads;

    if (tid == 0) {
        int acc = 0;
        for (int k = 0; k < nq; ++k) { s_slot[k] = (short)acc; acc += info->quad_pass[k] ? 1 : 0; }
        s_nh = acc;
    }
    __syncthreads();
    for (int q = tid; q < nq; q += PNT) {
        if (!info->quad_pass[q]) continue;
        Hyp hk = a.quad_hyp[(size_t)frame * MANTIS_MAX_QUADS + q];
        const int base = 4 * s_slot[q];
        hyps[base] = hk;
        for (int k = 1; k < 4; ++k) {
            Tf w2c;
            pm::tf_mul(a.rotz, hk.w2c, w2c);
            pm::hyp_set_w2c(hk, w2c);
            hyps[base + k] = hk;
        }
    }
    __syncthreads();
    if (nq == 0) {
        if (tid == 0) { info->n_hyps_generated = 0; info->n_sel = 0; }
        return;
    }
    const int nh = 4 * s_nh;

    // ---- Euler angles (tf::Matrix3x3(q).getRPY, stored as float like cv::Point3f) ----
    for (int i = tid; i < nh; i += PNT) {
        double m[9], rpy[3];
        pm::quat_to_mat(hyps[i].q, m);
        pm::get_rpy(m, rpy);
        s_ang[i][0] = (float)rpy[0];
        s_ang[i][1] = (float)rpy[1];
        s_ang[i][2] = (float)rpy[2];
        s_taken[i] = 0;
        s_best[i] = 0;
    }
    __syncthreads();

    // ---- BFcluster + keepLargestCluster ----
    const double rad = a.cluster_radius;
    int best_n = -1;
    for (int i = 0; i < nh; ++i) {
        if (s_taken[i]) continue;  // uniform: shared memory, read after a barrier
        for (int j = tid; j < nh; j += PNT) {
            bool m = false;
            if (!s_taken[j]) {
                double dx = (double)(s_ang[i][0] - s_ang[j][0]), dy = (double)(s_ang[i][1] - s_ang[j][1]),
                       dz = (double)(s_ang[i][2] - s_ang[j][2]);
                m = sqrt(dx * dx + dy * dy + dz * dz) <= rad * 1.5;
            }
            s_member[j] = m;
        }
        __syncthreads();
        if (tid == 0) {  // float accumulation in index order, as getClusterCenter does
            float c0 = 0.f, c1 = 0.f, c2 = 0.f;
            int cnt = 0;
            for (int j = 0; j < nh; ++j)
                if (s_member[j]) { c0 += s_ang[j][0]; c1 += s_ang[j][1]; c2 += s_ang[j][2]; ++cnt; }
            s_center[0] = (float)((double)c0 / (double)cnt);
            s_center[1] = (float)((double)c1 / (double)cnt);
            s_center[2] = (float)((double)c2 / (double)cnt);
            s_cnt = 0;
        }
        __syncthreads();
        int local = 0;
        for (int j = tid; j < nh; j += PNT) {
            bool m = false;
            if (!s_taken[j]) {
                double dx = (double)(s_center[0] - s_ang[j][0]), dy = (double)(s_center[1] - s_ang[j][1]),
                       dz = (double)(s_center[2] - s_ang[j][2]);
                m = sqrt(dx * dx + dy * dy + dz * dz) <= rad;
            }
            s_member[j] = m;
            local += m;
        }
        if (local) atomicAdd(&s_cnt, local);
        __syncthreads();
        const int cnt = s_cnt;
        const bool better = cnt > best_n;  // first-largest wins
        for (int j = tid; j < nh; j += PNT) {
            if (better) s_best[j] = s_member[j];
            if (s_member[j]) s_taken[j] = 1;
        }
        if (better) best_n = cnt;
        __syncthreads();
    }
    if (tid == 0) {
        int n = 0;
        for (int j = 0; j < nh; ++j)
            if (s_best[j]) sel[n++] = (short)j;
        info->n_hyps_generated = nh;
        info->n_sel = n;
    }
}

}  // namespace

int launch_pose(mantis_ctx* c, const FrameQuads* d_quads, const pm::Camera* d_cams, int n, Hyp* d_hyps, short* d_sel,
                FramePose* d_info)
{
    PoseArgs a;
    a.quads = d_quads; a.cams = d_cams; a.hyps = d_hyps; a.sel = d_sel; a.info = d_info;
    a.half_spacing = c->cfg.grid_spacing / 2;
    a.max_quad_error = c->cfg.max_quad_error;
    a.cluster_radius = c->cfg.max_angle_difference;
    a.rotz = c->rotz;
    int rc;
    if ((rc = ctx_reserve(c, c->quad_hyp, (size_t)n * MANTIS_MAX_QUADS * sizeof(Hyp))) ||
        (rc = ctx_reserve(c, c->pose_offs, (size_t)(n + 1) * sizeof(int))))
        return rc;
    a.quad_hyp = (Hyp*)c->quad_hyp.p;
    int* offs = (int*)c->pose_offs.p;
    k_pose_index<<<1, 32, 0, c->stream>>>(d_quads, n, offs);
    // the item count is only known on the device: launch for the capacity and let spare threads exit
    // (upper bound tightened with the previous stage's counts would need a sync)
    const int max_items = n * MANTIS_MAX_QUADS;
    k_pose_solve<<<(max_items + PNT - 1) / PNT, PNT, 0, c->stream>>>(a, offs, n);
    k_pose_cluster<<<n, PNT, 0, c->stream>>>(a);
    c->launches += 3;
    MCHECK(c, cudaGetLastError());
    return MANTIS_OK;
}
