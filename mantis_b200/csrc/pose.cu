// pose.cu -- K7: per-quad planar pose -> hypotheses -> rotation clustering, one CTA per frame.
//
//   undistortAndNormalizeQuadTestPoints   include/mantis3/QuadDetection.h:289-298
//   generateHypotheses / generateCentralHypotheses   HypothesisGeneration.h:23-109
//   PoseClusterer::clusterByAngle / BFcluster / keepLargestCluster / convert2Hypotheses
//                                          PoseClusterer.cpp:33-116, PoseClusterer.h:59-115
// RPP (rpp.cuh, rpp_fast.cuh) runs as a queue of independent ObjPose runs over all quads of the
// batch (k_pose_run) for corner ordering 0; ordering 1 is its mirror (k_pose_prep); passing quads emit four
// yaw-rotated hypotheses in quad order.  The greedy Euler-space clustering is sequential over
// seeds (as in the reference) and parallel over members.
#include <stdlib.h>
#include "ctx.h"
#include "rpp_fast.cuh"

namespace {

constexpr int PNT = 128;

struct RunOut {   // result of one ObjPose run
    double R[9], t[3], obj_err, img_err;
    int calls, pad_;  // AbsKernel calls of the run (the reference's `it`, RPP.cpp:66-208)
};

struct PoseCounters {   // zeroed before every batch
    int n_t1[2];      // (quad, ordering) tasks of round 0 / 1
    int n_t2[2];      // second-pose refinement runs of round 0 / 1
    int fetch1[2], fetch2[2];
};

struct PoseArgs {
    const FrameQuads* quads;
    const pm::Camera* cams;  // per frame
    int n_frames;
    Hyp* quad_hyp;           // [n][MANTIS_MAX_QUADS] RPP result of each passing quad
    Hyp* hyps;               // [n][MANTIS_MAX_HYPS]
    short* sel;              // [n][MANTIS_MAX_HYPS] members of the kept cluster
    FramePose* info;         // [n]
    // workspace of the solve (cap = n * MANTIS_MAX_QUADS quads; an entry is quad * 2 + ordering)
    PoseCounters* cnt;
    double* rays;            // [cap][8] undistorted test points
    int* list1;              // [2][cap] entries of round 0 / 1
    int* list2;              // [2][cap * 4] entry * 4 + candidate
    RunOut* run1;            // [cap * 2]
    int* ncand;              // [cap * 2]
    double* candR;           // [cap * 2 * 4][9]
    RunOut* run2;            // [cap * 2 * 4]
    int cap;
    int cams_per_unit;       // 1, or the cameras of a rig (k_pose_cluster pools their hypotheses)
    const Tf* b2c_inv;       // rigs only: per frame, base <- camera inverted; null = hypotheses stay camera poses
    const double* models;    // stage-level entry (mantis_rpp_batch) only: [cap][8] planar model (x, y) per quad; null = the grid square
    long long* dbg;          // tuning aid (null = off): slots 10..15 = AbsKernel steps spent in runs of <16, <64, <256, <1024, <4096, more steps
    double half_spacing, max_quad_error, cluster_radius;
    Tf rotz;                 // tf::Transform(tf::Quaternion(0, 0, 1/sqrt2, 1/sqrt2))
};

// gridSquarePossibilities, Mantis3Params.h:102-123
__device__ __forceinline__ void model_xy(double s, int ordering, double* m)
{
    const double sx[2][4] = {{1, -1, -1, 1}, {1, -1, -1, 1}}, sy[2][4] = {{1, 1, -1, -1}, {-1, -1, 1, 1}};
    for (int k = 0; k < 4; ++k) { m[2 * k] = sx[ordering][k] * s; m[2 * k + 1] = sy[ordering][k] * s; }
}

// ---- 1. undistort the test points, queue every quad ---------------------------------------------
// The reference solves the second corner ordering when the first one puts the camera under the floor
// (HypothesisGeneration.h:62-80), which is most quads.  The two models differ by the sign of y
// (gridSquarePossibilities, Mantis3Params.h:102-123) and the model is planar, so (R, t) solves ordering 0 exactly
// when (R diag(1, -1, -1), t) solves ordering 1: same points R P_i + t, same errors, the camera mirrored through
// the floor.  Every step of RPP keeps that symmetry (sums, polar factor, the second-pose quartic), so ordering 1
// is not run at all: k_pose_select mirrors the result of ordering 0.  Measured on the quads of 1024 rendered
// frames (tools/rpp_emul/mirror.py, 56 395 quads): the independently iterated ordering 1 and the mirror agree to 4e-16 (device code) /
// 7e-15 (oracle) with the same iteration counts in every run, a tenth of the distance between the device's and
// the oracle's RPP anyway.
__global__ void __launch_bounds__(PNT) k_pose_prep(PoseArgs a)
{
    __shared__ int s_base;
    const int frame = blockIdx.x, tid = threadIdx.x;
    const FrameQuads* fq = a.quads + frame;
    const int nq = fq->flags ? 0 : fq->n_quads;
    if (tid == 0) s_base = nq ? atomicAdd(&a.cnt->n_t1[0], nq) : 0;
    __syncthreads();
    const pm::Camera cam = a.cams[frame];
    for (int q = tid; q < nq; q += PNT) {
        const int id = frame * MANTIS_MAX_QUADS + q;
        for (int k = 0; k < 4; ++k)
            pm::fisheye_undistort(cam, fq->test_px[q][2 * k], fq->test_px[q][2 * k + 1], a.rays[(size_t)id * 8 + 2 * k],
                                  a.rays[(size_t)id * 8 + 2 * k + 1]);
        a.list1[s_base + q] = id * 2;   // (entry = quad * 2 + ordering; only ordering 0 is queued)
    }
}

// ---- 2. ObjPose runs, many per warp -------------------------------------------------------------
// A lane owns one run at a time and advances it by one AbsKernel step per loop trip; finished
// lanes pick up the next queued run every few trips, so lanes with short runs do not idle behind
// long ones (runs take 2 .. several hundred steps).
constexpr int RNT = 128;
#ifndef RUN_CTAS
#define RUN_CTAS 2
#endif
template <bool HAS_INIT>
__global__ void __launch_bounds__(RNT, RUN_CTAS) k_pose_run(PoseArgs a, int round)
{
    extern __shared__ double s_ctx[];  // [rppf::CTX_WORDS][RNT]
    const double TOL = 1e-5, EPSILON = 1e-8;
    const int lane = threadIdx.x & 31;
    double* ctx = s_ctx + threadIdx.x;
    const int* list = HAS_INIT ? a.list2 + (size_t)round * a.cap * 4 : a.list1 + (size_t)round * a.cap;
    const int ntask = HAS_INIT ? a.cnt->n_t2[round] : a.cnt->n_t1[round];
    int* fetch = HAS_INIT ? &a.cnt->fetch2[round] : &a.cnt->fetch1[round];
    RunOut* out = HAS_INIT ? a.run2 : a.run1;
    rppf::Lane L;
    rppf::Run S;
    S.it = 0; S.pending = 0; S.old_err = S.new_err = 0;
    int task = -1;
    bool exhausted = false;
    for (int trip = 0;; ++trip) {
        if ((trip & 3) == 0) {
            const bool want = task < 0 && !exhausted;
            const unsigned need = __ballot_sync(0xffffffffu, want);
            if (need) {
                const int leader = __ffs(need) - 1;
                int base = 0;
                if (lane == leader) base = atomicAdd(fetch, __popc(need));
                base = __shfl_sync(0xffffffffu, base, leader);
                if (want) {
                    const int my = base + __popc(need & ((1u << lane) - 1u));
                    if (my < ntask) {
                        task = list[my];
                        const int entry = HAS_INIT ? (task >> 2) : task;
                        double m[8];
                        if (a.models) { for (int k = 0; k < 8; ++k) m[k] = a.models[(size_t)(entry >> 1) * 8 + k]; }
                        else model_xy(a.half_spacing, entry & 1, m);
                        rppf::setup(ctx, RNT, a.rays + (size_t)(entry >> 1) * 8, m, L);
                        rppf::run_begin(ctx, RNT, L, S, HAS_INIT ? a.candR + (size_t)task * 9 : nullptr);
                    } else {
                        exhausted = true;
                    }
                }
            }
            if (__ballot_sync(0xffffffffu, task >= 0) == 0) break;
        }
        if (task >= 0 && rppf::run_step(ctx, RNT, L, S, TOL, EPSILON)) {
            if (a.dbg && !HAS_INIT) atomicAdd((unsigned long long*)&a.dbg[10 + (S.it >= 16) + (S.it >= 64) + (S.it >= 256) + (S.it >= 1024) + (S.it >= 4096)], (unsigned long long)S.it);
            RunOut o;
            rppf::finalize(ctx, RNT, L, o.R, o.t, o.obj_err, o.img_err);
            o.calls = S.it + (HAS_INIT ? 1 : 2);
            o.pad_ = 0;
            out[task] = o;
            task = -1;
        }
    }
}

// ---- 3. second-pose candidates (RPP::Get2ndPose_Exact), one thread per queued quad --------------
#ifndef SECOND_CTAS
#define SECOND_CTAS 4  // 128 registers: the spills sit in L1 (no shared memory here) and twice the warps hide the fp64 latency (pose 5.24 -> 4.92 ms per 4096 frames)
#endif
__global__ void __launch_bounds__(PNT, SECOND_CTAS) k_pose_second(PoseArgs a, int round)
{
    const int ntask = a.cnt->n_t1[round];
    const int* list = a.list1 + (size_t)round * a.cap;
    for (int i = blockIdx.x * PNT + threadIdx.x; i < ntask; i += gridDim.x * PNT) {
        const int entry = list[i];
        double m[8];
        if (a.models) { for (int k = 0; k < 8; ++k) m[k] = a.models[(size_t)(entry >> 1) * 8 + k]; }
        else model_xy(a.half_spacing, entry & 1, m);
        pm::V3 model[4], ray[4];
        for (int k = 0; k < 4; ++k) {
            model[k].v[0] = m[2 * k]; model[k].v[1] = m[2 * k + 1]; model[k].v[2] = 0;
            ray[k].v[0] = a.rays[(size_t)(entry >> 1) * 8 + 2 * k];
            ray[k].v[1] = a.rays[(size_t)(entry >> 1) * 8 + 2 * k + 1];
            ray[k].v[2] = 1.0;
        }
        pm::M3 R, solR[4];
        pm::V3 solT[4];
        for (int k = 0; k < 9; ++k) R.m[k] = a.run1[entry].R[k];
        const int ns = rpp::second_poses<4>(model, ray, R, solR, solT);
        a.ncand[entry] = ns;
        if (ns <= 0) continue;
        const int base = atomicAdd(&a.cnt->n_t2[round], ns);
        int* list2 = a.list2 + (size_t)round * a.cap * 4;
        for (int c = 0; c < ns; ++c) {
            for (int k = 0; k < 9; ++k) a.candR[((size_t)entry * 4 + c) * 9 + k] = solR[c].m[k];
            list2[base + c] = entry * 4 + c;
        }
    }
}

// ---- 4. best candidate -> hypothesis; a wrong-side solution of ordering 0 hands over to (the mirror that is) ordering 1 ----
__global__ void __launch_bounds__(PNT) k_pose_select(PoseArgs a, int round)
{
    const int ntask = a.cnt->n_t1[round];
    const int* list = a.list1 + (size_t)round * a.cap;
    for (int i = blockIdx.x * PNT + threadIdx.x; i < ntask; i += gridDim.x * PNT) {
        const int entry = list[i], id = entry >> 1;   // one thread per quad: the entry of ordering 0
        const int frame = id / MANTIS_MAX_QUADS, q = id - frame * MANTIS_MAX_QUADS;
        FramePose* info = a.info + frame;
        const int ns = a.ncand[entry];
        int best = -1;
        double lowest = 1e6;
        for (int c = 0; c < ns; ++c) {
            const double oe = a.run2[(size_t)entry * 4 + c].obj_err;
            if (oe < lowest) { lowest = oe; best = c; }
        }
        if (best < 0) {  // RPP::Rpp returned false: no further ordering is tried (HypothesisGeneration.h:62-66)
            info->quad_img_err[q] = 1.7976931348623157e308;
            info->quad_pass[q] = 0;
            continue;
        }
        const RunOut& r = a.run2[(size_t)entry * 4 + best];
        Tf c2w;
        for (int k = 0; k < 9; ++k) c2w.R[k] = r.R[k];
        for (int k = 0; k < 3; ++k) c2w.t[k] = r.t[k];
        Hyp h0;
        pm::hyp_set_c2w(h0, c2w);
        if (!(h0.w2c.t[2] >= 0)) {
            // camera under the floor: the other ordering decides, and its solution is this one mirrored (see k_pose_prep)
            for (int k = 0; k < 3; ++k) { c2w.R[3 * k + 1] = -c2w.R[3 * k + 1]; c2w.R[3 * k + 2] = -c2w.R[3 * k + 2]; }
            pm::hyp_set_c2w(h0, c2w);
        }
        const double error = r.img_err;
        const bool pass = !(error > a.max_quad_error);
        info->quad_img_err[q] = error;
        info->quad_pass[q] = pass;
        if (pass) {
            h0.error = 0;
            a.quad_hyp[(size_t)id] = h0;
        }
    }
}

// Per frame: hypotheses of the passing quads in quad order (4 yaw copies each), Euler angles,
// greedy clustering, largest cluster.
// A unit is one frame, or -- for a rig (mantis_localize_rig_batch) -- the a.cams_per_unit consecutive frames of one rig:
// then every camera's hypotheses are moved to the base frame (W2B = W2C_k * b2c_k^-1, the inverse of
// Mantis2Error.h:67-68's setW2C(hyp.getW2C() * b2c)) and pooled in camera order before the clustering.  The unit's
// pool / selection live in the slots of its first frame onwards (capacity cams_per_unit * MANTIS_MAX_HYPS); its
// totals (n_hyps_generated, n_sel) go to the FramePose of its first frame.
#ifndef CLUSTER_CTAS
#define CLUSTER_CTAS 8   // 64 registers: the kernel waits on barriers (one greedy seed after the other), more resident CTAs hide them
#endif
__global__ void __launch_bounds__(PNT, CLUSTER_CTAS) k_pose_cluster(PoseArgs a)
{
    extern __shared__ __align__(16) unsigned char s_dyn[];
    const int C = a.cams_per_unit, cap = C * MANTIS_MAX_HYPS;
    float (*s_ang)[3] = reinterpret_cast<float (*)[3]>(s_dyn);
    unsigned char* s_taken = s_dyn + (size_t)cap * 3 * sizeof(float);
    unsigned char* s_member = s_taken + cap;
    unsigned char* s_best = s_member + cap;
    __shared__ short s_slot[MANTIS_MAX_QUADS];
    __shared__ float s_center[3];
    __shared__ int s_cnt, s_nh;

    const int unit = blockIdx.x, tid = threadIdx.x;
    Hyp* hyps = a.hyps + (size_t)unit * cap;
    short* sel = a.sel + (size_t)unit * cap;
    FramePose* info = a.info + (size_t)unit * C;   // the unit's totals
    int nq_total = 0, nh_total = 0;
    for (int cam = 0; cam < C; ++cam) {
        const int frame = unit * C + cam;
        const FrameQuads* fq = a.quads + frame;
        const FramePose* finfo = a.info + frame;
        const int nq = fq->flags ? 0 : fq->n_quads;
        __syncthreads();
        if (tid == 0) {
            int acc = 0;
            for (int k = 0; k < nq; ++k) { s_slot[k] = (short)acc; acc += finfo->quad_pass[k] ? 1 : 0; }
            s_nh = acc;
        }
        __syncthreads();
        for (int q = tid; q < nq; q += PNT) {
            if (!finfo->quad_pass[q]) continue;
            Hyp hk = a.quad_hyp[(size_t)frame * MANTIS_MAX_QUADS + q];
            const int base = nh_total + 4 * s_slot[q];
            for (int k = 0; k < 4; ++k) {
                if (k > 0) {
                    Tf w2c;
                    pm::tf_mul(a.rotz, hk.w2c, w2c);
                    pm::hyp_set_w2c(hk, w2c);
                }
                if (a.b2c_inv) {   // camera pose -> base pose
                    Hyp hb = hk;
                    Tf w2b;
                    pm::tf_mul(hk.w2c, a.b2c_inv[frame], w2b);
                    pm::hyp_set_w2c(hb, w2b);
                    hyps[base + k] = hb;
                } else {
                    hyps[base + k] = hk;
                }
            }
        }
        nq_total += nq;
        nh_total += 4 * s_nh;
    }
    __syncthreads();
    if (nq_total == 0) {
        if (tid == 0) { info->n_hyps_generated = 0; info->n_sel = 0; }
        return;
    }
    const int nh = nh_total;

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
        if (tid < 3) {  // float accumulation in index order, as getClusterCenter does: one lane per coordinate,
                        // branch-free (x + 0.0f == x), so that the loads run ahead of the dependent additions
            float c = 0.f;
            int cnt = 0;
#pragma unroll 8
            for (int j = 0; j < nh; ++j) {
                const bool m = s_member[j];
                c += m ? s_ang[j][tid] : 0.0f;
                cnt += m ? 1 : 0;
            }
            s_center[tid] = (float)((double)c / (double)cnt);
            if (tid == 0) s_cnt = 0;
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

// ---- stage-level entry (mantis_rpp_batch): n planar 4-point problems through the kernels above -------------
// One queue entry per problem (entry = 2 * problem: "ordering 0" slot, the model comes from a.models).
__global__ void __launch_bounds__(PNT) k_rpp_stage_prep(PoseArgs a, const double* model_xyz, int model_shared, const double* image_xy,
                                                        int n, double* models_out)
{
    const int i = blockIdx.x * PNT + threadIdx.x;
    if (i == 0) a.cnt->n_t1[0] = n;
    if (i >= n) return;
    const double* mp = model_xyz + (model_shared ? 0 : (size_t)i * 12);
    for (int k = 0; k < 4; ++k) {
        models_out[(size_t)i * 8 + 2 * k] = mp[3 * k];
        models_out[(size_t)i * 8 + 2 * k + 1] = mp[3 * k + 1];
        a.rays[(size_t)i * 8 + 2 * k] = image_xy[(size_t)i * 8 + 2 * k];
        a.rays[(size_t)i * 8 + 2 * k + 1] = image_xy[(size_t)i * 8 + 2 * k + 1];
    }
    a.list1[i] = 2 * i;
}

// RPP::Rpp's choice among the refined candidates (RPP.cpp:36-52): lowest object-space error, first wins ties.
__global__ void __launch_bounds__(PNT) k_rpp_stage_gather(PoseArgs a, int n, mantis_rpp_result* out)
{
    const int i = blockIdx.x * PNT + threadIdx.x;
    if (i >= n) return;
    const int entry = 2 * i, ns = a.ncand[entry];
    mantis_rpp_result r;
    for (int k = 0; k < 9; ++k) r.R[k] = 0;
    for (int k = 0; k < 3; ++k) r.t[k] = 0;
    r.obj_err = r.img_err = 0;
    r.iters_first = a.run1[entry].calls;
    r.iters_best = 0;
    r.n_candidates = ns > 0 ? ns : 0;
    r.status = MANTIS_RPP_FAILED;
    int best = -1;
    double lowest = 1e6;
    for (int cnd = 0; cnd < ns; ++cnd) {
        const double oe = a.run2[(size_t)entry * 4 + cnd].obj_err;
        if (oe < lowest) { lowest = oe; best = cnd; }
    }
    if (best >= 0) {
        const RunOut& b = a.run2[(size_t)entry * 4 + best];
        for (int k = 0; k < 9; ++k) r.R[k] = b.R[k];
        for (int k = 0; k < 3; ++k) r.t[k] = b.t[k];
        r.obj_err = b.obj_err; r.img_err = b.img_err;
        r.iters_best = b.calls;
        r.status = MANTIS_RPP_OK;
    }
    out[i] = r;
}

// Workspace of the ObjPose queue for `cap` quads, carved out of one context buffer.
int pose_workspace(mantis_ctx* c, size_t cap, PoseArgs& a)
{
    a.cap = (int)cap;
    a.models = nullptr;
    a.cams_per_unit = 1;
    a.b2c_inv = nullptr;
    size_t o = 0;
    auto take = [&](size_t bytes) { size_t at = o; o += (bytes + 255) & ~(size_t)255; return at; };
    const size_t o_cnt = take(sizeof(PoseCounters)), o_rays = take(cap * 8 * sizeof(double)), o_l1 = take(cap * 2 * sizeof(int)),
                 o_l2 = take(cap * 8 * sizeof(int)), o_run1 = take(cap * 2 * sizeof(RunOut)), o_nc = take(cap * 2 * sizeof(int)),
                 o_cr = take(cap * 8 * 9 * sizeof(double)), o_run2 = take(cap * 8 * sizeof(RunOut));
    int rc;
    if ((rc = ctx_reserve(c, c->pose_offs, o))) return rc;
    char* ws = (char*)c->pose_offs.p;
    a.cnt = (PoseCounters*)(ws + o_cnt); a.rays = (double*)(ws + o_rays); a.list1 = (int*)(ws + o_l1);
    a.list2 = (int*)(ws + o_l2); a.run1 = (RunOut*)(ws + o_run1); a.ncand = (int*)(ws + o_nc);
    a.candR = (double*)(ws + o_cr); a.run2 = (RunOut*)(ws + o_run2);
    return MANTIS_OK;
}

}  // namespace

int launch_pose(mantis_ctx* c, const FrameQuads* d_quads, const pm::Camera* d_cams, int n, Hyp* d_hyps, short* d_sel,
                FramePose* d_info, int cams_per_unit, const Tf* d_b2c_inv)
{
    PoseArgs a;
    a.quads = d_quads; a.cams = d_cams; a.n_frames = n; a.hyps = d_hyps; a.sel = d_sel; a.info = d_info;
    a.half_spacing = c->cfg.grid_spacing / 2;
    a.max_quad_error = c->cfg.max_quad_error;
    a.cluster_radius = c->cfg.max_angle_difference;
    a.rotz = c->rotz;
    int rc;
    if ((rc = pose_workspace(c, (size_t)n * MANTIS_MAX_QUADS, a))) return rc;
    if ((rc = ctx_reserve(c, c->quad_hyp, (size_t)a.cap * sizeof(Hyp)))) return rc;
    a.quad_hyp = (Hyp*)c->quad_hyp.p;
    a.dbg = (long long*)c->dbg_clocks;
    MCHECK(c, cudaMemsetAsync(a.cnt, 0, sizeof(PoseCounters), c->stream));
    k_pose_prep<<<n, PNT, 0, c->stream>>>(a);
    static const int rb = []() { const char* e = getenv("MANTIS_RUN_BLOCKS"); int v = e ? atoi(e) : 0; return v > 0 ? v : 4; }();
    const int run_blocks = c->sm_count * rb, flat_blocks = c->sm_count * 4;
    const int run_smem = rppf::CTX_WORDS * RNT * (int)sizeof(double);
    MCHECK(c, cudaFuncSetAttribute(k_pose_run<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, run_smem));
    MCHECK(c, cudaFuncSetAttribute(k_pose_run<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, run_smem));
    k_pose_run<false><<<run_blocks, RNT, run_smem, c->stream>>>(a, 0);
    k_pose_second<<<c->sm_count * SECOND_CTAS * 2, PNT, 0, c->stream>>>(a, 0);
    k_pose_run<true><<<run_blocks, RNT, run_smem, c->stream>>>(a, 0);
    k_pose_select<<<flat_blocks, PNT, 0, c->stream>>>(a, 0);
    a.cams_per_unit = cams_per_unit;
    a.b2c_inv = d_b2c_inv;
    const int cluster_smem = cams_per_unit * MANTIS_MAX_HYPS * (3 * (int)sizeof(float) + 3);
    MCHECK(c, cudaFuncSetAttribute(k_pose_cluster, cudaFuncAttributeMaxDynamicSharedMemorySize, cluster_smem));
    k_pose_cluster<<<n / cams_per_unit, PNT, cluster_smem, c->stream>>>(a);
    c->launches += 6;
    MCHECK(c, cudaGetLastError());
    return MANTIS_OK;
}

// n planar 4-point problems (device pointers) through k_pose_run / k_pose_second, the kernels of the localization path.
int launch_rpp_planar4(mantis_ctx* c, const double* d_model_xyz, int model_shared, const double* d_image_xy, int n,
                       mantis_rpp_result* d_out)
{
    PoseArgs a;
    a.quads = nullptr; a.cams = nullptr; a.n_frames = 0; a.hyps = nullptr; a.sel = nullptr; a.info = nullptr; a.quad_hyp = nullptr;
    a.half_spacing = c->cfg.grid_spacing / 2;
    a.max_quad_error = c->cfg.max_quad_error;
    a.cluster_radius = c->cfg.max_angle_difference;
    a.rotz = c->rotz;
    int rc;
    if ((rc = pose_workspace(c, (size_t)n, a))) return rc;
    if ((rc = ctx_reserve(c, c->quad_hyp, (size_t)n * 8 * sizeof(double)))) return rc;   // (free here: holds the models)
    a.models = (const double*)c->quad_hyp.p;
    a.dbg = nullptr;
    MCHECK(c, cudaMemsetAsync(a.cnt, 0, sizeof(PoseCounters), c->stream));
    const int flat = (n + PNT - 1) / PNT;
    k_rpp_stage_prep<<<flat, PNT, 0, c->stream>>>(a, d_model_xyz, model_shared, d_image_xy, n, (double*)c->quad_hyp.p);
    const int run_blocks = c->sm_count * 4;
    const int run_smem = rppf::CTX_WORDS * RNT * (int)sizeof(double);
    MCHECK(c, cudaFuncSetAttribute(k_pose_run<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, run_smem));
    MCHECK(c, cudaFuncSetAttribute(k_pose_run<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, run_smem));
    k_pose_run<false><<<run_blocks, RNT, run_smem, c->stream>>>(a, 0);
    k_pose_second<<<c->sm_count * SECOND_CTAS * 2, PNT, 0, c->stream>>>(a, 0);
    k_pose_run<true><<<run_blocks, RNT, run_smem, c->stream>>>(a, 0);
    k_rpp_stage_gather<<<flat, PNT, 0, c->stream>>>(a, n, d_out);
    c->launches += 5;
    MCHECK(c, cudaGetLastError());
    return MANTIS_OK;
}
