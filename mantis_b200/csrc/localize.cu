// localize.cu -- mantis_localize_batch(): the whole quadDetection() callback (src/mantis3.cpp:68-135)
// for a batch of frames, plus the per-context tables it needs (world map, refinement
// perturbations, cell shifts).
#include <stdlib.h>
#include <new>
#include <thread>
#include <map>
#include <vector>
#include "ctx.h"
#include "cvrng.h"
#include "posemath.cuh"

static int upload(mantis_ctx* c, DevBuf& b, const void* src, size_t bytes)
{
    int rc = ctx_reserve(c, b, bytes ? bytes : 1);
    if (rc) return rc;
    if (bytes) MCHECK(c, cudaMemcpyAsync(b.p, src, bytes, cudaMemcpyHostToDevice, c->stream));
    MCHECK(c, cudaStreamSynchronize(c->stream));
    return MANTIS_OK;
}

// The map of params/map.yaml, as params/matGen_dense.m prints it ("%f") and
// parseCoordinatesFromString (Mantis3Params.h:125-152) reads it back (atof).
static double yaml_value(double v)
{
    char buf[64];
    snprintf(buf, sizeof(buf), "%f", v);
    return atof(buf);
}

static void default_map(double spacing, std::vector<double>& xyz, int& nw, int& nr, int& ng)
{
    auto push = [&](double x, double y) { xyz.push_back(yaml_value(x)); xyz.push_back(yaml_value(y)); xyz.push_back(0.0); };
    nw = nr = ng = 0;
    for (int iy = 0; iy < 8; ++iy)                       // 8 inner x-lines, 37 points each
        for (int ix = 0; ix < 37; ++ix) { push(spacing * (-4.5 + 0.25 * ix), spacing * (-3.5 + iy)); ++nw; }
    for (int ix = 0; ix < 10; ++ix)                      // 10 y-lines, 35 points each
        for (int iy = 0; iy < 35; ++iy) { push(spacing * (-4.5 + ix), spacing * (-4.25 + 0.25 * iy)); ++nw; }
    for (int ix = 0; ix < 37; ++ix) { push(spacing * (-4.5 + 0.25 * ix), -4.5 * spacing); ++nr; }  // red edge
    for (int ix = 0; ix < 37; ++ix) { push(spacing * (-4.5 + 0.25 * ix), 4.5 * spacing); ++ng; }   // green edge
}

int ctx_init_tables(mantis_ctx* c)
{
    // +90 degrees about z: tf::Transform(tf::Quaternion(0, 0, 1/sqrt(2), 1/sqrt(2)))
    double q[4] = {0, 0, 1 / sqrt(2.0), 1 / sqrt(2.0)};
    pm::quat_to_mat(q, c->rotz.R);
    c->rotz.t[0] = c->rotz.t[1] = c->rotz.t[2] = 0;

    // generateRandomHypothesis (PoseAdjustment.h:13-23) for every (iteration, particle), in draw
    // order; argument lists evaluated right to left (GCC, x86-64): yaw, pitch, roll, then z, y, x.
    const int total = c->cfg.iterations * c->cfg.particles;
    std::vector<Tf> rnd((size_t)(total > 0 ? total : 1));
    CvRng rng(c->cfg.rng_seed);
    for (int i = 0; i < total; ++i) {
        double yaw = rng.gaussian(c->cfg.rot_sigma), pitch = rng.gaussian(c->cfg.rot_sigma), roll = rng.gaussian(c->cfg.rot_sigma);
        double z = rng.gaussian(c->cfg.trans_sigma), y = rng.gaussian(c->cfg.trans_sigma), x = rng.gaussian(c->cfg.trans_sigma);
        pm::set_rpy(roll, pitch, yaw, rnd[i].R);
        rnd[i].t[0] = x; rnd[i].t[1] = y; rnd[i].t[2] = z;
    }
    int rc = upload(c, c->d_rand, rnd.data(), sizeof(Tf) * (size_t)total);
    if (rc) return rc;

    // computeAllShiftedHypothesesFAST (HypothesisGeneration.h:125-140): the loop variables are
    // accumulated in double exactly as written
    std::vector<double> sh;
    const double gs = c->cfg.grid_spacing, half = ((double)c->cfg.grid_size / 2.0) * gs;
    for (double x = -half + gs / 2.0; x < half; x += gs)
        for (double y = -half + gs / 2.0; y < half; y += gs) { sh.push_back(x); sh.push_back(y); }
    c->n_shift = (int)(sh.size() / 2);
    c->host_shifts = sh;
    if ((rc = upload(c, c->d_shifts, sh.data(), sizeof(double) * sh.size()))) return rc;

    {   // table of the fast atan in search.cu, rounded from long double
        std::vector<double> tab(2 * 128);
        const long double step = 3.14159265358979323846264338327950288L / 256.0L;
        for (int i = 0; i < 128; ++i) {
            tab[2 * i] = (double)(step * i);
            tab[2 * i + 1] = (double)tanl(step * i);
        }
        if ((rc = upload(c, c->d_atan, tab.data(), sizeof(double) * tab.size()))) return rc;
    }

    std::vector<double> xyz;
    int nw, nr, ng;
    default_map(c->cfg.grid_spacing, xyz, nw, nr, ng);
    return mantis_set_map(c, xyz.data(), nw, nr, ng);
}

// The shift hypotheses see map point p at p - shift (search.cu: score_points).  On a grid map most of those
// n_shift x n_map points coincide: list the distinct ones (to 1e-10 m; representatives are the first p - shift met,
// computed in double like any of them) and, per (shift, map point), the index of its representative, padded to
// whole warps with the index of an always-empty table entry.  A map that does not fold (more distinct points
// than the kernel's table holds) leaves n_shift_pts = 0 and every shift is projected on its own.
static int build_shift_tables(mantis_ctx* c)
{
    c->n_shift_pts = 0;
    const int n = c->n_map, ns = c->n_shift;
    if (n <= 0 || ns <= 0 || (int)c->host_shifts.size() < 2 * ns) return MANTIS_OK;
    constexpr int CAP = 6399;   // search.cu: SHIFT_PT_CAP
    const int chunks = (n + 31) / 32;
    struct Key { long long x, y, z; bool operator<(const Key& o) const { return x != o.x ? x < o.x : y != o.y ? y < o.y : z < o.z; } };
    std::map<Key, int> seen;
    std::vector<double> pts;
    std::vector<unsigned short> idx((size_t)ns * chunks * 32);
    for (int k = 0; k < ns; ++k) {
        const double sx = c->host_shifts[2 * k], sy = c->host_shifts[2 * k + 1];
        for (int i = 0; i < n; ++i) {
            const double x = c->host_map[3 * i] - sx, y = c->host_map[3 * i + 1] - sy, z = c->host_map[3 * i + 2];
            if (!(fabs(x) < 1e6 && fabs(y) < 1e6 && fabs(z) < 1e6)) return MANTIS_OK;
            const Key key{llround(x * 1e10), llround(y * 1e10), llround(z * 1e10)};
            auto it = seen.find(key);
            int q;
            if (it == seen.end()) {
                q = (int)(pts.size() / 3);
                if (q >= CAP) return MANTIS_OK;
                seen.emplace(key, q);
                pts.push_back(x); pts.push_back(y); pts.push_back(z);
            } else {
                q = it->second;
            }
            idx[((size_t)k * chunks) * 32 + i] = (unsigned short)q;
        }
    }
    const int nq = (int)(pts.size() / 3);
    // not worth a table unless it folds at least 4 : 1
    if ((long long)nq * 4 > (long long)ns * n) return MANTIS_OK;
    for (int k = 0; k < ns; ++k)
        for (int i = n; i < chunks * 32; ++i) idx[((size_t)k * chunks) * 32 + i] = (unsigned short)nq;
    int rc;
    if ((rc = upload(c, c->d_shift_pts, pts.data(), sizeof(double) * pts.size())) ||
        (rc = upload(c, c->d_shift_idx, idx.data(), sizeof(unsigned short) * idx.size())))
        return rc;
    c->n_shift_pts = nq;
    return MANTIS_OK;
}

extern "C" int mantis_set_map(mantis_ctx* c, const double* xyz, int n_white, int n_red, int n_green)
{
    if (!c || !xyz || n_white < 0 || n_red < 0 || n_green < 0) return MANTIS_ERR_INVALID;
    const int n = n_white + n_red + n_green;
    if (n <= 0 || n > 1024 || n_green > 40) return MANTIS_ERR_INVALID;
    MCHECK(c, cudaSetDevice(c->device));
    int rc = upload(c, c->d_map, xyz, sizeof(double) * 3 * (size_t)n);
    if (rc) return rc;
    c->n_map = n;
    c->n_green = n_green;
    c->n_red = n_red;
    c->host_map.assign(xyz, xyz + 3 * (size_t)n);
    for (int i = 0; i < mantis_ctx::MAX_LANES; ++i)
        if (c->lane[i] && (rc = mantis_set_map(c->lane[i], xyz, n_white, n_red, n_green))) return rc;
    c->map_planar = 1;
    for (int i = 0; i < n; ++i)
        if (xyz[3 * i + 2] != 0.0) c->map_planar = 0;
    return build_shift_tables(c);
}

// quadDetection() for n frames.  rig_cams == 0: every frame on its own (mantis_localize_batch).  rig_cams >= 1: the frames
// are rigs of rig_cams cameras localized jointly through their cam_to_base (mantis_localize_rig_batch); out / trace then
// hold one record per rig.
static int localize_impl(mantis_ctx* c, const mantis_frame* frames, int n, mantis_result* out, mantis_frame_trace* trace,
                         int mem, int rig_cams, bool async = false)
{
    if (!c || n < 0 || (n > 0 && (!frames || !out)) || (mem != MANTIS_MEM_HOST && mem != MANTIS_MEM_DEVICE))
        return MANTIS_ERR_INVALID;
    if (n == 0) return MANTIS_OK;
    const bool rig = rig_cams > 0;
    const int UC = rig ? rig_cams : 1;   // frames per unit of output
    const int W = frames[0].width, H = frames[0].height, stride0 = frames[0].stride_bytes;
    if (W <= 0 || H <= 0 || W > 32767 || H > 32767 || stride0 < 3 * W) return MANTIS_ERR_INVALID;
    for (int i = 0; i < n; ++i)
        if (!frames[i].bgr || frames[i].width != W || frames[i].height != H || frames[i].stride_bytes != stride0)
            return MANTIS_ERR_INVALID;
    // the camera model has no skew term (cv::fisheye applies alpha = K(0,1) / fx; the reference's cameras have 0):
    // refuse it instead of silently ignoring it
    for (int i = 0; i < n; ++i)
        if (frames[i].K[1] != 0.0f || !(frames[i].K[0] > 0.0f) || !(frames[i].K[4] > 0.0f)) {
            snprintf(c->err, sizeof(c->err), "frame %d: K must be upper triangular without skew (K[1] == 0) with positive focal lengths", i);
            return MANTIS_ERR_INVALID;
        }
    size_t dev_pitch = (size_t)stride0 * H;
    if (mem == MANTIS_MEM_DEVICE && n > 1) {
        dev_pitch = (size_t)(frames[1].bgr - frames[0].bgr);
        for (int i = 1; i < n; ++i)
            if (frames[i].bgr != frames[0].bgr + (size_t)i * dev_pitch) {
                snprintf(c->err, sizeof(c->err), "device-resident frames of one call must be equally spaced");
                return MANTIS_ERR_INVALID;
            }
    }
    MCHECK(c, cudaSetDevice(c->device));

    // Large device-resident batches CAN be split over a few lanes (child contexts: own stream, own scratch;
    // MANTIS_LANES >= 2) driven by one host thread each, so that the latency-bound kernels of one part (border
    // walks, RPP iterations) share the SMs with whatever the other parts are running.  That paid while those
    // kernels were slow; with the current ones a single lane is the fastest (29.3 / 30.1 / 30.5 ms per 4096
    // frames for 1 / 2 / 3 lanes), so the default is 1.  Per-stage profiling always uses the single-lane path.
    static const int lanes_env = []() { const char* e = getenv("MANTIS_LANES"); return e ? atoi(e) : 1; }();
    const int L = lanes_env > mantis_ctx::MAX_LANES ? mantis_ctx::MAX_LANES : lanes_env;
    if (!rig && !async && mem == MANTIS_MEM_DEVICE && !c->is_lane && !c->profiling && L >= 2 && n >= 512 * L) {
        for (int i = 0; i < L; ++i)
            if (!c->lane[i]) {
                int rc0 = mantis_create(&c->cfg, c->device, &c->lane[i]);
                if (rc0) { snprintf(c->err, sizeof(c->err), "cannot create execution lane %d", i); return rc0; }
                c->lane[i]->is_lane = true;
                if (!c->host_map.empty() &&
                    (rc0 = mantis_set_map(c->lane[i], c->host_map.data(), c->n_map - c->n_red - c->n_green, c->n_red, c->n_green)))
                    return rc0;
            }
        int rcs[mantis_ctx::MAX_LANES] = {};
        auto run = [&](int i) {
            const int f0 = (int)((long long)n * i / L), f1 = (int)((long long)n * (i + 1) / L);
            rcs[i] = mantis_localize_batch(c->lane[i], frames + f0, f1 - f0, out + f0, trace ? trace + f0 : nullptr, mem);
        };
        std::thread th[mantis_ctx::MAX_LANES];
        for (int i = 1; i < L; ++i) th[i] = std::thread(run, i);
        run(0);
        for (int i = 1; i < L; ++i) th[i].join();
        for (int i = 0; i < L; ++i)
            if (rcs[i]) { snprintf(c->err, sizeof(c->err), "lane %d: %s", i, c->lane[i]->err); return rcs[i]; }
        return MANTIS_OK;
    }

    // frames per pass of the kernel chain: large enough that the latency-bound tails (longest border
    // walk, slowest RPP run) are amortised over many frames, small enough that the per-frame scratch
    // (about 1.8 MB at 640x480, dominated by the contour buffers) stays within a few GB
    static const int chunk_env = []() { const char* e = getenv("MANTIS_CHUNK"); return e ? atoi(e) : 0; }();
    int CHUNK = chunk_env > 0 ? chunk_env : (int)(4096.0 * (640.0 * 480.0) / ((double)W * H));
    if (chunk_env <= 0) CHUNK = CHUNK < 256 ? 256 : CHUNK > 4096 ? 4096 : CHUNK;
    // Host frames: the step is bound by the H2D copy (3.8 GB per 4096 frames at 640x480), so what counts is that the
    // kernels of a chunk finish before the next copy does, and that the LAST chunk -- the only one whose kernels are
    // not hidden -- is short: 512-frame chunks, the final one split once more.
    if (mem == MANTIS_MEM_HOST && chunk_env <= 0 && CHUNK > 512) CHUNK = 512;
    CHUNK = CHUNK / UC * UC;             // whole rigs per chunk
    if (CHUNK < UC) CHUNK = UC;
    const int WW = words_per_row(W);
    const int nc_max = n < CHUNK ? n : CHUNK;
    const size_t plane_bytes = (size_t)nc_max * H * WW * sizeof(uint32_t);
    int rc;
    if ((rc = ctx_reserve(c, c->canny_bits, plane_bytes)) || (rc = ctx_reserve(c, c->morph_bits, plane_bytes)) ||
        (rc = ctx_reserve(c, c->mask_bits, plane_bytes)) ||
        (rc = ctx_reserve(c, c->frame_quads, (size_t)nc_max * sizeof(FrameQuads))) ||
        (rc = ctx_reserve(c, c->cams, (size_t)nc_max * sizeof(pm::Camera))) ||
        (rc = ctx_reserve(c, c->hyps, (size_t)nc_max * MANTIS_MAX_HYPS * sizeof(Hyp))) ||
        (rc = ctx_reserve(c, c->sel, (size_t)nc_max * MANTIS_MAX_HYPS * sizeof(short))) ||
        (rc = ctx_reserve(c, c->pose_info, (size_t)nc_max * sizeof(FramePose))) ||
        (rc = ctx_reserve(c, c->results, (size_t)nc_max * sizeof(mantis_result))) ||
        (rc = ctx_reserve(c, c->traces, (size_t)nc_max * sizeof(mantis_frame_trace))))
        return rc;
    // Host-resident frames: chunks are staged through two device buffers on a copy stream, so the
    // H2D transfer of chunk k+1 overlaps the kernels of chunk k.
    const bool host = mem == MANTIS_MEM_HOST;
    const size_t tight = (size_t)3 * W * H;
    bool contiguous = host && stride0 == 3 * W;
    if (host) {
        for (int i = 1; i < n && contiguous; ++i) contiguous = frames[i].bgr == frames[0].bgr + (size_t)i * tight;
        if ((rc = ctx_reserve(c, c->stage_in, (size_t)nc_max * tight)) ||
            (n > CHUNK && (rc = ctx_reserve(c, c->stage_in2, (size_t)nc_max * tight))))
            return rc;
        if (!c->copy_stream) {
            MCHECK(c, cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
            for (int i = 0; i < 2; ++i) {
                MCHECK(c, cudaEventCreateWithFlags(&c->ev_copied[i], cudaEventDisableTiming));
                MCHECK(c, cudaEventCreateWithFlags(&c->ev_free[i], cudaEventDisableTiming));
            }
        }
    }
    // chunk boundaries
    std::vector<int> start;
    for (int f0 = 0; f0 < n; f0 += CHUNK) {
        start.push_back(f0);
        const int nc = (n - f0 < CHUNK) ? (n - f0) : CHUNK;
        if (mem == MANTIS_MEM_HOST && f0 + CHUNK >= n && f0 > 0 && nc >= 256) start.push_back(f0 + nc / 2 / UC * UC);
    }
    start.push_back(n);
    const int n_chunks = (int)start.size() - 1;
    auto issue_copy = [&](int k) -> int {
        const int f0 = start[k], nc = start[k + 1] - f0, buf = k & 1;
        uint8_t* dst = (uint8_t*)(buf ? c->stage_in2.p : c->stage_in.p);
        if (k >= 2) MCHECK(c, cudaStreamWaitEvent(c->copy_stream, c->ev_free[buf], 0));
        if (contiguous) {
            MCHECK(c, cudaMemcpyAsync(dst, frames[f0].bgr, (size_t)nc * tight, cudaMemcpyHostToDevice, c->copy_stream));
        } else {
            for (int i = 0; i < nc; ++i)
                MCHECK(c, cudaMemcpy2DAsync(dst + (size_t)i * tight, 3 * W, frames[f0 + i].bgr, stride0, (size_t)3 * W, H,
                                            cudaMemcpyHostToDevice, c->copy_stream));
        }
        MCHECK(c, cudaEventRecord(c->ev_copied[buf], c->copy_stream));
        return MANTIS_OK;
    };
    std::vector<pm::Camera> cams((size_t)nc_max);
    std::vector<Tf> rig_tf;
    // The chunk loop as one unit: whatever makes it stop early, no copy may still be reading the caller's frames
    // (or writing the caller's results) once this call has returned.
    auto run_chunks = [&]() -> int {
    if (host && (rc = issue_copy(0))) return rc;
    for (int k = 0; k < n_chunks; ++k) {
        const int f0 = start[k], nc = start[k + 1] - f0;
        const uint8_t* d_in;
        size_t pitch;
        int stride;
        prof_mark(c, 0);
        if (host) {
            if (k + 1 < n_chunks && (rc = issue_copy(k + 1))) return rc;
            MCHECK(c, cudaStreamWaitEvent(c->stream, c->ev_copied[k & 1], 0));
            stride = 3 * W;
            pitch = tight;
            d_in = (const uint8_t*)((k & 1) ? c->stage_in2.p : c->stage_in.p);
        } else {
            stride = stride0;
            pitch = dev_pitch;
            d_in = frames[f0].bgr;
        }
        for (int i = 0; i < nc; ++i) {
            const mantis_frame& f = frames[f0 + i];
            pm::Camera& cm = cams[i];
            cm.fx = (double)f.K[0]; cm.fy = (double)f.K[4]; cm.cx = (double)f.K[2]; cm.cy = (double)f.K[5];
            for (int k = 0; k < 4; ++k) cm.D[k] = f.D[k];
        }
        MCHECK(c, cudaMemcpyAsync(c->cams.p, cams.data(), sizeof(pm::Camera) * (size_t)nc, cudaMemcpyHostToDevice, c->stream));
        const Tf* d_b2c = nullptr;
        const Tf* d_b2c_inv = nullptr;
        if (rig) {   // per frame: the camera in the base frame (MantisRequest.h:14-26's b2c) and its inverse
            rig_tf.resize(2 * (size_t)nc);
            for (int i = 0; i < nc; ++i) {
                const double* q = frames[f0 + i].cam_to_base;
                Tf& b = rig_tf[i];
                if (q[3] == 0 && q[4] == 0 && q[5] == 0 && q[6] == 0) {   // not given: the camera is the base
                    for (int k = 0; k < 9; ++k) b.R[k] = (k % 4 == 0) ? 1.0 : 0.0;
                    b.t[0] = b.t[1] = b.t[2] = 0;
                } else {
                    pm::quat_to_mat(q + 3, b.R);
                    for (int k = 0; k < 3; ++k) b.t[k] = q[k];
                }
                pm::tf_inv(b, rig_tf[(size_t)nc + i]);
            }
            if ((rc = ctx_reserve(c, c->rig_tf, sizeof(Tf) * 2 * (size_t)nc_max))) return rc;
            MCHECK(c, cudaMemcpyAsync(c->rig_tf.p, rig_tf.data(), sizeof(Tf) * 2 * (size_t)nc, cudaMemcpyHostToDevice, c->stream));
            d_b2c = (const Tf*)c->rig_tf.p;
            d_b2c_inv = d_b2c + nc;
        }

        prof_mark(c, 1);
        EdgeOutputs eo;
        eo.canny_bits = (uint32_t*)c->canny_bits.p;
        eo.morph_bits = (uint32_t*)c->morph_bits.p;
        eo.mask_bits = (uint32_t*)c->mask_bits.p;
        if ((rc = run_edge_stage(c, d_in, nc, W, H, stride, pitch, eo))) return rc;
        prof_mark(c, 3);
        if ((rc = launch_quads(c, eo.morph_bits, nc, W, H, (FrameQuads*)c->frame_quads.p))) return rc;
        prof_mark(c, 4);
        if ((rc = launch_pose(c, (const FrameQuads*)c->frame_quads.p, (const pm::Camera*)c->cams.p, nc, (Hyp*)c->hyps.p,
                              (short*)c->sel.p, (FramePose*)c->pose_info.p, UC, d_b2c_inv)))
            return rc;
        prof_mark(c, 5);
        if (rig)
            rc = launch_search_rig(c, d_in, pitch, stride, eo.mask_bits, nc / UC, UC, W, H, (const pm::Camera*)c->cams.p,
                                   (const FrameQuads*)c->frame_quads.p, (const Hyp*)c->hyps.p, (const short*)c->sel.p,
                                   (const FramePose*)c->pose_info.p, d_b2c, (mantis_result*)c->results.p,
                                   trace ? (mantis_frame_trace*)c->traces.p : nullptr);
        else
            rc = launch_search(c, d_in, pitch, stride, eo.mask_bits, nc, W, H, (const pm::Camera*)c->cams.p,
                               (const FrameQuads*)c->frame_quads.p, (const Hyp*)c->hyps.p, (const short*)c->sel.p,
                               (const FramePose*)c->pose_info.p, (mantis_result*)c->results.p,
                               trace ? (mantis_frame_trace*)c->traces.p : nullptr);
        if (rc) return rc;
        prof_mark(c, 6);
        if (host) MCHECK(c, cudaEventRecord(c->ev_free[k & 1], c->stream));
        MCHECK(c, cudaMemcpyAsync(out + f0 / UC, c->results.p, sizeof(mantis_result) * (size_t)(nc / UC), cudaMemcpyDeviceToHost,
                                  c->stream));
        if (trace)
            MCHECK(c, cudaMemcpyAsync(trace + f0 / UC, c->traces.p, sizeof(mantis_frame_trace) * (size_t)(nc / UC),
                                      cudaMemcpyDeviceToHost, c->stream));
        prof_mark(c, 7);
        // Everything above is ordered by the streams (scratch is reused chunk after chunk on c->stream, the staging
        // buffers through events), and the runtime has staged the pageable host sources (camera tables) before each
        // cudaMemcpyAsync returned: the host only has to wait when the caller wants the results now.
        if (!async) {
            MCHECK(c, cudaStreamSynchronize(c->stream));
            prof_collect(c, 0, MANTIS_N_STAGES);
        }
    }
    return MANTIS_OK;
    };
    rc = run_chunks();
    if (rc != MANTIS_OK) {  // drain both streams; the first error text is the one worth keeping
        if (host && c->copy_stream) cudaStreamSynchronize(c->copy_stream);
        cudaStreamSynchronize(c->stream);
    }
    return rc;
}

extern "C" int mantis_localize_batch(mantis_ctx* c, const mantis_frame* frames, int n, mantis_result* out,
                                     mantis_frame_trace* trace, int mem)
{
    return localize_impl(c, frames, n, out, trace, mem, 0);
}

extern "C" int mantis_localize_rig_batch(mantis_ctx* c, const mantis_frame* frames, int n_rigs, int cams_per_rig,
                                         mantis_result* out, mantis_frame_trace* trace, int mem)
{
    if (cams_per_rig < 1 || cams_per_rig > MANTIS_MAX_RIG_CAMERAS || n_rigs < 0 || (long long)n_rigs * cams_per_rig > 0x7fffffffLL)
        return MANTIS_ERR_INVALID;
    return localize_impl(c, frames, n_rigs * cams_per_rig, out, trace, mem, cams_per_rig);
}

// Asynchronous form: returns once the work is queued on the context's streams.  out / trace (and host frames) must
// stay valid, and the context must not be used for anything else, until mantis_wait() returns.
extern "C" int mantis_localize_batch_async(mantis_ctx* c, const mantis_frame* frames, int n, mantis_result* out,
                                           mantis_frame_trace* trace, int mem)
{
    if (c && c->profiling) return MANTIS_ERR_INVALID;   // per-stage timing needs the synchronous call
    return localize_impl(c, frames, n, out, trace, mem, 0, true);
}

extern "C" int mantis_wait(mantis_ctx* c)
{
    if (!c) return MANTIS_ERR_INVALID;
    MCHECK(c, cudaSetDevice(c->device));
    if (c->copy_stream) MCHECK(c, cudaStreamSynchronize(c->copy_stream));
    MCHECK(c, cudaStreamSynchronize(c->stream));
    return MANTIS_OK;
}
