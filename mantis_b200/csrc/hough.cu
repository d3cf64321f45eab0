// hough.cu -- K6: line voting, peak picking, line clustering and grid intersections (row a6').
//
// The reference has no Hough transform (its grid primitives are the cell quads of quads.cu,
// include/mantis3/QuadDetection.h:216-226); this stage realises the north star's wording with the
// semantics of cv::HoughLines(edges, 1, pi/180, threshold) of the OpenCV the reference links:
//   r = cvRound(x * cos_f32[n] + y * sin_f32[n]) + (numrho - 1) / 2,  accum[(n+1)(numrho+2) + r + 1]++
//   peaks: > threshold, > left, >= right (rho direction), > previous angle, >= next angle
//   sorted by votes descending, accumulator index ascending.
// Clustering / intersections are defined by oracle/hough.c (no third-party counterpart).
//
// k_hough_vote   one CTA per (frame, slab of angles).  The slab's accumulator rows (+1 halo angle on
//                each side, recomputed) live in shared memory; edge pixels are compacted from the
//                bit plane in chunks; a warp votes 32 pixels for one angle at a time and aggregates
//                equal rho cells with __match_any_sync so that one lane issues the shared atomic.
// k_hough_lines  one CTA per frame: bitonic sort of the peaks, then one warp does the sequential
//                greedy clustering (lanes search the seeds in parallel) and the pencil crossings.
#include <math.h>
#include "ctx.h"

namespace {

constexpr int HNT = 512;          // threads per voting CTA
constexpr int LIST_CAP = 8192;    // compacted edge pixels per chunk (= 256 words)
constexpr int PEAK_CAP = 4096;    // peaks kept per frame
constexpr int MAX_ANGLES = 180;
constexpr double HPI = 3.1415926535897932384626433832795;

struct HoughGeom {
    int W, H, WW, numangle, numrho, accw;  // accw = numrho + 2
    int slab, nslabs, rows_cap;            // angles per slab, slabs per frame, accumulator rows in smem
    int threshold;
};

__global__ void __launch_bounds__(HNT)
k_hough_vote(const uint32_t* __restrict__ bits, HoughGeom g, const float* __restrict__ trig /* cos[180], sin[180] */,
             int* __restrict__ peak_count, uint2* __restrict__ peaks)
{
    extern __shared__ uint32_t hsm[];
    uint32_t* acc = hsm;                                  // rows_cap x accw
    uint32_t* list = hsm + (size_t)g.rows_cap * g.accw;   // LIST_CAP packed (x | y << 16)
    __shared__ int s_count;
    __shared__ float s_cos[MAX_ANGLES + 2], s_sin[MAX_ANGLES + 2];

    const int frame = blockIdx.y, slab = blockIdx.x;
    const int a0 = slab * g.slab;                          // first interior angle
    const int a1 = min(a0 + g.slab, g.numangle);           // one past the last interior angle
    const int lo = a0 - 1;                                 // angle of accumulator row 0
    const int nrows = a1 - a0 + 2;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t* plane = bits + (size_t)frame * g.H * g.WW;

    for (int i = tid; i < nrows * g.accw; i += HNT) acc[i] = 0;
    for (int i = tid; i < nrows; i += HNT) {
        int n = lo + i;
        bool ok = n >= 0 && n < g.numangle;
        s_cos[i] = ok ? trig[n] : 0.f;
        s_sin[i] = ok ? trig[MAX_ANGLES + n] : 0.f;
    }
    const int total_words = g.H * g.WW;
    const int words_per_chunk = LIST_CAP / 32;
    const int half = (g.numrho - 1) / 2;

    for (int w0 = 0; w0 < total_words; w0 += words_per_chunk) {
        if (tid == 0) s_count = 0;
        __syncthreads();
        // compaction: each warp takes 32 words at a time, lane = word
        for (int wb = w0 + warp * 32; wb < min(w0 + words_per_chunk, total_words); wb += (HNT / 32) * 32) {
            int wi = wb + lane;
            uint32_t v = (wi < min(w0 + words_per_chunk, total_words)) ? plane[wi] : 0u;
            int c = __popc(v);
            int incl = c;
            for (int d = 1; d < 32; d <<= 1) {
                int t = __shfl_up_sync(0xffffffffu, incl, d);
                if (lane >= d) incl += t;
            }
            int tot = __shfl_sync(0xffffffffu, incl, 31);
            int base = 0;
            if (lane == 31 && tot) base = atomicAdd(&s_count, tot);
            base = __shfl_sync(0xffffffffu, base, 31);
            int o = base + incl - c;
            int y = wi / g.WW, xw = (wi - y * g.WW) << 5;
            while (v) {
                int b = __ffs(v) - 1;
                v &= v - 1;
                list[o++] = (uint32_t)(xw + b) | ((uint32_t)y << 16);
            }
        }
        __syncthreads();
        const int npx = s_count;
        const int ngroups = (npx + 31) >> 5;
        // votes: work item = (pixel group, accumulator row); halo rows outside [0, numangle) stay zero
        for (int it = warp; it < ngroups * nrows; it += HNT / 32) {
            int grp = it / nrows, k = it - grp * nrows;
            int n = lo + k;
            if (n < 0 || n >= g.numangle) continue;
            int pi = grp * 32 + lane;
            bool valid = pi < npx;
            uint32_t p = valid ? list[pi] : 0u;
            float a = __fmul_rn((float)(p & 0xffffu), s_cos[k]);
            float b = __fmul_rn((float)(p >> 16), s_sin[k]);
            int r = __float2int_rn(__fadd_rn(a, b)) + half;
            if (!valid) r = -1 - lane;  // unique keys: never aggregated with a real vote
            uint32_t peers = __match_any_sync(0xffffffffu, r);
            if (valid && lane == __ffs(peers) - 1) atomicAdd(&acc[k * g.accw + r + 1], (uint32_t)__popc(peers));
        }
        __syncthreads();
    }

    // peaks of the interior rows
    int* pc = peak_count + frame;
    uint2* pk = peaks + (size_t)frame * PEAK_CAP;
    const int ni = a1 - a0;
    for (int i = tid; i < ni * g.numrho; i += HNT) {
        int k = i / g.numrho + 1, r = i - (k - 1) * g.numrho;
        const uint32_t* c = acc + k * g.accw + r + 1;
        int v = (int)c[0];
        if (v > g.threshold && v > (int)c[-1] && v >= (int)c[1] && v > (int)c[-g.accw] && v >= (int)c[g.accw]) {
            int slot = atomicAdd(pc, 1);
            if (slot < PEAK_CAP) pk[slot] = make_uint2((uint32_t)v, (uint32_t)((lo + k + 1) * g.accw + r + 1));
        }
    }
}

__device__ __forceinline__ double ang_dist(double a, double b)
{
    double d = fabs(a - b);
    d = fmod(d, HPI);
    return d > HPI / 2 ? HPI - d : d;
}

struct LineParams {
    double rho_tol, theta_tol, pencil_tol, theta_step;
};

__global__ void __launch_bounds__(256)
k_hough_lines(const int* __restrict__ peak_count, const uint2* __restrict__ peaks, HoughGeom g, LineParams lp,
              mantis_frame_lines* __restrict__ out)
{
    __shared__ unsigned long long s_key[PEAK_CAP];
    __shared__ int s_seed_a[MANTIS_HOUGH_MAX_CLUSTERS];
    __shared__ long long s_seed_r[MANTIS_HOUGH_MAX_CLUSTERS];
    __shared__ long long s_v[MANTIS_HOUGH_MAX_CLUSTERS], s_r[MANTIS_HOUGH_MAX_CLUSTERS], s_a[MANTIS_HOUGH_MAX_CLUSTERS];
    __shared__ int s_members[MANTIS_HOUGH_MAX_CLUSTERS];
    __shared__ signed char s_pen[MANTIS_HOUGH_MAX_CLUSTERS];

    const int frame = blockIdx.x, tid = threadIdx.x;
    mantis_frame_lines* o = out + frame;
    const int total = peak_count[frame];
    const int n = min(total, PEAK_CAP);
    // key: votes descending, accumulator index ascending
    int m = 1;
    while (m < n) m <<= 1;
    for (int i = tid; i < m; i += blockDim.x) {
        unsigned long long k = ~0ull;
        if (i < n) {
            uint2 p = peaks[(size_t)frame * PEAK_CAP + i];
            k = ((unsigned long long)(0x7fffffffu - p.x) << 32) | p.y;
        }
        s_key[i] = k;
    }
    __syncthreads();
    for (int size = 2; size <= m; size <<= 1)
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            for (int i = tid; i < (m >> 1); i += blockDim.x) {
                int lo_i = ((i & ~(stride - 1)) << 1) | (i & (stride - 1));
                int hi_i = lo_i | stride;
                bool up = (lo_i & size) == 0;
                unsigned long long x = s_key[lo_i], y = s_key[hi_i];
                if ((x > y) == up) { s_key[lo_i] = y; s_key[hi_i] = x; }
            }
            __syncthreads();
        }
    const int nl = min(n, MANTIS_HOUGH_MAX_LINES);
    for (int i = tid; i < nl; i += blockDim.x) {
        unsigned long long k = s_key[i];
        int idx = (int)(uint32_t)k, votes = (int)(0x7fffffffu - (uint32_t)(k >> 32));
        int a = idx / g.accw - 1;
        int r = idx - (a + 1) * g.accw - 1;
        o->lines[i].rho = __fmul_rn((float)r - __fmul_rn((float)(g.numrho - 1), 0.5f), 1.0f);
        o->lines[i].theta = __fmul_rn((float)a, (float)lp.theta_step);
        o->lines[i].votes = votes;
    }
    if (tid >= 32) return;

    // ---- greedy clustering: sequential over lines, lanes search the seeds ----
    const int lane = tid;
    int nc = 0;
    const int numangle = g.numangle;
    const int half = (g.numrho - 1) / 2;
    for (int i = 0; i < nl; ++i) {
        unsigned long long k = s_key[i];
        int idx = (int)(uint32_t)k, votes = (int)(0x7fffffffu - (uint32_t)(k >> 32));
        int ai = idx / g.accw - 1;
        long long r2 = 2ll * ((idx - (ai + 1) * g.accw - 1) - half);
        int found = -1;
        for (int c0 = 0; c0 < nc && found < 0; c0 += 32) {
            int c = c0 + lane;
            bool hit = false;
            if (c < nc) {
                int as = s_seed_a[c];
                long long rs2 = s_seed_r[c];
                int a = ai;
                long long r = r2;
                if (abs(a - as) * 2 > numangle) { a = a > as ? a - numangle : a + numangle; r = -r; }
                hit = fabs((double)(a - as)) * lp.theta_step <= lp.theta_tol &&
                      fabs((double)(r - rs2)) * 0.5 <= lp.rho_tol;
            }
            uint32_t bal = __ballot_sync(0xffffffffu, hit);
            if (bal) found = c0 + __ffs(bal) - 1;
        }
        if (lane == 0) {
            int a_use = ai;
            long long r_use = r2;
            if (found < 0) {
                if (nc < MANTIS_HOUGH_MAX_CLUSTERS) {
                    found = nc;
                    s_seed_a[found] = ai; s_seed_r[found] = r2;
                    s_v[found] = 0; s_r[found] = 0; s_a[found] = 0; s_members[found] = 0;
                }
            } else {
                int as = s_seed_a[found];
                if (abs(ai - as) * 2 > numangle) { a_use = ai > as ? ai - numangle : ai + numangle; r_use = -r2; }
            }
            if (found >= 0) {
                s_v[found] += votes;
                s_r[found] += (long long)votes * r_use;
                s_a[found] += (long long)votes * a_use;
                s_members[found]++;
            }
        }
        found = __shfl_sync(0xffffffffu, found, 0);
        if (found == nc && nc < MANTIS_HOUGH_MAX_CLUSTERS) nc++;
        __syncwarp();
    }
    for (int c = lane; c < nc; c += 32) {
        o->clusters[c].rho = (float)((double)s_r[c] / (2.0 * (double)s_v[c]));
        o->clusters[c].theta = (float)(((double)s_a[c] / (double)s_v[c]) * lp.theta_step);
        o->clusters[c].votes = (int)s_v[c];
        o->clusters[c].members = s_members[c];
    }
    __syncwarp();

    // ---- pencils and their crossings (a-major order) ----
    int npts = 0;
    if (nc > 0) {
        int sa = 0, sb = -1;
        for (int i = 1; i < nc; ++i)
            if (s_v[i] > s_v[sa]) sa = i;
        const double th_a = (double)o->clusters[sa].theta;
        for (int i = 0; i < nc; ++i)
            if (ang_dist((double)o->clusters[i].theta, th_a) >= HPI / 4 && (sb < 0 || s_v[i] > s_v[sb])) sb = i;
        const double th_b = sb >= 0 ? (double)o->clusters[sb].theta : 0.0;
        for (int i = lane; i < nc; i += 32) {
            double th = (double)o->clusters[i].theta;
            int p = -1;
            if (ang_dist(th, th_a) <= lp.pencil_tol) p = 0;
            else if (sb >= 0 && ang_dist(th, th_b) <= lp.pencil_tol) p = 1;
            s_pen[i] = (signed char)p;
        }
        __syncwarp();
        for (int a = 0; a < nc; ++a) {
            if (s_pen[a] != 0) continue;
            const double tha = (double)o->clusters[a].theta, ra = (double)o->clusters[a].rho;
            const double ca = cos(tha), sna = sin(tha);
            for (int b0 = 0; b0 < nc; b0 += 32) {
                int b = b0 + lane;
                bool ok = false;
                double x = 0, y = 0;
                if (b < nc && s_pen[b] == 1) {
                    const double thb = (double)o->clusters[b].theta, rb = (double)o->clusters[b].rho;
                    const double cb = cos(thb), snb = sin(thb);
                    double det = ca * snb - sna * cb;
                    if (!(fabs(det) < 1e-9)) {
                        x = (ra * snb - rb * sna) / det;
                        y = (rb * ca - ra * cb) / det;
                        ok = (x >= 0 && x < g.W && y >= 0 && y < g.H);
                    }
                }
                uint32_t bal = __ballot_sync(0xffffffffu, ok);
                int pos = npts + __popc(bal & ((1u << lane) - 1u));
                if (ok && pos < MANTIS_HOUGH_MAX_POINTS) { o->points[pos][0] = (float)x; o->points[pos][1] = (float)y; }
                npts += __popc(bal);
            }
        }
    }
    if (lane == 0) {
        o->n_peaks = total;
        o->n_lines = nl;
        o->n_clusters = nc;
        o->n_points = min(npts, MANTIS_HOUGH_MAX_POINTS);
        o->flags = total > PEAK_CAP ? MANTIS_HOUGH_FLAG_PEAK_OVERFLOW : 0;
        o->pad_ = 0;
    }
}

// cv::HoughLines geometry: numangle for [0, pi) at `theta`, numrho = round((2 (W + H) + 1) / rho)
int hough_numangle(double theta)
{
    int numangle = (int)floor(HPI / theta) + 1;
    if (numangle > 1 && fabs(HPI - (numangle - 1) * theta) < theta / 2) --numangle;
    return numangle;
}

}  // namespace

extern "C" int mantis_hough_batch(mantis_ctx* c, const uint8_t* bgr, int n, int width, int height, int stride_bytes,
                                  size_t frame_pitch_bytes, const mantis_hough_params* prm, mantis_frame_lines* out,
                                  int mem)
{
    int rc = check_frames(c, bgr, n, width, height, stride_bytes, frame_pitch_bytes, mem);
    if (rc) return rc;
    if (!out || !prm || prm->threshold < 0) return MANTIS_ERR_INVALID;
    if (n == 0) return MANTIS_OK;
    MCHECK(c, cudaSetDevice(c->device));

    const double theta = HPI / 180;
    HoughGeom g;
    g.W = width; g.H = height; g.WW = words_per_row(width);
    g.numangle = hough_numangle(theta);
    g.numrho = (int)lrint((double)(2 * (width + height) + 1));
    g.accw = g.numrho + 2;
    g.threshold = prm->threshold;
    if (g.numangle > MAX_ANGLES) return MANTIS_ERR_UNSUPPORTED;
    // accumulator rows that fit beside the pixel list in one CTA's shared memory
    const size_t budget = 200 * 1024 - LIST_CAP * sizeof(uint32_t);
    g.rows_cap = (int)(budget / ((size_t)g.accw * sizeof(uint32_t)));
    if (g.rows_cap > 32) g.rows_cap = 32;
    if (g.rows_cap < 3) {
        snprintf(c->err, sizeof(c->err), "frame %dx%d too large for the on-chip Hough accumulator", width, height);
        return MANTIS_ERR_UNSUPPORTED;
    }
    g.slab = g.rows_cap - 2;
    g.nslabs = (g.numangle + g.slab - 1) / g.slab;
    // balance the slabs
    g.slab = (g.numangle + g.nslabs - 1) / g.nslabs;
    g.rows_cap = g.slab + 2;
    const size_t smem = ((size_t)g.rows_cap * g.accw + LIST_CAP) * sizeof(uint32_t);

    // trig table exactly as cv::HoughLines builds it (float angle accumulated in float)
    if (!c->hough_trig.p) {
        float tab[2 * MAX_ANGLES];
        float ang = 0.f;
        for (int i = 0; i < g.numangle; ang += (float)theta, i++) {
            tab[i] = (float)(cos((double)ang) * 1.0f);
            tab[MAX_ANGLES + i] = (float)(sin((double)ang) * 1.0f);
        }
        if ((rc = ctx_reserve(c, c->hough_trig, sizeof(tab)))) return rc;
        MCHECK(c, cudaMemcpyAsync(c->hough_trig.p, tab, sizeof(tab), cudaMemcpyHostToDevice, c->stream));
        MCHECK(c, cudaStreamSynchronize(c->stream));  // tab is a stack array
    }

    const uint8_t* d_in;
    if ((rc = stage_input(c, bgr, n, width, height, stride_bytes, frame_pitch_bytes, mem, &d_in))) return rc;
    const size_t plane_bytes = (size_t)n * height * g.WW * sizeof(uint32_t);
    EdgeOutputs eo;
    if (prm->use_morph) {
        if ((rc = ctx_reserve(c, c->morph_bits, plane_bytes))) return rc;
        eo.morph_bits = (uint32_t*)c->morph_bits.p;
    } else {
        if ((rc = ctx_reserve(c, c->canny_bits, plane_bytes))) return rc;
        eo.canny_bits = (uint32_t*)c->canny_bits.p;
    }
    if ((rc = run_edge_stage(c, d_in, n, width, height, stride_bytes, frame_pitch_bytes, eo))) return rc;
    const uint32_t* d_bits = prm->use_morph ? eo.morph_bits : eo.canny_bits;

    if ((rc = ctx_reserve(c, c->hough_count, (size_t)n * sizeof(int)))) return rc;
    if ((rc = ctx_reserve(c, c->hough_peaks, (size_t)n * PEAK_CAP * sizeof(uint2)))) return rc;
    mantis_frame_lines* d_out = out;
    if (mem == MANTIS_MEM_HOST) {
        if ((rc = ctx_reserve(c, c->hough_lines, (size_t)n * sizeof(mantis_frame_lines)))) return rc;
        d_out = (mantis_frame_lines*)c->hough_lines.p;
    }
    MCHECK(c, cudaMemsetAsync(c->hough_count.p, 0, (size_t)n * sizeof(int), c->stream));
    MCHECK(c, cudaFuncSetAttribute(k_hough_vote, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    LineParams lp;
    lp.rho_tol = prm->rho_tol; lp.theta_tol = prm->theta_tol; lp.pencil_tol = prm->pencil_tol; lp.theta_step = theta;
    for (int f0 = 0; f0 < n; f0 += 32768) {
        int nf = (n - f0 < 32768) ? (n - f0) : 32768;
        k_hough_vote<<<dim3(g.nslabs, nf), HNT, smem, c->stream>>>(
            d_bits + (size_t)f0 * height * g.WW, g, (const float*)c->hough_trig.p, (int*)c->hough_count.p + f0,
            (uint2*)c->hough_peaks.p + (size_t)f0 * PEAK_CAP);
        c->launches++;
        MCHECK(c, cudaGetLastError());
        k_hough_lines<<<nf, 256, 0, c->stream>>>((const int*)c->hough_count.p + f0,
                                                 (const uint2*)c->hough_peaks.p + (size_t)f0 * PEAK_CAP, g, lp,
                                                 d_out + f0);
        c->launches++;
        MCHECK(c, cudaGetLastError());
    }
    if (mem == MANTIS_MEM_HOST)
        MCHECK(c, cudaMemcpyAsync(out, d_out, (size_t)n * sizeof(mantis_frame_lines), cudaMemcpyDeviceToHost, c->stream));
    MCHECK(c, cudaStreamSynchronize(c->stream));
    return MANTIS_OK;
}
