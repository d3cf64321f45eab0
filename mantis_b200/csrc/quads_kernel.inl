// quads_kernel.inl -- the per-frame quad kernel and its launcher (included by quads.cu).
namespace {

struct QuadScratch {   // per resident CTA
    short2* pts;       // MANTIS_MAX_PTS
    short2* dst;       // MANTIS_MAX_PTS
    int2* stack;       // MANTIS_MAX_PTS + 2 * MAXB
    uint16_t* owner;   // W * H
};

struct QuadArgs {
    const uint32_t* morph_bits;
    int W, H, WW, n_frames;
    short2* pts; short2* dst; int2* stack; uint16_t* owner;  // slot-strided scratch
    FrameQuads* out;
    double eps, stretch, radius_mult;
    long long* phase_clocks;  // optional: 8 accumulated clock64 deltas (debug / tuning)
};

__device__ __forceinline__ uint32_t bits_at(const uint32_t* p, int WW, int H, int y, int j)
{
    if (y < 0 || y >= H || j < 0 || j >= WW) return 0u;
    return __ldg(p + (size_t)y * WW + j);
}

__global__ void __launch_bounds__(QNT) k_quads(QuadArgs a)
{
    __shared__ uint32_t s_cand[MAXCAND];
    __shared__ uint32_t s_ukey[MAXB];     // unsorted starts
    __shared__ uint32_t s_unp[MAXB];
    __shared__ uint32_t s_key[MAXB];      // sorted by discovery order
    __shared__ int s_np[MAXB], s_off[MAXB];
    __shared__ short s_an[MAXB];          // approxPolyDP vertex count
    __shared__ short s_parent[MAXB];      // holes: rank of the owning outer border
    __shared__ short s_pos2b[MAXB];       // CCOMP output position -> border
    short2 (*s_quad)[4] = reinterpret_cast<short2(*)[4]>(s_cand);  // the candidate queue is dead by then
    static_assert(sizeof(short2) * 4 * MAXB <= sizeof(uint32_t) * MAXCAND, "alias");
    __shared__ int s_ncand, s_nstart, s_flags;

    const int W = a.W, H = a.H, WW = a.WW;
    const int slot = blockIdx.x;
    short2* pts = a.pts + (size_t)slot * MANTIS_MAX_PTS;
    short2* dst = a.dst + (size_t)slot * MANTIS_MAX_PTS;
    int2* stack = a.stack + (size_t)slot * (MANTIS_MAX_PTS + 2 * MAXB);
    uint16_t* owner = a.owner + (size_t)slot * W * H;
    const int tid = threadIdx.x;

    for (int frame = blockIdx.x; frame < a.n_frames; frame += gridDim.x) {
        BitImg im;
        im.bits = a.morph_bits + (size_t)frame * H * WW;
        im.W = W; im.H = H; im.WW = WW;
        FrameQuads* out = a.out + frame;
        __syncthreads();
        if (tid == 0) { s_ncand = 0; s_nstart = 0; s_flags = 0; }
        __syncthreads();

        auto try_candidate = [&](uint32_t cand) {
            const bool is_hole = cand >> 31;
            const int y = (cand >> 15) & 0xffff, x = cand & 0x7fff;
            int np;
            bool ok;
            if (!is_hole) {
                OuterProbe v{x, y, 0};
                ok = follow_border(im, x, y, false, v);
                np = v.npts;
            } else {
                HoleProbe v{x, y, W, H, 0};
                ok = follow_border(im, x - 1, y, true, v);
                np = v.npts;
            }
            if (!ok) return;
            int k = atomicAdd(&s_nstart, 1);
            if (k >= MAXB) { atomicOr(&s_flags, FRAME_FLAG_BORDER_OVERFLOW); return; }
            s_ukey[k] = ((uint32_t)(y * (W + 1) + x) << 1) | (is_hole ? 1u : 0u);
            s_unp[k] = np;
        };

        long long t_prev = clock64();
        auto lap = [&](int k) { if (a.phase_clocks && tid == 0) { long long t = clock64(); atomicAdd((unsigned long long*)&a.phase_clocks[k], (unsigned long long)(t - t_prev)); t_prev = t; } };
        // 1. candidates (whole words at a time)
        for (int i = tid; i < H * WW; i += QNT) {
            int y = i / WW, j = i - y * WW;
            uint32_t c = bits_at(im.bits, WW, H, y, j);
            uint32_t cp = bits_at(im.bits, WW, H, y, j - 1);
            uint32_t u = bits_at(im.bits, WW, H, y - 1, j), up = bits_at(im.bits, WW, H, y - 1, j - 1),
                     un = bits_at(im.bits, WW, H, y - 1, j + 1);
            uint32_t left = (c << 1) | (cp >> 31);
            uint32_t u_l = (u << 1) | (up >> 31), u_r = (u >> 1) | (un << 31);
            uint32_t valid = (j == WW - 1 && (W & 31)) ? ((1u << (W & 31)) - 1u) : 0xffffffffu;
            uint32_t outer = c & ~left & ~u & ~u_l & ~u_r;
            uint32_t hole = ~c & left & u & valid;
            while (outer | hole) {
                bool is_hole = outer == 0;
                uint32_t& m = is_hole ? hole : outer;
                int bit = __ffs(m) - 1;
                m &= m - 1;
                uint32_t cand = (is_hole ? 0x80000000u : 0u) | ((uint32_t)y << 15) | (uint32_t)(j * 32 + bit);
                int k = atomicAdd(&s_ncand, 1);
                if (k < MAXCAND) s_cand[k] = cand;
                else try_candidate(cand);
            }
        }
        __syncthreads();
        lap(0);
        // 2. probe walks
        {
            int nc = min(s_ncand, MAXCAND);
            for (int i = tid; i < nc; i += QNT) try_candidate(s_cand[i]);
        }
        __syncthreads();
        lap(1);
        const int nb = min(s_nstart, MAXB);
        // 3. discovery order (rank by counting) and point offsets
        for (int i = tid; i < nb; i += QNT) {
            uint32_t k = s_ukey[i];
            int r = 0;
            for (int j = 0; j < nb; ++j) r += s_ukey[j] < k;
            s_key[r] = k;
            s_np[r] = (int)s_unp[i];
        }
        __syncthreads();
        if (tid == 0) {
            int acc = 0;
            for (int i = 0; i < nb; ++i) { s_off[i] = acc; acc += s_np[i]; }
            if (acc > MANTIS_MAX_PTS) s_flags |= FRAME_FLAG_POINT_OVERFLOW;
        }
        __syncthreads();
        lap(2);
        const bool overflow = s_flags != 0;
        int nraw = 0;
        if (!overflow) {
            // 4. extraction + left-edge ownership
            for (int b = tid; b < nb; b += QNT) {
                uint32_t k = s_key[b];
                bool is_hole = k & 1u;
                int p = (int)(k >> 1), y = p / (W + 1), x = p - y * (W + 1);
                Emit v{pts + s_off[b], owner, W, 0, (uint16_t)b};
                follow_border(im, is_hole ? x - 1 : x, y, is_hole, v);
            }
            __syncthreads();
            lap(3);
            // 5. approxPolyDP, 6. hole -> outer border
            for (int b = tid; b < nb; b += QNT) {
                int n4 = approx_poly_closed(pts + s_off[b], s_np[b], a.eps, dst + s_off[b], stack + s_off[b] + 2 * b);
                s_an[b] = (short)min(n4, 32767);
                if (n4 == 4)
                    for (int k = 0; k < 4; ++k) s_quad[b][k] = dst[s_off[b] + k];
                int cur = b;
                while (s_key[cur] & 1u) {  // climb: left end of the foreground run left of the hole start
                    int p = (int)(s_key[cur] >> 1), y = p / (W + 1), x = p - y * (W + 1);
                    int xl = x - 1;
                    while (im.at(xl - 1, y)) --xl;
                    cur = owner[(size_t)y * W + xl];
                }
                s_parent[b] = (short)cur;
            }
            __syncthreads();
            lap(4);
            // 7. CCOMP output order: outers by descending rank, each followed by its holes, descending
            for (int b = tid; b < nb; b += QNT) {
                bool hole_b = s_key[b] & 1u;
                int top_b = s_parent[b];
                int pos = 0;
                for (int o = 0; o < nb; ++o) {
                    if (o == b) continue;
                    bool hole_o = s_key[o] & 1u;
                    int top_o = s_parent[o];
                    bool before;
                    if (top_o != top_b) before = top_o > top_b;
                    else if (hole_o != hole_b) before = !hole_o;
                    else before = o > b;
                    pos += before;
                }
                s_pos2b[pos] = (short)b;
            }
            __syncthreads();
        }
        lap(5);
        // 8. quads: constructor, de-duplication, output (sequential, tiny)
        if (tid == 0) {
            out->n_contours = nb;
            out->flags = s_flags;
            if (!overflow) {
                float cx[MANTIS_MAX_QUADS], cy[MANTIS_MAX_QUADS];
                short qb[MANTIS_MAX_QUADS];
                unsigned char neighbor[MANTIS_MAX_QUADS];
                for (int p = 0; p < nb; ++p) {
                    int b = s_pos2b[p];
                    if (s_an[b] != 4) continue;
                    if (nraw >= MANTIS_MAX_QUADS) { out->flags |= FRAME_FLAG_QUAD_OVERFLOW; break; }
                    float xs = 0.f, ys = 0.f;
                    for (int k = 0; k < 4; ++k) { xs += (float)s_quad[b][k].x; ys += (float)s_quad[b][k].y; }
                    cx[nraw] = xs / 4.f;
                    cy[nraw] = ys / 4.f;
                    qb[nraw] = (short)b;
                    neighbor[nraw] = 0;
                    nraw++;
                }
                // removeDuplicateQuads: exact radius search, 4 results, unfilled slots alias quad 0
                for (int e = 0; e < nraw; ++e) {
                    if (neighbor[e]) continue;
                    int b = qb[e];
                    int ddx = s_quad[b][0].x - s_quad[b][1].x, ddy = s_quad[b][0].y - s_quad[b][1].y;
                    double side = sqrt((double)(ddx * ddx + ddy * ddy));
                    float radius = (float)(a.radius_mult * side);
                    int idx[4] = {0, 0, 0, 0};
                    float dd[4];
                    int found = 0;
                    for (int j = 0; j < nraw; ++j) {
                        float fx = cx[e] - cx[j], fy = cy[e] - cy[j];
                        float d = __fadd_rn(__fmul_rn(fx, fx), __fmul_rn(fy, fy));
                        if (!(d <= radius)) continue;
                        int pos = found < 4 ? found : 4;
                        while (pos > 0 && dd[pos - 1] > d) --pos;
                        if (pos >= 4) continue;
                        int last = found < 4 ? found : 3;
                        for (int k = last; k > pos; --k) { dd[k] = dd[k - 1]; idx[k] = idx[k - 1]; }
                        dd[pos] = d;
                        idx[pos] = j;
                        if (found < 4) ++found;
                    }
                    for (int k = 1; k < 4; ++k) neighbor[idx[k]] = 1;
                }
                int nq = 0;
                for (int e = 0; e < nraw; ++e) {
                    if (neighbor[e]) continue;
                    int b = qb[e];
                    for (int k = 0; k < 4; ++k) {
                        int px = s_quad[b][k].x, py = s_quad[b][k].y;
                        out->corners[nq][2 * k] = px;
                        out->corners[nq][2 * k + 1] = py;
                        float ex = (float)px - cx[e], ey = (float)py - cy[e];
                        float mx = (float)((double)ex * a.stretch), my = (float)((double)ey * a.stretch);
                        out->test_px[nq][2 * k] = (double)__fadd_rn(mx, cx[e]);
                        out->test_px[nq][2 * k + 1] = (double)__fadd_rn(my, cy[e]);
                    }
                    nq++;
                }
                out->n_quads_raw = nraw;
                out->n_quads = nq;
            } else {
                out->n_quads_raw = 0;
                out->n_quads = 0;
            }
        }
        lap(6);
    }
}

}  // namespace

int launch_quads(mantis_ctx* c, const uint32_t* d_morph_bits, int n, int w, int h, FrameQuads* d_out)
{
    const int slots = n < 2 * c->sm_count ? n : 2 * c->sm_count;
    int rc;
    if ((rc = ctx_reserve(c, c->q_pts, (size_t)slots * MANTIS_MAX_PTS * sizeof(short2))) ||
        (rc = ctx_reserve(c, c->q_dst, (size_t)slots * MANTIS_MAX_PTS * sizeof(short2))) ||
        (rc = ctx_reserve(c, c->q_stack, (size_t)slots * (MANTIS_MAX_PTS + 2 * MAXB) * sizeof(int2))) ||
        (rc = ctx_reserve(c, c->q_owner, (size_t)slots * w * h * sizeof(uint16_t))))
        return rc;
    QuadArgs a;
    a.morph_bits = d_morph_bits;
    a.W = w; a.H = h; a.WW = words_per_row(w); a.n_frames = n;
    a.pts = (short2*)c->q_pts.p; a.dst = (short2*)c->q_dst.p; a.stack = (int2*)c->q_stack.p;
    a.owner = (uint16_t*)c->q_owner.p;
    a.out = d_out;
    a.eps = (double)c->cfg.polygon_epsilon;
    a.stretch = c->cfg.quad_stretch_factor;
    a.radius_mult = c->cfg.search_radius_multiplier;
    a.phase_clocks = (long long*)c->dbg_clocks;
    k_quads<<<slots, QNT, 0, c->stream>>>(a);
    c->launches++;
    MCHECK(c, cudaGetLastError());
    return MANTIS_OK;
}
