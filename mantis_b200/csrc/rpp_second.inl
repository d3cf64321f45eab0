// rpp_second.inl -- second-pose search and the RPP driver (included by rpp.cuh).
namespace rpp {

// RPP::GetRotationY_wrtT (RPP.cpp:947-1222): minima of the object-space error over the rotation
// beta about y, and the optimal translation for each.
template <int N>
__device__ inline int rotation_y_wrt_t(const V3* v, const V3* p, const M3& Rz, double* bl, V3* tnew, int n = N)
{
    const double PI = 3.14159265358979323846;
    M3 Vp[N], G, I = pm::eye();
    for (int k = 0; k < 9; ++k) G.m[k] = 0;
    for (int i = 0; i < n; ++i) {
        double vv = v[i].v[0] * v[i].v[0] + v[i].v[1] * v[i].v[1] + v[i].v[2] * v[i].v[2];
        for (int a = 0; a < 3; ++a)
            for (int b = 0; b < 3; ++b) {
                Vp[i].m[a * 3 + b] = (v[i].v[a] * v[i].v[b]) / vv;
                G.m[a * 3 + b] += Vp[i].m[a * 3 + b];
            }
    }
    for (int k = 0; k < 9; ++k) G.m[k] = I.m[k] - G.m[k] / n;
    G = pm::inverse(G);
    for (int k = 0; k < 9; ++k) G.m[k] /= n;

    M3 opt_t;
    for (int k = 0; k < 9; ++k) opt_t.m[k] = 0;
    for (int i = 0; i < n; ++i) {
        double px = p[i].v[0], py = p[i].v[1], pz = p[i].v[2];
        M3 Rpi = {{-px, 2 * pz, px, py, 0, py, -pz, -2 * px, pz}};
        M3 A;
        for (int k = 0; k < 9; ++k) A.m[k] = Vp[i].m[k] - I.m[k];
        M3 B = pm::mul(pm::mul(A, Rz), Rpi);
        for (int k = 0; k < 9; ++k) opt_t.m[k] += B.m[k];
    }
    opt_t = pm::mul(G, opt_t);

    double E2[5] = {0, 0, 0, 0, 0};
    for (int i = 0; i < n; ++i) {
        double px = p[i].v[0], py = p[i].v[1], pz = p[i].v[2];
        M3 Rpi = {{-px, 2 * pz, px, py, 0, py, -pz, -2 * px, pz}};
        M3 A, S = pm::mul(Rz, Rpi);
        for (int k = 0; k < 9; ++k) { A.m[k] = I.m[k] - Vp[i].m[k]; S.m[k] += opt_t.m[k]; }
        M3 E = pm::mul(A, S);
        for (int k = 0; k < 3; ++k) {
            double e2 = E.m[k * 3 + 0], e1 = E.m[k * 3 + 1], e0 = E.m[k * 3 + 2];
            E2[0] += e2 * e2;
            E2[1] += 2 * e1 * e2;
            E2[2] += 2 * e0 * e2 + e1 * e1;
            E2[3] += 2 * e0 * e1;
            E2[4] += e0 * e0;
        }
    }
    double e4 = E2[0], e3 = E2[1], e2 = E2[2], e1 = E2[3], e0 = E2[4];
    double a4 = -e3, a3 = 4 * e4 - 2 * e2, a2 = -3 * e1 + 3 * e3, a1 = -4 * e0 + 2 * e2, a0 = e1;
    double coeffs[5] = {a4, a3, a2, a1, a0}, roots[4];
    int nr = poly_real_roots(coeffs, 4, roots);
    int nsol = 0;
    for (int i = 0; i < nr; ++i) {
        double at = roots[i];
        double q1 = 1.0 + at * at;
        if (!(fabs(q1 * q1 * q1) > 0.1)) continue;
        double sa = (2.0 * at) / (1.0 + at * at), ca = (1.0 - at * at) / (1.0 + at * at);
        double al_deg = atan2(sa, ca) * 180 / PI;
        double tmm = 4 * a4 * at * at * at + 3 * a3 * at * at + 2 * a2 * at + a1;
        if (!(tmm > 0)) continue;
        double beta = al_deg * PI / 180;  // the reference round-trips through degrees
        M3 R = pm::mul(Rz, rpy_mat(0, beta, 0));
        V3 topt = {{0, 0, 0}};
        for (int k = 0; k < n; ++k) {
            V3 rp = pm::mulv(R, p[k]);
            V3 vrp = pm::mulv(Vp[k], rp);
            for (int q = 0; q < 3; ++q) topt.v[q] += vrp.v[q] - rp.v[q];
        }
        bl[nsol] = (al_deg / 180) * PI;
        tnew[nsol] = pm::mulv(G, topt);
        nsol++;
    }
    return nsol;
}

// RPP::Get2ndPose_Exact + GetRfor2ndPose_V_Exact + DecomposeR: candidate second poses.
// Returns the number of candidates, or -1 where the reference returns false.
template <int N>
__device__ inline int second_poses(const V3* model, const V3* ray, const M3& R, M3* solR, V3* solT, int n = N)
{
    const double PI = 3.14159265358979323846;
    V3 cent = {{0, 0, 0}};
    for (int i = 0; i < n; ++i) {
        double x = ray[i].v[0], y = ray[i].v[1], z = ray[i].v[2];
        double m = 1.0 / sqrt(x * x + y * y + z * z);
        cent.v[0] += x * m; cent.v[1] += y * m; cent.v[2] += z * m;
    }
    for (int k = 0; k < 3; ++k) cent.v[k] /= 3;  // RPP::Mean divides by cols (= 3); harmless before NormRv
    {
        double m = 1.0 / sqrt(cent.v[0] * cent.v[0] + cent.v[1] * cent.v[1] + cent.v[2] * cent.v[2]);
        for (int k = 0; k < 3; ++k) cent.v[k] *= m;
    }
    // Rim = GetRotationbyVector((0,0,1), cent): rotation by acos(cent.z) about cent x z
    double winkel = acos(cent.v[2]);
    double ax = cent.v[1] * 1.0 - cent.v[2] * 0.0, ay = cent.v[2] * 0.0 - cent.v[0] * 1.0, az = 0.0;
    double an = sqrt(ax * ax + ay * ay + az * az);
    double sh = sin(winkel * 0.5), qa = cos(winkel * 0.5);
    double qb = ax / an * sh, qc = ay / an * sh, qd = az / an * sh;
    double qn = 1.0 / sqrt(qa * qa + qb * qb + qc * qc + qd * qd);
    qa *= qn; qb *= qn; qc *= qn; qd *= qn;
    M3 Rim;
    Rim.m[0] = qa * qa + qb * qb - qc * qc - qd * qd; Rim.m[1] = 2 * (qb * qc - qa * qd); Rim.m[2] = 2 * (qb * qd + qa * qc);
    Rim.m[3] = 2 * (qb * qc + qa * qd); Rim.m[4] = qa * qa + qc * qc - qb * qb - qd * qd; Rim.m[5] = 2 * (qc * qd - qa * qb);
    Rim.m[6] = 2 * (qb * qd - qa * qc); Rim.m[7] = 2 * (qc * qd + qa * qb); Rim.m[8] = qa * qa + qd * qd - qb * qb - qc * qc;

    V3 v_[N];
    for (int i = 0; i < n; ++i) v_[i] = pm::mulv(Rim, ray[i]);
    M3 R_ = pm::mul(Rim, R);

    // DecomposeR
    double cl = atan2(R_.m[7], R_.m[6]);
    M3 RzN = rpy_mat(0, 0, cl);
    M3 Rt = pm::mul(R_, RzN);
    double ang[3];
    if (Rt.m[7] > 1e-3) return -1;
    if (!rpy_ang_x(Rt, ang)) return -1;
    if (fabs(ang[0]) > 1e-3) return -1;
    M3 Rz2 = pm::mul(RzN, rpy_mat(0, 0, PI));
    M3 Rt2 = pm::mul(R_, Rz2);
    if (Rt2.m[7] > 1e-3) return -1;
    if (!rpy_ang_x(Rt2, ang)) return -1;

    M3 RzNt = pm::transpose(RzN);
    V3 P_[N];
    for (int i = 0; i < n; ++i) P_[i] = pm::mulv(RzNt, model[i]);
    if (!rpy_ang_x(Rt, ang)) return -1;
    M3 Rz = rpy_mat(0, 0, ang[2]);

    double bl[4];
    V3 tnew[4];
    int ns = rotation_y_wrt_t<N>(v_, P_, Rz, bl, tnew, n);
    if (ns == 0) return -1;
    M3 Rimt = pm::transpose(Rim);
    for (int j = 0; j < ns; ++j) {
        M3 Rs = pm::mul(pm::mul(Rz, rpy_mat(0, bl[j], 0)), RzNt);
        solR[j] = pm::mul(Rimt, Rs);
        solT[j] = pm::mulv(Rimt, tnew[j]);
    }
    return ns;
}

// RPP::Rpp.  Returns false where the reference returns false (or indexes sol[-1]).
// its (optional): AbsKernel calls of the first run, of the winning refinement run, and the number of candidates.
template <int N>
__device__ inline bool solve(const V3* model, const V3* ray, M3& R, V3& t, double& obj_err, double& img_err, int* its = nullptr, int n = N)
{
    ObjCtx<N> c;
    setup<N>(c, model, ray, n);
    M3 Rlu;
    V3 tlu;
    double oe, ie;
    int calls = 0;
    obj_pose<N>(c, false, pm::eye(), Rlu, tlu, oe, ie, &calls, n);
    if (its) { its[0] = calls; its[1] = 0; its[2] = 0; }
    M3 solR[4];
    V3 solT[4];
    int ns = second_poses<N>(model, ray, Rlu, solR, solT, n);
    if (ns <= 0) return false;
    if (its) its[2] = ns;
    int best = -1;
    double lowest = 1e6;
    for (int i = 0; i < ns; ++i) {
        M3 Ri;
        V3 ti;
        obj_pose<N>(c, true, solR[i], Ri, ti, oe, ie, &calls, n);
        if (oe < lowest) {
            lowest = oe; best = i; R = Ri; t = ti; obj_err = oe; img_err = ie;
            if (its) its[1] = calls;
        }
    }
    return best >= 0;
}

}  // namespace rpp
