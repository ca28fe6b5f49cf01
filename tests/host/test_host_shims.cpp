// tests/host/test_host_shims.cpp -- TEST DRIVER (links the oracle; never shipped).
// Exercises the C++ host shims exactly the way the reference's callers would:
//   * Ceres' calling convention: PrepareForEvaluation() once, then Evaluate() from 4 threads
//     with independently nullable Jacobian pointers (program_evaluator.h:155-258), Huber
//     correction applied by the caller (residual_block.cc:161-195)
//   * FeatureTracker::DetectNewFeatures / TrackOpticalFlow on cv::Mat-shaped inputs
// and compares with the oracle.  Prints HOST_SHIM_TEST PASS on success.
#include <array>
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <thread>
#include <vector>

#include "../../360_visual_inertial_odometry_b200/host/FeatureTrackerCuda.h"
#include "../../360_visual_inertial_odometry_b200/host/GpuBAEvaluator.h"
#include "../../oracle/ba_ref.h"
#include "../../oracle/fe_ref.h"

static unsigned long long g_seed = 88172645463325252ull;
static double urand() { g_seed ^= g_seed << 13; g_seed ^= g_seed >> 7; g_seed ^= g_seed << 17; return (g_seed >> 11) * (1.0 / 9007199254740992.0); }
static double nrand() { double u = urand() + 1e-12, v = urand(); return std::sqrt(-2 * std::log(u)) * std::cos(6.283185307179586 * v); }

static int fails = 0;
#define EXPECT(cond, ...) do { if (!(cond)) { std::printf("FAIL %s:%d: ", __FILE__, __LINE__); std::printf(__VA_ARGS__); std::printf("\n"); ++fails; } } while (0)

static double rel(const double* a, const double* b, size_t n) {
    double md = 0, mb = 0;
    for (size_t i = 0; i < n; ++i) { md = std::fmax(md, std::fabs(a[i] - b[i])); mb = std::fmax(mb, std::fabs(b[i])); }
    return md / std::fmax(mb, 1e-300);
}

static void test_ba() {
    const int K = 6, M = 150;
    const double cols = 960, rows = 480;
    const double info[4] = {1, 0, 0, 1};
    // T_cb: a proper rotation rounded through float (as Frame::GetTCB().cast<double>() is)
    double Tcb[16] = {0, -1, 0, 0.01, 0, 0, -1, 0.006, 1, 0, 0, 0.017, 0, 0, 0, 1};
    for (double& v : Tcb) v = (double)(float)(v + 0.0);
    std::vector<std::vector<double>> pose_params(K, std::vector<double>(6, 0.0)), point_params(M, std::vector<double>(3));
    std::vector<double> Twb(K * 16, 0.0);
    for (int k = 0; k < K; ++k) {
        const double a = 0.2 * k - 0.5, c = std::cos(a), s = std::sin(a);
        const double T[16] = {c, 0, s, std::cos(a), 0, 1, 0, 0.01 * k, -s, 0, c, std::sin(a), 0, 0, 0, 1};
        for (int i = 0; i < 16; ++i) Twb[k * 16 + i] = (double)(float)T[i];
        if (k > 0) for (int i = 0; i < 6; ++i) pose_params[k][i] = 1e-2 * nrand();
    }
    for (int j = 0; j < M; ++j) {
        double d[3] = {nrand(), 0.4 * nrand(), nrand()};
        const double n = std::sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]), r = 2 + 6 * urand();
        for (int i = 0; i < 3; ++i) point_params[j][i] = (double)(float)(d[i] / n * r);
    }
    vio_360::GpuBAEvaluator ev(cols, rows, Tcb, info, 1.0, 0);
    for (int k = 0; k < K; ++k) ev.AddPose(pose_params[k].data(), &Twb[k * 16]);
    for (int j = 0; j < M; ++j) ev.AddPoint(point_params[j].data());
    // observations: oracle projection of the (noisy) state + pixel noise
    std::vector<vio_360::GpuBackedBAFactor*> factors;
    std::vector<vo_factor> ofac;
    std::vector<int> fp, fl;
    for (int j = 0; j < M; ++j)
        for (int k = 0; k < K; ++k) {
            if (urand() < 0.2) continue;
            vo_factor f; std::memset(&f, 0, sizeof(f));
            f.cols = cols; f.rows = rows; std::memcpy(f.Tcb, Tcb, sizeof(Tcb)); std::memcpy(f.Twb_init, &Twb[k * 16], 128);
            std::memcpy(f.info, info, sizeof(info));
            // the factor clamps errors beyond 100 px to (100,100) (Factors.cpp:424-441): probe on a
            // 150-px grid until the probe pixel is within 100 px of the projection
            double r[2] = {100.0, 100.0};
            for (double pu = 75; pu < cols && r[0] == 100.0 && r[1] == 100.0; pu += 150)
                for (double pv = 75; pv < rows && r[0] == 100.0 && r[1] == 100.0; pv += 150) {
                    f.obs[0] = pu; f.obs[1] = pv;
                    vo_ba_evaluate(&f, pose_params[k].data(), point_params[j].data(), r, 0, nullptr, nullptr);
                }
            if (r[0] == 100.0 && r[1] == 100.0) continue;
            const double u = f.obs[0] - r[0] + nrand(), v = f.obs[1] - r[1] + nrand();
            if (u < 20 || u > cols - 20 || v < 20 || v > rows - 20) continue;
            f.obs[0] = (double)(float)u; f.obs[1] = (double)(float)v;
            factors.push_back(ev.AddFactor(k, j, f.obs[0], f.obs[1]));
            ofac.push_back(f); fp.push_back(k); fl.push_back(j);
        }
    ev.SetPoseConstant(0);
    for (int j = 0; j < M; j += 9) ev.SetPointConstant(j);
    ev.Finalize();
    const int O = (int)factors.size();
    EXPECT(O > 500, "too few factors: %d", O);

    ev.PrepareForEvaluation(true, true);
    std::vector<double> r(O * 2), Jp(O * 12), Jl(O * 6);
    auto worker = [&](int t) {
        for (int i = t; i < O; i += 4) {
            double* jac[2] = {(i % 3 == 0) ? nullptr : &Jp[12 * i], (i % 5 == 0) ? nullptr : &Jl[6 * i]};
            const double* params[2] = {pose_params[fp[i]].data(), point_params[fl[i]].data()};
            const bool ok = factors[i]->Evaluate(params, &r[2 * i], (i % 7 == 0) ? nullptr : jac);
            if (!ok) ++fails;
        }
    };
    std::vector<std::thread> th;
    for (int t = 0; t < 4; ++t) th.emplace_back(worker, t);
    for (auto& t : th) t.join();
    double worst_r = 0, worst_J = 0;
    for (int i = 0; i < O; ++i) {
        double ro[2], Jpo[12], Jlo[6];
        vo_ba_evaluate(&ofac[i], pose_params[fp[i]].data(), point_params[fl[i]].data(), ro, 1, Jpo, Jlo);
        worst_r = std::fmax(worst_r, std::fmax(std::fabs(ro[0] - r[2 * i]), std::fabs(ro[1] - r[2 * i + 1])));
        if (i % 7 != 0 && i % 3 != 0) worst_J = std::fmax(worst_J, rel(&Jp[12 * i], Jpo, 12) * (Jpo[0] != 0 || Jpo[3] != 0));
        if (i % 7 != 0 && i % 5 != 0) worst_J = std::fmax(worst_J, rel(&Jl[6 * i], Jlo, 6) * (Jlo[0] != 0 || Jlo[1] != 0));
    }
    EXPECT(worst_r < 1e-9, "residual mismatch %.3e", worst_r);
    EXPECT(worst_J < 1e-9, "jacobian mismatch %.3e", worst_J);

    // reduced camera system vs the oracle's Schur eliminator on Huber-corrected blocks
    std::vector<int> pidx(fp), lidx(fl);
    std::vector<double> ou(O), ov(O), pd(K * 6), pts(M * 3);
    std::vector<unsigned char> pfix(K, 0), lfix(M, 0);
    pfix[0] = 1; for (int j = 0; j < M; j += 9) lfix[j] = 1;
    for (int i = 0; i < O; ++i) { ou[i] = ofac[i].obs[0]; ov[i] = ofac[i].obs[1]; }
    for (int k = 0; k < K; ++k) std::memcpy(&pd[6 * k], pose_params[k].data(), 48);
    for (int j = 0; j < M; ++j) std::memcpy(&pts[3 * j], point_params[j].data(), 24);
    vo_window w; std::memset(&w, 0, sizeof(w));
    w.K = K; w.M = M; w.O = O; w.pose_idx = pidx.data(); w.point_idx = lidx.data(); w.obs_u = ou.data(); w.obs_v = ov.data();
    w.Twb_init = Twb.data(); std::memcpy(w.Tcb, Tcb, sizeof(Tcb)); w.cols = cols; w.rows = rows; std::memcpy(w.info, info, sizeof(info));
    w.huber_a = 1.0; w.pose_fixed = pfix.data(); w.point_fixed = lfix.data();
    std::vector<double> r2(O * 2), Jp2(O * 12), Jl2(O * 6), D(6 * K + 3 * M);
    double cost_o = 0;
    vo_window_evaluate(&w, pd.data(), pts.data(), 1, r2.data(), Jp2.data(), Jl2.data(), &cost_o, 2);
    vo_lm_diagonal(&w, Jp2.data(), Jl2.data(), 1e4, D.data());
    const int n = ev.ReducedDim();
    EXPECT(n == 6 * (K - 1), "reduced dim %d", n);
    std::vector<double> S(n * n), rhs(n), So(n * n), rhso(n);
    double cost = 0;
    ev.BuildReducedSystem(D.data(), S.data(), rhs.data(), &cost);
    vo_schur_eliminate(&w, r2.data(), Jp2.data(), Jl2.data(), D.data(), So.data(), rhso.data(), nullptr);
    EXPECT(rel(S.data(), So.data(), S.size()) < 1e-9, "S mismatch %.3e", rel(S.data(), So.data(), S.size()));
    EXPECT(rel(rhs.data(), rhso.data(), rhs.size()) < 1e-9, "rhs mismatch");
    EXPECT(std::fabs(cost - cost_o) < 1e-9 * cost_o, "cost mismatch %.17g %.17g", cost, cost_o);

    // chi-square pass + outlier flags, as Optimizer.cpp:898-915 does after the solve
    std::vector<double> chi2; std::vector<uint8_t> flags;
    ev.ComputeChiSquare(5.991, chi2, flags);
    int n_out = 0;
    for (int i = 0; i < O; ++i) {
        const double* params[2] = {pose_params[fp[i]].data(), point_params[fl[i]].data()};
        const double c = factors[i]->compute_chi_square(params);
        const double co = vo_ba_chi2(&ofac[i], params[0], params[1]);
        EXPECT(std::fabs(c - co) <= 1e-9 * std::fmax(1.0, co), "chi2 mismatch at %d", i);
        if (c > 5.991) { factors[i]->set_outlier(true); ++n_out; }
    }
    ev.PrepareForEvaluation(false, false);   // outlier flags changed: must re-evaluate even at the same point
    int checked = 0;
    for (int i = 0; i < O; ++i)
        if (factors[i]->is_outlier()) {
            double rr[2]; const double* params[2] = {pose_params[fp[i]].data(), point_params[fl[i]].data()};
            factors[i]->Evaluate(params, rr, nullptr);
            EXPECT(rr[0] == 640.0 && rr[1] == 480.0, "outlier residual");
            ++checked;
        }
    EXPECT(checked == n_out && n_out > 0, "outlier count %d %d", checked, n_out);
    // a factor evaluated at a point other than the uploaded one must not serve stale rows
    {
        double moved[6]; std::memcpy(moved, pose_params[fp[1]].data(), 48); moved[0] += 1e-3;
        const double* params[2] = {moved, point_params[fl[1]].data()};
        double rr[2]; bool threw = false;
        try { factors[1]->Evaluate(params, rr, nullptr); } catch (const std::runtime_error&) { threw = true; }
        EXPECT(threw, "Evaluate at a stale point must throw");
    }
    // ceres::Solve replacement (Optimizer.cpp:883-886): the device LM loop from the user's parameter blocks, back into them
    {
        const std::vector<double> p0_before = pose_params[0];
        const std::vector<double> fixed_pt_before = point_params[0];
        const vio360_lm_summary sm = ev.SolveLM();
        EXPECT(sm.termination >= 1 && sm.termination <= 5, "LM termination %d", sm.termination);
        EXPECT(sm.iterations >= 2 && sm.final_cost < sm.initial_cost, "LM did not reduce the cost: %.6g -> %.6g", sm.initial_cost, sm.final_cost);
        EXPECT(pose_params[0] == p0_before && point_params[0] == fixed_pt_before, "constant blocks moved");
        for (int k = 0; k < K; ++k) std::memcpy(&pd[6 * k], pose_params[k].data(), 48);
        for (int j = 0; j < M; ++j) std::memcpy(&pts[3 * j], point_params[j].data(), 24);
        std::vector<unsigned char> outl(O, 0);
        for (int i = 0; i < O; ++i) outl[i] = factors[i]->is_outlier() ? 1 : 0;
        w.outlier = outl.data();
        double cost_sol = 0;
        vo_window_evaluate(&w, pd.data(), pts.data(), 1, nullptr, nullptr, nullptr, &cost_sol, 2);
        EXPECT(std::fabs(cost_sol - sm.final_cost) < 1e-9 * cost_sol, "LM final cost %.17g vs oracle cost at the returned point %.17g", sm.final_cost, cost_sol);
        std::printf("  LM: %d iterations, cost %.6g -> %.6g, termination %d\n", sm.iterations, sm.initial_cost, sm.final_cost, sm.termination);
    }
    for (auto* f : factors) delete f;   // (a ceres::Problem would own them)
    std::printf("  BA seam: %d factors, |dr| %.2e, |dJ| %.2e, %d chi2 outliers\n", O, worst_r, worst_J, n_out);
}

static void test_fe() {
    const int W = 640, H = 320;
    std::vector<unsigned char> f0((size_t)W * H), f1((size_t)W * H);
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            auto tex = [](double xx, double yy) {
                return 127.5 + 50 * std::sin(0.11 * xx + 0.05 * yy) + 40 * std::sin(0.23 * yy - 0.07 * xx) + 30 * std::sin(0.37 * xx) * std::cos(0.29 * yy);
            };
            f0[(size_t)y * W + x] = (unsigned char)std::fmin(255.0, std::fmax(0.0, tex(x, y)));
            f1[(size_t)y * W + x] = (unsigned char)std::fmin(255.0, std::fmax(0.0, tex(x - 1.6, y + 0.7)));
        }
    struct Cam { int GetWidth() const { return 640; } int GetHeight() const { return 320; } } cam;
    vio_360::FeatureTrackerCuda tracker(&cam, 300, 30.0f, 0.01f, 20, 0);
    cv::Mat m0(H, W, CV_8UC1, f0.data(), (size_t)W), m1(H, W, CV_8UC1, f1.data(), (size_t)W);
    std::vector<cv::Point2f> corners = tracker.DetectNewFeatures(m0);
    std::vector<unsigned char> mask((size_t)W * H);
    vf_base_mask(W, H, 0.15f, 20, mask.data());
    std::vector<float> ref(600);
    const int nref = vf_good_features(f0.data(), W, H, W, mask.data(), W, 300, (double)0.01f, (double)30.0f, ref.data());
    EXPECT((int)corners.size() == nref && nref > 20, "corner count %zu vs %d", corners.size(), nref);
    for (int i = 0; i < nref && i < (int)corners.size(); ++i)
        EXPECT(corners[i].x == ref[2 * i] && corners[i].y == ref[2 * i + 1], "corner %d differs", i);
    std::vector<cv::Point2f> cur; std::vector<uchar> status;
    const int ntr = tracker.TrackOpticalFlow(m0, m1, corners, cur, status);
    std::vector<float> oxy(2 * corners.size()); std::vector<unsigned char> ost(corners.size()); std::vector<float> oerr(corners.size());
    const int nor = vf_lk_track(f0.data(), W, f1.data(), W, W, H, 3, 21, 30, (double)0.01f, (double)0.01f,
                                reinterpret_cast<const float*>(corners.data()), (int)corners.size(), oxy.data(), ost.data(), oerr.data());
    EXPECT(ntr == nor, "tracked %d vs %d", ntr, nor);
    double worst = 0;
    for (size_t i = 0; i < corners.size(); ++i) {
        EXPECT(status[i] == ost[i], "status %zu", i);
        if (ost[i]) worst = std::fmax(worst, std::fmax(std::fabs(cur[i].x - oxy[2 * i]), std::fabs(cur[i].y - oxy[2 * i + 1])));
    }
    EXPECT(worst < 0.05, "position mismatch %.4f", worst);
    std::vector<cv::Point2f> none, out; std::vector<uchar> st2;
    EXPECT(tracker.TrackOpticalFlow(m0, m1, none, out, st2) == 0 && out.empty() && st2.empty(), "empty input must be a no-op");
    // rotation RANSAC through the shim (same mt19937 draw as the reference, fixed seed): the tracked
    // pairs of a small pure shift are nearly all consistent with one rotation
    std::vector<cv::Point2f> gp, gc;
    for (size_t i = 0; i < corners.size(); ++i)
        if (status[i]) { gp.push_back(corners[i]); gc.push_back(cur[i]); }
    std::vector<uchar> inl;
    tracker.SetRansacSeed(1234u);
    tracker.SetRansacParams(200, 2.0f);
    const int nin = tracker.RejectOutliersRotationRANSAC(gp, gc, inl);
    int msum = 0;
    for (uchar m : inl) msum += m;
    EXPECT(inl.size() == gp.size() && nin == msum && nin > (int)gp.size() * 8 / 10, "ransac inliers %d of %zu (mask %d)", nin, gp.size(), msum);
    std::vector<cv::Point2f> two(gp.begin(), gp.begin() + 2), two_c(gc.begin(), gc.begin() + 2);
    std::vector<uchar> inl2;
    EXPECT(tracker.RejectOutliersRotationRANSAC(two, two_c, inl2) == 2 && inl2.size() == 2 && inl2[0] == 1 && inl2[1] == 1, "n < 3 keeps everything");
    // feature-list filters through the shim against the oracle (tracked points + a tight artificial cluster)
    std::vector<cv::Point2f> fl(gc);
    for (int i = 0; i < 6; ++i) fl.push_back(cv::Point2f(300.f + 0.5f * i, 200.f + 0.25f * i));
    std::vector<int> tc(fl.size());
    for (size_t i = 0; i < fl.size(); ++i) tc[i] = 1 + (int)((i * 7u) % 5u);
    const std::vector<uchar> kc = tracker.RemoveClusteredFeatures(fl), kg = tracker.LimitFeaturesPerGrid(fl, tc, 2);
    std::vector<unsigned char> okc(fl.size()), okg(fl.size());
    vf_cluster_filter(reinterpret_cast<const float*>(fl.data()), (int)fl.size(), W, H, 20, 10, 0.25f, okc.data());
    vf_grid_cap(reinterpret_cast<const float*>(fl.data()), tc.data(), (int)fl.size(), W, H, 20, 10, 2, okg.data());
    int nclu = 0, ncap = 0;
    for (size_t i = 0; i < fl.size(); ++i) {
        EXPECT(kc[i] == okc[i], "cluster filter %zu", i);
        EXPECT(kg[i] == okg[i], "grid cap %zu", i);
        nclu += !kc[i]; ncap += !kg[i];
    }
    EXPECT(nclu >= 6 && ncap > 0, "filters removed %d / %d features", nclu, ncap);
    // the whole TrackFeatures chain as one device-resident call per frame: first frame = detection + grid cap,
    // second frame = tracked features keep their ids and gain a track count, new ones fill the gaps
    vio_360::FeatureTrackerCuda seqtr(&cam, 300, 30.0f, 0.01f, 20, 0);
    seqtr.SetRansacSeed(99u);
    seqtr.SetRansacParams(100, 2.0f);
    vio_360::FeatureTrackerCuda::FrameFeatures fa, fb;
    seqtr.TrackFeatures(m0, fa);
    EXPECT(fa.num_tracked == 0 && fa.num_detected == nref && !fa.ids.empty() && fa.ids.size() <= (size_t)nref, "first frame: %d detected, %zu kept", fa.num_detected, fa.ids.size());
    for (size_t i = 0; i < fa.ids.size(); ++i) EXPECT(fa.track_counts[i] == 0, "fresh feature with a track count");
    seqtr.TrackFeatures(m1, fb);
    int carried = 0;
    for (size_t i = 0; i < fb.ids.size(); ++i)
        if (fb.track_counts[i] == 1) { ++carried; EXPECT(std::find(fa.ids.begin(), fa.ids.end(), fb.ids[i]) != fa.ids.end(), "carried id %d unknown", fb.ids[i]); }
        else EXPECT(fb.track_counts[i] == 0 && fb.ids[i] >= nref, "new feature id %d", fb.ids[i]);
    EXPECT(carried > (int)fa.ids.size() / 2 && fb.num_tracked >= carried, "second frame: %d carried of %zu (tracked %d)", carried, fa.ids.size(), fb.num_tracked);
    // frame buffers page-locked in place (RegisterFrameBuffer / auto registration): same results, and the
    // bookkeeping keeps at most max_buffers ranges registered
    {
        vio_360::FeatureTrackerCuda reg(&cam, 300, 30.0f, 0.01f, 20, 0);
        reg.SetRansacSeed(99u);
        reg.SetRansacParams(100, 2.0f);
        reg.SetAutoRegisterFrames(true, 1);
        vio_360::FeatureTrackerCuda::FrameFeatures ra, rb;
        reg.TrackFeatures(m0, ra);
        reg.TrackFeatures(m1, rb);       // evicts m0's registration (max_buffers = 1)
        EXPECT(ra.ids == fa.ids && rb.ids == fb.ids && rb.track_counts == fb.track_counts, "registered frame buffers change the chain's result");
        bool same = rb.points.size() == fb.points.size();
        for (size_t i = 0; same && i < rb.points.size(); ++i) same = rb.points[i].x == fb.points[i].x && rb.points[i].y == fb.points[i].y;
        EXPECT(same, "registered frame buffers change the tracked positions");
        reg.RegisterFrameBuffer(m0); reg.RegisterFrameBuffer(m0);     // idempotent
        reg.UnregisterFrameBuffer(m0); reg.UnregisterFrameBuffer(m1);
    }
    std::printf("  front-end: %d corners, %d tracked, |dxy| %.2e px, ransac %d/%zu inliers, cluster filter -%d, grid cap -%d\n", nref, ntr, worst, nin, gp.size(), nclu, ncap);
}

int main() {
    try {
        test_ba();
        test_fe();
    } catch (const std::exception& e) {
        std::printf("HOST_SHIM_TEST ERROR: %s\n", e.what());
        return 2;
    }
    std::printf(fails ? "HOST_SHIM_TEST FAIL (%d)\n" : "HOST_SHIM_TEST PASS\n", fails);
    return fails ? 1 : 0;
}
