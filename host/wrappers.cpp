// wrappers.cpp — the three run loops of the reference (wrappers/slamwrapper.cpp, ParticleSLAMWrapper.cpp,
// fastslam1wrapper.cpp, fastslam2wrapper.cpp, ekfslamwrapper.cpp) on top of the C ABI.  The loop body, the order of
// the rand() draws and the quirks are the reference's; the particle set / EKF state is a device handle.
#include <sys/time.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>

#include "../include/slamgpu.h"
#include "slam_host.hpp"

namespace slamhost {

namespace {

struct Sim {  // SLAMWrapper's state (slamwrapper.h / slamwrapper.cpp:8-55)
    const Conf &conf;
    const Map &map;
    float Vtrue, Gtrue = 0, Vnoisy = 0, Gnoisy = 0;
    float Q[4], R[4], Qe[4], Re[4];
    int nLoop;
    float dt, dtSum = 0;
    float xTrue[3] = {0, 0, 0};
    int iwp = 0;  // indexOfFirstWaypoint
    int currentIteration = 0;
    long mark = 0;

    Sim(const Conf &c, const Map &m) : conf(c), map(m) {
        Vtrue = c.V;
        const float q0 = (float) std::pow(c.sigmaV, 2), q3 = (float) std::pow(c.sigmaG, 2);
        const float r0 = (float) std::pow(c.sigmaR, 2), r3 = (float) std::pow(c.sigmaB, 2);
        const float k = (c.SWITCH_INFLATE_NOISE == 1) ? 2.0f : 1.0f;
        // SWITCH_INFLATE_NOISE = 1 doubles Q and R and leaves Qe / Re UNSET in the reference (slamwrapper.cpp:31-37:
        // an empty matrix reaches the filter).  Here the filter gets the inflated matrices.
        const float q[4] = {k * q0, 0, 0, k * q3}, r[4] = {k * r0, 0, 0, k * r3};
        for (int i = 0; i < 4; i++) {
            Q[i] = Qe[i] = q[i];
            R[i] = Re[i] = r[i];
        }
        nLoop = c.NUMBER_LOOPS;
        dt = c.DT_CONTROLS;
        if (c.SWITCH_SEED_RANDOM != 0) srand(c.SWITCH_SEED_RANDOM);
        else srand((unsigned) time(NULL));
    }

    // SLAMWrapper::control, slamwrapper.cpp:173-238 (waypoint mode)
    int control() {
        if (iwp == -1) return -1;
        update_steering(xTrue, map, &iwp, conf.AT_WAYPOINT, &Gtrue, conf.RATEG, conf.MAXG, dt);
        if (iwp == -1 && nLoop > 1) {
            iwp = 0;
            nLoop--;
        }
        if (iwp == -1 && nLoop == 1) return -1;
        predict_true_position(xTrue, Vtrue, Gtrue, conf.WHEELBASE, dt);
        if (conf.SWITCH_CONTROL_NOISE) add_control_noise(Vtrue, Gtrue, Q, &Vnoisy, &Gnoisy);
        return 1;
    }

    uint32_t lap() {  // updateMicrotimeMark
        struct timeval tv;
        gettimeofday(&tv, NULL);
        const long now = (long) tv.tv_sec * 1000000 + tv.tv_usec;
        const uint32_t d = (uint32_t) (now - mark);
        mark = now;
        return d;
    }
};

}  // namespace

RunResult run_fastslam(int method, const Conf &conf, const Map &map, PlotSink *plot, const RunOptions &opt) {
    RunResult res;
    if (map.wp_x.empty()) {
        res.error = "the map holds no waypoints";
        return res;
    }
    Sim sim(conf, map);
    const int64_t P = conf.NPARTICLES;
    const int n_lm = (int) map.lm_x.size();
    slamgpu_config cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.n_particles = P;
    cfg.nf_cap = opt.nf_cap > 0 ? opt.nf_cap : (n_lm > 0 ? n_lm : 1);
    cfg.method = method == 1 ? SLAMGPU_FASTSLAM1 : SLAMGPU_FASTSLAM2;
    cfg.wheel_base = conf.WHEELBASE;
    cfg.sigma_phi = conf.sigmaT;
    cfg.n_effective = conf.NEFFECTIVE;
    cfg.use_heading = conf.SWITCH_HEADING_KNOWN == 1;
    cfg.do_resample = conf.SWITCH_RESAMPLE == 1;
    cfg.scan_mode = opt.scan_mode;
    cfg.math_mode = opt.math_mode;
    cfg.device = opt.device;
    slamgpu_t *h = nullptr;
    if (slamgpu_create(&cfg, &h) != SLAMGPU_OK) {
        res.error = slamgpu_last_error();
        return res;
    }
    // FastSLAM1 always samples the controls per particle (fastslam1wrapper.cpp:20); FastSLAM2 only on request
    const bool predict_noise = (method == 1) || (conf.SWITCH_PREDICT_NOISE == 1);
    std::vector<float> g2(2 * (size_t) P), g3(3 * (size_t) P), z, zf, zn, xv_dec, lm_dec;
    std::vector<double> unif((size_t) P);
    std::vector<int> ids, idf, table(n_lm, -1);
    if (plot) plot->begin(conf, map, method == 1 ? "FASTSLAM1" : "FASTSLAM2");
    int rc = SLAMGPU_OK, n_obs = 0;
    while (rc == SLAMGPU_OK) {
        sim.lap();
        const int cs = sim.control();
        if (cs == -1) break;
        if (cs == 0) continue;
        const float *gp = nullptr;
        if (predict_noise) {  // predictState draws randn(2,1) per particle, in particle order (fastslam1.cpp:37-48)
            for (int64_t p = 0; p < P; p++) {
                float g[2];
                randn(2, g);
                g2[p] = g[0];
                g2[P + p] = g[1];
            }
            gp = g2.data();
        }
        rc = slamgpu_predict(h, sim.Vnoisy, sim.Gnoisy, sim.Qe, sim.dt, sim.xTrue[2], gp);
        if (rc != SLAMGPU_OK) break;
        sim.dtSum += sim.dt;
        if (sim.dtSum >= conf.DT_OBSERVE) {
            sim.dtSum = 0;
            get_observations(map, sim.xTrue, conf.MAX_RANGE, &ids, &z);
            if (conf.SWITCH_SENSOR_NOISE) add_observation_noise(&z, sim.R);
            if (plot) plot->laser(z.data(), (int) ids.size(), sim.xTrue);  // plines = makeLaserLines(...)
            const int Nf = slamgpu_landmark_count(h);
            data_association_known(z, ids, &table, Nf, &zf, &idf, &zn);
            const int K = (int) idf.size(), Kn = (int) zn.size() / 2;
            n_obs = K + Kn;  // the laser lines of the last observation stay on the plot until the next one
            const float *g3p = nullptr;
            if (method == 2 && (K > 0 || Kn > 0)) {  // sampleProposal / the zn-only branch: randn(3,1) per particle
                for (int64_t p = 0; p < P; p++) {
                    float g[3];
                    randn(3, g);
                    g3[p] = g[0];
                    g3[P + p] = g[1];
                    g3[2 * P + p] = g[2];
                }
                g3p = g3.data();
            }
            for (int64_t p = 0; p < P; p++) unif[p] = unif_rand();  // stratifiedRandom, on every update
            int32_t did = 0;
            float neff = 0;
            rc = slamgpu_update(h, K ? zf.data() : nullptr, K ? idf.data() : nullptr, K, Kn ? zn.data() : nullptr, Kn,
                                sim.Re, g3p, unif.data(), &did, &neff);
            if (rc != SLAMGPU_OK) break;
            static const bool diag = getenv("SLAM_HOST_DIAG") != nullptr;  // nEff / decision per update (debug aid)
            if (diag) {
                fprintf(stderr, "[diag] it %u K %d Kn %d nEff %g did %d\n", sim.currentIteration, K, Kn, neff, did);
                static bool reported = false;
                slamgpu_debug_capture(h, 1);
                if (neff != neff && !reported) {  // the weights this resample saw: all zero, or a NaN somewhere?
                    reported = true;
                    std::vector<float> wb((size_t) P);
                    if (slamgpu_debug_read(h, SLAMGPU_DEBUG_W_BEFORE, wb.data()) == SLAMGPU_OK) {
                        int zeros = 0, nans = 0, infs = 0;
                        float mx = 0;
                        for (float v : wb) {
                            zeros += v == 0.0f;
                            nans += v != v;
                            infs += (v - v) != 0.0f && v == v;
                            if (v == v && v > mx) mx = v;
                        }
                        fprintf(stderr, "[diag] weights entering the resample: %d zero, %d NaN, %d inf, max %g of %lld\n",
                                zeros, nans, infs, mx, (long long) P);
                    }
                }
            }
            res.observations++;
            res.resamples += did;
        }
        sim.currentIteration++;
        const uint32_t us = sim.lap();
        double est[3];
        rc = slamgpu_estimate(h, &est[0], &est[1], &est[2]);
        if (rc != SLAMGPU_OK) break;
        if (plot) {
            if (opt.plot_stride > 0) {  // drawParticles / drawFeatureParticles, decimated on the device
                const int64_t n = (P + opt.plot_stride - 1) / opt.plot_stride;
                const int Nf = slamgpu_landmark_count(h);
                xv_dec.resize(3 * (size_t) n);
                lm_dec.resize(2 * (size_t) n * (size_t) (Nf > 0 ? Nf : 1));
                rc = slamgpu_download_decimated(h, opt.plot_stride, xv_dec.data(), Nf > 0 ? lm_dec.data() : nullptr);
                if (rc != SLAMGPU_OK) break;
                plot->particles(xv_dec.data(), n, lm_dec.data(), Nf);
            }
            plot->tick(sim.currentIteration, sim.xTrue, est, n_obs, us);
        }
        for (int i = 0; i < 3; i++) res.x_est[i] = est[i];
        if (opt.max_iterations > 0 && sim.currentIteration >= opt.max_iterations) break;
    }
    if (rc != SLAMGPU_OK) res.error = slamgpu_last_error();
    res.iterations = sim.currentIteration;
    res.landmarks = slamgpu_landmark_count(h);
    for (int i = 0; i < 3; i++) res.x_true[i] = sim.xTrue[i];
    if (plot) plot->end();
    slamgpu_destroy(h);
    return res;
}

RunResult run_ekfslam(const Conf &conf, const Map &map, PlotSink *plot, const RunOptions &opt) {
    RunResult res;
    if (map.wp_x.empty()) {
        res.error = "the map holds no waypoints";
        return res;
    }
    Sim sim(conf, map);
    const int n_lm = (int) map.lm_x.size();
    slamgpu_ekf_config cfg;
    memset(&cfg, 0, sizeof(cfg));
    // with SWITCH_ASSOCIATION_KNOWN the reference's table is passed by value and never learns (ekfslamwrapper.cpp:80-83):
    // every sighting appends a landmark, so the map outgrows the landmark count
    cfg.nf_cap = opt.nf_cap > 0 ? opt.nf_cap : (conf.SWITCH_ASSOCIATION_KNOWN ? 64 * (n_lm + 1) : 2 * n_lm + 8);
    cfg.m_cap = n_lm > 0 ? n_lm : 1;
    cfg.wheel_base = conf.WHEELBASE;
    cfg.sigma_phi = conf.sigmaT;
    cfg.gate_reject = conf.GATE_REJECT;
    cfg.gate_augment = conf.GATE_AUGMENT;
    cfg.use_heading = conf.SWITCH_HEADING_KNOWN == 1;
    cfg.batch_update = conf.SWITCH_BATCH_UPDATE == 1;
    cfg.association_known = conf.SWITCH_ASSOCIATION_KNOWN;
    cfg.device = opt.device;
    slamgpu_ekf_t *h = nullptr;
    if (slamgpu_ekf_create(&cfg, &h) != SLAMGPU_OK) {
        res.error = slamgpu_last_error();
        return res;
    }
    std::vector<float> z, xs, Ps;
    std::vector<int> ids, table(n_lm, -1);
    if (plot) plot->begin(conf, map, "EKF1");
    int rc = SLAMGPU_OK, n_obs = 0;
    while (rc == SLAMGPU_OK) {
        sim.lap();
        const int cs = sim.control();
        if (cs == -1) break;
        if (cs == 0) continue;
        sim.dtSum += sim.dt;
        bool observe = false;
        if (sim.dtSum >= conf.DT_OBSERVE) {
            observe = true;
            sim.dtSum = 0;
            get_observations(map, sim.xTrue, conf.MAX_RANGE, &ids, &z);
            if (conf.SWITCH_SENSOR_NOISE) add_observation_noise(&z, sim.R);
            if (plot) plot->laser(z.data(), (int) ids.size(), sim.xTrue);
        }
        // landmarksRangeBearing keeps its last value between observations in the reference; sim only looks at it when
        // observe is set
        const float phi = (float) (sim.xTrue[2] + conf.sigmaT * unif_rand());
        const int m = observe ? (int) ids.size() : 0;
        int32_t nzf = 0, nzn = 0;
        rc = slamgpu_ekf_sim(h, sim.Vnoisy, sim.Gnoisy, sim.Qe, sim.dt, phi, observe ? 1 : 0, m ? z.data() : nullptr,
                             m ? ids.data() : nullptr, m, table.data(), n_lm, sim.Re, sim.R, nullptr, &nzf, &nzn);
        if (rc != SLAMGPU_OK) break;
        if (observe) {
            res.observations++;
            n_obs = m;
        }
        sim.currentIteration++;
        const uint32_t us = sim.lap();
        float xv[3];
        rc = slamgpu_ekf_get_pose(h, xv, nullptr);
        if (rc != SLAMGPU_OK) break;
        const double est[3] = {xv[0], xv[1], xv[2]};
        if (plot && plot->wants_ellipses()) {  // drawCovarianceEllipseLines, ekfslamwrapper.cpp:117-136
            const int n = slamgpu_ekf_size(h);
            xs.resize(n);
            Ps.resize((size_t) n * n);
            rc = slamgpu_ekf_get_state(h, xs.data(), Ps.data());
            if (rc != SLAMGPU_OK) break;
            for (int b = 0; b < 1 + (n - 3) / 2; b++) {
                const int o = b == 0 ? 0 : 3 + 2 * (b - 1);
                const double c[2] = {xs[o], xs[o + 1]};
                const float blk[4] = {Ps[(size_t) o * n + o], Ps[(size_t) o * n + o + 1], Ps[(size_t) (o + 1) * n + o],
                                      Ps[(size_t) (o + 1) * n + o + 1]};
                plot->ellipse(c, blk, b);
            }
        }
        if (plot) plot->tick(sim.currentIteration, sim.xTrue, est, n_obs, us);
        for (int i = 0; i < 3; i++) res.x_est[i] = est[i];
        if (opt.max_iterations > 0 && sim.currentIteration >= opt.max_iterations) break;
    }
    if (rc != SLAMGPU_OK) res.error = slamgpu_last_error();
    res.iterations = sim.currentIteration;
    res.landmarks = (slamgpu_ekf_size(h) - 3) / 2;
    for (int i = 0; i < 3; i++) res.x_true[i] = sim.xTrue[i];
    if (plot) plot->end();
    slamgpu_ekf_destroy(h);
    return res;
}

// ---- trace sink ---------------------------------------------------------------------------------------------------------
namespace {
struct TracePlot : PlotSink {
    FILE *fp;
    explicit TracePlot(FILE *f) : fp(f) {}
    ~TracePlot() override {
        if (fp) fclose(fp);
    }
    void begin(const Conf &, const Map &m, const std::string &method) override {
        fprintf(fp, "# %s landmarks=%zu waypoints=%zu\n# iteration x_true y_true t_true x_est y_est t_est n_obs\n",
                method.c_str(), m.lm_x.size(), m.wp_x.size());
    }
    void tick(int it, const float xt[3], const double xe[3], int n_obs, uint32_t) override {
        fprintf(fp, "%d %.9g %.9g %.9g %.12g %.12g %.12g %d\n", it, xt[0], xt[1], xt[2], xe[0], xe[1], xe[2], n_obs);
    }
    void end() override { fflush(fp); }
};
}  // namespace

namespace {
struct TeePlot : PlotSink {
    PlotSink *a, *b;
    TeePlot(PlotSink *x, PlotSink *y) : a(x), b(y) {}
    ~TeePlot() override {
        delete a;
        delete b;
    }
    void begin(const Conf &c, const Map &m, const std::string &s) override { a->begin(c, m, s); b->begin(c, m, s); }
    void tick(int i, const float t[3], const double e[3], int n, uint32_t us) override { a->tick(i, t, e, n, us); b->tick(i, t, e, n, us); }
    void particles(const float *x, int64_t n, const float *l, int nf) override { a->particles(x, n, l, nf); b->particles(x, n, l, nf); }
    void laser(const float *z, int n, const float x[3]) override { a->laser(z, n, x); b->laser(z, n, x); }
    bool wants_ellipses() override { return a->wants_ellipses() || b->wants_ellipses(); }
    void ellipse(const double xy[2], const float P[4], int i) override { a->ellipse(xy, P, i); b->ellipse(xy, P, i); }
    void end() override { a->end(); b->end(); }
};
}  // namespace
PlotSink *make_tee_plot(PlotSink *a, PlotSink *b) { return new TeePlot(a, b); }

PlotSink *make_trace_plot(const std::string &path) {
    FILE *f = fopen(path.c_str(), "w");
    return f ? new TracePlot(f) : nullptr;
}

}  // namespace slamhost
