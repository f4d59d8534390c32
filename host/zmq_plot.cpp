// zmq_plot.cpp — the GUI feed: a PlotSink that speaks the reference's NetworkPlot wire protocol
// (plotting/NetworkPlot.cpp:7-218 over zmqpp) so that the reference's `slam-gui` can sit on the other end unchanged.
//
// Protocol, as the reference emits it: a ZMQ_PAIR socket CONNECTS to tcp://127.0.0.1:4242; every call is one multipart
// message whose first frame is the method name (no terminator) and every further value is its own frame in zmqpp's
// encoding (libs/zmqpp/message.cpp:225-328): 32-bit integers in network byte order, float / double byte-swapped to
// big-endian, strings raw.  setLandmarks / setWaypoints / setParticles / setFeatureParticles carry
// [int32 n][n doubles][int32 n][n doubles] and are sent BLOCKING (NetworkPlot.cpp:33); everything else is sent with
// ZMQ_DONTWAIT and silently dropped when the peer is not ready (NetworkPlot.cpp:80 ...).  setCurrentIteration is
// disabled in the reference (NetworkPlot.cpp:180-183) and therefore never sent.
//
// libzmq is loaded at run time (no headers on the box): $SLAMGPU_LIBZMQ, else libzmq.so.5 / libzmq.so.
#include <dlfcn.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "slam_host.hpp"

namespace slamhost {

namespace {

constexpr int kZmqPair = 0, kZmqDontWait = 1, kZmqSndMore = 2, kZmqLinger = 17;

struct ZmqApi {
    void *lib = nullptr;
    void *(*ctx_new)() = nullptr;
    int (*ctx_term)(void *) = nullptr;
    void *(*socket)(void *, int) = nullptr;
    int (*close)(void *) = nullptr;
    int (*connect)(void *, const char *) = nullptr;
    int (*setsockopt)(void *, int, const void *, size_t) = nullptr;
    int (*send)(void *, const void *, size_t, int) = nullptr;
};

bool load_zmq(ZmqApi *a, std::string *err) {
    const char *env = getenv("SLAMGPU_LIBZMQ");
    const char *names[] = {env, "libzmq.so.5", "libzmq.so"};
    for (const char *n : names) {
        if (!n || !n[0]) continue;
        a->lib = dlopen(n, RTLD_NOW | RTLD_LOCAL);
        if (a->lib) break;
    }
    if (!a->lib) {
        *err = "libzmq not found (set SLAMGPU_LIBZMQ to its path)";
        return false;
    }
#define SYM(field, name)                                            \
    a->field = (decltype(a->field)) dlsym(a->lib, name);            \
    if (!a->field) {                                                \
        *err = std::string("libzmq lacks ") + name;                 \
        return false;                                               \
    }
    SYM(ctx_new, "zmq_ctx_new") SYM(ctx_term, "zmq_ctx_term") SYM(socket, "zmq_socket") SYM(close, "zmq_close")
    SYM(connect, "zmq_connect") SYM(setsockopt, "zmq_setsockopt") SYM(send, "zmq_send")
#undef SYM
    return true;
}

// one multipart message under construction: frames in zmqpp's encoding
struct Msg {
    std::vector<std::string> parts;
    Msg &str(const std::string &s) {
        parts.push_back(s);
        return *this;
    }
    Msg &be(const void *p, size_t n) {  // n bytes, reversed to big-endian (x86 host)
        std::string b((const char *) p, n);
        for (size_t i = 0; i < n / 2; i++) std::swap(b[i], b[n - 1 - i]);
        parts.push_back(b);
        return *this;
    }
    Msg &i32(int32_t v) { return be(&v, 4); }
    Msg &u32(uint32_t v) { return be(&v, 4); }
    Msg &f32(float v) { return be(&v, 4); }
    Msg &f64(double v) { return be(&v, 8); }
};

struct ZmqPlot : PlotSink {
    ZmqApi z;
    void *ctx = nullptr, *sock = nullptr;
    std::string name;
    std::vector<float> lines;  // laser lines [4][n] of the last observation (core.cpp:329-356)
    int n_lines = 0;
    std::vector<double> px, py, fx, fy;
    bool have_particles = false;

    ~ZmqPlot() override {
        if (sock) z.close(sock);      // NetworkPlot::~NetworkPlot, NetworkPlot.cpp:17-20
        if (ctx) z.ctx_term(ctx);
        if (z.lib) dlclose(z.lib);
    }

    bool send(const Msg &m, bool dont_block) {
        for (size_t i = 0; i < m.parts.size(); i++) {
            int flags = (i + 1 < m.parts.size()) ? kZmqSndMore : 0;
            if (dont_block) flags |= kZmqDontWait;
            if (z.send(sock, m.parts[i].data(), m.parts[i].size(), flags) < 0) return false;  // dropped, like zmqpp
        }
        return true;
    }
    void xy(const char *method, const std::vector<double> &xs, const std::vector<double> &ys) {
        Msg m;
        m.str(method).i32((int32_t) xs.size());
        for (double v : xs) m.f64(v);
        m.i32((int32_t) ys.size());
        for (double v : ys) m.f64(v);
        send(m, false);
    }
    void matrix(Msg &m, const float *a, int rows, int cols) {
        m.u32((uint32_t) rows).u32((uint32_t) cols);
        for (int i = 0; i < rows * cols; i++) m.f32(a[i]);
    }

    // SLAMBackendApplication.cpp:20 + SLAMWrapper::configurePlot (slamwrapper.cpp:96-138, setPlotRange :140-171)
    void begin(const Conf &conf, const Map &map, const std::string &) override {
        send(Msg().str("setSimulationName").str(name), true);
        send(Msg().str("setCarSize").f64(conf.WHEELBASE).u32(0), true);
        send(Msg().str("setCarSize").f64(conf.WHEELBASE).u32(1), true);
        std::vector<double> wx(map.wp_x.begin(), map.wp_x.end()), wy(map.wp_y.begin(), map.wp_y.end());
        std::vector<double> lx(map.lm_x.begin(), map.lm_x.end()), ly(map.lm_y.begin(), map.lm_y.end());
        xy("setWaypoints", wx, wy);
        xy("setLandmarks", lx, ly);
        double xmin = 1e30, xmax = -1e30, ymin = 1e30, ymax = -1e30;
        for (size_t i = 0; i < wx.size(); i++) {
            xmax = std::max(xmax, wx[i]); xmin = std::min(xmin, wx[i]);
            ymax = std::max(ymax, wy[i]); ymin = std::min(ymin, wy[i]);
        }
        for (size_t i = 0; i < lx.size(); i++) {
            xmax = std::max(xmax, lx[i]); xmin = std::min(xmin, lx[i]);
            ymax = std::max(ymax, ly[i]); ymin = std::min(ymin, ly[i]);
        }
        send(Msg().str("setPlotRange").f64(xmin - (xmax - xmin) * 0.05).f64(xmax + (xmax - xmin) * 0.05)
                 .f64(ymin - (ymax - ymin) * 0.05).f64(ymax + (ymax - ymin) * 0.05), true);
        send(Msg().str("addTruePosition").f64(0).f64(0), true);
        send(Msg().str("setCarTruePosition").f64(0).f64(0).f64(0), true);
        send(Msg().str("addEstimatedPosition").f64(0).f64(0), true);
        send(Msg().str("setCarEstimatedPosition").f64(0).f64(0).f64(0), true);
        send(Msg().str("plot"), true);
    }

    // makeLaserLines (core.cpp:329-356): each observation as a segment from the vehicle to the sighted point
    void laser(const float *zz, int n, const float x[3]) override {
        n_lines = n;
        lines.assign(4 * (size_t) n, 0.0f);
        const float c = std::cos(x[2]), s = std::sin(x[2]);
        for (int j = 0; j < n; j++) {
            const float gx = zz[2 * j] * std::cos(zz[2 * j + 1]), gy = zz[2 * j] * std::sin(zz[2 * j + 1]);
            lines[0 * n + j] = x[0];
            lines[1 * n + j] = x[1];
            lines[2 * n + j] = (c * gx + -s * gy) + x[0];
            lines[3 * n + j] = (s * gx + c * gy) + x[1];
        }
    }

    void particles(const float *xv, int64_t n, const float *lm_mean, int nf) override {
        px.assign(xv, xv + n);
        py.assign(xv + n, xv + 2 * n);
        fx.clear();
        fy.clear();
        for (int64_t p = 0; p < n; p++)  // particle-major, landmark-minor like drawFeatureParticles
            for (int l = 0; l < nf; l++) {
                fx.push_back(lm_mean[((size_t) l * n + p) * 2]);
                fy.push_back(lm_mean[((size_t) l * n + p) * 2 + 1]);
            }
        have_particles = true;
    }

    bool wants_ellipses() override { return true; }
    void ellipse(const double xy2[2], const float P[4], int idx) override {
        // makeCovarianceEllipse (core.cpp:359-381): 2-sigma ellipse through the matrix square root of the 2x2 block
        const double a = P[0], b = 0.5 * ((double) P[1] + P[2]), d = P[3];
        const double det = std::max(0.0, a * d - b * b), sdet = std::sqrt(det);
        const double t = std::sqrt(std::max(1e-300, a + d + 2 * sdet));
        const double r00 = (a + sdet) / t, r01 = b / t, r11 = (d + sdet) / t;
        const int N = 16;
        std::vector<float> L(2 * (N + 1));
        for (int i = 0; i <= N; i++) {
            const float p = (float) (i * (2.0 * M_PI / N)), cp = std::cos(p), sp = std::sin(p);
            L[i] = (float) (xy2[0] + 2.0 * (r00 * cp + r01 * sp));
            L[N + 1 + i] = (float) (xy2[1] + 2.0 * (r01 * cp + r11 * sp));
        }
        Msg m;
        m.str("setCovEllipse");
        matrix(m, L.data(), 2, N + 1);
        m.i32(idx);
        send(m, true);
    }

    // the per-iteration tail of the run loops (fastslam2wrapper.cpp:93-117, ekfslamwrapper.cpp:86-105)
    void tick(int, const float xt[3], const double xe[3], int, uint32_t loop_us) override {
        send(Msg().str("loopTime").u32(loop_us), true);
        if (have_particles) {
            xy("setParticles", px, py);
            xy("setFeatureParticles", fx, fy);
        }
        send(Msg().str("addTruePosition").f64(xt[0]).f64(xt[1]), true);
        send(Msg().str("addEstimatedPosition").f64(xe[0]).f64(xe[1]), true);
        send(Msg().str("setCarTruePosition").f64(xt[0]).f64(xt[1]).f64(xt[2]), true);
        send(Msg().str("setCarEstimatedPosition").f64(xe[0]).f64(xe[1]).f64(xe[2]), true);
        Msg m;
        m.str("setLaserLines");
        matrix(m, lines.data(), n_lines ? 4 : 0, n_lines);  // MatrixXf(0, 0) before the first observation
        send(m, true);
        send(Msg().str("plot"), true);
    }

    void end() override { send(Msg().str("endPlot"), true); }
};

}  // namespace

PlotSink *make_zmq_plot(const std::string &endpoint, const std::string &simulation_name, std::string *err) {
    ZmqPlot *p = new ZmqPlot();
    p->name = simulation_name;
    if (!load_zmq(&p->z, err)) {
        delete p;
        return nullptr;
    }
    p->ctx = p->z.ctx_new();
    p->sock = p->ctx ? p->z.socket(p->ctx, kZmqPair) : nullptr;
    const int linger = 200;  // ms: do not hang on exit when no GUI ever connected
    if (!p->sock || p->z.setsockopt(p->sock, kZmqLinger, &linger, sizeof(linger)) != 0 ||
        p->z.connect(p->sock, endpoint.c_str()) != 0) {
        *err = "cannot connect the plot socket to " + endpoint;
        delete p;
        return nullptr;
    }
    printf("Connected to plotting server\n");
    return p;
}

}  // namespace slamhost
