// TEST INFRASTRUCTURE ONLY — a NetworkPlot (reference class, plotting/NetworkPlot.h) that writes the per-iteration
// true / estimated pose to the file named by $SLAMREF_TRACE instead of talking zmq to the GUI.  Linked with the
// UNMODIFIED reference wrappers by oracle/Makefile into oracle/_ref/slam-backend-ref, it turns the reference binary
// into a golden-trace generator for the host-loop parity tests (same line format as host/wrappers.cpp's trace sink).
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include "src/backend/plotting/NetworkPlot.h"

static FILE *g_fp = nullptr;
static uint32_t g_it = 0;
static double g_true[3], g_est[3];
static int g_nobs = 0;

NetworkPlot::NetworkPlot() : socket(nullptr), context(nullptr) {
    const char *p = getenv("SLAMREF_TRACE");
    g_fp = p ? fopen(p, "w") : nullptr;
}
NetworkPlot::~NetworkPlot() {
    if (g_fp) fclose(g_fp);
    g_fp = nullptr;
}
void NetworkPlot::setLandmarks(std::vector<double> &, std::vector<double> &) {}
void NetworkPlot::setWaypoints(std::vector<double> &, std::vector<double> &) {}
void NetworkPlot::setParticles(std::vector<double> &, std::vector<double> &) {}
void NetworkPlot::setFeatureParticles(std::vector<double> &, std::vector<double> &) {}
void NetworkPlot::setLaserLines(Eigen::MatrixXf &l) { g_nobs = (int) l.cols(); }
void NetworkPlot::setCovEllipse(Eigen::MatrixXf &, int) {}
void NetworkPlot::addTruePosition(double, double) {}
void NetworkPlot::addEstimatedPosition(double, double) {}
void NetworkPlot::setCarSize(double, uint32_t) {}
void NetworkPlot::setCarTruePosition(double x, double y, double t) { g_true[0] = x; g_true[1] = y; g_true[2] = t; }
void NetworkPlot::setCarEstimatedPosition(double x, double y, double t) { g_est[0] = x; g_est[1] = y; g_est[2] = t; }
void NetworkPlot::setPlotRange(double, double, double, double) {}
void NetworkPlot::clear(void) {}
void NetworkPlot::setSimulationName(std::string) {}
void NetworkPlot::plot() {
    if (g_fp && g_it > 0)
        fprintf(g_fp, "%u %.9g %.9g %.9g %.12g %.12g %.12g %d\n", g_it, g_true[0], g_true[1], g_true[2], g_est[0],
                g_est[1], g_est[2], g_nobs);
}
void NetworkPlot::endPlot() {
    if (g_fp) fflush(g_fp);
}
void NetworkPlot::setCurrentIteration(uint32_t it) { g_it = it; }
void NetworkPlot::covEllipseAdd(uint32_t) {}
void NetworkPlot::loopTime(uint32_t) {}
