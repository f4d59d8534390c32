// slam_host.hpp — the C++ host side of the B200 backend: the reference's wrapper loop, map / configuration loading and
// world simulation helpers, rebuilt on top of the C ABI (include/slamgpu.h).  It is what `slam-backend` is in the
// reference (src/backend/slam-backend.cpp, SLAMBackendApplication.cpp, wrappers/*.cpp, the host parts of core.cpp):
// same command line, same .ini keys and precedence, same .mat text format, same consumption order of rand(), with
// the particle set / EKF state held by libslamgpu.so instead of vector<Particle> / Eigen matrices.
// No Eigen: everything here is a handful of floats per tick.
#pragma once

#include <cstdint>
#include <map>
#include <string>
#include <vector>

namespace slamhost {

// ---- Conf (core.h:17-90, core.cpp:971-1073; CParamArray utils.cpp:504-566,1032-1046) ---------------------------------
struct Conf {
    // control parameters
    float V, MAXG, RATEG, WHEELBASE, DT_CONTROLS;
    float sigmaV, sigmaG;
    float MAX_RANGE, DT_OBSERVE;
    float sigmaR, sigmaB, sigmaT;
    float GATE_REJECT, GATE_AUGMENT;
    float AT_WAYPOINT;
    int NUMBER_LOOPS, NPARTICLES, NEFFECTIVE;
    int SWITCH_CONTROL_NOISE, SWITCH_SENSOR_NOISE, SWITCH_INFLATE_NOISE, SWITCH_PREDICT_NOISE, SWITCH_SAMPLE_PROPOSAL,
        SWITCH_HEADING_KNOWN, SWITCH_RESAMPLE, SWITCH_PROFILE, SWITCH_SEED_RANDOM, SWITCH_ASSOCIATION_KNOWN,
        SWITCH_BATCH_UPDATE, SWITCH_USE_IEKF;

    // raw KEY = value items in load / set order (a later item with the same key wins, like CParamArray::set_s)
    std::vector<std::pair<std::string, std::string>> items;

    int load(const std::string &ini_path);          // 0 ok, -1 when the file cannot be opened (like the reference: ignored)
    void set_args(int argc, char **argv);           // every "-KEY value" pair, any key (utils.cpp:1032-1046)
    void set_s(const std::string &k, const std::string &v);
    std::string s(const std::string &k) const;      // "" when absent
    void parse();                                   // defaults <- items (core.cpp:971-1073)
    std::string dump() const;                       // one "KEY = value" line per parsed field (test hook)
};

// ---- map file (core.cpp:855-962): "lm <rows> <cols>" / "wp <rows> <cols>" headers, one column per line, '#' comments ---
struct Map {
    std::vector<float> lm_x, lm_y, wp_x, wp_y;
};
// Returns "" on success, else the error text (the reference prints it and exit()s; the caller decides here).
std::string read_input_file(const std::string &path, Map *out);

// ---- random numbers: everything goes through std::rand() in the reference's order (core.cpp:383-449,751-777) --------
void randn(int count, float *out);                  // nRandMat::randn(m, n) with m*n = count: consumes count + 1 rand()
double unif_rand();                                 // rand() / double(RAND_MAX)
float trig_offset(float a);                         // trigonometricOffset, core.cpp:460-477

// ---- world simulation (core.cpp:22-89, 176-273, 438-449) ------------------------------------------------------------
void add_control_noise(float V, float G, const float Q[4], float *Vn, float *Gn);
void predict_true_position(float x[3], float V, float G, float wheel_base, float dt);
void update_steering(const float x[3], const Map &map, int *index_of_first_waypoint, float minimum_distance, float *G,
                     float max_rate, float max_angle, float dt);
// Visible landmarks (semi-circle in front of the vehicle) and their range / bearing; ids_out = landmark identifiers.
void get_observations(const Map &map, const float x[3], float max_range, std::vector<int> *ids_out,
                      std::vector<float> *z_out /* [k][2] */);
void add_observation_noise(std::vector<float> *z, const float R[4]);
// dataAssociationKnown (core.cpp:91-120); the table holds floats in the reference (VectorXf), ints here.
void data_association_known(const std::vector<float> &z, const std::vector<int> &idz, std::vector<int> *table, int Nf,
                            std::vector<float> *zf, std::vector<int> *idf, std::vector<float> *zn);

// ---- plot sink: the calls the wrappers make on NetworkPlot (plotting/NetworkPlot.h) ----------------------------------
struct PlotSink {
    virtual ~PlotSink() {}
    virtual void begin(const Conf &, const Map &, const std::string &method) {}
    // one call per loop iteration, after the algorithm step
    virtual void tick(int iteration, const float x_true[3], const double x_est[3], int n_obs, uint32_t loop_us) {}
    virtual void particles(const float *xv, int64_t n, const float *lm_mean, int nf) {}  // decimated feed (xv [3][n])
    virtual void laser(const float *z, int n, const float x_true[3]) {}  // range / bearing of the last observation
    virtual bool wants_ellipses() { return false; }                      // EKF: covariance ellipses cost a state download
    virtual void ellipse(const double xy[2], const float P[4], int idx) {}
    virtual void end() {}
};
// Writes "iteration x_true y_true t_true x_est y_est t_est n_obs" lines; the GUI-less trace the tests read.
PlotSink *make_trace_plot(const std::string &path);
// The reference's NetworkPlot wire protocol (plotting/NetworkPlot.cpp) over a ZMQ_PAIR socket connected to `endpoint`
// (the reference hard-codes tcp://127.0.0.1:4242); libzmq is dlopen()ed.  nullptr + *err on failure.
PlotSink *make_zmq_plot(const std::string &endpoint, const std::string &simulation_name, std::string *err);
// Both at once (trace file + GUI).
PlotSink *make_tee_plot(PlotSink *a, PlotSink *b);

// ---- the run loops (wrappers/fastslam1wrapper.cpp, fastslam2wrapper.cpp, ekfslamwrapper.cpp) -------------------------
struct RunOptions {
    int device = 0;
    int max_iterations = 0;      // 0 = until the waypoints are done
    int64_t plot_stride = 0;     // 0 = no particle feed; k = every k-th particle to the sink
    int scan_mode = 0;           // SLAMGPU_SCAN_REFERENCE: bit-identical resampling at the reference's particle counts
    int math_mode = 0;           // SLAMGPU_MATH_FAST
    int nf_cap = 0;              // 0 = number of landmarks in the map
};
struct RunResult {
    int iterations = 0, observations = 0, resamples = 0, landmarks = 0;
    double x_est[3] = {0, 0, 0};
    float x_true[3] = {0, 0, 0};
    std::string error;           // "" = ok
};
RunResult run_fastslam(int method /*1|2*/, const Conf &conf, const Map &map, PlotSink *plot, const RunOptions &opt);
RunResult run_ekfslam(const Conf &conf, const Map &map, PlotSink *plot, const RunOptions &opt);

}  // namespace slamhost
