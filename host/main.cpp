// slam-backend-gpu — command-line twin of the reference's `slam-backend` (src/backend/slam-backend.cpp,
// SLAMBackendApplication.cpp:59-89): same options, same <map>.ini next to the map, same precedence
// (defaults <- ini <- "-KEY value" for any key), method fall-through to EKF.  Extra options of this build:
//   -plot ENDPOINT  feed the reference's slam-gui over its zmq protocol ("default" = tcp://127.0.0.1:4242, the address
//                   the reference hard-codes; the reference always connects, here it is opt-in)
//   -trace FILE     write the per-iteration true / estimated pose (the plot feed without a GUI)
//   -max_iter N     stop after N iterations
//   -device D       CUDA device ordinal
//   -plot_stride K  feed every K-th particle to the sink (0 = none)
//   -scan MODE      reference | blocked   (summation order behind the resampling CDF, include/slamgpu.h)
//   -math MODE      fast | reference      (per-particle arithmetic: the B200 kernels' own / the Eigen call order)
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>

#include "slam_host.hpp"

using namespace slamhost;

static void print_usage(char **argv) {
    printf("%s\n", argv[0]);
    printf("    -m                  [s] input map file name\n");
    printf("    -n                  [s] experiment name\n");
    printf("    -mode               [s] running mode\n");
    printf("        waypoints   : following given waypoints\n");
    printf("        interactive : use keyboard to control movement\n");
    printf("    -method             [s] SLAM method\n");
    printf("        EKF1        : EKF SLAM 1\n");
    printf("        FASTSLAM1       : FastSLAM 1\n");
    printf("        FASTSLAM2       : FastSLAM 2\n");
    printf("    -h  (print usage)\n");
    printf("\n");
}

int main(int argc, char **argv) {
    std::string map_file = "example_webmap.mat", name = "simulation";
    for (int i = 1; i < argc; i++) {
        if (strcmp(argv[i], "-m") == 0 && i + 1 < argc) map_file = argv[i + 1];
        if (strcmp(argv[i], "-n") == 0 && i + 1 < argc) name = argv[i + 1];
        if (strcmp(argv[i], "-h") == 0) {
            print_usage(argv);
            return 0;
        }
    }
    printf("map: %s\n", map_file.c_str());
    const size_t dot = map_file.find_last_of('.');
    const std::string ini = (dot == std::string::npos ? map_file : map_file.substr(0, dot)) + ".ini";
    Conf conf;
    conf.load(ini);  // a missing .ini leaves the defaults, like the reference
    conf.set_args(argc, argv);
    conf.parse();
    conf.set_s("simulation_name", name);
    if (conf.s("mode") == "interactive") {
        fprintf(stderr, "interactive mode needs the GUI's key feed, which is commented out in the reference too "
                        "(slamwrapper.cpp:192-224); running the waypoints\n");
    }
    Map map;
    const std::string err = read_input_file(map_file, &map);
    if (!err.empty()) {
        fprintf(stderr, "%s\n", err.c_str());
        return EXIT_FAILURE;
    }
    RunOptions opt;
    opt.device = atoi(conf.s("device").c_str());
    opt.max_iterations = atoi(conf.s("max_iter").c_str());
    opt.plot_stride = atoll(conf.s("plot_stride").c_str());
    const std::string scan = conf.s("scan");
    // the reference order is one device thread: right for the reference's own particle counts, hopeless at 10^6
    opt.scan_mode = scan == "blocked" ? 1 : (scan == "reference" ? 0 : (conf.NPARTICLES > 20000 ? 1 : 0));
    opt.math_mode = conf.s("math") == "reference" ? 1 : 0;
    std::unique_ptr<PlotSink> sink;
    if (!conf.s("trace").empty()) {
        sink.reset(make_trace_plot(conf.s("trace")));
        if (!sink) {
            fprintf(stderr, "cannot open trace file %s\n", conf.s("trace").c_str());
            return EXIT_FAILURE;
        }
    }
    if (!conf.s("plot").empty()) {
        std::string e;
        const std::string ep = conf.s("plot") == "default" ? "tcp://127.0.0.1:4242" : conf.s("plot");
        PlotSink *zp = make_zmq_plot(ep, name, &e);
        if (!zp) {
            fprintf(stderr, "%s\n", e.c_str());
            return EXIT_FAILURE;
        }
        if (sink) sink.reset(make_tee_plot(sink.release(), zp));
        else sink.reset(zp);
        if (opt.plot_stride == 0) opt.plot_stride = conf.NPARTICLES > 2000 ? conf.NPARTICLES / 1000 : 1;
    }
    const std::string method = conf.s("method");
    RunResult r;
    if (method == "FASTSLAM1") {
        printf("FastSLAM 1\n\n");
        r = run_fastslam(1, conf, map, sink.get(), opt);
    } else if (method == "FASTSLAM2") {
        printf("FastSLAM 2\n\n");
        r = run_fastslam(2, conf, map, sink.get(), opt);
    } else {  // anything else is EKF (SLAMBackendApplication.cpp:26-29)
        printf("EKFSLAM\n\n");
        r = run_ekfslam(conf, map, sink.get(), opt);
    }
    if (!r.error.empty()) {
        fprintf(stderr, "slam-backend-gpu: %s\n", r.error.c_str());
        return EXIT_FAILURE;
    }
    printf("iterations %d observations %d resamples %d landmarks %d\n", r.iterations, r.observations, r.resamples,
           r.landmarks);
    printf("true  %.6f %.6f %.6f\nest   %.6f %.6f %.6f\n", r.x_true[0], r.x_true[1], r.x_true[2], r.x_est[0], r.x_est[1],
           r.x_est[2]);
    return 0;
}
