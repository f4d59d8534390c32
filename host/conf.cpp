// conf.cpp — configuration and map loading of the host (reference: core.cpp:855-1073, utils.cpp:504-566,1032-1046).
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>

#include "slam_host.hpp"

namespace slamhost {

static std::string trim(const std::string &s) {
    const char *ws = " \t\r\n";
    const size_t a = s.find_first_not_of(ws);
    if (a == std::string::npos) return "";
    const size_t b = s.find_last_not_of(ws);
    return s.substr(a, b - a + 1);
}

// CParamArray::load (utils.cpp:504-566): "KEY = value" lines, '#' or ':' starts a comment line, the text between the
// first and the second '=' is the value.
int Conf::load(const std::string &path) {
    FILE *fp = fopen(path.c_str(), "rt");
    if (!fp) return -1;
    items.clear();
    char buf[4096];
    while (fgets(buf, sizeof(buf), fp)) {
        const std::string line = trim(buf);
        if (line.empty() || line[0] == '#' || line[0] == ':') continue;
        std::vector<std::string> parts;
        size_t pos = 0;
        while (true) {
            const size_t e = line.find('=', pos);
            if (e == std::string::npos) {
                parts.push_back(line.substr(pos));
                break;
            }
            parts.push_back(line.substr(pos, e - pos));
            pos = e + 1;
        }
        if (parts.size() >= 2) items.push_back({trim(parts[0]), trim(parts[1])});
    }
    fclose(fp);
    parse();
    return 0;
}

void Conf::set_s(const std::string &k, const std::string &v) {
    for (auto &it : items)
        if (it.first == k) {
            it.second = v;
            return;
        }
    items.push_back({k, v});
}

std::string Conf::s(const std::string &k) const {
    for (const auto &it : items)
        if (it.first == k) return it.second;
    return "";
}

// CParamArray::set_args (utils.cpp:1032-1046): any "-KEY value" pair becomes an item, whatever the key.
void Conf::set_args(int argc, char **argv) {
    for (int i = 1; i < argc; i++)
        if (argv[i][0] == '-' && i + 1 < argc) {
            const char *k = argv[i] + 1;
            set_s(k, argv[++i]);
        }
}

void Conf::parse() {
    // defaults, core.cpp:977-1028
    V = 3.0f;
    MAXG = (float) (30 * M_PI / 180);
    RATEG = (float) (20 * M_PI / 180);
    WHEELBASE = 4;
    DT_CONTROLS = 0.025f;
    sigmaV = 0.3f;
    sigmaG = (float) (3.0 * M_PI / 180);
    MAX_RANGE = 30.0f;
    DT_OBSERVE = 8 * DT_CONTROLS;
    sigmaR = 0.1f;
    sigmaB = (float) (1.0 * M_PI / 180);
    sigmaT = (float) (1.0 * M_PI / 180);
    GATE_REJECT = 4.0f;
    GATE_AUGMENT = 25.0f;
    AT_WAYPOINT = 1.0f;
    NUMBER_LOOPS = 2;
    NPARTICLES = 100;
    NEFFECTIVE = (int) (0.75 * NPARTICLES);  // NOT recomputed when NPARTICLES is overridden (core.cpp:1011-1012)
    SWITCH_CONTROL_NOISE = 1;
    SWITCH_SENSOR_NOISE = 1;
    SWITCH_INFLATE_NOISE = 0;
    SWITCH_PREDICT_NOISE = 0;
    SWITCH_SAMPLE_PROPOSAL = 1;
    SWITCH_HEADING_KNOWN = 1;
    SWITCH_RESAMPLE = 1;
    SWITCH_PROFILE = 1;
    SWITCH_SEED_RANDOM = 0;
    SWITCH_ASSOCIATION_KNOWN = 0;
    SWITCH_BATCH_UPDATE = 1;
    SWITCH_USE_IEKF = 0;
    auto f = [&](const char *k, float &v) {
        const std::string t = s(k);
        if (!t.empty()) v = (float) atof(t.c_str());
    };
    auto i = [&](const char *k, int &v) {
        const std::string t = s(k);
        if (!t.empty()) v = atoi(t.c_str());
    };
    f("Vtrue", V);  // the speed's key is Vtrue (core.cpp:1033)
    f("MAXG", MAXG);
    f("RATEG", RATEG);
    f("WHEELBASE", WHEELBASE);
    f("DT_CONTROLS", DT_CONTROLS);
    f("sigmaV", sigmaV);
    f("sigmaG", sigmaG);
    f("MAX_RANGE", MAX_RANGE);
    f("DT_OBSERVE", DT_OBSERVE);
    f("sigmaR", sigmaR);
    f("sigmaB", sigmaB);
    f("sigmaT", sigmaT);
    f("GATE_REJECT", GATE_REJECT);
    f("GATE_AUGMENT", GATE_AUGMENT);
    f("AT_WAYPOINT", AT_WAYPOINT);
    i("NUMBER_LOOPS", NUMBER_LOOPS);
    i("NPARTICLES", NPARTICLES);
    i("NEFFECTIVE", NEFFECTIVE);
    i("SWITCH_CONTROL_NOISE", SWITCH_CONTROL_NOISE);
    i("SWITCH_SENSOR_NOISE", SWITCH_SENSOR_NOISE);
    i("SWITCH_INFLATE_NOISE", SWITCH_INFLATE_NOISE);
    i("SWITCH_PREDICT_NOISE", SWITCH_PREDICT_NOISE);
    i("SWITCH_SAMPLE_PROPOSAL", SWITCH_SAMPLE_PROPOSAL);
    i("SWITCH_HEADING_KNOWN", SWITCH_HEADING_KNOWN);
    i("SWITCH_RESAMPLE", SWITCH_RESAMPLE);
    i("SWITCH_PROFILE", SWITCH_PROFILE);
    i("SWITCH_SEED_RANDOM", SWITCH_SEED_RANDOM);
    i("SWITCH_ASSOCIATION_KNOWN", SWITCH_ASSOCIATION_KNOWN);
    i("SWITCH_BATCH_UPDATE", SWITCH_BATCH_UPDATE);
    i("SWITCH_USE_IEKF", SWITCH_USE_IEKF);
}

std::string Conf::dump() const {
    char b[2048];
    snprintf(b, sizeof(b),
             "Vtrue=%.9g MAXG=%.9g RATEG=%.9g WHEELBASE=%.9g DT_CONTROLS=%.9g sigmaV=%.9g sigmaG=%.9g MAX_RANGE=%.9g "
             "DT_OBSERVE=%.9g sigmaR=%.9g sigmaB=%.9g sigmaT=%.9g GATE_REJECT=%.9g GATE_AUGMENT=%.9g AT_WAYPOINT=%.9g "
             "NUMBER_LOOPS=%d NPARTICLES=%d NEFFECTIVE=%d SWITCH_CONTROL_NOISE=%d SWITCH_SENSOR_NOISE=%d "
             "SWITCH_INFLATE_NOISE=%d SWITCH_PREDICT_NOISE=%d SWITCH_SAMPLE_PROPOSAL=%d SWITCH_HEADING_KNOWN=%d "
             "SWITCH_RESAMPLE=%d SWITCH_PROFILE=%d SWITCH_SEED_RANDOM=%d SWITCH_ASSOCIATION_KNOWN=%d "
             "SWITCH_BATCH_UPDATE=%d SWITCH_USE_IEKF=%d",
             V, MAXG, RATEG, WHEELBASE, DT_CONTROLS, sigmaV, sigmaG, MAX_RANGE, DT_OBSERVE, sigmaR, sigmaB, sigmaT,
             GATE_REJECT, GATE_AUGMENT, AT_WAYPOINT, NUMBER_LOOPS, NPARTICLES, NEFFECTIVE, SWITCH_CONTROL_NOISE,
             SWITCH_SENSOR_NOISE, SWITCH_INFLATE_NOISE, SWITCH_PREDICT_NOISE, SWITCH_SAMPLE_PROPOSAL,
             SWITCH_HEADING_KNOWN, SWITCH_RESAMPLE, SWITCH_PROFILE, SWITCH_SEED_RANDOM, SWITCH_ASSOCIATION_KNOWN,
             SWITCH_BATCH_UPDATE, SWITCH_USE_IEKF);
    return b;
}

// readInputFile (core.cpp:855-962).  Headers are "lm <rows> <cols>" and "wp <rows> <cols>"; the <cols> following
// lines hold one column each (its first <rows> tokens).  Like the reference the waypoint rows are read with the
// LANDMARK row count (core.cpp:948), which only matters for a file whose two blocks disagree.
std::string read_input_file(const std::string &path, Map *out) {
    std::ifstream in(path.c_str());
    if (!in) return "Unable to read input file " + path;
    out->lm_x.clear(); out->lm_y.clear(); out->wp_x.clear(); out->wp_y.clear();
    int lineno = 0;
    unsigned long lm_rows = 0;
    std::string str;
    auto tokens_of = [](const std::string &l) {
        std::istringstream is(l);
        std::vector<std::string> t;
        std::string w;
        while (is >> w) t.push_back(w);
        return t;
    };
    while (std::getline(in, str)) {
        lineno++;
        std::vector<std::string> tk = tokens_of(str);
        if (tk.empty() || tk[0][0] == '#') continue;
        const bool is_lm = tk[0] == "lm", is_wp = tk[0] == "wp";
        if (!is_lm && !is_wp) return "Unknown command " + tk[0] + " on line " + std::to_string(lineno);
        if (tk.size() != 3) return std::string("Wrong args for ") + tk[0] + " on line " + std::to_string(lineno);
        const unsigned long rows = (unsigned long) strtof(tk[1].c_str(), NULL);
        const unsigned long cols = (unsigned long) strtof(tk[2].c_str(), NULL);
        if (is_lm) lm_rows = rows;
        const unsigned long need = is_lm ? rows : lm_rows;
        if (need != 2) return "only 2-row lm / wp blocks are supported (x; y) on line " + std::to_string(lineno);
        for (unsigned long c = 0; c < cols; c++) {
            lineno++;
            if (!std::getline(in, str)) return "EOF after reading line " + std::to_string(lineno);
            std::vector<std::string> t = tokens_of(str);
            if (t.size() < rows) return "invalid line for " + tk[0] + " coordinate on line " + std::to_string(lineno);
            const float a = strtof(t[0].c_str(), NULL), b = strtof(t[1].c_str(), NULL);
            if (is_lm) {
                out->lm_x.push_back(a);
                out->lm_y.push_back(b);
            } else {
                out->wp_x.push_back(a);
                out->wp_y.push_back(b);
            }
        }
    }
    return "";
}

}  // namespace slamhost
