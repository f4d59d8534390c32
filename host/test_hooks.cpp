// test_hooks.cpp — extern "C" windows onto the host functions for the ctypes parity tests (tests/test_host_*.py).
// Part of libslamhost.so only; the executable does not need it.
#include <cstdlib>
#include <cstring>

#include "slam_host.hpp"

using namespace slamhost;

extern "C" {

// Parses ini (may be "") then argv-style overrides "k1\0v1\0k2\0v2..." (n pairs); writes Conf::dump() into out.
int slamhost_conf_dump(const char *ini, const char **keys, const char **vals, int n, char *out, int cap) {
    Conf c;
    int rc = 0;
    if (ini && ini[0]) rc = c.load(ini);
    for (int i = 0; i < n; i++) c.set_s(keys[i], vals[i]);
    c.parse();
    strncpy(out, c.dump().c_str(), cap - 1);
    out[cap - 1] = 0;
    return rc;
}

// Returns 0 and the counts; a second call with buffers copies the coordinates.  -1 on a parse error (text in err).
int slamhost_read_map(const char *path, int *n_lm, int *n_wp, float *lm, float *wp, char *err, int cap) {
    Map m;
    const std::string e = read_input_file(path, &m);
    if (!e.empty()) {
        strncpy(err, e.c_str(), cap - 1);
        err[cap - 1] = 0;
        return -1;
    }
    *n_lm = (int) m.lm_x.size();
    *n_wp = (int) m.wp_x.size();
    if (lm)
        for (size_t i = 0; i < m.lm_x.size(); i++) {
            lm[2 * i] = m.lm_x[i];
            lm[2 * i + 1] = m.lm_y[i];
        }
    if (wp)
        for (size_t i = 0; i < m.wp_x.size(); i++) {
            wp[2 * i] = m.wp_x[i];
            wp[2 * i + 1] = m.wp_y[i];
        }
    return 0;
}

void slamhost_srand(unsigned s) { srand(s); }
void slamhost_randn(int count, float *out) { randn(count, out); }
double slamhost_unif_rand() { return unif_rand(); }
float slamhost_trig_offset(float a) { return trig_offset(a); }
void slamhost_add_control_noise(float V, float G, const float *Q, float *Vn, float *Gn) {
    add_control_noise(V, G, Q, Vn, Gn);
}
void slamhost_predict_true_position(float *x, float V, float G, float wb, float dt) {
    predict_true_position(x, V, G, wb, dt);
}

static Map g_map;
void slamhost_set_map(const float *lm, int n_lm, const float *wp, int n_wp) {
    g_map = Map();
    for (int i = 0; i < n_lm; i++) {
        g_map.lm_x.push_back(lm[2 * i]);
        g_map.lm_y.push_back(lm[2 * i + 1]);
    }
    for (int i = 0; i < n_wp; i++) {
        g_map.wp_x.push_back(wp[2 * i]);
        g_map.wp_y.push_back(wp[2 * i + 1]);
    }
}
void slamhost_update_steering(const float *x, int *iwp, float min_d, float *G, float rate, float maxg, float dt) {
    update_steering(x, g_map, iwp, min_d, G, rate, maxg, dt);
}
// Visible landmarks of the current map; add_noise draws the observation noise (two randn calls).  Returns the count.
int slamhost_get_observations(const float *x, float max_range, int add_noise, const float *R, int *ids, float *z) {
    std::vector<int> id;
    std::vector<float> zz;
    get_observations(g_map, x, max_range, &id, &zz);
    if (add_noise) add_observation_noise(&zz, R);
    for (size_t i = 0; i < id.size(); i++) {
        ids[i] = id[i];
        z[2 * i] = zz[2 * i];
        z[2 * i + 1] = zz[2 * i + 1];
    }
    return (int) id.size();
}
// dataAssociationKnown on a caller-held table; returns K, writes Kn.
int slamhost_data_association_known(const float *z, const int *idz, int m, int *table, int ntab, int Nf, float *zf,
                                    int *idf, float *zn, int *Kn) {
    std::vector<float> zv(z, z + 2 * m), zfv, znv;
    std::vector<int> iv(idz, idz + m), tv(table, table + ntab), idfv;
    data_association_known(zv, iv, &tv, Nf, &zfv, &idfv, &znv);
    for (int i = 0; i < ntab; i++) table[i] = tv[i];
    for (size_t i = 0; i < zfv.size(); i++) zf[i] = zfv[i];
    for (size_t i = 0; i < idfv.size(); i++) idf[i] = idfv[i];
    for (size_t i = 0; i < znv.size(); i++) zn[i] = znv[i];
    *Kn = (int) znv.size() / 2;
    return (int) idfv.size();
}

// Emits a fixed script through the zmq plot sink (no GPU involved): begin on the current map, one observation with two
// laser lines, a 3-particle / 2-landmark feed, one covariance ellipse, one tick, end.  0 ok, -1 + err text otherwise.
int slamhost_zmq_selftest(const char *endpoint, char *err, int cap) {
    std::string e;
    PlotSink *p = make_zmq_plot(endpoint, "selftest", &e);
    if (!p) {
        strncpy(err, e.c_str(), cap - 1);
        err[cap - 1] = 0;
        return -1;
    }
    Conf c;
    c.parse();
    p->begin(c, g_map, "FASTSLAM2");
    const float xt[3] = {1.0f, 2.0f, 0.5f};
    const float z[4] = {5.0f, 0.25f, 7.5f, -1.0f};
    p->laser(z, 2, xt);
    const float xv[9] = {0.1f, 0.2f, 0.3f, 1.1f, 1.2f, 1.3f, 0, 0, 0};                    // [3][n = 3]
    const float lm[12] = {10, 11, 12, 13, 14, 15, 20, 21, 22, 23, 24, 25};                  // [nf = 2][n = 3][2]
    p->particles(xv, 3, lm, 2);
    const double cxy[2] = {1.0, 2.0};
    const float P[4] = {0.04f, 0.01f, 0.01f, 0.09f};
    p->ellipse(cxy, P, 0);
    const double xe[3] = {1.25, 1.75, 0.45};
    p->tick(1, xt, xe, 2, 1234u);
    p->end();
    delete p;
    return 0;
}

}  // extern "C"
