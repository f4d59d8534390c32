/* TEST INFRASTRUCTURE ONLY — CPU restatement of the reference's EKF-SLAM (see ekf_oracle.c).  Row-major P with
 * leading dimension ld. */
#ifndef EKF_ORACLE_H
#define EKF_ORACLE_H
#ifdef __cplusplus
extern "C" {
#endif

typedef struct so_ekf_config {
    float wheel_base, sigma_phi, gate_reject, gate_augment;
    int use_heading, batch_update, association_known;
} so_ekf_config;

void so_ekf_predict(int n, int ld, float *x, float *P, float V, float G, const float *Q, float wheel_base, float dt);
void so_ekf_observe_heading(int n, int ld, float *x, float *P, float phi, float sigma_phi);
void so_ekf_observe_model(const float *x, int idf, float *z, float *Hs);
void so_ekf_compute_association(int ld, const float *x, const float *P, const float *z, const float *R, int idf,
                                float *nis, float *nd);
void so_ekf_data_associate(int n, int ld, const float *x, const float *P, const float *z, int m, const float *R,
                           float gate1, float gate2, int *idf, int *zf_idx, int *n_zf, int *zn_idx, int *n_zn);
void so_ekf_data_associate_known(int n, const int *idz, int m, int *table, int *idf, int *zf_idx, int *n_zf,
                                 int *zn_idx, int *n_zn);
void so_ekf_batch_update(int n, int ld, float *x, float *P, const float *zf, const int *idf, int m, const float *R);
int so_ekf_add_one_z(int n, int ld, float *x, float *P, const float *z, const float *Re);
int so_ekf_sim(const so_ekf_config *cfg, int n, int ld, float *x, float *P, float V, float G, const float *Qe, float dt,
               float phi, int observe, const float *z, const int *idz, int m, int *table, const float *Re,
               const float *R, int *idf_out, int *n_zf_out, int *n_zn_out);

#ifdef __cplusplus
}
#endif
#endif
