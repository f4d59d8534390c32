// ekf.cu — C ABI of the EKF-SLAM entry point (include/slamgpu.h, slamgpu_ekf_*): sequencing of the K9 kernels
// (ekf_kernels.cuh) behind EKFSLAM::sim (algorithms/ekfslam.cpp:18-43).  The joint state lives on the device; the
// host keeps only its size and the tiny association lists.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstring>
#include <string>
#include <vector>

#include "ekf_kernels.cuh"
#include "handle.h"

using namespace slam;

struct slamgpu_ekf {
    slamgpu_ekf_config cfg;
    int n = 3, n_cap = 0, ld = 0, m_cap = 0;
    cudaStream_t stream = nullptr;
    float *x = nullptr, *P = nullptr;
    float *W = nullptr, *row2 = nullptr, *col2 = nullptr;        // heading update scratch [n_cap]
    float *nis = nullptr, *nd = nullptr, *outer = nullptr;       // association [m_cap][nf_cap], [m_cap]
    int *jbest = nullptr;
    float2 *d_z = nullptr;                                       // [m_cap] observations / associated observations
    int *d_idf = nullptr;
    float *Hs = nullptr, *v = nullptr, *PHt = nullptr, *U = nullptr, *Ui = nullptr, *W1 = nullptr;
    void *h_pin = nullptr;                                       // pinned staging: z, idf, jbest, outer
    int64_t launches = 0;

    EkfState st() const { return EkfState{x, P, ld}; }
};

#define KCHECK()                                                                                             \
    do {                                                                                                     \
        cudaError_t e_ = cudaGetLastError();                                                                 \
        if (e_ != cudaSuccess) return slam_fail(SLAMGPU_ECUDA, std::string("kernel launch: ") + cudaGetErrorString(e_)); \
    } while (0)

static void ekf_free(slamgpu_ekf *h) {
    if (!h) return;
    cudaSetDevice(h->cfg.device);
    void *dev[] = {h->x, h->P, h->W, h->row2, h->col2, h->nis, h->nd, h->outer, h->jbest, h->d_z, h->d_idf,
                   h->Hs, h->v, h->PHt, h->U, h->Ui, h->W1};
    for (void *p : dev)
        if (p) cudaFree(p);
    if (h->h_pin) cudaFreeHost(h->h_pin);
    if (h->stream) cudaStreamDestroy(h->stream);
    delete h;
}

extern "C" int slamgpu_ekf_create(const slamgpu_ekf_config *cfg, slamgpu_ekf_t **out) {
    if (!cfg || !out) return slam_fail(SLAMGPU_EINVAL, "null argument");
    if (cfg->nf_cap <= 0) return slam_fail(SLAMGPU_EINVAL, "nf_cap must be positive");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0 || cfg->device >= ndev)
        return slam_fail(SLAMGPU_ENODEVICE, "no usable CUDA device: slamgpu has no CPU path");
    CU(cudaSetDevice(cfg->device));
    slamgpu_ekf *h = new slamgpu_ekf();
    h->cfg = *cfg;
    h->m_cap = cfg->m_cap > 0 ? cfg->m_cap : 64;
    h->n_cap = 3 + 2 * cfg->nf_cap;
    h->ld = (h->n_cap + 31) / 32 * 32;
    const size_t nc = h->n_cap, q = 2 * (size_t) h->m_cap;
    int rc = SLAMGPU_OK;
    do {
#define ALLOC(p, bytes)                                          \
    if (cudaMalloc(&(p), (bytes)) != cudaSuccess) {              \
        rc = slam_fail(SLAMGPU_ENOMEM, "EKF device allocation"); \
        break;                                                   \
    }
        if (cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking) != cudaSuccess) {
            rc = slam_fail(SLAMGPU_ECUDA, "stream creation");
            break;
        }
        ALLOC(h->x, sizeof(float) * nc);
        ALLOC(h->P, sizeof(float) * nc * h->ld);
        ALLOC(h->W, sizeof(float) * nc);
        ALLOC(h->row2, sizeof(float) * nc);
        ALLOC(h->col2, sizeof(float) * nc);
        ALLOC(h->nis, sizeof(float) * h->m_cap * (size_t) cfg->nf_cap);
        ALLOC(h->nd, sizeof(float) * h->m_cap * (size_t) cfg->nf_cap);
        ALLOC(h->outer, sizeof(float) * h->m_cap);
        ALLOC(h->jbest, sizeof(int) * h->m_cap);
        ALLOC(h->d_z, sizeof(float2) * h->m_cap);
        ALLOC(h->d_idf, sizeof(int) * h->m_cap);
        ALLOC(h->Hs, sizeof(float) * 10 * h->m_cap);
        ALLOC(h->v, sizeof(float) * q);
        ALLOC(h->PHt, sizeof(float) * nc * q);
        ALLOC(h->U, sizeof(float) * q * q);
        ALLOC(h->Ui, sizeof(float) * q * q);
        ALLOC(h->W1, sizeof(float) * nc * q);
#undef ALLOC
        if (cudaHostAlloc(&h->h_pin, 32 * (size_t) h->m_cap, cudaHostAllocDefault) != cudaSuccess) {
            rc = slam_fail(SLAMGPU_ENOMEM, "EKF pinned allocation");
            break;
        }
        // xEstimated = 0, P = zeros(3,3): slamwrapper / ekfslamwrapper.cpp:40-41
        if (cudaMemsetAsync(h->x, 0, sizeof(float) * nc, h->stream) != cudaSuccess ||
            cudaMemsetAsync(h->P, 0, sizeof(float) * nc * h->ld, h->stream) != cudaSuccess ||
            cudaStreamSynchronize(h->stream) != cudaSuccess) {
            rc = slam_fail(SLAMGPU_ECUDA, "EKF state initialisation");
            break;
        }
    } while (0);
    if (rc != SLAMGPU_OK) {
        ekf_free(h);
        return rc;
    }
    *out = h;
    return SLAMGPU_OK;
}

extern "C" void slamgpu_ekf_destroy(slamgpu_ekf_t *h) { ekf_free(h); }
extern "C" int slamgpu_ekf_size(slamgpu_ekf_t *h) { return h ? h->n : SLAMGPU_EINVAL; }
extern "C" int64_t slamgpu_ekf_launch_count(slamgpu_ekf_t *h) { return h ? h->launches : 0; }

extern "C" int slamgpu_ekf_set_state(slamgpu_ekf_t *h, int32_t n, const float *x, const float *P) {
    if (!h || !x || !P) return slam_fail(SLAMGPU_EINVAL, "null argument");
    if (n < 3 || n > h->n_cap || ((n - 3) & 1)) return slam_fail(SLAMGPU_ECAP, "state size outside [3, 3 + 2*nf_cap]");
    CU(cudaSetDevice(h->cfg.device));
    CU(cudaMemcpyAsync(h->x, x, sizeof(float) * n, cudaMemcpyHostToDevice, h->stream));
    CU(cudaMemcpy2DAsync(h->P, sizeof(float) * h->ld, P, sizeof(float) * n, sizeof(float) * n, n, cudaMemcpyHostToDevice,
                         h->stream));
    CU(cudaStreamSynchronize(h->stream));
    h->n = n;
    return SLAMGPU_OK;
}

extern "C" int slamgpu_ekf_get_state(slamgpu_ekf_t *h, float *x, float *P) {
    if (!h) return slam_fail(SLAMGPU_EINVAL, "null handle");
    CU(cudaSetDevice(h->cfg.device));
    const int n = h->n;
    if (x) CU(cudaMemcpyAsync(x, h->x, sizeof(float) * n, cudaMemcpyDeviceToHost, h->stream));
    if (P)
        CU(cudaMemcpy2DAsync(P, sizeof(float) * n, h->P, sizeof(float) * h->ld, sizeof(float) * n, n,
                             cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    return SLAMGPU_OK;
}

// The per-tick plot feed of ekfslamwrapper.cpp:91-101: the pose and its 3x3 covariance only.
extern "C" int slamgpu_ekf_get_pose(slamgpu_ekf_t *h, float xv[3], float Pvv[9]) {
    if (!h || !xv) return slam_fail(SLAMGPU_EINVAL, "null argument");
    CU(cudaSetDevice(h->cfg.device));
    CU(cudaMemcpyAsync(xv, h->x, sizeof(float) * 3, cudaMemcpyDeviceToHost, h->stream));
    if (Pvv)
        CU(cudaMemcpy2DAsync(Pvv, sizeof(float) * 3, h->P, sizeof(float) * h->ld, sizeof(float) * 3, 3,
                             cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    return SLAMGPU_OK;
}

static inline dim3 tiles(int cols, int rows) { return dim3((cols + 31) / 32, (rows + 7) / 8); }

// EKFSLAM::sim (ekfslam.cpp:18-43).
extern "C" int slamgpu_ekf_sim(slamgpu_ekf_t *h, float V, float G, const float Qe[4], float dt, float phi,
                               int32_t observe, const float *z, const int32_t *idz, int32_t m, const int32_t *table,
                               int32_t ntab, const float Re[4], const float R[4], int32_t *idf_out, int32_t *n_zf_out,
                               int32_t *n_zn_out) {
    if (!h || !Qe) return slam_fail(SLAMGPU_EINVAL, "null argument");
    if (observe && (m < 0 || (m > 0 && !z) || !Re || !R)) return slam_fail(SLAMGPU_EINVAL, "bad observation set");
    if (observe && m > h->m_cap) return slam_fail(SLAMGPU_ECAP, "more observations than m_cap");
    if (observe && h->cfg.association_known && m > 0 && (!idz || !table))
        return slam_fail(SLAMGPU_EINVAL, "known association needs idz and the association table");
    CU(cudaSetDevice(h->cfg.device));
    cudaStream_t st = h->stream;
    const EkfState s = h->st();
    int n = h->n;
    // predict position & heading (ekfslam.cpp:24), then the compass (ekfslam.cpp:27-29)
    k_ekf_predict<<<1, 256, 0, st>>>(s, n, V, G, make_float4(Qe[0], Qe[1], Qe[2], Qe[3]), h->cfg.wheel_base, dt);
    KCHECK();
    h->launches++;
    if (h->cfg.use_heading) {
        const float Rphi = (float) ((double) h->cfg.sigma_phi * (double) h->cfg.sigma_phi);  // pow(sigmaPhi, 2)
        k_ekf_heading_prep<<<1, 256, 0, st>>>(s, n, phi, Rphi, h->W, h->row2, h->col2);
        KCHECK();
        k_ekf_heading_apply<<<tiles(n, n), 256, 0, st>>>(s, n, Rphi, h->W, h->row2, h->col2);
        KCHECK();
        h->launches += 2;
    }
    int n_zf = 0, n_zn = 0;
    if (observe && m > 0) {
        float2 *hz = (float2 *) h->h_pin;                    // [m_cap]
        int *hidf = (int *) (hz + h->m_cap);                 // [m_cap]
        int *hj = hidf + h->m_cap;                           // [m_cap]
        float *ho = (float *) (hj + h->m_cap);               // [m_cap]
        std::vector<int> zf_idx, zn_idx, idf;
        const int Nf = (n - 3) / 2;
        if (h->cfg.association_known) {  // dataAssociateKnown, ekfslam.cpp:201-236 (the table is the caller's: by value)
            for (int i = 0; i < m; i++) {
                const int ii = idz[i];
                if (ii < 0 || ii >= ntab) return slam_fail(SLAMGPU_EINVAL, "landmark identifier outside the table");
                if (table[ii] == -1) zn_idx.push_back(i);
                else {
                    if (table[ii] < 0 || table[ii] >= Nf) return slam_fail(SLAMGPU_EINVAL, "table entry outside the map");
                    zf_idx.push_back(i);
                    idf.push_back(table[ii]);
                }
            }
        } else if (Nf == 0) {  // no landmark to compare with: jbest = -1 and outer stays +inf > gate2
            for (int i = 0; i < m; i++) zn_idx.push_back(i);
        } else {               // dataAssociate, ekfslam.cpp:151-189
            for (int i = 0; i < m; i++) hz[i] = make_float2(z[2 * i], z[2 * i + 1]);
            CU(cudaMemcpyAsync(h->d_z, hz, sizeof(float2) * m, cudaMemcpyHostToDevice, st));
            k_ekf_assoc_pairs<<<dim3((Nf + 127) / 128, m), 128, 0, st>>>(s, Nf, m, h->d_z,
                                                                          make_float4(Re[0], Re[1], Re[2], Re[3]), h->nis,
                                                                          h->nd);
            KCHECK();
            k_ekf_assoc_reduce<<<m, 256, 0, st>>>(Nf, h->nis, h->nd, h->cfg.gate_reject, h->jbest, h->outer);
            KCHECK();
            h->launches += 2;
            CU(cudaMemcpyAsync(hj, h->jbest, sizeof(int) * m, cudaMemcpyDeviceToHost, st));
            CU(cudaMemcpyAsync(ho, h->outer, sizeof(float) * m, cudaMemcpyDeviceToHost, st));
            CU(cudaStreamSynchronize(st));
            for (int i = 0; i < m; i++) {
                if (hj[i] > -1) {
                    zf_idx.push_back(i);
                    idf.push_back(hj[i]);
                } else if (ho[i] > h->cfg.gate_augment) {
                    zn_idx.push_back(i);
                }
            }
        }
        n_zf = (int) zf_idx.size();
        n_zn = (int) zn_idx.size();
        if (n + 2 * n_zn > h->n_cap) return slam_fail(SLAMGPU_ECAP, "landmark capacity (nf_cap) exceeded");
        if (h->cfg.batch_update && n_zf > 0) {  // batchUpdate, ekfslam.cpp:238-267
            for (int i = 0; i < n_zf; i++) {
                hz[i] = make_float2(z[2 * zf_idx[i]], z[2 * zf_idx[i] + 1]);
                hidf[i] = idf[i];
            }
            CU(cudaMemcpyAsync(h->d_z, hz, sizeof(float2) * n_zf, cudaMemcpyHostToDevice, st));
            CU(cudaMemcpyAsync(h->d_idf, hidf, sizeof(int) * n_zf, cudaMemcpyHostToDevice, st));
            EkfBatch b{h->Hs, h->v, h->PHt, h->U, h->Ui, h->W1, h->d_idf, h->d_z};
            k_ekf_innovation<<<(n_zf + 63) / 64, 64, 0, st>>>(s, b, n_zf);
            KCHECK();
            k_ekf_pht<<<tiles(2 * n_zf, n), 256, 0, st>>>(s, b, n, n_zf);
            KCHECK();
            k_ekf_factor<<<1, 256, 0, st>>>(b, n_zf, make_float4(R[0], R[1], R[2], R[3]));
            KCHECK();
            k_ekf_gain<<<(n + 127) / 128, 128, 0, st>>>(s, b, n, n_zf);
            KCHECK();
            k_ekf_downdate<<<tiles(n, n), 256, 0, st>>>(s, b, n, n_zf);
            KCHECK();
            h->launches += 5;
        }
        for (int i = 0; i < n_zn; i++) {  // augment, ekfslam.cpp:318-322
            k_ekf_add_one<<<1, 256, 0, st>>>(s, n, z[2 * zn_idx[i]], z[2 * zn_idx[i] + 1],
                                             make_float4(Re[0], Re[1], Re[2], Re[3]));
            KCHECK();
            h->launches++;
            n += 2;
        }
        if (idf_out)
            for (int i = 0; i < n_zf; i++) idf_out[i] = idf[i];
        // the pinned staging area is reused by the next call
        CU(cudaStreamSynchronize(st));
    }
    h->n = n;
    if (n_zf_out) *n_zf_out = n_zf;
    if (n_zn_out) *n_zn_out = n_zn;
    return SLAMGPU_OK;
}
