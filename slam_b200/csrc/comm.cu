// comm.cu — exchange steps of the sharded resample, one process per device.
//
//   * the small exchanges of the resample (per-1024-block weight totals; CDF totals, w^2 totals and in-block maxima;
//     the 2-slot uniform halo; the tail record) go over peer memory: every rank stores its chunk into every peer's
//     copy and waits on arrival counters (xchg.cuh).  The two exchanges of the weights pass run INSIDE the cooperative
//     weights kernel (kernels.cuh k_weights_fused, CTA 0 between the arrival and the release of a grid-wide barrier);
//     the tail record and the occasional barrier are the stand-alone 1-CTA kernel k_xchg below.  The CDF itself is
//     never moved, the search reads it through the peer mappings;
//   * particle migration is NOT a collective: every rank maps its peers' pose and landmark buffers (CUDA IPC over
//     NVLink / NVSwitch) and the gather / update kernels (kernels.cuh K7, K3 REMAP) read particles[keep[i]] straight
//     from the owning device, with a barrier only where the row tags of capi.cu order_rows ask for one;
//   * NCCL remains for the bootstrap (the IPC handles travel in one all-gather) and the all-gather of keep[] and of the
//     coarse ancestor maps into the log (its own stream, off the critical path).  It is loaded at run time
//     from libnccl.so.2 (the copy torch already mapped when the caller is a torchrun rank, else the system one) so
//     that a single-device build has no NCCL dependency.
#include <dlfcn.h>
#include <nccl.h>

#include <cstdlib>
#include <cstring>

#include "handle.h"
#include "xchg.cuh"

using namespace slam;


struct NcclApi {
    void *lib = nullptr;
    decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
    decltype(&ncclCommInitRank) CommInitRank = nullptr;
    decltype(&ncclCommDestroy) CommDestroy = nullptr;
    decltype(&ncclAllGather) AllGather = nullptr;
    decltype(&ncclGetErrorString) GetErrorString = nullptr;
};

static int load_nccl(NcclApi *api, std::string *err) {
    if (api->lib) return 0;
    void *lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) {
        *err = std::string("dlopen(libnccl.so.2): ") + dlerror();
        return -1;
    }
#define SYM(name)                                                              \
    api->name = (decltype(api->name)) dlsym(lib, "nccl" #name);                \
    if (!api->name) {                                                          \
        *err = "libnccl.so.2 lacks nccl" #name;                                \
        return -1;                                                             \
    }
    SYM(GetUniqueId) SYM(CommInitRank) SYM(CommDestroy) SYM(AllGather) SYM(GetErrorString)
#undef SYM
    api->lib = lib;
    return 0;
}

static NcclApi g_nccl;

struct SlamComm {
    int rank = 0, world = 1, device = 0;
    ncclComm_t comm = nullptr;
    bool peers_open = false;
    PeerMaps peers;
    void *opened[SLAM_MAX_RANKS][9];
    unsigned long long epoch = 0;  // exchanges issued so far (k_xchg arrival counter)
    unsigned char *d_small = nullptr;  // scratch for handle exchange, barriers, estimate
    unsigned char *h_small = nullptr;
};

#define NC(call)                                                                                         \
    do {                                                                                                 \
        ncclResult_t r_ = (call);                                                                        \
        if (r_ != ncclSuccess)                                                                           \
            return slam_fail(SLAMGPU_ECOMM, std::string(#call) + ": " + g_nccl.GetErrorString(r_));      \
    } while (0)

int comm_unique_id(void *id128, std::string *err) {
    if (load_nccl(&g_nccl, err) != 0) return -1;
    ncclUniqueId id;
    ncclResult_t r = g_nccl.GetUniqueId(&id);
    if (r != ncclSuccess) {
        *err = std::string("ncclGetUniqueId: ") + g_nccl.GetErrorString(r);
        return -1;
    }
    static_assert(sizeof(id) == 128, "ncclUniqueId is 128 bytes");
    memcpy(id128, &id, 128);
    return 0;
}

SlamComm *comm_create(int rank, int world, const void *id128, int device, std::string *err) {
    if (world > SLAM_MAX_RANKS) {
        *err = "at most 8 devices (one NVSwitch box)";
        return nullptr;
    }
    if (load_nccl(&g_nccl, err) != 0) return nullptr;
    SlamComm *c = new SlamComm();
    c->rank = rank;
    c->world = world;
    c->device = device;
    memset(&c->peers, 0, sizeof(c->peers));
    memset(c->opened, 0, sizeof(c->opened));
    ncclUniqueId id;
    memcpy(&id, id128, 128);
    ncclResult_t r = g_nccl.CommInitRank(&c->comm, world, id, rank);
    if (r != ncclSuccess) {
        *err = std::string("ncclCommInitRank: ") + g_nccl.GetErrorString(r);
        delete c;
        return nullptr;
    }
    if (cudaMalloc(&c->d_small, 8192) != cudaSuccess || cudaHostAlloc(&c->h_small, 8192, cudaHostAllocDefault) != cudaSuccess) {
        *err = "comm scratch allocation failed";
        g_nccl.CommDestroy(c->comm);
        delete c;
        return nullptr;
    }
    return c;
}

void comm_destroy(SlamComm *c) {
    if (!c) return;
    for (int r = 0; r < c->world; r++)
        for (int j = 0; j < 9; j++)
            if (c->opened[r][j]) cudaIpcCloseMemHandle(c->opened[r][j]);
    if (c->comm) g_nccl.CommDestroy(c->comm);
    if (c->d_small) cudaFree(c->d_small);
    if (c->h_small) cudaFreeHost(c->h_small);
    delete c;
}



static int open_peers(SlamComm *c, slamgpu *h) {
    if (c->peers_open) return SLAMGPU_OK;
    cudaIpcMemHandle_t mine[9];
    void *ptrs[9] = {h->pose[0], h->pose[1], h->lm_mean, h->lm_cov, h->cdf, h->mloc, h->tot_slab, h->xflags, h->unif_base};
    for (int j = 0; j < 9; j++) CU(cudaIpcGetMemHandle(&mine[j], ptrs[j]));
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle is 64 bytes");
    const size_t per = sizeof(mine);  // 576 B per rank
    CU(cudaMemcpyAsync(c->d_small + (size_t) c->rank * per, mine, per, cudaMemcpyHostToDevice, h->stream));
    NC(g_nccl.AllGather(c->d_small + (size_t) c->rank * per, c->d_small, per, ncclChar, c->comm, h->stream));
    CU(cudaMemcpyAsync(c->h_small, c->d_small, per * c->world, cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    for (int r = 0; r < c->world; r++) {
        void *p[9];
        if (r == c->rank) {
            for (int j = 0; j < 9; j++) p[j] = ptrs[j];
        } else {
            cudaIpcMemHandle_t hd[9];
            memcpy(hd, c->h_small + (size_t) r * per, per);
            for (int j = 0; j < 9; j++) {
                CU(cudaIpcOpenMemHandle(&p[j], hd[j], cudaIpcMemLazyEnablePeerAccess));
                c->opened[r][j] = p[j];
            }
        }
        c->peers.pose[0][r] = (const float *) p[0];
        c->peers.pose[1][r] = (const float *) p[1];
        c->peers.mean[r] = (const float2 *) p[2];
        c->peers.cov[r] = (const float4 *) p[3];
        c->peers.cdf[r] = (const float *) p[4];
        c->peers.mloc[r] = (const float *) p[5];
        c->peers.slab[r] = (float *) p[6];
        c->peers.flags[r] = (unsigned long long *) p[7];
        c->peers.unif[r] = (double *) p[8];
    }
    c->peers.world = c->world;
    c->peers_open = true;
    return SLAMGPU_OK;
}

int comm_peer_maps(SlamComm *c, slamgpu *h, PeerMaps *out) {
    RET(open_peers(c, h));
    *out = c->peers;
    return SLAMGPU_OK;
}


// ---- small exchanges over peer memory (xchg.cuh) ---------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_xchg(const XchgArgs a) { xchg_body(a); }

int comm_fill_xchg(SlamComm *c, slamgpu *h, const int *arrays, int narr, bool halo, const void *blob, int blob_bytes,
                   int blob_slot, XchgArgs *out) {
    RET(open_peers(c, h));
    XchgArgs &a = *out;
    memset(&a, 0, sizeof(a));
    for (int r = 0; r < c->world; r++) {
        a.slab[r] = c->peers.slab[r];
        a.flags[r] = c->peers.flags[r];
        a.unif[r] = c->peers.unif[r];
    }
    a.rank = c->rank;
    a.world = c->world;
    a.epoch = ++c->epoch;
    a.narr = narr;
    for (int k = 0; k < narr; k++) a.arr_off[k] = arrays[k] * (int) h->nbg;
    a.b0 = (int) (h->slot0 / SLAM_SCAN_BLOCK);
    a.count = (int) h->nb;
    a.halo = halo ? 1 : 0;
    a.P = h->P;
    a.error = &h->d_sc->comm_error;
    if (blob) {
        a.blob_src = (const unsigned *) blob;
        a.blob_words = blob_bytes / 4;
        a.blob_off = (int) SLAM_BLOB_BASE(h->nbg) + blob_slot * SLAM_MAX_RANKS * SLAM_XCHG_BLOB_WORDS;
    }
    return SLAMGPU_OK;
}

int comm_exchange(SlamComm *c, slamgpu *h, const int *arrays, int narr, bool halo, const void *blob, int blob_bytes,
                  int blob_slot) {
    XchgArgs a;
    RET(comm_fill_xchg(c, h, arrays, narr, halo, blob, blob_bytes, blob_slot, &a));
    k_xchg<<<1, 256, 0, h->stream>>>(a);
    if (cudaGetLastError() != cudaSuccess) return slam_fail(SLAMGPU_ECOMM, "peer-memory exchange launch failed");
    h->launches++;
    return SLAMGPU_OK;
}

// The one NCCL call on the data path: keep[] of every rank into every rank's log (capi.cu resample_state, own stream).
int comm_allgather_i32(SlamComm *c, int *buf, size_t count, cudaStream_t st) {
    NC(g_nccl.AllGather(buf + (size_t) c->rank * count, buf, count, ncclInt32, c->comm, st));
    return SLAMGPU_OK;
}
