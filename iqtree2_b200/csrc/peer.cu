// peer.cu -- host side of the in-kernel exchange between pattern shards in different processes (peer.cuh).
#include "peer.cuh"

#define PEER_BUF_BYTES (sizeof(double) * 2 * B200_PEER_MAX * 4 + sizeof(unsigned long long) * 2 * B200_PEER_MAX)

// This rank's exchange buffer and the 64-byte CUDA IPC handle the other ranks open it with.
extern "C" int b200lk_peer_handle(b200lk_ctx *c, void *handle64) {
    if (!c || !handle64) return B200LK_ERR_ARG;
    SHARD_REFUSE(c, "the cross-process peer exchange");
    CUDA_TRY(c, cudaSetDevice(c->device));
    if (!c->peer_buf) {
        CUDA_TRY(c, cudaMalloc(&c->peer_buf, PEER_BUF_BYTES));
        CUDA_TRY(c, cudaMemset(c->peer_buf, 0, PEER_BUF_BYTES));
    }
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handle size");
    cudaIpcMemHandle_t h;
    CUDA_TRY(c, cudaIpcGetMemHandle(&h, c->peer_buf));
    memcpy(handle64, &h, 64);
    return B200LK_OK;
}

// handles: [world][64], rank-ordered, as gathered from every rank's b200lk_peer_handle.  After this call
// b200lk_evaluate* leaves the sums over ALL ranks in the result buffer of every rank (no collective call needed);
// every rank must issue the same sequence of evaluations.
extern "C" int b200lk_peer_connect(b200lk_ctx *c, int32_t rank, int32_t world, const void *handles) {
    if (!c || !handles || world < 1 || world > B200_PEER_MAX || rank < 0 || rank >= world) return B200LK_ERR_ARG;
    SHARD_REFUSE(c, "the cross-process peer exchange");
    if (!c->peer_buf || c->peer_world) {
        snprintf(c->err, sizeof(c->err), "b200lk_peer_connect: call b200lk_peer_handle first, and connect once");
        return B200LK_ERR_STATE;
    }
    CUDA_TRY(c, cudaSetDevice(c->device));
    std::vector<double *> bufs(world);
    for (int r = 0; r < world; r++) {
        if (r == rank) { bufs[r] = c->peer_buf; continue; }
        cudaIpcMemHandle_t h;
        memcpy(&h, (const char *)handles + (size_t)r * 64, 64);
        void *p = nullptr;
        CUDA_TRY(c, cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
        c->peer_mapped[r] = p;
        bufs[r] = (double *)p;
    }
    CUDA_TRY(c, cudaMalloc(&c->d_peer_bufs, sizeof(double *) * world));
    CUDA_TRY(c, cudaMemcpy(c->d_peer_bufs, bufs.data(), sizeof(double *) * world, cudaMemcpyHostToDevice));
    c->peer_rank = rank;
    c->peer_world = world;
    c->peer_seq = 0;
    return B200LK_OK;
}

// Back to purely local sums (e.g. when another rank could not map the buffers and everybody falls back to a collective).
extern "C" int b200lk_peer_disconnect(b200lk_ctx *c) {
    if (!c) return B200LK_ERR_ARG;
    SHARD_REFUSE(c, "the cross-process peer exchange");
    CUDA_TRY(c, cudaSetDevice(c->device));
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    for (int r = 0; r < B200_PEER_MAX; r++)
        if (c->peer_mapped[r]) {
            cudaIpcCloseMemHandle(c->peer_mapped[r]);
            c->peer_mapped[r] = nullptr;
        }
    if (c->d_peer_bufs) cudaFree(c->d_peer_bufs);
    c->d_peer_bufs = nullptr;
    c->peer_world = 0;
    c->peer_rank = 0;
    c->peer_seq = 0;
    return B200LK_OK;
}

void peer_release(b200lk_ctx *c) {
    for (int r = 0; r < B200_PEER_MAX; r++)
        if (c->peer_mapped[r]) cudaIpcCloseMemHandle(c->peer_mapped[r]);
    if (c->d_peer_bufs) cudaFree(c->d_peer_bufs);
    if (c->peer_buf) cudaFree(c->peer_buf);
}

PeerArgs peer_args_next(b200lk_ctx *c) {
    PeerArgs pa;
    pa.bufs = c->d_peer_bufs;
    pa.rank = c->peer_rank;
    pa.world = c->peer_world;
    pa.seq = c->peer_world > 1 ? ++c->peer_seq : 0;
    return pa;
}

__global__ void peer_exchange_kernel(double *result, int count, PeerArgs pa) {
    if (threadIdx.x == 0 && blockIdx.x == 0) peer_exchange(result, count, pa);
}

// the exchange as a launch of its own, behind kernels that do not finish with it themselves
int launch_peer_exchange(b200lk_ctx *c, int count) {
    if (c->peer_world <= 1) return B200LK_OK;
    peer_exchange_kernel<<<1, 32, 0, c->stream>>>(c->d_result, count, peer_args_next(c));
    CUDA_TRY(c, cudaGetLastError());
    c->n_launches++;
    return B200LK_OK;
}
