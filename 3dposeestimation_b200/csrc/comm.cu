/*
 * comm.cu - multi-GPU entry points: one process per GPU, NCCL over NVLink / NVSwitch.
 *
 * New relative to the reference, which is single-process.  The template DB is
 * row-sharded; every rank scans its shard for the same queries, the per-shard candidate lists
 * (k x {global index, exact distance} per query) are all-gathered and merged identically on
 * every rank, ties -> lower global index, so the result equals the unsharded
 * `knnMatch` (reference src/networkevaluator.cpp:267).
 *
 * NCCL is bound at run time (dlopen of libnccl.so.2) so the library loads, and every
 * single-GPU entry point works, on a machine without NCCL.
 */
#include <dlfcn.h>
#include <nccl.h>

#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

#include "common.h"

namespace pe {
namespace {

struct NcclApi {
    void* handle = nullptr;
    decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
    decltype(&ncclCommInitRank) CommInitRank = nullptr;
    decltype(&ncclCommDestroy) CommDestroy = nullptr;
    decltype(&ncclAllGather) AllGather = nullptr;
    decltype(&ncclAllReduce) AllReduce = nullptr;
    decltype(&ncclGetErrorString) GetErrorString = nullptr;
};

NcclApi g_nccl;
std::once_flag g_nccl_once;

int nccl_load() {
    std::call_once(g_nccl_once, [] {
        void* h = nullptr;
        for (const char* name : {"libnccl.so.2", "libnccl.so"}) {
            h = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
            if (h) break;
        }
        if (!h) return;
        NcclApi a;
        a.handle = h;
        a.GetUniqueId = reinterpret_cast<decltype(a.GetUniqueId)>(dlsym(h, "ncclGetUniqueId"));
        a.CommInitRank = reinterpret_cast<decltype(a.CommInitRank)>(dlsym(h, "ncclCommInitRank"));
        a.CommDestroy = reinterpret_cast<decltype(a.CommDestroy)>(dlsym(h, "ncclCommDestroy"));
        a.AllGather = reinterpret_cast<decltype(a.AllGather)>(dlsym(h, "ncclAllGather"));
        a.AllReduce = reinterpret_cast<decltype(a.AllReduce)>(dlsym(h, "ncclAllReduce"));
        a.GetErrorString = reinterpret_cast<decltype(a.GetErrorString)>(dlsym(h, "ncclGetErrorString"));
        if (a.GetUniqueId && a.CommInitRank && a.CommDestroy && a.AllGather && a.AllReduce && a.GetErrorString) g_nccl = a;
    });
    if (!g_nccl.handle) {
        set_error("NCCL (libnccl.so.2) could not be loaded: %s", dlerror() ? dlerror() : "missing symbols");
        return PE_ERR_NCCL;
    }
    return PE_OK;
}

#define PE_NCCL(expr)                                                                                   \
    do {                                                                                                \
        ncclResult_t _r = (expr);                                                                       \
        if (_r != ncclSuccess) {                                                                        \
            set_error("%s:%d: %s failed: %s", __FILE__, __LINE__, #expr, g_nccl.GetErrorString(_r));    \
            return PE_ERR_NCCL;                                                                         \
        }                                                                                               \
    } while (0)

}  // namespace
}  // namespace pe

using namespace pe;

/* Peer window (NVLink / NVSwitch peer memory): every rank owns [flags | 2 parities x world slots x kWinSlot bytes] in its
 * HBM and maps the windows of all peers (CUDA IPC).  An all-gather is then ONE kernel per rank: push my block into
 * slot[rank] of every peer with 16-byte stores over NVLink, publish the call's epoch in every peer's flag word,
 * wait for the epochs of all peers in my own flags, copy my window out.  The messages of this path are 0.26 - 1.3 MB
 * per rank: latency bound, which is where a single kernel beats a collective library call. */
constexpr size_t kWinSlot = 2u << 20;         /* bytes per rank and call */
constexpr size_t kWinFlags = 4096;
constexpr int kWinMaxWorld = 16;

struct PeerWin {
    unsigned char* base[kWinMaxWorld];        /* base[r] = rank r's window as mapped here */
};

struct pe_comm {
    int rank = 0, world = 1, device = 0;
    ncclComm_t comm = nullptr;
    cudaStream_t stream = nullptr;
    pe::DevBuf send, recv;
    bool p2p = false;                         /* all ranks mapped all windows */
    unsigned char* win = nullptr;             /* my window */
    PeerWin peers{};
    unsigned int* done = nullptr;             /* CTA counter of the push phase */
    unsigned int epoch = 0;                   /* calls so far: identical on every rank */
    int* p2p_err = nullptr;                   /* mapped host flag: a flag wait timed out */
};

namespace pe_p2p {

__device__ __forceinline__ void st_release_sys(unsigned int* p, unsigned int v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

/* grid <= number of SMs (all CTAs resident: the waiting phase must not keep a pushing CTA off the machine) */
__global__ void __launch_bounds__(256) p2p_allgather_kernel(PeerWin peers, int rank, int world, unsigned int epoch,
                                                            const uint4* __restrict__ send, size_t n16, uint4* __restrict__ recv,
                                                            unsigned int* done, int* err) {
    const size_t slot16 = kWinSlot / 16;
    const size_t par_off = kWinFlags / 16 + (size_t)(epoch & 1u) * world * slot16;
    /* 1. push: my block -> slot[rank] of every window (my own included) */
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, nth = (size_t)gridDim.x * blockDim.x;
    for (size_t i = tid; i < n16 * world; i += nth) {          /* one (peer, 16 bytes) store per thread and step */
        const size_t r = i / n16, j = i - r * n16;
        reinterpret_cast<uint4*>(peers.base[r])[par_off + (size_t)rank * slot16 + j] = send[j];
    }
    __threadfence_system();
    __syncthreads();
    /* 2. the last CTA to finish publishes the epoch in flag[rank] of every window */
    __shared__ bool last;
    if (threadIdx.x == 0) {
        last = atomicAdd(done, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (last) {
        if (threadIdx.x == 0) *done = 0;
        __threadfence_system();
        if ((int)threadIdx.x < world) st_release_sys(reinterpret_cast<unsigned int*>(peers.base[threadIdx.x]) + rank, epoch);
    }
    /* 3. wait for every rank's epoch in my flags (bounded: a missing peer becomes an error, not a hung GPU) */
    if ((int)threadIdx.x < world) {
        const unsigned int* f = reinterpret_cast<const unsigned int*>(peers.base[rank]) + threadIdx.x;
        bool ok = false;
        for (unsigned int spin = 0; spin < (1u << 24); ++spin) {
            if ((int)(ld_acquire_sys(f) - epoch) >= 0) { ok = true; break; }
            if (spin > 64) __nanosleep(32);
        }
        if (!ok) atomicExch(err, 1);
    }
    __syncthreads();
    /* 4. my window -> recv (rank-major, n16 per rank) */
    const uint4* w = reinterpret_cast<const uint4*>(peers.base[rank]) + par_off;
    for (size_t i = tid; i < n16 * world; i += nth) {
        const size_t r = i / n16, j = i - r * n16;
        recv[i] = __ldcg(w + r * slot16 + j);      /* L2: the lines were written by peers */
    }
}

}  // namespace pe_p2p
using pe_p2p::p2p_allgather_kernel;

namespace {

/* all-gather of `bytes` per rank on `stream`: the peer-window kernel when it applies, NCCL otherwise */
int comm_allgather(pe_comm* comm, const void* d_send, void* d_recv, size_t bytes, cudaStream_t stream, int num_sms) {
    if (comm->p2p && g_opt_p2p.load() != 0 && bytes % 16 == 0 && bytes <= kWinSlot &&
        (reinterpret_cast<uintptr_t>(d_send) & 15) == 0 && (reinterpret_cast<uintptr_t>(d_recv) & 15) == 0) {
        ++comm->epoch;
        size_t ctas = (bytes / 16 * comm->world + 255) / 256;
        if (ctas > (size_t)num_sms) ctas = num_sms;
        if (ctas < 1) ctas = 1;
        p2p_allgather_kernel<<<(unsigned)ctas, 256, 0, stream>>>(comm->peers, comm->rank, comm->world, comm->epoch,
                                                               static_cast<const uint4*>(d_send), bytes / 16, static_cast<uint4*>(d_recv),
                                                               comm->done, comm->p2p_err);
        PE_LAUNCHED();
        return PE_OK;
    }
    PE_NCCL(g_nccl.AllGather(d_send, d_recv, bytes, ncclInt8, comm->comm, stream));
    return PE_OK;
}

/* local top-k -> all-gather of [idx block | dist block] per rank -> merge.  T is float or int.
 * Every rank scans its shard for all nq queries; the merged neighbours are written for queries
 * [q_first, q_first + q_count) only (q_count < 0: all of them), which is what a rank that owns a slice of the query
 * batch needs (pe_match_patches*). */
template <typename T>
int sharded_knn(pe_db* db, pe_comm* comm, const void* d_queries, int64_t nq, int k, int32_t* d_out_idx, T* d_out_dist,
                bool hamming, int64_t q_first = 0, int64_t q_count = -1) {
    PE_REQUIRE(db != nullptr, "null db");
    PE_REQUIRE(nq >= 0 && k >= 1 && k <= 16, "bad k / nq");
    if (nq == 0) return PE_OK;
    if (q_count < 0) { q_first = 0; q_count = nq; }
    const int world = comm ? comm->world : 1;
    if (world == 1) {
        const size_t row = hamming ? (size_t)db->words : (size_t)db->dim;
        const void* q = static_cast<const char*>(d_queries) + (size_t)q_first * row * 4;
        return hamming ? knn_hamming_local(db, static_cast<const uint32_t*>(q), q_count, k, d_out_idx, reinterpret_cast<int32_t*>(d_out_dist))
                       : knn_l2_local(db, static_cast<const float*>(q), q_count, k, d_out_idx, reinterpret_cast<float*>(d_out_dist));
    }
    PE_REQUIRE(comm->device == db->device, "communicator and DB live on different devices");
    PE_CUDA(cudaSetDevice(db->device));
    db->last_comm = comm;
    const size_t block = (size_t)nq * k; /* entries per rank and per half */
    PE_CHECK(db->gather.reserve((size_t)(world + 1) * 2 * block * 4));
    int32_t* send = db->gather.as<int32_t>();          /* [idx block | dist block] of this rank */
    int32_t* recv = send + 2 * block;                  /* world x the same */
    if (hamming)
        PE_CHECK(knn_hamming_local(db, static_cast<const uint32_t*>(d_queries), nq, k, send, send + block));
    else
        PE_CHECK(knn_l2_local(db, static_cast<const float*>(d_queries), nq, k, send, reinterpret_cast<float*>(send + block)));
    PE_CHECK(comm_allgather(comm, send, recv, 2 * block * 4, db->stream, db->num_sms));
    /* candidate (q, shard, p): idx at shard*2*block + q*k + p, dist `block` entries further */
    const int32_t* ci = recv + (size_t)q_first * k;
    if (hamming)
        return knn_merge_i32(db, ci + block, ci, 2 * (long long)block, world, q_count, k, d_out_idx, reinterpret_cast<int32_t*>(d_out_dist));
    return knn_merge_f32(db, reinterpret_cast<const float*>(ci + block), ci, 2 * (long long)block, world, q_count, k, d_out_idx,
                         reinterpret_cast<float*>(d_out_dist));
}

/* one stage of the fused test pass on the DB's stream: descriptors of this rank's m patches (already queued on the
 * net's stream, `ready` recorded behind them) -> [all-gather ->] scan -> select [-> all-gather -> merge] -> the m x k
 * neighbours of this rank's patches.  Both collectives run on the DB's stream. */
int match_stage(pe_net* net, pe_db* db, pe_comm* comm, const float* d_desc, int64_t m, int k, int32_t* r_i, float* r_d) {
    const int world = comm ? comm->world : 1, rank = comm ? comm->rank : 0;
    const int D = db->dim;
    if (db->stream != net->stream) {
        PE_CUDA(cudaEventRecord(db->ev_match, net->stream));
        PE_CUDA(cudaStreamWaitEvent(db->stream, db->ev_match, 0));
    }
    if (world == 1) return knn_l2_local(db, d_desc, m, k, r_i, r_d);
    PE_CHECK(db->m_q.reserve((size_t)world * m * D * 4));
    PE_CHECK(comm_allgather(comm, d_desc, db->m_q.p, (size_t)m * D * 4, db->stream, db->num_sms));
    return sharded_knn<float>(db, comm, db->m_q.p, world * m, k, r_i, r_d, false, rank * m, m);
}

}  // namespace

namespace pe {
int comm_check(pe_comm* comm) {
    if (!comm || !comm->p2p_err) return PE_OK;
    if (*reinterpret_cast<volatile int*>(comm->p2p_err) == 0) return PE_OK;
    *comm->p2p_err = 0;
    set_error("peer-window all-gather: a rank did not publish its epoch in time (results of the call are invalid)");
    return PE_ERR_NCCL;
}
}  // namespace pe

extern "C" {

int pe_comm_unique_id(char* id128) {
    PE_REQUIRE(id128 != nullptr, "null pointer");
    PE_CHECK(nccl_load());
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    ncclUniqueId id;
    PE_NCCL(g_nccl.GetUniqueId(&id));
    memcpy(id128, &id, 128);
    return PE_OK;
}

int pe_comm_create(int rank, int world, const char* id128, int device, pe_comm** out) {
    PE_REQUIRE(out && id128, "null pointer");
    *out = nullptr;
    PE_REQUIRE(world >= 1 && rank >= 0 && rank < world, "rank %d outside world %d", rank, world);
    PE_CHECK(nccl_load());
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || device < 0 || device >= n) {
        cudaGetLastError();
        set_error("device %d not available", device);
        return PE_ERR_CUDA;
    }
    PE_CUDA(cudaSetDevice(device));
    pe_comm* c = new pe_comm();
    c->rank = rank; c->world = world; c->device = device;
    ncclUniqueId id;
    memcpy(&id, id128, 128);
    ncclResult_t r = g_nccl.CommInitRank(&c->comm, world, id, rank);
    if (r != ncclSuccess) {
        set_error("ncclCommInitRank failed: %s", g_nccl.GetErrorString(r));
        delete c;
        return PE_ERR_NCCL;
    }
    if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess) {
        g_nccl.CommDestroy(c->comm);
        delete c;
        set_error("cudaStreamCreate failed");
        return PE_ERR_CUDA;
    }
    /* peer windows: allocate, exchange the IPC handles through NCCL, map.  Any failure on any rank leaves every rank on
     * the NCCL all-gather (the outcome is agreed with an all-reduce), it is not an error. */
    if (world > 1 && world <= kWinMaxWorld && getenv("PE_NO_P2P") == nullptr) {
        const size_t win_bytes = kWinFlags + 2 * (size_t)world * kWinSlot;
        int ok = 1;
        cudaIpcMemHandle_t mine;
        memset(&mine, 0, sizeof(mine));
        if (cudaMalloc(reinterpret_cast<void**>(&c->win), win_bytes) != cudaSuccess) { ok = 0; c->win = nullptr; }
        if (ok && cudaMemset(c->win, 0, kWinFlags) != cudaSuccess) ok = 0;
        if (ok && cudaIpcGetMemHandle(&mine, c->win) != cudaSuccess) ok = 0;
        if (ok && cudaMalloc(reinterpret_cast<void**>(&c->done), 4) != cudaSuccess) ok = 0;
        if (ok && cudaMemset(c->done, 0, 4) != cudaSuccess) ok = 0;
        if (ok && cudaHostAlloc(reinterpret_cast<void**>(&c->p2p_err), sizeof(int), cudaHostAllocMapped) != cudaSuccess) ok = 0;
        if (ok) *c->p2p_err = 0;
        cudaGetLastError();
        /* [handle 64 B | ok 4 B | pad] per rank */
        constexpr size_t rec = 128;
        std::vector<unsigned char> h_all((size_t)world * rec, 0), h_me(rec, 0);
        memcpy(h_me.data(), &mine, sizeof(mine));
        memcpy(h_me.data() + 64, &ok, 4);
        unsigned char* d_x = nullptr;
        bool exchanged = false;
        if (cudaMalloc(reinterpret_cast<void**>(&d_x), (size_t)(world + 1) * rec) == cudaSuccess) {
            if (cudaMemcpy(d_x, h_me.data(), rec, cudaMemcpyHostToDevice) == cudaSuccess &&
                g_nccl.AllGather(d_x, d_x + rec, rec, ncclInt8, c->comm, c->stream) == ncclSuccess &&
                cudaStreamSynchronize(c->stream) == cudaSuccess &&
                cudaMemcpy(h_all.data(), d_x + rec, (size_t)world * rec, cudaMemcpyDeviceToHost) == cudaSuccess)
                exchanged = true;
        }
        int all_ok = exchanged ? 1 : 0;
        for (int r = 0; r < world && all_ok; ++r) {
            int okr = 0;
            memcpy(&okr, h_all.data() + (size_t)r * rec + 64, 4);
            if (!okr) all_ok = 0;
        }
        if (all_ok) {
            for (int r = 0; r < world; ++r) {
                if (r == rank) { c->peers.base[r] = c->win; continue; }
                cudaIpcMemHandle_t hnd;
                memcpy(&hnd, h_all.data() + (size_t)r * rec, sizeof(hnd));
                void* ptr = nullptr;
                if (cudaIpcOpenMemHandle(&ptr, hnd, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { all_ok = 0; cudaGetLastError(); break; }
                c->peers.base[r] = static_cast<unsigned char*>(ptr);
            }
        }
        /* second agreement round: did everybody map everything? */
        int agreed = 0;
        if (exchanged) {
            memset(h_me.data(), 0, rec);
            memcpy(h_me.data(), &all_ok, 4);
            if (cudaMemcpy(d_x, h_me.data(), rec, cudaMemcpyHostToDevice) == cudaSuccess &&
                g_nccl.AllGather(d_x, d_x + rec, rec, ncclInt8, c->comm, c->stream) == ncclSuccess &&
                cudaStreamSynchronize(c->stream) == cudaSuccess &&
                cudaMemcpy(h_all.data(), d_x + rec, (size_t)world * rec, cudaMemcpyDeviceToHost) == cudaSuccess) {
                agreed = 1;
                for (int r = 0; r < world; ++r) {
                    int okr = 0;
                    memcpy(&okr, h_all.data() + (size_t)r * rec, 4);
                    if (!okr) agreed = 0;
                }
            }
        }
        if (d_x) cudaFree(d_x);
        cudaGetLastError();
        c->p2p = agreed != 0;
    }
    *out = c;
    return PE_OK;
}

void pe_comm_destroy(pe_comm* comm) {
    if (!comm) return;
    cudaSetDevice(comm->device);
    if (comm->stream) { cudaStreamSynchronize(comm->stream); cudaStreamDestroy(comm->stream); }
    cudaDeviceSynchronize();
    for (int r = 0; r < comm->world && r < kWinMaxWorld; ++r)
        if (r != comm->rank && comm->peers.base[r]) cudaIpcCloseMemHandle(comm->peers.base[r]);
    if (comm->win) cudaFree(comm->win);
    if (comm->done) cudaFree(comm->done);
    if (comm->p2p_err) cudaFreeHost(comm->p2p_err);
    cudaGetLastError();
    comm->send.release();
    comm->recv.release();
    if (comm->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(comm->comm);
    delete comm;
}

int pe_comm_info(const pe_comm* comm, int* rank, int* world) {
    PE_REQUIRE(comm != nullptr, "null comm");
    if (rank) *rank = comm->rank;
    if (world) *world = comm->world;
    return PE_OK;
}

int pe_comm_p2p(const pe_comm* comm, int* active) {
    PE_REQUIRE(comm != nullptr, "null comm");
    if (active) *active = comm->p2p && g_opt_p2p.load() != 0 ? 1 : 0;
    if (comm->p2p_err && *reinterpret_cast<volatile int*>(comm->p2p_err) != 0) {
        set_error("peer-window all-gather: a rank did not publish its epoch in time");
        return PE_ERR_NCCL;
    }
    return PE_OK;
}

int pe_comm_allgather_f32(pe_comm* comm, const float* send, int64_t n, float* recv) {
    PE_REQUIRE(comm && recv && n >= 0 && (send || n == 0), "bad arguments");
    if (n == 0) return PE_OK;
    PE_CUDA(cudaSetDevice(comm->device));
    PE_CHECK(comm->send.reserve((size_t)n * 4));
    PE_CHECK(comm->recv.reserve((size_t)n * 4 * comm->world));
    PE_CUDA(cudaMemcpyAsync(comm->send.p, send, (size_t)n * 4, cudaMemcpyHostToDevice, comm->stream));
    PE_NCCL(g_nccl.AllGather(comm->send.p, comm->recv.p, (size_t)n, ncclFloat32, comm->comm, comm->stream));
    PE_CUDA(cudaMemcpyAsync(recv, comm->recv.p, (size_t)n * 4 * comm->world, cudaMemcpyDeviceToHost, comm->stream));
    PE_CUDA(cudaStreamSynchronize(comm->stream));
    return PE_OK;
}

int pe_comm_allgather_f32_dev(pe_comm* comm, pe_net* net, const float* d_send, int64_t n, float* d_recv) {
    PE_REQUIRE(comm && net && n >= 0 && ((d_send && d_recv) || n == 0), "bad arguments");
    PE_REQUIRE(comm->device == net->device, "communicator and net live on different devices");
    if (n == 0) return PE_OK;
    PE_CHECK(net_ensure_device(net));
    /* always NCCL: the peer-window protocol (epochs, window parity) is only sound for all-gathers that follow each other on
     * ONE stream, and that stream is the DB's (matcher exchanges); this call runs on the net's */
    PE_NCCL(g_nccl.AllGather(d_send, d_recv, (size_t)n, ncclFloat32, comm->comm, net->stream));
    return PE_OK;
}

int pe_comm_allreduce_sum_f64(pe_comm* comm, double* buf, int64_t n) {
    PE_REQUIRE(comm && n >= 0 && (buf || n == 0), "bad arguments");
    if (n == 0) return PE_OK;
    PE_CUDA(cudaSetDevice(comm->device));
    PE_CHECK(comm->send.reserve((size_t)n * 8));
    PE_CUDA(cudaMemcpyAsync(comm->send.p, buf, (size_t)n * 8, cudaMemcpyHostToDevice, comm->stream));
    PE_NCCL(g_nccl.AllReduce(comm->send.p, comm->send.p, (size_t)n, ncclFloat64, ncclSum, comm->comm, comm->stream));
    PE_CUDA(cudaMemcpyAsync(buf, comm->send.p, (size_t)n * 8, cudaMemcpyDeviceToHost, comm->stream));
    PE_CUDA(cudaStreamSynchronize(comm->stream));
    return PE_OK;
}

int pe_knn_sharded_dev(pe_db* db, pe_comm* comm, const float* d_queries, int64_t nq, int k, int32_t* d_out_idx,
                       float* d_out_dist) {
    PE_REQUIRE(db && !db->hamming, "need an L2 DB");
    return sharded_knn<float>(db, comm, d_queries, nq, k, d_out_idx, d_out_dist, false);
}

int pe_knn_sharded(pe_db* db, pe_comm* comm, const float* queries, int64_t nq, int k, int32_t* out_idx, float* out_dist) {
    PE_REQUIRE(db && !db->hamming, "need an L2 DB");
    PE_REQUIRE(nq >= 0 && k >= 1 && k <= 16, "bad k / nq");
    if (nq == 0) return PE_OK;
    PE_REQUIRE(queries && out_idx && out_dist, "null pointer");
    PE_CUDA(cudaSetDevice(db->device));
    PE_CHECK(db->q_in.reserve((size_t)nq * db->dim * 4));
    PE_CHECK(db->out_i.reserve((size_t)nq * k * 4));
    PE_CHECK(db->out_d.reserve((size_t)nq * k * 4));
    PE_CUDA(cudaMemcpyAsync(db->q_in.p, queries, (size_t)nq * db->dim * 4, cudaMemcpyHostToDevice, db->stream));
    PE_CHECK(sharded_knn<float>(db, comm, db->q_in.p, nq, k, db->out_i.as<int32_t>(), db->out_d.as<float>(), false));
    PE_CUDA(cudaMemcpyAsync(out_idx, db->out_i.p, (size_t)nq * k * 4, cudaMemcpyDeviceToHost, db->stream));
    PE_CUDA(cudaMemcpyAsync(out_dist, db->out_d.p, (size_t)nq * k * 4, cudaMemcpyDeviceToHost, db->stream));
    PE_CUDA(cudaStreamSynchronize(db->stream));
    PE_CHECK(comm_check(comm));
    return knn_tc_check(db);
}

/* Fused test pass of networkSolver::computeKNN / networkEvaluator::computeKNNAccuracy's first half
 * (reference src/networksolver.cpp:456-472, src/networkevaluator.cpp:182-197): descriptors of n host patches, then
 * their k nearest templates.  The batch is cut into 512-patch stages; the H2D copy of stage i+1 runs under the forward
 * and the DB scan of stage i, so on a PCIe-bound box the call takes the copy time plus one stage.  With a
 * communicator every rank must pass the same n; the DB is row-sharded and each stage's descriptors are all-gathered. */
static int match_patches_once(pe_net* net, pe_db* db, pe_comm* comm, const float* patches_hwc, int64_t n, int patch_channels,
                              int k, int32_t* out_idx, float* out_dist, float* out_desc) {
    const int D = net->def.desc_dim;
    const size_t patch_floats = (size_t)4096 * patch_channels;
    const int64_t sub = 512;
    const int64_t stage_n = n < sub ? n : sub;
    PE_CHECK(net->in_stage[0].reserve(stage_n * patch_floats * sizeof(float)));
    if (n > sub) PE_CHECK(net->in_stage[1].reserve(stage_n * patch_floats * sizeof(float)));
    PE_CHECK(net->out_stage.reserve((size_t)n * D * sizeof(float)));
    PE_CHECK(db->m_ri.reserve((size_t)n * k * 4));
    PE_CHECK(db->m_rd.reserve((size_t)n * k * 4));
    if (!db->ev_match) PE_CUDA(cudaEventCreateWithFlags(&db->ev_match, cudaEventDisableTiming));
    int it = 0;
    for (int64_t first = 0; first < n; first += sub, ++it) {
        const int64_t m = n - first < sub ? n - first : sub;
        const int b = it & 1;
        if (it >= 2) PE_CUDA(cudaStreamWaitEvent(net->copy_stream, net->ev_done[b], 0));
        PE_CUDA(cudaMemcpyAsync(net->in_stage[b].p, patches_hwc + first * patch_floats, m * patch_floats * sizeof(float),
                                cudaMemcpyHostToDevice, net->copy_stream));
        PE_CUDA(cudaEventRecord(net->ev_in[b], net->copy_stream));
        PE_CUDA(cudaStreamWaitEvent(net->stream, net->ev_in[b], 0));
        float* d_desc = net->out_stage.as<float>() + (size_t)first * D;       /* every stage has its own rows: no reuse hazard */
        PE_CHECK(net_forward_dev(net, net->in_stage[b].as<float>(), m, patch_channels, d_desc, D));
        PE_CUDA(cudaEventRecord(net->ev_done[b], net->stream));
        PE_CHECK(match_stage(net, db, comm, d_desc, m, k, db->m_ri.as<int32_t>() + (size_t)first * k, db->m_rd.as<float>() + (size_t)first * k));
    }
    PE_CUDA(cudaMemcpyAsync(out_idx, db->m_ri.p, (size_t)n * k * 4, cudaMemcpyDeviceToHost, db->stream));
    PE_CUDA(cudaMemcpyAsync(out_dist, db->m_rd.p, (size_t)n * k * 4, cudaMemcpyDeviceToHost, db->stream));
    if (out_desc) PE_CUDA(cudaMemcpyAsync(out_desc, net->out_stage.p, (size_t)n * D * sizeof(float), cudaMemcpyDeviceToHost, db->stream));
    PE_CUDA(cudaStreamSynchronize(db->stream));
    PE_CUDA(cudaStreamSynchronize(net->stream));
    PE_CHECK(forward_tc_check(net));
    PE_CHECK(comm_check(comm));
    return knn_tc_check(db);
}

int pe_match_patches(pe_net* net, pe_db* db, pe_comm* comm, const float* patches_hwc, int64_t n, int patch_channels,
                     int k, int32_t* out_idx, float* out_dist, float* out_desc) {
    PE_REQUIRE(net && db, "null net / db");
    PE_REQUIRE(!db->hamming, "pe_match_patches needs an L2 DB");
    PE_REQUIRE(n >= 0 && k >= 1 && k <= 16, "bad n / k");
    PE_REQUIRE(patch_channels == net->def.channels, "patches have %d channels, net input has %d", patch_channels, net->def.channels);
    PE_REQUIRE(net->def.desc_dim == db->dim, "descriptor dim %d != DB dim %d", net->def.desc_dim, db->dim);
    PE_REQUIRE(net->device == db->device && (!comm || comm->device == db->device), "net, DB and communicator must share a device");
    PE_CHECK(net_ensure_device(net));
    if (n == 0) return PE_OK;
    PE_REQUIRE(patches_hwc && out_idx && out_dist, "null pointer");
    int st = match_patches_once(net, db, comm, patches_hwc, n, patch_channels, k, out_idx, out_dist, out_desc);
    /* opt-in e4m3 cross terms withdrawn by an out-of-range batch (every rank sees its own patches, so a rank may replay
     * alone only without a communicator; with one the error is reported) */
    if (st == PE_RETRY_FP16 && !comm) st = match_patches_once(net, db, comm, patches_hwc, n, patch_channels, k, out_idx, out_dist, out_desc);
    return st == PE_RETRY_FP16 ? PE_ERR_INVALID : st;
}

int pe_match_patches_dev(pe_net* net, pe_db* db, pe_comm* comm, const float* d_patches_hwc, int64_t n, int patch_channels,
                         int k, int32_t* d_out_idx, float* d_out_dist, float* d_out_desc) {
    PE_REQUIRE(net && db, "null net / db");
    PE_REQUIRE(!db->hamming, "pe_match_patches_dev needs an L2 DB");
    PE_REQUIRE(n >= 0 && k >= 1 && k <= 16, "bad n / k");
    PE_REQUIRE(patch_channels == net->def.channels, "patches have %d channels, net input has %d", patch_channels, net->def.channels);
    PE_REQUIRE(net->def.desc_dim == db->dim, "descriptor dim %d != DB dim %d", net->def.desc_dim, db->dim);
    PE_REQUIRE(net->device == db->device && (!comm || comm->device == db->device), "net, DB and communicator must share a device");
    PE_CHECK(net_ensure_device(net));
    if (n == 0) return PE_OK;
    PE_REQUIRE(d_patches_hwc && d_out_idx && d_out_dist, "null pointer");
    const int D = net->def.desc_dim;
    if (!db->ev_match) PE_CUDA(cudaEventCreateWithFlags(&db->ev_match, cudaEventDisableTiming));
    for (int i = 0; i < 2; ++i)
        if (!db->ev_free[i]) PE_CUDA(cudaEventCreateWithFlags(&db->ev_free[i], cudaEventDisableTiming));
    const int b = (int)(db->match_calls++ & 1);
    const bool two_streams = db->stream != net->stream;
    if (db->m_desc[b].cap < (size_t)n * D * 4) {
        /* growing a buffer frees it: nothing queued may still read the old one */
        PE_CUDA(cudaStreamSynchronize(db->stream));
        PE_CHECK(db->m_desc[b].reserve((size_t)n * D * 4));
    } else if (two_streams) {
        PE_CUDA(cudaStreamWaitEvent(net->stream, db->ev_free[b], 0));   /* the call before last has finished reading buffer b */
    }
    float* d_desc = db->m_desc[b].as<float>();
    PE_CHECK(net_forward_dev(net, d_patches_hwc, n, patch_channels, d_desc, D));
    PE_CHECK(match_stage(net, db, comm, d_desc, n, k, d_out_idx, d_out_dist));
    if (d_out_desc) PE_CUDA(cudaMemcpyAsync(d_out_desc, d_desc, (size_t)n * D * 4, cudaMemcpyDeviceToDevice, db->stream));
    PE_CUDA(cudaEventRecord(db->ev_free[b], db->stream));
    /* per-kernel profiling brackets kernels with events on their own stream: keep the two streams from overlapping
     * while it is on, so that a bracket never includes time spent waiting for the other stream's CTAs to leave the SMs */
    if (two_streams && (net->profile || db->profile)) PE_CUDA(cudaStreamWaitEvent(net->stream, db->ev_free[b], 0));
    return PE_OK;
}

int pe_knn_hamming_sharded_dev(pe_db* db, pe_comm* comm, const uint32_t* d_qcodes, int64_t nq, int k, int32_t* d_out_idx,
                               int32_t* d_out_dist) {
    PE_REQUIRE(db && db->hamming, "need a Hamming DB");
    PE_REQUIRE(nq == 0 || (d_qcodes && d_out_idx && d_out_dist), "null pointer");
    return sharded_knn<int32_t>(db, comm, d_qcodes, nq, k, d_out_idx, d_out_dist, true);
}

/* computeHistogram's body over a query set that is split across the ranks (every rank passes its own nq_local queries
 * with their merged global neighbours): bin counts and the error sum are all-reduced, the per-query errors all-gathered
 * for the median, so every rank returns the histogram of the whole set (reference src/networkevaluator.cpp:269-355). */
int pe_histogram_sharded_dev(pe_db* db, pe_comm* comm, const int32_t* d_knn_idx, const float* d_q_labels, int64_t nq_local,
                             int k, const int* rot_inv, int n_rot, const float* bins, int n_bins, float* out) {
    PE_REQUIRE(db && bins && out && (rot_inv || n_rot == 0), "null pointer");
    PE_REQUIRE(nq_local >= 0 && (nq_local == 0 || (d_knn_idx && d_q_labels)), "null pointer");
    const int world = comm ? comm->world : 1;
    if (world == 1) return pose_histogram_dev(db, d_knn_idx, d_q_labels, nq_local, k, rot_inv, n_rot, bins, n_bins, out);
    PE_REQUIRE(comm->device == db->device, "communicator and DB live on different devices");
    db->last_comm = comm;
    unsigned long long* d_counts = nullptr;
    double* d_sum = nullptr;
    float *d_med = nullptr, *d_err = nullptr;
    /* err: [own nq_local | gathered world x nq_local] */
    PE_CHECK(pose_hist_local(db, d_knn_idx, d_q_labels, nq_local, k, rot_inv, n_rot, bins, n_bins, (world + 1) * (nq_local > 0 ? nq_local : 1),
                             &d_counts, &d_sum, &d_med, &d_err));
    PE_NCCL(g_nccl.AllReduce(d_counts, d_counts, 16, ncclUint64, ncclSum, comm->comm, db->stream));
    PE_NCCL(g_nccl.AllReduce(d_sum, d_sum, 1, ncclFloat64, ncclSum, comm->comm, db->stream));
    float* d_all = d_err + (nq_local > 0 ? nq_local : 1);
    if (nq_local > 0) PE_NCCL(g_nccl.AllGather(d_err, d_all, (size_t)nq_local, ncclFloat32, comm->comm, db->stream));
    return pose_hist_finish(db, d_counts, d_sum, d_med, d_all, world * nq_local, world * nq_local, n_bins, out);
}

/* computeConfusionMatrix's body over a query set split across the ranks: counts all-reduced
 * (reference src/networkevaluator.cpp:199-245) */
int pe_confusion_sharded_dev(pe_db* db, pe_comm* comm, const int32_t* d_knn_idx, const float* d_q_labels, int64_t nq_local,
                             int k, const int* class_to_local, int n_classes, int n_local, int nr_test_poses, float* out) {
    PE_REQUIRE(db && class_to_local && out, "null pointer");
    PE_REQUIRE(nq_local >= 0 && (nq_local == 0 || (d_knn_idx && d_q_labels)), "null pointer");
    PE_REQUIRE(k >= 1 && k <= 32, "k=%d outside 1..32", k);
    unsigned int* d_mat = nullptr;
    PE_CHECK(pose_confusion_local(db, d_knn_idx, d_q_labels, nq_local, k, class_to_local, n_classes, n_local, &d_mat));
    if (comm && comm->world > 1) {
        PE_REQUIRE(comm->device == db->device, "communicator and DB live on different devices");
        PE_NCCL(g_nccl.AllReduce(d_mat, d_mat, (size_t)n_local * n_local + 1, ncclUint32, ncclSum, comm->comm, db->stream));
    }
    return pose_confusion_finish(db, d_mat, n_local, nr_test_poses, out);
}

int pe_knn_hamming_sharded(pe_db* db, pe_comm* comm, const uint32_t* qcodes, int64_t nq, int k, int32_t* out_idx,
                           int32_t* out_dist) {
    PE_REQUIRE(db && db->hamming, "need a Hamming DB");
    PE_REQUIRE(nq >= 0 && k >= 1 && k <= 16, "bad k / nq");
    if (nq == 0) return PE_OK;
    PE_REQUIRE(qcodes && out_idx && out_dist, "null pointer");
    PE_CUDA(cudaSetDevice(db->device));
    PE_CHECK(db->q_in.reserve((size_t)nq * db->words * 4));
    PE_CHECK(db->out_i.reserve((size_t)nq * k * 4));
    PE_CHECK(db->out_d.reserve((size_t)nq * k * 4));
    PE_CUDA(cudaMemcpyAsync(db->q_in.p, qcodes, (size_t)nq * db->words * 4, cudaMemcpyHostToDevice, db->stream));
    PE_CHECK(sharded_knn<int32_t>(db, comm, db->q_in.p, nq, k, db->out_i.as<int32_t>(), db->out_d.as<int32_t>(), true));
    PE_CUDA(cudaMemcpyAsync(out_idx, db->out_i.p, (size_t)nq * k * 4, cudaMemcpyDeviceToHost, db->stream));
    PE_CUDA(cudaMemcpyAsync(out_dist, db->out_d.p, (size_t)nq * k * 4, cudaMemcpyDeviceToHost, db->stream));
    PE_CUDA(cudaStreamSynchronize(db->stream));
    PE_CHECK(comm_check(comm));
    return knn_tc_check(db);
}

}  // extern "C"
