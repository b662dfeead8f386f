// Communication between the ranks of a multi-GPU run (one process per GPU, the prior sample sharded over them):
// the optional NCCL communicator (dlopen) and the NVLink peer-memory all-reduce kernel that also finishes a swarm
// evaluation.  Entry points declared in include/kirstine_b200.h: kst_comm_*.
#include <dlfcn.h>
#include <stddef.h>
#include <math.h>
#include <string.h>

#include <mutex>

#include "kst_common.cuh"
#include "kst_pso_dev.cuh"

// ---------------------------------------------------------------------------------------
// NCCL through dlopen: no link-time dependency, single-GPU users never load it
// ---------------------------------------------------------------------------------------
namespace {
typedef struct { char internal[128]; } kst_nccl_id;
typedef int (*fn_get_unique_id)(kst_nccl_id *);
typedef int (*fn_comm_init_rank)(void **, int, kst_nccl_id, int);
typedef int (*fn_comm_destroy)(void *);
typedef int (*fn_all_reduce)(const void *, void *, size_t, int, int, void *, cudaStream_t);
typedef const char *(*fn_get_error_string)(int);
struct NcclApi {
    void *lib = nullptr;
    fn_get_unique_id get_unique_id = nullptr;
    fn_comm_init_rank comm_init_rank = nullptr;
    fn_comm_destroy comm_destroy = nullptr;
    fn_all_reduce all_reduce = nullptr;
    fn_get_error_string get_error_string = nullptr;
} g_nccl;
const int KST_NCCL_FLOAT64 = 8, KST_NCCL_SUM = 0;  // ncclDataType_t / ncclRedOp_t values (stable ABI)

std::mutex g_nccl_mutex;

int load_nccl()
{
    std::lock_guard<std::mutex> lock(g_nccl_mutex);  // handles on different threads may attach communicators at once
    if (g_nccl.lib) return KST_OK;
    const char *names[] = {"libnccl.so.2", "libnccl.so"};
    void *lib = nullptr;
    for (const char *n : names) {
        lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (lib) break;
    }
    if (!lib) {
        kst_set_error("NCCL not found: %s", dlerror());
        return KST_ENCCL;
    }
    g_nccl.get_unique_id = (fn_get_unique_id)dlsym(lib, "ncclGetUniqueId");
    g_nccl.comm_init_rank = (fn_comm_init_rank)dlsym(lib, "ncclCommInitRank");
    g_nccl.comm_destroy = (fn_comm_destroy)dlsym(lib, "ncclCommDestroy");
    g_nccl.all_reduce = (fn_all_reduce)dlsym(lib, "ncclAllReduce");
    g_nccl.get_error_string = (fn_get_error_string)dlsym(lib, "ncclGetErrorString");
    if (!g_nccl.get_unique_id || !g_nccl.comm_init_rank || !g_nccl.comm_destroy || !g_nccl.all_reduce) {
        kst_set_error("NCCL symbols missing");
        return KST_ENCCL;
    }
    g_nccl.lib = lib;
    return KST_OK;
}
int nccl_fail(const char *what, int rc)
{
    kst_set_error("%s failed: %s", what, g_nccl.get_error_string ? g_nccl.get_error_string(rc) : "?");
    return KST_ENCCL;
}
}  // namespace

extern "C" int kst_comm_unique_id(void *id128)
{
    int rc = load_nccl();
    if (rc) return rc;
    kst_nccl_id id;
    rc = g_nccl.get_unique_id(&id);
    if (rc) return nccl_fail("ncclGetUniqueId", rc);
    memcpy(id128, &id, sizeof(id));
    return KST_OK;
}

extern "C" int kst_comm_init(kst_problem *h, int nranks, int rank, const void *id128)
{
    if (!h || nranks < 1 || rank < 0 || rank >= nranks) {
        kst_set_error("comm_init: bad arguments");
        return KST_EINVAL;
    }
    if (nranks == 1) return KST_OK;
    int rc = load_nccl();
    if (rc) return rc;
    KST_CUDA(cudaSetDevice(h->device));
    kst_nccl_id id;
    memcpy(&id, id128, sizeof(id));
    rc = g_nccl.comm_init_rank(&h->comm, nranks, id, rank);
    if (rc) return nccl_fail("ncclCommInitRank", rc);
    h->nranks = nranks;
    h->rank = rank;
    return KST_OK;
}

// ---------------------------------------------------------------------------------------
// Small all-reduce over NVLink peer memory, fused into one single-CTA kernel:
//   publish this rank's vector in its exchange block -> release flag -> wait for every peer's flag ->
//   sum all vectors in RANK ORDER (P2P loads), so every rank ends with identical bits.
// Replaces ncclAllReduce for the S partial objectives of a swarm evaluation (8 KB at S = 1024): one launch,
// a few microseconds of NVLink latency, no proxy thread.  (A push version -- every rank writes its vector into its slot
// of every peer's block, one system fence, flags, local polling and local loads -- was measured on two GPUs and is not
// faster: the tail kernel takes 34 ... 43 us either way, 20 us of which are the finish and the best update.)  Two buffers alternate by sequence parity: a rank can
// only reach collective n + 2 after every peer has published n + 1, i.e. finished reading n.
// The flags carry the sequence number, so nothing has to be reset between calls.  A peer that never shows up
// trips a 10 s timeout (error word set, result NaN) instead of hanging the GPU.
// ---------------------------------------------------------------------------------------
#define KST_TAIL_NT 128   // particles per CTA of the multi-CTA tail kernel (k_pso_tail)
struct KstPeerBlock {
    double data[2][KST_PEER_CAP];
    unsigned long long flag[2];                                // single-CTA collective (k_peer_allreduce)
    unsigned long long flag_s[2][KST_PEER_CAP / KST_TAIL_NT];  // one per slice of KST_TAIL_NT elements (k_pso_tail)
    int error;
};
struct KstPeerPtrs {
    KstPeerBlock *blk[KST_PEER_MAX];
};

__device__ __forceinline__ unsigned long long kst_ld_acquire_sys(const unsigned long long *p)
{
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void kst_st_release_sys(unsigned long long *p, unsigned long long v)
{
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ double kst_ld_peer(const double *p)
{
    double v;
    asm volatile("ld.relaxed.sys.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
    return v;
}

// partials != nullptr fuses the last step of the objective evaluation into the collective: entry i of this rank's
// vector is the fixed-order sum of partials[i][0..n_blk) plus `offset` (exactly what k_obj_finish computes), so a
// swarm evaluation on several GPUs is k_objective + this one kernel.
__device__ __forceinline__ void kst_peer_allreduce_cta(double *buf, int count, const KstPeerPtrs &peers, int nranks,
                                                       int rank, unsigned long long seq,
                                                       const double *__restrict__ partials, int n_blk, double offset,
                                                       int nan_fill)
{
    __shared__ int s_fail;
    const int par = (int)(seq & 1ull);
    KstPeerBlock *mine = peers.blk[rank];
    for (int i = threadIdx.x; i < count; i += blockDim.x) {
        double v;
        if (partials) {
            double s = 0.0;
#pragma unroll 8
            for (int b = 0; b < n_blk; b++) s += partials[(size_t)b * count + i];  // eight loads in flight, added in order
            v = nan_fill ? __longlong_as_double(0x7ff8000000000000LL) : s + offset;
        } else {
            v = buf[i];
        }
        mine->data[par][i] = v;
    }
    if (threadIdx.x == 0) s_fail = 0;
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) kst_st_release_sys(&mine->flag[par], seq);
    // one waiting thread per peer
    if ((int)threadIdx.x < nranks && (int)threadIdx.x != rank) {
        const unsigned long long *f = &peers.blk[threadIdx.x]->flag[par];
        unsigned long long t0;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        while (kst_ld_acquire_sys(f) < seq) {
            unsigned long long t1;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
            if (t1 - t0 > 10000000000ull) {  // 10 s
                s_fail = 1;
                break;
            }
        }
    }
    __syncthreads();
    if (s_fail) {
        if (threadIdx.x == 0) mine->error = 1;
        for (int i = threadIdx.x; i < count; i += blockDim.x) buf[i] = __longlong_as_double(0x7ff8000000000000LL);
        __syncthreads();
        return;
    }
    for (int i = threadIdx.x; i < count; i += blockDim.x) {
        double v[KST_PEER_MAX];
#pragma unroll
        for (int r = 0; r < KST_PEER_MAX; r++)  // all loads in flight before the first use
            if (r < nranks) v[r] = kst_ld_peer(&peers.blk[r]->data[par][i]);
        double s = 0.0;
#pragma unroll
        for (int r = 0; r < KST_PEER_MAX; r++)
            if (r < nranks) s += v[r];
        buf[i] = s;
    }
    __syncthreads();
}

__global__ void __launch_bounds__(1024) k_peer_allreduce(double *buf, int count, KstPeerPtrs peers, int nranks, int rank,
                                                         unsigned long long seq, const double *__restrict__ partials,
                                                         int n_blk, double offset)
{
    kst_peer_allreduce_cta(buf, count, peers, nranks, rank, seq, partials, n_blk, offset, 0);
}

// The tail of a swarm iteration on several GPUs in ONE launch (after the evaluation kernel), spread over CTAs of
// KST_TAIL_NT particles like k_pso_best:
//   1. fx = fixed-order sum of the evaluation's per-CTA partial sums (+ offset), summed over the ranks through the
//      peer-memory exchange -- per slice: the CTA publishes its KST_TAIL_NT values in this rank's block, releases the
//      slice's flag, waits for the same slice's flag of every peer and sums the peers' values in rank order (the
//      protocol of kst_peer_allreduce_cta with one flag per slice; the parity argument holds slice by slice, because
//      a rank's launch n + 2 starts after its launch n + 1 has completed, i.e. after every slice has seen every
//      peer's flag of n + 1, i.e. after every peer has finished launch n);
//   2. pso_update_best!: personal bests per particle, the swarm-wide part by the CTA that arrives last
//      (kst_pso_best_grid).
// An iteration is then move + evaluation + this: three launches, as on one GPU.  (One CTA of 1024 threads did all of
// this before: 46 us on 8 GPUs.  Folding the next iteration's move in as well was measured with that version and
// dropped: about 100 us against 40 us for the separate launches.)
__global__ void __launch_bounds__(KST_TAIL_NT) k_pso_tail(KstPsoDev s, KstPeerPtrs peers, int nranks, int rank,
                                                          unsigned long long seq, const double *__restrict__ partials,
                                                          int n_blk, double offset, int nan_fill, double *tr_fx,
                                                          double *tr_pts, double *tr_w, int slot, KstBestScratch sc)
{
    __shared__ int s_fail;
    const int tid = threadIdx.x, c = blockIdx.x, par = (int)(seq & 1ull);
    const int i = c * KST_TAIL_NT + tid;
    const bool live = i < s.S;
    KstPeerBlock *mine = nullptr;
#pragma unroll
    for (int r = 0; r < KST_PEER_MAX; r++)
        if (r == rank) mine = peers.blk[r];
    double fg_old = -KST_DINF;
    if (tid == 0) fg_old = *s.fg;   // requested first, needed last (by the last CTA only)
    const double fpi = live ? kst_pso_best_prefetch(s, i, 0) : 0.0;
    if (live) mine->data[par][i] = kst_obj_finish_value(partials, n_blk, s.S, i, offset, nan_fill);
    if (tid == 0) s_fail = 0;
    __syncthreads();
    // the fence orders the whole CTA's stores (they happen before it through the barrier) ahead of the flag
    if (tid == 0) {
        __threadfence_system();
        kst_st_release_sys(&mine->flag_s[par][c], seq);
    }
    // one waiting thread per peer
#pragma unroll
    for (int r = 0; r < KST_PEER_MAX; r++)
        if (r < nranks && r != rank && tid == r) {
            const unsigned long long *f = &peers.blk[r]->flag_s[par][c];
            unsigned long long t0;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
            while (kst_ld_acquire_sys(f) < seq) {
                unsigned long long t1;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
                if (t1 - t0 > 10000000000ull) {  // 10 s
                    s_fail = 1;
                    break;
                }
            }
        }
    __syncthreads();
    double best = -KST_DINF;
    int bi = -1;
    bool notnan = false;
    if (s_fail && tid == 0) mine->error = 1;
    if (live) {
        double f = __longlong_as_double(0x7ff8000000000000LL);
        if (!s_fail) {
            double v[KST_PEER_MAX];
#pragma unroll
            for (int r = 0; r < KST_PEER_MAX; r++)  // all loads in flight before the first use
                if (r < nranks) v[r] = kst_ld_peer(&peers.blk[r]->data[par][i]);
            f = 0.0;
#pragma unroll
            for (int r = 0; r < KST_PEER_MAX; r++)
                if (r < nranks) f += v[r];
        }
        s.fx[i] = f;
        bool improved;
        kst_pso_best_update(s, i, 0, f, fpi, improved);
        notnan = f == f;
        if (improved) { best = f; bi = i; }
    }
    kst_pso_best_grid(s, best, bi, notnan, fg_old, sc, tr_fx, tr_pts, tr_w, slot, nullptr);
}

int kst_launch_pso_tail(kst_problem *h, int trace_slot, int n_blk, double offset, int nan_fill)
{
    KstPso &p = h->pso;
    if (h->nranks <= 1 || !(h->peer_ready && p.S <= KST_PEER_CAP)) return 1;
    KstPsoDev s;
    kst_pso_make_dev(h, &s);
    KstPeerPtrs pp;
    for (int r = 0; r < KST_PEER_MAX; r++) pp.blk[r] = (KstPeerBlock *)h->peer_map[r];
    const int rcp = kst_prof_begin(h, KST_PROF_BEST);
    if (rcp) return rcp;
    KstBestScratch sc;
    sc.val = p.best_val; sc.idx = p.best_idx; sc.counter = p.best_counter;
    k_pso_tail<<<(p.S + KST_TAIL_NT - 1) / KST_TAIL_NT, KST_TAIL_NT, 0, h->stream>>>(
        s, pp, h->nranks, h->rank, ++h->peer_seq, h->partials, n_blk, offset, nan_fill, trace_slot >= 0 ? p.tr_fx : nullptr,
        trace_slot >= 0 ? p.tr_pts : nullptr, trace_slot >= 0 ? p.tr_w : nullptr, trace_slot >= 0 ? trace_slot : 0, sc);
    KST_CUDA(cudaGetLastError());
    return kst_prof_end(h, KST_PROF_BEST);
}

int kst_allreduce_sum(kst_problem *h, double *d_buf, size_t count)
{
    if (h->nranks <= 1) return KST_OK;
    if (h->peer_ready && count <= KST_PEER_CAP) {
        KstPeerPtrs pp;
        for (int r = 0; r < KST_PEER_MAX; r++) pp.blk[r] = (KstPeerBlock *)h->peer_map[r];
        k_peer_allreduce<<<1, 1024, 0, h->stream>>>(d_buf, (int)count, pp, h->nranks, h->rank, ++h->peer_seq, nullptr, 0,
                                                    0.0);
        KST_CUDA(cudaGetLastError());
        return KST_OK;
    }
    if (!h->comm) { kst_set_error("all-reduce of %zu doubles needs the NCCL communicator (kst_comm_init)", count); return KST_ESTATE; }
    const int rc = g_nccl.all_reduce(d_buf, d_buf, count, KST_NCCL_FLOAT64, KST_NCCL_SUM, h->comm, h->stream);
    if (rc) return nccl_fail("ncclAllReduce", rc);
    return KST_OK;
}

// Swarm evaluation + sum over the ranks.  On the peer-memory path the per-CTA partial sums of k_objective go straight
// into the exchange kernel (no k_obj_finish launch in between).
int kst_objective_allreduce(kst_problem *h, int S, int K, const double *d_pts, const double *d_w, double *d_fx)
{
    const bool fuse = h->nranks > 1 && h->peer_ready && S <= KST_PEER_CAP;
    int n_blk = 0;
    int rc = kst_launch_objective(h, S, K, d_pts, d_w, d_fx, fuse ? &n_blk : nullptr);
    if (rc) return rc;
    if (!fuse) return kst_allreduce_sum(h, d_fx, (size_t)S);
    KstPeerPtrs pp;
    for (int r = 0; r < KST_PEER_MAX; r++) pp.blk[r] = (KstPeerBlock *)h->peer_map[r];
    k_peer_allreduce<<<1, 1024, 0, h->stream>>>(d_fx, S, pp, h->nranks, h->rank, ++h->peer_seq, h->partials, n_blk,
                                                h->offset);
    KST_CUDA(cudaGetLastError());
    return KST_OK;
}

extern "C" int kst_comm_peer_handle(kst_problem *h, int nranks, int rank, void *handle64)
{
    if (!h || !handle64 || nranks < 1 || nranks > KST_PEER_MAX || rank < 0 || rank >= nranks) {
        kst_set_error("comm_peer_handle: need 1 <= nranks <= %d and 0 <= rank < nranks", KST_PEER_MAX);
        return KST_EINVAL;
    }
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    KST_CUDA(cudaSetDevice(h->device));
    // setting a group up again on a live handle: unmap the previous group's blocks first
    if (h->peer_ready) {
        const int rc = kst_comm_peer_close(h);
        if (rc) return rc;
    }
    if (!h->peer_local) KST_CUDA(cudaMalloc(&h->peer_local, sizeof(KstPeerBlock)));
    // a fresh exchange starts at sequence 0 on every rank: stale flags from an earlier group must not satisfy a wait
    KST_CUDA(cudaStreamSynchronize(h->stream));
    KST_CUDA(cudaMemset(h->peer_local, 0, sizeof(KstPeerBlock)));
    h->peer_ready = 0;
    cudaIpcMemHandle_t ipc;
    KST_CUDA(cudaIpcGetMemHandle(&ipc, h->peer_local));
    memcpy(handle64, &ipc, sizeof(ipc));
    h->nranks = nranks;
    h->rank = rank;
    return KST_OK;
}

extern "C" int kst_comm_peer_open(kst_problem *h, const void *handles)
{
    if (!h || !handles || !h->peer_local) { kst_set_error("comm_peer_open before comm_peer_handle"); return KST_ESTATE; }
    KST_CUDA(cudaSetDevice(h->device));
    for (int r = 0; r < KST_PEER_MAX; r++) {  // a second open without a close in between
        if (r != h->rank && h->peer_map[r]) cudaIpcCloseMemHandle(h->peer_map[r]);
        h->peer_map[r] = nullptr;
    }
    for (int r = 0; r < h->nranks; r++) {
        if (r == h->rank) {
            h->peer_map[r] = h->peer_local;
            continue;
        }
        cudaIpcMemHandle_t ipc;
        memcpy(&ipc, (const char *)handles + 64 * r, sizeof(ipc));
        KST_CUDA(cudaIpcOpenMemHandle(&h->peer_map[r], ipc, cudaIpcMemLazyEnablePeerAccess));
    }
    h->peer_ready = 1;
    h->peer_seq = 0;
    return KST_OK;
}

extern "C" int kst_comm_peer_close(kst_problem *h)
{
    if (!h) { kst_set_error("NULL problem handle"); return KST_EINVAL; }
    KST_CUDA(cudaSetDevice(h->device));
    KST_CUDA(cudaStreamSynchronize(h->stream));
    for (int r = 0; r < KST_PEER_MAX; r++) {
        if (r != h->rank && h->peer_map[r]) cudaIpcCloseMemHandle(h->peer_map[r]);
        h->peer_map[r] = nullptr;
    }
    h->peer_ready = 0;
    return KST_OK;
}

extern "C" int kst_comm_peer_error(kst_problem *h, int *error)
{
    if (!h || !error) { kst_set_error("comm_peer_error: bad arguments"); return KST_EINVAL; }
    *error = 0;
    if (!h->peer_local) return KST_OK;
    KST_CUDA(cudaSetDevice(h->device));
    KST_CUDA(cudaStreamSynchronize(h->stream));
    KST_CUDA(cudaMemcpy(error, (const char *)h->peer_local + offsetof(KstPeerBlock, error), sizeof(int), cudaMemcpyDeviceToHost));
    return KST_OK;
}

// called by kst_problem_destroy: communicator, mapped peer blocks, own exchange block
void kst_comm_release(kst_problem *h)
{
    if (h->comm && g_nccl.comm_destroy) g_nccl.comm_destroy(h->comm);
    h->comm = nullptr;
    for (int r = 0; r < KST_PEER_MAX; r++) {
        if (r != h->rank && h->peer_map[r]) cudaIpcCloseMemHandle(h->peer_map[r]);
        h->peer_map[r] = nullptr;
    }
    if (h->peer_local) cudaFree(h->peer_local);
    h->peer_local = nullptr;
    h->peer_ready = 0;
}
