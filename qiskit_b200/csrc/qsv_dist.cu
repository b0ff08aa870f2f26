// Sharded states: one process per GPU, 2^g ranks, rank r holds the slab whose top g index bits
// equal r.  Gates on local qubits need no communication; gates that are block
// diagonal in a global qubit (diagonal gates, controls) are applied per rank with the block its
// rank bits select; everything else first makes the qubit local with a GLOBAL<->LOCAL QUBIT SWAP.
//
// The swap is an all-to-all between the 2^k ranks that differ in the k swapped global bits.
// Two transports, same semantics:
//   * P2P (default when CUDA IPC peer mapping works): ONE kernel per rank exchanges, in place and
//     without scratch, its sub-slabs with the peers' over NVLink (peer loads + peer stores; each
//     pair of ranks splits a sub-slab in halves so every element pair is moved by exactly one
//     thread).  Cross-rank ordering comes from 1-element NCCL all-reduces on the same stream.
//   * NCCL: pack kernel -> grouped ncclSend/ncclRecv through a bounded scratch ring -> unpack.
// NCCL (dlopen'ed, so a single-GPU process never needs it) also carries the expectation-value
// all-reduce and the sampling all-gather.
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>

#include "qsv_common.cuh"

#define MAX_RANKS 16
#define MAX_G 4

struct NcclApi {
    void* lib;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*);
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int);
    ncclResult_t (*CommDestroy)(ncclComm_t);
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t);
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t);
    ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
    ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
    ncclResult_t (*GroupStart)();
    ncclResult_t (*GroupEnd)();
    const char* (*GetErrorString)(ncclResult_t);
};
static NcclApi g_nccl;

static int nccl_load() {
    if (g_nccl.lib) return QSV_OK;
    const char* names[] = {getenv("QSV_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
    void* lib = nullptr;
    for (const char* nm : names) {
        if (!nm || !*nm) continue;
        lib = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
        if (lib) break;
    }
    if (!lib) QSV_FAIL(QSV_ERR_NCCL, "cannot dlopen libnccl.so.2: %s", dlerror());
#define LOAD(field, sym)                                                        \
    do {                                                                        \
        *(void**)(&g_nccl.field) = dlsym(lib, sym);                             \
        if (!g_nccl.field) QSV_FAIL(QSV_ERR_NCCL, "libnccl lacks symbol %s", sym); \
    } while (0)
    LOAD(GetUniqueId, "ncclGetUniqueId");
    LOAD(CommInitRank, "ncclCommInitRank");
    LOAD(CommDestroy, "ncclCommDestroy");
    LOAD(AllReduce, "ncclAllReduce");
    LOAD(AllGather, "ncclAllGather");
    LOAD(Send, "ncclSend");
    LOAD(Recv, "ncclRecv");
    LOAD(GroupStart, "ncclGroupStart");
    LOAD(GroupEnd, "ncclGroupEnd");
    LOAD(GetErrorString, "ncclGetErrorString");
#undef LOAD
    g_nccl.lib = lib;
    return QSV_OK;
}

#define QSV_NCCL(expr)                                                                          \
    do {                                                                                        \
        ncclResult_t r__ = (expr);                                                              \
        if (r__ != ncclSuccess) {                                                               \
            qsv_set_error("NCCL error at %s:%d: %s", __FILE__, __LINE__, g_nccl.GetErrorString(r__)); \
            return QSV_ERR_NCCL;                                                                \
        }                                                                                       \
    } while (0)

struct qsv_dist {
    ncclComm_t comm;
    int g;                         // log2(world)
    int p2p;                       // peer slabs mapped through CUDA IPC
    int use_p2p;                   // transport choice (option "dist_p2p")
    double2* peer_amp[MAX_RANKS];  // peer_amp[rank] = own slab
    void* ipc_base[MAX_RANKS];     // pointers returned by cudaIpcOpenMemHandle (to close)
    double* d_tok;                 // small device buffer: barrier token, gathers
    void* xbuf;                    // NCCL transport scratch
    size_t xbuf_bytes;
    int min_lpos;                  // lowest local bit position the planner may evict
    // device time of the exchanges: event pairs recorded around every swap, folded into
    // ctr.swap_ns when somebody reads the counters (no synchronisation on the hot path)
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> timers;
    // fused swap pass (k_fused_swap_pass, qsv_fused.cuh): per-tile arrival counters of every rank
    // (own allocation + CUDA-IPC mappings of the peers'), the running value they reach, an error word
    uint32_t* flags;               // own counters: 2^(n_local - 12) words + 1 error word at the end
    size_t flag_words;
    uint32_t* peer_flags[MAX_RANKS];
    void* ipc_flag_base[MAX_RANKS];
    uint32_t flag_target;          // identical on every rank: all ranks run the same fused swap passes
    int use_fuse;                  // option "dist_fuse_swap"
    int use_push;                  // option "dist_push_swap": standalone exchanges as gate-less scatter passes
    QsvPendingSwap pending;        // storage behind qsv_state::pending_swap
    int pending_gpos[MAX_G], pending_k;
    cudaEvent_t fused_e0;
};

void qsv_dist_collect_times(qsv_state* s) {
    qsv_dist* d = s->dist;
    if (!d) return;
    for (auto& pr : d->timers) {
        float ms = 0.f;
        if (cudaEventSynchronize(pr.second) == cudaSuccess && cudaEventElapsedTime(&ms, pr.first, pr.second) == cudaSuccess)
            s->ctr.swap_ns += (uint64_t)((double)ms * 1e6);
        cudaEventDestroy(pr.first);
        cudaEventDestroy(pr.second);
    }
    d->timers.clear();
}

void qsv_dist_free(qsv_state* s) {
    qsv_dist* d = s->dist;
    if (!d) return;
    cudaStreamSynchronize(s->stream);
    qsv_dist_collect_times(s);
    for (int r = 0; r < s->world; ++r) {
        if (d->ipc_base[r]) cudaIpcCloseMemHandle(d->ipc_base[r]);
        if (d->ipc_flag_base[r]) cudaIpcCloseMemHandle(d->ipc_flag_base[r]);
    }
    if (d->flags) cudaFree(d->flags);
    if (d->d_tok) cudaFree(d->d_tok);
    if (d->xbuf) cudaFree(d->xbuf);
    if (d->comm) g_nccl.CommDestroy(d->comm);
    delete d;
    s->dist = nullptr;
}

// stream-ordered barrier across ranks: nobody's later work starts before everybody's earlier
// work on the handle's stream has finished
static int dist_barrier(qsv_state* s) {
    qsv_dist* d = s->dist;
    QSV_NCCL(g_nccl.AllReduce(d->d_tok, d->d_tok, 1, ncclDouble, ncclSum, d->comm, s->stream));
    return QSV_OK;
}

int qsv_dist_allreduce_sum(qsv_state* s, double* dev_buf, size_t count) {
    if (!s->dist) QSV_FAIL(QSV_ERR_INVALID, "state is not sharded");
    QSV_NCCL(g_nccl.AllReduce(dev_buf, dev_buf, count, ncclDouble, ncclSum, s->dist->comm, s->stream));
    return QSV_OK;
}

int qsv_dist_allreduce_sum_u64(qsv_state* s, uint64_t* dev_buf, size_t count) {
    if (!s->dist) QSV_FAIL(QSV_ERR_INVALID, "state is not sharded");
    QSV_NCCL(g_nccl.AllReduce(dev_buf, dev_buf, count, ncclUint64, ncclSum, s->dist->comm, s->stream));
    return QSV_OK;
}

// out[world * count] <- every rank's in[count]
int qsv_dist_allgather(qsv_state* s, const double* dev_in, double* dev_out, size_t count) {
    if (!s->dist) QSV_FAIL(QSV_ERR_INVALID, "state is not sharded");
    QSV_NCCL(g_nccl.AllGather(dev_in, dev_out, count, ncclDouble, s->dist->comm, s->stream));
    return QSV_OK;
}

extern "C" int qsv_dist_unique_id(void* out128) {
    if (!out128) QSV_FAIL(QSV_ERR_INVALID, "qsv_dist_unique_id: out is NULL");
    QSV_TRY(nccl_load());
    ncclUniqueId id;
    QSV_NCCL(g_nccl.GetUniqueId(&id));
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId size");
    memcpy(out128, &id, 128);
    return QSV_OK;
}

extern "C" int qsv_create_dist(qsv_state** out, int n_qubits, int device, int rank, int world,
                               const void* unique_id128) {
    if (!out || !unique_id128) QSV_FAIL(QSV_ERR_INVALID, "qsv_create_dist: NULL argument");
    *out = nullptr;
    int g = 0;
    while ((1 << g) < world) ++g;
    if (world < 2 || (1 << g) != world || world > MAX_RANKS)
        QSV_FAIL(QSV_ERR_INVALID, "qsv_create_dist: world = %d must be a power of two in [2, %d]", world, MAX_RANKS);
    if (rank < 0 || rank >= world) QSV_FAIL(QSV_ERR_INVALID, "qsv_create_dist: rank %d outside [0, %d)", rank, world);
    if (n_qubits - g < g + 1)
        QSV_FAIL(QSV_ERR_INVALID, "qsv_create_dist: %d qubits are too few to shard over %d ranks", n_qubits, world);
    QSV_TRY(nccl_load());
    qsv_state* s = nullptr;
    QSV_TRY(qsv_state_alloc(&s, n_qubits, n_qubits - g, device, nullptr));
    s->rank = rank;
    s->world = world;
    qsv_dist* d = new qsv_dist();  // value-initialised: every field zero, the timer list empty
    d->g = g;
    s->dist = d;
    d->min_lpos = s->n_local >= 16 ? 5 : 0;
    ncclUniqueId id;
    memcpy(&id, unique_id128, 128);
    int rc = QSV_OK;
    do {
        ncclResult_t r = g_nccl.CommInitRank(&d->comm, world, id, rank);
        if (r != ncclSuccess) {
            qsv_set_error("ncclCommInitRank failed: %s", g_nccl.GetErrorString(r));
            rc = QSV_ERR_NCCL;
            break;
        }
        // token + gather area: [0] token, [8 ...] handles / totals
        if (cudaMalloc((void**)&d->d_tok, 4096) != cudaSuccess) {
            qsv_set_error("cudaMalloc of the sharding token buffer failed");
            rc = QSV_ERR_NOMEM;
            break;
        }
        cudaMemsetAsync(d->d_tok, 0, 4096, s->stream);
        // per-tile arrival counters of the fused swap pass (+ one error word), zeroed once: the counters
        // only ever grow, every fused swap pass waits for the next multiple
        d->flag_words = ((size_t)1 << (s->n_local > 12 ? s->n_local - 12 : 0)) + 1;
        if (cudaMalloc((void**)&d->flags, d->flag_words * sizeof(uint32_t)) != cudaSuccess) {
            qsv_set_error("cudaMalloc of the fused-swap counters failed");
            rc = QSV_ERR_NOMEM;
            break;
        }
        cudaMemsetAsync(d->flags, 0, d->flag_words * sizeof(uint32_t), s->stream);
        d->peer_flags[rank] = d->flags;
        // exchange CUDA IPC handles of the slabs and the counters through NCCL itself (no other channel needed)
        cudaIpcMemHandle_t mine[2];
        static_assert(sizeof(cudaIpcMemHandle_t) == 64, "ipc handle size");
        d->peer_amp[rank] = s->amp;
        d->p2p = 0;
        const char* env = getenv("QSV_DIST_P2P");
        const int want_p2p = env && *env ? atoi(env) : 1;
        int ok = want_p2p && cudaIpcGetMemHandle(&mine[0], s->amp) == cudaSuccess &&
                 cudaIpcGetMemHandle(&mine[1], d->flags) == cudaSuccess;
        if (!ok) {
            cudaGetLastError();
            memset(mine, 0, sizeof mine);
        }
        char* d_h = (char*)d->d_tok + 64;            // world * 128 bytes of handles, then this rank's pair
        char* d_mine = d_h + MAX_RANKS * 128;
        cudaMemcpyAsync(d_mine, mine, 128, cudaMemcpyHostToDevice, s->stream);
        r = g_nccl.AllGather(d_mine, d_h, 128, ncclChar, d->comm, s->stream);
        if (r != ncclSuccess) {
            qsv_set_error("ncclAllGather of IPC handles failed: %s", g_nccl.GetErrorString(r));
            rc = QSV_ERR_NCCL;
            break;
        }
        cudaIpcMemHandle_t all[MAX_RANKS][2];
        cudaMemcpyAsync(all, d_h, 128 * world, cudaMemcpyDeviceToHost, s->stream);
        if (cudaStreamSynchronize(s->stream) != cudaSuccess) {
            qsv_set_error("sharded setup: stream failed: %s", cudaGetErrorString(cudaGetLastError()));
            rc = QSV_ERR_CUDA;
            break;
        }
        if (ok) {
            for (int p = 0; p < world && ok; ++p) {
                if (p == rank) continue;
                void* base = nullptr;
                if (cudaIpcOpenMemHandle(&base, all[p][0], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
                    cudaGetLastError();
                    ok = 0;
                    break;
                }
                d->ipc_base[p] = base;
                d->peer_amp[p] = (double2*)base;
                base = nullptr;
                if (cudaIpcOpenMemHandle(&base, all[p][1], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
                    cudaGetLastError();
                    ok = 0;
                    break;
                }
                d->ipc_flag_base[p] = base;
                d->peer_flags[p] = (uint32_t*)base;
            }
        }
        // every rank must take the same transport: min over ranks of `ok`
        double flag = ok ? 1.0 : 0.0, *d_flag = d->d_tok + 1;
        cudaMemcpyAsync(d_flag, &flag, 8, cudaMemcpyHostToDevice, s->stream);
        r = g_nccl.AllReduce(d_flag, d_flag, 1, ncclDouble, ncclMin, d->comm, s->stream);
        if (r != ncclSuccess) {
            qsv_set_error("ncclAllReduce failed: %s", g_nccl.GetErrorString(r));
            rc = QSV_ERR_NCCL;
            break;
        }
        cudaMemcpyAsync(&flag, d_flag, 8, cudaMemcpyDeviceToHost, s->stream);
        cudaStreamSynchronize(s->stream);
        d->p2p = flag > 0.5;
        d->use_p2p = d->p2p;
        // bit 0: gather (first pass after the swap pulls), bit 1: scatter (last pass before it pushes; preferred)
        const char* fenv = getenv("QSV_DIST_FUSE_SWAP");
        d->use_fuse = d->p2p ? (fenv && *fenv ? atoi(fenv) & 3 : 3) : 0;
        const char* penv = getenv("QSV_DIST_PUSH_SWAP");
        d->use_push = penv && *penv ? atoi(penv) != 0 : 1;
    } while (0);
    if (rc != QSV_OK) {
        std::string msg = qsv_last_error();
        qsv_destroy(s);
        qsv_set_error("%s", msg.c_str());
        return rc;
    }
    QSV_TRY(qsv_init_basis(s, 0));
    *out = s;
    return QSV_OK;
}

extern "C" int qsv_dist_info(const qsv_state* s, int* rank, int* world, int* n_local) {
    if (!s) QSV_FAIL(QSV_ERR_INVALID, "qsv_dist_info: state is NULL");
    if (rank) *rank = s->rank;
    if (world) *world = s->world;
    if (n_local) *n_local = s->n_local;
    return QSV_OK;
}

int qsv_dist_set_option(qsv_state* s, const char* name, int64_t value) {
    if (!s->dist) QSV_FAIL(QSV_ERR_INVALID, "option '%s' needs a sharded state", name);
    if (!strcmp(name, "dist_p2p")) {
        if (value && !s->dist->p2p) QSV_FAIL(QSV_ERR_UNSUPPORTED, "peer slabs are not mapped (CUDA IPC unavailable)");
        s->dist->use_p2p = value != 0;
        return QSV_OK;
    }
    if (!strcmp(name, "dist_fuse_swap")) {
        if (value && !s->dist->p2p) QSV_FAIL(QSV_ERR_UNSUPPORTED, "peer slabs are not mapped (CUDA IPC unavailable)");
        s->dist->use_fuse = (int)(value & 3);  // 0 off, 1 gather, 2 scatter, 3 both (scatter where it applies)
        return QSV_OK;
    }
    if (!strcmp(name, "dist_push_swap")) {
        s->dist->use_push = value != 0;
        return QSV_OK;
    }
    if (!strcmp(name, "dist_min_lpos")) {
        if (value < 0 || value > s->n_local - s->dist->g - 1) QSV_FAIL(QSV_ERR_INVALID, "dist_min_lpos out of range");
        s->dist->min_lpos = (int)value;
        return QSV_OK;
    }
    QSV_FAIL(QSV_ERR_INVALID, "qsv_set_option: unknown option '%s'", name);
}

// ------------------------------------------------------------------ the qubit swap ---------
// Swap k global bits with k local bits.  Let a = this rank's value on the swapped global bits and
// j = an amplitude's value on the chosen local bits: amplitude (G = a, L = j) trades places with
// (G = j, L = a).  Rank a therefore exchanges its sub-slab L = j with peer j's sub-slab L = a,
// for every j != a; sub-slab L = a stays.
struct SwapArgs {
    double2* peer[1 << MAX_G];  // slab of the rank whose swapped global bits read j (peer[a] unused)
    int k;
    int lsorted[MAX_G];         // local bit positions, ascending (zero insertion)
    int lpos[MAX_G];            // local bit paired with swapped global bit i
    int a;
    uint64_t rest;              // 2^(n_local - k) amplitudes per sub-slab
};

__device__ __forceinline__ uint64_t swap_expand(uint64_t idx, const SwapArgs& A) {
    for (int i = 0; i < A.k; ++i) idx = insert_zero(idx, A.lsorted[i]);
    return idx;
}

// P2P transport.  grid = (blocks, 2^k - 1); pair (a, j) splits the sub-slab: the lower rank value
// moves the first half, the higher one the second half, so no element is touched twice.
__global__ void __launch_bounds__(256) k_swap_p2p(double2* __restrict__ mine, const __grid_constant__ SwapArgs A) {
    const int j = (int)blockIdx.y < A.a ? (int)blockIdx.y : (int)blockIdx.y + 1;
    double2* __restrict__ peer = A.peer[j];
    uint64_t my_off = 0, peer_off = 0;
    for (int i = 0; i < A.k; ++i) {
        my_off |= (uint64_t)((j >> i) & 1) << A.lpos[i];
        peer_off |= (uint64_t)((A.a >> i) & 1) << A.lpos[i];
    }
    const uint64_t half = A.rest >> 1;
    const uint64_t lo = A.a < j ? 0 : half, hi = lo + half;
    const uint64_t stride = (uint64_t)gridDim.x * 256;
    uint64_t idx = lo + blockIdx.x * 256ull + threadIdx.x;
    // four independent element pairs in flight per thread (peer loads have ~2 us latency)
    for (; idx + 3 * stride < hi; idx += 4 * stride) {
        uint64_t b[4];
        double2 x[4], y[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) b[u] = swap_expand(idx + u * stride, A);
#pragma unroll
        for (int u = 0; u < 4; ++u) y[u] = peer[b[u] | peer_off];
#pragma unroll
        for (int u = 0; u < 4; ++u) x[u] = mine[b[u] | my_off];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            mine[b[u] | my_off] = y[u];
            peer[b[u] | peer_off] = x[u];
        }
    }
    for (; idx < hi; idx += stride) {
        const uint64_t b = swap_expand(idx, A);
        const double2 y = peer[b | peer_off], x = mine[b | my_off];
        mine[b | my_off] = y;
        peer[b | peer_off] = x;
    }
}

// NCCL transport: gather sub-slab elements [c0, c0 + cnt) for every peer into a send buffer ...
__global__ void __launch_bounds__(256) k_swap_pack(const double2* __restrict__ mine, const __grid_constant__ SwapArgs A,
                                                   uint64_t c0, uint64_t cnt, double2* __restrict__ sendbuf) {
    const int jj = blockIdx.y, j = jj < A.a ? jj : jj + 1;
    uint64_t my_off = 0;
    for (int i = 0; i < A.k; ++i) my_off |= (uint64_t)((j >> i) & 1) << A.lpos[i];
    const uint64_t stride = (uint64_t)gridDim.x * 256;
    for (uint64_t i = blockIdx.x * 256ull + threadIdx.x; i < cnt; i += stride)
        sendbuf[(uint64_t)jj * cnt + i] = mine[swap_expand(c0 + i, A) | my_off];
}
// ... and scatter what the peers sent into the same places
__global__ void __launch_bounds__(256) k_swap_unpack(double2* __restrict__ mine, const __grid_constant__ SwapArgs A,
                                                     uint64_t c0, uint64_t cnt, const double2* __restrict__ recvbuf) {
    const int jj = blockIdx.y, j = jj < A.a ? jj : jj + 1;
    uint64_t my_off = 0;
    for (int i = 0; i < A.k; ++i) my_off |= (uint64_t)((j >> i) & 1) << A.lpos[i];
    const uint64_t stride = (uint64_t)gridDim.x * 256;
    for (uint64_t i = blockIdx.x * 256ull + threadIdx.x; i < cnt; i += stride)
        mine[swap_expand(c0 + i, A) | my_off] = recvbuf[(uint64_t)jj * cnt + i];
}

// Exchange physical global bits gpos[i] (>= n_local) with physical local bits lpos[i], i < k.
static int swap_args(qsv_state* s, int k, const int* gpos, const int* lpos, SwapArgs& A, int* peers_rank) {
    qsv_dist* d = s->dist;
    if (k < 1 || k > d->g) QSV_FAIL(QSV_ERR_INVALID, "internal: swap of %d bits", k);
    memset(&A, 0, sizeof A);
    A.k = k;
    A.rest = 1ull << (s->n_local - k);
    for (int i = 0; i < k; ++i) {
        A.lpos[i] = lpos[i];
        A.lsorted[i] = lpos[i];
        A.a |= ((s->rank >> (gpos[i] - s->n_local)) & 1) << i;
    }
    std::sort(A.lsorted, A.lsorted + k);
    for (int j = 0; j < (1 << k); ++j) {
        int r = s->rank;
        for (int i = 0; i < k; ++i) {
            const int bit = gpos[i] - s->n_local;
            r = (r & ~(1 << bit)) | (((j >> i) & 1) << bit);
        }
        peers_rank[j] = r;
        A.peer[j] = d->peer_amp[r];
    }
    return QSV_OK;
}

// the logical qubits that sit on those physical bits trade places
static void swap_commit_perm(qsv_state* s, int k, const int* gpos, const int* lpos) {
    for (int q = 0; q < s->n; ++q) {
        for (int i = 0; i < k; ++i) {
            if (s->perm[q] == gpos[i]) { s->perm[q] = lpos[i]; break; }
            if (s->perm[q] == lpos[i]) { s->perm[q] = gpos[i]; break; }
        }
    }
}

// the standalone exchange (data movement only; perm is the caller's business)
static int swap_execute(qsv_state* s, int k, const int* gpos, const int* lpos) {
    QSV_RANGE("qsv:qubit_swap");
    qsv_dist* d = s->dist;
    SwapArgs A;
    int peers_rank[1 << MAX_G];
    QSV_TRY(swap_args(s, k, gpos, lpos, A, peers_rank));
    const int npeers = (1 << k) - 1;
    static const bool trace = getenv("QSV_TRACE") != nullptr;
    cudaEvent_t te0 = nullptr, te1 = nullptr;
    QSV_CUDA(cudaEventCreate(&te0));
    QSV_CUDA(cudaEventCreate(&te1));
    QSV_CUDA(cudaEventRecord(te0, s->stream));
    if (d->use_p2p) {
        QSV_TRY(dist_barrier(s));
        uint64_t blocks = ((A.rest >> 1) + 256 * 4 - 1) / (256 * 4);
        const uint64_t cap = (uint64_t)s->num_sms * 8 / npeers + 1;
        if (blocks > cap) blocks = cap;
        if (blocks < 1) blocks = 1;
        k_swap_p2p<<<dim3((unsigned)blocks, npeers), 256, 0, s->stream>>>(s->amp, A);
        QSV_CUDA(cudaGetLastError());
        s->ctr.kernel_launches++;
        QSV_TRY(dist_barrier(s));
    } else {
        // scratch: send + recv, npeers * chunk elements each
        if (!d->xbuf) {
            size_t free_b = 0, total_b = 0;
            QSV_CUDA(cudaMemGetInfo(&free_b, &total_b));
            size_t want = std::min<size_t>((size_t)4 << 30, free_b / 2);
            want = std::min<size_t>(want, std::max<size_t>((size_t)32 << s->n_local, (size_t)1 << 20));  // <= two slabs
            want &= ~(size_t)4095;
            if (want < (1u << 16)) QSV_FAIL(QSV_ERR_NOMEM, "no device memory left for the exchange scratch");
            QSV_CUDA(cudaMalloc(&d->xbuf, want));
            d->xbuf_bytes = want;
        }
        uint64_t chunk = d->xbuf_bytes / (2ull * npeers * 16);
        if (chunk > A.rest) chunk = A.rest;
        double2* sendbuf = (double2*)d->xbuf;
        double2* recvbuf = sendbuf + (uint64_t)npeers * chunk;
        for (uint64_t c0 = 0; c0 < A.rest; c0 += chunk) {
            const uint64_t cnt = std::min<uint64_t>(chunk, A.rest - c0);
            uint64_t blocks = (cnt + 255) / 256;
            const uint64_t cap = (uint64_t)s->num_sms * 8 / npeers + 1;
            if (blocks > cap) blocks = cap;
            k_swap_pack<<<dim3((unsigned)blocks, npeers), 256, 0, s->stream>>>(s->amp, A, c0, cnt, sendbuf);
            QSV_CUDA(cudaGetLastError());
            QSV_NCCL(g_nccl.GroupStart());
            for (int jj = 0; jj < npeers; ++jj) {
                const int j = jj < A.a ? jj : jj + 1;
                QSV_NCCL(g_nccl.Send(sendbuf + (uint64_t)jj * cnt, cnt * 2, ncclDouble, peers_rank[j], d->comm, s->stream));
                QSV_NCCL(g_nccl.Recv(recvbuf + (uint64_t)jj * cnt, cnt * 2, ncclDouble, peers_rank[j], d->comm, s->stream));
            }
            QSV_NCCL(g_nccl.GroupEnd());
            k_swap_unpack<<<dim3((unsigned)blocks, npeers), 256, 0, s->stream>>>(s->amp, A, c0, cnt, recvbuf);
            QSV_CUDA(cudaGetLastError());
            s->ctr.kernel_launches += 2;
        }
    }
    QSV_CUDA(cudaEventRecord(te1, s->stream));
    d->timers.emplace_back(te0, te1);
    if (trace) {
        float ms = 0.f;
        QSV_CUDA(cudaEventSynchronize(te1));
        QSV_CUDA(cudaEventElapsedTime(&ms, te0, te1));
        const double gb = (double)npeers * A.rest * 16 / 1e9;
        if (s->rank == 0)
            fprintf(stderr, "[swap] k=%d %s sent %.2f GB/rank in %.3f ms = %.1f GB/s per direction\n", k,
                    d->use_p2p ? "p2p" : "nccl", gb, ms, gb / (ms / 1e3));
    }
    s->ctr.swap_count++;
    s->ctr.swap_bytes += (uint64_t)npeers * A.rest * 16;
    return QSV_OK;
}

static int swap_defer(qsv_state* s, int k, const int* gpos, const int* lpos, int mode);

// The standalone exchange.  Preferred: a gate-less scatter pass (remote writes only, same handshake as
// the fused swap passes); otherwise the pairwise in-place kernel / the NCCL transport.
static int dist_swap_bits(qsv_state* s, int k, const int* gpos, const int* lpos) {
    qsv_dist* d = s->dist;
    if (d->use_p2p && (d->use_fuse & 2) && d->use_push && s->n_local >= 12 && !s->pending_swap) {
        QSV_TRY(swap_defer(s, k, gpos, lpos, 2));
        QSV_TRY(qsv_fused_identity_pass(s));
        if (!s->pending_swap) {  // carried
            s->ctr.fused_swaps--;  // counted as an exchange of its own, not as one that rode inside a gate pass
            swap_commit_perm(s, k, gpos, lpos);
            return QSV_OK;
        }
        s->pending_swap = nullptr;
    }
    QSV_TRY(swap_execute(s, k, gpos, lpos));
    swap_commit_perm(s, k, gpos, lpos);
    return QSV_OK;
}

// ---- fused swap pass plumbing (the kernel and its descriptors live in qsv_fused.cuh / qsv_gate.cu) ----
// Plan the swap, update perm, and leave the data movement to the first fused pass that follows.
static int swap_defer(qsv_state* s, int k, const int* gpos, const int* lpos, int mode) {
    qsv_dist* d = s->dist;
    SwapArgs A;
    int peers_rank[1 << MAX_G];
    QSV_TRY(swap_args(s, k, gpos, lpos, A, peers_rank));
    QsvPendingSwap& W = d->pending;
    memset(&W, 0, sizeof W);
    W.mode = mode;
    W.k = k;
    W.a = A.a;
    for (int i = 0; i < k; ++i) {
        W.lpos[i] = lpos[i];
        d->pending_gpos[i] = gpos[i];
    }
    d->pending_k = k;
    for (int j = 0; j < (1 << k); ++j) {
        W.peer[j] = d->peer_amp[peers_rank[j]];
        W.flag[j] = d->peer_flags[peers_rank[j]];
    }
    W.err = d->flags + (d->flag_words - 1);
    if (mode == 1) swap_commit_perm(s, k, gpos, lpos);  // gather: the run that follows already sees the new layout
    s->pending_swap = &W;
    return QSV_OK;
}

// nobody could carry the pending swap: run it as the standalone exchange (perm is already updated,
// the data is not -- the kernel only needs the bit positions)
int qsv_dist_flush_swap(qsv_state* s) {
    qsv_dist* d = s->dist;
    if (!d || !s->pending_swap) return QSV_OK;
    s->pending_swap = nullptr;
    return swap_execute(s, d->pending_k, d->pending_gpos, d->pending.lpos);
}

int qsv_dist_fused_swap_begin(qsv_state* s, uint32_t signals_per_tile, uint32_t* target_out) {
    qsv_dist* d = s->dist;
    if (!d || !s->pending_swap) QSV_FAIL(QSV_ERR_INVALID, "internal: no pending swap");
    d->flag_target += signals_per_tile;
    *target_out = d->flag_target;
    cudaEvent_t e0 = nullptr;
    QSV_CUDA(cudaEventCreate(&e0));
    QSV_CUDA(cudaEventRecord(e0, s->stream));
    d->fused_e0 = e0;
    // gather: every rank's earlier work on its slab must be complete before anybody reads from it.
    // (scatter needs nothing here: nobody writes into a region before its owner has signalled that it
    // has loaded it, and the owner's own earlier work precedes its loads in stream order.)
    if (d->pending.mode == 1) return dist_barrier(s);
    return QSV_OK;
}

int qsv_dist_fused_swap_end(qsv_state* s) {
    qsv_dist* d = s->dist;
    // scatter: a rank's kernel ends when IT has stored everything, its peers may still be writing into
    // its slab: nobody touches its slab again before everybody's scatter pass has completed
    if (d->pending.mode == 2) QSV_TRY(dist_barrier(s));
    cudaEvent_t e1 = nullptr;
    QSV_CUDA(cudaEventCreate(&e1));
    QSV_CUDA(cudaEventRecord(e1, s->stream));
    d->timers.emplace_back(d->fused_e0, e1);
    const int k = d->pending_k;
    const uint64_t npeers = (1ull << k) - 1;
    s->ctr.swap_count++;
    s->ctr.fused_swaps++;
    s->ctr.swap_bytes += npeers * (16ull << (s->n_local - k));
    // a handshake that timed out leaves wrong amplitudes behind: find out now (one word, one sync per swap)
    QSV_TRY(qsv_ensure_pinned(s, 64));
    uint32_t* h = (uint32_t*)s->pinned + 8;
    QSV_CUDA(cudaMemcpyAsync(h, d->flags + (d->flag_words - 1), sizeof(uint32_t), cudaMemcpyDeviceToHost, s->stream));
    QSV_CUDA(cudaStreamSynchronize(s->stream));
    static const bool trace = getenv("QSV_TRACE") != nullptr;
    if (trace && s->rank == 0) {
        float ms = 0.f;
        cudaEventElapsedTime(&ms, d->fused_e0, e1);
        const double gb = (double)npeers * (16ull << (s->n_local - k)) / 1e9;
        fprintf(stderr, "[fused swap pass] k=%d %s %.2f GB/rank in %.3f ms = %.1f GB/s per direction\n", k,
                d->pending.mode == 1 ? "gathered" : "scattered", gb, ms, gb / (ms / 1e3));
    }
    if (*h) QSV_FAIL(QSV_ERR_NCCL, "fused swap pass: a peer did not arrive in time (handshake timeout)");
    return QSV_OK;
}

// ------------------------------------------------------------------ planner (host only) -----
// A gate needs qubit slot j LOCAL unless its matrix is block diagonal in that slot (the slot's bit
// is the same in the row and column index of every non-zero entry): diagonal gates need nothing,
// a controlled gate needs only its targets.
static uint32_t need_local_mask(const HGate& g) {
    if (g.diag) return 0;
    const int dim = 1 << g.k;
    uint32_t need = 0;
    for (int r = 0; r < dim; ++r)
        for (int c = 0; c < dim; ++c)
            if (g.m[(size_t)r * dim + c] != cplx(0.0, 0.0)) need |= (uint32_t)(r ^ c);
    return need;
}

struct PlanGate {
    int k;
    int q[QSV_MAX_GATE_QUBITS];  // logical qubits
    uint32_t need;               // slots that must be local
};
struct PlanStep {
    int kind;                  // 0 = apply gates, 1 = swap
    std::vector<int> gates;    // kind 0: gate indices, in application order
    int k;                     // kind 1
    int gin[MAX_G], lout[MAX_G];  // kind 1: logical qubits coming in (global now) / going out (local now)
};

// Greedy schedule: run every gate whose needed qubits are local (gates on other qubits may
// overtake a waiting gate); when none is left, ONE swap brings in every global qubit that is
// needed sooner than the local qubit it evicts (Belady: evict the qubits needed farthest ahead).
// A k-bit swap moves (1 - 2^-k) of the slab, so wide swaps are cheap compared with k single ones.
static int dist_plan(int n, int n_local, const int* perm_in, const std::vector<PlanGate>& gates, int min_lpos,
                     std::vector<PlanStep>& steps) {
    const int g = n - n_local;
    int perm[64];
    memcpy(perm, perm_in, sizeof(int) * 64);
    const size_t G = gates.size();
    std::vector<char> done(G, 0);
    size_t first = 0;
    const uint64_t all = n >= 64 ? ~0ull : ((1ull << n) - 1);
    while (true) {
        while (first < G && done[first]) ++first;
        if (first >= G) break;
        PlanStep run;
        run.kind = 0;
        run.k = 0;
        uint64_t blocked = 0;
        for (size_t i = first; i < G; ++i) {
            if (done[i]) continue;
            const PlanGate& pg = gates[i];
            uint64_t qm = 0;
            bool local = true;
            for (int j = 0; j < pg.k; ++j) {
                qm |= 1ull << pg.q[j];
                if (((pg.need >> j) & 1u) && perm[pg.q[j]] >= n_local) local = false;
            }
            if ((qm & blocked) || !local) {
                blocked |= qm;
                if ((blocked & all) == all) break;
                continue;
            }
            run.gates.push_back((int)i);
            done[i] = 1;
        }
        if (!run.gates.empty()) steps.push_back(std::move(run));
        while (first < G && done[first]) ++first;
        if (first >= G) break;
        // next time each logical qubit is needed local
        const size_t INF = (size_t)-1;
        std::vector<size_t> next_need(n, INF);
        for (size_t i = G; i-- > first;) {
            if (done[i]) continue;
            for (int j = 0; j < gates[i].k; ++j)
                if ((gates[i].need >> j) & 1u) next_need[gates[i].q[j]] = i;
        }
        std::vector<int> incoming, victims;
        for (int q = 0; q < n; ++q)
            if (perm[q] >= n_local && next_need[q] != INF) incoming.push_back(q);
        std::sort(incoming.begin(), incoming.end(), [&](int a, int b) { return next_need[a] < next_need[b]; });
        for (int q = 0; q < n; ++q)
            if (perm[q] < n_local && perm[q] >= min_lpos) victims.push_back(q);
        std::sort(victims.begin(), victims.end(), [&](int a, int b) {
            if (next_need[a] != next_need[b]) return next_need[a] > next_need[b];
            return perm[a] > perm[b];  // ties: evict high positions (longer contiguous runs)
        });
        PlanStep sw;
        sw.kind = 1;
        sw.k = 0;
        for (size_t i = 0; i < incoming.size() && i < victims.size() && sw.k < g; ++i) {
            if (i > 0 && next_need[victims[i]] <= next_need[incoming[i]]) break;
            sw.gin[sw.k] = incoming[i];
            sw.lout[sw.k] = victims[i];
            ++sw.k;
        }
        if (sw.k == 0 || next_need[sw.lout[0]] <= next_need[sw.gin[0]]) {
            qsv_set_error("sharded schedule: gate %zu needs more local qubits than a slab holds", first);
            return QSV_ERR_UNSUPPORTED;
        }
        for (int i = 0; i < sw.k; ++i) std::swap(perm[sw.gin[i]], perm[sw.lout[i]]);
        steps.push_back(sw);
    }
    return QSV_OK;
}

// Restrict a gate to this rank: slots on global bits are fixed to the rank's bit values.
static void restrict_to_rank(const HGate& g, const int* phys, int n_local, int rank, HGate& out, cplx& scalar,
                             bool& is_scalar) {
    int keep[QSV_MAX_GATE_QUBITS], nk = 0;
    int fixed = 0;
    for (int j = 0; j < g.k; ++j) {
        if (phys[j] >= n_local) fixed |= ((rank >> (phys[j] - n_local)) & 1) << j;
        else keep[nk++] = j;
    }
    auto expand = [&](int sub) {
        int full = fixed;
        for (int t = 0; t < nk; ++t) full |= ((sub >> t) & 1) << keep[t];
        return full;
    };
    const int dim = 1 << g.k, sdim = 1 << nk;
    is_scalar = nk == 0;
    if (is_scalar) {
        scalar = g.diag ? g.m[fixed] : g.m[(size_t)fixed * dim + fixed];
        return;
    }
    out.k = nk;
    memset(out.q, 0, sizeof out.q);
    for (int t = 0; t < nk; ++t) out.q[t] = phys[keep[t]];
    out.diag = g.diag;
    if (g.diag) {
        out.m.resize(sdim);
        for (int i = 0; i < sdim; ++i) out.m[i] = g.m[expand(i)];
    } else {
        out.m.resize((size_t)sdim * sdim);
        for (int r = 0; r < sdim; ++r)
            for (int c = 0; c < sdim; ++c) out.m[(size_t)r * sdim + c] = g.m[(size_t)expand(r) * dim + expand(c)];
    }
}

static int dist_do_swap(qsv_state* s, const PlanStep& st) {
    int gpos[MAX_G], lpos[MAX_G];
    for (int i = 0; i < st.k; ++i) {
        gpos[i] = s->perm[st.gin[i]];
        lpos[i] = s->perm[st.lout[i]];
    }
    return dist_swap_bits(s, st.k, gpos, lpos);
}

// gates carry LOGICAL qubits in q[]
int qsv_dist_apply(qsv_state* s, std::vector<HGate>& gates, int flags) {
    QSV_RANGE("qsv:dist_plan_and_apply");
    qsv_dist* d = s->dist;
    if (!d) QSV_FAIL(QSV_ERR_INVALID, "state is not sharded");
    std::vector<PlanGate> pg(gates.size());
    for (size_t i = 0; i < gates.size(); ++i) {
        pg[i].k = gates[i].k;
        memcpy(pg[i].q, gates[i].q, sizeof pg[i].q);
        pg[i].need = need_local_mask(gates[i]);
    }
    std::vector<PlanStep> steps;
    QSV_TRY(dist_plan(s->n, s->n_local, s->perm, pg, d->min_lpos, steps));
    cplx scalar(1.0, 0.0);
    const bool may_fuse = d->use_p2p && !(flags & QSV_FLAG_NO_FUSION) && s->n_local >= 12;
    std::vector<char> swap_done(steps.size(), 0);
    for (size_t si = 0; si < steps.size(); ++si) {
        const PlanStep& st = steps[si];
        if (st.kind == 1) {
            if (swap_done[si]) continue;  // rode inside the last pass of the previous run (scatter)
            // Fused swap pass, gather: when every gate of the following run is fully local after the swap,
            // all ranks hold the same gate list, plan the same passes, and the first fused pass can read its
            // tiles from the peers (qsv_gate.cu takes it from there, or falls back to the standalone kernel)
            bool defer = may_fuse && (d->use_fuse & 1) && si + 1 < steps.size() && steps[si + 1].kind == 0;
            if (defer) {
                int perm2[64];
                memcpy(perm2, s->perm, sizeof perm2);
                for (int i = 0; i < st.k; ++i) std::swap(perm2[st.gin[i]], perm2[st.lout[i]]);
                for (int id : steps[si + 1].gates)
                    for (int j = 0; j < gates[id].k && defer; ++j) defer = perm2[gates[id].q[j]] < s->n_local;
            }
            if (defer) {
                int gpos[MAX_G], lpos[MAX_G];
                for (int i = 0; i < st.k; ++i) {
                    gpos[i] = s->perm[st.gin[i]];
                    lpos[i] = s->perm[st.lout[i]];
                }
                QSV_TRY(swap_defer(s, st.k, gpos, lpos, 1));
            } else {
                QSV_TRY(dist_do_swap(s, st));
            }
            continue;
        }
        std::vector<HGate> local;
        local.reserve(st.gates.size());
        bool all_local = true;
        for (int id : st.gates) {
            const HGate& g = gates[id];
            int phys[QSV_MAX_GATE_QUBITS];
            bool any_global = false;
            for (int j = 0; j < g.k; ++j) {
                phys[j] = s->perm[g.q[j]];
                any_global |= phys[j] >= s->n_local;
            }
            if (!any_global) {
                HGate h = g;
                for (int j = 0; j < g.k; ++j) h.q[j] = phys[j];
                local.push_back(std::move(h));
                continue;
            }
            all_local = false;
            HGate h;
            cplx sc;
            bool is_scalar;
            restrict_to_rank(g, phys, s->n_local, s->rank, h, sc, is_scalar);
            if (is_scalar) {
                scalar *= sc;
                continue;
            }
            // the block this rank applies is often the identity (control bit 0 on this rank)
            bool ident = true;
            const int hd = 1 << h.k;
            for (int r = 0; r < hd && ident; ++r) {
                if (h.diag) ident = h.m[r] == cplx(1.0, 0.0);
                else
                    for (int c = 0; c < hd && ident; ++c)
                        ident = h.m[(size_t)r * hd + c] == (r == c ? cplx(1.0, 0.0) : cplx(0.0, 0.0));
            }
            if (!ident) local.push_back(std::move(h));
        }
        // Fused swap pass, scatter (preferred: remote writes are posted, remote reads are not): the swap
        // that FOLLOWS this run rides inside the run's last fused pass, which stores its tiles straight
        // into the slabs that own them after the swap.  Same condition: all ranks plan the same passes.
        const bool scatter = may_fuse && (d->use_fuse & 2) && all_local && !local.empty() && !s->pending_swap &&
                             si + 1 < steps.size() && steps[si + 1].kind == 1;
        if (scatter) {
            const PlanStep& sw = steps[si + 1];
            int gpos[MAX_G], lpos[MAX_G];
            for (int i = 0; i < sw.k; ++i) {
                gpos[i] = s->perm[sw.gin[i]];
                lpos[i] = s->perm[sw.lout[i]];
            }
            QSV_TRY(swap_defer(s, sw.k, gpos, lpos, 2));
        }
        // a rank-dependent scalar must not change what the planner sees while a swap is pending
        // (every rank has to cut the same passes): it waits for the next run or the final qsv_scale
        if (scalar != cplx(1.0, 0.0) && !local.empty() && !s->pending_swap) {
            for (cplx& v : local[0].m) v *= scalar;
            scalar = cplx(1.0, 0.0);
        }
        if (!local.empty()) QSV_TRY(qsv_apply_hgates(s, local, flags));
        else if (s->pending_swap) QSV_TRY(qsv_dist_flush_swap(s));
        if (scatter) {
            const PlanStep& sw = steps[si + 1];
            if (s->pending_swap) {  // the last pass could not carry it: standalone exchange
                s->pending_swap = nullptr;
                QSV_TRY(dist_do_swap(s, sw));
            } else {
                swap_commit_perm(s, d->pending_k, d->pending_gpos, d->pending.lpos);
            }
            swap_done[si + 1] = 1;
        }
    }
    if (s->pending_swap) QSV_TRY(qsv_dist_flush_swap(s));
    if (scalar != cplx(1.0, 0.0)) QSV_TRY(qsv_scale(s, scalar.real(), scalar.imag()));
    return QSV_OK;
}

int qsv_dist_min_lpos(const qsv_state* s) { return s->dist ? s->dist->min_lpos : 0; }

// Bring the listed logical qubits (bit mask) to local positions, evicting qubits outside `keep`.
int qsv_dist_localize(qsv_state* s, uint64_t want_local, uint64_t keep) {
    qsv_dist* d = s->dist;
    PlanStep sw;
    sw.kind = 1;
    sw.k = 0;
    std::vector<int> victims;
    // preferred victims sit at or above min_lpos (long contiguous runs); the low positions are taken
    // only when a term's X / Y factors leave nothing else
    for (int q = s->n - 1; q >= 0; --q)
        if (s->perm[q] < s->n_local && !((keep >> q) & 1ull)) victims.push_back(q);
    std::sort(victims.begin(), victims.end(), [&](int a, int b) {
        const bool ha = s->perm[a] >= d->min_lpos, hb = s->perm[b] >= d->min_lpos;
        if (ha != hb) return ha;
        return s->perm[a] > s->perm[b];
    });
    size_t vi = 0;
    for (int q = 0; q < s->n; ++q) {
        if (!((want_local >> q) & 1ull) || s->perm[q] < s->n_local) continue;
        if (vi >= victims.size())
            QSV_FAIL(QSV_ERR_UNSUPPORTED,
                     "sharded state: a Pauli term has X / Y factors on more qubits than a slab holds (%d local qubits)",
                     s->n_local);
        sw.gin[sw.k] = q;
        sw.lout[sw.k] = victims[vi++];
        ++sw.k;
    }
    if (sw.k == 0) return QSV_OK;
    return dist_do_swap(s, sw);
}

// Restore the identity layout (logical qubit q on physical bit q): global<->local exchanges for
// the global positions, then SWAP gates inside the slab.
int qsv_dist_canonicalize(qsv_state* s) {
    if (!s->dist) return QSV_OK;
    const int n = s->n, nl = s->n_local;
    auto at = [&](int pos) {
        for (int q = 0; q < n; ++q)
            if (s->perm[q] == pos) return q;
        return -1;
    };
    for (int p = nl; p < n; ++p) {
        if (s->perm[p] == p) continue;
        if (s->perm[p] >= nl) {
            // logical p sits on another global bit: park it on a local bit first
            int lp = nl - 1;
            int gp = s->perm[p];
            QSV_TRY(dist_swap_bits(s, 1, &gp, &lp));
        }
        int gp = p, lp = s->perm[p];
        QSV_TRY(dist_swap_bits(s, 1, &gp, &lp));
    }
    // local part: a sequence of physical SWAP gates
    std::vector<HGate> swaps;
    for (int pos = 0; pos < nl; ++pos) {
        const int q = at(pos);
        if (q == pos) continue;
        const int other = s->perm[pos];  // where logical `pos` sits now
        HGate h;
        h.k = 2;
        memset(h.q, 0, sizeof h.q);
        h.q[0] = pos;
        h.q[1] = other;
        h.diag = false;
        h.m.assign(16, cplx(0.0, 0.0));
        h.m[0] = h.m[6] = h.m[9] = h.m[15] = cplx(1.0, 0.0);
        swaps.push_back(h);
        s->perm[q] = other;
        s->perm[pos] = pos;
    }
    if (!swaps.empty()) QSV_TRY(qsv_apply_hgates(s, swaps, 0));
    return QSV_OK;
}

// ---- host-only view of the scheduler for tests (no device work) ---------------------------
// need[i]: bit j set = slot j of gate i must be local.  Steps are written as int32 words:
//   0, count, gate indices...      apply these gates (in this order)
//   1, k, (in_qubit, out_qubit) * k   swap: logical in_qubit becomes local, out_qubit global
extern "C" int qsv_dist_plan(int n_qubits, int world, const int* perm_in, uint64_t n_gates, const int32_t* ks,
                             const int32_t* qargs /* n_gates x QSV_MAX_GATE_QUBITS */, const uint32_t* need,
                             int min_lpos, int32_t* out_words, uint64_t cap_words, uint64_t* n_words,
                             int* perm_out) {
    if (!ks || !qargs || !need || !n_words) QSV_FAIL(QSV_ERR_INVALID, "qsv_dist_plan: NULL argument");
    int g = 0;
    while ((1 << g) < world) ++g;
    if ((1 << g) != world || n_qubits - g < 1) QSV_FAIL(QSV_ERR_INVALID, "qsv_dist_plan: bad world / qubit count");
    int perm[64];
    for (int q = 0; q < 64; ++q) perm[q] = perm_in && q < n_qubits ? perm_in[q] : q;
    std::vector<PlanGate> pg(n_gates);
    for (uint64_t i = 0; i < n_gates; ++i) {
        pg[i].k = ks[i];
        if (ks[i] < 0 || ks[i] > QSV_MAX_GATE_QUBITS) QSV_FAIL(QSV_ERR_INVALID, "qsv_dist_plan: bad k");
        for (int j = 0; j < QSV_MAX_GATE_QUBITS; ++j) pg[i].q[j] = qargs[i * QSV_MAX_GATE_QUBITS + j];
        pg[i].need = need[i];
    }
    std::vector<PlanStep> steps;
    QSV_TRY(dist_plan(n_qubits, n_qubits - g, perm, pg, min_lpos, steps));
    uint64_t w = 0;
    auto put = [&](int32_t v) {
        if (out_words && w < cap_words) out_words[w] = v;
        ++w;
    };
    for (const PlanStep& st : steps) {
        put(st.kind);
        if (st.kind == 0) {
            put((int32_t)st.gates.size());
            for (int id : st.gates) put(id);
        } else {
            put(st.k);
            for (int i = 0; i < st.k; ++i) {
                put(st.gin[i]);
                put(st.lout[i]);
                std::swap(perm[st.gin[i]], perm[st.lout[i]]);
            }
        }
    }
    *n_words = w;
    if (perm_out)
        for (int q = 0; q < n_qubits; ++q) perm_out[q] = perm[q];
    if (out_words && w > cap_words) QSV_FAIL(QSV_ERR_INVALID, "qsv_dist_plan: output buffer too small");
    return QSV_OK;
}

// need-local slot mask of one gate (same rule the engine uses), for the tests above
extern "C" int qsv_dist_need_mask(const qsv_gate* gate, uint32_t* out) {
    if (!gate || !out || !gate->data) QSV_FAIL(QSV_ERR_INVALID, "qsv_dist_need_mask: NULL argument");
    if (gate->kind == QSV_GATE_DIAGONAL || gate->kind == QSV_GATE_PHASE) {
        *out = 0;
        return QSV_OK;
    }
    if (gate->k < 0 || gate->k > QSV_MAX_GATE_QUBITS) QSV_FAIL(QSV_ERR_INVALID, "qsv_dist_need_mask: bad k");
    HGate h;
    h.k = gate->k;
    h.diag = false;
    const size_t dim = (size_t)1 << gate->k;
    const cplx* src = reinterpret_cast<const cplx*>(gate->data);
    h.m.assign(src, src + dim * dim);
    *out = need_local_mask(h);
    return QSV_OK;
}
