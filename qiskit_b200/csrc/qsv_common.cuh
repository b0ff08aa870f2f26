// Internal definitions shared by the translation units of libqsv_b200.so.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdarg>
#include <cstdio>
#include <complex>
#include <cstring>
#include <string>
#include <vector>

#include <nvtx3/nvToolsExt.h>

#include "../../include/qsv.h"

// NVTX range for the enclosing scope (header-only NVTX v3: a no-op unless a profiler injects itself).
// One range per entry point and per kernel family, so a timeline shows plan / launch / exchange phases.
struct QsvNvtxRange {
    explicit QsvNvtxRange(const char* name) { nvtxRangePushA(name); }
    ~QsvNvtxRange() { nvtxRangePop(); }
};
#define QSV_RANGE(name) QsvNvtxRange qsv_nvtx_range_##__LINE__(name)

// ---------------------------------------------------------------- errors --
void qsv_set_error(const char* fmt, ...);

#define QSV_FAIL(code, ...)         \
    do {                            \
        qsv_set_error(__VA_ARGS__); \
        return (code);              \
    } while (0)

#define QSV_CUDA(expr)                                                                   \
    do {                                                                                 \
        cudaError_t err__ = (expr);                                                      \
        if (err__ != cudaSuccess) {                                                      \
            qsv_set_error("CUDA error %s at %s:%d: %s", cudaGetErrorName(err__),         \
                          __FILE__, __LINE__, cudaGetErrorString(err__));                \
            return err__ == cudaErrorMemoryAllocation ? QSV_ERR_NOMEM : QSV_ERR_CUDA;    \
        }                                                                                \
    } while (0)

#define QSV_TRY(expr)                \
    do {                             \
        int rc__ = (expr);           \
        if (rc__ != QSV_OK) return rc__; \
    } while (0)

// ----------------------------------------------------------------- state --
struct qsv_dist;  // defined in qsv_dist.cu

// A global<->local qubit swap that has been planned (and already applied to `perm`) but not executed:
// the first fused pass of the next qsv_apply_hgates gathers its tiles from the peers' slabs
// (k_fused_swap_pass); any other consumer runs the standalone exchange first (qsv_dist_flush_swap).
struct QsvPendingSwap {
    int mode;                 // 1: gather -- the first fused pass AFTER the swap reads its tiles from the peers
                              // 2: scatter -- the last fused pass BEFORE the swap writes its tiles to the peers
    int k;
    int lpos[4];              // local bit paired with swapped global bit i
    int a;                    // this rank's value on the swapped global bits
    const double2* peer[16];  // slab of the rank whose swapped global bits read j
    uint32_t* flag[16];       // that rank's per-tile arrival counters
    uint32_t* err;            // device error word
};

struct qsv_state {
    int n;        // total qubits
    int n_local;  // qubits held on this device (n - log2(world))
    int device;
    cudaStream_t stream;
    double2* amp;  // 2^n_local amplitudes
    // reusable device scratch (reduction partials, op tables, cdf levels ...)
    void* scratch;
    size_t scratch_bytes;
    // pinned host staging for small results
    void* pinned;
    size_t pinned_bytes;
    // L2 flush buffer
    void* flush_buf;
    size_t flush_bytes;
    // timer
    cudaEvent_t ev0, ev1;
    cudaEvent_t tev0, tev1;  // QSV_TRACE per-pass timing (separate from the qsv_timer_begin / _end pair)
    // tunables
    int tile_bits;  // T: amplitudes per tile = 2^T
    int low_bits;   // L: lowest L index bits are always part of a tile (coalescing)
    int tile_threads;
    int use_direct;  // use register-resident kernels for single-gate passes
    int use_cluster; // fused passes through k_cluster_pass (register phases) instead of k_tile_pass
    int use_fused;   // fused passes through k_fused_pass (persistent, pipelined) when the tile is 2^12
    int use_direct_store;  // k_fused_pass: last phase stores registers straight to global memory
    int use_warp_local;    // k_fused_pass: warp-private sub-tiles when three tile bits stay untouched
    int use_linear_perm;   // fused pass: X/CX/SWAP-like permutations are absorbed by the tile-index -> slot map
    int use_diag_split;    // fused pass: separable diagonal tables -> uniform register factor + one lookup per thread
    int num_sms;
    qsv_counters ctr;
    // sharding
    int rank, world;
    qsv_dist* dist;
    // physical position of each logical qubit (identity unless sharded swaps happened)
    int perm[64];
    // scratch copy of the state used by basis-change expectation values (lazily allocated,
    // shares this handle's stream; owned by this handle)
    qsv_state* aux;
    int owns_stream;
    // lazy |index>: qsv_init_basis only records the index; the buffer is written by the first entry
    // point that needs amplitudes (qsv_materialize) or, better, by the product-state prologue of
    // qsv_apply_gates, which then never has to read the zeros
    // out-of-place request: the next pass reads its tiles from this buffer instead of `amp`
    const double2* read_from;
    // basis-change expectation values: the caller only needs |a|^2 of the result, so the last fused
    // pass of the next qsv_apply_hgates may store probabilities (FpPass.store_prob).  prob_layout
    // reports whether it did; the buffer then holds no amplitudes until it is rewritten.
    int prob_store_req, prob_layout;
    // product-state prologue (qsv_gate.cu): the half-register tables stay here until the first fused
    // pass generates its tiles from them (prod_pending) -- the state is then never written and read
    // back before its first gates; any other consumer calls qsv_flush_product first
    double2* prod_tab;
    size_t prod_tab_bytes;
    int prod_pending, prod_h;
    double prod_vec[64][4];  // per qubit: (re, im) of the amplitudes of |0> and |1>
    int basis_pending;
    QsvPendingSwap* pending_swap;  // sharded states: see QsvPendingSwap (owned by qsv_dist)
    int lazy_init;  // option: 0 = qsv_init_basis writes the buffer immediately
    int use_product_init;  // option: product-state prologue in qsv_apply_gates
    int use_product_fold;  // option: the first fused pass generates its tiles from the product tables
    uint64_t basis_index;
};

int qsv_state_alloc(qsv_state** out, int n_qubits, int n_local, int device, cudaStream_t share_stream);

static inline uint64_t qsv_local_size(const qsv_state* s) { return 1ull << s->n_local; }

// a gate after parsing: matrix on physical bit positions
typedef std::complex<double> cplx;
struct HGate {
    int k;
    int q[QSV_MAX_GATE_QUBITS];  // physical bit of matrix index bit j
    std::vector<cplx> m;         // 2^k x 2^k row-major (dense / mono) or 2^k (diag)
    bool diag;
};
// apply gates whose qubits are all local physical bits (qsv_gate.cu)
int qsv_apply_hgates(qsv_state* s, std::vector<HGate>& gates, int flags);

// sharded states (qsv_dist.cu)
int qsv_dist_flush_swap(qsv_state* s);                    // run a pending swap as a standalone exchange
int qsv_dist_fused_swap_begin(qsv_state* s, uint32_t signals_per_tile, uint32_t* target_out);  // (gather: barrier) + timer start
int qsv_dist_fused_swap_end(qsv_state* s);                // (scatter: barrier), timer stop, counters, handshake error check

int qsv_fused_identity_pass(qsv_state* s);               // qsv_gate.cu: gate-less scatter pass = the exchange itself

int qsv_materialize(qsv_state* s);  // write a pending |index> into the amplitude buffer
int qsv_flush_product(qsv_state* s);  // write a pending product state into the amplitude buffer (qsv_gate.cu)
int qsv_ensure_scratch(qsv_state* s, size_t bytes);
int qsv_ensure_pinned(qsv_state* s, size_t bytes);

// --------------------------------------------------------- device helpers --
__device__ __forceinline__ double2 cmul(double2 a, double2 b) {
    return make_double2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}
__device__ __forceinline__ double2 cfma(double2 a, double2 b, double2 c) {
    // c + a*b
    return make_double2(fma(a.x, b.x, fma(-a.y, b.y, c.x)), fma(a.x, b.y, fma(a.y, b.x, c.y)));
}
__device__ __forceinline__ double2 cadd(double2 a, double2 b) {
    return make_double2(a.x + b.x, a.y + b.y);
}

// insert a zero bit at position p of v
__host__ __device__ __forceinline__ uint64_t insert_zero(uint64_t v, int p) {
    uint64_t low = v & ((1ull << p) - 1);
    return ((v >> p) << (p + 1)) | low;
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
