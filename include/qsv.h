/*
 * qsv.h -- C ABI of libqsv_b200.so, a B200 (sm_100a) dense statevector engine.
 *
 * This is the drop-in boundary for ONE hot path of Qiskit 2.6.0.dev0: dense
 * statevector simulation behind Statevector.evolve / expectation_value /
 * probabilities / sample_memory and the StatevectorEstimator / Sampler.  The
 * reference has no FFI for this path (its C API, crates/cext, exposes no
 * simulation function), so each entry point below names the
 * reference function it replaces (paths relative to the reference checkout).
 * The Python side binds these with ctypes (qiskit_b200/_lib.py); the stub a
 * Qiskit maintainer would add is in INTEGRATION.md.
 *
 * Conventions
 *   - every function returns an int status: QSV_OK (0) or a QSV_ERR_* code;
 *     qsv_last_error() returns a thread-local, NUL-terminated message for the
 *     last failure on the calling thread;
 *   - no C++ exception, Python object or torch type crosses this boundary:
 *     plain pointers and sizes only;
 *   - host pointers are caller-owned and only touched during the call; device
 *     memory is owned by the opaque qsv_state and released by qsv_destroy;
 *   - complex128 data is passed as interleaved doubles (re, im);
 *   - qubit q is bit q of the amplitude index (statevector.py:937); a k-qubit
 *     matrix is 2^k x 2^k row-major with qargs[0] as the least significant bit
 *     of its row/column index (crates/circuit/src/gate_matrix.rs:30-50);
 *   - a handle may be used by one thread at a time, from any thread (every
 *     entry point selects the handle's device); different handles may be used
 *     concurrently.
 */
#ifndef QSV_H
#define QSV_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define QSV_OK 0
#define QSV_ERR_INVALID 1     /* bad argument (Python side raises ValueError / QiskitError) */
#define QSV_ERR_CUDA 2        /* CUDA runtime failure */
#define QSV_ERR_NOMEM 3       /* device allocation failed */
#define QSV_ERR_NCCL 4        /* NCCL failure or NCCL not loadable */
#define QSV_ERR_UNSUPPORTED 5 /* valid request outside what the engine implements */
#define QSV_ERR_OVERFLOW 6    /* num_qubits too large (pauli_exp_val.rs:44-48 -> OverflowError) */

#define QSV_MAX_GATE_QUBITS 12 /* widest dense matrix qsv_apply_matrix accepts */

/* gate kinds for qsv_apply_gates */
#define QSV_GATE_MATRIX 0   /* data: 2^k x 2^k row-major complex128 */
#define QSV_GATE_DIAGONAL 1 /* data: 2^k complex128 diagonal entries */
#define QSV_GATE_PHASE 2    /* global phase exp(i*params[0]); k = 0 (statevector.py:1026-1028) */

/* flags for qsv_apply_gates */
#define QSV_FLAG_NO_FUSION 1 /* one pass over the state per gate (roofline measurement mode) */

typedef struct qsv_state qsv_state;

typedef struct qsv_gate {
    int32_t kind;                        /* QSV_GATE_* */
    int32_t k;                           /* number of qubits */
    int32_t qargs[QSV_MAX_GATE_QUBITS];  /* qargs[0] = LSB of the matrix index */
    const double* data;                  /* host pointer, see kind */
    double params[4];
} qsv_gate;

/* cumulative per-handle counters (qsv_get_counters) */
typedef struct qsv_counters {
    uint64_t kernel_launches; /* kernels this library launched on the handle's stream */
    uint64_t state_passes;    /* full read+write sweeps over the amplitude array */
    uint64_t read_passes;     /* read-only sweeps (expval, probabilities, scan) */
    uint64_t gates_applied;   /* gates received through the API (before fusion) */
    uint64_t h2d_bytes;
    uint64_t d2h_bytes;
    uint64_t phases;          /* register-cluster phases executed inside fused passes */
    uint64_t swap_count;      /* sharded states: global<->local qubit exchanges */
    uint64_t swap_bytes;      /* bytes this rank sent over NVLink in those exchanges */
    uint64_t write_passes;    /* write-only sweeps (basis / product-state initialisation) */
    uint64_t swap_ns;         /* device time spent in those exchanges (CUDA events, nanoseconds) */
    uint64_t fused_swaps;     /* exchanges that rode inside the following fused gate pass (no standalone kernel) */
} qsv_counters;

/* ---- library ---------------------------------------------------------- */
int qsv_version(void);
const char* qsv_last_error(void);
int qsv_device_count(int* out);
/* free / total device memory in bytes */
int qsv_device_memory(int device, uint64_t* free_bytes, uint64_t* total_bytes);

/* ---- state lifetime: type Statevector (statevector.py:44-113), _data (:88) */
int qsv_create(qsv_state** out, int n_qubits, int device);
int qsv_destroy(qsv_state* s);
/* deep copy on the same device: QuantumState.copy (quantum_state.py:74-76) */
int qsv_clone(const qsv_state* s, qsv_state** out);
/* dst <- src (device-to-device); both handles must have the same size and device */
int qsv_copy_state(qsv_state* dst, const qsv_state* src);
int qsv_num_qubits(const qsv_state* s, int* out);
/* Statevector.from_int (statevector.py:790-816): |index> */
int qsv_init_basis(qsv_state* s, uint64_t index);
/* host <-> device copies of amplitudes [offset, offset+count) */
int qsv_upload(qsv_state* s, const double* c128, uint64_t offset, uint64_t count);
int qsv_download(const qsv_state* s, double* c128, uint64_t offset, uint64_t count);
int qsv_sync(const qsv_state* s);

/* ---- gate application: Statevector._evolve_operator (statevector.py:925-963) */
int qsv_apply_matrix(qsv_state* s, const double* mat_rowmajor_c128, int k, const int* qargs);
int qsv_apply_diagonal(qsv_state* s, const double* diag_c128, int k, const int* qargs);
/* a whole unrolled instruction stream, i.e. the leaf calls of
 * Statevector._evolve_instruction (statevector.py:965-1041); enables fusion */
int qsv_apply_gates(qsv_state* s, const qsv_gate* gates, uint64_t n_gates, int flags);
/* Statevector._multiply (statevector.py:415-435) and the global-phase update */
int qsv_scale(qsv_state* s, double re, double im);

/* ---- expectation values: crates/accelerate/src/pauli_exp_val.rs ---------- */
/* expval_pauli_no_x (:39-65) when x_mask == 0, else expval_pauli_with_x (:70-123).
 * phase = (-i)^(#Y); x_max is accepted for signature parity and ignored (any set
 * bit of x_mask splits the pairs). */
int qsv_expval_pauli(const qsv_state* s, uint64_t z_mask, uint64_t x_mask, double phase_re,
                     double phase_im, int x_max, double* out);
/* all terms of a SparsePauliOp (statevector.py:579-583): out[t] = <P_t>, with
 * phases[t] = (-i)^(#Y_t) as interleaved complex128 */
int qsv_expval_pauli_batch(const qsv_state* s, uint64_t n_terms, const uint64_t* z_masks,
                           const uint64_t* x_masks, const double* phases_c128, double* out);
/* <s|s> and <a|b> = sum conj(a_i) b_i (Statevector.inner, statevector.py:399-413) */
int qsv_norm2(const qsv_state* s, double* out);
int qsv_inner(const qsv_state* a, const qsv_state* b, double* out_c128);
/* <a|b> like qsv_inner, for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops,
 * harnesses that forbid allocation).  NO COUNTERPART IN THE REFERENCE.
 * The call allocates nothing (no cudaMalloc / cudaFree / pinned host memory / pooled handle):
 * the per-block partial sums live in one file-scope `static __device__ __align__(256)` array of 16640 bytes
 * (1024 block partials x {re, im} + one 256-byte line for the arrival ticket) that every GPU loading the
 * library holds until the process exits (csrc/qsv_noalloc.cu; its device address is looked up once per
 * device with cudaGetSymbolAddress).
 *   a_c128, b_c128: DEVICE pointers, `count` interleaved complex128 each, 16-byte aligned, on `device`;
 *   out_c128: DEVICE pointer to 2 doubles (re, im); stream: a cudaStream_t of `device` (NULL = default).
 *   Asynchronous: the result is valid once the work queued on `stream` has completed.
 * Limit: count <= 2^30 amplitudes (16 GiB per operand); larger counts return QSV_ERR_INVALID with a
 * message naming the limit -- use qsv_inner there.  Every call shares the one array: calls that overlap on
 * the GPU (different streams, two host threads on one device) race on it and must be serialised by the
 * caller; calls on one stream are safe back to back.  Makes `device` the calling thread's current device. */
int qsv_inner_static_scratch(int device, const double* a_c128, const double* b_c128, uint64_t count,
                             double* out_c128, void* stream);

/* ---- probabilities and sampling ---------------------------------------- */
/* Statevector.probabilities (statevector.py:663-668) + _subsystem_probabilities
 * (quantum_state.py:437-468): host_out[2^m], output bit j = qubit qargs[j];
 * qargs == NULL means all qubits in order.  Values are clipped to [0, 1]. */
int qsv_probabilities(const qsv_state* s, const int* qargs, int m, double* host_out);
/* QuantumState.sample_memory (quantum_state.py:236-265) with the uniforms made
 * explicit: out_indices[j] = inverse-cdf sample for uniforms[j] (the arithmetic of
 * numpy.random.Generator.choice(p=...)), under the summation tree documented in
 * DESIGN.md.  total_out (optional) receives the probability mass sum. */
int qsv_sample(const qsv_state* s, const int* qargs, int m, const double* uniforms,
               uint64_t shots, uint64_t* out_indices, double* total_out);
/* projector used by Statevector.reset (statevector.py:694-719) and
 * QuantumState.measure (quantum_state.py:314-331): keep the branch where the
 * qargs bits equal `outcome`, multiply it by `scale`, zero the rest; with
 * rotate_to_zero the kept branch is moved to the qargs = 0...0 branch. */
int qsv_collapse(qsv_state* s, const int* qargs, int m, uint64_t outcome, double scale,
                 int rotate_to_zero);
/* probabilities(qargs)[outcome], clipped to [0, 1]: the single entry of the marginal that
 * Statevector.reset (statevector.py:705-712) and QuantumState.measure (quantum_state.py:317-324)
 * normalise by -- one read pass, 8 bytes back instead of the whole 2^m array */
int qsv_outcome_probability(const qsv_state* s, const int* qargs, int m, uint64_t outcome, double* out);
/* StatevectorSampler._samples_to_packed_array (statevector_sampler.py:235-251):
 * indices[shots] -> out[shots * num_bytes]; clbit c takes bit src_bit[c] of the
 * index (src_bit[c] < 0: constant 0); big-endian bytes, clbit 0 = LSB of the last
 * byte.  Host-side helper (no device work). */
int qsv_pack_samples(const uint64_t* indices, uint64_t shots, const int* src_bit, int num_bits,
                     uint8_t* out);

/* ---- index-gather ops on whole states (single-GPU handles) ---- */
/* dst[i] = src[j] where bit src_bit[k] of j equals bit k of i (k < n).  Covers
 * Statevector.reverse_qargs (statevector.py:512-531), the layout permutation of
 * Statevector.from_circuit (:115-180) and DensityMatrix.reverse_qargs /
 * partial_transpose (densitymatrix.py:344-364, 839-866).  dst != src. */
int qsv_permute_qubits(qsv_state* dst, const qsv_state* src, const int* src_bit);
/* dst[i] = a[ia] * (conj_b ? conj(b[ib]) : b[ib]), bit j of ia = bit a_pos[j] of i
 * and bit j of ib = bit b_pos[j] of i.  Statevector.tensor / expand
 * (statevector.py:339-397: np.kron), DensityMatrix.tensor / expand
 * (densitymatrix.py:229-265) and the projector |v><v| of DensityMatrix.__init__
 * (densitymatrix.py:104: np.outer(v, conj(v))). */
int qsv_tensor(qsv_state* dst, const qsv_state* a, const qsv_state* b, const int* a_pos,
               const int* b_pos, int conj_b);
/* dst <- dst + (re + i im) * src: Statevector._add / DensityMatrix._add
 * (statevector.py:399-413, densitymatrix.py:267-285) without a host round trip */
int qsv_axpy(qsv_state* dst, const qsv_state* src, double re, double im);
/* element-wise complex conjugate in place: Statevector.conjugate (statevector.py:327-329),
 * DensityMatrix.conjugate (densitymatrix.py:213-215) */
int qsv_conjugate(qsv_state* s);

/* ---- density matrices ------------------------------------ */
/* A DensityMatrix of n qubits is held as the 2n-qubit handle whose amplitude
 * index is (row << n) | col, i.e. the C-order ravel of rho (densitymatrix.py
 * keeps rho as a 2^n x 2^n array).  Evolution rho -> U rho U^dagger
 * (densitymatrix.py:720-750) is qsv_apply_gates with U on qubits q + n and
 * conj(U) on qubits q.  All entry points below need an even-qubit handle. */
/* density_expval_pauli_no_x / density_expval_pauli_with_x
 * (crates/accelerate/src/pauli_exp_val.rs:128-154, :159-194), one value per term:
 * x_mask == 0: sum_i s_i Re rho[i][i];  else 2 * sum_{i0} s_i0 Re(phase * rho[i0 ^ x][i0])
 * over the i0 with bit x_max(x_mask) clear; s = (-1)^popcount(i & z_mask);
 * phases[t] = (-i)^(#Y_t).  The F-order ravel at densitymatrix.py:388 is folded
 * into the index arithmetic. */
int qsv_density_expval_pauli_batch(const qsv_state* s, uint64_t n_terms, const uint64_t* z_masks,
                                   const uint64_t* x_masks, const double* phases_c128, double* out);
/* DensityMatrix.probabilities (densitymatrix.py:498-506): |diag| -> marginal with
 * output bit j = qubit qargs[j] -> clip to [0, 1]; qargs == NULL: all qubits */
int qsv_density_probabilities(const qsv_state* s, const int* qargs, int m, double* host_out);
/* DensityMatrix.trace (:217-219) and purity (:221-227: trace(rho . rho)) */
int qsv_density_trace(const qsv_state* s, double* out_c128);
int qsv_density_purity(const qsv_state* s, double* out_c128);
/* DensityMatrix.reset(qargs) (densitymatrix.py:508-537): the reset SuperOp, i.e.
 * rho' = sum_k (|0><k| on qargs) rho (|k><0| on qargs) */
int qsv_density_reset(qsv_state* s, const int* qargs, int m);

/* ---- counts post-processing (handle-less: device id + host arrays) ---- */
/* bitstring_expval of crates/accelerate/src/sampled_exp_val.rs:46-67 for every term of a
 * diagonal operator over {I, Z, 0, 1}: out[t] = sum_k w[k] * prod_q d_q(bit q of keys[k]) / sum_k w[k]
 * with d = (+1, -1) on z_mask bits, (1, 0) on zero_mask bits, (0, 1) on one_mask bits.
 * keys = outcomes as integers (bit q = clbit / qubit q), weights = their counts or
 * quasi-probabilities (NULL: every key is one shot).  Used by sampled_expectation_value
 * (qiskit/result/sampled_expval.py:37-91) and BitArray.expectation_values
 * (primitives/containers/bit_array.py:613-656). */
int qsv_sampled_expval(int device, const uint64_t* keys, const double* weights, uint64_t n_keys,
                       uint64_t n_terms, const uint64_t* z_masks, const uint64_t* zero_masks,
                       const uint64_t* one_masks, double* out);
/* out[i] bit j = keys[i] bit src_bit[j] (src_bit[j] < 0: 0): marginal_counts / marginal_memory
 * (crates/accelerate/src/results/marginalization.rs) and BitArray.slice_bits on integer outcomes */
int qsv_marginalize_samples(int device, const uint64_t* keys, uint64_t n, const int* src_bit, int m,
                            uint64_t* out);
/* np.unique(samples, return_counts=True) of QuantumState.sample_counts (quantum_state.py:291-293)
 * and BitArray.get_int_counts: out_keys ascending, out_counts their multiplicities; both sized n */
int qsv_sample_counts(int device, const uint64_t* keys, uint64_t n, uint64_t* out_keys, uint64_t* out_counts,
                      uint64_t* n_unique);

/* ---- measurement hooks -------------------------------------------------- */
int qsv_get_counters(const qsv_state* s, qsv_counters* out);
int qsv_reset_counters(qsv_state* s);
/* the same counters summed over every handle this process has had (live, destroyed or reset):
 * what bench.py reports for a public-API call that creates its handles internally */
int qsv_process_counters(qsv_counters* out);
/* put every tunable of qsv_set_option back to the library default (a pooled handle must not carry
 * one user's experiment switches over to the next) */
int qsv_restore_defaults(qsv_state* s);
/* CUDA-event timer on the handle's stream: begin records an event, end records a
 * second one, synchronises and returns the elapsed milliseconds between them. */
int qsv_timer_begin(qsv_state* s);
int qsv_timer_end(qsv_state* s, float* ms_out);
/* engine tunables, by name; for experiments and A/B measurements, the defaults are the product:
 *   tile_bits, low_bits, tile_threads      tile geometry of the unfused tile pass
 *   use_direct, use_cluster, use_fused     which kernel takes a pass (k_direct*, k_cluster_pass, k_fused_pass)
 *   direct_store, warp_local, linear_perm, diag_split   k_fused_pass variants
 *   lazy_init, product_init, product_fold  deferred |index>, product-state prologue, prologue folded
 *                                          into the first fused pass
 *   dist_p2p, dist_fuse_swap, dist_min_lpos  sharded states: peer-memory transport, swap carried by the
 *                                          next fused pass (k_fused_swap_pass), lowest evictable local bit
 * Environment switches read once per process: QSV_TRACE (per-pass plan + time on stderr),
 * QSV_NO_ZYZ, QSV_NO_LIFT (planner rewrites), QSV_NO_PROB_STORE (basis-change expval keeps amplitudes). */
int qsv_set_option(qsv_state* s, const char* name, int64_t value);
/* write a buffer larger than L2 (flush between timed iterations) */
int qsv_flush_l2(qsv_state* s);
/* Host-only view of the single-GPU pass planner (no device work, usable without a GPU): one
 * line per planned pass (kernel kind, tile bits, phases, op counts) is written to `path`. */
int qsv_debug_plan(int n_qubits, const qsv_gate* gates, uint64_t n_gates, int flags, const char* path);

/* Host-only emulation of the remote addressing of the fused swap pass (k_fused_swap_pass).  mode 1
 * (gather): the first fused pass after a global<->local qubit swap reads its tiles from the peers' slabs;
 * modes 2 / 3 (scatter): the last fused pass before the swap stores its tiles into the peers' slabs through
 * its copy-out (2) or direct-store (3) mapping.  For that pass of `gates` on a slab of n_local qubits and
 * the swap (k, lpos[], a = this rank's value on the swapped global bits) every own index x gets src_j[x]
 * (the other rank as a value of the swapped global bits), src_addr[x] (local index there) and tile_of[x]
 * (rank-relative tile number);
 * sig[tile * 16 + i] = ranks tile `tile` signals (0xffffffff: unused); *nin = swapped bits that are
 * tile bits; *ok = 0 when the first pass is not a fused pass.  No device work (CPU tests). */
int qsv_debug_fused_swap(int mode, int n_local, int k, const int* lpos, int a, const qsv_gate* gates,
                         uint64_t n_gates, uint32_t* src_j, uint64_t* src_addr, uint32_t* tile_of, uint32_t* sig,
                         int* nin, int* ok);

/* ---- sharded states (one process per GPU; NCCL over NVLink) -------------- */
/* 128-byte ncclUniqueId for the bootstrap; rank 0 creates it and the host side
 * distributes it (torch.distributed / any channel) */
int qsv_dist_unique_id(void* out128);
/* n_qubits total; this rank holds the slab whose top log2(world) index bits are
 * `rank`; world must be a power of two */
int qsv_create_dist(qsv_state** out, int n_qubits, int device, int rank, int world,
                    const void* unique_id128);
int qsv_dist_info(const qsv_state* s, int* rank, int* world, int* n_local);
/* Host-only view of the sharded gate scheduler (no device work; used by the CPU tests that
 * replay the schedule on NumPy slabs over gloo).  Gate i acts on logical qubits
 * qargs[i*QSV_MAX_GATE_QUBITS ...] (ks[i] of them); need[i] has bit j set when slot j must be
 * local (qsv_dist_need_mask).  perm_in[q] = physical bit of logical qubit q (NULL: identity).
 * The schedule is written as int32 words:  0, count, gate indices...  (apply in this order) or
 * 1, k, (in_qubit, out_qubit) x k  (swap: logical in_qubit becomes local, out_qubit global). */
int qsv_dist_plan(int n_qubits, int world, const int* perm_in, uint64_t n_gates, const int32_t* ks,
                  const int32_t* qargs, const uint32_t* need, int min_lpos, int32_t* out_words,
                  uint64_t cap_words, uint64_t* n_words, int* perm_out);
/* slots of a gate that are not block diagonal (must be local before the gate can be applied) */
int qsv_dist_need_mask(const qsv_gate* gate, uint32_t* out);

#ifdef __cplusplus
}
#endif
#endif /* QSV_H */
