// Index-gather ops on whole states and the density-matrix entry points.
//
// A DensityMatrix of n qubits lives in a 2n-qubit handle: amplitude index = (row << n) | col, the
// C-order ravel of rho.  Evolution reuses the gate kernels (U on the row qubits, conj(U) on the
// column qubits); what is new here is the strided read-out of the reference's density kernels
// (crates/accelerate/src/pauli_exp_val.rs:128-194), the diagonal marginals, trace / purity, the
// reset channel and the bit-permutation / Kronecker kernels shared with Statevector.
#include "qsv_common.cuh"

#define DN_THREADS 256

struct BitMap {
    int8_t pos[64];
};

__device__ __forceinline__ double dn_block_sum(double v, double* sh) {
    v = warp_sum(v);
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    __syncthreads();
    if (lane == 0) sh[w] = v;
    __syncthreads();
    double r = 0.0;
    if (threadIdx.x == 0)
        for (int i = 0; i < (int)(blockDim.x >> 5); ++i) r += sh[i];  // fixed order
    return r;
}

// ---- permutation / Kronecker / axpy -----------------------------------------------------
__global__ void __launch_bounds__(DN_THREADS) k_permute_bits(double2* __restrict__ dst, const double2* __restrict__ src,
                                                             uint64_t count, int n, int first_moved, const BitMap M) {
    const uint64_t stride = (uint64_t)gridDim.x * DN_THREADS;
    const uint64_t keep = (1ull << first_moved) - 1;  // low bits that map to themselves
    for (uint64_t i = blockIdx.x * (uint64_t)DN_THREADS + threadIdx.x; i < count; i += stride) {
        uint64_t j = i & keep;
        for (int k = first_moved; k < n; ++k) j |= ((i >> k) & 1ull) << M.pos[k];
        dst[i] = src[j];
    }
}

__global__ void __launch_bounds__(DN_THREADS) k_tensor(double2* __restrict__ dst, const double2* __restrict__ a,
                                                       const double2* __restrict__ b, uint64_t count, int na, int nb,
                                                       const BitMap A, const BitMap B, int conj_b) {
    const uint64_t stride = (uint64_t)gridDim.x * DN_THREADS;
    for (uint64_t i = blockIdx.x * (uint64_t)DN_THREADS + threadIdx.x; i < count; i += stride) {
        uint64_t ia = 0, ib = 0;
        for (int j = 0; j < na; ++j) ia |= ((i >> A.pos[j]) & 1ull) << j;
        for (int j = 0; j < nb; ++j) ib |= ((i >> B.pos[j]) & 1ull) << j;
        const double2 x = a[ia];
        double2 y = b[ib];
        if (conj_b) y.y = -y.y;
        dst[i] = make_double2(x.x * y.x - x.y * y.y, x.x * y.y + x.y * y.x);
    }
}

__global__ void __launch_bounds__(DN_THREADS) k_axpy(double2* __restrict__ dst, const double2* __restrict__ src,
                                                     uint64_t count, double2 c) {
    const uint64_t stride = (uint64_t)gridDim.x * DN_THREADS;
    for (uint64_t i = blockIdx.x * (uint64_t)DN_THREADS + threadIdx.x; i < count; i += stride) {
        const double2 x = src[i], y = dst[i];
        dst[i] = make_double2(fma(c.x, x.x, fma(-c.y, x.y, y.x)), fma(c.x, x.y, fma(c.y, x.x, y.y)));
    }
}

__global__ void __launch_bounds__(DN_THREADS) k_conj(double2* __restrict__ a, uint64_t count) {
    const uint64_t stride = (uint64_t)gridDim.x * DN_THREADS;
    for (uint64_t i = blockIdx.x * (uint64_t)DN_THREADS + threadIdx.x; i < count; i += stride) a[i].y = -a[i].y;
}

// ---- density-matrix read-outs ---------------------------------------------------------------
// one block per Pauli term; the per-term loop is the literal arithmetic of pauli_exp_val.rs:147-153
// (no x) and :181-188 (with x) on rho[row][col] at (row << n) | col
__global__ void __launch_bounds__(DN_THREADS) k_density_expval(const double2* __restrict__ rho, int n,
                                                               const uint64_t* __restrict__ zs,
                                                               const uint64_t* __restrict__ xs,
                                                               const double2* __restrict__ phases,
                                                               double* __restrict__ out) {
    __shared__ double sh[DN_THREADS / 32];
    const uint64_t z = zs[blockIdx.x], x = xs[blockIdx.x];
    const uint64_t rows = 1ull << n;
    double acc = 0.0;
    if (x == 0) {
        for (uint64_t i = threadIdx.x; i < rows; i += DN_THREADS) {
            double val = rho[(i << n) | i].x;
            if (__popcll(i & z) & 1) val = -val;
            acc += val;
        }
    } else {
        const int x_max = 63 - __clzll((long long)x);
        const double2 ph = phases[blockIdx.x];
        for (uint64_t i = threadIdx.x; i < (rows >> 1); i += DN_THREADS) {
            const uint64_t iv = insert_zero(i, x_max);
            const double2 e = rho[((iv ^ x) << n) | iv];
            double val = 2.0 * (ph.x * e.x - ph.y * e.y);
            if (__popcll(iv & z) & 1) val = -val;
            acc += val;
        }
    }
    const double r = dn_block_sum(acc, sh);
    if (threadIdx.x == 0) out[blockIdx.x] = r;
}

// block j sums |rho[i][i]| over the rows i whose qargs bits spell j; QB = positions of the m output
// bits, OB = positions of the n - m summed bits
__global__ void __launch_bounds__(DN_THREADS) k_density_probs(const double2* __restrict__ rho, int n, int m,
                                                              const BitMap QB, const BitMap OB,
                                                              double* __restrict__ out) {
    __shared__ double sh[DN_THREADS / 32];
    uint64_t base = 0;
    for (int j = 0; j < m; ++j) base |= (uint64_t)((blockIdx.x >> j) & 1u) << QB.pos[j];
    const uint64_t others = 1ull << (n - m);
    double acc = 0.0;
    for (uint64_t o = threadIdx.x; o < others; o += DN_THREADS) {
        uint64_t i = base;
        for (int j = 0; j < n - m; ++j) i |= ((o >> j) & 1ull) << OB.pos[j];
        const double2 e = rho[(i << n) | i];
        acc += hypot(e.x, e.y);
    }
    const double r = dn_block_sum(acc, sh);
    if (threadIdx.x == 0) out[blockIdx.x] = fmin(fmax(r, 0.0), 1.0);
}

__global__ void __launch_bounds__(DN_THREADS) k_density_probs_full(const double2* __restrict__ rho, int n,
                                                                   const BitMap QB, double* __restrict__ out) {
    const uint64_t j = blockIdx.x * (uint64_t)DN_THREADS + threadIdx.x;
    if (j >> n) return;
    uint64_t i = 0;
    for (int b = 0; b < n; ++b) i |= ((j >> b) & 1ull) << QB.pos[b];
    const double2 e = rho[(i << n) | i];
    out[j] = fmin(fmax(hypot(e.x, e.y), 0.0), 1.0);
}

// partial sums of the trace (mode 0: sum rho[i][i]) or of trace(rho . rho) (mode 1: sum rho[r][c] rho[c][r])
__global__ void __launch_bounds__(DN_THREADS) k_density_trace(const double2* __restrict__ rho, int n, int mode,
                                                              double* __restrict__ partials) {
    __shared__ double sh[DN_THREADS / 32];
    const uint64_t rows = 1ull << n, count = mode ? rows * rows : rows;
    const uint64_t stride = (uint64_t)gridDim.x * DN_THREADS;
    double re = 0.0, im = 0.0;
    for (uint64_t i = blockIdx.x * (uint64_t)DN_THREADS + threadIdx.x; i < count; i += stride) {
        if (mode) {
            const uint64_t r = i >> n, c = i & (rows - 1);
            const double2 a = rho[i], b = rho[(c << n) | r];
            re += a.x * b.x - a.y * b.y;
            im += a.x * b.y + a.y * b.x;
        } else {
            const double2 a = rho[(i << n) | i];
            re += a.x;
            im += a.y;
        }
    }
    const double r0 = dn_block_sum(re, sh);
    const double r1 = dn_block_sum(im, sh);
    if (threadIdx.x == 0) {
        partials[blockIdx.x] = r0;
        partials[gridDim.x + blockIdx.x] = r1;
    }
}

__global__ void __launch_bounds__(DN_THREADS) k_dn_reduce(const double* __restrict__ partials, int nblocks,
                                                          double* __restrict__ out) {
    __shared__ double sh[DN_THREADS / 32];
    const double* row = partials + (size_t)blockIdx.x * nblocks;
    double v = 0.0;
    for (int b = threadIdx.x; b < nblocks; b += DN_THREADS) v += row[b];
    const double r = dn_block_sum(v, sh);
    if (threadIdx.x == 0) out[blockIdx.x] = r;
}

// reset channel on m qubits: one thread per (row, col) pair with the qargs bits clear
__global__ void __launch_bounds__(DN_THREADS) k_density_reset(double2* __restrict__ rho, int n, int m, const BitMap QB,
                                                              const BitMap OB) {
    const int free_bits = 2 * (n - m);
    const uint64_t g = blockIdx.x * (uint64_t)DN_THREADS + threadIdx.x;
    if (g >> free_bits) return;
    uint64_t r0 = 0, c0 = 0;
    for (int j = 0; j < n - m; ++j) {
        c0 |= ((g >> j) & 1ull) << OB.pos[j];
        r0 |= ((g >> (j + n - m)) & 1ull) << OB.pos[j];
    }
    double re = 0.0, im = 0.0;
    for (uint32_t k = 0; k < (1u << m); ++k) {
        uint64_t kb = 0;
        for (int j = 0; j < m; ++j) kb |= (uint64_t)((k >> j) & 1u) << QB.pos[j];
        const double2 e = rho[((r0 | kb) << n) | (c0 | kb)];
        re += e.x;
        im += e.y;
    }
    for (uint32_t kr = 0; kr < (1u << m); ++kr) {
        uint64_t rb = 0;
        for (int j = 0; j < m; ++j) rb |= (uint64_t)((kr >> j) & 1u) << QB.pos[j];
        for (uint32_t kc = 0; kc < (1u << m); ++kc) {
            uint64_t cb = 0;
            for (int j = 0; j < m; ++j) cb |= (uint64_t)((kc >> j) & 1u) << QB.pos[j];
            rho[((r0 | rb) << n) | (c0 | cb)] = (kr | kc) ? make_double2(0.0, 0.0) : make_double2(re, im);
        }
    }
}

// ---- host side ------------------------------------------------------------------------------
static int dn_grid(const qsv_state* s, uint64_t count) {
    const uint64_t want = (count + DN_THREADS - 1) / DN_THREADS;
    const uint64_t cap = (uint64_t)s->num_sms * 8;
    return (int)std::max<uint64_t>(1, std::min(want, cap));
}

static int dn_single(const char* fn, const qsv_state* s) {
    if (!s) QSV_FAIL(QSV_ERR_INVALID, "%s: NULL handle", fn);
    if (s->world > 1) QSV_FAIL(QSV_ERR_UNSUPPORTED, "%s: not available on sharded handles", fn);
    return QSV_OK;
}

static int dn_density(const char* fn, const qsv_state* s) {
    QSV_TRY(dn_single(fn, s));
    if (s->n & 1) QSV_FAIL(QSV_ERR_INVALID, "%s: a density matrix needs an even-qubit handle (got %d)", fn, s->n);
    return QSV_OK;
}

extern "C" int qsv_permute_qubits(qsv_state* dst, const qsv_state* csrc, const int* src_bit) {
    qsv_state* src = const_cast<qsv_state*>(csrc);
    QSV_TRY(dn_single("qsv_permute_qubits", dst));
    QSV_TRY(dn_single("qsv_permute_qubits", src));
    if (!src_bit || dst == src) QSV_FAIL(QSV_ERR_INVALID, "qsv_permute_qubits: NULL table or dst == src");
    if (dst->n != src->n || dst->device != src->device)
        QSV_FAIL(QSV_ERR_INVALID, "qsv_permute_qubits: handles differ in size or device");
    BitMap M;
    memset(&M, 0, sizeof M);
    uint64_t seen = 0;
    int first_moved = dst->n;
    for (int k = 0; k < dst->n; ++k) {
        if (src_bit[k] < 0 || src_bit[k] >= dst->n || ((seen >> src_bit[k]) & 1ull))
            QSV_FAIL(QSV_ERR_INVALID, "qsv_permute_qubits: src_bit is not a permutation of 0..%d", dst->n - 1);
        seen |= 1ull << src_bit[k];
        M.pos[k] = (int8_t)src_bit[k];
        if (src_bit[k] != k && first_moved == dst->n) first_moved = k;
    }
    QSV_CUDA(cudaSetDevice(dst->device));
    QSV_TRY(qsv_materialize(src));
    QSV_CUDA(cudaStreamSynchronize(src->stream));
    dst->basis_pending = 0;
    dst->prod_pending = 0;
    const uint64_t count = qsv_local_size(dst);
    k_permute_bits<<<dn_grid(dst, count), DN_THREADS, 0, dst->stream>>>(dst->amp, src->amp, count, dst->n, first_moved, M);
    QSV_CUDA(cudaGetLastError());
    dst->ctr.kernel_launches++;
    dst->ctr.state_passes++;
    QSV_CUDA(cudaStreamSynchronize(dst->stream));  // src may be recycled by its owner right after
    return QSV_OK;
}

extern "C" int qsv_tensor(qsv_state* dst, const qsv_state* ca, const qsv_state* cb, const int* a_pos,
                          const int* b_pos, int conj_b) {
    qsv_state* a = const_cast<qsv_state*>(ca);
    qsv_state* b = const_cast<qsv_state*>(cb);
    QSV_TRY(dn_single("qsv_tensor", dst));
    QSV_TRY(dn_single("qsv_tensor", a));
    QSV_TRY(dn_single("qsv_tensor", b));
    if (!a_pos || !b_pos || dst == a || dst == b) QSV_FAIL(QSV_ERR_INVALID, "qsv_tensor: NULL table or dst aliases an input");
    if (dst->n != a->n + b->n || dst->device != a->device || dst->device != b->device)
        QSV_FAIL(QSV_ERR_INVALID, "qsv_tensor: dst must hold %d + %d qubits on the inputs' device", a->n, b->n);
    BitMap A, B;
    memset(&A, 0, sizeof A);
    memset(&B, 0, sizeof B);
    uint64_t seen = 0;
    for (int j = 0; j < a->n + b->n; ++j) {
        const int p = j < a->n ? a_pos[j] : b_pos[j - a->n];
        if (p < 0 || p >= dst->n || ((seen >> p) & 1ull))
            QSV_FAIL(QSV_ERR_INVALID, "qsv_tensor: a_pos and b_pos must partition the bits of dst");
        seen |= 1ull << p;
        if (j < a->n) A.pos[j] = (int8_t)p;
        else B.pos[j - a->n] = (int8_t)p;
    }
    QSV_CUDA(cudaSetDevice(dst->device));
    QSV_TRY(qsv_materialize(a));
    QSV_TRY(qsv_materialize(b));
    QSV_CUDA(cudaStreamSynchronize(a->stream));
    QSV_CUDA(cudaStreamSynchronize(b->stream));
    dst->basis_pending = 0;
    dst->prod_pending = 0;
    const uint64_t count = qsv_local_size(dst);
    k_tensor<<<dn_grid(dst, count), DN_THREADS, 0, dst->stream>>>(dst->amp, a->amp, b->amp, count, a->n, b->n, A, B, conj_b);
    QSV_CUDA(cudaGetLastError());
    dst->ctr.kernel_launches++;
    dst->ctr.write_passes++;
    QSV_CUDA(cudaStreamSynchronize(dst->stream));
    return QSV_OK;
}

extern "C" int qsv_axpy(qsv_state* dst, const qsv_state* csrc, double re, double im) {
    qsv_state* src = const_cast<qsv_state*>(csrc);
    QSV_TRY(dn_single("qsv_axpy", dst));
    QSV_TRY(dn_single("qsv_axpy", src));
    if (dst->n != src->n || dst->device != src->device) QSV_FAIL(QSV_ERR_INVALID, "qsv_axpy: handles differ in size or device");
    QSV_CUDA(cudaSetDevice(dst->device));
    QSV_TRY(qsv_materialize(dst));
    QSV_TRY(qsv_materialize(src));
    if (src != dst) QSV_CUDA(cudaStreamSynchronize(src->stream));
    const uint64_t count = qsv_local_size(dst);
    k_axpy<<<dn_grid(dst, count), DN_THREADS, 0, dst->stream>>>(dst->amp, src->amp, count, make_double2(re, im));
    QSV_CUDA(cudaGetLastError());
    dst->ctr.kernel_launches++;
    dst->ctr.state_passes++;
    QSV_CUDA(cudaStreamSynchronize(dst->stream));
    return QSV_OK;
}

extern "C" int qsv_conjugate(qsv_state* s) {
    if (!s) QSV_FAIL(QSV_ERR_INVALID, "qsv_conjugate: NULL handle");
    QSV_CUDA(cudaSetDevice(s->device));
    QSV_TRY(qsv_materialize(s));
    const uint64_t count = qsv_local_size(s);  // element-wise: valid on every rank of a sharded state
    k_conj<<<dn_grid(s, count), DN_THREADS, 0, s->stream>>>(s->amp, count);
    QSV_CUDA(cudaGetLastError());
    s->ctr.kernel_launches++;
    s->ctr.state_passes++;
    return QSV_OK;
}

extern "C" int qsv_density_expval_pauli_batch(const qsv_state* cs, uint64_t n_terms, const uint64_t* z_masks,
                                              const uint64_t* x_masks, const double* phases_c128, double* out) {
    qsv_state* s = const_cast<qsv_state*>(cs);
    QSV_TRY(dn_density("qsv_density_expval_pauli_batch", s));
    if (n_terms == 0) return QSV_OK;
    if (!z_masks || !x_masks || !phases_c128 || !out) QSV_FAIL(QSV_ERR_INVALID, "qsv_density_expval_pauli_batch: NULL argument");
    const int n = s->n / 2;
    for (uint64_t t = 0; t < n_terms; ++t)
        if (n < 64 && ((z_masks[t] | x_masks[t]) >> n))
            QSV_FAIL(QSV_ERR_INVALID, "qsv_density_expval_pauli_batch: term %llu touches a qubit >= %d",
                     (unsigned long long)t, n);
    QSV_CUDA(cudaSetDevice(s->device));
    QSV_TRY(qsv_materialize(s));
    const size_t per = (size_t)n_terms * 8;
    QSV_TRY(qsv_ensure_scratch(s, 5 * per + 1024));
    char* base = (char*)s->scratch;
    uint64_t* d_z = (uint64_t*)base;
    uint64_t* d_x = (uint64_t*)(base + per);
    double2* d_ph = (double2*)(base + 2 * per);
    double* d_out = (double*)(base + 4 * per);
    QSV_CUDA(cudaMemcpyAsync(d_z, z_masks, per, cudaMemcpyHostToDevice, s->stream));
    QSV_CUDA(cudaMemcpyAsync(d_x, x_masks, per, cudaMemcpyHostToDevice, s->stream));
    QSV_CUDA(cudaMemcpyAsync(d_ph, phases_c128, 2 * per, cudaMemcpyHostToDevice, s->stream));
    for (uint64_t t0 = 0; t0 < n_terms; t0 += 1u << 20) {  // grid.x limit is far away; keep launches bounded anyway
        const unsigned cnt = (unsigned)std::min<uint64_t>(n_terms - t0, 1u << 20);
        k_density_expval<<<cnt, DN_THREADS, 0, s->stream>>>(s->amp, n, d_z + t0, d_x + t0, d_ph + t0, d_out + t0);
        s->ctr.kernel_launches++;
    }
    QSV_CUDA(cudaGetLastError());
    QSV_CUDA(cudaMemcpyAsync(out, d_out, per, cudaMemcpyDeviceToHost, s->stream));
    QSV_CUDA(cudaStreamSynchronize(s->stream));
    s->ctr.h2d_bytes += 4 * per;
    s->ctr.d2h_bytes += per;
    return QSV_OK;
}

static int dn_split_bits(const char* fn, int n, const int* qargs, int m, BitMap& QB, BitMap& OB) {
    memset(&QB, 0, sizeof QB);
    memset(&OB, 0, sizeof OB);
    uint64_t seen = 0;
    for (int j = 0; j < m; ++j) {
        const int q = qargs ? qargs[j] : j;
        if (q < 0 || q >= n || ((seen >> q) & 1ull)) QSV_FAIL(QSV_ERR_INVALID, "%s: qargs must be distinct qubits below %d", fn, n);
        seen |= 1ull << q;
        QB.pos[j] = (int8_t)q;
    }
    for (int q = 0, j = 0; q < n; ++q)
        if (!((seen >> q) & 1ull)) OB.pos[j++] = (int8_t)q;
    return QSV_OK;
}

extern "C" int qsv_density_probabilities(const qsv_state* cs, const int* qargs, int m, double* host_out) {
    qsv_state* s = const_cast<qsv_state*>(cs);
    QSV_TRY(dn_density("qsv_density_probabilities", s));
    if (!host_out) QSV_FAIL(QSV_ERR_INVALID, "qsv_density_probabilities: NULL output");
    const int n = s->n / 2;
    if (!qargs) m = n;
    if (m < 0 || m > n) QSV_FAIL(QSV_ERR_INVALID, "qsv_density_probabilities: %d qargs for %d qubits", m, n);
    BitMap QB, OB;
    QSV_TRY(dn_split_bits("qsv_density_probabilities", n, qargs, m, QB, OB));
    QSV_CUDA(cudaSetDevice(s->device));
    QSV_TRY(qsv_materialize(s));
    const size_t bytes = (size_t)8 << m;
    QSV_TRY(qsv_ensure_scratch(s, bytes));
    double* d_out = (double*)s->scratch;
    if (m == n) {
        const unsigned blocks = (unsigned)(((1ull << n) + DN_THREADS - 1) / DN_THREADS);
        k_density_probs_full<<<blocks, DN_THREADS, 0, s->stream>>>(s->amp, n, QB, d_out);
    } else {
        k_density_probs<<<1u << m, DN_THREADS, 0, s->stream>>>(s->amp, n, m, QB, OB, d_out);
    }
    QSV_CUDA(cudaGetLastError());
    QSV_CUDA(cudaMemcpyAsync(host_out, d_out, bytes, cudaMemcpyDeviceToHost, s->stream));
    QSV_CUDA(cudaStreamSynchronize(s->stream));
    s->ctr.kernel_launches++;
    s->ctr.d2h_bytes += bytes;
    return QSV_OK;
}

static int dn_trace_like(const char* fn, const qsv_state* cs, int mode, double* out_c128) {
    qsv_state* s = const_cast<qsv_state*>(cs);
    QSV_TRY(dn_density(fn, s));
    if (!out_c128) QSV_FAIL(QSV_ERR_INVALID, "%s: NULL output", fn);
    const int n = s->n / 2;
    QSV_CUDA(cudaSetDevice(s->device));
    QSV_TRY(qsv_materialize(s));
    const int blocks = dn_grid(s, mode ? (1ull << s->n) : (1ull << n));
    QSV_TRY(qsv_ensure_scratch(s, (size_t)(2 * blocks + 2) * 8));
    double* d_part = (double*)s->scratch;
    double* d_res = d_part + 2 * blocks;
    k_density_trace<<<blocks, DN_THREADS, 0, s->stream>>>(s->amp, n, mode, d_part);
    k_dn_reduce<<<2, DN_THREADS, 0, s->stream>>>(d_part, blocks, d_res);
    QSV_CUDA(cudaGetLastError());
    QSV_CUDA(cudaMemcpyAsync(out_c128, d_res, 16, cudaMemcpyDeviceToHost, s->stream));
    QSV_CUDA(cudaStreamSynchronize(s->stream));
    s->ctr.kernel_launches += 2;
    if (mode) s->ctr.read_passes++;
    return QSV_OK;
}

extern "C" int qsv_density_trace(const qsv_state* s, double* out_c128) { return dn_trace_like("qsv_density_trace", s, 0, out_c128); }
extern "C" int qsv_density_purity(const qsv_state* s, double* out_c128) { return dn_trace_like("qsv_density_purity", s, 1, out_c128); }

extern "C" int qsv_density_reset(qsv_state* s, const int* qargs, int m) {
    QSV_TRY(dn_density("qsv_density_reset", s));
    const int n = s->n / 2;
    if (!qargs || m < 1 || m > n) QSV_FAIL(QSV_ERR_INVALID, "qsv_density_reset: need 1..%d qargs", n);
    if (m > 10) QSV_FAIL(QSV_ERR_UNSUPPORTED, "qsv_density_reset: at most 10 qubits per call");
    BitMap QB, OB;
    QSV_TRY(dn_split_bits("qsv_density_reset", n, qargs, m, QB, OB));
    QSV_CUDA(cudaSetDevice(s->device));
    QSV_TRY(qsv_materialize(s));
    const uint64_t groups = 1ull << (2 * (n - m));
    const unsigned blocks = (unsigned)((groups + DN_THREADS - 1) / DN_THREADS);
    k_density_reset<<<blocks, DN_THREADS, 0, s->stream>>>(s->amp, n, m, QB, OB);
    QSV_CUDA(cudaGetLastError());
    s->ctr.kernel_launches++;
    s->ctr.state_passes++;
    return QSV_OK;
}
