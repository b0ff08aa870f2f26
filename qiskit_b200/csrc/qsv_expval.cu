// Pauli expectation values, norms and inner products.
// Replaces crates/accelerate/src/pauli_exp_val.rs (expval_pauli_no_x :39-65,
// expval_pauli_with_x :70-123) and the per-term Python loop of
// Statevector.expectation_value (statevector.py:579-583).
//
// Every kernel is a streaming read of the amplitude array (16 bytes per
// amplitude) reduced with warp shuffles, then per block through shared memory
// into a [term][block] array of partials; a second small kernel sums the
// partials in a fixed order, so results are run-to-run deterministic.
#include <algorithm>

#include "qsv_common.cuh"

#define EV_THREADS 256
#define EV_MAX_BLOCKS 2048
#define EV_TERMS 8  // terms sharing an x_mask evaluated per launch

struct TermBatch {
    uint64_t x_mask;
    int split;  // bit of x_mask used to enumerate pairs (lowest set bit); -1 when x_mask == 0
    int nterms;
    uint64_t z_mask[EV_TERMS];
    int phase_sel[EV_TERMS];  // (-i)^(#Y): 0 -> 1, 1 -> -i, 2 -> -1, 3 -> i
};

__device__ __forceinline__ double block_sum(double v, double* sh) {
    v = warp_sum(v);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __syncthreads();
    if (lane == 0) sh[warp] = v;
    __syncthreads();
    double r = 0.0;
    if (warp == 0) {
        r = lane < (EV_THREADS / 32) ? sh[lane] : 0.0;
        r = warp_sum(r);
    }
    return r;  // valid in warp 0
}

// Z-type terms (x_mask == 0):  sum_i (-1)^popc(i & z) |a_i|^2        (pauli_exp_val.rs:52-58)
// X-type terms:  sum over pairs (i0, i1 = i0 ^ x) of
//   s(i0) Re(ph a[i1] conj(a[i0])) + s(i1) Re(ph a[i0] conj(a[i1]))    (pauli_exp_val.rs:88-119)
// with c = a[i1] conj(a[i0]):  ph=1: Re c, ph=-i: Im c, ph=-1: -Re c, ph=i: -Im c; the second
// product is conj(c), so its imaginary part flips sign.
__global__ void __launch_bounds__(EV_THREADS) k_expval_batch(const double2* __restrict__ amp,
                                                             uint64_t count, uint64_t index_base,
                                                             const TermBatch B,
                                                             double* __restrict__ partials,
                                                             int partial_stride) {
    __shared__ double sh[EV_THREADS / 32];
    double acc[EV_TERMS];
#pragma unroll
    for (int t = 0; t < EV_TERMS; ++t) acc[t] = 0.0;
    const uint64_t stride = (uint64_t)gridDim.x * EV_THREADS;
    if (B.x_mask == 0) {
        for (uint64_t i = blockIdx.x * (uint64_t)EV_THREADS + threadIdx.x; i < count; i += stride) {
            double2 a = amp[i];
            double p = a.x * a.x + a.y * a.y;
            uint64_t gi = i | index_base;
#pragma unroll
            for (int t = 0; t < EV_TERMS; ++t)
                if (t < B.nterms) acc[t] += (__popcll(gi & B.z_mask[t]) & 1) ? -p : p;
        }
    } else {
        const uint64_t pairs = count >> 1;
        for (uint64_t i = blockIdx.x * (uint64_t)EV_THREADS + threadIdx.x; i < pairs; i += stride) {
            uint64_t i0 = insert_zero(i, B.split);
            uint64_t i1 = i0 ^ B.x_mask;
            double2 a0 = amp[i0], a1 = amp[i1];
            double cre = a1.x * a0.x + a1.y * a0.y;
            double cim = a1.y * a0.x - a1.x * a0.y;
            uint64_t g0 = i0 | index_base, g1 = i1 | index_base;
#pragma unroll
            for (int t = 0; t < EV_TERMS; ++t) {
                if (t < B.nterms) {
                    const int sel = B.phase_sel[t];
                    double v0, v1;
                    if (sel == 0) { v0 = cre; v1 = cre; }
                    else if (sel == 1) { v0 = cim; v1 = -cim; }
                    else if (sel == 2) { v0 = -cre; v1 = -cre; }
                    else { v0 = -cim; v1 = cim; }
                    if (__popcll(g0 & B.z_mask[t]) & 1) v0 = -v0;
                    if (__popcll(g1 & B.z_mask[t]) & 1) v1 = -v1;
                    acc[t] += v0 + v1;
                }
            }
        }
    }
#pragma unroll
    for (int t = 0; t < EV_TERMS; ++t) {
        if (t < B.nterms) {
            double r = block_sum(acc[t], sh);
            if (threadIdx.x == 0) partials[(size_t)t * partial_stride + blockIdx.x] = r;
        }
    }
}

// out[t] = sum_b partials[t][b], one block per term, fixed order
__global__ void __launch_bounds__(EV_THREADS) k_reduce_partials(const double* __restrict__ partials,
                                                                int nblocks, int partial_stride,
                                                                double* __restrict__ out) {
    __shared__ double sh[EV_THREADS / 32];
    const double* row = partials + (size_t)blockIdx.x * partial_stride;
    double v = 0.0;
    for (int b = threadIdx.x; b < nblocks; b += EV_THREADS) v += row[b];
    double r = block_sum(v, sh);
    if (threadIdx.x == 0) out[blockIdx.x] = r;
}

// <a|b> = sum conj(a_i) b_i ; with a == b this is the squared norm
__global__ void __launch_bounds__(EV_THREADS) k_inner(const double2* __restrict__ a,
                                                      const double2* __restrict__ b, uint64_t count,
                                                      double* __restrict__ partials, int partial_stride) {
    __shared__ double sh[EV_THREADS / 32];
    double re = 0.0, im = 0.0;
    const uint64_t stride = (uint64_t)gridDim.x * EV_THREADS;
    for (uint64_t i = blockIdx.x * (uint64_t)EV_THREADS + threadIdx.x; i < count; i += stride) {
        double2 x = a[i], y = b[i];
        re += x.x * y.x + x.y * y.y;
        im += x.x * y.y - x.y * y.x;
    }
    double r0 = block_sum(re, sh);
    double r1 = block_sum(im, sh);
    if (threadIdx.x == 0) {
        partials[blockIdx.x] = r0;
        partials[partial_stride + blockIdx.x] = r1;
    }
}

static int grid_for(const qsv_state* s, uint64_t items) {
    uint64_t blocks = (items + EV_THREADS - 1) / EV_THREADS;
    uint64_t cap = (uint64_t)s->num_sms * 8;
    if (cap > EV_MAX_BLOCKS) cap = EV_MAX_BLOCKS;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (int)blocks;
}

int qsv_dist_allreduce_sum(qsv_state* s, double* dev_buf, size_t count);  // qsv_dist.cu
int qsv_dist_localize_xmask(qsv_state* s, uint64_t phys_x_mask);          // qsv_dist.cu

static uint64_t map_mask(const qsv_state* s, uint64_t logical) {
    uint64_t phys = 0;
    for (int q = 0; q < s->n; ++q)
        if ((logical >> q) & 1ull) phys |= 1ull << s->perm[q];
    return phys;
}

static int phase_selector(double re, double im, int* sel) {
    // (-i)^k for k = 0..3
    const double tol = 1e-12;
    if (fabs(re - 1.0) < tol && fabs(im) < tol) *sel = 0;
    else if (fabs(re) < tol && fabs(im + 1.0) < tol) *sel = 1;
    else if (fabs(re + 1.0) < tol && fabs(im) < tol) *sel = 2;
    else if (fabs(re) < tol && fabs(im - 1.0) < tol) *sel = 3;
    else QSV_FAIL(QSV_ERR_INVALID, "expval: phase (%g, %g) is not a power of -i", re, im);
    return QSV_OK;
}

extern "C" int qsv_expval_pauli_batch(const qsv_state* cs, uint64_t n_terms, const uint64_t* z_masks,
                                      const uint64_t* x_masks, const double* phases, double* out) {
    qsv_state* s = const_cast<qsv_state*>(cs);
    if (!s || (n_terms && (!z_masks || !x_masks || !out)))
        QSV_FAIL(QSV_ERR_INVALID, "qsv_expval_pauli_batch: NULL argument");
    if (n_terms == 0) return QSV_OK;
    QSV_CUDA(cudaSetDevice(s->device));
    const uint64_t full = s->n >= 64 ? ~0ull : ((1ull << s->n) - 1);
    // group terms by x_mask (physical bits), preserving order inside a group
    struct Item { uint64_t x, z; int sel; uint64_t idx; };
    std::vector<Item> items(n_terms);
    for (uint64_t t = 0; t < n_terms; ++t) {
        if ((z_masks[t] | x_masks[t]) & ~full)
            QSV_FAIL(QSV_ERR_INVALID, "qsv_expval_pauli_batch: term %llu has mask bits beyond qubit %d",
                     (unsigned long long)t, s->n - 1);
        int sel = 0;
        if (phases) QSV_TRY(phase_selector(phases[2 * t], phases[2 * t + 1], &sel));
        items[t] = Item{map_mask(s, x_masks[t]), map_mask(s, z_masks[t]), sel, t};
    }
    std::vector<uint64_t> order(n_terms);
    for (uint64_t t = 0; t < n_terms; ++t) order[t] = t;
    std::stable_sort(order.begin(), order.end(),
                     [&](uint64_t a, uint64_t b) { return items[a].x < items[b].x; });
    // batches
    std::vector<TermBatch> batches;
    std::vector<std::vector<uint64_t>> batch_terms;
    for (uint64_t p = 0; p < n_terms;) {
        TermBatch B;
        memset(&B, 0, sizeof B);
        B.x_mask = items[order[p]].x;
        B.split = B.x_mask ? __builtin_ctzll(B.x_mask) : -1;
        std::vector<uint64_t> ids;
        while (p < n_terms && items[order[p]].x == B.x_mask && B.nterms < EV_TERMS) {
            B.z_mask[B.nterms] = items[order[p]].z;
            B.phase_sel[B.nterms] = items[order[p]].sel;
            ids.push_back(items[order[p]].idx);
            ++B.nterms;
            ++p;
        }
        batches.push_back(B);
        batch_terms.push_back(ids);
    }
    const uint64_t count = qsv_local_size(s);
    const uint64_t local_mask = count - 1;
    const uint64_t index_base = s->world > 1 ? ((uint64_t)s->rank << s->n_local) : 0;
    const int blocks = grid_for(s, count);
    // scratch: partials [EV_TERMS][blocks] per batch is reused; results [n_terms]
    const size_t part_bytes = (size_t)EV_TERMS * blocks * sizeof(double);
    const size_t res_off = (part_bytes + 255) & ~(size_t)255;
    QSV_TRY(qsv_ensure_scratch(s, res_off + n_terms * sizeof(double) + 256));
    double* d_part = (double*)s->scratch;
    double* d_res = (double*)((char*)s->scratch + res_off);
    std::vector<uint64_t> result_slot(n_terms);
    uint64_t slot = 0;
    for (size_t b = 0; b < batches.size(); ++b) {
        TermBatch& B = batches[b];
        if (B.x_mask & ~local_mask) {
            // a pair partner lives on another rank: bring the X qubits local first
            QSV_TRY(qsv_dist_localize_xmask(s, B.x_mask));
            QSV_FAIL(QSV_ERR_UNSUPPORTED, "internal: x_mask with global bits after localisation");
        }
        k_expval_batch<<<blocks, EV_THREADS, 0, s->stream>>>(s->amp, count, index_base, B, d_part, blocks);
        k_reduce_partials<<<B.nterms, EV_THREADS, 0, s->stream>>>(d_part, blocks, blocks, d_res + slot);
        QSV_CUDA(cudaGetLastError());
        s->ctr.kernel_launches += 2;
        s->ctr.read_passes += 1;
        for (int t = 0; t < B.nterms; ++t) result_slot[batch_terms[b][t]] = slot + t;
        slot += B.nterms;
    }
    if (s->world > 1) QSV_TRY(qsv_dist_allreduce_sum(s, d_res, n_terms));
    QSV_TRY(qsv_ensure_pinned(s, n_terms * sizeof(double)));
    QSV_CUDA(cudaMemcpyAsync(s->pinned, d_res, n_terms * sizeof(double), cudaMemcpyDeviceToHost, s->stream));
    QSV_CUDA(cudaStreamSynchronize(s->stream));
    s->ctr.d2h_bytes += n_terms * sizeof(double);
    const double* h = (const double*)s->pinned;
    for (uint64_t t = 0; t < n_terms; ++t) out[t] = h[result_slot[t]];
    return QSV_OK;
}

extern "C" int qsv_expval_pauli(const qsv_state* s, uint64_t z_mask, uint64_t x_mask, double phase_re,
                                double phase_im, int x_max, double* out) {
    (void)x_max;
    if (!s || !out) QSV_FAIL(QSV_ERR_INVALID, "qsv_expval_pauli: NULL argument");
    if (s->n >= 64)
        QSV_FAIL(QSV_ERR_OVERFLOW, "The value for num_qubits, %d, is too large and would overflow", s->n);
    double ph[2] = {phase_re, phase_im};
    if (x_mask == 0) { ph[0] = 1.0; ph[1] = 0.0; }
    return qsv_expval_pauli_batch(s, 1, &z_mask, &x_mask, ph, out);
}

static int inner_impl(qsv_state* s, const double2* a, const double2* b, double* out2) {
    const uint64_t count = qsv_local_size(s);
    const int blocks = grid_for(s, count);
    const size_t part_bytes = (size_t)2 * blocks * sizeof(double);
    const size_t res_off = (part_bytes + 255) & ~(size_t)255;
    QSV_TRY(qsv_ensure_scratch(s, res_off + 256));
    double* d_part = (double*)s->scratch;
    double* d_res = (double*)((char*)s->scratch + res_off);
    k_inner<<<blocks, EV_THREADS, 0, s->stream>>>(a, b, count, d_part, blocks);
    k_reduce_partials<<<2, EV_THREADS, 0, s->stream>>>(d_part, blocks, blocks, d_res);
    QSV_CUDA(cudaGetLastError());
    s->ctr.kernel_launches += 2;
    s->ctr.read_passes += 1;
    if (s->world > 1) QSV_TRY(qsv_dist_allreduce_sum(s, d_res, 2));
    QSV_TRY(qsv_ensure_pinned(s, 16));
    QSV_CUDA(cudaMemcpyAsync(s->pinned, d_res, 16, cudaMemcpyDeviceToHost, s->stream));
    QSV_CUDA(cudaStreamSynchronize(s->stream));
    out2[0] = ((double*)s->pinned)[0];
    out2[1] = ((double*)s->pinned)[1];
    return QSV_OK;
}

extern "C" int qsv_norm2(const qsv_state* cs, double* out) {
    qsv_state* s = const_cast<qsv_state*>(cs);
    if (!s || !out) QSV_FAIL(QSV_ERR_INVALID, "qsv_norm2: NULL argument");
    QSV_CUDA(cudaSetDevice(s->device));
    double r[2];
    QSV_TRY(inner_impl(s, s->amp, s->amp, r));
    *out = r[0];
    return QSV_OK;
}

extern "C" int qsv_inner(const qsv_state* ca, const qsv_state* cb, double* out_c128) {
    qsv_state* a = const_cast<qsv_state*>(ca);
    qsv_state* b = const_cast<qsv_state*>(cb);
    if (!a || !b || !out_c128) QSV_FAIL(QSV_ERR_INVALID, "qsv_inner: NULL argument");
    if (a->n != b->n || a->device != b->device || a->world != b->world)
        QSV_FAIL(QSV_ERR_INVALID, "qsv_inner: states differ in size, device or sharding");
    for (int q = 0; q < a->n; ++q)
        if (a->perm[q] != b->perm[q]) QSV_FAIL(QSV_ERR_UNSUPPORTED, "qsv_inner: states have different qubit layouts");
    QSV_CUDA(cudaSetDevice(a->device));
    QSV_CUDA(cudaStreamSynchronize(b->stream));
    return inner_impl(a, a->amp, b->amp, out_c128);
}
