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

// All Z-type terms of an observable in ONE streaming pass.
//   <Z^m> = sum_i (-1)^popc(i & m) p_i  is the Walsh-Hadamard transform of p at m.
// A warp takes 256 consecutive amplitudes (lane l holds elements r*32 + l, r = 0..7: coalesced
// 512-byte loads), transforms them in place (3 register stages + 5 shuffle stages), parks the
// 256 coefficients in shared memory, and then every lane serves its own terms
// t = jj*32 + lane: coefficient index = low 8 bits of z_t, sign = parity of the remaining mask
// bits against the chunk index.  Cost per amplitude ~30 instructions, independent of the
// number of terms (up to WHT_JJ*32 per launch).
#define WHT_JJ 16
#define WHT_TERMS (WHT_JJ * 32)
#define WHT_WARPS 8

// PROB: the buffer holds |a|^2 in the layout k_fused_pass's store_prob leaves behind (probability
// of amplitude i at double (i >> 4) * 32 + (i & 15)): half the bytes of the amplitude read.
// JJ: terms per lane (the launch serves up to 32 * JJ terms); fewer terms -> fewer registers -> a
// third resident block per SM.
template <bool PROB, int JJ>
__global__ void __launch_bounds__(32 * WHT_WARPS, JJ <= 12 ? 3 : 2) k_expval_z_wht(const double2* __restrict__ amp,
                                                                 uint64_t nchunks, uint64_t chunk_base,
                                                                 const uint64_t* __restrict__ zmasks,
                                                                 int nterms, double* __restrict__ partials,
                                                                 int partial_stride) {
    __shared__ double wsm[WHT_WARPS][256];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // per term: byte offset of its coefficient in the chunk's table (low 8 mask bits) and the mask
    // bits above them (at most 32: states have at most 40 qubits, checked by the host)
    uint32_t zlo[JJ], zhi[JJ];
    double acc[JJ];
#pragma unroll
    for (int jj = 0; jj < JJ; ++jj) {
        const int t = jj * 32 + lane;
        const uint64_t z = t < nterms ? zmasks[t] : 0ull;
        zlo[jj] = (uint32_t)(z & 255ull) << 3;
        zhi[jj] = (uint32_t)(z >> 8);
        acc[jj] = 0.0;
    }
    for (uint64_t chunk = blockIdx.x * (uint64_t)WHT_WARPS + warp; chunk < nchunks;
         chunk += (uint64_t)gridDim.x * WHT_WARPS) {
        double p[8];
        if (PROB) {
            const double* src = reinterpret_cast<const double*>(amp) + chunk * 512 + (lane & 15) + ((lane >> 4) << 5);
#pragma unroll
            for (int r = 0; r < 8; ++r) p[r] = src[r * 64];
        } else {
            const double2* src = amp + chunk * 256 + lane;
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                const double2 a = src[r * 32];
                p[r] = a.x * a.x + a.y * a.y;
            }
        }
#pragma unroll
        for (int sft = 1; sft < 8; sft <<= 1) {
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                if (!(r & sft)) {
                    const double a = p[r], b = p[r | sft];
                    p[r] = a + b;
                    p[r | sft] = a - b;
                }
            }
        }
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                const double other = __shfl_xor_sync(0xffffffffu, p[r], o);
                p[r] = fma((lane & o) ? -1.0 : 1.0, p[r], other);  // other - p / other + p, exactly
            }
        }
#pragma unroll
        for (int r = 0; r < 8; ++r) wsm[warp][r * 32 + lane] = p[r];
        __syncwarp();
        const uint32_t gchunk = (uint32_t)(chunk | chunk_base);
        const char* coef = reinterpret_cast<const char*>(wsm[warp]);
#pragma unroll
        for (int jj = 0; jj < JJ; ++jj) {
            if (jj * 32 < nterms) {
                const double v = *reinterpret_cast<const double*>(coef + zlo[jj]);
                const int flip = (__popc(zhi[jj] & gchunk) & 1) << 31;  // sign bit of the coefficient
                acc[jj] += __hiloint2double(__double2hiint(v) ^ flip, __double2loint(v));
            }
        }
        __syncwarp();
    }
    // reduce the WHT_WARPS warps of the block, term by term
    __syncthreads();
#pragma unroll
    for (int jj = 0; jj < JJ; ++jj) {
        if (jj * 32 < nterms) {
            wsm[warp][lane] = acc[jj];
            __syncthreads();
            if (warp == 0) {
                double v = 0.0;
#pragma unroll
                for (int w = 0; w < WHT_WARPS; ++w) v += wsm[w][lane];
                const int t = jj * 32 + lane;
                if (t < nterms) partials[(size_t)t * partial_stride + blockIdx.x] = v;
            }
            __syncthreads();
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
int qsv_dist_localize(qsv_state* s, uint64_t want_local, uint64_t keep);   // qsv_dist.cu
int qsv_dist_min_lpos(const qsv_state* s);

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

struct TermItem {
    uint64_t lx, lz;  // logical masks (as given by the caller)
    uint64_t x, z;  // physical masks
    int sel;        // phase selector: phase = (-i)^sel
    int ny;         // number of Y factors: popc(x & z)
    uint64_t idx;   // position in the caller's arrays
};

// Z-type terms of `st` (masks in zlist) -> d_res[slot0 ...]; one WHT pass per WHT_TERMS terms.
static int run_wht_terms(qsv_state* st, const std::vector<uint64_t>& zlist, double* d_res,
                         qsv_counters* ctr, bool prob_layout = false) {
    if (st->n > 40) QSV_FAIL(QSV_ERR_UNSUPPORTED, "expectation values: more than 40 qubits");
    const uint64_t count = qsv_local_size(st);
    const uint64_t nchunks = count >> 8;
    const uint64_t chunk_base = st->world > 1 ? ((uint64_t)st->rank << (st->n_local - 8)) : 0;
    uint64_t blocks64 = (nchunks + WHT_WARPS - 1) / WHT_WARPS;
    // grid-stride kernel: whole waves only (3 resident blocks per SM up to 384 terms, else 2)
    const size_t most = std::min<size_t>(WHT_TERMS, zlist.size());
    uint64_t cap = (uint64_t)st->num_sms * (most <= 384 ? 3 : 4);
    if (blocks64 > cap) blocks64 = cap;
    const int blocks = (int)blocks64;
    // scratch (of st): [partials WHT_TERMS x blocks][z masks WHT_TERMS]
    const size_t part_bytes = (size_t)WHT_TERMS * blocks * sizeof(double);
    const size_t z_off = (part_bytes + 255) & ~(size_t)255;
    QSV_TRY(qsv_ensure_scratch(st, z_off + WHT_TERMS * sizeof(uint64_t) + 256));
    double* d_part = (double*)st->scratch;
    uint64_t* d_z = (uint64_t*)((char*)st->scratch + z_off);
    for (size_t p0 = 0; p0 < zlist.size(); p0 += WHT_TERMS) {
        const int nt = (int)std::min<size_t>(WHT_TERMS, zlist.size() - p0);
        // pageable source: the copy is staged by the runtime before the call returns
        QSV_CUDA(cudaMemcpyAsync(d_z, zlist.data() + p0, nt * sizeof(uint64_t), cudaMemcpyHostToDevice,
                                 st->stream));
#define QSV_WHT_LAUNCH(PROB, JJ) \
    k_expval_z_wht<PROB, JJ><<<blocks, 32 * WHT_WARPS, 0, st->stream>>>(st->amp, nchunks, chunk_base, d_z, nt, d_part, blocks)
        const int jj = (nt + 31) / 32;
        if (prob_layout) {
            if (jj <= 4) QSV_WHT_LAUNCH(true, 4);
            else if (jj <= 8) QSV_WHT_LAUNCH(true, 8);
            else if (jj <= 12) QSV_WHT_LAUNCH(true, 12);
            else QSV_WHT_LAUNCH(true, 16);
        } else {
            if (jj <= 4) QSV_WHT_LAUNCH(false, 4);
            else if (jj <= 8) QSV_WHT_LAUNCH(false, 8);
            else if (jj <= 12) QSV_WHT_LAUNCH(false, 12);
            else QSV_WHT_LAUNCH(false, 16);
        }
#undef QSV_WHT_LAUNCH
        k_reduce_partials<<<nt, EV_THREADS, 0, st->stream>>>(d_part, blocks, blocks, d_res + p0);
        QSV_CUDA(cudaGetLastError());
        ctr->kernel_launches += 2;
        ctr->read_passes += 1;
        ctr->h2d_bytes += nt * sizeof(uint64_t);
    }
    return QSV_OK;
}

// x-type terms evaluated directly: one pass per (x_mask, <= EV_TERMS terms)
static int run_direct_terms(qsv_state* s, const std::vector<TermItem>& items,
                            const std::vector<size_t>& which, double* d_res_base,
                            std::vector<uint64_t>& result_slot, uint64_t& slot) {
    std::vector<size_t> order(which);
    std::stable_sort(order.begin(), order.end(),
                     [&](size_t a, size_t b) { return items[a].x < items[b].x; });
    const uint64_t count = qsv_local_size(s);
    const uint64_t local_mask = count - 1;
    const uint64_t index_base = s->world > 1 ? ((uint64_t)s->rank << s->n_local) : 0;
    const int blocks = grid_for(s, count);
    double* d_part = (double*)s->scratch;
    for (size_t p = 0; p < order.size();) {
        TermBatch B;
        memset(&B, 0, sizeof B);
        B.x_mask = items[order[p]].x;
        B.split = B.x_mask ? __builtin_ctzll(B.x_mask) : -1;
        if (B.x_mask & ~local_mask)
            QSV_FAIL(QSV_ERR_UNSUPPORTED, "expval: X/Y factor on a sharded (global) qubit");
        size_t first = p;
        while (p < order.size() && items[order[p]].x == B.x_mask && B.nterms < EV_TERMS) {
            B.z_mask[B.nterms] = items[order[p]].z;
            B.phase_sel[B.nterms] = items[order[p]].sel;
            ++B.nterms;
            ++p;
        }
        k_expval_batch<<<blocks, EV_THREADS, 0, s->stream>>>(s->amp, count, index_base, B, d_part, blocks);
        k_reduce_partials<<<B.nterms, EV_THREADS, 0, s->stream>>>(d_part, blocks, blocks, d_res_base + slot);
        QSV_CUDA(cudaGetLastError());
        s->ctr.kernel_launches += 2;
        s->ctr.read_passes += 1;
        for (int t = 0; t < B.nterms; ++t) result_slot[items[order[first + t]].idx] = slot + t;
        slot += B.nterms;
    }
    return QSV_OK;
}

// Qubit-wise-commuting family: every member is, on each qubit, either I or the family's basis
struct QwcGroup {
    uint64_t xb, zb;  // basis per qubit: (x,z) = (1,0) X, (1,1) Y, (0,1) Z, (0,0) unset
    std::vector<size_t> members;
    int distinct_x;
};

extern "C" int qsv_expval_pauli_batch(const qsv_state* cs, uint64_t n_terms, const uint64_t* z_masks,
                                      const uint64_t* x_masks, const double* phases, double* out) {
    QSV_RANGE("qsv_expval_pauli_batch");
    qsv_state* s = const_cast<qsv_state*>(cs);
    if (!s || (n_terms && (!z_masks || !x_masks || !out)))
        QSV_FAIL(QSV_ERR_INVALID, "qsv_expval_pauli_batch: NULL argument");
    if (n_terms == 0) return QSV_OK;
    QSV_CUDA(cudaSetDevice(s->device));
    QSV_TRY(qsv_materialize(s));
    const uint64_t full = s->n >= 64 ? ~0ull : ((1ull << s->n) - 1);
    std::vector<TermItem> items(n_terms);
    for (uint64_t t = 0; t < n_terms; ++t) {
        if ((z_masks[t] | x_masks[t]) & ~full)
            QSV_FAIL(QSV_ERR_INVALID, "qsv_expval_pauli_batch: term %llu has mask bits beyond qubit %d",
                     (unsigned long long)t, s->n - 1);
        int sel = 0;
        if (phases) QSV_TRY(phase_selector(phases[2 * t], phases[2 * t + 1], &sel));
        TermItem it;
        it.lx = x_masks[t];
        it.lz = z_masks[t];
        it.x = map_mask(s, x_masks[t]);
        it.z = map_mask(s, z_masks[t]);
        it.sel = sel;
        it.ny = __builtin_popcountll(it.x & it.z);
        it.idx = t;
        items[t] = it;
    }
    const uint64_t count = qsv_local_size(s);
    const int blocks = grid_for(s, count);
    // device results: one double per term, slots assigned as work is issued
    const size_t part_bytes = (size_t)EV_TERMS * blocks * sizeof(double);
    const size_t wht_bytes = (size_t)WHT_TERMS * (s->num_sms * 4) * sizeof(double) + WHT_TERMS * 8 + 512;
    const size_t res_off = ((part_bytes > wht_bytes ? part_bytes : wht_bytes) + 255) & ~(size_t)255;
    QSV_TRY(qsv_ensure_scratch(s, res_off + n_terms * sizeof(double) + 256));
    double* d_res = (double*)((char*)s->scratch + res_off);
    std::vector<uint64_t> result_slot(n_terms);
    std::vector<double> factor(n_terms, 1.0);  // multiplies the device value on the host
    uint64_t slot = 0;
    const bool wht_ok = s->n_local >= 8;

    // --- 1. Z-type terms: one WHT pass
    std::vector<size_t> zterms, xterms;
    for (size_t t = 0; t < n_terms; ++t) (items[t].x == 0 && wht_ok ? zterms : xterms).push_back(t);
    if (!zterms.empty()) {
        std::vector<uint64_t> zl(zterms.size());
        for (size_t i = 0; i < zterms.size(); ++i) {
            zl[i] = items[zterms[i]].z;
            result_slot[items[zterms[i]].idx] = slot + i;
        }
        QSV_TRY(run_wht_terms(s, zl, d_res + slot, &s->ctr));
        slot += zterms.size();
    }

    // --- 2. X-type terms: qubit-wise-commuting families
    std::vector<QwcGroup> groups;
    std::vector<size_t> direct;
    if (wht_ok && s->world == 1) {
        for (size_t t : xterms) {
            const TermItem& it = items[t];
            const uint64_t supp = it.x | it.z;
            bool placed = false;
            for (QwcGroup& g : groups) {
                const uint64_t set = (g.xb | g.zb) & supp;  // qubits where both have an opinion
                if (((g.xb ^ it.x) & set) || ((g.zb ^ it.z) & set)) continue;
                g.xb |= it.x;
                g.zb |= it.z;
                g.members.push_back(t);
                placed = true;
                break;
            }
            if (!placed) {
                QwcGroup g;
                g.xb = it.x;
                g.zb = it.z;
                g.members.push_back(t);
                groups.push_back(g);
            }
        }
    } else {
        direct = xterms;
    }
    // decide per family: rotate a scratch copy into the Z basis (cost ~ 1 copy + gate passes +
    // 1 read pass) or evaluate every distinct x_mask directly (one read pass each)
    std::vector<QwcGroup*> rotated;
    for (QwcGroup& g : groups) {
        std::vector<uint64_t> xs;
        for (size_t t : g.members) xs.push_back(items[t].x);
        std::sort(xs.begin(), xs.end());
        g.distinct_x = (int)(std::unique(xs.begin(), xs.end()) - xs.begin());
        const int nrot = __builtin_popcountll(g.xb);
        const double cost_rot = 2.0 + 2.0 * ((nrot + 7) / 8) + 1.0;  // in read-pass units
        if ((double)g.distinct_x > cost_rot) rotated.push_back(&g);
        else direct.insert(direct.end(), g.members.begin(), g.members.end());
    }
    if (!rotated.empty() && !s->aux) {
        qsv_state* aux = nullptr;
        int rc = qsv_state_alloc(&aux, s->n, s->n_local, s->device, s->stream);
        if (rc == QSV_OK) {
            aux->tile_bits = s->tile_bits;
            aux->low_bits = s->low_bits;
            aux->tile_threads = s->tile_threads;
            aux->use_direct = s->use_direct;
            aux->use_cluster = s->use_cluster;
            aux->use_fused = s->use_fused;
            aux->use_direct_store = s->use_direct_store;
            aux->use_warp_local = s->use_warp_local;
            aux->use_linear_perm = s->use_linear_perm;
            aux->use_diag_split = s->use_diag_split;
            aux->lazy_init = 0;
            s->aux = aux;
        } else if (rc == QSV_ERR_NOMEM) {
            for (QwcGroup* g : rotated) direct.insert(direct.end(), g->members.begin(), g->members.end());
            rotated.clear();  // no room for a scratch copy: fall back to direct passes
        } else {
            return rc;
        }
    }
    if (!rotated.empty()) {
        qsv_state* aux = s->aux;
        // current per-qubit basis of the scratch copy (starts as the state itself: all Z)
        uint64_t cur_x = 0, cur_z = 0;
        bool have_copy = false;
        const double r2 = 0.70710678118654752440;
        // Which families start again from the state itself is decided up front: the family before
        // such a restart (and the last one) leaves nothing anyone reads but |a|^2, so its final
        // rotation pass stores probabilities only -- half the bytes written, half the bytes the
        // Walsh-Hadamard pass reads.
        static const bool prob_store = getenv("QSV_NO_PROB_STORE") == nullptr;
        auto frame_at = [](uint64_t xb, uint64_t zb, uint64_t b) { return (xb & b) ? ((zb & b) ? 2 : 1) : 0; };
        auto passes_needed = [](int nq) { return nq <= 0 ? 0 : (nq <= 12 ? 1 : 1 + (nq - 12 + 7) / 8); };
        std::vector<char> restart(rotated.size(), 0);
        {
            uint64_t fx = 0, fz = 0;
            for (size_t gi = 0; gi < rotated.size(); ++gi) {
                int n_cur = 0, n_psi = 0;
                for (int q = 0; q < s->n_local; ++q) {
                    const uint64_t b = 1ull << q;
                    const int want = frame_at(rotated[gi]->xb, rotated[gi]->zb, b);
                    n_cur += want != frame_at(fx, fz, b);
                    n_psi += want != 0;
                }
                restart[gi] = gi == 0 || passes_needed(n_psi) <= passes_needed(n_cur);
                fx = rotated[gi]->xb;  // every local qubit ends in the family's frame (Z where it has no X)
                fz = rotated[gi]->zb & rotated[gi]->xb;
            }
        }
        for (size_t gi = 0; gi < rotated.size(); ++gi) {
            QwcGroup* g = rotated[gi];
            const bool last_use = gi + 1 == rotated.size() || restart[gi + 1];
            // Two ways to reach this family's frame: keep rotating the scratch copy in place, or start
            // again from the state itself (out of place, no copy).  From the Z frame every rotation is
            // a real H, preceded by a diagonal S^dagger for Y -- half the FP64 work of the complex
            // X <-> Y change of frame -- so restart whenever that does not take more passes.
            auto frame_of = [](uint64_t xb, uint64_t zb, uint64_t b) { return (xb & b) ? ((zb & b) ? 2 : 1) : 0; };
            int n_cur = 0, n_psi = 0;
            for (int q = 0; q < s->n_local; ++q) {
                const uint64_t b = 1ull << q;
                const int want = frame_of(g->xb, g->zb, b);
                n_cur += want != frame_of(cur_x, cur_z, b);
                n_psi += want != 0;
            }
            auto passes_for = [](int nq) { return nq <= 0 ? 0 : (nq <= 12 ? 1 : 1 + (nq - 12 + 7) / 8); };
            if (!have_copy || restart[gi]) {
                // no copy: the first rotation pass reads the state and writes the scratch buffer
                aux->read_from = s->amp;
                have_copy = true;
                cur_x = cur_z = 0;
            }
            // per rotated qubit: V_new * V_cur^dagger with V_X = H, V_Y = H S^dagger, V_Z = I
            // From the Z frame, when only |a|^2 of the result is read, the butterflies run without
            // their 1/sqrt(2): the factor 2^-k of k of them is exact and goes into the term values.
            static const bool allow_unnorm = getenv("QSV_NO_UNNORM") == nullptr;
            const bool unnorm = allow_unnorm && last_use && cur_x == 0 && s->n_local <= 60;
            int n_butterflies = 0;
            std::vector<HGate> layer;
            for (int q = 0; q < s->n_local; ++q) {
                const uint64_t b = 1ull << q;
                const int want = frame_of(g->xb, g->zb, b);
                const int have = frame_of(cur_x, cur_z, b);
                if (want == have) continue;
                const cplx I(0.0, 1.0);
                cplx vx[4] = {r2, r2, r2, -r2};                       // H
                cplx vy[4] = {r2, -I * r2, r2, I * r2};               // H S^dagger
                cplx id[4] = {1.0, 0.0, 0.0, 1.0};
                HGate hg;
                hg.k = 1;
                memset(hg.q, 0, sizeof hg.q);
                hg.q[0] = q;
                static const bool fold_sdg = getenv("QSV_NO_D1RY") == nullptr;
                if (have == 0 && want == 2 && !fold_sdg) {  // S^dagger as a diagonal gate (a phase table), then H
                    hg.diag = true;
                    hg.m = {cplx(1.0, 0.0), -I};
                    layer.push_back(hg);
                }
                hg.diag = false;
                hg.m.resize(4);
                if (have == 0) {
                    // Y: H S^dagger as ONE gate -- real H applied to (a0, -i a1), and -i is a swap of
                    // components (FP_D1RY): no phase table, the FP64 work of the X family
                    const cplx* v0 = (want == 2 && fold_sdg) ? vy : vx;
                    for (int i = 0; i < 4; ++i) hg.m[i] = unnorm ? v0[i] / r2 : v0[i];
                    if (unnorm) {  // exactly (1 1 ; 1 -1) or (1 -i ; 1 i), whatever the division rounded to
                        for (int i = 0; i < 4; ++i)
                            hg.m[i] = cplx(std::round(hg.m[i].real()), std::round(hg.m[i].imag()));
                        ++n_butterflies;
                    }
                } else {
                    const cplx* vn = want == 1 ? vx : (want == 2 ? vy : id);
                    const cplx* vc = have == 1 ? vx : vy;
                    for (int r = 0; r < 2; ++r)
                        for (int c = 0; c < 2; ++c)  // (vn * vc^dagger)[r][c] = sum_k vn[r][k] conj(vc[c][k])
                            hg.m[r * 2 + c] = vn[r * 2 + 0] * std::conj(vc[c * 2 + 0]) +
                                              vn[r * 2 + 1] * std::conj(vc[c * 2 + 1]);
                }
                layer.push_back(hg);
                cur_x = (cur_x & ~b) | ((want ? b : 0));
                cur_z = (cur_z & ~b) | ((want == 2 ? b : 0));
            }
            if (layer.empty() && aux->read_from) {
                QSV_CUDA(cudaMemcpyAsync(aux->amp, s->amp, (size_t)16 << s->n_local, cudaMemcpyDeviceToDevice,
                                         s->stream));
                s->ctr.state_passes++;
                aux->read_from = nullptr;
            }
            if (!layer.empty()) {
                const uint64_t before_l = aux->ctr.kernel_launches, before_p = aux->ctr.state_passes;
                aux->prob_store_req = prob_store && last_use;
                QSV_TRY(qsv_apply_hgates(aux, layer, 0));
                s->ctr.kernel_launches += aux->ctr.kernel_launches - before_l;
                s->ctr.state_passes += aux->ctr.state_passes - before_p;
            }
            // in the rotated frame every member is a Z string on its support
            std::vector<uint64_t> zl(g->members.size());
            for (size_t i = 0; i < g->members.size(); ++i) {
                const TermItem& it = items[g->members[i]];
                zl[i] = it.x | it.z;
                result_slot[it.idx] = slot + i;
                // Re((-i)^(sel - ny)) corrects a caller-supplied phase that differs from (-i)^ny
                const int d = ((it.sel - it.ny) % 4 + 4) % 4;
                factor[it.idx] = (d == 0 ? 1.0 : (d == 2 ? -1.0 : 0.0)) * std::ldexp(1.0, -n_butterflies);
            }
            const bool probs = !layer.empty() && aux->prob_layout;
            QSV_TRY(run_wht_terms(aux, zl, d_res + slot, &s->ctr, probs));
            if (probs) {
                aux->prob_layout = 0;
                have_copy = false;  // the buffer holds no amplitudes any more
            }
            // run_wht_terms wrote through aux's scratch into this handle's d_res: same stream
            slot += g->members.size();
        }
    }
    if (s->world > 1) {
        // sharded: a term whose X/Y factors sit on global qubits needs them local first.  Rounds:
        // evaluate what is ready, then ONE qubit swap for the next batch of terms.
        const uint64_t local_mask = qsv_local_size(s) - 1;
        std::vector<size_t> remaining(direct);
        int g = 0;
        while ((1 << g) < s->world) ++g;
        while (!remaining.empty()) {
            std::vector<size_t> ready, later;
            for (size_t t : remaining) {
                items[t].x = map_mask(s, items[t].lx);
                items[t].z = map_mask(s, items[t].lz);
                ((items[t].x & ~local_mask) ? later : ready).push_back(t);
            }
            if (!ready.empty()) QSV_TRY(run_direct_terms(s, items, ready, d_res, result_slot, slot));
            remaining.swap(later);
            if (remaining.empty()) break;
            // the first waiting term decides; further terms join while their global X qubits fit
            // in the same swap and enough evictable local qubits remain outside their X masks
            uint64_t want = 0, keep = 0;
            auto globals_of = [&](uint64_t lx) {
                uint64_t w = 0;
                for (int q = 0; q < s->n; ++q)
                    if (((lx >> q) & 1ull) && s->perm[q] >= s->n_local) w |= 1ull << q;
                return w;
            };
            auto evictable = [&](uint64_t kp) {
                int c = 0;
                for (int q = 0; q < s->n; ++q)
                    if (s->perm[q] < s->n_local && !((kp >> q) & 1ull)) ++c;
                return c;
            };
            for (size_t t : remaining) {
                const uint64_t w2 = want | globals_of(items[t].lx);
                const uint64_t k2 = keep | items[t].lx;
                if (want && (__builtin_popcountll(w2) > g || evictable(k2) < __builtin_popcountll(w2))) continue;
                want = w2;
                keep = k2;
            }
            QSV_TRY(qsv_dist_localize(s, want, keep));
        }
    } else if (!direct.empty()) {
        QSV_TRY(run_direct_terms(s, items, direct, d_res, result_slot, slot));
    }

    if (s->world > 1) QSV_TRY(qsv_dist_allreduce_sum(s, d_res, n_terms));
    QSV_TRY(qsv_ensure_pinned(s, n_terms * sizeof(double)));
    QSV_CUDA(cudaMemcpyAsync(s->pinned, d_res, n_terms * sizeof(double), cudaMemcpyDeviceToHost, s->stream));
    QSV_CUDA(cudaStreamSynchronize(s->stream));
    s->ctr.d2h_bytes += n_terms * sizeof(double);
    const double* h = (const double*)s->pinned;
    for (uint64_t t = 0; t < n_terms; ++t) out[t] = factor[t] * h[result_slot[t]];
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
    QSV_TRY(qsv_materialize(s));
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
    QSV_TRY(qsv_materialize(a));
    QSV_TRY(qsv_materialize(b));
    QSV_CUDA(cudaStreamSynchronize(b->stream));
    return inner_impl(a, a->amp, b->amp, out_c128);
}
