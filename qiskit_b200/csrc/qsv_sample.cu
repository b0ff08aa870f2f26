// Probabilities, marginals, inverse-cdf sampling and collapse.
// Replaces Statevector.probabilities (statevector.py:663-673),
// QuantumState._subsystem_probabilities (quantum_state.py:437-468), the
// cumsum / searchsorted arithmetic of numpy.random.Generator.choice called from
// QuantumState.sample_memory (quantum_state.py:258-265), and the projector update
// of reset / measure (statevector.py:694-719, quantum_state.py:314-331).
//
// Sampling contract (stated in NumPy in oracle/statevector_ref.py, "engine
// summation tree"): p_i = fl(fl(re*re) + fl(im*im)); level 0 = the probabilities,
// level k+1 = block totals of level k, block = 32 segments x 32 items; inside a
// block w[seg][k] = fl(E[seg] + s[seg][k]) with s the sequential in-segment prefix
// and E the sequential chain of segment totals; a uniform u descends from the top
// block with exclusive prefix X: first item j with fl(fl(X + w_j) / total) > u,
// clamped to the block's last valid item; X <- fl(X + w_{j-1}).
#include "qsv_common.cuh"

#define SEG 32
#define BLK 1024
#define WARPS_PER_CTA 4
#define PAD 33

// level-0 item i of the probability source: amplitudes (|a|^2) or an explicit array
struct ProbSource {
    const double2* amp;  // used when probs == nullptr
    const double* probs;
    uint64_t count;
};

__device__ __forceinline__ double prob_at(const ProbSource& src, uint64_t i) {
    if (i >= src.count) return 0.0;
    if (src.probs) return src.probs[i];
    double2 a = src.amp[i];
    return __dadd_rn(__dmul_rn(a.x, a.x), __dmul_rn(a.y, a.y));
}

// Per-warp: load block `blk` of a level (coalesced), leave lane l with its segment's
// sequential inclusive prefix s[0..31] and the chained exclusive prefix E of segment totals.
__device__ __forceinline__ void warp_block_prefix(const ProbSource& src, uint64_t blk, double* wsm,
                                                  double (&s)[SEG], double& E) {
    const int lane = threadIdx.x & 31;
    const uint64_t base = blk * BLK;
#pragma unroll 4
    for (int r = 0; r < SEG; ++r) wsm[r * PAD + lane] = prob_at(src, base + (uint64_t)r * 32 + lane);
    __syncwarp();
    double run = 0.0;
#pragma unroll
    for (int k = 0; k < SEG; ++k) {
        double v = wsm[lane * PAD + k];
        run = k == 0 ? v : __dadd_rn(run, v);
        s[k] = run;
    }
    __syncwarp();
    // chain of segment totals: E_l = ((t_0 + t_1) + ...) + t_{l-1}
    double chain = 0.0;
    E = 0.0;
#pragma unroll
    for (int l = 0; l < 32; ++l) {
        double t = __shfl_sync(0xffffffffu, run, l);
        if (lane == l) E = chain;
        chain = l == 0 ? t : __dadd_rn(chain, t);
    }
}

// totals[blk] = w[31][31] of block blk; optionally also the full in-block prefix w
__global__ void __launch_bounds__(32 * WARPS_PER_CTA) k_block_totals(ProbSource src, uint64_t nblocks,
                                                                     double* __restrict__ totals,
                                                                     double* __restrict__ w_out,
                                                                     double* __restrict__ seg_out) {
    __shared__ double sm[WARPS_PER_CTA][SEG * PAD];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (uint64_t blk = blockIdx.x * (uint64_t)WARPS_PER_CTA + warp; blk < nblocks;
         blk += (uint64_t)gridDim.x * WARPS_PER_CTA) {
        double s[SEG], E;
        warp_block_prefix(src, blk, sm[warp], s, E);
        if (lane == 31) totals[blk] = __dadd_rn(E, s[SEG - 1]);
        // w[lane][31] of every segment (level 0): lets the search skip 31 of the block's 32 segments
        if (seg_out) seg_out[blk * SEG + lane] = __dadd_rn(E, s[SEG - 1]);
        if (w_out) {
#pragma unroll
            for (int k = 0; k < SEG; ++k) w_out[blk * BLK + (uint64_t)lane * SEG + k] = __dadd_rn(E, s[k]);
        }
        __syncwarp();
    }
}

struct LevelDesc {
    const double* w;  // stored in-block prefixes of this level (levels >= 1)
    uint64_t n_valid;
};

#define MAX_LEVELS 5

#define MAX_SHARDS 16
struct SearchParams {
    ProbSource src;  // level 0
    const double* seg_w;  // level 0: w[seg][31] of every block (nullptr: recompute whole blocks)
    int nlevels;     // total number of levels including level 0
    LevelDesc lv[MAX_LEVELS];  // lv[k] for k >= 1
    double total;
    // sharded states: the ranks form one more level above the local trees, chained sequentially:
    // chain[r] = fl(chain[r-1] + t_{r-1}), chain[0] = 0, total = chain[world]; a uniform belongs to
    // the first rank r with fl(chain[r+1] / total) > u (clamped) and descends with X = chain[r]
    int world, rank;
    uint64_t index_base;
    double chain[MAX_SHARDS + 1];
};

// one warp per shot
__global__ void __launch_bounds__(32 * WARPS_PER_CTA) k_search(SearchParams P, const double* __restrict__ u_in,
                                                               uint64_t shots, uint64_t* __restrict__ out) {
    __shared__ double sm[WARPS_PER_CTA][SEG * PAD];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (uint64_t sh = blockIdx.x * (uint64_t)WARPS_PER_CTA + warp; sh < shots;
         sh += (uint64_t)gridDim.x * WARPS_PER_CTA) {
        const double u = u_in[sh];
        double X = 0.0;
        uint64_t blk = 0;
        if (P.world > 1) {
            int owner = P.world - 1;
            for (int r = P.world - 2; r >= 0; --r)
                if (__ddiv_rn(P.chain[r + 1], P.total) > u) owner = r;
            if (owner != P.rank) {
                if (lane == 0) out[sh] = 0;
                continue;
            }
            X = P.chain[owner];
        }
        for (int level = P.nlevels - 1; level >= 1; --level) {
            const double* w = P.lv[level].w + blk * BLK;
            const uint64_t base = blk * BLK;
            uint64_t nv = P.lv[level].n_valid - base;
            const int last = (int)(nv < BLK ? nv : BLK) - 1;
            // step 1: first segment whose last entry exceeds u
            double wl = w[lane * SEG + SEG - 1];
            bool hit = __ddiv_rn(__dadd_rn(X, wl), P.total) > u;
            unsigned m = __ballot_sync(0xffffffffu, hit);
            int seg = m ? __ffs(m) - 1 : 31;
            // step 2: first entry inside that segment
            double we = w[seg * SEG + lane];
            hit = __ddiv_rn(__dadd_rn(X, we), P.total) > u;
            m = __ballot_sync(0xffffffffu, hit);
            int j = seg * SEG + (m ? __ffs(m) - 1 : 31);
            if (j > last) j = last;
            if (j > 0) X = __dadd_rn(X, w[j - 1]);
            blk = base + j;
        }
        if (P.seg_w) {
            // level 0 through the stored segment ends: w[l][31] = fl(E_l + t_l) is both the value the
            // full scan compares for item (l, 31) and E_{l+1}; items are non-decreasing, so the first
            // segment whose end exceeds u holds the answer and only its 32 items are re-derived
            const uint64_t base = blk * BLK;
            uint64_t nv = P.src.count - base;
            const int last = (int)(nv < BLK ? nv : BLK) - 1;
            const double wl = P.seg_w[blk * SEG + lane];
            const bool hit = __ddiv_rn(__dadd_rn(X, wl), P.total) > u;
            const unsigned m = __ballot_sync(0xffffffffu, hit);
            int j = BLK - 1;
            if (m) {
                const int l = __ffs(m) - 1;
                const double prev = __shfl_sync(0xffffffffu, wl, l > 0 ? l - 1 : 0);
                const double E = l > 0 ? prev : 0.0;
                const double pk = prob_at(P.src, base + (uint64_t)l * SEG + lane);
                double run = 0.0;
#pragma unroll
                for (int i = 0; i < SEG; ++i) {  // sequential in-segment prefix: lane k ends with s[k]
                    const double v = __shfl_sync(0xffffffffu, pk, i);
                    if (i == 0) run = v;
                    else if (i <= lane) run = __dadd_rn(run, v);
                }
                const bool h2 = __ddiv_rn(__dadd_rn(X, __dadd_rn(E, run)), P.total) > u;
                const unsigned m2 = __ballot_sync(0xffffffffu, h2);
                j = l * SEG + (m2 ? __ffs(m2) - 1 : SEG - 1);
            }
            if (j > last) j = last;
            if (lane == 0) out[sh] = (base + j) | P.index_base;
            __syncwarp();
            continue;
        }
        // level 0: recompute the in-block prefix from the source
        double s[SEG], E;
        warp_block_prefix(P.src, blk, sm[warp], s, E);
        int first = SEG;
#pragma unroll
        for (int k = SEG - 1; k >= 0; --k) {
            double c = __ddiv_rn(__dadd_rn(X, __dadd_rn(E, s[k])), P.total);
            if (c > u) first = k;
        }
        unsigned m = __ballot_sync(0xffffffffu, first < SEG);
        int j;
        if (m) {
            int l = __ffs(m) - 1;
            j = l * SEG + __shfl_sync(0xffffffffu, first, l);
        } else {
            j = BLK - 1;
        }
        const uint64_t base = blk * BLK;
        uint64_t nv = P.src.count - base;
        const int last = (int)(nv < BLK ? nv : BLK) - 1;
        if (j > last) j = last;
        if (lane == 0) out[sh] = (base + j) | P.index_base;
        __syncwarp();
    }
}

// p[i] = |a_i|^2 for the identity qargs; marginal: out[gather(i)] += |a_i|^2
__global__ void k_probs_full(const double2* __restrict__ amp, uint64_t count, double* __restrict__ out) {
    uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (; i < count; i += stride) {
        double2 a = amp[i];
        out[i] = __dadd_rn(__dmul_rn(a.x, a.x), __dmul_rn(a.y, a.y));
    }
}

struct GatherBits {
    int m;
    int bit[64];  // physical bit feeding output bit j
};

// Deterministic marginal: one thread per (output index, slice); slices are reduced in order.
// out index o fixes the m measured bits; the thread loops over the other n-m bits.
__global__ void k_probs_marginal(const double2* __restrict__ amp, int n, GatherBits G,
                                 uint64_t rest_count, double* __restrict__ out) {
    // grid-stride over outputs; each thread sums its rest-indices sequentially
    const uint64_t outs = 1ull << G.m;
    uint64_t o = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    // sorted measured bits for zero insertion
    int sorted[64];
    for (int j = 0; j < G.m; ++j) sorted[j] = G.bit[j];
    for (int a = 1; a < G.m; ++a) {
        int v = sorted[a], b = a - 1;
        while (b >= 0 && sorted[b] > v) { sorted[b + 1] = sorted[b]; --b; }
        sorted[b + 1] = v;
    }
    for (; o < outs; o += stride) {
        uint64_t fixed = 0;
        for (int j = 0; j < G.m; ++j) fixed |= ((o >> j) & 1ull) << G.bit[j];
        double acc = 0.0;
        for (uint64_t r = 0; r < rest_count; ++r) {
            uint64_t idx = r;
            for (int j = 0; j < G.m; ++j) idx = insert_zero(idx, sorted[j]);
            double2 a = amp[idx | fixed];
            acc += a.x * a.x + a.y * a.y;
        }
        out[o] = acc;
    }
    (void)n;
}

// Marginal for few outputs / many summed bits: block-parallel over the rest index with
// per-output shared-memory accumulators, then one atomicAdd per (block, output).
__global__ void __launch_bounds__(256) k_probs_marginal_small(const double2* __restrict__ amp,
                                                              uint64_t count, GatherBits G,
                                                              double* __restrict__ out) {
    extern __shared__ double acc_sm[];
    const uint32_t outs = 1u << G.m;
    for (uint32_t o = threadIdx.x; o < outs; o += blockDim.x) acc_sm[o] = 0.0;
    __syncthreads();
    uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (; i < count; i += stride) {
        double2 a = amp[i];
        double p = a.x * a.x + a.y * a.y;
        uint32_t o = 0;
        for (int j = 0; j < G.m; ++j) o |= (uint32_t)((i >> G.bit[j]) & 1ull) << j;
        atomicAdd(&acc_sm[o], p);
    }
    __syncthreads();
    for (uint32_t o = threadIdx.x; o < outs; o += blockDim.x) atomicAdd(&out[o], acc_sm[o]);
}

__global__ void k_clip01(double* p, uint64_t n) {
    uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) p[i] = fmin(fmax(p[i], 0.0), 1.0);
}

// keep the branch where the masked index bits equal `keep`, scaled; zero the rest.  One thread per
// amplitude, any number of measured bits.  rotate: the kept branch moves to the masked bits = 0
// branch; the pair (i, i | keep) is then owned by the thread whose masked bits are zero.
__global__ void __launch_bounds__(256) k_collapse(double2* __restrict__ amp, uint64_t count, uint64_t mask,
                                                  uint64_t keep, double scale, int rotate) {
    uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    const double2 zero = make_double2(0.0, 0.0);
    for (; i < count; i += stride) {
        const uint64_t c = i & mask;
        if (!rotate || keep == 0) {
            if (c == keep) {
                double2 v = amp[i];
                amp[i] = make_double2(v.x * scale, v.y * scale);
            } else {
                amp[i] = zero;
            }
        } else if (c == 0) {
            double2 v = amp[i | keep];
            amp[i] = make_double2(v.x * scale, v.y * scale);
            amp[i | keep] = zero;
        } else if (c != keep) {
            amp[i] = zero;
        }
    }
}

// partial sums of |a_i|^2 over the amplitudes whose masked index bits equal `keep`
__global__ void __launch_bounds__(256) k_masked_norm(const double2* __restrict__ amp, uint64_t count, uint64_t mask,
                                                     uint64_t keep, double* __restrict__ partials) {
    __shared__ double sh[8];
    double acc = 0.0;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < count; i += stride) {
        if ((i & mask) != keep) continue;
        const double2 a = amp[i];
        acc += a.x * a.x + a.y * a.y;
    }
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x < 32) {
        double v = threadIdx.x < 8 ? sh[threadIdx.x] : 0.0;
        v = warp_sum(v);
        if (threadIdx.x == 0) partials[blockIdx.x] = v;
    }
}

__global__ void __launch_bounds__(256) k_sum_partials(const double* __restrict__ partials, int n, double* __restrict__ out) {
    __shared__ double sh[8];
    double acc = 0.0;
    for (int i = threadIdx.x; i < n; i += 256) acc += partials[i];
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x < 32) {
        double v = threadIdx.x < 8 ? sh[threadIdx.x] : 0.0;
        v = warp_sum(v);
        if (threadIdx.x == 0) out[0] = v;
    }
}

// ----------------------------------------------------------------- host ---
int qsv_dist_canonicalize(qsv_state* s);
int qsv_dist_allgather(qsv_state* s, const double* dev_in, double* dev_out, size_t count);
int qsv_dist_allreduce_sum_u64(qsv_state* s, uint64_t* dev_buf, size_t count);

static int resolve_qargs(const qsv_state* s, const int* qargs, int m, GatherBits& G, bool& identity,
                         const char* who) {
    if (m < 0 || m > s->n) QSV_FAIL(QSV_ERR_INVALID, "%s: m = %d outside [0, %d]", who, m, s->n);
    G.m = m;
    identity = (m == s->n);
    for (int j = 0; j < m; ++j) {
        int q = qargs ? qargs[j] : j;
        if (q < 0 || q >= s->n) QSV_FAIL(QSV_ERR_INVALID, "%s: qarg %d out of range", who, q);
        for (int t = 0; t < j; ++t)
            if ((qargs ? qargs[t] : t) == q) QSV_FAIL(QSV_ERR_INVALID, "%s: duplicate qarg %d", who, q);
        G.bit[j] = s->perm[q];
        identity &= (G.bit[j] == j);
    }
    return QSV_OK;
}

static int cap_blocks(const qsv_state* s, uint64_t items, int threads) {
    uint64_t b = (items + threads - 1) / threads;
    uint64_t cap = (uint64_t)s->num_sms * 16;
    if (b > cap) b = cap;
    if (b < 1) b = 1;
    return (int)b;
}

// device array of marginal probabilities (unclipped) at d_out[2^m]
static int marginal_to_device(qsv_state* s, const GatherBits& G, double* d_out) {
    const uint64_t count = qsv_local_size(s);
    const int n = s->n_local;
    if (G.m <= 11) {
        QSV_CUDA(cudaMemsetAsync(d_out, 0, sizeof(double) << G.m, s->stream));
        int blocks = cap_blocks(s, count, 256);
        k_probs_marginal_small<<<blocks, 256, sizeof(double) << G.m, s->stream>>>(s->amp, count, G, d_out);
    } else {
        int blocks = cap_blocks(s, 1ull << G.m, 128);
        k_probs_marginal<<<blocks, 128, 0, s->stream>>>(s->amp, n, G, 1ull << (n - G.m), d_out);
    }
    QSV_CUDA(cudaGetLastError());
    s->ctr.kernel_launches++;
    s->ctr.read_passes++;
    return QSV_OK;
}

extern "C" int qsv_probabilities(const qsv_state* cs, const int* qargs, int m, double* host_out) {
    QSV_RANGE("qsv_probabilities");
    qsv_state* s = const_cast<qsv_state*>(cs);
    if (!s || !host_out) QSV_FAIL(QSV_ERR_INVALID, "qsv_probabilities: NULL argument");
    if (!qargs) m = s->n;
    QSV_CUDA(cudaSetDevice(s->device));
    QSV_TRY(qsv_materialize(const_cast<qsv_state*>(s)));
    if (s->world > 1) QSV_FAIL(QSV_ERR_UNSUPPORTED, "qsv_probabilities: sharded states (use qsv_sample / expval)");
    GatherBits G;
    bool identity;
    QSV_TRY(resolve_qargs(s, qargs, m, G, identity, "qsv_probabilities"));
    const uint64_t outs = 1ull << m;
    QSV_TRY(qsv_ensure_scratch(s, outs * sizeof(double)));
    double* d_out = (double*)s->scratch;
    if (identity) {
        k_probs_full<<<cap_blocks(s, outs, 256), 256, 0, s->stream>>>(s->amp, outs, d_out);
        s->ctr.kernel_launches++;
        s->ctr.read_passes++;
    } else {
        QSV_TRY(marginal_to_device(s, G, d_out));
    }
    k_clip01<<<cap_blocks(s, outs, 256), 256, 0, s->stream>>>(d_out, outs);
    QSV_CUDA(cudaGetLastError());
    s->ctr.kernel_launches++;
    QSV_CUDA(cudaMemcpyAsync(host_out, d_out, outs * sizeof(double), cudaMemcpyDeviceToHost, s->stream));
    QSV_CUDA(cudaStreamSynchronize(s->stream));
    s->ctr.d2h_bytes += outs * sizeof(double);
    return QSV_OK;
}

extern "C" int qsv_sample(const qsv_state* cs, const int* qargs, int m, const double* uniforms,
                          uint64_t shots, uint64_t* out_indices, double* total_out) {
    QSV_RANGE("qsv_sample");
    qsv_state* s = const_cast<qsv_state*>(cs);
    if (!s || (shots && (!uniforms || !out_indices))) QSV_FAIL(QSV_ERR_INVALID, "qsv_sample: NULL argument");
    if (!qargs) m = s->n;
    QSV_CUDA(cudaSetDevice(s->device));
    QSV_TRY(qsv_materialize(const_cast<qsv_state*>(s)));
    if (s->world > 1) {
        if (m != s->n) QSV_FAIL(QSV_ERR_UNSUPPORTED, "qsv_sample: sharded states sample all qubits only");
        for (int j = 0; j < m; ++j)
            if (qargs && qargs[j] != j) QSV_FAIL(QSV_ERR_UNSUPPORTED, "qsv_sample: sharded states need qargs in order");
        QSV_TRY(qsv_dist_canonicalize(s));  // the summation tree is defined on the identity layout
    }
    GatherBits G;
    bool identity;
    QSV_TRY(resolve_qargs(s, qargs, m, G, identity, "qsv_sample"));
    if (s->world > 1) identity = true;
    const uint64_t outs = s->world > 1 ? qsv_local_size(s) : 1ull << m;
    // scratch layout: [marginal probs (if any)] [level totals / prefixes ...] [uniforms] [indices]
    uint64_t nblk[MAX_LEVELS + 1];
    int nlevels = 1;
    nblk[0] = (outs + BLK - 1) / BLK;  // number of level-0 blocks = items of level 1
    while (nblk[nlevels - 1] > 1) {
        nblk[nlevels] = (nblk[nlevels - 1] + BLK - 1) / BLK;
        ++nlevels;
        if (nlevels > MAX_LEVELS) QSV_FAIL(QSV_ERR_UNSUPPORTED, "qsv_sample: too many levels");
    }
    // level k >= 1 has nblk[k-1] items, stored padded to nblk[k]*BLK prefixes
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off += (bytes + 255) & ~(size_t)255; return o; };
    size_t o_marg = identity ? 0 : take(outs * sizeof(double));
    size_t o_tot[MAX_LEVELS + 1], o_w[MAX_LEVELS + 1];
    for (int k = 0; k < nlevels; ++k) o_tot[k] = take(nblk[k] * sizeof(double));  // totals of level-k blocks
    for (int k = 1; k < nlevels; ++k) o_w[k] = take(nblk[k] * BLK * sizeof(double));
    // stored segment ends of level 0 (1/32 of the items): worth it once the shots would otherwise
    // re-read more than the extra array costs to write
    const bool use_seg = shots >= 4096 && nblk[0] * SEG * sizeof(double) <= ((size_t)4 << 30);
    size_t o_seg = use_seg ? take(nblk[0] * SEG * sizeof(double)) : 0;
    size_t o_all = take(MAX_SHARDS * sizeof(double));
    size_t o_u = take(shots * sizeof(double));
    size_t o_idx = take(shots * sizeof(uint64_t));
    QSV_TRY(qsv_ensure_scratch(s, off + 256));
    char* base = (char*)s->scratch;
    ProbSource src;
    src.amp = s->amp;
    src.probs = nullptr;
    src.count = outs;
    if (!identity) {
        QSV_TRY(marginal_to_device(s, G, (double*)(base + o_marg)));
        src.probs = (const double*)(base + o_marg);
    }
    const int threads = 32 * WARPS_PER_CTA;
    {
        uint64_t ctas = (nblk[0] + WARPS_PER_CTA - 1) / WARPS_PER_CTA;
        uint64_t cap = (uint64_t)s->num_sms * 8;
        if (ctas > cap) ctas = cap;
        k_block_totals<<<(unsigned)ctas, threads, 0, s->stream>>>(src, nblk[0], (double*)(base + o_tot[0]), nullptr,
                                                                  use_seg ? (double*)(base + o_seg) : nullptr);
        s->ctr.kernel_launches++;
        s->ctr.read_passes++;
    }
    SearchParams P;
    memset(&P, 0, sizeof P);
    P.src = src;
    P.seg_w = use_seg ? (const double*)(base + o_seg) : nullptr;
    P.nlevels = nlevels;
    for (int k = 1; k < nlevels; ++k) {
        ProbSource lsrc;
        lsrc.amp = nullptr;
        lsrc.probs = (const double*)(base + o_tot[k - 1]);
        lsrc.count = nblk[k - 1];
        uint64_t ctas = (nblk[k] + WARPS_PER_CTA - 1) / WARPS_PER_CTA;
        k_block_totals<<<(unsigned)ctas, threads, 0, s->stream>>>(lsrc, nblk[k], (double*)(base + o_tot[k]),
                                                                  (double*)(base + o_w[k]), nullptr);
        s->ctr.kernel_launches++;
        P.lv[k].w = (const double*)(base + o_w[k]);
        P.lv[k].n_valid = nblk[k - 1];
    }
    QSV_CUDA(cudaGetLastError());
    // total = the single top-level block total
    QSV_TRY(qsv_ensure_pinned(s, 64));
    QSV_CUDA(cudaMemcpyAsync(s->pinned, base + o_tot[nlevels - 1], sizeof(double), cudaMemcpyDeviceToHost, s->stream));
    QSV_CUDA(cudaStreamSynchronize(s->stream));
    double total = *(const double*)s->pinned;
    if (s->world > 1) {
        double* d_tot = (double*)(base + o_tot[nlevels - 1]);
        double* d_all = (double*)(base + o_all);
        QSV_TRY(qsv_dist_allgather(s, d_tot, d_all, 1));
        double tot[MAX_SHARDS];
        QSV_CUDA(cudaMemcpyAsync(s->pinned, d_all, s->world * sizeof(double), cudaMemcpyDeviceToHost, s->stream));
        QSV_CUDA(cudaStreamSynchronize(s->stream));
        memcpy(tot, s->pinned, s->world * sizeof(double));
        P.world = s->world;
        P.rank = s->rank;
        P.index_base = (uint64_t)s->rank << s->n_local;
        P.chain[0] = 0.0;
        for (int r = 0; r < s->world; ++r) P.chain[r + 1] = r == 0 ? tot[0] : P.chain[r] + tot[r];
        total = P.chain[s->world];
    } else {
        P.world = 1;
    }
    if (total_out) *total_out = total;
    if (shots == 0) return QSV_OK;
    if (!(total > 0.0)) QSV_FAIL(QSV_ERR_INVALID, "qsv_sample: probabilities sum to %g", total);
    P.total = total;
    QSV_CUDA(cudaMemcpyAsync(base + o_u, uniforms, shots * sizeof(double), cudaMemcpyHostToDevice, s->stream));
    s->ctr.h2d_bytes += shots * sizeof(double);
    {
        uint64_t ctas = (shots + WARPS_PER_CTA - 1) / WARPS_PER_CTA;
        uint64_t cap = (uint64_t)s->num_sms * 32;
        if (ctas > cap) ctas = cap;
        k_search<<<(unsigned)ctas, threads, 0, s->stream>>>(P, (const double*)(base + o_u), shots,
                                                            (uint64_t*)(base + o_idx));
        QSV_CUDA(cudaGetLastError());
        s->ctr.kernel_launches++;
    }
    if (s->world > 1) QSV_TRY(qsv_dist_allreduce_sum_u64(s, (uint64_t*)(base + o_idx), shots));
    QSV_CUDA(cudaMemcpyAsync(out_indices, base + o_idx, shots * sizeof(uint64_t), cudaMemcpyDeviceToHost, s->stream));
    QSV_CUDA(cudaStreamSynchronize(s->stream));
    s->ctr.d2h_bytes += shots * sizeof(uint64_t);
    return QSV_OK;
}

extern "C" int qsv_collapse(qsv_state* s, const int* qargs, int m, uint64_t outcome, double scale,
                            int rotate_to_zero) {
    QSV_RANGE("qsv_collapse");
    if (!s || (!qargs && m)) QSV_FAIL(QSV_ERR_INVALID, "qsv_collapse: NULL argument");
    QSV_CUDA(cudaSetDevice(s->device));
    QSV_TRY(qsv_materialize(const_cast<qsv_state*>(s)));
    if (s->world > 1) QSV_FAIL(QSV_ERR_UNSUPPORTED, "qsv_collapse: sharded states");
    GatherBits G;
    bool identity;
    QSV_TRY(resolve_qargs(s, qargs, m, G, identity, "qsv_collapse"));
    if (m < 64 && outcome >> m) QSV_FAIL(QSV_ERR_INVALID, "qsv_collapse: outcome out of range");
    uint64_t mask = 0, keep = 0;
    for (int j = 0; j < m; ++j) {
        mask |= 1ull << G.bit[j];
        keep |= ((outcome >> j) & 1ull) << G.bit[j];
    }
    const uint64_t count = qsv_local_size(s);
    k_collapse<<<cap_blocks(s, count, 256), 256, 0, s->stream>>>(s->amp, count, mask, keep, scale, rotate_to_zero);
    QSV_CUDA(cudaGetLastError());
    s->ctr.kernel_launches++;
    s->ctr.state_passes++;
    return QSV_OK;
}

// probability of ONE outcome of the marginal on qargs: the value Statevector.reset / measure divide by
// (statevector.py:705-712, quantum_state.py:317-324 read it out of the full marginal array)
extern "C" int qsv_outcome_probability(const qsv_state* cs, const int* qargs, int m, uint64_t outcome, double* out) {
    qsv_state* s = const_cast<qsv_state*>(cs);
    if (!s || !out || (!qargs && m)) QSV_FAIL(QSV_ERR_INVALID, "qsv_outcome_probability: NULL argument");
    QSV_CUDA(cudaSetDevice(s->device));
    QSV_TRY(qsv_materialize(s));
    if (s->world > 1) QSV_FAIL(QSV_ERR_UNSUPPORTED, "qsv_outcome_probability: sharded states");
    GatherBits G;
    bool identity;
    QSV_TRY(resolve_qargs(s, qargs, m, G, identity, "qsv_outcome_probability"));
    if (m < 64 && outcome >> m) QSV_FAIL(QSV_ERR_INVALID, "qsv_outcome_probability: outcome out of range");
    uint64_t mask = 0, keep = 0;
    for (int j = 0; j < m; ++j) {
        mask |= 1ull << G.bit[j];
        keep |= ((outcome >> j) & 1ull) << G.bit[j];
    }
    const uint64_t count = qsv_local_size(s);
    const int blocks = cap_blocks(s, count, 256);
    QSV_TRY(qsv_ensure_scratch(s, (size_t)(blocks + 1) * sizeof(double) + 256));
    QSV_TRY(qsv_ensure_pinned(s, 64));
    double* d_part = (double*)s->scratch;
    k_masked_norm<<<blocks, 256, 0, s->stream>>>(s->amp, count, mask, keep, d_part + 1);
    k_sum_partials<<<1, 256, 0, s->stream>>>(d_part + 1, blocks, d_part);
    QSV_CUDA(cudaGetLastError());
    s->ctr.kernel_launches += 2;
    s->ctr.read_passes++;
    QSV_CUDA(cudaMemcpyAsync(s->pinned, d_part, sizeof(double), cudaMemcpyDeviceToHost, s->stream));
    QSV_CUDA(cudaStreamSynchronize(s->stream));
    s->ctr.d2h_bytes += sizeof(double);
    const double p = *(const double*)s->pinned;
    *out = p < 0.0 ? 0.0 : (p > 1.0 ? 1.0 : p);  // np.clip of Statevector.probabilities (statevector.py:668)
    return QSV_OK;
}
