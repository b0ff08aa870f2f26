// Counts post-processing on the device: the step after the sampler.
//
// Integer / bit work over the shot array, HBM bound: diagonal-operator expectation values from
// samples (crates/accelerate/src/sampled_exp_val.rs:46-67), marginalisation of samples
// (crates/accelerate/src/results/marginalization.rs, BitArray.slice_bits) and the counts
// histogram (BitArray.get_int_counts, quantum_state.py:291-293) through an open-addressing
// hash table in HBM.  Handle-less entry points: they take a device id and host arrays.
#include "qsv_common.cuh"

#include <algorithm>
#include <utility>

#define CT_THREADS 256
#define CT_TERMS 8  // terms evaluated per pass over the keys

struct CtTerms {
    uint64_t z[CT_TERMS], zero[CT_TERMS], one[CT_TERMS];
    int n;
};

// partials[t][block] = sum over this block's keys of w * (+1 | -1 | 0)
__global__ void __launch_bounds__(CT_THREADS) k_sampled_expval(const uint64_t* __restrict__ keys,
                                                               const double* __restrict__ weights, uint64_t n_keys,
                                                               const CtTerms T, double* __restrict__ partials) {
    __shared__ double sh[CT_THREADS / 32];
    double acc[CT_TERMS];
#pragma unroll
    for (int t = 0; t < CT_TERMS; ++t) acc[t] = 0.0;
    const uint64_t stride = (uint64_t)gridDim.x * CT_THREADS;
    for (uint64_t i = blockIdx.x * (uint64_t)CT_THREADS + threadIdx.x; i < n_keys; i += stride) {
        const uint64_t k = keys[i];
        const double w = weights ? weights[i] : 1.0;
#pragma unroll
        for (int t = 0; t < CT_TERMS; ++t) {
            if (t >= T.n) break;
            // '0' wants the bit clear, '1' wants it set (OPERS table, sampled_exp_val.rs:27-33)
            const bool keep = ((k & T.zero[t]) == 0) && ((k & T.one[t]) == T.one[t]);
            const double sgn = (__popcll(k & T.z[t]) & 1) ? -w : w;
            acc[t] += keep ? sgn : 0.0;
        }
    }
    for (int t = 0; t < T.n; ++t) {
        double v = warp_sum(acc[t]);
        __syncthreads();
        if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
        __syncthreads();
        if (threadIdx.x == 0) {
            double r = 0.0;
            for (int i = 0; i < CT_THREADS / 32; ++i) r += sh[i];
            partials[(size_t)t * gridDim.x + blockIdx.x] = r;
        }
    }
}

__global__ void __launch_bounds__(CT_THREADS) k_weight_sum(const double* __restrict__ weights, uint64_t n,
                                                           double* __restrict__ partials) {
    __shared__ double sh[CT_THREADS / 32];
    double acc = 0.0;
    const uint64_t stride = (uint64_t)gridDim.x * CT_THREADS;
    for (uint64_t i = blockIdx.x * (uint64_t)CT_THREADS + threadIdx.x; i < n; i += stride) acc += weights[i];
    double v = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0) {
        double r = 0.0;
        for (int i = 0; i < CT_THREADS / 32; ++i) r += sh[i];
        partials[blockIdx.x] = r;
    }
}

// out[row] = sum_b partials[row][b] in a fixed order (one block per row)
__global__ void __launch_bounds__(CT_THREADS) k_ct_reduce(const double* __restrict__ partials, int nblocks,
                                                          double* __restrict__ out) {
    __shared__ double sh[CT_THREADS / 32];
    const double* row = partials + (size_t)blockIdx.x * nblocks;
    double acc = 0.0;
    for (int b = threadIdx.x; b < nblocks; b += CT_THREADS) acc += row[b];
    double v = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0) {
        double r = 0.0;
        for (int i = 0; i < CT_THREADS / 32; ++i) r += sh[i];
        out[blockIdx.x] = r;
    }
}

struct CtBits {
    int8_t src[64];
};

// out[i] bit j = keys[i] bit src[j] (src[j] < 0: constant 0)
__global__ void __launch_bounds__(CT_THREADS) k_marginalize(const uint64_t* __restrict__ keys, uint64_t n, int m,
                                                            const CtBits B, uint64_t* __restrict__ out) {
    const uint64_t stride = (uint64_t)gridDim.x * CT_THREADS;
    for (uint64_t i = blockIdx.x * (uint64_t)CT_THREADS + threadIdx.x; i < n; i += stride) {
        const uint64_t k = keys[i];
        uint64_t r = 0;
        for (int j = 0; j < m; ++j)
            if (B.src[j] >= 0) r |= ((k >> B.src[j]) & 1ull) << j;
        out[i] = r;
    }
}

// ---- counts: open-addressing hash table (key + 1 stored so that 0 marks an empty slot) --------
__device__ __forceinline__ uint64_t ct_hash(uint64_t k) {
    k ^= k >> 33;
    k *= 0xff51afd7ed558ccdull;
    k ^= k >> 33;
    k *= 0xc4ceb9fe1a85ec53ull;
    k ^= k >> 33;
    return k;
}

__global__ void __launch_bounds__(CT_THREADS) k_hist_insert(const uint64_t* __restrict__ keys, uint64_t n,
                                                            unsigned long long* __restrict__ slots,
                                                            unsigned long long* __restrict__ counts,
                                                            uint64_t mask) {
    const uint64_t stride = (uint64_t)gridDim.x * CT_THREADS;
    for (uint64_t i = blockIdx.x * (uint64_t)CT_THREADS + threadIdx.x; i < n; i += stride) {
        const unsigned long long tag = keys[i] + 1ull;  // key 2^64-1 is rejected on the host
        uint64_t h = ct_hash(keys[i]) & mask;
        while (true) {
            const unsigned long long prev = atomicCAS(&slots[h], 0ull, tag);
            if (prev == 0ull || prev == tag) {
                atomicAdd(&counts[h], 1ull);
                break;
            }
            h = (h + 1) & mask;
        }
    }
}

__global__ void __launch_bounds__(CT_THREADS) k_hist_compact(const unsigned long long* __restrict__ slots,
                                                             const unsigned long long* __restrict__ counts,
                                                             uint64_t table, uint64_t* __restrict__ out_keys,
                                                             uint64_t* __restrict__ out_counts,
                                                             unsigned long long* __restrict__ n_out) {
    const uint64_t stride = (uint64_t)gridDim.x * CT_THREADS;
    for (uint64_t i = blockIdx.x * (uint64_t)CT_THREADS + threadIdx.x; i < table; i += stride) {
        if (slots[i]) {
            const unsigned long long pos = atomicAdd(n_out, 1ull);
            out_keys[pos] = slots[i] - 1ull;
            out_counts[pos] = counts[i];
        }
    }
}

// ---- host side ------------------------------------------------------------------------------
namespace {
struct DevBuf {  // cudaMalloc'd scratch released on scope exit
    void* p = nullptr;
    ~DevBuf() {
        if (p) cudaFree(p);
    }
    int alloc(size_t bytes) {
        QSV_CUDA(cudaMalloc(&p, bytes ? bytes : 1));
        return QSV_OK;
    }
    template <class T>
    T* as() const {
        return static_cast<T*>(p);
    }
};

int ct_device(const char* fn, int device, int* sms) {
    int count = 0;
    QSV_CUDA(cudaGetDeviceCount(&count));
    if (device < 0 || device >= count) QSV_FAIL(QSV_ERR_INVALID, "%s: device %d out of range (%d visible)", fn, device, count);
    QSV_CUDA(cudaSetDevice(device));
    QSV_CUDA(cudaDeviceGetAttribute(sms, cudaDevAttrMultiProcessorCount, device));
    return QSV_OK;
}

int ct_grid(int sms, uint64_t n) {
    const uint64_t want = (n + CT_THREADS - 1) / CT_THREADS;
    return (int)std::max<uint64_t>(1, std::min<uint64_t>(want, (uint64_t)sms * 8));
}
}  // namespace

extern "C" int qsv_sampled_expval(int device, const uint64_t* keys, const double* weights, uint64_t n_keys,
                                  uint64_t n_terms, const uint64_t* z_masks, const uint64_t* zero_masks,
                                  const uint64_t* one_masks, double* out) {
    if (n_terms == 0) return QSV_OK;
    if (!keys || !z_masks || !zero_masks || !one_masks || !out) QSV_FAIL(QSV_ERR_INVALID, "qsv_sampled_expval: NULL argument");
    if (n_keys == 0) QSV_FAIL(QSV_ERR_INVALID, "qsv_sampled_expval: empty distribution");
    for (uint64_t t = 0; t < n_terms; ++t)
        if ((z_masks[t] & zero_masks[t]) | (z_masks[t] & one_masks[t]) | (zero_masks[t] & one_masks[t]))
            QSV_FAIL(QSV_ERR_INVALID, "qsv_sampled_expval: term %llu has overlapping Z/0/1 masks", (unsigned long long)t);
    int sms = 0;
    QSV_TRY(ct_device("qsv_sampled_expval", device, &sms));
    const int blocks = ct_grid(sms, n_keys);
    DevBuf d_keys, d_w, d_part, d_res;
    QSV_TRY(d_keys.alloc(n_keys * 8));
    QSV_CUDA(cudaMemcpy(d_keys.p, keys, n_keys * 8, cudaMemcpyHostToDevice));
    if (weights) {
        QSV_TRY(d_w.alloc(n_keys * 8));
        QSV_CUDA(cudaMemcpy(d_w.p, weights, n_keys * 8, cudaMemcpyHostToDevice));
    }
    QSV_TRY(d_part.alloc((size_t)CT_TERMS * blocks * 8));
    QSV_TRY(d_res.alloc((n_terms + 1) * 8));
    double denom = (double)n_keys;
    if (weights) {  // denominator: sum of the distribution's values (sampled_exp_val.rs:52)
        k_weight_sum<<<blocks, CT_THREADS>>>(d_w.as<double>(), n_keys, d_part.as<double>());
        k_ct_reduce<<<1, CT_THREADS>>>(d_part.as<double>(), blocks, d_res.as<double>() + n_terms);
        QSV_CUDA(cudaMemcpy(&denom, d_res.as<double>() + n_terms, 8, cudaMemcpyDeviceToHost));
    }
    for (uint64_t t0 = 0; t0 < n_terms; t0 += CT_TERMS) {
        CtTerms T;
        memset(&T, 0, sizeof T);
        T.n = (int)std::min<uint64_t>(CT_TERMS, n_terms - t0);
        for (int t = 0; t < T.n; ++t) {
            T.z[t] = z_masks[t0 + t];
            T.zero[t] = zero_masks[t0 + t];
            T.one[t] = one_masks[t0 + t];
        }
        k_sampled_expval<<<blocks, CT_THREADS>>>(d_keys.as<uint64_t>(), weights ? d_w.as<double>() : nullptr, n_keys, T,
                                                 d_part.as<double>());
        k_ct_reduce<<<T.n, CT_THREADS>>>(d_part.as<double>(), blocks, d_res.as<double>() + t0);
    }
    QSV_CUDA(cudaGetLastError());
    QSV_CUDA(cudaMemcpy(out, d_res.p, n_terms * 8, cudaMemcpyDeviceToHost));
    for (uint64_t t = 0; t < n_terms; ++t) out[t] /= denom;  // exp_val / denom (:66)
    return QSV_OK;
}

extern "C" int qsv_marginalize_samples(int device, const uint64_t* keys, uint64_t n, const int* src_bit, int m,
                                       uint64_t* out) {
    if (n == 0) return QSV_OK;
    if (!keys || !out || (m > 0 && !src_bit)) QSV_FAIL(QSV_ERR_INVALID, "qsv_marginalize_samples: NULL argument");
    if (m < 0 || m > 64) QSV_FAIL(QSV_ERR_INVALID, "qsv_marginalize_samples: %d output bits", m);
    CtBits B;
    memset(&B, 0, sizeof B);
    for (int j = 0; j < m; ++j) {
        if (src_bit[j] >= 64) QSV_FAIL(QSV_ERR_INVALID, "qsv_marginalize_samples: source bit %d", src_bit[j]);
        B.src[j] = (int8_t)(src_bit[j] < 0 ? -1 : src_bit[j]);
    }
    int sms = 0;
    QSV_TRY(ct_device("qsv_marginalize_samples", device, &sms));
    DevBuf d_in, d_out;
    QSV_TRY(d_in.alloc(n * 8));
    QSV_TRY(d_out.alloc(n * 8));
    QSV_CUDA(cudaMemcpy(d_in.p, keys, n * 8, cudaMemcpyHostToDevice));
    k_marginalize<<<ct_grid(sms, n), CT_THREADS>>>(d_in.as<uint64_t>(), n, m, B, d_out.as<uint64_t>());
    QSV_CUDA(cudaGetLastError());
    QSV_CUDA(cudaMemcpy(out, d_out.p, n * 8, cudaMemcpyDeviceToHost));
    return QSV_OK;
}

extern "C" int qsv_sample_counts(int device, const uint64_t* keys, uint64_t n, uint64_t* out_keys, uint64_t* out_counts,
                                 uint64_t* n_unique) {
    if (!n_unique) QSV_FAIL(QSV_ERR_INVALID, "qsv_sample_counts: NULL argument");
    *n_unique = 0;
    if (n == 0) return QSV_OK;
    if (!keys || !out_keys || !out_counts) QSV_FAIL(QSV_ERR_INVALID, "qsv_sample_counts: NULL argument");
    for (uint64_t i = 0; i < n; ++i)
        if (keys[i] == ~0ull) QSV_FAIL(QSV_ERR_UNSUPPORTED, "qsv_sample_counts: key 2^64-1 is reserved");
    uint64_t table = 1024;
    while (table < 2 * n) table <<= 1;
    int sms = 0;
    QSV_TRY(ct_device("qsv_sample_counts", device, &sms));
    DevBuf d_keys, d_slots, d_counts, d_ok, d_oc, d_n;
    QSV_TRY(d_keys.alloc(n * 8));
    QSV_TRY(d_slots.alloc(table * 8));
    QSV_TRY(d_counts.alloc(table * 8));
    QSV_TRY(d_ok.alloc(n * 8));
    QSV_TRY(d_oc.alloc(n * 8));
    QSV_TRY(d_n.alloc(8));
    QSV_CUDA(cudaMemcpy(d_keys.p, keys, n * 8, cudaMemcpyHostToDevice));
    QSV_CUDA(cudaMemset(d_slots.p, 0, table * 8));
    QSV_CUDA(cudaMemset(d_counts.p, 0, table * 8));
    QSV_CUDA(cudaMemset(d_n.p, 0, 8));
    k_hist_insert<<<ct_grid(sms, n), CT_THREADS>>>(d_keys.as<uint64_t>(), n, d_slots.as<unsigned long long>(),
                                                   d_counts.as<unsigned long long>(), table - 1);
    k_hist_compact<<<ct_grid(sms, table), CT_THREADS>>>(d_slots.as<unsigned long long>(), d_counts.as<unsigned long long>(),
                                                        table, d_ok.as<uint64_t>(), d_oc.as<uint64_t>(),
                                                        d_n.as<unsigned long long>());
    QSV_CUDA(cudaGetLastError());
    unsigned long long nu = 0;
    QSV_CUDA(cudaMemcpy(&nu, d_n.p, 8, cudaMemcpyDeviceToHost));
    std::vector<std::pair<uint64_t, uint64_t>> kv(nu);
    std::vector<uint64_t> hk(nu), hc(nu);
    QSV_CUDA(cudaMemcpy(hk.data(), d_ok.p, nu * 8, cudaMemcpyDeviceToHost));
    QSV_CUDA(cudaMemcpy(hc.data(), d_oc.p, nu * 8, cudaMemcpyDeviceToHost));
    for (uint64_t i = 0; i < nu; ++i) kv[i] = {hk[i], hc[i]};
    std::sort(kv.begin(), kv.end());  // np.unique order (quantum_state.py:292)
    for (uint64_t i = 0; i < nu; ++i) {
        out_keys[i] = kv[i].first;
        out_counts[i] = kv[i].second;
    }
    *n_unique = nu;
    return QSV_OK;
}
