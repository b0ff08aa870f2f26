// State lifetime, transfers, counters, timers.  C ABI: include/qsv.h.
#include "qsv_common.cuh"

#include <mutex>

static thread_local std::string g_last_error;

// process-wide view of the per-handle counters: live handles + what retired / reset handles had
static std::mutex g_reg_mutex;
static std::vector<qsv_state*> g_live;
static qsv_counters g_retired;

static void ctr_add(qsv_counters& a, const qsv_counters& b) {
    a.kernel_launches += b.kernel_launches;
    a.state_passes += b.state_passes;
    a.read_passes += b.read_passes;
    a.gates_applied += b.gates_applied;
    a.h2d_bytes += b.h2d_bytes;
    a.d2h_bytes += b.d2h_bytes;
    a.phases += b.phases;
    a.swap_count += b.swap_count;
    a.swap_bytes += b.swap_bytes;
    a.write_passes += b.write_passes;
    a.swap_ns += b.swap_ns;
    a.fused_swaps += b.fused_swaps;
}

void qsv_set_error(const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_last_error = buf;
}

extern "C" const char* qsv_last_error(void) { return g_last_error.c_str(); }
extern "C" int qsv_version(void) { return 100; }

extern "C" int qsv_device_count(int* out) {
    if (!out) QSV_FAIL(QSV_ERR_INVALID, "qsv_device_count: out is NULL");
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
        cudaGetLastError();
        n = 0;
    }
    *out = n;
    return QSV_OK;
}

extern "C" int qsv_device_memory(int device, uint64_t* free_bytes, uint64_t* total_bytes) {
    QSV_CUDA(cudaSetDevice(device));
    size_t f = 0, t = 0;
    QSV_CUDA(cudaMemGetInfo(&f, &t));
    if (free_bytes) *free_bytes = f;
    if (total_bytes) *total_bytes = t;
    return QSV_OK;
}

int qsv_ensure_scratch(qsv_state* s, size_t bytes) {
    if (s->scratch_bytes >= bytes) return QSV_OK;
    if (s->scratch) {
        QSV_CUDA(cudaStreamSynchronize(s->stream));
        QSV_CUDA(cudaFree(s->scratch));
        s->scratch = nullptr;
        s->scratch_bytes = 0;
    }
    size_t want = bytes < (1u << 20) ? (1u << 20) : bytes;
    QSV_CUDA(cudaMalloc(&s->scratch, want));
    s->scratch_bytes = want;
    return QSV_OK;
}

int qsv_ensure_pinned(qsv_state* s, size_t bytes) {
    if (s->pinned_bytes >= bytes) return QSV_OK;
    if (s->pinned) {
        QSV_CUDA(cudaStreamSynchronize(s->stream));
        QSV_CUDA(cudaFreeHost(s->pinned));
        s->pinned = nullptr;
        s->pinned_bytes = 0;
    }
    size_t want = bytes < (1u << 16) ? (1u << 16) : bytes;
    QSV_CUDA(cudaMallocHost(&s->pinned, want));
    s->pinned_bytes = want;
    return QSV_OK;
}

static int env_int(const char* name, int dflt) {
    const char* v = getenv(name);
    return v && *v ? atoi(v) : dflt;
}

void qsv_state_defaults(qsv_state* s) {
    s->tile_bits = env_int("QSV_TILE_BITS", 12);
    s->low_bits = env_int("QSV_LOW_BITS", 4);
    s->tile_threads = env_int("QSV_TILE_THREADS", 256);
    s->use_direct = env_int("QSV_USE_DIRECT", 1);
    s->use_cluster = env_int("QSV_USE_CLUSTER", 1);
    s->use_fused = env_int("QSV_USE_FUSED", 1);
    s->use_direct_store = env_int("QSV_DIRECT_STORE", 1);
    s->use_linear_perm = env_int("QSV_LINEAR_PERM", 1);
    s->use_diag_split = env_int("QSV_DIAG_SPLIT", 1);
    s->lazy_init = env_int("QSV_LAZY_INIT", 1);
    s->use_product_init = env_int("QSV_PRODUCT_INIT", 1);
    s->use_product_fold = env_int("QSV_PRODUCT_FOLD", 1);
    s->use_warp_local = env_int("QSV_WARP_LOCAL", 0);  // measured: no gain (barriers are not the limiter), and its 32-byte store segments cost 20 %
}

int qsv_state_alloc(qsv_state** out, int n_qubits, int n_local, int device, cudaStream_t share_stream) {
    if (!out) QSV_FAIL(QSV_ERR_INVALID, "qsv_create: out is NULL");
    *out = nullptr;
    if (n_qubits < 0) QSV_FAIL(QSV_ERR_INVALID, "qsv_create: n_qubits = %d < 0", n_qubits);
    if (n_qubits >= 64)
        QSV_FAIL(QSV_ERR_OVERFLOW, "The value for num_qubits, %d, is too large and would overflow",
                 n_qubits);
    int ndev = 0;
    qsv_device_count(&ndev);
    if (ndev == 0) QSV_FAIL(QSV_ERR_CUDA, "no CUDA device visible (this engine has no CPU path)");
    if (device < 0 || device >= ndev)
        QSV_FAIL(QSV_ERR_INVALID, "qsv_create: device %d out of range [0, %d)", device, ndev);
    QSV_CUDA(cudaSetDevice(device));
    if (n_local > 40) QSV_FAIL(QSV_ERR_NOMEM, "qsv_create: 2^%d amplitudes do not fit a device", n_local);
    size_t free_b = 0, total_b = 0;
    QSV_CUDA(cudaMemGetInfo(&free_b, &total_b));
    size_t need = (size_t)16 << n_local;
    if (need > free_b)
        QSV_FAIL(QSV_ERR_NOMEM, "qsv_create: state needs %zu bytes, device %d has %zu free", need,
                 device, free_b);
    qsv_state* s = new qsv_state();
    memset(s, 0, sizeof *s);
    s->n = n_qubits;
    s->n_local = n_local;
    s->device = device;
    s->rank = 0;
    s->world = 1;
    for (int q = 0; q < 64; ++q) s->perm[q] = q;
    cudaError_t e = cudaMalloc((void**)&s->amp, need);
    if (e != cudaSuccess) {
        cudaGetLastError();
        delete s;
        QSV_FAIL(QSV_ERR_NOMEM, "qsv_create: cudaMalloc(%zu) failed: %s", need, cudaGetErrorString(e));
    }
    if (share_stream) {
        s->stream = share_stream;
        s->owns_stream = 0;
    } else {
        QSV_CUDA(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
        s->owns_stream = 1;
    }
    QSV_CUDA(cudaEventCreate(&s->ev0));
    QSV_CUDA(cudaEventCreate(&s->ev1));
    QSV_CUDA(cudaEventCreate(&s->tev0));
    QSV_CUDA(cudaEventCreate(&s->tev1));
    cudaDeviceProp prop;
    QSV_CUDA(cudaGetDeviceProperties(&prop, device));
    s->num_sms = prop.multiProcessorCount;
    qsv_state_defaults(s);
    if (!share_stream) {  // scratch copies on a shared stream report through their owner's counters
        std::lock_guard<std::mutex> lk(g_reg_mutex);
        g_live.push_back(s);
    }
    *out = s;
    return QSV_OK;
}

extern "C" int qsv_create(qsv_state** out, int n_qubits, int device) {
    QSV_TRY(qsv_state_alloc(out, n_qubits, n_qubits, device, nullptr));
    return qsv_init_basis(*out, 0);
}

void qsv_dist_free(qsv_state* s);  // qsv_dist.cu
void qsv_dist_collect_times(qsv_state* s);
int qsv_dist_set_option(qsv_state* s, const char* name, int64_t value);

extern "C" int qsv_destroy(qsv_state* s) {
    if (!s) return QSV_OK;
    cudaSetDevice(s->device);
    cudaStreamSynchronize(s->stream);
    {
        std::lock_guard<std::mutex> lk(g_reg_mutex);
        for (size_t i = 0; i < g_live.size(); ++i)
            if (g_live[i] == s) {
                g_live.erase(g_live.begin() + i);
                ctr_add(g_retired, s->ctr);
                break;
            }
    }
    if (s->dist) qsv_dist_free(s);
    if (s->aux) {
        qsv_destroy(s->aux);
        s->aux = nullptr;
    }
    if (s->amp) cudaFree(s->amp);
    if (s->scratch) cudaFree(s->scratch);
    if (s->flush_buf) cudaFree(s->flush_buf);
    if (s->prod_tab) cudaFree(s->prod_tab);
    if (s->pinned) cudaFreeHost(s->pinned);
    cudaEventDestroy(s->ev0);
    cudaEventDestroy(s->ev1);
    cudaEventDestroy(s->tev0);
    cudaEventDestroy(s->tev1);
    if (s->owns_stream) cudaStreamDestroy(s->stream);
    delete s;
    return QSV_OK;
}

extern "C" int qsv_clone(const qsv_state* s, qsv_state** out) {
    if (!s) QSV_FAIL(QSV_ERR_INVALID, "qsv_clone: state is NULL");
    if (s->world != 1) QSV_FAIL(QSV_ERR_UNSUPPORTED, "qsv_clone: sharded states cannot be cloned");
    QSV_TRY(qsv_state_alloc(out, s->n, s->n_local, s->device, nullptr));
    qsv_state* c = *out;
    c->tile_bits = s->tile_bits;
    c->low_bits = s->low_bits;
    c->tile_threads = s->tile_threads;
    c->use_direct = s->use_direct;
    c->use_cluster = s->use_cluster;
    c->use_fused = s->use_fused;
    c->use_direct_store = s->use_direct_store;
    c->use_warp_local = s->use_warp_local;
    c->use_linear_perm = s->use_linear_perm;
    c->use_diag_split = s->use_diag_split;
    c->lazy_init = s->lazy_init;
    c->use_product_init = s->use_product_init;
    c->use_product_fold = s->use_product_fold;
    QSV_TRY(qsv_materialize(const_cast<qsv_state*>(s)));
    c->basis_pending = 0;
    QSV_CUDA(cudaStreamSynchronize(s->stream));
    QSV_CUDA(cudaMemcpyAsync(c->amp, s->amp, (size_t)16 << s->n_local, cudaMemcpyDeviceToDevice,
                             c->stream));
    QSV_CUDA(cudaStreamSynchronize(c->stream));
    return QSV_OK;
}

extern "C" int qsv_copy_state(qsv_state* dst, const qsv_state* src) {
    if (!dst || !src) QSV_FAIL(QSV_ERR_INVALID, "qsv_copy_state: NULL argument");
    if (dst->n != src->n || dst->n_local != src->n_local || dst->device != src->device ||
        dst->world != src->world)
        QSV_FAIL(QSV_ERR_INVALID, "qsv_copy_state: handles differ in size, device or sharding");
    QSV_CUDA(cudaSetDevice(src->device));
    QSV_TRY(qsv_materialize(const_cast<qsv_state*>(src)));
    dst->basis_pending = 0;
    dst->prod_pending = 0;
    QSV_CUDA(cudaStreamSynchronize(src->stream));
    QSV_CUDA(cudaMemcpyAsync(dst->amp, src->amp, (size_t)16 << src->n_local, cudaMemcpyDeviceToDevice,
                             dst->stream));
    memcpy(dst->perm, src->perm, sizeof dst->perm);
    QSV_CUDA(cudaStreamSynchronize(dst->stream));
    return QSV_OK;
}

extern "C" int qsv_num_qubits(const qsv_state* s, int* out) {
    if (!s || !out) QSV_FAIL(QSV_ERR_INVALID, "qsv_num_qubits: NULL argument");
    *out = s->n;
    return QSV_OK;
}

__global__ void k_set_one(double2* amp, uint64_t idx) { amp[idx] = make_double2(1.0, 0.0); }

static int write_basis(qsv_state* s, uint64_t index) {
    QSV_CUDA(cudaMemsetAsync(s->amp, 0, (size_t)16 << s->n_local, s->stream));
    uint64_t owner = s->world > 1 ? index >> s->n_local : 0;
    if ((int)owner == s->rank) {
        uint64_t local = index & (qsv_local_size(s) - 1);
        k_set_one<<<1, 1, 0, s->stream>>>(s->amp, local);
        s->ctr.kernel_launches++;
    }
    s->ctr.write_passes++;
    QSV_CUDA(cudaGetLastError());
    return QSV_OK;
}

int qsv_materialize(qsv_state* s) {
    if (s->prod_pending) return qsv_flush_product(s);  // only after a failed qsv_apply_gates
    if (!s->basis_pending) return QSV_OK;
    s->basis_pending = 0;
    return write_basis(s, s->basis_index);
}

extern "C" int qsv_init_basis(qsv_state* s, uint64_t index) {
    if (!s) QSV_FAIL(QSV_ERR_INVALID, "qsv_init_basis: state is NULL");
    if (s->n < 64 && index >> s->n)
        QSV_FAIL(QSV_ERR_INVALID, "qsv_init_basis: index %llu out of range for %d qubits",
                 (unsigned long long)index, s->n);
    QSV_CUDA(cudaSetDevice(s->device));
    for (int q = 0; q < 64; ++q) s->perm[q] = q;
    s->prod_pending = 0;
    if (s->world == 1 && s->lazy_init) {  // deferred: see qsv_state::basis_pending
        s->basis_pending = 1;
        s->basis_index = index;
        return QSV_OK;
    }
    s->basis_pending = 0;
    return write_basis(s, index);
}

int qsv_dist_canonicalize(qsv_state* s);  // qsv_dist.cu: restore the identity qubit layout

extern "C" int qsv_upload(qsv_state* s, const double* c128, uint64_t offset, uint64_t count) {
    if (!s || (!c128 && count)) QSV_FAIL(QSV_ERR_INVALID, "qsv_upload: NULL argument");
    if (offset + count > qsv_local_size(s))
        QSV_FAIL(QSV_ERR_INVALID, "qsv_upload: [%llu, %llu) exceeds the local slab of %llu amplitudes",
                 (unsigned long long)offset, (unsigned long long)(offset + count),
                 (unsigned long long)qsv_local_size(s));
    QSV_CUDA(cudaSetDevice(s->device));
    if (count == qsv_local_size(s)) s->basis_pending = s->prod_pending = 0;  // a full overwrite needs no zeros first
    QSV_TRY(qsv_materialize(s));
    if (s->world > 1) QSV_TRY(qsv_dist_canonicalize(s));
    QSV_CUDA(cudaMemcpyAsync(s->amp + offset, c128, count * 16, cudaMemcpyHostToDevice, s->stream));
    QSV_CUDA(cudaStreamSynchronize(s->stream));
    s->ctr.h2d_bytes += count * 16;
    return QSV_OK;
}

extern "C" int qsv_download(const qsv_state* cs, double* c128, uint64_t offset, uint64_t count) {
    qsv_state* s = const_cast<qsv_state*>(cs);
    if (!s || (!c128 && count)) QSV_FAIL(QSV_ERR_INVALID, "qsv_download: NULL argument");
    if (offset + count > qsv_local_size(s))
        QSV_FAIL(QSV_ERR_INVALID, "qsv_download: range exceeds the local slab");
    QSV_CUDA(cudaSetDevice(s->device));
    QSV_TRY(qsv_materialize(s));
    if (s->world > 1) QSV_TRY(qsv_dist_canonicalize(s));
    QSV_CUDA(cudaMemcpyAsync(c128, s->amp + offset, count * 16, cudaMemcpyDeviceToHost, s->stream));
    QSV_CUDA(cudaStreamSynchronize(s->stream));
    s->ctr.d2h_bytes += count * 16;
    return QSV_OK;
}

extern "C" int qsv_sync(const qsv_state* s) {
    if (!s) QSV_FAIL(QSV_ERR_INVALID, "qsv_sync: state is NULL");
    QSV_CUDA(cudaSetDevice(s->device));
    QSV_CUDA(cudaStreamSynchronize(s->stream));
    return QSV_OK;
}

extern "C" int qsv_get_counters(const qsv_state* s, qsv_counters* out) {
    if (!s || !out) QSV_FAIL(QSV_ERR_INVALID, "qsv_get_counters: NULL argument");
    if (s->dist) qsv_dist_collect_times(const_cast<qsv_state*>(s));  // folds finished swap timers into ctr.swap_ns
    *out = s->ctr;
    return QSV_OK;
}

extern "C" int qsv_reset_counters(qsv_state* s) {
    if (!s) QSV_FAIL(QSV_ERR_INVALID, "qsv_reset_counters: state is NULL");
    std::lock_guard<std::mutex> lk(g_reg_mutex);
    if (s->owns_stream) ctr_add(g_retired, s->ctr);
    memset(&s->ctr, 0, sizeof s->ctr);
    return QSV_OK;
}

extern "C" int qsv_process_counters(qsv_counters* out) {
    if (!out) QSV_FAIL(QSV_ERR_INVALID, "qsv_process_counters: out is NULL");
    std::lock_guard<std::mutex> lk(g_reg_mutex);
    *out = g_retired;
    for (const qsv_state* s : g_live) ctr_add(*out, s->ctr);
    return QSV_OK;
}

extern "C" int qsv_restore_defaults(qsv_state* s) {
    if (!s) QSV_FAIL(QSV_ERR_INVALID, "qsv_restore_defaults: state is NULL");
    qsv_state_defaults(s);  // a pending lazy |index> stays pending: qsv_materialize does not consult the option
    if (s->aux) qsv_state_defaults(s->aux), s->aux->lazy_init = 0;
    return QSV_OK;
}

extern "C" int qsv_timer_begin(qsv_state* s) {
    if (!s) QSV_FAIL(QSV_ERR_INVALID, "qsv_timer_begin: state is NULL");
    QSV_CUDA(cudaSetDevice(s->device));
    QSV_CUDA(cudaEventRecord(s->ev0, s->stream));
    return QSV_OK;
}

extern "C" int qsv_timer_end(qsv_state* s, float* ms_out) {
    if (!s || !ms_out) QSV_FAIL(QSV_ERR_INVALID, "qsv_timer_end: NULL argument");
    QSV_CUDA(cudaSetDevice(s->device));
    QSV_CUDA(cudaEventRecord(s->ev1, s->stream));
    QSV_CUDA(cudaEventSynchronize(s->ev1));
    QSV_CUDA(cudaEventElapsedTime(ms_out, s->ev0, s->ev1));
    return QSV_OK;
}

extern "C" int qsv_set_option(qsv_state* s, const char* name, int64_t value) {
    if (!s || !name) QSV_FAIL(QSV_ERR_INVALID, "qsv_set_option: NULL argument");
    if (!strcmp(name, "tile_bits")) {
        if (value < 10 || value > 13) QSV_FAIL(QSV_ERR_INVALID, "tile_bits must be in [10, 13]");
        s->tile_bits = (int)value;
    } else if (!strcmp(name, "low_bits")) {
        if (value < 0 || value > 6) QSV_FAIL(QSV_ERR_INVALID, "low_bits must be in [0, 6]");
        s->low_bits = (int)value;
    } else if (!strcmp(name, "tile_threads")) {
        if (value != 128 && value != 256 && value != 512)
            QSV_FAIL(QSV_ERR_INVALID, "tile_threads must be 128, 256 or 512");
        s->tile_threads = (int)value;
    } else if (!strcmp(name, "use_direct")) {
        s->use_direct = value != 0;
    } else if (!strcmp(name, "use_cluster")) {
        s->use_cluster = value != 0;
    } else if (!strcmp(name, "use_fused")) {
        s->use_fused = value != 0;
    } else if (!strcmp(name, "direct_store")) {
        s->use_direct_store = value != 0;
    } else if (!strcmp(name, "warp_local")) {
        s->use_warp_local = value != 0;
    } else if (!strcmp(name, "linear_perm")) {
        s->use_linear_perm = value != 0;
    } else if (!strcmp(name, "diag_split")) {
        s->use_diag_split = value != 0;
    } else if (!strcmp(name, "product_init")) {
        s->use_product_init = value != 0;
    } else if (!strcmp(name, "product_fold")) {
        s->use_product_fold = value != 0;
    } else if (!strcmp(name, "lazy_init")) {
        QSV_TRY(qsv_materialize(s));
        s->lazy_init = value != 0;
    } else if (!strncmp(name, "dist_", 5)) {
        return qsv_dist_set_option(s, name, value);
    } else {
        QSV_FAIL(QSV_ERR_INVALID, "qsv_set_option: unknown option '%s'", name);
    }
    return QSV_OK;
}

__global__ void k_fill(uint4* p, size_t n) {
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    size_t stride = (size_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) p[i] = make_uint4(1, 2, 3, 4);
}

extern "C" int qsv_flush_l2(qsv_state* s) {
    if (!s) QSV_FAIL(QSV_ERR_INVALID, "qsv_flush_l2: state is NULL");
    QSV_CUDA(cudaSetDevice(s->device));
    const size_t bytes = (size_t)256 << 20;  // 256 MiB > 126 MB L2
    if (!s->flush_buf) {
        QSV_CUDA(cudaMalloc(&s->flush_buf, bytes));
        s->flush_bytes = bytes;
    }
    k_fill<<<s->num_sms * 8, 256, 0, s->stream>>>((uint4*)s->flush_buf, bytes / 16);
    QSV_CUDA(cudaGetLastError());
    return QSV_OK;
}

__global__ void k_scale(double2* amp, uint64_t n, double2 c) {
    uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) amp[i] = cmul(amp[i], c);
}

extern "C" int qsv_scale(qsv_state* s, double re, double im) {
    if (!s) QSV_FAIL(QSV_ERR_INVALID, "qsv_scale: state is NULL");
    QSV_CUDA(cudaSetDevice(s->device));
    QSV_TRY(qsv_materialize(s));
    uint64_t n = qsv_local_size(s);
    uint64_t blocks = (n + 255) / 256;
    uint64_t cap = (uint64_t)s->num_sms * 16;
    if (blocks > cap) blocks = cap;
    k_scale<<<(unsigned)blocks, 256, 0, s->stream>>>(s->amp, n, make_double2(re, im));
    QSV_CUDA(cudaGetLastError());
    s->ctr.kernel_launches++;
    s->ctr.state_passes++;
    return QSV_OK;
}

extern "C" int qsv_pack_samples(const uint64_t* indices, uint64_t shots, const int* src_bit,
                                int num_bits, uint8_t* out) {
    if ((!indices && shots) || (!src_bit && num_bits) || (!out && shots && num_bits))
        QSV_FAIL(QSV_ERR_INVALID, "qsv_pack_samples: NULL argument");
    if (num_bits < 0) QSV_FAIL(QSV_ERR_INVALID, "qsv_pack_samples: num_bits < 0");
    int num_bytes = (num_bits + 7) / 8;
    for (uint64_t sidx = 0; sidx < shots; ++sidx) {
        uint64_t v = indices[sidx];
        uint8_t* row = out + sidx * (uint64_t)num_bytes;
        for (int b = 0; b < num_bytes; ++b) row[b] = 0;
        for (int c = 0; c < num_bits; ++c) {
            int sb = src_bit[c];
            if (sb < 0) continue;
            if ((v >> sb) & 1ull) row[num_bytes - 1 - (c >> 3)] |= (uint8_t)(1u << (c & 7));
        }
    }
    return QSV_OK;
}
