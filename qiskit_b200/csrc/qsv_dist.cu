// Sharded states: one process per GPU, NCCL over NVLink (loaded with dlopen so the
// single-GPU path has no NCCL dependency).  Placeholder until the exchange kernels land.
#include "qsv_common.cuh"

#include <complex>


void qsv_dist_free(qsv_state* s) { (void)s; }

int qsv_dist_canonicalize(qsv_state* s) {
    for (int q = 0; q < s->n; ++q)
        if (s->perm[q] != q) QSV_FAIL(QSV_ERR_UNSUPPORTED, "sharded layout restore not implemented");
    return QSV_OK;
}

int qsv_dist_apply(qsv_state* s, std::vector<HGate>& gates, int flags) {
    (void)s; (void)gates; (void)flags;
    QSV_FAIL(QSV_ERR_UNSUPPORTED, "sharded gate application not implemented");
}

int qsv_dist_allreduce_sum(qsv_state* s, double* dev_buf, size_t count) {
    (void)s; (void)dev_buf; (void)count;
    QSV_FAIL(QSV_ERR_UNSUPPORTED, "sharded reductions not implemented");
}

int qsv_dist_localize_xmask(qsv_state* s, uint64_t phys_x_mask) {
    (void)s; (void)phys_x_mask;
    QSV_FAIL(QSV_ERR_UNSUPPORTED, "sharded expval with global X bits not implemented");
}

extern "C" int qsv_dist_unique_id(void* out128) {
    (void)out128;
    QSV_FAIL(QSV_ERR_UNSUPPORTED, "sharded states not implemented");
}

extern "C" int qsv_create_dist(qsv_state** out, int n_qubits, int device, int rank, int world,
                               const void* unique_id128) {
    (void)out; (void)n_qubits; (void)device; (void)rank; (void)world; (void)unique_id128;
    QSV_FAIL(QSV_ERR_UNSUPPORTED, "sharded states not implemented");
}

extern "C" int qsv_dist_info(const qsv_state* s, int* rank, int* world, int* n_local) {
    if (!s) QSV_FAIL(QSV_ERR_INVALID, "qsv_dist_info: state is NULL");
    if (rank) *rank = s->rank;
    if (world) *world = s->world;
    if (n_local) *n_local = s->n_local;
    return QSV_OK;
}
