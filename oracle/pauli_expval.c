/*
 * C restatement of crates/accelerate/src/pauli_exp_val.rs (TEST INFRASTRUCTURE --
 * see oracle/__init__.py; never linked into the product).
 *
 * The reference sums with rayon `into_par_iter().map().sum()` for >= 19 qubits
 * (pauli_exp_val.rs:60-64, :120-122); OpenMP `reduction(+)` plays that role here so
 * the CPU baseline keeps the reference's multithreaded speed.  Formulas are the
 * reference's, term for term.
 *
 * Build: make -C oracle     (gcc -O3 -fopenmp -shared -fPIC)
 */
#include <stddef.h>
#include <stdint.h>

/* pauli_exp_val.rs:39-65 */
double oracle_expval_pauli_no_x(const double *data, unsigned num_qubits, uint64_t z_mask) {
    const uint64_t size = (uint64_t)1 << num_qubits;
    double sum = 0.0;
#pragma omp parallel for reduction(+ : sum) schedule(static)
    for (uint64_t i = 0; i < size; ++i) {
        double re = data[2 * i], im = data[2 * i + 1];
        double val = re * re + im * im;
        if (__builtin_popcountll(i & z_mask) & 1) val *= -1.0;
        sum += val;
    }
    return sum;
}

/* pauli_exp_val.rs:70-123 */
double oracle_expval_pauli_with_x(const double *data, unsigned num_qubits, uint64_t z_mask,
                                  uint64_t x_mask, double phase_re, double phase_im, unsigned x_max) {
    const uint64_t mask_u = ~((((uint64_t)1) << (x_max + 1)) - 1);
    const uint64_t mask_l = (((uint64_t)1) << x_max) - 1;
    const uint64_t size = (uint64_t)1 << (num_qubits - 1);
    double sum = 0.0;
#pragma omp parallel for reduction(+ : sum) schedule(static)
    for (uint64_t i = 0; i < size; ++i) {
        uint64_t i0 = ((i << 1) & mask_u) | (i & mask_l);
        uint64_t i1 = i0 ^ x_mask;
        double a0r = data[2 * i0], a0i = data[2 * i0 + 1];
        double a1r = data[2 * i1], a1i = data[2 * i1 + 1];
        /* val_0 = Re(phase * (a1 * conj(a0))) */
        double c0r = a1r * a0r + a1i * a0i, c0i = a1i * a0r - a1r * a0i;
        double val0 = phase_re * c0r - phase_im * c0i;
        /* val_1 = Re(phase * (a0 * conj(a1))) */
        double c1r = a0r * a1r + a0i * a1i, c1i = a0i * a1r - a0r * a1i;
        double val1 = phase_re * c1r - phase_im * c1i;
        double val = val0;
        if (__builtin_popcountll(i0 & z_mask) & 1) val *= -1.0;
        if (__builtin_popcountll(i1 & z_mask) & 1) val -= val1;
        else val += val1;
        sum += val;
    }
    return sum;
}

/* 1- and 2-qubit gate application in place: NOT the reference's algorithm (that is
 * NumPy transpose + dot, restated in statevector_ref.evolve_operator) -- a plain strided
 * loop used only to build large test states quickly. */
void oracle_apply_1q(double *data, unsigned num_qubits, unsigned q, const double *m) {
    const uint64_t half = (uint64_t)1 << (num_qubits - 1);
    const uint64_t bit = (uint64_t)1 << q;
#pragma omp parallel for schedule(static)
    for (uint64_t g = 0; g < half; ++g) {
        uint64_t i0 = ((g >> q) << (q + 1)) | (g & (bit - 1));
        uint64_t i1 = i0 | bit;
        double a0r = data[2 * i0], a0i = data[2 * i0 + 1], a1r = data[2 * i1], a1i = data[2 * i1 + 1];
        data[2 * i0] = m[0] * a0r - m[1] * a0i + m[2] * a1r - m[3] * a1i;
        data[2 * i0 + 1] = m[0] * a0i + m[1] * a0r + m[2] * a1i + m[3] * a1r;
        data[2 * i1] = m[4] * a0r - m[5] * a0i + m[6] * a1r - m[7] * a1i;
        data[2 * i1 + 1] = m[4] * a0i + m[5] * a0r + m[6] * a1i + m[7] * a1r;
    }
}
