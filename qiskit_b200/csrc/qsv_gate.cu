// Gate application: Statevector._evolve_operator (statevector.py:925-963) and the
// leaf calls of Statevector._evolve_instruction (:965-1041), re-designed for B200.
//
// The reference makes three full out-of-place copies of the state per gate
// (transpose, zgemm, transpose).  Here a gate stream is partitioned into PASSES:
// one pass reads and writes every amplitude exactly once (32 bytes of HBM traffic
// per amplitude) and applies every gate of the pass on a tile of 2^T amplitudes
// staged in shared memory.  A tile spans T "active" index bits: the lowest L bits
// (contiguous 16*2^L-byte runs, so global accesses stay coalesced) plus the bits
// the pass's gates act on.  Everything is in place: no second state buffer.
#include <algorithm>
#include <cmath>
#include <complex>

#include "qsv_common.cuh"
#include "qsv_fused.cuh"

// ------------------------------------------------------------ device ops --
enum {
    OP_DENSE1 = 0,   // 2x2 complex
    OP_DENSE1R = 1,  // 2x2 real
    OP_DENSE2 = 2,   // 4x4 complex
    OP_DENSE2R = 3,  // 4x4 real
    OP_DIAG = 4,     // phase table over <= DIAG_MAX_BITS tile bits
    OP_MONO = 5,     // generalised permutation on k <= 3 bits: out[r] = coef[r] * in[src[r]]
};

#define DIAG_MAX_BITS 10
#define MAX_TILE_BITS 13

struct DevOp {
    int32_t kind;
    int32_t k;
    int32_t bits[DIAG_MAX_BITS];  // tile-local bit positions (matrix bit j -> bits[j])
    uint32_t data_off;            // offset into the pass data buffer, in double2 units
    uint32_t aux;                 // MONO: src rows packed 3 bits each; bit 24+r set = row r untouched
    uint32_t flags;               // MONO: 1 = all coefficients are exactly 1
    uint32_t pad;
};

struct PassParams {
    int T;          // tile bits
    int nops;
    int pos[MAX_TILE_BITS];  // ascending global bit position of tile bit b
    const DevOp* ops;
    const double2* data;
};

__device__ __forceinline__ uint32_t insert_zero32(uint32_t v, int p) {
    uint32_t low = v & ((1u << p) - 1u);
    return ((v >> p) << (p + 1)) | low;
}

template <int NT>
__device__ __forceinline__ void op_dense1(double2* tile, uint32_t tsize, const DevOp* __restrict__ opp,
                                          const double2* __restrict__ data, bool real) {
    const int p = opp->bits[0];
    const uint32_t doff = opp->data_off;
    const double2 m00 = __ldg(data + doff + 0), m01 = __ldg(data + doff + 1);
    const double2 m10 = __ldg(data + doff + 2), m11 = __ldg(data + doff + 3);
    const uint32_t half = tsize >> 1;
    if (real) {
#pragma unroll 4
        for (uint32_t g = threadIdx.x; g < half; g += NT) {
            uint32_t i0 = insert_zero32(g, p), i1 = i0 | (1u << p);
            double2 a0 = tile[i0], a1 = tile[i1];
            tile[i0] = make_double2(fma(m00.x, a0.x, m01.x * a1.x), fma(m00.x, a0.y, m01.x * a1.y));
            tile[i1] = make_double2(fma(m10.x, a0.x, m11.x * a1.x), fma(m10.x, a0.y, m11.x * a1.y));
        }
    } else {
#pragma unroll 4
        for (uint32_t g = threadIdx.x; g < half; g += NT) {
            uint32_t i0 = insert_zero32(g, p), i1 = i0 | (1u << p);
            double2 a0 = tile[i0], a1 = tile[i1];
            tile[i0] = cfma(m01, a1, cmul(m00, a0));
            tile[i1] = cfma(m11, a1, cmul(m10, a0));
        }
    }
}

template <int NT>
__device__ __forceinline__ void op_dense2(double2* tile, uint32_t tsize, const DevOp* __restrict__ opp,
                                          const double2* __restrict__ data, bool real) {
    const int p0 = opp->bits[0], p1 = opp->bits[1];
    const uint32_t doff = opp->data_off;
    const int lo = min(p0, p1), hi = max(p0, p1);
    const uint32_t o1 = 1u << p0, o2 = 1u << p1;
    const uint32_t quarter = tsize >> 2;
    if (real) {
        double m[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) m[i] = __ldg(&data[doff + i].x);
#pragma unroll 1
        for (uint32_t g = threadIdx.x; g < quarter; g += NT) {
            uint32_t b = insert_zero32(insert_zero32(g, lo), hi);
            double2 a0 = tile[b], a1 = tile[b | o1], a2 = tile[b | o2], a3 = tile[b | o1 | o2];
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                double re = m[4 * r] * a0.x, im = m[4 * r] * a0.y;
                re = fma(m[4 * r + 1], a1.x, re); im = fma(m[4 * r + 1], a1.y, im);
                re = fma(m[4 * r + 2], a2.x, re); im = fma(m[4 * r + 2], a2.y, im);
                re = fma(m[4 * r + 3], a3.x, re); im = fma(m[4 * r + 3], a3.y, im);
                tile[b | ((r & 1) ? o1 : 0u) | ((r & 2) ? o2 : 0u)] = make_double2(re, im);
            }
        }
    } else {
        double2 m[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) m[i] = __ldg(data + doff + i);
#pragma unroll 1
        for (uint32_t g = threadIdx.x; g < quarter; g += NT) {
            uint32_t b = insert_zero32(insert_zero32(g, lo), hi);
            double2 a0 = tile[b], a1 = tile[b | o1], a2 = tile[b | o2], a3 = tile[b | o1 | o2];
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                double2 acc = cmul(m[4 * r], a0);
                acc = cfma(m[4 * r + 1], a1, acc);
                acc = cfma(m[4 * r + 2], a2, acc);
                acc = cfma(m[4 * r + 3], a3, acc);
                tile[b | ((r & 1) ? o1 : 0u) | ((r & 2) ? o2 : 0u)] = acc;
            }
        }
    }
}

template <int NT>
__device__ __forceinline__ void op_diag(double2* tile, uint32_t tsize, const DevOp* __restrict__ opp,
                                        const double2* __restrict__ data) {
    const int d = opp->k;
    const double2* __restrict__ tab = data + opp->data_off;
    for (uint32_t j = threadIdx.x; j < tsize; j += NT) {
        uint32_t idx = 0;
        for (int b = 0; b < d; ++b) idx |= ((j >> opp->bits[b]) & 1u) << b;
        tile[j] = cmul(tile[j], __ldg(tab + idx));
    }
}

template <int NT>
__device__ __forceinline__ void op_mono(double2* tile, uint32_t tsize, const DevOp* __restrict__ opp,
                                        const double2* __restrict__ data) {
    const int k = opp->k;
    const uint32_t dim = 1u << k;
    const uint32_t groups = tsize >> k;
    const int b0 = opp->bits[0], b1 = opp->bits[1], b2 = opp->bits[2];
    const uint32_t aux = opp->aux, doff = opp->data_off;
    // sorted insertion order for the group base
    int s0 = b0, s1 = k > 1 ? b1 : 31, s2 = k > 2 ? b2 : 31;
    if (s0 > s1) { int t = s0; s0 = s1; s1 = t; }
    if (s1 > s2) { int t = s1; s1 = s2; s2 = t; }
    if (s0 > s1) { int t = s0; s0 = s1; s1 = t; }
    const bool unit = opp->flags & 1u;
    for (uint32_t g = threadIdx.x; g < groups; g += NT) {
        uint32_t b = insert_zero32(g, s0);
        if (k > 1) b = insert_zero32(b, s1);
        if (k > 2) b = insert_zero32(b, s2);
        double2 v[8];
#pragma unroll
        for (uint32_t r = 0; r < 8; ++r) {
            if (r < dim) {
                uint32_t off = ((r & 1u) << b0) | (k > 1 ? ((r >> 1) & 1u) << b1 : 0u) |
                               (k > 2 ? ((r >> 2) & 1u) << b2 : 0u);
                v[r] = tile[b | off];
            }
        }
#pragma unroll
        for (uint32_t r = 0; r < 8; ++r) {
            if (r < dim && !((aux >> (24 + r)) & 1u)) {
                uint32_t src = (aux >> (3 * r)) & 7u;
                uint32_t off = ((r & 1u) << b0) | (k > 1 ? ((r >> 1) & 1u) << b1 : 0u) |
                               (k > 2 ? ((r >> 2) & 1u) << b2 : 0u);
                double2 x = v[0];
#pragma unroll
                for (uint32_t c = 1; c < 8; ++c)
                    if (c == src) x = v[c];
                tile[b | off] = unit ? x : cmul(__ldg(data + doff + r), x);
            }
        }
    }
}

// One pass over the state.  Grid = number of tiles; dynamic smem = 16 * 2^T bytes.
template <int NT>
__global__ void __launch_bounds__(NT, 512 / NT) k_tile_pass(double2* __restrict__ amp, const PassParams P) {
    extern __shared__ double2 tile[];
    __shared__ uint64_t hi_off[1 << (MAX_TILE_BITS - 7)];
    const int T = P.T;
    const uint32_t tsize = 1u << T;
    constexpr int LOGNT = NT == 128 ? 7 : (NT == 256 ? 8 : 9);

    // tile base: blockIdx deposited on the non-active bits
    uint64_t base = blockIdx.x;
    for (int b = 0; b < T; ++b) base = insert_zero(base, P.pos[b]);
    // per-thread offset of tile-local index bits [0, LOGNT) and a table for the rest
    uint64_t lo_off = 0;
    for (int b = 0; b < T && b < LOGNT; ++b) lo_off |= (uint64_t)((threadIdx.x >> b) & 1u) << P.pos[b];
    if (T > LOGNT && threadIdx.x < (1u << (T - LOGNT))) {
        uint64_t o = 0;
        for (int b = LOGNT; b < T; ++b) o |= (uint64_t)((threadIdx.x >> (b - LOGNT)) & 1u) << P.pos[b];
        hi_off[threadIdx.x] = o;
    }
    __syncthreads();
    double2* __restrict__ src = amp + base + lo_off;
    if (T > LOGNT) {
        const uint32_t R = 1u << (T - LOGNT);
#pragma unroll 8
        for (uint32_t r = 0; r < R; ++r) tile[threadIdx.x + r * NT] = src[hi_off[r]];
    } else if (threadIdx.x < tsize) {
        tile[threadIdx.x] = src[0];
    }
    __syncthreads();

    for (int i = 0; i < P.nops; ++i) {
        const DevOp* __restrict__ op = P.ops + i;
        switch (op->kind) {
            case OP_DENSE1: op_dense1<NT>(tile, tsize, op, P.data, false); break;
            case OP_DENSE1R: op_dense1<NT>(tile, tsize, op, P.data, true); break;
            case OP_DENSE2: op_dense2<NT>(tile, tsize, op, P.data, false); break;
            case OP_DENSE2R: op_dense2<NT>(tile, tsize, op, P.data, true); break;
            case OP_DIAG: op_diag<NT>(tile, tsize, op, P.data); break;
            case OP_MONO: op_mono<NT>(tile, tsize, op, P.data); break;
        }
        __syncthreads();
    }

    if (T > LOGNT) {
        const uint32_t R = 1u << (T - LOGNT);
#pragma unroll 8
        for (uint32_t r = 0; r < R; ++r) src[hi_off[r]] = tile[threadIdx.x + r * NT];
    } else if (threadIdx.x < tsize) {
        src[0] = tile[threadIdx.x];
    }
}

// ---- cluster pass: the fused-pass kernel ---------------------------------------------------
// k_tile_pass above sweeps the whole tile through shared memory once per gate; ncu shows it
// bound by shared-memory wavefronts and barriers (profiles/r1_ncu_tile.txt).  Here a pass is
// cut into PHASES.  In a phase every thread owns the 2^4 amplitudes that differ in four
// "register bits" of the tile index and applies every gate of the phase that acts inside
// those bits without touching shared memory; diagonal gates act on any bits.  Between
// phases the tile is re-dealt through shared memory (one write + one read per amplitude).
// Shared-memory slots are XOR-swizzled so that both the coalesced global<->shared copy and the
// per-phase gathers are free of bank conflicts (host picks the lane bits accordingly).
#define CL_R 4
#define CL_REGS (1 << CL_R)
enum { CL_DENSE1 = 0, CL_DENSE1R = 1, CL_DENSE2 = 2, CL_DENSE2R = 3, CL_DIAG = 4, CL_MONO = 5, CL_DIAGU = 6 };

struct ClSub {                 // 64 bytes
    uint8_t kind, k;
    uint8_t b[3];              // register-bit indices, ascending (dense / mono)
    uint8_t flags;             // MONO: 1 = all coefficients are exactly 1
    uint8_t nbits;             // DIAG: table index bits
    int8_t bits[10];           // DIAG: tile bit feeding table index bit j
    uint8_t pad0[3];
    uint32_t data_off;         // into the pass data buffer, double2 units
    uint32_t aux;              // MONO: source row of output row r at bits [3r, 3r+3)
    uint16_t regpart[CL_REGS]; // DIAG: table index contribution of register r
    uint32_t pad1[1];
};
static_assert(sizeof(ClSub) == 64, "ClSub layout");

#define CL_PERM_ID 0xFEDCBA9876543210ull
#define CL_NOCOEF 0xffffffffu
struct ClPhase {               // 48 bytes
    int8_t rpos[CL_R];         // tile bit of register bit j (ascending)
    int8_t tpos[9];            // tile bit of thread-index bit j
    uint8_t nsub;
    uint16_t sub_first;
    // generalised permutations inside the register bits that open / close the phase are folded
    // into the shared-memory addresses: register r is READ from slot nibble r of rperm (times
    // data[rcoef + r]) and register q is WRITTEN to slot nibble q of wperm (times data[wcoef + q])
    uint32_t rperm_lo, rperm_hi, wperm_lo, wperm_hi;
    uint32_t rcoef, wcoef;     // CL_NOCOEF = all ones
    uint32_t pad[2];
};
static_assert(sizeof(ClPhase) == 48, "ClPhase layout");

struct ClPass {
    int T, nphases;
    uint64_t index_base;       // sharded states: rank << n_local (bits of diagonal gates above the slab)
    int pos[MAX_TILE_BITS];    // ascending global bit position of tile bit b
    const ClPhase* phases;
    const ClSub* subs;
    const double2* data;
};

__host__ __device__ __forceinline__ uint32_t cl_swz(uint32_t i) {
    return i ^ (((i >> 3) ^ (i >> 6) ^ (i >> 9) ^ (i >> 12)) & 7u);
}

// matrix entry load that the compiler may not hoist or keep live across groups (register budget)
__device__ __forceinline__ double2 cl_ld(const double2* p) {
    double2 r;
    asm volatile("ld.global.nc.v2.f64 {%0, %1}, [%2];" : "=d"(r.x), "=d"(r.y) : "l"(p));
    return r;
}

template <int B>
__device__ __forceinline__ void cl_dense1(double2 (&v)[CL_REGS], const double2* __restrict__ m, bool real) {
    const double2 m00 = __ldg(m), m01 = __ldg(m + 1), m10 = __ldg(m + 2), m11 = __ldg(m + 3);
#pragma unroll
    for (int i = 0; i < CL_REGS; ++i) {
        if (i & (1 << B)) continue;
        const double2 a0 = v[i], a1 = v[i | (1 << B)];
        if (real) {
            v[i] = make_double2(fma(m00.x, a0.x, m01.x * a1.x), fma(m00.x, a0.y, m01.x * a1.y));
            v[i | (1 << B)] = make_double2(fma(m10.x, a0.x, m11.x * a1.x), fma(m10.x, a0.y, m11.x * a1.y));
        } else {
            v[i] = cfma(m01, a1, cmul(m00, a0));
            v[i | (1 << B)] = cfma(m11, a1, cmul(m10, a0));
        }
    }
}

template <int B0, int B1>
__device__ __forceinline__ void cl_dense2(double2 (&v)[CL_REGS], const double2* __restrict__ m, bool real) {
#pragma unroll
    for (int i = 0; i < CL_REGS; ++i) {
        if (i & ((1 << B0) | (1 << B1))) continue;
        const double2 a0 = v[i], a1 = v[i | (1 << B0)], a2 = v[i | (1 << B1)],
                      a3 = v[i | (1 << B0) | (1 << B1)];
        double2 o[4];
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const double2 c0 = __ldg(m + 4 * r), c1 = __ldg(m + 4 * r + 1), c2 = __ldg(m + 4 * r + 2),
                          c3 = __ldg(m + 4 * r + 3);
            if (real) {
                double re = c0.x * a0.x, im = c0.x * a0.y;
                re = fma(c1.x, a1.x, re); im = fma(c1.x, a1.y, im);
                re = fma(c2.x, a2.x, re); im = fma(c2.x, a2.y, im);
                re = fma(c3.x, a3.x, re); im = fma(c3.x, a3.y, im);
                o[r] = make_double2(re, im);
            } else {
                o[r] = cfma(c3, a3, cfma(c2, a2, cfma(c1, a1, cmul(c0, a0))));
            }
        }
        v[i] = o[0];
        v[i | (1 << B0)] = o[1];
        v[i | (1 << B1)] = o[2];
        v[i | (1 << B0) | (1 << B1)] = o[3];
    }
}

// generalised permutation on K register bits: out[r] = coef[r] * in[src[r]]
template <int K, int B0, int B1, int B2>
__device__ __forceinline__ void cl_mono(double2 (&v)[CL_REGS], uint32_t aux, const double2* __restrict__ coef,
                                        bool unit) {
    constexpr int DIM = 1 << K;
    constexpr int MASK = (1 << B0) | (K > 1 ? (1 << B1) : 0) | (K > 2 ? (1 << B2) : 0);
#pragma unroll
    for (int i = 0; i < CL_REGS; ++i) {
        if (i & MASK) continue;
        double2 in[DIM], out[DIM];
#pragma unroll
        for (int r = 0; r < DIM; ++r)
            in[r] = v[i | ((r & 1) << B0) | (K > 1 ? ((r >> 1) & 1) << B1 : 0) | (K > 2 ? ((r >> 2) & 1) << B2 : 0)];
#pragma unroll
        for (int r = 0; r < DIM; ++r) {
            const uint32_t src = (aux >> (3 * r)) & 7u;
            double2 x = in[0];
#pragma unroll
            for (int c = 1; c < DIM; ++c)
                if (src == (uint32_t)c) x = in[c];
            out[r] = unit ? x : cmul(__ldg(coef + r), x);
        }
#pragma unroll
        for (int r = 0; r < DIM; ++r)
            v[i | ((r & 1) << B0) | (K > 1 ? ((r >> 1) & 1) << B1 : 0) | (K > 2 ? ((r >> 2) & 1) << B2 : 0)] = out[r];
    }
}

__device__ __forceinline__ void cl_apply(double2 (&v)[CL_REGS], const ClSub* __restrict__ sp,
                                         const double2* __restrict__ data, uint32_t base, uint64_t tile_base) {
    const uint32_t w0 = __ldg(reinterpret_cast<const uint32_t*>(sp));  // kind, k, b0, b1
    const uint32_t w1 = __ldg(reinterpret_cast<const uint32_t*>(sp) + 1);  // b2, flags, nbits, bits[0]
    const int kind = w0 & 255, k = (w0 >> 8) & 255, b0 = (w0 >> 16) & 255, b1 = (w0 >> 24) & 255;
    const int b2 = w1 & 255;
    const bool unit = (w1 >> 8) & 1;
    const double2* __restrict__ d = data + __ldg(&sp->data_off);
#ifndef CL_NO_D1
    if (kind == CL_DENSE1 || kind == CL_DENSE1R) {
        const bool real = kind == CL_DENSE1R;
        switch (b0) {
            case 0: cl_dense1<0>(v, d, real); break;
            case 1: cl_dense1<1>(v, d, real); break;
            case 2: cl_dense1<2>(v, d, real); break;
            default: cl_dense1<3>(v, d, real); break;
        }
    } else
#endif
#ifndef CL_NO_D2
    if (kind == CL_DENSE2 || kind == CL_DENSE2R) {
        const bool real = kind == CL_DENSE2R;
        switch (b0 * 4 + b1) {
            case 1: cl_dense2<0, 1>(v, d, real); break;
            case 2: cl_dense2<0, 2>(v, d, real); break;
            case 3: cl_dense2<0, 3>(v, d, real); break;
            case 6: cl_dense2<1, 2>(v, d, real); break;
            case 7: cl_dense2<1, 3>(v, d, real); break;
            default: cl_dense2<2, 3>(v, d, real); break;
        }
    } else
#endif
#ifndef CL_NO_DIAG
    if (kind == CL_DIAG) {
        const int nb = (w1 >> 16) & 255;
        uint32_t thr = 0;
        for (int j = 0; j < nb; ++j) {
            const int bj = sp->bits[j];  // tile bit, or 64 + global bit for a bit outside the tile
            const uint32_t bit = bj >= 64 ? (uint32_t)(tile_base >> (bj - 64)) & 1u : (base >> bj) & 1u;
            thr |= bit << j;
        }
#pragma unroll
        for (int r = 0; r < CL_REGS; ++r) v[r] = cmul(v[r], __ldg(d + (thr | sp->regpart[r])));
    } else
#endif
    {
#ifndef CL_NO_MONO
        const uint32_t aux = __ldg(&sp->aux);
        if (k == 1) {
            switch (b0) {
                case 0: cl_mono<1, 0, 0, 0>(v, aux, d, unit); break;
                case 1: cl_mono<1, 1, 0, 0>(v, aux, d, unit); break;
                case 2: cl_mono<1, 2, 0, 0>(v, aux, d, unit); break;
                default: cl_mono<1, 3, 0, 0>(v, aux, d, unit); break;
            }
        } else if (k == 2) {
            switch (b0 * 4 + b1) {
                case 1: cl_mono<2, 0, 1, 0>(v, aux, d, unit); break;
                case 2: cl_mono<2, 0, 2, 0>(v, aux, d, unit); break;
                case 3: cl_mono<2, 0, 3, 0>(v, aux, d, unit); break;
                case 6: cl_mono<2, 1, 2, 0>(v, aux, d, unit); break;
                case 7: cl_mono<2, 1, 3, 0>(v, aux, d, unit); break;
                default: cl_mono<2, 2, 3, 0>(v, aux, d, unit); break;
            }
        } else {
            switch (b0 * 16 + b1 * 4 + b2) {
                case 6: cl_mono<3, 0, 1, 2>(v, aux, d, unit); break;
                case 7: cl_mono<3, 0, 1, 3>(v, aux, d, unit); break;
                case 11: cl_mono<3, 0, 2, 3>(v, aux, d, unit); break;
                default: cl_mono<3, 1, 2, 3>(v, aux, d, unit); break;
            }
        }
#endif
    }
}

// One fused pass.  Tile = 2^T amplitudes, T = log2(NT) + log2(SETS) + CL_R: NT threads, each
// owning SETS groups of 16 amplitudes per phase (processed one after the other, so the register
// budget is that of one group: 254 registers at NT = 128, two CTAs per SM, no spills).
// Grid = number of tiles; dynamic smem = 16 * 2^T bytes.
template <int NT, int SETS>
__global__ void __launch_bounds__(NT, 2) k_cluster_pass(double2* __restrict__ amp, const __grid_constant__ ClPass P) {
    extern __shared__ double2 tile[];
    constexpr int LOGNT = NT == 64 ? 6 : 7;
    constexpr int LOGS = SETS == 1 ? 0 : (SETS == 2 ? 1 : 2);
    constexpr int LOGV = LOGNT + LOGS;        // tile bits spanned by (thread, set)
    constexpr int T = LOGV + CL_R;
    constexpr int NCOPY = CL_REGS * SETS;      // 16-byte elements per thread in the global copies
    const uint32_t t = threadIdx.x;
    const uint32_t st = cl_swz(t);
    uint64_t tile_base = blockIdx.x;
    for (int b = 0; b < T; ++b) tile_base = insert_zero(tile_base, P.pos[b]);
    uint64_t gbase = tile_base;
#pragma unroll
    for (int b = 0; b < LOGNT; ++b) gbase |= (uint64_t)((t >> b) & 1u) << P.pos[b];
    double2* __restrict__ g = amp + gbase;
    tile_base |= P.index_base;

    // global -> shared: element e = t + r * NT of the tile, 16-byte async copies
#pragma unroll
    for (int r = 0; r < NCOPY; ++r) {
        uint64_t off = 0;
#pragma unroll
        for (int j = 0; j < T - LOGNT; ++j)
            if ((r >> j) & 1) off |= 1ull << P.pos[LOGNT + j];
        const uint32_t dst = (uint32_t)__cvta_generic_to_shared(tile + (st ^ cl_swz((uint32_t)r << LOGNT)));
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(dst), "l"(g + off));
    }
    asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;\n" ::: "memory");
    __syncthreads();

    for (int p = 0; p < P.nphases; ++p) {
        const uint4* __restrict__ ph = reinterpret_cast<const uint4*>(P.phases + p);
        const uint4 pw = __ldg(ph);       // bytes 0..3 rpos, 4..12 tpos, 13 nsub, 14..15 sub_first
        const uint4 pq = __ldg(ph + 1);   // rperm, wperm
        const uint4 pc = __ldg(ph + 2);   // rcoef, wcoef
        uint32_t s[CL_R];
#pragma unroll
        for (int j = 0; j < CL_R; ++j) s[j] = cl_swz(1u << ((pw.x >> (8 * j)) & 255));
        const int nsub = (pw.w >> 8) & 255;
        const ClSub* __restrict__ sp = P.subs + (pw.w >> 16);
#pragma unroll 1
        for (int set = 0; set < SETS; ++set) {
            const uint32_t tv = t | ((uint32_t)set << LOGNT);
            uint32_t base = 0;
#pragma unroll
            for (int j = 0; j < LOGV; ++j) {
                const int byte = 4 + j;
                const uint32_t word = byte < 8 ? pw.y : (byte < 12 ? pw.z : pw.w);
                const int tp = (word >> (8 * (byte & 3))) & 255;
                base |= ((tv >> j) & 1u) << tp;
            }
            const uint32_t sb = cl_swz(base);
            double2 v[CL_REGS];
            if (pq.x == (uint32_t)CL_PERM_ID && pq.y == (uint32_t)(CL_PERM_ID >> 32)) {
#pragma unroll
                for (int r = 0; r < CL_REGS; ++r)
                    v[r] = tile[sb ^ ((r & 1) ? s[0] : 0) ^ ((r & 2) ? s[1] : 0) ^ ((r & 4) ? s[2] : 0) ^ ((r & 8) ? s[3] : 0)];
            } else {
#pragma unroll
                for (int r = 0; r < CL_REGS; ++r) {
                    const uint32_t q = ((r < 8 ? pq.x : pq.y) >> (4 * (r & 7))) & 15u;
                    v[r] = tile[sb ^ ((q & 1) ? s[0] : 0) ^ ((q & 2) ? s[1] : 0) ^ ((q & 4) ? s[2] : 0) ^ ((q & 8) ? s[3] : 0)];
                }
            }
            if (pc.x != CL_NOCOEF) {
                const double2* __restrict__ c = P.data + pc.x;
#pragma unroll
                for (int r = 0; r < CL_REGS; ++r) v[r] = cmul(__ldg(c + r), v[r]);
            }
            for (int i = 0; i < nsub; ++i) cl_apply(v, sp + i, P.data, base, tile_base);
            if (pc.y != CL_NOCOEF) {
                const double2* __restrict__ c = P.data + pc.y;
#pragma unroll
                for (int r = 0; r < CL_REGS; ++r) v[r] = cmul(__ldg(c + r), v[r]);
            }
            if (pq.z == (uint32_t)CL_PERM_ID && pq.w == (uint32_t)(CL_PERM_ID >> 32)) {
#pragma unroll
                for (int r = 0; r < CL_REGS; ++r)
                    tile[sb ^ ((r & 1) ? s[0] : 0) ^ ((r & 2) ? s[1] : 0) ^ ((r & 4) ? s[2] : 0) ^ ((r & 8) ? s[3] : 0)] = v[r];
            } else {
#pragma unroll
                for (int r = 0; r < CL_REGS; ++r) {
                    const uint32_t q = ((r < 8 ? pq.z : pq.w) >> (4 * (r & 7))) & 15u;
                    tile[sb ^ ((q & 1) ? s[0] : 0) ^ ((q & 2) ? s[1] : 0) ^ ((q & 4) ? s[2] : 0) ^ ((q & 8) ? s[3] : 0)] = v[r];
                }
            }
        }
        __syncthreads();
    }

#pragma unroll
    for (int r = 0; r < NCOPY; ++r) {
        uint64_t off = 0;
#pragma unroll
        for (int j = 0; j < T - LOGNT; ++j)
            if ((r >> j) & 1) off |= 1ull << P.pos[LOGNT + j];
        g[off] = tile[st ^ cl_swz((uint32_t)r << LOGNT)];
    }
}

// ---- register-resident kernels for a pass that holds a single dense 1- or 2-qubit gate ----
// No shared-memory staging: each thread loads its 2 or 4 amplitudes, multiplies by the matrix
// (kernel parameter, so the entries are constant-bank operands of the DFMAs) and stores.
// Consecutive threads take consecutive groups, so for targets >= bit 5 every warp access is
// one contiguous 512-byte run; for low targets the paired loads of a thread share sectors.
struct Mat2 { double2 m[4]; };
struct Mat4 { double2 m[16]; };

template <bool REAL>
__global__ void __launch_bounds__(256) k_direct1(double2* __restrict__ amp, uint64_t half, int p,
                                                 const __grid_constant__ Mat2 M) {
    const uint64_t bit = 1ull << p;
    const uint64_t stride = (uint64_t)gridDim.x * 256;
#pragma unroll 2
    for (uint64_t g = blockIdx.x * 256ull + threadIdx.x; g < half; g += stride) {
        const uint64_t i0 = insert_zero(g, p), i1 = i0 | bit;
        const double2 a0 = amp[i0], a1 = amp[i1];
        double2 r0, r1;
        if (REAL) {
            r0 = make_double2(fma(M.m[0].x, a0.x, M.m[1].x * a1.x), fma(M.m[0].x, a0.y, M.m[1].x * a1.y));
            r1 = make_double2(fma(M.m[2].x, a0.x, M.m[3].x * a1.x), fma(M.m[2].x, a0.y, M.m[3].x * a1.y));
        } else {
            r0 = cfma(M.m[1], a1, cmul(M.m[0], a0));
            r1 = cfma(M.m[3], a1, cmul(M.m[2], a0));
        }
        amp[i0] = r0;
        amp[i1] = r1;
    }
}

template <bool REAL>
__global__ void __launch_bounds__(256) k_direct2(double2* __restrict__ amp, uint64_t quarter, int lo,
                                                 int hi, uint64_t o1, uint64_t o2,
                                                 const __grid_constant__ Mat4 M) {
    const uint64_t stride = (uint64_t)gridDim.x * 256;
#pragma unroll 2
    for (uint64_t g = blockIdx.x * 256ull + threadIdx.x; g < quarter; g += stride) {
        const uint64_t b = insert_zero(insert_zero(g, lo), hi);
        const double2 a0 = amp[b], a1 = amp[b | o1], a2 = amp[b | o2], a3 = amp[b | o1 | o2];
        double2 r[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            if (REAL) {
                double re = M.m[4 * i].x * a0.x, im = M.m[4 * i].x * a0.y;
                re = fma(M.m[4 * i + 1].x, a1.x, re); im = fma(M.m[4 * i + 1].x, a1.y, im);
                re = fma(M.m[4 * i + 2].x, a2.x, re); im = fma(M.m[4 * i + 2].x, a2.y, im);
                re = fma(M.m[4 * i + 3].x, a3.x, re); im = fma(M.m[4 * i + 3].x, a3.y, im);
                r[i] = make_double2(re, im);
            } else {
                double2 acc = cmul(M.m[4 * i], a0);
                acc = cfma(M.m[4 * i + 1], a1, acc);
                acc = cfma(M.m[4 * i + 2], a2, acc);
                acc = cfma(M.m[4 * i + 3], a3, acc);
                r[i] = acc;
            }
        }
        amp[b] = r[0];
        amp[b | o1] = r[1];
        amp[b | o2] = r[2];
        amp[b | o1 | o2] = r[3];
    }
}

// Generic dense k-qubit gate, 3 <= k <= 12 (user-supplied unitaries, projectors).
// One CTA per group of 2^k amplitudes: gather to smem, one warp per output row
// (coalesced matrix reads, shuffle reduction), scatter back.  In place.
__global__ void __launch_bounds__(256) k_dense_big(double2* __restrict__ amp, int k,
                                                   const int* __restrict__ qsorted,
                                                   const int* __restrict__ qbits,
                                                   const double2* __restrict__ mat) {
    extern __shared__ double2 sm[];
    const uint32_t dim = 1u << k;
    double2* vec = sm;
    double2* outv = sm + dim;
    uint64_t base = blockIdx.x;
    for (int i = 0; i < k; ++i) base = insert_zero(base, qsorted[i]);
    for (uint32_t c = threadIdx.x; c < dim; c += blockDim.x) {
        uint64_t off = 0;
        for (int i = 0; i < k; ++i) off |= (uint64_t)((c >> i) & 1u) << qbits[i];
        vec[c] = amp[base | off];
    }
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
    for (uint32_t r = warp; r < dim; r += nwarps) {
        const double2* row = mat + (size_t)r * dim;
        double re = 0.0, im = 0.0;
        for (uint32_t c = lane; c < dim; c += 32) {
            double2 m = __ldg(row + c), v = vec[c];
            re = fma(m.x, v.x, fma(-m.y, v.y, re));
            im = fma(m.x, v.y, fma(m.y, v.x, im));
        }
        re = warp_sum(re);
        im = warp_sum(im);
        if (lane == 0) outv[r] = make_double2(re, im);
    }
    __syncthreads();
    for (uint32_t c = threadIdx.x; c < dim; c += blockDim.x) {
        uint64_t off = 0;
        for (int i = 0; i < k; ++i) off |= (uint64_t)((c >> i) & 1u) << qbits[i];
        amp[base | off] = outv[c];
    }
}

// ------------------------------------------------------------- host side --
static bool mat_is_diag(const std::vector<cplx>& m, int dim) {
    for (int r = 0; r < dim; ++r)
        for (int c = 0; c < dim; ++c)
            if (r != c && m[(size_t)r * dim + c] != cplx(0.0, 0.0)) return false;
    return true;
}

static bool mat_is_real(const std::vector<cplx>& m) {
    for (const cplx& v : m)
        if (v.imag() != 0.0) return false;
    return true;
}

// exactly one nonzero per row and per column
static bool mat_is_mono(const std::vector<cplx>& m, int dim, int* src) {
    std::vector<int> used(dim, 0);
    for (int r = 0; r < dim; ++r) {
        int nz = -1;
        for (int c = 0; c < dim; ++c)
            if (m[(size_t)r * dim + c] != cplx(0.0, 0.0)) {
                if (nz >= 0) return false;
                nz = c;
            }
        if (nz < 0 || used[nz]) return false;
        used[nz] = 1;
        src[r] = nz;
    }
    return true;
}

static double gate_cost(const HGate& g) {
    const int dim = 1 << g.k;
    if (g.diag) return 1.0;
    int src[1 << 3];
    if (g.k <= 3 && mat_is_mono(g.m, dim, src)) {
        // a pure permutation is folded into addressing (free); one with phases costs a complex
        // multiply per amplitude, more than the shared phase table a diagonal gate joins
        bool unit = true;
        for (int r = 0; r < dim; ++r) unit &= (g.m[(size_t)r * dim + src[r]] == cplx(1.0, 0.0));
        return unit ? 0.25 : 1.5;
    }
    double c = 4.0 * dim;  // DFMA per amplitude of a dense complex block
    if (mat_is_real(g.m)) c *= 0.5;
    return c;
}

// expand gate g (on its qubits) to the qubit list of `host` (superset), as a dense matrix
static std::vector<cplx> expand_to(const HGate& g, const int* hq, int hk) {
    const int hdim = 1 << hk, gdim = 1 << g.k;
    int where[QSV_MAX_GATE_QUBITS];
    for (int j = 0; j < g.k; ++j) {
        where[j] = -1;
        for (int i = 0; i < hk; ++i)
            if (hq[i] == g.q[j]) where[j] = i;
    }
    std::vector<cplx> out((size_t)hdim * hdim, cplx(0.0, 0.0));
    for (int r = 0; r < hdim; ++r)
        for (int c = 0; c < hdim; ++c) {
            int gr = 0, gc = 0, rest_r = r, rest_c = c;
            for (int j = 0; j < g.k; ++j) {
                gr |= ((r >> where[j]) & 1) << j;
                gc |= ((c >> where[j]) & 1) << j;
                rest_r &= ~(1 << where[j]);
                rest_c &= ~(1 << where[j]);
            }
            if (rest_r != rest_c) continue;
            out[(size_t)r * hdim + c] = g.diag ? (gr == gc ? g.m[gr] : cplx(0.0, 0.0)) : g.m[(size_t)gr * gdim + gc];
        }
    return out;
}

static void normalise_diag(HGate& g) {
    if (g.diag) return;
    const int dim = 1 << g.k;
    if (!mat_is_diag(g.m, dim)) return;
    std::vector<cplx> d(dim);
    for (int i = 0; i < dim; ++i) d[i] = g.m[(size_t)i * dim + i];
    g.m.swap(d);
    g.diag = true;
}

static bool subset_of(const HGate& a, const HGate& b) {
    for (int j = 0; j < a.k; ++j) {
        bool f = false;
        for (int i = 0; i < b.k; ++i) f |= (b.q[i] == a.q[j]);
        if (!f) return false;
    }
    return true;
}

// later * earlier on the qubits of `host` (the wider of the two)
static HGate merge_gates(const HGate& earlier, const HGate& later, const HGate& host) {
    HGate out;
    out.k = host.k;
    memcpy(out.q, host.q, sizeof out.q);
    const int dim = 1 << host.k;
    if (earlier.diag && later.diag) {
        out.diag = true;
        out.m.assign(dim, cplx(1.0, 0.0));
        for (int i = 0; i < dim; ++i) {
            for (const HGate* g : {&earlier, &later}) {
                int gi = 0;
                for (int j = 0; j < g->k; ++j)
                    for (int t = 0; t < host.k; ++t)
                        if (host.q[t] == g->q[j]) gi |= ((i >> t) & 1) << j;
                out.m[i] *= g->m[gi];
            }
        }
        return out;
    }
    std::vector<cplx> a = expand_to(earlier, host.q, host.k);
    std::vector<cplx> b = expand_to(later, host.q, host.k);
    out.diag = false;
    out.m.assign((size_t)dim * dim, cplx(0.0, 0.0));
    for (int r = 0; r < dim; ++r)
        for (int t = 0; t < dim; ++t) {
            cplx v = b[(size_t)r * dim + t];
            if (v == cplx(0.0, 0.0)) continue;
            for (int c = 0; c < dim; ++c) out.m[(size_t)r * dim + c] += v * a[(size_t)t * dim + c];
        }
    normalise_diag(out);
    return out;
}

// Peephole fusion: fold a gate into the previous gate on its qubits (or the
// previous gate into it) when one's qubits contain the other's and the merged
// block is not more expensive than the two apart.
static void fuse_gates(std::vector<HGate>& gates, int nbits) {
    std::vector<int> last(nbits, -1);  // index into `out` of the last gate on each bit
    std::vector<HGate> out;
    std::vector<char> dead;
    for (HGate& g : gates) {
        int cand = -2;
        bool same = true;
        for (int j = 0; j < g.k; ++j) {
            int l = last[g.q[j]];
            if (cand == -2) cand = l;
            else if (l != cand) same = false;
        }
        bool merged = false;
        if (g.k > 0 && same && cand >= 0 && !dead[cand] && g.k <= 2 && out[cand].k <= 2) {
            HGate& h = out[cand];
            if (subset_of(g, h)) {
                // h is the most recent gate on all of g's qubits: g folds into h in place
                HGate mg = merge_gates(h, g, h);
                if (gate_cost(mg) <= gate_cost(h) + gate_cost(g)) {
                    h = mg;
                    merged = true;
                }
            }
        }
        if (!merged && g.k > 0 && g.k <= 2) {
            // every earlier gate that sits entirely inside g's qubits and is the last gate on
            // all of its own qubits can be pulled forward into g
            for (int j = 0; j < g.k; ++j) {
                int l = last[g.q[j]];
                if (l < 0 || dead[l]) continue;
                HGate& h = out[l];
                if (h.k > g.k || !subset_of(h, g)) continue;
                bool is_last = true;
                for (int t = 0; t < h.k; ++t) is_last &= (last[h.q[t]] == l);
                if (!is_last) continue;
                HGate mg = merge_gates(h, g, g);
                if (gate_cost(mg) <= gate_cost(h) + gate_cost(g)) {
                    g = mg;
                    dead[l] = 1;
                    for (int t = 0; t < h.k; ++t) last[h.q[t]] = -1;
                }
            }
        }
        if (!merged) {
            out.push_back(g);
            dead.push_back(0);
            for (int j = 0; j < g.k; ++j) last[g.q[j]] = (int)out.size() - 1;
        }
    }
    gates.clear();
    for (size_t i = 0; i < out.size(); ++i)
        if (!dead[i]) gates.push_back(std::move(out[i]));
}

// A dense complex one-qubit unitary costs 8 DFMA per amplitude; written as D2 . R . D1 with R real
// (|U00| -|U10| ; |U10| |U00|) and D1, D2 diagonal phases it costs 4, and the diagonals join the
// phase tables of the pass (shared by all register bits of a phase, merged with neighbouring RZ /
// RZZ / CP gates).  RX = S^dagger RY S is the typical case (Trotter circuits).  Only exact
// reconstructions are taken (unitary input, both |U00| and |U10| well away from 0).
// R * diag(1, -i) with R real: first column real, second column purely imaginary (FP_D1RY)
static bool mat_is_real_times_sdg(const std::vector<cplx>& m) {
    return m.size() == 4 && m[0].imag() == 0.0 && m[2].imag() == 0.0 && m[1].real() == 0.0 && m[3].real() == 0.0;
}

static void split_complex_1q(std::vector<HGate>& gates) {
    static const bool enabled = !getenv("QSV_NO_ZYZ");
    if (!enabled || gates.size() < 2) return;
    std::vector<HGate> out;
    out.reserve(gates.size() + 8);
    for (HGate& g : gates) {
        bool done = false;
        if (g.k == 1 && !g.diag && g.m.size() == 4 && !mat_is_real(g.m) && !mat_is_real_times_sdg(g.m)) {
            const cplx u00 = g.m[0], u01 = g.m[1], u10 = g.m[2], u11 = g.m[3];
            const double c = std::abs(u00), sn = std::abs(u10), c2 = std::abs(u11);
            if (c > 1e-3 && sn > 1e-3 && c2 > 1e-3) {
                const cplx q0 = u00 / c, q1 = u10 / sn, p1 = (u11 / c2) * std::conj(q1);
                // reconstruction: (q0 c, -q0 s p1 ; q1 s, q1 c p1)
                const double err = std::max(std::max(std::abs(q0 * c - u00), std::abs(-q0 * sn * p1 - u01)),
                                            std::max(std::abs(q1 * sn - u10), std::abs(q1 * c * p1 - u11)));
                if (err < 4e-16) {
                    HGate d1, r, d2;
                    d1.k = r.k = d2.k = 1;
                    memset(d1.q, 0, sizeof d1.q);
                    memset(r.q, 0, sizeof r.q);
                    memset(d2.q, 0, sizeof d2.q);
                    d1.q[0] = r.q[0] = d2.q[0] = g.q[0];
                    d1.diag = d2.diag = true;
                    r.diag = false;
                    d1.m = {cplx(1.0, 0.0), p1};
                    r.m = {cplx(c, 0.0), cplx(-sn, 0.0), cplx(sn, 0.0), cplx(c, 0.0)};
                    d2.m = {q0, q1};
                    out.push_back(std::move(d1));
                    out.push_back(std::move(r));
                    out.push_back(std::move(d2));
                    done = true;
                }
            }
        }
        if (!done) out.push_back(std::move(g));
    }
    gates.swap(out);
}

struct PassPlan {
    std::vector<int> gate_ids;
    uint64_t qmask;
};

static int popc64(uint64_t v) { return __builtin_popcountll(v); }

// Greedy partition into passes.  A gate may join the open pass if its qubits are
// not blocked by an earlier gate that was left for later, and the pass still
// fits in T bits together with the forced low bits.
static void plan_passes(const std::vector<HGate>& gates, int n_local, int T, int L, bool no_fusion,
                        bool free_diag, std::vector<PassPlan>& passes) {
    const size_t G = gates.size();
    std::vector<char> done(G, 0);
    size_t first = 0;
    const uint64_t lowmask = L >= 64 ? ~0ull : ((1ull << L) - 1);
    const int cap = std::min(T, n_local);
    while (first < G) {
        while (first < G && done[first]) ++first;
        if (first >= G) break;
        PassPlan p;
        p.qmask = 0;
        uint64_t blocked = 0;
        const uint64_t all = n_local >= 64 ? ~0ull : ((1ull << n_local) - 1);
        const size_t window = 4096;
        for (size_t i = first; i < G && i < first + window; ++i) {
            if (done[i]) continue;
            uint64_t gm = 0;
            for (int j = 0; j < gates[i].k; ++j) gm |= 1ull << gates[i].q[j];
            bool ok = !(gm & blocked);
            // the cluster kernel reads the bits of a diagonal gate from the tile's position when
            // they are not tile bits: diagonal gates never claim tile bits
            const uint64_t tm = (free_diag && gates[i].diag) ? 0ull : gm;
            if (ok) {
                uint64_t u = p.qmask | tm;
                int need = popc64((u | lowmask) & all);
                if (need > cap) ok = false;
                if (no_fusion && !p.gate_ids.empty()) ok = false;
            }
            if (ok) {
                p.qmask |= tm;
                p.gate_ids.push_back((int)i);
                done[i] = 1;
            } else {
                blocked |= gm;
                if ((blocked & all) == all) break;
            }
            if (no_fusion && !p.gate_ids.empty()) break;
        }
        passes.push_back(std::move(p));
    }
}

static int launch_tile_pass(qsv_state* s, const PassParams& P, uint64_t ntiles) {
    const size_t smem = (size_t)16 << P.T;
    if (ntiles > 0x7fffffffull) QSV_FAIL(QSV_ERR_UNSUPPORTED, "too many tiles for one launch");
    switch (s->tile_threads) {
        case 128:
            QSV_CUDA(cudaFuncSetAttribute(k_tile_pass<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            k_tile_pass<128><<<(unsigned)ntiles, 128, smem, s->stream>>>(s->amp, P);
            break;
        case 512:
            QSV_CUDA(cudaFuncSetAttribute(k_tile_pass<512>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            k_tile_pass<512><<<(unsigned)ntiles, 512, smem, s->stream>>>(s->amp, P);
            break;
        default:
            QSV_CUDA(cudaFuncSetAttribute(k_tile_pass<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            k_tile_pass<256><<<(unsigned)ntiles, 256, smem, s->stream>>>(s->amp, P);
            break;
    }
    QSV_CUDA(cudaGetLastError());
    s->ctr.kernel_launches++;
    s->ctr.state_passes++;
    return QSV_OK;
}

static int apply_dense_big(qsv_state* s, const HGate& g) {
    QSV_RANGE("qsv:dense_big");
    const int k = g.k;
    const size_t dim = (size_t)1 << k;
    // scratch layout: [matrix dim*dim double2][qsorted k ints][qbits k ints]
    size_t mat_bytes = dim * dim * 16;
    size_t bytes = mat_bytes + 2 * QSV_MAX_GATE_QUBITS * sizeof(int);
    QSV_TRY(qsv_ensure_scratch(s, bytes));
    QSV_TRY(qsv_ensure_pinned(s, bytes));
    QSV_CUDA(cudaStreamSynchronize(s->stream));  // pinned staging reuse
    char* h = (char*)s->pinned;
    std::vector<cplx> dense;
    const cplx* src = g.m.data();
    if (g.diag) {
        dense.assign(dim * dim, cplx(0.0, 0.0));
        for (size_t i = 0; i < dim; ++i) dense[i * dim + i] = g.m[i];
        src = dense.data();
    }
    memcpy(h, src, mat_bytes);
    int* hs = (int*)(h + mat_bytes);
    int* hb = hs + QSV_MAX_GATE_QUBITS;
    for (int i = 0; i < k; ++i) hs[i] = hb[i] = g.q[i];
    std::sort(hs, hs + k);
    QSV_CUDA(cudaMemcpyAsync(s->scratch, h, bytes, cudaMemcpyHostToDevice, s->stream));
    const size_t smem = 2 * dim * 16;
    QSV_CUDA(cudaFuncSetAttribute(k_dense_big, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    uint64_t groups = 1ull << (s->n_local - k);
    if (groups > 0x7fffffffull) QSV_FAIL(QSV_ERR_UNSUPPORTED, "too many groups for one launch");
    const int* ds = (const int*)((char*)s->scratch + mat_bytes);
    k_dense_big<<<(unsigned)groups, 256, smem, s->stream>>>(s->amp, k, ds, ds + QSV_MAX_GATE_QUBITS,
                                                             (const double2*)s->scratch);
    QSV_CUDA(cudaGetLastError());
    s->ctr.kernel_launches++;
    s->ctr.state_passes++;
    return QSV_OK;
}

// Build the device ops of one pass into the staging area (no launch).
static int build_pass(qsv_state* s, const std::vector<HGate>& gates, const PassPlan& plan, int T,
                      std::vector<char>& stage, PassParams& P, size_t& ops_off, size_t& data_off) {
    const int n = s->n_local;
    const int cap = std::min(T, n);
    uint64_t active = plan.qmask;
    for (int b = 0; b < n && popc64(active) < cap; ++b) active |= 1ull << b;  // low bits first
    memset(&P, 0, sizeof P);
    P.T = popc64(active);
    int local_of[64];
    for (int b = 0, t = 0; b < 64; ++b) {
        local_of[b] = -1;
        if ((active >> b) & 1ull) {
            P.pos[t] = b;
            local_of[b] = t++;
        }
    }
    std::vector<DevOp> ops;
    std::vector<double2> data;
    size_t gi = 0;
    const std::vector<int>& ids = plan.gate_ids;
    while (gi < ids.size()) {
        const HGate& g = gates[ids[gi]];
        DevOp op;
        memset(&op, 0, sizeof op);
        if (g.diag) {
            // runs of consecutive diagonal gates collapse into one phase table
            uint64_t um = 0;
            size_t gj = gi;
            while (gj < ids.size() && gates[ids[gj]].diag) {
                uint64_t gm = 0;
                for (int j = 0; j < gates[ids[gj]].k; ++j) gm |= 1ull << gates[ids[gj]].q[j];
                if (popc64(um | gm) > DIAG_MAX_BITS) break;
                um |= gm;
                ++gj;
            }
            if (gj == gi)
                QSV_FAIL(QSV_ERR_UNSUPPORTED, "diagonal gate wider than %d qubits", DIAG_MAX_BITS);
            int d = 0;
            int bit_of[64];
            for (int b = 0; b < 64; ++b)
                if ((um >> b) & 1ull) {
                    bit_of[b] = d;
                    op.bits[d++] = local_of[b];
                }
            op.kind = OP_DIAG;
            op.k = d;
            op.data_off = (uint32_t)data.size();
            std::vector<cplx> tab((size_t)1 << d, cplx(1.0, 0.0));
            for (size_t t = gi; t < gj; ++t) {
                const HGate& h = gates[ids[t]];
                for (size_t i = 0; i < tab.size(); ++i) {
                    int hi = 0;
                    for (int j = 0; j < h.k; ++j) hi |= (int)((i >> bit_of[h.q[j]]) & 1) << j;
                    tab[i] *= h.m[hi];
                }
            }
            for (const cplx& v : tab) data.push_back(make_double2(v.real(), v.imag()));
            ops.push_back(op);
            gi = gj;
            continue;
        }
        const int dim = 1 << g.k;
        int srcrow[8];
        if (g.k <= 3 && mat_is_mono(g.m, dim, srcrow)) {
            op.kind = OP_MONO;
            op.k = g.k;
            for (int j = 0; j < g.k; ++j) op.bits[j] = local_of[g.q[j]];
            op.data_off = (uint32_t)data.size();
            bool unit = true;
            for (int r = 0; r < dim; ++r) {
                cplx c = g.m[(size_t)r * dim + srcrow[r]];
                unit &= (c == cplx(1.0, 0.0));
                op.aux |= (uint32_t)srcrow[r] << (3 * r);
                if (srcrow[r] == r && c == cplx(1.0, 0.0)) op.aux |= 1u << (24 + r);
                data.push_back(make_double2(c.real(), c.imag()));
            }
            op.flags = unit ? 1u : 0u;
            ops.push_back(op);
        } else if (g.k == 1 || g.k == 2) {
            bool real = mat_is_real(g.m);
            op.kind = g.k == 1 ? (real ? OP_DENSE1R : OP_DENSE1) : (real ? OP_DENSE2R : OP_DENSE2);
            op.k = g.k;
            for (int j = 0; j < g.k; ++j) op.bits[j] = local_of[g.q[j]];
            op.data_off = (uint32_t)data.size();
            for (const cplx& v : g.m) data.push_back(make_double2(v.real(), v.imag()));
            ops.push_back(op);
        } else {
            QSV_FAIL(QSV_ERR_UNSUPPORTED, "internal: dense %d-qubit gate reached the tile path", g.k);
        }
        ++gi;
    }
    const size_t ops_bytes = ops.size() * sizeof(DevOp);
    const size_t data_bytes = data.size() * sizeof(double2);
    ops_off = (stage.size() + 15) & ~(size_t)15;
    data_off = ops_off + ops_bytes;
    stage.resize(data_off + data_bytes);
    memcpy(stage.data() + ops_off, ops.data(), ops_bytes);
    memcpy(stage.data() + data_off, data.data(), data_bytes);
    P.nops = (int)ops.size();
    return QSV_OK;
}

// ---- cluster pass: host side ----------------------------------------------------------------
struct TileOp {
    int kind;                  // CL_*
    int k;                     // qubits (dense / mono) or table bits (diag)
    int bits[DIAG_MAX_BITS];   // tile-local bit of matrix / table index bit j
    std::vector<cplx> m;       // dense: 2^k x 2^k; mono: coef[2^k]; diag: table[2^k]
    int src[8];                // mono
    bool unit;
};

// classify the gates of one pass into tile ops, pass order kept: consecutive diagonal gates share one table
static int make_tile_ops_runs(const std::vector<HGate>& gates, const PassPlan& plan, const int* local_of,
                         std::vector<TileOp>& ops) {
    const std::vector<int>& ids = plan.gate_ids;
    size_t gi = 0;
    while (gi < ids.size()) {
        const HGate& g = gates[ids[gi]];
        TileOp op;
        memset(op.bits, 0, sizeof op.bits);
        memset(op.src, 0, sizeof op.src);
        op.unit = false;
        if (g.diag) {
            uint64_t um = 0;
            size_t gj = gi;
            while (gj < ids.size() && gates[ids[gj]].diag) {
                uint64_t gm = 0;
                for (int j = 0; j < gates[ids[gj]].k; ++j) gm |= 1ull << gates[ids[gj]].q[j];
                if (popc64(um | gm) > DIAG_MAX_BITS) break;
                um |= gm;
                ++gj;
            }
            if (gj == gi) QSV_FAIL(QSV_ERR_UNSUPPORTED, "diagonal gate wider than %d qubits", DIAG_MAX_BITS);
            int d = 0;
            int bit_of[64];
            for (int b = 0; b < 64; ++b)
                if ((um >> b) & 1ull) {
                    bit_of[b] = d;
                    op.bits[d++] = local_of[b] >= 0 ? local_of[b] : 64 + b;  // 64 + global bit: outside the tile
                }
            op.kind = CL_DIAG;
            op.k = d;
            op.m.assign((size_t)1 << d, cplx(1.0, 0.0));
            for (size_t t = gi; t < gj; ++t) {
                const HGate& h = gates[ids[t]];
                for (size_t i = 0; i < op.m.size(); ++i) {
                    int hi = 0;
                    for (int j = 0; j < h.k; ++j) hi |= (int)((i >> bit_of[h.q[j]]) & 1) << j;
                    op.m[i] *= h.m[hi];
                }
            }
            ops.push_back(std::move(op));
            gi = gj;
            continue;
        }
        const int dim = 1 << g.k;
        for (int j = 0; j < g.k && j < DIAG_MAX_BITS; ++j) op.bits[j] = local_of[g.q[j]];
        op.k = g.k;
        if (g.k <= 3 && mat_is_mono(g.m, dim, op.src)) {
            op.kind = CL_MONO;
            op.unit = true;
            op.m.resize(dim);
            for (int r = 0; r < dim; ++r) {
                op.m[r] = g.m[(size_t)r * dim + op.src[r]];
                op.unit &= (op.m[r] == cplx(1.0, 0.0));
            }
        } else if (g.k == 1 || g.k == 2) {
            const bool real = mat_is_real(g.m);
            op.kind = g.k == 1 ? (real ? CL_DENSE1R : CL_DENSE1) : (real ? CL_DENSE2R : CL_DENSE2);
            op.m = g.m;
        } else {
            QSV_FAIL(QSV_ERR_UNSUPPORTED, "internal: dense %d-qubit gate reached the tile path", g.k);
        }
        ops.push_back(std::move(op));
        ++gi;
    }
    return QSV_OK;
}

// classify the gates of one pass into tile ops.
//
// Diagonal gates commute with each other and with every gate on other qubits, so each one may
// run anywhere between the non-diagonal gates that share a qubit with it.  The gates are collected
// at as few points of the pass as possible and merged there into as few phase tables of <= DIAG_MAX_BITS index bits as a greedy cover finds:
// a table grows by the bit that completes the most not-yet-covered gates.  (A QFT pass has a
// controlled-phase ladder from each of its 8 targets to all lower qubits: consecutive-run merging
// gave 29 tables, this gives 7 in-tile tables + the 8 x 21 cross pairs in 5-by-5 tables.)
static int make_tile_ops_regrouped(const std::vector<HGate>& gates, const PassPlan& plan, const int* local_of,
                                   std::vector<TileOp>& ops, bool stab) {
    const std::vector<int>& ids = plan.gate_ids;
    const size_t N = ids.size();
    std::vector<uint64_t> qm(N, 0);
    for (size_t i = 0; i < N; ++i)
        for (int j = 0; j < gates[ids[i]].k; ++j) qm[i] |= 1ull << gates[ids[i]].q[j];
    // slot s = "just before gate s" (s = N: end of the pass).  A diagonal gate may run in any slot
    // after the last earlier and up to the first later non-diagonal gate that shares a qubit with it.
    // Fewest slots that stab every interval (earliest-deadline sweep) -> largest groups to merge.
    std::vector<std::vector<size_t>> pend(N + 1);
    std::vector<size_t> lo(N, 0), hi(N, N), diag_ids;
    for (size_t i = 0; i < N; ++i) {
        if (!gates[ids[i]].diag) continue;
        if (popc64(qm[i]) > DIAG_MAX_BITS) QSV_FAIL(QSV_ERR_UNSUPPORTED, "diagonal gate wider than %d qubits", DIAG_MAX_BITS);
        for (size_t j = i + 1; j < N; ++j)
            if (!gates[ids[j]].diag && (qm[j] & qm[i])) {
                hi[i] = j;
                break;
            }
        for (size_t j = i; j-- > 0;)
            if (!gates[ids[j]].diag && (qm[j] & qm[i])) {
                lo[i] = j + 1;
                break;
            }
        diag_ids.push_back(i);
    }
    std::sort(diag_ids.begin(), diag_ids.end(), [&](size_t a, size_t b) { return hi[a] != hi[b] ? hi[a] < hi[b] : a < b; });
    {
        std::vector<char> placed(N, 0);
        for (size_t a : diag_ids) {
            if (placed[a]) continue;
            const size_t slot = hi[a];
            for (size_t b : diag_ids)
                if (!placed[b] && (stab ? lo[b] <= slot : hi[b] == slot)) {  // hi[b] >= slot: diag_ids is sorted by hi
                    placed[b] = 1;
                    pend[slot].push_back(b);
                }
        }
    }
    auto emit_table = [&](uint64_t um, const std::vector<size_t>& members) {
        TileOp op;
        memset(op.bits, 0, sizeof op.bits);
        memset(op.src, 0, sizeof op.src);
        op.unit = false;
        int d = 0;
        int bit_of[64];
        for (int bb = 0; bb < 64; ++bb)
            if ((um >> bb) & 1ull) {
                bit_of[bb] = d;
                op.bits[d++] = local_of[bb] >= 0 ? local_of[bb] : 64 + bb;  // 64 + global bit: outside the tile
            }
        op.kind = CL_DIAG;
        op.k = d;
        op.m.assign((size_t)1 << d, cplx(1.0, 0.0));
        for (size_t t : members) {
            const HGate& h = gates[ids[t]];
            for (size_t i = 0; i < op.m.size(); ++i) {
                int hi = 0;
                for (int j = 0; j < h.k; ++j) hi |= (int)((i >> bit_of[h.q[j]]) & 1) << j;
                op.m[i] *= h.m[hi];
            }
        }
        ops.push_back(std::move(op));
    };
    auto emit_group = [&](std::vector<size_t>& rest) {
        while (!rest.empty()) {
            uint64_t um = qm[rest[0]];
            while (popc64(um) < DIAG_MAX_BITS) {
                uint64_t cand = 0;
                for (size_t t : rest) cand |= qm[t] & ~um;
                if (!cand) break;
                int best = -1, best_gain = 0;
                for (int bb = 0; bb < 64; ++bb) {
                    if (!((cand >> bb) & 1ull)) continue;
                    const uint64_t with = um | (1ull << bb);
                    int gain = 0;
                    for (size_t t : rest) gain += (qm[t] & ~um) && !(qm[t] & ~with);
                    if (gain > best_gain) best_gain = gain, best = bb;
                }
                if (best >= 0) {
                    um |= 1ull << best;
                    continue;
                }
                // no single bit completes a gate: take the gate that misses the fewest bits, if it fits
                size_t pick = rest.size();
                int miss = 65;
                for (size_t x = 0; x < rest.size(); ++x) {
                    const int m = popc64(qm[rest[x]] & ~um);
                    if (m > 0 && m < miss) miss = m, pick = x;
                }
                if (pick == rest.size() || popc64(um) + miss > DIAG_MAX_BITS) break;
                um |= qm[rest[pick]];
            }
            std::vector<size_t> members, left;
            for (size_t t : rest) (qm[t] & ~um ? left : members).push_back(t);
            emit_table(um, members);
            rest.swap(left);
        }
    };
    for (size_t gi = 0; gi <= N; ++gi) {
        if (!pend[gi].empty()) emit_group(pend[gi]);
        if (gi == N || gates[ids[gi]].diag) continue;
        const HGate& g = gates[ids[gi]];
        TileOp op;
        memset(op.bits, 0, sizeof op.bits);
        memset(op.src, 0, sizeof op.src);
        op.unit = false;
        const int dim = 1 << g.k;
        for (int j = 0; j < g.k && j < DIAG_MAX_BITS; ++j) op.bits[j] = local_of[g.q[j]];
        op.k = g.k;
        if (g.k <= 3 && mat_is_mono(g.m, dim, op.src)) {
            op.kind = CL_MONO;
            op.unit = true;
            op.m.resize(dim);
            for (int r = 0; r < dim; ++r) {
                op.m[r] = g.m[(size_t)r * dim + op.src[r]];
                op.unit &= (op.m[r] == cplx(1.0, 0.0));
            }
        } else if (g.k == 1 || g.k == 2) {
            const bool real = mat_is_real(g.m);
            op.kind = g.k == 1 ? (real ? CL_DENSE1R : CL_DENSE1) : (real ? CL_DENSE2R : CL_DENSE2);
            op.m = g.m;
        } else {
            QSV_FAIL(QSV_ERR_UNSUPPORTED, "internal: dense %d-qubit gate reached the tile path", g.k);
        }
        ops.push_back(std::move(op));
    }
    return QSV_OK;
}

// a regrouped order is used only when it needs fewer phase tables than the pass order (a tie keeps
// the pass order, whose tables line up with the gate layers: EfficientSU2's RZ layers stay as they were)
static int make_tile_ops(const std::vector<HGate>& gates, const PassPlan& plan, const int* local_of,
                         std::vector<TileOp>& ops) {
    std::vector<TileOp> runs, late, stabbed;
    QSV_TRY(make_tile_ops_runs(gates, plan, local_of, runs));
    QSV_TRY(make_tile_ops_regrouped(gates, plan, local_of, late, false));    // every gate at its latest slot
    QSV_TRY(make_tile_ops_regrouped(gates, plan, local_of, stabbed, true));  // fewest slots
    auto tables = [](const std::vector<TileOp>& v) {
        size_t c = 0;
        for (const TileOp& o : v) c += o.kind == CL_DIAG;
        return c;
    };
    std::vector<TileOp>* best = &runs;
    if (tables(late) < tables(*best)) best = &late;
    if (tables(stabbed) < tables(*best)) best = &stabbed;
    ops = std::move(*best);
    return QSV_OK;
}

static inline int rank_in(uint32_t set, int bit) { return __builtin_popcount(set & ((1u << bit) - 1u)); }

// Build phases / sub-ops / data of one cluster pass into the staging area (no launch).
static int build_cluster_pass(qsv_state* s, const std::vector<HGate>& gates, const PassPlan& plan, int T,
                              std::vector<char>& stage, ClPass& P, size_t& phases_off, size_t& subs_off,
                              size_t& data_off) {
    const int n = s->n_local;
    const int cap = std::min(T, n);
    uint64_t active = plan.qmask;
    for (int b = 0; b < n && popc64(active) < cap; ++b) active |= 1ull << b;  // low bits first
    memset(&P, 0, sizeof P);
    P.T = popc64(active);
    int local_of[64];
    for (int b = 0, t = 0; b < 64; ++b) {
        local_of[b] = -1;
        if ((active >> b) & 1ull) {
            P.pos[t] = b;
            local_of[b] = t++;
        }
    }
    std::vector<TileOp> ops;
    QSV_TRY(make_tile_ops(gates, plan, local_of, ops));

    std::vector<ClPhase> phases;
    std::vector<ClSub> subs;
    std::vector<double2> data;
    std::vector<char> done(ops.size(), 0);
    size_t ndone = 0;
    const int TT = P.T;
    while (ndone < ops.size()) {
        uint32_t S = 0, blocked = 0;
        std::vector<int> sel;
        for (size_t i = 0; i < ops.size(); ++i) {
            if (done[i]) continue;
            uint32_t m = 0;
            for (int j = 0; j < ops[i].k; ++j)
                if (ops[i].bits[j] < 64) m |= 1u << ops[i].bits[j];
            if (m & blocked) {
                blocked |= m;
                continue;
            }
            if (ops[i].kind != CL_DIAG) {
                if (__builtin_popcount(S | m) > CL_R) {
                    blocked |= m;
                    continue;
                }
                S |= m;
            }
            sel.push_back((int)i);
            done[i] = 1;
            ++ndone;
        }
        for (int b = TT - 1; b >= 0 && __builtin_popcount(S) < CL_R; --b) S |= 1u << b;
        ClPhase ph;
        memset(&ph, 0, sizeof ph);
        for (int b = 0, j = 0; b < TT; ++b)
            if ((S >> b) & 1u) ph.rpos[j++] = (int8_t)b;
        // thread bits: lanes 0..2 on three non-register tile bits with distinct (bit mod 3), so that
        // a quarter-warp's 16-byte accesses fall into distinct swizzled slots
        uint32_t rest = ((1u << TT) - 1u) & ~S;
        int nt = 0;
        for (int c = 0; c < 3; ++c)
            for (int b = 0; b < TT; ++b)
                if (((rest >> b) & 1u) && b % 3 == c) {
                    ph.tpos[nt++] = (int8_t)b;
                    rest &= ~(1u << b);
                    break;
                }
        for (int b = 0; b < TT; ++b)
            if ((rest >> b) & 1u) ph.tpos[nt++] = (int8_t)b;
        // Generalised permutations that can be moved to the end (start) of the phase -- they act on
        // bits no later (earlier) remaining op touches -- are composed into one permutation of the
        // 16 register slots and folded into the shared-memory write (read) addresses.
        std::vector<char> role(sel.size(), 0);  // 0 in-register, 1 leading, 2 trailing
        auto op_mask = [&](int id) {
            uint32_t m = 0;
            for (int j = 0; j < ops[id].k; ++j)
                if (ops[id].bits[j] < 64) m |= 1u << ops[id].bits[j];
            return m;
        };
        {
            uint32_t blk = 0;
            for (size_t i = sel.size(); i-- > 0;) {
                const uint32_t m = op_mask(sel[i]);
                if (ops[sel[i]].kind == CL_MONO && !(m & blk)) role[i] = 2;
                else blk |= m;
            }
            blk = 0;
            for (size_t i = 0; i < sel.size(); ++i) {
                if (role[i]) continue;
                const uint32_t m = op_mask(sel[i]);
                if (ops[sel[i]].kind == CL_MONO && !(m & blk)) role[i] = 1;
                else blk |= m;
            }
        }
        auto compose = [&](int which, int* src16, cplx* c16) {
            bool any = false;
            for (int r = 0; r < CL_REGS; ++r) {
                src16[r] = r;
                c16[r] = cplx(1.0, 0.0);
            }
            for (size_t i = 0; i < sel.size(); ++i) {
                if (role[i] != which) continue;
                any = true;
                const TileOp& op = ops[sel[i]];
                int rb[3];
                for (int j = 0; j < op.k; ++j) rb[j] = rank_in(S, op.bits[j]);
                int nsrc[CL_REGS];
                cplx nc[CL_REGS];
                for (int r = 0; r < CL_REGS; ++r) {
                    int rr = 0;
                    for (int j = 0; j < op.k; ++j) rr |= ((r >> rb[j]) & 1) << j;
                    int x = r;
                    for (int j = 0; j < op.k; ++j) x = (x & ~(1 << rb[j])) | (((op.src[rr] >> j) & 1) << rb[j]);
                    nsrc[r] = src16[x];
                    nc[r] = op.m[rr] * c16[x];
                }
                memcpy(src16, nsrc, sizeof nsrc);
                for (int r = 0; r < CL_REGS; ++r) c16[r] = nc[r];
            }
            return any;
        };
        uint64_t rperm = CL_PERM_ID, wperm = CL_PERM_ID;
        ph.rcoef = ph.wcoef = CL_NOCOEF;
        {
            int src16[CL_REGS];
            cplx c16[CL_REGS];
            if (compose(1, src16, c16)) {  // v[r] = c[r] * slot[src[r]]
                rperm = 0;
                bool unit = true;
                for (int r = 0; r < CL_REGS; ++r) {
                    rperm |= (uint64_t)src16[r] << (4 * r);
                    unit &= (c16[r] == cplx(1.0, 0.0));
                }
                if (!unit) {
                    ph.rcoef = (uint32_t)data.size();
                    for (int r = 0; r < CL_REGS; ++r) data.push_back(make_double2(c16[r].real(), c16[r].imag()));
                }
            }
            if (compose(2, src16, c16)) {  // slot[r] = c[r] * v[src[r]]  ->  register q goes to slot dst[q]
                wperm = 0;
                bool unit = true;
                cplx cq[CL_REGS];
                for (int r = 0; r < CL_REGS; ++r) {
                    wperm |= (uint64_t)r << (4 * src16[r]);
                    cq[src16[r]] = c16[r];
                    unit &= (c16[r] == cplx(1.0, 0.0));
                }
                if (!unit) {
                    ph.wcoef = (uint32_t)data.size();
                    for (int q = 0; q < CL_REGS; ++q) data.push_back(make_double2(cq[q].real(), cq[q].imag()));
                }
            }
        }
        ph.rperm_lo = (uint32_t)rperm;
        ph.rperm_hi = (uint32_t)(rperm >> 32);
        ph.wperm_lo = (uint32_t)wperm;
        ph.wperm_hi = (uint32_t)(wperm >> 32);
        std::vector<int> mid;
        for (size_t i = 0; i < sel.size(); ++i)
            if (!role[i]) mid.push_back(sel[i]);
        ph.sub_first = (uint16_t)subs.size();
        ph.nsub = (uint8_t)mid.size();
        if (mid.size() > 255 || subs.size() + mid.size() > 65535)
            QSV_FAIL(QSV_ERR_UNSUPPORTED, "internal: too many ops in one pass");
        for (int id : mid) {
            const TileOp& op = ops[id];
            ClSub sb;
            memset(&sb, 0, sizeof sb);
            sb.kind = (uint8_t)op.kind;
            sb.k = (uint8_t)op.k;
            sb.data_off = (uint32_t)data.size();
            if (op.kind == CL_DIAG) {
                sb.nbits = (uint8_t)op.k;
                for (int j = 0; j < op.k; ++j) sb.bits[j] = (int8_t)op.bits[j];
                for (int r = 0; r < CL_REGS; ++r) {
                    uint32_t roff = 0;
                    for (int j = 0; j < CL_R; ++j) roff |= (uint32_t)((r >> j) & 1) << ph.rpos[j];
                    uint32_t part = 0;
                    for (int j = 0; j < op.k; ++j)
                        if (op.bits[j] < 64) part |= ((roff >> op.bits[j]) & 1u) << j;
                    sb.regpart[r] = (uint16_t)part;
                }
                for (const cplx& v : op.m) data.push_back(make_double2(v.real(), v.imag()));
            } else {
                // canonical order: register bits ascending; order[j'] = matrix bit feeding canonical bit j'
                int rb[3], order[3] = {0, 1, 2};
                for (int j = 0; j < op.k; ++j) rb[j] = rank_in(S, op.bits[j]);
                std::sort(order, order + op.k, [&](int a, int b) { return rb[a] < rb[b]; });
                for (int j = 0; j < op.k; ++j) sb.b[j] = (uint8_t)rb[order[j]];
                const int dim = 1 << op.k;
                auto to_orig = [&](int rc) {
                    int r = 0;
                    for (int j = 0; j < op.k; ++j) r |= ((rc >> j) & 1) << order[j];
                    return r;
                };
                auto to_canon = [&](int r) {
                    int rc = 0;
                    for (int j = 0; j < op.k; ++j) rc |= ((r >> order[j]) & 1) << j;
                    return rc;
                };
                if (op.kind == CL_MONO) {
                    sb.flags = op.unit ? 1 : 0;
                    for (int rc = 0; rc < dim; ++rc) {
                        const int r = to_orig(rc);
                        sb.aux |= (uint32_t)to_canon(op.src[r]) << (3 * rc);
                        data.push_back(make_double2(op.m[r].real(), op.m[r].imag()));
                    }
                } else {
                    for (int rc = 0; rc < dim; ++rc)
                        for (int cc = 0; cc < dim; ++cc) {
                            const cplx v = op.m[(size_t)to_orig(rc) * dim + to_orig(cc)];
                            data.push_back(make_double2(v.real(), v.imag()));
                        }
                }
            }
            subs.push_back(sb);
        }
        phases.push_back(ph);
    }
    const size_t ph_bytes = phases.size() * sizeof(ClPhase);
    const size_t sub_bytes = subs.size() * sizeof(ClSub);
    const size_t data_bytes = data.size() * sizeof(double2);
    phases_off = (stage.size() + 63) & ~(size_t)63;
    subs_off = (phases_off + ph_bytes + 63) & ~(size_t)63;
    data_off = (subs_off + sub_bytes + 15) & ~(size_t)15;
    stage.resize(data_off + data_bytes);
    memcpy(stage.data() + phases_off, phases.data(), ph_bytes);
    memcpy(stage.data() + subs_off, subs.data(), sub_bytes);
    memcpy(stage.data() + data_off, data.data(), data_bytes);
    P.nphases = (int)phases.size();
    s->ctr.phases += phases.size();
    return QSV_OK;
}

template <int NT, int SETS>
static int launch_cluster_nt(qsv_state* s, const ClPass& P, uint64_t ntiles) {
    const size_t smem = (size_t)16 << P.T;
    QSV_CUDA(cudaFuncSetAttribute(k_cluster_pass<NT, SETS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_cluster_pass<NT, SETS><<<(unsigned)ntiles, NT, smem, s->stream>>>(s->amp, P);
    return QSV_OK;
}

static int launch_cluster_pass(qsv_state* s, const ClPass& P, uint64_t ntiles) {
    if (ntiles > 0x7fffffffull) QSV_FAIL(QSV_ERR_UNSUPPORTED, "too many tiles for one launch");
    switch (P.T) {
        case 10: QSV_TRY((launch_cluster_nt<64, 1>(s, P, ntiles))); break;
        case 11: QSV_TRY((launch_cluster_nt<128, 1>(s, P, ntiles))); break;
        case 12: QSV_TRY((launch_cluster_nt<128, 2>(s, P, ntiles))); break;
        case 13: QSV_TRY((launch_cluster_nt<128, 4>(s, P, ntiles))); break;
        default: QSV_FAIL(QSV_ERR_UNSUPPORTED, "internal: cluster pass with %d tile bits", P.T);
    }
    QSV_CUDA(cudaGetLastError());
    s->ctr.kernel_launches++;
    s->ctr.state_passes++;
    return QSV_OK;
}

// ---- fused pass (k_fused_pass, qsv_fused.cuh): host side -------------------------------------
struct FusedBuilt {
    FpPass P;
    size_t tabs_off, diags_off, gdata_off;
};

// Build the descriptors of one fused pass.  `ok` = false when the pass does not fit the kernel's
// fixed capacities (the caller falls back to k_cluster_pass).
static int build_fused_pass(qsv_state* s, const std::vector<HGate>& gates, const PassPlan& plan,
                            std::vector<char>& stage, FusedBuilt& out, bool& ok) {
    ok = false;
    const int n = s->n_local;
    if (n < FP_T) return QSV_OK;
    uint64_t active = plan.qmask;
    for (int b = 0; b < n && popc64(active) < FP_T; ++b) active |= 1ull << b;  // low bits first
    if (popc64(active) != FP_T) return QSV_OK;
    static const bool use_lifting = !getenv("QSV_NO_LIFT");
    static const bool use_d1l4 = !(getenv("QSV_NO_D1L4") && atoi(getenv("QSV_NO_D1L4")));
    static const bool use_bilinear = getenv("QSV_BILINEAR") && atoi(getenv("QSV_BILINEAR"));  // measured slower (profiles/r2_bilinear_tables_ab.txt): opt-in
    FpPass& P = out.P;
    memset(&P, 0, sizeof P);
    int local_of[64];
    for (int b = 0, t = 0; b < 64; ++b) {
        local_of[b] = -1;
        if ((active >> b) & 1ull) {
            P.pos[t] = b;
            local_of[b] = t++;
        }
    }
    std::vector<TileOp> ops;
    QSV_TRY(make_tile_ops(gates, plan, local_of, ops));

    // Permutation gates whose index map is affine over GF(2) and whose coefficients are all 1 (X, CX,
    // SWAP and products of them) never need register bits: the tile stays where it is in shared
    // memory and the map "tile index -> slot" (`A`, affine) absorbs them.  Because the swizzle is
    // XOR-linear too, slot = table[thread] ^ mask[register] keeps its form, so the kernel is unchanged.
    std::vector<char> lin(ops.size(), 0);
    if (s->use_linear_perm)
        for (size_t i = 0; i < ops.size(); ++i) {
            const TileOp& op = ops[i];
            if (op.kind != CL_MONO || !op.unit || op.k > 3) continue;
            bool ok_lin = true;
            const int dim = 1 << op.k;
            for (int x = 0; x < dim && ok_lin; ++x)
                for (int y = 0; y < dim; ++y)
                    if ((op.src[x ^ y] ^ op.src[0]) != ((op.src[x] ^ op.src[0]) ^ (op.src[y] ^ op.src[0]))) {
                        ok_lin = false;
                        break;
                    }
            lin[i] = ok_lin;
        }
    // ---- phase selection: [leading permutations][dense / diagonal][trailing permutations]
    struct PhaseSel {
        uint32_t S;
        std::vector<int> sel;
        std::vector<char> role;  // 0 mid, 1 leading permutation, 2 trailing permutation (both in register bits),
                                 // 3 leading / 4 trailing affine permutation absorbed by the slot map
    };
    std::vector<PhaseSel> sels;
    {
        std::vector<char> done(ops.size(), 0);
        size_t ndone = 0;
        while (ndone < ops.size()) {
            PhaseSel ps;
            ps.S = 0;
            uint32_t blocked = 0;
            uint32_t reg_lead = 0;    // bits touched by a leading permutation held in register bits
            uint32_t free_trail = 0;  // bits touched by a trailing permutation absorbed by the slot map
            int state[FP_T];  // per tile bit: 0 untouched, 1 dense/diag seen, 2 closed by a trailing permutation
            for (int b = 0; b < FP_T; ++b) state[b] = 0;
            for (size_t i = 0; i < ops.size(); ++i) {
                if (done[i]) continue;
                uint32_t m = 0;
                for (int j = 0; j < ops[i].k; ++j)
                    if (ops[i].bits[j] < 64) m |= 1u << ops[i].bits[j];
                if (m & blocked) {
                    blocked |= m;
                    continue;
                }
                bool closed = false, fresh = true;
                for (int b = 0; b < FP_T; ++b)
                    if ((m >> b) & 1u) {
                        closed |= state[b] == 2;
                        fresh &= state[b] == 0;
                    }
                char role = 0;
                if (ops[i].kind == CL_DIAG) {
                    if (closed) {
                        blocked |= m;
                        continue;
                    }
                    for (int b = 0; b < FP_T; ++b)
                        if (((m >> b) & 1u) && state[b] == 0) state[b] = 1;
                } else if (lin[i] && !(fresh && (m & reg_lead))) {
                    // order on shared bits: slot map first, register permutations after it (read side);
                    // register permutations first, slot map after them (write side)
                    role = fresh ? 3 : 4;
                    if (role == 4) {
                        free_trail |= m;
                        for (int b = 0; b < FP_T; ++b)
                            if ((m >> b) & 1u) state[b] = 2;
                    }
                } else {
                    if (__builtin_popcount(ps.S | m) > 4 || (ops[i].kind != CL_MONO && closed) ||
                        (ops[i].kind == CL_MONO && !fresh && (m & free_trail))) {
                        blocked |= m;
                        continue;
                    }
                    ps.S |= m;
                    if (ops[i].kind == CL_MONO) {
                        role = fresh ? 1 : 2;
                        if (role == 1) reg_lead |= m;
                        if (role == 2)
                            for (int b = 0; b < FP_T; ++b)
                                if ((m >> b) & 1u) state[b] = 2;
                    } else {
                        for (int b = 0; b < FP_T; ++b)
                            if ((m >> b) & 1u) state[b] = 1;
                    }
                }
                ps.sel.push_back((int)i);
                ps.role.push_back(role);
                done[i] = 1;
                ++ndone;
            }
            sels.push_back(std::move(ps));
        }
    }
    if (sels.empty() && plan.gate_ids.empty()) {
        // a pass without gates: one empty phase (registers loaded and stored unchanged).  Used as the
        // standalone qubit exchange of sharded states (qsv_fused_identity_pass: a scatter pass with no arithmetic)
        PhaseSel ps;
        ps.S = 0;
        sels.push_back(std::move(ps));
    }
    if (sels.empty() || sels.size() > FP_MAX_PHASES) return QSV_OK;
    // Warp-local passes: three tile bits that no dense / permutation op of the pass touches become
    // the warp index in EVERY phase.  A warp then owns the same 512 amplitudes from the first
    // phase to the last and phases only need __syncwarp().
    uint32_t W = 0;
    {
        uint32_t touched = 0;
        for (const PhaseSel& ps : sels) touched |= ps.S;
        const uint32_t untouched = ((1u << FP_T) - 1u) & ~touched;
        if (s->use_warp_local && __builtin_popcount(untouched) >= 3) {
            for (int b = FP_T - 1; b >= 0 && __builtin_popcount(W) < 3; --b)
                if ((untouched >> b) & 1u) W |= 1u << b;
        }
        for (PhaseSel& ps : sels)
            for (int b = FP_T - 1; b >= 0 && __builtin_popcount(ps.S) < 4; --b)
                if (!((W >> b) & 1u)) ps.S |= 1u << b;
    }
    P.warp_local = W ? 1 : 0;

    std::vector<uint16_t> tabs;
    std::vector<FpDiag> diags;
    std::vector<double2> gdata;
    int nsubs = 0, nmat = 0;
    bool low4 = true;  // tile bits 0..3 are index bits 0..3 (contiguous 256-byte runs)
    for (int b = 0; b < 4; ++b) low4 &= (P.pos[b] == b);

    // slot map: tile index x (as the circuit sees it at this point of the pass) lives in slot
    // fp_swz(A(x)), A(x) = off ^ XOR of col[b] over the set bits of x; identity after the copy-in
    struct Aff {
        uint32_t col[FP_T], off;
        uint32_t operator()(uint32_t x) const {
            uint32_t y = off;
            for (int b = 0; b < FP_T; ++b)
                if ((x >> b) & 1u) y ^= col[b];
            return y;
        }
    } A;
    A.off = 0;
    for (int b = 0; b < FP_T; ++b) A.col[b] = 1u << b;
    // after permutation `op` (new[x] = old[sub(x)], sub replaces the op's bits by src[...]) : A <- A o sub
    auto absorb = [&](const TileOp& op) {
        auto sub = [&](uint32_t x) {
            int rr = 0;
            for (int j = 0; j < op.k; ++j) rr |= (int)((x >> op.bits[j]) & 1u) << j;
            uint32_t y = x;
            for (int j = 0; j < op.k; ++j)
                y = (y & ~(1u << op.bits[j])) | ((uint32_t)((op.src[rr] >> j) & 1) << op.bits[j]);
            return y;
        };
        Aff B;
        B.off = A(sub(0));
        for (int b = 0; b < FP_T; ++b) B.col[b] = A(sub(1u << b)) ^ B.off;
        A = B;
    };
    auto is_identity = [&]() {
        bool id = A.off == 0;
        for (int b = 0; b < FP_T; ++b) id &= A.col[b] == (1u << b);
        return id;
    };

    for (size_t pi = 0; pi < sels.size(); ++pi) {
        const PhaseSel& ps = sels[pi];
        const uint32_t S = ps.S;
        FpPhase& ph = P.phases[pi];
        int rpos[4];
        for (int b = 0, j = 0; b < FP_T; ++b)
            if ((S >> b) & 1u) rpos[j++] = b;
        bool has_trail = false;
        for (size_t i = 0; i < ps.sel.size(); ++i) {
            if (ps.role[i] == 3) absorb(ops[ps.sel[i]]);
            has_trail |= ps.role[i] == 4;
        }
        const bool direct = (pi + 1 == sels.size()) && low4 && !(S & 15u) && s->use_direct_store && !has_trail;
        // thread bits
        int tpos[8], nt = 0;
        uint32_t rest = ((1u << FP_T) - 1u) & ~S & ~W;
        if (!direct) {
            // lanes 0..2 on three tile bits whose slots differ in linearly independent swizzled
            // columns (slot bits 0..2): a quarter-warp's 16-byte accesses are then conflict free.
            // With the identity slot map these are three bits with distinct (bit mod 3).
            auto colv = [&](int b) { return fp_swz(A.col[b]) & 7u; };
            bool found = false;
            for (int b0 = 0; b0 < FP_T && !found; ++b0) {
                if (!((rest >> b0) & 1u) || !colv(b0)) continue;
                for (int b1 = b0 + 1; b1 < FP_T && !found; ++b1) {
                    if (!((rest >> b1) & 1u) || !colv(b1) || colv(b1) == colv(b0)) continue;
                    for (int b2 = b1 + 1; b2 < FP_T && !found; ++b2) {
                        if (!((rest >> b2) & 1u) || !colv(b2)) continue;
                        const uint32_t c2 = colv(b2);
                        if (c2 == colv(b0) || c2 == colv(b1) || c2 == (colv(b0) ^ colv(b1))) continue;
                        tpos[nt++] = b0;
                        tpos[nt++] = b1;
                        tpos[nt++] = b2;
                        rest &= ~((1u << b0) | (1u << b1) | (1u << b2));
                        found = true;
                    }
                }
            }
        }
        for (int b = 0; b < FP_T; ++b)
            if ((rest >> b) & 1u) tpos[nt++] = b;
        for (int b = 0; b < FP_T; ++b)
            if ((W >> b) & 1u) tpos[nt++] = b;  // thread bits 5..7 = warp index
        if (nt != 8) QSV_FAIL(QSV_ERR_UNSUPPORTED, "internal: fused pass thread bits");
        uint16_t base_of[FP_NT];
        for (int t = 0; t < FP_NT; ++t) {
            uint32_t base = 0;
            for (int j = 0; j < 8; ++j) base |= (uint32_t)((t >> j) & 1) << tpos[j];
            base_of[t] = (uint16_t)base;
        }
        if (tabs.size() / FP_NT >= FP_MAX_TABS) return QSV_OK;
        ph.tab = (uint16_t)(tabs.size() / FP_NT);
        for (int t = 0; t < FP_NT; ++t) tabs.push_back((uint16_t)fp_swz(A(base_of[t])));
        auto slot_mask = [&](int nib) {
            uint32_t m = 0;
            for (int j = 0; j < 4; ++j)
                if ((nib >> j) & 1) m ^= fp_swz(A.col[rpos[j]]);
            return (uint16_t)(m << 4);
        };
        // compose the leading / trailing permutations into one map of the 16 register slots
        auto compose = [&](int which, int* src16, cplx* c16) {
            bool any = false;
            for (int r = 0; r < 16; ++r) {
                src16[r] = r;
                c16[r] = cplx(1.0, 0.0);
            }
            for (size_t i = 0; i < ps.sel.size(); ++i) {
                if (ps.role[i] != which) continue;
                any = true;
                const TileOp& op = ops[ps.sel[i]];
                int rb[3];
                for (int j = 0; j < op.k; ++j) rb[j] = rank_in(S, op.bits[j]);
                int nsrc[16];
                cplx nc[16];
                for (int r = 0; r < 16; ++r) {
                    int rr = 0;
                    for (int j = 0; j < op.k; ++j) rr |= ((r >> rb[j]) & 1) << j;
                    int x = r;
                    for (int j = 0; j < op.k; ++j) x = (x & ~(1 << rb[j])) | (((op.src[rr] >> j) & 1) << rb[j]);
                    nsrc[r] = src16[x];
                    nc[r] = op.m[rr] * c16[x];
                }
                memcpy(src16, nsrc, sizeof nsrc);
                for (int r = 0; r < 16; ++r) c16[r] = nc[r];
            }
            return any;
        };
        int src16[16], dst16[16];
        cplx c16[16];
        ph.rcoef = ph.wcoef = FP_NOCOEF;
        compose(1, src16, c16);  // v[r] = c[r] * slot[src[r]]
        {
            bool unit = true;
            for (int r = 0; r < 16; ++r) {
                ph.rd[r] = slot_mask(src16[r]);
                unit &= (c16[r] == cplx(1.0, 0.0));
            }
            if (!unit) {
                ph.rcoef = (uint32_t)gdata.size();
                for (int r = 0; r < 16; ++r) gdata.push_back(make_double2(c16[r].real(), c16[r].imag()));
            }
        }
        compose(2, src16, c16);  // slot[r] = c[r] * v[src[r]]: register q = src[r] goes to slot r
        {
            bool unit = true;
            cplx cq[16];
            for (int r = 0; r < 16; ++r) {
                dst16[src16[r]] = r;
                cq[src16[r]] = c16[r];
                unit &= (c16[r] == cplx(1.0, 0.0));
            }
            for (int q = 0; q < 16; ++q) ph.wr[q] = slot_mask(dst16[q]);
            if (!unit) {
                ph.wcoef = (uint32_t)gdata.size();
                for (int q = 0; q < 16; ++q) gdata.push_back(make_double2(cq[q].real(), cq[q].imag()));
            }
        }
        ph.direct = direct ? 1 : 0;
        if (direct) {
            for (int j = 0; j < 8; ++j) P.st_tpos[j] = (uint8_t)tpos[j];
            for (int q = 0; q < 16; ++q) {
                uint64_t off = 0;
                for (int j = 0; j < 4; ++j)
                    if ((dst16[q] >> j) & 1) off |= 1ull << P.pos[rpos[j]];
                P.st_off[q] = off * sizeof(double2);
            }
        }
        ph.sub_first = (uint16_t)nsubs;
        int cnt = 0;
        // ops of this phase; a diagonal table that factors into (register bits) x (all other bits) is
        // split: the register factor is 16 uniform coefficients from the constant bank, the rest is ONE
        // table lookup per thread -- instead of 16 scattered 16-byte loads per thread through L1
        std::vector<TileOp> pops;
        for (size_t i = 0; i < ps.sel.size(); ++i) {
            if (ps.role[i]) continue;
            const TileOp& op = ops[ps.sel[i]];
            uint32_t regm = 0;  // table index bits fed by register bits
            if (op.kind == CL_DIAG && s->use_diag_split)
                for (int j = 0; j < op.k; ++j)
                    if (op.bits[j] < 64 && ((S >> op.bits[j]) & 1u)) regm |= 1u << j;
            bool split = regm != 0 && op.m[0] != cplx(0.0, 0.0);
            if (split) {
                const cplx t0 = op.m[0];
                for (size_t x = 0; x < op.m.size() && split; ++x) {
                    const cplx lhs = op.m[x] * t0, rhs = op.m[x & regm] * op.m[x & ~(size_t)regm];
                    split = std::abs(lhs - rhs) <= 1e-14 * std::max(1.0, std::abs(lhs));
                }
            }
            if (!split) {
                pops.push_back(op);
                continue;
            }
            TileOp u;
            memset(u.bits, 0, sizeof u.bits);
            memset(u.src, 0, sizeof u.src);
            u.unit = false;
            u.kind = CL_DIAGU;
            u.k = 4;
            u.m.assign(16, cplx(1.0, 0.0));
            for (int r = 0; r < 16; ++r) {
                size_t x = 0;
                for (int j = 0; j < op.k; ++j)
                    if ((regm >> j) & 1u) x |= (size_t)((r >> rank_in(S, op.bits[j])) & 1) << j;
                u.m[r] = op.m[x];
            }
            pops.push_back(std::move(u));
            TileOp rest;
            memset(rest.bits, 0, sizeof rest.bits);
            memset(rest.src, 0, sizeof rest.src);
            rest.unit = false;
            rest.kind = CL_DIAG;
            rest.k = 0;
            int jmap[DIAG_MAX_BITS];
            for (int j = 0; j < op.k; ++j)
                if (!((regm >> j) & 1u)) {
                    jmap[rest.k] = j;
                    rest.bits[rest.k++] = op.bits[j];
                }
            rest.m.assign((size_t)1 << rest.k, cplx(1.0, 0.0));
            bool trivial = true;
            for (size_t y = 0; y < rest.m.size(); ++y) {
                size_t x = 0;
                for (int jj = 0; jj < rest.k; ++jj) x |= ((y >> jj) & 1) << jmap[jj];
                rest.m[y] = op.m[x] / op.m[0];
                trivial &= rest.m[y] == cplx(1.0, 0.0);
            }
            if (!trivial) pops.push_back(std::move(rest));
        }
        for (size_t i = 0; i < pops.size(); ++i) {
            const TileOp& op = pops[i];
            if (nsubs >= FP_MAX_SUBS || cnt >= 255) return QSV_OK;
            FpSub& sb = P.subs[nsubs];
            memset(&sb, 0, sizeof sb);
            if (op.kind == CL_DIAGU) {
                if (nmat + 32 > FP_MAX_MAT) return QSV_OK;
                sb.kind = FP_DIAGU;
                sb.mat = (uint16_t)nmat;
                for (int r = 0; r < 16; ++r) {
                    P.mats[nmat++] = op.m[r].real();
                    P.mats[nmat++] = op.m[r].imag();
                }
            } else if (op.kind == CL_DIAG) {
                if (tabs.size() / FP_NT >= FP_MAX_TABS || diags.size() >= FP_MAX_DIAGS) return QSV_OK;
                // Bilinear factorisation (FP_DIAGB, qsv_fused.cuh): T[r, o] = T0[o] * U[r] * prod_{a in r} C_a[o].
                // Holds for every product of controlled-phase / RZZ / RZ gates (QFT, Trotter steps); verified
                // entry by entry, so any other table simply keeps the generic path below.
                if (use_bilinear) {
                    int rj[4], na = 0, oj[DIAG_MAX_BITS], no = 0;   // table index bits fed by register bits / by the rest
                    for (int j = 0; j < op.k; ++j) {
                        if (op.bits[j] < 64 && ((S >> op.bits[j]) & 1u)) rj[na++] = j;
                        else oj[no++] = j;
                    }
                    bool ok_b = na >= 1 && na < op.k && std::abs(op.m[0]) > 0.5;
                    const size_t nrow = (size_t)1 << no;
                    std::vector<cplx> t0(nrow), ca((size_t)na * nrow), ureg((size_t)1 << na);
                    auto full = [&](size_t rr, size_t oo) {   // table index from (register part, other part)
                        size_t x = 0;
                        for (int a = 0; a < na; ++a) x |= ((rr >> a) & 1) << rj[a];
                        for (int b = 0; b < no; ++b) x |= ((oo >> b) & 1) << oj[b];
                        return x;
                    };
                    for (size_t oo = 0; oo < nrow && ok_b; ++oo) {
                        t0[oo] = op.m[full(0, oo)];
                        ok_b = std::abs(t0[oo]) > 0.5;
                        for (int a = 0; a < na && ok_b; ++a) ca[(size_t)a * nrow + oo] = op.m[full((size_t)1 << a, oo)] / t0[oo];
                    }
                    for (size_t rr = 0; rr < ureg.size() && ok_b; ++rr) {
                        cplx den = t0[0];
                        for (int a = 0; a < na; ++a)
                            if ((rr >> a) & 1) den *= ca[(size_t)a * nrow];
                        ureg[rr] = op.m[full(rr, 0)] / den;
                    }
                    for (size_t rr = 0; rr < ureg.size() && ok_b; ++rr)
                        for (size_t oo = 0; oo < nrow && ok_b; ++oo) {
                            cplx model = t0[oo] * ureg[rr];
                            for (int a = 0; a < na; ++a)
                                if ((rr >> a) & 1) model *= ca[(size_t)a * nrow + oo];
                            ok_b = std::abs(model - op.m[full(rr, oo)]) <= 2e-14;
                        }
                    bool u_trivial = true;
                    for (const cplx& u : ureg) u_trivial &= std::abs(u - cplx(1.0, 0.0)) <= 1e-15;
                    if (ok_b && (u_trivial || (nsubs + 1 < FP_MAX_SUBS && nmat + 32 <= FP_MAX_MAT && cnt < 254))) {
                        if (!u_trivial) {   // the register-only factor: 16 uniform coefficients
                            FpSub& su = P.subs[nsubs];
                            memset(&su, 0, sizeof su);
                            su.kind = FP_DIAGU;
                            su.mat = (uint16_t)nmat;
                            for (int r = 0; r < 16; ++r) {
                                size_t rr = 0;
                                for (int a = 0; a < na; ++a) rr |= (size_t)((r >> rank_in(S, op.bits[rj[a]])) & 1) << a;
                                P.mats[nmat++] = ureg[rr].real();
                                P.mats[nmat++] = ureg[rr].imag();
                            }
                            ++nsubs;
                            ++cnt;
                        }
                        FpSub& sb2 = P.subs[nsubs];
                        memset(&sb2, 0, sizeof sb2);
                        FpDiag d;
                        memset(&d, 0, sizeof d);
                        d.table_off = (uint32_t)gdata.size();
                        d.tab = (uint16_t)(tabs.size() / FP_NT);
                        d.regpart[0] = (uint16_t)na;
                        for (int a = 0; a < na; ++a) d.regpart[1 + a] = (uint16_t)rank_in(S, op.bits[rj[a]]);
                        d.regpart[5] = (uint16_t)(1 + na);
                        uint32_t tb[DIAG_MAX_BITS];
                        for (int b = 0; b < no; ++b) {
                            tb[b] = 0xffffffffu;
                            const int bit = op.bits[oj[b]];
                            if (bit >= 64) {
                                d.gbit[d.nglob] = (uint8_t)(bit - 64);
                                d.gj[d.nglob] = (uint8_t)b;
                                ++d.nglob;
                            } else {
                                tb[b] = (uint32_t)bit;
                            }
                        }
                        for (int t = 0; t < FP_NT; ++t) {
                            uint32_t thr = 0;
                            for (int b = 0; b < no; ++b)
                                if (tb[b] != 0xffffffffu) thr |= ((base_of[t] >> tb[b]) & 1u) << b;
                            tabs.push_back((uint16_t)thr);
                        }
                        for (size_t oo = 0; oo < nrow; ++oo) {
                            gdata.push_back(make_double2(t0[oo].real(), t0[oo].imag()));
                            for (int a = 0; a < na; ++a) {
                                const cplx c = ca[(size_t)a * nrow + oo];
                                gdata.push_back(make_double2(c.real(), c.imag()));
                            }
                        }
                        sb2.kind = FP_DIAGB;
                        sb2.mat = (uint16_t)diags.size();
                        diags.push_back(d);
                        ++nsubs;
                        ++cnt;
                        continue;
                    }
                }
                FpDiag d;
                memset(&d, 0, sizeof d);
                d.table_off = (uint32_t)gdata.size();
                d.tab = (uint16_t)(tabs.size() / FP_NT);
                bool any_reg = false;
                uint32_t tmask_bits[DIAG_MAX_BITS];
                for (int j = 0; j < op.k; ++j) {
                    tmask_bits[j] = 0xffffffffu;
                    if (op.bits[j] >= 64) {
                        d.gbit[d.nglob] = (uint8_t)(op.bits[j] - 64);
                        d.gj[d.nglob] = (uint8_t)j;
                        ++d.nglob;
                    } else if ((S >> op.bits[j]) & 1u) {
                        any_reg = true;
                    } else {
                        tmask_bits[j] = (uint32_t)op.bits[j];
                    }
                }
                for (int r = 0; r < 16; ++r) {
                    uint32_t roff = 0;
                    for (int j = 0; j < 4; ++j) roff |= (uint32_t)((r >> j) & 1) << rpos[j];
                    uint32_t part = 0;
                    for (int j = 0; j < op.k; ++j)
                        if (op.bits[j] < 64 && ((S >> op.bits[j]) & 1u)) part |= ((roff >> op.bits[j]) & 1u) << j;
                    d.regpart[r] = (uint16_t)part;
                }
                d.all_same = any_reg ? 0 : 1;
                for (int t = 0; t < FP_NT; ++t) {
                    uint32_t thr = 0;
                    for (int j = 0; j < op.k; ++j)
                        if (tmask_bits[j] != 0xffffffffu) thr |= ((base_of[t] >> tmask_bits[j]) & 1u) << j;
                    tabs.push_back((uint16_t)thr);
                }
                for (const cplx& v : op.m) gdata.push_back(make_double2(v.real(), v.imag()));
                sb.kind = FP_DIAG;
                sb.mat = (uint16_t)diags.size();
                diags.push_back(d);
            } else {
                // canonical order: register bits ascending; order[j'] = matrix bit feeding canonical bit j'
                int rb[2], order[2] = {0, 1};
                for (int j = 0; j < op.k; ++j) rb[j] = rank_in(S, op.bits[j]);
                if (op.k == 2 && rb[0] > rb[1]) std::swap(order[0], order[1]);
                const int dim = 1 << op.k;
                auto to_orig = [&](int rc) {
                    int r = 0;
                    for (int j = 0; j < op.k; ++j) r |= ((rc >> j) & 1) << order[j];
                    return r;
                };
                const bool real = op.kind == CL_DENSE1R || op.kind == CL_DENSE2R;
                const int need = dim * dim * (real ? 1 : 2);
                if (nmat + need > FP_MAX_MAT) return QSV_OK;
                sb.kind = op.k == 1 ? (real ? FP_D1R : FP_D1C) : (real ? FP_D2R : FP_D2C);
                sb.b = op.k == 1 ? (uint8_t)rb[0] : (uint8_t)(rb[order[0]] | (rb[order[1]] << 2));
                sb.mat = (uint16_t)nmat;
                if (op.k == 1 && op.m[0] == cplx(1.0, 0.0) && op.m[2] == cplx(1.0, 0.0) &&
                    ((op.m[1] == cplx(1.0, 0.0) && op.m[3] == cplx(-1.0, 0.0)) ||
                     (op.m[1] == cplx(0.0, -1.0) && op.m[3] == cplx(0.0, 1.0)))) {
                    sb.kind = op.m[1].imag() == 0.0 ? FP_D1H : FP_D1HY;  // unnormalised butterfly: additions only
                    ++nsubs;
                    ++cnt;
                    continue;
                }
                if (op.k == 1 && !real && mat_is_real_times_sdg(op.m)) {
                    sb.kind = FP_D1RY;  // (a0, a1) -> R (a0, -i a1): the cost of the real gate
                    P.mats[nmat++] = op.m[0].real();
                    P.mats[nmat++] = -op.m[1].imag();
                    P.mats[nmat++] = op.m[2].real();
                    P.mats[nmat++] = -op.m[3].imag();
                    ++nsubs;
                    ++cnt;
                    continue;
                }
                if (op.k == 1 && real && use_lifting) {
                    // a plane rotation (c -s ; s c) with c >= 0 (every RY, the real part of a split
                    // RX / U3) runs as three shears: 3 instead of 4 FP64 operations per amplitude
                    const double c = op.m[0].real(), ms = op.m[1].real(), sn = op.m[2].real(), c1 = op.m[3].real();
                    if (c == c1 && ms == -sn && c >= 0.0 && std::abs(c * c + sn * sn - 1.0) <= 4.5e-16) {  // the bound of split_complex_1q: the shears reproduce the matrix to ~1 ulp
                        sb.kind = FP_D1L;
                        P.mats[nmat++] = -sn / (1.0 + c);
                        P.mats[nmat++] = sn;
                        ++nsubs;
                        ++cnt;
                        continue;
                    }
                }
                for (int rc = 0; rc < dim; ++rc)
                    for (int cc = 0; cc < dim; ++cc) {
                        const cplx v = op.m[(size_t)to_orig(rc) * dim + to_orig(cc)];
                        P.mats[nmat++] = v.real();
                        if (!real) P.mats[nmat++] = v.imag();
                    }
            }
            ++nsubs;
            ++cnt;
        }
        // peephole: four consecutive lifted rotations on the four different register bits (they commute)
        // become one FP_D1L4 sub-op -- the RY layer of a hardware-efficient ansatz, every phase of it
        if (use_d1l4) {
            const int first = ph.sub_first;
            for (int i = first; i + 3 < nsubs; ++i) {
                int seen = 0;
                bool ok4 = true;
                for (int j = 0; j < 4 && ok4; ++j) {
                    const FpSub& q = P.subs[i + j];
                    ok4 = q.kind == FP_D1L && !((seen >> q.b) & 1);
                    seen |= 1 << q.b;
                }
                if (!ok4 || seen != 15) {
                    // the same for four butterflies of one kind (basis change of an X / Y family)
                    const int k0 = P.subs[i].kind;
                    seen = 0;
                    bool okh = k0 == FP_D1H || k0 == FP_D1HY;
                    for (int j = 0; j < 4 && okh; ++j) {
                        const FpSub& q = P.subs[i + j];
                        okh = q.kind == k0 && !((seen >> q.b) & 1);
                        seen |= 1 << q.b;
                    }
                    if (!okh || seen != 15) continue;
                    FpSub m;
                    memset(&m, 0, sizeof m);
                    m.kind = FP_D1H4;
                    m.b = k0 == FP_D1HY ? 1 : 0;
                    P.subs[i] = m;
                    for (int j = i + 4; j < nsubs; ++j) P.subs[j - 3] = P.subs[j];
                    nsubs -= 3;
                    cnt -= 3;
                    continue;
                }
                if (nmat + 8 > FP_MAX_MAT) continue;
                double c8[8];
                for (int j = 0; j < 4; ++j) {
                    const FpSub& q = P.subs[i + j];
                    c8[2 * q.b] = P.mats[q.mat];
                    c8[2 * q.b + 1] = P.mats[q.mat + 1];
                }
                FpSub m;
                memset(&m, 0, sizeof m);
                m.kind = FP_D1L4;
                m.mat = (uint16_t)nmat;
                for (int j = 0; j < 8; ++j) P.mats[nmat++] = c8[j];
                P.subs[i] = m;
                for (int j = i + 4; j < nsubs; ++j) P.subs[j - 3] = P.subs[j];
                nsubs -= 3;
                cnt -= 3;
            }
        }
        if (use_d1l4) {
            // ... a uniform register-bit diagonal next to a per-thread-scalar table (the two halves of a
            // separable diagonal such as an RZ layer; diagonals commute, either order)
            for (int i = ph.sub_first; i + 1 < nsubs; ++i) {
                const FpSub a = P.subs[i], b2 = P.subs[i + 1];
                const bool au = a.kind == FP_DIAGU, bu = b2.kind == FP_DIAGU;
                const FpSub& u = au ? a : b2;
                const FpSub& d = au ? b2 : a;
                if (au == bu || d.kind != FP_DIAG || !diags[d.mat].all_same) continue;
                FpSub m;
                memset(&m, 0, sizeof m);
                m.kind = FP_DIAGUS;
                m.mat = u.mat;
                m.pad = d.mat;
                P.subs[i] = m;
                for (int j = i + 2; j < nsubs; ++j) P.subs[j - 1] = P.subs[j];
                nsubs -= 1;
                cnt -= 1;
            }
            // ... four real 2x2 gates on the four register bits (Hadamard / RY layers that are not lifted)
            for (int i = ph.sub_first; i + 3 < nsubs; ++i) {
                int seen = 0;
                bool ok4 = nmat + 16 <= FP_MAX_MAT;
                for (int j = 0; j < 4 && ok4; ++j) {
                    const FpSub& q = P.subs[i + j];
                    ok4 = q.kind == FP_D1R && !((seen >> q.b) & 1);
                    seen |= 1 << q.b;
                }
                if (!ok4 || seen != 15) continue;
                double c16[16];
                for (int j = 0; j < 4; ++j) {
                    const FpSub& q = P.subs[i + j];
                    for (int e = 0; e < 4; ++e) c16[4 * q.b + e] = P.mats[q.mat + e];
                }
                FpSub m;
                memset(&m, 0, sizeof m);
                m.kind = FP_D1R4;
                m.mat = (uint16_t)nmat;
                for (int j = 0; j < 16; ++j) P.mats[nmat++] = c16[j];
                P.subs[i] = m;
                for (int j = i + 4; j < nsubs; ++j) P.subs[j - 3] = P.subs[j];
                nsubs -= 3;
                cnt -= 3;
            }
        }
        ph.nsub = (uint8_t)cnt;
        for (size_t i = 0; i < ps.sel.size(); ++i)
            if (ps.role[i] == 4) absorb(ops[ps.sel[i]]);  // the data stays put; later readers use the new map
    }
    for (int r = 0; r < 16; ++r) {
        uint64_t o = 0;
        for (int j = 0; j < 4; ++j)
            if ((r >> j) & 1) o |= 1ull << P.pos[8 + j];
        P.row_off[r] = o * sizeof(double2);
    }
    // copy-out (passes without a direct-store phase): tile index (row r, thread t) is read from
    // slot fp_swz(A(t)) ^ fp_swz(A.col-part of r << 8)
    for (int r = 0; r < 16; ++r) P.out_row[r] = (uint16_t)(fp_swz(A((uint32_t)r << 8) ^ A.off) << 4);
    P.out_tab = 0xffff;
    if (!is_identity()) {
        if (tabs.size() / FP_NT >= FP_MAX_TABS) return QSV_OK;
        P.out_tab = (uint16_t)(tabs.size() / FP_NT);
        for (int t = 0; t < FP_NT; ++t) tabs.push_back((uint16_t)fp_swz(A((uint32_t)t)));
    }
    // runs of consecutive non-tile index bits: tile number bits [src, src+len) land at [dst, dst+len)
    P.nseg = 0;
    for (int b = 0, src = 0; b < n;) {
        if ((active >> b) & 1ull) {
            ++b;
            continue;
        }
        int e = b;
        while (e < n && !((active >> e) & 1ull)) ++e;
        P.seg_src[P.nseg] = (uint8_t)src;
        P.seg_dst[P.nseg] = (uint8_t)b;
        P.seg_mask[P.nseg] = (e - b) >= 32 ? 0xffffffffu : ((1u << (e - b)) - 1u);
        ++P.nseg;
        src += e - b;
        b = e;
    }
    P.nphases = (int)sels.size();
    P.ntabs = (int)(tabs.size() / FP_NT);
    P.ntiles = (uint32_t)(1ull << (n - FP_T));
    P.index_base = s->world > 1 ? ((uint64_t)s->rank << s->n_local) : 0;
    const size_t tabs_bytes = tabs.size() * sizeof(uint16_t);
    const size_t diag_bytes = diags.size() * sizeof(FpDiag);
    const size_t gdata_bytes = gdata.size() * sizeof(double2);
    out.tabs_off = (stage.size() + 63) & ~(size_t)63;
    out.diags_off = (out.tabs_off + tabs_bytes + 63) & ~(size_t)63;
    out.gdata_off = (out.diags_off + 63) & ~(size_t)63;
    stage.resize(out.gdata_off + gdata_bytes);
    memcpy(stage.data() + out.tabs_off, tabs.data(), tabs_bytes);
    if (diag_bytes) memcpy(P.diags, diags.data(), diag_bytes);
    if (gdata_bytes) memcpy(stage.data() + out.gdata_off, gdata.data(), gdata_bytes);
    s->ctr.phases += sels.size();
    ok = true;
    return QSV_OK;
}

static void trace_fused_pass(const FpPass& P, FILE* f) {
    int cnt[16] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    fprintf(f, "[fused] T bits:");
    for (int b = 0; b < FP_T; ++b) fprintf(f, " %d", P.pos[b]);
    fprintf(f, " | %sphases %d:", P.warp_local ? "warp-local " : "", P.nphases);
    for (int p = 0; p < P.nphases; ++p) {
        fprintf(f, " [%d%s%s%s]", P.phases[p].nsub, P.phases[p].direct ? "d" : "",
                P.phases[p].rcoef != FP_NOCOEF ? "r" : "", P.phases[p].wcoef != FP_NOCOEF ? "w" : "");
        for (int i = 0; i < P.phases[p].nsub; ++i) cnt[P.subs[P.phases[p].sub_first + i].kind]++;
    }
    int same = 0;
    for (int p = 0; p < P.nphases; ++p)
        for (int i = 0; i < P.phases[p].nsub; ++i) {
            const FpSub& sb = P.subs[P.phases[p].sub_first + i];
            if (sb.kind == FP_DIAG && P.diags[sb.mat].all_same) ++same;
        }
    fprintf(f, " | d1c %d d1r %d d1l %d (+%d x4) d1ry %d d1h %d d2c %d d2r %d diag %d (per-thread scalar %d) diagb %d diagu %d | tabs %d\n",
            cnt[0], cnt[1] + 4 * cnt[14], cnt[6], cnt[11], cnt[7], cnt[8] + cnt[9] + 4 * cnt[12], cnt[2] + 2 * cnt[13], cnt[3], cnt[4] + cnt[15], same, cnt[10], cnt[5] + cnt[15], P.ntabs);
}

static int upload_product_tables(qsv_state* s, uint64_t neutral, const std::vector<cplx>& extra);

// Descriptor of the remote gather of a fused swap pass: where the swapped local bits sit relative to
// the pass's tile (row bits, copy-in thread bits, or outside the tile), see FpRemote in qsv_fused.cuh.
static void build_fp_remote(const FpPass& P, const QsvPendingSwap& W, FpRemote& R) {
    memset(&R, 0, sizeof R);
    R.k = W.k;
    R.a = W.a;
    R.err = W.err;
    for (int j = 0; j < (1 << W.k); ++j) {
        R.peer[j] = W.peer[j];
        R.flag[j] = W.flag[j];
    }
    int tile_bit_of[64];
    for (int b = 0; b < 64; ++b) tile_bit_of[b] = -1;
    for (int b = 0; b < FP_T; ++b) tile_bit_of[P.pos[b]] = b;
    for (int i = 0; i < W.k; ++i) {
        const int lp = W.lpos[i];
        const uint64_t abit = (uint64_t)((W.a >> i) & 1);
        R.lmask |= 1ull << lp;
        R.a_l |= abit << lp;
        R.thr_bit[i] = -1;
        const int tb = tile_bit_of[lp];
        if (tb < 0) {
            R.out_j[R.nout] = (uint8_t)i;
            R.out_pos[R.nout] = (uint8_t)lp;
            ++R.nout;
            R.out_xor |= abit << lp;
        } else {
            R.in_j[R.nin++] = (uint8_t)i;
            if (tb < 8) R.thr_bit[i] = (int8_t)tb;
        }
    }
    for (int r = 0; r < 16; ++r) {
        uint32_t jr = 0;
        uint64_t clean = 0;
        for (int jb = 0; jb < 4; ++jb) {
            if (!((r >> jb) & 1)) continue;
            const int pos = P.pos[8 + jb];
            int which = -1;
            for (int i = 0; i < W.k; ++i)
                if (W.lpos[i] == pos) which = i;
            if (which >= 0) jr |= 1u << which;
            else clean |= 1ull << pos;
        }
        R.jrow[r] = (uint8_t)jr;
        R.row_clean[r] = clean * sizeof(double2);
    }
    // tile order: bring the tile-number bits of the non-tile swapped positions down to counter bits 9, 10, ...
    // (just above the grid of 2 x 148 CTAs) so that local and remote tiles alternate within every CTA
    {
        int ntbits = 0;
        while ((1u << ntbits) < P.ntiles) ++ntbits;
        int lo = 9;
        for (int i = 0; i < R.nout; ++i) {
            int q = 0;  // tile-number bit of index bit out_pos[i] = non-tile index bits below it
            for (int b = 0; b < R.out_pos[i]; ++b)
                if (tile_bit_of[b] < 0) ++q;
            if (q <= lo || lo >= ntbits) continue;
            bool clash = false;  // `lo` must not itself be the home of another swapped bit
            for (int i2 = 0; i2 < R.nout; ++i2) {
                int q2 = 0;
                for (int b = 0; b < R.out_pos[i2]; ++b)
                    if (tile_bit_of[b] < 0) ++q2;
                clash |= q2 == lo;
            }
            if (clash) { ++lo; --i; continue; }
            R.tswap_lo[R.ntswap] = (uint8_t)lo;
            R.tswap_hi[R.ntswap] = (uint8_t)q;
            ++R.ntswap;
            ++lo;
        }
    }
    // scatter mode, direct-store phase: the same decomposition for the store mapping
    for (int i = 0; i < W.k; ++i) {
        R.st_thr_bit[i] = -1;
        for (int j = 0; j < 8; ++j)
            if (P.pos[P.st_tpos[j]] == W.lpos[i]) R.st_thr_bit[i] = (int8_t)j;
    }
    for (int r = 0; r < 16; ++r) {
        uint32_t jr = 0;
        const uint64_t off = P.st_off[r] / sizeof(double2);
        for (int i = 0; i < W.k; ++i)
            if ((off >> W.lpos[i]) & 1ull) jr |= 1u << i;
        R.st_jreg[r] = (uint8_t)jr;
        R.st_clean[r] = (off & ~R.lmask) * sizeof(double2);
    }
    static const long long to_ms = getenv("QSV_FUSED_SWAP_TIMEOUT_MS") ? atoll(getenv("QSV_FUSED_SWAP_TIMEOUT_MS")) : 4000;
    R.timeout = to_ms * 2000000ll;  // ~2 GHz SM clock
}

// a fused pass can carry the pending swap when it reads its own buffer in the plain way
static bool fused_swap_capable(const qsv_state* s, const FpPass& P) {
    return s->pending_swap && !s->read_from && !s->prod_pending && !P.store_prob;
}

static int launch_fused_swap_pass(qsv_state* s, const FpPass& P) {
    const int mode = s->pending_swap->mode;
    QSV_RANGE("qsv:fused_swap_pass");
    // two CTAs per SM: each may use half of the SM's shared memory minus the 1 KB the system reserves per CTA
    static int smem_cta = 0;
    if (!smem_cta) {
        int per_sm = 0;
        QSV_CUDA(cudaDeviceGetAttribute(&per_sm, cudaDevAttrMaxSharedMemoryPerMultiprocessor, s->device));
        int per_block = 0;
        QSV_CUDA(cudaDeviceGetAttribute(&per_block, cudaDevAttrMaxSharedMemoryPerBlockOptin, s->device));
        smem_cta = std::min(per_sm / 2 - 1024, per_block);
        QSV_CUDA(cudaFuncSetAttribute(k_fused_swap_pass<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_cta));
        QSV_CUDA(cudaFuncSetAttribute(k_fused_swap_pass<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_cta));
    }
    const size_t tabs_bytes = (size_t)P.ntabs * FP_NT * sizeof(uint16_t);
    FpRemote R;
    build_fp_remote(P, *s->pending_swap, R);
    static const int stage_max = getenv("QSV_FUSED_SWAP_STAGE") ? atoi(getenv("QSV_FUSED_SWAP_STAGE")) : 15;
    const long spare = (long)smem_cta - (long)(FP_TILE * sizeof(double2)) - (long)tabs_bytes;
    R.stage_rows = mode == 1 ? (int)std::max(0l, std::min<long>(std::min(stage_max, 15), spare / 4096)) : 0;
    const size_t smem = FP_TILE * sizeof(double2) + (size_t)R.stage_rows * 4096 + tabs_bytes;
    static const bool trace = getenv("QSV_TRACE") != nullptr;
    if (trace) {
        trace_fused_pass(P, stderr);
        fprintf(stderr, "        (fused swap pass, %s: k=%d, %d of the swapped bits are tile bits, %d rows of the next tile staged)\n",
                mode == 1 ? "gather" : "scatter", R.k, R.nin, R.stage_rows);
    }
    QSV_TRY(qsv_dist_fused_swap_begin(s, 1u << R.nin, &R.target));
    const unsigned grid = (unsigned)std::min<uint64_t>(P.ntiles, (uint64_t)s->num_sms * 2);
    if (mode == 1) k_fused_swap_pass<1><<<grid, FP_NT, smem, s->stream>>>(s->amp, P, R);
    else k_fused_swap_pass<2><<<grid, FP_NT, smem, s->stream>>>(s->amp, P, R);
    QSV_CUDA(cudaGetLastError());
    s->pending_swap = nullptr;
    s->ctr.kernel_launches++;
    s->ctr.state_passes++;
    return qsv_dist_fused_swap_end(s);
}

static int launch_fused_pass(qsv_state* s, const FusedBuilt& fb, bool is_last) {
    QSV_RANGE("qsv:fused_pass");
    const FpPass& P = fb.P;
    if (s->pending_swap && s->pending_swap->mode == 1) {  // gather: the first pass after the swap, or nobody
        if (fused_swap_capable(s, P)) return launch_fused_swap_pass(s, P);
        QSV_TRY(qsv_dist_flush_swap(s));
    }
    // scatter: only the last pass of the run may carry it (the caller runs the standalone exchange otherwise)
    if (s->pending_swap && s->pending_swap->mode == 2 && is_last && fused_swap_capable(s, P))
        return launch_fused_swap_pass(s, P);
    const size_t smem = FP_TILE * sizeof(double2) + (size_t)P.ntabs * FP_NT * sizeof(uint16_t);
    static bool attr_set = false;
    if (!attr_set) {
        QSV_CUDA(cudaFuncSetAttribute(k_fused_pass, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)(FP_TILE * sizeof(double2) + FP_MAX_TABS * FP_NT * sizeof(uint16_t))));
        attr_set = true;
    }
    static const bool trace = getenv("QSV_TRACE") != nullptr;
    if (trace) trace_fused_pass(P, stderr);
    if (trace) QSV_CUDA(cudaEventRecord(s->tev0, s->stream));
    const double2* src = s->read_from ? s->read_from : s->amp;
    s->read_from = nullptr;  // only the first pass after the request reads the other buffer
    if (s->prod_pending && !s->use_product_fold) QSV_TRY(qsv_flush_product(s));
    if (s->prod_pending) {
        // first pass after the product-state prologue: nothing is read, phase 0 builds its registers
        // from the per-qubit vectors (FpPass.gen_*)
        s->prod_pending = 0;
        FpPass G = P;
        uint64_t tile_qubits = 0;
        for (int b = 0; b < FP_T; ++b) tile_qubits |= 1ull << P.pos[b];
        std::vector<cplx> tmpl(FP_TILE);  // the tile every tile is a multiple of
        for (uint32_t y = 0; y < FP_TILE; ++y) {
            cplx v(1.0, 0.0);
            for (int b = 0; b < FP_T; ++b) {
                const double* pv = s->prod_vec[P.pos[b]];
                v *= ((y >> b) & 1u) ? cplx(pv[2], pv[3]) : cplx(pv[0], pv[1]);
            }
            tmpl[y] = v;
        }
        QSV_TRY(upload_product_tables(s, tile_qubits, tmpl));
        G.gen_h = s->prod_h;
        G.gen_lo = s->prod_tab;
        G.gen_hi = G.gen_lo + ((size_t)1 << s->prod_h);
        G.gen_tile = G.gen_hi + ((size_t)1 << (s->n_local - s->prod_h));
        unsigned grid = (unsigned)std::min<uint64_t>(G.ntiles, (uint64_t)s->num_sms * 2);
        if (trace) fprintf(stderr, "        (product state: the copy-in streams the 64 KB template tile, no state read)\n");
        k_fused_pass<<<grid, FP_NT, smem, s->stream>>>(s->amp, s->amp, G);
        QSV_CUDA(cudaGetLastError());
        if (trace) {
            float ms = 0.f;
            QSV_CUDA(cudaEventRecord(s->tev1, s->stream));
            QSV_CUDA(cudaEventSynchronize(s->tev1));
            QSV_CUDA(cudaEventElapsedTime(&ms, s->tev0, s->tev1));
            fprintf(stderr, "        -> %.3f ms\n", ms);
        }
        s->ctr.kernel_launches++;
        s->ctr.write_passes++;  // one sweep written, none read
        return QSV_OK;
    }
    {
        unsigned grid = (unsigned)std::min<uint64_t>(P.ntiles, (uint64_t)s->num_sms * 2);
        k_fused_pass<<<grid, FP_NT, smem, s->stream>>>(s->amp, src, P);
    }
    QSV_CUDA(cudaGetLastError());
    if (trace) {
        float ms = 0.f;
        QSV_CUDA(cudaEventRecord(s->tev1, s->stream));
        QSV_CUDA(cudaEventSynchronize(s->tev1));
        QSV_CUDA(cudaEventElapsedTime(&ms, s->tev0, s->tev1));
        fprintf(stderr, "        -> %.3f ms\n", ms);
    }
    s->ctr.kernel_launches++;
    s->ctr.state_passes++;
    return QSV_OK;
}

// One fused pass with no gates over the low 12 index bits.  With a pending scatter swap it IS the
// standalone global<->local qubit exchange: every tile is loaded from the own slab and stored into the
// slabs that own it afterwards -- remote WRITES only (k_swap_p2p also reads remotely, and remote reads
// cap at ~340 GB/s per GPU on this machine).  Returns with s->pending_swap still set when the pass
// could not be built (the caller falls back to k_swap_p2p).
int qsv_fused_identity_pass(qsv_state* s) {
    if (!s->pending_swap || s->pending_swap->mode != 2 || s->n_local < FP_T) return QSV_OK;
    std::vector<HGate> none;
    PassPlan plan;
    plan.qmask = 0;
    std::vector<char> stage;
    FusedBuilt fb;
    bool ok = false;
    const uint64_t phases_before = s->ctr.phases;
    QSV_TRY(build_fused_pass(s, none, plan, stage, fb, ok));
    s->ctr.phases = phases_before;
    if (!ok) return QSV_OK;
    const size_t used = stage.size();
    if (used) {
        QSV_TRY(qsv_ensure_scratch(s, used + 256));
        QSV_TRY(qsv_ensure_pinned(s, used + 256));
        QSV_CUDA(cudaStreamSynchronize(s->stream));
        memcpy(s->pinned, stage.data(), used);
        QSV_CUDA(cudaMemcpyAsync(s->scratch, s->pinned, used, cudaMemcpyHostToDevice, s->stream));
        s->ctr.h2d_bytes += used;
    }
    fb.P.tabs = (const uint16_t*)((char*)s->scratch + fb.tabs_off);
    fb.P.gdata = (const double2*)((char*)s->scratch + fb.gdata_off);
    if (!fused_swap_capable(s, fb.P)) return QSV_OK;
    const uint64_t passes_before = s->ctr.state_passes;
    QSV_TRY(launch_fused_swap_pass(s, fb.P));
    s->ctr.state_passes = passes_before;  // an exchange, not a gate pass
    return QSV_OK;
}

static bool direct_capable(const HGate& g) {
    if (g.diag || g.k < 1 || g.k > 2) return false;
    int src[4];
    return !mat_is_mono(g.m, 1 << g.k, src);
}

static int launch_direct(qsv_state* s, const HGate& g) {
    QSV_RANGE("qsv:direct_gate");
    const bool real = mat_is_real(g.m);
    const uint64_t groups = qsv_local_size(s) >> g.k;
    uint64_t blocks = (groups + 255) / 256;
    const uint64_t cap = (uint64_t)s->num_sms * 16;
    if (blocks > cap) blocks = cap;
    if (g.k == 1) {
        Mat2 M;
        for (int i = 0; i < 4; ++i) M.m[i] = make_double2(g.m[i].real(), g.m[i].imag());
        if (real) k_direct1<true><<<(unsigned)blocks, 256, 0, s->stream>>>(s->amp, groups, g.q[0], M);
        else k_direct1<false><<<(unsigned)blocks, 256, 0, s->stream>>>(s->amp, groups, g.q[0], M);
    } else {
        Mat4 M;
        for (int i = 0; i < 16; ++i) M.m[i] = make_double2(g.m[i].real(), g.m[i].imag());
        const int lo = std::min(g.q[0], g.q[1]), hi = std::max(g.q[0], g.q[1]);
        const uint64_t o1 = 1ull << g.q[0], o2 = 1ull << g.q[1];
        if (real) k_direct2<true><<<(unsigned)blocks, 256, 0, s->stream>>>(s->amp, groups, lo, hi, o1, o2, M);
        else k_direct2<false><<<(unsigned)blocks, 256, 0, s->stream>>>(s->amp, groups, lo, hi, o1, o2, M);
    }
    QSV_CUDA(cudaGetLastError());
    s->ctr.kernel_launches++;
    s->ctr.state_passes++;
    return QSV_OK;
}

static bool tile_capable(const HGate& g) {
    if (g.diag) return g.k <= DIAG_MAX_BITS;
    if (g.k <= 2) return true;
    int src[8];
    return g.k == 3 && mat_is_mono(g.m, 8, src);
}

// Apply classified gates whose qubits are all local (physical bits < n_local).
static thread_local FILE* g_plan_dump = nullptr;
static thread_local FpPass* g_capture_first = nullptr;   // qsv_debug_fused_swap: copy of the first (gather) / last (scatter) fused pass
static thread_local int g_capture_state = 0;             // 0 nothing seen, 1 captured, 2 that pass is not a fused pass
static thread_local int g_capture_last = 0;              // capture the last pass of the stream instead of the first

int qsv_apply_hgates(qsv_state* s, std::vector<HGate>& gates, int flags) {
    QSV_RANGE("qsv:plan_and_apply");
    const bool no_fusion = flags & QSV_FLAG_NO_FUSION;
    const bool want_prob = s->prob_store_req != 0;  // one-shot request (qsv_expval.cu)
    s->prob_store_req = 0;
    s->prob_layout = 0;
    if (!no_fusion) {
        fuse_gates(gates, s->n_local);
        split_complex_1q(gates);
    }
    size_t start = 0;
    while (start < gates.size()) {
        size_t end = start;
        while (end < gates.size() && tile_capable(gates[end])) ++end;
        if (end > start) {
            std::vector<HGate> seg(gates.begin() + start, gates.begin() + end);
            std::vector<PassPlan> passes;
            const int T = s->tile_bits;
            const bool cluster = s->use_cluster && std::min(T, s->n_local) >= 10;
            plan_passes(seg, s->n_local, T, s->low_bits, no_fusion, cluster, passes);
            std::vector<PassParams> params(passes.size());
            std::vector<ClPass> cparams(passes.size());
            std::vector<size_t> ops_off(passes.size()), data_off(passes.size()), subs_off(passes.size());
            std::vector<char> kind(passes.size(), 0);  // 0 legacy tile, 1 cluster, 2 direct, 3 fused
            std::vector<FusedBuilt> fparams(s->use_fused && cluster && T == FP_T ? passes.size() : 0);
            std::vector<char> stage;
            for (size_t i = 0; i < passes.size(); ++i) {
                if (s->use_direct && passes[i].gate_ids.size() == 1 && direct_capable(seg[passes[i].gate_ids[0]])) {
                    kind[i] = 2;
                    continue;
                }
                if (!fparams.empty()) {
                    bool ok = false;
                    QSV_TRY(build_fused_pass(s, seg, passes[i], stage, fparams[i], ok));
                    if (ok) {
                        kind[i] = 3;
                        continue;
                    }
                }
                if (cluster) {
                    kind[i] = 1;
                    QSV_TRY(build_cluster_pass(s, seg, passes[i], T, stage, cparams[i], ops_off[i], subs_off[i],
                                               data_off[i]));
                } else {
                    QSV_TRY(build_pass(s, seg, passes[i], T, stage, params[i], ops_off[i], data_off[i]));
                }
            }
            if (g_plan_dump) {  // host-only planning (qsv_debug_plan): report and skip the device work
                if (g_capture_first && !passes.empty() && (g_capture_last || g_capture_state == 0)) {
                    const size_t which = g_capture_last ? passes.size() - 1 : 0;
                    if (kind[which] == 3 && (!g_capture_last || end == gates.size())) {
                        *g_capture_first = fparams[which].P;
                        g_capture_state = 1;
                    } else {
                        g_capture_state = 2;
                    }
                }
                for (size_t i = 0; i < passes.size(); ++i) {
                    fprintf(g_plan_dump, "pass kind=%d gates=%zu ", (int)kind[i], passes[i].gate_ids.size());
                    if (kind[i] == 3) trace_fused_pass(fparams[i].P, g_plan_dump);
                    else fprintf(g_plan_dump, "\n");
                }
                if (end < gates.size()) {
                    fprintf(g_plan_dump, "pass kind=big k=%d\n", gates[end].k);
                    ++end;
                }
                start = end;
                continue;
            }
            const size_t used = stage.size();
            if (used) {
                QSV_TRY(qsv_ensure_scratch(s, used + 256));
                QSV_TRY(qsv_ensure_pinned(s, used + 256));
                // the pinned staging area and the device op buffer are reused across calls
                QSV_CUDA(cudaStreamSynchronize(s->stream));
                memcpy(s->pinned, stage.data(), used);
                QSV_CUDA(cudaMemcpyAsync(s->scratch, s->pinned, used, cudaMemcpyHostToDevice, s->stream));
                s->ctr.h2d_bytes += used;
            }
            if (!passes.empty() && kind[0] != 3) QSV_TRY(qsv_flush_product(s));
            if (s->pending_swap && s->pending_swap->mode == 1 && (passes.empty() || kind[0] != 3))
                QSV_TRY(qsv_dist_flush_swap(s));
            if (s->read_from && kind[0] != 3) {  // out-of-place request, but the first pass cannot do it
                QSV_CUDA(cudaMemcpyAsync(s->amp, s->read_from, (size_t)16 << s->n_local, cudaMemcpyDeviceToDevice,
                                         s->stream));
                s->ctr.state_passes++;
                s->read_from = nullptr;
            }
            if (want_prob && end == gates.size() && !passes.empty() && kind.back() == 3) {
                FpPass& last = fparams.back().P;
                if (last.nphases > 0 && last.phases[last.nphases - 1].direct) {
                    last.store_prob = 1;
                    s->prob_layout = 1;
                }
            }
            for (size_t i = 0; i < passes.size(); ++i) {
                char* base = (char*)s->scratch;
                if (kind[i] == 2) {
                    QSV_TRY(launch_direct(s, seg[passes[i].gate_ids[0]]));
                } else if (kind[i] == 3) {
                    fparams[i].P.tabs = (const uint16_t*)(base + fparams[i].tabs_off);
                    fparams[i].P.gdata = (const double2*)(base + fparams[i].gdata_off);
                    QSV_TRY(launch_fused_pass(s, fparams[i], end == gates.size() && i + 1 == passes.size()));
                } else if (kind[i] == 1) {
                    cparams[i].phases = (const ClPhase*)(base + ops_off[i]);
                    cparams[i].subs = (const ClSub*)(base + subs_off[i]);
                    cparams[i].data = (const double2*)(base + data_off[i]);
                    QSV_TRY(launch_cluster_pass(s, cparams[i], 1ull << (s->n_local - cparams[i].T)));
                } else {
                    params[i].ops = (const DevOp*)(base + ops_off[i]);
                    params[i].data = (const double2*)(base + data_off[i]);
                    QSV_TRY(launch_tile_pass(s, params[i], 1ull << (s->n_local - params[i].T)));
                }
            }
        }
        if (end < gates.size()) {
            QSV_TRY(qsv_flush_product(s));
            if (s->pending_swap && s->pending_swap->mode == 1) QSV_TRY(qsv_dist_flush_swap(s));
            if (s->read_from) {
                QSV_CUDA(cudaMemcpyAsync(s->amp, s->read_from, (size_t)16 << s->n_local, cudaMemcpyDeviceToDevice,
                                         s->stream));
                s->ctr.state_passes++;
                s->read_from = nullptr;
            }
            QSV_TRY(apply_dense_big(s, gates[end]));
            ++end;
        }
        start = end;
    }
    if (s->pending_swap && s->pending_swap->mode == 1) QSV_TRY(qsv_dist_flush_swap(s));  // nothing consumed it (empty stream)
    return QSV_OK;
}

static int validate_qargs(const qsv_state* s, int k, const int* qargs, const char* who) {
    if (k < 0 || k > QSV_MAX_GATE_QUBITS)
        QSV_FAIL(QSV_ERR_INVALID, "%s: k = %d outside [0, %d]", who, k, QSV_MAX_GATE_QUBITS);
    if (k > s->n) QSV_FAIL(QSV_ERR_INVALID, "%s: %d-qubit gate on a %d-qubit state", who, k, s->n);
    for (int i = 0; i < k; ++i) {
        if (qargs[i] < 0 || qargs[i] >= s->n)
            QSV_FAIL(QSV_ERR_INVALID, "%s: qarg %d out of range for %d qubits", who, qargs[i], s->n);
        for (int j = 0; j < i; ++j)
            if (qargs[i] == qargs[j]) QSV_FAIL(QSV_ERR_INVALID, "%s: duplicate qarg %d", who, qargs[i]);
    }
    return QSV_OK;
}

// sharded states: make every qubit the stream needs local (qsv_dist.cu)
int qsv_dist_apply(qsv_state* s, std::vector<HGate>& gates, int flags);

static int qsv_apply_gates_impl(qsv_state* s, const qsv_gate* gates, uint64_t n_gates, int flags);
void qsv_state_defaults(qsv_state* s);  // qsv_state.cu

// amp[i] = lo[i mod 2^h] * hi[i div 2^h]: a product state from two half-register tables
__global__ void __launch_bounds__(256) k_product_state(double2* __restrict__ amp, uint64_t count, int h,
                                                       const double2* __restrict__ lo,
                                                       const double2* __restrict__ hi) {
    const uint64_t stride = (uint64_t)gridDim.x * 256;
    const uint64_t mask = (1ull << h) - 1;
    for (uint64_t i = blockIdx.x * 256ull + threadIdx.x; i < count; i += stride)
        amp[i] = cmul(__ldg(lo + (i & mask)), __ldg(hi + (i >> h)));
}

// Product-state prologue (single GPU, state = pending |index>): every one-qubit gate that comes
// before the first multi-qubit gate on its qubit acts on a product state, so it is applied to
// that qubit's 2-vector on the host; the state is then written once as the tensor product.
// Consumed gates are removed from `hg`.  Does nothing (basis stays pending) if no gate qualifies.
static int product_prologue(qsv_state* s, std::vector<HGate>& hg) {
    const int n = s->n_local;
    cplx vec[64][2];
    for (int q = 0; q < n; ++q) {
        const int bit = (int)((s->basis_index >> q) & 1ull);
        vec[q][0] = bit ? 0.0 : 1.0;
        vec[q][1] = bit ? 1.0 : 0.0;
    }
    const uint64_t all = n >= 64 ? ~0ull : ((1ull << n) - 1);
    uint64_t blocked = 0;
    std::vector<char> consumed(hg.size(), 0);
    size_t count = 0;
    for (size_t i = 0; i < hg.size() && blocked != all; ++i) {
        const HGate& g = hg[i];
        uint64_t m = 0;
        for (int j = 0; j < g.k; ++j) m |= 1ull << g.q[j];
        if (g.k == 1 && !(m & blocked)) {
            cplx* v = vec[g.q[0]];
            if (g.diag) {
                v[0] *= g.m[0];
                v[1] *= g.m[1];
            } else {
                const cplx a = g.m[0] * v[0] + g.m[1] * v[1], b = g.m[2] * v[0] + g.m[3] * v[1];
                v[0] = a;
                v[1] = b;
            }
            consumed[i] = 1;
            ++count;
        } else {
            blocked |= m;
        }
    }
    if (!count) return QSV_OK;
    // the sweep is deferred: the first fused pass builds its registers from the per-qubit vectors
    // (launch_fused_pass); whoever else touches the buffer first calls qsv_flush_product
    for (int q = 0; q < n; ++q) {
        s->prod_vec[q][0] = vec[q][0].real();
        s->prod_vec[q][1] = vec[q][0].imag();
        s->prod_vec[q][2] = vec[q][1].real();
        s->prod_vec[q][3] = vec[q][1].imag();
    }
    s->prod_pending = 1;
    s->basis_pending = 0;
    std::vector<HGate> rest;
    rest.reserve(hg.size() - count);
    for (size_t i = 0; i < hg.size(); ++i)
        if (!consumed[i]) rest.push_back(std::move(hg[i]));
    hg.swap(rest);
    return QSV_OK;
}

// Upload the half-register tables of the pending product state: lo[i] = prod_{q < h} vec[q][bit q of i],
// hi likewise for the qubits from h up; qubits in `neutral` contribute the factor 1 (they are
// supplied elsewhere).  `extra` complex numbers are appended after the two tables.
static int upload_product_tables(qsv_state* s, uint64_t neutral, const std::vector<cplx>& extra) {
    const int n = s->n_local;
    const int h = (n + 1) / 2;
    const size_t nlo = (size_t)1 << h, nhi = (size_t)1 << (n - h);
    const size_t bytes = (nlo + nhi + extra.size()) * sizeof(double2);
    if (s->prod_tab_bytes < bytes) {
        if (s->prod_tab) QSV_CUDA(cudaFree(s->prod_tab));
        s->prod_tab = nullptr;
        s->prod_tab_bytes = 0;
        QSV_CUDA(cudaMalloc(&s->prod_tab, bytes));
        s->prod_tab_bytes = bytes;
    }
    std::vector<cplx> tab(nlo + nhi + extra.size());
    auto fill = [&](cplx* t, int q0, int nq) {  // t[i] = prod_j vec[q0 + j][bit j of i]
        t[0] = 1.0;
        for (int j = 0; j < nq; ++j) {
            const double* pv = s->prod_vec[q0 + j];
            const bool skip = (neutral >> (q0 + j)) & 1ull;
            const cplx v0 = skip ? cplx(1.0, 0.0) : cplx(pv[0], pv[1]);
            const cplx v1 = skip ? cplx(1.0, 0.0) : cplx(pv[2], pv[3]);
            const size_t half = (size_t)1 << j;
            for (size_t i = 0; i < half; ++i) {
                t[i + half] = t[i] * v1;
                t[i] = t[i] * v0;
            }
        }
    };
    fill(tab.data(), 0, h);
    fill(tab.data() + nlo, h, n - h);
    std::copy(extra.begin(), extra.end(), tab.begin() + nlo + nhi);
    // pageable source: staged by the runtime before the call returns (the pinned area may still be
    // feeding the pass descriptors to the device)
    QSV_CUDA(cudaMemcpyAsync(s->prod_tab, tab.data(), bytes, cudaMemcpyHostToDevice, s->stream));
    s->ctr.h2d_bytes += bytes;
    s->prod_h = h;
    return QSV_OK;
}

int qsv_flush_product(qsv_state* s) {
    if (!s->prod_pending) return QSV_OK;
    s->prod_pending = 0;
    QSV_TRY(upload_product_tables(s, 0, {}));
    const uint64_t total = qsv_local_size(s);
    uint64_t blocks = (total + 255) / 256;
    const uint64_t cap = (uint64_t)s->num_sms * 16;
    if (blocks > cap) blocks = cap;
    const double2* d_lo = s->prod_tab;
    k_product_state<<<(unsigned)blocks, 256, 0, s->stream>>>(s->amp, total, s->prod_h, d_lo,
                                                             d_lo + ((size_t)1 << s->prod_h));
    QSV_CUDA(cudaGetLastError());
    s->ctr.kernel_launches++;
    s->ctr.write_passes++;
    return QSV_OK;
}

extern "C" int qsv_apply_gates(qsv_state* s, const qsv_gate* gates, uint64_t n_gates, int flags) {
    QSV_RANGE("qsv_apply_gates");
    if (!s || (!gates && n_gates)) QSV_FAIL(QSV_ERR_INVALID, "qsv_apply_gates: NULL argument");
    QSV_CUDA(cudaSetDevice(s->device));
    return qsv_apply_gates_impl(s, gates, n_gates, flags);
}

static int qsv_apply_gates_impl(qsv_state* s, const qsv_gate* gates, uint64_t n_gates, int flags) {
    std::vector<HGate> hg;
    hg.reserve(n_gates);
    cplx gphase(1.0, 0.0);
    for (uint64_t i = 0; i < n_gates; ++i) {
        const qsv_gate& g = gates[i];
        if (g.kind == QSV_GATE_PHASE) {
            gphase *= std::polar(1.0, g.params[0]);
            continue;
        }
        if (g.kind != QSV_GATE_MATRIX && g.kind != QSV_GATE_DIAGONAL)
            QSV_FAIL(QSV_ERR_INVALID, "qsv_apply_gates: gate %llu has unknown kind %d",
                     (unsigned long long)i, g.kind);
        QSV_TRY(validate_qargs(s, g.k, g.qargs, "qsv_apply_gates"));
        if (!g.data) QSV_FAIL(QSV_ERR_INVALID, "qsv_apply_gates: gate %llu has no data", (unsigned long long)i);
        HGate h;
        h.k = g.k;
        memset(h.q, 0, sizeof h.q);
        for (int j = 0; j < g.k; ++j) h.q[j] = g.qargs[j];  // logical; mapped to physical below
        const size_t dim = (size_t)1 << g.k;
        const cplx* src = reinterpret_cast<const cplx*>(g.data);
        if (g.kind == QSV_GATE_DIAGONAL) {
            h.diag = true;
            h.m.assign(src, src + dim);
        } else {
            h.diag = false;
            h.m.assign(src, src + dim * dim);
            normalise_diag(h);
        }
        if (h.k == 0) {  // 0-qubit "gate": a scalar
            gphase *= h.m[0];
            continue;
        }
        hg.push_back(std::move(h));
    }
    s->ctr.gates_applied += n_gates;
    if (gphase != cplx(1.0, 0.0)) {
        if (hg.empty()) return qsv_scale(s, gphase.real(), gphase.imag());
        for (cplx& v : hg[0].m) v *= gphase;
    }
    if (s->basis_pending && !g_plan_dump) {
        // |index> is still only recorded: fold the leading one-qubit gates into a product state and
        // write it in one sweep (no zero fill, no read); otherwise write |index> now
        if (!(flags & QSV_FLAG_NO_FUSION) && s->use_product_init) QSV_TRY(product_prologue(s, hg));
        if (!s->prod_pending) QSV_TRY(qsv_materialize(s));
    }
    if (hg.empty() || s->world > 1) QSV_TRY(qsv_flush_product(s));
    if (hg.empty()) return QSV_OK;
    if (s->world > 1) return qsv_dist_apply(s, hg, flags);
    const int rc = qsv_apply_hgates(s, hg, flags);
    if (rc != QSV_OK) return rc;
    return qsv_flush_product(s);  // no pass took it (cannot happen with a non-empty stream; cheap to guarantee)
}

// Host-only view of the single-GPU pass planner (no device work): writes one line per pass to
// `path`.  Used to study fusion offline; not part of the simulation path.
extern "C" int qsv_debug_plan(int n_qubits, const qsv_gate* gates, uint64_t n_gates, int flags, const char* path) {
    if (n_qubits < 1 || n_qubits > 40 || !path) QSV_FAIL(QSV_ERR_INVALID, "qsv_debug_plan: bad argument");
    FILE* f = fopen(path, "w");
    if (!f) QSV_FAIL(QSV_ERR_INVALID, "qsv_debug_plan: cannot open %s", path);
    qsv_state st;
    memset(&st, 0, sizeof st);
    st.n = st.n_local = n_qubits;
    st.world = 1;
    st.num_sms = 148;
    for (int q = 0; q < 64; ++q) st.perm[q] = q;
    qsv_state_defaults(&st);
    g_plan_dump = f;
    cudaSetDevice(0);  // may fail without a device: irrelevant here
    cudaGetLastError();
    int rc = qsv_apply_gates_impl(&st, gates, n_gates, flags);
    g_plan_dump = nullptr;
    fprintf(f, "counters: passes=%llu phases=%llu\n", (unsigned long long)st.ctr.state_passes,
            (unsigned long long)st.ctr.phases);
    fclose(f);
    return rc;
}

extern "C" int qsv_apply_matrix(qsv_state* s, const double* mat, int k, const int* qargs) {
    if (!s || !mat || (!qargs && k)) QSV_FAIL(QSV_ERR_INVALID, "qsv_apply_matrix: NULL argument");
    qsv_gate g;
    memset(&g, 0, sizeof g);
    g.kind = QSV_GATE_MATRIX;
    g.k = k;
    if (k < 0 || k > QSV_MAX_GATE_QUBITS)
        QSV_FAIL(QSV_ERR_INVALID, "qsv_apply_matrix: k = %d outside [0, %d]", k, QSV_MAX_GATE_QUBITS);
    for (int i = 0; i < k; ++i) g.qargs[i] = qargs[i];
    g.data = mat;
    return qsv_apply_gates(s, &g, 1, 0);
}

extern "C" int qsv_apply_diagonal(qsv_state* s, const double* diag, int k, const int* qargs) {
    if (!s || !diag || (!qargs && k)) QSV_FAIL(QSV_ERR_INVALID, "qsv_apply_diagonal: NULL argument");
    qsv_gate g;
    memset(&g, 0, sizeof g);
    g.kind = QSV_GATE_DIAGONAL;
    g.k = k;
    if (k < 0 || k > QSV_MAX_GATE_QUBITS)
        QSV_FAIL(QSV_ERR_INVALID, "qsv_apply_diagonal: k = %d outside [0, %d]", k, QSV_MAX_GATE_QUBITS);
    for (int i = 0; i < k; ++i) g.qargs[i] = qargs[i];
    g.data = diag;
    return qsv_apply_gates(s, &g, 1, 0);
}

// Host-only emulation of the address arithmetic of k_fused_swap_pass (tests/test_host.py): for the
// first fused pass of `gates` on a slab of n_local qubits and a pending swap (k, lpos, a), every
// destination index x of this rank gets the source it is gathered from -- src_j[x] (value of the
// swapped global bits of the source rank) and src_addr[x] (local index there) -- and tile_of[x], the
// RELATIVE tile number that owns x; sig[tile * 16 + i] lists the ranks (as j values) tile `tile`
// signals (0xffffffff = unused).  Uses the same FpPass / FpRemote builders and the same
// __host__ __device__ helpers as the kernel.  *ok = 0 when the first pass is not a fused pass.
extern "C" int qsv_debug_fused_swap(int mode, int n_local, int k, const int* lpos, int a, const qsv_gate* gates,
                                    uint64_t n_gates, uint32_t* src_j, uint64_t* src_addr, uint32_t* tile_of,
                                    uint32_t* sig, int* nin_out, int* ok) {
    if (mode < 1 || mode > 3) QSV_FAIL(QSV_ERR_INVALID, "qsv_debug_fused_swap: mode must be 1 (gather), 2 (scatter, copy-out) or 3 (scatter, direct store)");
    if (n_local < FP_T || n_local > 30 || k < 1 || k > FPR_MAXK || !lpos || !src_j || !src_addr || !tile_of || !sig || !ok)
        QSV_FAIL(QSV_ERR_INVALID, "qsv_debug_fused_swap: bad argument");
    qsv_state st;
    memset(&st, 0, sizeof st);
    st.n = st.n_local = n_local;
    st.world = 1;
    st.num_sms = 148;
    for (int q = 0; q < 64; ++q) st.perm[q] = q;
    qsv_state_defaults(&st);
    FILE* f = fopen("/dev/null", "w");
    if (!f) QSV_FAIL(QSV_ERR_INVALID, "qsv_debug_fused_swap: cannot open /dev/null");
    FpPass* P = new FpPass();
    g_plan_dump = f;
    g_capture_first = P;
    g_capture_state = 0;
    g_capture_last = mode != 1;
    cudaSetDevice(0);
    cudaGetLastError();
    int rc = qsv_apply_gates_impl(&st, gates, n_gates, 0);
    g_plan_dump = nullptr;
    g_capture_first = nullptr;
    fclose(f);
    g_capture_last = 0;
    *ok = (rc == QSV_OK && g_capture_state == 1) ? 1 : 0;
    if (*ok && mode == 3 && !(P->nphases > 0 && P->phases[P->nphases - 1].direct)) *ok = 0;  // no direct-store phase
    if (!*ok) {
        delete P;
        return rc;
    }
    QsvPendingSwap W;
    memset(&W, 0, sizeof W);
    W.k = k;
    W.a = a;
    for (int i = 0; i < k; ++i) W.lpos[i] = lpos[i];
    FpRemote R;
    build_fp_remote(*P, W, R);
    if (nin_out) *nin_out = R.nin;
    const uint64_t count = 1ull << n_local;
    for (uint64_t x = 0; x < count; ++x) tile_of[x] = 0xffffffffu;
    for (uint32_t tile = 0; tile < P->ntiles; ++tile) {
        const uint64_t base = fp_tile_base(*P, fpr_tile_number(R, tile)) ^ R.out_xor;
        const uint32_t jt = fpr_j_tile(R, base);
        for (uint32_t x = 0; x < 16; ++x) sig[(size_t)tile * 16 + x] = x < (1u << R.nin) ? fpr_signal_peer(R, jt, x) : 0xffffffffu;
        for (uint32_t t = 0; t < FP_NT && mode == 3; ++t) {   // scatter through the direct-store mapping
            const uint64_t stt = fp_st_thread(*P, t);
            const uint32_t jth = fpr_j_store_thread(R, t);
            for (int r = 0; r < 16; ++r) {
                const uint64_t own = base + stt + P->st_off[r] / sizeof(double2);
                if (own >= count || tile_of[own] != 0xffffffffu) {
                    delete P;
                    QSV_FAIL(QSV_ERR_INVALID, "qsv_debug_fused_swap: element %llu covered twice or out of range",
                             (unsigned long long)own);
                }
                tile_of[own] = tile;
                src_j[own] = jt | jth | R.st_jreg[r];
                src_addr[own] = ((base & ~R.lmask) | R.a_l) + (stt & ~R.lmask) + R.st_clean[r] / sizeof(double2);
            }
        }
        for (uint32_t t = 0; t < FP_NT && mode != 3; ++t) {   // copy-in (gather) / copy-out (scatter) mapping
            const uint64_t cp = fp_cp_off(*P, t);
            const uint32_t jth = fpr_j_thread(R, t);
            for (int r = 0; r < 16; ++r) {
                const uint64_t dest = base + cp + P->row_off[r] / sizeof(double2);
                if (dest >= count || tile_of[dest] != 0xffffffffu) {
                    delete P;
                    QSV_FAIL(QSV_ERR_INVALID, "qsv_debug_fused_swap: destination %llu covered twice or out of range",
                             (unsigned long long)dest);
                }
                tile_of[dest] = tile;
                src_j[dest] = jt | jth | R.jrow[r];
                src_addr[dest] = ((base & ~R.lmask) | R.a_l) + (cp & ~R.lmask) + R.row_clean[r] / sizeof(double2);
            }
        }
    }
    delete P;
    return QSV_OK;
}
