// k_fused_pass: the fused-pass kernel (second generation of k_cluster_pass, qsv_gate.cu).
//
// ncu on k_cluster_pass (profiles/r1_ncu_cluster.txt) showed 8 resident warps per SM stalled on
// the tile load, 145-245 thread instructions per amplitude of which only 32 were FP64 (64
// register moves per sub-op from the dynamic dispatch, per-thread loads of uniform descriptors,
// bit-by-bit index arithmetic) and instruction-cache misses.  This kernel keeps the idea -- a
// pass = tile of 2^12 amplitudes in shared memory, cut into phases that work on 4 "register
// bits" -- and changes the execution model:
//   * persistent CTAs (one per SM, 256 threads, 16 amplitudes per thread) walk the tiles with a
//     two-deep cp.async pipeline: tile i+1 streams into the second buffer while tile i computes;
//   * everything uniform (phase and sub-op descriptors, gate matrices) lives in the kernel
//     parameter (constant bank), everything per-thread (slot of the thread in a phase, phase-table
//     index of a diagonal op) is a 16-bit table built on the host and cached in shared memory;
//   * permutation gates (CX, SWAP, X, CCX, ...) are never executed: a phase is
//     [permutations] [dense / diagonal ops] [permutations] and the permutations are folded into
//     the shared-memory slot each register is read from / written to (16 XOR masks per phase);
//   * dense ops update the register file in place (no copies across the dispatch loop);
//   * when the last phase leaves the four lowest tile bits to the threads, registers are stored
//     straight to global memory (256-byte runs) instead of through shared memory.
#pragma once

#include "qsv_common.cuh"

#define FP_T 12
#define FP_NT 256
#define FP_TILE (1 << FP_T)
#define FP_MAX_PHASES 12
#define FP_MAX_SUBS 112
#define FP_MAX_MAT 1408  // doubles of matrix data per pass
#define FP_MAX_TABS 48   // per-thread u16 tables (one per phase + one per diagonal op)
#define FP_MAX_DIAGS 40  // diagonal ops per pass
#define FP_NOCOEF 0xffffffffu

enum { FP_D1C = 0, FP_D1R = 1, FP_D2C = 2, FP_D2R = 3, FP_DIAG = 4, FP_DIAGU = 5, FP_D1L = 6, FP_D1RY = 7, FP_D1H = 8, FP_D1HY = 9,
       FP_DIAGB = 10, FP_D1L4 = 11, FP_D1H4 = 12, FP_D2C2 = 13, FP_D1R4 = 14, FP_DIAGUS = 15 };

struct FpPhase {           // 80 bytes
    uint16_t rd[16];       // register r is read from byte offset (thread offset ^ rd[r]) of the tile buffer
    uint16_t wr[16];       // ... and written to (thread offset ^ wr[r]); slot masks pre-multiplied by 16
    uint32_t rcoef, wcoef; // 16 complex factors applied after the read / before the write (gdata offset)
    uint16_t sub_first;
    uint8_t nsub;
    uint8_t direct;        // 1: registers go straight to global memory (last phase only)
    uint16_t tab;          // per-thread table: swizzled slot of the thread in this phase
    uint16_t pad;
};

struct FpSub {             // 8 bytes
    uint8_t kind;          // FP_*
    uint8_t b;             // register bits: b0 | b1 << 2 (ascending)
    uint16_t mat;          // offset into mats (doubles); FP_DIAG: index into diags
    uint32_t pad;
};

struct FpDiag {            // 64 bytes, kernel parameter (uniform: the off-tile bits are gathered once per warp)
    uint32_t table_off;    // into gdata (double2 units)
    uint16_t tab;          // per-thread table: table index bits contributed by the thread
    uint8_t nglob;         // table index bits fed by index bits outside the tile
    uint8_t all_same;      // 1: no register bit feeds the table (one lookup per thread)
    uint16_t regpart[16];  // table index bits contributed by register r
    uint8_t gbit[10];      // global index bit ...
    uint8_t gj[10];        // ... feeding table index bit gj
    uint32_t pad;
};
static_assert(sizeof(FpDiag) == 64, "FpDiag layout");

struct FpPass {
    uint64_t index_base;   // sharded states: rank << n_local
    uint32_t ntiles;
    int32_t nphases;
    int32_t ntabs;
    int32_t warp_local;    // phases exchange data inside warps only (__syncwarp between phases)
    // gen_h > 0: product-state prologue folded into this (first) pass.  No tile is read: the state
    // is a product over qubits, so every tile is the same 2^12 amplitudes (the product over the tile's
    // qubits, gen_tile, in tile-index order) times one scalar,
    //   gen_lo[base mod 2^gen_h] * gen_hi[base div 2^gen_h]  (qubits outside the tile; base = tile base).
    // The copy-in streams gen_tile (64 KB, L2-resident) instead of the state; phase 0 applies the scalar.
    int32_t gen_h;
    const double2* gen_lo;
    const double2* gen_hi;
    const double2* gen_tile;
    int32_t store_prob;    // direct-store last phase writes |a|^2 only: amplitude i -> double (i >> 4) * 32 + (i & 15),
                           // i.e. the first half of the 256-byte run the tile itself owns (in place, no hazard)
    int32_t pos[FP_T];     // ascending global bit position of tile bit b
    const uint16_t* tabs;  // [ntabs][256]
    const double2* gdata;  // diagonal tables and coefficient vectors
    uint64_t st_off[16];   // direct store: global BYTE offset of register r
    uint64_t row_off[16];  // copy-in / copy-out: global BYTE offset of tile row r (256 elements per row)
    int32_t nseg;          // tile index -> tile base: base |= ((tile >> src) & mask) << dst per segment
    uint8_t seg_src[16], seg_dst[16];
    uint32_t seg_mask[16];
    uint8_t st_tpos[8];    // direct store: tile bit of thread bit j
    uint16_t out_tab;      // copy-out: per-thread slot table (0xffff: slot = fp_swz(t), the copy-in layout)
    uint16_t out_row[16];  // copy-out: slot mask of tile row r (pre-multiplied by 16)
    FpPhase phases[FP_MAX_PHASES];
    FpSub subs[FP_MAX_SUBS];
    double mats[FP_MAX_MAT];
    FpDiag diags[FP_MAX_DIAGS];
};

__host__ __device__ __forceinline__ uint32_t fp_swz(uint32_t i) {
    return i ^ (((i >> 3) ^ (i >> 6) ^ (i >> 9)) & 7u);
}

// ---- fused swap pass: a global<->local qubit swap riding inside the copy-in of a fused pass ------
//
// Sharded states (qsv_dist.cu).  Swapping k global bits with the local bits lpos[0..k) means: the
// amplitude that ends up at local index x of this rank (swapped global bits = a) sits, before the
// swap, on the rank whose swapped global bits equal the bits of x at lpos[] (call that value j), at
// the local index x with those bits replaced by a.  Instead of moving the data first and sweeping
// it again, the first fused pass after the swap gathers its tiles straight from the peers' slabs
// (CUDA-IPC mappings, NVLink) and stores them -- in place -- into its own slab.
//
// In place across ranks needs an ordering: the region a tile of rank a overwrites is exactly what the
// matching tiles of the ranks it read from are reading.  "Matching" tiles carry the same tile NUMBER
// on every rank because tiles are enumerated relative to the rank: the swapped local bits that are
// not tile bits are XORed with the rank's own bits (out_xor).  Every CTA therefore (1) waits for its
// tile to land, (2) adds 1 to the arrival counter of that tile number on every rank it read from,
// (3) computes, (4) spins until its own counter for the tile has reached `target` (all readers of its
// region have their data), (5) stores.  All ranks launch the same grid over the same tile numbers, all
// CTAs of the grid are resident (2 per SM), a CTA signals tile i before it waits for tile i, and loads
// never wait: no cycle.  A clock64 timeout turns a would-be hang into an error word.
#define FPR_MAXK 4
struct FpRemote {
    const double2* peer[1 << FPR_MAXK];  // slab of the rank whose swapped global bits read j (peer[a]: own)
    uint32_t* flag[1 << FPR_MAXK];       // arrival counters (one per tile number) of that rank
    uint32_t* err;                       // set to 1 when a handshake times out
    uint32_t target;                     // value every counter of this rank reaches during this pass
    int32_t a;                           // this rank's value on the swapped global bits
    int32_t k, nin, nout;
    uint8_t in_j[FPR_MAXK];              // swapped local bits that are tile bits: their bit in j
    uint8_t out_j[FPR_MAXK];             // swapped local bits that are not: their bit in j ...
    uint8_t out_pos[FPR_MAXK];           // ... and their position in the local index
    int8_t thr_bit[FPR_MAXK];            // [i]: copy-in thread bit (0..7) carrying swapped local bit i, or -1
    uint8_t jrow[16];                    // bits of j contributed by tile row r
    uint64_t row_clean[16];              // byte offset of tile row r with the swapped bits cleared
    uint64_t lmask;                      // local-index mask of the swapped local bits
    uint64_t a_l;                        // a scattered to those positions
    uint64_t out_xor;                    // a's bits at the non-tile swapped positions (relative tile order)
    long long timeout;                   // clock64 ticks a CTA waits before it reports an error
    // Rows (4 KB each) of the NEXT tile that are prefetched into a staging area behind the tile buffer
    // while the current tile computes: with one buffer per CTA every CTA loads, then computes, and all
    // CTAs of all ranks do so in step (the handshake keeps them together), so the links idle during
    // the arithmetic (measured: 327 GB/s per direction at k = 1).  The staged rows keep ~2/3 of a tile
    // per CTA in flight during the phases; they are moved into the tile buffer (each thread moves the
    // elements it loaded itself) once the current tile has left it.
    int32_t stage_rows;
    // ---- scatter mode (k_fused_swap_pass<2>): the LAST fused pass before the swap reads its own slab and
    // stores every tile into the slabs of the ranks that own it after the swap.  Remote WRITES are posted
    // (measured: a B200 sustains ~340 GB/s of remote reads but about twice that of remote writes), and the
    // same handshake orders them: a rank signals that it has loaded tile i before anybody may overwrite it.
    // Direct-store phase: register r of thread t goes to rank j = j_tile | j_thread | st_jreg[r].
    // Tile order.  The swapped bits outside the tile are high index bits, i.e. the top bits of the tile
    // number: walked in plain order a rank would first do all its local tiles (links idle) and then all
    // remote ones (links saturated, SMs waiting).  The tile COUNTER therefore has those bits exchanged
    // with counter bits just above the grid size (tswap_hi[i] <-> tswap_lo[i]), so every CTA alternates
    // between local and remote tiles.  The same bijection on every rank: counters still match.
    int32_t ntswap;
    uint8_t tswap_lo[FPR_MAXK], tswap_hi[FPR_MAXK];
    int8_t st_thr_bit[FPR_MAXK];         // [i]: direct-store thread bit (0..7) carrying swapped local bit i, or -1
    uint8_t st_jreg[16];                 // bits of j contributed by register r
    uint64_t st_clean[16];               // P.st_off[r] with the swapped bits cleared
};

__host__ __device__ __forceinline__ uint64_t fp_tile_base(const FpPass& P, uint32_t tile) {
    uint64_t base = 0;
    for (int sgm = 0; sgm < P.nseg; ++sgm)
        base |= (uint64_t)((tile >> P.seg_src[sgm]) & P.seg_mask[sgm]) << P.seg_dst[sgm];
    return base;
}
// tile counter -> tile number (see FpRemote::ntswap)
__host__ __device__ __forceinline__ uint32_t fpr_tile_number(const FpRemote& R, uint32_t counter) {
    uint32_t tnum = counter;
    for (int i = 0; i < R.ntswap; ++i) {
        const uint32_t x = ((tnum >> R.tswap_lo[i]) ^ (tnum >> R.tswap_hi[i])) & 1u;
        tnum ^= (x << R.tswap_lo[i]) | (x << R.tswap_hi[i]);
    }
    return tnum;
}
// bits of j fixed by the tile (swapped local bits outside the tile)
__host__ __device__ __forceinline__ uint32_t fpr_j_tile(const FpRemote& R, uint64_t tile_base) {
    uint32_t j = 0;
    for (int i = 0; i < R.nout; ++i) j |= (uint32_t)((tile_base >> R.out_pos[i]) & 1ull) << R.out_j[i];
    return j;
}
// bits of j carried by copy-in thread t
__host__ __device__ __forceinline__ uint32_t fpr_j_thread(const FpRemote& R, uint32_t t) {
    uint32_t j = 0;
    for (int i = 0; i < R.k; ++i)
        if (R.thr_bit[i] >= 0) j |= ((t >> R.thr_bit[i]) & 1u) << i;
    return j;
}
// bits of j carried by direct-store thread t (scatter mode)
__host__ __device__ __forceinline__ uint32_t fpr_j_store_thread(const FpRemote& R, uint32_t t) {
    uint32_t j = 0;
    for (int i = 0; i < R.k; ++i)
        if (R.st_thr_bit[i] >= 0) j |= ((t >> R.st_thr_bit[i]) & 1u) << i;
    return j;
}
// direct-store thread offset (elements) of thread t
__host__ __device__ __forceinline__ uint64_t fp_st_thread(const FpPass& P, uint32_t t) {
    uint64_t o = 0;
    for (int b = 0; b < 8; ++b) o |= (uint64_t)((t >> b) & 1u) << P.pos[P.st_tpos[b]];
    return o;
}
// copy-in thread offset (elements) of thread t: tile bits 0..7 scattered to their index positions
__host__ __device__ __forceinline__ uint64_t fp_cp_off(const FpPass& P, uint32_t t) {
    uint64_t o = 0;
    for (int b = 0; b < 8; ++b) o |= (uint64_t)((t >> b) & 1u) << P.pos[b];
    return o;
}
// rank (as a value j of the swapped global bits) signalled by lane x of a tile's 2^nin signals
__host__ __device__ __forceinline__ uint32_t fpr_signal_peer(const FpRemote& R, uint32_t j_tile, uint32_t x) {
    uint32_t j = j_tile;
    for (int b = 0; b < R.nin; ++b) j |= ((x >> b) & 1u) << R.in_j[b];
    return j;
}

// ---- dense ops on the 16 registers of a thread, in place --------------------------------
template <int B>
__device__ __forceinline__ void fp_d1c(double2 (&v)[16], const FpPass& P, int mo) {
    const double m00r = P.mats[mo + 0], m00i = P.mats[mo + 1], m01r = P.mats[mo + 2], m01i = P.mats[mo + 3];
    const double m10r = P.mats[mo + 4], m10i = P.mats[mo + 5], m11r = P.mats[mo + 6], m11i = P.mats[mo + 7];
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        if (i & (1 << B)) continue;
        const double2 a0 = v[i], a1 = v[i | (1 << B)];
        double tr = m00r * a0.x;
        tr = fma(-m00i, a0.y, tr);
        tr = fma(m01r, a1.x, tr);
        tr = fma(-m01i, a1.y, tr);
        double ti = m00r * a0.y;
        ti = fma(m00i, a0.x, ti);
        ti = fma(m01r, a1.y, ti);
        ti = fma(m01i, a1.x, ti);
        double ur = m10r * a0.x;
        ur = fma(-m10i, a0.y, ur);
        ur = fma(m11r, a1.x, ur);
        ur = fma(-m11i, a1.y, ur);
        double ui = m10r * a0.y;
        ui = fma(m10i, a0.x, ui);
        ui = fma(m11r, a1.y, ui);
        ui = fma(m11i, a1.x, ui);
        v[i] = make_double2(tr, ti);
        v[i | (1 << B)] = make_double2(ur, ui);
    }
}

template <int B>
__device__ __forceinline__ void fp_d1r(double2 (&v)[16], const FpPass& P, int mo) {
    const double m00 = P.mats[mo + 0], m01 = P.mats[mo + 1], m10 = P.mats[mo + 2], m11 = P.mats[mo + 3];
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        if (i & (1 << B)) continue;
        const double2 a0 = v[i], a1 = v[i | (1 << B)];
        v[i] = make_double2(fma(m01, a1.x, m00 * a0.x), fma(m01, a1.y, m00 * a0.y));
        v[i | (1 << B)] = make_double2(fma(m11, a1.x, m10 * a0.x), fma(m11, a1.y, m10 * a0.y));
    }
}

// R * diag(1, -i) with R real (H S^dagger, the change to the Y basis): -i only swaps and negates the
// components of the second amplitude, so the gate costs what the real R costs
template <int B>
__device__ __forceinline__ void fp_d1ry(double2 (&v)[16], const FpPass& P, int mo) {
    const double m00 = P.mats[mo + 0], m01 = P.mats[mo + 1], m10 = P.mats[mo + 2], m11 = P.mats[mo + 3];
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        if (i & (1 << B)) continue;
        const double2 a0 = v[i], a1 = make_double2(v[i | (1 << B)].y, -v[i | (1 << B)].x);
        v[i] = make_double2(fma(m01, a1.x, m00 * a0.x), fma(m01, a1.y, m00 * a0.y));
        v[i | (1 << B)] = make_double2(fma(m11, a1.x, m10 * a0.x), fma(m11, a1.y, m10 * a0.y));
    }
}

// unnormalised Hadamard butterfly (1 1 ; 1 -1): additions only.  Basis-change expectation values
// use it when only |a|^2 of the result is read: the factor 2^-k of k butterflies is exact and is
// applied to the term values on the host.  Y = true: on (a0, -i a1), the Y-basis change.
template <int B, bool Y>
__device__ __forceinline__ void fp_d1h(double2 (&v)[16]) {
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        if (i & (1 << B)) continue;
        const double2 a0 = v[i];
        const double2 a1 = Y ? make_double2(v[i | (1 << B)].y, -v[i | (1 << B)].x) : v[i | (1 << B)];
        v[i] = make_double2(a0.x + a1.x, a0.y + a1.y);
        v[i | (1 << B)] = make_double2(a0.x - a1.x, a0.y - a1.y);
    }
}

// plane rotation (c -s ; s c), c >= 0, as three shears (lifting): x += t y; y += s x; x += t y with
// t = -s / (1 + c) -- 3 FMA per real pair instead of 4 operations; |t| <= 1, so no amplification
template <int B>
__device__ __forceinline__ void fp_d1l(double2 (&v)[16], const FpPass& P, int mo) {
    const double t = P.mats[mo + 0], s = P.mats[mo + 1];
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        if (i & (1 << B)) continue;
        double2 a0 = v[i], a1 = v[i | (1 << B)];
        a0.x = fma(t, a1.x, a0.x);
        a0.y = fma(t, a1.y, a0.y);
        a1.x = fma(s, a0.x, a1.x);
        a1.y = fma(s, a0.y, a1.y);
        a0.x = fma(t, a1.x, a0.x);
        a0.y = fma(t, a1.y, a0.y);
        v[i] = a0;
        v[i | (1 << B)] = a1;
    }
}

template <int B0, int B1>
__device__ __forceinline__ void fp_d2c(double2 (&v)[16], const FpPass& P, int mo) {
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        if (i & ((1 << B0) | (1 << B1))) continue;
        const double2 a[4] = {v[i], v[i | (1 << B0)], v[i | (1 << B1)], v[i | (1 << B0) | (1 << B1)]};
        double2 o[4];
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            double re = 0.0, im = 0.0;
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const double mr = P.mats[mo + 8 * r + 2 * c], mi = P.mats[mo + 8 * r + 2 * c + 1];
                if (c == 0) {
                    re = mr * a[0].x;
                    im = mr * a[0].y;
                } else {
                    re = fma(mr, a[c].x, re);
                    im = fma(mr, a[c].y, im);
                }
                re = fma(-mi, a[c].y, re);
                im = fma(mi, a[c].x, im);
            }
            o[r] = make_double2(re, im);
        }
        v[i] = o[0];
        v[i | (1 << B0)] = o[1];
        v[i | (1 << B1)] = o[2];
        v[i | (1 << B0) | (1 << B1)] = o[3];
    }
}

template <int B0, int B1>
__device__ __forceinline__ void fp_d2r(double2 (&v)[16], const FpPass& P, int mo) {
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        if (i & ((1 << B0) | (1 << B1))) continue;
        const double2 a[4] = {v[i], v[i | (1 << B0)], v[i | (1 << B1)], v[i | (1 << B0) | (1 << B1)]};
        double2 o[4];
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const double m0 = P.mats[mo + 4 * r], m1 = P.mats[mo + 4 * r + 1], m2 = P.mats[mo + 4 * r + 2],
                         m3 = P.mats[mo + 4 * r + 3];
            o[r] = make_double2(fma(m3, a[3].x, fma(m2, a[2].x, fma(m1, a[1].x, m0 * a[0].x))),
                                fma(m3, a[3].y, fma(m2, a[2].y, fma(m1, a[1].y, m0 * a[0].y))));
        }
        v[i] = o[0];
        v[i | (1 << B0)] = o[1];
        v[i | (1 << B1)] = o[2];
        v[i | (1 << B0) | (1 << B1)] = o[3];
    }
}

// v *= c without a register copy: the two products by c.y go to temporaries, the FMAs land in place
__device__ __forceinline__ void fp_cmul_inplace(double2& v, const double2 c) {
    const double t = c.y * v.y, u = c.y * v.x;
    v.x = fma(c.x, v.x, -t);
    v.y = fma(c.x, v.y, u);
}

__device__ __forceinline__ void fp_diag(double2 (&v)[16], const FpPass& P, int di, const uint16_t* __restrict__ tabs_sm,
                                        uint64_t tile_base, uint32_t t) {
    const FpDiag& d = P.diags[di];
    uint32_t glob = 0;  // uniform: table index bits fed by index bits outside the tile
    for (int j = 0; j < d.nglob; ++j) glob |= (uint32_t)((tile_base >> d.gbit[j]) & 1ull) << d.gj[j];
    const uint32_t thr = tabs_sm[d.tab * FP_NT + t] | glob;
    const double2* __restrict__ tab = P.gdata + d.table_off + thr;
    if (d.all_same) {
        const double2 c = __ldg(tab);
#pragma unroll
        for (int r = 0; r < 16; ++r) fp_cmul_inplace(v[r], c);
    } else {
#pragma unroll
        for (int r = 0; r < 16; ++r) fp_cmul_inplace(v[r], __ldg(tab + d.regpart[r]));
    }
}

// v[r] *= c for the registers whose bit B is set
template <int B>
__device__ __forceinline__ void fp_mul_bit(double2 (&v)[16], const double2 c) {
#pragma unroll
    for (int r = 0; r < 16; ++r)
        if (r & (1 << B)) fp_cmul_inplace(v[r], c);
}
__device__ __forceinline__ void fp_mul_bit_dyn(double2 (&v)[16], int b, const double2 c) {
    switch (b) {
        case 0: fp_mul_bit<0>(v, c); break;
        case 1: fp_mul_bit<1>(v, c); break;
        case 2: fp_mul_bit<2>(v, c); break;
        default: fp_mul_bit<3>(v, c); break;
    }
}

// FP_DIAGB, a "bilinear" phase table.  A product of controlled-phase / RZZ-like gates has a quadratic
// log-phase, so its table over (register bits r, other bits o) factors as
//     T[r, o] = T0[o] * U[r] * prod_{a in r} C_a[o]
// (checked numerically on the host, entry by entry).  U is the uniform 16-entry factor (an FP_DIAGU op);
// this op applies T0 and the C_a: the thread reads 1 + na CONSECUTIVE entries of row o instead of the 16
// scattered 16-byte entries FP_DIAG needs, and spends 16 + 8 (na - 1) + 1 complex multiplications.
// FpDiag fields reused: regpart[0] = na, regpart[1 + a] = register bit of factor a, regpart[5] = row stride.
__device__ __forceinline__ void fp_diagb(double2 (&v)[16], const FpPass& P, int di, const uint16_t* __restrict__ tabs_sm,
                                         uint64_t tile_base, uint32_t t) {
    const FpDiag& d = P.diags[di];
    uint32_t glob = 0;
    for (int j = 0; j < d.nglob; ++j) glob |= (uint32_t)((tile_base >> d.gbit[j]) & 1ull) << d.gj[j];
    const uint32_t row = tabs_sm[d.tab * FP_NT + t] | glob;
    const double2* __restrict__ tab = P.gdata + d.table_off + row * d.regpart[5];
    const int na = d.regpart[0];
    const double2 c0 = __ldg(tab);
    const double2 c1 = cmul(c0, __ldg(tab + 1));   // na >= 1 by construction
    const int b0 = d.regpart[1];
    // registers with bit b0 clear take T0, the others T0 * C_0: one multiplication per register
    switch (b0) {
        case 0:
#pragma unroll
            for (int r = 0; r < 16; ++r) fp_cmul_inplace(v[r], (r & 1) ? c1 : c0);
            break;
        case 1:
#pragma unroll
            for (int r = 0; r < 16; ++r) fp_cmul_inplace(v[r], (r & 2) ? c1 : c0);
            break;
        case 2:
#pragma unroll
            for (int r = 0; r < 16; ++r) fp_cmul_inplace(v[r], (r & 4) ? c1 : c0);
            break;
        default:
#pragma unroll
            for (int r = 0; r < 16; ++r) fp_cmul_inplace(v[r], (r & 8) ? c1 : c0);
            break;
    }
    for (int a = 1; a < na; ++a) fp_mul_bit_dyn(v, d.regpart[1 + a], __ldg(tab + 1 + a));
}

__device__ __forceinline__ void fp_apply(double2 (&v)[16], const FpPass& P, int si, const uint16_t* __restrict__ tabs_sm,
                                         uint64_t tile_base, uint32_t t) {
    const int kind = P.subs[si].kind, b = P.subs[si].b, mo = P.subs[si].mat;
    switch (kind * 16 + b) {
        case FP_D1C * 16 + 0: fp_d1c<0>(v, P, mo); break;
        case FP_D1C * 16 + 1: fp_d1c<1>(v, P, mo); break;
        case FP_D1C * 16 + 2: fp_d1c<2>(v, P, mo); break;
        case FP_D1C * 16 + 3: fp_d1c<3>(v, P, mo); break;
        case FP_D1R * 16 + 0: fp_d1r<0>(v, P, mo); break;
        case FP_D1R * 16 + 1: fp_d1r<1>(v, P, mo); break;
        case FP_D1R * 16 + 2: fp_d1r<2>(v, P, mo); break;
        case FP_D1R * 16 + 3: fp_d1r<3>(v, P, mo); break;
        case FP_D1RY * 16 + 0: fp_d1ry<0>(v, P, mo); break;
        case FP_D1RY * 16 + 1: fp_d1ry<1>(v, P, mo); break;
        case FP_D1RY * 16 + 2: fp_d1ry<2>(v, P, mo); break;
        case FP_D1RY * 16 + 3: fp_d1ry<3>(v, P, mo); break;
        case FP_D1H * 16 + 0: fp_d1h<0, false>(v); break;
        case FP_D1H * 16 + 1: fp_d1h<1, false>(v); break;
        case FP_D1H * 16 + 2: fp_d1h<2, false>(v); break;
        case FP_D1H * 16 + 3: fp_d1h<3, false>(v); break;
        case FP_D1HY * 16 + 0: fp_d1h<0, true>(v); break;
        case FP_D1HY * 16 + 1: fp_d1h<1, true>(v); break;
        case FP_D1HY * 16 + 2: fp_d1h<2, true>(v); break;
        case FP_D1HY * 16 + 3: fp_d1h<3, true>(v); break;
        case FP_D1L * 16 + 0: fp_d1l<0>(v, P, mo); break;
        case FP_D1L * 16 + 1: fp_d1l<1>(v, P, mo); break;
        case FP_D1L * 16 + 2: fp_d1l<2>(v, P, mo); break;
        case FP_D1L * 16 + 3: fp_d1l<3>(v, P, mo); break;
        case FP_D1L4 * 16:  // four lifted rotations, one per register bit, behind ONE dispatch: a rotation is only
                            // 48 DFMA per thread, and every dispatch costs a dependent descriptor -> branch -> constants chain
            fp_d1l<0>(v, P, mo);
            fp_d1l<1>(v, P, mo + 2);
            fp_d1l<2>(v, P, mo + 4);
            fp_d1l<3>(v, P, mo + 6);
            break;
        case FP_D1H4 * 16:  // unnormalised Hadamard butterflies on all four register bits (X-basis change of a phase)
            fp_d1h<0, false>(v);
            fp_d1h<1, false>(v);
            fp_d1h<2, false>(v);
            fp_d1h<3, false>(v);
            break;
        case FP_D1H4 * 16 + 1:  // ... the Y-basis change
            fp_d1h<0, true>(v);
            fp_d1h<1, true>(v);
            fp_d1h<2, true>(v);
            fp_d1h<3, true>(v);
            break;
        case FP_D1R4 * 16:  // four real 2x2 gates, one per register bit
            fp_d1r<0>(v, P, mo);
            fp_d1r<1>(v, P, mo + 4);
            fp_d1r<2>(v, P, mo + 8);
            fp_d1r<3>(v, P, mo + 12);
            break;
        case FP_D2C * 16 + 4: fp_d2c<0, 1>(v, P, mo); break;
        case FP_D2C * 16 + 8: fp_d2c<0, 2>(v, P, mo); break;
        case FP_D2C * 16 + 12: fp_d2c<0, 3>(v, P, mo); break;
        case FP_D2C * 16 + 9: fp_d2c<1, 2>(v, P, mo); break;
        case FP_D2C * 16 + 13: fp_d2c<1, 3>(v, P, mo); break;
        case FP_D2C * 16 + 14: fp_d2c<2, 3>(v, P, mo); break;
        case FP_D2R * 16 + 4: fp_d2r<0, 1>(v, P, mo); break;
        case FP_D2R * 16 + 8: fp_d2r<0, 2>(v, P, mo); break;
        case FP_D2R * 16 + 12: fp_d2r<0, 3>(v, P, mo); break;
        case FP_D2R * 16 + 9: fp_d2r<1, 2>(v, P, mo); break;
        case FP_D2R * 16 + 13: fp_d2r<1, 3>(v, P, mo); break;
        case FP_D2R * 16 + 14: fp_d2r<2, 3>(v, P, mo); break;
        case FP_DIAGU * 16: {  // register-bit factor of a separable diagonal: 16 uniform coefficients
#pragma unroll
            for (int r = 0; r < 16; ++r) fp_cmul_inplace(v[r], make_double2(P.mats[mo + 2 * r], P.mats[mo + 2 * r + 1]));
            break;
        }
        case FP_DIAGUS * 16: {
            // a separable diagonal in one dispatch: the per-thread scalar of its non-register bits (table
            // `b2`, all_same) is requested first, the 16 uniform register-bit coefficients are applied while
            // the look-up is in flight, then the scalar
            const FpDiag& d = P.diags[P.subs[si].pad];
            uint32_t glob = 0;
            for (int j = 0; j < d.nglob; ++j) glob |= (uint32_t)((tile_base >> d.gbit[j]) & 1ull) << d.gj[j];
            const double2 c = __ldg(P.gdata + d.table_off + (tabs_sm[d.tab * FP_NT + t] | glob));
#pragma unroll
            for (int r = 0; r < 16; ++r) fp_cmul_inplace(v[r], make_double2(P.mats[mo + 2 * r], P.mats[mo + 2 * r + 1]));
#pragma unroll
            for (int r = 0; r < 16; ++r) fp_cmul_inplace(v[r], c);
            break;
        }
        case FP_DIAGB * 16: fp_diagb(v, P, mo, tabs_sm, tile_base, t); break;
        default: fp_diag(v, P, mo, tabs_sm, tile_base, t); break;
    }
}

// spin until this rank's arrival counter of tile `tile` has reached the pass's target
__device__ __forceinline__ void fpr_wait(const FpRemote& R, uint32_t tile) {
    const volatile uint32_t* f = R.flag[R.a] + tile;
    const long long t0 = clock64();
    uint32_t spins = 0;
    while (true) {
        uint32_t v;
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];\n" : "=r"(v) : "l"(f) : "memory");
        if ((int32_t)(v - R.target) >= 0) break;
        if ((++spins & 63u) == 0 && (*reinterpret_cast<const volatile uint32_t*>(R.err) || clock64() - t0 > R.timeout)) {
            atomicExch(R.err, 1u);  // one late CTA is enough: everybody else stops waiting too
            break;
        }
    }
}

// dynamic smem: [4096] double2 tile buffer, then [ntabs][256] u16 tables.  Two CTAs per SM: while
// one waits for its tile or at a barrier the other computes; in a direct-store last phase the
// buffer is free as soon as every thread has read its registers, so the next tile's cp.async
// copies are issued before that phase's arithmetic.
// `src` = `amp` for the usual in-place pass; a different `src` makes the pass out of place (tiles are
// read from src and written to amp), which saves the scratch copy of basis-change expectation values.
// MODE 0: plain pass.  MODE 1: fused swap pass, gather (tiles read from the peers' slabs, see FpRemote).
// MODE 2: fused swap pass, scatter (tiles stored into the peers' slabs).
template <int MODE>
__device__ __forceinline__ void fp_pass_body(double2* amp, const double2* src, const FpPass& P, const FpRemote& R) {
    constexpr bool REMOTE = MODE != 0, GATHER = MODE == 1, SCATTER = MODE == 2;
    extern __shared__ __align__(16) unsigned char fp_smem[];
    unsigned char* const tilebuf = fp_smem;
    const int stage_rows = GATHER ? R.stage_rows : 0;
    unsigned char* const stagebuf = fp_smem + FP_TILE * sizeof(double2);  // REMOTE only: [stage_rows][4 KB]
    uint16_t* const tabs_sm = reinterpret_cast<uint16_t*>(fp_smem + FP_TILE * sizeof(double2) + (size_t)stage_rows * 4096);
    const uint32_t t = threadIdx.x;

    for (int i = t; i < P.ntabs * (FP_NT / 2); i += FP_NT)   // two u16 per word
        reinterpret_cast<uint32_t*>(tabs_sm)[i] = __ldg(reinterpret_cast<const uint32_t*>(P.tabs) + i);

    // global offset of this thread's element of a 256-element row (copy-in / copy-out mapping)
    uint64_t cp_off = 0;
#pragma unroll
    for (int b = 0; b < 8; ++b) cp_off |= (uint64_t)((t >> b) & 1u) << P.pos[b];
    const uint32_t st16 = fp_swz(t) << 4;
    // direct-store mapping of the last phase
    uint64_t st_thread = 0;
#pragma unroll
    for (int b = 0; b < 8; ++b) st_thread |= (uint64_t)((t >> b) & 1u) << P.pos[P.st_tpos[b]];
    // fused swap pass: this thread's share of the source-rank selector and its cleaned offset
    const uint32_t rj_thr = REMOTE ? fpr_j_thread(R, t) : 0u;
    const uint64_t rcp_clean = REMOTE ? (cp_off & ~R.lmask) : 0ull;
    const uint32_t sj_thr = SCATTER ? fpr_j_store_thread(R, t) : 0u;
    const uint64_t sst_clean = SCATTER ? (st_thread & ~R.lmask) : 0ull;

    auto tile_base_of = [&](uint32_t tile) {
        const uint64_t base = fp_tile_base(P, REMOTE ? fpr_tile_number(R, tile) : tile);
        return REMOTE ? (base ^ R.out_xor) : base;
    };
    // REMOTE: rows [r0, r1) of a tile, gathered from the ranks that hold them, into `buf`
    auto issue_remote_rows = [&](uint64_t tile_base, int r0, int r1, unsigned char* buf) {
        const uint32_t dst0 = (uint32_t)__cvta_generic_to_shared(buf);
        const uint32_t jt = fpr_j_tile(R, tile_base) | rj_thr;
        const uint64_t off = ((tile_base & ~R.lmask) | R.a_l) + rcp_clean;
#pragma unroll
        for (int r = 0; r < 16; ++r) {
            if (r < r0 || r >= r1) continue;
            const uint32_t dst = dst0 + (st16 ^ (fp_swz((uint32_t)r << 8) << 4));
            const char* g = reinterpret_cast<const char*>(R.peer[jt | R.jrow[r]] + off) + R.row_clean[r];
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(dst), "l"(g));
        }
        asm volatile("cp.async.commit_group;\n" ::: "memory");
    };
    // REMOTE, buffer free: the rows of the next tile that were not staged go straight into the tile
    // buffer; the staged ones (complete once at most that newest group is pending) are moved over
    auto finish_remote_load = [&](uint64_t next_base) {
        issue_remote_rows(next_base, stage_rows, 16, tilebuf);
        asm volatile("cp.async.wait_group 1;\n" ::: "memory");
#pragma unroll
        for (int r = 0; r < 16; ++r) {
            if (r >= stage_rows) continue;
            const uint32_t o = st16 ^ (fp_swz((uint32_t)r << 8) << 4);
            *reinterpret_cast<double2*>(tilebuf + o) = *reinterpret_cast<const double2*>(stagebuf + o);
        }
    };
    auto issue_load = [&](uint64_t tile_base) {
        const uint32_t dst0 = (uint32_t)__cvta_generic_to_shared(tilebuf);
        if (GATHER) {
            issue_remote_rows(tile_base, 0, 16, tilebuf);
            return;
        } else if (!REMOTE && P.gen_h > 0) {
            const char* g = reinterpret_cast<const char*>(P.gen_tile + t);
#pragma unroll
            for (int r = 0; r < 16; ++r) {
                const uint32_t dst = dst0 + (st16 ^ (fp_swz((uint32_t)r << 8) << 4));
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(dst), "l"(g + r * (FP_NT * 16)));
            }
        } else {
            const char* g = reinterpret_cast<const char*>(src + tile_base + cp_off);
#pragma unroll
            for (int r = 0; r < 16; ++r) {
                const uint32_t dst = dst0 + (st16 ^ (fp_swz((uint32_t)r << 8) << 4));
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(dst), "l"(g + P.row_off[r]));
            }
        }
        asm volatile("cp.async.commit_group;\n" ::: "memory");
    };

    uint32_t tile = blockIdx.x;
    if (tile >= P.ntiles) return;
    uint64_t tile_base = tile_base_of(tile);
    issue_load(tile_base);
    while (true) {
        asm volatile("cp.async.wait_group 0;\n" ::: "memory");
        __syncthreads();  // tile (and, first time, the tables) visible to every thread
        if (REMOTE && t < (1u << R.nin)) {
            // every load of this CTA has landed: tell the ranks it read from (their tile `tile`)
            __threadfence_system();
            atomicAdd_system(R.flag[fpr_signal_peer(R, fpr_j_tile(R, tile_base), t)] + tile, 1u);
        }
        const uint32_t next = tile + gridDim.x;
        const bool has_next = next < P.ntiles;
        const uint64_t next_base = has_next ? tile_base_of(next) : 0;
        if (GATHER && has_next && stage_rows > 0) issue_remote_rows(next_base, 0, stage_rows, stagebuf);
        const uint64_t diag_base = tile_base | P.index_base;
        if (!REMOTE && P.gen_h > 0 && has_next) {  // the next tile's two table entries: into L1 ahead of their use
            asm volatile("prefetch.global.L1 [%0];\n" ::"l"(P.gen_lo + (next_base & ((1ull << P.gen_h) - 1))));
            asm volatile("prefetch.global.L1 [%0];\n" ::"l"(P.gen_hi + (next_base >> P.gen_h)));
        }
        bool stored = false;
#pragma unroll 1
        for (int p = 0; p < P.nphases; ++p) {
            const FpPhase& ph = P.phases[p];
            const uint32_t sb = (uint32_t)tabs_sm[ph.tab * FP_NT + t] << 4;
            double2 v[16];
            if (!REMOTE && P.gen_h > 0 && p == 0) {
                const double2 cb = cmul(__ldg(P.gen_lo + (tile_base & ((1ull << P.gen_h) - 1))),
                                        __ldg(P.gen_hi + (tile_base >> P.gen_h)));
#pragma unroll
                for (int r = 0; r < 16; ++r)
                    v[r] = cmul(cb, *reinterpret_cast<const double2*>(tilebuf + (sb ^ ph.rd[r])));
            } else {
#pragma unroll
                for (int r = 0; r < 16; ++r) v[r] = *reinterpret_cast<const double2*>(tilebuf + (sb ^ ph.rd[r]));
            }
            if (ph.direct) {
                // every thread has its registers: the buffer can take the next tile now
                __syncthreads();
                if (has_next) {
                    if (GATHER) finish_remote_load(next_base);
                    else issue_load(next_base);
                }
                if (REMOTE) {
                    // the next tile is on its way while thread 0 waits for the peers to have read what
                    // this tile overwrites (one NVLink latency when the pass has no arithmetic to hide it)
                    if (t == 0) fpr_wait(R, tile);
                    __syncthreads();
                }
            }
            if (ph.rcoef != FP_NOCOEF) {
                const double2* __restrict__ c = P.gdata + ph.rcoef;
#pragma unroll
                for (int r = 0; r < 16; ++r) fp_cmul_inplace(v[r], __ldg(c + r));
            }
            const int s0 = ph.sub_first, s1 = s0 + ph.nsub;
#pragma unroll 1
            for (int si = s0; si < s1; ++si) fp_apply(v, P, si, tabs_sm, diag_base, t);
            if (ph.wcoef != FP_NOCOEF) {
                const double2* __restrict__ c = P.gdata + ph.wcoef;
#pragma unroll
                for (int r = 0; r < 16; ++r) fp_cmul_inplace(v[r], __ldg(c + r));
            }
            if (ph.direct) {
                char* g = reinterpret_cast<char*>(amp + tile_base + st_thread);
                if (!REMOTE && P.store_prob) {
                    g -= (st_thread & 15ull) << 3;
#pragma unroll
                    for (int r = 0; r < 16; ++r)
                        *reinterpret_cast<double*>(g + P.st_off[r] - (((P.st_off[r] >> 4) & 15ull) << 3)) =
                            v[r].x * v[r].x + v[r].y * v[r].y;
                } else if (SCATTER) {
                    const uint32_t jt = fpr_j_tile(R, tile_base) | sj_thr;
                    const uint64_t off = ((tile_base & ~R.lmask) | R.a_l) + sst_clean;
#pragma unroll
                    for (int r = 0; r < 16; ++r)
                        *reinterpret_cast<double2*>(reinterpret_cast<char*>(const_cast<double2*>(R.peer[jt | R.st_jreg[r]]) + off) +
                                                    R.st_clean[r]) = v[r];
                } else {
#pragma unroll
                    for (int r = 0; r < 16; ++r) *reinterpret_cast<double2*>(g + P.st_off[r]) = v[r];
                }
                stored = true;
            } else {
#pragma unroll
                for (int r = 0; r < 16; ++r) *reinterpret_cast<double2*>(tilebuf + (sb ^ ph.wr[r])) = v[r];
                if (P.warp_local && p + 1 < P.nphases) __syncwarp();
                else __syncthreads();
            }
        }
        if (!stored) {
            if (REMOTE) {
                if (t == 0) fpr_wait(R, tile);
                __syncthreads();
            }
            const uint32_t so16 = P.out_tab == 0xffff ? st16 : ((uint32_t)tabs_sm[P.out_tab * FP_NT + t] << 4);
            if (SCATTER) {
                const uint32_t jt = fpr_j_tile(R, tile_base) | rj_thr;
                const uint64_t off = ((tile_base & ~R.lmask) | R.a_l) + rcp_clean;
#pragma unroll
                for (int r = 0; r < 16; ++r)
                    *reinterpret_cast<double2*>(reinterpret_cast<char*>(const_cast<double2*>(R.peer[jt | R.jrow[r]]) + off) +
                                                R.row_clean[r]) =
                        *reinterpret_cast<const double2*>(tilebuf + (so16 ^ P.out_row[r]));
            } else {
                char* g = reinterpret_cast<char*>(amp + tile_base + cp_off);
#pragma unroll
                for (int r = 0; r < 16; ++r)
                    *reinterpret_cast<double2*>(g + P.row_off[r]) =
                        *reinterpret_cast<const double2*>(tilebuf + (so16 ^ P.out_row[r]));
            }
            __syncthreads();
            if (has_next) {
                if (GATHER) finish_remote_load(next_base);
                else issue_load(next_base);
            }
        }
        if (!has_next) break;
        tile = next;
        tile_base = next_base;
    }
}

__global__ void __launch_bounds__(FP_NT, 2) k_fused_pass(double2* amp, const double2* src, const __grid_constant__ FpPass P) {
    fp_pass_body<0>(amp, src, P, *reinterpret_cast<const FpRemote*>(&P));  // R is never read
}

// the same pass carrying a global<->local qubit swap: MODE 1 gathers its tiles from the peers' slabs
// (first pass after the swap), MODE 2 scatters them into the peers' slabs (last pass before it)
template <int MODE>
__global__ void __launch_bounds__(FP_NT, 2) k_fused_swap_pass(double2* amp, const __grid_constant__ FpPass P,
                                                              const __grid_constant__ FpRemote R) {
    fp_pass_body<MODE>(amp, amp, P, R);
}

// (k_fused_pass3 -- one CTA per SM, two 256-thread groups over three tile buffers -- was measured in round 1:
// QuantumVolume 30 q -7 %, EfficientSU2 28 q +6 %; removed in round 2, the A/B is profiles/r1k_stagger_ab.txt and
// DESIGN.md section 3.)
