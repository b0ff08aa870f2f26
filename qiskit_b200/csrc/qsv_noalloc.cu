// qsv_noalloc.cu -- one entry point that allocates nothing.
//
// qsv_inner_static_scratch computes <a|b> = sum conj(a_i) b_i like qsv_inner (qsv_expval.cu), but for
// callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
// that forbid allocation): the per-block partial sums that qsv_inner keeps in a cudaMalloc'ed,
// per-handle scratch buffer live here in ONE file-scope __device__ array of fixed size.  Every GPU
// that loads the library holds those QSV_NA_SCRATCH_BYTES bytes until the process exits.
//
// No counterpart in the reference (see include/qsv.h).
//
// Rules this file keeps:
//   * no cudaMalloc / cudaFree / cudaMallocHost / allocating library call anywhere on the path;
//   * in HOST code `g_na_scratch` is only a stand-in for the symbol, not a device address: the address
//     handed to the kernel comes from cudaGetSymbolAddress, looked up once per device with that device
//     current, return code checked;
//   * the array is shared by every call and keeps the last call's contents: the arrival ticket (the only
//     word that must start at zero) is cleared on the caller's stream before every launch; the partial
//     sums are fully overwritten before they are read;
//   * two calls that overlap on the GPU (different streams, or two host threads on one device) race on
//     the array -- callers serialise them (one stream, or an event between streams).
#include <mutex>

#include "qsv_common.cuh"

#define NA_THREADS 256
#define NA_MAX_BLOCKS 1024                       // partial sums: 2 doubles (re, im) per block
#define NA_MAX_AMPS_LOG2 30                      // 2^30 amplitudes = 16 GiB per operand; 2^20 per block at the cap
#define NA_PART_BYTES (2 * NA_MAX_BLOCKS * 8)    // 16 KiB
#define QSV_NA_SCRATCH_BYTES (NA_PART_BYTES + 256)  // + one 256-byte line for the arrival ticket
#define NA_MAX_DEVICES 64

static __device__ __align__(256) unsigned char g_na_scratch[QSV_NA_SCRATCH_BYTES];

__device__ __forceinline__ double na_block_sum(double v, double* sh) {
    v = warp_sum(v);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __syncthreads();
    if (lane == 0) sh[warp] = v;
    __syncthreads();
    double r = 0.0;
    if (warp == 0) {
        r = lane < (NA_THREADS / 32) ? sh[lane] : 0.0;
        r = warp_sum(r);
    }
    return r;  // valid in warp 0
}

// Same per-block arithmetic as k_inner (qsv_expval.cu); the second launch of the allocating path
// (k_reduce_partials) is replaced by "the last block to arrive sums the partials in a fixed order", so
// the result does not depend on which block that is.
__global__ void __launch_bounds__(NA_THREADS) k_inner_static_scratch(const double2* __restrict__ a,
                                                                      const double2* __restrict__ b,
                                                                      uint64_t count, unsigned char* scratch,
                                                                      double* __restrict__ out) {
    __shared__ double sh[NA_THREADS / 32];
    __shared__ unsigned int s_last;
    double* part = reinterpret_cast<double*>(scratch);
    unsigned int* ticket = reinterpret_cast<unsigned int*>(scratch + NA_PART_BYTES);
    double re = 0.0, im = 0.0;
    const uint64_t stride = (uint64_t)gridDim.x * NA_THREADS;
    for (uint64_t i = blockIdx.x * (uint64_t)NA_THREADS + threadIdx.x; i < count; i += stride) {
        double2 x = a[i], y = b[i];
        re += x.x * y.x + x.y * y.y;
        im += x.x * y.y - x.y * y.x;
    }
    double r0 = na_block_sum(re, sh);
    double r1 = na_block_sum(im, sh);
    if (threadIdx.x == 0) {
        part[blockIdx.x] = r0;
        part[NA_MAX_BLOCKS + blockIdx.x] = r1;
        __threadfence();  // partials visible device-wide before the ticket is taken
        s_last = (atomicAdd(ticket, 1u) == gridDim.x - 1) ? 1u : 0u;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    double v0 = 0.0, v1 = 0.0;
    for (unsigned int blk = threadIdx.x; blk < gridDim.x; blk += NA_THREADS) {
        v0 += __ldcg(part + blk);  // L2: written by other SMs during this launch
        v1 += __ldcg(part + NA_MAX_BLOCKS + blk);
    }
    r0 = na_block_sum(v0, sh);
    r1 = na_block_sum(v1, sh);
    if (threadIdx.x == 0) {
        out[0] = r0;
        out[1] = r1;
        *ticket = 0u;  // leave the array ready; the host also clears it before every launch
    }
}

namespace {
struct NaDevice {
    unsigned char* scratch = nullptr;  // cudaGetSymbolAddress(g_na_scratch) on this device
    int num_sms = 0;
};
std::mutex g_na_mutex;
NaDevice g_na_dev[NA_MAX_DEVICES];
}  // namespace

// set-up, once per device, with that device current
static int na_setup(int device, NaDevice* out) {
    std::lock_guard<std::mutex> lock(g_na_mutex);
    NaDevice& d = g_na_dev[device];
    if (!d.scratch) {
        void* ptr = nullptr;
        QSV_CUDA(cudaGetSymbolAddress(&ptr, g_na_scratch));
        if (!ptr) QSV_FAIL(QSV_ERR_CUDA, "qsv_inner_static_scratch: cudaGetSymbolAddress returned a NULL address");
        int sms = 0;
        QSV_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
        d.num_sms = sms > 0 ? sms : 1;
        d.scratch = static_cast<unsigned char*>(ptr);
    }
    *out = d;
    return QSV_OK;
}

extern "C" int qsv_inner_static_scratch(int device, const double* a_c128, const double* b_c128, uint64_t count,
                                        double* out_c128, void* stream) {
    if (!a_c128 || !b_c128 || !out_c128) QSV_FAIL(QSV_ERR_INVALID, "qsv_inner_static_scratch: NULL argument");
    if (device < 0 || device >= NA_MAX_DEVICES)
        QSV_FAIL(QSV_ERR_INVALID, "qsv_inner_static_scratch: device %d out of range", device);
    if (count > (1ull << NA_MAX_AMPS_LOG2))
        QSV_FAIL(QSV_ERR_INVALID,
                 "qsv_inner_static_scratch: count %llu exceeds the maximum of 2^%d amplitudes that the static "
                 "scratch array (%d bytes, %d block partials) is sized for; use qsv_inner",
                 (unsigned long long)count, NA_MAX_AMPS_LOG2, (int)QSV_NA_SCRATCH_BYTES, NA_MAX_BLOCKS);
    if (((uintptr_t)a_c128 & 15) || ((uintptr_t)b_c128 & 15) || ((uintptr_t)out_c128 & 7))
        QSV_FAIL(QSV_ERR_INVALID, "qsv_inner_static_scratch: operands must be 16-byte aligned, the result 8-byte aligned");
    QSV_CUDA(cudaSetDevice(device));
    NaDevice d;
    QSV_TRY(na_setup(device, &d));
    uint64_t blocks = (count + NA_THREADS - 1) / NA_THREADS;
    uint64_t cap = (uint64_t)d.num_sms * 8;
    if (cap > NA_MAX_BLOCKS) cap = NA_MAX_BLOCKS;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    // the ticket must start at zero even if an earlier launch was aborted half way
    QSV_CUDA(cudaMemsetAsync(d.scratch + NA_PART_BYTES, 0, 256, st));
    k_inner_static_scratch<<<(unsigned int)blocks, NA_THREADS, 0, st>>>(
        reinterpret_cast<const double2*>(a_c128), reinterpret_cast<const double2*>(b_c128), count, d.scratch,
        out_c128);
    QSV_CUDA(cudaGetLastError());
    return QSV_OK;
}
