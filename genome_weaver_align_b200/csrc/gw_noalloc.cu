// gw_noalloc.cu — gw_verify_pairs_noalloc: the pair verifier (BitParallelSmithWaterman.alignBlockDetailed, as gw_verify_pairs) for
// callers that cannot allocate device memory (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc). It has no
// counterpart in the reference.
//
// The only scratch memory the verifier needs is the DP history (16 bytes of VP/VN per window column, 64-row block and pair, gw_verify.cuh). Here it
// lives in a file-scope __device__ array of GW_NOALLOC_SCRATCH_BYTES, so every GPU that loads libgwalign holds that much until the
// process exits. Inputs and outputs are the caller's DEVICE buffers. This file calls no cudaMalloc / cudaFree and no allocating
// library routine; the launch is asynchronous on the caller's stream.
//
// The array's NAME is a host-side stand-in in host code, not a device address: the address a kernel can use comes from
// cudaGetSymbolAddress, taken once per device with that device current (scratch_of_current_device below).
//
// All calls share the array: it keeps the previous call's contents, and two calls whose kernels overlap on the GPU (different
// streams, or two host threads) race on it. Nothing in it has to start at zero — every history element a traceback reads was written by
// the same pair in the same launch (gw_dp_begin writes column 0, gw_dp_columns writes every block of every later column, zeros included,
// before gw_dp_traceback reads any) — but the OUTPUT cigar rows are cleared to their full stride by the kernel, which is what
// gw_verify_pairs' memset does for its own device buffer.
#include <cuda_runtime.h>

#include <mutex>

#include "gw_job.cuh"
#include "gw_kernels.h"

#define GW_NOALLOC_SCRATCH_BYTES_ (64u << 20)
static_assert(GW_NOALLOC_SCRATCH_BYTES_ == GW_NOALLOC_SCRATCH_BYTES, "gw_kernels.h and gw_noalloc.cu disagree on the scratch size");

static __device__ __align__(256) unsigned char gw_noalloc_scratch[GW_NOALLOC_SCRATCH_BYTES_];

// gw_verify_pairs_kernel (gw_kernels.cu) with the history in the static scratch array, a per-pair length guard (the host cannot see
// the offsets, they are in device memory) and the cigar rows cleared to the stride
__global__ void __launch_bounds__(128) gw_verify_pairs_noalloc_kernel(const char* refs, const u64* ref_off, const char* queries, const u64* query_off,
                                                                      uint32_t n, uint32_t max_cols, uint32_t max_m, int band_k, int32_t* result,
                                                                      uint32_t* cigar_ops, uint32_t cigar_stride, GwU64x2* hist) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const char* r = refs + ref_off[i];
    const char* qs = queries + query_off[i];
    const u64 ncols64 = ref_off[i + 1] - ref_off[i], m64 = query_off[i + 1] - query_off[i];
    uint32_t* row = cigar_ops + (size_t)i * cigar_stride;
    if (ncols64 > max_cols || m64 > max_m) {   // longer than the caller stated: no history room for it
        result[i * 4 + 0] = GW_NOALLOC_PAIR_TOO_LONG; result[i * 4 + 1] = 0; result[i * 4 + 2] = 0; result[i * 4 + 3] = 0;
        for (uint32_t j = 0; j < cigar_stride; ++j) row[j] = 0;
        return;
    }
    const int Ncols = (int)ncols64, m = (int)m64;
    auto qsym = [&](int j) -> int { return gw_code_of((unsigned char)qs[j]); };
    auto refsym = [&](int j) -> int { return gw_code_of((unsigned char)r[j]); };
    u64 peq[4][GW_MAX_NB];
    for (int a = 0; a < 4; ++a) for (int b = 0; b < GW_MAX_NB; ++b) peq[a][b] = 0;
    for (int j = 0; j < m; ++j) {
        int cde = qsym(j);
        if (cde == 4) { for (int a = 0; a < 4; ++a) peq[a][j >> 6] |= 1ULL << (j & 63); }  // QueryMask.java:50-55: N matches every base
        else peq[cde][j >> 6] |= 1ULL << (j & 63);
    }
    uint32_t cig[GW_CIG_CAP];
    GwVerifyOut out;
    gw_verify(m, Ncols, band_k, peq, refsym, qsym, hist + i, (size_t)n, cig, &out);
    result[i * 4 + 0] = out.valid; result[i * 4 + 1] = out.pos; result[i * 4 + 2] = out.diff;
    result[i * 4 + 3] = out.cig_overflow ? -1 : out.cig_n;
    for (uint32_t j = 0; j < cigar_stride; ++j) row[j] = (int)j < out.cig_n ? cig[j] : 0u;
}

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_mu;
void* g_scratch[kMaxDevices];
bool g_have[kMaxDevices];

// device address of the scratch array on the CURRENT device (each device has its own instance of a __device__ variable)
cudaError_t scratch_of_current_device(GwU64x2** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= kMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lk(g_mu);
    if (!g_have[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, gw_noalloc_scratch);
        if (e != cudaSuccess) return e;
        if (!p) return cudaErrorInvalidSymbol;
        g_scratch[dev] = p;
        g_have[dev] = true;
    }
    *out = static_cast<GwU64x2*>(g_scratch[dev]);
    return cudaSuccess;
}
}  // namespace

// what gw_verify_pairs allocates for the same batch: (columns + 2) x blocks x 16 B per pair
size_t gwk_verify_pairs_noalloc_need(uint32_t n, uint32_t max_cols, uint32_t max_m) {
    return (size_t)n * ((size_t)max_cols + 2) * (size_t)((max_m + 63) / 64 > 0 ? (max_m + 63) / 64 : 1) * sizeof(GwU64x2);
}

// the caller (gw_api_components.inc) has made the device current and checked the arguments, including need <= GW_NOALLOC_SCRATCH_BYTES
cudaError_t gwk_verify_pairs_noalloc(cudaStream_t st, const char* refs, const u64* ref_off, const char* queries, const u64* query_off, uint32_t n,
                                     uint32_t max_cols, uint32_t max_m, int band_k, int32_t* result, uint32_t* cigar_ops, uint32_t cigar_stride) {
    GwU64x2* hist = nullptr;
    cudaError_t e = scratch_of_current_device(&hist);
    if (e != cudaSuccess) return e;
    if (!n) return cudaSuccess;
    gw_verify_pairs_noalloc_kernel<<<(n + 127) / 128, 128, 0, st>>>(refs, ref_off, queries, query_off, n, max_cols, max_m, band_k, result,
                                                                    cigar_ops, cigar_stride, hist);
    return cudaGetLastError();
}
