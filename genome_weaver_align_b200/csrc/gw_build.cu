// gw_build.cu — `genome-weaver bwt` on the device: suffix array of the CYCLIC rotations of a text (no sentinel), its BWT and the
// suffix-array samples. Replaces BWTransform.pac2bwt / bwt and CyclicSAIS.SAIS of the reference
// (align/src/main/java/org/utgenome/weaver/align/BWTransform.java:110-151, :163-179; align/sais/CyclicSAIS.java:223-382) with
// prefix doubling: rotations are first sorted by their 27-symbol prefix packed base 5 into one 64-bit key, then by
// (rank[i], rank[i+h]) with h doubling until all ranks are distinct. Remaining ties (a text that is a power of a shorter
// string) are broken by position. Radix sorting and the max-scan are CUB (library plumbing; this is the index builder,
// not the alignment hot path). Symbol order A<C<G<T<N as in the reference.
#include <cuda_runtime.h>
#include <stdio.h>

#include <cub/cub.cuh>
#include <string>
#include <vector>

#include "gw_host.h"

#define GWB_CHECK(x)                                                                                         \
    do {                                                                                                     \
        cudaError_t e_ = (x);                                                                                \
        if (e_ != cudaSuccess) throw GwIoError(std::string("CUDA error in index build: ") + cudaGetErrorString(e_)); \
    } while (0)

namespace {

typedef unsigned long long ull;

__global__ void gwb_init_keys(const uint8_t* code, ull n, ull* key, uint32_t* idx) {
    ull i = (ull)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    ull v = 0, p = i;
    for (int j = 0; j < 27; ++j) {
        v = v * 5ull + code[p];
        if (++p == n) p = 0;
    }
    key[i] = v;
    idx[i] = (uint32_t)i;
}

// head[j] = j where a new key group starts, else 0; counts the groups
__global__ void gwb_group_heads(const ull* key, ull n, uint32_t* head, ull* n_groups) {
    ull j = (ull)blockIdx.x * blockDim.x + threadIdx.x;
    bool is_head = false;
    if (j < n) {
        is_head = j == 0 || key[j] != key[j - 1];
        head[j] = is_head ? (uint32_t)j : 0u;
    }
    unsigned b = __ballot_sync(0xFFFFFFFFu, is_head);
    if ((threadIdx.x & 31) == 0 && b) atomicAdd(n_groups, (ull)__popc(b));
}

__global__ void gwb_scatter_rank(const uint32_t* grp, const uint32_t* idx, ull n, uint32_t* rank) {
    ull j = (ull)blockIdx.x * blockDim.x + threadIdx.x;
    if (j < n) rank[idx[j]] = grp[j];
}

__global__ void gwb_pair_keys(const uint32_t* rank, ull n, ull h, ull* key, uint32_t* idx) {
    ull i = (ull)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    ull p = (i + h) % n;
    key[i] = ((ull)rank[i] << 32) | (ull)rank[p];
    idx[i] = (uint32_t)i;
}

// bwt[j] = text[(SA[j] - 1 + N) % N]  (BWTransform.java:172-179); sample[j / 32] = SA[j] for j % 32 == 0
__global__ void gwb_bwt_and_samples(const uint8_t* code, const uint32_t* sa, ull n, uint8_t* bwt, uint32_t* ssa) {
    ull j = (ull)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    ull s = sa[j];
    bwt[j] = code[s == 0 ? n - 1 : s - 1];
    if ((j & 31ull) == 0) ssa[j >> 5] = (uint32_t)s;
}

template <typename T>
struct DevBuf {
    T* p = nullptr;
    explicit DevBuf(size_t n) { GWB_CHECK(cudaMalloc((void**)&p, std::max<size_t>(n, 1) * sizeof(T))); }
    ~DevBuf() { cudaFree(p); }
    DevBuf(const DevBuf&) = delete;
};

inline unsigned grid_of(ull n) { return (unsigned)((n + 255) / 256); }

}  // namespace

// suffix array of the cyclic rotations of code[0..n), on `device`; optionally also returns the SA itself
void gw_build_bwt_device(int device, const uint8_t* codes, uint64_t n, std::vector<uint8_t>& bwt, std::vector<uint32_t>& ssa,
                         std::vector<uint32_t>* sa_out) {
    bwt.assign((size_t)n, 0);
    ssa.assign((size_t)((n + 31) / 32), 0);
    if (sa_out) sa_out->assign((size_t)n, 0);
    if (n == 0) return;
    if (n >= 0xFFFFFFFFull) throw GwIoError("texts of 2^32-1 symbols or more are not supported");
    GWB_CHECK(cudaSetDevice(device));
    DevBuf<uint8_t> d_code(n), d_bwt(n);
    DevBuf<ull> d_key(n), d_key_alt(n), d_groups(1);
    DevBuf<uint32_t> d_idx(n), d_idx_alt(n), d_rank(n), d_head(n), d_ssa((n + 31) / 32);
    GWB_CHECK(cudaMemcpy(d_code.p, codes, n, cudaMemcpyHostToDevice));

    size_t tmp_sort = 0, tmp_scan = 0;
    {
        cub::DoubleBuffer<ull> k(d_key.p, d_key_alt.p);
        cub::DoubleBuffer<uint32_t> v(d_idx.p, d_idx_alt.p);
        GWB_CHECK(cub::DeviceRadixSort::SortPairs(nullptr, tmp_sort, k, v, (long long)n, 0, 64));
        GWB_CHECK(cub::DeviceScan::InclusiveScan(nullptr, tmp_scan, d_head.p, d_head.p, cub::Max(), (long long)n));
    }
    DevBuf<uint8_t> d_tmp(std::max(tmp_sort, tmp_scan));

    const uint32_t* sa_sorted = nullptr;
    auto sort_and_rank = [&](int end_bit) -> ull {
        cub::DoubleBuffer<ull> k(d_key.p, d_key_alt.p);
        cub::DoubleBuffer<uint32_t> v(d_idx.p, d_idx_alt.p);
        size_t bytes = tmp_sort;
        GWB_CHECK(cub::DeviceRadixSort::SortPairs(d_tmp.p, bytes, k, v, (long long)n, 0, end_bit));
        GWB_CHECK(cudaMemset(d_groups.p, 0, sizeof(ull)));
        gwb_group_heads<<<grid_of(n), 256>>>(k.Current(), n, d_head.p, d_groups.p);
        bytes = tmp_scan;
        GWB_CHECK(cub::DeviceScan::InclusiveScan(d_tmp.p, bytes, d_head.p, d_head.p, cub::Max(), (long long)n));
        gwb_scatter_rank<<<grid_of(n), 256>>>(d_head.p, v.Current(), n, d_rank.p);
        ull g = 0;
        GWB_CHECK(cudaMemcpy(&g, d_groups.p, sizeof(ull), cudaMemcpyDeviceToHost));
        sa_sorted = v.Current();
        return g;
    };

    gwb_init_keys<<<grid_of(n), 256>>>(d_code.p, n, d_key.p, d_idx.p);
    ull groups = sort_and_rank(63);   // 5^27 < 2^63
    int rank_bits = 1;
    while ((1ull << rank_bits) < n) ++rank_bits;
    for (ull covered = 27; groups < n && covered < n; covered *= 2) {
        gwb_pair_keys<<<grid_of(n), 256>>>(d_rank.p, n, covered % n, d_key.p, d_idx.p);
        groups = sort_and_rank(32 + rank_bits);
    }
    gwb_bwt_and_samples<<<grid_of(n), 256>>>(d_code.p, sa_sorted, n, d_bwt.p, d_ssa.p);
    GWB_CHECK(cudaGetLastError());
    GWB_CHECK(cudaMemcpy(bwt.data(), d_bwt.p, n, cudaMemcpyDeviceToHost));
    GWB_CHECK(cudaMemcpy(ssa.data(), d_ssa.p, ssa.size() * sizeof(uint32_t), cudaMemcpyDeviceToHost));
    if (sa_out) GWB_CHECK(cudaMemcpy(sa_out->data(), sa_sorted, n * sizeof(uint32_t), cudaMemcpyDeviceToHost));
    GWB_CHECK(cudaDeviceSynchronize());
}

// FASTA / code array -> complete host index (both BWTs and sample arrays)
void gw_build_host_index(int device, GwHostIndex& ix) {
    ix.N = ix.text.size();
    gw_build_bwt_device(device, ix.text.data(), ix.N, ix.bwt[0], ix.ssa[0], nullptr);
    std::vector<uint8_t> rev(ix.text.rbegin(), ix.text.rend());
    gw_build_bwt_device(device, rev.data(), ix.N, ix.bwt[1], ix.ssa[1], nullptr);
}
