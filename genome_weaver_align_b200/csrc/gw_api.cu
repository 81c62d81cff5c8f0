// gw_api.cu — the C-ABI of libgwalign.so (include/gw_align.h): index lifecycle, aligner lifecycle, the batch alignment
// driver (waves -> quick scan -> search / locate / verify rounds -> emit) and the component entry points used by the
// parity tests. Everything computes on the CUDA device; there is no CPU fallback.
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <map>
#include <string>
#include <thread>
#include <vector>

#include "../../include/gw_align.h"
#include "gw_host.h"
#include "gw_image.h"
#include "gw_kernels.h"

static_assert(sizeof(gw_hit_t) == sizeof(GwHitRec), "gw_hit_t and the device record must have the same layout");
static_assert(GW_VERIFY_PAIRS_NOALLOC_MAX_BYTES == GW_NOALLOC_SCRATCH_BYTES, "header and gw_noalloc.cu disagree on the scratch size");
static_assert(GW_CHR_NULL == GW_CHR_NULL_ && GW_CHR_EMPTY == GW_CHR_EMPTY_ && GW_CHR_STAR == GW_CHR_STAR_, "chr id constants");

bool gw_sam_one(const std::vector<GwContig>& ctg, const std::string& name, const char* bases, size_t m, const char* qual, size_t qlen,
                const gw_hit_t* chain, int n, const uint32_t* ops, std::string& out);

namespace {

thread_local std::string g_err;

struct CudaFail { cudaError_t e; const char* what; };
#define CK(x)                                         \
    do {                                              \
        cudaError_t e_ = (x);                         \
        if (e_ != cudaSuccess) throw CudaFail{e_, #x}; \
    } while (0)

int fail(int code, const std::string& msg) { g_err = msg; return code; }

// runs body, mapping exceptions to error codes
template <typename F>
int guarded(F body) {
    try {
        return body();
    } catch (CudaFail& f) {
        return fail(GW_E_CUDA, std::string("CUDA: ") + cudaGetErrorString(f.e) + " at " + f.what);
    } catch (GwIoError& e) {
        std::string w = e.what();
        return fail(w.compare(0, 4, "CUDA") == 0 ? GW_E_CUDA : GW_E_IO, w);
    } catch (std::bad_alloc&) {
        return fail(GW_E_ARG, "out of host memory");
    } catch (std::exception& e) {
        return fail(GW_E_ARG, e.what());
    }
}

void use_device(int device) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) throw CudaFail{e == cudaSuccess ? cudaErrorNoDevice : e, "no CUDA device (libgwalign has no CPU path)"};
    if (device < 0 || device >= n) throw CudaFail{cudaErrorInvalidDevice, "device index out of range"};
    CK(cudaSetDevice(device));
}

// growable device buffer
struct DBuf {
    void* p = nullptr;
    size_t cap = 0;
    void ensure(size_t bytes, bool keep = false, cudaStream_t st = 0) {
        if (bytes <= cap) return;
        // large buffers are sized from a memory budget (align_device): they grow to exactly what is asked for, and the old block is
        // given back first when its contents are not needed
        size_t ncap = bytes >= ((size_t)1 << 30) ? bytes : std::max(bytes, cap + cap / 2);
        if (!keep && p) { cudaFree(p); p = nullptr; cap = 0; }
        void* np = nullptr;
        CK(cudaMalloc(&np, ncap));
        if (keep && p && cap) { CK(cudaMemcpyAsync(np, p, cap, cudaMemcpyDeviceToDevice, st)); CK(cudaStreamSynchronize(st)); }
        if (p) cudaFree(p);
        p = np; cap = ncap;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
    template <typename T> T* as() const { return (T*)p; }
};

}  // namespace

// ---------------------------------------------------------------------------------------------------------------------
struct gw_index {
    int device = 0;
    uint64_t N = 0;
    std::vector<GwContig> contigs;
    GwIndexView view;
    DBuf loc[2];   // densified locate table (empty: the .ssa samples serve)
    DBuf occ_all;  // both occurrence tables in ONE allocation, so that one L2 access-policy window can cover them
    size_t occ_bytes = 0;
    DBuf ssa[2], ref_lo, ref_hi, ref_nf, ctg_off, ctg_id, ctg_rank;
    ~gw_index() {
        occ_all.release();
        for (int s = 0; s < 2; ++s) { ssa[s].release(); loc[s].release(); }
        ref_lo.release(); ref_hi.release(); ref_nf.release(); ctg_off.release(); ctg_id.release(); ctg_rank.release();
    }
};

namespace {

void upload(DBuf& b, const void* src, size_t bytes) {
    b.ensure(std::max<size_t>(bytes, 16));
    if (bytes) CK(cudaMemcpy(b.p, src, bytes, cudaMemcpyHostToDevice));
}

// host index -> HBM image (layout: gw_common.h; the arrays are laid out on the host by gw_image.h, then copied)
void make_device_index(gw_index* ix, const GwHostIndex& h, int device) {
    use_device(device);
    ix->device = device;
    ix->N = h.N;
    ix->contigs = h.contigs;
    GwIndexImage img;
    gw_layout_index(h, img);
    GwIndexView& v = ix->view;
    v = img.view;
    const size_t occ0 = (img.occ[0].size() * sizeof(GwOccBlock) + 255) & ~(size_t)255, occ1 = img.occ[1].size() * sizeof(GwOccBlock);
    ix->occ_all.ensure(std::max<size_t>(occ0 + occ1, 16));
    ix->occ_bytes = occ0 + occ1;
    for (int s = 0; s < 2; ++s) {
        GwOccBlock* dst = reinterpret_cast<GwOccBlock*>(ix->occ_all.as<char>() + (s ? occ0 : 0));
        if (!img.occ[s].empty()) CK(cudaMemcpy(dst, img.occ[s].data(), img.occ[s].size() * sizeof(GwOccBlock), cudaMemcpyHostToDevice));
        v.occ[s] = dst;
        upload(ix->ssa[s], h.ssa[s].data(), h.ssa[s].size() * sizeof(uint32_t));
        v.ssa[s] = ix->ssa[s].as<uint32_t>();
        v.loc[s] = v.ssa[s];
    }
    v.loc_step = GW_SSA_L;
    upload(ix->ref_lo, img.ref_lo.data(), img.ref_lo.size() * 8); upload(ix->ref_hi, img.ref_hi.data(), img.ref_hi.size() * 8);
    upload(ix->ref_nf, img.ref_nf.data(), img.ref_nf.size() * 8);
    v.ref_lo = ix->ref_lo.as<u64>(); v.ref_hi = ix->ref_hi.as<u64>(); v.ref_nf = ix->ref_nf.as<u64>();
    upload(ix->ctg_off, img.ctg_off.data(), img.ctg_off.size() * 4); upload(ix->ctg_id, img.ctg_id.data(), img.ctg_id.size() * 4);
    upload(ix->ctg_rank, img.ctg_rank.data(), img.ctg_rank.size() * 4);
    v.ctg_off = ix->ctg_off.as<uint32_t>(); v.ctg_id = ix->ctg_id.as<int32_t>(); v.ctg_rank = ix->ctg_rank.as<int32_t>();
    CK(cudaDeviceSynchronize());
}

// Locate table with the suffix-array value of every `step`-th row (1, 2, 4, 8 or 16; GW_SSA_L drops the table), built on the device
// from the index itself. step 0: the densest table that takes at most a fifth of the free device memory.
void densify_index(gw_index* ix, int step) {
    CK(cudaSetDevice(ix->device));
    if (step == 0) {
        size_t free_b = 0, total_b = 0;
        CK(cudaMemGetInfo(&free_b, &total_b));
        step = 1;
        while (step < GW_SSA_L && 2.0 * 4.0 * (double)((ix->N + step - 1) / step) > 0.2 * (double)free_b) step *= 2;
    }
    GwIndexView& v = ix->view;
    for (int s = 0; s < 2; ++s) { v.loc[s] = v.ssa[s]; ix->loc[s].release(); }
    v.loc_step = GW_SSA_L;
    if (step >= GW_SSA_L || ix->N == 0) return;
    DBuf ctr;
    ctr.ensure(16);
    for (int s = 0; s < 2; ++s) {
        const size_t entries = (size_t)((ix->N + step - 1) / step);
        ix->loc[s].ensure(entries * 4);
        CK(cudaMemset(ix->loc[s].p, 0xFF, entries * 4));
        CK(cudaMemset(ctr.p, 0, 16));
        gwk_densify(0, v, s, (uint32_t)step, ix->loc[s].as<uint32_t>(), v.n_ssa, ctr.as<uint32_t>());
        CK(cudaDeviceSynchronize());
    }
    ctr.release();
    CK(cudaDeviceSynchronize());
    CK(cudaGetLastError());
    for (int s = 0; s < 2; ++s) v.loc[s] = ix->loc[s].as<uint32_t>();
    v.loc_step = (uint32_t)step;
}

std::vector<uint8_t> codes_of(const char* bases, uint64_t n) {
    std::vector<uint8_t> c((size_t)n);
    for (uint64_t i = 0; i < n; ++i) c[(size_t)i] = (uint8_t)gw_code_of((unsigned char)bases[i]);
    return c;
}

}  // namespace

extern "C" {

const char* gw_last_error(void) { return g_err.c_str(); }

int gw_index_build(const char* fasta_path, int device) {
    return guarded([&]() -> int {
        if (!fasta_path) return fail(GW_E_ARG, "fasta_path is null");
        use_device(device);
        GwHostIndex h;
        gw_pack_fasta(fasta_path, h.text, h.contigs);
        gw_build_host_index(device, h);
        gw_save_index_files(fasta_path, h);
        return GW_OK;
    });
}

int gw_index_open(const char* fasta_path, int device, gw_index_t** out) {
    return guarded([&]() -> int {
        if (!fasta_path || !out) return fail(GW_E_ARG, "null argument");
        *out = nullptr;
        use_device(device);
        GwHostIndex h;
        gw_load_index_files(fasta_path, h);
        gw_index* ix = new gw_index();
        try { make_device_index(ix, h, device); densify_index(ix, 0); } catch (...) { delete ix; throw; }
        *out = ix;
        return GW_OK;
    });
}

int gw_index_locate_table(gw_index_t* ix, int step) {
    return guarded([&]() -> int {
        if (!ix || step < 0 || step > GW_SSA_L || (step & (step - 1)) != 0) return fail(GW_E_ARG, "step must be 0 (automatic), 1, 2, 4, 8, 16 or 32");
        densify_index(ix, step);
        return GW_OK;
    });
}
int gw_index_locate_step(const gw_index_t* ix) { return ix ? (int)ix->view.loc_step : -1; }

int gw_index_from_sequence(const char* name, const char* bases, int device, gw_index_t** out) {
    return guarded([&]() -> int {
        if (!name || !bases || !out) return fail(GW_E_ARG, "null argument");
        *out = nullptr;
        use_device(device);
        GwHostIndex h;
        std::string s;
        for (const char* p = bases; *p; ++p) if (*p != ' ') s += *p;   // ACGTSequence(String) skips blanks (ACGTSequence.java:86-98)
        h.text = codes_of(s.data(), s.size());
        GwContig c;
        c.name = name; c.desc = name; c.length = (int64_t)s.size(); c.offset = 0;
        h.contigs.push_back(c);
        gw_build_host_index(device, h);
        gw_index* ix = new gw_index();
        try { make_device_index(ix, h, device); densify_index(ix, 0); } catch (...) { delete ix; throw; }
        *out = ix;
        return GW_OK;
    });
}

void gw_index_close(gw_index_t* ix) {
    if (!ix) return;
    cudaSetDevice(ix->device);
    delete ix;
}
int64_t gw_index_text_size(const gw_index_t* ix) { return ix ? (int64_t)ix->N : -1; }
int32_t gw_index_num_contigs(const gw_index_t* ix) { return ix ? (int32_t)ix->contigs.size() : -1; }
const char* gw_index_contig_name(const gw_index_t* ix, int32_t i) { return (ix && i >= 0 && i < (int32_t)ix->contigs.size()) ? ix->contigs[(size_t)i].name.c_str() : nullptr; }
int64_t gw_index_contig_length(const gw_index_t* ix, int32_t i) { return (ix && i >= 0 && i < (int32_t)ix->contigs.size()) ? ix->contigs[(size_t)i].length : -1; }
int64_t gw_index_contig_offset(const gw_index_t* ix, int32_t i) { return (ix && i >= 0 && i < (int32_t)ix->contigs.size()) ? ix->contigs[(size_t)i].offset : -1; }
int gw_index_device(const gw_index_t* ix) { return ix ? ix->device : -1; }

}  // extern "C"

#include "gw_api_align.inc"
#include "gw_api_components.inc"
