// capi_refine.cu -- back-check and refinement-estimate entry points of the C ABI (include/speq_cuda.h).
#include <cuda_runtime.h>

#include <cstring>
#include <vector>

#include "capi_util.h"
#include "kernels.h"

using namespace speqb;

// accessors into speqg_scan (defined in capi_scan.cu)
namespace speqb
{
struct scan_refine_view
{
    int device;
    uint32_t k, V;
    uint64_t n;
    DevIndex view;
    cudaStream_t stream;
    unsigned long long * mult;
    const unsigned long long * unique_n;   // closed form of the unrecorded windows (product kernels), else null
    double * d_percent, * d_tkpg, * d_partial;
    int32_t * d_gc;
    int refine_grid;
};
scan_refine_view refine_view(speqg_scan * s);   // capi_scan.cu
}   // namespace speqb

extern "C" {

int speqg_backcheck(speqg_index * idx, int device, uint32_t k, uint64_t * unique_ref, uint64_t * total_ref)
{
    SPEQG_TRY
    need(idx && unique_ref && total_ref, "NULL argument");
    need(k >= 1 && k <= kMaxK, "k must be in [1, 127]");
    DeviceIndex & di = device_index(idx, device);
    device_guard guard(device);
    HostIndex const & h = idx->host;
    uint32_t V = h.V;
    LaunchCfg cfg{};
    CK(scan_launch_cfg(M_BACKCHECK, 0, V, &cfg));

    // cut every text (scaffold and reverse complement, src/fm_scanner.cpp:1533-1538) into segments of windows
    uint64_t total_windows = 0;
    for (uint32_t i = 0; i < h.S; ++i)
        if (h.t_len[i] >= k) total_windows += h.t_len[i] - k + 1;
    for (uint32_t g = 0; g < V; ++g) unique_ref[g] = total_ref[g] = 0;
    if (total_windows == 0) return SPEQG_OK;
    uint64_t lanes = (uint64_t)cfg.grid * cfg.block;
    uint64_t seg = total_windows / (8 * lanes);
    seg = std::min<uint64_t>(std::max<uint64_t>(seg, 32), 8192);
    std::vector<SegmentX> segs;
    segs.reserve(total_windows / seg + h.S);
    for (uint32_t i = 0; i < h.S; ++i)
    {
        if (h.t_len[i] < k) continue;
        uint32_t nw = h.t_len[i] - k + 1;
        for (uint32_t st = 0; st < nw; st += (uint32_t)seg)
        {
            SegmentX sx{};
            sx.g0 = h.t_grp_off[i];
            sx.start = st;
            sx.nwin = (uint32_t)std::min<uint64_t>(seg, nw - st);
            sx.group = (uint32_t)h.group_of_text[i];
            segs.push_back(sx);
        }
    }
    SegmentX * d_segs = nullptr;
    KBlock * d_blocks_k = nullptr;
    unsigned long long * d_cnt = nullptr;
    cudaStream_t st = nullptr;
    try
    {
        CK(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
        d_segs = dev_upload(segs.data(), segs.size());
        d_blocks_k = dev_alloc<KBlock>(h.blocks.size());
        CK(launch_fill_e(di.blocks, d_blocks_k, di.lcp, (uint32_t)h.n, k, st));
        DevIndex view = di.view;
        view.blocks = d_blocks_k;
        d_cnt = dev_alloc<unsigned long long>(2 * (size_t)V + SC_COUNT);
        CK(cudaMemsetAsync(d_cnt, 0, (2 * (size_t)V + SC_COUNT) * 8, st));
        ScanArgs a{};
        a.k = k;
        a.V = V;
        a.unique = d_cnt;
        a.total_g = d_cnt + V;
        a.scalars = d_cnt + 2 * (size_t)V;
        a.segs = d_segs;
        a.n_segs = segs.size();
        BatchView v{};
        v.bases = di.t_bases; v.inv = di.t_inv; v.qual = nullptr; v.grp_off = di.t_grp_off; v.len = di.t_len;
        v.read_group = di.t_group; v.n_reads = h.S;
        CK(launch_scan_roll(view, v, a, M_BACKCHECK, cfg, st));
        std::vector<unsigned long long> out(2 * (size_t)V);
        CK(cudaMemcpyAsync(out.data(), d_cnt, out.size() * 8, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        for (uint32_t g = 0; g < V; ++g)
        {
            unique_ref[g] = out[g];
            total_ref[g] = out[V + g];
        }
    }
    catch (...)
    {
        cudaFree(d_segs); cudaFree(d_cnt); cudaFree(d_blocks_k);
        if (st) cudaStreamDestroy(st);
        throw;
    }
    cudaFree(d_segs); cudaFree(d_cnt); cudaFree(d_blocks_k);
    cudaStreamDestroy(st);
    SPEQG_CATCH
}

int speqg_refine(speqg_scan * scan, const double * percent, const int32_t * group_counts, double * tkpg)
{
    SPEQG_TRY
    need(scan && percent && group_counts && tkpg, "NULL argument");
    scan_refine_view r = refine_view(scan);
    if (!r.mult) throw api_error(SPEQG_ESTATE, "speqg_refine needs a scan opened with record_multiplicity = 1");
    device_guard guard(r.device);
    // pageable host sources: cudaMemcpyAsync stages them before returning
    CK(cudaMemcpyAsync(r.d_percent, percent, r.V * 8, cudaMemcpyHostToDevice, r.stream));
    CK(cudaMemcpyAsync(r.d_gc, group_counts, r.V * 4, cudaMemcpyHostToDevice, r.stream));
    CK(launch_refine(r.view, r.k, r.mult, r.d_percent, r.d_gc, r.V, r.d_partial, r.refine_grid, 256, r.stream));
    CK(launch_refine_reduce(r.d_partial, r.refine_grid, r.V, r.d_tkpg, r.unique_n, r.d_percent, r.stream));
    CK(cudaMemcpyAsync(tkpg, r.d_tkpg, r.V * 8, cudaMemcpyDeviceToHost, r.stream));
    CK(cudaStreamSynchronize(r.stream));
    SPEQG_CATCH
}

int speqg_scan_multiplicity_count(speqg_scan * scan, uint64_t * n_entries)
{
    SPEQG_TRY
    need(scan && n_entries, "NULL argument");
    scan_refine_view r = refine_view(scan);
    if (!r.mult) throw api_error(SPEQG_ESTATE, "scan was opened without record_multiplicity");
    device_guard guard(r.device);
    unsigned long long * d = dev_alloc<unsigned long long>(1);
    unsigned long long hres = 0;
    cudaError_t e = cudaMemsetAsync(d, 0, 8, r.stream);
    if (e == cudaSuccess) e = launch_count_nonzero(r.mult, (uint32_t)r.n, d, r.stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(&hres, d, 8, cudaMemcpyDeviceToHost, r.stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(r.stream);
    cudaFree(d);
    CK(e);
    *n_entries = hres;
    SPEQG_CATCH
}

}   // extern "C"
