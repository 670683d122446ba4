// capi.cu -- the C ABI of libspeq_cuda.so (include/speq_cuda.h): handles, device memory, streams, launches.
//
// No exception crosses the ABI: every entry point catches, stores the message in a thread-local string and returns a
// negative SPEQG_E* code.  There is no CPU path: every compute entry point needs a CUDA device.
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstring>
#include <string>
#include <vector>

#include "capi_util.h"
#include "kernels.h"

using namespace speqb;

namespace speqb
{
DeviceIndex & device_index(speqg_index * idx, int device)
{
    auto it = idx->dev.find(device);
    if (it != idx->dev.end()) return it->second;
    device_guard guard(device);
    HostIndex const & h = idx->host;
    DeviceIndex d;
    d.device = device;
    try
    {
        // query code -> text symbol: A1 C2 G3 T5, N4 (fm_index.hpp:302: rank + 1; dna5 ranks A0 C1 G2 N3 T4)
        static const uint32_t sym_of_code[5] = {1, 2, 3, 5, 4};
        uint32_t ctab[16] = {0};
        for (int c = 0; c < 5; ++c)
        {
            ctab[c] = (uint32_t)h.C[sym_of_code[c]];
            ctab[kCtabEnd + c] = (uint32_t)h.C[sym_of_code[c] + 1];
        }
        d.blocks = dev_upload(h.blocks.data(), h.blocks.size(), &d.bytes);
        d.grp = dev_upload(h.grp.data(), h.grp.size(), &d.bytes);
        d.lcp = dev_upload(h.lcp.data(), h.lcp.size(), &d.bytes);
        d.pn = dev_upload(h.pn.data(), h.pn.size(), &d.bytes);
        d.up = dev_alloc<UpRec>(h.pn.size());
        d.bytes += h.pn.size() * sizeof(UpRec);
        CK(launch_build_up(d.lcp, d.pn, d.up, h.pn.size(), nullptr));
        CK(cudaStreamSynchronize(nullptr));
        d.n_pos = dev_upload(h.n_pos.data(), h.n_pos.size(), &d.bytes);
        d.ctab = dev_upload(ctab, 16, &d.bytes);
        d.t_bases = dev_upload(h.t_bases.data(), h.t_bases.size(), &d.bytes);
        d.t_inv = dev_upload(h.t_inv.data(), h.t_inv.size(), &d.bytes);
        d.t_grp_off = dev_upload(h.t_grp_off.data(), h.t_grp_off.size(), &d.bytes);
        d.t_len = dev_upload(h.t_len.data(), h.t_len.size(), &d.bytes);
        d.t_group = dev_upload(h.group_of_text.data(), h.group_of_text.size(), &d.bytes);
    }
    catch (...)
    {
        free_device_index(d);
        throw;
    }
    DevIndex & v = d.view;
    v.blocks = d.blocks;
    v.lcp = d.lcp;
    v.pn = d.pn;
    v.up = d.up;
    v.grp = d.grp;
    v.n_pos = d.n_pos;
    v.ctab = d.ctab;
    v.n_npos = (uint32_t)h.n_pos.size();
    v.n = (uint32_t)h.n;
    return idx->dev.emplace(device, d).first->second;
}
}   // namespace speqb

extern "C" {

const char * speqg_last_error(void) { return g_err.c_str(); }
int speqg_abi_version(void) { return SPEQG_ABI_VERSION; }

int speqg_device_count(int * count)
{
    SPEQG_TRY
    need(count != nullptr, "count is NULL");
    *count = 0;
    cudaError_t e = cudaGetDeviceCount(count);
    if (e != cudaSuccess)
    {
        *count = 0;
        throw api_error(SPEQG_ECUDA, std::string("no usable CUDA device (there is no CPU path): cudaGetDeviceCount: ") + cudaGetErrorString(e));
    }
    SPEQG_CATCH
}

int speqg_index_build(const uint8_t * const * texts, const uint64_t * lens, uint32_t n_texts, const int32_t * group_of_text,
                      uint32_t n_groups, uint32_t flags, speqg_index ** out)
{
    SPEQG_TRY
    need(texts && lens && group_of_text && out, "NULL argument");
    auto * idx = new speqg_index();
    try
    {
        build_host_index(texts, lens, n_texts, group_of_text, n_groups, flags, idx->host);
    }
    catch (...)
    {
        delete idx;
        throw;
    }
    *out = idx;
    SPEQG_CATCH
}

int speqg_index_save(const speqg_index * idx, const char * path)
{
    SPEQG_TRY
    need(idx && path, "NULL argument");
    save_host_index(idx->host, path);
    SPEQG_CATCH
}

int speqg_index_save_at(const speqg_index * idx, const char * path, uint64_t offset)
{
    SPEQG_TRY
    need(idx && path, "NULL argument");
    save_host_index(idx->host, path, offset);
    SPEQG_CATCH
}

int speqg_index_load_at(const char * path, uint64_t offset, speqg_index ** out)
{
    SPEQG_TRY
    need(path && out, "NULL argument");
    auto * idx = new speqg_index();
    try
    {
        load_host_index(path, idx->host, offset);
    }
    catch (...)
    {
        delete idx;
        throw;
    }
    *out = idx;
    SPEQG_CATCH
}

int speqg_host_alloc(size_t bytes, void ** out)
{
    SPEQG_TRY
    need(out != nullptr, "NULL argument");
    *out = nullptr;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) throw api_error(SPEQG_ECUDA, "no usable CUDA device (there is no CPU path)");
    CK(cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocPortable));
    SPEQG_CATCH
}

void speqg_host_free(void * p)
{
    if (p) cudaFreeHost(p);
}

int speqg_index_load(const char * path, speqg_index ** out)
{
    SPEQG_TRY
    need(path && out, "NULL argument");
    auto * idx = new speqg_index();
    try
    {
        load_host_index(path, idx->host);
    }
    catch (...)
    {
        delete idx;
        throw;
    }
    *out = idx;
    SPEQG_CATCH
}

int speqg_index_get_info(const speqg_index * idx, speqg_index_info * info)
{
    SPEQG_TRY
    need(idx && info, "NULL argument");
    HostIndex const & h = idx->host;
    info->n = h.n;
    info->n_texts = h.S;
    info->n_groups = h.V;
    for (int c = 0; c < 8; ++c) info->C[c] = h.C[c];
    info->n_exceptions = h.n_exc;
    info->device_bytes = h.blocks.size() * sizeof(KBlock) + h.grp.size() * 2 + h.lcp.size() + h.pn.size() * 24 + 64 +
                         h.n_pos.size() * 4 + h.t_bases.size() * 8 + h.t_inv.size() * 4 + h.t_grp_off.size() * 8 +
                         h.t_len.size() * 4 + h.group_of_text.size() * 4;
    SPEQG_CATCH
}

int speqg_index_export(const speqg_index * idx, uint32_t * sa, uint8_t * bwt, uint32_t * group_of_sa, uint8_t * lcp)
{
    SPEQG_TRY
    need(idx != nullptr, "NULL argument");
    HostIndex const & h = idx->host;
    if (sa)
    {
        if (h.sa.size() != h.n) throw api_error(SPEQG_ESTATE, "index was built without SPEQG_BUILD_KEEP_SA");
        std::copy(h.sa.begin(), h.sa.end(), sa);
    }
    if (bwt)
    {
        // decode the 2-bit planes; exceptions come back as text symbols: N = 4, delimiter / sentinel reported as 6 / 0
        // cannot be told apart here, so they are reported as 0xFF except N
        static const uint8_t sym_of_code[4] = {1, 2, 3, 5};
        for (uint64_t i = 0; i < h.n; ++i)
        {
            KBlock const & blk = h.blocks[i / kBlockSyms];
            uint32_t j = (uint32_t)(i % kBlockSyms);
            uint32_t c = ((blk.lo[j / 32] >> (j % 32)) & 1u) | (((blk.hi[j / 32] >> (j % 32)) & 1u) << 1);
            bwt[i] = ((blk.exc[j / 32] >> (j % 32)) & 1u) ? 0xFF : sym_of_code[c];
        }
        for (uint32_t p : h.n_pos) bwt[p] = (uint8_t)kSymN;
    }
    if (group_of_sa)
        for (uint64_t i = 0; i < h.n; ++i) group_of_sa[i] = h.grp[i] == kNoGroup16 ? kNoGroup : h.grp[i];
    if (lcp) std::copy(h.lcp.begin(), h.lcp.begin() + h.n, lcp);
    SPEQG_CATCH
}

int speqg_index_upload(speqg_index * idx, int device)
{
    SPEQG_TRY
    need(idx != nullptr, "NULL argument");
    device_index(idx, device);
    SPEQG_CATCH
}

void speqg_index_free(speqg_index * idx)
{
    if (!idx) return;
    for (auto & kv : idx->dev)
    {
        int prev = -1;
        cudaGetDevice(&prev);
        cudaSetDevice(kv.first);
        free_device_index(kv.second);
        if (prev >= 0) cudaSetDevice(prev);
    }
    delete idx;
}

uint64_t speqg_pack_groups(const uint64_t * offsets, uint64_t n_reads)
{
    if (!offsets) return 0;
    return pack_groups(offsets, n_reads);
}

int speqg_pack_reads(const uint8_t * seq, const uint8_t * qual, const uint64_t * offsets, uint64_t n_reads, int encoding,
                     uint32_t phred_cutoff, int n_only, int threads, speqg_batch * out)
{
    SPEQG_TRY
    need(offsets && out, "NULL argument");
    need(n_reads == 0 || seq != nullptr || offsets[n_reads] == 0, "seq is NULL");
    need(out->bases && out->inv && out->grp_off && (out->len || n_reads == 0), "output arrays must be preallocated");
    pack_reads(seq, qual, offsets, n_reads, encoding, phred_cutoff, n_only != 0, threads, *out);
    SPEQG_CATCH
}

int speqg_unique_to_percent(const double * unique_in_reads, uint64_t total_in_reads, const double * unique_in_refs,
                            const double * total_in_refs, uint32_t n_groups, double * out)
{
    SPEQG_TRY
    need(unique_in_reads && unique_in_refs && total_in_refs && out, "NULL argument");
    // exact operation order of src/fm_scanner.cpp:1455-1474
    for (uint32_t i = 0; i < n_groups; ++i)
    {
        out[i] = 0.0;
        if (total_in_refs[i] > 0.0)
        {
            double percent_uniques = unique_in_refs[i] / total_in_refs[i];
            out[i] = 100.0 * unique_in_reads[i] / static_cast<double>(total_in_reads) / percent_uniques;
        }
    }
    SPEQG_CATCH
}

}   // extern "C"
