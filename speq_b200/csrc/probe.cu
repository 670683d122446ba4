// probe.cu -- measurement aid behind speqg_probe_random_sectors (include/speq_cuda.h): the rate at which this device serves
// INDEPENDENT random 32-byte sector loads from a table of a given size.  That is the roofline of the scan kernels -- every
// probe of the k-mer table and every rank step of the FM search is one data-dependent 32-byte sector -- and bench.py
// measures it in the same process, for the size of the structure the timed kernel probes (L2-resident or HBM-resident).
// Not part of the scan path.
#include <cuda_runtime.h>

#include <algorithm>

#include "capi_util.h"

namespace speqb
{
namespace
{
template <int ILP>
__global__ void __launch_bounds__(256) probe_gather_kernel(const uint8_t * __restrict__ base, const uint64_t sectors, const int iters,
                                                           uint32_t * __restrict__ out)
{
    uint64_t x = (blockIdx.x * 256ull + threadIdx.x) * 0x9E3779B97F4A7C15ull + 12345ull;
    uint32_t acc = 0u;
    for (int it = 0; it < iters; ++it)
    {
        uint32_t v[ILP];
#pragma unroll
        for (int j = 0; j < ILP; ++j)
        {
            x ^= x << 13; x ^= x >> 7; x ^= x << 17;                 // xorshift64: independent addresses, no load feeds the next
            const uint8_t * p = base + (x % sectors) * 32ull;
            uint32_t w[8];
            asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                         : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]), "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7]) : "l"(p));
            v[j] = w[0] ^ w[7];
        }
#pragma unroll
        for (int j = 0; j < ILP; ++j) acc ^= v[j];
    }
    if (acc == 0x12345678u) out[0] = acc;
}

template <int ILP> double run_probe(const uint8_t * d, uint64_t bytes, int ctas_per_sm, int sms, uint32_t * out, cudaStream_t st)
{
    const int grid = sms * ctas_per_sm, iters = (2048 / ILP);
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    probe_gather_kernel<ILP><<<grid, 256, 0, st>>>(d, bytes / 32, iters / 4, out);      // warm-up: the table into L2 where it fits
    CK(cudaEventRecord(e0, st));
    probe_gather_kernel<ILP><<<grid, 256, 0, st>>>(d, bytes / 32, iters, out);
    CK(cudaEventRecord(e1, st));
    CK(cudaEventSynchronize(e1));
    float ms = 0.f;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    CK(cudaGetLastError());
    return (double)grid * 256.0 * iters * ILP / (ms * 1e-3) / 1e9;
}
}   // namespace
}   // namespace speqb

using namespace speqb;

extern "C" int speqg_probe_random_sectors(int device, uint64_t table_bytes, double * gsectors_per_s)
{
    SPEQG_TRY
    need(gsectors_per_s != nullptr, "NULL argument");
    need(table_bytes >= 4096 && table_bytes <= (64ull << 30), "table_bytes must be in [4 KB, 64 GB]");
    device_guard guard(device);
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    uint8_t * d = nullptr;
    uint32_t * out = nullptr;
    cudaStream_t st = nullptr;
    double best = 0.0;
    try
    {
        CK(cudaMalloc(reinterpret_cast<void **>(&d), table_bytes));
        CK(cudaMalloc(reinterpret_cast<void **>(&out), 64));
        CK(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
        CK(cudaMemsetAsync(d, 1, table_bytes, st));
        // the ceiling is the best of a few shapes of the same access pattern: loads in flight per lane x resident CTAs per SM
        best = std::max(best, run_probe<8>(d, table_bytes, 8, sms, out, st));
        best = std::max(best, run_probe<16>(d, table_bytes, 8, sms, out, st));
        best = std::max(best, run_probe<16>(d, table_bytes, 4, sms, out, st));
        best = std::max(best, run_probe<4>(d, table_bytes, 8, sms, out, st));
    }
    catch (...)
    {
        cudaFree(d); cudaFree(out);
        if (st) cudaStreamDestroy(st);
        throw;
    }
    cudaFree(d); cudaFree(out);
    cudaStreamDestroy(st);
    *gsectors_per_s = best;
    SPEQG_CATCH
}
