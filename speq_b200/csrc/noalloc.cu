// noalloc.cu -- speqg_refine_noalloc: one refinement estimate that allocates nothing (include/speq_cuda.h).
//
// No counterpart in the reference; added for callers that cannot call cudaMalloc (code
// under CUDA graph capture, tight serving loops, harnesses that forbid it).  speqg_refine stages its inputs through buffers of
// the scan handle and reduces per-block partial sums that live in a buffer allocated at speqg_scan_open; here the inputs and
// the output are the caller's device arrays and the partial sums live in a file-scope __device__ array of a fixed size, which
// every GPU that loads the library holds until the process exits.  The kernels are the ones speqg_refine launches
// (refine_kernel + refine_reduce_kernel, kernels.cu) on the same grid, so the result is the same sum in the same order.
//
// The array's NAME is a host-side stand-in in host code, not a device address: its address is taken once per device with
// cudaGetSymbolAddress (that device current) and that pointer goes to the kernels.
#include <cuda_runtime.h>

#include <mutex>

#include "capi_util.h"
#include "kernels.h"

using namespace speqb;

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
    const unsigned long long * unique_n;
    double * d_percent, * d_tkpg, * d_partial;
    int32_t * d_gc;
    int refine_grid;
};
bool refine_view_ready(speqg_scan * s, scan_refine_view * out);   // capi_scan.cu
}   // namespace speqb

namespace
{
// per-block partial sums of the estimate: [refine grid][V] doubles.  1 MiB per GPU.
constexpr size_t kNoAllocDoubles = SPEQG_NOALLOC_MAX_DOUBLES;
static __device__ __align__(256) double g_refine_scratch[kNoAllocDoubles];

constexpr int kMaxDevices = 64;
std::mutex g_scratch_mutex;
double * g_scratch_ptr[kMaxDevices] = {};

// the device address of the array on `device` (which must be current): resolved once per device
double * scratch_on(int device)
{
    need(device >= 0 && device < kMaxDevices, "speqg_refine_noalloc: device index beyond 63");
    std::lock_guard<std::mutex> lock(g_scratch_mutex);
    if (g_scratch_ptr[device] == nullptr)
    {
        void * p = nullptr;
        CK(cudaGetSymbolAddress(&p, g_refine_scratch));
        need(p != nullptr, "speqg_refine_noalloc: the scratch array has no device address");
        g_scratch_ptr[device] = static_cast<double *>(p);
    }
    return g_scratch_ptr[device];
}
}   // namespace

extern "C" {

int speqg_refine_noalloc(speqg_scan * scan, const double * d_percent, const int32_t * d_group_counts, double * d_tkpg)
{
    SPEQG_TRY
    need(scan && d_percent && d_group_counts && d_tkpg, "NULL argument");
    scan_refine_view r{};
    if (!refine_view_ready(scan, &r))
        throw api_error(SPEQG_ESTATE, "speqg_refine_noalloc: work is pending on the handle (a deferred FASTQ region, multiplicities not "
                                      "yet folded): call speqg_refine or speqg_scan_multiplicity_count once after the last submit");
    if (!r.mult) throw api_error(SPEQG_ESTATE, "speqg_refine_noalloc needs a scan opened with record_multiplicity = 1");
    if ((size_t)r.refine_grid * (size_t)r.V > kNoAllocDoubles)
        throw api_error(SPEQG_EINVAL, "speqg_refine_noalloc: variants x refinement grid (" + std::to_string(r.V) + " x " +
                                          std::to_string(r.refine_grid) + ") exceeds the fixed scratch of " +
                                          std::to_string(kNoAllocDoubles) + " doubles (SPEQG_NOALLOC_MAX_DOUBLES); use speqg_refine");
    device_guard guard(r.device);
    double * partial = scratch_on(r.device);
    // every block writes its whole row of the scratch and the reduction reads exactly the rows written: nothing has to start at
    // zero, and what the previous call left is overwritten before it is read
    CK(launch_refine(r.view, r.k, r.mult, d_percent, d_group_counts, r.V, partial, r.refine_grid, 256, r.stream));
    CK(launch_refine_reduce(partial, r.refine_grid, r.V, d_tkpg, r.unique_n, d_percent, r.stream));
    SPEQG_CATCH
}

}   // extern "C"
