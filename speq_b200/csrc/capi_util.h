// capi_util.h -- helpers shared by the translation units that implement the C ABI (capi.cu, capi_scan.cu).
#pragma once
#include <cuda_runtime.h>

#include <algorithm>
#include <string>

#include "speq_internal.h"

namespace speqb
{
inline thread_local std::string g_err;

inline int fail(int code, std::string const & msg)
{
    g_err = msg;
    return code;
}

#define SPEQG_TRY try {
#define SPEQG_CATCH                                                                      \
    }                                                                                    \
    catch (api_error const & e) { return fail(e.code, e.what()); }                       \
    catch (std::bad_alloc const &) { return fail(SPEQG_ENOMEM, "out of host memory"); }  \
    catch (std::exception const & e) { return fail(SPEQG_EINVAL, e.what()); }            \
    return SPEQG_OK;

inline void ck(cudaError_t e, const char * what)
{
    if (e != cudaSuccess)
        throw api_error(e == cudaErrorMemoryAllocation ? SPEQG_ENOMEM : SPEQG_ECUDA,
                        std::string(what) + ": " + cudaGetErrorName(e) + " (" + cudaGetErrorString(e) + ")");
}
#define CK(call) ck((call), #call)

inline void need(bool cond, const char * msg)
{
    if (!cond) throw api_error(SPEQG_EINVAL, msg);
}

struct device_guard
{
    int prev = -1;
    explicit device_guard(int dev)
    {
        int n = 0;
        cudaError_t e = cudaGetDeviceCount(&n);
        if (e != cudaSuccess || n == 0)
            throw api_error(SPEQG_ECUDA, std::string("no usable CUDA device (there is no CPU path): ") +
                                             (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0"));
        if (dev < 0 || dev >= n) throw api_error(SPEQG_EINVAL, "device index out of range");
        CK(cudaGetDevice(&prev));
        if (prev != dev) CK(cudaSetDevice(dev));
        else prev = -1;
    }
    ~device_guard()
    {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

template <class T> T * dev_alloc(size_t count, uint64_t * bytes = nullptr)
{
    T * p = nullptr;
    size_t b = std::max<size_t>(count, 1) * sizeof(T);
    CK(cudaMalloc(reinterpret_cast<void **>(&p), b));
    if (bytes) *bytes += b;
    return p;
}
template <class T> T * dev_upload(T const * src, size_t count, uint64_t * bytes = nullptr)
{
    T * p = dev_alloc<T>(count, bytes);
    if (count) CK(cudaMemcpy(p, src, count * sizeof(T), cudaMemcpyHostToDevice));
    return p;
}

inline void free_device_index(DeviceIndex & d)
{
    cudaFree(d.blocks); cudaFree(d.grp); cudaFree(d.lcp); cudaFree(d.pn); cudaFree(d.up); cudaFree(d.n_pos); cudaFree(d.ctab);
    cudaFree(d.t_bases); cudaFree(d.t_inv); cudaFree(d.t_grp_off); cudaFree(d.t_len); cudaFree(d.t_group);
    d = DeviceIndex{};
}


DeviceIndex & device_index(speqg_index * idx, int device);   // capi.cu: upload on first use
}   // namespace speqb
