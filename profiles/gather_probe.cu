// gather_probe.cu -- measurement aid (not product code): the rate at which a B200 serves independent random 32-byte
// sector reads from a table much larger than L2 -- the "HBM random-access roofline" the scan kernel's C4 numbers are
// compared with.  Build + run on the GPU box:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/gather_probe profiles/gather_probe.cu && /tmp/gather_probe
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

template <int BYTES, int ILP>
__global__ void __launch_bounds__(256) gather(const uint8_t * __restrict__ base, uint64_t units, uint32_t stride, int iters, uint32_t * out)
{
    uint64_t x = (blockIdx.x * 256ull + threadIdx.x) * 0x9E3779B97F4A7C15ull + 12345;
    uint32_t acc = 0;
    for (int it = 0; it < iters; ++it)
    {
        uint32_t v[ILP];
#pragma unroll
        for (int j = 0; j < ILP; ++j)
        {
            x ^= x << 13; x ^= x >> 7; x ^= x << 17;
            const uint8_t * p = base + (x % units) * stride;
            if (BYTES == 32)
            {
                uint32_t w[8];
                asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                             : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]), "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7]) : "l"(p));
                v[j] = w[0] ^ w[7];
            }
            else if (BYTES == 64)   // both sectors of one 64-byte burst
            {
                uint32_t w[8], z[8];
                asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                             : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]), "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7]) : "l"(p));
                asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                             : "=r"(z[0]), "=r"(z[1]), "=r"(z[2]), "=r"(z[3]), "=r"(z[4]), "=r"(z[5]), "=r"(z[6]), "=r"(z[7]) : "l"(p + 32));
                v[j] = w[0] ^ z[7];
            }
            else
            {
                uint32_t w;
                asm volatile("ld.global.nc.u32 %0, [%1];" : "=r"(w) : "l"(p));
                v[j] = w;
            }
        }
#pragma unroll
        for (int j = 0; j < ILP; ++j) acc ^= v[j];
    }
    if (acc == 0x12345678u) out[0] = acc;
}

template <int BYTES, int ILP> void run(const char * name, const uint8_t * d, uint64_t bytes, uint32_t stride, uint32_t * out)
{
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const int grid = sms * 8, iters = 512 / ILP * 4;
    const uint64_t units = bytes / stride;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    gather<BYTES, ILP><<<grid, 256>>>(d, units, stride, iters / 4, out);
    cudaEventRecord(e0);
    gather<BYTES, ILP><<<grid, 256>>>(d, units, stride, iters, out);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    const double acc = (double)grid * 256 * iters * ILP;
    printf("%-34s table %5.1f GB stride %3u: %7.2f G accesses/s  %8.1f GB/s useful  (%.2f ms)\n", name, bytes / 1e9, stride, acc / ms / 1e6,
           acc * BYTES / ms / 1e6, ms);
}

int main()
{
    const uint64_t big = 4ull << 30, small = 64ull << 20;
    uint8_t * d = nullptr;
    uint32_t * out = nullptr;
    cudaMalloc(&d, big);
    cudaMalloc(&out, 64);
    cudaMemset(d, 1, big);
    run<32, 4>("32 B sector, ILP 4", d, big, 32, out);
    run<32, 8>("32 B sector, ILP 8", d, big, 32, out);
    run<32, 8>("32 B sector at 64 B stride, ILP 8", d, big, 64, out);
    run<32, 8>("32 B sector at 128 B stride, ILP 8", d, big, 128, out);
    run<64, 4>("64 B burst (2 sectors), ILP 4", d, big, 64, out);
    run<4, 8>("4 B word, ILP 8", d, big, 4, out);
    for (uint64_t sz = 256ull << 20; sz <= big; sz <<= 1) run<32, 8>("32 B sector, table size sweep", d, sz, 32, out);
    for (uint64_t sz = 256ull << 20; sz <= big; sz <<= 1) run<64, 4>("64 B burst, table size sweep", d, sz, 64, out);
    run<32, 8>("32 B sector, L2-resident table", d, small, 32, out);
    run<64, 4>("64 B burst, L2-resident table", d, small, 64, out);
    cudaError_t e = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(e));
    return e == cudaSuccess ? 0 : 1;
}
