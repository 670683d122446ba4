// sa_device.cu -- suffix array of the indexed text on the GPU (SPEQG_BUILD_DEVICE_SA): prefix doubling.
//
// The reference sorts suffixes with divsufsort on one core (sdsl construct_sa.hpp:80-130); on the 2e8-symbol
// text of the bacterial config the host SA-IS here needs ~19 s of the 38 s index build.  The suffix array of a text is
// unique (the sentinel is the smallest symbol and occurs once), so any correct construction yields the array the
// reference's index is built on; tests compare this one with the host SA-IS element by element.
//
// Round 0 sorts the suffixes by their first 21 symbols (3 bits each in one 63-bit key).  Round h sorts by the pair
// (rank of suffix i, rank of suffix i + h), which orders by the first 2h symbols, and re-ranks; the depth doubles until
// every rank is unique.  The sort and the scans are the CUDA toolkit's library primitives (cub::DeviceRadixSort,
// cub::DeviceScan, cub::DeviceReduce -- a plain sort, like calling cuBLAS for a plain GEMM); the key building and
// ranking kernels are below.  This is index construction, not the timed scan path.
#include <cub/cub.cuh>
#include <cuda_runtime.h>

#include <cstdint>

namespace speqb
{
namespace
{
constexpr int kBlk = 256;
constexpr uint32_t kSyms0 = 21;   // symbols in the round-0 key (3 bits each)

__global__ void sa_key0_kernel(const uint8_t * __restrict__ T, uint32_t n, uint64_t * __restrict__ key, uint32_t * __restrict__ sa)
{
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    {
        uint64_t k = 0;
        for (uint32_t j = 0; j < kSyms0; ++j) k = (k << 3) | (i + j < n ? (uint64_t)T[i + j] : 0ull);
        key[i] = k;
        sa[i] = i;
    }
}

// head[p] = p when the sorted key at p differs from its predecessor (a new group starts), else 0
__global__ void sa_heads_kernel(const uint64_t * __restrict__ key, uint32_t n, uint32_t * __restrict__ head)
{
    for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < n; p += gridDim.x * blockDim.x)
        head[p] = (p == 0 || key[p] != key[p - 1]) ? p : 0u;
}

// after the inclusive max-scan head[p] is the first position of p's group: that is the rank of suffix sa[p];
// singles counts the positions that start a group (all n of them once the order is total)
__global__ void sa_rank_kernel(const uint32_t * __restrict__ sa, const uint32_t * __restrict__ head, uint32_t n, uint32_t * __restrict__ rank)
{
    for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < n; p += gridDim.x * blockDim.x) rank[sa[p]] = head[p];
}

__global__ void sa_count_heads_kernel(const uint32_t * __restrict__ head, uint32_t n, unsigned long long * __restrict__ out)
{
    unsigned long long c = 0;
    for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < n; p += gridDim.x * blockDim.x) c += (p == 0 || head[p] != 0u) ? 1u : 0u;
    c = __reduce_add_sync(0xFFFFFFFFu, (unsigned)c);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(out, c);
}

// key of round h for the suffix at sorted position p: (rank[i], rank[i + h] + 1), 0 when i + h runs off the text
__global__ void sa_keyh_kernel(const uint32_t * __restrict__ sa, const uint32_t * __restrict__ rank, uint32_t n, uint32_t h,
                               uint64_t * __restrict__ key)
{
    for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < n; p += gridDim.x * blockDim.x)
    {
        const uint32_t i = sa[p];
        const uint64_t hi = rank[i];
        const uint64_t lo = (uint64_t)i + h < n ? (uint64_t)rank[i + h] + 1ull : 0ull;
        key[p] = (hi << 32) | lo;
    }
}
// ---- arrays derived from the suffix array: the random gathers and the suffix comparisons the host would do one by one
// text as 4-bit symbols, 16 per word (symbol j of a word at bits [4j, 4j+4)), zero padded
__global__ void sa_nibbles_kernel(const uint8_t * __restrict__ T, uint32_t n, uint64_t * __restrict__ nib, uint32_t n_words)
{
    for (uint32_t w = blockIdx.x * blockDim.x + threadIdx.x; w < n_words; w += gridDim.x * blockDim.x)
    {
        uint64_t v = 0;
        for (uint32_t j = 0; j < 16u; ++j)
        {
            const uint64_t p = (uint64_t)w * 16u + j;
            if (p < n) v |= (uint64_t)(T[p] & 15u) << (4u * j);
        }
        nib[w] = v;
    }
}

__device__ __forceinline__ uint64_t nib16(const uint64_t * __restrict__ nib, uint32_t pos)
{
    const uint32_t w = pos >> 4, off = pos & 15u;
    const uint64_t lo = nib[w] >> (4u * off);
    return off ? lo | (nib[w + 1] << (64u - 4u * off)) : lo;
}

// bwt[i] = T[SA[i] - 1] (wrapping), own[i] = owner of text position SA[i], lcp[i] = min(cap, lcp(SA[i-1], SA[i])), lcp[0] = 0.
// The sentinel is unique, so two different suffixes differ at or before the sentinel of the shorter one: the zero padding
// behind the text is never compared as equal.
__global__ void sa_derive_kernel(const uint8_t * __restrict__ T, const uint16_t * __restrict__ owner, const uint64_t * __restrict__ nib,
                                 const uint32_t * __restrict__ sa, uint32_t n, uint32_t cap, uint8_t * __restrict__ bwt,
                                 uint16_t * __restrict__ own, uint8_t * __restrict__ lcp)
{
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    {
        const uint32_t b = sa[i];
        bwt[i] = T[b ? b - 1u : n - 1u];
        own[i] = owner[b];
        uint32_t l = 0;
        if (i)
        {
            const uint32_t a = sa[i - 1];
            while (l < cap)
            {
                const uint64_t x = nib16(nib, a + l) ^ nib16(nib, b + l);
                if (x) { l += (uint32_t)(__ffsll((long long)x) - 1) >> 2; break; }
                l += 16u;
            }
            l = l < cap ? l : cap;
        }
        lcp[i] = (uint8_t)l;
    }
}
// ---- previous / next strictly smaller value over lcp[0, n] (lcp[n] = 0): two levels of block minima, so a search walks
// at most one block of values, one superblock of block minima and the superblock minima
constexpr uint32_t kPnB = 256;    // values per block, blocks per superblock

__global__ void sa_blockmin_kernel(const uint8_t * __restrict__ v, uint32_t count, uint8_t * __restrict__ out, uint32_t n_out)
{
    for (uint32_t b = blockIdx.x * blockDim.x + threadIdx.x; b < n_out; b += gridDim.x * blockDim.x)
    {
        uint32_t m = 255u;
        const uint64_t lo = (uint64_t)b * kPnB, hi = lo + kPnB < count ? lo + kPnB : count;
        for (uint64_t i = lo; i < hi; ++i) m = v[i] < m ? v[i] : m;
        out[b] = (uint8_t)m;
    }
}

// pn[i] = {largest j < i with lcp[j] < lcp[i] (0 when there is none), smallest j > i with lcp[j] < lcp[i] (n when none)}
__global__ void sa_pn_kernel(const uint8_t * __restrict__ lcp, const uint8_t * __restrict__ bmin, const uint8_t * __restrict__ smin,
                             uint32_t n, uint2 * __restrict__ pn)
{
    const uint32_t count = n + 1u;                                 // positions 0..n
    const uint32_t nb = (count + kPnB - 1u) / kPnB, ns = (nb + kPnB - 1u) / kPnB;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x)
    {
        const uint32_t v = lcp[i];
        uint32_t x = 0u, y = n;
        if (v)
        {
            // ---- to the left
            int64_t j = (int64_t)i - 1;
            const int64_t blo = (int64_t)(i / kPnB) * kPnB;
            while (j >= blo && lcp[j] >= v) --j;
            if (j < blo && blo > 0)
            {
                int64_t b = (int64_t)(i / kPnB) - 1;                // blocks of the same superblock
                const int64_t slo = (b + 1) / kPnB * kPnB;
                while (b >= slo && bmin[b] >= v) --b;
                if (b < slo)
                {
                    int64_t sb = slo / kPnB - 1;                    // superblocks
                    while (sb >= 0 && smin[sb] >= v) --sb;
                    b = -1;
                    if (sb >= 0)
                    {
                        b = (sb + 1) * kPnB - 1;
                        while (bmin[b] >= v) --b;
                    }
                }
                j = -1;
                if (b >= 0)
                {
                    j = (b + 1) * (int64_t)kPnB - 1;
                    while (lcp[j] >= v) --j;
                }
            }
            x = j >= 0 ? (uint32_t)j : 0u;
            // ---- to the right (lcp[n] = 0 < v ends every search)
            uint64_t r = (uint64_t)i + 1u;
            const uint64_t bhi = ((uint64_t)(i / kPnB) + 1u) * kPnB < count ? ((uint64_t)(i / kPnB) + 1u) * kPnB : count;
            while (r < bhi && lcp[r] >= v) ++r;
            if (r == bhi && r < count)
            {
                uint64_t b = i / kPnB + 1u;
                const uint64_t shi = (b / kPnB + 1u) * kPnB < nb ? (b / kPnB + 1u) * kPnB : nb;
                while (b < shi && bmin[b] >= v) ++b;
                if (b == shi)
                {
                    uint64_t sb = shi / kPnB;
                    while (sb < ns && smin[sb] >= v) ++sb;
                    b = sb * kPnB;
                    while (bmin[b] >= v) ++b;
                }
                r = b * kPnB;
                while (lcp[r] >= v) ++r;
            }
            y = (uint32_t)r;
        }
        pn[i] = make_uint2(x, y);
    }
}
}   // namespace

#define SA_CK(x)                                     \
    do                                               \
    {                                                \
        err = (x);                                   \
        if (err != cudaSuccess) goto done;           \
    } while (0)

// T: n symbols on the host (values < 8, T[n-1] = 0 the unique smallest); SA: n entries on the host
// owner / bwt / own / lcp: optional (all or none) -- owner[p] = u16 group of text position p on the host; outputs n entries each
// pn: n + 1 pairs {previous, next strictly smaller lcp position}
cudaError_t device_suffix_array(const uint8_t * T, uint64_t n64, int32_t * SA, const uint16_t * owner, uint32_t lcp_cap, uint8_t * bwt,
                                uint16_t * own, uint8_t * lcp, uint32_t * pn)
{
    const uint32_t n = (uint32_t)n64;
    cudaError_t err = cudaSuccess;
    uint8_t * d_T = nullptr;
    uint64_t * d_key = nullptr, * d_key2 = nullptr;
    uint32_t * d_sa = nullptr, * d_sa2 = nullptr, * d_rank = nullptr, * d_head = nullptr;
    unsigned long long * d_cnt = nullptr;
    uint16_t * d_owner = nullptr;
    uint64_t * d_nib = nullptr;
    uint8_t * d_min = nullptr;
    uint2 * d_pn = nullptr;
    void * d_tmp = nullptr;
    size_t tmp_sort = 0, tmp_scan = 0, tmp_bytes = 0;
    int sms = 148, dev = 0;
    cudaStream_t st = nullptr;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int grid = sms * 8;
    int rank_bits = 1;
    while ((1ull << rank_bits) < (uint64_t)n + 1) ++rank_bits;

    SA_CK(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    SA_CK(cudaMalloc(&d_T, n));
    SA_CK(cudaMalloc(&d_key, (size_t)n * 8));
    SA_CK(cudaMalloc(&d_key2, (size_t)n * 8));
    SA_CK(cudaMalloc(&d_sa, (size_t)n * 4));
    SA_CK(cudaMalloc(&d_sa2, (size_t)n * 4));
    SA_CK(cudaMalloc(&d_rank, (size_t)n * 4));
    SA_CK(cudaMalloc(&d_head, (size_t)n * 4));
    SA_CK(cudaMalloc(&d_cnt, 8));
    SA_CK(cub::DeviceRadixSort::SortPairs(nullptr, tmp_sort, d_key, d_key2, d_sa, d_sa2, (int)n, 0, 64, st));
    SA_CK(cub::DeviceScan::InclusiveScan(nullptr, tmp_scan, d_head, d_head, cub::Max(), (int)n, st));
    tmp_bytes = tmp_sort > tmp_scan ? tmp_sort : tmp_scan;
    SA_CK(cudaMalloc(&d_tmp, tmp_bytes));
    SA_CK(cudaMemcpyAsync(d_T, T, n, cudaMemcpyHostToDevice, st));

    sa_key0_kernel<<<grid, kBlk, 0, st>>>(d_T, n, d_key, d_sa);
    for (uint32_t h = kSyms0, round = 0;; ++round)
    {
        // sort (key, suffix); round 0 keys use 63 bits, later keys (rank, rank + 1) use rank_bits + 1 bits in each half
        const int end_bit = round == 0 ? 63 : 32 + rank_bits;
        SA_CK(cub::DeviceRadixSort::SortPairs(d_tmp, tmp_bytes, d_key, d_key2, d_sa, d_sa2, (int)n, 0, end_bit, st));
        {
            uint64_t * tk = d_key; d_key = d_key2; d_key2 = tk;
            uint32_t * ts = d_sa; d_sa = d_sa2; d_sa2 = ts;
        }
        sa_heads_kernel<<<grid, kBlk, 0, st>>>(d_key, n, d_head);
        SA_CK(cudaMemsetAsync(d_cnt, 0, 8, st));
        sa_count_heads_kernel<<<grid, kBlk, 0, st>>>(d_head, n, d_cnt);
        unsigned long long groups = 0;
        SA_CK(cudaMemcpyAsync(&groups, d_cnt, 8, cudaMemcpyDeviceToHost, st));
        SA_CK(cudaStreamSynchronize(st));
        if (groups == n) break;                                   // every suffix has its own rank: the order is total
        if (round > 40 || (uint64_t)h >= 2ull * n) { err = cudaErrorUnknown; goto done; }
        SA_CK(cub::DeviceScan::InclusiveScan(d_tmp, tmp_bytes, d_head, d_head, cub::Max(), (int)n, st));
        sa_rank_kernel<<<grid, kBlk, 0, st>>>(d_sa, d_head, n, d_rank);
        sa_keyh_kernel<<<grid, kBlk, 0, st>>>(d_sa, d_rank, n, h, d_key);
        h = h > 0x40000000u ? 0x80000000u : h * 2u;
    }
    SA_CK(cudaGetLastError());
    SA_CK(cudaMemcpyAsync(SA, d_sa, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
    SA_CK(cudaStreamSynchronize(st));
    if (owner)
    {
        // the sort buffers are free now: d_key holds the three output arrays (4 n + 1 bytes <= 8 n)
        const uint32_t n_words = n / 16u + 10u;                    // + padding read by the last comparisons (cap + 16 symbols)
        uint8_t * d_bwt = reinterpret_cast<uint8_t *>(d_key);
        uint8_t * d_lcp = d_bwt + n;
        uint16_t * d_own = reinterpret_cast<uint16_t *>(d_bwt + 2 * (size_t)n + (n & 1u));
        SA_CK(cudaMalloc(&d_nib, (size_t)n_words * 8));
        SA_CK(cudaMalloc(&d_owner, (size_t)n * 2));
        SA_CK(cudaMemcpyAsync(d_owner, owner, (size_t)n * 2, cudaMemcpyHostToDevice, st));
        sa_nibbles_kernel<<<grid, kBlk, 0, st>>>(d_T, n, d_nib, n_words);
        sa_derive_kernel<<<grid, kBlk, 0, st>>>(d_T, d_owner, d_nib, d_sa, n, lcp_cap, d_bwt, d_own, d_lcp);
        SA_CK(cudaGetLastError());
        SA_CK(cudaMemcpyAsync(bwt, d_bwt, n, cudaMemcpyDeviceToHost, st));
        SA_CK(cudaMemcpyAsync(lcp, d_lcp, n, cudaMemcpyDeviceToHost, st));
        SA_CK(cudaMemcpyAsync(own, d_own, (size_t)n * 2, cudaMemcpyDeviceToHost, st));
        SA_CK(cudaStreamSynchronize(st));
        if (pn)
        {
            // lcp[0, n] with lcp[n] = 0 in d_key2 (free), minima behind it; the links go to d_key ((n + 1) * 8 bytes needs its
            // own buffer: the outputs above have been copied out, d_key has n * 8)
            const uint32_t count = n + 1u, nb = (count + kPnB - 1u) / kPnB, ns = (nb + kPnB - 1u) / kPnB;
            uint8_t * d_l = reinterpret_cast<uint8_t *>(d_key2);
            SA_CK(cudaMalloc(&d_min, (size_t)nb + ns + 16));
            SA_CK(cudaMalloc(&d_pn, (size_t)count * 8));
            SA_CK(cudaMemcpyAsync(d_l, d_lcp, n, cudaMemcpyDeviceToDevice, st));
            SA_CK(cudaMemsetAsync(d_l + n, 0, 1, st));
            sa_blockmin_kernel<<<grid, kBlk, 0, st>>>(d_l, count, d_min, nb);
            sa_blockmin_kernel<<<grid, kBlk, 0, st>>>(d_min, nb, d_min + nb, ns);
            sa_pn_kernel<<<grid, kBlk, 0, st>>>(d_l, d_min, d_min + nb, n, d_pn);
            SA_CK(cudaGetLastError());
            SA_CK(cudaMemcpyAsync(pn, d_pn, (size_t)count * 8, cudaMemcpyDeviceToHost, st));
            SA_CK(cudaStreamSynchronize(st));
        }
    }
done:
    cudaFree(d_owner); cudaFree(d_nib); cudaFree(d_min); cudaFree(d_pn);
    cudaFree(d_T); cudaFree(d_key); cudaFree(d_key2); cudaFree(d_sa); cudaFree(d_sa2); cudaFree(d_rank); cudaFree(d_head);
    cudaFree(d_cnt); cudaFree(d_tmp);
    if (st) cudaStreamDestroy(st);
    return err;
}
}   // namespace speqb

extern "C" int speqb_device_suffix_array(const uint8_t * T, uint64_t n, int32_t * SA, const uint16_t * owner, uint32_t lcp_cap,
                                         uint8_t * bwt, uint16_t * own, uint8_t * lcp, uint32_t * pn, const char ** msg)
{
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e == cudaSuccess && count == 0) e = cudaErrorNoDevice;
    if (e == cudaSuccess) e = speqb::device_suffix_array(T, n, SA, owner, lcp_cap, bwt, own, lcp, pn);
    if (msg) *msg = cudaGetErrorString(e);
    return (int)e;
}
