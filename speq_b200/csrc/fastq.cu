// fastq.cu -- the FASTQ record grammar as kernels (speqg_scan_submit_fastq): raw file bytes in device memory -> a packed batch.
//
// What the reference gets from seqan3::sequence_file_input (record grammar: S3/include/seqan3/io/sequence_file/
// format_fastq.hpp:93-205: '@' id line, sequence, '+' line, |sequence| quality characters) for the common STRICT form of a
// FASTQ file -- four lines per record, LF or CRLF -- is done here on the raw bytes, so the host only has to bring file regions
// to the device:
//   fq_count_kernel      newlines per 4 KB chunk
//   fq_scan_kernel       exclusive prefix sum of the chunk counts (one CTA; a region has a few 10^4 chunks) and their total
//   fq_lines_kernel      line_start[l + 1] = position behind the l-th newline
//   fq_records_kernel    record r = lines 4r .. 4r+3: '@' / '+' line markers, |sequence| == |quality| (a trailing CR is not part
//                        of a line); where its sequence and quality start and how long they are; its first 32-base group
//   fq_pack_kernel       the packer of kernels.cu (dna5 / phred42 rules, 2-bit codes + invalid-base bits) reading the two lines
//                        where they lie in the raw bytes
//   fq_finalize_kernel   publishes the number of reads for the scan kernel that follows -- 0 when the region is not in the strict
//                        form (a marker or length check failed, or a character outside the IUPAC set was met: seqan3 skips blanks
//                        inside sequence lines and accepts multi-line records): such a region is NOT scanned and is reported
//                        (speqg_scan_fastq_rejected), so that the caller reads it again with the full grammar on the host.
#include <cuda_runtime.h>

#include <algorithm>

#include "kernels.h"

namespace speqb
{
namespace
{
constexpr uint32_t kFqChunk = 4096;     // bytes per chunk = 256 threads x 16 bytes
constexpr int kFqBlock = 256;

__device__ __forceinline__ uint32_t nl_mask16(const uint8_t * __restrict__ raw, uint64_t pos, uint64_t n)
{
    // bit j set: raw[pos + j] == '\n' (positions at and behind n never count); raw is 16-byte aligned and padded
    if (pos >= n) return 0u;
    const uint4 v = *reinterpret_cast<const uint4 *>(raw + pos);
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
    uint32_t m = 0u;
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int b = 0; b < 4; ++b)
            if (((w[i] >> (8 * b)) & 0xFFu) == 0x0Au) m |= 1u << (4 * i + b);
    const uint64_t left = n - pos;
    return left >= 16u ? m : (m & ((1u << left) - 1u));
}

__global__ void __launch_bounds__(kFqBlock) fq_count_kernel(const uint8_t * __restrict__ raw, const uint64_t n, uint32_t * __restrict__ chunk_cnt,
                                                            const uint32_t n_chunks)
{
    __shared__ uint32_t warp_sum[kFqBlock / 32];
    for (uint32_t c = blockIdx.x; c < n_chunks; c += gridDim.x)
    {
        uint32_t cnt = (uint32_t)__popc(nl_mask16(raw, (uint64_t)c * kFqChunk + threadIdx.x * 16u, n));
        cnt = __reduce_add_sync(0xFFFFFFFFu, cnt);
        if ((threadIdx.x & 31u) == 0u) warp_sum[threadIdx.x >> 5] = cnt;
        __syncthreads();
        if (threadIdx.x == 0)
        {
            uint32_t s = 0;
            for (int i = 0; i < kFqBlock / 32; ++i) s += warp_sum[i];
            chunk_cnt[c] = s;
        }
        __syncthreads();
    }
}

// in place: chunk_cnt[c] -> number of newlines in front of chunk c; total[0] = all newlines.  One CTA of 1024 threads.
__global__ void __launch_bounds__(1024) fq_scan_kernel(uint32_t * __restrict__ chunk_cnt, const uint32_t n_chunks, unsigned long long * __restrict__ total)
{
    __shared__ uint32_t warp_tot[32];
    __shared__ uint32_t carry;
    if (threadIdx.x == 0) carry = 0u;
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
    for (uint32_t base = 0; base < n_chunks; base += 1024u)
    {
        const uint32_t i = base + threadIdx.x;
        const uint32_t v = i < n_chunks ? chunk_cnt[i] : 0u;
        uint32_t x = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1)
        {
            const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, x, o);
            if ((int)lane >= o) x += y;
        }
        if (lane == 31u) warp_tot[wid] = x;
        __syncthreads();
        if (wid == 0u)
        {
            uint32_t t = warp_tot[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1)
            {
                const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, t, o);
                if ((int)lane >= o) t += y;
            }
            warp_tot[lane] = t;           // inclusive over the warps
        }
        __syncthreads();
        const uint32_t before = carry + (wid ? warp_tot[wid - 1u] : 0u) + (x - v);
        if (i < n_chunks) chunk_cnt[i] = before;
        __syncthreads();
        if (threadIdx.x == 1023u) carry = before + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) total[0] = carry;
}

__global__ void __launch_bounds__(kFqBlock) fq_lines_kernel(const uint8_t * __restrict__ raw, const uint64_t n, const uint32_t * __restrict__ chunk_off,
                                                            const uint32_t n_chunks, uint64_t * __restrict__ line_start)
{
    __shared__ uint32_t warp_tot[kFqBlock / 32];
    const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
    if (blockIdx.x == 0 && threadIdx.x == 0) line_start[0] = 0ull;
    for (uint32_t c = blockIdx.x; c < n_chunks; c += gridDim.x)
    {
        const uint64_t pos = (uint64_t)c * kFqChunk + threadIdx.x * 16u;
        uint32_t m = nl_mask16(raw, pos, n);
        const uint32_t v = (uint32_t)__popc(m);
        uint32_t x = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1)
        {
            const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, x, o);
            if ((int)lane >= o) x += y;
        }
        if (lane == 31u) warp_tot[wid] = x;
        __syncthreads();
        uint32_t before = chunk_off[c] + (x - v);
        for (uint32_t w = 0; w < wid; ++w) before += warp_tot[w];
        while (m)
        {
            const uint32_t j = (uint32_t)__ffs((int)m) - 1u;
            m &= m - 1u;
            line_start[(uint64_t)before + 1ull] = pos + j + 1ull;
            ++before;
        }
        __syncthreads();
    }
}

// one thread per record.  n_lines = number of lines (a last line without newline counts); line l = [line_start[l], end), where
// end = line_start[l + 1] - 1 (the newline), or n for an unterminated last line.
__global__ void __launch_bounds__(kFqBlock) fq_records_kernel(const uint8_t * __restrict__ raw, const uint64_t n, const uint64_t * __restrict__ line_start,
                                                              const uint64_t n_newlines, const uint64_t n_rec, uint64_t * __restrict__ soff,
                                                              uint64_t * __restrict__ qoff, uint32_t * __restrict__ len, uint64_t * __restrict__ grp_off,
                                                              const uint32_t stride, const uint32_t slot, const uint64_t gbase,
                                                              unsigned int * __restrict__ reject)
{
    // stride / slot: record r becomes read r * stride + slot of the batch (mates of a pair interleave: stride 2, slot = mate);
    // gbase: first 32-base group this file's reads may use (the second mate's groups lie behind all of the first's)
    for (uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n_rec; r += (uint64_t)gridDim.x * blockDim.x)
    {
        const uint64_t s0 = line_start[4 * r], s1 = line_start[4 * r + 1], s2 = line_start[4 * r + 2], s3 = line_start[4 * r + 3];
        const uint64_t e3 = 4 * r + 4 <= n_newlines ? line_start[4 * r + 4] - 1ull : n;      // end of the quality line
        uint64_t ls = (s2 - 1ull) - s1, lq = e3 - s3;
        if (ls && raw[s1 + ls - 1ull] == '\r') --ls;
        if (lq && raw[s3 + lq - 1ull] == '\r') --lq;
        // '@' opens the id line, '+' the separator line (format_fastq.hpp:112-125,172-186); equal lengths (:196-202)
        const bool ok = raw[s0] == '@' && raw[s2] == '+' && ls == lq && ls <= 0xFFFFFFFFull;    // (an empty line holds its newline there)
        if (!ok) atomicOr(reject, 1u);
        const uint64_t o = r * stride + slot;
        soff[o] = s1;
        qoff[o] = s3;
        len[o] = ok ? (uint32_t)ls : 0u;
        grp_off[o] = gbase + (s1 >> 5) + r;      // monotone, never overlapping: consecutive sequence lines are >= len bytes apart
        if (r + 1 == n_rec && slot + 1u == stride) grp_off[n_rec * stride] = gbase + (n >> 5) + n_rec + 1ull;
    }
}

__global__ void __launch_bounds__(kFqBlock) fq_pack_kernel(const uint8_t * __restrict__ raw, const uint64_t * __restrict__ soff,
                                                           const uint64_t * __restrict__ qoff, const uint32_t * __restrict__ len,
                                                           const uint64_t * __restrict__ grp_off, const uint64_t n_rec, const uint32_t stride,
                                                           const uint32_t slot, const uint32_t cutoff,
                                                           uint64_t * __restrict__ bases, uint32_t * __restrict__ inv, uint8_t * __restrict__ qual_out,
                                                           unsigned int * __restrict__ reject)
{
    // letters: dna5 (S3/include/seqan3/alphabet/nucleotide/dna5.hpp:87-119) over the legal IUPAC set of the sequence file
    // (S3/include/seqan3/io/sequence_file/input.hpp:145-162); anything else -> the region is handed back to the host reader
    __shared__ uint8_t code[256];
    for (int c = threadIdx.x; c < 256; c += blockDim.x)
    {
        uint8_t v = 255;
        switch (c)
        {
            case 'A': case 'a': v = 0; break;
            case 'C': case 'c': v = 1; break;
            case 'G': case 'g': v = 2; break;
            case 'T': case 't': case 'U': case 'u': v = 3; break;
            case 'N': case 'n': case 'R': case 'r': case 'Y': case 'y': case 'S': case 's': case 'W': case 'w':
            case 'K': case 'k': case 'M': case 'm': case 'B': case 'b': case 'D': case 'd': case 'H': case 'h':
            case 'V': case 'v': v = 4; break;
            default: break;
        }
        code[c] = v;
    }
    __syncthreads();
    if (*reinterpret_cast<volatile unsigned int *>(reject)) return;       // the records kernel found the region irregular
    const uint32_t lane = threadIdx.x & 31u;
    const uint64_t nwarps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
    bool any_bad = false;
    for (uint64_t r = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; r < n_rec; r += nwarps)
    {
        const uint64_t o = r * stride + slot;
        const uint64_t so = soff[o], qo = qoff[o], g0 = grp_off[o];
        const uint32_t L = len[o];
        const uint32_t ng = (L + 31u) >> 5;
        for (uint32_t g = lane; g < ng; g += 32u)
        {
            uint64_t w = 0;
            uint32_t m = 0;
            const uint32_t lim = min(32u, L - g * 32u);
            for (uint32_t j = 0; j < lim; ++j)
            {
                uint8_t c = code[raw[so + g * 32u + j]];
                if (c == 255) { any_bad = true; c = 4; }
                uint32_t iv = c == 4;
                if (c == 4) c = 0;
                // phred42: clamp(char - 33, 0, 41) (S3/include/seqan3/alphabet/quality/quality_base.hpp:139-160)
                const uint32_t q = raw[qo + g * 32u + j];
                const uint32_t ph = (q < 33u || q >= 128u) ? 0u : (q >= 75u ? 41u : q - 33u);
                if (!(ph > cutoff)) iv = 1;
                if (qual_out) qual_out[(g0 + g) * 32 + j] = (uint8_t)ph;
                w |= (uint64_t)c << (2u * j);
                m |= iv << j;
            }
            if (qual_out) for (uint32_t j = lim; j < 32u; ++j) qual_out[(g0 + g) * 32 + j] = 0;
            bases[g0 + g] = w;
            inv[g0 + g] = m;
        }
    }
    if (any_bad) atomicOr(reject, 1u);
}

__global__ void fq_finalize_kernel(const unsigned int * __restrict__ reject, const uint64_t n_rec, unsigned long long * __restrict__ n_reads_dev,
                                   unsigned int * __restrict__ region_flag)
{
    const unsigned int bad = *reject;
    *n_reads_dev = bad ? 0ull : n_rec;
    *region_flag = bad ? 1u : 0u;
}

int fq_sms()
{
    static int sms = 0;
    if (sms == 0)
    {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (sms <= 0) sms = 148;
    }
    return sms;
}
}   // namespace

uint32_t fastq_chunks(uint64_t n_bytes) { return (uint32_t)((n_bytes + kFqChunk - 1) / kFqChunk); }

cudaError_t launch_fastq_count(const uint8_t * raw, uint64_t n_bytes, uint32_t * chunk_cnt, unsigned long long * total, cudaStream_t st)
{
    const uint32_t nc = fastq_chunks(n_bytes);
    if (nc == 0) return cudaMemsetAsync(total, 0, 8, st);
    fq_count_kernel<<<std::min<uint32_t>(nc, (uint32_t)fq_sms() * 16u), kFqBlock, 0, st>>>(raw, n_bytes, chunk_cnt, nc);
    fq_scan_kernel<<<1, 1024, 0, st>>>(chunk_cnt, nc, total);
    return cudaGetLastError();
}

// lines, records and packing of ONE raw region into reads r * stride + slot of the batch; *reject is set when the region is not
// in the strict form (it must have been zeroed before the first region of the batch: launch_fastq_begin)
cudaError_t launch_fastq_records(const uint8_t * raw, uint64_t n_bytes, const uint32_t * chunk_off, uint64_t n_newlines, uint64_t n_rec,
                                 uint64_t * line_start, uint64_t * soff, uint64_t * qoff, uint32_t * len, uint64_t * grp_off, uint32_t stride,
                                 uint32_t slot, uint64_t gbase, uint32_t cutoff, uint64_t * bases, uint32_t * inv, uint8_t * qual_out,
                                 unsigned int * reject, cudaStream_t st)
{
    const uint32_t nc = fastq_chunks(n_bytes);
    if (nc && n_rec)
    {
        fq_lines_kernel<<<std::min<uint32_t>(nc, (uint32_t)fq_sms() * 16u), kFqBlock, 0, st>>>(raw, n_bytes, chunk_off, nc, line_start);
        const uint64_t want = (n_rec + kFqBlock - 1) / kFqBlock;
        fq_records_kernel<<<(int)std::min<uint64_t>(want, (uint64_t)fq_sms() * 16), kFqBlock, 0, st>>>(raw, n_bytes, line_start, n_newlines, n_rec, soff, qoff,
                                                                                                      len, grp_off, stride, slot, gbase, reject);
        const uint64_t wantp = (n_rec * 32 + kFqBlock - 1) / kFqBlock;
        fq_pack_kernel<<<(int)std::min<uint64_t>(wantp, (uint64_t)fq_sms() * 16), kFqBlock, 0, st>>>(raw, soff, qoff, len, grp_off, n_rec, stride, slot, cutoff,
                                                                                                    bases, inv, qual_out, reject);
    }
    return cudaGetLastError();
}

cudaError_t launch_fastq_begin(unsigned int * reject, cudaStream_t st) { return cudaMemsetAsync(reject, 0, 4, st); }

// n_reads: reads of the batch when the region(s) were regular
cudaError_t launch_fastq_finalize(const unsigned int * reject, uint64_t n_reads, unsigned long long * n_reads_dev, unsigned int * region_flag,
                                  cudaStream_t st)
{
    fq_finalize_kernel<<<1, 1, 0, st>>>(reject, n_reads, n_reads_dev, region_flag);
    return cudaGetLastError();
}
}   // namespace speqb
