// scanh.cu -- the k-mer table scan (SPEQG_ALGO_KTABLE): the FM search memoised per (index, k).
//
// What the reference computes per window is a function of the window's k-mer alone: the SA interval of the k-mer in the
// index (fm_index_cursor.hpp:110-146, k steps) and, after locate(), whether every occurrence lies in one variant and
// which (src/fm_scanner.cpp:166-177).  For k <= 31 that function is materialised once per scan handle:
//
//   ktable_build_kernel   walks the index's own texts (both strands of every scaffold, already resident as a packed batch)
//                         and inserts every N-free k-mer into an open-addressing table keyed by its 2-bit code:
//                         value = the variant of the text, or kValMulti once two variants have been seen.
//   ktable_lb_kernel      (record_multiplicity only) resolves each stored k-mer to the start of its SA interval with k
//                         rank steps on the blocked BWT, so multiplicities recorded per slot reach the refinement kernel
//                         as mult[lb] exactly as the rolling kernel records them.
//   scanh_kernel          one lane per read (pair): rolls the 2-bit k-mer of the window (one shift per base) and probes
//                         ONE 32-byte bucket per valid window -- one sector per window instead of two per base and no
//                         instruction spent on interval arithmetic.  A k-mer is absent from the index iff the probe
//                         reaches an empty slot, which is "the backward search failed" of the reference.
//
// Buckets are 32 bytes = 2 slots of {key, value}; a probe loads the bucket with one LDG.256 and moves to the next
// bucket only when both slots hold other keys.  Units are claimed in chunks of kClaimChunk per warp (one global atomic
// per chunk), tallies are the shared-memory ones of scan_dev.cuh.
#include <algorithm>

#include "kernels.h"
#include "scan_dev.cuh"

namespace speqb
{
namespace
{
constexpr int kBlockH = 256;
constexpr uint32_t kFullH = 0xFFFFFFFFu;
constexpr uint32_t kClaimChunk = 64;
constexpr uint64_t kKeyEmpty = ~0ull;
constexpr uint64_t kValEmpty = ~0ull;
constexpr uint64_t kValMulti = 0xFFFFFFFEull;      // low word; a group is < 65535
constexpr uint64_t kHashMul = 0x9E3779B97F4A7C15ull;

__device__ __forceinline__ uint32_t hbucket(uint64_t key, uint32_t nb)
{
    return __umulhi((uint32_t)((key * kHashMul) >> 32), nb);
}

// ---------------------------------------------------------------------------------------------------------------
// build: one thread per 32-base group of a text; the 32 windows that start in the group
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBlockH) ktable_build_kernel(const uint64_t * __restrict__ bases, const uint32_t * __restrict__ inv,
                                                               const uint64_t * __restrict__ grp_off, const uint32_t * __restrict__ len,
                                                               const int32_t * __restrict__ text_group, HSlot * __restrict__ tab,
                                                               const uint32_t nb, const uint32_t k, unsigned long long * __restrict__ n_keys)
{
    const uint32_t text = blockIdx.y;
    const uint32_t L = len[text];
    if (L < k) return;
    const uint32_t nwin = L - k + 1u;
    const uint64_t g0 = grp_off[text];
    const uint32_t grp = (uint32_t)text_group[text];
    const uint64_t mask = (k == 32u) ? ~0ull : ((1ull << (2u * k)) - 1ull);
    const uint64_t vmask = (1ull << k) - 1ull;
    for (uint32_t gi = blockIdx.x * blockDim.x + threadIdx.x; gi * 32u < nwin; gi += gridDim.x * blockDim.x)
    {
        // bases [32 gi, 32 gi + 64): every window that starts in this group (k <= 32)
        const uint32_t ngroups = (L + 31u) >> 5;
        const uint64_t b0 = bases[g0 + gi], b1 = gi + 1u < ngroups ? bases[g0 + gi + 1u] : 0ull;
        const uint64_t v = (uint64_t)inv[g0 + gi] | ((gi + 1u < ngroups ? (uint64_t)inv[g0 + gi + 1u] : 0ull) << 32);
        for (uint32_t j = 0; j < 32u; ++j)
        {
            const uint32_t p = gi * 32u + j;
            if (p >= nwin) break;
            if ((v >> j) & vmask) continue;                       // the window holds an N: no ACGT query matches it
            const uint64_t key = (j == 0u ? b0 : (b0 >> (2u * j)) | (b1 << (64u - 2u * j))) & mask;
            uint32_t bk = hbucket(key, nb);
            while (true)
            {
                HSlot * s = tab + 2ull * bk;
                bool placed = false;
                for (int w = 0; w < 2 && !placed; ++w)
                {
                    unsigned long long * kp = reinterpret_cast<unsigned long long *>(&s[w].key);
                    unsigned long long old = *reinterpret_cast<volatile unsigned long long *>(kp);
                    if (old == kKeyEmpty)
                    {
                        old = atomicCAS(kp, kKeyEmpty, key);
                        if (old == kKeyEmpty) atomicAdd(n_keys, 1ull);
                    }
                    if (old == kKeyEmpty || old == key)
                    {
                        unsigned long long * vp = reinterpret_cast<unsigned long long *>(&s[w].val);
                        const unsigned long long was = atomicCAS(vp, kValEmpty, (unsigned long long)grp);
                        if (was != kValEmpty && was != grp && was != kValMulti) atomicExch(vp, kValMulti);
                        placed = true;
                    }
                }
                if (placed) break;
                bk = bk + 1u == nb ? 0u : bk + 1u;
            }
        }
    }
}

// slot -> start of the k-mer's SA interval (k rank steps from the root); kNoGroup for empty slots
__global__ void __launch_bounds__(kBlockH) ktable_lb_kernel(const HSlot * __restrict__ tab, const uint64_t n_slots, const DevIndex ix,
                                                            const uint32_t k, uint32_t * __restrict__ slot_lb)
{
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_slots; i += (uint64_t)gridDim.x * blockDim.x)
    {
        const uint64_t key = tab[i].key;
        uint32_t out = kNoGroup;
        if (key != kKeyEmpty)
        {
            uint32_t lb = 0u, rb = ix.n - 1u;
            bool ok = true;
            for (uint32_t j = 0; j < k && ok; ++j) ok = fm_step(ix, ix.ctab, lb, rb, (uint32_t)(key >> (2u * j)) & 3u);
            if (ok) out = lb;
        }
        slot_lb[i] = out;
    }
}

// mult[lb] += per-slot multiplicities (and clear them): the refinement kernel reads mult[] as the rolling scan leaves it
__global__ void __launch_bounds__(kBlockH) ktable_fold_mult_kernel(unsigned long long * __restrict__ slot_mult, const uint32_t * __restrict__ slot_lb,
                                                                   const uint64_t n_slots, unsigned long long * __restrict__ mult)
{
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_slots; i += (uint64_t)gridDim.x * blockDim.x)
    {
        const unsigned long long m = slot_mult[i];
        if (m == 0ull) continue;
        slot_mult[i] = 0ull;
        const uint32_t lb = slot_lb[i];
        if (lb != kNoGroup) atomicAdd(mult + lb, m);
    }
}

// ---------------------------------------------------------------------------------------------------------------
// scan
// ---------------------------------------------------------------------------------------------------------------
enum : uint32_t { H_FETCH = 0, H_FIRST = 1, H_RUN = 2, H_PROBE = 3, H_DONE = 4 };

template <int MODE>
__global__ void __launch_bounds__(kBlockH, (MODE & M_LOCAL) ? 4 : 6) scanh_kernel(const HSlot * __restrict__ tab, const uint32_t nb, unsigned long long * __restrict__ slot_mult,
                                                           const BatchView b, const ScanArgs a, const int paired)
{
    constexpr bool LOCAL = (MODE & M_LOCAL) != 0;
    constexpr bool RECORD = (MODE & M_RECORD) != 0;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    Smem32 sm = smem32_view(smem_raw, a.V);
    smem32_clear(sm, a.V);
    Tally32<MODE> env{b, a, sm};

    const uint32_t k = a.k;
    const uint32_t sh = 2u * (k - 1u);
    const uint32_t units = (uint32_t)(paired ? b.n_reads / 2 : b.n_reads);
    const uint32_t lane = threadIdx.x & 31u;

    // lane state
    uint32_t phase = H_FETCH;
    bool mate2 = false;
    uint32_t u = 0u;                             // unit (read or pair)
    uint64_t g0 = 0ull;                          // first 32-base group of the current read
    uint32_t t = 0u, pend = 0u, run = 0u;        // next base, end of the read, valid bases behind t
    uint64_t key = 0ull;                         // the last k bases, base i of the window at bits [2i, 2i+2)
    uint32_t wlo = 0u, whi = 0u, winv = 0u;      // current 32-base group, base t in the low bits
    uint32_t nlo = 0u, nhi = 0u, ninv = 0u;      // the following group (prefetched)
    uint32_t pb = 0u;                            // bucket being probed
    uint32_t gmin = kNoGroup, gmax = 0u, c_total = 0u;
    uint32_t agg_g = kNoGroup, agg_n = 0u;
    double agg_w = 0.0;
    // warp-uniform chunk of claimed units
    uint32_t wcur = 0u, wend = 0u;

    while (true)
    {
        __syncwarp();
        const uint32_t ph = phase;
        if (!__any_sync(kFullH, ph != H_DONE)) break;
        const uint32_t m_fe = __ballot_sync(kFullH, ph == H_FETCH);
        if (m_fe)
        {
            // ---- next read: claim (first mate only), then length + first group offset
            const bool claim = ph == H_FETCH && !mate2;
            const uint32_t mc = __ballot_sync(kFullH, claim);
            if (mc)
            {
                const uint32_t cnt = (uint32_t)__popc(mc), left = wend - wcur;
                uint32_t nbase = 0u;
                if (cnt > left)
                {
                    if (lane == 0u) nbase = (uint32_t)atomicAdd(a.work, (unsigned long long)kClaimChunk);
                    nbase = __shfl_sync(kFullH, nbase, 0);
                }
                const uint32_t r = (uint32_t)__popc(mc & ((1u << lane) - 1u));
                if (claim) u = r < left ? wcur + r : nbase + (r - left);
                if (cnt > left) { wcur = nbase + (cnt - left); wend = nbase + kClaimChunk; }
                else wcur += cnt;
            }
            if (ph == H_FETCH)
            {
                if (u >= units) phase = H_DONE;
                else
                {
                    // issued now, consumed by H_FIRST in the next iteration
                    const uint64_t fr = paired ? 2ull * u + (mate2 ? 1u : 0u) : (uint64_t)u;
                    pend = load_ro(b.len + fr);
                    g0 = load_ro(b.grp_off + fr);
                    phase = H_FIRST;
                }
            }
        }
        if (ph != H_FETCH && ph != H_DONE)
        {
            // ---- loads of this iteration
            uint32_t w[8];
            bool look = false;
            if (ph == H_FIRST)
            {
                const uint32_t L = pend;
                pend = L >= k ? L : 0u;                           // L < k: the sliding view is empty (sliding.hpp:129-142)
                env.count_fetch(L >= k ? L - k + 1u : 0u);
                t = 0u; run = 0u; key = 0ull;
                if (pend)
                {
                    const uint64_t v = load_ro(b.bases + g0);
                    wlo = (uint32_t)v; whi = (uint32_t)(v >> 32);
                    winv = load_ro(b.inv + g0);
                }
                phase = H_RUN;
            }
            else if (ph == H_RUN)
            {
                if ((t & 31u) == 0u && ((t >> 5) + 1u) * 32u < pend)
                {
                    const uint64_t v = load_ro(b.bases + g0 + (t >> 5) + 1u);
                    nlo = (uint32_t)v; nhi = (uint32_t)(v >> 32);
                    ninv = load_ro(b.inv + g0 + (t >> 5) + 1u);
                }
                if (t < pend)
                {
                    const uint32_t c = wlo & 3u;
                    key = (key >> 2) | ((uint64_t)c << sh);
                    run = (winv & 1u) ? 0u : run + 1u;            // N / low quality: no window may contain this base
                    t += 1u;
                    wlo = (wlo >> 2) | (whi << 30); whi >>= 2; winv >>= 1;
                    if ((t & 31u) == 0u) { wlo = nlo; whi = nhi; winv = ninv; }
                    if (run >= k)
                    {
                        c_total += 1u;                            // ++total_kmers (src/fm_scanner.cpp:164)
                        pb = hbucket(key, nb);
                        look = true;
                    }
                }
            }
            else if (ph == H_PROBE) look = true;
            if (look)
            {
                const uint32_t * p = reinterpret_cast<const uint32_t *>(tab + 2ull * pb);
                asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                             : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]), "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7]) : "l"(p));
                const uint64_t k0 = u64of(w[0], w[1]), k1 = u64of(w[4], w[5]);
                const bool h0 = k0 == key, h1 = k1 == key;
                if (h0 || h1)
                {
                    const uint32_t g = h0 ? w[2] : w[6];
                    if (RECORD) atomicAdd(slot_mult + 2ull * pb + (h0 ? 0u : 1u), 1ull);
                    if (g != (uint32_t)kValMulti)
                    {
                        // ++unique_kmers[which_hit] and the read-level ambiguity state (src/fm_scanner.cpp:178-191)
                        gmin = g < gmin ? g : gmin;
                        gmax = g > gmax ? g : gmax;
                        if (g != agg_g)
                        {
                            if (agg_g != kNoGroup) env.unique(agg_g, agg_n, agg_w);
                            agg_g = g; agg_n = 0u; agg_w = 0.0;
                        }
                        if (LOCAL) agg_w += env.weight(g0, t, k);
                        else agg_n += 1u;
                    }
                    phase = H_RUN;
                }
                else if (k0 == kKeyEmpty || k1 == kKeyEmpty) phase = H_RUN;      // the k-mer does not occur in the index
                else
                {
                    pb = pb + 1u == nb ? 0u : pb + 1u;
                    phase = H_PROBE;
                }
            }
            if (phase == H_RUN && t >= pend)
            {
                // the read is finished
                if (paired && !mate2) mate2 = true;
                else
                {
                    // is_ambiguous: unique k-mers of two different variants in one read / pair (src/fm_scanner.cpp:183-190,213)
                    env.count_unit(c_total, gmin != kNoGroup && gmin != gmax);
                    c_total = 0u; gmin = kNoGroup; gmax = 0u;
                    mate2 = false;
                }
                phase = H_FETCH;
            }
        }
    }
    if (agg_g != kNoGroup) env.unique(agg_g, agg_n, agg_w);
    smem32_flush<LOCAL>(sm, a);
}

int g_sm_count_h = 0;
int sm_count_h()
{
    if (g_sm_count_h == 0)
    {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&g_sm_count_h, cudaDevAttrMultiProcessorCount, dev);
        if (g_sm_count_h <= 0) g_sm_count_h = 148;
    }
    return g_sm_count_h;
}
}   // namespace

#define SPEQB_SCANH_DISPATCH(M, EXPR)                                                   \
    switch (M)                                                                          \
    {                                                                                   \
        case 0: { auto kern = scanh_kernel<0>; EXPR; break; }                           \
        case M_RECORD: { auto kern = scanh_kernel<M_RECORD>; EXPR; break; }             \
        case M_LOCAL: { auto kern = scanh_kernel<M_LOCAL>; EXPR; break; }               \
        case M_LOCAL | M_RECORD: { auto kern = scanh_kernel<M_LOCAL | M_RECORD>; EXPR; break; } \
        default: return cudaErrorInvalidValue;                                          \
    }

cudaError_t scanh_launch_cfg(int mode, uint32_t V, LaunchCfg * cfg)
{
    const size_t smem = scan32_smem_bytes(V);
    int bps = 0;
    cudaError_t e = cudaSuccess;
    const int m = mode & (M_LOCAL | M_RECORD);
    SPEQB_SCANH_DISPATCH(m, {
        if (smem > 48 * 1024) e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, kern, kBlockH, smem);
    })
    if (e != cudaSuccess) return e;
    if (bps <= 0) return cudaErrorLaunchOutOfResources;
    cfg->block = kBlockH;
    cfg->grid = sm_count_h() * bps;
    return cudaSuccess;
}

cudaError_t launch_scanh(const HSlot * tab, uint32_t n_buckets, unsigned long long * slot_mult, BatchView const & b, ScanArgs const & a,
                         int mode, LaunchCfg cfg, cudaStream_t st)
{
    const size_t smem = scan32_smem_bytes(a.V);
    const int paired = (mode & M_PAIRED) ? 1 : 0;
    const int m = mode & (M_LOCAL | M_RECORD);
    cudaError_t e = cudaMemsetAsync(a.work, 0, sizeof(unsigned long long), st);
    if (e != cudaSuccess) return e;
    SPEQB_SCANH_DISPATCH(m, (kern<<<cfg.grid, cfg.block, smem, st>>>(tab, n_buckets, slot_mult, b, a, paired)))
    return cudaGetLastError();
}

cudaError_t launch_ktable_build(const uint64_t * t_bases, const uint32_t * t_inv, const uint64_t * t_grp_off, const uint32_t * t_len,
                                const int32_t * t_group, uint32_t n_texts, uint32_t max_len, HSlot * tab, uint32_t n_buckets, uint32_t k,
                                unsigned long long * n_keys, cudaStream_t st)
{
    cudaError_t e = cudaMemsetAsync(tab, 0xFF, (size_t)n_buckets * 2 * sizeof(HSlot), st);
    if (e != cudaSuccess) return e;
    if (n_texts == 0 || max_len < k) return cudaSuccess;
    const uint32_t groups = (max_len + 31u) / 32u;
    const uint32_t gx = std::min<uint32_t>((groups + kBlockH - 1) / kBlockH, (uint32_t)sm_count_h() * 8u);
    // grid.y <= 65535 texts per launch
    for (uint32_t t0 = 0; t0 < n_texts; t0 += 65535u)
    {
        const uint32_t nt = std::min<uint32_t>(65535u, n_texts - t0);
        ktable_build_kernel<<<dim3(gx, nt), kBlockH, 0, st>>>(t_bases, t_inv, t_grp_off + t0, t_len + t0, t_group + t0, tab, n_buckets, k, n_keys);
    }
    return cudaGetLastError();
}

cudaError_t launch_ktable_lb(const HSlot * tab, uint64_t n_slots, DevIndex const & ix, uint32_t k, uint32_t * slot_lb, cudaStream_t st)
{
    ktable_lb_kernel<<<sm_count_h() * 8, kBlockH, 0, st>>>(tab, n_slots, ix, k, slot_lb);
    return cudaGetLastError();
}

cudaError_t launch_ktable_fold_mult(unsigned long long * slot_mult, const uint32_t * slot_lb, uint64_t n_slots, unsigned long long * mult,
                                    cudaStream_t st)
{
    ktable_fold_mult_kernel<<<sm_count_h() * 8, kBlockH, 0, st>>>(slot_mult, slot_lb, n_slots, mult);
    return cudaGetLastError();
}
}   // namespace speqb
