// scan32.cu -- the product scan kernel for sm_100a and the per-k index specialisation it runs on.
//
//   fill32_kernel   builds the 32-byte SBlocks (fm_layout.h) for one k from the stored index: rank counts and bit planes
//                   from the KBlocks, e bits (lcp >= k-1) and uq bits (the depth-k lcp interval of the suffix lies in one
//                   variant) from lcp, the previous/next-smaller links and the group-change counts.
//   scan32_kernel   one lane per read (pair): the flat state machine of scan32_core.h.  Every trip of the loop issues
//                   all loads of the iteration back to back (two LDG.256 for the blocks at lb and rb1; the read group,
//                   parent-jump and fetch loads only in the iterations that need them) and then consumes them, so a
//                   window costs one round trip to L2 / HBM.  Units are claimed with one warp-aggregated atomic, tallies
//                   live in shared memory and reach global memory with one atomic per variant and CTA.
//
// Reference behaviour covered: worker / do_a_count of speq::scan::_async_{one,two}_with_{global,local}_error_rate
// (src/fm_scanner.cpp:145-216, 418-491, 662-732, 911-998).
#include <algorithm>

#include "kernels.h"
#include "scan32_core.h"
#include "scan_dev.cuh"

namespace speqb
{
namespace
{
constexpr int kBlock32 = 256;
constexpr uint32_t kFull = 0xFFFFFFFFu;

template <int MODE, bool G8> struct Env32 : Tally32<MODE>
{
    const FBlock * fblk;
    DevIndex const & ix;
    __device__ __forceinline__ Env32(const FBlock * f, DevIndex const & i, BatchView const & b_, ScanArgs const & a_, Smem32 s)
        : Tally32<MODE>{b_, a_, s}, fblk(f), ix(i) {}
    __device__ __forceinline__ uint32_t load_grp(uint32_t i) const
    {
        // G8: the variant byte sits in the 64-byte burst of the suffix's own block
        uint32_t v;
        if (G8) asm volatile("ld.global.nc.u8 %0, [%1];" : "=r"(v) : "l"(&fblk[i >> 5].g[i & 31u]));
        else asm volatile("ld.global.nc.u16 %0, [%1];" : "=r"(v) : "l"(ix.grp + i));
        return v;
    }
    __device__ __forceinline__ uint32_t exc_word(uint32_t blk) const { return load_ro(&ix.blocks[blk >> 1].exc[blk & 1u]); }
    __device__ __forceinline__ uint64_t qual8(uint64_t g0, uint32_t pos) const
    {
        return load_ro(reinterpret_cast<const uint64_t *>(this->b.qual + 32ull * g0 + pos));
    }
    __device__ __forceinline__ uint32_t qual1(uint64_t g0, uint32_t pos) const { return load_ro(this->b.qual + 32ull * g0 + pos); }
    __device__ __forceinline__ double den(uint32_t q) const { return this->sm.den[q]; }
    __device__ __forceinline__ double rden(uint32_t q) const { return this->sm.rden[q]; }
    __device__ __forceinline__ void record(uint32_t lb) const { atomicAdd(this->a.mult + lb, 1ull); }
};

#ifndef SPEQB_RMINB_LOCAL
#define SPEQB_RMINB_LOCAL 4
#endif
template <int MODE, bool G8, bool MASK>
__global__ void __launch_bounds__(kBlock32, (MODE & M_LOCAL) ? SPEQB_RMINB_LOCAL : 4) scan32_kernel(const FBlock * __restrict__ fblk, const DevIndex ix, const BatchView b,
                                                             const ScanArgs a, const int paired)
{
    constexpr bool LOCAL = (MODE & M_LOCAL) != 0;
    constexpr bool RECORD = (MODE & M_RECORD) != 0;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    Smem32 sm = smem32_view(smem_raw, a.V);
    smem32_clear(sm, a.V);
    smem32_tables(sm, a, LOCAL);

    Cfg32 cfg;
    cfg.k = a.k;
    cfg.n = ix.n;
    const uint64_t n_reads_b = b.n_reads_dev ? *b.n_reads_dev : b.n_reads;
    cfg.units = paired ? n_reads_b / 2 : n_reads_b;
    cfg.paired = paired;
    cfg.local = LOCAL;
    cfg.record = RECORD;
    cfg.mask64 = MASK;
    cfg.skip_unique = 1;
    Env32<MODE, G8> env(fblk, ix, b, a, sm);
    const uint32_t lane = threadIdx.x & 31u;
    const uint64_t nthreads = (uint64_t)gridDim.x * blockDim.x;

    Lane32 s;
    lane32_init(s, (uint64_t)blockIdx.x * blockDim.x + threadIdx.x, ix.n);
    Loads32 ld;
#pragma unroll
    for (int i = 0; i < 8; ++i) ld.A[i] = ld.B[i] = 0u;

    while (true)
    {
        __syncwarp();
        const uint32_t ph = s.phase;
        const uint32_t m_up = __ballot_sync(kFull, ph == P_UP);
        const uint32_t m_fe = __ballot_sync(kFull, ph == P_FETCH);
        if (!__any_sync(kFull, ph != P_DONE)) break;
        // ---- every load of this iteration, issued before anything is consumed
        const bool stp = ph == P_STEP;
        ld256_if(ld.A, fblk + (s.lb >> 5), stp);
        ld256_if(ld.B, fblk + (s.rb1 >> 5), stp);
        {
            // the next 32 bases of the read: straight into the lane's window registers (all 32 old bases are consumed)
            const bool needw = stp && s.t < s.pend && (s.t & 31u) == 0u;
            ld32x2_if(s.wlo, s.whi, b.bases + s.g0 + (s.t >> 5), needw);
            ld32_if(s.winv, b.inv + s.g0 + (s.t >> 5), needw);
            if (LOCAL && stp) lane32_load_qual(s, cfg, env);
        }
        if (m_up)
        {
            const bool up = ph == P_UP;
            ld128_if(ld.A, ix.up + s.lb, up);
            ld128_if(ld.B, ix.up + s.rb1, up);
        }
        if (m_fe)
        {
            const bool fetch = ph == P_FETCH && s.u < cfg.units;
            const uint64_t fr = lane32_fetch_read(s, cfg);
            ld32_if(ld.B[4], b.len + fr, fetch);
            ld32x2_if(ld.B[5], ld.B[6], b.grp_off + fr, fetch);
            // claim the unit after this one: one atomic per warp and iteration
            const bool claim = fetch && !(s.flags & F_MATE2);
            const uint32_t mc = __ballot_sync(kFull, claim);
            if (mc)
            {
                const int leader = __ffs((int)mc) - 1;
                unsigned long long base = 0ull;
                if ((int)lane == leader) base = atomicAdd(a.work, (unsigned long long)__popc(mc));
                base = __shfl_sync(kFull, base, leader);
                if (claim) s.unext = nthreads + base + (uint64_t)__popc(mc & ((1u << lane) - 1u));
            }
        }
        // ---- consume
        if (stp) lane32_step(s, ld, cfg, env);
        else if (ph == P_UP) lane32_up(s, ld, cfg);
        else if (ph == P_FETCH) lane32_fetch(s, ld, cfg, env);
    }
    lane32_flush_agg(s, env);
    smem32_flush<LOCAL>(sm, a);
}

// ---------------------------------------------------------------------------------------------------------------
// Per-k specialisation: one warp per SBlock, one lane per suffix.
//   e bit   lcp[i] >= k-1
//   uq bit  the maximal run of suffixes around i whose neighbours share >= k symbols (the SA interval of the k-mer that
//           prefixes suffix i) holds no group change.  Its ends are found by hopping over the previous / next-smaller
//           links (every hop lowers the lcp value, so at most kLcpCap hops), the changes are counted with the KBlocks'
//           chg bits and cumulative counts.  Suffixes shorter than k are never the lb of a depth-k hit.
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t chg_upto(const KBlock * kb, uint32_t j)
{
    const KBlock * p = kb + (j >> 6);
    return load_ro(&p->chgcum) + popc64(u64of(load_ro(&p->chg[0]), load_ro(&p->chg[1])) & upto64(j & 63u));
}

__global__ void __launch_bounds__(kBlock32) fill32_kernel(const KBlock * __restrict__ kb, const uint8_t * __restrict__ lcp,
                                                          const uint2x * __restrict__ pn, const uint16_t * __restrict__ grp,
                                                          const uint32_t * __restrict__ ctab, FBlock * __restrict__ out,
                                                          const uint32_t n, const uint32_t k)
{
    const uint32_t nblocks = n / 32u + 1u;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t blk = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; blk < nblocks; blk += nwarps)
    {
        const uint32_t i = blk * 32u + lane;
        bool e = false, uq = false;
        if (i < n)
        {
            e = (uint32_t)lcp[i] + 1u >= k;
            uint32_t s = i;
            while ((uint32_t)lcp[s] >= k) s = pn[s].x;
            uint32_t t = i + 1u;
            while ((uint32_t)lcp[t] >= k) t = pn[t].y;
            uq = chg_upto(kb, t - 1u) == chg_upto(kb, s);
        }
        const uint32_t we = __ballot_sync(kFull, e), wu = __ballot_sync(kFull, uq);
        const KBlock * p = kb + (blk >> 1);
        const uint32_t half = blk & 1u;
        uint32_t v = 0u;
        if (lane < 4u)
        {
            v = ctab[lane] + p->cnt[lane];     // C[c] folded into the counts
            if (half)
            {
                const uint32_t xl = (lane & 1u) ? 0u : ~0u, xh = (lane & 2u) ? 0u : ~0u;
                v += popc32((p->lo[0] ^ xl) & (p->hi[0] ^ xh) & ~p->exc[0]);
            }
            if (lane == 0u && p->exc[half] != 0u) v |= kSFlagExc;
        }
        else if (lane == 4u) v = p->lo[half];
        else if (lane == 5u) v = p->hi[half];
        else if (lane == 6u) v = we;
        else if (lane == 7u) v = wu;
        if (lane < 8u) reinterpret_cast<uint32_t *>(out + blk)[lane] = v;
        // variants as bytes (meaningful when the index has <= 255 variants; 0xFFFF -> 0xFF)
        out[blk].g[lane] = i < n ? (uint8_t)grp[i] : (uint8_t)0xFF;
    }
}

__global__ void __launch_bounds__(kBlock32) build_up_kernel(const uint8_t * __restrict__ lcp, const uint2x * __restrict__ pn,
                                                            UpRec * __restrict__ out, const uint64_t count)
{
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (uint64_t)gridDim.x * blockDim.x)
    {
        const uint2x v = pn[i];
        UpRec r;
        r.psv = v.x; r.nsv = v.y; r.lcp = lcp[i]; r.pad = 0u;
        out[i] = r;
    }
}

int g_sm_count32 = 0;
int sm_count32()
{
    if (g_sm_count32 == 0)
    {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&g_sm_count32, cudaDevAttrMultiProcessorCount, dev);
        if (g_sm_count32 <= 0) g_sm_count32 = 148;
    }
    return g_sm_count32;
}
}   // namespace

size_t scan32_smem_bytes(uint32_t V) { return 32 + (size_t)V * 12 + 2 * 42 * 8; }

// G8: variants as bytes inside the blocks (V <= 255); MASK (implies G8): the unit's variants as a 64-bit mask (V <= 64)
#define SPEQB_SCAN32_CASE(M, G, MK, EXPR) { auto kern = scan32_kernel<M, G, MK>; EXPR; break; }
#define SPEQB_SCAN32_DISPATCH(M, V_, EXPR)                                                        \
    switch ((M) | ((V_) <= kMaxGroups8 ? 16 : 0) | ((V_) <= kMaxGroupsMask ? 32 : 0))             \
    {                                                                                             \
        case 0: SPEQB_SCAN32_CASE(0, false, false, EXPR)                                          \
        case M_RECORD: SPEQB_SCAN32_CASE(M_RECORD, false, false, EXPR)                            \
        case M_LOCAL: SPEQB_SCAN32_CASE(M_LOCAL, false, false, EXPR)                              \
        case M_LOCAL | M_RECORD: SPEQB_SCAN32_CASE(M_LOCAL | M_RECORD, false, false, EXPR)        \
        case 16: SPEQB_SCAN32_CASE(0, true, false, EXPR)                                          \
        case 16 | M_RECORD: SPEQB_SCAN32_CASE(M_RECORD, true, false, EXPR)                        \
        case 16 | M_LOCAL: SPEQB_SCAN32_CASE(M_LOCAL, true, false, EXPR)                          \
        case 16 | M_LOCAL | M_RECORD: SPEQB_SCAN32_CASE(M_LOCAL | M_RECORD, true, false, EXPR)    \
        case 48: SPEQB_SCAN32_CASE(0, true, true, EXPR)                                           \
        case 48 | M_RECORD: SPEQB_SCAN32_CASE(M_RECORD, true, true, EXPR)                         \
        case 48 | M_LOCAL: SPEQB_SCAN32_CASE(M_LOCAL, true, true, EXPR)                           \
        case 48 | M_LOCAL | M_RECORD: SPEQB_SCAN32_CASE(M_LOCAL | M_RECORD, true, true, EXPR)     \
        default: return cudaErrorInvalidValue;                                                    \
    }

cudaError_t scan32_launch_cfg(int mode, uint32_t V, LaunchCfg * cfg)
{
    const size_t smem = scan32_smem_bytes(V);
    int bps = 0;
    cudaError_t e = cudaSuccess;
    const int m = mode & (M_LOCAL | M_RECORD);
    SPEQB_SCAN32_DISPATCH(m, V, {
        if (smem > 48 * 1024) e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, kern, kBlock32, smem);
    })
    if (e != cudaSuccess) return e;
    if (bps <= 0) return cudaErrorLaunchOutOfResources;
    cfg->block = kBlock32;
    cfg->grid = sm_count32() * bps;
    return cudaSuccess;
}

cudaError_t launch_scan32(const FBlock * fblk, DevIndex const & ix, BatchView const & b, ScanArgs const & a, int mode, LaunchCfg cfg,
                          cudaStream_t st)
{
    const size_t smem = scan32_smem_bytes(a.V);
    const int paired = (mode & M_PAIRED) ? 1 : 0;
    const int m = mode & (M_LOCAL | M_RECORD);
    cudaError_t e = cudaMemsetAsync(a.work, 0, sizeof(unsigned long long), st);
    if (e != cudaSuccess) return e;
    SPEQB_SCAN32_DISPATCH(m, a.V, (kern<<<cfg.grid, cfg.block, smem, st>>>(fblk, ix, b, a, paired)))
    return cudaGetLastError();
}

cudaError_t launch_fill32(const KBlock * kb, const uint8_t * lcp, const uint2x * pn, const uint16_t * grp, const uint32_t * ctab,
                          FBlock * out, uint32_t n, uint32_t k, cudaStream_t st)
{
    fill32_kernel<<<sm_count32() * 8, kBlock32, 0, st>>>(kb, lcp, pn, grp, ctab, out, n, k);
    return cudaGetLastError();
}

cudaError_t launch_build_up(const uint8_t * lcp, const uint2x * pn, UpRec * out, uint64_t count, cudaStream_t st)
{
    build_up_kernel<<<sm_count32() * 8, kBlock32, 0, st>>>(lcp, pn, out, count);
    return cudaGetLastError();
}
}   // namespace speqb
