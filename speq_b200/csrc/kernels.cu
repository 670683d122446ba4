// kernels.cu -- sm_100a kernels of the scan hot path.
//
//   scan_roll_kernel    the product kernel: one lane per read (pair / text segment), rolling k-mer window over the
//                       blocked FM index (scan_core.h), group-uniqueness test on the SA-order group runs, per-variant
//                       tallies in shared memory, one global atomic per variant and CTA at the end.
//   scan_direct_kernel  every window searched from the root, one warp per read: the literal trip counts of the
//                       reference (seqan3 search_trivial -> extend_right) -- used for the roofline's work counters and
//                       as an independent cross-check of the rolling kernel.
//   refine_kernel       the refinement estimate over the recorded k-mer multiplicities.
//
// Reference behaviour covered: worker / do_a_count (src/fm_scanner.cpp:145-216, 418-491, 662-732, 911-998),
// search_and_count (src/fm_scanner.cpp:1503-1528), estimate lambdas (src/fm_scanner.cpp:1087-1125, 1176-1219).
#include <algorithm>

#include "kernels.h"
#include "scan_core.h"

namespace speqb
{
namespace
{
constexpr uint32_t kNone = 0xFFFFFFFFu;
constexpr int kBlock = 256;

__device__ __forceinline__ unsigned long long warp_sum(unsigned long long v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xFFFFFFFFu, v, o);
    return v;
}

struct SmemView
{
    double * w;        // [V] local-mode weighted unique tallies
    uint32_t * cnt;    // [V] unique tallies (global mode, back-check)
    uint32_t * tot;    // [V] back-check: windows per group
};
__device__ __forceinline__ SmemView smem_view(unsigned char * raw, uint32_t V)
{
    SmemView s;
    s.w = reinterpret_cast<double *>(raw);
    s.cnt = reinterpret_cast<uint32_t *>(raw + (size_t)V * 8);
    s.tot = s.cnt + V;
    return s;
}

// ---------------------------------------------------------------------------------------------------------------
// Rolling scan.  A lane owns one unit (read, read pair or text segment) at a time and runs the flat state machine of
// scan_core.h: every trip of the loop is ONE iteration of lane_plan -> loads -> lane_consume.  All loads of an
// iteration are issued back to back before anything is consumed, so an iteration costs one round trip to L2 / HBM
// whatever phase each lane is in, and the warp is re-converged at the top of every trip.
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void ld256_if(uint32_t * w, const void * p, bool pred)
{
    asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.s32 q, %9, 0;\n\t@q ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];\n\t}"
                 : "+r"(w[0]), "+r"(w[1]), "+r"(w[2]), "+r"(w[3]), "+r"(w[4]), "+r"(w[5]), "+r"(w[6]), "+r"(w[7])
                 : "l"(p), "r"((int)pred));
}
__device__ __forceinline__ void ld64_if(uint64_t & v, const void * p, bool pred)
{
    asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.s32 q, %2, 0;\n\t@q ld.global.nc.u64 %0, [%1];\n\t}" : "+l"(v) : "l"(p), "r"((int)pred));
}
__device__ __forceinline__ void ld32_if(uint32_t & v, const void * p, bool pred)
{
    asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.s32 q, %2, 0;\n\t@q ld.global.nc.u32 %0, [%1];\n\t}" : "+r"(v) : "l"(p), "r"((int)pred));
}

__device__ __forceinline__ void ld8_if(uint32_t & v, const void * p, bool pred)
{
    asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.s32 q, %2, 0;\n\t@q ld.global.nc.u8 %0, [%1];\n\t}" : "+r"(v) : "l"(p), "r"((int)pred));
}
__device__ __forceinline__ void ld64x2_if(uint32_t & x, uint32_t & y, const void * p, bool pred)
{
    asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.s32 q, %3, 0;\n\t@q ld.global.nc.v2.u32 {%0,%1}, [%2];\n\t}" : "+r"(x), "+r"(y) : "l"(p), "r"((int)pred));
}

template <int MODE> struct DevEnv
{
    static constexpr bool LOCAL = (MODE & M_LOCAL) != 0;
    DevIndex const & ix;
    BatchView const & b;
    ScanArgs const & a;
    SmemView sm;
    __device__ __forceinline__ uint32_t load_grp(uint32_t i) const
    {
        uint32_t v;
        asm volatile("ld.global.nc.u16 %0, [%1];" : "=r"(v) : "l"(ix.grp + i));
        return v;
    }
    __device__ __forceinline__ uint32_t npos_rank(uint32_t i) const { return lower_bound_u32(ix.n_pos, ix.n_npos, i); }
    __device__ __forceinline__ double weight(uint64_t g0, uint32_t t_end, uint32_t k) const
    {
        // window = positions [t_end-k, t_end); the left fold of src/fm_scanner.cpp:454, a / (1 - 1/pow(10, q/10)), with
        // the denominators tabulated by the host's libm: IEEE division keeps every weight bit-exact
        const uint8_t * q = b.qual + 32u * g0 + (t_end - k);
        double w = 1.0;
        for (uint32_t i = 0; i < k; ++i) w = w / a.qden[load_ro(q + i)];
        return w;
    }
    __device__ __forceinline__ void unique(uint32_t g, uint32_t n, double w) const
    {
        if (LOCAL) atomicAdd(&sm.w[g], w);
        else atomicAdd(&sm.cnt[g], n);
    }
    __device__ __forceinline__ void record(uint32_t lb) const { atomicAdd(a.mult + lb, 1ull); }
    __device__ __forceinline__ void backtotal(uint32_t g, uint32_t n) const { atomicAdd(&sm.tot[g], n); }
};

template <int MODE>
__global__ void __launch_bounds__(kBlock, 3) scan_roll_kernel(const DevIndex ix, const BatchView b, const ScanArgs a, const int paired)
{
    constexpr bool LOCAL = (MODE & M_LOCAL) != 0;
    constexpr bool RECORD = (MODE & M_RECORD) != 0;
    constexpr bool BACK = (MODE & M_BACKCHECK) != 0;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    SmemView sm = smem_view(smem_raw, a.V);
    __shared__ uint32_t ctab[16];
    if (threadIdx.x < 16) ctab[threadIdx.x] = ix.ctab[threadIdx.x];
    for (uint32_t g = threadIdx.x; g < a.V; g += blockDim.x)
    {
        sm.w[g] = 0.0;
        sm.cnt[g] = 0u;
        sm.tot[g] = 0u;
    }
    __syncthreads();

    ScanCfg cfg;
    cfg.k = a.k;
    cfg.units = BACK ? a.n_segs : (paired ? b.n_reads / 2 : b.n_reads);
    cfg.stride = (uint64_t)gridDim.x * blockDim.x;
    cfg.paired = BACK ? 0 : paired;
    cfg.local = LOCAL;
    cfg.record = RECORD;
    cfg.back = BACK;
    DevEnv<MODE> env{ix, b, a, sm};

    Lane s;
    lane_init(s, (uint64_t)blockIdx.x * blockDim.x + threadIdx.x);
    Loads ld;
#pragma unroll
    for (int i = 0; i < 16; ++i) ld.A.w[i] = ld.B.w[i] = 0u;
    ld.rdw = 0; ld.rdi = 0; ld.flen = ld.fstart = ld.fgroup = 0; ld.fg0 = 0;
    uint32_t seg[8] = {0, 0, 0, 0, 0, 0, 0, 0};

    while (true)
    {
        __syncwarp();
        if (!__any_sync(0xFFFFFFFFu, s.phase != PH_DONE)) break;
        Plan p;
        lane_plan(s, ix, b, cfg, p);
        // ---- every load of this iteration, issued before anything is consumed
        ld256_if(ld.A.w, p.pa, p.blocks);
        ld256_if(ld.B.w, p.pb, p.blocks);
        ld256_if(ld.A.w + 8, reinterpret_cast<const char *>(p.pa) + 32, p.blocks);
        ld256_if(ld.B.w + 8, reinterpret_cast<const char *>(p.pb) + 32, p.blocks);
        ld8_if(ld.A.w[0], ix.lcp + s.lb, p.up);
        ld8_if(ld.B.w[0], ix.lcp + s.rb + 1u, p.up);
        ld64x2_if(ld.A.w[2], ld.A.w[3], ix.pn + s.lb, p.up);
        ld64x2_if(ld.B.w[2], ld.B.w[3], ix.pn + s.rb + 1u, p.up);
        ld64_if(ld.rdw, p.pw, p.word);
        ld32_if(ld.rdi, p.pi, p.word);
        if (BACK)
        {
            ld256_if(seg, a.segs + (p.fetch ? p.fr : 0), p.fetch);
            ld.fg0 = u64of(seg[0], seg[1]);
            ld.fstart = seg[2];
            ld.flen = seg[3];
            ld.fgroup = seg[4];
        }
        else
        {
            ld32_if(ld.flen, b.len + p.fr, p.fetch);
            ld64_if(ld.fg0, b.grp_off + p.fr, p.fetch);
        }
        // ---- consume
        lane_consume(s, ld, ix, ctab, cfg, env);
    }
    lane_flush_agg(s, cfg, env);
    __syncwarp();
    unsigned long long v0 = warp_sum(s.c_total), v1 = warp_sum(s.c_amb), v2 = warp_sum(s.c_look), v3 = warp_sum(s.c_reads);
    if ((threadIdx.x & 31) == 0 && !BACK)
    {
        if (v0) atomicAdd(a.scalars + SC_TOTAL, v0);
        if (v1) atomicAdd(a.scalars + SC_AMBIG, v1);
        if (v2) atomicAdd(a.scalars + SC_LOOKUPS, v2);
        if (v3) atomicAdd(a.scalars + SC_READS, v3);
    }
    __syncthreads();
    for (uint32_t g = threadIdx.x; g < a.V; g += blockDim.x)
    {
        if (LOCAL) { if (sm.w[g] != 0.0) atomicAdd(a.unique_w + g, sm.w[g]); }
        else if (sm.cnt[g]) atomicAdd(a.unique + g, (unsigned long long)sm.cnt[g]);
        if (BACK && sm.tot[g]) atomicAdd(a.total_g + g, (unsigned long long)sm.tot[g]);
    }
}

// ---------------------------------------------------------------------------------------------------------------
// Direct scan: one warp per read (pair); lane j searches windows j, j+32, ... from the root.
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ bool window_has_inv(const uint32_t * inv, uint64_t g0, uint32_t s, uint32_t k)
{
    uint32_t e = s + k;   // exclusive
    for (uint32_t w = s >> 5; w <= ((e - 1u) >> 5); ++w)
    {
        uint32_t m = load_ro(inv + g0 + w);
        uint32_t lo = w == (s >> 5) ? (s & 31u) : 0u;
        uint32_t hi = w == ((e - 1u) >> 5) ? ((e - 1u) & 31u) : 31u;   // inclusive
        uint32_t mask = (hi == 31u ? 0xFFFFFFFFu : ((1u << (hi + 1u)) - 1u)) & ~((1u << lo) - 1u);
        if (m & mask) return true;
    }
    return false;
}

template <int STATS>
__global__ void __launch_bounds__(kBlock) scan_direct_kernel(const DevIndex ix, const BatchView b, const ScanArgs a, const int paired,
                                                             const int record)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    SmemView sm = smem_view(smem_raw, a.V);
    __shared__ uint32_t ctab[16];
    if (threadIdx.x < 16) ctab[threadIdx.x] = ix.ctab[threadIdx.x];
    for (uint32_t g = threadIdx.x; g < a.V; g += blockDim.x)
    {
        sm.w[g] = 0.0;
        sm.cnt[g] = 0u;
        sm.tot[g] = 0u;
    }
    __syncthreads();
    const uint32_t k = a.k;
    const uint32_t lane = threadIdx.x & 31u;
    const uint64_t units = paired ? b.n_reads / 2 : b.n_reads;
    const uint64_t nwarps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
    unsigned long long c_total = 0, c_look = 0, c_steps = 0, c_hit = 0, c_occ = 0, c_occsec = 0;
    uint32_t c_amb = 0, c_reads = 0;
    for (uint64_t u = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; u < units; u += nwarps)
    {
        uint32_t gmin = kNone, gmax = 0;
        for (int mate = 0; mate < (paired ? 2 : 1); ++mate)
        {
            uint64_t r = paired ? 2 * u + mate : u;
            uint32_t L = b.len[r];
            uint64_t g0 = b.grp_off[r];
            uint32_t nw = L >= k ? L - k + 1u : 0u;
            if (lane == 0)
            {
                ++c_reads;
                c_look += nw;
            }
            for (uint32_t s = lane; s < nw; s += 32u)
            {
                uint32_t lb, rb, steps = 0;
                if (STATS)
                {
                    bool ok = direct_window(ix, ctab, b.bases, b.inv, g0, s, k, lb, rb, steps);
                    c_steps += steps;
                    if (ok)
                    {
                        uint32_t occ = rb - lb + 1u;
                        ++c_hit;
                        c_occ += occ;
                        c_occsec += (occ + 31u) >> 5;
                    }
                }
                else
                {
                    if (window_has_inv(b.inv, g0, s, k)) continue;
                    ++c_total;
                    if (!direct_window(ix, ctab, b.bases, b.inv, g0, s, k, lb, rb, steps)) continue;
                    if (record) atomicAdd(a.mult + lb, 1ull);
                    uint32_t grp = interval_group(ix, lb, rb);
                    if (grp != kNoGroup)
                    {
                        gmin = min(gmin, grp);
                        gmax = max(gmax, grp);
                        atomicAdd(&sm.cnt[grp], 1u);
                    }
                }
            }
        }
        __syncwarp();
#pragma unroll
        for (int o = 16; o > 0; o >>= 1)
        {
            gmin = min(gmin, __shfl_xor_sync(0xFFFFFFFFu, gmin, o));
            gmax = max(gmax, __shfl_xor_sync(0xFFFFFFFFu, gmax, o));
        }
        if (lane == 0 && gmin != kNone && gmin != gmax) ++c_amb;
    }
    __syncwarp();
    unsigned long long v0 = warp_sum(c_total), v1 = warp_sum(c_amb), v2 = warp_sum(c_look), v3 = warp_sum(c_reads);
    unsigned long long v4 = warp_sum(c_steps), v5 = warp_sum(c_hit), v6 = warp_sum(c_occ), v7 = warp_sum(c_occsec);
    if (lane == 0)
    {
        if (v2) atomicAdd(a.scalars + SC_LOOKUPS, v2);
        if (v3) atomicAdd(a.scalars + SC_READS, v3);
        if (STATS)
        {
            if (v4) atomicAdd(a.scalars + SC_STEPS, v4);
            if (v5) atomicAdd(a.scalars + SC_HIT, v5);
            if (v6) atomicAdd(a.scalars + SC_OCC, v6);
            if (v7) atomicAdd(a.scalars + SC_OCCSEC, v7);
        }
        else
        {
            if (v0) atomicAdd(a.scalars + SC_TOTAL, v0);
            if (v1) atomicAdd(a.scalars + SC_AMBIG, v1);
        }
    }
    __syncthreads();
    if (!STATS)
        for (uint32_t g = threadIdx.x; g < a.V; g += blockDim.x)
            if (sm.cnt[g]) atomicAdd(a.unique + g, (unsigned long long)sm.cnt[g]);
}

// ---------------------------------------------------------------------------------------------------------------
// Refinement estimate (src/fm_scanner.cpp:1087-1125): for a valid window with hit histogram h[g] (all occurrences),
// a[g] = h[g] * percent[g] / group_counts[g], s = sum(a), tkpg[g] += a[g] / s when s > 0.  The per-window weight the
// reference multiplies h by is the same for every g of one window and cancels in a[g]/s.  Windows with the same
// k-mer share one SA interval, so the sum over windows is a sum over recorded interval starts times multiplicity.
// ---------------------------------------------------------------------------------------------------------------
// smallest p' > p with a group change at p' (chg bit), capped at `limit` (exclusive upper bound of the search)
__device__ __forceinline__ uint32_t next_change(const KBlock * blocks, uint32_t p, uint32_t limit)
{
    uint32_t q = p + 1u;
    while (q < limit)
    {
        const KBlock * kb = blocks + (q >> 6);
        uint64_t m = u64of(load_ro(&kb->chg[0]), load_ro(&kb->chg[1])) & from64(q & 63u);
        if (m) return min(limit, (q & ~63u) + ctz64(m));
        q = (q & ~63u) + 64u;
    }
    return limit;
}

__global__ void __launch_bounds__(kBlock) refine_kernel(const DevIndex ix, const uint32_t k, const unsigned long long * __restrict__ mult,
                                                        const double * __restrict__ percent, const int32_t * __restrict__ group_counts,
                                                        const uint32_t V, double * __restrict__ partial)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double * s_t = reinterpret_cast<double *>(smem_raw);
    double * s_p = s_t + V;
    double * s_c = s_p + V;
    for (uint32_t g = threadIdx.x; g < V; g += blockDim.x)
    {
        s_t[g] = 0.0;
        s_p[g] = percent[g];
        s_c[g] = (double)group_counts[g];
    }
    __syncthreads();
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < ix.n; i += stride)
    {
        unsigned long long m = mult[i];
        if (m == 0ull) continue;
        // the k-mer's interval [i, rb]: it ends before the first lcp < k
        uint32_t rb = (uint32_t)i;
        while (rb + 1u < ix.n && load_ro(ix.lcp + rb + 1u) >= k) ++rb;
        double s = 0.0;
        for (uint32_t p = (uint32_t)i; p <= rb;)
        {
            uint32_t e = next_change(ix.blocks, p, rb + 1u);   // run of one group = [p, e)
            uint32_t g = load_ro(ix.grp + p);
            s += (double)(e - p) * s_p[g] / s_c[g];
            p = e;
        }
        if (!(s > 0.0)) continue;
        for (uint32_t p = (uint32_t)i; p <= rb;)
        {
            uint32_t e = next_change(ix.blocks, p, rb + 1u);
            uint32_t g = load_ro(ix.grp + p);
            double av = (double)(e - p) * s_p[g] / s_c[g];
            atomicAdd(&s_t[g], (double)m * (av / s));
            p = e;
        }
    }
    __syncthreads();
    for (uint32_t g = threadIdx.x; g < V; g += blockDim.x) partial[(size_t)blockIdx.x * V + g] = s_t[g];
}

// copy the rank blocks and fill the e bits for one k: bit j of block b = (lcp[64b+j] >= k-1)
__global__ void __launch_bounds__(kBlock) fill_e_kernel(const KBlock * __restrict__ src, KBlock * __restrict__ dst, const uint8_t * __restrict__ lcp,
                                                        const uint32_t n, const uint32_t k)
{
    const uint32_t nblocks = n / kBlockSyms + 1u;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t blk = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; blk < nblocks; blk += nwarps)
    {
        const uint32_t * s32 = reinterpret_cast<const uint32_t *>(src + blk);
        uint32_t * d32 = reinterpret_cast<uint32_t *>(dst + blk);
        uint32_t i0 = blk * 64u + lane, i1 = i0 + 32u;
        bool e0 = i0 < n && (uint32_t)lcp[i0] + 1u >= k;
        bool e1 = i1 < n && (uint32_t)lcp[i1] + 1u >= k;
        uint32_t w0 = __ballot_sync(0xFFFFFFFFu, e0), w1 = __ballot_sync(0xFFFFFFFFu, e1);
        if (lane < 16)
        {
            uint32_t v = s32[lane];
            if (lane == 8) v = w0;
            if (lane == 9) v = w1;
            d32[lane] = v;
        }
    }
}

__global__ void refine_reduce_kernel(const double * __restrict__ partial, int nparts, uint32_t V, double * __restrict__ tkpg,
                                     const unsigned long long * __restrict__ unique_n, const double * __restrict__ percent)
{
    uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= V) return;
    double acc = 0.0;
    for (int p = 0; p < nparts; ++p) acc += partial[(size_t)p * V + g];
    // windows whose k-mer lies in variant g only were not recorded: each adds a[g] / a[g] = 1 while a[g] > 0, i.e. percent[g] > 0
    if (unique_n != nullptr && percent[g] > 0.0) acc += (double)unique_n[g];
    tkpg[g] = acc;
}

__global__ void count_nonzero_kernel(const unsigned long long * __restrict__ mult, uint32_t n, unsigned long long * out)
{
    unsigned long long c = 0;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) c += mult[i] != 0ull;
    c = warp_sum(c);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(out, c);
}

// ---------------------------------------------------------------------------------------------------------------
// Device packer: one warp per read, lane j packs 32-base groups j, j+32, ...  Alphabet rules as in pack.cpp.
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBlock) pack_kernel(const uint8_t * __restrict__ seq, const uint8_t * __restrict__ qual, const uint64_t off0,
                                                      const uint64_t * __restrict__ offsets, const uint64_t n_reads, const int encoding,
                                                      const uint32_t cutoff, const int n_only, uint64_t * __restrict__ bases,
                                                      uint32_t * __restrict__ inv, uint8_t * __restrict__ qual_out,
                                                      uint64_t * __restrict__ grp_off, const int make_grp_off, uint32_t * __restrict__ len,
                                                      unsigned long long * bad)
{
    __shared__ uint8_t code[256];
    for (int c = threadIdx.x; c < 256; c += blockDim.x)
    {
        uint8_t v = 255;
        if (encoding == 0)
        {
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
        }
        else if (c <= 4)
            v = c == 3 ? 4 : (c == 4 ? 3 : c);
        code[c] = v;
    }
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31u;
    const uint64_t nwarps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
    bool any_bad = false, any_long = false;
    for (uint64_t r = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; r < n_reads; r += nwarps)
    {
        // make_grp_off: read r starts at 32-base group floor(o / 32) + r -- monotone, never overlapping (at most one group
        // of slack per read), and computable without a prefix sum over the read lengths
        const uint64_t o = offsets[r] - off0, L = offsets[r + 1] - offsets[r];
        const uint64_t g0 = make_grp_off ? (o >> 5) + r : grp_off[r];
        const uint64_t ng = (L >> 32) ? 0 : (L + 31) >> 5;
        if (L >> 32) any_long = true;            // reported at finish; the read is presented as empty
        if (lane == 0)
        {
            len[r] = (L >> 32) ? 0u : (uint32_t)L;
            if (make_grp_off) grp_off[r] = g0;
        }
        for (uint64_t g = lane; g < ng; g += 32u)
        {
            uint64_t w = 0;
            uint32_t m = 0;
            const uint32_t lim = (uint32_t)min((uint64_t)32, L - g * 32);
            for (uint32_t j = 0; j < lim; ++j)
            {
                uint8_t c = code[seq[o + g * 32 + j]];
                if (c == 255) { any_bad = true; c = 4; }
                uint32_t iv = c == 4;
                if (c == 4) c = 0;
                if (qual)
                {
                    uint32_t q = qual[o + g * 32 + j];
                    uint32_t ph = encoding == 0 ? ((q < 33u || q >= 128u) ? 0u : (q >= 75u ? 41u : q - 33u)) : (q > 41u ? 41u : q);
                    if (!n_only && !(ph > cutoff)) iv = 1;
                    if (qual_out) qual_out[(g0 + g) * 32 + j] = (uint8_t)ph;
                }
                w |= (uint64_t)c << (2u * j);
                m |= iv << j;
            }
            if (qual_out) for (uint32_t j = lim; j < 32u; ++j) qual_out[(g0 + g) * 32 + j] = 0;
            bases[g0 + g] = w;
            inv[g0 + g] = m;
        }
    }
    if (any_bad) atomicOr(bad, 1ull);
    if (any_long) atomicOr(bad, 2ull);
}

int g_sm_count = 0;
int sm_count()
{
    if (g_sm_count == 0)
    {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&g_sm_count, cudaDevAttrMultiProcessorCount, dev);
        if (g_sm_count <= 0) g_sm_count = 148;
    }
    return g_sm_count;
}

template <class K> cudaError_t prep_kernel(K kern, size_t smem, int * blocks_per_sm)
{
    cudaError_t e = cudaSuccess;
    if (smem > 48 * 1024) e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    return cudaOccupancyMaxActiveBlocksPerMultiprocessor(blocks_per_sm, kern, kBlock, smem);
}
}   // namespace

size_t scan_smem_bytes(uint32_t V) { return (size_t)V * 16; }

#define SPEQB_ROLL_DISPATCH(M, EXPR)                                                                   \
    switch (M)                                                                                         \
    {                                                                                                  \
        case 0: { auto kern = scan_roll_kernel<0>; EXPR; break; }                                      \
        case M_RECORD: { auto kern = scan_roll_kernel<M_RECORD>; EXPR; break; }                        \
        case M_LOCAL: { auto kern = scan_roll_kernel<M_LOCAL>; EXPR; break; }                          \
        case M_LOCAL | M_RECORD: { auto kern = scan_roll_kernel<M_LOCAL | M_RECORD>; EXPR; break; }    \
        case M_BACKCHECK: { auto kern = scan_roll_kernel<M_BACKCHECK>; EXPR; break; }                  \
        default: return cudaErrorInvalidValue;                                                         \
    }

cudaError_t scan_launch_cfg(int mode, int direct, uint32_t V, LaunchCfg * cfg)
{
    size_t smem = scan_smem_bytes(V);
    int bps = 0;
    cudaError_t e = cudaSuccess;
    if (direct == 1) e = prep_kernel(scan_direct_kernel<0>, smem, &bps);
    else if (direct == 2) e = prep_kernel(scan_direct_kernel<1>, smem, &bps);
    else
    {
        int m = mode & (M_LOCAL | M_RECORD | M_BACKCHECK);
        SPEQB_ROLL_DISPATCH(m, e = prep_kernel(kern, smem, &bps))
    }
    if (e != cudaSuccess) return e;
    if (bps <= 0) return cudaErrorLaunchOutOfResources;
    cfg->block = kBlock;
    cfg->grid = sm_count() * bps;
    return cudaSuccess;
}

cudaError_t launch_scan_roll(DevIndex const & ix, BatchView const & b, ScanArgs const & a, int mode, LaunchCfg cfg, cudaStream_t st)
{
    size_t smem = scan_smem_bytes(a.V);
    int paired = (mode & M_PAIRED) ? 1 : 0;
    int m = mode & (M_LOCAL | M_RECORD | M_BACKCHECK);
    SPEQB_ROLL_DISPATCH(m, (kern<<<cfg.grid, cfg.block, smem, st>>>(ix, b, a, paired)))
    return cudaGetLastError();
}

cudaError_t launch_scan_direct(DevIndex const & ix, BatchView const & b, ScanArgs const & a, int mode, int stats, LaunchCfg cfg, cudaStream_t st)
{
    size_t smem = scan_smem_bytes(a.V);
    int paired = (mode & M_PAIRED) ? 1 : 0;
    int record = (mode & M_RECORD) ? 1 : 0;
    if (stats) scan_direct_kernel<1><<<cfg.grid, cfg.block, smem, st>>>(ix, b, a, paired, 0);
    else scan_direct_kernel<0><<<cfg.grid, cfg.block, smem, st>>>(ix, b, a, paired, record);
    return cudaGetLastError();
}

cudaError_t launch_refine(DevIndex const & ix, uint32_t k, const unsigned long long * mult, const double * percent,
                          const int32_t * group_counts, uint32_t V, double * partial, int grid, int block, cudaStream_t st)
{
    size_t smem = (size_t)V * 24;
    if (smem > 48 * 1024)
    {
        cudaError_t e = cudaFuncSetAttribute(refine_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    refine_kernel<<<grid, block, smem, st>>>(ix, k, mult, percent, group_counts, V, partial);
    return cudaGetLastError();
}

cudaError_t launch_refine_reduce(const double * partial, int nparts, uint32_t V, double * tkpg, const unsigned long long * unique_n,
                                 const double * percent, cudaStream_t st)
{
    refine_reduce_kernel<<<(V + 127) / 128, 128, 0, st>>>(partial, nparts, V, tkpg, unique_n, percent);
    return cudaGetLastError();
}

cudaError_t launch_pack(const uint8_t * seq, const uint8_t * qual, uint64_t off0, const uint64_t * offsets, uint64_t n_reads,
                        int encoding, uint32_t phred_cutoff, int n_only, uint64_t * bases, uint32_t * inv, uint8_t * qual_out,
                        uint64_t * grp_off, int make_grp_off, uint32_t * len, unsigned long long * bad, cudaStream_t st)
{
    if (n_reads == 0) return cudaSuccess;
    uint64_t want = (n_reads * 32 + kBlock - 1) / kBlock;
    int grid = (int)std::min<uint64_t>(want, (uint64_t)sm_count() * 16);
    pack_kernel<<<grid, kBlock, 0, st>>>(seq, qual, off0, offsets, n_reads, encoding, phred_cutoff, n_only, bases, inv, qual_out, grp_off, make_grp_off, len, bad);
    return cudaGetLastError();
}

cudaError_t launch_fill_e(const KBlock * src, KBlock * dst, const uint8_t * lcp, uint32_t n, uint32_t k, cudaStream_t st)
{
    fill_e_kernel<<<sm_count() * 8, kBlock, 0, st>>>(src, dst, lcp, n, k);
    return cudaGetLastError();
}

cudaError_t launch_count_nonzero(const unsigned long long * mult, uint32_t n, unsigned long long * out, cudaStream_t st)
{
    count_nonzero_kernel<<<sm_count() * 4, 256, 0, st>>>(mult, n, out);
    return cudaGetLastError();
}
}   // namespace speqb
