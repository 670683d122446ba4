// scan_dev.cuh -- device helpers shared by the scan kernels (scan32.cu, scanh.cu): predicated read-only loads and the
// CTA's shared-memory tallies (do_a_count's counters, src/fm_scanner.cpp:153-196 / :426-471 of the reference).
#pragma once
#include "kernels.h"
#include "scan_core.h"

namespace speqb
{
namespace
{
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
__device__ __forceinline__ void ld32x2_if(uint32_t & x, uint32_t & y, const void * p, bool pred)
{
    asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.s32 q, %3, 0;\n\t@q ld.global.nc.v2.u32 {%0,%1}, [%2];\n\t}" : "+r"(x), "+r"(y) : "l"(p), "r"((int)pred));
}
__device__ __forceinline__ void ld32_if(uint32_t & v, const void * p, bool pred)
{
    asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.s32 q, %2, 0;\n\t@q ld.global.nc.u32 %0, [%1];\n\t}" : "+r"(v) : "l"(p), "r"((int)pred));
}
__device__ __forceinline__ void ld128_if(uint32_t * w, const void * p, bool pred)
{
    asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.s32 q, %5, 0;\n\t@q ld.global.nc.v4.u32 {%0,%1,%2,%3}, [%4];\n\t}"
                 : "+r"(w[0]), "+r"(w[1]), "+r"(w[2]), "+r"(w[3])
                 : "l"(p), "r"((int)pred));
}

// shared-memory tallies of one CTA
struct Smem32
{
    double * w;                   // [V] local-mode weighted unique tallies
    uint32_t * cnt;               // [V] unique tallies (global mode; local mode: windows behind the weights)
    unsigned long long * scal;    // [4] total, ambiguous, lookups, reads
    double * den;                 // [kQTab] 1 - 10^(-q/10) and
    double * rden;                // [kQTab] its reciprocal, by phred value (local mode)
};
constexpr uint32_t kQTab = 42;    // phred42: 0..41
__device__ __forceinline__ Smem32 smem32_view(unsigned char * raw, uint32_t V)
{
    Smem32 s;
    s.scal = reinterpret_cast<unsigned long long *>(raw);
    s.w = reinterpret_cast<double *>(raw + 32);
    s.den = s.w + V;
    s.rden = s.den + kQTab;
    s.cnt = reinterpret_cast<uint32_t *>(s.rden + kQTab);
    return s;
}

// Local-error mode: the weight of a window is the product over its k bases of 1 / (1 - 10^(-q/10)) (the reference folds the
// divisions left to right from 1.0, src/fm_scanner.cpp:454).  Recomputing it per unique hit costs k dependent fp64 divisions
// and k byte loads; here it is a ROLLING product along the read: times 1/den[q] for the base that enters, times den[q] for the
// base that leaves (both tabulated by the host's libm), reset behind an invalid base, and recomputed from scratch every 256
// bases so that the rounding of at most ~2 * 256 + k multiplications separates it from the reference's k divisions: below
// 1e-13 relative per window, an order of magnitude inside the 1e-12 the local-mode tallies are held to.  The quality bytes of
// the entering and the leaving base come from two 8-byte streams (one aligned load per 8 bases each, prefetched one ahead).
struct QRoll
{
    double P = 1.0;
    uint64_t qn = 0ull, qnn = 0ull, qo = 0ull, qon = 0ull;
    const uint8_t * qb = nullptr;     // quality bytes of the read (32-byte aligned: reads start at 32-base groups)
    uint32_t lim = 0u;                // loadable bytes: the read's length rounded up to its last 32-base group
    __device__ __forceinline__ uint64_t ld8(uint32_t pos8) const
    {
        uint64_t v = 0ull;
        if (pos8 < lim) v = load_ro(reinterpret_cast<const uint64_t *>(qb + pos8));
        return v;
    }
    // the lane starts consuming bases of a read of L bases at position t
    __device__ __forceinline__ void start(const uint8_t * q, uint32_t L, uint32_t t, uint32_t k)
    {
        qb = q;
        lim = (L + 31u) & ~31u;
        P = 1.0;
        const uint32_t u = (t > k ? t : k) - k;        // first position of the leaving stream (read when t' = u + k)
        qn = ld8(t & ~7u);  qnn = ld8((t & ~7u) + 8u);
        qo = ld8(u & ~7u);  qon = ld8((u & ~7u) + 8u);
    }
    // the base at position t has been consumed; run_after = valid bases ending at t (0: the base itself is invalid)
    __device__ __forceinline__ void step(uint32_t t, uint32_t k, uint32_t run_after, const double * den, const double * rden)
    {
        const uint32_t q = (uint32_t)(qn >> (8u * (t & 7u))) & 0xFFu;
        if ((t & 7u) == 7u) { qn = qnn; qnn = ld8((t & ~7u) + 16u); }
        uint32_t qold = 1u;
        if (t >= k)
        {
            const uint32_t u = t - k;
            qold = (uint32_t)(qo >> (8u * (u & 7u))) & 0xFFu;
            if ((u & 7u) == 7u) { qo = qon; qon = ld8((u & ~7u) + 16u); }
        }
        if (run_after == 0u) P = 1.0;
        else
        {
            P *= rden[q];
            if (run_after > k) P *= den[qold];
            if ((t & 255u) == 255u && run_after >= k)
            {
                double e = 1.0;                                       // periodic exact restart (long reads)
                for (uint32_t i = t + 1u - k; i <= t; ++i) e *= rden[load_ro(qb + i)];
                P = e;
            }
        }
    }
};


// The tally side of a scan kernel's Env: weighted / counted unique hits per variant, total / ambiguous / lookups / reads.
template <int MODE> struct Tally32
{
    static constexpr bool LOCAL = (MODE & M_LOCAL) != 0;
    BatchView const & b;
    ScanArgs const & a;
    Smem32 sm;
    // n windows whose k-mer lies in variant g only; w: the sum of their phred weights (local mode)
    __device__ __forceinline__ void unique(uint32_t g, uint32_t n, double w) const
    {
        if (LOCAL) atomicAdd(&sm.w[g], w);
        atomicAdd(&sm.cnt[g], n);
    }
    // 64-bit shared counters as two native 32-bit atomics (a 64-bit shared atomicAdd is a CAS loop)
    __device__ __forceinline__ void add64(int i, uint32_t v) const
    {
        uint32_t * w = reinterpret_cast<uint32_t *>(sm.scal + i);
        const uint32_t old = atomicAdd(w, v);
        if (old + v < old) atomicAdd(w + 1, 1u);
    }
    __device__ __forceinline__ void count_fetch(uint32_t nwin) const
    {
        add64(2, nwin);
        add64(3, 1u);
    }
    __device__ __forceinline__ void count_unit(uint32_t total, bool amb) const
    {
        if (total) add64(0, total);
        if (amb) add64(1, 1u);
    }
    // an ambiguous unit (is_ambiguous, src/fm_scanner.cpp:183-190): which variants met -- the key of the reference's
    // fusion_totals map (src/fm_scanner.cpp:983-995); one insert-or-increment in the device map per ambiguous unit
#ifndef SPEQB_FUSION_NOINLINE
#define SPEQB_FUSION_NOINLINE 0
#endif
#if SPEQB_FUSION_NOINLINE
    __device__ __noinline__ void fusion(unsigned long long key) const
#else
    __device__ __forceinline__ void fusion(unsigned long long key) const
#endif
    {
        if (a.fkeys == nullptr) return;
        uint64_t x = key * 0x9E3779B97F4A7C15ull;
        uint32_t h = (uint32_t)(x >> 32) & a.fcap_mask;
        for (uint32_t probe = 0; probe < kFusionMaxProbe; ++probe)
        {
            unsigned long long old = *reinterpret_cast<volatile unsigned long long *>(a.fkeys + h);
            if (old == 0ull) old = atomicCAS(a.fkeys + h, 0ull, key);
            if (old == 0ull || old == key)
            {
                atomicAdd(a.fvals + h, 1ull);
                return;
            }
            h = (h + 1u) & a.fcap_mask;
        }
        atomicAdd(a.fvals + a.fcap_mask + 1u, 1ull);      // the map is full around this key: counted, not keyed
    }
};

// The variants with a unique k-mer in the current unit (read / pair): the ambiguity state of do_a_count
// (src/fm_scanner.cpp:183-190) and the key of the fusion map.  MASK: one bit per variant (V <= 64); else smallest / largest
// variant and a "more in between" flag.
template <bool MASK> struct UnitSet;
template <> struct UnitSet<true>
{
    uint64_t m = 0ull;
    __device__ __forceinline__ void add(uint32_t g) { m |= 1ull << g; }
    __device__ __forceinline__ bool ambiguous() const { return (m & (m - 1ull)) != 0ull; }
    __device__ __forceinline__ unsigned long long key() const { return m; }
    __device__ __forceinline__ void reset() { m = 0ull; }
    __device__ __forceinline__ void warp_merge()
    {
        const uint32_t lo = __reduce_or_sync(0xFFFFFFFFu, (uint32_t)m), hi = __reduce_or_sync(0xFFFFFFFFu, (uint32_t)(m >> 32));
        m = ((uint64_t)hi << 32) | lo;
    }
};
template <> struct UnitSet<false>
{
    uint32_t gmin = kNoGroup, gmax = 0u;
    bool more = false;
    __device__ __forceinline__ void add(uint32_t g)
    {
        if (gmin == kNoGroup) { gmin = g; gmax = g; }
        else if (g < gmin) { more = more || gmin != gmax; gmin = g; }
        else if (g > gmax) { more = more || gmin != gmax; gmax = g; }
        else more = more || (g != gmin && g != gmax);
    }
    __device__ __forceinline__ bool ambiguous() const { return gmin != kNoGroup && gmin != gmax; }
    __device__ __forceinline__ unsigned long long key() const { return fusion_key_minmax(gmin, gmax, more); }
    __device__ __forceinline__ void reset() { gmin = kNoGroup; gmax = 0u; more = false; }
    __device__ __forceinline__ void warp_merge()
    {
        const uint32_t lo = __reduce_min_sync(0xFFFFFFFFu, gmin), hi = __reduce_max_sync(0xFFFFFFFFu, gmax);
        // a variant other than the warp's smallest / largest: inside one lane's range (its flag) or one of a lane's ends
        const bool in = more || (gmin != kNoGroup && ((gmin != lo && gmin != hi) || (gmax != lo && gmax != hi)));
        more = __any_sync(0xFFFFFFFFu, in);
        gmin = lo; gmax = hi;
    }
};

__device__ __forceinline__ void smem32_clear(Smem32 const & sm, uint32_t V)
{
    if (threadIdx.x < 4) sm.scal[threadIdx.x] = 0ull;
    for (uint32_t g = threadIdx.x; g < V; g += blockDim.x)
    {
        sm.w[g] = 0.0;
        sm.cnt[g] = 0u;
    }
}
// ... and the phred tables of local mode (host libm values, ScanArgs::qden / qrden)
__device__ __forceinline__ void smem32_tables(Smem32 const & sm, ScanArgs const & a, bool local)
{
    if (local)
        for (uint32_t q = threadIdx.x; q < kQTab; q += blockDim.x)
        {
            sm.den[q] = a.qden[q];
            sm.rden[q] = a.qrden[q];
        }
    __syncthreads();
}

template <bool LOCAL>
__device__ __forceinline__ void smem32_flush(Smem32 const & sm, ScanArgs const & a)
{
    __syncthreads();
    if (threadIdx.x == 0)
    {
        if (sm.scal[0]) atomicAdd(a.scalars + SC_TOTAL, sm.scal[0]);
        if (sm.scal[1]) atomicAdd(a.scalars + SC_AMBIG, sm.scal[1]);
        if (sm.scal[2]) atomicAdd(a.scalars + SC_LOOKUPS, sm.scal[2]);
        if (sm.scal[3]) atomicAdd(a.scalars + SC_READS, sm.scal[3]);
    }
    for (uint32_t g = threadIdx.x; g < a.V; g += blockDim.x)
    {
        if (LOCAL) { if (sm.w[g] != 0.0) atomicAdd(a.unique_w + g, sm.w[g]); }
        else if (sm.cnt[g]) atomicAdd(a.unique + g, (unsigned long long)sm.cnt[g]);
        if (a.unique_n && sm.cnt[g]) atomicAdd(a.unique_n + g, (unsigned long long)sm.cnt[g]);
    }
}
}   // namespace
}   // namespace speqb
