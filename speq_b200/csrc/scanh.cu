// scanh.cu -- the k-mer table scan (SPEQG_ALGO_KTABLE): the FM search memoised per (index, k).
//
// What the reference computes per window is a function of the window's k-mer alone: the SA interval of the k-mer in the
// index (fm_index_cursor.hpp:110-146, k steps) and, after locate(), whether every occurrence lies in one variant and
// which (src/fm_scanner.cpp:166-177).  For k <= 31 that function is materialised once per scan handle:
//
//   ktable_build_kernel   walks the index's own texts (both strands of every scaffold, already resident as a packed batch)
//                         and inserts every N-free k-mer: value = the variant of the text, or kValMulti once texts of two
//                         variants hold it.
//   ktable_fold_mult_kernel (record_multiplicity only) resolves every k-mer that was hit to the start of its SA interval with k
//                         rank steps on the blocked BWT, so multiplicities recorded per slot reach the refinement kernel
//                         as mult[lb] exactly as the rolling kernel records them.
//   scanh_kernel          one lane per read (pair): rolls the 2-bit k-mer of the window (one shift per base) and probes
//                         ONE 32-byte bucket per valid window -- one sector per window instead of two per base and no
//                         instruction spent on interval arithmetic.  A k-mer is absent from the index iff the probe
//                         reaches a bucket with a free slot, which is "the backward search failed" of the reference.
//
// Table layout (KTabView, kernels.h): 2^q buckets of four 8-byte slots = one 32-byte sector per probe.  The key is hashed
// with an invertible multiplication mod 4^k; the top q bits choose the home bucket and only the remaining rbits are
// stored (quotienting), next to the displacement d <= 6 from the home bucket and the 16-bit value:
//     slot = value << 48 | d << rbits | remainder,   all ones = free.
// Insertion takes the first free slot from the home bucket on, so a lookup stops at the first bucket whose last slot is
// free.  Units are claimed in chunks of kClaimChunk per warp (one global atomic per chunk); tallies go to the shared-memory
// counters of scan_dev.cuh, per-read scalars through per-lane registers.
#include <algorithm>

#include <cub/device/device_scan.cuh>

#include "kernels.h"
#include "scan_dev.cuh"

namespace speqb
{
namespace
{
constexpr int kBlockH = 256;
constexpr uint32_t kFullH = 0xFFFFFFFFu;
constexpr uint32_t kClaimChunk = 64;
#if defined(__CUDACC__)
// units a warp claims at a time: kClaimChunk, or for batches of fewer than kClaimChunk units per warp their even share (lanes
// at and above a share below 32 stay idle, so at most `chunk` lanes claim at once)
__device__ __forceinline__ uint32_t claim_chunk(uint32_t units)
{
    const uint32_t nwarps = gridDim.x * (blockDim.x >> 5);
    const uint32_t per_warp = (units + nwarps - 1u) / nwarps;
    return per_warp >= kClaimChunk ? kClaimChunk : (per_warp ? per_warp : 1u);
}
#endif
constexpr uint64_t kSlotFree = ~0ull;
constexpr uint32_t kValMulti = 0xFFFFu;            // value of a k-mer held by texts of several variants; a variant is < 0x8000
constexpr uint32_t kMaxDisp = 6;
constexpr uint32_t kNoRun = 0xFFFFFFFFu;
// bit 63 of a slot = "several variants".  In a table with positions (KTabView::pshift != 0, record_multiplicity) such a slot
// holds, above the displacement, the smallest position in the packed texts where its k-mer starts (ktable_canon_kernel):
// consecutive windows of a read that hit a conserved stretch then have consecutive positions, and the scan records a RUN of
// hits with two atomics on a difference array (ScanArgs::diff) instead of one atomic per window.
__device__ __forceinline__ bool slot_multi(uint32_t hi) { return (hi >> 31) != 0u; }
// Addressing.  Narrow table (wide = 0): bucket b = slots[4b, 4b + 4), one 32-byte sector; the record-mode counters live in a
// side array.  Wide table (wide = 1, tables beyond L2 with record_multiplicity): bucket b = a 64-byte record, slots in its first
// sector and the four counters of those slots in its second -- the counter atomic of a hit then goes to the line (and page)
// the probe has just brought in, instead of paying a second address translation and DRAM access somewhere else (section 8).
__device__ __forceinline__ const uint64_t * kt_bucket(KTabView const & kt, uint32_t bk) { return kt.slots + ((uint64_t)bk << (2u + kt.wide)); }
__device__ __forceinline__ uint64_t * kt_slot(KTabView const & kt, uint64_t i) { return kt.slots + ((i >> 2) << (2u + kt.wide)) + (i & 3ull); }
__device__ __forceinline__ unsigned long long * kt_counter(KTabView const & kt, unsigned long long * side, uint64_t i)
{
    return kt.wide ? reinterpret_cast<unsigned long long *>(kt_slot(kt, i) + 4) : side + i;
}

// ---------------------------------------------------------------------------------------------------------------
// build: one thread per 32-base group of a text; the 32 windows that start in the group
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ bool ktable_insert(KTabView const & kt, uint64_t key, uint32_t grp, unsigned long long * n_keys)
{
    const uint64_t h = (key * kt.mul) & kt.mask_b;
    const uint32_t home = (uint32_t)(h >> kt.rbits);
    const uint64_t r = h & kt.rmask;
    const uint64_t kmask = (kt.rmask << 3) | 7ull;        // remainder + displacement
    for (uint32_t d = 0; d <= kMaxDisp; ++d)
    {
        const uint32_t bk = (home + d) & kt.nbmask;
        const uint64_t want = r | ((uint64_t)d << kt.rbits);
        for (int j = 0; j < 4; ++j)
        {
            unsigned long long * p = reinterpret_cast<unsigned long long *>(kt_slot(kt, 4ull * bk + j));
            unsigned long long old = *reinterpret_cast<volatile unsigned long long *>(p);
            while (true)
            {
                if (old == kSlotFree)
                {
                    const unsigned long long prev = atomicCAS(p, kSlotFree, want | ((unsigned long long)grp << 48));
                    if (prev == kSlotFree)
                    {
                        atomicAdd(n_keys, 1ull);
                        return true;
                    }
                    old = prev;
                }
                if ((old & kmask) != want) break;              // another k-mer lives here: next slot
                const uint32_t v = (uint32_t)(old >> 48);
                if (v == grp || v == kValMulti) return true;
                const unsigned long long prev = atomicCAS(p, old, old | ((unsigned long long)kValMulti << 48));
                if (prev == old) return true;
                old = prev;
            }
        }
    }
    return false;
}

__global__ void __launch_bounds__(kBlockH) ktable_build_kernel(const uint64_t * __restrict__ bases, const uint32_t * __restrict__ inv,
                                                               const uint64_t * __restrict__ grp_off, const uint32_t * __restrict__ len,
                                                               const int32_t * __restrict__ text_group, const KTabView kt, const uint32_t k,
                                                               unsigned long long * __restrict__ n_keys)
{
    const uint32_t text = blockIdx.y;
    const uint32_t L = len[text];
    if (L < k) return;
    const uint32_t nwin = L - k + 1u;
    const uint64_t g0 = grp_off[text];
    const uint32_t grp = (uint32_t)text_group[text];
    for (uint32_t gi = blockIdx.x * blockDim.x + threadIdx.x; (uint64_t)gi * 32u < nwin; gi += gridDim.x * blockDim.x)
    {
        // bases [32 gi, 32 gi + 64): every window that starts in this group (k <= 31).  The k-mer code is rolled exactly as
        // the scan rolls it: key = key << 2 | base, i.e. the first base of the window in the highest of its k digits.
        const uint32_t ngroups = (L + 31u) >> 5;
        const uint64_t b0 = bases[g0 + gi], b1 = gi + 1u < ngroups ? bases[g0 + gi + 1u] : 0ull;
        const uint64_t v = (uint64_t)inv[g0 + gi] | ((gi + 1u < ngroups ? (uint64_t)inv[g0 + gi + 1u] : 0ull) << 32);
        uint64_t key = 0ull;
        uint32_t run = 0u;
        for (uint32_t j = 0; j < 31u + k; ++j)
        {
            const uint32_t pos = gi * 32u + j;
            if (pos >= L) break;
            const uint32_t c = (uint32_t)((j < 32u ? b0 >> (2u * j) : b1 >> (2u * (j - 32u))) & 3ull);
            key = (key << 2) | c;
            run = ((v >> j) & 1ull) ? 0u : run + 1u;              // an N: no ACGT query matches a window that holds it
            if (j + 1u >= k && run >= k && !ktable_insert(kt, key & kt.mask_b, grp, n_keys))
                atomicExch(n_keys + 1, 1ull);                     // displacement > kMaxDisp: the host regrows
        }
    }
}

// Multiplicities are recorded per table slot.  Before the refinement reads mult[] they are moved to the start of each
// k-mer's SA interval -- what the rolling scan records: the hash is inverted to get the k-mer back, k rank steps from the root
// give the interval (only slots that were hit pay for the search), and the slot counter is cleared.
__global__ void __launch_bounds__(kBlockH) ktable_fold_mult_kernel(const KTabView kt, const DevIndex ix, const uint32_t k,
                                                                   unsigned long long * __restrict__ slot_mult, unsigned long long * __restrict__ mult)
{
    const uint64_t n_slots = 4ull * ((uint64_t)kt.nbmask + 1ull);
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_slots; i += (uint64_t)gridDim.x * blockDim.x)
    {
        unsigned long long * cnt = kt_counter(kt, slot_mult, i);
        const unsigned long long m = *cnt;
        if (m == 0ull) continue;
        *cnt = 0ull;
        const uint64_t w = *kt_slot(kt, i);
        const uint32_t d = (uint32_t)(w >> kt.rbits) & 7u;
        const uint32_t home = ((uint32_t)(i >> 2) - d) & kt.nbmask;
        const uint64_t h = ((uint64_t)home << kt.rbits) | (w & kt.rmask);
        const uint64_t key = (h * kt.mul_inv) & kt.mask_b;          // k bases, the first in the highest digit
        uint32_t lb = 0u, rb = ix.n - 1u;
        bool ok = true;
        for (uint32_t j = 0; j < k && ok; ++j) ok = fm_step(ix, ix.ctab, lb, rb, (uint32_t)(key >> (2u * (k - 1u - j))) & 3u);
        if (ok) atomicAdd(mult + lb, m);
    }
}

// where `key` is stored: slot index, or ~0 when it is not in the table
__device__ __forceinline__ uint64_t ktable_slot_of(KTabView const & kt, uint64_t key)
{
    const uint64_t h = (key * kt.mul) & kt.mask_b;
    uint32_t bk = (uint32_t)(h >> kt.rbits);
    uint64_t want = h & kt.rmask;
    const uint64_t kmask = (kt.rmask << 3) | 7ull;
    for (uint32_t d = 0; d <= kMaxDisp; ++d)
    {
        bool last_free = false;
        for (int j = 0; j < 4; ++j)
        {
            const uint64_t w = *reinterpret_cast<const volatile uint64_t *>(kt_slot(kt, 4ull * bk + j));
            if (((w ^ want) & kmask) == 0ull) return 4ull * bk + j;
            if (j == 3) last_free = w == kSlotFree;
        }
        if (last_free) break;
        bk = (bk + 1u) & kt.nbmask;
        want += 1ull << kt.rbits;
    }
    return ~0ull;
}

// Second pass of the build for a table with positions: every N-free window of the texts whose k-mer is held by several
// variants lowers the slot's position field to its own start in the packed texts (32 * group + offset).  The first pass left
// the field of such slots at a value above every position (variant field 0x7FFF, rest 0), so atomicMin on the whole word works.
//
// Walk form (tflags != NULL; the slots were prepared by ktable_prep_pos_kernel): EVERY window lowers its slot's position, and the
// thread writes what the walking scan reads per 32-base group of the texts: which windows that end in the group are k-mers of
// one text, which of those are held by several variants, and the variant of the text.
template <bool WALK>
__global__ void __launch_bounds__(kBlockH) ktable_canon_kernel(const uint64_t * __restrict__ bases, const uint32_t * __restrict__ inv,
                                                               const uint64_t * __restrict__ grp_off, const uint32_t * __restrict__ len,
                                                               const int32_t * __restrict__ text_group, const KTabView kt, const uint32_t k,
                                                               uint64_t * __restrict__ tflags, uint16_t * __restrict__ gvar)
{
    const uint32_t text = blockIdx.y;
    const uint32_t L = len[text];
    if (L < k) return;
    const uint32_t nwin = L - k + 1u;
    const uint64_t g0 = grp_off[text];
    const uint64_t low_mask = (1ull << kt.pshift) - 1ull;
    for (uint32_t gi = blockIdx.x * blockDim.x + threadIdx.x; (uint64_t)gi * 32u < nwin + k - 1u; gi += gridDim.x * blockDim.x)
    {
        // windows that END in bases [32 gi, 32 gi + 32): the k-1 bases before the group are rolled in first
        const uint32_t first = gi * 32u >= k - 1u ? gi * 32u - (k - 1u) : 0u;
        uint64_t key = 0ull;
        uint32_t run = 0u, vbits = 0u, mbits = 0u;
        for (uint32_t pos = first; pos < L && pos < gi * 32u + 32u; ++pos)
        {
            const uint64_t bw = bases[g0 + (pos >> 5)];
            const uint32_t iw = inv[g0 + (pos >> 5)];
            key = (key << 2) | ((bw >> (2u * (pos & 31u))) & 3ull);
            run = ((iw >> (pos & 31u)) & 1u) ? 0u : run + 1u;
            if (pos < gi * 32u || run < k) continue;
            const uint64_t si = ktable_slot_of(kt, key & kt.mask_b);
            if (si == ~0ull) continue;
            const uint64_t w = *kt_slot(kt, si);
            if (WALK)
            {
                vbits |= 1u << (pos & 31u);
                if (w >> 63) mbits |= 1u << (pos & 31u);
            }
            else if (!(w >> 63)) continue;
            const uint64_t start = 32ull * g0 + (pos + 1u - k);
            if (WALK && kt.pshift == 0u) atomicMin(kt.pos_side + si, (uint32_t)start);
            else atomicMin(reinterpret_cast<unsigned long long *>(kt_slot(kt, si)), (w & (1ull << 63)) | (start << kt.pshift) | (w & low_mask));
        }
        if (WALK)
        {
            tflags[g0 + gi] = (uint64_t)vbits | ((uint64_t)mbits << 32);
            gvar[g0 + gi] = (uint16_t)text_group[text];
        }
    }
}

// walk form, between the build and the canon pass: the variant leaves every stored slot -- bit 63 = held by several variants,
// the position field = all ones (above every position), remainder and displacement stay
__global__ void __launch_bounds__(kBlockH) ktable_prep_pos_kernel(const KTabView kt)
{
    const uint64_t n_slots = 4ull * ((uint64_t)kt.nbmask + 1ull);
    const uint64_t low_mask = (1ull << kt.pshift) - 1ull;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_slots; i += (uint64_t)gridDim.x * blockDim.x)
    {
        const uint64_t w = *kt_slot(kt, i);
        if (w == kSlotFree) continue;
        const uint64_t multi = (w >> 48) == kValMulti ? 1ull << 63 : 0ull;
        const uint64_t var = kt.vshift && !multi ? (w >> 48) << kt.vshift : 0ull;
        *kt_slot(kt, i) = multi | (((1ull << 63) - 1ull) & ~low_mask) | (w & low_mask) | var;
    }
}

// Walk form of the fold, after the prefix sum: the walking scan records whole stretches of text positions, whatever their k-mers.
// A position whose k-mer belongs to one variant drops its count (the refinement takes those from the per-variant counts); one
// whose k-mer has its smallest position elsewhere moves its count there, so that kpos_fold_kernel searches every k-mer once.
__global__ void __launch_bounds__(kBlockH) kpos_canon_counts_kernel(const KTabView kt, const uint32_t k, unsigned long long * __restrict__ cnt,
                                                                    const uint64_t npos)
{
    for (uint64_t p = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; p < npos; p += (uint64_t)gridDim.x * blockDim.x)
    {
        const unsigned long long m = cnt[p];
        if (m == 0ull) continue;
        const uint64_t e = p + k - 1u;
        const uint64_t fl = e < 32ull * kt.tgroups ? kt.tflags[e >> 5] : 0ull;
        if (!((fl >> (32u + (uint32_t)(e & 31ull))) & 1ull))
        {
            cnt[p] = 0ull;
            continue;
        }
        uint64_t key = 0ull;
        for (uint32_t j = 0; j < k; ++j)
        {
            const uint64_t q = p + j;
            key = (key << 2) | ((kt.tbases[q >> 5] >> (2u * (uint32_t)(q & 31ull))) & 3ull);
        }
        const uint64_t si = ktable_slot_of(kt, key & kt.mask_b);
        if (si == ~0ull) continue;
        const uint64_t c = kt.pshift ? (*kt_slot(kt, si) & ~(1ull << 63)) >> kt.pshift : (uint64_t)kt.pos_side[si];
        if (c != p)
        {
            atomicAdd(cnt + c, m);
            cnt[p] = 0ull;
        }
    }
}

// Recorded runs -> multiplicities.  cnt[p] (the inclusive prefix sum of the difference array) = valid windows whose k-mer has
// its smallest text position at p.  The k-mer is read back from the packed texts, k rank steps from the root give the start of
// its SA interval, and the count moves to mult[lb] -- what the rolling scan records.  Only positions that were hit pay.
__global__ void __launch_bounds__(kBlockH) kpos_fold_kernel(const uint64_t * __restrict__ t_bases, const DevIndex ix, const uint32_t k,
                                                            unsigned long long * __restrict__ cnt, const uint64_t npos,
                                                            unsigned long long * __restrict__ mult)
{
    for (uint64_t p = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; p < npos; p += (uint64_t)gridDim.x * blockDim.x)
    {
        const unsigned long long m = cnt[p];
        cnt[p] = 0ull;
        if (m == 0ull) continue;
        uint32_t lb = 0u, rb = ix.n - 1u;
        bool ok = true;
        for (uint32_t j = 0; j < k && ok; ++j)
        {
            const uint64_t q = p + j;
            ok = fm_step(ix, ix.ctab, lb, rb, (uint32_t)(t_bases[q >> 5] >> (2u * (q & 31u))) & 3u);
        }
        if (ok) atomicAdd(mult + lb, m);
    }
}

// presence filter of a built table: the hash of every stored k-mer is rebuilt from its slot (home bucket + remainder)
__global__ void __launch_bounds__(kBlockH) kfilter_build_kernel(const KTabView kt)
{
    const uint64_t n_slots = 4ull * ((uint64_t)kt.nbmask + 1ull);
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_slots; i += (uint64_t)gridDim.x * blockDim.x)
    {
        const uint64_t w = *kt_slot(kt, i);
        if (w == ~0ull) continue;
        const uint32_t d = (uint32_t)(w >> kt.rbits) & 7u;
        const uint32_t home = ((uint32_t)(i >> 2) - d) & kt.nbmask;
        const uint64_t h = ((uint64_t)home << kt.rbits) | (w & kt.rmask);
        const uint32_t fi = (uint32_t)(h >> kt.fshift);
        atomicOr(kt.filt + (fi >> 5), 1u << (fi & 31u));
    }
}

// ---------------------------------------------------------------------------------------------------------------
// scan
// ---------------------------------------------------------------------------------------------------------------
enum : uint32_t { H_FETCH = 0, H_FIRST = 1, H_RUN = 2, H_DONE = 3, H_WAIT = 4 };
constexpr uint32_t kClaimChunkW = 4;   // reads a warp claims at a time when it works on one read together
// windows per lane and loop trip (their probes are independent, so all loads are issued together) and resident CTAs per SM
#ifndef SPEQB_KWIN
#define SPEQB_KWIN 3
#endif
#ifndef SPEQB_HMINB
#define SPEQB_HMINB 4
#endif
// record-mode instantiations: 3 CTAs per SM (85 registers: no spills; C3 19.04 -> 18.70 ms per 25 M reads, C2 unchanged).
// SPEQB_RUN_DEFER = 1 parks a closed run in two more registers and writes it at one place per loop trip instead of where it
// closes: measured slower at 4 CTAs/SM (C2 7.96 vs 6.80 ms, spills) and equal at 3 (profiles/r2_sweep.txt) -- off.
#ifndef SPEQB_HMINB_REC
#define SPEQB_HMINB_REC 3
#endif
#ifndef SPEQB_RUN_DEFER
#define SPEQB_RUN_DEFER 0
#endif
#ifndef SPEQB_HMINB_LOCAL
#define SPEQB_HMINB_LOCAL 2
#endif
#ifndef SPEQB_EXP_FREE_INPUT
#define SPEQB_EXP_FREE_INPUT 0
#endif
constexpr int kWin = SPEQB_KWIN;

// WPR (long reads): the 32 lanes of a warp share one read -- lane l takes windows [l * per, (l + 1) * per), per = ceil(nwin / 32),
// pre-rolling the k-1 bases before its first window; the ambiguity state of the read is the warp's min / max variant.
template <int MODE, bool WPR, bool MASK>
__global__ void __launch_bounds__(kBlockH, (MODE & M_LOCAL) ? SPEQB_HMINB_LOCAL : (MODE & M_RECORD) ? SPEQB_HMINB_REC : SPEQB_HMINB) scanh_kernel(const KTabView kt, unsigned long long * __restrict__ slot_mult,
                                                                                  const BatchView b, const ScanArgs a, const int paired)
{
    constexpr bool LOCAL = (MODE & M_LOCAL) != 0;
    constexpr bool RECORD = (MODE & M_RECORD) != 0;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    Smem32 sm = smem32_view(smem_raw, a.V);
    smem32_clear(sm, a.V);
    smem32_tables(sm, a, LOCAL);
    Tally32<MODE> env{b, a, sm};

    const uint32_t k = a.k;
    const uint64_t n_reads_b = b.n_reads_dev ? *b.n_reads_dev : b.n_reads;
    const uint32_t units = (uint32_t)(paired ? n_reads_b / 2 : n_reads_b);
    const uint32_t lane = threadIdx.x & 31u;
    const uint64_t kmask = (kt.rmask << 3) | 7ull;
    const uint32_t kmask_lo = (uint32_t)kmask, kmask_hi = (uint32_t)(kmask >> 32);

    // A small batch is spread over all warps -- ceil(units / warps) lanes of each take one unit -- instead of filling the first
    // warps' lanes twice: the kernel then lasts one read, not two, and every SM takes part.
    const uint32_t chunk = claim_chunk(units);

    // lane state
    uint32_t phase = (!WPR && lane >= chunk) ? H_DONE : H_FETCH;
    bool mate2 = false;
    uint32_t u = 0u;                             // unit (read or pair)
    uint64_t g0 = 0ull;                          // first 32-base group of the current read
    uint32_t t = 0u, pend = 0u, run = 0u;        // next base, end of the read, valid bases behind t
    uint64_t key = 0ull;                         // key << 2 | base per base: the low 2k bits are the window (first base highest);
                                                 // the bits above are stale bases -- the hash is taken mod 4^k, which ignores them
    uint64_t cb = 0ull, nb = 0ull;               // the 32-base group that holds base t, and the following one (prefetched)
    uint32_t ci = 0u, ni = 0u;                   // their invalid-base bits
    UnitSet<MASK> uset;                          // variants with a unique k-mer in the current unit
    uint32_t c_total = 0u;
    uint32_t agg_g = kNoGroup, agg_n = 0u;
    double agg_w = 0.0;
    QRoll qr;                                    // LOCAL: the rolling phred weight of the window that ends at the cursor
    // RECORD with positions: the open run [rs, rl] of consecutive text positions, and one closed run [cs, cl] waiting for the
    // single place per loop trip where runs are written (two atomics on the difference array) -- the windows of a trip only
    // compare and move registers, so the divergent atomic code exists once, not once per unrolled window
    uint32_t rs = kNoRun, rl = 0xFFFFFFFDu, cs = kNoRun, cl = 0u;
    auto run_write = [&](uint32_t s0, uint32_t s1) {
        atomicAdd(a.diff + s0, 1ull);
        atomicAdd(a.diff + s1 + 1u, ~0ull);                       // -1
    };
    // per-lane scalars (flushed at the end, or before they could wrap)
    uint32_t n_total = 0u, n_amb = 0u, n_look = 0u, n_reads = 0u, n_probe = 0u;
    // warp-uniform chunk of claimed units
    uint32_t wcur = 0u, wend = 0u;

    // t has reached a multiple of 32: the prefetched group becomes the current one and the one after it is requested
    auto next_group = [&]() {
        cb = nb; ci = ni;
        if (((t >> 5) + 1u) * 32u < pend)
        {
            nb = load_ro(b.bases + g0 + (t >> 5) + 1u);
            ni = load_ro(b.inv + g0 + (t >> 5) + 1u);
        }
    };
    // one base (start of a read / behind an N, where no window is due)
    auto roll1 = [&]() {
        const uint32_t o = t & 31u;
        key = (key << 2) | ((uint32_t)(cb >> (2u * o)) & 3u);
        run = ((ci >> o) & 1u) ? 0u : run + 1u;                   // N / low quality: no window may contain this base
        if (LOCAL) qr.step(t, k, run, sm.den, sm.rden);
        t += 1u;
        if ((t & 31u) == 0u) next_group();
    };

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
            if (WPR)
            {
                if (mc)
                {
                    if (wcur == wend)
                    {
                        uint32_t nbase = 0u;
                        if (lane == 0u) nbase = (uint32_t)atomicAdd(a.work, (unsigned long long)kClaimChunkW);
                        wcur = __shfl_sync(kFullH, nbase, 0);
                        wend = wcur + kClaimChunkW;
                    }
                    u = wcur;
                    wcur += 1u;
                }
            }
            else if (mc)
            {
                const uint32_t cnt = (uint32_t)__popc(mc), left = wend - wcur;
                uint32_t nbase = 0u;
                if (cnt > left)
                {
                    if (lane == 0u) nbase = (uint32_t)atomicAdd(a.work, (unsigned long long)chunk);
                    nbase = __shfl_sync(kFullH, nbase, 0);
                }
                const uint32_t r = (uint32_t)__popc(mc & ((1u << lane) - 1u));
                if (claim) u = r < left ? wcur + r : nbase + (r - left);
                if (cnt > left) { wcur = nbase + (cnt - left); wend = nbase + chunk; }
                else wcur += cnt;
            }
            if (ph == H_FETCH)
            {
                if (u >= units) phase = H_DONE;
                else
                {
                    // issued now, consumed by H_FIRST in the next iteration
#if SPEQB_EXP_FREE_INPUT
                    // measurement only (profiles/): every lane of a warp reads the SAME read, so the input loads of a warp are one
                    // broadcast wavefront each -- the upper bound of what staging the reads in shared memory could save; the
                    // probes stay spread because the hash below is salted per lane.  Results are meaningless in this build.
                    const uint64_t fr = (paired ? 2ull * u + (mate2 ? 1u : 0u) : (uint64_t)u) & ~31ull;
#else
                    const uint64_t fr = paired ? 2ull * u + (mate2 ? 1u : 0u) : (uint64_t)u;
#endif
                    pend = load_ro(b.len + fr);
                    g0 = load_ro(b.grp_off + fr);
                    phase = H_FIRST;
                }
            }
        }
        if (ph == H_FIRST)
        {
            const uint32_t L = pend;
            run = 0u; key = 0ull;
            if (WPR)
            {
                const uint32_t nwin = L >= k ? L - k + 1u : 0u, per = (nwin + 31u) >> 5;
                const uint32_t s0 = lane * per < nwin ? lane * per : nwin, s1 = nwin - s0 < per ? nwin : s0 + per;
                if (lane == 0u) { n_look += nwin; n_reads += 1u; }
                t = s0;
                pend = s1 > s0 ? s1 + k - 1u : s0;                // this lane's bases: its windows and the k-1 bases before
            }
            else
            {
                pend = L >= k ? L : 0u;                           // L < k: the sliding view is empty (sliding.hpp:129-142)
                n_look += L >= k ? L - k + 1u : 0u;
                n_reads += 1u;
                t = 0u;
            }
            if (t < pend)
            {
                cb = load_ro(b.bases + g0 + (t >> 5));
                ci = load_ro(b.inv + g0 + (t >> 5));
                if (((t >> 5) + 1u) * 32u < pend)
                {
                    nb = load_ro(b.bases + g0 + (t >> 5) + 1u);
                    ni = load_ro(b.inv + g0 + (t >> 5) + 1u);
                }
                if (LOCAL) qr.start(b.qual + 32ull * g0, L, t, k);
            }
            phase = H_RUN;
        }
        else if (ph == H_RUN)
        {
            // ---- up to kWin windows: roll to each, issue every probe, then consume them in window order
            while (run + 1u < k && t < pend) roll1();             // start of a read / behind an N: no window is due yet
            uint32_t w[kWin][8];
            uint32_t pb[kWin], want_lo[kWin], want_hi[kWin];
            bool due[kWin];
            double pw[kWin];                                      // LOCAL: weight of the window that ends with base j
            const uint32_t t0 = t, o = t & 31u;
            const uint32_t nbase = pend - t < (uint32_t)kWin ? pend - t : (uint32_t)kWin;
            // the next kWin bases in one register each (they may straddle into the prefetched group)
            uint32_t bits = (uint32_t)(cb >> (2u * o)), ibits = ci >> o;
            if (o + (uint32_t)kWin > 32u)
            {
                bits |= (uint32_t)nb << (2u * (32u - o));
                ibits |= ni << (32u - o);
            }
#pragma unroll
            for (int j = 0; j < kWin; ++j)
            {
                due[j] = false;
                if ((uint32_t)j < nbase)
                {
                    key = (key << 2) | ((bits >> (2 * j)) & 3u);
                    run = ((ibits >> j) & 1u) ? 0u : run + 1u;    // N / low quality: no window may contain this base
                    if (LOCAL) { qr.step(t0 + (uint32_t)j, k, run, sm.den, sm.rden); pw[j] = qr.P; }
                    due[j] = run >= k;
                }
                if (due[j])
                {
#if SPEQB_EXP_FREE_INPUT
                    const uint64_t h = ((key ^ ((uint64_t)lane * 0x9E3779B1ull)) * kt.mul) & kt.mask_b;
#else
                    const uint64_t h = (key * kt.mul) & kt.mask_b;
#endif
                    pb[j] = (uint32_t)(h >> kt.rbits);
                    const uint64_t want = h & kt.rmask;
                    want_lo[j] = (uint32_t)want; want_hi[j] = (uint32_t)(want >> 32);
                    asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                                 : "=r"(w[j][0]), "=r"(w[j][1]), "=r"(w[j][2]), "=r"(w[j][3]), "=r"(w[j][4]), "=r"(w[j][5]), "=r"(w[j][6]), "=r"(w[j][7])
                                 : "l"(kt.slots + 4ull * pb[j]));
                }
            }
            t += nbase;
            if (o + nbase >= 32u) next_group();
#pragma unroll
            for (int j = 0; j < kWin; ++j)
            {
                if (!due[j]) continue;
                c_total += 1u;                                    // ++total_kmers (src/fm_scanner.cpp:164)
                n_probe += 1u;
                uint32_t d = 0u;
                while (true)
                {
                    const bool h0 = (((w[j][0] ^ want_lo[j]) & kmask_lo) | ((w[j][1] ^ want_hi[j]) & kmask_hi)) == 0u;
                    const bool h1 = (((w[j][2] ^ want_lo[j]) & kmask_lo) | ((w[j][3] ^ want_hi[j]) & kmask_hi)) == 0u;
                    const bool h2 = (((w[j][4] ^ want_lo[j]) & kmask_lo) | ((w[j][5] ^ want_hi[j]) & kmask_hi)) == 0u;
                    const bool h3 = (((w[j][6] ^ want_lo[j]) & kmask_lo) | ((w[j][7] ^ want_hi[j]) & kmask_hi)) == 0u;
                    if (h0 || h1 || h2 || h3)
                    {
                        const uint32_t hi = h0 ? w[j][1] : h1 ? w[j][3] : h2 ? w[j][5] : w[j][7];
                        if (!slot_multi(hi))
                        {
                            // ++unique_kmers[which_hit] and the read-level ambiguity state (src/fm_scanner.cpp:178-191).  The
                            // run-length tally is closed at the end of every unit, so the variant of an open run is already
                            // in the unit's set.  Nothing is recorded for the refinement: a k-mer of one variant contributes
                            // a[g] / a[g] = 1 per window (src/fm_scanner.cpp:1099-1119), i.e. the variant's unique count.
                            uint32_t g = hi >> 16;
                            if (kt.vshift != 0u)
                            {
                                // table in walk form with room for the variant in the low word
                                const uint32_t lo = h0 ? w[j][0] : h1 ? w[j][2] : h2 ? w[j][4] : w[j][6];
                                g = (lo >> kt.vshift) & 0x7FFFu;
                            }
                            else if (kt.walk && kt.pshift != 0u)
                            {
                                // table in walk form: the slot holds the k-mer's position, the variant is that of its text
                                uint32_t p = hi & 0x7FFFFFFFu;
                                if (kt.pshift != 32u)
                                {
                                    const uint32_t lo = h0 ? w[j][0] : h1 ? w[j][2] : h2 ? w[j][4] : w[j][6];
                                    p = (uint32_t)((u64of(lo, hi) & ~(1ull << 63)) >> kt.pshift);
                                }
                                g = kt.gvar[p >> 5];
                            }
                            if (g != agg_g)
                            {
                                if (agg_g != kNoGroup) env.unique(agg_g, agg_n, agg_w);
                                agg_g = g; agg_n = 0u; agg_w = 0.0;
                                uset.add(g);
                            }
                            if (LOCAL) agg_w += pw[j];
                            agg_n += 1u;
                        }
                        else if (RECORD)
                        {
                            if (kt.pshift != 0u || kt.walk)
                            {
                                // the k-mer's smallest text position: extends the open run or starts a new one
                                // (pshift == 32 whenever remainder + displacement fit the low word: the position is the high
                                //  word without its top bit; pshift == 0 in walk form: the side array)
                                uint32_t p = hi & 0x7FFFFFFFu;
                                if (kt.pshift == 0u) p = kt.pos_side[4ull * pb[j] + (h0 ? 0u : h1 ? 1u : h2 ? 2u : 3u)];
                                else if (kt.pshift != 32u)
                                {
                                    const uint32_t lo = h0 ? w[j][0] : h1 ? w[j][2] : h2 ? w[j][4] : w[j][6];
                                    p = (uint32_t)((u64of(lo, hi) & ~(1ull << 63)) >> kt.pshift);
                                }
                                if (p == rl + 1u) rl = p;
                                else
                                {
#if SPEQB_RUN_DEFER
                                    if (cs != kNoRun) run_write(cs, cl);      // a second run closed within one trip (rare)
                                    cs = rs; cl = rl;
#else
                                    if (rs != kNoRun) run_write(rs, rl);
#endif
                                    rs = p; rl = p;
                                }
                            }
                            else atomicAdd(slot_mult + 4ull * pb[j] + (h0 ? 0u : h1 ? 1u : h2 ? 2u : 3u), 1ull);   // (this kernel never runs on a wide table)
                        }
                        break;
                    }
                    // a free last slot ends the probe sequence: the k-mer does not occur in the index
                    if ((w[j][6] & w[j][7]) == 0xFFFFFFFFu || d == kMaxDisp) break;
                    // rare: the home bucket is full of other k-mers; the k-mer may sit in the next bucket
                    d += 1u;
                    n_probe += 1u;
                    pb[j] = (pb[j] + 1u) & kt.nbmask;
                    const uint64_t want = u64of(want_lo[j], want_hi[j]) + (1ull << kt.rbits);
                    want_lo[j] = (uint32_t)want; want_hi[j] = (uint32_t)(want >> 32);
                    asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                                 : "=r"(w[j][0]), "=r"(w[j][1]), "=r"(w[j][2]), "=r"(w[j][3]), "=r"(w[j][4]), "=r"(w[j][5]), "=r"(w[j][6]), "=r"(w[j][7])
                                 : "l"(kt.slots + 4ull * pb[j]));
                }
            }
            if (RECORD && cs != kNoRun)
            {
                run_write(cs, cl);
                cs = kNoRun;
            }
        }
        if (WPR)
        {
            // the lane's share of the read is finished: wait for the warp, then merge the read's ambiguity state
            if (phase == H_RUN && t >= pend) phase = H_WAIT;
            if (__all_sync(kFullH, phase == H_WAIT))
            {
                if (paired && !mate2) mate2 = true;
                else
                {
                    uset.warp_merge();
                    n_total += c_total;
                    if (lane == 0u && uset.ambiguous()) { n_amb += 1u; env.fusion(uset.key()); }
                    c_total = 0u; uset.reset();
                    if (agg_g != kNoGroup) env.unique(agg_g, agg_n, agg_w);
                    agg_g = kNoGroup; agg_n = 0u; agg_w = 0.0;
                    mate2 = false;
                    if ((n_total | n_look) & 0x80000000u)
                    {
                        env.add64(0, n_total); env.add64(2, n_look);
                        n_total = 0u; n_look = 0u;
                    }
                }
                phase = H_FETCH;
            }
        }
        else if (phase == H_RUN && t >= pend)
        {
            // the read is finished
            if (paired && !mate2) mate2 = true;
            else
            {
                // is_ambiguous: unique k-mers of two different variants in one read / pair (src/fm_scanner.cpp:183-190,213)
                n_total += c_total;
                if (uset.ambiguous()) { n_amb += 1u; env.fusion(uset.key()); }
                c_total = 0u; uset.reset();
                if (agg_g != kNoGroup) env.unique(agg_g, agg_n, agg_w);
                agg_g = kNoGroup; agg_n = 0u; agg_w = 0.0;
                mate2 = false;
                if ((n_total | n_look) & 0x80000000u)
                {
                    env.add64(0, n_total); env.add64(2, n_look);
                    n_total = 0u; n_look = 0u;
                }
            }
            phase = H_FETCH;
        }
    }
    if (agg_g != kNoGroup) env.unique(agg_g, agg_n, agg_w);
    if (RECORD)
    {
        if (cs != kNoRun) run_write(cs, cl);
        if (rs != kNoRun) run_write(rs, rl);
    }
    if (n_total) env.add64(0, n_total);
    if (n_amb) env.add64(1, n_amb);
    if (n_look) env.add64(2, n_look);
    if (n_reads) env.add64(3, n_reads);
    {
        // sectors probed (table loads issued): one global atomic per warp
        const uint32_t wp = __reduce_add_sync(kFullH, n_probe);
        if (lane == 0u && wp) atomicAdd(a.scalars + SC_STEPS, (unsigned long long)wp);
    }
    smem32_flush<LOCAL>(sm, a);
}

// L2 eviction hints (pair-table scan): evict_first for the table probes (each sector is used once), evict_last for the filter
#ifndef SPEQB_L2HINT
#define SPEQB_L2HINT 1
#endif
__device__ __forceinline__ uint64_t l2_policy_keep()
{
    uint64_t p = 0ull;
#if SPEQB_L2HINT
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
#endif
    return p;
}
__device__ __forceinline__ uint64_t l2_policy_stream()
{
    uint64_t p = 0ull;
#if SPEQB_L2HINT
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
#endif
    return p;
}
__device__ __forceinline__ void ldg256_stream(uint32_t (&w)[8], const void * ptr, uint64_t pol)
{
#if SPEQB_L2HINT
    asm volatile("ld.global.nc.L2::cache_hint.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8], %9;"
                 : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]), "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7])
                 : "l"(ptr), "l"(pol));
#else
    asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]), "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7])
                 : "l"(ptr));
#endif
}
// record-mode probes of the pair-table scan ask L2 to fetch 64 bytes: the bucket's counter sector rides along with its slots,
// so the counter atomic that follows finds its sector in L2 (C4, 20 M reads: 56.0 -> 52.1 ms, profiles/r2_sweep.txt)
#ifndef SPEQB_WIDE_PF
#define SPEQB_WIDE_PF 1
#endif
__device__ __forceinline__ void ldg256_stream_pf64(uint32_t (&w)[8], const void * ptr, uint64_t pol)
{
    asm volatile("ld.global.nc.L2::cache_hint.L2::64B.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8], %9;"
                 : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]), "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7])
                 : "l"(ptr), "l"(pol));
}
__device__ __forceinline__ uint32_t ldg32_keep(const uint32_t * ptr, uint64_t pol)
{
    uint32_t v;
#if SPEQB_L2HINT
    asm volatile("ld.global.nc.L2::cache_hint.u32 %0, [%1], %2;" : "=r"(v) : "l"(ptr), "l"(pol));
#else
    asm volatile("ld.global.nc.u32 %0, [%1];" : "=r"(v) : "l"(ptr));
#endif
    return v;
}

// ---------------------------------------------------------------------------------------------------------------
// The walking scan (global mode, one lane per read; table in walk form).  Two facts make most probes unnecessary:
//   * a window of a read that was found at text position p, followed by a read base that equals the text's next base, IS the
//     k-mer at p + 1 -- whose answer (a k-mer of one text? held by several variants?) was computed when the table was built
//     and lies in two bits of KTabView::tflags; its variant, when it has one, is that of the text;
//   * reads follow some text for most of their length: a sequencing error or a variant site ends a stretch, k windows later
//     the next one begins.
// So a lane alternates between two steps.  PROBE: exactly scanh_kernel's trip (kWin windows, one bucket each); when the last
// window of the trip was found, its slot gives a text position and the lane starts to WALK: up to 32 read bases are compared
// with the text in one XOR, the flag words say how many of the windows count for the variant of the text and how many are held
// by several variants (two popcounts), and the whole stretch is recorded as one run of text positions (the fold drops the
// positions of single-variant k-mers and moves the others to their k-mer's smallest position, kpos_canon_counts_kernel).
// Tallies are the same integers as scanh_kernel's: the answer per window is the same function of the window's k-mer.
// ---------------------------------------------------------------------------------------------------------------
#ifndef SPEQB_WMINB
#define SPEQB_WMINB 3
#endif
// -DSPEQB_BOUNDS=1: a checking build (compute-sanitizer is not available on the pool) -- every index the walking scan derives
// from table contents is checked before use; a violation is counted in scalars[SC_OCC] and speqg_scan_finish fails on it
#ifndef SPEQB_BOUNDS
#define SPEQB_BOUNDS 0
#endif
#if SPEQB_BOUNDS
#define SPEQB_CHECK(cond) do { if (!(cond)) atomicAdd(a.scalars + SC_OCC, 1ull); } while (0)
#else
#define SPEQB_CHECK(cond) do { } while (0)
#endif
#ifndef SPEQB_WKWIN
#define SPEQB_WKWIN 3
#endif
// 1: the table lookups of a probe trip are spread over all 32 lanes through a queue in shared memory (needs SPEQB_WKWIN <= 3).
// Measured SLOWER and therefore off (profiles/r2_sweep_walk_coop.txt: C2 3.53 -> 3.80 ms per 10 M reads, C3 11.28 -> 12.60 ms
// per 25 M, C4 13.78 -> 16.42 ms per 20 M): every lane works in a probe trip then (75 registers, no bucket arrays), but the
// queue's two warp syncs, the shared-memory round trip and one lookup in flight per lane instead of three cost more than the
// idle lanes did.
#ifndef SPEQB_WCOOP
#define SPEQB_WCOOP 0
#endif
constexpr size_t kWalkQueueBytes = SPEQB_WCOOP ? (size_t)(kBlockH / 32) * 32u * SPEQB_WKWIN * 8u : 0u;   // per CTA
__host__ __device__ inline size_t scanw_queue_offset(uint32_t V) { return (32u + (size_t)V * 12u + 2u * 42u * 8u + 7u) & ~(size_t)7u; }
// which step a warp runs next: the one with the largest (lanes waiting for it) x weight
// (3 / 2 / 1 against 1 / 1 / 1: C2 3.62 -> 3.52 ms per 10 M reads, C3 11.64 -> 11.26 ms per 25 M, C4 14.31 -> 13.74 ms per 20 M)
#ifndef SPEQB_WGT_F
#define SPEQB_WGT_F 3
#endif
#ifndef SPEQB_WGT_W
#define SPEQB_WGT_W 2
#endif
#ifndef SPEQB_WGT_P
#define SPEQB_WGT_P 1
#endif
// the 32 bases in front of the cursor, as a key (digit i = base t-1-i): 2-bit digits reversed
__device__ __forceinline__ uint64_t rev_digits(uint64_t s)
{
    const uint64_t r = __brevll(s);
    return ((r & 0x5555555555555555ull) << 1) | ((r >> 1) & 0x5555555555555555ull);
}

template <bool RECORD, bool MASK>
__global__ void __launch_bounds__(kBlockH, SPEQB_WMINB) scanw_kernel(const KTabView kt, const BatchView b, const ScanArgs a, const int paired)
{
    constexpr int MODE = RECORD ? M_RECORD : 0;
    constexpr int kW = SPEQB_WKWIN;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    Smem32 sm = smem32_view(smem_raw, a.V);
    smem32_clear(sm, a.V);
    smem32_tables(sm, a, false);
    Tally32<MODE> env{b, a, sm};
#if SPEQB_WCOOP
    static_assert(SPEQB_WKWIN <= 3, "the queue offsets are summed from two ballots");
    // the warp's queue of keys / answers: 32 lanes x kW entries
    uint64_t * const wq = reinterpret_cast<uint64_t *>(smem_raw + scanw_queue_offset(a.V)) + (threadIdx.x >> 5) * (32u * SPEQB_WKWIN);
#endif

    const uint32_t k = a.k;
    const uint64_t n_reads_b = b.n_reads_dev ? *b.n_reads_dev : b.n_reads;
    const uint32_t units = (uint32_t)(paired ? n_reads_b / 2 : n_reads_b);
    const uint32_t lane = threadIdx.x & 31u;
    const uint64_t kmask = (kt.rmask << 3) | 7ull;
    const uint32_t kmask_lo = (uint32_t)kmask, kmask_hi = (uint32_t)(kmask >> 32);
    const uint32_t chunk = claim_chunk(units);

    // A lane is in one of three states: it needs a read (H_FETCH), it walks (H_RUN with a text position), it probes (H_RUN
    // without).  Lanes change state at different times, so each loop trip runs ONE of the three steps -- the one most lanes
    // wait for -- and the others wait a trip: every step then runs with most of the warp instead of a third of it.
    uint32_t phase = lane >= chunk ? H_DONE : H_FETCH;
    bool mate2 = false;
    uint32_t u = 0u;
    uint64_t g0 = 0ull;
    uint32_t t = 0u, pend = 0u, run = 0u;
    uint64_t key = 0ull;
    uint64_t cb = 0ull, nb = 0ull, pvb = 0ull;   // the 32-base group that holds base t, the following one, the one before
    uint32_t ci = 0u, ni = 0u, pvi = 0u;
    UnitSet<MASK> uset;
    uint32_t c_total = 0u;
    uint32_t agg_g = kNoGroup, agg_n = 0u;
    // the walk: text position of the LAST base of the window that ends at read base t - 1, when that window was found there
    uint32_t we = kNoRun, cur_g = 0u;
    uint32_t rs = kNoRun, rl = 0xFFFFFFFDu;
    auto run_write = [&](uint32_t s0, uint32_t s1) {
        SPEQB_CHECK(s0 <= s1 && (uint64_t)s1 + 1u < 32ull * kt.tgroups + 2ull);
        atomicAdd(a.diff + s0, 1ull);
        atomicAdd(a.diff + s1 + 1u, ~0ull);
    };
    uint32_t n_total = 0u, n_amb = 0u, n_look = 0u, n_reads = 0u, n_probe = 0u;
    uint32_t n_wwin = 0u, n_wstep = 0u;          // windows answered by the walk, and the steps (<= 32 windows each) that did it
    uint32_t wcur = 0u, wend = 0u;

    auto next_group = [&]() {
        pvb = cb; pvi = ci;
        cb = nb; ci = ni;
        if (((t >> 5) + 1u) * 32u < pend)
        {
            nb = load_ro(b.bases + g0 + (t >> 5) + 1u);
            ni = load_ro(b.inv + g0 + (t >> 5) + 1u);
        }
    };
    // key and run of the cursor from the 32 bases in front of it (the two groups that hold them), instead of rolling them in
    // one by one: lanes start their reads at different times here, and a per-base loop would run for one lane at a time
    auto seek = [&]() {
        const uint32_t o2 = t & 31u;
        key = rev_digits(o2 ? (pvb >> (2u * o2)) | (cb << (64u - 2u * o2)) : pvb);
        const uint32_t iv = __funnelshift_r(pvi, ci, o2);         // bit 31 = base t - 1
        const uint32_t z = (uint32_t)__clz((int)iv);
        run = z < t ? z : t;
    };
    // a unique hit of variant g (n windows)
    auto count_unique = [&](uint32_t g, uint32_t n) {
        if (g != agg_g)
        {
            if (agg_g != kNoGroup) env.unique(agg_g, agg_n, 0.0);
            agg_g = g; agg_n = 0u;
            uset.add(g);
        }
        agg_n += n;
    };

    while (true)
    {
        __syncwarp();
        const bool is_f = phase == H_FETCH, is_w = phase == H_RUN && we != kNoRun, is_p = phase == H_RUN && we == kNoRun;
        const uint32_t m_f = __ballot_sync(kFullH, is_f), m_w = __ballot_sync(kFullH, is_w), m_p = __ballot_sync(kFullH, is_p);
        if (!(m_f | m_w | m_p)) break;
        // (weights: a probe trip costs about three walk trips, so it pays to let lanes gather in front of it)
        const int c_f = __popc(m_f) * SPEQB_WGT_F, c_w = __popc(m_w) * SPEQB_WGT_W, c_p = __popc(m_p) * SPEQB_WGT_P;
        if (c_f >= c_w && c_f >= c_p)
        {
            // ---- next read: claim (first mate only), length + first group, the first groups of bases, the cursor at base k - 1
            const bool claim = is_f && !mate2;
            const uint32_t mc = __ballot_sync(kFullH, claim);
            if (mc)
            {
                const uint32_t cnt = (uint32_t)__popc(mc), left = wend - wcur;
                uint32_t nbase = 0u;
                if (cnt > left)
                {
                    if (lane == 0u) nbase = (uint32_t)atomicAdd(a.work, (unsigned long long)chunk);
                    nbase = __shfl_sync(kFullH, nbase, 0);
                }
                const uint32_t r = (uint32_t)__popc(mc & ((1u << lane) - 1u));
                if (claim) u = r < left ? wcur + r : nbase + (r - left);
                if (cnt > left) { wcur = nbase + (cnt - left); wend = nbase + chunk; }
                else wcur += cnt;
            }
            if (is_f)
            {
                if (u >= units) phase = H_DONE;
                else
                {
                    const uint64_t fr = paired ? 2ull * u + (mate2 ? 1u : 0u) : (uint64_t)u;
                    const uint32_t L = load_ro(b.len + fr);
                    g0 = load_ro(b.grp_off + fr);
                    run = 0u; key = 0ull; we = kNoRun; pvb = 0ull; pvi = 0u;
                    pend = L >= k ? L : 0u;                       // L < k: the sliding view is empty (sliding.hpp:129-142)
                    n_look += L >= k ? L - k + 1u : 0u;
                    n_reads += 1u;
                    t = 0u;
                    if (pend)
                    {
                        cb = load_ro(b.bases + g0);
                        ci = load_ro(b.inv + g0);
                        if (32u < pend)
                        {
                            nb = load_ro(b.bases + g0 + 1u);
                            ni = load_ro(b.inv + g0 + 1u);
                        }
                        t = k - 1u;                               // (k <= 31: inside the first group) the first window ends here
                        seek();
                    }
                    phase = H_RUN;
                }
            }
        }
        else if (c_w >= c_p)
        {
            if (is_w)
            {
                // ---- walk: the read bases from t on against the text bases from we + 1 on
                // (tflags ends with two zero words: behind the last text no window is valid, whatever the clamped base words hold)
                const uint32_t e1 = we + 1u, ot = e1 & 31u, gt = e1 >> 5, glast = kt.tgroups - 1u;
                const uint64_t T0 = load_ro(kt.tbases + (gt < glast ? gt : glast)), T1 = load_ro(kt.tbases + (gt + 1u < glast ? gt + 1u : glast));
                SPEQB_CHECK(gt + 1u < kt.tgroups + 2u && t < pend);
                const uint64_t F0 = load_ro(kt.tflags + gt), F1 = load_ro(kt.tflags + gt + 1u);
                const uint32_t o = t & 31u;
                const uint64_t TB = ot ? (T0 >> (2u * ot)) | (T1 << (64u - 2u * ot)) : T0;
                const uint64_t RB = o ? (cb >> (2u * o)) | (nb << (64u - 2u * o)) : cb;
                const uint32_t TV = __funnelshift_r((uint32_t)F0, (uint32_t)F1, ot);          // windows that end at e1 + j: k-mers of one text
                const uint32_t TM = __funnelshift_r((uint32_t)(F0 >> 32), (uint32_t)(F1 >> 32), ot);   // ... held by several variants
                const uint32_t RI = __funnelshift_r(ci, ni, o);                                 // read bases that are no ACGT
                uint64_t x = RB ^ TB;
                x = (x | (x >> 1)) & 0x5555555555555555ull;
                const uint32_t lim = pend - t < 32u ? pend - t : 32u;
                uint32_t m = x ? ((uint32_t)__ffsll((long long)x) - 1u) >> 1 : 32u;             // bases that agree
                const uint32_t stop = RI | ~TV;
                const uint32_t m2 = stop ? (uint32_t)__ffs((int)stop) - 1u : 32u;
                m = m < m2 ? m : m2;
                m = m < lim ? m : lim;
                if (m != 0u)
                {
                    const uint32_t mm = m == 32u ? 0xFFFFFFFFu : (1u << m) - 1u;
                    const uint32_t nm = (uint32_t)__popc(TM & mm), nu = m - nm;
                    c_total += m;                                 // every one of the m windows is a k-mer of the index
                    n_wwin += m;
                    n_wstep += 1u;
                    if (nu) count_unique(cur_g, nu);
                    if (RECORD && nm)
                    {
                        const uint32_t s0 = e1 - (k - 1u), s1 = s0 + m - 1u;   // where the m windows START in the texts
                        if (s0 == rl + 1u) rl = s1;
                        else
                        {
                            if (rs != kNoRun) run_write(rs, rl);
                            rs = s0; rl = s1;
                        }
                    }
                    we += m;
                    t += m;
                    if (o + m >= 32u) next_group();
                }
                if (m < lim)
                {
                    // the stretch ends in front of base t (a base that differs, an N, the end of the text): the table again.
                    // Key and run of the cursor are rebuilt from the two groups that hold the bases before t.
                    we = kNoRun;
                    seek();
                }
            }
        }
#if SPEQB_WCOOP
        else
        {
            // ---- probe, the whole warp working: the lanes that wait for a probe roll up to kW windows each and put the keys of
            //      the valid ones into the warp's queue in shared memory; then ALL 32 lanes look keys up (hash, bucket, four
            //      compares), whatever step they wait for themselves; then the probing lanes take their answers in window
            //      order up to the FIRST window that was found -- its slot gives a text position, the walk goes on from there.
            uint64_t keys[kW];
            bool due[kW];
            uint32_t nd = 0u, nbase = 0u, o = 0u;
            if (is_p)
            {
                // (behind an N: the next window that can be valid ends k - 1 - run bases further on)
                while (run + 1u < k && t < pend)
                {
                    const uint32_t o3 = t & 31u, skip = k - 1u - run < pend - t ? k - 1u - run : pend - t;
                    t += skip;
                    if (o3 + skip >= 32u) next_group();
                    seek();
                }
                o = t & 31u;
                nbase = pend - t < (uint32_t)kW ? pend - t : (uint32_t)kW;
                uint32_t bits = (uint32_t)(cb >> (2u * o)), ibits = ci >> o;
                if (o + (uint32_t)kW > 32u)
                {
                    bits |= (uint32_t)nb << (2u * (32u - o));
                    ibits |= ni << (32u - o);
                }
#pragma unroll
                for (int j = 0; j < kW; ++j)
                {
                    due[j] = false;
                    if ((uint32_t)j < nbase)
                    {
                        key = (key << 2) | ((bits >> (2 * j)) & 3u);
                        run = ((ibits >> j) & 1u) ? 0u : run + 1u;
                        due[j] = run >= k;
                    }
                    keys[j] = key;
                    nd += due[j] ? 1u : 0u;
                }
            }
            // where this lane's keys go: the sum of nd (<= 3) over the lanes below, from two ballots
            const uint32_t b0 = __ballot_sync(kFullH, (nd & 1u) != 0u), b1 = __ballot_sync(kFullH, (nd & 2u) != 0u);
            const uint32_t below = (1u << lane) - 1u;
            const uint32_t off = (uint32_t)__popc(b0 & below) + 2u * (uint32_t)__popc(b1 & below);
            const uint32_t Q = (uint32_t)__popc(b0) + 2u * (uint32_t)__popc(b1);
            if (nd)
            {
                uint32_t qi = off;
#pragma unroll
                for (int j = 0; j < kW; ++j)
                    if (due[j]) wq[qi++] = keys[j];
            }
            __syncwarp();
            for (uint32_t i = lane; i < Q; i += 32u)
            {
                const uint64_t h = (wq[i] * kt.mul) & kt.mask_b;
                uint32_t bk = (uint32_t)(h >> kt.rbits);
                uint64_t want = h & kt.rmask;
                uint64_t res = kSlotFree;                         // not in the index
                bool ask = true;
                if (kt.filt)
                {
                    // table beyond L2: the presence filter (L2-resident) answers most windows that hold a sequencing error
                    const uint32_t fi = (uint32_t)(h >> kt.fshift);
                    ask = ((ldg32_keep(kt.filt + (fi >> 5), l2_policy_keep()) >> (fi & 31u)) & 1u) != 0u;
                }
                if (ask)
                {
                    for (uint32_t d = 0u; d <= kMaxDisp; ++d)
                    {
                        uint32_t w[8];
                        if (kt.filt) ldg256_stream(w, kt.slots + 4ull * bk, l2_policy_stream());
                        else
                            asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                                         : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]), "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7])
                                         : "l"(kt.slots + 4ull * bk));
                        n_probe += 1u;
                        const uint32_t want_lo = (uint32_t)want, want_hi = (uint32_t)(want >> 32);
                        const bool h0 = (((w[0] ^ want_lo) & kmask_lo) | ((w[1] ^ want_hi) & kmask_hi)) == 0u;
                        const bool h1 = (((w[2] ^ want_lo) & kmask_lo) | ((w[3] ^ want_hi) & kmask_hi)) == 0u;
                        const bool h2 = (((w[4] ^ want_lo) & kmask_lo) | ((w[5] ^ want_hi) & kmask_hi)) == 0u;
                        const bool h3 = (((w[6] ^ want_lo) & kmask_lo) | ((w[7] ^ want_hi) & kmask_hi)) == 0u;
                        if (h0 || h1 || h2 || h3)
                        {
                            const uint32_t hi = h0 ? w[1] : h1 ? w[3] : h2 ? w[5] : w[7];
                            // positions in a side array: "several variants" + variant on top, the slot's index below
                            if (kt.pshift == 0u) res = ((uint64_t)(hi >> 16) << 48) | (4ull * bk + (h0 ? 0u : h1 ? 1u : h2 ? 2u : 3u));
                            else res = u64of(h0 ? w[0] : h1 ? w[2] : h2 ? w[4] : w[6], hi);
                            break;
                        }
                        // a free last slot ends the probe sequence: the k-mer does not occur in the index
                        if ((w[6] & w[7]) == 0xFFFFFFFFu) break;
                        // rare: the home bucket is full of other k-mers; the k-mer may sit in the next bucket
                        bk = (bk + 1u) & kt.nbmask;
                        want += 1ull << kt.rbits;
                    }
                }
                wq[i] = res;
            }
            __syncwarp();
            if (is_p)
            {
                int jh = -1;                                      // the first window of the trip that was found
                uint64_t rh = 0ull;
                uint32_t qi = off;
#pragma unroll
                for (int j = 0; j < kW; ++j)
                {
                    if (!due[j] || jh >= 0) continue;
                    c_total += 1u;                                // ++total_kmers (src/fm_scanner.cpp:164)
                    const uint64_t r = wq[qi++];
                    if (r != kSlotFree) { jh = j; rh = r; }
                }
                // the cursor moves behind the window that was found, or over the whole trip
                const uint32_t adv = jh >= 0 ? (uint32_t)jh + 1u : nbase;
                // (the windows behind a hit belong to the walk: key and run are rebuilt where it ends)
                t += adv;
                if (o + adv >= 32u) next_group();
                if (jh >= 0)
                {
                    const bool multi = (rh >> 63) != 0ull;
                    uint32_t p;                                   // where the k-mer starts in the texts (its smallest position)
                    if (kt.pshift == 0u) p = load_ro(kt.pos_side + (rh & 0x3FFFFFFFFull));
                    else if (kt.pshift == 32u) p = (uint32_t)(rh >> 32) & 0x7FFFFFFFu;
                    else p = (uint32_t)((rh & ~(1ull << 63)) >> kt.pshift);
                    we = p + k - 1u;
                    cur_g = kt.pshift == 0u && !multi ? (uint32_t)(rh >> 48) & 0x7FFFu : (uint32_t)load_ro(kt.gvar + (p >> 5));
                    if (!multi) count_unique(cur_g, 1u);          // ++unique_kmers[which_hit] (src/fm_scanner.cpp:178-191)
                    else if (RECORD)
                    {
                        if (p == rl + 1u) rl = p;
                        else
                        {
                            if (rs != kNoRun) run_write(rs, rl);
                            rs = p; rl = p;
                        }
                    }
                }
            }
            __syncwarp();                                         // the queue is reused by the next probe trip
        }
#else
        else if (is_p)
        {
            // ---- probe: up to kW windows, every probe issued, then consumed in window order up to the FIRST window that is
            //      found -- its slot gives a text position, the walk goes on from there and answers the windows behind it
            // (behind an N: the next window that can be valid ends k - 1 - run bases further on)
            while (run + 1u < k && t < pend)
            {
                const uint32_t o3 = t & 31u, skip = k - 1u - run < pend - t ? k - 1u - run : pend - t;
                t += skip;
                if (o3 + skip >= 32u) next_group();
                seek();
            }
            uint32_t w[kW][8];
            uint32_t pb[kW], want_lo[kW], want_hi[kW];
            bool due[kW], ask[kW];                                // a valid window; one whose bucket is read
            const uint32_t o = t & 31u;
            const uint32_t nbase = pend - t < (uint32_t)kW ? pend - t : (uint32_t)kW;
            uint32_t bits = (uint32_t)(cb >> (2u * o)), ibits = ci >> o;
            if (o + (uint32_t)kW > 32u)
            {
                bits |= (uint32_t)nb << (2u * (32u - o));
                ibits |= ni << (32u - o);
            }
#pragma unroll
            for (int j = 0; j < kW; ++j)
            {
                due[j] = false;
                if ((uint32_t)j < nbase)
                {
                    key = (key << 2) | ((bits >> (2 * j)) & 3u);
                    run = ((ibits >> j) & 1u) ? 0u : run + 1u;
                    due[j] = run >= k;
                }
                ask[j] = due[j];
                if (due[j])
                {
                    const uint64_t h = (key * kt.mul) & kt.mask_b;
                    pb[j] = (uint32_t)(h >> kt.rbits);
                    const uint64_t want = h & kt.rmask;
                    want_lo[j] = (uint32_t)want; want_hi[j] = (uint32_t)(want >> 32);
                    if (kt.filt)
                    {
                        // table beyond L2: the presence filter (L2-resident) answers most windows that hold a sequencing
                        // error; its word is parked in w[j][0] until the bucket is asked for
                        const uint32_t fi = (uint32_t)(h >> kt.fshift);
                        w[j][0] = ldg32_keep(kt.filt + (fi >> 5), l2_policy_keep());
                        w[j][1] = fi & 31u;
                    }
                    else
                        asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                                     : "=r"(w[j][0]), "=r"(w[j][1]), "=r"(w[j][2]), "=r"(w[j][3]), "=r"(w[j][4]), "=r"(w[j][5]), "=r"(w[j][6]), "=r"(w[j][7])
                                     : "l"(kt.slots + 4ull * pb[j]));
                }
            }
            if (kt.filt)
            {
#pragma unroll
                for (int j = 0; j < kW; ++j)
                {
                    if (!due[j]) continue;
                    ask[j] = ((w[j][0] >> w[j][1]) & 1u) != 0u;
                    if (ask[j]) ldg256_stream(w[j], kt.slots + 4ull * pb[j], l2_policy_stream());
                }
            }
            int jh = -1;                                          // the first window of the trip that was found
            uint32_t h_hi = 0u, h_lo = 0u;                        // its slot (h_lo: the low word, or with side positions the slot's index)
#pragma unroll
            for (int j = 0; j < kW; ++j)
            {
                if (!due[j] || jh >= 0) continue;
                c_total += 1u;                                    // ++total_kmers (src/fm_scanner.cpp:164)
                if (!ask[j]) continue;                            // not in the filter: not in the index
                n_probe += 1u;
                uint32_t d = 0u;
                while (true)
                {
                    const bool h0 = (((w[j][0] ^ want_lo[j]) & kmask_lo) | ((w[j][1] ^ want_hi[j]) & kmask_hi)) == 0u;
                    const bool h1 = (((w[j][2] ^ want_lo[j]) & kmask_lo) | ((w[j][3] ^ want_hi[j]) & kmask_hi)) == 0u;
                    const bool h2 = (((w[j][4] ^ want_lo[j]) & kmask_lo) | ((w[j][5] ^ want_hi[j]) & kmask_hi)) == 0u;
                    const bool h3 = (((w[j][6] ^ want_lo[j]) & kmask_lo) | ((w[j][7] ^ want_hi[j]) & kmask_hi)) == 0u;
                    if (h0 || h1 || h2 || h3)
                    {
                        jh = j;
                        h_hi = h0 ? w[j][1] : h1 ? w[j][3] : h2 ? w[j][5] : w[j][7];
                        h_lo = kt.pshift == 0u ? 4u * pb[j] + (h0 ? 0u : h1 ? 1u : h2 ? 2u : 3u) : h0 ? w[j][0] : h1 ? w[j][2] : h2 ? w[j][4] : w[j][6];
                        break;
                    }
                    // a free last slot ends the probe sequence: the k-mer does not occur in the index
                    if ((w[j][6] & w[j][7]) == 0xFFFFFFFFu || d == kMaxDisp) break;
                    // rare: the home bucket is full of other k-mers; the k-mer may sit in the next bucket
                    d += 1u;
                    n_probe += 1u;
                    pb[j] = (pb[j] + 1u) & kt.nbmask;
                    const uint64_t want = u64of(want_lo[j], want_hi[j]) + (1ull << kt.rbits);
                    want_lo[j] = (uint32_t)want; want_hi[j] = (uint32_t)(want >> 32);
                    asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                                 : "=r"(w[j][0]), "=r"(w[j][1]), "=r"(w[j][2]), "=r"(w[j][3]), "=r"(w[j][4]), "=r"(w[j][5]), "=r"(w[j][6]), "=r"(w[j][7])
                                 : "l"(kt.slots + 4ull * pb[j]));
                }
            }
            // the cursor moves behind the window that was found, or over the whole trip
            const uint32_t adv = jh >= 0 ? (uint32_t)jh + 1u : nbase;
            t += adv;
            if (o + adv >= 32u) next_group();
            if (jh >= 0)
            {
                // where the k-mer starts in the texts (its smallest position): in the slot, or in the side array
                uint32_t p = h_hi & 0x7FFFFFFFu;
                SPEQB_CHECK(kt.pshift != 0u || h_lo < 4ull * ((uint64_t)kt.nbmask + 1ull));
                if (kt.pshift == 0u) p = load_ro(kt.pos_side + h_lo);
                else if (kt.pshift != 32u) p = (uint32_t)((u64of(h_lo, h_hi) & ~(1ull << 63)) >> kt.pshift);
                SPEQB_CHECK((uint64_t)p + k <= 32ull * kt.tgroups);
                we = p + k - 1u;
                cur_g = kt.pshift == 0u && !slot_multi(h_hi) ? h_hi >> 16 : (uint32_t)load_ro(kt.gvar + (p >> 5));
                if (!slot_multi(h_hi)) count_unique(cur_g, 1u);   // ++unique_kmers[which_hit] (src/fm_scanner.cpp:178-191)
                else if (RECORD)
                {
                    if (p == rl + 1u) rl = p;
                    else
                    {
                        if (rs != kNoRun) run_write(rs, rl);
                        rs = p; rl = p;
                    }
                }
            }
        }
#endif
        if (phase == H_RUN && t >= pend)
        {
            // the read is finished
            if (paired && !mate2) mate2 = true;
            else
            {
                // is_ambiguous: unique k-mers of two different variants in one read / pair (src/fm_scanner.cpp:183-190,213)
                n_total += c_total;
                if (uset.ambiguous()) { n_amb += 1u; env.fusion(uset.key()); }
                c_total = 0u; uset.reset();
                if (agg_g != kNoGroup) env.unique(agg_g, agg_n, 0.0);
                agg_g = kNoGroup; agg_n = 0u;
                mate2 = false;
                if ((n_total | n_look) & 0x80000000u)
                {
                    env.add64(0, n_total); env.add64(2, n_look);
                    n_total = 0u; n_look = 0u;
                }
            }
            we = kNoRun;
            phase = H_FETCH;
        }
    }
    if (agg_g != kNoGroup) env.unique(agg_g, agg_n, 0.0);
    if (RECORD && rs != kNoRun) run_write(rs, rl);
    if (n_total) env.add64(0, n_total);
    if (n_amb) env.add64(1, n_amb);
    if (n_look) env.add64(2, n_look);
    if (n_reads) env.add64(3, n_reads);
    {
        // sectors probed (table loads issued): one global atomic per warp
        const uint32_t wp = __reduce_add_sync(kFullH, n_probe);
        if (lane == 0u && wp) atomicAdd(a.scalars + SC_STEPS, (unsigned long long)wp);
        // (per-lane counts stay far below 2^32 / 32 for any batch that fits a device: the warp sums do not wrap)
        const uint32_t ww = __reduce_add_sync(kFullH, n_wwin), ws = __reduce_add_sync(kFullH, n_wstep);
        if (lane == 0u && ww) atomicAdd(a.scalars + SC_HIT, (unsigned long long)ww);
        if (lane == 0u && ws) atomicAdd(a.scalars + SC_OCCSEC, (unsigned long long)ws);
    }
    smem32_flush<false>(sm, a);
}

// ---------------------------------------------------------------------------------------------------------------
// Two windows per probe (SPEQG_ALGO_KPAIR).  A second table holds every N-free (k+1)-mer of the texts with the values of
// its two k-mers (one byte each: variant < 255, 0xFF = several variants); a (k+1)-mer of the index contains two k-mers of
// the index, so a hit answers two consecutive windows with one sector.  A miss (a sequencing error somewhere in the k+1
// bases, or a junction the index does not have) says nothing about the two k-mers: they are then looked up one by one in
// the k-mer table, exactly as scanh_kernel does.  Windows that cannot be paired (odd one at the end of a read, next to an
// N) take the same single-window path.
// ---------------------------------------------------------------------------------------------------------------
constexpr uint32_t kAbsent = 0x10000u;

// value of `key` in a table (synchronous probe sequence), kAbsent when it is not stored; bucket / j = where it is stored
struct KFound
{
    uint32_t val, bucket, j, probes;
};
__device__ __forceinline__ KFound ktable_find(KTabView const & kt, uint64_t key)
{
    KFound f;
    f.val = kAbsent; f.bucket = 0u; f.j = 0u; f.probes = 0u;
    const uint64_t h = (key * kt.mul) & kt.mask_b;
    uint32_t bk = (uint32_t)(h >> kt.rbits);
    uint64_t want = h & kt.rmask;
    const uint64_t kmask = (kt.rmask << 3) | 7ull;
    for (uint32_t d = 0; d <= kMaxDisp; ++d)
    {
        uint32_t w[8];
        asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]), "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7])
                     : "l"(kt_bucket(kt, bk)));
        f.probes += 1u;
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (f.val == kAbsent && ((u64of(w[2 * j], w[2 * j + 1]) ^ want) & kmask) == 0ull)
            {
                f.val = w[2 * j + 1] >> 16; f.bucket = bk; f.j = (uint32_t)j;
            }
        if (f.val != kAbsent || (w[6] & w[7]) == 0xFFFFFFFFu) break;
        bk = (bk + 1u) & kt.nbmask;
        want += 1ull << kt.rbits;
    }
    return f;
}

// value byte of a k-mer for the pair table: variant (< 255) or 0xFF
__device__ __forceinline__ uint32_t pair_byte(uint32_t v) { return v == kValMulti ? 0xFFu : v; }

__global__ void __launch_bounds__(kBlockH) kpair_build_kernel(const uint64_t * __restrict__ bases, const uint32_t * __restrict__ inv,
                                                              const uint64_t * __restrict__ grp_off, const uint32_t * __restrict__ len,
                                                              const KTabView kt, const KTabView kp, const uint32_t k,
                                                              unsigned long long * __restrict__ n_keys)
{
    const uint32_t text = blockIdx.y;
    const uint32_t L = len[text];
    const uint32_t k2 = k + 1u;
    if (L < k2) return;
    const uint32_t nwin = L - k2 + 1u;
    const uint64_t g0 = grp_off[text];
    for (uint32_t gi = blockIdx.x * blockDim.x + threadIdx.x; (uint64_t)gi * 32u < nwin; gi += gridDim.x * blockDim.x)
    {
        const uint32_t ngroups = (L + 31u) >> 5;
        const uint64_t b0 = bases[g0 + gi], b1 = gi + 1u < ngroups ? bases[g0 + gi + 1u] : 0ull;
        const uint64_t v = (uint64_t)inv[g0 + gi] | ((gi + 1u < ngroups ? (uint64_t)inv[g0 + gi + 1u] : 0ull) << 32);
        uint64_t key = 0ull;
        uint32_t run = 0u;
        for (uint32_t j = 0; j < 31u + k2; ++j)
        {
            const uint32_t pos = gi * 32u + j;
            if (pos >= L) break;
            const uint32_t c = (uint32_t)((j < 32u ? b0 >> (2u * j) : b1 >> (2u * (j - 32u))) & 3ull);
            key = (key << 2) | c;
            run = ((v >> j) & 1ull) ? 0u : run + 1u;
            if (run < k2) continue;
            // both k-mers are in the k-mer table (it was built from the same windows)
            const uint32_t v1 = ktable_find(kt, (key >> 2) & kt.mask_b).val, v2 = ktable_find(kt, key & kt.mask_b).val;
            if (v1 == kAbsent || v2 == kAbsent) { atomicExch(n_keys + 1, 2ull); continue; }
            if (!ktable_insert(kp, key & kp.mask_b, pair_byte(v1) | (pair_byte(v2) << 8), n_keys)) atomicExch(n_keys + 1, 1ull);
        }
    }
}

// the same for the pair table: a hit counted both k-mers of the stored (k+1)-mer
__global__ void __launch_bounds__(kBlockH) kpair_fold_mult_kernel(const KTabView kp, const DevIndex ix, const uint32_t k,
                                                                  unsigned long long * __restrict__ slot_mult, unsigned long long * __restrict__ mult)
{
    const uint64_t n_slots = 4ull * ((uint64_t)kp.nbmask + 1ull);
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_slots; i += (uint64_t)gridDim.x * blockDim.x)
    {
        unsigned long long * cnt = kt_counter(kp, slot_mult, i);
        const unsigned long long m = *cnt;
        if (m == 0ull) continue;
        *cnt = 0ull;
        const uint64_t w = *kt_slot(kp, i);
        const uint32_t d = (uint32_t)(w >> kp.rbits) & 7u;
        const uint32_t home = ((uint32_t)(i >> 2) - d) & kp.nbmask;
        const uint64_t h = ((uint64_t)home << kp.rbits) | (w & kp.rmask);
        const uint64_t key = (h * kp.mul_inv) & kp.mask_b;          // k+1 bases, the first in the highest digit
        const uint32_t vv = (uint32_t)(w >> 48);                    // value bytes of the two k-mers: variant, or 0xFF
        for (uint32_t which = 0; which < 2u; ++which)
        {
            if (((vv >> (8u * which)) & 0xFFu) != 0xFFu) continue;  // held by one variant: not recorded (closed form)
            uint32_t lb = 0u, rb = ix.n - 1u;
            bool ok = true;
            for (uint32_t j = 0; j < k && ok; ++j) ok = fm_step(ix, ix.ctab, lb, rb, (uint32_t)(key >> (2u * (k - (j + which)))) & 3u);
            if (ok) atomicAdd(mult + lb, m);
        }
    }
}

#ifndef SPEQB_PWIN
#define SPEQB_PWIN 1
#endif
#ifndef SPEQB_PMINB
#define SPEQB_PMINB 4
#endif
constexpr int kPair = SPEQB_PWIN;     // pairs of windows per lane and loop trip

// WPR: a warp per read, as in scanh_kernel (long reads)
template <int MODE, bool WPR, bool MASK>
__global__ void __launch_bounds__(kBlockH, (MODE & M_LOCAL) ? 2 : SPEQB_PMINB) scanp_kernel(const KTabView kt, const KTabView kp, unsigned long long * __restrict__ slot_mult,
                                                                                  unsigned long long * __restrict__ pair_mult, const BatchView b, const ScanArgs a,
                                                                                  const int paired)
{
    constexpr bool LOCAL = (MODE & M_LOCAL) != 0;
    constexpr bool RECORD = (MODE & M_RECORD) != 0;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    Smem32 sm = smem32_view(smem_raw, a.V);
    smem32_clear(sm, a.V);
    smem32_tables(sm, a, LOCAL);
    Tally32<MODE> env{b, a, sm};

    const uint32_t k = a.k;
    const uint64_t n_reads_b = b.n_reads_dev ? *b.n_reads_dev : b.n_reads;
    const uint32_t units = (uint32_t)(paired ? n_reads_b / 2 : n_reads_b);
    const uint32_t lane = threadIdx.x & 31u;
    const uint64_t pmask = (kp.rmask << 3) | 7ull;
    const uint32_t pmask_lo = (uint32_t)pmask, pmask_hi = (uint32_t)(pmask >> 32);
    const uint64_t tmask = (kt.rmask << 3) | 7ull;
    const uint32_t tmask_lo = (uint32_t)tmask, tmask_hi = (uint32_t)(tmask >> 32);

    // the tables stream from HBM and must not push the presence filter out of L2
    const uint64_t pol_keep = l2_policy_keep(), pol_stream = l2_policy_stream();
    const uint32_t chunk = claim_chunk(units);                    // see scanh_kernel
    auto probe256 = [&](uint32_t (&dst)[8], const void * ptr, uint64_t pol) {
        if (SPEQB_WIDE_PF && RECORD) ldg256_stream_pf64(dst, ptr, pol);
        else ldg256_stream(dst, ptr, pol);
    };

    uint32_t phase = (!WPR && lane >= chunk) ? H_DONE : H_FETCH;
    bool mate2 = false;
    uint32_t u = 0u;
    uint64_t g0 = 0ull;
    uint32_t t = 0u, pend = 0u, run = 0u;
    uint64_t key = 0ull;                         // key << 2 | base per base (see scanh_kernel)
    uint64_t cb = 0ull, nb = 0ull;
    uint32_t ci = 0u, ni = 0u;
    UnitSet<MASK> uset;
    uint32_t c_total = 0u;
    uint32_t agg_g = kNoGroup, agg_n = 0u;
    double agg_w = 0.0;
    uint32_t n_total = 0u, n_amb = 0u, n_look = 0u, n_reads = 0u, n_probe = 0u;
    uint32_t wcur = 0u, wend = 0u;
    QRoll qr;                                    // LOCAL: the rolling phred weight of the window that ends at the cursor

    auto next_group = [&]() {
        cb = nb; ci = ni;
        if (((t >> 5) + 1u) * 32u < pend)
        {
            nb = load_ro(b.bases + g0 + (t >> 5) + 1u);
            ni = load_ro(b.inv + g0 + (t >> 5) + 1u);
        }
    };
    auto roll1 = [&]() {
        const uint32_t o = t & 31u;
        key = (key << 2) | ((uint32_t)(cb >> (2u * o)) & 3u);
        run = ((ci >> o) & 1u) ? 0u : run + 1u;
        if (LOCAL) qr.step(t, k, run, sm.den, sm.rden);
        t += 1u;
        if ((t & 31u) == 0u) next_group();
    };
    // ++unique_kmers[which_hit] and the read-level ambiguity state (src/fm_scanner.cpp:178-191) for a window whose k-mer
    // lies in variant g only; t_end: end of the window (local mode weight)
    // wgt: the window's phred weight (local mode; the rolling product of scanh_kernel, QRoll)
    auto tally = [&](uint32_t g, double wgt) {
        if (g != agg_g)
        {
            if (agg_g != kNoGroup) env.unique(agg_g, agg_n, agg_w);
            agg_g = g; agg_n = 0u; agg_w = 0.0;
            uset.add(g);
        }
        if (LOCAL) agg_w += wgt;
        agg_n += 1u;
    };
    while (true)
    {
        __syncwarp();
        const uint32_t ph = phase;
        if (!__any_sync(kFullH, ph != H_DONE)) break;
        const uint32_t m_fe = __ballot_sync(kFullH, ph == H_FETCH);
        if (m_fe)
        {
            const bool claim = ph == H_FETCH && !mate2;
            const uint32_t mc = __ballot_sync(kFullH, claim);
            if (WPR)
            {
                if (mc)
                {
                    if (wcur == wend)
                    {
                        uint32_t nb0 = 0u;
                        if (lane == 0u) nb0 = (uint32_t)atomicAdd(a.work, (unsigned long long)kClaimChunkW);
                        wcur = __shfl_sync(kFullH, nb0, 0);
                        wend = wcur + kClaimChunkW;
                    }
                    u = wcur;
                    wcur += 1u;
                }
            }
            else if (mc)
            {
                const uint32_t cnt = (uint32_t)__popc(mc), left = wend - wcur;
                uint32_t cbase = 0u;
                if (cnt > left)
                {
                    if (lane == 0u) cbase = (uint32_t)atomicAdd(a.work, (unsigned long long)chunk);
                    cbase = __shfl_sync(kFullH, cbase, 0);
                }
                const uint32_t r = (uint32_t)__popc(mc & ((1u << lane) - 1u));
                if (claim) u = r < left ? wcur + r : cbase + (r - left);
                if (cnt > left) { wcur = cbase + (cnt - left); wend = cbase + chunk; }
                else wcur += cnt;
            }
            if (ph == H_FETCH)
            {
                if (u >= units) phase = H_DONE;
                else
                {
                    const uint64_t fr = paired ? 2ull * u + (mate2 ? 1u : 0u) : (uint64_t)u;
                    pend = load_ro(b.len + fr);
                    g0 = load_ro(b.grp_off + fr);
                    phase = H_FIRST;
                }
            }
        }
        if (ph == H_FIRST)
        {
            const uint32_t L = pend;
            run = 0u; key = 0ull;
            if (WPR)
            {
                const uint32_t nwin = L >= k ? L - k + 1u : 0u, per = (nwin + 31u) >> 5;
                const uint32_t s0 = lane * per < nwin ? lane * per : nwin, s1 = nwin - s0 < per ? nwin : s0 + per;
                if (lane == 0u) { n_look += nwin; n_reads += 1u; }
                t = s0;
                pend = s1 > s0 ? s1 + k - 1u : s0;                // this lane's bases: its windows and the k-1 bases before
            }
            else
            {
                pend = L >= k ? L : 0u;                           // L < k: the sliding view is empty (sliding.hpp:129-142)
                n_look += L >= k ? L - k + 1u : 0u;
                n_reads += 1u;
                t = 0u;
            }
            if (t < pend)
            {
                cb = load_ro(b.bases + g0 + (t >> 5));
                ci = load_ro(b.inv + g0 + (t >> 5));
                if (((t >> 5) + 1u) * 32u < pend)
                {
                    nb = load_ro(b.bases + g0 + (t >> 5) + 1u);
                    ni = load_ro(b.inv + g0 + (t >> 5) + 1u);
                }
                if (LOCAL) qr.start(b.qual + 32ull * g0, L, t, k);
            }
            phase = H_RUN;
        }
        else if (ph == H_RUN)
        {
            while (run + 1u < k && t < pend) roll1();             // start of a read / behind an N: no window is due yet
            uint32_t w[kPair][8];
            uint32_t pb[kPair], want_lo[kPair], want_hi[kPair];
            uint64_t key_a[kPair], key_b[kPair];                  // the k-mer register after the first / second base of each pair
            uint32_t mode[kPair];                                 // bit 0: first window due, bit 1: second window due
            double pw[2 * kPair];                                 // LOCAL: weight of the window that ends with base i of the trip
            const uint32_t t0 = t, o = t & 31u;
            const uint32_t nbase = pend - t < 2u * (uint32_t)kPair ? pend - t : 2u * (uint32_t)kPair;
            uint32_t bits = (uint32_t)(cb >> (2u * o)), ibits = ci >> o;
            if (o + 2u * (uint32_t)kPair > 32u && o != 0u)
            {
                bits |= (uint32_t)nb << (2u * (32u - o));
                ibits |= ni << (32u - o);
            }
#pragma unroll
            for (int j = 0; j < kPair; ++j)
            {
                mode[j] = 0u;
                if ((uint32_t)(2 * j) < nbase)
                {
                    key = (key << 2) | ((bits >> (4 * j)) & 3u);
                    run = ((ibits >> (2 * j)) & 1u) ? 0u : run + 1u;
                    if (LOCAL) { qr.step(t0 + 2u * (uint32_t)j, k, run, sm.den, sm.rden); pw[2 * j] = qr.P; }
                    mode[j] = run >= k ? 1u : 0u;
                }
                key_a[j] = key;
                if ((uint32_t)(2 * j + 1) < nbase)
                {
                    key = (key << 2) | ((bits >> (4 * j + 2)) & 3u);
                    run = ((ibits >> (2 * j + 1)) & 1u) ? 0u : run + 1u;
                    if (LOCAL) { qr.step(t0 + 2u * (uint32_t)j + 1u, k, run, sm.den, sm.rden); pw[2 * j + 1] = qr.P; }
                    mode[j] |= run >= k ? 2u : 0u;
                }
                key_b[j] = key;
            }
            t += nbase;
            if (o + nbase >= 32u) next_group();
            // ---- presence filter (L2): a window whose k-mer is not in the index -- a sequencing error -- is answered here and
            //      costs no HBM sector; a pair is probed only when both of its k-mers may exist
            if (kt.filt != nullptr)
            {
                uint32_t fw[2 * kPair], fb[2 * kPair];
#pragma unroll
                for (int i = 0; i < 2 * kPair; ++i)
                {
                    if (!((mode[i >> 1] >> (i & 1)) & 1u)) continue;
                    const uint64_t h = (((i & 1) ? key_b[i >> 1] : key_a[i >> 1]) * kt.mul) & kt.mask_b;
                    fb[i] = (uint32_t)(h >> kt.fshift);
                    fw[i] = ldg32_keep(kt.filt + (fb[i] >> 5), pol_keep);
                }
#pragma unroll
                for (int i = 0; i < 2 * kPair; ++i)
                {
                    if (!((mode[i >> 1] >> (i & 1)) & 1u)) continue;
                    c_total += 1u;                                // ++total_kmers (src/fm_scanner.cpp:164)
                    if (!((fw[i] >> (fb[i] & 31u)) & 1u)) mode[i >> 1] &= ~(1u << (i & 1));
                }
            }
            else
            {
#pragma unroll
                for (int j = 0; j < kPair; ++j) c_total += (mode[j] & 1u) + (mode[j] >> 1);
            }
#pragma unroll
            for (int j = 0; j < kPair; ++j)
            {
                if (mode[j] == 3u)
                {
                    const uint64_t h = (key_b[j] * kp.mul) & kp.mask_b;
                    pb[j] = (uint32_t)(h >> kp.rbits);
                    const uint64_t want = h & kp.rmask;
                    want_lo[j] = (uint32_t)want; want_hi[j] = (uint32_t)(want >> 32);
                    probe256(w[j], kt_bucket(kp, pb[j]), pol_stream);
                }
            }
            // ---- pairs: one probe answers both windows
#pragma unroll
            for (int j = 0; j < kPair; ++j)
            {
                if (mode[j] != 3u) continue;
                n_probe += 1u;
                uint32_t d = 0u;
                while (true)
                {
                    const bool h0 = (((w[j][0] ^ want_lo[j]) & pmask_lo) | ((w[j][1] ^ want_hi[j]) & pmask_hi)) == 0u;
                    const bool h1 = (((w[j][2] ^ want_lo[j]) & pmask_lo) | ((w[j][3] ^ want_hi[j]) & pmask_hi)) == 0u;
                    const bool h2 = (((w[j][4] ^ want_lo[j]) & pmask_lo) | ((w[j][5] ^ want_hi[j]) & pmask_hi)) == 0u;
                    const bool h3 = (((w[j][6] ^ want_lo[j]) & pmask_lo) | ((w[j][7] ^ want_hi[j]) & pmask_hi)) == 0u;
                    if (h0 || h1 || h2 || h3)
                    {
                        const uint32_t vv = (h0 ? w[j][1] : h1 ? w[j][3] : h2 ? w[j][5] : w[j][7]) >> 16;
                        const uint32_t v1 = vv & 0xFFu, v2 = vv >> 8;
                        // a k-mer of one variant needs no record (see scanh_kernel); the fold skips that half of the pair
                        if (RECORD && (v1 == 0xFFu || v2 == 0xFFu)) atomicAdd(kt_counter(kp, pair_mult, 4ull * pb[j] + (h0 ? 0u : h1 ? 1u : h2 ? 2u : 3u)), 1ull);
                        if (v1 != 0xFFu) tally(v1, LOCAL ? pw[2 * j] : 0.0);
                        if (v2 != 0xFFu) tally(v2, LOCAL ? pw[2 * j + 1] : 0.0);
                        mode[j] = 0u;                             // answered
                        break;
                    }
                    if ((w[j][6] & w[j][7]) == 0xFFFFFFFFu || d == kMaxDisp) break;   // not a (k+1)-mer of the index
                    d += 1u;
                    n_probe += 1u;
                    pb[j] = (pb[j] + 1u) & kp.nbmask;
                    const uint64_t want = u64of(want_lo[j], want_hi[j]) + (1ull << kp.rbits);
                    want_lo[j] = (uint32_t)want; want_hi[j] = (uint32_t)(want >> 32);
                    probe256(w[j], kt_bucket(kp, pb[j]), pol_stream);
                }
            }
            // ---- what is left: unpaired windows and the halves of missed pairs go through the k-mer table, all first probes
            //      issued together (w[] of the pairs is dead by now)
            {
                uint32_t sw[2 * kPair][8];
                uint32_t sb[2 * kPair], sl[2 * kPair], sh_[2 * kPair];
#pragma unroll
                for (int i = 0; i < 2 * kPair; ++i)
                {
                    if (!((mode[i >> 1] >> (i & 1)) & 1u)) continue;
                    const uint64_t h = (((i & 1) ? key_b[i >> 1] : key_a[i >> 1]) * kt.mul) & kt.mask_b;
                    sb[i] = (uint32_t)(h >> kt.rbits);
                    const uint64_t want = h & kt.rmask;
                    sl[i] = (uint32_t)want; sh_[i] = (uint32_t)(want >> 32);
                    probe256(sw[i], kt_bucket(kt, sb[i]), pol_stream);
                }
#pragma unroll
                for (int i = 0; i < 2 * kPair; ++i)
                {
                    if (!((mode[i >> 1] >> (i & 1)) & 1u)) continue;
                    n_probe += 1u;
                    uint32_t d = 0u;
                    while (true)
                    {
                        const bool h0 = (((sw[i][0] ^ sl[i]) & tmask_lo) | ((sw[i][1] ^ sh_[i]) & tmask_hi)) == 0u;
                        const bool h1 = (((sw[i][2] ^ sl[i]) & tmask_lo) | ((sw[i][3] ^ sh_[i]) & tmask_hi)) == 0u;
                        const bool h2 = (((sw[i][4] ^ sl[i]) & tmask_lo) | ((sw[i][5] ^ sh_[i]) & tmask_hi)) == 0u;
                        const bool h3 = (((sw[i][6] ^ sl[i]) & tmask_lo) | ((sw[i][7] ^ sh_[i]) & tmask_hi)) == 0u;
                        if (h0 || h1 || h2 || h3)
                        {
                            const uint32_t hi = h0 ? sw[i][1] : h1 ? sw[i][3] : h2 ? sw[i][5] : sw[i][7];
                            if (!slot_multi(hi)) tally(hi >> 16, LOCAL ? pw[i] : 0.0);
                            else if (RECORD) atomicAdd(kt_counter(kt, slot_mult, 4ull * sb[i] + (h0 ? 0u : h1 ? 1u : h2 ? 2u : 3u)), 1ull);
                            break;
                        }
                        if ((sw[i][6] & sw[i][7]) == 0xFFFFFFFFu || d == kMaxDisp) break;   // the k-mer does not occur in the index
                        d += 1u;
                        n_probe += 1u;
                        sb[i] = (sb[i] + 1u) & kt.nbmask;
                        const uint64_t want = u64of(sl[i], sh_[i]) + (1ull << kt.rbits);
                        sl[i] = (uint32_t)want; sh_[i] = (uint32_t)(want >> 32);
                        probe256(sw[i], kt_bucket(kt, sb[i]), pol_stream);
                    }
                }
            }
        }
        if (WPR)
        {
            // the lane's share of the read is finished: wait for the warp, then merge the read's ambiguity state
            if (phase == H_RUN && t >= pend) phase = H_WAIT;
            if (__all_sync(kFullH, phase == H_WAIT))
            {
                if (paired && !mate2) mate2 = true;
                else
                {
                    uset.warp_merge();
                    n_total += c_total;
                    if (lane == 0u && uset.ambiguous()) { n_amb += 1u; env.fusion(uset.key()); }
                    c_total = 0u; uset.reset();
                    if (agg_g != kNoGroup) env.unique(agg_g, agg_n, agg_w);
                    agg_g = kNoGroup; agg_n = 0u; agg_w = 0.0;
                    mate2 = false;
                    if ((n_total | n_look) & 0x80000000u)
                    {
                        env.add64(0, n_total); env.add64(2, n_look);
                        n_total = 0u; n_look = 0u;
                    }
                }
                phase = H_FETCH;
            }
        }
        else if (phase == H_RUN && t >= pend)
        {
            if (paired && !mate2) mate2 = true;
            else
            {
                // is_ambiguous: unique k-mers of two different variants in one read / pair (src/fm_scanner.cpp:183-190,213)
                n_total += c_total;
                if (uset.ambiguous()) { n_amb += 1u; env.fusion(uset.key()); }
                c_total = 0u; uset.reset();
                if (agg_g != kNoGroup) env.unique(agg_g, agg_n, agg_w);
                agg_g = kNoGroup; agg_n = 0u; agg_w = 0.0;
                mate2 = false;
                if ((n_total | n_look) & 0x80000000u)
                {
                    env.add64(0, n_total); env.add64(2, n_look);
                    n_total = 0u; n_look = 0u;
                }
            }
            phase = H_FETCH;
        }
    }
    if (agg_g != kNoGroup) env.unique(agg_g, agg_n, agg_w);
    if (n_total) env.add64(0, n_total);
    if (n_amb) env.add64(1, n_amb);
    if (n_look) env.add64(2, n_look);
    if (n_reads) env.add64(3, n_reads);
    {
        // sectors probed (table loads issued): one global atomic per warp
        const uint32_t wp = __reduce_add_sync(kFullH, n_probe);
        if (lane == 0u && wp) atomicAdd(a.scalars + SC_STEPS, (unsigned long long)wp);
    }
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

constexpr int kWprBit = 16;   // dispatch selector: the warp-per-read instantiation
constexpr int kMaskBit = 32;  // dispatch selector: the unit's variants as a 64-bit mask (V <= kMaxGroupsMask)
#define SPEQB_SCANH_CASES(K, B, MK, EXPR)                                               \
        case (B): { auto kern = K<0, false, MK>; EXPR; break; }                         \
        case (B) | M_RECORD: { auto kern = K<M_RECORD, false, MK>; EXPR; break; }       \
        case (B) | M_LOCAL: { auto kern = K<M_LOCAL, false, MK>; EXPR; break; }         \
        case (B) | M_LOCAL | M_RECORD: { auto kern = K<M_LOCAL | M_RECORD, false, MK>; EXPR; break; } \
        case (B) | kWprBit: { auto kern = K<0, true, MK>; EXPR; break; }                \
        case (B) | kWprBit | M_RECORD: { auto kern = K<M_RECORD, true, MK>; EXPR; break; } \
        case (B) | kWprBit | M_LOCAL: { auto kern = K<M_LOCAL, true, MK>; EXPR; break; } \
        case (B) | kWprBit | M_LOCAL | M_RECORD: { auto kern = K<M_LOCAL | M_RECORD, true, MK>; EXPR; break; }
#define SPEQB_SCANH_DISPATCH(M, EXPR)                                                   \
    switch (M)                                                                          \
    {                                                                                   \
        SPEQB_SCANH_CASES(scanh_kernel, 0, false, EXPR)                                 \
        SPEQB_SCANH_CASES(scanh_kernel, kMaskBit, true, EXPR)                           \
        default: return cudaErrorInvalidValue;                                          \
    }

cudaError_t scanh_launch_cfg(int mode, uint32_t V, int walk, LaunchCfg * cfg)
{
    const size_t smem = scan32_smem_bytes(V);
    int bps = 0;
    cudaError_t e = cudaSuccess;
    // both instantiations (lane per read, warp per read) are launched with one grid: the smaller occupancy
    for (int wpr = 0; wpr <= kWprBit; wpr += kWprBit)
    {
        const int m = (mode & (M_LOCAL | M_RECORD)) | wpr | (V <= kMaxGroupsMask ? kMaskBit : 0);
        int bw = 0;
        SPEQB_SCANH_DISPATCH(m, {
            if (smem > 48 * 1024) e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bw, kern, kBlockH, smem);
        })
        if (e != cudaSuccess) return e;
        if (bw <= 0) return cudaErrorLaunchOutOfResources;
        bps = wpr == 0 || bw < bps ? bw : bps;
    }
    if (walk && !(mode & M_LOCAL))
    {
        // walk != 0: the grid of the walking scan alone (it has its own launch configuration: the probing kernels keep theirs)
        int bw = 0;
        bps = 1 << 30;
        const bool rec = (mode & M_RECORD) != 0, mask = V <= kMaxGroupsMask;
        const size_t smem_w = scanw_queue_offset(V) + kWalkQueueBytes;
        auto occ = [&](auto kern) {
            if (smem_w > 48 * 1024) e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_w);
            if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bw, kern, kBlockH, smem_w);
        };
        if (rec && mask) occ(scanw_kernel<true, true>);
        else if (rec) occ(scanw_kernel<true, false>);
        else if (mask) occ(scanw_kernel<false, true>);
        else occ(scanw_kernel<false, false>);
        if (e != cudaSuccess) return e;
        if (bw <= 0) return cudaErrorLaunchOutOfResources;
        bps = bw < bps ? bw : bps;
    }
    cfg->block = kBlockH;
    cfg->grid = sm_count_h() * bps;
    return cudaSuccess;
}

// Geometry of the table for `keys` distinct k-mers (an upper bound is fine): load factor in (0.2, 0.4] (down to 0.1 for tables
// of at most kSmallTableBytes), grown by `grow` doublings; the remainder must fit below the displacement and value fields.
constexpr uint64_t kSmallTableBytes = 32ull << 20;
bool ktable_geometry(uint32_t k, uint64_t keys, uint32_t grow, KTabView * kt)
{
    const uint32_t B = 2u * k;
    uint32_t q = 1u;
    while (q < B && (1ull << q) * 8ull < keys * 5ull) ++q;       // 4 * 2^q * 0.4 >= keys
    // while the table stays far below L2 anyway, spend memory on a load factor <= 0.1: a k-mer displaced from a full home
    // bucket costs a second, dependent round trip, and one such lane holds up its whole warp
    while (q < B && (64ull << q) <= kSmallTableBytes && (4ull << q) < keys * 10ull) ++q;
    q = std::min(B, q + grow);
    if (B > 45u && q < B - 45u) q = B - 45u;
    if (q > 31u) return false;
    kt->q = q;
    kt->rbits = B - q;
    kt->nbmask = (1u << q) - 1u;
    kt->mask_b = B == 64u ? ~0ull : (1ull << B) - 1ull;
    kt->rmask = kt->rbits == 0u ? 0ull : (1ull << kt->rbits) - 1ull;
    kt->mul = 0x9E3779B97F4A7C15ull;
    uint64_t x = kt->mul;                                         // Newton: the inverse mod 2^64 (and so mod 2^B)
    for (int i = 0; i < 6; ++i) x *= 2ull - kt->mul * x;
    kt->mul_inv = x;
    kt->filt = nullptr;
    kt->fshift = 0u;
    kt->pshift = 0u;
    kt->wide = 0u;
    kt->walk = 0u;
    kt->tgroups = 0u;
    kt->tbases = nullptr;
    kt->tflags = nullptr;
    kt->gvar = nullptr;
    kt->pos_side = nullptr;
    kt->vshift = 0u;
    return true;
}

// Presence filter: one bit per value of the top fbits bits of the k-mer hash, about 4 bits per stored k-mer, at most
// 2^SPEQB_FBITS_MAX bits (32 MB: it has to stay in L2 while the tables stream from HBM).
#ifndef SPEQB_FBITS_MAX
#define SPEQB_FBITS_MAX 28
#endif
uint32_t kfilter_bits(uint32_t k, uint64_t keys)
{
    uint32_t f = 10u;
    while (f < (uint32_t)SPEQB_FBITS_MAX && (1ull << f) < keys * 4ull) ++f;
    return std::min(f, 2u * k);
}

cudaError_t launch_kfilter_build(KTabView const & kt, uint32_t fbits, cudaStream_t st)
{
    cudaError_t e = cudaMemsetAsync(kt.filt, 0, std::max<uint64_t>(4, (1ull << fbits) / 8), st);
    if (e != cudaSuccess) return e;
    kfilter_build_kernel<<<sm_count_h() * 8, kBlockH, 0, st>>>(kt);
    return cudaGetLastError();
}

cudaError_t launch_scanh(KTabView const & kt, unsigned long long * slot_mult, BatchView const & b, ScanArgs const & a, int mode,
                         int warp_per_read, LaunchCfg cfg, cudaStream_t st)
{
    const size_t smem = scan32_smem_bytes(a.V);
    const int paired = (mode & M_PAIRED) ? 1 : 0;
    const int m = (mode & (M_LOCAL | M_RECORD)) | (warp_per_read ? kWprBit : 0) | (a.V <= kMaxGroupsMask ? kMaskBit : 0);
    cudaError_t e = cudaMemsetAsync(a.work, 0, sizeof(unsigned long long), st);
    if (e != cudaSuccess) return e;
    if (kt.walk && !warp_per_read && !(mode & M_LOCAL))
    {
        // global mode, one lane per read, table in walk form: the walking scan
        const bool rec = (mode & M_RECORD) != 0, mask = a.V <= kMaxGroupsMask;
        const size_t smem_w = scanw_queue_offset(a.V) + kWalkQueueBytes;
        if (rec && mask) scanw_kernel<true, true><<<cfg.grid, cfg.block, smem_w, st>>>(kt, b, a, paired);
        else if (rec) scanw_kernel<true, false><<<cfg.grid, cfg.block, smem_w, st>>>(kt, b, a, paired);
        else if (mask) scanw_kernel<false, true><<<cfg.grid, cfg.block, smem_w, st>>>(kt, b, a, paired);
        else scanw_kernel<false, false><<<cfg.grid, cfg.block, smem_w, st>>>(kt, b, a, paired);
        return cudaGetLastError();
    }
    SPEQB_SCANH_DISPATCH(m, (kern<<<cfg.grid, cfg.block, smem, st>>>(kt, slot_mult, b, a, paired)))
    return cudaGetLastError();
}

#define SPEQB_SCANP_DISPATCH(M, EXPR)                                                   \
    switch (M)                                                                          \
    {                                                                                   \
        SPEQB_SCANH_CASES(scanp_kernel, 0, false, EXPR)                                 \
        SPEQB_SCANH_CASES(scanp_kernel, kMaskBit, true, EXPR)                           \
        default: return cudaErrorInvalidValue;                                          \
    }

cudaError_t scanp_launch_cfg(int mode, uint32_t V, LaunchCfg * cfg)
{
    const size_t smem = scan32_smem_bytes(V);
    int bps = 0;
    cudaError_t e = cudaSuccess;
    for (int wpr = 0; wpr <= kWprBit; wpr += kWprBit)
    {
        const int m = (mode & (M_LOCAL | M_RECORD)) | wpr | (V <= kMaxGroupsMask ? kMaskBit : 0);
        int bw = 0;
        SPEQB_SCANP_DISPATCH(m, {
            if (smem > 48 * 1024) e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bw, kern, kBlockH, smem);
        })
        if (e != cudaSuccess) return e;
        if (bw <= 0) return cudaErrorLaunchOutOfResources;
        bps = wpr == 0 || bw < bps ? bw : bps;
    }
    cfg->block = kBlockH;
    cfg->grid = sm_count_h() * bps;
    return cudaSuccess;
}

cudaError_t launch_scanp(KTabView const & kt, KTabView const & kp, unsigned long long * slot_mult, unsigned long long * pair_mult,
                         BatchView const & b, ScanArgs const & a, int mode, int warp_per_read, LaunchCfg cfg, cudaStream_t st)
{
    const size_t smem = scan32_smem_bytes(a.V);
    const int paired = (mode & M_PAIRED) ? 1 : 0;
    const int m = (mode & (M_LOCAL | M_RECORD)) | (warp_per_read ? kWprBit : 0) | (a.V <= kMaxGroupsMask ? kMaskBit : 0);
    cudaError_t e = cudaMemsetAsync(a.work, 0, sizeof(unsigned long long), st);
    if (e != cudaSuccess) return e;
    SPEQB_SCANP_DISPATCH(m, (kern<<<cfg.grid, cfg.block, smem, st>>>(kt, kp, slot_mult, pair_mult, b, a, paired)))
    return cudaGetLastError();
}

// wide tables: slots_too != 0: all slots free (all ones) and the counters zero; else only the counters back to zero
__global__ void __launch_bounds__(kBlockH) ktable_wide_clear_kernel(uint64_t * __restrict__ slots, const uint64_t buckets, const int slots_too)
{
    // one 16-byte store per thread: 4 threads per 64-byte record
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < buckets * 4ull; i += (uint64_t)gridDim.x * blockDim.x)
    {
        const uint32_t part = (uint32_t)(i & 3ull);
        if (part < 2u && !slots_too) continue;
        const uint32_t v = part < 2u ? 0xFFFFFFFFu : 0u;
        reinterpret_cast<uint4 *>(slots)[i] = make_uint4(v, v, v, v);
    }
}
// all slots free; wide tables: their counters zero
cudaError_t launch_ktable_clear(KTabView const & kt, cudaStream_t st)
{
    const size_t buckets = (size_t)kt.nbmask + 1;
    if (!kt.wide) return cudaMemsetAsync(kt.slots, 0xFF, buckets * 32, st);
    ktable_wide_clear_kernel<<<sm_count_h() * 8, kBlockH, 0, st>>>(kt.slots, buckets, 1);
    return cudaGetLastError();
}
// wide tables: the counters back to zero (speqg_scan_reset)
cudaError_t launch_ktable_zero_counters(KTabView const & kt, cudaStream_t st)
{
    if (!kt.wide) return cudaSuccess;
    ktable_wide_clear_kernel<<<sm_count_h() * 8, kBlockH, 0, st>>>(kt.slots, (size_t)kt.nbmask + 1, 0);
    return cudaGetLastError();
}

// the pair table: every N-free (k+1)-mer of the texts -> the values of its two k-mers (looked up in the k-mer table kt)
cudaError_t launch_kpair_build(const uint64_t * t_bases, const uint32_t * t_inv, const uint64_t * t_grp_off, const uint32_t * t_len,
                               uint32_t n_texts, uint32_t max_len, KTabView const & kt, KTabView const & kp, uint32_t k,
                               unsigned long long * n_keys, cudaStream_t st)
{
    cudaError_t e = launch_ktable_clear(kp, st);
    if (e == cudaSuccess) e = cudaMemsetAsync(n_keys, 0, 16, st);
    if (e != cudaSuccess) return e;
    if (n_texts == 0 || max_len < k + 1) return cudaSuccess;
    const uint32_t groups = (max_len + 31u) / 32u;
    const uint32_t gx = std::min<uint32_t>((groups + kBlockH - 1) / kBlockH, (uint32_t)sm_count_h() * 8u);
    for (uint32_t t0 = 0; t0 < n_texts; t0 += 65535u)
    {
        const uint32_t nt = std::min<uint32_t>(65535u, n_texts - t0);
        kpair_build_kernel<<<dim3(gx, nt), kBlockH, 0, st>>>(t_bases, t_inv, t_grp_off + t0, t_len + t0, kt, kp, k, n_keys);
    }
    return cudaGetLastError();
}

cudaError_t launch_kpair_fold_mult(KTabView const & kp, DevIndex const & ix, uint32_t k, unsigned long long * slot_mult,
                                   unsigned long long * mult, cudaStream_t st)
{
    kpair_fold_mult_kernel<<<sm_count_h() * 8, kBlockH, 0, st>>>(kp, ix, k, slot_mult, mult);
    return cudaGetLastError();
}

// n_keys: two u64 -- [0] distinct k-mers stored, [1] set when a k-mer could not be placed within kMaxDisp buckets
cudaError_t launch_ktable_build(const uint64_t * t_bases, const uint32_t * t_inv, const uint64_t * t_grp_off, const uint32_t * t_len,
                                const int32_t * t_group, uint32_t n_texts, uint32_t max_len, KTabView const & kt, uint32_t k,
                                unsigned long long * n_keys, cudaStream_t st)
{
    cudaError_t e = launch_ktable_clear(kt, st);
    if (e == cudaSuccess) e = cudaMemsetAsync(n_keys, 0, 16, st);
    if (e != cudaSuccess) return e;
    if (n_texts == 0 || max_len < k) return cudaSuccess;
    const uint32_t groups = (max_len + 31u) / 32u;
    const uint32_t gx = std::min<uint32_t>((groups + kBlockH - 1) / kBlockH, (uint32_t)sm_count_h() * 8u);
    // grid.y <= 65535 texts per launch
    for (uint32_t t0 = 0; t0 < n_texts; t0 += 65535u)
    {
        const uint32_t nt = std::min<uint32_t>(65535u, n_texts - t0);
        ktable_build_kernel<<<dim3(gx, nt), kBlockH, 0, st>>>(t_bases, t_inv, t_grp_off + t0, t_len + t0, t_group + t0, kt, k, n_keys);
    }
    return cudaGetLastError();
}

// positions fit above the displacement (and below the variant field's all-ones pattern the first pass leaves in such slots)
bool ktable_positions_fit(KTabView const & kt, uint64_t npos)
{
    uint32_t bits = 1u;
    while (bits < 63u && (1ull << bits) < npos + 2ull) ++bits;
    return kt.rbits + bits <= 59u && npos < 0xFFFFFFF0ull;
}

cudaError_t launch_ktable_canon(const uint64_t * t_bases, const uint32_t * t_inv, const uint64_t * t_grp_off, const uint32_t * t_len,
                                const int32_t * t_group, uint32_t n_texts, uint32_t max_len, KTabView const & kt, uint32_t k,
                                uint64_t * tflags, uint16_t * gvar, cudaStream_t st)
{
    if (n_texts == 0 || max_len < k) return cudaSuccess;
    const uint32_t groups = (max_len + 31u) / 32u;
    const uint32_t gx = std::min<uint32_t>((groups + kBlockH - 1) / kBlockH, (uint32_t)sm_count_h() * 8u);
    if (tflags && kt.pshift != 0u) ktable_prep_pos_kernel<<<sm_count_h() * 8, kBlockH, 0, st>>>(kt);
    if (tflags && kt.pshift == 0u)
    {
        cudaError_t e = cudaMemsetAsync(kt.pos_side, 0xFF, 16ull * ((uint64_t)kt.nbmask + 1ull), st);
        if (e != cudaSuccess) return e;
    }
    for (uint32_t t0 = 0; t0 < n_texts; t0 += 65535u)
    {
        const uint32_t nt = std::min<uint32_t>(65535u, n_texts - t0);
        if (tflags)
            ktable_canon_kernel<true><<<dim3(gx, nt), kBlockH, 0, st>>>(t_bases, t_inv, t_grp_off + t0, t_len + t0, t_group + t0, kt, k, tflags, gvar);
        else
            ktable_canon_kernel<false><<<dim3(gx, nt), kBlockH, 0, st>>>(t_bases, t_inv, t_grp_off + t0, t_len + t0, nullptr, kt, k, nullptr, nullptr);
    }
    return cudaGetLastError();
}

// diff -> inclusive prefix sum in place (cub) -> multiplicities at the interval starts; diff is left zeroed
cudaError_t launch_kpos_fold(const uint64_t * t_bases, DevIndex const & ix, uint32_t k, unsigned long long * diff, uint64_t npos,
                             void * tmp, size_t tmp_bytes, unsigned long long * mult, KTabView const & kt, cudaStream_t st)
{
    size_t need_bytes = 0;
    cudaError_t e = cub::DeviceScan::InclusiveSum(nullptr, need_bytes, diff, diff, (int64_t)(npos + 2), st);
    if (e != cudaSuccess) return e;
    if (need_bytes > tmp_bytes) return cudaErrorInvalidValue;
    e = cub::DeviceScan::InclusiveSum(tmp, tmp_bytes, diff, diff, (int64_t)(npos + 2), st);
    if (e != cudaSuccess) return e;
    if (kt.walk) kpos_canon_counts_kernel<<<sm_count_h() * 8, kBlockH, 0, st>>>(kt, k, diff, npos);
    kpos_fold_kernel<<<sm_count_h() * 8, kBlockH, 0, st>>>(t_bases, ix, k, diff, npos + 2, mult);
    return cudaGetLastError();
}

size_t kpos_fold_tmp_bytes(uint64_t npos)
{
    size_t need_bytes = 0;
    unsigned long long * d = nullptr;
    cub::DeviceScan::InclusiveSum(nullptr, need_bytes, d, d, (int64_t)(npos + 2), (cudaStream_t)0);
    return need_bytes + 256;
}

cudaError_t launch_ktable_fold_mult(KTabView const & kt, DevIndex const & ix, uint32_t k, unsigned long long * slot_mult,
                                    unsigned long long * mult, cudaStream_t st)
{
    ktable_fold_mult_kernel<<<sm_count_h() * 8, kBlockH, 0, st>>>(kt, ix, k, slot_mult, mult);
    return cudaGetLastError();
}
}   // namespace speqb
