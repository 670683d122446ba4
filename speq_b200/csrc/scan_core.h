// scan_core.h -- per-lane arithmetic of the scan hot path: rank on the blocked BWT, one backward-search step, the
// LCP contraction, the rolling k-mer window and the from-the-root ("direct") window search.
//
// Everything here is __host__ __device__: kernels.cu runs it on the GPU (the product path); tests/sim compiles the same
// header with g++ as a white-box debugging aid (never shipped, never linked into libspeq_cuda.so).
//
// Reference arithmetic reproduced (paths relative to the reference checkout, S3 = deps/seqan3):
//   step        fm_index_cursor::backward_search, S3/include/seqan3/search/fm_index/fm_index_cursor.hpp:110-146:
//               whole interval -> [C[c], C[c+1]-1]; else l' = C[c]+rank(l,c), r' = C[c]+rank(r+1,c)-1; fail if r'<l'
//   direct      extend_right(range), fm_index_cursor.hpp:301-330: k steps from the root, stop at the first failure
//   window rule ranges::views::sliding(k) -> windows [s, s+k), s in [0, L-k] (range-v3 view/sliding.hpp:129-142);
//               valid iff no base of the window is N or has phred <= cutoff (src/fm_scanner.cpp:162)
// The rolling search visits every read position once: the state is the SA interval of the longest suffix of
// read[0..t] that occurs in the text, capped at k symbols.  Because the SA interval of a pattern is unique, the interval
// it reports for a window is bit-identical to the one k independent steps from the root produce.
#pragma once
#include "fm_layout.h"

namespace speqb
{
#if defined(__CUDA_ARCH__)
SPEQB_HD uint32_t popc32(uint32_t x) { return (uint32_t)__popc(x); }
SPEQB_HD uint32_t clz64(uint64_t x) { return (uint32_t)__clzll((long long)x); }
SPEQB_HD uint32_t ctz64(uint64_t x) { return (uint32_t)__ffsll((long long)x) - 1u; }
SPEQB_HD RankBlock load_block(const RankBlock * p)
{
    RankBlock b;
    asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(b.cnt[0]), "=r"(b.cnt[1]), "=r"(b.cnt[2]), "=r"(b.cnt[3]), "=r"(b.lo[0]), "=r"(b.lo[1]),
                   "=r"(b.hi[0]), "=r"(b.hi[1])
                 : "l"(p));
    return b;
}
template <class T> SPEQB_HD T load_ro(const T * p) { return __ldg(p); }
SPEQB_HD GroupRun load_gr(const GroupRun * p)
{
    uint2 v = __ldg(reinterpret_cast<const uint2 *>(p));
    GroupRun g;
    g.run_end = v.x;
    g.group = v.y;
    return g;
}
#else
SPEQB_HD uint32_t popc32(uint32_t x) { return (uint32_t)__builtin_popcount(x); }
SPEQB_HD uint32_t clz64(uint64_t x) { return (uint32_t)__builtin_clzll(x); }
SPEQB_HD uint32_t ctz64(uint64_t x) { return (uint32_t)__builtin_ctzll(x); }
SPEQB_HD RankBlock load_block(const RankBlock * p) { return *p; }
template <class T> SPEQB_HD T load_ro(const T * p) { return *p; }
SPEQB_HD GroupRun load_gr(const GroupRun * p) { return *p; }
#endif

// first index in sorted a[0, n) whose value is >= x
SPEQB_HD uint32_t lower_bound_u32(const uint32_t * a, uint32_t n, uint32_t x)
{
    uint32_t lo = 0, hi = n;
    while (lo < hi)
    {
        uint32_t mid = (lo + hi) >> 1;
        if (load_ro(a + mid) < x) lo = mid + 1; else hi = mid;
    }
    return lo;
}

// symbols with 2-bit code c among the first j (0..63) symbols of a block; exceptions are coded 0 and counted as A here
SPEQB_HD uint32_t block_count(RankBlock const & b, uint32_t c, uint32_t j)
{
    uint32_t m0 = j >= 32u ? 0xFFFFFFFFu : ((1u << j) - 1u);
    uint32_t m1 = j <= 32u ? 0u : ((1u << (j - 32u)) - 1u);
    uint32_t xl = (c & 1u) ? 0u : 0xFFFFFFFFu;
    uint32_t xh = (c & 2u) ? 0u : 0xFFFFFFFFu;
    uint32_t w0 = (b.lo[0] ^ xl) & (b.hi[0] ^ xh) & m0;
    uint32_t w1 = (b.lo[1] ^ xl) & (b.hi[1] ^ xh) & m1;
    return popc32(w0) + popc32(w1);
}

// rank(i, c): occurrences of query code c (A0 C1 G2 T3) in bwt[0, i), 0 <= i <= n, given the block that holds i
SPEQB_HD uint32_t rank_in(DevIndex const & ix, RankBlock const & b, uint32_t i, uint32_t c)
{
    uint32_t base = b.cnt[c];
    uint32_t r = block_count(b, c, i & (kBlockSyms - 1u));
    if (c == 0u)
    {
        if (base & kExcFlag)
        {
            base &= ~kExcFlag;
            uint32_t start = i & ~(kBlockSyms - 1u);
            if (i != start)
                r -= lower_bound_u32(ix.exc_pos, ix.n_exc, i) - lower_bound_u32(ix.exc_pos, ix.n_exc, start);
        }
    }
    return base + r;
}

// One backward-search step with query code c (0..3, or kCodeN).  Returns false (interval untouched) when the extended
// pattern does not occur.
// ctab: C[c] at ctab[c], Cend[c] at ctab[kCtabEnd + c] (a shared-memory copy on the device: indexing the kernel
// parameter struct dynamically would spill it to local memory).
SPEQB_HD bool fm_step(DevIndex const & ix, const uint32_t * ctab, uint32_t & lb, uint32_t & rb, uint32_t c)
{
    uint32_t nl, nr;   // nr exclusive
    const uint32_t cc = ctab[c];
    if (lb == 0u && rb + 1u == ix.n)
    {
        nl = cc;
        nr = ctab[kCtabEnd + c];
    }
    else if (c == kCodeN)
    {
        nl = cc + lower_bound_u32(ix.n_pos, ix.n_npos, lb);
        nr = cc + lower_bound_u32(ix.n_pos, ix.n_npos, rb + 1u);
    }
    else
    {
        uint32_t bl = lb >> 6, br = (rb + 1u) >> 6;
        RankBlock B = load_block(ix.blocks + bl);
        nl = cc + rank_in(ix, B, lb, c);
        if (br != bl) B = load_block(ix.blocks + br);
        nr = cc + rank_in(ix, B, rb + 1u, c);
    }
    if (nr <= nl) return false;
    lb = nl;
    rb = nr - 1u;
    return true;
}

// high bit of each byte set iff that byte < dd (bytes <= 127, 1 <= dd <= 127)
SPEQB_HD uint64_t lcp_lt_mask(uint64_t w, uint32_t dd)
{
    const uint64_t H = 0x8080808080808080ull, L = 0x0101010101010101ull;
    return ~((w | H) - (uint64_t)dd * L) & H;
}

// Shorten the pattern of interval [lb, rb] to depth dd (drop query symbols from the LEFT of the window): the result is
// the maximal range around [lb, rb] whose inner lcp values are all >= dd.
SPEQB_HD void fm_contract(DevIndex const & ix, uint32_t & lb, uint32_t & rb, uint32_t dd)
{
    if (dd == 0u)
    {
        lb = 0u;
        rb = ix.n - 1u;
        return;
    }
    const uint64_t * lw = reinterpret_cast<const uint64_t *>(ix.lcp);
    uint32_t wl = lb >> 3, wr = (rb + 1u) >> 3;
    uint64_t first = load_ro(lw + wl);
    uint64_t second = (wr == wl) ? first : load_ro(lw + wr);
    // left boundary: largest i <= lb with lcp[i] < dd (lcp[0] = 0 stops the scan)
    uint64_t m = lcp_lt_mask(first, dd) & (~0ull >> (8u * (7u - (lb & 7u))));
    while (m == 0ull)
    {
        --wl;
        m = lcp_lt_mask(load_ro(lw + wl), dd);
    }
    lb = wl * 8u + ((63u - clz64(m)) >> 3);
    // right boundary: smallest j > rb with lcp[j] < dd (zero padding behind n stops the scan)
    m = lcp_lt_mask(second, dd) & (~0ull << (8u * ((rb + 1u) & 7u)));
    while (m == 0ull)
    {
        ++wr;
        m = lcp_lt_mask(load_ro(lw + wr), dd);
    }
    rb = wr * 8u + (ctz64(m) >> 3) - 1u;
}

// ---- rolling window --------------------------------------------------------------------------------------------
struct RollState
{
    uint32_t lb, rb;      // SA interval of the current pattern
    uint32_t d;           // its length (<= k)
    uint32_t run;         // consecutive searchable positions ending at the previous position (window validity)
    uint32_t retry;       // last extension failed: contract before trying the same position again
};

SPEQB_HD void roll_reset(DevIndex const & ix, RollState & s)
{
    s.lb = 0u;
    s.rb = ix.n - 1u;
    s.d = 0u;
    s.run = 0u;
    s.retry = 0u;
}

// Result of one roll_advance call
enum : uint32_t
{
    ROLL_CONSUMED = 1u,   // the position was consumed (advance t)
    ROLL_WINDOW = 2u,     // a valid window [t-k+1, t] ends at this position (counts towards total_kmers)
    ROLL_HIT = 4u         // ... and it occurs in the text: s.lb, s.rb is its SA interval
};

// Feed read position t (query code c, `inv` = position cannot be part of a counted window and resets the search).
// One call performs at most one contraction and one step, so lanes of a warp stay convergent; a failed extension
// returns 0 and must be called again with the same position.
SPEQB_HD uint32_t roll_advance(DevIndex const & ix, const uint32_t * ctab, uint32_t k, RollState & s, uint32_t c, bool inv)
{
    if (inv)
    {
        roll_reset(ix, s);
        return ROLL_CONSUMED;
    }
    if (s.retry || s.d == k)
    {
        s.d -= 1u;
        fm_contract(ix, s.lb, s.rb, s.d);
        s.retry = 0u;
    }
    bool ok = fm_step(ix, ctab, s.lb, s.rb, c);
    if (ok) s.d += 1u;
    else if (s.d != 0u)
    {
        s.retry = 1u;
        return 0u;
    }
    s.run += 1u;
    uint32_t r = ROLL_CONSUMED;
    if (s.run >= k)
    {
        r |= ROLL_WINDOW;
        if (s.d == k) r |= ROLL_HIT;
    }
    return r;
}

// ---- packed read access ----------------------------------------------------------------------------------------
SPEQB_HD uint32_t code_at(const uint64_t * bases, uint64_t g0, uint32_t t)
{
    return (uint32_t)(load_ro(bases + g0 + (t >> 5)) >> (2u * (t & 31u))) & 3u;
}
SPEQB_HD uint32_t inv_at(const uint32_t * inv, uint64_t g0, uint32_t t)
{
    return (load_ro(inv + g0 + (t >> 5)) >> (t & 31u)) & 1u;
}

// ---- direct window search (reference trip counts) ------------------------------------------------------------
// k steps from the root for window [s, s+k) of a read; N positions (inv bit, batch packed with n_only) are searched
// as the text symbol N exactly like the reference does.  *steps += backward_search calls executed, including the
// failing one (fm_index_cursor.hpp:316-324).  Returns true and the interval when the window occurs.
SPEQB_HD bool direct_window(DevIndex const & ix, const uint32_t * ctab, const uint64_t * bases, const uint32_t * inv, uint64_t g0, uint32_t s,
                            uint32_t k, uint32_t & lb, uint32_t & rb, uint32_t & steps)
{
    lb = 0u;
    rb = ix.n - 1u;
    for (uint32_t i = 0; i < k; ++i)
    {
        uint32_t t = s + i;
        uint32_t c = inv_at(inv, g0, t) ? kCodeN : code_at(bases, g0, t);
        ++steps;
        if (!fm_step(ix, ctab, lb, rb, c)) return false;
    }
    return true;
}
}   // namespace speqb
