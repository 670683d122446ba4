// scan32_core.h -- per-lane arithmetic of the product scan kernel (scan32.cu): the rolling k-mer window over the
// 32-byte SBlocks (fm_layout.h), as a flat state machine in which every iteration costs one round trip to memory.
//
// __host__ __device__ like scan_core.h: scan32.cu runs it on the GPU; tests/sim compiles the same header with g++ as a
// white-box check of the state machine against the oracle in the CPU-only test tier (never shipped, never linked into
// libspeq_cuda.so).
//
// Reference arithmetic reproduced (S3 = deps/seqan3 of the reference checkout):
//   step        fm_index_cursor::backward_search, S3/include/seqan3/search/fm_index/fm_index_cursor.hpp:110-146
//   window rule ranges::views::sliding(k) (range-v3 view/sliding.hpp:129-142); valid iff no base of the window is N or
//               has phred <= cutoff (src/fm_scanner.cpp:162)
//   tallies     do_a_count, src/fm_scanner.cpp:153-196 (global), :426-471 (local), :665-708 / :916-962 (paired)
// State of a lane: the SA interval [lb, rb1) of the longest suffix of read[0..t) that occurs in the text, capped at k
// symbols (depth d).  The SA interval of a pattern is unique, so the interval reported for a window is bit-identical to
// the one k backward-search steps from the root produce.
//   P_STEP   loads the SBlocks at lb and rb1.  Resolves the previous window's hit (uq bit at lb -> group load), slides
//            the window (d == k: contract to depth k-1 with the e bits), and extends by the next base (rank step).
//   P_UP     parent jump (failed extension, or the e-bit scan left the two blocks): loads lcp + previous/next-smaller
//            links at lb and rb1 and moves to the enclosing lcp interval.
//   P_FETCH  next unit (read or pair): loads its length and first 32-base group.
#pragma once
#include "scan_core.h"

namespace speqb
{
enum : uint32_t { P_FETCH = 0, P_STEP = 1, P_UP = 2, P_DONE = 3 };
enum : uint32_t { F_PENDING = 1u, F_HAVE_G = 2u, F_MATE2 = 4u };

struct Lane32
{
    uint32_t phase, flags;
    uint32_t lb, rb1, d, run;     // interval [lb, rb1), depth, length of the current run of valid bases
    uint32_t t, pend;             // next base, end of the read (0 when the read is shorter than k)
    uint32_t wlo, whi, winv;      // the current 32-base group, shifted so that base t is in the low bits
    uint64_t g0;                  // first 32-base group of the read
    uint64_t u, unext;            // unit being scanned / claimed for afterwards
    uint32_t gval, gt;            // group of the last unique hit and the window end it belongs to
    uint32_t gmin, gmax;          // unit-level ambiguity state: smallest / largest variant with a unique k-mer in the unit,
    uint32_t gmore;               //   whether further variants lie between them (Cfg32::mask64 == 0),
    uint64_t gset;                //   or one bit per variant (Cfg32::mask64 != 0: at most 64 variants)
    uint32_t agg_g, agg_n;        // run-length aggregation of the lane's unique tallies
    double agg_w;
    uint32_t c_total;             // valid windows of the current unit
    // local-error mode: the phred weight of the window that ends at the cursor, as a rolling product (times 1/den[q] for the
    // base that enters, times den[q] for the base that leaves, 1 behind an invalid base; see QRoll in scan_dev.cuh for the
    // error bound), the weights of the hit being resolved / taken, and the 8 quality bytes around base t and base t - k
    double P, pw0, pw1;
    uint64_t qn, qo;
};

// What one iteration loads.  A lane is in exactly one phase, so the phases share registers:
//   P_STEP   A, B = SBlocks at lb / rb1   (the read group goes straight into Lane32::wlo / whi / winv)
//   P_UP     A[0..3] = UpRec of lb, B[0..3] = UpRec of rb1 ({psv, nsv, lcp, -})
//   P_FETCH  B[4] = length of the read, B[5], B[6] = its first 32-base group (low, high word)
struct Loads32
{
    uint32_t A[8], B[8];
};

struct Cfg32
{
    uint32_t k, n;
    uint64_t units;
    int paired, local, record;
    int mask64;                   // the unit's variants are kept as a bit mask (V <= 64), else as (smallest, largest, more)
    int skip_unique;              // record only windows whose k-mer is held by several variants: the others add exactly 1 to
                                  // their variant's refinement estimate (src/fm_scanner.cpp:1099-1119), i.e. the unique tally
};

SPEQB_HD void lane32_init(Lane32 & s, uint64_t first_unit, uint32_t n)
{
    s.phase = P_FETCH; s.flags = 0; s.lb = 0; s.rb1 = n; s.d = 0; s.run = 0; s.t = 0; s.pend = 0; s.wlo = s.whi = s.winv = 0;
    s.g0 = 0; s.u = first_unit; s.unext = first_unit; s.gval = 0; s.gt = 0; s.gmin = kNoGroup; s.gmax = 0; s.gmore = 0; s.gset = 0;
    s.agg_g = kNoGroup; s.agg_n = 0; s.agg_w = 0.0; s.c_total = 0;
    s.P = 1.0; s.pw0 = s.pw1 = 0.0; s.qn = s.qo = 0;
}

// Env supplies the side effects:
//   uint32_t env.load_grp(i)               group of suffix i
//   uint32_t env.exc_word(blk)             exception mask of SBlock blk (slow path)
//   uint64_t env.qual8(g0, pos)            local mode: the 8 phred bytes of the read at positions [pos, pos + 8), pos % 8 == 0
//   uint32_t env.qual1(g0, pos)            ... one phred byte (periodic exact restart of the rolling weight)
//   double   env.den(q) / env.rden(q)      1 - 10^(-q/10) and its reciprocal (host libm tables)
//   void     env.unique(group, n, w)       add n (global) or w (local) to the variant's tally
//   void     env.record(lb)                one more valid window whose k-mer interval starts at lb
//   void     env.count_fetch(nwin)         a read with nwin windows was presented
//   void     env.count_unit(total, amb)    a unit ended with `total` valid windows; amb: it was ambiguous
//   void     env.fusion(key)               an ambiguous unit ended: key of the set of its unique variants (kernels.h)
template <class Env>
SPEQB_HD void lane32_flush_agg(Lane32 & s, Env & env)
{
    if (s.agg_g != kNoGroup) env.unique(s.agg_g, s.agg_n, s.agg_w);
    s.agg_g = kNoGroup;
    s.agg_n = 0;
    s.agg_w = 0.0;
}

template <class Env>
SPEQB_HD void lane32_take_group(Lane32 & s, Cfg32 const & cfg, Env & env)
{
    // ++unique_kmers[which_hit] and the read-level ambiguity state (src/fm_scanner.cpp:178-191)
    const uint32_t g = s.gval;
    s.flags &= ~F_HAVE_G;
    if (cfg.mask64) s.gset |= 1ull << g;
    else
    {
        if (s.gmin == kNoGroup) { s.gmin = g; s.gmax = g; }
        else if (g < s.gmin) { s.gmore |= s.gmin != s.gmax ? 1u : 0u; s.gmin = g; }
        else if (g > s.gmax) { s.gmore |= s.gmin != s.gmax ? 1u : 0u; s.gmax = g; }
        else s.gmore |= (g != s.gmin && g != s.gmax) ? 1u : 0u;
    }
    if (g != s.agg_g)
    {
        lane32_flush_agg(s, env);
        s.agg_g = g;
    }
    if (cfg.local) s.agg_w += s.pw1;
    s.agg_n += 1u;
}

#if defined(__CUDA_ARCH__)
SPEQB_HD uint32_t popc32(uint32_t x) { return (uint32_t)__popc(x); }
#else
SPEQB_HD uint32_t popc32(uint32_t x) { return (uint32_t)__builtin_popcount(x); }
#endif
SPEQB_HD uint32_t below32(uint32_t j) { return (1u << j) - 1u; }   // bits [0, j), j in [0, 31]
SPEQB_HD uint32_t upto32(uint32_t j) { return (2u << j) - 1u; }    // bits [0, j], j in [0, 31]
SPEQB_HD uint32_t from32(uint32_t j) { return ~0u << j; }          // bits [j, 31]

// index of the read a lane in P_FETCH loads
SPEQB_HD uint64_t lane32_fetch_read(Lane32 const & s, Cfg32 const & cfg)
{
    return cfg.paired ? 2 * s.u + ((s.flags & F_MATE2) ? 1u : 0u) : s.u;
}

SPEQB_HD void lane32_advance(Lane32 & s)
{
    s.t += 1u;
    s.wlo = (s.wlo >> 2) | (s.whi << 30);
    s.whi >>= 2;
    s.winv >>= 1;
}

// Loads of a P_STEP iteration in local mode: the quality bytes around the base about to be consumed (t) and around the base
// that leaves the window with it (t - k), one aligned 8-byte load per stream every 8 bases.  Idempotent: an iteration that
// does not consume its base (parent jump, failed extension) loads the same bytes again.
template <class Env>
SPEQB_HD void lane32_load_qual(Lane32 & s, Cfg32 const & cfg, Env & env)
{
    if (!cfg.local || s.t >= s.pend) return;
    if ((s.t & 7u) == 0u) s.qn = env.qual8(s.g0, s.t);
    if (s.t >= cfg.k && ((s.t - cfg.k) & 7u) == 0u) s.qo = env.qual8(s.g0, s.t - cfg.k);
}

// P_STEP.  ld.A / ld.B: the SBlocks at lb / rb1.  The read group of base t has been loaded into s.wlo / s.whi / s.winv by
// the caller when t is a multiple of 32.  Written select-style so that the common case is one straight run of code.
template <class Env>
SPEQB_HD void lane32_step(Lane32 & s, Loads32 const & ld, Cfg32 const & cfg, Env & env)
{
    const uint32_t k = cfg.k;
    if (s.flags & F_HAVE_G) lane32_take_group(s, cfg, env);
    if (s.flags & F_PENDING)
    {
        // every occurrence of the window's k-mer in one variant <=> uq bit of any suffix of its interval
        s.flags &= ~F_PENDING;
        if ((ld.A[7] >> (s.lb & 31u)) & 1u)
        {
            s.gval = env.load_grp(s.lb);
            s.gt = s.t;
            s.pw1 = s.pw0;
            s.flags |= F_HAVE_G;
        }
        else if (cfg.record && cfg.skip_unique) env.record(s.lb);
    }
    if (s.t >= s.pend)
    {
        // the read is finished (its last hit, if any, was resolved above)
        if (s.flags & F_HAVE_G) lane32_take_group(s, cfg, env);
        if (cfg.paired && !(s.flags & F_MATE2)) s.flags |= F_MATE2;
        else
        {
            // is_ambiguous: unique k-mers of two different variants in one read / pair (src/fm_scanner.cpp:183-190,213)
            const bool amb = cfg.mask64 ? (s.gset & (s.gset - 1ull)) != 0ull : (s.gmin != kNoGroup && s.gmin != s.gmax);
            env.count_unit(s.c_total, amb);
            if (amb) env.fusion(cfg.mask64 ? s.gset : fusion_key_minmax(s.gmin, s.gmax, s.gmore != 0u));
            s.c_total = 0u;
            s.gmin = kNoGroup; s.gmax = 0u; s.gmore = 0u; s.gset = 0ull;
            s.flags &= ~F_MATE2;
            s.u = s.unext;
        }
        s.phase = P_FETCH;
        return;
    }
    const uint32_t c = s.wlo & 3u;
    const bool c1 = (c & 1u) != 0u, c2 = (c & 2u) != 0u;
    // N / low quality: no window may contain this base; the search restarts behind it (root interval, run = 0)
    const bool inv = (s.winv & 1u) != 0u;
    // the window slides (d == k): drop its first base = contract the interval to depth k-1 with the e bits
    const bool slide = s.d == k;
    uint32_t jl = s.lb & 31u, jr = s.rb1 & 31u;
    const uint32_t zl = ~ld.A[6] & upto32(jl);
    const uint32_t zr = ~ld.B[6] & from32(jr);
    if (slide && !inv && (zl == 0u || zr == 0u))
    {
        // the widened interval leaves the two blocks: it is the parent lcp interval (depth k-1); jump there
        // (k == 1: every e bit is set, so each base takes this jump to the root)
        s.d = k - 1u;
        s.phase = P_UP;
        return;
    }
    if (slide)
    {
        jl = 31u - clz32(zl | 1u);            // (| 1, | 2^31 only matter on an invalid base, whose values are unused)
        jr = ctz32(zr | 0x80000000u);
    }
    const uint32_t lbc = (s.lb & ~31u) + jl, rbc = (s.rb1 & ~31u) + jr;
    const uint32_t dd = slide ? k - 1u : s.d;
    // rank step (fm_index_cursor.hpp:110-146); the blocks' counts include C[c], and the root interval [0, n) needs no
    // special case: rank(0, c) = 0 and rank(n, c) = #c give [C[c], C[c+1])
    uint32_t ml = below32(jl), mr = below32(jr);
    uint32_t a0 = ld.A[0], b0 = ld.B[0];
    if ((a0 | b0) & kSFlagExc)
    {
        // N / delimiter / sentinel in the BWT of one of the blocks: they match no query symbol
        if (a0 & kSFlagExc) ml &= ~env.exc_word(s.lb >> 5);
        if (b0 & kSFlagExc) mr &= ~env.exc_word(s.rb1 >> 5);
        a0 &= ~kSFlagExc;
        b0 &= ~kSFlagExc;
    }
    const uint32_t xl = c1 ? 0u : ~0u, xh = c2 ? 0u : ~0u;
    const uint32_t ca = c2 ? (c1 ? ld.A[3] : ld.A[2]) : (c1 ? ld.A[1] : a0);
    const uint32_t cb = c2 ? (c1 ? ld.B[3] : ld.B[2]) : (c1 ? ld.B[1] : b0);
    const uint32_t nl = ca + popc32((ld.A[4] ^ xl) & (ld.A[5] ^ xh) & ml);
    const uint32_t nr = cb + popc32((ld.B[4] ^ xl) & (ld.B[5] ^ xh) & mr);   // exclusive
    const bool ok = nr > nl;
    if (!ok && !inv && dd != 0u)
    {
        // read[t-d..t] does not occur: shorten it from the left (to the parent interval) and try the same base again
        if (dd == 1u) { s.lb = 0u; s.rb1 = cfg.n; s.d = 0u; }
        else { s.lb = lbc; s.rb1 = rbc; s.d = dd; s.phase = P_UP; }
        return;
    }
    const bool keep = ok && !inv;
    s.lb = keep ? nl : 0u;
    s.rb1 = keep ? nr : cfg.n;
    s.d = keep ? dd + 1u : 0u;
    s.run = inv ? 0u : s.run + 1u;
    if (cfg.local)
    {
        // the base at position t enters the window; once more than k valid bases are behind the cursor, base t - k leaves
        if (inv) s.P = 1.0;
        else
        {
            s.P *= env.rden((uint32_t)(s.qn >> (8u * (s.t & 7u))) & 0xFFu);
            if (s.run > k) s.P *= env.den((uint32_t)(s.qo >> (8u * ((s.t - k) & 7u))) & 0xFFu);
            if ((s.t & 255u) == 255u && s.run >= k)
            {
                double e = 1.0;                                       // periodic exact restart (long reads)
                for (uint32_t i = s.t + 1u - k; i <= s.t; ++i) e *= env.rden(env.qual1(s.g0, i));
                s.P = e;
            }
        }
    }
    lane32_advance(s);
    if (s.run >= k)
    {
        s.c_total += 1u;                         // ++total_kmers (src/fm_scanner.cpp:164)
        if (s.d == k)
        {
            s.flags |= F_PENDING;
            s.pw0 = s.P;
            if (cfg.record && !cfg.skip_unique) env.record(s.lb);
        }
    }
}

// P_UP: parent of the lcp interval [lb, rb1): string depth l = max(lcp[lb], lcp[rb1]); the side(s) whose lcp equals l move
// to the previous / next position with a smaller lcp value.
SPEQB_HD void lane32_up(Lane32 & s, Loads32 const & ld, Cfg32 const & cfg)
{
    const uint32_t ul = ld.A[2], ur = ld.B[2];
    const uint32_t l = ul > ur ? ul : ur;
    if (l == 0u) { s.lb = 0u; s.rb1 = cfg.n; }
    else
    {
        if (ul == l) s.lb = ld.A[0];
        if (ur == l) s.rb1 = ld.B[1];
    }
    s.d = l;
    s.phase = P_STEP;
}

// P_FETCH.  The caller has set s.unext (the unit after s.u) when this is the first mate of the unit.
template <class Env>
SPEQB_HD void lane32_fetch(Lane32 & s, Loads32 const & ld, Cfg32 const & cfg, Env & env)
{
    if (s.u >= cfg.units)
    {
        s.phase = P_DONE;
        return;
    }
    const uint32_t L = ld.B[4];
    s.g0 = u64of(ld.B[5], ld.B[6]);
    s.lb = 0u; s.rb1 = cfg.n; s.d = 0u; s.run = 0u; s.t = 0u; s.P = 1.0;
    s.pend = L >= cfg.k ? L : 0u;                  // L < k: the sliding view is empty
    env.count_fetch(L >= cfg.k ? L - cfg.k + 1u : 0u);
    s.phase = P_STEP;
}
}   // namespace speqb
