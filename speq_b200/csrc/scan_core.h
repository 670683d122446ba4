// scan_core.h -- per-lane arithmetic of the scan hot path: rank on the blocked BWT, one backward-search step, the
// LCP contraction, and the rolling k-mer window as a flat, one-round-trip-per-iteration state machine.
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
SPEQB_HD uint32_t popc64(uint64_t x) { return (uint32_t)__popcll(x); }
SPEQB_HD uint32_t clz64(uint64_t x) { return (uint32_t)__clzll((long long)x); }
SPEQB_HD uint32_t ctz64(uint64_t x) { return (uint32_t)__ffsll((long long)x) - 1u; }
SPEQB_HD uint32_t clz32(uint32_t x) { return (uint32_t)__clz((int)x); }
SPEQB_HD uint32_t ctz32(uint32_t x) { return (uint32_t)__ffs((int)x) - 1u; }
template <class T> SPEQB_HD T load_ro(const T * p) { return __ldg(p); }
#else
SPEQB_HD uint32_t popc64(uint64_t x) { return (uint32_t)__builtin_popcountll(x); }
SPEQB_HD uint32_t clz64(uint64_t x) { return (uint32_t)__builtin_clzll(x); }
SPEQB_HD uint32_t ctz64(uint64_t x) { return (uint32_t)__builtin_ctzll(x); }
SPEQB_HD uint32_t clz32(uint32_t x) { return (uint32_t)__builtin_clz(x); }
SPEQB_HD uint32_t ctz32(uint32_t x) { return (uint32_t)__builtin_ctz(x); }
template <class T> SPEQB_HD T load_ro(const T * p) { return *p; }
#endif

SPEQB_HD uint64_t u64of(uint32_t lo, uint32_t hi) { return (uint64_t)lo | ((uint64_t)hi << 32); }
SPEQB_HD uint64_t below64(uint32_t j) { return (1ull << j) - 1ull; }   // bits [0, j), j in [0, 63]
SPEQB_HD uint64_t upto64(uint32_t j) { return (2ull << j) - 1ull; }    // bits [0, j], j in [0, 63]
SPEQB_HD uint64_t from64(uint32_t j) { return ~0ull << j; }            // bits [j, 63]

// One KBlock in registers: w[0..3] cnt, w[4..5] lo, w[6..7] hi, w[8..9] e, w[10..11] chg, w[12..13] exc, w[14] chgcum.
// (Indexed with compile-time constants only, so it stays in registers.)
struct BlkRegs
{
    uint32_t w[16];
};

SPEQB_HD BlkRegs load_kblock(const KBlock * p)
{
    BlkRegs b;
#if defined(__CUDA_ARCH__)
    asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(b.w[0]), "=r"(b.w[1]), "=r"(b.w[2]), "=r"(b.w[3]), "=r"(b.w[4]), "=r"(b.w[5]), "=r"(b.w[6]), "=r"(b.w[7])
                 : "l"(p));
    asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(b.w[8]), "=r"(b.w[9]), "=r"(b.w[10]), "=r"(b.w[11]), "=r"(b.w[12]), "=r"(b.w[13]), "=r"(b.w[14]), "=r"(b.w[15])
                 : "l"(reinterpret_cast<const char *>(p) + 32));
#else
    const uint32_t * s = reinterpret_cast<const uint32_t *>(p);
    for (int i = 0; i < 16; ++i) b.w[i] = s[i];
#endif
    return b;
}

// occurrences of query code c (A0 C1 G2 T3) in bwt[0, 64*block + j), j in [0, 63]
SPEQB_HD uint32_t blk_rank(BlkRegs const & b, uint32_t c, uint32_t j)
{
    uint64_t lo = u64of(b.w[4], b.w[5]), hi = u64of(b.w[6], b.w[7]), exc = u64of(b.w[12], b.w[13]);
    uint64_t xl = (c & 1u) ? 0ull : ~0ull, xh = (c & 2u) ? 0ull : ~0ull;
    uint64_t m = (lo ^ xl) & (hi ^ xh) & ~exc & below64(j);
    uint32_t base = c == 0u ? b.w[0] : (c == 1u ? b.w[1] : (c == 2u ? b.w[2] : b.w[3]));
    return base + popc64(m);
}
// number of group changes in [0, 64*block + j), j in [0, 63]
SPEQB_HD uint32_t blk_chg_before(BlkRegs const & b, uint32_t j) { return b.w[14] + popc64(u64of(b.w[10], b.w[11]) & below64(j)); }
// number of group changes in [0, 64*block + j], j in [0, 63]
SPEQB_HD uint32_t blk_chg_upto(BlkRegs const & b, uint32_t j) { return b.w[14] + popc64(u64of(b.w[10], b.w[11]) & upto64(j)); }

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

// One backward-search step with query code c (0..3, or kCodeN), loading its blocks itself (direct search, host tools).
// Returns false (interval untouched) when the extended pattern does not occur.
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
        BlkRegs B = load_kblock(ix.blocks + bl);
        nl = cc + blk_rank(B, c, lb & 63u);
        if (br != bl) B = load_kblock(ix.blocks + br);
        nr = cc + blk_rank(B, c, (rb + 1u) & 63u);
    }
    if (nr <= nl) return false;
    lb = nl;
    rb = nr - 1u;
    return true;
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

// "every occurrence in [lb, rb] belongs to one variant" + that variant (kNoGroup when not)
SPEQB_HD uint32_t interval_group(DevIndex const & ix, uint32_t lb, uint32_t rb)
{
    BlkRegs A = load_kblock(ix.blocks + (lb >> 6));
    uint32_t cl = blk_chg_upto(A, lb & 63u);
    BlkRegs B = load_kblock(ix.blocks + ((rb + 1u) >> 6));
    uint32_t cr = blk_chg_before(B, (rb + 1u) & 63u);
    return cl == cr ? (uint32_t)load_ro(ix.grp + lb) : kNoGroup;
}

// ================================================================================================================
// Rolling window as a flat state machine.  One call of lane_plan + (loads) + lane_consume = one ITERATION; every
// iteration needs at most ONE round trip to memory, whatever the lane is doing, so the 32 lanes of a warp run it in
// lock step (the kernel re-converges them with __syncwarp at the top of the loop):
//   PH_FETCH  pick the next unit (read, pair or text segment): loads its length / first group
//   PH_STEP   loads the two KBlocks at lb and rb+1; resolves the uniqueness of the previous window's hit; contracts
//             d == k to k-1 with the e bits; performs the rank step for the next base
//   PH_UP     parent jump (after a failed extension, or when the e-bit scan leaves the two blocks): loads lcp and the
//             previous/next-smaller-value links at lb and rb+1 and moves to the enclosing lcp interval in one iteration.
//             Shortening the pattern to any depth above the parent's leaves the interval -- and therefore the outcome of
//             the failed extension -- unchanged, so jumping straight to the parent's depth max(lcp[lb], lcp[rb+1]) visits
//             exactly the intervals that matter and ends in the same state as shortening one symbol at a time.
// ================================================================================================================
enum : uint32_t { PH_FETCH = 0, PH_STEP = 1, PH_UP = 2, PH_DONE = 3 };
enum : uint32_t
{
    LF_PENDING = 1u,   // the last step produced a hit whose group uniqueness is not yet known
    LF_HAVE_G = 2u,    // a group load is in flight / waiting in `gval`
    LF_MATE2 = 4u      // paired: the second mate of the unit is being scanned
};

struct Lane
{
    uint32_t phase, flags;
    uint32_t lb, rb, d, run;
    uint32_t t, pend;
    uint32_t cw, winv;
    uint64_t wbase, g0;
    uint64_t u;
    uint32_t rgroup;        // back-check: variant of the text being scanned
    uint32_t gval, gt;      // group of the last unique hit (load result) and the window end it belongs to
    uint32_t gmin, gmax;    // unit-level ambiguity state
    uint32_t agg_g, agg_n;  // run-length aggregation of this lane's unique tallies
    double agg_w;
    unsigned long long c_total, c_look;
    uint32_t c_amb, c_reads;
};

struct Plan
{
    bool blocks, word, up, fetch;
    const KBlock * pa;
    const KBlock * pb;
    const uint64_t * pw;
    const uint32_t * pi;
    uint64_t fr;   // read (or segment) index to fetch
};

struct Loads
{
    BlkRegs A, B;           // KBlocks at lb / rb+1; in PH_UP: A.w[0] = lcp[lb], A.w[2] = psv[lb]; B.w[0] = lcp[rb+1], B.w[3] = nsv[rb+1]
    uint64_t rdw;
    uint32_t rdi;
    uint32_t flen, fstart, fgroup;
    uint64_t fg0;
};

struct ScanCfg
{
    uint32_t k;
    uint64_t units, stride;   // units to process, distance between two units of one lane
    int paired, local, record, back;
};

SPEQB_HD void lane_init(Lane & s, uint64_t first_unit)
{
    s.phase = PH_FETCH; s.flags = 0; s.lb = 0; s.rb = 0; s.d = 0; s.run = 0; s.t = 0; s.pend = 0; s.cw = 0xFFFFFFFFu; s.winv = 0;
    s.wbase = 0; s.g0 = 0; s.u = first_unit; s.rgroup = 0; s.gval = 0; s.gt = 0;
    s.gmin = kNoGroup; s.gmax = 0; s.agg_g = kNoGroup; s.agg_n = 0; s.agg_w = 0.0; s.c_total = s.c_look = 0; s.c_amb = s.c_reads = 0;
}

SPEQB_HD void lane_plan(Lane const & s, DevIndex const & ix, BatchView const & b, ScanCfg const & cfg, Plan & p)
{
    p.blocks = p.word = p.up = p.fetch = false;
    p.pa = p.pb = ix.blocks; p.pw = b.bases; p.pi = b.inv; p.fr = 0;
    if (s.phase == PH_STEP)
    {
        bool stepping = s.t < s.pend;
        p.blocks = (s.flags & LF_PENDING) || (stepping && s.d != 0u);
        p.pa = ix.blocks + (s.lb >> 6);
        p.pb = ix.blocks + ((s.rb + 1u) >> 6);
        p.word = stepping && (s.t >> 5) != s.cw;
        p.pw = b.bases + s.g0 + (s.t >> 5);
        p.pi = b.inv + s.g0 + (s.t >> 5);
    }
    else if (s.phase == PH_UP)
        p.up = true;   // lcp[lb], lcp[rb+1], pn[lb], pn[rb+1]: addressed from s.lb / s.rb by the caller
    else if (s.phase == PH_FETCH)
    {
        p.fetch = s.u < cfg.units;
        p.fr = cfg.back ? s.u : (cfg.paired ? 2 * s.u + ((s.flags & LF_MATE2) ? 1u : 0u) : s.u);
    }
}

// Env supplies the side effects: group / quality loads, tallies, multiplicity recording.
//   uint32_t env.load_grp(i)                      group of suffix i
//   uint32_t env.npos_rank(i)                     # N in bwt[0, i)
//   double   env.weight(g0, t_end, k)             local-mode weight of the window ending (exclusive) at t_end
//   void     env.unique(group, n, w)              add n (global / back-check) or w (local) to the variant's tally
//   void     env.record(lb)                       one more valid window whose k-mer interval starts at lb
//   void     env.backtotal(group, nwin)           back-check: windows presented for a variant
template <class Env>
SPEQB_HD void lane_flush_agg(Lane & s, ScanCfg const & cfg, Env & env)
{
    if (s.agg_g != kNoGroup) env.unique(s.agg_g, s.agg_n, s.agg_w);
    s.agg_g = kNoGroup;
    s.agg_n = 0;
    s.agg_w = 0.0;
}

template <class Env>
SPEQB_HD void lane_take_group(Lane & s, ScanCfg const & cfg, Env & env)
{
    // consume the group of the last unique hit (src/fm_scanner.cpp:178-191: ++unique_kmers[which_hit]; ambiguity state)
    uint32_t g = s.gval;
    s.flags &= ~LF_HAVE_G;
    s.gmin = g < s.gmin ? g : s.gmin;
    s.gmax = g > s.gmax ? g : s.gmax;
    if (g != s.agg_g)
    {
        lane_flush_agg(s, cfg, env);
        s.agg_g = g;
    }
    if (cfg.local) s.agg_w += env.weight(s.g0, s.gt, cfg.k);
    else s.agg_n += 1u;
}

template <class Env>
SPEQB_HD void lane_consume(Lane & s, Loads const & ld, DevIndex const & ix, const uint32_t * ctab, ScanCfg const & cfg, Env & env)
{
    const uint32_t k = cfg.k;
    if (s.phase == PH_STEP)
    {
        if (s.flags & LF_HAVE_G) lane_take_group(s, cfg, env);
        if (s.flags & LF_PENDING)
        {
            // every occurrence in one variant <=> no group change in (lb, rb]
            s.flags &= ~LF_PENDING;
            if (blk_chg_upto(ld.A, s.lb & 63u) == blk_chg_before(ld.B, (s.rb + 1u) & 63u))
            {
                s.gval = env.load_grp(s.lb);
                s.gt = s.t;
                s.flags |= LF_HAVE_G;
            }
        }
        if (s.t < s.pend)
        {
            if ((s.t >> 5) != s.cw)
            {
                s.cw = s.t >> 5;
                s.wbase = ld.rdw;
                s.winv = ld.rdi;
            }
            uint32_t c = (uint32_t)(s.wbase >> (2u * (s.t & 31u))) & 3u;
            bool ib = (s.winv >> (s.t & 31u)) & 1u;
            if (cfg.back) { if (ib) c = kCodeN; ib = false; }
            if (ib)
            {
                // N / low quality: no window may contain this base; the search restarts behind it
                s.lb = 0u; s.rb = ix.n - 1u; s.d = 0u; s.run = 0u;
                s.t += 1u;
                return;
            }
            uint32_t jl = s.lb & 63u, jr = (s.rb + 1u) & 63u;
            if (s.d == k)
            {
                // the window slides: drop its first base = contract the interval to depth k-1
                if (k == 1u) { s.lb = 0u; s.rb = ix.n - 1u; s.d = 0u; }
                else
                {
                    uint64_t zl = ~u64of(ld.A.w[8], ld.A.w[9]) & upto64(jl);
                    uint64_t zr = ~u64of(ld.B.w[8], ld.B.w[9]) & from64(jr);
                    s.d = k - 1u;
                    if (zl == 0ull || zr == 0ull)
                    {
                        // the widened interval leaves the two blocks: it is the parent lcp interval (depth k-1); jump there
                        s.phase = PH_UP;
                        return;
                    }
                    jl = 63u - clz64(zl);
                    jr = ctz64(zr);
                    s.lb = (s.lb & ~63u) + jl;
                    s.rb = ((s.rb + 1u) & ~63u) + jr - 1u;
                }
            }
            uint32_t nl, nr;
            const uint32_t cc = ctab[c];
            if (s.d == 0u) { nl = cc; nr = ctab[kCtabEnd + c]; }
            else if (c == kCodeN) { nl = cc + env.npos_rank(s.lb); nr = cc + env.npos_rank(s.rb + 1u); }
            else { nl = cc + blk_rank(ld.A, c, jl); nr = cc + blk_rank(ld.B, c, jr); }
            if (nr > nl)
            {
                s.lb = nl; s.rb = nr - 1u; s.d += 1u;
            }
            else if (s.d != 0u)
            {
                // read[t-d..t] does not occur: shorten it from the left (to the parent interval) and try the same base again
                if (s.d == 1u) { s.lb = 0u; s.rb = ix.n - 1u; s.d = 0u; }
                else s.phase = PH_UP;
                return;
            }
            s.run += 1u;
            s.t += 1u;
            if (s.run >= k)
            {
                s.c_total += 1ull;                       // ++total_kmers (src/fm_scanner.cpp:164)
                if (s.d == k)
                {
                    s.flags |= LF_PENDING;
                    if (cfg.record) env.record(s.lb);
                }
            }
            return;
        }
        // the read is finished (its last hit, if any, was resolved above)
        if (s.flags & LF_HAVE_G) lane_take_group(s, cfg, env);
        if (!cfg.back && cfg.paired && !(s.flags & LF_MATE2)) s.flags |= LF_MATE2;
        else
        {
            if (!cfg.back && s.gmin != kNoGroup && s.gmin != s.gmax) s.c_amb += 1u;   // is_ambiguous (src/fm_scanner.cpp:183-190,213)
            s.gmin = kNoGroup; s.gmax = 0u;
            s.flags &= ~LF_MATE2;
            s.u += cfg.stride;
        }
        s.phase = PH_FETCH;
        return;
    }
    if (s.phase == PH_UP)
    {
        // parent of the lcp interval [lb, rb]: string depth l = max(lcp[lb], lcp[rb+1]); the side(s) whose lcp equals l move to
        // the previous / next position with a smaller lcp value
        uint32_t ll = ld.A.w[0], lr = ld.B.w[0];
        uint32_t l = ll > lr ? ll : lr;
        if (l == 0u) { s.lb = 0u; s.rb = ix.n - 1u; }
        else
        {
            if (ll == l) s.lb = ld.A.w[2];
            if (lr == l) s.rb = ld.B.w[3] - 1u;
        }
        s.d = l;
        s.phase = PH_STEP;
        return;
    }
    if (s.phase == PH_FETCH)
    {
        if (s.u >= cfg.units)
        {
            s.phase = PH_DONE;
            return;
        }
        uint32_t L = ld.flen;
        s.g0 = ld.fg0;
        s.cw = 0xFFFFFFFFu;
        s.lb = 0u; s.rb = ix.n - 1u; s.d = 0u; s.run = 0u;
        if (cfg.back)
        {
            // a segment of a text: windows [start, start + nwin)
            s.t = ld.fstart;
            s.pend = ld.fstart + L + k - 1u;
            s.rgroup = ld.fgroup;
            env.backtotal(ld.fgroup, L);
        }
        else
        {
            s.t = 0u;
            s.pend = L >= k ? L : 0u;                   // L < k: the sliding view is empty
            s.c_reads += 1u;
            if (L >= k) s.c_look += L - k + 1u;
        }
        s.phase = PH_STEP;
        return;
    }
}
}   // namespace speqb
