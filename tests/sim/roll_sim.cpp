// roll_sim.cpp -- TEST-ONLY white-box harness: runs the per-lane state machine of speq_b200/csrc/scan_core.h (the very
// header the CUDA kernels compile) on the host, lane by lane, so that index layout / rolling-window logic can be
// checked against the oracle in the CPU-only test tier.  It is never linked into libspeq_cuda.so and is not a product
// path: the product has no CPU fallback.  Built by tests/helpers.py with g++.
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

#include "../../speq_b200/csrc/scan32_core.h"
#include "../../speq_b200/csrc/speq_internal.h"

using namespace speqb;

namespace
{
struct HostView
{
    std::vector<KBlock> blocks_k;
    uint32_t ctab[16];
    DevIndex ix;
};

void make_view(HostIndex const & h, uint32_t k, HostView & hv)
{
    hv.blocks_k = h.blocks;
    for (uint64_t i = 0; i < h.n; ++i)
        if ((uint32_t)h.lcp[i] + 1u >= k) hv.blocks_k[i / 64].e[(i % 64) / 32] |= 1u << (i % 32);
    static const uint32_t sym_of_code[5] = {1, 2, 3, 5, 4};
    memset(hv.ctab, 0, sizeof hv.ctab);
    for (int c = 0; c < 5; ++c)
    {
        hv.ctab[c] = (uint32_t)h.C[sym_of_code[c]];
        hv.ctab[kCtabEnd + c] = (uint32_t)h.C[sym_of_code[c] + 1];
    }
    hv.ix.blocks = hv.blocks_k.data();
    hv.ix.lcp = h.lcp.data();
    hv.ix.pn = h.pn.data();
    hv.ix.grp = h.grp.data();
    hv.ix.n_pos = h.n_pos.data();
    hv.ix.ctab = hv.ctab;
    hv.ix.n_npos = (uint32_t)h.n_pos.size();
    hv.ix.n = (uint32_t)h.n;
}

// Host restatement of fill32_kernel (scan32.cu): the per-k 32-byte blocks of the product kernel.
uint32_t chg_upto_host(HostIndex const & h, uint32_t j)
{
    KBlock const & b = h.blocks[j >> 6];
    return b.chgcum + popc64(u64of(b.chg[0], b.chg[1]) & upto64(j & 63u));
}
void make_sblocks(HostIndex const & h, uint32_t k, const uint32_t * ctab, std::vector<SBlock> & out)
{
    const uint32_t n = (uint32_t)h.n;
    out.assign(n / 32 + 1, SBlock{});
    for (uint32_t blk = 0; blk < out.size(); ++blk)
    {
        KBlock const & p = h.blocks[blk >> 1];
        const uint32_t half = blk & 1u;
        SBlock & o = out[blk];
        for (uint32_t c = 0; c < 4; ++c)
        {
            o.cnt[c] = ctab[c] + p.cnt[c];
            const uint32_t xl = (c & 1u) ? 0u : ~0u, xh = (c & 2u) ? 0u : ~0u;
            if (half) o.cnt[c] += popc32((p.lo[0] ^ xl) & (p.hi[0] ^ xh) & ~p.exc[0]);
        }
        if (p.exc[half]) o.cnt[0] |= kSFlagExc;
        o.lo = p.lo[half];
        o.hi = p.hi[half];
        for (uint32_t j = 0; j < 32; ++j)
        {
            const uint32_t i = blk * 32 + j;
            if (i >= n) break;
            if ((uint32_t)h.lcp[i] + 1u >= k) o.e |= 1u << j;
            uint32_t s0 = i;
            while ((uint32_t)h.lcp[s0] >= k) s0 = h.pn[s0].x;
            uint32_t t0 = i + 1;
            while ((uint32_t)h.lcp[t0] >= k) t0 = h.pn[t0].y;
            if (chg_upto_host(h, t0 - 1) == chg_upto_host(h, s0)) o.uq |= 1u << j;
        }
    }
}

struct HostEnv32
{
    HostIndex const & h;
    const uint8_t * qual;
    const double * qden;
    std::vector<uint64_t> & uniq;
    std::vector<double> & uniq_w;
    uint64_t * mult;
    bool local;
    uint64_t & total, & amb, & look;
    uint32_t load_grp(uint32_t i) const { return h.grp[i]; }
    uint32_t exc_word(uint32_t blk) const { return h.blocks[blk >> 1].exc[blk & 1u]; }
    uint64_t qual8(uint64_t g0, uint32_t pos) const
    {
        uint64_t v;
        memcpy(&v, qual + 32u * g0 + pos, 8);
        return v;
    }
    uint32_t qual1(uint64_t g0, uint32_t pos) const { return qual[32u * g0 + pos]; }
    double den(uint32_t q) const { return qden[q]; }
    double rden(uint32_t q) const { return 1.0 / qden[q]; }
    void unique(uint32_t g, uint32_t n, double w) const
    {
        if (local) uniq_w[g] += w;
        else uniq[g] += n;
    }
    void record(uint32_t lb) const { if (mult) ++mult[lb]; }
    void count_fetch(uint32_t nwin) const { look += nwin; }
    void count_unit(uint32_t t, bool a) const { total += t; amb += a ? 1 : 0; }
    void fusion(uint64_t) const {}
};

struct HostEnv
{
    DevIndex const & ix;
    const uint8_t * qual;
    const double * qden;
    std::vector<uint64_t> & uniq;
    std::vector<double> & uniq_w;
    std::vector<uint64_t> & totg;
    uint64_t * mult;
    bool local;
    uint32_t load_grp(uint32_t i) const { return ix.grp[i]; }
    uint32_t npos_rank(uint32_t i) const { return lower_bound_u32(ix.n_pos, ix.n_npos, i); }
    double weight(uint64_t g0, uint32_t t_end, uint32_t k) const
    {
        const uint8_t * q = qual + 32u * g0 + (t_end - k);
        double w = 1.0;
        for (uint32_t i = 0; i < k; ++i) w = w / qden[q[i]];
        return w;
    }
    void unique(uint32_t g, uint32_t n, double w) const
    {
        if (local) uniq_w[g] += w;
        else uniq[g] += n;
    }
    void record(uint32_t lb) const { if (mult) ++mult[lb]; }
    void backtotal(uint32_t g, uint32_t n) const { totg[g] += n; }
};

void do_loads(Plan const & p, Loads & ld, Lane const & s, DevIndex const & ix, BatchView const & b, std::vector<uint32_t> const & seg_tab, bool back)
{
    if (p.blocks)
    {
        memcpy(ld.A.w, p.pa, 64);
        memcpy(ld.B.w, p.pb, 64);
    }
    if (p.up)
    {
        ld.A.w[0] = ix.lcp[s.lb]; ld.B.w[0] = ix.lcp[s.rb + 1u];
        ld.A.w[2] = ix.pn[s.lb].x; ld.A.w[3] = ix.pn[s.lb].y;
        ld.B.w[2] = ix.pn[s.rb + 1u].x; ld.B.w[3] = ix.pn[s.rb + 1u].y;
    }
    if (p.word) { ld.rdw = *p.pw; ld.rdi = *p.pi; }
    if (p.fetch)
    {
        if (back)
        {
            const uint32_t * sg = &seg_tab[p.fr * 8];
            ld.fg0 = u64of(sg[0], sg[1]); ld.fstart = sg[2]; ld.flen = sg[3]; ld.fgroup = sg[4];
        }
        else { ld.flen = b.len[p.fr]; ld.fg0 = b.grp_off[p.fr]; }
    }
}
}   // namespace

extern "C" {
// out: unique[V], then total, ambiguous, lookups (back-check: unique_ref[V], total_ref[V]).  out_w (may be NULL): local
// mode weighted tallies.  mult (n entries, may be NULL) receives window multiplicities per lb.
// algo 0 = rolling state machine (nlanes lanes striding over the units), 1 = direct.  qual: per-base phred at 32*g0+j.
int sim_scan(const uint8_t * const * texts, const uint64_t * lens, uint32_t S, const int32_t * group_of_text, uint32_t V,
             const uint64_t * bases, const uint32_t * inv, const uint8_t * qual, const uint64_t * grp_off, const uint32_t * len,
             uint64_t n_reads, uint32_t k, int paired, int algo, int backcheck, int local, uint32_t nlanes, uint32_t seg_len,
             uint64_t * out, double * out_w, uint64_t * mult, uint64_t * phase_counts)
{
    try
    {
        HostIndex h;
        build_host_index(texts, lens, S, group_of_text, V, 0, h);
        HostView hv;
        make_view(h, k, hv);
        DevIndex const & ix = hv.ix;
        std::vector<uint64_t> uniq(V, 0), totg(V, 0);
        std::vector<double> uniq_w(V, 0.0);
        uint64_t total = 0, amb = 0, look = 0;
        double qden[42];
        for (int q = 0; q < 42; ++q) { double b = q; qden[q] = 1.0 - 1.0 / pow(10.0, b / 10.0); }
        BatchView bv{};
        bv.bases = bases; bv.inv = inv; bv.qual = qual; bv.grp_off = grp_off; bv.len = len; bv.n_reads = n_reads;
        std::vector<uint32_t> seg_tab;
        uint64_t n_segs = 0;
        if (backcheck)
        {
            bv.bases = h.t_bases.data(); bv.inv = h.t_inv.data(); bv.grp_off = h.t_grp_off.data(); bv.len = h.t_len.data(); bv.n_reads = S;
            paired = 0;
            for (uint32_t i = 0; i < S; ++i)
            {
                if (h.t_len[i] < k) continue;
                uint32_t nw = h.t_len[i] - k + 1;
                for (uint32_t st = 0; st < nw; st += seg_len)
                {
                    uint32_t e[8] = {(uint32_t)h.t_grp_off[i], (uint32_t)(h.t_grp_off[i] >> 32), st, nw - st < seg_len ? nw - st : seg_len,
                                     (uint32_t)group_of_text[i], 0, 0, 0};
                    seg_tab.insert(seg_tab.end(), e, e + 8);
                    ++n_segs;
                }
            }
        }
        if (algo == 0)
        {
            ScanCfg cfg;
            cfg.k = k;
            cfg.units = backcheck ? n_segs : (paired ? n_reads / 2 : n_reads);
            cfg.stride = nlanes;
            cfg.paired = paired; cfg.local = local; cfg.record = mult != nullptr; cfg.back = backcheck;
            HostEnv env{ix, qual, qden, uniq, uniq_w, totg, mult, local != 0};
            std::vector<Lane> lanes(nlanes);
            std::vector<Loads> loads(nlanes);
            for (uint32_t l = 0; l < nlanes; ++l) { lane_init(lanes[l], l); memset(&loads[l], 0, sizeof(Loads)); }
            bool any = true;
            uint64_t guard = 0;
            while (any)
            {
                any = false;
                for (uint32_t l = 0; l < nlanes; ++l)
                {
                    Lane & s = lanes[l];
                    if (s.phase == PH_DONE) continue;
                    any = true;
                    if (phase_counts) ++phase_counts[s.phase];
                    Plan p;
                    lane_plan(s, ix, bv, cfg, p);
                    do_loads(p, loads[l], s, ix, bv, seg_tab, backcheck != 0);
                    lane_consume(s, loads[l], ix, hv.ctab, cfg, env);
                }
                if (++guard > (1ull << 40)) return -2;
            }
            for (auto & s : lanes)
            {
                lane_flush_agg(s, cfg, env);
                total += s.c_total; amb += s.c_amb; look += s.c_look;
            }
        }
        else if (algo == 2)
        {
            // the product kernel's state machine (scan32_core.h) over host-built SBlocks; lanes claim units in turn
            if (backcheck) return -3;
            std::vector<SBlock> sb;
            make_sblocks(h, k, hv.ctab, sb);
            Cfg32 cfg;
            cfg.k = k; cfg.n = (uint32_t)h.n; cfg.units = paired ? n_reads / 2 : n_reads;
            cfg.paired = paired; cfg.local = local; cfg.record = mult != nullptr;
            cfg.mask64 = V <= 64 ? 1 : 0;
            cfg.skip_unique = 0;      // the tests compare the recorded multiplicities with the from-the-root search's
            HostEnv32 env{h, qual, qden, uniq, uniq_w, mult, local != 0, total, amb, look};
            std::vector<Lane32> lanes(nlanes);
            std::vector<Loads32> loads(nlanes);
            uint64_t work = nlanes;
            for (uint32_t l = 0; l < nlanes; ++l) { lane32_init(lanes[l], l, cfg.n); memset(&loads[l], 0, sizeof(Loads32)); }
            bool any = true;
            uint64_t guard = 0;
            while (any)
            {
                any = false;
                for (uint32_t l = 0; l < nlanes; ++l)
                {
                    Lane32 & s = lanes[l];
                    Loads32 & ld = loads[l];
                    if (s.phase == P_DONE) continue;
                    any = true;
                    if (phase_counts) ++phase_counts[s.phase];
                    if (s.phase == P_STEP)
                    {
                        memcpy(ld.A, &sb[s.lb >> 5], 32);
                        memcpy(ld.B, &sb[s.rb1 >> 5], 32);
                        if (s.t < s.pend && (s.t & 31u) == 0u)
                        {
                            uint64_t w = bases[s.g0 + (s.t >> 5)];
                            s.wlo = (uint32_t)w; s.whi = (uint32_t)(w >> 32);
                            s.winv = inv[s.g0 + (s.t >> 5)];
                        }
                        lane32_load_qual(s, cfg, env);
                        lane32_step(s, ld, cfg, env);
                    }
                    else if (s.phase == P_UP)
                    {
                        ld.A[0] = h.pn[s.lb].x; ld.A[1] = h.pn[s.lb].y; ld.A[2] = h.lcp[s.lb];
                        ld.B[0] = h.pn[s.rb1].x; ld.B[1] = h.pn[s.rb1].y; ld.B[2] = h.lcp[s.rb1];
                        lane32_up(s, ld, cfg);
                    }
                    else
                    {
                        if (s.u < cfg.units)
                        {
                            uint64_t fr = lane32_fetch_read(s, cfg);
                            ld.B[4] = len[fr]; ld.B[5] = (uint32_t)grp_off[fr]; ld.B[6] = (uint32_t)(grp_off[fr] >> 32);
                            if (!(s.flags & F_MATE2)) s.unext = work++;
                        }
                        lane32_fetch(s, ld, cfg, env);
                    }
                }
                if (++guard > (1ull << 40)) return -2;
            }
            for (auto & s : lanes) lane32_flush_agg(s, env);
        }
        else
        {
            uint64_t units = paired ? n_reads / 2 : n_reads;
            for (uint64_t u = 0; u < units; ++u)
            {
                uint32_t gmin = 0xFFFFFFFFu, gmax = 0;
                for (int mate = 0; mate < (paired ? 2 : 1); ++mate)
                {
                    uint64_t r = paired ? 2 * u + mate : u;
                    uint32_t L = len[r];
                    uint64_t g0 = grp_off[r];
                    if (L < k) continue;
                    look += L - k + 1;
                    for (uint32_t s0 = 0; s0 + k <= L; ++s0)
                    {
                        bool bad = false;
                        for (uint32_t i = 0; i < k; ++i) bad |= inv_at(inv, g0, s0 + i) != 0;
                        if (bad) continue;
                        ++total;
                        uint32_t lb, rb, steps = 0;
                        if (!direct_window(ix, hv.ctab, bases, inv, g0, s0, k, lb, rb, steps)) continue;
                        if (mult) ++mult[lb];
                        uint32_t g = interval_group(ix, lb, rb);
                        if (g != kNoGroup)
                        {
                            ++uniq[g];
                            gmin = gmin < g ? gmin : g;
                            gmax = gmax > g ? gmax : g;
                        }
                    }
                }
                if (gmin != 0xFFFFFFFFu && gmin != gmax) ++amb;
            }
        }
        for (uint32_t g = 0; g < V; ++g) out[g] = uniq[g];
        if (out_w) for (uint32_t g = 0; g < V; ++g) out_w[g] = uniq_w[g];
        if (backcheck)
            for (uint32_t g = 0; g < V; ++g) out[V + g] = totg[g];
        else
        {
            out[V] = total;
            out[V + 1] = amb;
            out[V + 2] = look;
        }
        return 0;
    }
    catch (...)
    {
        return -1;
    }
}
}
