// roll_sim.cpp -- TEST-ONLY white-box harness: runs the per-lane arithmetic of speq_b200/csrc/scan_core.h (the very
// header the CUDA kernels compile) on the host, one "lane" at a time, so that index layout / rolling-window logic can be
// checked against the oracle in the CPU-only test tier.  It is never linked into libspeq_cuda.so and is not a product
// path: the product has no CPU fallback.  Built by tests/conftest.py with g++.
#include <cstdint>
#include <cstring>
#include <vector>

#include "../../speq_b200/csrc/scan_core.h"
#include "../../speq_b200/csrc/speq_internal.h"

using namespace speqb;

static DevIndex host_view(HostIndex const & h)
{
    DevIndex v{};
    v.blocks = reinterpret_cast<const RankBlock *>(h.blocks.data());
    v.lcp = h.lcp.data();
    v.gr = h.gr.data();
    v.exc_pos = h.exc_pos.data();
    v.n_pos = h.n_pos.data();
    v.n_exc = (uint32_t)h.exc_pos.size();
    v.n_npos = (uint32_t)h.n_pos.size();
    v.n = (uint32_t)h.n;
    static const uint32_t sym_of_code[5] = {1, 2, 3, 5, 4};
    for (int c = 0; c < 5; ++c)
    {
        v.ctab[c] = (uint32_t)h.C[sym_of_code[c]];
        v.ctab[kCtabEnd + c] = (uint32_t)h.C[sym_of_code[c] + 1];
    }
    return v;
}

extern "C" {
// out: unique[V], then total, ambiguous, lookups.  mult (n entries, may be NULL) receives window multiplicities per lb.
// algo 0 = rolling, 1 = direct.  backcheck != 0: batch = the index's own texts, out = unique_ref[V], total_ref[V].
int sim_scan(const uint8_t * const * texts, const uint64_t * lens, uint32_t S, const int32_t * group_of_text, uint32_t V,
             const uint64_t * bases, const uint32_t * inv, const uint64_t * grp_off, const uint32_t * len, uint64_t n_reads,
             uint32_t k, int paired, int algo, int backcheck, uint64_t * out, uint64_t * mult)
{
    try
    {
        HostIndex h;
        build_host_index(texts, lens, S, group_of_text, V, 0, h);
        DevIndex ix = host_view(h);
        const uint32_t * ctab = ix.ctab;
        std::vector<uint64_t> uniq(V, 0), totg(V, 0);
        uint64_t total = 0, amb = 0, look = 0;
        if (backcheck)
        {
            bases = h.t_bases.data(); inv = h.t_inv.data(); grp_off = h.t_grp_off.data(); len = h.t_len.data(); n_reads = S;
            paired = 0;
        }
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
                if (backcheck) totg[group_of_text[r]] += L - k + 1;
                if (algo == 0)
                {
                    RollState s;
                    roll_reset(ix, s);
                    uint32_t t = 0;
                    while (t < L)
                    {
                        uint32_t c = code_at(bases, g0, t);
                        bool ib = inv_at(inv, g0, t);
                        if (backcheck && ib) c = kCodeN;
                        uint32_t res = roll_advance(ix, ctab, k, s, c, backcheck ? false : ib);
                        if (res & ROLL_CONSUMED) ++t;
                        if (res & ROLL_WINDOW) ++total;
                        if (res & ROLL_HIT)
                        {
                            if (mult) ++mult[s.lb];
                            GroupRun g = ix.gr[s.lb];
                            if (g.run_end >= s.rb)
                            {
                                ++uniq[g.group];
                                gmin = gmin < g.group ? gmin : g.group;
                                gmax = gmax > g.group ? gmax : g.group;
                            }
                        }
                    }
                }
                else
                {
                    for (uint32_t s0 = 0; s0 + k <= L; ++s0)
                    {
                        bool bad = false;
                        for (uint32_t i = 0; i < k; ++i) bad |= inv_at(inv, g0, s0 + i) != 0;
                        if (bad && !backcheck) continue;
                        ++total;
                        uint32_t lb, rb, steps = 0;
                        if (!direct_window(ix, ctab, bases, inv, g0, s0, k, lb, rb, steps)) continue;
                        if (mult) ++mult[lb];
                        GroupRun g = ix.gr[lb];
                        if (g.run_end >= rb)
                        {
                            ++uniq[g.group];
                            gmin = gmin < g.group ? gmin : g.group;
                            gmax = gmax > g.group ? gmax : g.group;
                        }
                    }
                }
            }
            if (gmin != 0xFFFFFFFFu && gmin != gmax) ++amb;
        }
        for (uint32_t g = 0; g < V; ++g) out[g] = uniq[g];
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
