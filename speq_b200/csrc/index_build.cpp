// index_build.cpp -- host construction of the blocked FM index (see fm_layout.h for the layout).
//
// Reference behaviour reproduced (not its code): speq::fm::generate_fm_index (src/fm_indexer.cpp:8-52) hands
// seqan3::fm_index the collection {scaffold_0, revcomp(scaffold_0), scaffold_1, ...}; seqan3 joins rank+1 symbols with
// delimiter 6, reverses, and sdsl appends the 0 sentinel and builds SA/BWT/C
// (deps/seqan3/include/seqan3/search/fm_index/fm_index.hpp:247-316).  locate() maps SA[i] back to a text with
// loc = (n - k - 1) - SA[i] (fm_index_cursor.hpp:103-107,521-538).  Because an ACGT(N) query can never span a
// delimiter, the text that owns occurrence i is the text that owns the FIRST symbol of suffix SA[i] in the reversed
// text -- independent of k -- so the whole locate() walk collapses into the per-suffix group array built here.
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <thread>

#include <chrono>
#include <cstdio>
#include <cstdlib>

#include "sais.hpp"
#include "speq_internal.h"

namespace speqb
{
// suffix array of T[0, n) (symbols < 8, unique smallest sentinel last) by prefix doubling on the current device
// (sa_device.cu); 0 or a CUDA error code, *msg its text.  Weak: this file is also compiled on its own by the CPU-tier
// white-box tests (tests/sim), which have no CUDA objects to link.
// With owner != null it also returns bwt[i] = T[SA[i]-1], own[i] = owner[SA[i]] and lcp[i] = min(cap, lcp(SA[i-1], SA[i])).
extern "C" int speqb_device_suffix_array(const uint8_t * T, uint64_t n, int32_t * SA, const uint16_t * owner, uint32_t lcp_cap,
                                         uint8_t * bwt, uint16_t * own, uint8_t * lcp, uint32_t * pn, const char ** msg) __attribute__((weak));

static const uint8_t kCodeOfRank[5] = {0, 1, 2, 0, 3};   // dna5 rank -> 2-bit code (N coded 0, flagged in inv)

static void pack_text_as_read(uint8_t const * ranks, uint64_t len, uint64_t * bases, uint32_t * inv)
{
    uint64_t groups = (len + 31) / 32;
    for (uint64_t g = 0; g < groups; ++g)
    {
        uint64_t w = 0;
        uint32_t m = 0;
        uint64_t lim = std::min<uint64_t>(32, len - g * 32);
        for (uint64_t j = 0; j < lim; ++j)
        {
            uint8_t r = ranks[g * 32 + j];
            w |= uint64_t(kCodeOfRank[r]) << (2 * j);
            if (r == 3) m |= 1u << j;
        }
        bases[g] = w;
        inv[g] = m;
    }
}

void build_host_index(uint8_t const * const * texts, uint64_t const * lens, uint32_t S, int32_t const * group_of_text,
                      uint32_t V, uint32_t flags, HostIndex & h)
{
    if (S == 0) throw api_error(SPEQG_EINVAL, "The text that is indexed cannot be empty.");
    if (V == 0 || V > kMaxGroups) throw api_error(SPEQG_EINVAL, "n_groups must be in [1, 65535]");
    uint64_t text_size = 0;
    bool all_empty = true;
    for (uint32_t i = 0; i < S; ++i)
    {
        if (lens[i]) all_empty = false;
        text_size += 1 + lens[i];
        if (group_of_text[i] < 0 || (uint32_t)group_of_text[i] >= V)
            throw api_error(SPEQG_EINVAL, "text " + std::to_string(i) + " has group " + std::to_string(group_of_text[i]) +
                                              " outside [0, n_groups): every scaffold must be assigned to a variant "
                                              "(the reference indexes vectors with -1 for unassigned scaffolds)");
    }
    if (all_empty) throw api_error(SPEQG_EINVAL, "A text collection that only contains empty texts cannot be indexed.");
    uint64_t m = text_size - (S > 1 ? 1 : 0);   // fm_index.hpp:283
    uint64_t n = m + 1;                          // + sentinel
    if (n >= (uint64_t(1) << 31)) throw api_error(SPEQG_EINVAL, "collection too large: indexed text must be < 2^31 symbols");

    h = HostIndex{};
    h.n = n; h.S = S; h.V = V;
    h.group_of_text.assign(group_of_text, group_of_text + S);

    // T = reverse(join(rank+1, 6)) + {0}; owner[p] = group owning T[p] (kNoGroup for delimiter / sentinel)
    std::vector<uint8_t> T(n);
    std::vector<uint32_t> owner(n, kNoGroup);
    {
        uint64_t f = 0;   // position in the forward concatenation; T position = m-1-f
        for (uint32_t i = 0; i < S; ++i)
        {
            uint8_t const * t = texts[i];
            uint32_t g = (uint32_t)group_of_text[i];
            for (uint64_t j = 0; j < lens[i]; ++j, ++f)
            {
                if (t[j] > 4) throw api_error(SPEQG_EINVAL, "text symbols must be dna5 ranks 0..4");
                T[m - 1 - f] = (uint8_t)(t[j] + 1);
                owner[m - 1 - f] = g;
            }
            if (i + 1 < S || S == 1) { T[m - 1 - f] = (uint8_t)kSymDelim; ++f; }   // fm_index.hpp:302,310-311
        }
        T[m] = 0;
    }

    const bool timing = getenv("SPEQB_TIMING") != nullptr;
    auto tick = std::chrono::steady_clock::now();
    auto lap = [&](const char * what) {
        auto now = std::chrono::steady_clock::now();
        if (timing) fprintf(stderr, "[index build] %-10s %.3f s\n", what, std::chrono::duration<double>(now - tick).count());
        tick = now;
    };
    std::vector<int32_t> SA(n);
    // filled by the device path: the gathers through SA and the suffix comparisons (random accesses on the host)
    std::vector<uint8_t> d_bwt, d_lcp;
    std::vector<uint16_t> d_own;
    if (flags & SPEQG_BUILD_DEVICE_SA)
    {
        // prefix doubling on the current CUDA device (sa_device.cu); the suffix array is unique, so it equals SA-IS's
        if (!speqb_device_suffix_array) throw api_error(SPEQG_ECUDA, "SPEQG_BUILD_DEVICE_SA: built without the device suffix sorter");
        const char * msg = "";
        std::vector<uint16_t> owner16(n);
        for (uint64_t i = 0; i < n; ++i) owner16[i] = owner[i] == kNoGroup ? kNoGroup16 : (uint16_t)owner[i];
        d_bwt.resize(n); d_own.resize(n); d_lcp.resize(n);
        h.pn.resize(n + 1);
        static_assert(sizeof(uint2x) == 8, "pn entries are two u32");
        if (speqb_device_suffix_array(T.data(), n, SA.data(), owner16.data(), kLcpCap, d_bwt.data(), d_own.data(), d_lcp.data(),
                                      reinterpret_cast<uint32_t *>(h.pn.data()), &msg) != 0)
            throw api_error(SPEQG_ECUDA, std::string("suffix array on the device failed: ") + msg);
    }
    else suffix_array(T.data(), SA.data(), (int32_t)n, 7);
    lap("sa-is");

    // C[] (deps/seqan3/include/seqan3/search/fm_index/detail/csa_alphabet_strategy.hpp:84-117)
    uint64_t cnt[8] = {0};
    for (uint64_t i = 0; i < n; ++i) ++cnt[T[i]];
    h.C[0] = 0;
    for (int c = 1; c < 8; ++c) h.C[c] = h.C[c - 1] + cnt[c - 1];

    // BWT -> KBlocks.  bwt[i] = T[SA[i]-1], wrapping (sdsl construct_bwt.hpp:35-73).  N / delimiter / sentinel are coded 0
    // in the planes and flagged in the block's exc bits; cnt[] counts A,C,G,T only.
    // Group of every suffix + change bits: chg[i] = group(i) != group(i-1), chg[0] = 1.
    uint64_t nblocks = n / kBlockSyms + 1;
    h.blocks.assign(nblocks, KBlock{});
    h.grp.resize(n);
    static const int8_t code_of_sym[7] = {-1, 0, 1, 2, -1, 3, -1};
    uint32_t run[4] = {0, 0, 0, 0};
    uint32_t chgcum = 0;
    uint32_t prev_group = 0;
    for (uint64_t i = 0; i < n; ++i)
    {
        uint64_t b = i / kBlockSyms, j = i % kBlockSyms;
        KBlock & blk = h.blocks[b];
        if (j == 0) { blk.cnt[0] = run[0]; blk.cnt[1] = run[1]; blk.cnt[2] = run[2]; blk.cnt[3] = run[3]; blk.chgcum = chgcum; }
        uint8_t sym = !d_bwt.empty() ? d_bwt[i] : T[SA[i] ? SA[i] - 1 : n - 1];
        int8_t c2 = code_of_sym[sym];
        if (c2 < 0)
        {
            blk.exc[j / 32] |= 1u << (j % 32);
            ++h.n_exc;
            if (sym == kSymN) h.n_pos.push_back((uint32_t)i);
            c2 = 0;
        }
        else
            ++run[c2];
        blk.lo[j / 32] |= uint32_t(c2 & 1) << (j % 32);
        blk.hi[j / 32] |= uint32_t((c2 >> 1) & 1) << (j % 32);
        uint32_t g = !d_own.empty() ? (d_own[i] == kNoGroup16 ? kNoGroup : (uint32_t)d_own[i]) : owner[SA[i]];
        h.grp[i] = g == kNoGroup ? kNoGroup16 : (uint16_t)g;
        if (i == 0 || g != prev_group)
        {
            blk.chg[j / 32] |= 1u << (j % 32);
            ++chgcum;
        }
        prev_group = g;
    }
    if (n % kBlockSyms == 0)
    {
        KBlock & blk = h.blocks[nblocks - 1];
        blk.cnt[0] = run[0]; blk.cnt[1] = run[1]; blk.cnt[2] = run[2]; blk.cnt[3] = run[3]; blk.chgcum = chgcum;
    }

    lap("bwt+grp");
    // capped LCP array (Kasai et al. 2001, with the cap folded in: lcp[rank[i+1]] >= lcp[rank[i]] - 1 also holds for
    // min(cap, lcp)).  The sentinel is unique, so the symbol comparison stops before either suffix runs off the text.
    if (!d_lcp.empty())
    {
        h.lcp.assign(lcp_padded_size(n), 0);
        memcpy(h.lcp.data(), d_lcp.data(), n);
    }
    else
    {
        std::vector<int32_t> rank(n);
        for (uint64_t i = 0; i < n; ++i) rank[SA[i]] = (int32_t)i;
        h.lcp.assign(lcp_padded_size(n), 0);
        uint32_t hh = 0;
        for (uint64_t i = 0; i < n; ++i)
        {
            int32_t r = rank[i];
            if (r == 0) { hh = 0; continue; }
            uint64_t j = (uint64_t)SA[r - 1];
            while (hh < kLcpCap && T[i + hh] == T[j + hh]) ++hh;
            h.lcp[r] = (uint8_t)hh;
            if (hh) --hh;
        }
    }

    lap("lcp");
    // previous / next smaller value links over the capped lcp array (lcp[n] = 0 closes every interval)
    if (d_lcp.empty())
    {
        h.pn.assign(n + 1, uint2x{0u, (uint32_t)n});
        std::vector<uint32_t> stack;
        stack.reserve(256);
        for (uint64_t i = 0; i <= n; ++i)
        {
            uint8_t v = h.lcp[i];   // the padding makes lcp[n] = 0
            while (!stack.empty() && h.lcp[stack.back()] > v)
            {
                h.pn[stack.back()].y = (uint32_t)i;   // first position to the right with a smaller value
                stack.pop_back();
            }
            // stack top now holds a value <= v; an equal value shares its previous STRICTLY smaller position with i
            if (stack.empty()) h.pn[i].x = 0u;
            else if (h.lcp[stack.back()] == v) h.pn[i].x = h.pn[stack.back()].x;
            else h.pn[i].x = stack.back();
            stack.push_back((uint32_t)i);
        }
    }

    lap("psv/nsv");
    if (flags & SPEQG_BUILD_KEEP_SA)
    {
        h.sa.resize(n);
        for (uint64_t i = 0; i < n; ++i) h.sa[i] = (uint32_t)SA[i];
    }

    // texts as a packed read batch for the back-check
    h.t_grp_off.resize(S + 1);
    h.t_len.resize(S);
    uint64_t g = 0;
    for (uint32_t i = 0; i < S; ++i) { h.t_grp_off[i] = g; h.t_len[i] = (uint32_t)lens[i]; g += (lens[i] + 31) / 32; }
    h.t_grp_off[S] = g;
    for (uint32_t i = 0; i < S; ++i)
        if (lens[i] >= (uint64_t(1) << 32)) throw api_error(SPEQG_EINVAL, "text longer than 2^32");
    h.t_bases.assign(g + 1, 0);
    h.t_inv.assign(g + 1, 0);
    for (uint32_t i = 0; i < S; ++i) pack_text_as_read(texts[i], lens[i], &h.t_bases[h.t_grp_off[i]], &h.t_inv[h.t_grp_off[i]]);
}

// ---- cache file: own binary payload (the reference's .idx is a cereal+sdsl archive; byte compatibility is a later row)
static const char kMagic[8] = {'S', 'P', 'E', 'Q', 'B', '2', '0', '3'};

template <class T> static void wr(FILE * f, T const & v)
{
    if (fwrite(&v, sizeof(T), 1, f) != 1) throw api_error(SPEQG_EIO, "short write");
}
template <class T> static void wrv(FILE * f, std::vector<T> const & v)
{
    uint64_t n = v.size();
    wr(f, n);
    if (n && fwrite(v.data(), sizeof(T), n, f) != n) throw api_error(SPEQG_EIO, "short write");
}
template <class T> static void rd(FILE * f, T & v)
{
    if (fread(&v, sizeof(T), 1, f) != 1) throw api_error(SPEQG_EIO, "short read (truncated index payload)");
}
template <class T> static void rdv(FILE * f, std::vector<T> & v)
{
    uint64_t n;
    rd(f, n);
    if (n > (uint64_t(1) << 36)) throw api_error(SPEQG_EIO, "corrupt index payload");
    v.resize(n);
    if (n && fread(v.data(), sizeof(T), n, f) != n) throw api_error(SPEQG_EIO, "short read (truncated index payload)");
}

void save_host_index(HostIndex const & h, std::string const & path, uint64_t offset)
{
    FILE * f = fopen(path.c_str(), offset ? "r+b" : "wb");
    if (!f) throw api_error(SPEQG_EIO, "cannot open " + path + " for writing");
    try
    {
        if (offset && fseek(f, (long)offset, SEEK_SET) != 0) throw api_error(SPEQG_EIO, "cannot seek in " + path);
        if (fwrite(kMagic, 1, 8, f) != 8) throw api_error(SPEQG_EIO, "short write");
        wr(f, h.n); wr(f, h.S); wr(f, h.V);
        for (int c = 0; c < 8; ++c) wr(f, h.C[c]);
        wr(f, h.n_exc); wrv(f, h.blocks); wrv(f, h.grp); wrv(f, h.lcp); wrv(f, h.pn); wrv(f, h.n_pos); wrv(f, h.group_of_text);
        wrv(f, h.t_bases); wrv(f, h.t_inv); wrv(f, h.t_grp_off); wrv(f, h.t_len); wrv(f, h.sa);
    }
    catch (...) { fclose(f); throw; }
    if (fclose(f) != 0) throw api_error(SPEQG_EIO, "cannot close " + path);
}

void load_host_index(std::string const & path, HostIndex & h, uint64_t offset)
{
    FILE * f = fopen(path.c_str(), "rb");
    if (!f) throw api_error(SPEQG_EIO, "cannot open " + path);
    try
    {
        if (offset && fseek(f, (long)offset, SEEK_SET) != 0) throw api_error(SPEQG_EIO, "cannot seek in " + path);
        char mg[8];
        if (fread(mg, 1, 8, f) != 8 || memcmp(mg, kMagic, 8) != 0) throw api_error(SPEQG_EIO, path + " is not a speq-b200 index payload");
        h = HostIndex{};
        rd(f, h.n); rd(f, h.S); rd(f, h.V);
        for (int c = 0; c < 8; ++c) rd(f, h.C[c]);
        rd(f, h.n_exc); rdv(f, h.blocks); rdv(f, h.grp); rdv(f, h.lcp); rdv(f, h.pn); rdv(f, h.n_pos); rdv(f, h.group_of_text);
        rdv(f, h.t_bases); rdv(f, h.t_inv); rdv(f, h.t_grp_off); rdv(f, h.t_len); rdv(f, h.sa);
        if (h.blocks.size() != h.n / kBlockSyms + 1 || h.grp.size() != h.n || h.group_of_text.size() != h.S ||
            h.lcp.size() != lcp_padded_size(h.n) || h.pn.size() != h.n + 1)
            throw api_error(SPEQG_EIO, "inconsistent index payload");
    }
    catch (...) { fclose(f); throw; }
    fclose(f);
}
}   // namespace speqb
