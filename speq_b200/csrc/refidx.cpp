// refidx.cpp -- read-only loader for index files written by the REFERENCE's `speq index`, behind the
// speqg_refidx_* entry points of include/speq_cuda.h.  Host code; no CUDA.
//
// The reference archives (src/fm_indexer.cpp:39-50) with cereal::BinaryOutputArchive, in this order:
//     std::string                  path of the reference FASTA
//     file_time_type x 2           mtimes of the FASTA and of the groups file (cereal/types/chrono.hpp: the tick count, int64)
//     std::vector<std::string>     group_names         (u64 size, then per string u64 length + bytes)
//     std::vector<int>             group_scaffolds     (u64 size + raw int32)
//     std::vector<int>             group_counts
//     seqan3::fm_index             S3/include/seqan3/search/fm_index/fm_index.hpp:486-514: the sdsl csa_wt, text_begin, ...
// and the csa_wt (SDSL/csa_wt.hpp:396-404) starts with its wavelet tree, wt_blcd = wt_pc<balanced_shape, ...>
// (SDSL/wt_pc.hpp:654-663):
//     u64 m_size, u64 m_sigma
//     int_vector<1> m_bv           (SDSL/int_vector.hpp:1843-1849: u8 width, float growth_factor, u64 size in bits, then
//                                   ceil(size / 64) little-endian words)
//     rank_support_v m_bv_rank     (SDSL/rank_support_v.hpp:138-141: one int_vector<64>)
//     select_support_scan x 2      (nothing, SDSL/select_support_scan.hpp:81)
//     _byte_tree m_tree            (SDSL/wt_helper.hpp:339-345: vector of nodes {u64 bv_pos, u64 bv_pos_rank, u16 parent,
//                                   u16 child[2]} (:139-146), u16 m_c_to_leaf[256], u64 m_path[256])
// Nothing behind the tree is needed: the wavelet tree IS the BWT of the indexed text, and the indexed text determines
// everything else (SA samples, C array, text_begin).  So the loader
//   1. decodes bwt[0, n) from the tree: the bit segment of a node says, symbol by symbol, which child the symbol went to
//      (wt_pc::operator[], SDSL/wt_pc.hpp:294-310); the segments of the children are unshuffled into the parent's order;
//   2. inverts the BWT (LF walk from row 0, the sentinel's suffix): sdsl's text = reverse(T0' 6 T1' 6 ... ) 0
//      (fm_index.hpp:283-316; construct_bwt: bwt[i] = text[SA[i] - 1], SDSL/construct_bwt.hpp:35-73);
//   3. splits it at the delimiters into the texts the reference indexed (scaffold, reverse complement, scaffold, ...) and
//      hands them to this library's own index builder.
// The result is the index `speq index` of this build would write for the same FASTA and groups file.
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "capi_util.h"

using namespace speqb;

struct speqg_refidx
{
    std::string reference_path;
    int64_t scaffold_time = 0, group_time = 0;
    std::vector<std::string> group_names;
    std::vector<int32_t> group_scaffolds, group_counts;
    uint64_t n = 0;                                  // fm_index::size(): text length incl. delimiters and the sentinel
    std::vector<std::vector<uint8_t>> texts;         // dna5 ranks (A0 C1 G2 N3 T4), in the order the reference indexed them
};

namespace
{
struct reader
{
    FILE * f = nullptr;
    std::string path;
    explicit reader(std::string const & p) : f(fopen(p.c_str(), "rb")), path(p)
    {
        if (!f) throw api_error(SPEQG_EIO, "cannot open " + p);
    }
    ~reader() { if (f) fclose(f); }
    void raw(void * dst, size_t bytes)
    {
        if (bytes && fread(dst, 1, bytes, f) != bytes) throw api_error(SPEQG_EIO, path + ": truncated (not an index written by the reference?)");
    }
    template <class T> T get()
    {
        T v{};
        raw(&v, sizeof v);
        return v;
    }
    void skip(uint64_t bytes)
    {
        if (fseeko(f, static_cast<off_t>(bytes), SEEK_CUR) != 0) throw api_error(SPEQG_EIO, path + ": truncated");
    }
    std::string str()
    {
        const uint64_t len = get<uint64_t>();
        if (len > (1u << 20)) throw api_error(SPEQG_EIO, path + ": not an index written by the reference (string length)");
        std::string s(len, '\0');
        raw(s.data(), len);
        return s;
    }
    std::vector<int32_t> ints()
    {
        const uint64_t len = get<uint64_t>();
        if (len > (1u << 28)) throw api_error(SPEQG_EIO, path + ": not an index written by the reference (vector length)");
        std::vector<int32_t> v(len);
        raw(v.data(), len * 4);
        return v;
    }
    // sdsl::int_vector<w>: u8 width, float growth_factor, u64 size (bits), ceil(size / 64) words
    std::vector<uint64_t> int_vector(uint8_t want_width, uint64_t & bits)
    {
        const uint8_t width = get<uint8_t>();
        (void)get<float>();
        bits = get<uint64_t>();
        if (width != want_width || bits > (1ull << 40)) throw api_error(SPEQG_EIO, path + ": unexpected sdsl int_vector header");
        std::vector<uint64_t> w((bits + 63) >> 6);
        raw(w.data(), w.size() * 8);
        return w;
    }
};

struct wt_node
{
    uint64_t bv_pos, bv_pos_rank;
    uint16_t parent, child[2];
};
constexpr uint16_t kUndef = 0xFFFF;

// symbols of the `size` positions of node v, in the order of its parent's sequence restricted to it
void decode_node(std::vector<wt_node> const & nodes, std::vector<int> const & sym_of_leaf, std::vector<uint64_t> const & bv, uint16_t v,
                 uint64_t size, std::vector<uint8_t> & out)
{
    out.resize(size);
    wt_node const & nd = nodes[v];
    if (nd.child[0] == kUndef)
    {
        if (sym_of_leaf[v] < 0) throw api_error(SPEQG_EIO, "wavelet tree leaf without a symbol");
        memset(out.data(), sym_of_leaf[v], size);
        return;
    }
    uint64_t ones = 0;
    for (uint64_t j = 0; j < size; ++j)
    {
        const uint64_t p = nd.bv_pos + j;
        ones += (bv[p >> 6] >> (p & 63)) & 1ull;
    }
    std::vector<uint8_t> left, right;
    decode_node(nodes, sym_of_leaf, bv, nd.child[0], size - ones, left);
    decode_node(nodes, sym_of_leaf, bv, nd.child[1], ones, right);
    uint64_t l = 0, r = 0;
    for (uint64_t j = 0; j < size; ++j)
    {
        const uint64_t p = nd.bv_pos + j;
        out[j] = ((bv[p >> 6] >> (p & 63)) & 1ull) ? right[r++] : left[l++];
    }
}

void load(std::string const & path, speqg_refidx & R)
{
    reader in(path);
    R.reference_path = in.str();
    R.scaffold_time = in.get<int64_t>();
    R.group_time = in.get<int64_t>();
    const uint64_t n_names = in.get<uint64_t>();
    if (n_names > 65535) throw api_error(SPEQG_EIO, path + ": not an index written by the reference (group count)");
    for (uint64_t i = 0; i < n_names; ++i) R.group_names.push_back(in.str());
    R.group_scaffolds = in.ints();
    R.group_counts = in.ints();
    if (R.group_counts.size() != n_names) throw api_error(SPEQG_EIO, path + ": group_names and group_counts differ in length");
    // ---- csa_wt: the wavelet tree over the BWT
    const uint64_t n = in.get<uint64_t>(), sigma = in.get<uint64_t>();
    if (n < 2 || n > kMaxN32 || sigma < 2 || sigma > 7) throw api_error(SPEQG_EIO, path + ": unexpected wavelet tree header (size / sigma)");
    uint64_t bv_bits = 0, rank_words = 0;
    std::vector<uint64_t> bv = in.int_vector(1, bv_bits);
    {
        // rank_support_v::m_basic_block: int_vector<64>, not needed
        const uint8_t width = in.get<uint8_t>();
        (void)in.get<float>();
        rank_words = in.get<uint64_t>();
        if (width != 64 || rank_words > (1ull << 36)) throw api_error(SPEQG_EIO, path + ": unexpected rank support header");
        in.skip(((rank_words + 63) >> 6) * 8);
    }
    const uint64_t n_nodes = in.get<uint64_t>();
    if (n_nodes == 0 || n_nodes > 2 * 256) throw api_error(SPEQG_EIO, path + ": unexpected wavelet tree shape");
    std::vector<wt_node> nodes(n_nodes);
    for (auto & nd : nodes)
    {
        nd.bv_pos = in.get<uint64_t>();
        nd.bv_pos_rank = in.get<uint64_t>();
        nd.parent = in.get<uint16_t>();
        nd.child[0] = in.get<uint16_t>();
        nd.child[1] = in.get<uint16_t>();
    }
    uint16_t c_to_leaf[256];
    in.raw(c_to_leaf, sizeof c_to_leaf);
    std::vector<int> sym_of_leaf(n_nodes, -1);
    for (int c = 0; c < 256; ++c)
        if (c_to_leaf[c] != kUndef)
        {
            if (c_to_leaf[c] >= n_nodes || c > 6) throw api_error(SPEQG_EIO, path + ": the index is not over dna5 (symbol out of range)");
            sym_of_leaf[c_to_leaf[c]] = c;
        }
    for (auto const & nd : nodes)
        if (nd.child[0] != kUndef && (nd.child[0] >= n_nodes || nd.child[1] >= n_nodes || nd.bv_pos > bv_bits))
            throw api_error(SPEQG_EIO, path + ": corrupt wavelet tree");
    std::vector<uint8_t> bwt;
    decode_node(nodes, sym_of_leaf, bv, 0, n, bwt);
    bv.clear();
    bv.shrink_to_fit();
    // ---- invert: LF[i] = C[bwt[i]] + #{j < i : bwt[j] == bwt[i]}; row 0 is the suffix that starts with the sentinel
    uint64_t cnt[8] = {0};
    for (uint64_t i = 0; i < n; ++i) ++cnt[bwt[i]];
    if (cnt[0] != 1) throw api_error(SPEQG_EIO, path + ": the BWT does not hold exactly one sentinel");
    uint64_t C[8], run = 0;
    for (int c = 0; c < 8; ++c) { C[c] = run; run += cnt[c]; }
    std::vector<uint32_t> lf(n);
    {
        uint64_t seen[8] = {0};
        for (uint64_t i = 0; i < n; ++i) lf[i] = (uint32_t)(C[bwt[i]] + seen[bwt[i]]++);
    }
    std::vector<uint8_t> text(n - 1);       // sdsl's text without the sentinel = reverse(T0' 6 T1' 6 ... T(S-1)')
    {
        uint64_t i = 0;
        for (uint64_t p = n - 1; p-- > 0;)
        {
            const uint8_t c = bwt[i];
            if (c == 0) throw api_error(SPEQG_EIO, path + ": the BWT is not a single cycle");
            text[p] = c;
            i = lf[i];
        }
    }
    lf.clear();
    lf.shrink_to_fit();
    // ---- un-reverse and split at the delimiter (sigma + 1 = 6); symbols are dna5 rank + 1 (fm_index.hpp:285,302).
    //      One text: the reference appends a delimiter to it (fm_index.hpp:310-313); several: none behind the last.
    std::vector<uint8_t> cur;
    R.texts.clear();
    bool last_was_delim = false;
    for (uint64_t p = n - 1; p-- > 0;)
    {
        const uint8_t c = text[p];
        last_was_delim = c == kSymDelim;
        if (last_was_delim) { R.texts.push_back(std::move(cur)); cur.clear(); }
        else cur.push_back((uint8_t)(c - 1u));
    }
    if (!(last_was_delim && R.texts.size() == 1)) R.texts.push_back(std::move(cur));
    R.n = n;
}
}   // namespace

extern "C" {

int speqg_refidx_open(const char * path, speqg_refidx ** out)
{
    SPEQG_TRY
    need(path && out, "NULL argument");
    auto * r = new speqg_refidx();
    try
    {
        load(path, *r);
    }
    catch (...)
    {
        delete r;
        throw;
    }
    *out = r;
    SPEQG_CATCH
}

const char * speqg_refidx_reference_path(const speqg_refidx * r) { return r ? r->reference_path.c_str() : ""; }
int64_t speqg_refidx_file_time(const speqg_refidx * r, int which) { return !r ? 0 : which == 0 ? r->scaffold_time : r->group_time; }
uint64_t speqg_refidx_size(const speqg_refidx * r) { return r ? r->n : 0; }
uint32_t speqg_refidx_n_groups(const speqg_refidx * r) { return r ? (uint32_t)r->group_names.size() : 0; }
const char * speqg_refidx_group_name(const speqg_refidx * r, uint32_t g) { return r && g < r->group_names.size() ? r->group_names[g].c_str() : ""; }
int32_t speqg_refidx_group_count(const speqg_refidx * r, uint32_t g) { return r && g < r->group_counts.size() ? r->group_counts[g] : 0; }
uint32_t speqg_refidx_n_scaffolds(const speqg_refidx * r) { return r ? (uint32_t)r->group_scaffolds.size() : 0; }
int32_t speqg_refidx_group_of_scaffold(const speqg_refidx * r, uint32_t j) { return r && j < r->group_scaffolds.size() ? r->group_scaffolds[j] : -1; }
uint32_t speqg_refidx_n_texts(const speqg_refidx * r) { return r ? (uint32_t)r->texts.size() : 0; }

int speqg_refidx_text(const speqg_refidx * r, uint32_t i, const uint8_t ** ranks, uint64_t * len)
{
    SPEQG_TRY
    need(r && ranks && len, "NULL argument");
    need(i < r->texts.size(), "text index out of range");
    *ranks = r->texts[i].data();
    *len = r->texts[i].size();
    SPEQG_CATCH
}

int speqg_refidx_build_index(const speqg_refidx * r, uint32_t flags, speqg_index ** out)
{
    SPEQG_TRY
    need(r && out, "NULL argument");
    const size_t S = r->texts.size();
    need(S % 2 == 0 && S / 2 <= r->group_scaffolds.size(),
         "the archive's texts are not (scaffold, reverse complement) pairs covered by its group_scaffolds");
    std::vector<const uint8_t *> ptrs(S);
    std::vector<uint64_t> lens(S);
    std::vector<int32_t> grp(S);
    for (size_t i = 0; i < S; ++i)
    {
        ptrs[i] = r->texts[i].data();
        lens[i] = r->texts[i].size();
        grp[i] = r->group_scaffolds[i / 2];                   // double_group_scaffolds, src/fm_scanner.cpp:74-77
    }
    return speqg_index_build(ptrs.data(), lens.data(), (uint32_t)S, grp.data(), (uint32_t)r->group_names.size(), flags, out);
    SPEQG_CATCH
}

void speqg_refidx_close(speqg_refidx * r) { delete r; }

}   // extern "C"
