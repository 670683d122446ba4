// pack.cpp -- host packer: raw reads (ASCII or rank bytes) -> 2-bit packed, length-prefixed batch (fm_layout.h).
//
// Alphabet rules are the reference's (via seqan3): sequence letters are read over the dna15 legal alphabet
// (deps/seqan3/include/seqan3/io/sequence_file/input.hpp:145-162) and converted to dna5 -- ACGT / acgt, U->T, every
// other IUPAC letter -> N (deps/seqan3/include/seqan3/alphabet/nucleotide/dna5.hpp:87-119); a non-IUPAC letter is a
// parse error.  Qualities are phred42: clamp(char - 33, 0, 41)
// (deps/seqan3/include/seqan3/alphabet/quality/quality_base.hpp:139-160, phred42.hpp:77-80).
// Window validity (src/fm_scanner.cpp:162) is "min phred > cutoff and no N", i.e. no base of the window is N or has
// phred <= cutoff; both conditions fold into one `inv` bit per base.
#include <atomic>
#include <cstring>
#include <thread>

#include "speq_internal.h"

namespace speqb
{
namespace
{
struct tables
{
    uint8_t code[256];     // 0..3 code, 4 = N, 255 = illegal
    uint8_t phred[256];
    tables()
    {
        memset(code, 255, sizeof code);
        const char * iupac = "ACGTUNRYSWKMBDHVacgtunryswkmbdhv";
        for (const char * p = iupac; *p; ++p) code[(unsigned char)*p] = 4;
        code['A'] = code['a'] = 0;
        code['C'] = code['c'] = 1;
        code['G'] = code['g'] = 2;
        code['T'] = code['t'] = code['U'] = code['u'] = 3;
        for (int c = 0; c < 256; ++c) phred[c] = (c < 33 || c >= 128) ? 0 : (c >= 33 + 42 ? 41 : c - 33);
    }
};
const tables & tb()
{
    static tables t;
    return t;
}
const uint8_t kCodeOfRank[5] = {0, 1, 2, 4, 3};
}   // namespace

uint64_t pack_groups(uint64_t const * offsets, uint64_t n_reads)
{
    uint64_t g = 0;
    for (uint64_t i = 0; i < n_reads; ++i) g += (offsets[i + 1] - offsets[i] + 31) / 32;
    return g;
}

void pack_reads(uint8_t const * seq, uint8_t const * qual, uint64_t const * offsets, uint64_t n_reads, int encoding,
                uint32_t phred_cutoff, bool n_only, int threads, speqg_batch & out)
{
    if (encoding != SPEQG_ENC_ASCII && encoding != SPEQG_ENC_RANK) throw api_error(SPEQG_EINVAL, "unknown encoding");
    uint64_t g = 0;
    for (uint64_t i = 0; i < n_reads; ++i)
    {
        uint64_t L = offsets[i + 1] - offsets[i];
        if (L >= (uint64_t(1) << 32)) throw api_error(SPEQG_EINVAL, "read longer than 2^32");
        out.grp_off[i] = g;
        out.len[i] = (uint32_t)L;
        g += (L + 31) / 32;
    }
    out.grp_off[n_reads] = g;
    out.n_reads = n_reads;
    out.n_groups32 = g;

    tables const & t = tb();
    std::atomic<int> bad{0};
    auto work = [&](uint64_t lo, uint64_t hi)
    {
        for (uint64_t i = lo; i < hi; ++i)
        {
            uint8_t const * s = seq + offsets[i];
            uint8_t const * q = qual ? qual + offsets[i] : nullptr;
            uint64_t L = out.len[i], g0 = out.grp_off[i];
            for (uint64_t gi = 0; gi * 32 < L; ++gi)
            {
                uint64_t w = 0;
                uint32_t m = 0;
                uint64_t lim = L - gi * 32 < 32 ? L - gi * 32 : 32;
                for (uint64_t j = 0; j < lim; ++j)
                {
                    uint8_t ch = s[gi * 32 + j];
                    uint8_t c = encoding == SPEQG_ENC_ASCII ? t.code[ch] : (ch <= 4 ? kCodeOfRank[ch] : 255);
                    if (c == 255) { bad.store(1); c = 4; }
                    uint32_t inv = (c == 4);
                    if (c == 4) c = 0;
                    uint8_t ph = 0;
                    if (q)
                    {
                        ph = encoding == SPEQG_ENC_ASCII ? t.phred[q[gi * 32 + j]] : (q[gi * 32 + j] > 41 ? 41 : q[gi * 32 + j]);
                        if (!n_only && !(ph > phred_cutoff)) inv = 1;
                        if (out.qual) out.qual[(g0 + gi) * 32 + j] = ph;
                    }
                    w |= uint64_t(c) << (2 * j);
                    m |= inv << j;
                }
                if (out.qual) for (uint64_t j = lim; j < 32; ++j) out.qual[(g0 + gi) * 32 + j] = 0;
                out.bases[g0 + gi] = w;
                out.inv[g0 + gi] = m;
            }
        }
    };
    if (threads <= 1 || n_reads < 1024) work(0, n_reads);
    else
    {
        std::vector<std::thread> th;
        for (int k = 0; k < threads; ++k) th.emplace_back(work, n_reads * k / threads, n_reads * (k + 1) / threads);
        for (auto & x : th) x.join();
    }
    if (bad.load()) throw api_error(SPEQG_EINVAL, "parse_error: illegal (non-IUPAC) sequence character in a read");
}
}   // namespace speqb
