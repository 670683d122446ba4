// sais.hpp -- suffix array construction by induced sorting (Nong, Zhang & Chan, "Two efficient algorithms for linear
// time suffix array construction", 2009).  Own implementation; replaces the reference's libdivsufsort call
// (deps/seqan3/submodules/sdsl-lite/include/sdsl/construct_sa.hpp:80-130).  The suffix array of a text whose last
// symbol is a unique smallest sentinel is unique, so the result is identical to the reference's whatever the algorithm.
#pragma once
#include <cstdint>
#include <vector>

namespace speqb
{
namespace sais_detail
{
struct type_bits
{
    std::vector<uint64_t> w;
    explicit type_bits(size_t n) : w((n + 63) / 64, 0) {}
    bool get(int64_t i) const { return (w[(size_t)i >> 6] >> (i & 63)) & 1u; }
    void set(int64_t i, bool b)
    {
        uint64_t m = uint64_t(1) << (i & 63);
        if (b) w[(size_t)i >> 6] |= m; else w[(size_t)i >> 6] &= ~m;
    }
    bool lms(int64_t i) const { return i > 0 && get(i) && !get(i - 1); }
};

template <class I>
inline void buckets(std::vector<I> const & cnt, std::vector<I> & bkt, bool end)
{
    I sum = 0;
    for (size_t c = 0; c < cnt.size(); ++c)
    {
        sum += cnt[c];
        bkt[c] = end ? sum : sum - cnt[c];
    }
}

template <class C, class I>
inline void induce(type_bits const & t, I * SA, C const * s, std::vector<I> const & cnt, std::vector<I> & bkt, I n)
{
    buckets(cnt, bkt, false);
    for (I i = 0; i < n; ++i)
    {
        I j = SA[i] - 1;
        if (j >= 0 && !t.get(j)) SA[bkt[s[j]]++] = j;
    }
    buckets(cnt, bkt, true);
    for (I i = n - 1; i >= 0; --i)
    {
        I j = SA[i] - 1;
        if (j >= 0 && t.get(j)) SA[--bkt[s[j]]] = j;
    }
}

// s[0..n): symbols in [0, K); s[n-1] must be the unique smallest symbol.  SA has room for n entries.
template <class C, class I>
void sais(C const * s, I * SA, I n, I K)
{
    if (n == 1) { SA[0] = 0; return; }
    type_bits t((size_t)n);
    t.set(n - 1, true);
    t.set(n - 2, false);
    for (I i = n - 3; i >= 0; --i) t.set(i, s[i] < s[i + 1] || (s[i] == s[i + 1] && t.get(i + 1)));

    std::vector<I> cnt((size_t)K, 0), bkt((size_t)K, 0);
    for (I i = 0; i < n; ++i) ++cnt[s[i]];

    // stage 1: sort the LMS substrings
    buckets(cnt, bkt, true);
    for (I i = 0; i < n; ++i) SA[i] = -1;
    for (I i = 1; i < n; ++i) if (t.lms(i)) SA[--bkt[s[i]]] = i;
    induce(t, SA, s, cnt, bkt, n);

    I n1 = 0;
    for (I i = 0; i < n; ++i) if (t.lms(SA[i])) SA[n1++] = SA[i];
    for (I i = n1; i < n; ++i) SA[i] = -1;
    I name = 0, prev = -1;
    for (I i = 0; i < n1; ++i)
    {
        I pos = SA[i];
        bool diff = false;
        for (I d = 0; d < n; ++d)
        {
            if (prev == -1 || s[pos + d] != s[prev + d] || t.get(pos + d) != t.get(prev + d)) { diff = true; break; }
            if (d > 0 && (t.lms(pos + d) || t.lms(prev + d))) break;
        }
        if (diff) { ++name; prev = pos; }
        SA[n1 + pos / 2] = name - 1;
    }
    for (I i = n - 1, j = n - 1; i >= n1; --i) if (SA[i] >= 0) SA[j--] = SA[i];

    // stage 2: suffix array of the reduced string
    I * SA1 = SA;
    I * s1 = SA + n - n1;
    if (name < n1) sais<I, I>(s1, SA1, n1, name);
    else for (I i = 0; i < n1; ++i) SA1[s1[i]] = i;

    // stage 3: induce the full suffix array
    buckets(cnt, bkt, true);
    for (I i = 1, j = 0; i < n; ++i) if (t.lms(i)) s1[j++] = i;
    for (I i = 0; i < n1; ++i) SA1[i] = s1[SA1[i]];
    for (I i = n1; i < n; ++i) SA[i] = -1;
    for (I i = n1 - 1; i >= 0; --i)
    {
        I j = SA[i];
        SA[i] = -1;
        SA[--bkt[s[j]]] = j;
    }
    induce(t, SA, s, cnt, bkt, n);
}
}   // namespace sais_detail

// Suffix array of text[0..n) over bytes < K with text[n-1] the unique smallest symbol; n < 2^31.
inline void suffix_array(uint8_t const * text, int32_t * SA, int32_t n, int32_t K)
{
    sais_detail::sais<uint8_t, int32_t>(text, SA, n, K);
}
}   // namespace speqb
