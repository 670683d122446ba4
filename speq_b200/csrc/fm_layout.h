// fm_layout.h -- data layout of the GPU-resident FM index and of packed read batches (shared by host and device).
//
// The indexed text is the one seqan3 builds for speq (reference: deps/seqan3/include/seqan3/search/fm_index/
// fm_index.hpp:247-316): symbols 0 sentinel, 1 A, 2 C, 3 G, 4 N, 5 T, 6 delimiter; texts joined by the delimiter,
// reversed, sentinel appended.  The suffix array of that text is unique, so SA intervals here are identical to the
// reference's at every backward-search step.
//
// Rank structure ("blocks"): the BWT is cut into blocks of 64 symbols; each block is ONE 32-byte sector:
//     u32 cnt[4]   occurrences of A,C,G,T in bwt[0, 64*b)          (bit 31 of cnt[0] = block holds an exception)
//     u32 lo[2]    bit plane 0 of the 2-bit codes (A0 C1 G2 T3): symbol j at bit j%32 of lo[j/32]
//     u32 hi[2]    bit plane 1
// so rank(i, c) costs exactly one sector load (one LDG.256) and the in-block count is two LOP3 + POPC per word.
// BWT positions that hold N / delimiter / sentinel ("exceptions") are coded as 0 and listed in `exc_pos` (sorted);
// only blocks with the flag bit need the correction, and only for c = A.
//
// LCP structure: lcp[i] = min(127, longest common prefix of suffixes SA[i-1] and SA[i]) (lcp[0] = 0).  The SA interval
// of a pattern shortened by its FIRST query symbol (= last symbol of the reversed pattern) is the maximal range around
// the old interval whose inner lcp values are >= the new depth; this is what lets consecutive, overlapping k-mer
// windows of a read share their search work (scan_core.h) while producing the very same intervals as k independent
// backward-search steps per window.
//
// Group structure: gr[i] = {run_end, group}: group of the text that owns suffix SA[i] (kNoGroup for delimiter /
// sentinel suffixes) and the last index of the maximal run of equal groups starting at i.  "All occurrences of a
// k-mer lie in one variant" (reference: src/fm_scanner.cpp:166-177 after locate()) becomes gr[lb].run_end >= rb.
#pragma once
#include <cstdint>

#if defined(__CUDACC__)
#define SPEQB_HD __host__ __device__ __forceinline__
#else
#define SPEQB_HD inline
#endif

namespace speqb
{
constexpr uint32_t kBlockSyms = 64;
constexpr uint32_t kExcFlag = 0x80000000u;
constexpr uint32_t kNoGroup = 0xFFFFFFFFu;
constexpr uint32_t kSymN = 4;       // text symbol of N
constexpr uint32_t kSymDelim = 6;   // text symbol of the delimiter (sigma + 1, fm_index.hpp:285)
constexpr uint32_t kMaxK = 127;     // lcp is stored in 7 bits; depth k-1 <= 126 must be representable
constexpr uint32_t kLcpCap = 127;
constexpr uint32_t kCodeN = 4;      // query code of N (back-check only); A0 C1 G2 T3

struct __attribute__((aligned(32))) RankBlock
{
    uint32_t cnt[4];
    uint32_t lo[2];
    uint32_t hi[2];
};

struct GroupRun
{
    uint32_t run_end;
    uint32_t group;
};

// Device view of the index (passed to kernels by value).
struct DevIndex
{
    const RankBlock * blocks;     // n/64 + 1 blocks
    const uint8_t * lcp;          // n entries, padded to a multiple of 8 bytes (+8) with zeros
    const GroupRun * gr;          // n entries
    const uint32_t * exc_pos;     // sorted BWT positions holding a non-ACGT symbol
    const uint32_t * n_pos;       // sorted BWT positions holding N
    uint32_t n_exc;
    uint32_t n_npos;
    uint32_t n;                   // text length including the sentinel
    uint32_t ctab[16];            // [c]: C of query code c (A,C,G,T,N) = #text symbols smaller; [kCtabEnd + c]: C of the
                                  // next text symbol (exclusive end of the symbol's SA range)
};
constexpr uint32_t kCtabEnd = 8;

// Device/host view of one packed batch of reads.  Read i owns 32-base groups [grp_off[i], grp_off[i+1]).
//   bases[g]  u64: 32 x 2-bit codes     inv[g] u32: 1 bit per base (scan: N or phred <= cutoff; back-check: N)
//   qual      u8 phred per base at 32*grp_off[i] + j (local-error mode only, else null)
struct BatchView
{
    const uint64_t * bases;
    const uint32_t * inv;
    const uint8_t * qual;
    const uint64_t * grp_off;     // n_reads + 1
    const uint32_t * len;         // n_reads
    const int32_t * read_group;   // back-check only: group of the scaffold the read came from
    uint64_t n_reads;
};
}   // namespace speqb
