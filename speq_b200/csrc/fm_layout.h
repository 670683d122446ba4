// fm_layout.h -- data layout of the GPU-resident FM index and of packed read batches (shared by host and device).
//
// The indexed text is the one seqan3 builds for speq (reference: deps/seqan3/include/seqan3/search/fm_index/
// fm_index.hpp:247-316): symbols 0 sentinel, 1 A, 2 C, 3 G, 4 N, 5 T, 6 delimiter; texts joined by the delimiter,
// reversed, sentinel appended.  The suffix array of that text is unique, so SA intervals here are identical to the
// reference's at every backward-search step.
//
// KBlock: everything one rolling-search iteration needs about 64 consecutive suffixes, in ONE 64-byte half cache line
// (two 32-byte sectors, fetched by two LDG.256):
//   sector 0  cnt[4]   occurrences of A,C,G,T in bwt[0, 64*b)
//             lo, hi   the two bit planes of the 2-bit BWT codes (A0 C1 G2 T3), symbol j at bit j
//   sector 1  e        bit j: lcp[64b+j] >= k-1      (filled per scan for the scan's k; zero in the stored index)
//             chg      bit j: group of suffix 64b+j differs from the group of suffix 64b+j-1
//             exc      bit j: bwt[64b+j] is N / delimiter / sentinel (coded 0 in the planes, never counted)
//             chgcum   number of chg bits in [0, 64*b)
// rank(i, c) = cnt[c] + popc(match_c & ~exc & below(i % 64)) -- one block, no exception list.
// "The pattern shortened by its first symbol has the same SA interval" (the common case of a sliding k-mer window) is
// e[lb] == 0 && e[rb+1] == 0, and otherwise the widened interval is found by a bit scan in the very same two blocks the
// next rank step needs.  "All occurrences lie in one variant" (reference: src/fm_scanner.cpp:166-177 after locate())
// is "no chg bit in (lb, rb]" = chgcum/popc arithmetic on those same two blocks.  So a window costs one round trip to
// two half lines, plus one 2-byte group load when (and only when) the k-mer is unique to a variant.
//
// lcp[i] = min(127, longest common prefix of suffixes SA[i-1] and SA[i]) (lcp[0] = lcp[n] = 0), one byte per suffix, and
// pn[i] = {largest j < i with lcp[j] < lcp[i], smallest j > i with lcp[j] < lcp[i]} serve the uncommon moves (after a
// mismatch, or when the widened interval leaves the two blocks): the parent lcp interval in one round trip.
// grp[i]  = variant of the text that owns suffix SA[i] (u16; kNoGroup16 for delimiter / sentinel suffixes).
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
constexpr uint32_t kNoGroup = 0xFFFFFFFFu;
constexpr uint16_t kNoGroup16 = 0xFFFFu;
constexpr uint32_t kMaxGroups = 65535;
constexpr uint32_t kSymN = 4;       // text symbol of N
constexpr uint32_t kSymDelim = 6;   // text symbol of the delimiter (sigma + 1, fm_index.hpp:285)
constexpr uint32_t kMaxK = 127;     // lcp is stored in 7 bits; depth k-1 <= 126 must be representable
constexpr uint32_t kLcpCap = 127;
constexpr uint32_t kCodeN = 4;      // query code of N (back-check only); A0 C1 G2 T3
constexpr uint32_t kLcpChunk = 32;  // the lcp array is read in 32-byte chunks (one LDG.256)

struct __attribute__((aligned(8))) uint2x
{
    uint32_t x, y;
};

struct __attribute__((aligned(64))) KBlock
{
    uint32_t cnt[4];
    uint32_t lo[2];
    uint32_t hi[2];
    uint32_t e[2];
    uint32_t chg[2];
    uint32_t exc[2];
    uint32_t chgcum;
    uint32_t pad;
};
static_assert(sizeof(KBlock) == 64, "KBlock must be 64 bytes");

// SBlock: the scan kernel's per-k view of 32 consecutive suffixes in ONE 32-byte sector (one LDG.256 per lane):
//   cnt[4]  C[c] + occurrences of c = A,C,G,T in bwt[0, 32*b) -- the LF value of the block's first suffix, so a rank step
//           is cnt[c] + popc and the root interval [0, n) needs no special case; bit 31 of cnt[0] = the block holds an
//           exception (N / delimiter / sentinel in the BWT): the lane then fetches the exception mask from the KBlock
//           array (rare slow path)
//   lo, hi  the two bit planes of the 2-bit BWT codes, symbol j at bit j
//   e       bit j: lcp[32b+j] >= k-1
//   uq      bit j: every suffix of the depth-k lcp interval that contains suffix 32b+j belongs to one variant
// Built on the device from the KBlocks + lcp + pn for the scan's k (fill32_kernel).  A window then costs two sector
// loads (the blocks at lb and rb+1) and, when the k-mer is unique to a variant, one 2-byte group load.
struct __attribute__((aligned(32))) SBlock
{
    uint32_t cnt[4];
    uint32_t lo, hi, e, uq;
};
static_assert(sizeof(SBlock) == 32, "SBlock must be 32 bytes");
// FBlock: what the scan kernel actually addresses -- the SBlock of 32 suffixes and, in the other sector of the same
// 64-byte DRAM burst, their variants as bytes (used when the index has <= 255 variants; 0xFF = delimiter / sentinel
// suffix).  On indexes larger than L2 the group load of a unique hit then rides on the burst that brought the SBlock.
struct __attribute__((aligned(64))) FBlock
{
    SBlock s;
    uint8_t g[32];
};
static_assert(sizeof(FBlock) == 64, "FBlock must be 64 bytes");
constexpr uint32_t kMaxGroups8 = 255;
// UpRec: everything a parent jump needs about one suffix position in ONE 16-byte load.
struct __attribute__((aligned(16))) UpRec
{
    uint32_t psv, nsv;   // previous / next position with a smaller lcp value
    uint32_t lcp;        // min(127, lcp)
    uint32_t pad;
};
static_assert(sizeof(UpRec) == 16, "UpRec must be 16 bytes");
// the k-mer table scan (scanh.cu) memoises the search per (index, k) for k up to:
constexpr uint32_t kMaxKTable = 31;
constexpr uint32_t kMaxGroupsPair = 254;   // the pair table stores one byte per k-mer: variant, or 0xFF
constexpr uint32_t kSFlagExc = 0x80000000u;
constexpr uint64_t kMaxN32 = 0x7FFFFFFFull;   // cnt[0] keeps bit 31 for the exception flag

constexpr uint32_t kCtabEnd = 8;

// Key of an ambiguous unit in the fusion map when the index has more variants than a 64-bit mask holds: smallest and largest
// variant with a unique k-mer in the unit, and whether further variants lie between them.
SPEQB_HD uint64_t fusion_key_minmax(uint32_t gmin, uint32_t gmax, bool more)
{
    return (1ull << 63) | (more ? 1ull << 32 : 0ull) | ((uint64_t)gmax << 16) | (uint64_t)gmin;
}

// Device view of the index (passed to kernels by value).
struct DevIndex
{
    const KBlock * blocks;        // n/64 + 1 blocks (for a scan: the copy whose e bits are filled for the scan's k)
    const uint8_t * lcp;          // n + 1 entries (lcp[n] = 0), zero padded
    const uint2x * pn;            // n + 1 entries: {previous smaller, next smaller} lcp positions
    const UpRec * up;             // n + 1 entries: pn + lcp in one record (built on the device at upload)
    const uint16_t * grp;         // n entries
    const uint32_t * n_pos;       // sorted BWT positions holding N (rank of the symbol N; back-check only)
    const uint32_t * ctab;        // 16 entries: [c] = C of query code c (A,C,G,T,N) = #text symbols smaller;
                                  //             [kCtabEnd + c] = C of the next text symbol (exclusive end of the range)
    uint32_t n_npos;
    uint32_t n;                   // text length including the sentinel
};

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
    const unsigned long long * n_reads_dev;   // optional: the number of reads, decided on the device (FASTQ regions parsed there);
                                              // n_reads is then an upper bound
};

// One unit of lane work in the back-check: windows [start, start + nwin) of text `read`.
struct Segment
{
    uint32_t read;
    uint32_t start;
    uint32_t nwin;
};
}   // namespace speqb
