// kernels.h -- launchers of the sm_100a kernels (kernels.cu), called by the C ABI (capi.cu).
#pragma once
#include <cuda_runtime.h>

#include "fm_layout.h"

namespace speqb
{
enum : int
{
    M_LOCAL = 1,       // fp64 phred-weighted unique tallies
    M_RECORD = 2,      // record per-k-mer multiplicities for the refinement
    M_BACKCHECK = 4,   // reads are the index's own texts; N is searched; tallies are unique_ref/total_ref
    M_PAIRED = 8       // reads 2i, 2i+1 share one ambiguity state
};

// indices into the u64 scalar counter block
enum : int { SC_TOTAL = 0, SC_AMBIG = 1, SC_LOOKUPS = 2, SC_STEPS = 3, SC_HIT = 4, SC_OCC = 5, SC_OCCSEC = 6, SC_READS = 7, SC_COUNT = 8 };

// One unit of lane work in the back-check, as the kernel fetches it (one 32-byte load): windows [start, start + nwin) of the
// text whose packed bases start at 32-base group g0 and whose variant is `group`.
struct __attribute__((aligned(32))) SegmentX
{
    uint64_t g0;
    uint32_t start;
    uint32_t nwin;
    uint32_t group;
    uint32_t pad[3];
};

struct ScanArgs
{
    uint32_t k;
    uint32_t V;
    unsigned long long * unique;      // [V] (back-check: unique_ref)
    unsigned long long * total_g;     // [V] back-check: total_ref
    double * unique_w;                // [V] local mode
    unsigned long long * scalars;     // [SC_COUNT]
    unsigned long long * mult;        // [n] multiplicity of the valid k-mer windows whose interval starts at lb (k-mers held by
                                      //     several variants only: see unique_n)
    unsigned long long * unique_n;    // [V] or null: valid windows whose k-mer lies in variant g only.  Such a window adds
                                      //     a[g] / a[g] = 1 to the refinement estimate (src/fm_scanner.cpp:1099-1119), so the
                                      //     product kernels record no multiplicity for it and the estimate adds this count
    unsigned long long * diff;        // k-mer table with positions: difference array over the packed-text positions; a run of
                                      //     hits at consecutive positions [s, e] is recorded as diff[s] += 1, diff[e + 1] -= 1
    const double * qden;              // [42] 1 - 1/pow(10, q/10), host libm values   (local mode)
    const double * qrden;             // [42] 1 / qden (inf for q = 0: never inside a valid window, phred > cutoff >= 0)
    const SegmentX * segs;             // optional segment table (back-check); null: one segment per read / pair
    uint64_t n_segs;
    unsigned long long * work;        // scan32: next unclaimed unit minus the number of threads (zero at launch)
    // ambiguous units by the SET of variants that met in them (the key the reference's fusion_totals map was meant to hold,
    // src/fm_scanner.cpp:915,983-995): an open-addressing map in device memory, fkeys[i] = 0 when free.  Key of a unit:
    // V <= kMaxGroupsMask: bit g set for every variant g with a unique k-mer in the unit; else fusion_key_minmax().
    // fvals[fcap] counts the units that found no place within kFusionMaxProbe slots.
    unsigned long long * fkeys;       // [fcap] or null
    unsigned long long * fvals;       // [fcap + 1]
    uint32_t fcap_mask;               // fcap - 1 (fcap is a power of two)
};
constexpr uint32_t kMaxGroupsMask = 64;        // variant sets of indexes with at most this many variants are kept as bit masks
constexpr uint32_t kFusionMaxProbe = 128;

struct LaunchCfg
{
    int grid;     // CTAs (multiple of the SM count)
    int block;    // threads per CTA
};

// rolling-window scan (the product kernel)
cudaError_t launch_scan_roll(DevIndex const & ix, BatchView const & b, ScanArgs const & a, int mode, LaunchCfg cfg, cudaStream_t st);
// from-the-root search of every window, one warp per read.  stats != 0: only the reference-semantics work counters
// (SC_STEPS, SC_HIT, SC_OCC, SC_OCCSEC, SC_LOOKUPS) are produced; else the same tallies as launch_scan_roll (global mode).
cudaError_t launch_scan_direct(DevIndex const & ix, BatchView const & b, ScanArgs const & a, int mode, int stats, LaunchCfg cfg, cudaStream_t st);
cudaError_t scan_launch_cfg(int mode, int direct, uint32_t V, LaunchCfg * cfg);   // occupancy-sized persistent grid
size_t scan_smem_bytes(uint32_t V);

// refinement estimate: tkpg[g] = sum over recorded k-mers of mult * a[g] / sum(a), a[g] = h[g] * percent[g] / group_counts[g]
cudaError_t launch_refine(DevIndex const & ix, uint32_t k, const unsigned long long * mult, const double * percent,
                          const int32_t * group_counts, uint32_t V, double * partial, int grid, int block, cudaStream_t st);
// unique_n != null: tkpg[g] += unique_n[g] where percent[g] > 0 (the closed form of the windows that were not recorded)
cudaError_t launch_refine_reduce(const double * partial, int nparts, uint32_t V, double * tkpg, const unsigned long long * unique_n,
                                 const double * percent, cudaStream_t st);
// device packer: raw reads (ASCII or dna5 ranks) already in device memory -> packed batch.  make_grp_off != 0: the kernel
// lays read r out at 32-base group floor((offsets[r] - off0) / 32) + r and writes grp_off[r] itself (the batch then needs
// (offsets[n_reads] - off0) / 32 + n_reads + 1 groups); else grp_off is an input.
// *bad is set to 1 when an illegal sequence character is met.
// seq / qual hold the bytes [off0, offsets[n_reads]) of the caller's arrays (offsets are relative to the caller's).
cudaError_t launch_pack(const uint8_t * seq, const uint8_t * qual, uint64_t off0, const uint64_t * offsets, uint64_t n_reads,
                        int encoding, uint32_t phred_cutoff, int n_only, uint64_t * bases, uint32_t * inv, uint8_t * qual_out,
                        uint64_t * grp_off, int make_grp_off, uint32_t * len, unsigned long long * bad, cudaStream_t st);
// per-scan specialisation of the rank blocks: copy `src` to `dst` and fill the e bits (lcp >= k-1) for this k
cudaError_t launch_fill_e(const KBlock * src, KBlock * dst, const uint8_t * lcp, uint32_t n, uint32_t k, cudaStream_t st);
// the product scan kernel (scan32.cu): rolling window over the per-k 32-byte SBlocks; a.work must point to one u64
cudaError_t launch_scan32(const FBlock * fblk, DevIndex const & ix, BatchView const & b, ScanArgs const & a, int mode, LaunchCfg cfg,
                          cudaStream_t st);
cudaError_t scan32_launch_cfg(int mode, uint32_t V, LaunchCfg * cfg);
size_t scan32_smem_bytes(uint32_t V);
// per-scan specialisation: SBlocks (rank counts, bit planes, e = lcp >= k-1, uq = depth-k interval in one variant) for this k
cudaError_t launch_fill32(const KBlock * kb, const uint8_t * lcp, const uint2x * pn, const uint16_t * grp, const uint32_t * ctab,
                          FBlock * out, uint32_t n, uint32_t k, cudaStream_t st);
// parent-jump records {psv, nsv, lcp} (one 16-byte load per side) from the stored lcp and pn arrays; once per device copy
cudaError_t launch_build_up(const uint8_t * lcp, const uint2x * pn, UpRec * out, uint64_t count, cudaStream_t st);
// k-mer table scan (scanh.cu): the search memoised per (index, k), k <= kMaxKTable
struct KTabView
{
    uint64_t * slots;             // 4 << q slots: value << 48 | displacement << rbits | remainder; all ones = free
    uint32_t q, rbits, nbmask;    // 2^q buckets of 4 slots; rbits = 2k - q stored key bits; nbmask = 2^q - 1
    uint64_t mask_b, rmask;       // 4^k - 1, 2^rbits - 1
    uint64_t mul, mul_inv;        // invertible multiplicative hash mod 4^k and its inverse
    uint32_t * filt;              // optional presence filter (pair-table scan): bit (hash >> fshift) is set for every stored k-mer
    uint32_t fshift;
    uint32_t wide;                // 1: 64-byte buckets = four slots + their four record-mode counters (tables beyond L2)
    uint32_t pshift;              // != 0: slots of k-mers held by several variants carry, from this bit up to bit 62, the
                                  // smallest position (32 * group + offset in the packed texts) where the k-mer starts
    // walk != 0 (global mode, positions fit): EVERY slot carries its k-mer's smallest text position (bit 63 still says "several
    // variants"; the variant of a single-variant k-mer is that of the text the position lies in, gvar[position >> 5]), and the
    // scan follows the packed texts from a hit instead of probing the table for every window (scanw_kernel, scanh.cu)
    uint32_t walk;
    uint32_t tgroups;             // 32-base groups of the packed texts
    const uint64_t * tbases;      // the packed texts (2 bits per base, 32 per word)
    const uint64_t * tflags;      // per group: low word = bit j set when the window that ENDS at base 32 g + j is an N-free k-mer of
                                  // one text; high word = ... and texts of several variants hold that k-mer
    const uint16_t * gvar;        // variant of the text a group belongs to
    // walk != 0 with pshift == 0 (a position does not fit next to the remainder: large tables): the slots keep their variant
    // and the positions live in a side array, one per slot -- read only where the walk starts and where a hit is recorded
    uint32_t * pos_side;
    // walk != 0, pshift == 32 and 15 more bits free in the low word (rbits + 18 <= 32: small k): single-variant slots ALSO keep
    // their variant, from this bit up -- the warp-per-read scan (long reads, scanh_kernel) then needs no gvar load per hit
    uint32_t vshift;
};
bool ktable_positions_fit(KTabView const & kt, uint64_t npos);
cudaError_t launch_ktable_clear(KTabView const & kt, cudaStream_t st);
cudaError_t launch_ktable_zero_counters(KTabView const & kt, cudaStream_t st);
// t_group / tflags / gvar != NULL: the walk form (every slot gets its position; the per-group flags and variants are written)
cudaError_t launch_ktable_canon(const uint64_t * t_bases, const uint32_t * t_inv, const uint64_t * t_grp_off, const uint32_t * t_len,
                                const int32_t * t_group, uint32_t n_texts, uint32_t max_len, KTabView const & kt, uint32_t k,
                                uint64_t * tflags, uint16_t * gvar, cudaStream_t st);
size_t kpos_fold_tmp_bytes(uint64_t npos);
cudaError_t launch_kpos_fold(const uint64_t * t_bases, DevIndex const & ix, uint32_t k, unsigned long long * diff, uint64_t npos,
                             void * tmp, size_t tmp_bytes, unsigned long long * mult, KTabView const & kt, cudaStream_t st);
// presence filter of the k-mer table for the pair-table scan: 2^fbits bits, L2-resident next to the HBM-resident tables
uint32_t kfilter_bits(uint32_t k, uint64_t keys);
cudaError_t launch_kfilter_build(KTabView const & kt, uint32_t fbits, cudaStream_t st);
bool ktable_geometry(uint32_t k, uint64_t keys, uint32_t grow, KTabView * kt);
cudaError_t launch_ktable_build(const uint64_t * t_bases, const uint32_t * t_inv, const uint64_t * t_grp_off, const uint32_t * t_len,
                                const int32_t * t_group, uint32_t n_texts, uint32_t max_len, KTabView const & kt, uint32_t k,
                                unsigned long long * n_keys, cudaStream_t st);
cudaError_t launch_ktable_fold_mult(KTabView const & kt, DevIndex const & ix, uint32_t k, unsigned long long * slot_mult,
                                    unsigned long long * mult, cudaStream_t st);
// warp_per_read != 0: the 32 lanes of a warp split the windows of one read (long reads); else one read / pair per lane
cudaError_t launch_scanh(KTabView const & kt, unsigned long long * slot_mult, BatchView const & b, ScanArgs const & a, int mode,
                         int warp_per_read, LaunchCfg cfg, cudaStream_t st);
cudaError_t scanh_launch_cfg(int mode, uint32_t V, int walk, LaunchCfg * cfg);   // walk != 0: the configuration of scanw_kernel instead
// two windows per probe (SPEQG_ALGO_KPAIR): a second table over the (k+1)-mers of the texts, values = one byte per k-mer
cudaError_t launch_kpair_build(const uint64_t * t_bases, const uint32_t * t_inv, const uint64_t * t_grp_off, const uint32_t * t_len,
                               uint32_t n_texts, uint32_t max_len, KTabView const & kt, KTabView const & kp, uint32_t k,
                               unsigned long long * n_keys, cudaStream_t st);
cudaError_t launch_kpair_fold_mult(KTabView const & kp, DevIndex const & ix, uint32_t k, unsigned long long * slot_mult,
                                   unsigned long long * mult, cudaStream_t st);
cudaError_t launch_scanp(KTabView const & kt, KTabView const & kp, unsigned long long * slot_mult, unsigned long long * pair_mult,
                         BatchView const & b, ScanArgs const & a, int mode, int warp_per_read, LaunchCfg cfg, cudaStream_t st);
cudaError_t scanp_launch_cfg(int mode, uint32_t V, LaunchCfg * cfg);
// FASTQ record grammar on the device (fastq.cu): phase 1 counts the newlines of a raw region (chunk_cnt becomes the exclusive
// prefix sum per 4 KB chunk, *total the count); phase 2, once the host knows the count, finds the lines and records, packs the
// reads and publishes the number of reads (0 and *region_flag = 1 when the region is not in the strict 4-line form)
uint32_t fastq_chunks(uint64_t n_bytes);
cudaError_t launch_fastq_count(const uint8_t * raw, uint64_t n_bytes, uint32_t * chunk_cnt, unsigned long long * total, cudaStream_t st);
cudaError_t launch_fastq_begin(unsigned int * reject, cudaStream_t st);
cudaError_t launch_fastq_records(const uint8_t * raw, uint64_t n_bytes, const uint32_t * chunk_off, uint64_t n_newlines, uint64_t n_rec,
                                 uint64_t * line_start, uint64_t * soff, uint64_t * qoff, uint32_t * len, uint64_t * grp_off, uint32_t stride,
                                 uint32_t slot, uint64_t gbase, uint32_t cutoff, uint64_t * bases, uint32_t * inv, uint8_t * qual_out,
                                 unsigned int * reject, cudaStream_t st);
cudaError_t launch_fastq_finalize(const unsigned int * reject, uint64_t n_reads, unsigned long long * n_reads_dev, unsigned int * region_flag,
                                  cudaStream_t st);
cudaError_t launch_count_nonzero(const unsigned long long * mult, uint32_t n, unsigned long long * out, cudaStream_t st);
}   // namespace speqb
