/* speq_cuda.h -- C ABI of libspeq_cuda.so: the B200 (sm_100a) replacement for SPeQ's scan hot path.
 *
 * SPeQ (ibharvey/speq) has no plugin/FFI layer of its own: the hot path is reached from three C++ entry points called
 * by main() (reference src/main.cpp:21-35).  Every entry point below names the reference interface it replaces
 * (paths relative to the reference checkout; S3 = deps/seqan3, SDSL = S3/submodules/sdsl-lite/include/sdsl).
 * INTEGRATION.md shows the binding a maintainer of the reference would add.
 *
 * Conventions: plain pointers and sizes only; every function returns 0 on success and a negative SPEQG_E* code on
 * failure, with a human-readable message available from speqg_last_error() (thread-local).  No exception crosses the
 * ABI.  Handles are not thread-safe; use one speqg_scan per host thread / CUDA device.  There is NO CPU fallback:
 * every compute entry point fails with SPEQG_ECUDA when no CUDA device is usable.
 */
#ifndef SPEQ_CUDA_H
#define SPEQ_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SPEQG_ABI_VERSION 2   /* 2: fusion sets, FASTQ regions, copy tickets, reference .idx import, device packer, probe */

enum
{
    SPEQG_OK = 0,
    SPEQG_EINVAL = -1,   /* bad argument (message says which) */
    SPEQG_ECUDA = -2,    /* CUDA runtime error / no device */
    SPEQG_EIO = -3,      /* file could not be read / written / has the wrong format */
    SPEQG_ENOMEM = -4,
    SPEQG_ESTATE = -5    /* call made in the wrong state (e.g. refine without recorded multiplicities) */
};

typedef struct speqg_index speqg_index;   /* FM index over the variant collection (host copy + per-device copies) */
typedef struct speqg_scan speqg_scan;     /* one scan stream on one device */

const char * speqg_last_error(void);
int speqg_abi_version(void);
int speqg_device_count(int * count);

/* ------------------------------------------------------------------------------------------------------------------
 * Index.  Replaces seqan3::fm_index{fnr_ref_seqs} (reference src/fm_indexer.cpp:19-36; construction
 * S3/include/seqan3/search/fm_index/fm_index.hpp:233-316; sdsl::construct_im SDSL/construct.hpp:110-166) and the
 * text-position machinery behind locate() (S3/include/seqan3/search/fm_index/fm_index_cursor.hpp:521-538).
 *
 * texts[i] / lens[i]: the S texts of the collection, as dna5 RANKS (A0 C1 G2 N3 T4), in the order speq indexes them:
 *   text 2j = scaffold j, text 2j+1 = reverse complement of scaffold j (src/fm_indexer.cpp:25-33).
 * group_of_text[i]: variant of text i in [0, V) -- the reference's double_group_scaffolds (src/fm_scanner.cpp:74-77).
 *   Negative (unassigned) entries are rejected: the reference indexes vectors with -1 in its refinement loop
 *   (src/fm_scanner.cpp:1102), which is undefined behaviour.
 * flags: SPEQG_BUILD_KEEP_SA keeps the suffix array for speqg_index_export (tests).
 * ------------------------------------------------------------------------------------------------------------------ */
#define SPEQG_BUILD_KEEP_SA 1u
#define SPEQG_BUILD_DEVICE_SA 2u   /* sort the suffixes on the current CUDA device (prefix doubling) instead of SA-IS on one
                                      host core; same suffix array, SPEQG_ECUDA without a device */

typedef struct speqg_index_info
{
    uint64_t n;             /* indexed text length incl. delimiters and sentinel == seqan3 fm_index::size() */
    uint32_t n_texts;       /* S */
    uint32_t n_groups;      /* V */
    uint64_t C[8];          /* cumulative symbol counts: C[c] = #text symbols < c, c = 0 sentinel,1 A,2 C,3 G,4 N,5 T,6 delim */
    uint64_t n_exceptions;  /* BWT positions holding N / delimiter / sentinel */
    uint64_t device_bytes;  /* bytes one device copy occupies */
} speqg_index_info;

int speqg_index_build(const uint8_t * const * texts, const uint64_t * lens, uint32_t n_texts,
                      const int32_t * group_of_text, uint32_t n_groups, uint32_t flags, speqg_index ** out);
int speqg_index_save(const speqg_index * idx, const char * path);
int speqg_index_load(const char * path, speqg_index ** out);
/* same, with the payload starting `offset` bytes into an existing file (the `speq` binary keeps the reference's .idx
 * metadata -- reference path, mtimes, group names / scaffolds / counts, src/fm_indexer.cpp:39-50 -- in front of it). */
int speqg_index_save_at(const speqg_index * idx, const char * path, uint64_t offset);
int speqg_index_load_at(const char * path, uint64_t offset, speqg_index ** out);
int speqg_index_get_info(const speqg_index * idx, speqg_index_info * info);
/* White-box export for parity tests: any pointer may be NULL (n entries each).  sa needs SPEQG_BUILD_KEEP_SA.  bwt holds
 * text symbols (A1 C2 G3 N4 T5; delimiter and sentinel both 0xFF); lcp is capped at 127. */
int speqg_index_export(const speqg_index * idx, uint32_t * sa, uint8_t * bwt, uint32_t * group_of_sa, uint8_t * lcp);
int speqg_index_upload(speqg_index * idx, int device);
void speqg_index_free(speqg_index * idx);

/* ------------------------------------------------------------------------------------------------------------------
 * Index files written by the REFERENCE (`speq index` of ibharvey/speq): read-only import.
 * The file is a cereal::BinaryOutputArchive of (reference FASTA path, two file times, group_names, group_scaffolds,
 * group_counts, seqan3::fm_index) -- src/fm_indexer.cpp:39-50, loaded by src/fm_scanner.cpp:44-58; the fm_index starts with the
 * sdsl csa_wt, whose wavelet tree is the BWT of the indexed text (S3/include/seqan3/search/fm_index/fm_index.hpp:486-514,
 * SDSL/csa_wt.hpp:396-404, SDSL/wt_pc.hpp:654-663, SDSL/int_vector.hpp:1843-1849).  speqg_refidx_open parses the metadata,
 * decodes the BWT from the tree, inverts it and splits the text at the delimiters: the texts the reference indexed
 * (scaffold j, reverse complement of scaffold j, ...; src/fm_indexer.cpp:25-33) as dna5 ranks.  speqg_refidx_build_index
 * builds this library's index over them -- what speqg_index_build would produce from the original FASTA.
 * SPEQG_EIO when the file is not such an archive.
 * ------------------------------------------------------------------------------------------------------------------ */
typedef struct speqg_refidx speqg_refidx;
int speqg_refidx_open(const char * path, speqg_refidx ** out);
const char * speqg_refidx_reference_path(const speqg_refidx * r);             /* the archived in_file_references */
int64_t speqg_refidx_file_time(const speqg_refidx * r, int which);           /* 0: FASTA, 1: groups file (file_clock ticks) */
uint64_t speqg_refidx_size(const speqg_refidx * r);                          /* seqan3 fm_index::size() */
uint32_t speqg_refidx_n_groups(const speqg_refidx * r);
const char * speqg_refidx_group_name(const speqg_refidx * r, uint32_t g);
int32_t speqg_refidx_group_count(const speqg_refidx * r, uint32_t g);
uint32_t speqg_refidx_n_scaffolds(const speqg_refidx * r);                   /* length of group_scaffolds */
int32_t speqg_refidx_group_of_scaffold(const speqg_refidx * r, uint32_t j);
uint32_t speqg_refidx_n_texts(const speqg_refidx * r);
int speqg_refidx_text(const speqg_refidx * r, uint32_t i, const uint8_t ** ranks, uint64_t * len);   /* owned by r */
int speqg_refidx_build_index(const speqg_refidx * r, uint32_t flags, speqg_index ** out);             /* flags: SPEQG_BUILD_* */
void speqg_refidx_close(speqg_refidx * r);

/* ------------------------------------------------------------------------------------------------------------------
 * Packed read batches.  Replaces what seqan3::sequence_file_input hands the reference's workers (vector<dna5> +
 * vector<phred42> per record, src/fm_scanner.cpp:198-199) by a 2-bit packed, length-prefixed batch.
 * Read i owns 32-base groups [grp_off[i], grp_off[i+1]):
 *   bases[g]: 32 x 2-bit codes (A0 C1 G2 T3; N coded 0), base j of the read in bits [2*(j%32), +2) of group j/32
 *   inv[g]:   1 bit per base: base is N, or (scan) its phred <= phred_cutoff -- a window containing a set bit is not
 *             counted (src/fm_scanner.cpp:162: min(phred) > cutoff && no 'N')
 *   qual:     phred (0..41) per base at byte 32*grp_off[i] + j; only for local-error mode, else NULL
 * ------------------------------------------------------------------------------------------------------------------ */
typedef struct speqg_batch
{
    uint64_t * bases;
    uint32_t * inv;
    uint8_t * qual;
    uint64_t * grp_off;    /* n_reads + 1 entries */
    uint32_t * len;        /* n_reads entries */
    uint64_t n_reads;
    uint64_t n_groups32;   /* == grp_off[n_reads] */
} speqg_batch;

#define SPEQG_ENC_ASCII 0   /* seq: IUPAC letters (dna5 rules, S3/include/seqan3/alphabet/nucleotide/dna5.hpp:87-119);
                               qual: phred+33 chars (clamped to 0..41, S3/.../quality/quality_base.hpp:139-160) */
#define SPEQG_ENC_RANK 1    /* seq: dna5 ranks 0..4; qual: phred values */

/* offsets: n_reads+1 byte offsets into seq/qual (read i = [offsets[i], offsets[i+1])).  qual may be NULL (no quality
 * filter, e.g. reference scaffolds).  n_only != 0: inv marks N only (back-check / statistics).  The arrays of `out`
 * must be preallocated: n_groups32 = speqg_pack_groups(offsets, n_reads), bases n_groups32 u64, inv n_groups32 u32,
 * qual 32*n_groups32 bytes or NULL, grp_off n_reads+1, len n_reads.  Illegal (non-IUPAC) letters -> SPEQG_EINVAL
 * (the reference throws seqan3::parse_error). */
uint64_t speqg_pack_groups(const uint64_t * offsets, uint64_t n_reads);
int speqg_pack_reads(const uint8_t * seq, const uint8_t * qual, const uint64_t * offsets, uint64_t n_reads,
                     int encoding, uint32_t phred_cutoff, int n_only, int threads, speqg_batch * out);

/* ------------------------------------------------------------------------------------------------------------------
 * Scan.  Replaces the worker bodies of speq::scan::_async_{one,two}_with_{global,local}_error_rate
 * (src/fm_scanner.cpp:145-216, 418-491, 662-732, 911-998): sliding k-mer windows, seqan3::search() with
 * max_error{total{0}} (S3/include/seqan3/search/algorithm/search.hpp:166-192), and do_a_count.
 * ------------------------------------------------------------------------------------------------------------------ */
#define SPEQG_MODE_GLOBAL 0   /* --fixed-accuracy != 0: integer unique tallies (src/fm_scanner.cpp:180) */
#define SPEQG_MODE_LOCAL 1    /* fp64 tallies weighted by prod 1/(1-10^(-q/10)) (src/fm_scanner.cpp:454-455) */

#define SPEQG_ALGO_ROLLING 0  /* one pass per read: consecutive windows share their search state (any k) */
#define SPEQG_ALGO_DIRECT 1   /* every window searched from the root, k steps, like seqan3's search_trivial (global mode) */
#define SPEQG_ALGO_ROLLING64 2 /* the rolling search on the generic 64-byte blocks (the back-check's kernel); cross-checks */
#define SPEQG_ALGO_KTABLE 3   /* k <= 31: the search memoised per (index, k) as a k-mer table built from the index's own texts;
                                 one 32-byte probe per window.  Same tallies, bit for bit */
#define SPEQG_ALGO_AUTO 4     /* k <= 31: KTABLE, or KPAIR when the k-mer table is larger than 3/4 of L2 and the index has <= 254
                                 variants; else ROLLING */
#define SPEQG_ALGO_KPAIR 5    /* KTABLE plus a table over the (k+1)-mers of the texts: one probe answers two consecutive
                                 windows; a miss falls back to two k-mer probes; a presence filter of the k-mers (kept in L2)
                                 answers absent k-mers before either table is touched.  k <= 31, <= 254 variants */

typedef struct speqg_params
{
    uint32_t k;                    /* k-mer length, 1..128 (include/arg_parse.h:20) */
    uint32_t phred_cutoff;         /* include/arg_parse.h:21; used by the device packer of submit_ascii */
    int32_t mode;                  /* SPEQG_MODE_* */
    int32_t paired;                /* reads 2i and 2i+1 of every batch are mates sharing one ambiguity state */
    int32_t record_multiplicity;   /* keep per-k-mer multiplicities so speqg_refine can run without re-reading reads */
    uint32_t algorithm;            /* SPEQG_ALGO_* */
    void * stream;                 /* cudaStream_t to launch on; NULL = a stream owned by the handle */
} speqg_params;

typedef struct speqg_counts
{
    uint64_t total_kmers;       /* valid windows (src/fm_scanner.cpp:164) */
    uint64_t ambiguous_reads;   /* reads / pairs with unique k-mers of >= 2 variants (src/fm_scanner.cpp:183-190,213) */
    uint64_t lookups;           /* windows presented = sum max(0, L-k+1) */
    uint64_t reads;
    uint64_t kernel_launches;   /* kernels launched by this handle since open / reset */
    /* filled by speqg_scan_stats only (reference-semantics work counters for the roofline): */
    uint64_t steps_executed;    /* backward_search calls incl. the failing one (fm_index_cursor.hpp:316-324); from
                                   speqg_scan_finish with a table algorithm: 32-byte table probes issued */
    uint64_t kmers_hit;         /* k-mers with a non-empty final interval; from speqg_scan_finish with the walking table scan
                                   (global mode): windows answered from the packed texts, without a table probe */
    uint64_t occurrences;       /* sum of final interval sizes */
    uint64_t occ_sectors;       /* sum ceil(occ/32); from speqg_scan_finish with the walking table scan: walk steps (<= 32
                                   windows each) */
} speqg_counts;

int speqg_scan_open(speqg_index * idx, int device, const speqg_params * params, speqg_scan ** out);
/* host packed batch: H2D copy + kernel, asynchronous; the host arrays may be reused after speqg_scan_sync. */
int speqg_scan_submit_packed(speqg_scan * scan, const speqg_batch * host_batch);
/* host raw reads (ASCII or ranks): H2D copy of the bytes + device packer + kernel, asynchronous. */
int speqg_scan_submit_raw(speqg_scan * scan, const uint8_t * seq, const uint8_t * qual, const uint64_t * offsets,
                          uint64_t n_reads, int encoding);
/* Raw FASTQ bytes: a region of a (plain, uncompressed) FASTQ file that starts at a record start and ends behind a record,
 * single-end.  H2D copy of the bytes, then the record grammar of seqan3 (S3/include/seqan3/io/sequence_file/format_fastq.hpp:
 * 93-205) for the STRICT form -- four lines per record ('@' id, sequence, '+' line, |sequence| quality characters), LF or CRLF --
 * runs as kernels: line starts, record checks, the packer reading sequence and quality where they lie, the scan.  Asynchronous;
 * the host does nothing per record.  *region (may be NULL) = the number of this region within the scan, from 0.  A region that is
 * not in the strict form (multi-line records, blank lines, blanks inside a sequence line, letters outside the IUPAC set, length
 * mismatch) is NOT scanned: speqg_scan_fastq_rejected names it, and the caller reads that region again with the full grammar on
 * the host and submits it with speqg_scan_submit_raw (tallies add up, the order of regions does not matter).
 * *ticket (may be NULL): pass to speqg_scan_wait_ticket before `bytes` is overwritten. */
int speqg_scan_submit_fastq(speqg_scan * scan, const uint8_t * bytes, uint64_t n_bytes, uint64_t * region, uint64_t * ticket);
/* The same for paired input (scan opened with paired = 1): a region of the first mates' file and a region of the second mates'
 * file that hold the SAME number of whole records (record i of one pairs with record i of the other, what views::zip of the two
 * files yields, src/fm_scanner.cpp:651); the device interleaves them.  Regions whose record counts differ are rejected like
 * irregular ones. */
int speqg_scan_submit_fastq_pair(speqg_scan * scan, const uint8_t * bytes1, uint64_t n_bytes1, const uint8_t * bytes2,
                                 uint64_t n_bytes2, uint64_t * region, uint64_t * ticket);
/* Queue everything submitted so far on the scan's stream without waiting for it: the second phase of the last FASTQ region
 * (lines, records, pack, scan) is otherwise issued by the next call.  Call it before work of your own that must run BEHIND the
 * scan on that stream -- e.g. the in-place all-reduce of speqg_scan_device_counters. */
int speqg_scan_flush(speqg_scan * scan);
/* ticket of the most recent submit that copied host memory (submit_packed / submit_raw / submit_fastq); 0 before the first */
int speqg_scan_last_ticket(speqg_scan * scan, uint64_t * ticket);
/* wait for the host->device copy of ONE earlier submit (its host buffer may then be refilled; later copies and all kernels
 * keep running) */
int speqg_scan_wait_ticket(speqg_scan * scan, uint64_t ticket);
/* regions submitted since open / reset that were not scanned because they are not strict 4-line FASTQ.  Syncs the scan.
 * regions[i], i < min(*n_rejected, capacity). */
int speqg_scan_fastq_rejected(speqg_scan * scan, uint64_t * regions, uint64_t capacity, uint64_t * n_rejected);
/* batch whose arrays already live in device memory of the scan's device (pointers are device pointers). */
int speqg_scan_submit_device(speqg_scan * scan, const speqg_batch * device_batch);
int speqg_scan_sync(speqg_scan * scan);
/* wait only for the host->device copies submitted so far (the host buffers may be refilled; kernels keep running). */
int speqg_scan_wait_copies(speqg_scan * scan);
/* sync, then copy the tallies out.  unique_u64 / unique_f64: n_groups entries each, either may be NULL. */
int speqg_scan_finish(speqg_scan * scan, uint64_t * unique_u64, double * unique_f64, speqg_counts * counts);
int speqg_scan_reset(speqg_scan * scan);
/* Reference-semantics work counters for one device batch (every window searched from the root with no jump table,
 * N searched as a symbol, stop at the first empty interval).  Not part of the product path; used for the roofline. */
int speqg_scan_stats(speqg_scan * scan, const speqg_batch * device_batch, speqg_counts * counts);
/* What the handle resolved SPEQG_ALGO_AUTO to, and the size of its per-(index, k) structure: the k-mer table
 * (algorithm == SPEQG_ALGO_KTABLE: bytes, distinct k-mers stored) or the rank blocks of the rolling search (keys = 0). */
int speqg_scan_info(speqg_scan * scan, uint32_t * algorithm, uint64_t * structure_bytes, uint64_t * keys);
/* Multi-GPU (replaces the sum over the per-thread vectors after future.get(), src/fm_scanner.cpp:224-233, across devices):
 * the handle's counters in device memory, for a sum-all-reduce in place on the scan's stream between the last submit and
 * speqg_scan_finish, which then returns the global tallies.  u64_counters: unique[n_groups], back-check totals[n_groups],
 * then total_kmers, ambiguous_reads, lookups, steps_executed, kmers_hit, occurrences, occ_sectors, reads (*n_u64 entries);
 * f64_counters: the local-mode tallies (*n_f64 = n_groups entries).  Any output pointer may be NULL.  After
 * speqg_scan_submit_fastq, call speqg_scan_flush before the reduction (the last region's kernels are queued lazily). */
/* Ambiguous reads / pairs by the variants that met in them (fix-forward of the reference's fusion_totals map, whose key is
 * always the empty set because set_groups is passed by value, src/fm_scanner.cpp:915-916,983-995,1033): counts[a * n_groups
 * + b] = units whose smallest unique variant is a and largest is b (a < b); the entries sum to ambiguous_reads.  Syncs the
 * scan.  Derived from the set map below; not filled by SPEQG_ALGO_DIRECT / SPEQG_ALGO_ROLLING64. */
int speqg_scan_fusion_counts(speqg_scan * scan, uint64_t * counts);
/* The same by the full SET of variants: the key std::map<std::set<int>,int> fusion_totals of the reference was meant to hold
 * (src/fm_scanner.cpp:915, 947-952 inserts both sides of every change of variant, :983-995 counts the unit under its set).
 * keys[i] / counts[i], i < min(*n_sets, capacity): for indexes of at most 64 variants keys[i] has bit g set for every variant g
 * with a unique k-mer in the unit; above 64 variants keys[i] = 1 << 63 | more << 32 | largest << 16 | smallest, `more` = further
 * variants lie between the two.  *n_sets = distinct sets seen (call with capacity 0 to size the arrays).  The device map holds
 * 2^20 sets; units that found no place are counted in *unkeyed_units (0 in practice; they are still part of ambiguous_reads).
 * Syncs the scan.  Not filled by SPEQG_ALGO_DIRECT / SPEQG_ALGO_ROLLING64. */
int speqg_scan_fusion_sets(speqg_scan * scan, uint64_t * keys, uint64_t * counts, uint64_t capacity, uint64_t * n_sets,
                           uint64_t * unkeyed_units);
int speqg_scan_device_counters(speqg_scan * scan, uint64_t ** u64_counters, uint64_t * n_u64, double ** f64_counters, uint64_t * n_f64);
void speqg_scan_close(speqg_scan * scan);

/* The device packer as an entry point of its own: raw reads that already live in device memory -> a packed batch in device
 * memory (what speqg_scan_submit_raw runs behind its host-to-device copy; same rules as speqg_pack_reads).  d_seq / d_qual /
 * d_offsets: device pointers (read i = bytes [d_offsets[i], d_offsets[i+1]) of d_seq; d_qual may be NULL).  device_batch: device
 * pointers of preallocated arrays; grp_off[0..n_reads] is an INPUT (first 32-base group of every read, e.g. a running sum of
 * ceil(len / 32)), bases / inv / len (and qual when not NULL) are written.  Several calls may fill disjoint read ranges of one
 * large batch.  Asynchronous on `stream` (NULL: the default stream) except for the final check: an illegal (non-IUPAC) letter
 * -> SPEQG_EINVAL, like the reference's seqan3::parse_error. */
int speqg_pack_device(int device, const uint8_t * d_seq, const uint8_t * d_qual, const uint64_t * d_offsets, uint64_t n_reads,
                      int encoding, uint32_t phred_cutoff, int n_only, const speqg_batch * device_batch, void * stream);

/* Measurement aid (not on the scan path): giga (1e9) INDEPENDENT random 32-byte sector loads per second this device serves
 * from a table of table_bytes -- the roofline of kernels whose every access is one data-dependent sector (k-mer table probes,
 * rank steps of the FM search).  Allocates and frees the table; takes the best of a few launch shapes. */
int speqg_probe_random_sectors(int device, uint64_t table_bytes, double * gsectors_per_s);

/* page-locked host memory for read batches (replaces the heap vectors seqan3's record buffer holds,
 * deps/seqan3/include/seqan3/range/views/async_input_buffer.hpp:89-116). */
int speqg_host_alloc(size_t bytes, void ** out);
void speqg_host_free(void * p);

/* device-resident copy of a host batch (for speqg_scan_submit_device); free with speqg_batch_free_device. */
int speqg_batch_to_device(int device, const speqg_batch * host_batch, speqg_batch * device_batch);
void speqg_batch_free_device(int device, speqg_batch * device_batch);

/* ------------------------------------------------------------------------------------------------------------------
 * Back-check.  Replaces speq::scan::_async_count_unique_kmers_per_group (src/fm_scanner.cpp:1476-1576): for every
 * scaffold (text 2j) and its reverse complement (text 2j+1), every k-mer incl. those containing N; unique iff every
 * occurrence lies in the scaffold's own group.  unique_ref / total_ref: n_groups entries.
 * ------------------------------------------------------------------------------------------------------------------ */
int speqg_backcheck(speqg_index * idx, int device, uint32_t k, uint64_t * unique_ref, uint64_t * total_ref);

/* ------------------------------------------------------------------------------------------------------------------
 * Refinement estimate.  Replaces speq::scan::_async_{one,two}_{global,local}_estimate_kmer_per_group
 * (src/fm_scanner.cpp:1069-1453) WITHOUT re-reading the reads: uses the multiplicities recorded during the scan
 * (params.record_multiplicity).  percent, group_counts: n_groups entries in; tkpg: n_groups entries out.
 * The per-k-mer weight of the reference (1/pp-(1-pp), resp. the phred products) multiplies every group of one k-mer
 * alike and cancels in a[g]/sum(a) (src/fm_scanner.cpp:1099-1119), so it is not needed here.
 * ------------------------------------------------------------------------------------------------------------------ */
int speqg_refine(speqg_scan * scan, const double * percent, const int32_t * group_counts, double * tkpg);
/* export / import the recorded multiplicities (lb, rb, count) so several devices or ranks can be merged. */
int speqg_scan_multiplicity_count(speqg_scan * scan, uint64_t * n_entries);
/* The same estimate for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  Calls no cudaMalloc, cudaFree
 * or allocating library routine, copies nothing and does not synchronise: it launches speqg_refine's two kernels on the
 * scan's stream (speqg_params.stream, else the handle's own; speqg_scan_sync waits for it).
 *   d_percent, d_group_counts: n_groups entries in DEVICE memory of the scan's device; d_tkpg: n_groups entries out, DEVICE.
 * The per-block partial sums live in one file-scope __device__ array of SPEQG_NOALLOC_MAX_DOUBLES doubles (1 MiB, held by
 * every GPU that loads the library until the process exits): n_groups x (4 x multiprocessors) must not exceed it -- at most
 * 221 variants on a 148-SM B200 -- else SPEQG_EINVAL.  All calls (all handles, all streams) of one device share that array,
 * which keeps the last call's contents (every call overwrites what it reads, nothing has to start at zero): calls that
 * overlap on the GPU race -- keep them on one stream or order them with events.  SPEQG_ESTATE when work is still due before
 * the multiplicities can be read (reads submitted since the last speqg_refine / speqg_scan_multiplicity_count, a deferred
 * FASTQ region): call one of those two once after the last submit, then this as often as needed. */
#define SPEQG_NOALLOC_MAX_DOUBLES 131072
int speqg_refine_noalloc(speqg_scan * scan, const double * d_percent, const int32_t * d_group_counts, double * d_tkpg);

/* ------------------------------------------------------------------------------------------------------------------
 * Host-side solve (fp64, exact operation order of the reference).
 * speqg_unique_to_percent: speq::scan::unique_to_percent (src/fm_scanner.cpp:1455-1474).
 * speqg_solve: the loop of src/fm_scanner.cpp:238-279 driven by speqg_refine; returns the iteration count.
 * ------------------------------------------------------------------------------------------------------------------ */
int speqg_unique_to_percent(const double * unique_in_reads, uint64_t total_in_reads, const double * unique_in_refs,
                            const double * total_in_refs, uint32_t n_groups, double * out);

#ifdef __cplusplus
}
#endif
#endif /* SPEQ_CUDA_H */
