// seqio.h -- streaming FASTA / FASTQ (optionally gzip) reader that hands out raw, length-prefixed read batches
// (ASCII letters + phred+33 chars, flat byte arrays + offsets) ready for speqg_scan_submit_raw.
// Parsing rules follow what the reference gets from seqan3::sequence_file_input:
//   format by extension (.fa .fasta .fna .ffn .faa .frn / .fq .fastq, + .gz), FASTA ids start with '>' or ';',
//   sequence may span lines, whitespace (and digits in FASTA) skipped; FASTQ: '@' id, sequence lines up to '+',
//   then exactly |seq| quality characters.  Letters are validated on the device (non-IUPAC -> parse_error).
#pragma once
#include <cstdint>
#include <cstdio>
#include <filesystem>
#include <string>
#include <vector>

namespace speq::io
{
struct read_batch
{
    uint8_t * seq = nullptr;      // capacity bytes (pinned, owned by the caller)
    uint8_t * qual = nullptr;
    uint64_t * offsets = nullptr; // capacity reads + 1
    uint64_t cap_bytes = 0, cap_reads = 0;
    uint64_t n_reads = 0;         // filled
};

class sequence_reader
{
public:
    // start_offset: byte offset of a record start (plain files; used when the parallel FASTQ indexer hands over)
    explicit sequence_reader(std::filesystem::path const & p, uint64_t start_offset = 0);
    // FASTQ records held in memory: bytes [begin, end) of a mapped file (a region the device parser handed back)
    sequence_reader(const char * data, uint64_t begin, uint64_t end, std::string const & name);
    ~sequence_reader();
    sequence_reader(sequence_reader const &) = delete;
    // next record: returns false at end of file.  qual is filled with 'J' (phred 41) for FASTA records.
    bool next(std::string & seq, std::string & qual);
    bool is_fastq() const { return fastq_; }

private:
    bool getline(std::string & out);
    void * gz_ = nullptr;
    std::vector<char> buf_;
    size_t pos_ = 0, len_ = 0;
    bool eof_ = false, fastq_ = false, have_peek_ = false;
    const char * mem_ = nullptr;       // memory source: [mpos_, mend_)
    uint64_t mpos_ = 0, mend_ = 0;
    std::string peek_;
    std::string path_;
};

// ---- parallel ingest of plain (uncompressed) FASTQ in the strict 4-line form -------------------------------------
// The file is mapped; a region is cut into one piece per thread at record starts (a line that begins with '@' whose
// second successor begins with '+' -- a quality line can begin with '@', but then that line is a sequence line and
// cannot begin with '+'), every piece is indexed independently (memchr per line), and the records are then copied into
// the caller's pinned batch by the same threads.  Anything that is not the strict form (multi-line records, blanks
// inside a sequence line, length mismatch) stops the indexer in front of that record: the caller continues with the
// serial sequence_reader from that offset, which implements the full seqan3 grammar.
struct fq_rec
{
    uint64_t start;   // file offset of the '@'
    uint64_t seq;     // file offset of the sequence line
    uint64_t qual;    // file offset of the quality line
    uint64_t next;    // file offset of the following record
    uint32_t len;
};

class fastq_map
{
public:
    static bool eligible(std::filesystem::path const & p);     // .fq / .fastq, not compressed
    explicit fastq_map(std::filesystem::path const & p);
    ~fastq_map();
    fastq_map(fastq_map const &) = delete;
    // index records starting at `pos` (a record start) within about max_bytes; returns false when one is irregular
    // (out then holds the regular records before it).  out[i].next of the last record is where to go on.
    bool index(uint64_t pos, uint64_t max_bytes, int threads, std::vector<fq_rec> & out) const;
    const char * data() const { return data_; }
    uint64_t size() const { return size_; }
    // first record start at or after p (a line that begins with '@' whose second successor begins with '+'); size() if none
    uint64_t record_start_from(uint64_t p) const { return find_start(p); }
    // end of a region that starts at the record start `pos` and holds at most max_bytes: the last record start <= pos + max_bytes
    // (size() when the rest of the file fits); pos itself when not even one record fits
    uint64_t region_end(uint64_t pos, uint64_t max_bytes) const;
    // bytes [pos, pos + len) of the file into dst with `threads` concurrent pread()s: a plain kernel copy out of the page cache
    // (no page faults, unlike touching the mapping for the first time)
    // newlines != NULL: also the number of '\n' among the bytes, counted block by block while they are still in cache
    void read_into(void * dst, uint64_t pos, uint64_t len, int threads, uint64_t * newlines = nullptr) const;

private:
    uint64_t find_start(uint64_t p) const;
    const char * data_ = nullptr;
    uint64_t size_ = 0;
    int fd_ = -1;
};

// copy records [r0, r1) of `recs` (step `stride`, slot offset `slot`: mates of a pair interleave with stride 2) into the
// flat arrays; dst offsets come from `offsets` (already filled for every slot)
// newlines in [p, p + n), counted by `threads` threads; and the offset just behind the nth (>= 1) of them (n when there are fewer)
uint64_t count_newlines(const char * p, uint64_t n, int threads);
// the length of [p, p + n) without its last `drop` lines (the last one may lack its newline): a backward search, cheap when
// few lines go
uint64_t offset_without_last_lines(const char * p, uint64_t n, uint64_t drop);
uint64_t offset_after_newline(const char * p, uint64_t n, uint64_t nth, int threads);
// plain copy by `threads` threads (page cache -> pinned staging buffer)
void parallel_copy(void * dst, const void * src, uint64_t bytes, int threads);
void copy_records(const char * data, std::vector<fq_rec> const & recs, uint64_t n, uint8_t * seq, uint8_t * qual,
                  uint64_t const * offsets, uint64_t stride, uint64_t slot, int threads);

// whole FASTA file -> dna5 rank arrays (A0 C1 G2 N3 T4), one per record (the reference index input)
std::vector<std::vector<uint8_t>> read_reference_ranks(std::filesystem::path const & p);
}   // namespace speq::io
