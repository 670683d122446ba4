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
    explicit sequence_reader(std::filesystem::path const & p);
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
    std::string peek_;
    std::string path_;
};

// whole FASTA file -> dna5 rank arrays (A0 C1 G2 N3 T4), one per record (the reference index input)
std::vector<std::vector<uint8_t>> read_reference_ranks(std::filesystem::path const & p);
}   // namespace speq::io
