// args.h -- command line of the `speq` host binary: sub-commands index | scan | all.
// Same surface as the reference (src/arg_parse.cpp:4-126, include/arg_parse.h:10-28): flag names, defaults, validators
// and the "argument parsing error" lines; seqan3's argument_parser itself (and its version phone-home) is not used.
#pragma once
#include <filesystem>
#include <string>

namespace speq::args
{
struct cmd_arguments
{
    std::filesystem::path in_file_reads_path_1{};
    std::filesystem::path in_file_reads_path_2{};
    std::filesystem::path in_file_references{};
    std::filesystem::path in_file_references_groups{};
    std::filesystem::path io_file_index{"output.idx"};
    std::filesystem::path out_file_path{"output.txt"};
    bool is_force{};
    unsigned int threads{2};
    unsigned int chunk{10000};
    unsigned int kmer{70};
    unsigned int phred_cutoff{30};
    double fixed_accuracy{0.0};
    double precision{1e-6};
    bool is_indexer{false};
    bool is_scanner{false};
    bool is_parsed{false};
    // additions of the B200 build (absent from the reference)
    unsigned int gpus{1};                 // --gpus N: devices the reads are sharded over (index replicated)
    unsigned int batch_reads{1u << 18};   // --batch-reads: reads per C-ABI submit (2 x 2 pinned buffers of 80 MB per device)
    std::filesystem::path json_path{};    // --json FILE: full-precision results for parity checks
    bool write_output{false};             // --write-output: write the per-variant table to -o/--output (the reference parses -o,
                                          // src/arg_parse.cpp:20, and never writes it; off by default for that reason)
};

cmd_arguments initialize_argument_parser(int argc, char ** argv);
void check_out_file(std::filesystem::path & a_path, bool is_force);
void check_in_file(std::filesystem::path & a_path);
}   // namespace speq::args
