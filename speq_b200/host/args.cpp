// args.cpp -- see args.h.  Behaviour follows the reference's src/arg_parse.cpp; the parser is our own.
#include "args.h"

#include <cstdlib>
#include <iostream>
#include <map>
#include <sstream>
#include <stdexcept>
#include <thread>
#include <vector>

namespace speq::args
{
namespace
{
struct parser_error : std::runtime_error
{
    using std::runtime_error::runtime_error;
};

struct option
{
    char short_id;
    std::string long_id;
    bool is_flag;
};

// which options each sub-command accepts (src/arg_parse.cpp:15-29, 59-71, 102-109)
const std::vector<option> kAll = {{'1', "forward", false}, {'2', "reverse", false}, {'r', "reference", false}, {'g', "groups", false},
                                  {'x', "index", false}, {'o', "output", false}, {'f', "force", true}, {'k', "kmer", false},
                                  {0, "precision-cutoff", false}, {0, "phred-cutoff", false}, {0, "fixed-accuracy", false},
                                  {'t', "threads", false}, {0, "gpus", false}, {0, "batch-reads", false}, {0, "json", false},
                                  {0, "write-output", true}};
const std::vector<option> kScan = {{'1', "forward", false}, {'2', "reverse", false}, {'x', "index", false}, {'o', "output", false},
                                   {'f', "force", true}, {'k', "kmer", false}, {0, "precision-cutoff", false},
                                   {0, "phred-cutoff", false}, {0, "fixed-accuracy", false}, {'t', "threads", false},
                                   {0, "gpus", false}, {0, "batch-reads", false}, {0, "json", false}, {0, "write-output", true}};
const std::vector<option> kIndex = {{'r', "reference", false}, {'g', "groups", false}, {'x', "index", false}, {'f', "force", true},
                                    {'t', "threads", false}};

std::string opt_name(option const & o)
{
    return o.short_id ? std::string("-") + o.short_id + "/--" + o.long_id : "--" + o.long_id;
}

template <class T> T to_number(std::string const & v, option const & o)
{
    std::istringstream is(v);
    T x{};
    is >> x;
    if (is.fail() || !is.eof()) throw parser_error("Value parse failed for " + opt_name(o) + ": Argument " + v + " could not be parsed as type " +
                                                   (std::is_floating_point<T>::value ? "double." : "unsigned 32 bit integer."));
    return x;
}

void range_check(double v, double lo, double hi, option const & o)
{
    if (v < lo || v > hi)
        throw parser_error("Validation failed for option " + opt_name(o) + ": Value " + std::to_string(v) + " is not in range [" +
                           std::to_string(lo) + "," + std::to_string(hi) + "].");
}

void assign(cmd_arguments & a, option const & o, std::string const & v)
{
    std::string const & n = o.long_id;
    if (n == "forward") a.in_file_reads_path_1 = v;
    else if (n == "reverse") a.in_file_reads_path_2 = v;
    else if (n == "reference") a.in_file_references = v;
    else if (n == "groups") a.in_file_references_groups = v;
    else if (n == "index") a.io_file_index = v;
    else if (n == "output") a.out_file_path = v;
    else if (n == "force") a.is_force = true;
    else if (n == "kmer") a.kmer = to_number<unsigned int>(v, o);
    else if (n == "precision-cutoff") a.precision = to_number<double>(v, o);
    else if (n == "phred-cutoff") a.phred_cutoff = to_number<unsigned int>(v, o);
    else if (n == "fixed-accuracy")
    {
        a.fixed_accuracy = to_number<double>(v, o);
        range_check(a.fixed_accuracy, 0.0, 1.0, o);                                        // src/arg_parse.cpp:27
    }
    else if (n == "threads")
    {
        a.threads = to_number<unsigned int>(v, o);
        range_check(a.threads, 2, static_cast<double>(std::thread::hardware_concurrency()), o);   // src/arg_parse.cpp:29
    }
    else if (n == "gpus")
    {
        a.gpus = to_number<unsigned int>(v, o);
        range_check(a.gpus, 1, 64, o);
    }
    else if (n == "batch-reads")
    {
        a.batch_reads = to_number<unsigned int>(v, o);
        range_check(a.batch_reads, 2, 1u << 28, o);
    }
    else if (n == "json") a.json_path = v;
    else if (n == "write-output") a.write_output = true;
}

void parse_options(cmd_arguments & a, std::vector<option> const & opts, std::vector<std::string> const & argv)
{
    for (size_t i = 0; i < argv.size(); ++i)
    {
        std::string const & tok = argv[i];
        option const * hit = nullptr;
        std::string value;
        bool have_value = false;
        if (tok.rfind("--", 0) == 0)
        {
            std::string name = tok.substr(2);
            auto eq = name.find('=');
            if (eq != std::string::npos) { value = name.substr(eq + 1); name = name.substr(0, eq); have_value = true; }
            for (auto const & o : opts) if (o.long_id == name) hit = &o;
            if (!hit) throw parser_error("Unknown option " + tok + ". In case this is meant to be a non-option/argument/parameter, please specify the start of non-options with '--'. See -h/--help for program information.");
        }
        else if (tok.size() >= 2 && tok[0] == '-')
        {
            for (auto const & o : opts) if (o.short_id && o.short_id == tok[1]) hit = &o;
            if (!hit) throw parser_error("Unknown option " + tok.substr(0, 2) + ". In case this is meant to be a non-option/argument/parameter, please specify the start of non-options with '--'. See -h/--help for program information.");
            if (tok.size() > 2) { value = tok.substr(2); if (!value.empty() && value[0] == '=') value.erase(0, 1); have_value = true; }
        }
        else
            throw parser_error("Too many arguments provided. Please see -h/--help for more information.");
        if (hit->is_flag) { assign(a, *hit, ""); continue; }
        if (!have_value)
        {
            if (i + 1 >= argv.size()) throw parser_error("Missing value for option " + opt_name(*hit));
            value = argv[++i];
        }
        assign(a, *hit, value);
    }
}

void help(std::string const & sub)
{
    std::cerr << "speq " << sub << " [options]    (B200 build of SPeQ's scan hot path; flags as in the reference, version 0.1.0)\n"
                 "  -1/--forward FILE  -2/--reverse FILE  -r/--reference FASTA  -g/--groups FILE  -x/--index PREFIX  -o/--output FILE  -f/--force\n"
                 "  -k/--kmer N (70)  --precision-cutoff X (1e-06)  --phred-cutoff N (30)  --fixed-accuracy X (0 = use phred)  -t/--threads N (2)\n"
                 "  --gpus N (1)  --batch-reads N  --json FILE  --write-output (write the per-variant table to -o FILE)\n";
}
}   // namespace

void check_in_file(std::filesystem::path & a_path)
{
    namespace fs = std::filesystem;
    if (a_path.empty()) return;
    if (a_path.is_relative())
    {
        if (fs::exists(fs::current_path() / a_path)) a_path = fs::current_path() / a_path;
        else throw parser_error("The relative-path file \"" + a_path.string() + "\" was not found.");
    }
    else if (!fs::exists(a_path))
        throw parser_error("The full-path file \"" + a_path.string() + "\" was not found.");
}

void check_out_file(std::filesystem::path & a_path, bool is_force)
{
    namespace fs = std::filesystem;
    if (a_path.is_relative())
    {
        a_path = fs::current_path() / a_path;
        if (fs::exists(a_path) && !is_force)
            throw parser_error("Cowardly refusing to use an existing output file. Use '-f' to overwrite.");
    }
}

cmd_arguments initialize_argument_parser(int argc, char ** argv)
{
    cmd_arguments args{};
    std::string sub = argc >= 2 ? argv[1] : "";
    if (sub != "index" && sub != "scan" && sub != "all")
    {
        if (sub == "-h" || sub == "--help") { help("[index|scan|all]"); return args; }
        std::cerr << "SPeQ Error: " << (sub.empty() ? std::string("Please specify which sub program you want to use (one of [index,scan,all]). Use -h/--help for more information.")
                                                    : "You either forgot or misspelled the subcommand! Please specify which sub program you want to use (one of [index,scan,all]). Use -h/--help for more information.")
                  << "\n";
        return args;
    }
    std::vector<std::string> rest(argv + 2, argv + argc);
    for (auto const & r : rest)
        if (r == "-h" || r == "--help") { help(sub); return args; }
    args.is_indexer = (sub == "index" || sub == "all");
    args.is_scanner = (sub == "scan" || sub == "all");
    try
    {
        parse_options(args, sub == "all" ? kAll : (sub == "scan" ? kScan : kIndex), rest);
        if (sub == "index")
        {
            check_in_file(args.in_file_references);            // src/arg_parse.cpp:115-117
            check_in_file(args.in_file_references_groups);
            check_out_file(args.out_file_path, args.is_force);
        }
        else
        {
            check_in_file(args.in_file_reads_path_1);          // src/arg_parse.cpp:35-37, 77-80
            check_in_file(args.in_file_reads_path_2);
            if (sub == "scan") check_in_file(args.io_file_index);
            check_out_file(args.out_file_path, args.is_force);
        }
        args.is_parsed = true;
    }
    catch (parser_error const & e)
    {
        std::cerr << "speq " << sub << " | argument parsing error:  " << e.what() << "\n";
    }
    return args;
}
}   // namespace speq::args
