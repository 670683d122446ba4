#include "seqio.h"

#include <zlib.h>

#include <algorithm>
#include <cctype>
#include <cstring>
#include <stdexcept>

namespace speq::io
{
namespace
{
bool has_ext(std::string const & name, std::initializer_list<const char *> exts)
{
    for (auto e : exts)
    {
        size_t n = strlen(e);
        if (name.size() >= n && name.compare(name.size() - n, n, e) == 0) return true;
    }
    return false;
}
}   // namespace

sequence_reader::sequence_reader(std::filesystem::path const & p) : buf_(1 << 20), path_(p.string())
{
    std::string name = p.filename().string();
    std::transform(name.begin(), name.end(), name.begin(), [](unsigned char c) { return std::tolower(c); });
    if (has_ext(name, {".gz", ".bgzf"})) name = name.substr(0, name.rfind('.'));
    if (has_ext(name, {".fq", ".fastq"})) fastq_ = true;
    else if (has_ext(name, {".fa", ".fasta", ".fna", ".ffn", ".faa", ".frn", ".fas"})) fastq_ = false;
    else throw std::runtime_error("No valid format found for this extension: " + path_);   // seqan3::unhandled_extension_error
    gz_ = gzopen(path_.c_str(), "rb");   // zlib reads plain files transparently
    if (!gz_) throw std::runtime_error("Could not open file " + path_ + " for reading.");
    gzbuffer(static_cast<gzFile>(gz_), 1 << 20);
}

sequence_reader::~sequence_reader()
{
    if (gz_) gzclose(static_cast<gzFile>(gz_));
}

bool sequence_reader::getline(std::string & out)
{
    out.clear();
    if (have_peek_) { out.swap(peek_); have_peek_ = false; return true; }
    bool got = false;
    while (true)
    {
        if (pos_ == len_)
        {
            if (eof_) break;
            int r = gzread(static_cast<gzFile>(gz_), buf_.data(), static_cast<unsigned>(buf_.size()));
            if (r < 0) throw std::runtime_error("read error in " + path_);
            if (r == 0) { eof_ = true; break; }
            len_ = static_cast<size_t>(r);
            pos_ = 0;
        }
        char * b = buf_.data() + pos_;
        char * nl = static_cast<char *>(memchr(b, '\n', len_ - pos_));
        got = true;
        if (nl)
        {
            out.append(b, nl);
            pos_ = static_cast<size_t>(nl - buf_.data()) + 1;
            break;
        }
        out.append(b, buf_.data() + len_);
        pos_ = len_;
    }
    if (!out.empty() && out.back() == '\r') out.pop_back();
    return got;
}

bool sequence_reader::next(std::string & seq, std::string & qual)
{
    seq.clear();
    qual.clear();
    std::string line;
    do
    {
        if (!getline(line)) return false;
    } while (line.empty());
    if (fastq_)
    {
        if (line[0] != '@') throw std::runtime_error("parse_error: FASTQ record does not start with '@' in " + path_);
        while (true)
        {
            if (!getline(line)) throw std::runtime_error("parse_error: unexpected end of input in " + path_);
            if (!line.empty() && line[0] == '+') break;
            for (char c : line) if (!std::isspace(static_cast<unsigned char>(c))) seq.push_back(c);
        }
        while (qual.size() < seq.size())
        {
            if (!getline(line)) throw std::runtime_error("parse_error: unexpected end of input in " + path_);
            for (char c : line) if (!std::isspace(static_cast<unsigned char>(c))) qual.push_back(c);
        }
        if (qual.size() != seq.size()) throw std::runtime_error("parse_error: quality length differs from sequence length in " + path_);
    }
    else
    {
        if (line[0] != '>' && line[0] != ';') throw std::runtime_error("parse_error: FASTA record does not start with '>' in " + path_);
        while (getline(line))
        {
            if (!line.empty() && (line[0] == '>' || line[0] == ';')) { peek_.swap(line); have_peek_ = true; break; }
            for (char c : line)
                if (!std::isspace(static_cast<unsigned char>(c)) && !std::isdigit(static_cast<unsigned char>(c))) seq.push_back(c);
        }
        qual.assign(seq.size(), 'J');
    }
    return true;
}

std::vector<std::vector<uint8_t>> read_reference_ranks(std::filesystem::path const & p)
{
    // dna5 ranks by letter (deps/seqan3/include/seqan3/alphabet/nucleotide/dna5.hpp:87-119); the legal letters are the
    // IUPAC (dna15) set of the sequence file (deps/seqan3/include/seqan3/io/sequence_file/input.hpp:145-162)
    uint8_t rank[256];
    memset(rank, 255, sizeof rank);
    for (const char * q = "ACGTUNRYSWKMBDHVacgtunryswkmbdhv"; *q; ++q) rank[static_cast<unsigned char>(*q)] = 3;
    rank['A'] = rank['a'] = 0; rank['C'] = rank['c'] = 1; rank['G'] = rank['g'] = 2;
    rank['T'] = rank['t'] = rank['U'] = rank['u'] = 4;
    sequence_reader rd(p);
    std::vector<std::vector<uint8_t>> out;
    std::string s, q;
    while (rd.next(s, q))
    {
        std::vector<uint8_t> r(s.size());
        for (size_t i = 0; i < s.size(); ++i)
        {
            r[i] = rank[static_cast<unsigned char>(s[i])];
            if (r[i] == 255) throw std::runtime_error(std::string("parse_error: illegal sequence character '") + s[i] + "' in " + p.string());
        }
        out.push_back(std::move(r));
    }
    return out;
}
}   // namespace speq::io
