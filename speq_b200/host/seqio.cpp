#include "seqio.h"

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include <zlib.h>

#include <algorithm>
#include <cctype>
#include <cstring>
#include <stdexcept>
#include <thread>

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

sequence_reader::sequence_reader(std::filesystem::path const & p, uint64_t start_offset) : buf_(1 << 20), path_(p.string())
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
    if (start_offset && gzseek(static_cast<gzFile>(gz_), static_cast<z_off_t>(start_offset), SEEK_SET) < 0)
        throw std::runtime_error("Could not seek in " + path_);
}

sequence_reader::sequence_reader(const char * data, uint64_t begin, uint64_t end, std::string const & name)
    : fastq_(true), mem_(data), mpos_(begin), mend_(end), path_(name)
{
}

sequence_reader::~sequence_reader()
{
    if (gz_) gzclose(static_cast<gzFile>(gz_));
}

bool sequence_reader::getline(std::string & out)
{
    out.clear();
    if (have_peek_) { out.swap(peek_); have_peek_ = false; return true; }
    if (mem_)
    {
        if (mpos_ >= mend_) return false;
        const char * b = mem_ + mpos_;
        const char * nl = static_cast<const char *>(memchr(b, '\n', mend_ - mpos_));
        const uint64_t len = nl ? static_cast<uint64_t>(nl - b) : mend_ - mpos_;
        out.assign(b, len);
        mpos_ += len + (nl ? 1 : 0);
        if (!out.empty() && out.back() == '\r') out.pop_back();
        return true;
    }
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

// ---------------------------------------------------------------------------------------------------------------
// parallel FASTQ
// ---------------------------------------------------------------------------------------------------------------
bool fastq_map::eligible(std::filesystem::path const & p)
{
    std::string name = p.filename().string();
    std::transform(name.begin(), name.end(), name.begin(), [](unsigned char c) { return std::tolower(c); });
    if (!has_ext(name, {".fq", ".fastq"})) return false;
    // a gzip stream under a plain name is still handled by zlib in the serial reader
    FILE * f = fopen(p.c_str(), "rb");
    if (!f) return false;
    unsigned char m[2] = {0, 0};
    size_t got = fread(m, 1, 2, f);
    fclose(f);
    return !(got == 2 && m[0] == 0x1f && m[1] == 0x8b);
}

fastq_map::fastq_map(std::filesystem::path const & p)
{
    fd_ = open(p.c_str(), O_RDONLY);
    if (fd_ < 0) throw std::runtime_error("Could not open file " + p.string() + " for reading.");
    struct stat st;
    if (fstat(fd_, &st) != 0) { close(fd_); throw std::runtime_error("Could not stat " + p.string()); }
    size_ = static_cast<uint64_t>(st.st_size);
    if (size_)
    {
        void * m = mmap(nullptr, size_, PROT_READ, MAP_PRIVATE, fd_, 0);
        if (m == MAP_FAILED) { close(fd_); throw std::runtime_error("Could not map " + p.string()); }
        madvise(m, size_, MADV_SEQUENTIAL);
        data_ = static_cast<const char *>(m);
    }
}

fastq_map::~fastq_map()
{
    if (data_) munmap(const_cast<char *>(data_), size_);
    if (fd_ >= 0) close(fd_);
}

namespace
{
// end of the line that starts at p (position of '\n', or size when the last line is unterminated)
inline uint64_t line_end(const char * d, uint64_t size, uint64_t p)
{
    const void * nl = memchr(d + p, '\n', size - p);
    return nl ? static_cast<uint64_t>(static_cast<const char *>(nl) - d) : size;
}
}   // namespace

uint64_t fastq_map::find_start(uint64_t p) const
{
    if (p == 0) return 0;
    // first line start at or after p
    uint64_t q = data_[p - 1] == '\n' ? p : line_end(data_, size_, p) + 1;
    while (q < size_)
    {
        uint64_t e0 = line_end(data_, size_, q);
        if (data_[q] == '@' && e0 < size_)
        {
            uint64_t e1 = line_end(data_, size_, e0 + 1);
            if (e1 < size_ && e1 + 1 < size_ && data_[e1 + 1] == '+') return q;
        }
        q = e0 + 1;
    }
    return size_;
}

uint64_t fastq_map::region_end(uint64_t pos, uint64_t max_bytes) const
{
    if (pos >= size_ || size_ - pos <= max_bytes) return size_;
    const uint64_t limit = pos + max_bytes;
    // walk record starts forward from a point shortly before the limit; widen the window when no start lies in it
    for (uint64_t back = 1ull << 16; ; back <<= 3)
    {
        const uint64_t from = limit > pos + back ? limit - back : pos;
        uint64_t s = from == pos ? pos : find_start(from);
        if (s <= limit && (s > pos || from == pos))
        {
            while (true)
            {
                const uint64_t nx = find_start(s + 1);
                if (nx > limit) break;
                s = nx;
            }
            return s;                  // == pos: the record at pos alone is larger than max_bytes
        }
        if (from == pos) return pos;
    }
}

void fastq_map::read_into(void * dst, uint64_t pos, uint64_t len, int threads, uint64_t * newlines) const
{
    threads = std::max(1, std::min<int>(threads, static_cast<int>(len >> 20) + 1));
    std::vector<char> failed(static_cast<size_t>(threads), 0);
    std::vector<uint64_t> nl(static_cast<size_t>(threads), 0);
    auto work = [&](int i)
    {
        uint64_t a = len * static_cast<uint64_t>(i) / static_cast<uint64_t>(threads), b = len * (static_cast<uint64_t>(i) + 1) / static_cast<uint64_t>(threads);
        // with a line count wanted: blocks small enough to still be in this core's cache when they are counted
        const uint64_t block = newlines ? (256u << 10) : ~uint64_t(0);
        while (a < b)
        {
            ssize_t r = pread(fd_, static_cast<char *>(dst) + a, std::min(b - a, block), static_cast<off_t>(pos + a));
            if (r <= 0) { failed[static_cast<size_t>(i)] = 1; return; }
            if (newlines)
            {
                const char * q = static_cast<const char *>(dst) + a, * e = q + r;
                uint64_t c = 0;
                while (q < e && (q = static_cast<const char *>(memchr(q, '\n', static_cast<size_t>(e - q)))) != nullptr) { ++c; ++q; }
                nl[static_cast<size_t>(i)] += c;
            }
            a += static_cast<uint64_t>(r);
        }
    };
    if (threads == 1) work(0);
    else
    {
        std::vector<std::thread> th;
        for (int i = 0; i < threads; ++i) th.emplace_back(work, i);
        for (auto & t : th) t.join();
    }
    for (char f : failed)
        if (f) throw std::runtime_error("read error in the FASTQ file");
    if (newlines)
    {
        *newlines = 0;
        for (uint64_t c : nl) *newlines += c;
    }
}

namespace
{
uint64_t count_nl_serial(const char * p, uint64_t n)
{
    uint64_t c = 0;
    const char * e = p + n;
    while (p < e)
    {
        const char * q = static_cast<const char *>(memchr(p, '\n', static_cast<size_t>(e - p)));
        if (!q) break;
        ++c;
        p = q + 1;
    }
    return c;
}
std::vector<uint64_t> piece_counts(const char * p, uint64_t n, int threads, std::vector<uint64_t> & cut)
{
    threads = std::max(1, std::min<int>(threads, static_cast<int>(n >> 20) + 1));
    cut.assign(static_cast<size_t>(threads) + 1, 0);
    for (int i = 0; i <= threads; ++i) cut[static_cast<size_t>(i)] = n * static_cast<uint64_t>(i) / static_cast<uint64_t>(threads);
    std::vector<uint64_t> cnt(static_cast<size_t>(threads), 0);
    if (threads == 1) cnt[0] = count_nl_serial(p, n);
    else
    {
        std::vector<std::thread> th;
        for (int i = 0; i < threads; ++i)
            th.emplace_back([&, i]() { cnt[static_cast<size_t>(i)] = count_nl_serial(p + cut[static_cast<size_t>(i)], cut[static_cast<size_t>(i) + 1] - cut[static_cast<size_t>(i)]); });
        for (auto & t : th) t.join();
    }
    return cnt;
}
}   // namespace

uint64_t count_newlines(const char * p, uint64_t n, int threads)
{
    std::vector<uint64_t> cut;
    uint64_t total = 0;
    for (uint64_t c : piece_counts(p, n, threads, cut)) total += c;
    return total;
}

uint64_t offset_after_newline(const char * p, uint64_t n, uint64_t nth, int threads)
{
    std::vector<uint64_t> cut;
    std::vector<uint64_t> cnt = piece_counts(p, n, threads, cut);
    uint64_t before = 0;
    for (size_t i = 0; i < cnt.size(); ++i)
    {
        if (before + cnt[i] >= nth)
        {
            const char * q = p + cut[i];
            const char * e = p + cut[i + 1];
            for (uint64_t left = nth - before; left; --left)
            {
                q = static_cast<const char *>(memchr(q, '\n', static_cast<size_t>(e - q)));
                ++q;
            }
            return static_cast<uint64_t>(q - p);
        }
        before += cnt[i];
    }
    return n;
}

uint64_t offset_without_last_lines(const char * p, uint64_t n, uint64_t drop)
{
    uint64_t end = n;
    for (; drop && end; --drop)
    {
        const uint64_t upto = p[end - 1] == '\n' ? end - 1 : end;     // the line's own terminator is not the previous line's
        const void * q = memrchr(p, '\n', static_cast<size_t>(upto));
        end = q ? static_cast<uint64_t>(static_cast<const char *>(q) - p) + 1 : 0;
    }
    return end;
}

void parallel_copy(void * dst, const void * src, uint64_t bytes, int threads)
{
    threads = std::max(1, threads);
    if (threads == 1 || bytes < (8u << 20))
    {
        memcpy(dst, src, bytes);
        return;
    }
    std::vector<std::thread> th;
    for (int i = 0; i < threads; ++i)
    {
        const uint64_t a = bytes * static_cast<uint64_t>(i) / static_cast<uint64_t>(threads) & ~uint64_t(4095);
        const uint64_t b = i + 1 == threads ? bytes : (bytes * (static_cast<uint64_t>(i) + 1) / static_cast<uint64_t>(threads) & ~uint64_t(4095));
        th.emplace_back([=]() { memcpy(static_cast<char *>(dst) + a, static_cast<const char *>(src) + a, b - a); });
    }
    for (auto & t : th) t.join();
}

bool fastq_map::index(uint64_t pos, uint64_t max_bytes, int threads, std::vector<fq_rec> & out) const
{
    out.clear();
    if (pos >= size_) return true;
    const uint64_t end = std::min(size_, pos + max_bytes);
    threads = std::max(1, threads);
    // piece boundaries at record starts
    std::vector<uint64_t> cut(static_cast<size_t>(threads) + 1);
    cut[0] = pos;
    for (int i = 1; i < threads; ++i)
    {
        uint64_t want = pos + (end - pos) / static_cast<uint64_t>(threads) * static_cast<uint64_t>(i);
        cut[static_cast<size_t>(i)] = std::max(cut[static_cast<size_t>(i) - 1], find_start(want));
    }
    cut[static_cast<size_t>(threads)] = end == size_ ? size_ : std::max(cut[static_cast<size_t>(threads) - 1], find_start(end));
    std::vector<std::vector<fq_rec>> part(static_cast<size_t>(threads));
    std::vector<char> bad(static_cast<size_t>(threads), 0);
    auto work = [&](int ti)
    {
        const char * d = data_;
        uint64_t p = cut[static_cast<size_t>(ti)], stop = cut[static_cast<size_t>(ti) + 1];
        auto & v = part[static_cast<size_t>(ti)];
        v.reserve((stop - p) / 200 + 16);
        while (p < stop)
        {
            if (d[p] == '\n') { ++p; continue; }                 // blank line between records
            if (d[p] == '\r' && p + 1 < size_ && d[p + 1] == '\n') { p += 2; continue; }
            fq_rec r;
            r.start = p;
            if (d[p] != '@') { bad[static_cast<size_t>(ti)] = 1; return; }
            uint64_t e0 = line_end(d, size_, p);
            if (e0 >= size_) { bad[static_cast<size_t>(ti)] = 1; return; }
            r.seq = e0 + 1;
            uint64_t e1 = line_end(d, size_, r.seq);
            if (e1 >= size_ || e1 + 1 >= size_ || d[e1 + 1] != '+') { bad[static_cast<size_t>(ti)] = 1; return; }
            uint64_t e2 = line_end(d, size_, e1 + 1);
            if (e2 >= size_) { bad[static_cast<size_t>(ti)] = 1; return; }
            r.qual = e2 + 1;
            uint64_t e3 = line_end(d, size_, r.qual);
            uint64_t ls = e1 - r.seq, lq = e3 - r.qual;
            if (ls && d[e1 - 1] == '\r') --ls;
            if (lq && e3 > r.qual && d[e3 - 1] == '\r') --lq;
            // the strict form: equal lengths, no blanks inside the sequence or quality line
            if (ls != lq || ls > 0xFFFFFFFFull || memchr(d + r.seq, ' ', ls) || memchr(d + r.seq, '\t', ls) ||
                memchr(d + r.qual, ' ', lq) || memchr(d + r.qual, '\t', lq))
            {
                bad[static_cast<size_t>(ti)] = 1;
                return;
            }
            r.len = static_cast<uint32_t>(ls);
            r.next = std::min(size_, e3 + 1);
            v.push_back(r);
            p = r.next;
        }
    };
    if (threads == 1) work(0);
    else
    {
        std::vector<std::thread> th;
        for (int i = 0; i < threads; ++i) th.emplace_back(work, i);
        for (auto & t : th) t.join();
    }
    for (int i = 0; i < threads; ++i)
    {
        auto & v = part[static_cast<size_t>(i)];
        out.insert(out.end(), v.begin(), v.end());
        if (bad[static_cast<size_t>(i)]) return false;             // records of later pieces are dropped: the caller resumes serially
    }
    return true;
}

void copy_records(const char * data, std::vector<fq_rec> const & recs, uint64_t n, uint8_t * seq, uint8_t * qual,
                  uint64_t const * offsets, uint64_t stride, uint64_t slot, int threads)
{
    threads = std::max(1, threads);
    auto work = [&](int ti)
    {
        uint64_t a = n * static_cast<uint64_t>(ti) / static_cast<uint64_t>(threads), b = n * (static_cast<uint64_t>(ti) + 1) / static_cast<uint64_t>(threads);
        for (uint64_t i = a; i < b; ++i)
        {
            fq_rec const & r = recs[i];
            uint64_t o = offsets[i * stride + slot];
            memcpy(seq + o, data + r.seq, r.len);
            memcpy(qual + o, data + r.qual, r.len);
        }
    };
    if (threads == 1 || n < 4096) { for (int i = 0; i < threads; ++i) work(i); return; }
    std::vector<std::thread> th;
    for (int i = 0; i < threads; ++i) th.emplace_back(work, i);
    for (auto & t : th) t.join();
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
