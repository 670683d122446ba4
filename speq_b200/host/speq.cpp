// speq.cpp -- the `speq index | scan | all` host binary of the B200 build.
//
// Keeps the reference's surface for the scan hot path -- sub-commands, flags, `.idx` / `<stem>_<k>mer.dat` caching with
// the same staleness rules, and the stderr report -- and calls libspeq_cuda.so (include/speq_cuda.h) for everything that
// computes.  Reference entry points mirrored (paths relative to the reference checkout):
//   main                      src/main.cpp:15-40
//   speq::fm::index           src/fm_indexer.cpp:55-110   (generate_fm_index :8-52)
//   speq::scan::async_one/two src/fm_scanner.cpp:5-33, with the four _async_* bodies :35-279, 310-545, 548-792, 795-1065
//   back-check                src/fm_scanner.cpp:1476-1576
// There is no CPU path: without a CUDA device every compute call fails and the message is printed.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <unistd.h>
#include <cstring>
#include <fstream>
#include <iostream>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

#include "../../include/speq_cuda.h"
#include "args.h"
#include "groups.h"
#include "seqio.h"

namespace fs = std::filesystem;
using speq::args::cmd_arguments;

namespace
{
void ck(int rc, const char * what)
{
    if (rc != SPEQG_OK) throw std::runtime_error(std::string(what) + ": " + speqg_last_error());
}

// seqan3::debug_stream prints ranges as [a,b,c] and arithmetic values with the ostream defaults
template <class T> std::ostream & operator<<(std::ostream & os, std::vector<T> const & v)
{
    os << '[';
    for (size_t i = 0; i < v.size(); ++i) os << (i ? "," : "") << v[i];
    return os << ']';
}

int64_t mtime_of(fs::path const & p)
{
    return static_cast<int64_t>(fs::last_write_time(p).time_since_epoch().count());
}

// ---- .idx container: metadata the reference archives in front of the index (src/fm_indexer.cpp:39-50), then our payload
const char kIdxMagic[8] = {'S', 'P', 'Q', 'I', 'D', 'X', '0', '2'};

struct idx_meta
{
    std::string scaffold_file;
    int64_t scaffold_time = 0, group_time = 0;
    std::vector<std::string> group_names;
    std::vector<int> group_scaffolds, group_counts;
    uint64_t payload_offset = 0;
    bool reference_format = false;   // the file is the reference's own cereal + sdsl archive (read through speqg_refidx_*)
    std::shared_ptr<speqg_refidx> ref;   // ... and stays open (texts decoded once) until the index over them is built
};

void put_u64(std::ostream & os, uint64_t v) { os.write(reinterpret_cast<const char *>(&v), 8); }
uint64_t get_u64(std::istream & is)
{
    uint64_t v = 0;
    is.read(reinterpret_cast<char *>(&v), 8);
    if (!is) throw std::runtime_error("truncated index file");
    return v;
}
void put_str(std::ostream & os, std::string const & s) { put_u64(os, s.size()); os.write(s.data(), static_cast<std::streamsize>(s.size())); }
std::string get_str(std::istream & is)
{
    uint64_t n = get_u64(is);
    if (n > (1u << 20)) throw std::runtime_error("corrupt index file");
    std::string s(n, '\0');
    is.read(s.data(), static_cast<std::streamsize>(n));
    if (!is) throw std::runtime_error("truncated index file");
    return s;
}
void put_ints(std::ostream & os, std::vector<int> const & v)
{
    put_u64(os, v.size());
    os.write(reinterpret_cast<const char *>(v.data()), static_cast<std::streamsize>(v.size() * sizeof(int)));
}
std::vector<int> get_ints(std::istream & is)
{
    uint64_t n = get_u64(is);
    if (n > (1u << 28)) throw std::runtime_error("corrupt index file");
    std::vector<int> v(n);
    is.read(reinterpret_cast<char *>(v.data()), static_cast<std::streamsize>(n * sizeof(int)));
    if (!is) throw std::runtime_error("truncated index file");
    return v;
}

uint64_t write_meta(fs::path const & p, idx_meta const & m)
{
    std::ofstream os(p, std::ios::binary | std::ios::trunc);
    if (!os) throw std::runtime_error("cannot write " + p.string());
    os.write(kIdxMagic, 8);
    put_str(os, m.scaffold_file);
    put_u64(os, static_cast<uint64_t>(m.scaffold_time));
    put_u64(os, static_cast<uint64_t>(m.group_time));
    put_u64(os, m.group_names.size());
    for (auto const & s : m.group_names) put_str(os, s);
    put_ints(os, m.group_scaffolds);
    put_ints(os, m.group_counts);
    uint64_t off = static_cast<uint64_t>(os.tellp());
    off = (off + 8 + 63) / 64 * 64;   // payload starts 64-byte aligned; its offset is stored just before the padding
    put_u64(os, off);
    while (static_cast<uint64_t>(os.tellp()) < off) os.put('\0');
    return off;
}

idx_meta read_meta(fs::path const & p)
{
    std::ifstream is(p, std::ios::binary);
    if (!is) throw std::runtime_error("cannot open index file " + p.string());
    char mg[8];
    is.read(mg, 8);
    idx_meta m;
    if (!is || memcmp(mg, kIdxMagic, 8) != 0)
    {
        // not this build's container: an index written by the reference's `speq index` (src/fm_indexer.cpp:39-50)?
        is.close();
        speqg_refidx * r = nullptr;
        if (speqg_refidx_open(p.c_str(), &r) != SPEQG_OK)
            throw std::runtime_error(p.string() + " is neither an index of this build nor one written by the reference: " + speqg_last_error());
        m.scaffold_file = speqg_refidx_reference_path(r);
        m.scaffold_time = speqg_refidx_file_time(r, 0);
        m.group_time = speqg_refidx_file_time(r, 1);
        for (uint32_t g = 0; g < speqg_refidx_n_groups(r); ++g)
        {
            m.group_names.push_back(speqg_refidx_group_name(r, g));
            m.group_counts.push_back(speqg_refidx_group_count(r, g));
        }
        for (uint32_t j = 0; j < speqg_refidx_n_scaffolds(r); ++j) m.group_scaffolds.push_back(speqg_refidx_group_of_scaffold(r, j));
        m.reference_format = true;
        m.ref = std::shared_ptr<speqg_refidx>(r, speqg_refidx_close);
        return m;
    }
    m.scaffold_file = get_str(is);
    m.scaffold_time = static_cast<int64_t>(get_u64(is));
    m.group_time = static_cast<int64_t>(get_u64(is));
    uint64_t n = get_u64(is);
    for (uint64_t i = 0; i < n; ++i) m.group_names.push_back(get_str(is));
    m.group_scaffolds = get_ints(is);
    m.group_counts = get_ints(is);
    m.payload_offset = get_u64(is);
    return m;
}

struct index_handle
{
    speqg_index * h = nullptr;
    ~index_handle() { speqg_index_free(h); }
};

// ---- speq::fm::generate_fm_index (src/fm_indexer.cpp:8-52)
void generate_fm_index(cmd_arguments & args, int64_t scaffold_time, int64_t group_time, std::vector<std::string> const & group_names,
                       std::vector<int> const & group_scaffolds, std::vector<int> const & group_counts)
{
    auto scaffolds = speq::io::read_reference_ranks(args.in_file_references);
    // every sequence twice; odd copies <- complement(reverse)  (src/fm_indexer.cpp:19-33)
    static const uint8_t comp[5] = {4, 2, 1, 3, 0};
    std::vector<std::vector<uint8_t>> texts;
    std::vector<int32_t> group_of_text;
    for (size_t j = 0; j < scaffolds.size(); ++j)
    {
        std::vector<uint8_t> rc(scaffolds[j].rbegin(), scaffolds[j].rend());
        for (auto & c : rc) c = comp[c];
        int g = j < group_scaffolds.size() ? group_scaffolds[j] : -1;
        texts.push_back(std::move(scaffolds[j]));
        texts.push_back(std::move(rc));
        group_of_text.push_back(g);
        group_of_text.push_back(g);
    }
    std::vector<const uint8_t *> ptrs;
    std::vector<uint64_t> lens;
    for (auto const & t : texts) { ptrs.push_back(t.data()); lens.push_back(t.size()); }
    index_handle idx;
    // with a CUDA device the suffixes are sorted there (prefix doubling); without one SA-IS runs on the host -- same index
    int ndev = 0;
    if (speqg_device_count(&ndev) != SPEQG_OK) ndev = 0;
    ck(speqg_index_build(ptrs.data(), lens.data(), static_cast<uint32_t>(texts.size()), group_of_text.data(),
                         static_cast<uint32_t>(group_names.size()), ndev > 0 ? SPEQG_BUILD_DEVICE_SA : 0u, &idx.h), "speqg_index_build");
    args.io_file_index.replace_extension(".idx");
    idx_meta m{args.in_file_references.string(), scaffold_time, group_time, group_names, group_scaffolds, group_counts, 0};
    uint64_t off = write_meta(args.io_file_index, m);
    ck(speqg_index_save_at(idx.h, args.io_file_index.c_str(), off), "speqg_index_save_at");
}

// ---- speq::fm::index (src/fm_indexer.cpp:55-110)
int do_index(cmd_arguments & args)
{
    int64_t scaffold_time = mtime_of(args.in_file_references);
    int64_t group_time = mtime_of(args.in_file_references_groups);
    std::vector<int> group_scaffolds, group_counts;
    std::vector<std::string> group_names;
    speq::file_to_map(args.in_file_references_groups, group_names, group_scaffolds, group_counts);
    bool rebuild = true;
    if (fs::exists(args.io_file_index) && !args.is_force)
    {
        args.io_file_index.replace_extension(".idx");
        idx_meta old = read_meta(args.io_file_index);
        rebuild = fs::path(old.scaffold_file) != args.in_file_references || old.group_time != group_time || old.scaffold_time != scaffold_time;
    }
    if (rebuild) generate_fm_index(args, scaffold_time, group_time, group_names, group_scaffolds, group_counts);
    return 0;
}

// ---- pinned double buffers per device
struct host_batch
{
    uint8_t * seq = nullptr;
    uint8_t * qual = nullptr;
    uint64_t * offsets = nullptr;
    uint64_t cap_bytes = 0, cap_reads = 0, n_reads = 0, n_bytes = 0;
    // with_records = false: only the byte buffer (raw FASTQ regions for the device parser); the quality and offset arrays of
    // the record paths are pinned on first use (ensure_records) -- pinning is a large part of a short run's start-up
    void alloc(uint64_t bytes, uint64_t reads, bool with_records = true)
    {
        release();
        void * p = nullptr;
        ck(speqg_host_alloc(bytes, &p), "speqg_host_alloc"); seq = static_cast<uint8_t *>(p);
        cap_bytes = bytes; cap_reads = reads; n_reads = 0; n_bytes = 0;
        if (with_records) ensure_records();
    }
    void ensure_records()
    {
        if (qual) return;
        void * p = nullptr;
        ck(speqg_host_alloc(cap_bytes, &p), "speqg_host_alloc"); qual = static_cast<uint8_t *>(p);
        ck(speqg_host_alloc((cap_reads + 1) * 8, &p), "speqg_host_alloc"); offsets = static_cast<uint64_t *>(p);
        offsets[0] = 0;
    }
    void release()
    {
        speqg_host_free(seq); speqg_host_free(qual); speqg_host_free(offsets);
        seq = qual = nullptr; offsets = nullptr;
    }
    bool fits(uint64_t len, uint64_t n) const { return n_reads + n <= cap_reads && n_bytes + len <= cap_bytes; }
    void push(std::string const & s, std::string const & q)
    {
        memcpy(seq + n_bytes, s.data(), s.size());
        memcpy(qual + n_bytes, q.data(), q.size());
        n_bytes += s.size();
        offsets[++n_reads] = n_bytes;
    }
    void clear() { n_reads = 0; n_bytes = 0; if (offsets) offsets[0] = 0; }
};

struct device_lane
{
    speqg_scan * scan = nullptr;
    host_batch buf[2];
    int cur = 0;
    // raw FASTQ regions parsed on the device: copy ticket of each staging buffer, and where every submitted region lies in the file
    uint64_t ticket[2] = {0, 0};
    std::vector<std::pair<uint64_t, uint64_t>> regions;
    std::vector<std::pair<uint64_t, uint64_t>> regions2;     // paired input: the region of the second mates' file
};

struct scan_out
{
    std::vector<uint64_t> unique_int;
    std::vector<double> unique_w;
    uint64_t total_kmers = 0, ambiguous = 0, lookups = 0, reads = 0;
    double seconds = 0;
};

// worker bodies + reduction of the futures (src/fm_scanner.cpp:138-236 and the three sibling functions)
scan_out run_scan(cmd_arguments const & args, std::vector<device_lane> & lanes, uint32_t V, bool paired)
{
    auto t0 = std::chrono::steady_clock::now();
    const uint64_t per = paired ? 2 : 1;
    size_t dev = 0;
    auto flush = [&](device_lane & L)
    {
        host_batch & b = L.buf[L.cur];
        if (b.n_reads == 0) return;
        ck(speqg_scan_submit_raw(L.scan, b.seq, b.qual, b.offsets, b.n_reads, SPEQG_ENC_ASCII), "speqg_scan_submit_raw");
        ck(speqg_scan_last_ticket(L.scan, &L.ticket[L.cur]), "speqg_scan_last_ticket");
        L.cur ^= 1;
        // only the copy that last read the buffer about to be refilled has to be over; the one just queued keeps running
        // while the host parses the next batch
        ck(speqg_scan_wait_ticket(L.scan, L.ticket[L.cur]), "speqg_scan_wait_ticket");
        L.buf[L.cur].clear();
    };
    // ---- plain 4-line FASTQ: mapped file, indexed and copied into the pinned batches by the parser pool (-t); anything
    //      irregular hands over to the serial reader (full seqan3 grammar) at the offset of the first irregular record
    uint64_t off1 = 0, off2 = 0;
    bool serial = true;
    // ---- single-end plain FASTQ: the file's bytes go to the device region by region (cut at record starts) and the record
    //      grammar runs there (speqg_scan_submit_fastq); the host only copies page cache -> pinned staging.  Regions the device
    //      finds irregular (multi-line records, blanks, ...) are read again below with the full grammar; the rest of the file
    //      goes through the host paths when a single record exceeds the staging buffer.  SPEQ_HOST_FASTQ=1 keeps the host indexer.
    if (!paired && !getenv("SPEQ_HOST_FASTQ") && speq::io::fastq_map::eligible(args.in_file_reads_path_1))
    {
        const int T = std::max(1, static_cast<int>(args.threads) - 1);
        speq::io::fastq_map m1(args.in_file_reads_path_1);
        while (off1 < m1.size())
        {
            device_lane & L = lanes[dev];
            host_batch & b = L.buf[L.cur];
            const uint64_t end = m1.region_end(off1, b.cap_bytes);
            if (end == off1) break;                                   // a record larger than the staging buffer: host paths
            ck(speqg_scan_wait_ticket(L.scan, L.ticket[L.cur]), "speqg_scan_wait_ticket");   // this buffer's previous copy
            m1.read_into(b.seq, off1, end - off1, T);
            uint64_t region = 0;
            ck(speqg_scan_submit_fastq(L.scan, b.seq, end - off1, &region, &L.ticket[L.cur]), "speqg_scan_submit_fastq");
            if (region != L.regions.size()) throw std::runtime_error("internal error: FASTQ region numbering");
            L.regions.emplace_back(off1, end - off1);
            L.cur ^= 1;
            off1 = end;
            dev = (dev + 1) % lanes.size();
        }
        // regions the device handed back: the serial grammar over the mapped bytes, through the raw-read path
        for (auto & L : lanes)
        {
            uint64_t n_rej = 0;
            ck(speqg_scan_fastq_rejected(L.scan, nullptr, 0, &n_rej), "speqg_scan_fastq_rejected");
            if (n_rej == 0) continue;
            std::vector<uint64_t> rej(n_rej);
            ck(speqg_scan_fastq_rejected(L.scan, rej.data(), n_rej, &n_rej), "speqg_scan_fastq_rejected");
            ck(speqg_scan_wait_ticket(L.scan, L.ticket[0]), "speqg_scan_wait_ticket");
            ck(speqg_scan_wait_ticket(L.scan, L.ticket[1]), "speqg_scan_wait_ticket");
            L.buf[0].ensure_records(); L.buf[1].ensure_records();
            L.buf[0].clear(); L.buf[1].clear();
            for (uint64_t id : rej)
            {
                auto const & rg = L.regions.at(id);
                speq::io::sequence_reader rd(m1.data(), rg.first, rg.first + rg.second, args.in_file_reads_path_1.string());
                std::string sq, ql;
                while (rd.next(sq, ql))
                {
                    if (!L.buf[L.cur].fits(sq.size(), 1))
                    {
                        flush(L);
                        if (sq.size() > L.buf[L.cur].cap_bytes)
                        {
                            ck(speqg_scan_sync(L.scan), "speqg_scan_sync");
                            L.buf[L.cur].alloc(sq.size() * 2, L.buf[L.cur].cap_reads);
                        }
                    }
                    L.buf[L.cur].push(sq, ql);
                }
            }
            flush(L);
        }
        serial = off1 < m1.size();
    }
    // ---- paired plain FASTQ: a region of each file with the same number of records (cut by line count, verified to end at
    //      record starts) goes to the device, which parses and interleaves the mates (speqg_scan_submit_fastq_pair)
    bool paired_device = false;
    if (paired && !getenv("SPEQ_HOST_FASTQ") && speq::io::fastq_map::eligible(args.in_file_reads_path_1) &&
        speq::io::fastq_map::eligible(args.in_file_reads_path_2))
    {
        paired_device = true;
        const int T = std::max(1, static_cast<int>(args.threads) - 1);
        speq::io::fastq_map m1(args.in_file_reads_path_1), m2(args.in_file_reads_path_2);
        for (auto & L : lanes) { L.buf[0].ensure_records(); L.buf[1].ensure_records(); }   // the quality staging holds file 2's bytes
        while (off1 < m1.size() && off2 < m2.size())
        {
            device_lane & L = lanes[dev];
            host_batch & b = L.buf[L.cur];
            uint64_t n1 = m1.region_end(off1, b.cap_bytes) - off1, n2 = m2.region_end(off2, b.cap_bytes) - off2;
            if (n1 == 0 || n2 == 0) break;                            // a record larger than the staging buffer: host paths
            ck(speqg_scan_wait_ticket(L.scan, L.ticket[L.cur]), "speqg_scan_wait_ticket");
            uint64_t nl1 = 0, nl2 = 0;
            m1.read_into(b.seq, off1, n1, T, &nl1);
            m2.read_into(b.qual, off2, n2, T, &nl2);
            const char * p1 = reinterpret_cast<const char *>(b.seq), * p2 = reinterpret_cast<const char *>(b.qual);
            const uint64_t l1 = nl1 + (p1[n1 - 1] != '\n'), l2 = nl2 + (p2[n2 - 1] != '\n');
            const uint64_t recs = std::min(l1, l2) / 4;
            if (recs == 0) break;
            if (l1 != 4 * recs) n1 = speq::io::offset_without_last_lines(p1, n1, l1 - 4 * recs);
            if (l2 != 4 * recs) n2 = speq::io::offset_without_last_lines(p2, n2, l2 - 4 * recs);
            // both cuts must be record starts (4-line records all the way): else the files are not in the strict form here
            if ((off1 + n1 < m1.size() && m1.record_start_from(off1 + n1) != off1 + n1) ||
                (off2 + n2 < m2.size() && m2.record_start_from(off2 + n2) != off2 + n2))
                break;
            uint64_t region = 0;
            ck(speqg_scan_submit_fastq_pair(L.scan, b.seq, n1, b.qual, n2, &region, &L.ticket[L.cur]), "speqg_scan_submit_fastq_pair");
            if (region != L.regions.size()) throw std::runtime_error("internal error: FASTQ region numbering");
            L.regions.emplace_back(off1, n1);
            L.regions2.emplace_back(off2, n2);
            L.cur ^= 1;
            off1 += n1;
            off2 += n2;
            dev = (dev + 1) % lanes.size();
        }
        // region pairs the device handed back: both regions again with the serial grammar, zipped.  Regions were cut by LINE
        // count; where a region is not in the 4-line form the two files may hold different numbers of records in it, and then
        // every later region paired the wrong mates: the tallies are dropped and both files go through the serial reader from
        // the start (views::zip over two record streams, src/fm_scanner.cpp:651)
        bool lost_step = false;
        for (auto & L : lanes)
        {
            if (lost_step) break;
            uint64_t n_rej = 0;
            ck(speqg_scan_fastq_rejected(L.scan, nullptr, 0, &n_rej), "speqg_scan_fastq_rejected");
            if (n_rej == 0) continue;
            std::vector<uint64_t> rej(n_rej);
            ck(speqg_scan_fastq_rejected(L.scan, rej.data(), n_rej, &n_rej), "speqg_scan_fastq_rejected");
            ck(speqg_scan_wait_ticket(L.scan, L.ticket[0]), "speqg_scan_wait_ticket");
            ck(speqg_scan_wait_ticket(L.scan, L.ticket[1]), "speqg_scan_wait_ticket");
            L.buf[0].clear(); L.buf[1].clear();
            for (uint64_t id : rej)
            {
                if (lost_step) break;
                auto const & r1 = L.regions.at(id);
                auto const & r2 = L.regions2.at(id);
                speq::io::sequence_reader rd1(m1.data(), r1.first, r1.first + r1.second, args.in_file_reads_path_1.string());
                speq::io::sequence_reader rd2(m2.data(), r2.first, r2.first + r2.second, args.in_file_reads_path_2.string());
                std::string sa, qa, sb, qb;
                while (true)
                {
                    const bool g1 = rd1.next(sa, qa), g2 = rd2.next(sb, qb);
                    if (g1 != g2) { lost_step = true; break; }
                    if (!g1) break;
                    if (!L.buf[L.cur].fits(sa.size() + sb.size(), 2))
                    {
                        flush(L);
                        if (sa.size() + sb.size() > L.buf[L.cur].cap_bytes)
                        {
                            ck(speqg_scan_sync(L.scan), "speqg_scan_sync");
                            L.buf[L.cur].alloc((sa.size() + sb.size()) * 2, L.buf[L.cur].cap_reads);
                        }
                    }
                    L.buf[L.cur].push(sa, qa);
                    L.buf[L.cur].push(sb, qb);
                }
            }
            flush(L);
        }
        for (auto & L : lanes)
        {
            ck(speqg_scan_wait_ticket(L.scan, L.ticket[0]), "speqg_scan_wait_ticket");
            ck(speqg_scan_wait_ticket(L.scan, L.ticket[1]), "speqg_scan_wait_ticket");
            L.buf[0].clear(); L.buf[1].clear();
        }
        serial = off1 < m1.size() && off2 < m2.size();
        if (lost_step)
        {
            std::cerr << "speq: the paired FASTQ files lose step inside a region that is not in the 4-line form; reading both "
                         "files again with the serial grammar\n";
            for (auto & L : lanes)
            {
                ck(speqg_scan_sync(L.scan), "speqg_scan_sync");
                ck(speqg_scan_reset(L.scan), "speqg_scan_reset");
                L.regions.clear(); L.regions2.clear();
                L.cur = 0;
            }
            off1 = off2 = 0;
            dev = 0;
            serial = true;
            paired_device = false;
        }
    }
    if (serial || (paired && !paired_device) || getenv("SPEQ_HOST_FASTQ") || !speq::io::fastq_map::eligible(args.in_file_reads_path_1))
        for (auto & L : lanes) { L.buf[0].ensure_records(); L.buf[1].ensure_records(); }
    if (paired_device || (!paired && !getenv("SPEQ_HOST_FASTQ") && speq::io::fastq_map::eligible(args.in_file_reads_path_1))) {}
    else if (speq::io::fastq_map::eligible(args.in_file_reads_path_1) && (!paired || speq::io::fastq_map::eligible(args.in_file_reads_path_2)))
    {
        serial = false;
        const int T = std::max(1, static_cast<int>(args.threads) - 1);
        speq::io::fastq_map m1(args.in_file_reads_path_1);
        std::unique_ptr<speq::io::fastq_map> m2;
        if (paired) m2 = std::make_unique<speq::io::fastq_map>(args.in_file_reads_path_2);
        std::vector<speq::io::fq_rec> v1, v2;
        while (off1 < m1.size() && (!paired || off2 < m2->size()))
        {
            device_lane & L = lanes[dev];
            host_batch & b = L.buf[L.cur];
            // a region whose sequence bytes fit the batch: sequence <= half of the record bytes
            const uint64_t region = b.cap_bytes / per;
            bool ok = m1.index(off1, region, T, v1);
            if (paired) ok = m2->index(off2, region, T, v2) && ok;
            uint64_t n = paired ? std::min(v1.size(), v2.size()) : v1.size();
            n = std::min<uint64_t>(n, b.cap_reads / per);
            uint64_t bytes = 0, m = 0;
            for (; m < n; ++m)
            {
                uint64_t add = v1[m].len + (paired ? v2[m].len : 0u);
                if (bytes + add > b.cap_bytes) break;
                b.offsets[per * m] = bytes;
                if (paired) b.offsets[per * m + 1] = bytes + v1[m].len;
                bytes += add;
            }
            n = m;
            if (n == 0)
            {
                serial = true;            // an irregular (or oversized) record is next: the serial reader takes over
                break;
            }
            b.offsets[per * n] = bytes;
            speq::io::copy_records(m1.data(), v1, n, b.seq, b.qual, b.offsets, per, 0, T);
            if (paired) speq::io::copy_records(m2->data(), v2, n, b.seq, b.qual, b.offsets, per, 1, T);
            b.n_reads = per * n;
            b.n_bytes = bytes;
            off1 = v1[n - 1].next;
            if (paired) off2 = v2[n - 1].next;
            flush(L);
            dev = (dev + 1) % lanes.size();
            if (!ok && n == (paired ? std::min(v1.size(), v2.size()) : v1.size()))
            {
                serial = true;
                break;
            }
        }
        if (serial && (off1 >= m1.size() || (paired && off2 >= m2->size()))) serial = false;   // nothing left
    }
    if (serial)
    {
    speq::io::sequence_reader r1(args.in_file_reads_path_1, off1);
    std::unique_ptr<speq::io::sequence_reader> r2;
    if (paired) r2 = std::make_unique<speq::io::sequence_reader>(args.in_file_reads_path_2, off2);
    std::string s1, q1, s2, q2;
    while (r1.next(s1, q1))
    {
        if (paired && !r2->next(s2, q2)) break;   // views::zip stops at the shorter file (src/fm_scanner.cpp:651)
        uint64_t need = s1.size() + (paired ? s2.size() : 0);
        device_lane * L = &lanes[dev];
        if (!L->buf[L->cur].fits(need, per))
        {
            flush(*L);
            dev = (dev + 1) % lanes.size();
            L = &lanes[dev];
            if (!L->buf[L->cur].fits(need, per))
            {
                if (L->buf[L->cur].n_reads) flush(*L);
                if (need > L->buf[L->cur].cap_bytes)
                {
                    ck(speqg_scan_sync(L->scan), "speqg_scan_sync");
                    L->buf[L->cur].alloc(need * 2, L->buf[L->cur].cap_reads);
                }
            }
        }
        L->buf[L->cur].push(s1, q1);
        if (paired) L->buf[L->cur].push(s2, q2);
    }
    }
    for (auto & L : lanes) flush(L);
    scan_out out;
    out.unique_int.assign(V, 0);
    out.unique_w.assign(V, 0.0);
    for (auto & L : lanes)   // the per-device tallies are summed like the reference sums its futures (src/fm_scanner.cpp:224-233)
    {
        std::vector<uint64_t> u(V);
        std::vector<double> w(V);
        speqg_counts c{};
        ck(speqg_scan_finish(L.scan, u.data(), w.data(), &c), "speqg_scan_finish");
        for (uint32_t g = 0; g < V; ++g) { out.unique_int[g] += u[g]; out.unique_w[g] += w[g]; }
        out.total_kmers += c.total_kmers; out.ambiguous += c.ambiguous_reads; out.lookups += c.lookups; out.reads += c.reads;
    }
    out.seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    return out;
}

std::vector<double> unique_to_percent(std::vector<double> const & unique_in_reads, uint64_t total_in_reads,
                                      std::vector<double> const & unique_in_refs, std::vector<double> const & total_in_refs)
{
    std::vector<double> out(unique_in_refs.size(), 0.0);
    ck(speqg_unique_to_percent(unique_in_reads.data(), total_in_reads, unique_in_refs.data(), total_in_refs.data(),
                               static_cast<uint32_t>(out.size()), out.data()), "speqg_unique_to_percent");
    return out;
}

// --write-output: the per-variant result table as tab-separated text in -o/--output.  The reference declares the option
// (src/arg_parse.cpp:20) and reports on stderr only; this is the fix-forward of SURVEY.md 8f-4, off by default.
void write_table(fs::path const & p, idx_meta const & meta, scan_out const & so, std::vector<double> const & unique_totals,
                 std::vector<double> const & initial, std::vector<double> const & fin, std::map<uint64_t, uint64_t> const & fusion,
                 uint64_t fusion_unkeyed)
{
    FILE * f = fopen(p.c_str(), "w");
    if (!f) throw std::runtime_error("cannot write " + p.string());
    fprintf(f, "# total_kmers\t%llu\n# ambiguous_reads\t%llu\n", static_cast<unsigned long long>(so.total_kmers),
            static_cast<unsigned long long>(so.ambiguous));
    fprintf(f, "group\tisolates\tunique_kmers\tpercent_initial\tpercent_final\n");
    for (size_t g = 0; g < fin.size(); ++g)
        fprintf(f, "%s\t%d\t%.17g\t%.17g\t%.17g\n", meta.group_names[g].c_str(), meta.group_counts[g], unique_totals[g], initial[g], fin[g]);
    // the fix-forward of the reference's fusion_totals report (std::map<std::set<int>,int>, src/fm_scanner.cpp:915,947-952,
    // 983-995,1033): ambiguous reads / pairs by the SET of variants whose unique k-mers they hold.  Up to 64 variants the key is
    // the full set; above, the smallest and largest variant, with "..." between them when further variants met in the unit.
    const size_t V = fin.size();
    fprintf(f, "# ambiguous reads / pairs by the set of variants that met in them\nfusion\tgroups\tcount\n");
    for (auto const & kv : fusion)
    {
        std::string names;
        if (V <= 64)
        {
            for (size_t g = 0; g < V; ++g)
                if ((kv.first >> g) & 1ull) names += (names.empty() ? "" : ",") + meta.group_names[g];
        }
        else
        {
            names = meta.group_names[kv.first & 0xFFFFull];
            names += ((kv.first >> 32) & 1ull) ? ",...," : ",";
            names += meta.group_names[(kv.first >> 16) & 0xFFFFull];
        }
        fprintf(f, "fusion\t%s\t%llu\n", names.c_str(), static_cast<unsigned long long>(kv.second));
    }
    if (fusion_unkeyed) fprintf(f, "fusion\t(not keyed: device map full)\t%llu\n", static_cast<unsigned long long>(fusion_unkeyed));
    fclose(f);
}

void write_json(fs::path const & p, cmd_arguments const & args, bool local, bool paired, std::vector<uint64_t> const & uref,
                std::vector<uint64_t> const & tref, scan_out const & so, std::vector<double> const & unique_totals,
                std::vector<double> const & initial, std::vector<double> const & fin, size_t iters, double t_backcheck, double t_refine)
{
    FILE * f = fopen(p.c_str(), "w");
    if (!f) throw std::runtime_error("cannot write " + p.string());
    auto vu = [&](const char * k, std::vector<uint64_t> const & v)
    {
        fprintf(f, "\"%s\":[", k);
        for (size_t i = 0; i < v.size(); ++i) fprintf(f, "%s%llu", i ? "," : "", static_cast<unsigned long long>(v[i]));
        fprintf(f, "],");
    };
    auto vd = [&](const char * k, std::vector<double> const & v)
    {
        fprintf(f, "\"%s\":[", k);
        for (size_t i = 0; i < v.size(); ++i) fprintf(f, "%s%.17g", i ? "," : "", v[i]);
        fprintf(f, "],");
    };
    fprintf(f, "{\"k\":%u,\"mode\":\"%s\",\"paired\":%s,\"gpus\":%u,", args.kmer, local ? "local" : "global", paired ? "true" : "false", args.gpus);
    vu("unique_ref", uref); vu("total_ref", tref); vu("unique", so.unique_int); vd("unique_w", so.unique_w); vd("unique_totals", unique_totals);
    vd("percent_initial", initial); vd("percent_final", fin);
    fprintf(f, "\"total_kmers\":%llu,\"ambiguous\":%llu,\"lookups\":%llu,\"reads\":%llu,\"iterations\":%zu,",
            (unsigned long long)so.total_kmers, (unsigned long long)so.ambiguous, (unsigned long long)so.lookups, (unsigned long long)so.reads, iters);
    fprintf(f, "\"t_backcheck_s\":%.6f,\"t_scan_s\":%.6f,\"t_refine_s\":%.6f}\n", t_backcheck, so.seconds, t_refine);
    fclose(f);
}

// ---- speq::scan::async_one / async_two
int do_scan(cmd_arguments & args)
{
    const bool paired = !args.in_file_reads_path_2.empty();
    const bool local = (args.fixed_accuracy == 0.0);                                 // src/fm_scanner.cpp:8-17
    const double percent_perfect = local ? 0.0 : pow(args.fixed_accuracy, args.kmer);
    // SPEQ_TIMING=1: wall-clock stamps of the phases on stderr (profiles/cli_e2e.py)
    const auto T0 = std::chrono::steady_clock::now();
    const bool timing = getenv("SPEQ_TIMING") != nullptr;
    auto stamp = [&](const char * what)
    {
        if (timing) fprintf(stderr, "[timing] %-34s %.3f s\n", what, std::chrono::duration<double>(std::chrono::steady_clock::now() - T0).count());
    };
    // Start-up overlap: the CUDA context and the pinned staging buffers do not depend on the index, so a helper thread creates
    // them while this thread reads the .idx from disk.
    int ndev = 0;
    ck(speqg_device_count(&ndev), "speqg_device_count");
    if (ndev <= 0) throw std::runtime_error("no CUDA device: speq scan has no CPU path");
    const unsigned gpus = std::min<unsigned>(args.gpus, static_cast<unsigned>(ndev));
    std::vector<device_lane> lanes(gpus);
    const uint64_t cap_reads = std::max<uint64_t>(2, args.batch_reads & ~1u);
    const uint64_t cap_bytes = std::min<uint64_t>(cap_reads * 320, uint64_t(512) << 20);
    std::string prep_error;
    // (paired input always needs the second staging array -- the device parser takes the second file's bytes there, the host
    //  paths the qualities -- so it is pinned here as well instead of inside the scan phase)
    const bool records_now = !args.in_file_reads_path_2.empty();
    std::thread prep([&]()
    {
        try
        {
            for (unsigned d = 0; d < gpus; ++d)
            {
                lanes[d].buf[0].alloc(cap_bytes, cap_reads, records_now);
                lanes[d].buf[1].alloc(cap_bytes, cap_reads, records_now);
            }
        }
        catch (std::exception const & e) { prep_error = e.what(); }
    });
    struct joiner { std::thread & t; ~joiner() { if (t.joinable()) t.join(); } } prep_join{prep};
    // import the index (src/fm_scanner.cpp:37-58)
    args.io_file_index.replace_extension(".idx");
    idx_meta meta = read_meta(args.io_file_index);
    index_handle idx;
    if (meta.reference_format)
    {
        // an index the reference built: its texts are decoded from the archived wavelet tree and indexed again here
        int nd = 0;
        if (speqg_device_count(&nd) != SPEQG_OK) nd = 0;
        ck(speqg_refidx_build_index(meta.ref.get(), nd > 0 ? SPEQG_BUILD_DEVICE_SA : 0u, &idx.h), "speqg_refidx_build_index");
        meta.ref.reset();
    }
    else
        ck(speqg_index_load_at(args.io_file_index.c_str(), meta.payload_offset, &idx.h), "speqg_index_load_at");
    if (args.in_file_references != fs::path(meta.scaffold_file))                      // src/fm_scanner.cpp:60-72
    {
        if (fs::exists(meta.scaffold_file)) args.in_file_references = meta.scaffold_file;
        else throw std::logic_error{"SPeQ could not find the sequence file used to make the specified index file."};
    }
    const uint32_t V = static_cast<uint32_t>(meta.group_names.size());
    stamp("index file read");

    // <stem>_<k>mer.dat: back-check counts, cached per k, keyed by the .idx mtime (src/fm_scanner.cpp:79-135)
    int64_t idx_time = mtime_of(args.io_file_index);
    fs::path dat = args.io_file_index;
    dat.replace_filename(args.io_file_index.stem().string() + "_" + std::to_string(args.kmer) + "mer.dat");
    std::vector<uint64_t> unique_ref(V, 0), total_ref(V, 0);
    bool have_dat = false;
    double t_backcheck = 0;
    if (fs::exists(dat))
    {
        // the reference's archive, byte for byte (src/fm_scanner.cpp:1561-1571: cereal binary of the .idx file time and two
        // std::vector<std::size_t>): int64 tick count, then twice (u64 size, size x u64) -- files are interchangeable
        std::ifstream is(dat, std::ios::binary);
        try
        {
            if (static_cast<int64_t>(get_u64(is)) == idx_time && get_u64(is) == V)
            {
                is.read(reinterpret_cast<char *>(unique_ref.data()), V * 8);
                if (is && get_u64(is) == V)
                {
                    is.read(reinterpret_cast<char *>(total_ref.data()), V * 8);
                    have_dat = static_cast<bool>(is);
                }
            }
        }
        catch (std::exception const &) { have_dat = false; }      // truncated / foreign file: rebuilt below
    }
    if (!have_dat)
    {
        auto t0 = std::chrono::steady_clock::now();
        ck(speqg_backcheck(idx.h, 0, args.kmer, unique_ref.data(), total_ref.data()), "speqg_backcheck");
        t_backcheck = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        std::ofstream os(dat, std::ios::binary | std::ios::trunc);
        put_u64(os, static_cast<uint64_t>(idx_time));
        put_u64(os, V);
        os.write(reinterpret_cast<const char *>(unique_ref.data()), V * 8);
        put_u64(os, V);
        os.write(reinterpret_cast<const char *>(total_ref.data()), V * 8);
        std::cerr << unique_ref << "\n";                                              // src/fm_scanner.cpp:1572-1573
        std::cerr << total_ref << "\n";
    }

    stamp("back-check / .dat");
    // one scan handle per device, index replicated
    speqg_params prm{};
    prm.k = args.kmer;
    prm.phred_cutoff = args.phred_cutoff;
    prm.mode = local ? SPEQG_MODE_LOCAL : SPEQG_MODE_GLOBAL;
    prm.paired = paired ? 1 : 0;
    prm.record_multiplicity = 1;
    prm.algorithm = SPEQG_ALGO_AUTO;   // k <= 31: k-mer table scan; else the rolling FM search
    for (unsigned d = 0; d < gpus; ++d) ck(speqg_scan_open(idx.h, static_cast<int>(d), &prm, &lanes[d].scan), "speqg_scan_open");
    stamp("scan handles open");
    prep.join();
    if (!prep_error.empty()) throw std::runtime_error(prep_error);
    stamp("pinned staging ready");
    int rc = 1;
    try
    {
        scan_out so = run_scan(args, lanes, V, paired);
        stamp("scan done");
        std::vector<double> unique_totals(V, 0.0);
        if (local) unique_totals = so.unique_w;
        else for (uint32_t i = 0; i < V; ++i) unique_totals[i] += static_cast<double>(so.unique_int[i]) / percent_perfect;   // :234-236
        std::vector<double> d_uref(unique_ref.begin(), unique_ref.end()), d_tref(total_ref.begin(), total_ref.end());
        auto percent = unique_to_percent(unique_totals, so.total_kmers, d_uref, d_tref);                                    // :238-243
        auto initial = percent;
        // report (global single :245-247, local single :513-514, global paired :757-759, local paired :1030-1033)
        std::cerr << percent << "\n";
        if (!(local && !paired)) std::cerr << unique_totals << "\n";
        std::cerr << so.total_kmers << "\t" << so.ambiguous << "\n";
        if (local && paired)
        {
            // the reference's fusion map only ever holds the empty-set key (set_groups is passed by value, :916)
            if (so.ambiguous) std::cerr << "[([]," << so.ambiguous << ")]\n";
            else std::cerr << "[]\n";
        }
        // refinement loop (src/fm_scanner.cpp:248-279): the estimate runs on the recorded k-mer multiplicities
        auto t0 = std::chrono::steady_clock::now();
        std::vector<double> diff(V, 1.0);
        size_t iters = 0;
        while (V > 0 && *std::max_element(diff.begin(), diff.end()) > args.precision)
        {
            std::vector<double> next_tkpg(V, 0.0), part(V);
            for (auto & L : lanes)
            {
                ck(speqg_refine(L.scan, percent.data(), meta.group_counts.data(), part.data()), "speqg_refine");
                for (uint32_t g = 0; g < V; ++g) next_tkpg[g] += part[g];
            }
            if (local && !paired) std::cerr << "1: " << unique_totals << "\n2: " << so.total_kmers << "\n3: " << next_tkpg << "\n";   // :530
            auto next_percent = unique_to_percent(unique_totals, so.total_kmers, unique_totals, next_tkpg);
            for (uint32_t i = 0; i < V; ++i) diff[i] = std::abs(next_percent[i] - percent[i]);
            percent = next_percent;
            if (local && !paired)
            {
                std::cerr << "Percent of each group: " << percent << "\n";                                                  // :543-544
                std::cerr << "Total Kmers per group: " << next_tkpg << "\n\n";
            }
            else
            {
                std::cerr << percent << "\n";                                                                               // :277-278
                std::cerr << next_tkpg << "\n\n";
            }
            ++iters;
        }
        double t_refine = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        stamp("refinement done");
        if (args.write_output)
        {
            // ambiguous reads / pairs by the variants that met in them, summed over the devices
            std::map<uint64_t, uint64_t> fusion;
            uint64_t unkeyed = 0;
            for (auto & L : lanes)
            {
                uint64_t n = 0, lost = 0;
                ck(speqg_scan_fusion_sets(L.scan, nullptr, nullptr, 0, &n, &lost), "speqg_scan_fusion_sets");
                std::vector<uint64_t> keys(n), cnts(n);
                ck(speqg_scan_fusion_sets(L.scan, keys.data(), cnts.data(), n, &n, &lost), "speqg_scan_fusion_sets");
                for (uint64_t i = 0; i < n; ++i) fusion[keys[i]] += cnts[i];
                unkeyed += lost;
            }
            write_table(args.out_file_path, meta, so, unique_totals, initial, percent, fusion, unkeyed);
        }
        if (!args.json_path.empty())
            write_json(args.json_path, args, local, paired, unique_ref, total_ref, so, unique_totals, initial, percent, iters, t_backcheck, t_refine);
    }
    catch (...)
    {
        for (auto & L : lanes) { speqg_scan_close(L.scan); L.buf[0].release(); L.buf[1].release(); }
        throw;
    }
    for (auto & L : lanes) { speqg_scan_close(L.scan); L.buf[0].release(); L.buf[1].release(); }
    return rc;
}
}   // namespace

// `speq ingest-check <reads.fq> [threads]`: host-only self-check of the parallel FASTQ indexer against the serial reader
// (reads, bases, FNV-1a of all sequence and quality bytes in record order).  Not a reference sub-command.
int ingest_check(int argc, char ** argv)
{
    if (argc < 3) { std::cerr << "usage: speq ingest-check <reads.fq> [threads] [region_bytes]\n"; return 2; }
    fs::path p = argv[2];
    int T = argc > 3 ? std::max(1, atoi(argv[3])) : 4;
    uint64_t region = argc > 4 ? strtoull(argv[4], nullptr, 10) : (uint64_t(8) << 20);
    auto fnv = [](uint64_t h, const char * d, size_t n) { for (size_t i = 0; i < n; ++i) { h ^= static_cast<unsigned char>(d[i]); h *= 1099511628211ull; } return h; };
    uint64_t n_ser = 0, b_ser = 0, h_ser = 1469598103934665603ull;
    auto c0 = std::chrono::steady_clock::now();
    {
        speq::io::sequence_reader rd(p);
        std::string s, q;
        while (rd.next(s, q)) { ++n_ser; b_ser += s.size(); h_ser = fnv(fnv(h_ser, s.data(), s.size()), q.data(), q.size()); }
    }
    auto c1 = std::chrono::steady_clock::now();
    double t_copy = 0;
    uint64_t n_par = 0, b_par = 0, h_par = 1469598103934665603ull, off = 0, handed = 0;
    if (speq::io::fastq_map::eligible(p))
    {
        speq::io::fastq_map m(p);
        std::vector<speq::io::fq_rec> v;
        bool ok = true;
        while (ok && off < m.size())
        {
            auto k0 = std::chrono::steady_clock::now();
            ok = m.index(off, region, T, v);
            if (v.empty()) { ok = false; break; }
            std::vector<uint64_t> offs(v.size() + 1, 0);
            for (size_t i = 0; i < v.size(); ++i) offs[i + 1] = offs[i] + v[i].len;
            std::vector<uint8_t> sq(offs.back() + 1), ql(offs.back() + 1);
            speq::io::copy_records(m.data(), v, v.size(), sq.data(), ql.data(), offs.data(), 1, 0, T);
            t_copy += std::chrono::duration<double>(std::chrono::steady_clock::now() - k0).count();
            for (size_t i = 0; i < v.size(); ++i)
                h_par = fnv(fnv(h_par, reinterpret_cast<const char *>(sq.data()) + offs[i], v[i].len), reinterpret_cast<const char *>(ql.data()) + offs[i], v[i].len);
            n_par += v.size(); b_par += offs.back();
            off = v.back().next;
        }
        if (!ok && off < m.size())
        {
            handed = off;
            speq::io::sequence_reader rd(p, off);
            std::string s, q;
            while (rd.next(s, q)) { ++n_par; b_par += s.size(); h_par = fnv(fnv(h_par, s.data(), s.size()), q.data(), q.size()); }
        }
    }
    std::cerr << "serial reader " << std::chrono::duration<double>(c1 - c0).count() << " s (incl. hashing); parallel index + copy "
              << t_copy << " s with " << T << " threads\n";
    std::cout << "serial " << n_ser << " " << b_ser << " " << h_ser << "\nparallel " << n_par << " " << b_par << " " << h_par
              << " handover " << handed << "\n";
    return (n_ser == n_par && b_ser == b_par && h_ser == h_par) ? 0 : 1;
}

int main(int argc, char ** argv)
{
    if (argc > 1 && std::string(argv[1]) == "ingest-check")
    {
        try { return ingest_check(argc, argv); }
        catch (std::exception const & e) { std::cerr << e.what() << "\n"; return 3; }
    }
    auto args = speq::args::initialize_argument_parser(argc, argv);
    if (!args.is_parsed) return 0;   // the reference returns 0 after printing the parse error (src/main.cpp:17-40)
    try
    {
        if (args.is_indexer) do_index(args);
        if (args.is_scanner) do_scan(args);
    }
    catch (std::exception const & e)
    {
        // the reference lets exceptions escape (terminate); we print the same message and exit non-zero
        std::cerr << "terminate called after throwing an instance of an exception\n  what():  " << e.what() << "\n";
        return 134;
    }
    // every result is written: leave without tearing the CUDA context down handle by handle (about 0.3 s of a 1.5 s run)
    std::cout.flush();
    std::cerr.flush();
    fflush(nullptr);
    _exit(0);
}
