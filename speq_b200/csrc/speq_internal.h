// speq_internal.h -- host-side structures behind the opaque handles of include/speq_cuda.h.
#pragma once
#include <cstdint>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/speq_cuda.h"
#include "fm_layout.h"

namespace speqb
{
struct api_error : std::runtime_error
{
    int code;
    api_error(int c, std::string const & m) : std::runtime_error(m), code(c) {}
};

// Host copy of the index (what .idx stores).
struct HostIndex
{
    uint64_t n = 0;
    uint32_t S = 0, V = 0;
    uint64_t C[8] = {0};
    std::vector<KBlock> blocks;            // n/64 + 1 blocks; e bits zero (they are filled per scan, for the scan's k)
    std::vector<uint16_t> grp;             // n: variant of the text owning suffix SA[i] (kNoGroup16: delimiter / sentinel)
    std::vector<uint8_t> lcp;              // lcp_padded_size(n) entries (zero padded), capped at kLcpCap
    std::vector<uint2x> pn;                // n + 1: previous / next smaller lcp value positions (fm_layout.h)
    std::vector<uint32_t> n_pos;           // sorted BWT positions holding N
    uint64_t n_exc = 0;                    // BWT positions holding N / delimiter / sentinel
    std::vector<int32_t> group_of_text;    // S
    // the S texts packed as a read batch (inv = N) for the back-check
    std::vector<uint64_t> t_bases;
    std::vector<uint32_t> t_inv;
    std::vector<uint64_t> t_grp_off;
    std::vector<uint32_t> t_len;
    // optional (SPEQG_BUILD_KEEP_SA)
    std::vector<uint32_t> sa;
};

struct DeviceIndex
{
    int device = -1;
    KBlock * blocks = nullptr;
    uint16_t * grp = nullptr;
    uint8_t * lcp = nullptr;
    uint2x * pn = nullptr;
    UpRec * up = nullptr;
    uint32_t * n_pos = nullptr;
    uint32_t * ctab = nullptr;
    // texts batch
    uint64_t * t_bases = nullptr;
    uint32_t * t_inv = nullptr;
    uint64_t * t_grp_off = nullptr;
    uint32_t * t_len = nullptr;
    int32_t * t_group = nullptr;
    uint64_t bytes = 0;
    DevIndex view{};
};

inline uint64_t lcp_padded_size(uint64_t n) { return ((n + kLcpChunk) / kLcpChunk + 1) * kLcpChunk; }

void build_host_index(uint8_t const * const * texts, uint64_t const * lens, uint32_t S, int32_t const * group_of_text,
                      uint32_t V, uint32_t flags, HostIndex & out);
void save_host_index(HostIndex const & h, std::string const & path, uint64_t offset = 0);
void load_host_index(std::string const & path, HostIndex & h, uint64_t offset = 0);

// packer (pack.cpp)
uint64_t pack_groups(uint64_t const * offsets, uint64_t n_reads);
void pack_reads(uint8_t const * seq, uint8_t const * qual, uint64_t const * offsets, uint64_t n_reads, int encoding,
                uint32_t phred_cutoff, bool n_only, int threads, speqg_batch & out);
}   // namespace speqb

struct speqg_index
{
    speqb::HostIndex host;
    std::map<int, speqb::DeviceIndex> dev;
};
