// capi_scan.cu -- scan / back-check / refinement entry points of the C ABI (include/speq_cuda.h).
//
// One speqg_scan owns: the tallies in device memory, a compute stream (the caller's or its own), a copy stream, and a
// ring of device staging slots so the H2D copy of batch i+1 overlaps the kernel of batch i.
#include <cuda_runtime.h>

#include <cmath>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "capi_util.h"
#include "kernels.h"

using namespace speqb;

namespace
{
constexpr int kSlots = 3;
// slots of the fusion map (variant set of an ambiguous unit -> units): 16 MB of device memory per scan handle
constexpr uint32_t kFusionCap = 1u << 20;
// a k-mer table slot keeps a variant below 2^15 next to the "several variants" bit
constexpr uint32_t kMaxGroupsTable = 32767;
// per-CTA shared-memory tallies: 32 + 12 V bytes in the scan kernels, 24 V bytes in refine_kernel (227 KB per CTA)
constexpr uint32_t kMaxGroupsScan = 8192;

template <class T> struct dev_buf
{
    T * p = nullptr;
    size_t cap = 0;
    void reserve(size_t count)
    {
        if (count <= cap) return;
        if (p) CK(cudaFree(p));
        p = nullptr;
        cap = 0;
        size_t want = count + count / 8 + 64;
        CK(cudaMalloc(reinterpret_cast<void **>(&p), want * sizeof(T)));
        cap = want;
    }
    void release()
    {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
};

struct Slot
{
    dev_buf<uint64_t> bases, grp_off, raw_off;
    dev_buf<uint32_t> inv, len;
    dev_buf<uint8_t> qual, raw_seq, raw_qual;
    // FASTQ regions parsed on the device (speqg_scan_submit_fastq)
    dev_buf<uint32_t> fq_chunk, fq_chunk2;        // newlines per 4 KB chunk -> exclusive prefix sum (2: the second mate's file)
    dev_buf<uint64_t> fq_lines, fq_lines2, fq_soff, fq_qoff;
    dev_buf<unsigned long long> fq_small;         // [0] newline count, [1] reads published for the scan kernel, [2] reject flag,
                                                  // [3] newline count of the second file
    unsigned long long * h_total = nullptr;       // pinned [2]: the newline counts, read by the host between the two phases
    cudaEvent_t counted = nullptr;
    uint64_t fq_bytes = 0, fq_bytes2 = 0, fq_region = 0;
    bool fq_terminated = true, fq_terminated2 = true, fq_pair = false;
    cudaEvent_t copied = nullptr, done = nullptr;
    uint64_t ticket = 0;
    bool busy = false;
    void release()
    {
        bases.release(); grp_off.release(); raw_off.release(); inv.release(); len.release();
        qual.release(); raw_seq.release(); raw_qual.release();
        fq_chunk.release(); fq_chunk2.release(); fq_lines.release(); fq_lines2.release(); fq_soff.release(); fq_qoff.release(); fq_small.release();
        if (h_total) cudaFreeHost(h_total);
        h_total = nullptr;
        if (counted) cudaEventDestroy(counted);
        if (copied) cudaEventDestroy(copied);
        if (done) cudaEventDestroy(done);
        copied = done = counted = nullptr;
    }
};
// regions of one scan whose irregularity can be reported (speqg_scan_fastq_rejected); further regions are refused
constexpr uint32_t kMaxFastqRegions = 1u << 20;
}   // namespace

struct speqg_scan
{
    speqg_index * idx = nullptr;
    DeviceIndex * di = nullptr;
    DevIndex view{};               // di->view with `blocks` replaced by this scan's k-specialised copy
    KBlock * d_blocks_k = nullptr;   // SPEQG_ALGO_ROLLING64 only
    FBlock * d_sblk = nullptr;       // per-k blocks of the product kernel (SBlock + variant bytes)
    // SPEQG_ALGO_KTABLE: the per-(index, k) k-mer table and, with record_multiplicity, per-slot counts
    KTabView kt{};
    unsigned long long * d_kstat = nullptr;   // [2]: keys stored, placement failure flag
    unsigned long long * d_slot_mult = nullptr;
    bool slot_mult_dirty = false;
    uint64_t ktab_keys = 0;
    uint32_t filt_bits = 0;       // presence filter of the k-mer table (pair-table scan): 2^filt_bits bits
    // SPEQG_ALGO_KPAIR: the table over the (k+1)-mers (two windows per probe) next to the k-mer table
    KTabView kp{};
    unsigned long long * d_pair_mult = nullptr;
    uint64_t kpair_keys = 0;
    LaunchCfg cfg_p{};
    uint32_t algo = SPEQG_ALGO_ROLLING;   // resolved algorithm (AUTO -> KTABLE / ROLLING)
    int device = 0;
    speqg_params p{};
    int mode = 0;
    cudaStream_t compute = nullptr, copy = nullptr;
    bool own_compute = false;
    // the scan's counters live in one device block (one memset per reset, one copy per finish):
    // unique[V] total_g[V] scalars[SC_COUNT] unique_w[V] bad[1] work[1]; h_state is its pinned host mirror.  The integer
    // counters come first so that a multi-GPU driver sums them across ranks in place with one all-reduce
    // (speqg_scan_device_counters)
    // ambiguous units by the set of variants that met in them: open-addressing map in device memory (ScanArgs::fkeys)
    unsigned long long * d_fkeys = nullptr, * d_fvals = nullptr;
    // windows whose k-mer lies in one variant are not recorded for the refinement: their count per variant (closed form)
    unsigned long long * d_unique_n = nullptr;
    // k-mer table with positions: recorded runs as a difference array over the packed-text positions
    unsigned long long * d_diff = nullptr;
    uint64_t * d_tflags = nullptr;           // walk form of the k-mer table: per-group window flags and variants (KTabView)
    uint16_t * d_gvar = nullptr;
    uint32_t * d_pos_side = nullptr;
    void * d_scan_tmp = nullptr;
    size_t scan_tmp_bytes = 0;
    uint64_t npos = 0;
    bool diff_dirty = false;
    bool mult_dirty = true;         // mult[] holds counts (or was never cleared): reset has to zero it
    unsigned long long * d_state = nullptr, * h_state = nullptr;
    size_t state_words = 0;
    unsigned long long * d_unique = nullptr, * d_total_g = nullptr, * d_scalars = nullptr, * d_mult = nullptr, * d_bad = nullptr, * d_work = nullptr;
    double * d_unique_w = nullptr, * d_qden = nullptr, * d_percent = nullptr, * d_tkpg = nullptr, * d_partial = nullptr;
    int32_t * d_gc = nullptr;
    LaunchCfg cfg{}, cfg64{}, cfg_direct{}, cfg_stats{}, cfg_h{}, cfg_w{};
    int refine_grid = 0;
    Slot slots[kSlots];
    int next_slot = 0;
    uint64_t launches = 0;
    uint64_t tickets = 0;                  // submits that copied host memory, numbered from 1
    // FASTQ regions: the second phase of the last submitted region (lines, records, pack, scan) is issued by the next call
    int fq_pending = -1;
    uint64_t fq_regions = 0;
    unsigned int * d_fq_flags = nullptr;   // [kMaxFastqRegions] 1 = the region was not in the strict form and was not scanned
};

namespace speqb
{
void fastq_phase2(speqg_scan * s);   // the deferred second phase of the last FASTQ region (below)
struct scan_refine_view
{
    int device;
    uint32_t k, V;
    uint64_t n;
    DevIndex view;
    cudaStream_t stream;
    unsigned long long * mult;
    const unsigned long long * unique_n;
    double * d_percent, * d_tkpg, * d_partial;
    int32_t * d_gc;
    int refine_grid;
};
scan_refine_view refine_view(speqg_scan * s)
{
    {
        device_guard guard(s->device);
        fastq_phase2(s);
    }
    // k-mer table scan: multiplicities were recorded per table slot; move them to mult[lb] (what the rolling scan records)
    if ((s->slot_mult_dirty || s->diff_dirty) && s->d_mult) s->mult_dirty = true;
    if (s->slot_mult_dirty && s->d_mult)
    {
        device_guard guard(s->device);
        if (s->d_slot_mult || s->kt.wide) CK(launch_ktable_fold_mult(s->kt, s->view, s->p.k, s->d_slot_mult, s->d_mult, s->compute));
        if (s->d_pair_mult || s->kp.wide) CK(launch_kpair_fold_mult(s->kp, s->view, s->p.k, s->d_pair_mult, s->d_mult, s->compute));
        s->slot_mult_dirty = false;
    }
    // k-mer table with positions: recorded runs -> counts per text position -> mult[lb]
    if (s->diff_dirty && s->d_mult)
    {
        device_guard guard(s->device);
        CK(launch_kpos_fold(s->di->t_bases, s->view, s->p.k, s->d_diff, s->npos, s->d_scan_tmp, s->scan_tmp_bytes, s->d_mult, s->kt, s->compute));
        s->diff_dirty = false;
    }
    return scan_refine_view{s->device, s->p.k, s->idx->host.V, s->idx->host.n, s->view, s->compute, s->d_mult, s->d_unique_n,
                            s->d_percent, s->d_tkpg, s->d_partial, s->d_gc, s->refine_grid};
}
// The same view WITHOUT any work: false when work is still due before the refinement may read mult[] (a deferred FASTQ region,
// multiplicities not yet folded to the interval starts).  For speqg_refine_noalloc (noalloc.cu), which may only launch kernels.
bool refine_view_ready(speqg_scan * s, scan_refine_view * out)
{
    if (s->fq_pending >= 0 || ((s->slot_mult_dirty || s->diff_dirty) && s->d_mult)) return false;
    *out = scan_refine_view{s->device, s->p.k, s->idx->host.V, s->idx->host.n, s->view, s->compute, s->d_mult, s->d_unique_n,
                            s->d_percent, s->d_tkpg, s->d_partial, s->d_gc, s->refine_grid};
    return true;
}
}   // namespace speqb

namespace
{
void scan_release(speqg_scan * s)
{
    cudaFree(s->d_state); cudaFreeHost(s->h_state); cudaFree(s->d_fkeys); cudaFree(s->d_fvals); cudaFree(s->d_mult);
    cudaFree(s->d_unique_n); cudaFree(s->d_diff); cudaFree(s->d_scan_tmp); cudaFree(s->d_tflags); cudaFree(s->d_gvar); cudaFree(s->d_pos_side);
    cudaFree(s->d_qden); cudaFree(s->d_percent); cudaFree(s->d_tkpg); cudaFree(s->d_partial);
    cudaFree(s->d_gc); cudaFree(s->d_blocks_k); cudaFree(s->d_sblk);
    cudaFree(s->kt.slots); cudaFree(s->kt.filt); cudaFree(s->d_slot_mult); cudaFree(s->d_kstat);
    cudaFree(s->kp.slots); cudaFree(s->d_pair_mult); cudaFree(s->d_fq_flags);
    for (auto & sl : s->slots) sl.release();
    if (s->copy) cudaStreamDestroy(s->copy);
    if (s->own_compute && s->compute) cudaStreamDestroy(s->compute);
}

ScanArgs scan_args(speqg_scan * s)
{
    ScanArgs a{};
    a.k = s->p.k;
    a.V = s->idx->host.V;
    a.unique = s->d_unique;
    a.total_g = s->d_total_g;
    a.unique_w = s->d_unique_w;
    a.scalars = s->d_scalars;
    a.mult = s->d_mult;
    a.qden = s->d_qden;
    a.qrden = s->d_qden + 42;
    a.segs = nullptr;
    a.n_segs = 0;
    a.work = s->d_work;
    a.fkeys = s->d_fkeys;
    a.fvals = s->d_fvals;
    a.fcap_mask = kFusionCap - 1u;
    a.unique_n = s->d_unique_n;
    a.diff = s->d_diff;

    return a;
}

void check_batch(speqg_scan * s, speqg_batch const * b)
{
    need(b != nullptr, "batch is NULL");
    if (b->n_reads == 0) return;
    need(b->bases && b->inv && b->grp_off && b->len, "batch arrays must not be NULL");
    need(!(s->mode & M_LOCAL) || b->qual != nullptr, "local-error mode needs per-base qualities (batch.qual)");
    need(!(s->mode & M_PAIRED) || b->n_reads % 2 == 0, "paired mode needs an even number of reads per batch (mates interleaved)");
    need(b->n_reads < (uint64_t(1) << 32), "more than 2^32 reads in one batch");
}


// reads at least this long on average are scanned by a warp each (scanh.cu, WPR) instead of by a lane each
constexpr uint64_t kWarpPerReadBases = 1024;

// groups: 32-base groups of the batch (an upper bound is fine: only the average read length is taken from it)
void launch_on_view(speqg_scan * s, BatchView const & v, uint64_t groups)
{
    if (v.n_reads == 0) return;
    ScanArgs a = scan_args(s);
    static const uint64_t wpr_bases = getenv("SPEQB_WPR_BASES") ? strtoull(getenv("SPEQB_WPR_BASES"), nullptr, 10) : kWarpPerReadBases;
    const int wpr = groups * 32u >= v.n_reads * wpr_bases;   // table scans: a warp per read for batches of long reads
    if (s->algo == SPEQG_ALGO_DIRECT) CK(launch_scan_direct(s->view, v, a, s->mode, 0, s->cfg_direct, s->compute));
    else if (s->algo == SPEQG_ALGO_ROLLING64) CK(launch_scan_roll(s->view, v, a, s->mode, s->cfg64, s->compute));
    else if (s->algo == SPEQG_ALGO_KPAIR)
    {
        CK(launch_scanp(s->kt, s->kp, s->d_slot_mult, s->d_pair_mult, v, a, s->mode, wpr, s->cfg_p, s->compute));
        if (s->mode & M_RECORD) s->slot_mult_dirty = true;
    }
    else if (s->algo == SPEQG_ALGO_KTABLE)
    {
        // (launch_scanh runs the walking scan on a walk-form table in global mode with one lane per read)
        const bool walks = s->kt.walk && !wpr && !(s->mode & M_LOCAL);
        CK(launch_scanh(s->kt, s->d_slot_mult, v, a, s->mode, wpr, walks ? s->cfg_w : s->cfg_h, s->compute));
        if (s->mode & M_RECORD) (s->d_diff ? s->diff_dirty : s->slot_mult_dirty) = true;
    }
    else CK(launch_scan32(s->d_sblk, s->view, v, a, s->mode, s->cfg, s->compute));
    if ((s->mode & M_RECORD) && s->algo != SPEQG_ALGO_KTABLE && s->algo != SPEQG_ALGO_KPAIR) s->mult_dirty = true;   // these record into mult[] directly
    ++s->launches;
}

Slot & acquire_slot(speqg_scan * s)
{
    Slot & sl = s->slots[s->next_slot];
    sl.ticket = ++s->tickets;
    s->next_slot = (s->next_slot + 1) % kSlots;
    if (!sl.copied)
    {
        CK(cudaEventCreateWithFlags(&sl.copied, cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&sl.done, cudaEventDisableTiming));
    }
    // the staging buffers may be regrown (cudaFree) or overwritten only after the kernel that read them has finished
    if (sl.busy) CK(cudaStreamWaitEvent(s->copy, sl.done, 0));
    return sl;
}
}   // namespace

// Second phase of a FASTQ region: the newline count is on the host by now (or as soon as its event has fired), so the line /
// record arrays can be sized and the rest -- lines, records, pack, scan -- is queued behind the count kernels.
void speqb::fastq_phase2(speqg_scan * s)
{
    if (s->fq_pending < 0) return;
    Slot & sl = s->slots[s->fq_pending];
    s->fq_pending = -1;
    CK(cudaEventSynchronize(sl.counted));
    const uint32_t files = sl.fq_pair ? 2u : 1u;
    const uint64_t n[2] = {sl.fq_bytes, sl.fq_bytes2}, n_nl[2] = {sl.h_total[0], sl.h_total[1]};
    const uint64_t n_lines[2] = {n_nl[0] + (sl.fq_terminated ? 0u : 1u), n_nl[1] + (sl.fq_terminated2 ? 0u : 1u)};
    unsigned int * d_reject = reinterpret_cast<unsigned int *>(sl.fq_small.p + 2);
    unsigned long long * d_nreads = sl.fq_small.p + 1;
    bool whole = n_lines[0] % 4 == 0 && n_lines[0] / 4 < (uint64_t(1) << 31);
    if (sl.fq_pair) whole = whole && n_lines[1] == n_lines[0];          // mate i of file 1 pairs with record i of file 2
    if (!whole)
    {
        // not a whole number of 4-line records (or not as many mates as reads): nothing of the region is scanned; the caller
        // reads it with the host grammar
        unsigned int one = 1u;
        CK(cudaMemcpyAsync(s->d_fq_flags + sl.fq_region, &one, 4, cudaMemcpyHostToDevice, s->compute));
        CK(cudaEventRecord(sl.done, s->compute));
        sl.busy = true;
        return;
    }
    const uint64_t n_rec = n_lines[0] / 4, n_reads = n_rec * files;
    // first 32-base group of each file's reads: read r of a file sits at group base + floor(sequence offset / 32) + r
    const uint64_t gbase[2] = {0, (n[0] >> 5) + n_rec + 1};
    const uint64_t G = gbase[1] + (sl.fq_pair ? (n[1] >> 5) + n_rec + 1 : 0) + 1;
    sl.fq_lines.reserve(n_lines[0] + 2);
    if (sl.fq_pair) sl.fq_lines2.reserve(n_lines[1] + 2);
    sl.fq_soff.reserve(n_reads + 1); sl.fq_qoff.reserve(n_reads + 1);
    sl.len.reserve(n_reads + 1); sl.grp_off.reserve(n_reads + 2); sl.bases.reserve(G); sl.inv.reserve(G);
    const bool local = (s->mode & M_LOCAL) != 0;
    if (local) sl.qual.reserve(G * 32 + 32);
    CK(launch_fastq_begin(d_reject, s->compute));
    CK(launch_fastq_records(sl.raw_seq.p, n[0], sl.fq_chunk.p, n_nl[0], n_rec, sl.fq_lines.p, sl.fq_soff.p, sl.fq_qoff.p, sl.len.p, sl.grp_off.p,
                            files, 0u, gbase[0], s->p.phred_cutoff, sl.bases.p, sl.inv.p, local ? sl.qual.p : nullptr, d_reject, s->compute));
    if (sl.fq_pair)
        CK(launch_fastq_records(sl.raw_qual.p, n[1], sl.fq_chunk2.p, n_nl[1], n_rec, sl.fq_lines2.p, sl.fq_soff.p, sl.fq_qoff.p, sl.len.p,
                                sl.grp_off.p, files, 1u, gbase[1], s->p.phred_cutoff, sl.bases.p, sl.inv.p, local ? sl.qual.p : nullptr, d_reject,
                                s->compute));
    CK(launch_fastq_finalize(d_reject, n_reads, d_nreads, s->d_fq_flags + sl.fq_region, s->compute));
    s->launches += 1 + 3 * files;
    if (n_rec)
    {
        BatchView v{};
        v.bases = sl.bases.p; v.inv = sl.inv.p; v.qual = local ? sl.qual.p : nullptr;
        v.grp_off = sl.grp_off.p; v.len = sl.len.p; v.read_group = nullptr; v.n_reads = n_reads; v.n_reads_dev = d_nreads;
        // a warp per read for long reads: the sequences are at most half of the regions' bytes
        launch_on_view(s, v, ((n[0] + n[1]) / 2) / 32 + 1);
    }
    CK(cudaEventRecord(sl.done, s->compute));
    sl.busy = true;
}

extern "C" {

int speqg_scan_open(speqg_index * idx, int device, const speqg_params * params, speqg_scan ** out)
{
    SPEQG_TRY
    need(idx && params && out, "NULL argument");
    need(params->k >= 1 && params->k <= kMaxK, "k must be in [1, 127]");
    need(params->mode == SPEQG_MODE_GLOBAL || params->mode == SPEQG_MODE_LOCAL, "unknown mode");
    need(params->algorithm <= SPEQG_ALGO_KPAIR, "unknown algorithm");
    need(!(params->algorithm == SPEQG_ALGO_KPAIR && (params->k > kMaxKTable || idx->host.V > kMaxGroupsPair)),
         "the pair-table algorithm needs k <= 31 and at most 254 variants");
    need(!(params->algorithm == SPEQG_ALGO_KTABLE && params->k > kMaxKTable), "the k-mer table algorithm needs k <= 31");
    need(!(params->algorithm == SPEQG_ALGO_KTABLE && idx->host.V > kMaxGroupsTable), "the k-mer table algorithm needs at most 32767 variants");
    need(idx->host.n <= kMaxN32, "index too large: the indexed text must be shorter than 2^31 symbols");
    need(idx->host.V <= kMaxGroupsScan, "too many variants for the scan kernels: the per-CTA tallies (12 B per variant, 24 B in the "
                                       "refinement) live in shared memory; at most 8192 variants");
    need(!(params->algorithm == SPEQG_ALGO_DIRECT && params->mode == SPEQG_MODE_LOCAL), "the direct algorithm covers global mode only");
    DeviceIndex & di = device_index(idx, device);
    device_guard guard(device);
    auto * s = new speqg_scan();
    try
    {
        s->idx = idx;
        s->di = &di;
        s->device = device;
        s->p = *params;
        s->algo = params->algorithm != SPEQG_ALGO_AUTO ? params->algorithm
                  : params->k > kMaxKTable || idx->host.V > kMaxGroupsTable ? SPEQG_ALGO_ROLLING
                                                       : SPEQG_ALGO_KTABLE;   // upgraded to KPAIR below when the table misses L2
        const bool auto_algo = params->algorithm == SPEQG_ALGO_AUTO;
        s->mode = (params->mode == SPEQG_MODE_LOCAL ? M_LOCAL : 0) | (params->record_multiplicity ? M_RECORD : 0) |
                  (params->paired ? M_PAIRED : 0);
        uint32_t V = idx->host.V;
        if (params->stream) s->compute = static_cast<cudaStream_t>(params->stream);
        else
        {
            CK(cudaStreamCreateWithFlags(&s->compute, cudaStreamNonBlocking));
            s->own_compute = true;
        }
        CK(cudaStreamCreateWithFlags(&s->copy, cudaStreamNonBlocking));
        s->state_words = 3 * (size_t)V + SC_COUNT + 2;
        s->d_state = dev_alloc<unsigned long long>(s->state_words);
        CK(cudaMallocHost(&s->h_state, s->state_words * 8));
        s->d_unique = s->d_state;
        s->d_total_g = s->d_unique + V;
        s->d_scalars = s->d_total_g + V;
        s->d_unique_w = reinterpret_cast<double *>(s->d_scalars + SC_COUNT);
        s->d_bad = s->d_scalars + SC_COUNT + V;
        s->d_work = s->d_bad + 1;
        if (s->algo != SPEQG_ALGO_DIRECT && s->algo != SPEQG_ALGO_ROLLING64)
        {
            s->d_fkeys = dev_alloc<unsigned long long>(kFusionCap);
            s->d_fvals = dev_alloc<unsigned long long>(kFusionCap + 1);
            // the product kernels record multiplicities only for k-mers held by several variants; the rest is this count
            if (params->record_multiplicity) s->d_unique_n = dev_alloc<unsigned long long>(V);
        }
        s->d_percent = dev_alloc<double>(V);
        s->d_tkpg = dev_alloc<double>(V);
        s->d_gc = dev_alloc<int32_t>(V);
        if (params->record_multiplicity) s->d_mult = dev_alloc<unsigned long long>(idx->host.n);
        // local-mode denominators with the host's libm, exactly the reference's expression (src/fm_scanner.cpp:454)
        double qden[84];                     // [0, 42): the denominators; [42, 84): their reciprocals (rolling window weight)
        for (int q = 0; q < 42; ++q)
        {
            double b = q;
            qden[q] = 1.0 - 1.0 / pow(10.0, b / 10.0);
            qden[42 + q] = 1.0 / qden[q];
        }
        s->d_qden = dev_upload(qden, 84);
        // rank blocks specialised for this k: one pass over the index, once per scan handle
        s->view = di.view;
        if (s->algo == SPEQG_ALGO_KTABLE || s->algo == SPEQG_ALGO_KPAIR)
        try
        {
            // distinct k-mers of the index <= suffixes that start a depth-k lcp interval; load factor in (0.2, 0.4]
            HostIndex const & h = idx->host;
            uint64_t ub = 0;
            for (uint64_t i = 0; i < h.n; ++i) ub += h.lcp[i] < params->k ? 1u : 0u;
            uint32_t max_len = 0;
            for (uint32_t L : h.t_len) max_len = L > max_len ? L : max_len;
            s->d_kstat = dev_alloc<unsigned long long>(2);
            int l2 = 0;
            CK(cudaDeviceGetAttribute(&l2, cudaDevAttrL2CacheSize, device));
            s->npos = 32ull * h.t_bases.size();
            const bool record = params->record_multiplicity != 0, no_wide = getenv("SPEQB_NO_WIDE") != nullptr;
            // Global mode: the single-window table in walk form (KTabView::walk, scanw_kernel) -- the scan answers most windows
            // from the packed texts and two flag bits per text position instead of a probe, so it is also what AUTO picks for
            // tables beyond L2 (where the pair table used to halve the probes).  SPEQB_NO_WALK=1 keeps the probing scans.
            const bool walk_ok = !(s->mode & M_LOCAL) && !getenv("SPEQB_NO_WALK") && !getenv("SPEQB_NO_RUNS") && s->npos < 0x7FFF0000ull;
            for (uint32_t grow = 0;; ++grow)
            {
                need(grow < 4 && ktable_geometry(params->k, ub, grow, &s->kt), "k-mer table too large");
                // a table beyond L2 that has to count hits per slot keeps the counters inside its (then 64-byte) buckets
                const bool big = 32ull * ((uint64_t)s->kt.nbmask + 1) > (uint64_t)l2 / 4 * 3 || getenv("SPEQB_FORCE_WIDE");
                const bool pair = s->algo == SPEQG_ALGO_KPAIR || (auto_algo && idx->host.V <= kMaxGroupsPair && big && !walk_ok);
                // (only the pair-table scan reads wide tables: the single-window kernel's hot loop addresses 32-byte buckets)
                s->kt.wide = record && big && !no_wide && pair ? 1u : 0u;
                s->kt.slots = dev_alloc<uint64_t>((4ull << s->kt.wide) * ((uint64_t)s->kt.nbmask + 1));
                CK(launch_ktable_build(di.t_bases, di.t_inv, di.t_grp_off, di.t_len, di.t_group, h.S, max_len, s->kt, params->k,
                                       s->d_kstat, s->compute));
                unsigned long long st[2] = {0, 0};
                CK(cudaMemcpyAsync(st, s->d_kstat, 16, cudaMemcpyDeviceToHost, s->compute));
                CK(cudaStreamSynchronize(s->compute));
                s->ktab_keys = st[0];
                if (!st[1]) break;
                // a k-mer ended up more than 6 buckets from its home (very unlikely below load 0.4): double the table
                CK(cudaFree(s->kt.slots));
                s->kt.slots = nullptr;
            }
            if (auto_algo && idx->host.V <= kMaxGroupsPair)
            {
                // A table that does not stay in L2 is bound by the HBM random-sector rate: halving the probes (two windows
                // per probe) pays there (C4: +43 %), while on an L2-resident table the single-window kernel is faster.
                if (32ull * ((uint64_t)s->kt.nbmask + 1) > (uint64_t)l2 / 4 * 3 && !walk_ok) s->algo = SPEQG_ALGO_KPAIR;
            }
            const bool walk = s->algo == SPEQG_ALGO_KTABLE && !s->kt.wide && walk_ok;
            CK(scanh_launch_cfg(s->mode, V, 0, &s->cfg_h));
            if (walk) CK(scanh_launch_cfg(s->mode, V, 1, &s->cfg_w));
            if (walk)
            {
                // the position goes into the slot when it fits next to remainder and displacement (from bit 32 when those fit the
                // low word: one AND extracts it); else into a side array, one 32-bit position per slot
                const uint64_t n_slots = 4ull * ((uint64_t)s->kt.nbmask + 1);
                if (ktable_positions_fit(s->kt, s->npos) && !getenv("SPEQB_FORCE_SIDE"))
                {
                    s->kt.pshift = s->kt.rbits + 3u <= 32u ? 32u : s->kt.rbits + 3u;
                    if (s->kt.rbits + 18u <= 32u) s->kt.vshift = s->kt.rbits + 3u;       // room for the variant too (KTabView::vshift)
                }
                else
                {
                    s->d_pos_side = dev_alloc<uint32_t>(n_slots);
                    s->kt.pos_side = s->d_pos_side;
                }
                s->kt.tgroups = (uint32_t)h.t_bases.size();
                s->d_tflags = dev_alloc<uint64_t>((size_t)s->kt.tgroups + 2);
                s->d_gvar = dev_alloc<uint16_t>((size_t)s->kt.tgroups + 2);
                CK(cudaMemsetAsync(s->d_tflags, 0, ((size_t)s->kt.tgroups + 2) * 8, s->compute));
                CK(cudaMemsetAsync(s->d_gvar, 0, ((size_t)s->kt.tgroups + 2) * 2, s->compute));
                CK(launch_ktable_canon(di.t_bases, di.t_inv, di.t_grp_off, di.t_len, di.t_group, h.S, max_len, s->kt, params->k,
                                       s->d_tflags, s->d_gvar, s->compute));
                s->kt.walk = 1u;
                s->kt.tbases = di.t_bases;
                s->kt.tflags = s->d_tflags;
                s->kt.gvar = s->d_gvar;
                if (params->record_multiplicity)
                {
                    s->d_diff = dev_alloc<unsigned long long>(s->npos + 2);
                    s->scan_tmp_bytes = kpos_fold_tmp_bytes(s->npos);
                    CK(cudaMalloc(&s->d_scan_tmp, s->scan_tmp_bytes));
                }
                // a table beyond L2: the presence filter answers the windows with a sequencing error out of L2
                if ((32ull * ((uint64_t)s->kt.nbmask + 1) > (uint64_t)l2 / 4 * 3 || getenv("SPEQB_FORCE_FILTER")) && !getenv("SPEQB_NO_FILTER"))
                {
                    s->filt_bits = kfilter_bits(params->k, s->ktab_keys);
                    s->kt.fshift = 2u * params->k - s->filt_bits;
                    s->kt.filt = dev_alloc<uint32_t>(std::max<uint64_t>(1, (1ull << s->filt_bits) / 32));
                    CK(launch_kfilter_build(s->kt, s->filt_bits, s->compute));
                }
            }
            else if (params->record_multiplicity)
            {
                // Single-window table scan: hits on k-mers held by several variants are recorded as runs over the positions of
                // the packed texts (two atomics per run of consecutive hits) when a position fits into the slot next to the
                // remainder; else, and for the pair-table scan, one atomic per hit on a per-slot counter.
                if (s->algo == SPEQG_ALGO_KTABLE && !s->kt.wide && !getenv("SPEQB_NO_RUNS") && ktable_positions_fit(s->kt, s->npos))
                {
                    // the position starts at bit 32 when remainder + displacement fit the low word (one AND extracts it)
                    s->kt.pshift = (s->kt.rbits + 3u <= 32u && s->npos < 0x7FFF0000ull) ? 32u : s->kt.rbits + 3u;   // (below the 0x7FFF.. pattern pass 1 leaves)
                    CK(launch_ktable_canon(di.t_bases, di.t_inv, di.t_grp_off, di.t_len, nullptr, h.S, max_len, s->kt, params->k,
                                           nullptr, nullptr, s->compute));
                    s->d_diff = dev_alloc<unsigned long long>(s->npos + 2);
                    s->scan_tmp_bytes = kpos_fold_tmp_bytes(s->npos);
                    CK(cudaMalloc(&s->d_scan_tmp, s->scan_tmp_bytes));
                }
                else if (!s->kt.wide)
                {
                    const uint64_t n_slots = 4ull * ((uint64_t)s->kt.nbmask + 1);
                    s->d_slot_mult = dev_alloc<unsigned long long>(n_slots);
                    CK(cudaMemsetAsync(s->d_slot_mult, 0, n_slots * 8, s->compute));
                }
            }
            if (s->algo == SPEQG_ALGO_KPAIR)
            {
                uint64_t ub2 = 0;
                for (uint64_t i = 0; i < h.n; ++i) ub2 += h.lcp[i] < params->k + 1 ? 1u : 0u;
                for (uint32_t grow = 0;; ++grow)
                {
                    need(grow < 4 && ktable_geometry(params->k + 1, ub2, grow, &s->kp), "pair table too large");
                    s->kp.wide = record && !no_wide && (32ull * ((uint64_t)s->kp.nbmask + 1) > (uint64_t)l2 / 4 * 3 || getenv("SPEQB_FORCE_WIDE")) ? 1u : 0u;
                    s->kp.slots = dev_alloc<uint64_t>((4ull << s->kp.wide) * ((uint64_t)s->kp.nbmask + 1));
                    CK(launch_kpair_build(di.t_bases, di.t_inv, di.t_grp_off, di.t_len, h.S, max_len, s->kt, s->kp, params->k,
                                          s->d_kstat, s->compute));
                    unsigned long long st[2] = {0, 0};
                    CK(cudaMemcpyAsync(st, s->d_kstat, 16, cudaMemcpyDeviceToHost, s->compute));
                    CK(cudaStreamSynchronize(s->compute));
                    s->kpair_keys = st[0];
                    need(st[1] != 2, "internal error: a k-mer of the texts is missing from the k-mer table");
                    if (!st[1]) break;
                    CK(cudaFree(s->kp.slots));
                    s->kp.slots = nullptr;
                }
                if (params->record_multiplicity && !s->kp.wide)
                {
                    const uint64_t n_slots = 4ull * ((uint64_t)s->kp.nbmask + 1);
                    s->d_pair_mult = dev_alloc<unsigned long long>(n_slots);
                    CK(cudaMemsetAsync(s->d_pair_mult, 0, n_slots * 8, s->compute));
                }
                CK(scanp_launch_cfg(s->mode, V, &s->cfg_p));
                // presence filter of the k-mer table: answers the windows with a sequencing error out of L2
                if (!getenv("SPEQB_NO_FILTER"))
                {
                    s->filt_bits = kfilter_bits(params->k, s->ktab_keys);
                    s->kt.fshift = 2u * params->k - s->filt_bits;
                    s->kt.filt = dev_alloc<uint32_t>(std::max<uint64_t>(1, (1ull << s->filt_bits) / 32));
                    CK(launch_kfilter_build(s->kt, s->filt_bits, s->compute));
                }
            }
        }
        catch (api_error const & e)
        {
            // SPEQG_ALGO_AUTO: a table that cannot be built here (geometry beyond 2^31 buckets, not enough device memory for the
            // tables and their counters) is no reason to fail -- the rolling FM search works from the 1 + 2 bytes per suffix the
            // index already holds.  An explicitly requested table algorithm reports the error.
            if (!auto_algo || (e.code != SPEQG_ENOMEM && e.code != SPEQG_EINVAL)) throw;
            cudaGetLastError();
            cudaFree(s->kt.slots); cudaFree(s->kt.filt); cudaFree(s->kp.slots); cudaFree(s->d_slot_mult); cudaFree(s->d_pair_mult);
            cudaFree(s->d_diff); cudaFree(s->d_scan_tmp); cudaFree(s->d_kstat); cudaFree(s->d_tflags); cudaFree(s->d_gvar); cudaFree(s->d_pos_side);
            s->kt = KTabView{}; s->kp = KTabView{};
            s->d_slot_mult = s->d_pair_mult = s->d_diff = s->d_kstat = nullptr;
            s->d_tflags = nullptr; s->d_gvar = nullptr; s->d_pos_side = nullptr;
            s->d_scan_tmp = nullptr;
            s->ktab_keys = s->kpair_keys = 0;
            s->algo = SPEQG_ALGO_ROLLING;
        }
        if (s->algo == SPEQG_ALGO_KTABLE || s->algo == SPEQG_ALGO_KPAIR) {}
        else if (s->algo == SPEQG_ALGO_ROLLING64)
        {
            s->d_blocks_k = dev_alloc<KBlock>(idx->host.blocks.size());
            CK(launch_fill_e(di.blocks, s->d_blocks_k, di.lcp, (uint32_t)idx->host.n, params->k, s->compute));
            s->view.blocks = s->d_blocks_k;
            CK(scan_launch_cfg(s->mode, 0, V, &s->cfg64));
        }
        else
        {
            s->d_sblk = dev_alloc<FBlock>(idx->host.n / 32 + 1);
            CK(launch_fill32(di.blocks, di.lcp, di.pn, di.grp, di.ctab, s->d_sblk, (uint32_t)idx->host.n, params->k, s->compute));
        }
        CK(scan32_launch_cfg(s->mode, V, &s->cfg));
        CK(scan_launch_cfg(s->mode, 1, V, &s->cfg_direct));
        CK(scan_launch_cfg(s->mode, 2, V, &s->cfg_stats));
        int sms = 0;
        CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
        s->refine_grid = sms * 4;
        s->d_partial = dev_alloc<double>((size_t)s->refine_grid * V);
    }
    catch (...)
    {
        scan_release(s);
        delete s;
        throw;
    }
    *out = s;
    int rc = speqg_scan_reset(s);
    if (rc != SPEQG_OK)
    {
        scan_release(s);
        delete s;
        *out = nullptr;
        return rc;
    }
    SPEQG_CATCH
}

int speqg_scan_reset(speqg_scan * s)
{
    SPEQG_TRY
    need(s != nullptr, "NULL argument");
    device_guard guard(s->device);
    fastq_phase2(s);
    s->fq_regions = 0;
    if (s->d_fq_flags) CK(cudaMemsetAsync(s->d_fq_flags, 0, (size_t)kMaxFastqRegions * 4, s->compute));
    CK(cudaMemsetAsync(s->d_state, 0, s->state_words * 8, s->compute));
    if (s->d_fkeys)
    {
        CK(cudaMemsetAsync(s->d_fkeys, 0, (size_t)kFusionCap * 8, s->compute));
        CK(cudaMemsetAsync(s->d_fvals, 0, ((size_t)kFusionCap + 1) * 8, s->compute));
    }
    if (s->d_unique_n) CK(cudaMemsetAsync(s->d_unique_n, 0, (size_t)s->idx->host.V * 8, s->compute));
    if (s->d_diff) CK(cudaMemsetAsync(s->d_diff, 0, (s->npos + 2) * 8, s->compute));
    s->diff_dirty = false;
    // the table scans write mult[] only when the refinement folds their counters into it; a scan that was never refined
    // leaves it clean (8 bytes per suffix not rewritten per step)
    if (s->d_mult && s->mult_dirty)
    {
        CK(cudaMemsetAsync(s->d_mult, 0, s->idx->host.n * 8, s->compute));
        s->mult_dirty = false;
    }
    if (s->d_slot_mult) CK(cudaMemsetAsync(s->d_slot_mult, 0, 4ull * ((uint64_t)s->kt.nbmask + 1) * 8, s->compute));
    if (s->d_pair_mult) CK(cudaMemsetAsync(s->d_pair_mult, 0, 4ull * ((uint64_t)s->kp.nbmask + 1) * 8, s->compute));
    if (s->kt.slots && s->kt.wide) CK(launch_ktable_zero_counters(s->kt, s->compute));
    if (s->kp.slots && s->kp.wide) CK(launch_ktable_zero_counters(s->kp, s->compute));
    s->slot_mult_dirty = false;
    s->launches = 0;
    SPEQG_CATCH
}

int speqg_scan_submit_packed(speqg_scan * s, const speqg_batch * hb)
{
    SPEQG_TRY
    need(s != nullptr, "NULL argument");
    check_batch(s, hb);
    if (hb->n_reads == 0) return SPEQG_OK;
    device_guard guard(s->device);
    fastq_phase2(s);
    Slot & sl = acquire_slot(s);
    uint64_t R = hb->n_reads, G = hb->n_groups32;
    need(G == hb->grp_off[R], "n_groups32 != grp_off[n_reads]");
    sl.bases.reserve(G + 1); sl.inv.reserve(G + 1); sl.grp_off.reserve(R + 1); sl.len.reserve(R);
    CK(cudaMemcpyAsync(sl.bases.p, hb->bases, G * 8, cudaMemcpyHostToDevice, s->copy));
    CK(cudaMemcpyAsync(sl.inv.p, hb->inv, G * 4, cudaMemcpyHostToDevice, s->copy));
    CK(cudaMemcpyAsync(sl.grp_off.p, hb->grp_off, (R + 1) * 8, cudaMemcpyHostToDevice, s->copy));
    CK(cudaMemcpyAsync(sl.len.p, hb->len, R * 4, cudaMemcpyHostToDevice, s->copy));
    if (s->mode & M_LOCAL)
    {
        sl.qual.reserve(G * 32 + 32);
        CK(cudaMemcpyAsync(sl.qual.p, hb->qual, G * 32, cudaMemcpyHostToDevice, s->copy));
    }
    CK(cudaEventRecord(sl.copied, s->copy));
    CK(cudaStreamWaitEvent(s->compute, sl.copied, 0));
    BatchView v{};
    v.bases = sl.bases.p; v.inv = sl.inv.p; v.qual = (s->mode & M_LOCAL) ? sl.qual.p : nullptr;
    v.grp_off = sl.grp_off.p; v.len = sl.len.p; v.read_group = nullptr; v.n_reads = R;
    launch_on_view(s, v, G);
    CK(cudaEventRecord(sl.done, s->compute));
    sl.busy = true;
    SPEQG_CATCH
}

int speqg_scan_submit_raw(speqg_scan * s, const uint8_t * seq, const uint8_t * qual, const uint64_t * offsets, uint64_t n_reads,
                          int encoding)
{
    SPEQG_TRY
    need(s != nullptr && offsets != nullptr, "NULL argument");
    if (n_reads == 0) return SPEQG_OK;
    need(seq != nullptr, "seq is NULL");
    need(encoding == SPEQG_ENC_ASCII || encoding == SPEQG_ENC_RANK, "unknown encoding");
    need(!(s->mode & M_LOCAL) || qual != nullptr, "local-error mode needs qualities");
    need(!(s->mode & M_PAIRED) || n_reads % 2 == 0, "paired mode needs an even number of reads per batch (mates interleaved)");
    need(n_reads < (uint64_t(1) << 32), "more than 2^32 reads in one batch");
    device_guard guard(s->device);
    uint64_t R = n_reads, base = offsets[0], bytes = offsets[R] - base;
    need(offsets[R] >= base, "offsets must be non-decreasing");
    // the device packer lays the reads out itself (group floor(o / 32) + r): no pass over the reads on the host
    uint64_t G = (bytes >> 5) + R + 1;
    fastq_phase2(s);
    Slot & sl = acquire_slot(s);
    sl.bases.reserve(G + 1); sl.inv.reserve(G + 1); sl.grp_off.reserve(R + 1); sl.len.reserve(R); sl.raw_off.reserve(R + 1);
    sl.raw_seq.reserve(bytes + 1);
    if (qual) sl.raw_qual.reserve(bytes + 1);
    if (s->mode & M_LOCAL) sl.qual.reserve(G * 32 + 32);
    CK(cudaMemcpyAsync(sl.raw_seq.p, seq + base, bytes, cudaMemcpyHostToDevice, s->copy));
    if (qual) CK(cudaMemcpyAsync(sl.raw_qual.p, qual + base, bytes, cudaMemcpyHostToDevice, s->copy));
    CK(cudaMemcpyAsync(sl.raw_off.p, offsets, (R + 1) * 8, cudaMemcpyHostToDevice, s->copy));
    CK(cudaEventRecord(sl.copied, s->copy));
    CK(cudaStreamWaitEvent(s->compute, sl.copied, 0));
    // offsets are relative to `seq`; the device copy starts at offsets[0]
    CK(launch_pack(sl.raw_seq.p, qual ? sl.raw_qual.p : nullptr, base, sl.raw_off.p, R, encoding, s->p.phred_cutoff, 0,
                   sl.bases.p, sl.inv.p, (s->mode & M_LOCAL) ? sl.qual.p : nullptr, sl.grp_off.p, 1, sl.len.p, s->d_bad, s->compute));
    ++s->launches;
    BatchView v{};
    v.bases = sl.bases.p; v.inv = sl.inv.p; v.qual = (s->mode & M_LOCAL) ? sl.qual.p : nullptr;
    v.grp_off = sl.grp_off.p; v.len = sl.len.p; v.read_group = nullptr; v.n_reads = R;
    launch_on_view(s, v, G);
    CK(cudaEventRecord(sl.done, s->compute));
    sl.busy = true;
    SPEQG_CATCH
}

static int submit_fastq_impl(speqg_scan * s, const uint8_t * bytes, uint64_t n_bytes, const uint8_t * bytes2, uint64_t n_bytes2, bool pair,
                             uint64_t * region, uint64_t * ticket)
{
    SPEQG_TRY
    need(s != nullptr, "NULL argument");
    need(pair == ((s->mode & M_PAIRED) != 0), pair ? "speqg_scan_submit_fastq_pair needs a scan opened with paired = 1"
                                                   : "speqg_scan_submit_fastq takes single-end input (mates of a pair live in two files: "
                                                     "speqg_scan_submit_fastq_pair)");
    need(s->algo != SPEQG_ALGO_DIRECT && s->algo != SPEQG_ALGO_ROLLING64, "speqg_scan_submit_fastq runs the product kernels only");
    need(s->fq_regions < kMaxFastqRegions, "too many FASTQ regions in one scan (2^20)");
    need(n_bytes < (uint64_t(1) << 40) && n_bytes2 < (uint64_t(1) << 40), "FASTQ region too large");
    device_guard guard(s->device);
    if (!s->d_fq_flags)
    {
        s->d_fq_flags = dev_alloc<unsigned int>(kMaxFastqRegions);
        CK(cudaMemsetAsync(s->d_fq_flags, 0, (size_t)kMaxFastqRegions * 4, s->compute));
    }
    const uint64_t id = s->fq_regions++;
    if (region) *region = id;
    if (n_bytes == 0 && n_bytes2 == 0)
    {
        if (ticket) *ticket = 0;
        return SPEQG_OK;
    }
    need((n_bytes == 0 || bytes != nullptr) && (n_bytes2 == 0 || bytes2 != nullptr), "bytes is NULL");
    // the slot after the pending one: its copy is queued first, so it overlaps the pending region's second phase
    Slot & sl = acquire_slot(s);
    if (!sl.counted)
    {
        CK(cudaEventCreateWithFlags(&sl.counted, cudaEventDisableTiming));
        CK(cudaMallocHost(reinterpret_cast<void **>(&sl.h_total), 16));
    }
    if (ticket) *ticket = sl.ticket;
    sl.raw_seq.reserve(n_bytes + 64);
    sl.fq_chunk.reserve(fastq_chunks(n_bytes) + 1);
    sl.fq_small.reserve(8);
    if (n_bytes) CK(cudaMemcpyAsync(sl.raw_seq.p, bytes, n_bytes, cudaMemcpyHostToDevice, s->copy));
    if (pair)
    {
        sl.raw_qual.reserve(n_bytes2 + 64);                 // (the second file's bytes use the raw-read path's quality staging)
        sl.fq_chunk2.reserve(fastq_chunks(n_bytes2) + 1);
        if (n_bytes2) CK(cudaMemcpyAsync(sl.raw_qual.p, bytes2, n_bytes2, cudaMemcpyHostToDevice, s->copy));
    }
    CK(cudaEventRecord(sl.copied, s->copy));
    sl.fq_pair = pair;
    sl.fq_bytes = n_bytes;
    sl.fq_bytes2 = pair ? n_bytes2 : 0;
    sl.fq_region = id;
    sl.fq_terminated = n_bytes == 0 || bytes[n_bytes - 1] == '\n';
    sl.fq_terminated2 = !pair || n_bytes2 == 0 || bytes2[n_bytes2 - 1] == '\n';
    sl.h_total[1] = 0;
    fastq_phase2(s);                                   // the region submitted before this one
    CK(cudaStreamWaitEvent(s->compute, sl.copied, 0));
    CK(launch_fastq_count(sl.raw_seq.p, n_bytes, sl.fq_chunk.p, sl.fq_small.p, s->compute));
    s->launches += 2;
    CK(cudaMemcpyAsync(sl.h_total, sl.fq_small.p, 8, cudaMemcpyDeviceToHost, s->compute));
    if (pair)
    {
        CK(launch_fastq_count(sl.raw_qual.p, n_bytes2, sl.fq_chunk2.p, sl.fq_small.p + 3, s->compute));
        s->launches += 2;
        CK(cudaMemcpyAsync(sl.h_total + 1, sl.fq_small.p + 3, 8, cudaMemcpyDeviceToHost, s->compute));
    }
    CK(cudaEventRecord(sl.counted, s->compute));
    s->fq_pending = (int)(&sl - s->slots);
    SPEQG_CATCH
}

int speqg_scan_submit_fastq(speqg_scan * s, const uint8_t * bytes, uint64_t n_bytes, uint64_t * region, uint64_t * ticket)
{
    return submit_fastq_impl(s, bytes, n_bytes, nullptr, 0, false, region, ticket);
}

int speqg_scan_submit_fastq_pair(speqg_scan * s, const uint8_t * bytes1, uint64_t n_bytes1, const uint8_t * bytes2, uint64_t n_bytes2,
                                 uint64_t * region, uint64_t * ticket)
{
    return submit_fastq_impl(s, bytes1, n_bytes1, bytes2, n_bytes2, true, region, ticket);
}

int speqg_scan_flush(speqg_scan * s)
{
    SPEQG_TRY
    need(s != nullptr, "NULL argument");
    device_guard guard(s->device);
    fastq_phase2(s);
    SPEQG_CATCH
}

int speqg_scan_last_ticket(speqg_scan * s, uint64_t * ticket)
{
    SPEQG_TRY
    need(s != nullptr && ticket != nullptr, "NULL argument");
    *ticket = s->tickets;
    SPEQG_CATCH
}

int speqg_scan_wait_ticket(speqg_scan * s, uint64_t ticket)
{
    SPEQG_TRY
    need(s != nullptr, "NULL argument");
    if (ticket == 0) return SPEQG_OK;
    device_guard guard(s->device);
    // the staging ring reuses a slot only after the copy of its previous occupant: an evicted ticket has completed
    for (auto & sl : s->slots)
        if (sl.ticket == ticket && sl.copied) CK(cudaEventSynchronize(sl.copied));
    SPEQG_CATCH
}

int speqg_scan_fastq_rejected(speqg_scan * s, uint64_t * regions, uint64_t capacity, uint64_t * n_rejected)
{
    SPEQG_TRY
    need(s != nullptr && n_rejected != nullptr, "NULL argument");
    need(capacity == 0 || regions != nullptr, "regions is NULL");
    *n_rejected = 0;
    if (!s->d_fq_flags || s->fq_regions == 0) return SPEQG_OK;
    device_guard guard(s->device);
    fastq_phase2(s);
    std::vector<unsigned int> flags(s->fq_regions);
    CK(cudaMemcpyAsync(flags.data(), s->d_fq_flags, s->fq_regions * 4, cudaMemcpyDeviceToHost, s->compute));
    CK(cudaStreamSynchronize(s->compute));
    uint64_t n = 0;
    for (uint64_t i = 0; i < s->fq_regions; ++i)
        if (flags[i])
        {
            if (n < capacity) regions[n] = i;
            ++n;
        }
    *n_rejected = n;
    SPEQG_CATCH
}

int speqg_scan_submit_device(speqg_scan * s, const speqg_batch * db)
{
    SPEQG_TRY
    need(s != nullptr, "NULL argument");
    check_batch(s, db);
    if (db->n_reads == 0) return SPEQG_OK;
    device_guard guard(s->device);
    fastq_phase2(s);
    BatchView v{};
    v.bases = db->bases; v.inv = db->inv; v.qual = (s->mode & M_LOCAL) ? db->qual : nullptr;
    v.grp_off = db->grp_off; v.len = db->len; v.read_group = nullptr; v.n_reads = db->n_reads;
    launch_on_view(s, v, db->n_groups32);
    SPEQG_CATCH
}

int speqg_scan_sync(speqg_scan * s)
{
    SPEQG_TRY
    need(s != nullptr, "NULL argument");
    device_guard guard(s->device);
    fastq_phase2(s);
    CK(cudaStreamSynchronize(s->copy));
    CK(cudaStreamSynchronize(s->compute));
    for (auto & sl : s->slots) sl.busy = false;
    SPEQG_CATCH
}

int speqg_scan_wait_copies(speqg_scan * s)
{
    SPEQG_TRY
    need(s != nullptr, "NULL argument");
    device_guard guard(s->device);
    CK(cudaStreamSynchronize(s->copy));
    SPEQG_CATCH
}

int speqg_scan_finish(speqg_scan * s, uint64_t * unique_u64, double * unique_f64, speqg_counts * counts)
{
    SPEQG_TRY
    need(s != nullptr, "NULL argument");
    device_guard guard(s->device);
    fastq_phase2(s);
    // every kernel of the scan runs on the compute stream, behind the copies it waits for: one copy + one sync ends the job
    uint32_t V = s->idx->host.V;
    CK(cudaMemcpyAsync(s->h_state, s->d_state, s->state_words * 8, cudaMemcpyDeviceToHost, s->compute));
    CK(cudaStreamSynchronize(s->compute));
    for (auto & sl : s->slots) sl.busy = false;
    const unsigned long long * hs = s->h_state;
    const unsigned long long bad = hs[3 * (size_t)V + SC_COUNT];
    if (bad & 2ull) throw api_error(SPEQG_EINVAL, "read longer than 2^32 bases");
    if (bad) throw api_error(SPEQG_EINVAL, "parse_error: illegal (non-IUPAC) sequence character in a read");
#if defined(SPEQB_BOUNDS) && SPEQB_BOUNDS
    if (s->kt.walk && hs[2 * (size_t)V + SC_OCC])
        throw api_error(SPEQG_ESTATE, "checking build: " + std::to_string(hs[2 * (size_t)V + SC_OCC]) + " index checks failed in scanw_kernel");
#endif
    if (unique_u64) memcpy(unique_u64, hs, V * 8);
    if (unique_f64) memcpy(unique_f64, hs + 2 * (size_t)V + SC_COUNT, V * 8);
    if (counts)
    {
        const unsigned long long * sc = hs + 2 * (size_t)V;
        memset(counts, 0, sizeof *counts);
        counts->total_kmers = sc[SC_TOTAL];
        counts->ambiguous_reads = sc[SC_AMBIG];
        counts->lookups = sc[SC_LOOKUPS];
        counts->reads = sc[SC_READS];
        counts->kernel_launches = s->launches;
        counts->steps_executed = sc[SC_STEPS];   // table algorithms: 32-byte table probes issued (0 otherwise)
        if (s->kt.walk)
        {
            counts->kmers_hit = sc[SC_HIT];      // the walking scan: windows answered from the texts, without a probe ...
            counts->occ_sectors = sc[SC_OCCSEC]; // ... and the steps (<= 32 windows each) that answered them
        }
    }
    SPEQG_CATCH
}

int speqg_scan_device_counters(speqg_scan * s, uint64_t ** u64_counters, uint64_t * n_u64, double ** f64_counters, uint64_t * n_f64)
{
    SPEQG_TRY
    need(s != nullptr, "NULL argument");
    uint32_t V = s->idx->host.V;
    if (u64_counters) *u64_counters = reinterpret_cast<uint64_t *>(s->d_state);
    if (n_u64) *n_u64 = 2 * (uint64_t)V + SC_COUNT;
    if (f64_counters) *f64_counters = s->d_unique_w;
    if (n_f64) *n_f64 = V;
    SPEQG_CATCH
}

int speqg_scan_fusion_sets(speqg_scan * s, uint64_t * keys, uint64_t * counts, uint64_t capacity, uint64_t * n_sets,
                           uint64_t * unkeyed_units)
{
    SPEQG_TRY
    need(s != nullptr && n_sets != nullptr, "NULL argument");
    need(s->d_fkeys != nullptr, "fusion sets are kept by the product kernels only (not SPEQG_ALGO_DIRECT / SPEQG_ALGO_ROLLING64)");
    need(capacity == 0 || (keys != nullptr && counts != nullptr), "keys / counts are NULL");
    device_guard guard(s->device);
    fastq_phase2(s);
    std::vector<unsigned long long> hk(kFusionCap), hv((size_t)kFusionCap + 1);
    CK(cudaMemcpyAsync(hk.data(), s->d_fkeys, (size_t)kFusionCap * 8, cudaMemcpyDeviceToHost, s->compute));
    CK(cudaMemcpyAsync(hv.data(), s->d_fvals, ((size_t)kFusionCap + 1) * 8, cudaMemcpyDeviceToHost, s->compute));
    CK(cudaStreamSynchronize(s->compute));
    uint64_t n = 0;
    for (uint32_t i = 0; i < kFusionCap; ++i)
    {
        if (hk[i] == 0ull) continue;
        if (n < capacity) { keys[n] = hk[i]; counts[n] = hv[i]; }
        ++n;
    }
    *n_sets = n;
    if (unkeyed_units) *unkeyed_units = hv[kFusionCap];
    SPEQG_CATCH
}

int speqg_scan_fusion_counts(speqg_scan * s, uint64_t * counts)
{
    SPEQG_TRY
    need(s != nullptr && counts != nullptr, "NULL argument");
    need(s->d_fkeys != nullptr, "fusion counts are kept by the product kernels only (not SPEQG_ALGO_DIRECT / SPEQG_ALGO_ROLLING64)");
    uint64_t n = 0, lost = 0;
    int rc = speqg_scan_fusion_sets(s, nullptr, nullptr, 0, &n, &lost);
    if (rc != SPEQG_OK) return rc;
    need(lost == 0, "the fusion map overflowed: some ambiguous units were counted without their variant set");
    std::vector<uint64_t> k(n), c(n);
    rc = speqg_scan_fusion_sets(s, k.data(), c.data(), n, &n, &lost);
    if (rc != SPEQG_OK) return rc;
    const uint32_t V = s->idx->host.V;
    memset(counts, 0, (size_t)V * V * 8);
    for (uint64_t i = 0; i < n; ++i)
    {
        uint32_t a, b;
        if (V <= kMaxGroupsMask) { a = (uint32_t)__builtin_ctzll(k[i]); b = 63u - (uint32_t)__builtin_clzll(k[i]); }
        else { a = (uint32_t)(k[i] & 0xFFFFu); b = (uint32_t)((k[i] >> 16) & 0xFFFFu); }
        counts[(size_t)a * V + b] += c[i];
    }
    SPEQG_CATCH
}

int speqg_scan_stats(speqg_scan * s, const speqg_batch * db, speqg_counts * counts)
{
    SPEQG_TRY
    need(s && db && counts, "NULL argument");
    memset(counts, 0, sizeof *counts);
    if (db->n_reads == 0) return SPEQG_OK;
    need(db->bases && db->inv && db->grp_off && db->len, "batch arrays must not be NULL");
    device_guard guard(s->device);
    unsigned long long * sc_dev = dev_alloc<unsigned long long>(SC_COUNT);
    try
    {
        CK(cudaMemsetAsync(sc_dev, 0, SC_COUNT * 8, s->compute));
        ScanArgs a = scan_args(s);
        a.scalars = sc_dev;
        BatchView v{};
        v.bases = db->bases; v.inv = db->inv; v.qual = nullptr; v.grp_off = db->grp_off; v.len = db->len; v.n_reads = db->n_reads;
        CK(launch_scan_direct(s->view, v, a, s->mode & M_PAIRED, 1, s->cfg_stats, s->compute));
        unsigned long long sc[SC_COUNT];
        CK(cudaMemcpyAsync(sc, sc_dev, sizeof sc, cudaMemcpyDeviceToHost, s->compute));
        CK(cudaStreamSynchronize(s->compute));
        counts->lookups = sc[SC_LOOKUPS];
        counts->reads = sc[SC_READS];
        counts->steps_executed = sc[SC_STEPS];
        counts->kmers_hit = sc[SC_HIT];
        counts->occurrences = sc[SC_OCC];
        counts->occ_sectors = sc[SC_OCCSEC];
    }
    catch (...)
    {
        cudaFree(sc_dev);
        throw;
    }
    cudaFree(sc_dev);
    SPEQG_CATCH
}

int speqg_scan_info(speqg_scan * s, uint32_t * algorithm, uint64_t * structure_bytes, uint64_t * keys)
{
    SPEQG_TRY
    need(s != nullptr, "NULL argument");
    if (algorithm) *algorithm = s->algo;
    if (structure_bytes)
    {
        if (s->algo == SPEQG_ALGO_KTABLE) *structure_bytes = (32ull << s->kt.wide) * ((uint64_t)s->kt.nbmask + 1);
        else if (s->algo == SPEQG_ALGO_KPAIR)
            *structure_bytes = (32ull << s->kt.wide) * ((uint64_t)s->kt.nbmask + 1) + (32ull << s->kp.wide) * ((uint64_t)s->kp.nbmask + 1) +
                               (s->kt.filt ? (1ull << s->filt_bits) / 8 : 0);
        else if (s->algo == SPEQG_ALGO_ROLLING) *structure_bytes = sizeof(FBlock) * (s->idx->host.n / 32 + 1);
        else *structure_bytes = sizeof(KBlock) * s->idx->host.blocks.size();
    }
    if (keys) *keys = s->algo == SPEQG_ALGO_KTABLE ? s->ktab_keys : s->algo == SPEQG_ALGO_KPAIR ? s->ktab_keys + s->kpair_keys : 0;
    SPEQG_CATCH
}

void speqg_scan_close(speqg_scan * s)
{
    if (!s) return;
    int prev = -1;
    cudaGetDevice(&prev);
    cudaSetDevice(s->device);
    cudaStreamSynchronize(s->copy);
    cudaStreamSynchronize(s->compute);
    scan_release(s);
    if (prev >= 0) cudaSetDevice(prev);
    delete s;
}

int speqg_batch_to_device(int device, const speqg_batch * hb, speqg_batch * db)
{
    SPEQG_TRY
    need(hb && db, "NULL argument");
    need(hb->bases && hb->inv && hb->grp_off && (hb->len || hb->n_reads == 0), "batch arrays must not be NULL");
    device_guard guard(device);
    memset(db, 0, sizeof *db);
    uint64_t R = hb->n_reads, G = hb->n_groups32;
    try
    {
        db->bases = dev_upload(hb->bases, G);
        db->inv = dev_upload(hb->inv, G);
        db->grp_off = dev_upload(hb->grp_off, R + 1);
        db->len = dev_upload(hb->len, R);
        if (hb->qual) db->qual = dev_upload(hb->qual, G * 32);
    }
    catch (...)
    {
        cudaFree(db->bases); cudaFree(db->inv); cudaFree(db->grp_off); cudaFree(db->len); cudaFree(db->qual);
        memset(db, 0, sizeof *db);
        throw;
    }
    db->n_reads = R;
    db->n_groups32 = G;
    SPEQG_CATCH
}

int speqg_pack_device(int device, const uint8_t * d_seq, const uint8_t * d_qual, const uint64_t * d_offsets, uint64_t n_reads,
                      int encoding, uint32_t phred_cutoff, int n_only, const speqg_batch * db, void * stream)
{
    SPEQG_TRY
    need(db != nullptr && d_offsets != nullptr, "NULL argument");
    if (n_reads == 0) return SPEQG_OK;
    need(d_seq != nullptr, "d_seq is NULL");
    need(db->bases && db->inv && db->grp_off && db->len, "batch arrays must not be NULL");
    need(encoding == SPEQG_ENC_ASCII || encoding == SPEQG_ENC_RANK, "unknown encoding");
    device_guard guard(device);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    unsigned long long * d_bad = dev_alloc<unsigned long long>(1);
    unsigned long long bad = 0;
    cudaError_t e = cudaMemsetAsync(d_bad, 0, 8, st);
    if (e == cudaSuccess)
        e = launch_pack(d_seq, d_qual, 0, d_offsets, n_reads, encoding, phred_cutoff, n_only, db->bases, db->inv, db->qual, db->grp_off, 0,
                        db->len, d_bad, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(&bad, d_bad, 8, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    cudaFree(d_bad);
    CK(e);
    if (bad & 2ull) throw api_error(SPEQG_EINVAL, "read longer than 2^32 bases");
    if (bad) throw api_error(SPEQG_EINVAL, "parse_error: illegal (non-IUPAC) sequence character in a read");
    SPEQG_CATCH
}

void speqg_batch_free_device(int device, speqg_batch * db)
{
    if (!db) return;
    int prev = -1;
    cudaGetDevice(&prev);
    cudaSetDevice(device);
    cudaFree(db->bases); cudaFree(db->inv); cudaFree(db->grp_off); cudaFree(db->len); cudaFree(db->qual);
    memset(db, 0, sizeof *db);
    if (prev >= 0) cudaSetDevice(prev);
}

}   // extern "C"
