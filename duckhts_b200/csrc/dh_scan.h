// dh_scan.h — state of one record scan (BAM or BCF) over a BGZF image: the segment pipeline shared by dh_bam.cu and
// dh_bcf.cu (H2D -> inflate -> boundary walk -> per-batch columnar decode -> D2H).  Internal, not part of the C-ABI.
#pragma once
#include <deque>
#include <mutex>
#include <string>
#include <vector>

#include "bam_decode.cuh"
#include "dh_index.h"
#include "dh_internal.h"

namespace dh {

constexpr int NSEG = 3;
constexpr int MAX_SLOTS = 32, DEFAULT_SLOTS = 6;   // batch slots of the pipeline (dh_bam_set_slots)
constexpr uint64_t A256 = 255;

struct Batch {           // one output batch of the current segment
    uint32_t b_lo, b_hi; // walk-block index range
    uint64_t rows, heap, row0;
    uint64_t map = 0;    // AUXILIARY_TAGS key/value pairs
};

struct SegSlot {
    DevBuf comp, out, desc, status, work, rec;
    int64_t rec_bias = 0;
    PinBuf h_desc, h_status, h_blk_start, h_entry, h_todo, h_walk, h_row_base, h_heap_base, h_map_base;
    cudaEvent_t ev_mid = nullptr, ev_h2d = nullptr, ev_inf0 = nullptr, ev_inf1 = nullptr, ev_inf_done = nullptr, ev_free = nullptr, ev_walk0 = nullptr,
                ev_walk1 = nullptr, ev_walk2 = nullptr;
    size_t b0 = 0, b1 = 0;  // file block range
    uint64_t cbase = 0, obytes = 0;
    int issued_for = -1;    // segment index whose inflate was issued into this slot
};

enum { SLOT_FREE = 0, SLOT_INFLIGHT = 1, SLOT_HELD = 2 };
struct BatchSlot {
    DevBuf arena;
    PinBuf h_arena;
    cudaEvent_t ev_k0 = nullptr, ev_k1 = nullptr, ev_c0 = nullptr, ev_done = nullptr;
    dh_bam_batch desc;
    DecodeParams dp;          // device view of the batch (dh_bam_batch_checksum on device-resident batches)
    uint64_t dev_recoff = 0;  // arena offset of the compact layout's rec_off array
    bool timed = false;
    bool on_device = false;   // DH_BAM_DEVICE_OUTPUT batch: the host arena was not filled
    bool last = false;        // the scan ends with this batch: it carries the final status
    int state = SLOT_FREE;
};

struct Seg { size_t b0, b1; int stage = -1; };   // stage: file mode, which pinned staging buffer holds the segment's compressed bytes

}  // namespace dh

struct dh_bcf_state;
using namespace dh;   // internal header: the scan state below is made of dh:: types

struct dh_bam {
    int kind = 0;                  // 0 = BAM records, 1 = BCF records (dh_bcf.cu drives the same segment pipeline)
    dh_bcf_state *bcf = nullptr;
    dh_ctx *ctx = nullptr;
    const uint8_t *data = nullptr;
    uint64_t size = 0;
    // file mode (dh_bam_open_file): the image is never held whole; segments are read into two pinned staging buffers by parallel
    // pread and framed (A1) as they arrive, so the block table grows with the scan.  The table is one contiguous run of blocks
    // starting at bt.blocks[0].coffset (the header blocks, or the first block of the range being scanned).
    int fd = -1;
    uint64_t fsize = 0;
    uint64_t frame_pos = 0;        // file offset of the first block not framed yet
    bool frame_done = false;       // nothing more to frame for this scan (end of file, broken block, or past the range end)
    uint64_t limit_coff = UINT64_MAX;   // file mode: block address and in-block offset of the range end (UINT64_MAX: to EOF)
    uint32_t limit_uoff = 0;
    uint64_t first_voff = 0;       // virtual offset of the first alignment record
    std::vector<uint8_t> head;     // file mode: the leading bytes that hold the header blocks
    PinBuf stage[2];
    cudaEvent_t ev_stage[2] = {nullptr, nullptr};   // the last host-to-device copy out of each staging buffer
    bool stage_used[2] = {false, false};
    BlockTable bt;
    bool truncated = false;        // the image ends in a broken / cut block: the scan returns what lies in front of it, then status -2
    // header
    std::string text;
    std::vector<std::string> ref_names;
    std::vector<uint32_t> ref_lens;
    size_t first_blk = 0;
    uint32_t first_off = 0;
    // dictionaries (host strings must keep stable addresses: cells point at them)
    std::deque<std::string> dict;
    std::vector<StrCell> ref_cells;
    std::vector<RgEntry> rg;
    std::vector<std::string> rg_samples;   // SM value of every rg entry ("" = none)
    std::string rg_ids;
    DevBuf d_ref_cells, d_rg, d_rg_ids, d_tags;
    DevBuf d_image;                // whole compressed image staged in HBM (DH_BAM_DEVICE_INPUT)
    uint32_t mode = 0;
    // projection
    uint64_t col_mask = (1ull << BAM_CORE_COLS) - 1;
    std::vector<TagReq> tags;
    ScanTables tabs;
    bool aux_map = false;          // AUXILIARY_TAGS MAP column requested (A16)
    std::vector<uint32_t> aux_excl; // tags left out of it (bit set over t0 | t1 << 8), empty = none
    DevBuf d_aux_excl;
    // knobs
    uint64_t seg_comp = 1024ull << 20;
    uint32_t batch_rows = 1u << 20;
    uint64_t carry_cap = 8ull << 20;
    // scan state
    std::vector<Seg> segs;
    SegSlot ss[NSEG];
    BatchSlot bs[MAX_SLOTS];
    int n_slots = DEFAULT_SLOTS;
    int auto_slot = -1;            // slot handed out by dh_bam_next_batch: released by the next call
    mutable std::mutex mu;         // serialises the entry points that touch the scan state (worker threads of one query)
    bool scanning = false;
    int cur_seg = -1;              // segment whose batches are being issued
    std::vector<Batch> batches;    // of cur_seg
    size_t next_batch_i = 0;
    uint64_t issued = 0, returned = 0;  // batch counters
    std::deque<int> inflight;           // slots in issue order
    uint64_t rows_total = 0;
    int64_t carry = 0;             // bytes of the incomplete record at the end of the previous segment
    int final_status = 0;
    bool chain_done = false;       // range end / error / EOF reached: no further segments
    size_t end_blk = 0;            // first block NOT inflated for this scan
    size_t limit_blk = 0;          // block index of voff_end (records starting at/after (limit_blk, limit_off) are excluded)
    uint32_t limit_off = 0;
    bool to_eof = true;
    int64_t first_entry = 0;       // in-block offset of the first record of the scan
    dh_bam_scan_stats st;
    cudaEvent_t ev_scan0 = nullptr;
    // indexed scans (A8): overlap filter applied by the walk kernel + the virtual-offset ranges still to scan
    int filt_mode = 0;
    std::vector<RegIntv> filt;
    DevBuf d_filt;
    std::deque<IdxChunk> pending;
    // N3: record predicate evaluated by the walk kernel (dh_bam_set_filter)
    uint32_t flag_require = 0, flag_exclude = 0, min_mapq = 0;
    // dh_bam_batch_checksum: dictionary strings the device cannot follow (host pointers in cells) -> their hashes
    DevBuf d_dict_ptr, d_dict_hash, d_ref_hash, d_sample_hash, d_sums;
    PinBuf h_sums;
    bool checksum_tables = false;
    std::vector<uint64_t> h_ref_hash, h_sample_hash, h_dict_ptr, h_dict_hash;
};


namespace dh {
// shared pipeline steps (dh_bam.cu)
int scan_ensure_events(dh_bam *b);
int scan_begin_range(dh_bam *b, uint64_t voff_beg, uint64_t voff_end, bool fresh);
// 1 = a batch of the current segment is ready to be issued, 0 = scan exhausted, < 0 error
int scan_have_work(dh_bam *b);
// batch slots: a free slot or -1; give a held slot back
int scan_take_slot(dh_bam *b);
void scan_release_slot(dh_bam *b, int slot);
int scan_free_slots(const dh_bam *b);
// indexed scan over b->filt (intervals): plans the virtual-offset ranges, uploads the filter, begins the first range
int scan_start_filtered(dh_bam *b, const dh_index *idx, int mode);
// inflate the leading blocks of the image until `want` inflated bytes are available (or the file ends)
int scan_inflate_head(dh_bam *b, uint64_t want, std::vector<uint8_t> &u);
// BCF side (dh_bcf.cu)
cudaError_t bcf_launch_walk(dh_bam *b, const WalkParams &wp, cudaStream_t stream);
void bcf_state_free(dh_bcf_state *s);
}  // namespace dh
