// dh_index.h — BAI/CSI index model + region planning (host side of SURVEY §8 row A8); see dh_index.cu.
#pragma once
#include <stdint.h>

#include <string>
#include <vector>

#include "bam_decode.cuh"

namespace dh {

struct IdxChunk { uint64_t u, v; };        // [u, v) virtual offsets (hts_pair64_t)
struct IdxBin {
    uint32_t bin;
    uint64_t loff;                         // CSI: stored; BAI: derived from the linear index (update_loff)
    std::vector<IdxChunk> chunks;
};
struct IdxRef {
    std::vector<IdxBin> bins;              // sorted by bin number (htslib keeps a hash; lookups are the same)
    std::vector<uint64_t> lin;             // BAI linear index (16 kb windows), gaps filled like idx_read_core
    const IdxBin *find(uint32_t bin) const;
};
struct HtsIndex {
    bool csi = false;
    int min_shift = 14, n_lvls = 5;
    uint32_t n_bins = 37449;               // bins of levels 0..n_lvls; the meta bin is n_bins + 1
    std::vector<IdxRef> refs;
    uint64_t n_no_coor = 0;
};

// Parse an (uncompressed) BAI or CSI image.
bool index_parse(const uint8_t *data, uint64_t size, HtsIndex &ix, std::string &err);
// sam_itr_regarray's front half: region strings -> sorted, merged intervals.  Returns the number of intervals
// (0 = no region names a known contig), or -1 with err set.
int regions_to_intervals(const char *const *regions, int n_regions, const std::vector<std::string> &ref_names,
                         std::vector<RegIntv> &out, std::string &err);
void query_chunks(const HtsIndex &ix, int tid, int64_t beg, int64_t end, std::vector<IdxChunk> &chunks);
// chunk lists of all intervals reduced to sorted, disjoint virtual-offset ranges to scan
void plan_ranges(const HtsIndex &ix, const std::vector<RegIntv> &intervals, std::vector<IdxChunk> &ranges);
// [first chunk begin, last chunk end) of one contig; false if the contig has no indexed records
bool contig_span(const HtsIndex &ix, int tid, uint64_t &u, uint64_t &v);

}  // namespace dh

struct dh_index {
    dh::HtsIndex ix;
};
