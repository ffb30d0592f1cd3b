// host/profiles.hpp -- copy-number profile ingest: CSV -> per-cell copy-number vectors ->
// breakpoint vectors in the reference's position order, and the inverse transform
// (/root/reference/src/lazac.cxx:41-61 read_cn_profiles, src/copy_number.cxx:25-59
// convert_to_breakpoint_profile, :61-95 convert_to_copynumber_profile).
#pragma once
#include <cstdint>
#include <map>
#include <string>
#include <vector>

namespace lazac_host {

struct GenomicBin {
    std::string chromosome, allele;
    int start = 0, end = 0;
    bool operator<(const GenomicBin& o) const;  // (chromosome, allele, start, end), strings byte-wise
};

struct CellProfile {
    std::vector<int> cn;
    std::vector<GenomicBin> bins;
};

// node -> profile in file order; only the cn_a column is read, allele is always "cn_a"
// (src/lazac.cxx:49-51).  std::map: cells iterate in byte-lexicographic name order.
std::map<std::string, CellProfile> read_cn_profiles(const std::string& path);

struct BreakpointSet {
    std::vector<std::string> names;   // sorted (std::map order == std::sort order of src/lazac.cxx:69)
    int L = 0;
    std::vector<int32_t> bp;          // [names.size()][L]
    std::vector<GenomicBin> bins;     // position annotation of the LAST cell (src/lazac.cxx:146)
};

// convert_to_breakpoint_profile(cn, 2) for every cell.
BreakpointSet to_breakpoint_set(const std::map<std::string, CellProfile>& cells);

// Binary profile cache: the BreakpointSet of a CSV as one file next to it -- a header with the
// size and modification time of the CSV it was made from, the cell names, the bins, the int32 matrix -- so that a second
// run over the same input skips the CSV parse and the per-cell transform.  load returns false when the cache is missing,
// stale (the CSV changed) or malformed; nothing but `lazac --cache` reads or writes it.
bool save_breakpoint_cache(const std::string& cache_path, const std::string& csv_path, const BreakpointSet& set);
bool load_breakpoint_cache(const std::string& cache_path, const std::string& csv_path, BreakpointSet& set);

// Position order of convert_to_breakpoint_profile / convert_to_copynumber_profile: positions grouped
// by (chromosome, allele) in std::map order, each group stable-sorted by bin (src/copy_number.cxx:28-41).
struct SegmentOrder {
    std::vector<size_t> perm;      // output position -> input position
    std::vector<size_t> seg_end;   // end of every segment in output positions
};
SegmentOrder make_segment_order(const std::vector<GenomicBin>& bins);

// Per (chromosome, allele) segment: bp[0] = cn[0] - 2, bp[i] = cn[i] - cn[i-1], bp[last] = 2 - cn[last].
void breakpoint_transform(const CellProfile& in, std::vector<int>& bp, std::vector<GenomicBin>& bins);
// Inverse with the same first/last rules, for one labelled vector over `bins`.
void copynumber_transform(const std::vector<int>& bp, const std::vector<GenomicBin>& bins, std::vector<int>& cn,
                          std::vector<GenomicBin>& out_bins);

// The same two with a precomputed order (one order serves every cell / node over the same bins).
void breakpoint_transform(const CellProfile& in, const SegmentOrder& order, std::vector<int>& bp, std::vector<GenomicBin>& bins);
void copynumber_transform(const std::vector<int>& bp, const std::vector<GenomicBin>& bins, const SegmentOrder& order, std::vector<int>& cn,
                          std::vector<GenomicBin>* out_bins);

}  // namespace lazac_host
