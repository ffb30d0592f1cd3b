// host/profiles.cpp -- see profiles.hpp.
#include "profiles.hpp"

#include <algorithm>
#include <fstream>
#include <numeric>
#include <stdexcept>

namespace lazac_host {

bool GenomicBin::operator<(const GenomicBin& o) const {
    if (chromosome != o.chromosome) return chromosome < o.chromosome;
    if (allele != o.allele) return allele < o.allele;
    if (start != o.start) return start < o.start;
    return end < o.end;
}

namespace {

// One CSV record (RFC 4180 quoting, optional CR before LF).  Returns false at end of input.
bool next_record(std::istream& in, std::vector<std::string>& fields) {
    fields.clear();
    std::string cur;
    bool any = false, quoted = false;
    int ch;
    while ((ch = in.get()) != EOF) {
        any = true;
        char c = (char)ch;
        if (quoted) {
            if (c == '"') {
                if (in.peek() == '"') {
                    cur += '"';
                    in.get();
                } else {
                    quoted = false;
                }
            } else {
                cur += c;
            }
        } else if (c == '"' && cur.empty()) {
            quoted = true;
        } else if (c == ',') {
            fields.push_back(cur);
            cur.clear();
        } else if (c == '\n') {
            break;
        } else if (c != '\r') {
            cur += c;
        }
    }
    if (!any) return false;
    fields.push_back(cur);
    return true;
}

int to_int(const std::string& s, const char* column) {
    try {
        size_t used = 0;
        int v = std::stoi(s, &used);
        if (used != s.size()) throw std::invalid_argument(s);
        return v;
    } catch (const std::exception&) {
        throw std::runtime_error(std::string("column '") + column + "': not an integer: '" + s + "'");
    }
}

// Groups positions by (chromosome, allele) in std::map order and stable-sorts each group by bin;
// returns the permutation (output position -> input position) and the segment boundaries.
void segment_order(const std::vector<GenomicBin>& bins, std::vector<size_t>& perm, std::vector<size_t>& seg_end) {
    std::map<std::pair<std::string, std::string>, std::vector<size_t>> groups;
    for (size_t i = 0; i < bins.size(); i++) groups[{bins[i].chromosome, bins[i].allele}].push_back(i);
    perm.clear();
    seg_end.clear();
    for (auto& kv : groups) {
        std::vector<size_t>& idx = kv.second;
        std::stable_sort(idx.begin(), idx.end(), [&](size_t a, size_t b) { return bins[a] < bins[b]; });
        perm.insert(perm.end(), idx.begin(), idx.end());
        seg_end.push_back(perm.size());
    }
}

}  // namespace

std::map<std::string, CellProfile> read_cn_profiles(const std::string& path) {
    std::ifstream in(path, std::ios::binary);
    if (!in) throw std::runtime_error("Cannot open file " + path);
    std::vector<std::string> rec;
    if (!next_record(in, rec)) throw std::runtime_error("empty CSV: " + path);
    auto column = [&](const char* name) {
        auto it = std::find(rec.begin(), rec.end(), name);
        if (it == rec.end()) throw std::runtime_error(std::string("Can't find a column named ") + name);
        return (size_t)(it - rec.begin());
    };
    const size_t c_node = column("node"), c_chrom = column("chrom"), c_start = column("start"), c_end = column("end"),
                 c_cn = column("cn_a");
    const size_t width = rec.size();
    std::map<std::string, CellProfile> cells;
    while (next_record(in, rec)) {
        if (rec.size() == 1 && rec[0].empty()) continue;  // blank line
        if (rec.size() < width) throw std::runtime_error("short CSV row in " + path);
        CellProfile& cell = cells[rec[c_node]];
        GenomicBin b;
        b.chromosome = rec[c_chrom];
        b.allele = "cn_a";
        b.start = to_int(rec[c_start], "start");
        b.end = to_int(rec[c_end], "end");
        cell.cn.push_back(to_int(rec[c_cn], "cn_a"));
        cell.bins.push_back(b);
    }
    return cells;
}

void breakpoint_transform(const CellProfile& in, std::vector<int>& bp, std::vector<GenomicBin>& bins) {
    std::vector<size_t> perm, seg_end;
    segment_order(in.bins, perm, seg_end);
    bp.resize(perm.size());
    bins.resize(perm.size());
    size_t lo = 0;
    for (size_t hi : seg_end) {
        for (size_t k = lo; k < hi; k++) {
            const int v = in.cn[perm[k]];
            if (k == lo) bp[k] = v - 2;
            else if (k == hi - 1) bp[k] = 2 - v;
            else bp[k] = v - in.cn[perm[k - 1]];
            bins[k] = in.bins[perm[k]];
        }
        lo = hi;
    }
}

void copynumber_transform(const std::vector<int>& bp, const std::vector<GenomicBin>& bins, std::vector<int>& cn,
                          std::vector<GenomicBin>& out_bins) {
    std::vector<size_t> perm, seg_end;
    segment_order(bins, perm, seg_end);
    cn.resize(perm.size());
    out_bins.resize(perm.size());
    size_t lo = 0;
    for (size_t hi : seg_end) {
        for (size_t k = lo; k < hi; k++) {
            const int v = bp[perm[k]];
            if (k == lo) cn[k] = v + 2;
            else if (k == hi - 1) cn[k] = 2 - v;
            else cn[k] = v + cn[k - 1];
            out_bins[k] = bins[perm[k]];
        }
        lo = hi;
    }
}

BreakpointSet to_breakpoint_set(const std::map<std::string, CellProfile>& cells) {
    BreakpointSet out;
    std::vector<int> bp;
    for (const auto& kv : cells) {
        breakpoint_transform(kv.second, bp, out.bins);
        if (out.names.empty()) out.L = (int)bp.size();
        if ((int)bp.size() != out.L)
            throw std::runtime_error("cell '" + kv.first + "' has " + std::to_string(bp.size()) + " bins, expected " +
                                     std::to_string(out.L) + " (all cells must cover the same bins)");
        out.names.push_back(kv.first);
        out.bp.insert(out.bp.end(), bp.begin(), bp.end());
    }
    return out;
}

}  // namespace lazac_host
