// host/profiles.cpp -- see profiles.hpp.
#include "profiles.hpp"

#include <sys/stat.h>

#include <algorithm>
#include <fstream>
#include <cstring>
#include <numeric>
#include <sstream>
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

// The general reader: any RFC 4180 file, one character at a time.
static std::map<std::string, CellProfile> read_cn_profiles_stream(std::istream& in, const std::string& path) {
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

// Plain decimal integer in [p, e): the common case of std::stoi; anything else goes through to_int.
static int fast_int(const char* p, const char* e, const char* column) {
    const char* q = p;
    bool neg = false;
    if (q < e && (*q == '-' || *q == '+')) neg = *q++ == '-';
    long v = 0;
    const char* digits = q;
    while (q < e && *q >= '0' && *q <= '9' && q - digits < 9) v = v * 10 + (*q++ - '0');
    if (q == e && q > digits) return (int)(neg ? -v : v);
    return to_int(std::string(p, e), column);
}

std::map<std::string, CellProfile> read_cn_profiles(const std::string& path) {
    std::ifstream in(path, std::ios::binary);
    if (!in) throw std::runtime_error("Cannot open file " + path);
    std::string buf;
    in.seekg(0, std::ios::end);
    const std::streamoff size = in.tellg();
    in.seekg(0, std::ios::beg);
    if (size > 0) {
        buf.resize((size_t)size);
        in.read(&buf[0], size);
        buf.resize((size_t)in.gcount());
    }
    // Quotes (or a bare CR) are rare in these files: they take the general reader.  Everything else is
    // split in place -- one pass, no per-field strings, the cell looked up only when the node changes.
    bool plain = buf.find('"') == std::string::npos;
    for (size_t q = buf.find('\r'); plain && q != std::string::npos; q = buf.find('\r', q + 1))
        if (q + 1 >= buf.size() || buf[q + 1] != '\n') plain = false;
    if (!plain) {
        std::istringstream ss(buf);
        return read_cn_profiles_stream(ss, path);
    }
    const char* p = buf.data();
    const char* const end = p + buf.size();
    auto next_line = [&](const char*& lo, const char*& hi) {  // [lo, hi) without the line terminator
        if (p >= end) return false;
        lo = p;
        const char* nl = (const char*)memchr(p, '\n', (size_t)(end - p));
        hi = nl ? nl : end;
        p = nl ? nl + 1 : end;
        if (hi > lo && hi[-1] == '\r') hi--;
        return true;
    };
    const char *lo, *hi;
    if (!next_line(lo, hi)) throw std::runtime_error("empty CSV: " + path);
    std::vector<std::string> header;
    for (const char* f = lo;;) {
        const char* c = (const char*)memchr(f, ',', (size_t)(hi - f));
        header.emplace_back(f, c ? c : hi);
        if (!c) break;
        f = c + 1;
    }
    auto column = [&](const char* name) {
        auto it = std::find(header.begin(), header.end(), name);
        if (it == header.end()) throw std::runtime_error(std::string("Can't find a column named ") + name);
        return (size_t)(it - header.begin());
    };
    const size_t c_node = column("node"), c_chrom = column("chrom"), c_start = column("start"), c_end = column("end"),
                 c_cn = column("cn_a");
    const size_t width = header.size();
    std::map<std::string, CellProfile> cells;
    std::string last_name;
    CellProfile* cell = nullptr;
    std::vector<std::pair<const char*, const char*>> f(width);
    const std::string allele = "cn_a";
    size_t rows_per_cell = 0;
    while (p < end) {
        // one scan of the line: field boundaries and the terminator together (lines are ~25 bytes)
        size_t nf = 0;
        const char* s0 = p;
        const char* q = p;
        for (; q < end && *q != '\n'; q++) {
            if (*q == ',') {
                if (nf < width) f[nf] = {s0, q};
                nf++;
                s0 = q + 1;
            }
        }
        const char* hi2 = q;
        p = q < end ? q + 1 : end;
        if (hi2 > s0 && hi2[-1] == '\r') hi2--;
        if (nf == 0 && hi2 == s0) continue;  // blank line
        if (nf < width) f[nf] = {s0, hi2};
        nf++;
        if (nf < width) throw std::runtime_error("short CSV row in " + path);
        const size_t name_len = (size_t)(f[c_node].second - f[c_node].first);
        if (!cell || name_len != last_name.size() || memcmp(f[c_node].first, last_name.data(), name_len) != 0) {
            if (cell) rows_per_cell = std::max(rows_per_cell, cell->cn.size());
            last_name.assign(f[c_node].first, name_len);
            cell = &cells[last_name];
            if (cell->cn.empty() && rows_per_cell) {  // the cells of a file have the same number of bins
                cell->cn.reserve(rows_per_cell);
                cell->bins.reserve(rows_per_cell);
            }
        }
        cell->bins.emplace_back();
        GenomicBin& b = cell->bins.back();
        b.chromosome.assign(f[c_chrom].first, f[c_chrom].second);
        b.allele = allele;
        b.start = fast_int(f[c_start].first, f[c_start].second, "start");
        b.end = fast_int(f[c_end].first, f[c_end].second, "end");
        cell->cn.push_back(fast_int(f[c_cn].first, f[c_cn].second, "cn_a"));
    }
    return cells;
}

SegmentOrder make_segment_order(const std::vector<GenomicBin>& bins) {
    SegmentOrder o;
    segment_order(bins, o.perm, o.seg_end);
    return o;
}

void breakpoint_transform(const CellProfile& in, const SegmentOrder& o, std::vector<int>& bp, std::vector<GenomicBin>& bins) {
    const std::vector<size_t>& perm = o.perm;
    bp.resize(perm.size());
    bins.resize(perm.size());
    size_t lo = 0;
    for (size_t hi : o.seg_end) {
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

void breakpoint_transform(const CellProfile& in, std::vector<int>& bp, std::vector<GenomicBin>& bins) {
    breakpoint_transform(in, make_segment_order(in.bins), bp, bins);
}

void copynumber_transform(const std::vector<int>& bp, const std::vector<GenomicBin>& bins, const SegmentOrder& o, std::vector<int>& cn,
                          std::vector<GenomicBin>* out_bins) {
    const std::vector<size_t>& perm = o.perm;
    cn.resize(perm.size());
    if (out_bins) out_bins->resize(perm.size());
    size_t lo = 0;
    for (size_t hi : o.seg_end) {
        for (size_t k = lo; k < hi; k++) {
            const int v = bp[perm[k]];
            if (k == lo) cn[k] = v + 2;
            else if (k == hi - 1) cn[k] = 2 - v;
            else cn[k] = v + cn[k - 1];
            if (out_bins) (*out_bins)[k] = bins[perm[k]];
        }
        lo = hi;
    }
}

void copynumber_transform(const std::vector<int>& bp, const std::vector<GenomicBin>& bins, std::vector<int>& cn,
                          std::vector<GenomicBin>& out_bins) {
    copynumber_transform(bp, bins, make_segment_order(bins), cn, &out_bins);
}

static bool same_bins(const std::vector<GenomicBin>& a, const std::vector<GenomicBin>& b) {
    if (a.size() != b.size()) return false;
    for (size_t i = 0; i < a.size(); i++)
        if (a[i].start != b[i].start || a[i].end != b[i].end || a[i].chromosome != b[i].chromosome || a[i].allele != b[i].allele) return false;
    return true;
}

BreakpointSet to_breakpoint_set(const std::map<std::string, CellProfile>& cells) {
    BreakpointSet out;
    std::vector<int> bp;
    // cells normally list the same bins in the same order: the segment order (a map + stable sorts on
    // strings) is then computed once and reused
    const CellProfile* order_of = nullptr;
    SegmentOrder order;
    for (const auto& kv : cells) {
        if (!order_of || !same_bins(order_of->bins, kv.second.bins)) {
            order = make_segment_order(kv.second.bins);
            order_of = &kv.second;
        }
        breakpoint_transform(kv.second, order, bp, out.bins);
        if (out.names.empty()) out.L = (int)bp.size();
        if ((int)bp.size() != out.L)
            throw std::runtime_error("cell '" + kv.first + "' has " + std::to_string(bp.size()) + " bins, expected " +
                                     std::to_string(out.L) + " (all cells must cover the same bins)");
        out.names.push_back(kv.first);
        out.bp.insert(out.bp.end(), bp.begin(), bp.end());
    }
    return out;
}

// ---- binary profile cache -------------------------------------------------------------------------------
namespace {
const char kCacheMagic[8] = {'L', 'Z', 'B', 'P', 'v', '1', 0, 0};

bool csv_identity(const std::string& path, uint64_t& size, int64_t& mtime_ns) {
    struct stat st;
    if (stat(path.c_str(), &st) != 0) return false;
    size = (uint64_t)st.st_size;
    mtime_ns = (int64_t)st.st_mtim.tv_sec * 1000000000ll + st.st_mtim.tv_nsec;
    return true;
}
void put_str(std::vector<char>& out, const std::string& s) {
    const uint32_t n = (uint32_t)s.size();
    out.insert(out.end(), (const char*)&n, (const char*)&n + 4);
    out.insert(out.end(), s.begin(), s.end());
}
bool get_str(const char*& p, const char* end, std::string& s) {
    if (end - p < 4) return false;
    uint32_t n;
    memcpy(&n, p, 4);
    p += 4;
    if ((size_t)(end - p) < n) return false;
    s.assign(p, n);
    p += n;
    return true;
}
}  // namespace

bool save_breakpoint_cache(const std::string& cache_path, const std::string& csv_path, const BreakpointSet& set) {
    uint64_t size = 0;
    int64_t mtime = 0;
    if (!csv_identity(csv_path, size, mtime)) return false;
    std::vector<char> head(kCacheMagic, kCacheMagic + 8);
    const uint64_t n = set.names.size(), L = (uint64_t)set.L, nb = set.bins.size();
    for (const uint64_t v : {size, (uint64_t)mtime, n, L, nb}) head.insert(head.end(), (const char*)&v, (const char*)&v + 8);
    for (const auto& nm : set.names) put_str(head, nm);
    for (const auto& b : set.bins) {
        put_str(head, b.chromosome);
        put_str(head, b.allele);
        const int32_t se[2] = {b.start, b.end};
        head.insert(head.end(), (const char*)se, (const char*)se + 8);
    }
    while (head.size() % 8) head.push_back(0);
    const std::string tmp = cache_path + ".tmp";
    FILE* f = std::fopen(tmp.c_str(), "wb");
    if (!f) return false;
    bool ok = std::fwrite(head.data(), 1, head.size(), f) == head.size();
    ok = ok && std::fwrite(set.bp.data(), sizeof(int32_t), set.bp.size(), f) == set.bp.size();
    ok = (std::fclose(f) == 0) && ok;
    if (ok) ok = std::rename(tmp.c_str(), cache_path.c_str()) == 0;
    if (!ok) std::remove(tmp.c_str());
    return ok;
}

bool load_breakpoint_cache(const std::string& cache_path, const std::string& csv_path, BreakpointSet& set) {
    uint64_t size = 0;
    int64_t mtime = 0;
    if (!csv_identity(csv_path, size, mtime)) return false;
    FILE* f = std::fopen(cache_path.c_str(), "rb");
    if (!f) return false;
    std::fseek(f, 0, SEEK_END);
    const long total = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    std::vector<char> data((size_t)std::max(0l, total));
    const bool read_ok = total > 0 && std::fread(data.data(), 1, data.size(), f) == data.size();
    std::fclose(f);
    if (!read_ok || data.size() < 48 || memcmp(data.data(), kCacheMagic, 8) != 0) return false;
    uint64_t h[5];
    memcpy(h, data.data() + 8, 40);
    if (h[0] != size || (int64_t)h[1] != mtime) return false;  // the CSV changed since the cache was written
    const uint64_t n = h[2], L = h[3], nb = h[4];
    if (L > (1ull << 31) || n > (1ull << 31) || nb > (1ull << 31)) return false;
    const char* p = data.data() + 48;
    const char* end = data.data() + data.size();
    BreakpointSet out;
    out.L = (int)L;
    out.names.resize(n);
    for (auto& nm : out.names)
        if (!get_str(p, end, nm)) return false;
    out.bins.resize(nb);
    for (auto& b : out.bins) {
        if (!get_str(p, end, b.chromosome) || !get_str(p, end, b.allele) || end - p < 8) return false;
        int32_t se[2];
        memcpy(se, p, 8);
        p += 8;
        b.start = se[0], b.end = se[1];
    }
    size_t off = (size_t)(p - data.data());
    off = (off + 7) / 8 * 8;
    if (data.size() - off != n * L * sizeof(int32_t)) return false;
    out.bp.resize(n * L);
    memcpy(out.bp.data(), data.data() + off, n * L * sizeof(int32_t));
    set = std::move(out);
    return true;
}

}  // namespace lazac_host
