// host/main.cpp -- the `lazac` command line of the reference (subcommands `distance` and `nni`,
// same flags, same output files) over liblazac_b200.so.
// Follows /root/reference/src/lazac.cxx: do_distance :97-132, do_nni :134-338, main :340-431.
// Additive, not in the reference: `--device N` on both subcommands; the GPU-free helper
// subcommands `nj` (seed tree from a PREFIX_dist_matrix.txt) and `bp` (breakpoint matrix of a CSV),
// used by the CPU-side tests.
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <future>
#include <memory>
#include <ctime>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <map>
#include <optional>
#include <random>
#include <sstream>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

#include "lazac_b200.h"
#include "newick.hpp"
#include "profiles.hpp"
#include "search.hpp"
#include "seed_tree.hpp"

using namespace lazac_host;

namespace {

// "[2026-11-04 17:31:08.321] [lazac] [info] message" (the reference's spdlog default pattern)
void log_info(const std::string& msg) {
    using namespace std::chrono;
    const auto now = system_clock::now();
    const std::time_t tt = system_clock::to_time_t(now);
    const int ms = (int)(duration_cast<milliseconds>(now.time_since_epoch()).count() % 1000);
    std::tm tm_buf;
    localtime_r(&tt, &tm_buf);
    char stamp[64];
    std::strftime(stamp, sizeof stamp, "%Y-%m-%d %H:%M:%S", &tm_buf);
    std::printf("[%s.%03d] [lazac] [info] %s\n", stamp, ms, msg.c_str());
    std::fflush(stdout);
}

struct Args {
    std::string command, cn_profile, tree, output;
    double aggression = 1.0;
    int iterations = 100, seed = 0, device = 0;
    std::vector<int> devices;  // --devices: several GPUs of the box for the distance matrix (trees run on the first)
    bool greedy = false, have_output = false;
    bool cache = false;  // --cache: keep / use the binary profile cache next to the CSV
};

const char* USAGE =
    "Usage: lazac distance cn_profile -o PREFIX [--device N | --devices LIST]\n"
    "       lazac nni cn_profile [-t TREE] -o PREFIX [-a AGGRESSION] [-i ITERATIONS] [-g] [-s SEED] [--device N | --devices LIST]\n"
    "\n"
    "  -t, --tree        seed tree in Newick format. [default: \"\"]\n"
    "  -o, --output      prefix of the output files [required]\n"
    "  -a, --aggression  aggression of stochastic perturbation in (0, infinity) [default: 1]\n"
    "  -i, --iterations  number of iterations to perform without improvement before stopping [default: 100]\n"
    "  -g, --greedy      use greedy hill climbing strategy as opposed to full NNI neighborhood exploration\n"
    "  -s, --seed        seed for random number generator [default: 0]\n"
    "      --device      CUDA device index [default: 0]\n"
    "      --devices     GPUs for the distance matrix, e.g. 0-7 or 0,2,3 (the search runs on the first)\n"
    "      --cache       read the breakpoint profiles from CSV.lzbp when it matches the CSV, write it otherwise\n";

Args parse_args(int argc, char** argv) {
    Args a;
    if (argc < 2) throw std::runtime_error("missing subcommand");
    a.command = argv[1];
    if (a.command != "distance" && a.command != "nni" && a.command != "nj" && a.command != "bp" && a.command != "writebench")
        throw std::runtime_error("unknown subcommand: " + a.command);
    bool have_positional = false;
    for (int i = 2; i < argc; i++) {
        const std::string s = argv[i];
        auto value = [&]() -> std::string {
            if (i + 1 >= argc) throw std::runtime_error(s + ": expected 1 argument(s). 0 provided.");
            return argv[++i];
        };
        auto number = [&](const std::string& v, bool integer) -> double {
            size_t used = 0;
            double x = 0;
            try {
                x = integer ? (double)std::stoi(v, &used) : std::stod(v, &used);
            } catch (const std::exception&) {
                used = 0;
            }
            if (used != v.size() || v.empty()) throw std::runtime_error("pattern '" + v + "' not found");
            return x;
        };
        if (s == "-o" || s == "--output") {
            a.output = value();
            a.have_output = true;
        } else if (s == "-t" || s == "--tree") {
            a.tree = value();
        } else if (s == "-a" || s == "--aggression") {
            a.aggression = number(value(), false);
        } else if (s == "-i" || s == "--iterations") {
            a.iterations = (int)number(value(), true);
        } else if (s == "-s" || s == "--seed") {
            a.seed = (int)number(value(), true);
        } else if (s == "--device") {
            a.device = (int)number(value(), true);
        } else if (s == "--devices") {  // "0-7", "0,2,3", "1"
            const std::string v = value();
            size_t pos = 0;
            while (pos <= v.size()) {
                const size_t comma = std::min(v.find(',', pos), v.size());
                const std::string part = v.substr(pos, comma - pos);
                const size_t dash = part.find('-');
                const int lo = (int)number(dash == std::string::npos ? part : part.substr(0, dash), true);
                const int hi = dash == std::string::npos ? lo : (int)number(part.substr(dash + 1), true);
                if (lo < 0 || hi < lo || hi - lo > 1023) throw std::runtime_error("pattern '" + v + "' not found");
                for (int d = lo; d <= hi; d++) a.devices.push_back(d);
                pos = comma + 1;
            }
            if (a.devices.empty()) throw std::runtime_error("--devices: expected at least one device");
            a.device = a.devices[0];
        } else if (s == "--cache") {
            a.cache = true;
        } else if (s == "-g" || s == "--greedy") {
            a.greedy = true;
        } else if (s == "-h" || s == "--help") {
            std::cout << USAGE;
            std::exit(0);
        } else if (!s.empty() && s[0] == '-' && s.size() > 1 && !(s[1] >= '0' && s[1] <= '9')) {
            throw std::runtime_error("Unknown argument: " + s);
        } else if (!have_positional) {
            a.cn_profile = s;
            have_positional = true;
        } else {
            throw std::runtime_error("Maximum number of positional arguments exceeded");
        }
    }
    if (!have_positional) throw std::runtime_error("cn_profile: 1 argument(s) expected. 0 provided.");
    if (!a.have_output) throw std::runtime_error("-o: required.");
    return a;
}

struct Gpu {
    lzg_ctx* ctx = nullptr;
    lzg_multi* multi = nullptr;  // --devices: one context per GPU; ctx = the first one's
    explicit Gpu(const std::vector<int>& devices) {
        std::vector<int32_t> d(devices.begin(), devices.end());
        int rc = lzg_multi_init(d.data(), (int32_t)d.size(), &multi);
        if (rc == LZG_OK) rc = lzg_multi_context(multi, 0, &ctx);
        if (rc != LZG_OK)
            throw std::runtime_error(std::string("cannot open CUDA devices: ") + lzg_status_string(rc) + ": " + lzg_last_error(nullptr));
    }
    explicit Gpu(int device) {
        const auto t0 = std::chrono::steady_clock::now();
        int rc = lzg_init(device, &ctx);
        if (getenv("LAZAC_TIMING"))
            std::cerr << "[lazac timing] lzg_init (CUDA context) " << std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count()
                      << " ms" << std::endl;
        if (rc != LZG_OK)  // no CPU fallback: the hot path exists on the GPU only
            throw std::runtime_error(std::string("cannot open CUDA device: ") + lzg_status_string(rc) + ": " + lzg_last_error(nullptr));
    }
    ~Gpu() {
        if (multi) lzg_multi_shutdown(multi);
        else if (ctx) lzg_shutdown(ctx);
    }
    void check(int rc) const {
        if (rc != LZG_OK) throw_lzg(ctx, rc);
    }
};

// build_distance_matrix (src/lazac.cxx:63-95) through the C ABI; returns the device-resident profiles too.
std::vector<int32_t> build_distance_matrix(const Gpu& gpu, const BreakpointSet& set, lzg_profiles** keep) {
    const int n = (int)set.names.size();
    lzg_profiles* p = nullptr;
    std::vector<int32_t> D((size_t)n * n);
    log_info("Built 0 out of " + std::to_string(n) + " rows of distance matrix.");
    if (gpu.multi && lzg_multi_size(gpu.multi) > 1) {
        // several GPUs: row blocks uploaded per GPU, tile rows dealt to them, the matrix gathered here
        const int rc = lzg_distance_matrix_multi(gpu.multi, set.bp.data(), n, set.L, D.data(), nullptr);
        if (rc != LZG_OK) throw std::runtime_error(std::string(lzg_status_string(rc)) + ": " + lzg_multi_last_error(gpu.multi));
        if (keep) gpu.check(lzg_profiles_upload(gpu.ctx, set.bp.data(), n, set.L, &p));  // the trees of the search live on the first GPU
    } else {
        gpu.check(lzg_profiles_upload(gpu.ctx, set.bp.data(), n, set.L, &p));
        gpu.check(lzg_distance_matrix(gpu.ctx, p, D.data()));
    }
    log_info("Finished building " + std::to_string(n) + " x " + std::to_string(n) + " distance matrix.");
    if (keep) *keep = p;
    else if (p) lzg_profiles_free(p);
    return D;
}

// PREFIX_dist_matrix.csv / .txt (src/lazac.cxx:107-131): byte-identical to the reference's value-by-value streaming, written
// as blocks of rows formatted straight into one preallocated buffer with a hand-rolled integer formatter (SURVEY.md 8(f)
// rank 3: cfg3's matrix is 4e8 numbers, written twice; std::to_string + string append per value took minutes there).
static inline char* put_int(char* p, long long v) {
    if (v < 0) {
        *p++ = '-';
        v = -v;
    }
    char digits[20];
    int nd = 0;
    do {
        digits[nd++] = (char)('0' + v % 10);
        v /= 10;
    } while (v);
    while (nd) *p++ = digits[--nd];
    return p;
}

static void write_matrix(const std::string& path, const std::vector<std::string>& names, const std::vector<int32_t>& D, char sep, bool csv_header) {
    const auto t0 = std::chrono::steady_clock::now();
    const size_t n = names.size();
    FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) throw std::runtime_error("Cannot open file " + path);
    std::string head;
    if (csv_header) {
        for (size_t i = 0; i < n; i++) {
            if (i != 0) head += ',';
            head += names[i];
        }
        head += '\n';
    } else {
        head = std::to_string(n) + "\n";
    }
    std::fwrite(head.data(), 1, head.size(), f);
    size_t longest = 0;
    for (const auto& nm : names) longest = std::max(longest, nm.size());
    const size_t row_cap = longest + 1 + n * 12;                       // name, then n x (separator + up to 11 characters)
    // Blocks of rows are formatted by a few threads at once, each into its own buffer, and written in order: the formatting
    // (a hand-rolled integer formatter, ~4 ns a value on one core) is the cost, not the write.
    const size_t rows_per_block = std::max<size_t>(1, ((size_t)8 << 20) / row_cap);
    const unsigned hw = std::thread::hardware_concurrency();
    const size_t n_thr = std::max<size_t>(1, std::min<size_t>({(size_t)(hw ? hw : 1), (size_t)16, (n + rows_per_block - 1) / std::max<size_t>(rows_per_block, 1)}));
    std::vector<std::vector<char>> bufs(n_thr);
    for (auto& b : bufs) b.resize(row_cap * std::min(rows_per_block, std::max<size_t>(n, 1)));
    std::vector<size_t> used(n_thr, 0);
    auto format_block = [&](size_t k, size_t i0) {
        char* p = bufs[k].data();
        for (size_t i = i0; i < std::min(n, i0 + rows_per_block); i++) {
            std::memcpy(p, names[i].data(), names[i].size());
            p += names[i].size();
            const int32_t* row = D.data() + i * n;
            for (size_t j = 0; j < n; j++) {
                *p++ = sep;
                p = put_int(p, row[j]);
            }
            *p++ = '\n';
        }
        used[k] = (size_t)(p - bufs[k].data());
    };
    for (size_t w0 = 0; w0 < n; w0 += n_thr * rows_per_block) {
        std::vector<std::thread> pool;
        size_t k_used = 0;
        for (size_t k = 0; k < n_thr && w0 + k * rows_per_block < n; k++, k_used++) {
            if (n_thr == 1) format_block(0, w0);
            else pool.emplace_back(format_block, k, w0 + k * rows_per_block);
        }
        for (auto& th : pool) th.join();
        for (size_t k = 0; k < k_used; k++) std::fwrite(bufs[k].data(), 1, used[k], f);
    }
    std::fclose(f);
    if (getenv("LAZAC_TIMING"))
        std::cerr << "[lazac timing] " << path << ": " << std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count() << " ms" << std::endl;
}

void write_matrix_txt(const std::string& path, const std::vector<std::string>& names, const std::vector<int32_t>& D) {
    write_matrix(path, names, D, ' ', false);
}

void write_matrix_csv(const std::string& path, const std::vector<std::string>& names, const std::vector<int32_t>& D) {
    write_matrix(path, names, D, ',', true);
}

// CSV -> breakpoint set; LAZAC_TIMING=1 logs how long the two steps took.
BreakpointSet ingest(const std::string& path, bool cache = false) {
    const auto t0 = std::chrono::steady_clock::now();
    if (cache) {
        BreakpointSet cached;
        if (load_breakpoint_cache(path + ".lzbp", path, cached)) {
            if (getenv("LAZAC_TIMING"))
                fprintf(stderr, "[lazac timing] profile cache %s.lzbp %.0f ms\n", path.c_str(), std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
            return cached;
        }
    }
    const std::map<std::string, CellProfile> cells = read_cn_profiles(path);
    const auto t1 = std::chrono::steady_clock::now();
    BreakpointSet set = to_breakpoint_set(cells);
    const auto t2 = std::chrono::steady_clock::now();
    if (getenv("LAZAC_TIMING"))
        fprintf(stderr, "[lazac timing] read_cn_profiles %.0f ms, to_breakpoint_set %.0f ms\n", std::chrono::duration<double, std::milli>(t1 - t0).count(),
                std::chrono::duration<double, std::milli>(t2 - t1).count());
    if (cache && !save_breakpoint_cache(path + ".lzbp", path, set)) fprintf(stderr, "lazac: could not write %s.lzbp\n", path.c_str());
    return set;
}

// The CUDA context of a fresh process takes seconds (2.4 s measured on a B200 box): open it on a
// second thread while this one parses the CSV.
std::future<std::unique_ptr<Gpu>> open_gpu_async(const Args& a) {
    const int device = a.device;
    const std::vector<int> devices = a.devices;
    return std::async(std::launch::async, [device, devices]() { return devices.empty() ? std::make_unique<Gpu>(device) : std::make_unique<Gpu>(devices); });
}

void do_distance(const Args& a) {
    auto gpu_opening = open_gpu_async(a);
    const BreakpointSet set = ingest(a.cn_profile, a.cache);
    const std::unique_ptr<Gpu> gpu_owner = gpu_opening.get();
    Gpu& gpu = *gpu_owner;
    const std::vector<int32_t> D = build_distance_matrix(gpu, set, nullptr);
    write_matrix_csv(a.output + "_dist_matrix.csv", set.names, D);
    write_matrix_txt(a.output + "_dist_matrix.txt", set.names, D);
}

std::string slurp(const std::string& path) {
    std::ifstream in(path);
    std::stringstream buffer;
    buffer << in.rdbuf();
    return buffer.str();
}

// PREFIX_dist_matrix.txt -> names + integer matrix (helper subcommand `nj`)
void read_matrix_txt(const std::string& path, std::vector<std::string>& names, std::vector<int32_t>& D) {
    std::ifstream in(path);
    if (!in) throw std::runtime_error("Cannot open file " + path);
    size_t n = 0;
    in >> n;
    names.resize(n);
    D.resize(n * n);
    for (size_t i = 0; i < n; i++) {
        in >> names[i];
        for (size_t j = 0; j < n; j++) in >> D[i * n + j];
    }
    if (!in) throw std::runtime_error("Failed to parse distance matrix.");
}

struct Candidate {
    GpuTree t;
};

std::string scores_text(const std::vector<int>& v) {  // pprint compact: [a, b, c]
    std::string s = "[";
    for (size_t i = 0; i < v.size(); i++) {
        if (i) s += ", ";
        s += std::to_string(v[i]);
    }
    return s + "]";
}

void do_nni(const Args& a) {
    const auto t_start = std::chrono::steady_clock::now();
    struct Total {  // LAZAC_TIMING: wall time of the whole command, printed when do_nni returns
        std::chrono::steady_clock::time_point t0;
        ~Total() {
            if (getenv("LAZAC_TIMING"))
                std::cerr << "[lazac timing] do_nni total " << std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count()
                          << " ms" << std::endl;
        }
    } total{t_start};
    /* Load copy number profiles */
    auto gpu_opening = open_gpu_async(a);
    const BreakpointSet set = ingest(a.cn_profile, a.cache);
    const int n_cells = (int)set.names.size();
    const std::unique_ptr<Gpu> gpu_owner = gpu_opening.get();
    Gpu& gpu = *gpu_owner;
    lzg_profiles* prof = nullptr;

    /* Load/initialize seed tree */
    NewickTree nt;
    if (a.tree.empty()) {
        log_info("No seed tree provided for NNI inference, building tree using neighbor joining.");
        const std::vector<int32_t> D = build_distance_matrix(gpu, set, &prof);
        write_matrix_txt(a.output + "_dist_matrix.txt", set.names, D);
        const std::string output_tree = a.output + "_nj_tree.newick";
        log_info("Building NJ tree from distance matrix using neighbor joining.");
        for (const std::string& nm : set.names) {
            // the reference's matrix parser takes a token that starts like a number for a distance
            // (src/dmat.cxx:163-291) and then fails: keep that a clear error instead of a wrong tree
            if (nm.empty() || (nm[0] >= '0' && nm[0] <= '9') || nm[0] == '.' || nm[0] == '+' || nm[0] == '-')
                throw std::runtime_error("Failed to parse distance matrix.");
        }
        // LAZAC_NJ=host: the Clearcut restatement on the CPU (what the parity tests compare the GPU with)
        const char* nj_env = getenv("LAZAC_NJ");
        const SeedTree nj = nj_env && std::string(nj_env) == "host" ? neighbor_joining_clearcut(D, set.names, (uint32_t)a.seed)
                                                                   : neighbor_joining_gpu(gpu.ctx, D, set.names, (uint32_t)a.seed, prof);
        log_info("Outputting NJ tree to file: " + output_tree);
        {
            std::ofstream out(output_tree, std::ios::out);
            out << nj.newick();
        }
        nt = read_newick(slurp(output_tree));
    } else {
        log_info("Reading seed tree from file: " + a.tree);
        nt = read_newick(slurp(a.tree));
        gpu.check(lzg_profiles_upload(gpu.ctx, set.bp.data(), n_cells, set.L, &prof));
    }

    /* Creates rectilinear vertex data */
    const int n = nt.size();
    std::map<std::string, int> row_of;
    for (int i = 0; i < n_cells; i++) row_of[set.names[i]] = i;
    Topology topo;
    topo.c0.assign(n, -1);
    topo.c1.assign(n, -1);
    std::vector<int32_t> leaf_row(n, -1);
    for (int u = 0; u < n; u++) {
        const auto& ch = nt.children[u];
        if (ch.empty()) {
            auto it = row_of.find(nt.name[u]);
            if (it == row_of.end()) throw std::runtime_error("leaf '" + nt.name[u] + "' of the seed tree has no copy number profile");
            leaf_row[u] = it->second;
        } else if (ch.size() == 2) {
            topo.set_children(u, ch[0], ch[1]);
        } else {
            throw std::logic_error("every child must have exactly two children");
        }
    }
    topo.rebuild_parents();

    lzg_tree* base = nullptr;
    gpu.check(lzg_tree_create(gpu.ctx, prof, n, topo.c0.data(), topo.c1.data(), leaf_row.data(), &base));

    std::ranlux48_base gen(a.seed);
    const bool timing = getenv("LAZAC_TIMING") != nullptr;
    using clk = std::chrono::steady_clock;
    auto ms_since = [](clk::time_point t0) { return std::chrono::duration<double, std::milli>(clk::now() - t0).count(); };
    double t_init = 0, t_score = 0, t_clone = 0, t_draws = 0, t_climb = 0;
    auto t_mark = clk::now();

    /* Candidate tree set is obtained by randomly perturbing candidate trees. */
    std::vector<Candidate> candidates;
    const float aggressions[] = {0, 0.25, 0.50, 0.75, 1, 1.25, 1.5, 1.75};
    for (float aggression : aggressions) {
        Candidate c;
        c.t.topo = stochastic_nni(topo, gen, aggression);
        gpu.check(lzg_tree_clone(gpu.ctx, base, &c.t.h));
        gpu.check(lzg_tree_set_topology(gpu.ctx, c.t.h, c.t.topo.c0.data(), c.t.topo.c1.data()));
        candidates.push_back(c);
    }
    lzg_tree_free(base);
    t_init = ms_since(t_mark);

    auto score_all = [&]() {
        const auto t0 = clk::now();
        for (auto& c : candidates) {
            int32_t s = 0;
            gpu.check(lzg_tree_score(gpu.ctx, c.t.h, &s));
            c.t.score = s;
        }
        std::sort(candidates.begin(), candidates.end(), [](const Candidate& x, const Candidate& y) { return x.t.score > y.t.score; });
        t_score += ms_since(t0);
    };

    std::string progress;  // nlohmann dump(): keys in alphabetical order, no spaces
    bool any_progress = false;
    ClimbStats stats;
    int counter = 0, iteration = 0;
    for (; counter < a.iterations; iteration++) {
        score_all();
        std::vector<int> scores;
        for (auto& c : candidates) scores.push_back(c.t.score);

        progress += any_progress ? "," : "[";
        any_progress = true;
        progress += "{\"iteration\":" + std::to_string(iteration) + ",\"scores\":[";
        for (size_t i = 0; i < scores.size(); i++) progress += (i ? "," : "") + std::to_string(scores[i]);
        progress += "]}";
        log_info("Candidate tree scores @ iteration " + std::to_string(iteration) + ": " + scores_text(scores));

        // Select and perturb candidate tree.
        std::uniform_int_distribution<int> distrib(0, (int)candidates.size() - 1);
        const int candidate_tree_idx = distrib(gen);

        GpuTree work;
        t_mark = clk::now();
        work.topo = candidates[candidate_tree_idx].t.topo;
        work.score = candidates[candidate_tree_idx].t.score;
        gpu.check(lzg_tree_clone(gpu.ctx, candidates[candidate_tree_idx].t.h, &work.h));
        t_clone += ms_since(t_mark);
        t_mark = clk::now();
        std::uniform_real_distribution<double> aggression_distrib(0, a.aggression);
        // the reference perturbs a temporary and drops it (src/lazac.cxx:284): only the RNG draws matter
        stochastic_nni_draws(work.topo, gen, (float)aggression_distrib(gen));
        t_draws += ms_since(t_mark);

        t_mark = clk::now();
        hill_climb(gpu.ctx, work, gen, a.greedy, &stats);
        t_climb += ms_since(t_mark);
        if (work.score < candidates[0].t.score) {
            lzg_tree_free(candidates[0].t.h);
            candidates[0].t = work;
            log_info("Updated candidate tree set.");
            counter = 0;
            continue;
        }
        lzg_tree_free(work.h);
        counter++;
    }
    score_all();
    if (timing)
        std::cerr << "[lazac timing] search: candidate set " << t_init << " ms, scoring the set " << t_score << " ms, clones " << t_clone
                  << " ms, perturbation draws " << t_draws << " ms, hill climbs " << t_climb << " ms (" << stats.passes << " passes)" << std::endl;
    GpuTree& best = candidates.back().t;

    /* ancestral_labeling (src/copy_number.cxx:187-225): labels on the GPU, pre-order renumbering here */
    const int L = set.L;
    std::vector<int32_t> labels((size_t)n * L), branch(n);
    t_mark = clk::now();
    gpu.check(lzg_tree_ancestral(gpu.ctx, best.h, labels.data(), branch.data()));
    if (timing) std::cerr << "[lazac timing] ancestral labels (GPU + read-back of " << (size_t)n * L * 4 / 1000000 << " MB) " << ms_since(t_mark) << " ms" << std::endl;
    std::vector<int> old_id, new_parent;
    {
        std::vector<std::pair<int, int>> st;  // (node, new id of its parent)
        st.emplace_back(0, -1);
        while (!st.empty()) {
            const auto [u, p] = st.back();
            st.pop_back();
            const int k = (int)old_id.size();
            old_id.push_back(u);
            new_parent.push_back(p);
            if (!best.topo.leaf(u)) {
                st.emplace_back(best.topo.c0[u], k);
                st.emplace_back(best.topo.c1[u], k);
            }
        }
    }
    std::vector<std::string> names(n);
    std::vector<std::vector<int>> children(n);
    std::vector<std::optional<int>> length(n);
    for (int k = 0; k < n; k++) {
        names[k] = nt.name[old_id[k]];
        if (new_parent[k] >= 0) {
            children[new_parent[k]].push_back(k);
            length[k] = branch[old_id[k]];
        }
    }
    for (auto& c : children) std::sort(c.begin(), c.end());

    {
        std::ofstream out(a.output + "_tree.newick", std::ios::out);
        out << write_newick(names, children, length) << "\n";
    }
    {
        std::ofstream out(a.output + "_info.json", std::ios::out);
        out << (any_progress ? progress + "]" : std::string("null")) << "\n";
    }
    {
        std::ofstream out(a.output + "_ancestral_cn_profile.csv", std::ios::out);
        out << "node,chrom,allele,start,end,cn\n";
        // every node is labelled over the same positions: one segment order, one ",chrom,allele,start,end," per output row
        const SegmentOrder order = make_segment_order(set.bins);
        std::vector<std::string> mid(L);
        for (int i = 0; i < L; i++) {
            const GenomicBin& b = set.bins[order.perm[i]];
            mid[i] = "," + b.chromosome + "," + b.allele + "," + std::to_string(b.start) + "," + std::to_string(b.end) + ",";
        }
        std::vector<int> bp(L), cn;
        size_t mid_bytes = 0;
        for (const auto& m : mid) mid_bytes += m.size();
        std::vector<char> block;
        const auto t_out = std::chrono::steady_clock::now();
        for (int k = 0; k < n; k++) {
            const int32_t* row = labels.data() + (size_t)old_id[k] * L;
            bp.assign(row, row + L);
            copynumber_transform(bp, set.bins, order, cn, nullptr);
            // one buffer per node, rows laid down with memcpy and a hand-rolled integer formatter
            // (snprintf per row was most of the 2.5 s this file took at cfg4's size)
            const std::string& name = names[k];
            block.resize(mid_bytes + (size_t)L * (name.size() + 13));
            char* p = block.data();
            for (int i = 0; i < L; i++) {
                memcpy(p, name.data(), name.size());
                p += name.size();
                memcpy(p, mid[i].data(), mid[i].size());
                p += mid[i].size();
                long long v = cn[i];
                if (v < 0) {
                    *p++ = '-';
                    v = -v;
                }
                char digits[12];
                int nd = 0;
                do {
                    digits[nd++] = (char)('0' + v % 10);
                    v /= 10;
                } while (v);
                while (nd) *p++ = digits[--nd];
                *p++ = '\n';
            }
            out.write(block.data(), (std::streamsize)(p - block.data()));
        }
        if (timing)
            std::cerr << "[lazac timing] ancestral CSV " << std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_out).count()
                      << " ms" << std::endl;
    }
    log_info("search: " + std::to_string(stats.passes) + " neighbourhood passes, " + std::to_string(stats.candidates) +
             " candidates scored, " + std::to_string(stats.moves) + " moves accepted");
    for (auto& c : candidates) lzg_tree_free(c.t.h);
    lzg_profiles_free(prof);
}

void do_nj(const Args& a) {  // helper: PREFIX_dist_matrix.txt -> Newick; on the CPU unless LAZAC_NJ=gpu
    std::vector<std::string> names;
    std::vector<int32_t> D;
    read_matrix_txt(a.cn_profile, names, D);
    std::ofstream out(a.output, std::ios::out);
    const char* nj_env = getenv("LAZAC_NJ");
    if (nj_env && std::string(nj_env) == "gpu") {
        Gpu gpu(0);
        out << neighbor_joining_gpu(gpu.ctx, D, names, (uint32_t)a.seed).newick();
    } else {
        out << neighbor_joining_clearcut(D, names, (uint32_t)a.seed).newick();
    }
}

void do_writebench(const Args& a) {  // helper: time the two matrix writers on an N x N matrix of distance-sized values, no GPU
    const size_t n = (size_t)std::stoul(a.cn_profile);
    std::vector<std::string> names(n);
    for (size_t i = 0; i < n; i++) {
        char nm[32];
        std::snprintf(nm, sizeof nm, "cell_%06zu", i);
        names[i] = nm;
    }
    std::vector<int32_t> D(n * n);
    uint32_t x = 12345;
    for (auto& v : D) {
        x = x * 1664525u + 1013904223u;
        v = (int32_t)(x >> 20);  // 0 .. 4095, the size of cfg3's distances
    }
    using clk = std::chrono::steady_clock;
    auto t0 = clk::now();
    write_matrix_csv(a.output + "_dist_matrix.csv", names, D);
    auto t1 = clk::now();
    write_matrix_txt(a.output + "_dist_matrix.txt", names, D);
    auto t2 = clk::now();
    std::printf("{\"n\": %zu, \"values\": %zu, \"csv_s\": %.3f, \"txt_s\": %.3f}\n", n, n * n, std::chrono::duration<double>(t1 - t0).count(),
                std::chrono::duration<double>(t2 - t1).count());
}

void do_bp(const Args& a) {  // helper: CSV -> names + breakpoint matrix (text), no GPU
    const BreakpointSet set = ingest(a.cn_profile, a.cache);
    std::ofstream out(a.output, std::ios::out);
    out << set.names.size() << " " << set.L << "\n";
    for (size_t i = 0; i < set.names.size(); i++) {
        out << set.names[i];
        for (int k = 0; k < set.L; k++) out << " " << set.bp[i * set.L + k];
        out << "\n";
    }
    for (const auto& b : set.bins) out << b.chromosome << " " << b.allele << " " << b.start << " " << b.end << "\n";
}

}  // namespace

int main(int argc, char** argv) {
    if (argc >= 2 && std::string(argv[1]) == "selftest") {  // host-only checks, no GPU (tests/test_host_cpu.py)
        int bad = 0;
        for (int leaves : {2, 3, 4, 5, 8, 33, 200, 1000})
            for (unsigned seed : {1u, 2u, 3u})
                for (float aggression : {0.0f, 0.3f, 1.0f, 1.75f})
                    if (!stochastic_nni_selftest(leaves, seed, aggression)) {
                        std::cerr << "stochastic_nni mismatch: leaves " << leaves << " seed " << seed << " aggression " << aggression << std::endl;
                        bad++;
                    }
        std::cout << (bad ? "selftest FAILED" : "selftest ok") << std::endl;
        return bad ? 3 : 0;
    }
    Args a;
    try {
        a = parse_args(argc, argv);
    } catch (const std::runtime_error& err) {
        std::cerr << err.what() << std::endl << USAGE;
        return 1;
    }
    try {
        if (a.command == "nni") do_nni(a);
        else if (a.command == "distance") do_distance(a);
        else if (a.command == "nj") do_nj(a);
        else if (a.command == "writebench") do_writebench(a);
        else do_bp(a);
    } catch (const std::exception& e) {
        // the reference lets these terminate the process (std::terminate); we report and exit non-zero
        std::cerr << "lazac: " << e.what() << std::endl;
        return 2;
    }
    return 0;
}
