// spydrpick_main.cpp -- C++ host of the B200-native SpydrPick hot path: the reference's command line,
// option semantics and output listings, with the four accelerated seams of src/SpydrPick.cpp bound to the
// C-ABI of include/spydrpick_b200.h (this file is exactly the "reference-side binding" of INTEGRATION.md,
// written without Boost / TBB so that it builds with g++ alone).
//
//   main pipeline               ref: src/SpydrPick.cpp:45-510
//   options                     ref: src/SpydrPick_options.cpp:144-159, src/ARACNE_options.cpp:139-147,
//                                    externals/apegrunt/src/Apegrunt_options.cpp:246-303
//   FASTA ingest, state coding  ref: apegrunt/Alignment_parsers.hpp:73-144, src/StateVector_state_types.cpp:33-44
//   column filter               ref: apegrunt/Alignment_filter.hpp:250-339
//   writers                     ref: graph/Graph_output_format.hpp:49-74, include/SpydrPick.hpp:89-129
//
// Parsing, filtering and the quartile / threshold bookkeeping stay on the host in the reference's style; sample
// reweighting (K1), threshold estimation (K2b), the all-pairs scan (K2/K2r), the edge sort and ARACNE (K3) run on
// the GPU. There is no CPU fallback: without an sm_100 device spk_create fails and the program exits like the
// reference does on fatal errors (message + EXIT_FAILURE).
#include <algorithm>
#include <chrono>
#include <cinttypes>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <sstream>
#include <string>
#include <sys/stat.h>
#include <unordered_map>
#include <vector>

#include "spydrpick_b200.h"

namespace {

struct Options {
    // SpydrPick_options
    bool verbose = false;
    double mi_threshold = -1.0;
    uint64_t mi_values = 0;
    double mi_pseudocount = 0.5;
    uint64_t mi_threshold_iterations = 10;
    uint64_t mi_threshold_pairs = 0;
    uint64_t ld_threshold = 0;
    bool no_aracne = false;
    int threads = -1;                           // accepted, unused (the GPU does the work)
    // ARACNE_options
    double aracne_edge_threshold = 2.220446049250313e-16;
    uint64_t aracne_block_size = 16384;         // accepted for compatibility; the result does not depend on it
    // Apegrunt_options
    std::vector<std::string> alignment_files;
    std::string mapping_list, sample_weights_file, outfile;
    int input_indexing_base = 1, output_indexing_base = 1;
    bool no_filter_alignment = false;
    bool host_filter = false;              // column filter on the host (default: on the device)
    double maf_threshold = 0.01, gap_threshold = 0.15;
    bool no_sample_reweighting = false, rescale_sample_weights = false, output_sample_weights = false;
    double sample_reweighting_threshold = 0.9;
    uint64_t genome_size = 0;
    bool linear_genome = false;
    // additions of this implementation
    int device = -1;
    uint64_t seed = 0;                          // 0 = seed from the clock like the reference (SpydrPick.hpp:178)
    int engine = 0;
    std::string output_dir = ".";
};

[[noreturn]] void die(const std::string& msg) {
    std::cerr << "SpydrPick error: " << msg << "\n\n";
    std::exit(EXIT_FAILURE);
}

void print_help() {
    std::cout <<
        "SpydrPick: MI-ARACNE genome-wide co-evolution analysis (B200-native hot path).\n"
        "Usage: SpydrPick [options] <alignmentfile>\n\n"
        "SpydrPick options:\n"
        "  -h [ --help ]                      Print this help message.\n"
        "  --version                          Print version information.\n"
        "  -v [ --verbose ]                   Be verbose.\n"
        "  --mi-threshold arg (=-1)           The MI threshold value (0=no threshold; -1=determine automatically).\n"
        "  --mi-values arg (=0)               Approximate number of MI values to calculate from data.\n"
        "  --mi-pseudocount arg (=0.5)        The MI pseudocount value.\n"
        "  --mi-threshold-iterations arg (=10) Number of iterations for estimating saving threshold.\n"
        "  --mi-threshold-pairs arg (=0)      Number of sampled pairs for estimating saving threshold (0=auto).\n"
        "  --ld-threshold arg (=0)            Threshold distance for linkage disequilibrium (LD).\n"
        "  --no-aracne                        Skip ARACNE, only calculate MI.\n"
        "  -t [ --threads ] arg (=-1)         Accepted for compatibility (the hot path runs on the GPU).\n"
        "ARACNE options:\n"
        "  --aracne-edge-threshold arg        Equality tolerance of the ARACNE edge comparison (default DBL_EPSILON).\n"
        "  --aracne-block-size arg (=16384)   Accepted for compatibility (the result does not depend on it).\n"
        "Alignment options:\n"
        "  --alignmentfile arg                The input alignment filename (FASTA).\n"
        "  --mapping-list arg                 Name of file containing loci mappings.\n"
        "  --sample-weights arg               Name of file containing sample weights.\n"
        "  --input-indexing-base arg (=1)     Base index for input.\n"
        "  --output-indexing-base arg (=1)    Base index for output.\n"
        "  --no-filter-alignment              Do not filter alignment columns by applying MAF and GAP thresholds.\n"
        "  --host-filter                      Count state frequencies and extract the kept columns on the host (default: device).\n"
        "  --maf-threshold arg (=0.01)        Minor state frequency threshold.\n"
        "  --gap-threshold arg (=0.15)        Gap frequency threshold.\n"
        "  --no-sample-reweighting            Turn sample reweighting off.\n"
        "  --sample-reweighting-threshold arg (=0.9)\n"
        "  --rescale-sample-weights           Rescale sample weights so that n(effective) == n.\n"
        "  --output-sample-weights            Output sample weights.\n"
        "  --genome-size arg (=0)             Genome size, if different from input.\n"
        "  --linear-genome                    Assume linear genome in distance calculations.\n"
        "B200 options:\n"
        "  --device arg (=-1)                 CUDA device (-1 = current).\n"
        "  --seed arg (=0)                    Seed of the threshold estimator (0 = clock, like the reference).\n"
        "  --engine arg (=0)                  0/3 int8 compact tcgen05 filter, 2 dense bf16 filter, 1 fp64 exhaustive.\n"
        "  --output-dir arg (=.)              Directory for the output files.\n";
}

Options parse_options(int argc, char** argv) {
    Options o;
    auto need = [&](int& i, const std::string& name, const std::string& inl, bool has_inl) -> std::string {
        if (has_inl) return inl;
        if (i + 1 >= argc) die("the required argument for option '--" + name + "' is missing");
        return argv[++i];
    };
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        if (a.rfind("--", 0) != 0 && !(a.size() == 2 && a[0] == '-')) { o.alignment_files.push_back(a); continue; }
        std::string name, inl; bool has_inl = false;
        if (a.rfind("--", 0) == 0) {
            name = a.substr(2);
            const size_t eq = name.find('=');
            if (eq != std::string::npos) { inl = name.substr(eq + 1); name = name.substr(0, eq); has_inl = true; }
        } else {
            switch (a[1]) { case 'h': name = "help"; break; case 'v': name = "verbose"; break; case 't': name = "threads"; break;
                            case 'o': name = "outfile"; break; default: die("unrecognised option '" + a + "'"); }
        }
        auto U = [&](const std::string& s) { return (uint64_t)std::strtoull(s.c_str(), nullptr, 10); };
        auto D = [&](const std::string& s) { return std::strtod(s.c_str(), nullptr); };
        if (name == "help") { print_help(); std::exit(EXIT_SUCCESS); }
        else if (name == "version") { std::cout << "SpydrPick (" << spk_version() << ")\n"; std::exit(EXIT_SUCCESS); }
        else if (name == "verbose") o.verbose = true;
        else if (name == "mi-threshold") o.mi_threshold = D(need(i, name, inl, has_inl));
        else if (name == "mi-values") o.mi_values = U(need(i, name, inl, has_inl));
        else if (name == "mi-pseudocount") o.mi_pseudocount = D(need(i, name, inl, has_inl));
        else if (name == "mi-threshold-iterations") o.mi_threshold_iterations = U(need(i, name, inl, has_inl));
        else if (name == "mi-threshold-pairs") o.mi_threshold_pairs = U(need(i, name, inl, has_inl));
        else if (name == "ld-threshold") o.ld_threshold = U(need(i, name, inl, has_inl));
        else if (name == "no-aracne") o.no_aracne = true;
        else if (name == "threads") o.threads = std::atoi(need(i, name, inl, has_inl).c_str());
        else if (name == "aracne-edge-threshold") o.aracne_edge_threshold = D(need(i, name, inl, has_inl));
        else if (name == "aracne-block-size") o.aracne_block_size = U(need(i, name, inl, has_inl));
        else if (name == "alignmentfile") o.alignment_files.push_back(need(i, name, inl, has_inl));
        else if (name == "mapping-list") o.mapping_list = need(i, name, inl, has_inl);
        else if (name == "sample-weights") o.sample_weights_file = need(i, name, inl, has_inl);
        else if (name == "outfile") o.outfile = need(i, name, inl, has_inl);
        else if (name == "input-indexing-base") o.input_indexing_base = std::atoi(need(i, name, inl, has_inl).c_str());
        else if (name == "output-indexing-base") o.output_indexing_base = std::atoi(need(i, name, inl, has_inl).c_str());
        else if (name == "no-filter-alignment") o.no_filter_alignment = true;
        else if (name == "host-filter") o.host_filter = true;
        else if (name == "maf-threshold") o.maf_threshold = D(need(i, name, inl, has_inl));
        else if (name == "gap-threshold") o.gap_threshold = D(need(i, name, inl, has_inl));
        else if (name == "no-sample-reweighting") o.no_sample_reweighting = true;
        else if (name == "sample-reweighting-threshold") o.sample_reweighting_threshold = D(need(i, name, inl, has_inl));
        else if (name == "rescale-sample-weights") o.rescale_sample_weights = true;
        else if (name == "output-sample-weights") o.output_sample_weights = true;
        else if (name == "genome-size") o.genome_size = U(need(i, name, inl, has_inl));
        else if (name == "linear-genome") o.linear_genome = true;
        else if (name == "device") o.device = std::atoi(need(i, name, inl, has_inl).c_str());
        else if (name == "seed") o.seed = U(need(i, name, inl, has_inl));
        else if (name == "engine") o.engine = std::atoi(need(i, name, inl, has_inl).c_str());
        else if (name == "output-dir") o.output_dir = need(i, name, inl, has_inl);
        else die("unrecognised option '--" + name + "'");
    }
    if (o.alignment_files.empty()) { std::cerr << "SpydrPick error: No alignment file specified.\n\n"; print_help(); std::exit(EXIT_FAILURE); }
    if (o.alignment_files.size() > 1) die("exactly one alignment file is supported (inter-alignment mode is out of scope)");
    return o;
}

// ---- alignment ----------------------------------------------------------------------------------------------
struct HostAlignment {
    std::string id;                       // file stem (+ ".filter id" after filtering)
    uint32_t n = 0; uint64_t L = 0;       // samples, loci
    std::vector<std::string> names;
    uint8_t* states = nullptr;            // [n][L] codes a=0,t=1,c=2,g=3,GAP=4 (pinned host memory)
    std::vector<uint64_t> pos;            // loci translation: original position of every column (0-based)
    uint64_t n_original_positions = 0;
};

inline uint8_t char_to_state(char ch) {                     // StateVector_state_types.cpp:33-44
    switch (ch | 0x20) { case 'a': return 0; case 't': return 1; case 'c': return 2; case 'g': return 3; default: return 4; }
}

std::string file_stem(const std::string& path) {
    const size_t sl = path.find_last_of('/');
    std::string f = sl == std::string::npos ? path : path.substr(sl + 1);
    const size_t dot = f.find_last_of('.');
    return dot == std::string::npos || dot == 0 ? f : f.substr(0, dot);
}

HostAlignment read_fasta(const std::string& path) {
    std::ifstream in(path, std::ios::binary | std::ios::ate);
    if (!in) die("could not open alignment file \"" + path + "\"");
    const std::streamsize size = in.tellg();
    in.seekg(0);
    std::string buf((size_t)size, '\0');
    if (size > 0 && !in.read(&buf[0], size)) die("could not read alignment file \"" + path + "\"");
    if (buf.size() >= 2 && (unsigned char)buf[0] == 0x1f && (unsigned char)buf[1] == 0x8b) die("compressed FASTA is not supported: decompress \"" + path + "\" first");
    HostAlignment A; A.id = file_stem(path);
    // pass 1: record boundaries and the sequence length of every record
    struct Rec { size_t seq_begin, seq_end; };
    std::vector<Rec> recs;
    size_t p = 0;
    while (p < buf.size()) {
        if (buf[p] != '>') { ++p; continue; }
        size_t eol = buf.find('\n', p); if (eol == std::string::npos) eol = buf.size();
        std::string name = buf.substr(p + 1, eol - p - 1);
        while (!name.empty() && (name.back() == '\r' || name.back() == ' ')) name.pop_back();
        A.names.push_back(name);
        size_t nxt = buf.find("\n>", eol == buf.size() ? eol : eol);
        const size_t end = nxt == std::string::npos ? buf.size() : nxt + 1;
        recs.push_back({std::min(eol + 1, buf.size()), end});
        p = end;
    }
    if (recs.empty()) die("no FASTA records in \"" + path + "\"");
    auto seq_len = [&](const Rec& r) { size_t k = 0; for (size_t q = r.seq_begin; q < r.seq_end; ++q) { const char ch = buf[q]; if (ch != '\n' && ch != '\r' && ch != ' ' && ch != '\t') ++k; } return k; };
    A.n = (uint32_t)recs.size(); A.L = seq_len(recs[0]);
    if (A.L < 2) die("alignment has fewer than 2 columns");
    void* mem = nullptr;
    if (spk_host_alloc((size_t)A.n * A.L, &mem) != SPK_OK) die("could not allocate pinned host memory for the alignment");
    A.states = (uint8_t*)mem;
    for (uint32_t s = 0; s < A.n; ++s) {
        uint8_t* row = A.states + (size_t)s * A.L; size_t k = 0;
        for (size_t q = recs[s].seq_begin; q < recs[s].seq_end; ++q) {
            const char ch = buf[q];
            if (ch == '\n' || ch == '\r' || ch == ' ' || ch == '\t') continue;
            if (k < A.L) row[k] = char_to_state(ch);
            ++k;
        }
        if (k != A.L) die("sequence \"" + A.names[s] + "\" has " + std::to_string(k) + " positions, expected " + std::to_string(A.L));
    }
    A.pos.resize(A.L); for (uint64_t c = 0; c < A.L; ++c) A.pos[c] = c;
    A.n_original_positions = A.L;
    return A;
}

std::vector<uint64_t> read_index_list(const std::string& path, int base) {
    std::ifstream in(path);
    if (!in) die("could not open list file \"" + path + "\"");
    std::vector<uint64_t> v; std::string tok; char ch;
    auto flush = [&]() { if (!tok.empty()) { const long long x = std::atoll(tok.c_str()) - base; if (x < 0) die("index below the indexing base in \"" + path + "\""); v.push_back((uint64_t)x); tok.clear(); } };
    while (in.get(ch)) { if (ch >= '0' && ch <= '9') tok.push_back(ch); else flush(); }
    flush();
    return v;
}

// ref Alignment_filter.hpp:250-339: unweighted state frequencies; keep iff #(non-gap states with freq >= maf) > 1
// and gap frequency <= gap threshold. Returns the filter id string appended to the alignment id.
std::string filter_alignment(HostAlignment& A, const Options& o, bool verbose) {
    const uint32_t n = A.n; const uint64_t L = A.L;
    std::vector<uint32_t> cnt(L * 5, 0);
    for (uint32_t s = 0; s < n; ++s) { const uint8_t* row = A.states + (size_t)s * L; for (uint64_t c = 0; c < L; ++c) cnt[c * 5 + row[c]]++; }
    std::vector<uint64_t> keep;
    for (uint64_t c = 0; c < L; ++c) {
        int nsig = 0;
        for (int a = 0; a < 4; ++a) if (double(cnt[c * 5 + a]) / double(n) >= o.maf_threshold) ++nsig;
        if (nsig > 1 && double(cnt[c * 5 + 4]) / double(n) <= o.gap_threshold) keep.push_back(c);
    }
    if (keep.size() < 2) die("fewer than 2 columns pass the MAF / gap filter");
    if (verbose) std::cout << "SpydrPick: column filter keeps " << keep.size() << " of " << L << " loci\n";
    void* mem = nullptr;
    if (spk_host_alloc((size_t)n * keep.size(), &mem) != SPK_OK) die("could not allocate pinned host memory for the filtered alignment");
    uint8_t* f = (uint8_t*)mem;
    for (uint32_t s = 0; s < n; ++s) { const uint8_t* row = A.states + (size_t)s * L; uint8_t* out = f + (size_t)s * keep.size(); for (size_t k = 0; k < keep.size(); ++k) out[k] = row[keep[k]]; }
    std::vector<uint64_t> npos(keep.size()); for (size_t k = 0; k < keep.size(); ++k) npos[k] = A.pos[keep[k]];
    spk_host_free(A.states); A.states = f; A.L = keep.size(); A.pos.swap(npos);
    char id[128];
    std::snprintf(id, sizeof id, "filtered_ge%03zumaf_le%03zugf_gt01-lt05states", (size_t)(o.maf_threshold * 100.), (size_t)(o.gap_threshold * 100.));
    return id;
}

} // namespace

// ---- writers ------------------------------------------------------------------------------------------------
#include <charconv>
#include <thread>

namespace {

inline uint64_t genome_dist(uint64_t i1, uint64_t i2, bool linear, uint64_t circle) {        // apegrunt/Distance.hpp:28-37
    const uint64_t d = i1 > i2 ? i1 - i2 : i2 - i1;
    if (linear) return d;
    const uint64_t alt = circle - d;                                                          // size_t wrap like the reference
    return alt < d ? alt : d;
}
inline void put_u64(std::string& s, uint64_t v) { char t[24]; auto r = std::to_chars(t, t + sizeof t, v); s.append(t, r.ptr); }
inline void put_fixed(std::string& s, double v, int prec) {                                   // == printf("%.*f")
    char t[64]; auto r = std::to_chars(t, t + sizeof t, v, std::chars_format::fixed, prec);
    if (r.ec == std::errc()) s.append(t, r.ptr); else { std::snprintf(t, sizeof t, "%.*f", prec, v); s += t; }
}

// get_unique_ofstream (apegrunt/Apegrunt_IO_misc.hpp:63-78): append .1, .2, ... if the name exists
std::string unique_name(const std::string& name) {
    struct stat st;
    if (stat(name.c_str(), &st) != 0) return name;
    for (int k = 1;; ++k) { const std::string cand = name + "." + std::to_string(k); if (stat(cand.c_str(), &st) != 0) return cand; }
}

template <class F>
void write_parallel(const std::string& path, uint64_t n_items, F format_range) {
    const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    const uint64_t chunk = 1u << 18;
    std::ofstream out(path, std::ios::binary);
    if (!out) die("could not open output file \"" + path + "\"");
    for (uint64_t base = 0; base < n_items; base += (uint64_t)hw * chunk) {
        const unsigned nt = (unsigned)std::min<uint64_t>(hw, (n_items - base + chunk - 1) / chunk);
        std::vector<std::string> parts(nt);
        std::vector<std::thread> th;
        for (unsigned t = 0; t < nt; ++t)
            th.emplace_back([&, t]() { const uint64_t b = base + (uint64_t)t * chunk; format_range(b, std::min(n_items, b + chunk), parts[t]); });
        for (auto& x : th) x.join();
        for (auto& p : parts) out.write(p.data(), (std::streamsize)p.size());
    }
}

// Graph_output_format.hpp:49-74: "pos1 pos2 distance aracne_flag mi" (6 decimals)
void write_edges(const std::string& path, const spk_edge* e, const uint8_t* flags, uint64_t ne, const HostAlignment& A, const Options& o) {
    write_parallel(path, ne, [&](uint64_t b, uint64_t en, std::string& s) {
        s.reserve((en - b) * 40);
        for (uint64_t k = b; k < en; ++k) {
            const uint32_t n1 = std::min(e[k].i, e[k].j), n2 = std::max(e[k].i, e[k].j);
            const uint64_t i1 = A.pos[n1] + o.output_indexing_base, i2 = A.pos[n2] + o.output_indexing_base;
            put_u64(s, i1); s += ' '; put_u64(s, i2); s += ' '; put_u64(s, genome_dist(i1, i2, o.linear_genome, A.n_original_positions)); s += ' ';
            s += (flags && flags[k]) ? '1' : '0'; s += ' ';
            put_fixed(s, e[k].mi, 6); s += '\n';
        }
    });
}

// SpydrPick.hpp:89-129: outliers above the Tukey fence beyond the LD threshold, with the gap-free MI, the gap
// effect in percent and the extreme-outlier flag
uint64_t write_outliers(const std::string& path, const spk_edge* e, const uint8_t* flags, uint64_t ne, const spk_edge* wo, uint64_t nwo,
                        double outlier_thr, double extreme_thr, const HostAlignment& A, const Options& o) {
    auto eid = [](const spk_edge& x) { return ((uint64_t)std::max(x.i, x.j) << 32) | std::min(x.i, x.j); };
    std::vector<spk_edge> w(wo, wo + nwo);
    std::sort(w.begin(), w.end(), [&](const spk_edge& a, const spk_edge& b) { return eid(a) < eid(b); });
    std::string s; uint64_t lines = 0;
    for (uint64_t k = 0; k < ne; ++k) {
        if (e[k].mi < outlier_thr) break;
        auto it = std::lower_bound(w.begin(), w.end(), e[k], [&](const spk_edge& a, const spk_edge& b) { return eid(a) < eid(b); });
        const double w2 = (it != w.end() && eid(*it) == eid(e[k])) ? it->mi : e[k].mi;
        const uint32_t n1 = std::min(e[k].i, e[k].j), n2 = std::max(e[k].i, e[k].j);
        const uint64_t i1 = A.pos[n1] + o.output_indexing_base, i2 = A.pos[n2] + o.output_indexing_base;
        const uint64_t d = genome_dist(i1, i2, o.linear_genome, A.n_original_positions);
        if (d > o.ld_threshold) {
            put_u64(s, i1); s += ' '; put_u64(s, i2); s += ' '; put_u64(s, d); s += ' ';
            s += (flags && flags[k]) ? '1' : '0'; s += ' ';
            put_fixed(s, e[k].mi, 6); s += ' '; put_fixed(s, w2, 6); s += ' ';
            put_fixed(s, (1.0 - (w2 / e[k].mi)) * 100, 1); s += ' ';
            s += (e[k].mi > extreme_thr) ? '1' : '0'; s += '\n';
            ++lines;
        }
    }
    std::ofstream out(path, std::ios::binary);
    if (!out) die("could not open output file \"" + path + "\"");
    out.write(s.data(), (std::streamsize)s.size());
    return lines;
}

struct Timer {
    std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
    double lap() { auto t = std::chrono::steady_clock::now(); const double s = std::chrono::duration<double>(t - t0).count(); t0 = t; return s; }
};

void check(spk_ctx* ctx, int rc, const char* what) {
    if (rc != SPK_OK) die(std::string(what) + ": " + (spk_last_error(ctx) ? spk_last_error(ctx) : "unknown error"));
}

} // namespace

// ---- main: src/SpydrPick.cpp:45-510 -----------------------------------------------------------------------------
int main(int argc, char** argv) {
    Options o = parse_options(argc, argv);
    Timer total, lap;
    if (o.verbose) std::cout << "SpydrPick (" << spk_version() << ")\n";

    // -- ingest (src/SpydrPick.cpp:163; apegrunt/Alignment_utility.hpp:69-160)
    HostAlignment A = read_fasta(o.alignment_files[0]);
    if (o.verbose) std::cout << "SpydrPick: read " << A.n << " sequences x " << A.L << " loci from \"" << o.alignment_files[0] << "\" (" << lap.lap() << " s)\n";
    if (!o.mapping_list.empty()) {
        std::vector<uint64_t> m = read_index_list(o.mapping_list, o.input_indexing_base);
        if (m.size() != A.L) die("mapping list and alignment sizes are mismatched (" + std::to_string(m.size()) + " and " + std::to_string(A.L) + ")");
        A.pos.swap(m);                     // the mapping list only replaces the loci translation: n_original_positions stays the
    }                                      // FASTA column count unless --genome-size is given (Alignment_utility.hpp:128-160)
    if (o.genome_size) A.n_original_positions = o.genome_size;

    spk_ctx* ctx = nullptr;
    if (spk_create(&ctx, o.device, o.verbose ? 1 : 0) != SPK_OK) die(std::string("cannot initialise the CUDA context: ") + spk_last_error(nullptr));

    // -- column filter (src/SpydrPick.cpp:234-297) + alignment upload: the raw matrix goes up once, state frequencies are
    //    counted and the kept columns extracted on the device (spk_set_alignment_filtered); --host-filter keeps the host pass
    if (!o.no_filter_alignment && !o.host_filter) {
        if (A.L >= (1ull << 32)) die("too many loci");
        std::vector<uint8_t> keep(A.L); uint64_t n_kept = 0;
        check(ctx, spk_set_alignment_filtered(ctx, A.states, A.n, A.L, A.pos.data(), o.maf_threshold, o.gap_threshold, keep.data(), &n_kept),
              "column filter + alignment upload");
        if (o.verbose) std::cout << "SpydrPick: column filter keeps " << n_kept << " of " << A.L << " loci\n";
        std::vector<uint64_t> npos; npos.reserve(n_kept);
        for (uint64_t c = 0; c < A.L; ++c) if (keep[c]) npos.push_back(A.pos[c]);
        A.L = n_kept; A.pos.swap(npos);
        char id[128];
        std::snprintf(id, sizeof id, "filtered_ge%03zumaf_le%03zugf_gt01-lt05states", (size_t)(o.maf_threshold * 100.), (size_t)(o.gap_threshold * 100.));
        A.id += std::string(".") + id;
        if (o.verbose) std::cout << "SpydrPick: filtered alignment has " << A.L << " loci (" << lap.lap() << " s)\n";
        if (A.L >= (1ull << 29)) die("too many loci");
    } else {
        if (!o.no_filter_alignment) { A.id += "." + filter_alignment(A, o, o.verbose); if (o.verbose) std::cout << "SpydrPick: filtered alignment has " << A.L << " loci (" << lap.lap() << " s)\n"; }
        if (A.L >= (1ull << 29)) die("too many loci");
        check(ctx, spk_set_alignment(ctx, A.states, A.n, (uint32_t)A.L, A.pos.data()), "alignment upload");
    }
    const uint32_t n = A.n, L = (uint32_t)A.L;

    // -- sample weights (src/SpydrPick.cpp:321; apegrunt/Alignment_utility.hpp:186-258)
    std::vector<double> w(n, 1.0);
    if (!o.no_sample_reweighting) {
        if (!o.sample_weights_file.empty()) {
            std::ifstream in(o.sample_weights_file); if (!in) die("could not open sample weights file \"" + o.sample_weights_file + "\"");
            size_t k = 0; double x; while (k < n && (in >> x)) w[k++] = x;
            if (k != n) die("sample weights: expected " + std::to_string(n) + " values, got " + std::to_string(k));
        } else {
            std::vector<uint64_t> counts(n);
            const uint64_t min_identical = (uint64_t)(o.sample_reweighting_threshold * double(L));          // weights.hpp:228,251
            check(ctx, spk_neighbour_counts(ctx, min_identical, SPK_QUIRKS_REFERENCE, counts.data()), "sample reweighting");
            for (uint32_t s = 0; s < n; ++s) w[s] = 1.0 / double(counts[s]);                                   // weights.hpp:349
        }
        if (o.rescale_sample_weights) { double n_eff = 0; for (double x : w) n_eff += x; for (double& x : w) x = x * double(n) / n_eff; }   // (w * n) / n_eff, Alignment_utility.hpp:215,245
        check(ctx, spk_set_weights(ctx, w.data()), "sample weights");
    }
    double neff = 0; for (double x : w) neff += x;
    if (o.verbose) std::cout << "SpydrPick: effective sample size " << neff << " (" << lap.lap() << " s)\n";
    const std::string size_string = "L" + std::to_string(L) + "n" + std::to_string(n);                        // Alignment_utility.hpp:66
    const std::string prefix = o.output_dir + "/" + A.id + "." + size_string;
    if (o.output_sample_weights) {
        std::ofstream out(unique_name(prefix + ".weights"));
        for (double x : w) { std::string s; put_fixed(s, x, 6); out << s << "\n"; }
    }

    // -- saving threshold (src/SpydrPick.cpp:336-356; SpydrPick.hpp:285-343)
    double mi_threshold = o.mi_threshold;
    if (mi_threshold < 0.0) {
        const uint64_t top_pairs = o.mi_values ? o.mi_values : std::min<uint64_t>(10000000ull, 100ull * L);
        const uint64_t seed = o.seed ? o.seed : (uint64_t)std::chrono::system_clock::now().time_since_epoch().count();
        check(ctx, spk_determine_mi_threshold(ctx, top_pairs, o.mi_threshold_pairs, (uint32_t)o.mi_threshold_iterations, o.mi_pseudocount, seed, &mi_threshold), "MI threshold estimation");
        if (o.verbose) std::cout << "SpydrPick: MI saving threshold = " << mi_threshold << " (" << lap.lap() << " s)\n";
    }

    // -- all-pairs scan + sort (src/SpydrPick.cpp:384-398)
    spk_scan_params sp; std::memset(&sp, 0, sizeof sp);
    sp.threshold = mi_threshold; sp.pseudocount = o.mi_pseudocount; sp.ld_threshold = o.ld_threshold; sp.linear_genome = o.linear_genome ? 1 : 0;
    sp.circle_size = 0; sp.max_edges = o.mi_values ? 3 * o.mi_values + 65536 : 0; sp.sort_desc = 1; sp.engine = o.engine; sp.n_parts = 1;
    sp.keep_on_device = 2;       // the host copies of the edge lists run on a side stream while ARACNE works on the device copy
    spk_scan_result res;
    check(ctx, spk_mi_scan(ctx, &sp, &res), "MI scan");
    double outlier_thr = 0, extreme_thr = 0;
    check(ctx, spk_outlier_thresholds(res.colmax, L, &outlier_thr, &extreme_thr), "outlier thresholds");
    if (o.verbose) std::cout << "SpydrPick: " << res.n_edges << " edges stored (" << res.n_candidates << " candidates refined in fp64, filter error "
                             << res.max_filter_error << "); outlier threshold = " << outlier_thr << ", extreme outlier threshold = " << extreme_thr
                             << " (" << lap.lap() << " s)\n";

    // -- ARACNE (src/SpydrPick.cpp:406-421)
    void* flag_mem = nullptr;
    if (spk_host_alloc((size_t)std::max<uint64_t>(res.n_edges, 1), &flag_mem) != SPK_OK) die("could not allocate pinned host memory for the ARACNE flags");
    uint8_t* flags = (uint8_t*)flag_mem;
    std::memset(flags, 0, (size_t)res.n_edges);
    if (!o.no_aracne && res.n_edges) {
        check(ctx, spk_aracne(ctx, res.edges, res.n_edges, o.aracne_edge_threshold, 0, 1, flags), "ARACNE");
        if (o.verbose) { uint64_t nd = 0; for (uint64_t e = 0; e < res.n_edges; ++e) nd += flags[e]; std::cout << "SpydrPick: ARACNE keeps " << nd << " direct edges (" << lap.lap() << " s)\n"; }
    }
    check(ctx, spk_wait_download(ctx), "edge download");          // res.edges / res.edges_wo_gaps are complete from here on

    // -- outputs (src/SpydrPick.cpp:424-480)
    const std::string couplings = prefix + ".spydrpick_couplings." + std::to_string(o.output_indexing_base) + "-based";
    const std::string edges_file = unique_name(couplings + "." + std::to_string(res.n_edges) + "edges");
    write_edges(edges_file, res.edges, flags, res.n_edges, A, o);
    const std::string outliers_file = unique_name(couplings + ".outliers");
    const uint64_t n_out = write_outliers(outliers_file, res.edges, flags, res.n_edges, res.edges_wo_gaps, res.n_wo_gaps, outlier_thr, extreme_thr, A, o);
    if (o.verbose) std::cout << "SpydrPick: wrote \"" << edges_file << "\" and \"" << outliers_file << "\" (" << n_out << " outliers; " << lap.lap() << " s)\n";
    if (o.verbose) {
        spk_timings tm; if (spk_get_timings(ctx, &tm) == SPK_OK)
            std::cout << "SpydrPick: device ms: reweighting " << tm.hamming_ms << ", filter GEMM " << tm.gemm_ms << ", fp64 refine " << tm.refine_ms
                      << ", sort " << tm.sort_ms << ", ARACNE " << tm.aracne_build_ms + tm.aracne_ms << "\n";
        std::cout << "SpydrPick: total " << total.lap() << " s\n";
    }
    spk_destroy(ctx);
    spk_host_free(A.states);
    spk_host_free(flag_mem);
    return EXIT_SUCCESS;
}
