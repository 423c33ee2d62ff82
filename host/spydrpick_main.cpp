// spydrpick_main.cpp -- C++ host of the B200-native SpydrPick hot path: the reference's command line,
// option semantics and output listings, with the accelerated seams of src/SpydrPick.cpp bound to the
// C-ABI of include/spydrpick_b200.h (this file is exactly the "reference-side binding" of INTEGRATION.md,
// written without Boost / TBB so that it builds with g++ alone).
//
//   main pipeline               ref: src/SpydrPick.cpp:45-510
//   options                     ref: src/SpydrPick_options.cpp:144-159, src/ARACNE_options.cpp:139-156,
//                                    externals/apegrunt/src/Apegrunt_options.cpp:246-304
//   FASTA ingest, state coding  ref: apegrunt/Alignment_parsers.hpp:49-144, src/StateVector_state_types.cpp:33-44
//   loci lists (a, a-b)         ref: apegrunt/Loci_parser_grammar.hpp:65-84, apegrunt/Loci_parsers.hpp:60-85
//   include / exclude / sample  ref: src/SpydrPick.cpp:191-231, 299-318; apegrunt/Alignment_impl_block_compressed_storage.hpp:502-530
//   column filter               ref: apegrunt/Alignment_filter.hpp:250-339
//   alignment / loci export     ref: apegrunt/Alignment_utility.hpp:295-323, apegrunt/Alignment_generators.hpp:45-58, apegrunt/Loci_generators.hpp:31-44
//   writers                     ref: graph/Graph_output_format.hpp:49-74, include/SpydrPick.hpp:89-129
//   outlier-node export         ref: src/SpydrPick.cpp:483-510, graph/Graph_utility.hpp:36-55
//
// Parsing, filtering decisions and the quartile / threshold bookkeeping stay on the host in the reference's style; sample
// reweighting (K1), threshold estimation (K2b), the all-pairs scan (K2/K2r), the edge sort and ARACNE (K3) run on the GPU(s):
// --gpus N runs one host thread per device, every thread making the same (collective) C-ABI calls on its own context.
// There is no CPU fallback: without an sm_100 device spk_create fails and the program exits like the reference does on fatal
// errors (message + EXIT_FAILURE).
#include <algorithm>
#include <atomic>
#include <charconv>
#include <chrono>
#include <cinttypes>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <set>
#include <sstream>
#include <string>
#include <sys/mman.h>
#include <sys/stat.h>
#include <fcntl.h>
#include <thread>
#include <unistd.h>
#include <vector>
#include <zlib.h>

#include "spydrpick_b200.h"

namespace {

struct Options {
    // SpydrPick_options (src/SpydrPick_options.cpp:144-159)
    bool verbose = false;
    double mi_threshold = -1.0;
    uint64_t mi_values = 0;
    double mi_pseudocount = 0.5;
    uint64_t mi_threshold_iterations = 10;
    uint64_t mi_threshold_pairs = 0;
    uint64_t ld_threshold = 0;
    bool no_aracne = false;
    int threads = -1;                           // host threads for parsing / writing (-1 = all)
    // ARACNE_options (src/ARACNE_options.cpp:151-156)
    double aracne_edge_threshold = 2.220446049250313e-16;
    uint64_t aracne_block_size = 16384;         // accepted: the result does not depend on it (the reference streams its list in blocks)
    uint64_t aracne_node_grouping_size = 16;    // accepted: a scheduling knob of the reference's CPU implementation
    std::string aracne_outputfile = "aracne.out";   // accepted: never written by SpydrPick's main() either
    // Apegrunt_options (externals/apegrunt/src/Apegrunt_options.cpp:260-304)
    std::vector<std::string> alignment_files;
    std::string include_list, exclude_list, mapping_list, sample_list, sample_weights_file, sample_groups_file;
    int input_indexing_base = 1, output_indexing_base = 1;
    bool no_filter_alignment = false;
    double maf_threshold = 0.01, gap_threshold = 0.15;
    bool no_sample_reweighting = false, rescale_sample_weights = false;
    double sample_reweighting_threshold = 0.9;
    bool output_state_frequencies = false, output_sample_weights = false, output_sample_groups = false, output_sample_distance_matrix = false;
    bool output_alignment = false, output_filtered_alignment = false;
    uint64_t genome_size = 0;
    bool linear_genome = false;
    int begin_locus = 1, end_locus = -1;        // accepted: not read by the reference's SpydrPick path
    // additions of this implementation
    int device = -1, gpus = 1;
    uint64_t seed = 0;                          // 0 = seed from the clock like the reference (SpydrPick.hpp:178)
    std::string threshold_sampler = "device";   // device | reference
    int engine = 0;
    bool host_filter = false;
    std::string output_dir = ".";
};

[[noreturn]] void die(const std::string& msg) {
    std::cerr << "SpydrPick error: " << msg << "\n\n";
    std::cout.flush(); std::cerr.flush();
    std::_Exit(EXIT_FAILURE);               // (not exit(): the CUDA start-up thread may still be running; nothing to unwind here)
}

void print_help() {
    std::cout <<
        "SpydrPick: MI-ARACNE genome-wide co-evolution analysis (B200-native hot path).\n\n"
        "Usage: SpydrPick [options] <alignmentfile>\nOption '--help' will print a list of available options.\n"
        "  -h [ --help ]                         Print this help message.\n"
        "  --version                             Print version information.\n"
        "  -v [ --verbose ]                      Be verbose.\n"
        "  --mi-threshold arg (=-1)              The MI threshold value (0=no threshold; -1=determine automatically).\n"
        "  --mi-values arg (=0)                  Approximate number of MI values to calculate from data (default=#samples*100).\n"
        "  --mi-pseudocount arg (=0.5)           The MI pseudocount value.\n"
        "  --mi-threshold-iterations arg (=10)   Number of iterations for estimating saving threshold.\n"
        "  --mi-threshold-pairs arg (=0)         Number of sampled pairs for estimating saving threshold (0=determine automatically).\n"
        "  --ld-threshold arg (=0)               Threshold distance for linkage disequilibrium (LD).\n"
        "  --no-aracne                           Skip ARACNE, only calculate MI.\n"
        "  -t [ --threads ] arg (=-1)            Host threads for parsing and writing (-1=all); the hot path runs on the GPU(s).\n"
        "  --alignmentfile arg                   The input alignment filename (FASTA, optionally .gz).\n"
        "  --include-list arg                    Name of file containing list of loci to include in analysis.\n"
        "  --exclude-list arg                    Name of file containing list of loci to exclude from analysis.\n"
        "  --mapping-list arg                    Name of file containing loci mappings.\n"
        "  --sample-list arg                     The sample filter list input filename.\n"
        "  --sample-weights arg                  Name of file containing sample weights.\n"
        "  --sample-groups arg                   Name of file containing sample groups (accepted; unused by SpydrPick).\n"
        "  --input-indexing-base arg (=1)        Base index for input.\n"
        "  --no-filter-alignment                 Do not filter alignment columns by applying MAF and GAP thresholds.\n"
        "  --maf-threshold arg (=0.01)           Minor state frequency threshold. Loci with less than 2 states above threshold are removed from alignment.\n"
        "  --gap-threshold arg (=0.15)           Gap frequency threshold. Positions with a gap frequency above the threshold are excluded from the pair-analysis.\n"
        "  --no-sample-reweighting               Turn sample reweighting off, i.e. do not try to correct for population structure.\n"
        "  --sample-reweighting-threshold arg (=0.9)\n"
        "                                        Fraction of identical positions required for two samples to be considered identical.\n"
        "  --rescale-sample-weights              Rescale sample weights so that n(effective) == n.\n"
        "  --output-state-frequencies            Write column-wise state frequencies to file.\n"
        "  --output-sample-weights               Output sample weights.\n"
        "  --output-sample-groups                Output sample groups (accepted; unused by SpydrPick).\n"
        "  --output-sample-distance-matrix       Output triangular sample-sample Hamming distance matrix (not available: the reference\n"
        "                                        itself recurses without end in Alignment::distance_matrix() and crashes).\n"
        "  --output-indexing-base arg (=1)       Base index for output.\n"
        "  --output-alignment                    Write alignment to file.\n"
        "  --output-filtered-alignment           Write filtered alignment to file.\n"
        "  --genome-size arg                     Genome size, if different from input. Default = 0: detect size from input.\n"
        "  --linear-genome                       Assume linear genome in distance calculations.\n"
        "  --begin arg (=1)                      The first locus index to work on (accepted; unused by SpydrPick).\n"
        "  --end arg (=-1)                       The last locus index to work on (accepted; unused by SpydrPick).\n"
        "  -o [ --aracne-outputfile ] arg (=aracne.out)  Accepted for compatibility (SpydrPick never writes it).\n"
        "  --aracne-edge-threshold arg (=2.22045e-16)    Equality tolerance threshold. Edges differing by less than this value are considered equal in strength.\n"
        "  --aracne-block-size arg (=16384)      Accepted for compatibility (the result does not depend on it).\n"
        "  --aracne-node-grouping-size arg (=16) Accepted for compatibility.\n"
        "B200 options:\n"
        "  --gpus arg (=1)                       Number of GPUs of this box to use (tile space partitioned, edge lists all-gathered over NVLink).\n"
        "  --device arg (=-1)                    First CUDA device (-1 = current).\n"
        "  --seed arg (=0)                       Seed of the threshold estimator (0 = clock, like the reference).\n"
        "  --threshold-sampler arg (=device)     device: pairs sampled on the GPU; reference: the reference's mt19937 sampler (host).\n"
        "  --engine arg (=0)                     0/4 int8 tcgen05 filter with factorised byte weights, 3 int8 with 16-bit weights, 2 dense bf16 filter, 1 fp64 exhaustive.\n"
        "  --host-filter                         Count state frequencies and extract the kept columns on the host (default: device).\n"
        "  --output-dir arg (=.)                 Directory for the output files.\n";
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
        if (a.rfind("--", 0) != 0 && !(a.size() >= 2 && a[0] == '-' && !std::isdigit((unsigned char)a[1]))) { o.alignment_files.push_back(a); continue; }
        std::string name, inl; bool has_inl = false;
        if (a.rfind("--", 0) == 0) {
            name = a.substr(2);
            const size_t eq = name.find('=');
            if (eq != std::string::npos) { inl = name.substr(eq + 1); name = name.substr(0, eq); has_inl = true; }
        } else {
            switch (a[1]) { case 'h': name = "help"; break; case 'v': name = "verbose"; break; case 't': name = "threads"; break;
                            case 'o': name = "aracne-outputfile"; break; default: die("unrecognised option '" + a + "'"); }
            if (a.size() > 2) { inl = a.substr(2); has_inl = true; }
        }
        auto U = [&](const std::string& s) { return (uint64_t)std::strtoull(s.c_str(), nullptr, 10); };
        auto D = [&](const std::string& s) { return std::strtod(s.c_str(), nullptr); };
        auto I = [&](const std::string& s) { return std::atoi(s.c_str()); };
        auto arg = [&]() { return need(i, name, inl, has_inl); };
        if (name == "help") { print_help(); std::exit(EXIT_SUCCESS); }
        else if (name == "version") { std::cout << "SpydrPick (" << spk_version() << ")\n"; std::exit(EXIT_SUCCESS); }
        else if (name == "verbose") o.verbose = true;
        else if (name == "mi-threshold") o.mi_threshold = D(arg());
        else if (name == "mi-values") o.mi_values = U(arg());
        else if (name == "mi-pseudocount") o.mi_pseudocount = D(arg());
        else if (name == "mi-threshold-iterations") o.mi_threshold_iterations = U(arg());
        else if (name == "mi-threshold-pairs") o.mi_threshold_pairs = U(arg());
        else if (name == "ld-threshold") o.ld_threshold = U(arg());
        else if (name == "no-aracne") o.no_aracne = true;
        else if (name == "threads") o.threads = I(arg());
        else if (name == "aracne-edge-threshold") o.aracne_edge_threshold = D(arg());
        else if (name == "aracne-block-size") o.aracne_block_size = U(arg());
        else if (name == "aracne-node-grouping-size") o.aracne_node_grouping_size = U(arg());
        else if (name == "aracne-outputfile") o.aracne_outputfile = arg();
        else if (name == "alignmentfile") o.alignment_files.push_back(arg());
        else if (name == "include-list") o.include_list = arg();
        else if (name == "exclude-list") o.exclude_list = arg();
        else if (name == "mapping-list") o.mapping_list = arg();
        else if (name == "sample-list") o.sample_list = arg();
        else if (name == "sample-weights") o.sample_weights_file = arg();
        else if (name == "sample-groups") o.sample_groups_file = arg();
        else if (name == "input-indexing-base") o.input_indexing_base = I(arg());
        else if (name == "output-indexing-base") o.output_indexing_base = I(arg());
        else if (name == "no-filter-alignment") o.no_filter_alignment = true;
        else if (name == "host-filter") o.host_filter = true;
        else if (name == "maf-threshold") o.maf_threshold = D(arg());
        else if (name == "gap-threshold") o.gap_threshold = D(arg());
        else if (name == "no-sample-reweighting") o.no_sample_reweighting = true;
        else if (name == "sample-reweighting-threshold") o.sample_reweighting_threshold = D(arg());
        else if (name == "rescale-sample-weights") o.rescale_sample_weights = true;
        else if (name == "output-state-frequencies") o.output_state_frequencies = true;
        else if (name == "output-sample-weights") o.output_sample_weights = true;
        else if (name == "output-sample-groups") o.output_sample_groups = true;
        else if (name == "output-sample-distance-matrix") o.output_sample_distance_matrix = true;
        else if (name == "output-alignment") o.output_alignment = true;
        else if (name == "output-filtered-alignment") o.output_filtered_alignment = true;
        else if (name == "genome-size") o.genome_size = U(arg());
        else if (name == "linear-genome") o.linear_genome = true;
        else if (name == "begin") o.begin_locus = I(arg());
        else if (name == "end") o.end_locus = I(arg());
        else if (name == "device") o.device = I(arg());
        else if (name == "gpus") o.gpus = I(arg());
        else if (name == "seed") o.seed = U(arg());
        else if (name == "threshold-sampler") o.threshold_sampler = arg();
        else if (name == "engine") o.engine = I(arg());
        else if (name == "output-dir") o.output_dir = arg();
        else die("unrecognised option '--" + name + "'");
    }
    if (o.alignment_files.empty()) { std::cerr << "SpydrPick ERROR: No alignment file specified!\n"; print_help(); std::exit(EXIT_FAILURE); }
    if (o.alignment_files.size() > 1) die("will not accept more than 1 alignment (got " + std::to_string(o.alignment_files.size()) + " filenames)");
    if (o.gpus < 1 || o.gpus > 16) die("--gpus must be 1..16");
    if (o.threshold_sampler != "device" && o.threshold_sampler != "reference") die("--threshold-sampler must be 'device' or 'reference'");
    return o;
}

unsigned host_threads(const Options& o) {
    const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    return o.threads > 0 ? (unsigned)o.threads : hw;
}

template <class F> void parallel_chunks(unsigned nt, size_t n, F f) {         // f(begin, end, thread)
    nt = (unsigned)std::max<size_t>(1, std::min<size_t>(nt, n));
    if (nt == 1) { f(size_t(0), n, 0u); return; }
    std::vector<std::thread> th;
    for (unsigned t = 0; t < nt; ++t) th.emplace_back([=]() { f(n * t / nt, n * (t + 1) / nt, t); });
    for (auto& x : th) x.join();
}

// ---- alignment ----------------------------------------------------------------------------------------------
struct HostAlignment {
    std::string id;                       // file stem (+ list / filter ids, as the reference appends them)
    uint32_t n = 0; uint64_t L = 0;       // samples, loci
    std::vector<std::string> names;
    uint8_t* states = nullptr;            // [n][L_src] codes a=0,t=1,c=2,g=3,GAP=4 (pinned host memory)
    std::vector<uint64_t> pos;            // loci translation: original position of every column (0-based)
    uint64_t n_original_positions = 0;
    // the device-side column filter leaves the host matrix as it is: column k of the alignment is column col_src[k] of the
    // L_src-column host matrix (empty = identity); only the exports read it
    uint64_t L_src = 0; std::vector<uint64_t> col_src;
    void adopt(uint8_t* s) { if (states) std::free(states); states = s; }
    uint64_t src_col(uint64_t k) const { return col_src.empty() ? k : col_src[k]; }
};

// The alignment lives in ordinary (pageable) memory: the CUDA contexts are still being created on another thread while the
// FASTA file is parsed (context creation takes far longer than the parse), and the one upload of a few hundred MB through the
// driver's staging buffers costs less than page-locking them would.
uint8_t* alloc_states(size_t bytes) {
    void* mem = nullptr;
    if (posix_memalign(&mem, 4096, bytes ? bytes : 16) != 0) die("could not allocate " + std::to_string(bytes) + " bytes of host memory for the alignment");
    return (uint8_t*)mem;
}

struct StateLut { uint8_t v[256]; StateLut() { std::memset(v, 4, sizeof v); v['a'] = v['A'] = 0; v['t'] = v['T'] = 1; v['c'] = v['C'] = 2; v['g'] = v['G'] = 3; } };
const StateLut STATE_LUT;                 // StateVector_state_types.cpp:33-44: a,t,c,g (either case); everything else is a gap
inline bool is_ws(char ch) { return ch == '\n' || ch == '\r' || ch == ' ' || ch == '\t' || ch == '\v' || ch == '\f'; }

std::string file_stem(const std::string& path) {
    const size_t sl = path.find_last_of('/');
    std::string f = sl == std::string::npos ? path : path.substr(sl + 1);
    const size_t dot = f.find_last_of('.');
    return dot == std::string::npos || dot == 0 ? f : f.substr(0, dot);
}
bool ends_with(const std::string& s, const char* suf) { const size_t k = std::strlen(suf); return s.size() >= k && s.compare(s.size() - k, k, suf) == 0; }

// FASTA ingest (apegrunt/Alignment_parsers.hpp:72-144): the file is mapped, record starts are found by all host threads in
// parallel, and every record is transcoded straight into its row of the page-locked state matrix. .gz goes through zlib first.
HostAlignment read_fasta(const std::string& path, const Options& o) {
    HostAlignment A;
    const char* data = nullptr; size_t size = 0;
    std::string inflated; void* map = nullptr; size_t map_size = 0;
    if (ends_with(path, ".xz")) die("xz-compressed FASTA is not supported in this build: decompress \"" + path + "\" first");
    if (ends_with(path, ".gz")) {
        gzFile gz = gzopen(path.c_str(), "rb");
        if (!gz) die("could not open alignment file \"" + path + "\"");
        gzbuffer(gz, 1 << 20);
        std::vector<char> buf(1 << 22); int k;
        while ((k = gzread(gz, buf.data(), (unsigned)buf.size())) > 0) inflated.append(buf.data(), (size_t)k);
        if (k < 0) die("could not decompress \"" + path + "\"");
        gzclose(gz);
        data = inflated.data(); size = inflated.size();
    } else {
        const int fd = ::open(path.c_str(), O_RDONLY);
        if (fd < 0) die("could not open alignment file \"" + path + "\"");
        struct stat st; if (fstat(fd, &st) != 0) die("could not stat \"" + path + "\"");
        size = (size_t)st.st_size; map_size = size;
        if (size) { map = mmap(nullptr, size, PROT_READ, MAP_PRIVATE | MAP_POPULATE, fd, 0); if (map == MAP_FAILED) die("could not map \"" + path + "\""); }
        ::close(fd);
        data = (const char*)map;
    }
    A.id = file_stem(path);                           // filepath.stem(): "x.fasta.gz" -> "x.fasta", as the reference (:132)
    const unsigned nt = host_threads(o);
    // record starts: '>' at the beginning of the file (after leading whitespace) or right after a newline
    std::vector<std::vector<size_t>> found(nt);
    parallel_chunks(nt, size, [&](size_t b, size_t e, unsigned t) {
        for (size_t p = b; p < e; ++p) {
            const char* q = (const char*)std::memchr(data + p, '>', e - p);
            if (!q) break;
            p = (size_t)(q - data);
            if (p == 0 || data[p - 1] == '\n') found[t].push_back(p);
            else { size_t k = p; bool lead = true; while (k > 0) { --k; if (!is_ws(data[k])) { lead = false; break; } } if (lead) found[t].push_back(p); }
        }
    });
    std::vector<size_t> starts;
    for (auto& v : found) starts.insert(starts.end(), v.begin(), v.end());
    if (starts.empty()) die("no FASTA records in \"" + path + "\"");
    A.n = (uint32_t)starts.size();
    A.names.resize(A.n);
    std::vector<size_t> seq_begin(A.n), seq_end(A.n);
    for (uint32_t s = 0; s < A.n; ++s) {
        const size_t p = starts[s], end = s + 1 < A.n ? starts[s + 1] : size;
        const char* nl = (const char*)std::memchr(data + p, '\n', end - p);
        const size_t eol = nl ? (size_t)(nl - data) : end;
        A.names[s].assign(data + p + 1, eol - p - 1);
        while (!A.names[s].empty() && (A.names[s].back() == '\r')) A.names[s].pop_back();
        seq_begin[s] = std::min(eol + 1, end); seq_end[s] = end;
    }
    { size_t k = 0; for (size_t q = seq_begin[0]; q < seq_end[0]; ++q) k += !is_ws(data[q]); A.L = k; }
    if (A.L < 2) die("alignment has fewer than 2 columns");
    A.states = alloc_states((size_t)A.n * A.L);
    std::atomic<int64_t> bad(-1);
    parallel_chunks(nt, A.n, [&](size_t b, size_t e, unsigned) {
        for (size_t s = b; s < e; ++s) {
            uint8_t* row = A.states + s * A.L; size_t k = 0;
            for (size_t q = seq_begin[s]; q < seq_end[s]; ++q) { const char ch = data[q]; if (is_ws(ch)) continue; if (k < A.L) row[k] = STATE_LUT.v[(unsigned char)ch]; ++k; }
            if (k != A.L) bad.store((int64_t)s);
        }
    });
    if (bad.load() >= 0) die("sequence \"" + A.names[(size_t)bad.load()] + "\" does not have " + std::to_string(A.L) + " positions");
    if (map) munmap(map, map_size);
    A.pos.resize(A.L); for (uint64_t c = 0; c < A.L; ++c) A.pos[c] = c;
    A.L_src = A.L;
    A.n_original_positions = A.L;                       // Alignment_parsers.hpp:133 (genome size is applied by the caller)
    return A;
}

// whitespace-separated `a` or `a-b` (inclusive) entries minus the indexing base (Loci_parser_grammar.hpp:65-84,
// Loci_impl_default_storage.hpp:135-163); parsing stops at the first token that is neither, like the grammar does
std::vector<uint64_t> read_loci_list(const std::string& path, int base) {
    std::ifstream in(path, std::ios::binary);
    if (!in) die("could not open list file \"" + path + "\"");
    std::stringstream ss; ss << in.rdbuf();
    const std::string t = ss.str();
    std::vector<uint64_t> v; size_t p = 0;
    auto skip = [&]() { while (p < t.size() && std::isspace((unsigned char)t[p])) ++p; };
    auto num = [&](uint64_t& x) { if (p >= t.size() || !std::isdigit((unsigned char)t[p])) return false; x = 0; while (p < t.size() && std::isdigit((unsigned char)t[p])) x = x * 10 + (uint64_t)(t[p++] - '0'); return true; };
    for (;;) {
        skip(); uint64_t a = 0, b = 0;
        if (!num(a)) break;
        const size_t after = p; skip();
        if (p < t.size() && t[p] == '-') {
            ++p; skip();
            if (!num(b)) break;
            for (uint64_t x = a; x < b + 1; ++x) v.push_back(x - (uint64_t)base);
        } else { p = after; v.push_back(a - (uint64_t)base); }
    }
    return v;
}

// Alignment::subset (Alignment_impl_block_compressed_storage.hpp:502-530): `idx` are column INDICES of the current alignment;
// same number of entries as columns -> the alignment is returned unchanged (:514), whatever the entries are
void subset_columns(HostAlignment& A, const std::vector<uint64_t>& idx, const std::string& list_id, const Options& o) {
    if (idx.empty()) die("subset alignment is empty: nothing to do");
    if (idx.size() == A.L) return;
    for (uint64_t c : idx) if (c >= A.L) die("locus index " + std::to_string(c) + " out of range (alignment has " + std::to_string(A.L) + " loci)");
    const uint64_t Lk = idx.size();
    uint8_t* f = alloc_states((size_t)A.n * Lk);
    parallel_chunks(host_threads(o), A.n, [&](size_t b, size_t e, unsigned) {
        for (size_t s = b; s < e; ++s) { const uint8_t* row = A.states + s * A.L_src; uint8_t* out = f + s * Lk; for (uint64_t k = 0; k < Lk; ++k) out[k] = row[A.src_col(idx[k])]; }
    });
    std::vector<uint64_t> npos(Lk); for (uint64_t k = 0; k < Lk; ++k) npos[k] = A.pos[idx[k]];
    A.adopt(f); A.L = Lk; A.L_src = Lk; A.col_src.clear(); A.pos.swap(npos);
    if (!list_id.empty()) A.id += "." + list_id;                      // :527
}

// Alignment::subsample (Alignment_impl_base.hpp:120-140): rows in list order; weights are not transferred
void subsample_rows(HostAlignment& A, const std::vector<uint64_t>& rows, const std::string& list_id) {
    if (rows.empty()) die("sample list is empty: nothing to do");
    if (!A.col_src.empty()) die("internal: sample list on a lazily filtered alignment");
    for (uint64_t r : rows) if (r >= A.n) die("sample index " + std::to_string(r) + " out of range (alignment has " + std::to_string(A.n) + " samples)");
    uint8_t* f = alloc_states(rows.size() * (size_t)A.L);
    std::vector<std::string> names(rows.size());
    for (size_t k = 0; k < rows.size(); ++k) { std::memcpy(f + k * A.L, A.states + rows[k] * A.L, A.L); names[k] = A.names[rows[k]]; }
    A.adopt(f); A.n = (uint32_t)rows.size(); A.names.swap(names);
    if (!list_id.empty()) A.id += "." + list_id;
}

std::string filter_id(const Options& o) {
    char id[128];
    std::snprintf(id, sizeof id, "filtered_ge%03zumaf_le%03zugf_gt01-lt05states", (size_t)(o.maf_threshold * 100.), (size_t)(o.gap_threshold * 100.));
    return id;
}

// host variant of the column filter (--host-filter): Alignment_filter.hpp:250-339
std::vector<uint64_t> host_filter_list(const HostAlignment& A, const Options& o) {
    const uint32_t n = A.n; const uint64_t L = A.L;
    std::vector<uint32_t> cnt(L * 5, 0);
    for (uint32_t s = 0; s < n; ++s) { const uint8_t* row = A.states + (size_t)s * A.L_src; for (uint64_t c = 0; c < L; ++c) cnt[c * 5 + row[A.src_col(c)]]++; }
    std::vector<uint64_t> keep;
    for (uint64_t c = 0; c < L; ++c) {
        int nsig = 0;
        for (int a = 0; a < 4; ++a) if (double(cnt[c * 5 + a]) / double(n) >= o.maf_threshold) ++nsig;
        if (nsig > 1 && double(cnt[c * 5 + 4]) / double(n) <= o.gap_threshold) keep.push_back(c);
    }
    return keep;
}

// ---- writers ------------------------------------------------------------------------------------------------
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
void write_parallel(const std::string& path, uint64_t n_items, unsigned hw, F format_range) {
    const uint64_t chunk = 1u << 17;
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
    write_parallel(path, ne, host_threads(o), [&](uint64_t b, uint64_t en, std::string& s) {
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
// effect in percent and the extreme-outlier flag. The list is sorted, so the candidates are the leading run with
// mi >= outlier threshold; that run is formatted in parallel against the id-sorted MI-without-gaps list.
uint64_t write_outliers(const std::string& path, const spk_edge* e, const uint8_t* flags, uint64_t ne, const spk_edge* wo, uint64_t nwo,
                        double outlier_thr, double extreme_thr, const HostAlignment& A, const Options& o) {
    auto eid = [](const spk_edge& x) { return ((uint64_t)std::max(x.i, x.j) << 32) | std::min(x.i, x.j); };
    uint64_t n_run = 0;
    while (n_run < ne && !(e[n_run].mi < outlier_thr)) ++n_run;                      // the reference breaks at the first weight < threshold (:104)
    std::vector<spk_edge> w;
    if (n_run) {
        w.assign(wo, wo + nwo);
        std::sort(w.begin(), w.end(), [&](const spk_edge& a, const spk_edge& b) { return eid(a) < eid(b); });
    }
    std::atomic<uint64_t> lines(0);
    write_parallel(path, n_run, host_threads(o), [&](uint64_t b, uint64_t en, std::string& s) {
        uint64_t mine = 0;
        for (uint64_t k = b; k < en; ++k) {
            auto it = std::lower_bound(w.begin(), w.end(), e[k], [&](const spk_edge& a, const spk_edge& bb) { return eid(a) < eid(bb); });
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
                ++mine;
            }
        }
        lines += mine;
    });
    return lines.load();
}

// output_alignment (Alignment_utility.hpp:295-323): <id>.L<L>n<n>.fasta (one line of lower-case states per record,
// Alignment_generators.hpp:45-58) and <id>.L<L>n<n>.loci (original position + output base per line, Loci_generators.hpp:31-44).
// `cols` are column indices of the current alignment.
void write_alignment(const HostAlignment& A, const std::vector<uint64_t>& cols, const Options& o, bool verbose) {
    static const char CH[5] = {'a', 't', 'c', 'g', '-'};
    const uint64_t Lk = cols.size();
    const std::string stem = o.output_dir + "/" + A.id + ".L" + std::to_string(Lk) + "n" + std::to_string(A.n);
    const std::string ffile = unique_name(stem + ".fasta"), lfile = unique_name(stem + ".loci");
    if (verbose) std::cout << "apegrunt: write alignment to file " << ffile << "\n";
    std::vector<uint64_t> sc(Lk); for (uint64_t k = 0; k < Lk; ++k) sc[k] = A.src_col(cols[k]);
    write_parallel(ffile, A.n, host_threads(o), [&](uint64_t b, uint64_t e, std::string& s) {
        s.reserve((e - b) * (Lk + 32));
        for (uint64_t r = b; r < e; ++r) {
            s += '>'; s += A.names[r]; s += '\n';
            const uint8_t* row = A.states + r * A.L_src;
            for (uint64_t k = 0; k < Lk; ++k) s += CH[row[sc[k]] < 5 ? row[sc[k]] : 4];
            s += '\n';
        }
    });
    if (verbose) std::cout << "apegrunt: write original loci indices to file " << lfile << "\n";
    std::string s;
    for (uint64_t k = 0; k < Lk; ++k) { put_u64(s, A.pos[cols[k]] + (uint64_t)o.output_indexing_base); s += '\n'; }
    std::ofstream out(lfile, std::ios::binary); out.write(s.data(), (std::streamsize)s.size());
}

struct Timer {
    std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
    double lap() { auto t = std::chrono::steady_clock::now(); const double s = std::chrono::duration<double>(t - t0).count(); t0 = t; return s; }
};

void check(spk_ctx* ctx, int rc, const char* what) {
    if (rc != SPK_OK) die(std::string(what) + ": " + (spk_last_error(ctx) && *spk_last_error(ctx) ? spk_last_error(ctx) : "unknown error"));
}

} // namespace

// ---- main: src/SpydrPick.cpp:45-510 -----------------------------------------------------------------------------
int main(int argc, char** argv) {
    Options o = parse_options(argc, argv);
    Timer total, lap;
    const bool V = o.verbose;
    if (V) std::cout << "SpydrPick (" << spk_version() << ")\n";

    // -- contexts: one per GPU (SPMD); rank 0 is the one whose results are written. Created on a side thread: CUDA start-up
    //    (driver, primary contexts, NCCL communicator) takes seconds and overlaps the FASTA ingest below.
    const int G = o.gpus;
    std::vector<spk_ctx*> ctxs(G, nullptr);
    std::string ctx_error; double ctx_seconds = 0.0;
    std::thread ctx_thread([&]() {
        Timer t;
        // one thread per device: primary-context creation is per device and takes most of a second each
        std::vector<std::string> errs(G);
        std::vector<std::thread> th;
        for (int g = 0; g < G; ++g)
            th.emplace_back([&, g]() {
                const int dev = (G == 1) ? o.device : ((o.device < 0 ? 0 : o.device) + g);
                if (spk_create(&ctxs[g], dev, (V && g == 0) ? 1 : 0) != SPK_OK) errs[g] = std::string("cannot initialise the CUDA context: ") + spk_last_error(nullptr);
            });
        for (auto& x : th) x.join();
        for (int g = 0; g < G; ++g) if (!errs[g].empty()) { ctx_error = errs[g]; return; }
        if (G > 1 && spk_comm_init_all(ctxs.data(), G) != SPK_OK) ctx_error = std::string("NCCL communicator: ") + spk_last_error(ctxs[0]);
        ctx_seconds = t.lap();
    });
    struct Joiner { std::thread& t; ~Joiner() { if (t.joinable()) t.join(); } } ctx_joiner{ctx_thread};

    // -- ingest (src/SpydrPick.cpp:163; apegrunt/Alignment_utility.hpp:69-160)
    HostAlignment A = read_fasta(o.alignment_files[0], o);
    if (o.genome_size) A.n_original_positions = o.genome_size;             // Alignment_parsers.hpp:133
    if (V) std::cout << "SpydrPick: read " << A.n << " sequences x " << A.L << " loci from \"" << o.alignment_files[0] << "\" (" << lap.lap() << " s)\n";
    if (!o.mapping_list.empty()) {
        std::vector<uint64_t> m = read_loci_list(o.mapping_list, o.input_indexing_base);
        if (m.size() != A.L) die("mapping list and alignment sizes are mismatched (" + std::to_string(m.size()) + " and " + std::to_string(A.L) + ")");
        A.pos.swap(m);                     // the mapping list only replaces the loci translation: n_original_positions stays the
    }                                      // FASTA column count unless --genome-size is given (Alignment_utility.hpp:128-160)
    if (o.output_sample_distance_matrix) std::cerr << "SpydrPick: --output-sample-distance-matrix is not available in this build (ignored)\n";
    if (o.output_alignment) {
        std::vector<uint64_t> all(A.L); for (uint64_t c = 0; c < A.L; ++c) all[c] = c;
        write_alignment(A, all, o, V);
    }
    // -- include / exclude lists (src/SpydrPick.cpp:191-231)
    if (!o.include_list.empty()) {
        subset_columns(A, read_loci_list(o.include_list, o.input_indexing_base), file_stem(o.include_list), o);
        if (V) std::cout << "SpydrPick: include list leaves " << A.L << " loci\n";
    }
    if (!o.exclude_list.empty()) {
        // alignment->get_loci_translation() - exclude_list: translation VALUES not in the list (Loci_operations.hpp:37-46),
        // handed to subset() which reads them as column indices (identical things without a mapping list)
        const std::vector<uint64_t> ex = read_loci_list(o.exclude_list, o.input_indexing_base);
        const std::set<uint64_t> exs(ex.begin(), ex.end());
        std::vector<uint64_t> keep; keep.reserve(A.L);
        for (uint64_t c = 0; c < A.L; ++c) if (!exs.count(A.pos[c])) keep.push_back(A.pos[c]);
        subset_columns(A, keep, "", o);
        if (V) std::cout << "SpydrPick: exclude list leaves " << A.L << " loci\n";
    }

    ctx_thread.join();
    if (!ctx_error.empty()) die(ctx_error);
    if (V) std::cout << "SpydrPick: " << G << " CUDA context" << (G > 1 ? "s + NCCL communicator" : "") << " ready (" << ctx_seconds << " s of start-up, overlapped with the ingest; waited " << lap.lap() << " s)\n";
    spk_ctx* ctx = ctxs[0];
    auto on_all = [&](auto fn) {           // the same call on every rank's context, one host thread each (collective calls meet on the GPUs)
        if (G == 1) { fn(0); return; }
        std::vector<std::thread> th;
        for (int g = 0; g < G; ++g) th.emplace_back([&, g]() { fn(g); });
        for (auto& t : th) t.join();
    };

    // -- column filter (src/SpydrPick.cpp:234-297) + alignment upload: the raw matrix goes up once, state frequencies are
    //    counted and the kept columns extracted on the device (spk_set_alignment_filtered); --host-filter keeps the host pass
    const uint64_t L_unfiltered = A.L;
    bool uploaded = false;
    if (!o.no_filter_alignment) {
        if (A.L >= (1ull << 32)) die("too many loci");
        std::vector<uint64_t> kept;
        if (!o.sample_list.empty()) o.host_filter = true;      // the sample list is applied after the filter: the host copy has to be in step
        if (!o.host_filter) {
            std::vector<std::vector<uint8_t>> keeps(G, std::vector<uint8_t>(A.L));
            std::vector<uint64_t> nk(G, 0); std::vector<int> empty(G, 0);
            on_all([&](int g) {
                const int rc = spk_set_alignment_filtered(ctxs[g], A.states, A.n, A.L, A.pos.data(), o.maf_threshold, o.gap_threshold, keeps[g].data(), &nk[g]);
                if (rc != SPK_OK) {
                    if (std::string(spk_last_error(ctxs[g])).find("fewer than 2 columns") != std::string::npos) { empty[g] = 1; return; }
                    check(ctxs[g], rc, "column filter + alignment upload");
                }
            });
            if (V) std::cout << "SpydrPick: apply filter rules.. " << nk[0] << " positions fulfill filter criteria\n";
            if (empty[0]) { if (V) std::cout << "SpydrPick: nothing to do\n"; std::exit(EXIT_SUCCESS); }
            for (uint64_t c = 0; c < A.L; ++c) if (keeps[0][c]) kept.push_back(c);
            uploaded = true;
            // the host matrix stays as it is (only the exports read it, through col_src); positions and id follow the filter
            if (kept.size() != A.L) {
                std::vector<uint64_t> npos; npos.reserve(kept.size()); for (uint64_t c : kept) npos.push_back(A.pos[c]);
                A.pos.swap(npos); A.col_src = kept; A.L = kept.size(); A.id += "." + filter_id(o);
            }
        } else {
            kept = host_filter_list(A, o);
            if (V) std::cout << "SpydrPick: apply filter rules.. " << kept.size() << " positions fulfill filter criteria\n";
            if (kept.empty()) { if (V) std::cout << "SpydrPick: nothing to do\n"; std::exit(EXIT_SUCCESS); }
            if (kept.size() != A.L) subset_columns(A, kept, filter_id(o), o);
        }
        if (V) std::cout << "SpydrPick: filtered alignment has " << A.L << " of " << L_unfiltered << " loci (" << lap.lap() << " s)\n";
    }
    // -- sample list (src/SpydrPick.cpp:299-318)
    if (!o.sample_list.empty()) {
        subsample_rows(A, read_loci_list(o.sample_list, o.input_indexing_base), file_stem(o.sample_list));
        if (V) std::cout << "SpydrPick: sample list leaves " << A.n << " samples\n";
        uploaded = false;
    }
    if (A.L < 2) die("fewer than 2 loci left: nothing to do");
    if (A.L >= (1ull << 29)) die("too many loci");
    if (!uploaded) on_all([&](int g) { check(ctxs[g], spk_set_alignment(ctxs[g], A.states, A.n, (uint32_t)A.L, A.pos.data()), "alignment upload"); });
    const uint32_t n = A.n, L = (uint32_t)A.L;

    // -- sample weights (src/SpydrPick.cpp:321; apegrunt/Alignment_utility.hpp:186-258)
    std::vector<double> w(n, 1.0);
    if (!o.no_sample_reweighting) {
        if (!o.sample_weights_file.empty()) {
            std::ifstream in(o.sample_weights_file); if (!in) die("could not open sample weights file \"" + o.sample_weights_file + "\"");
            size_t k = 0; double x; while (k < n && (in >> x)) w[k++] = x;
            if (k != n) die("sample weights: expected " + std::to_string(n) + " values, got " + std::to_string(k));
            if (o.rescale_sample_weights) { double n_eff = 0; for (double x2 : w) n_eff += x2; for (double& x2 : w) x2 = x2 * double(n) / n_eff; }   // (w * n) / n_eff, Alignment_utility.hpp:215
            on_all([&](int g) { check(ctxs[g], spk_set_weights(ctxs[g], w.data()), "sample weights"); });
        } else {
            const uint64_t min_identical = (uint64_t)(o.sample_reweighting_threshold * double(L));          // weights.hpp:228,251
            on_all([&](int g) { check(ctxs[g], spk_reweight(ctxs[g], min_identical, SPK_QUIRKS_REFERENCE, o.rescale_sample_weights ? 1 : 0, g == 0 ? w.data() : nullptr, nullptr), "sample reweighting"); });
        }
    }
    double neff = 0; for (double x : w) neff += x;
    if (V) std::cout << "apegrunt: effective sample size = " << neff << " (" << lap.lap() << " s)\n";
    const std::string size_string = "L" + std::to_string(L) + "n" + std::to_string(n);                        // Alignment_utility.hpp:66
    const std::string prefix = o.output_dir + "/" + A.id + "." + size_string;
    if (o.output_sample_weights) {                                                                            // Alignment_utility.hpp:261-292: scientific, 8 digits
        std::ofstream out(unique_name(prefix + ".weights"));
        char t[64]; for (double x : w) { std::snprintf(t, sizeof t, "%.8e\n", x); out << t; }
    }
    if (o.output_filtered_alignment) {
        std::vector<uint64_t> all(A.L); for (uint64_t c = 0; c < A.L; ++c) all[c] = c;
        write_alignment(A, all, o, V);
    }
    if (o.output_state_frequencies) {
        // apegrunt::output_state_frequencies (Alignment_utility.hpp:431-449, 470-580): per column the fraction of samples in each of
        // the five states, "f f f f f \n" at the stream's default precision; with reweighting also the weight-normalised table.
        // The two *_distribution.txt summaries (apegrunt's Boost.Accumulators extension) are not written.
        std::vector<uint32_t> cnt(A.L * 5, 0);
        std::vector<double> wcnt(o.no_sample_reweighting ? 0 : A.L * 5, 0.0);
        parallel_chunks(host_threads(o), A.L, [&](size_t b, size_t e, unsigned) {
            for (uint32_t s = 0; s < n; ++s) {
                const uint8_t* row = A.states + (size_t)s * A.L_src; const double ws = w[s];
                for (size_t c = b; c < e; ++c) { const uint8_t x = row[A.src_col(c)]; cnt[c * 5 + x]++; if (!wcnt.empty()) wcnt[c * 5 + x] += ws; }
            }
        });
        auto write_table = [&](const std::string& path, auto value) {
            const std::string fn = unique_name(path);
            if (V) std::cout << "apegrunt: write " << (path.find("weighted") != std::string::npos ? "weighted " : "") << "columnwise state frequencies to file \"" << fn << "\"\n";
            std::ofstream out(fn);
            char t[32]; std::string line;
            for (uint64_t c = 0; c < A.L; ++c) {
                line.clear();
                for (int a = 0; a < 5; ++a) { std::snprintf(t, sizeof t, "%g ", value(c, a)); line += t; }     // operator<<(double): %g, 6 digits
                line += "\n"; out << line;
            }
        };
        write_table(prefix + ".frequencies.txt", [&](uint64_t c, int a) { return double(cnt[c * 5 + a]) / double(n); });
        if (!o.no_sample_reweighting) write_table(prefix + ".weighted_frequencies.txt", [&](uint64_t c, int a) { return wcnt[c * 5 + a] / neff; });
        if (V) std::cout << "SpydrPick: the frequency_distribution.txt summaries are not written by this build\n";
    }

    // -- saving threshold (src/SpydrPick.cpp:336-356; SpydrPick.hpp:285-343)
    double mi_threshold = o.mi_threshold;
    if (mi_threshold < 0.0) {
        const uint64_t top_pairs = o.mi_values ? o.mi_values : std::min<uint64_t>(10000000ull, 100ull * L);
        const uint64_t possible = (uint64_t)L * (L - 1) / 2;
        if (V) std::cout << "SpydrPick: determine MI threshold for saving approx. " << top_pairs << " top pairs\n";
        if (top_pairs >= possible) {
            mi_threshold = 0.0;             // every pair is wanted (the reference's estimator indexes out of range here)
            if (V) std::cout << "SpydrPick: fewer pairs than requested values: MI save threshold = 0\n";
        } else {
            const uint64_t seed = o.seed ? o.seed : (uint64_t)std::chrono::system_clock::now().time_since_epoch().count();
            const int sampler = o.threshold_sampler == "reference" ? SPK_SAMPLER_REFERENCE : SPK_SAMPLER_DEVICE;
            std::vector<double> thr(G, 0.0);
            on_all([&](int g) { check(ctxs[g], spk_determine_mi_threshold_ex(ctxs[g], top_pairs, o.mi_threshold_pairs, (uint32_t)o.mi_threshold_iterations, o.mi_pseudocount, seed, sampler, &thr[g]), "MI threshold estimation"); });
            mi_threshold = thr[0];
            std::cout.precision(6);
            if (V) std::cout << "SpydrPick: MI save threshold = " << mi_threshold << " (" << lap.lap() << " s)\n";
        }
    } else if (V) std::cout << "SpydrPick: user-defined MI save threshold = " << mi_threshold << "\n";

    // -- all-pairs scan + sort (src/SpydrPick.cpp:384-398)
    spk_scan_params sp; std::memset(&sp, 0, sizeof sp);
    sp.threshold = mi_threshold; sp.pseudocount = o.mi_pseudocount; sp.ld_threshold = o.ld_threshold; sp.linear_genome = o.linear_genome ? 1 : 0;
    sp.circle_size = 0; sp.max_edges = o.mi_values ? 3 * o.mi_values / (uint64_t)G + 65536 : 0; sp.sort_desc = 1; sp.engine = o.engine;
    sp.n_parts = (G > 1) ? 0 : 1;      // G > 1: the collective scan (partition + exchange)
    sp.keep_on_device = 2;             // the host copies of the edge lists run on a side stream while ARACNE works on the device copy
    std::vector<spk_scan_result> results(G);
    on_all([&](int g) { check(ctxs[g], spk_mi_scan(ctxs[g], &sp, &results[g]), "MI scan"); });
    const spk_scan_result& res = results[0];
    double outlier_thr = 0, extreme_thr = 0;
    check(ctx, spk_outlier_thresholds(res.colmax, L, &outlier_thr, &extreme_thr), "outlier thresholds");
    if (V) {
        std::cout.precision(6);
        std::cout << "SpydrPick: outlier threshold=" << outlier_thr << "\nSpydrPick: extreme outlier threshold=" << extreme_thr << "\n";
        std::cout << "SpydrPick: " << res.n_edges << " edges stored (" << res.n_candidates << " candidates refined in fp64 on rank 0, filter error "
                  << res.max_filter_error << ", audit " << res.audit_max_error << " over " << res.audit_pairs << " pairs, band " << res.band << ") (" << lap.lap() << " s)\n";
    }

    // -- ARACNE (src/SpydrPick.cpp:406-421)
    void* flag_mem = nullptr;
    if (spk_host_alloc((size_t)std::max<uint64_t>(res.n_edges, 1), &flag_mem) != SPK_OK) die("could not allocate pinned host memory for the ARACNE flags");
    uint8_t* flags = (uint8_t*)flag_mem;
    std::memset(flags, 0, (size_t)res.n_edges);
    if (!o.no_aracne && res.n_edges) {
        if (V) std::cout << "SpydrPick: run ARACNE\n";
        on_all([&](int g) {
            if (G == 1) check(ctxs[g], spk_aracne(ctxs[g], res.edges, res.n_edges, o.aracne_edge_threshold, 0, 1, flags), "ARACNE");
            else check(ctxs[g], spk_aracne(ctxs[g], g == 0 ? res.edges : nullptr, res.n_edges, o.aracne_edge_threshold, 0, 0, g == 0 ? flags : nullptr), "ARACNE");
        });
        if (V) { uint64_t nd = 0; for (uint64_t e = 0; e < res.n_edges; ++e) nd += flags[e]; std::cout << "SpydrPick: ARACNE keeps " << nd << " direct edges (" << lap.lap() << " s)\n"; }
    }
    check(ctx, spk_wait_download(ctx), "edge download");          // res.edges / res.edges_wo_gaps are complete from here on

    // -- outputs (src/SpydrPick.cpp:424-480)
    const std::string couplings = prefix + ".spydrpick_couplings." + std::to_string(o.output_indexing_base) + "-based";
    const std::string edges_file = unique_name(couplings + "." + std::to_string(res.n_edges) + "edges");
    if (V) std::cout << "SpydrPick: write network (" << res.n_edges << " edges) to file \"" << edges_file << "\"\n";
    write_edges(edges_file, res.edges, flags, res.n_edges, A, o);
    if (V) std::cout << "      done in " << lap.lap() << " s\n";
    const std::string outliers_file = unique_name(couplings + ".outliers");
    if (V) std::cout << "SpydrPick: write outlier network to file \"" << outliers_file << "\"\n";
    const uint64_t n_out = write_outliers(outliers_file, res.edges, flags, res.n_edges, res.edges_wo_gaps, res.n_wo_gaps, outlier_thr, extreme_thr, A, o);
    if (V) std::cout << "      " << n_out << " outliers, done in " << lap.lap() << " s\n";

    // -- nodes involved in outlier edges + their sub-alignment (src/SpydrPick.cpp:483-510; Graph_utility.hpp:36-55)
    {
        std::vector<uint64_t> nodes;
        for (uint64_t k = 0; k < res.n_edges; ++k) { if (!(res.edges[k].mi >= outlier_thr)) break; nodes.push_back(res.edges[k].i); nodes.push_back(res.edges[k].j); }
        std::sort(nodes.begin(), nodes.end()); nodes.erase(std::unique(nodes.begin(), nodes.end()), nodes.end());
        std::cout << "SpydrPick: extract nodes involved in outlier edges: found " << nodes.size() << " nodes\n";
        if (nodes.size() < A.L && !nodes.empty()) write_alignment(A, nodes, o, V);
    }
    if (V) {
        spk_timings tm; if (spk_get_timings(ctx, &tm) == SPK_OK)
            std::cout << "SpydrPick: device ms (rank 0): reweighting " << tm.hamming_ms << ", threshold " << tm.threshold_ms << ", filter GEMM " << tm.gemm_ms << ", audit " << tm.audit_ms
                      << ", fp64 refine " << tm.refine_ms << ", sort " << tm.sort_ms << ", exchange " << tm.exchange_ms << ", ARACNE " << tm.aracne_build_ms + tm.aracne_ms << "\n";
        std::cout << "SpydrPick: analysis completed\n      total " << total.lap() << " s\n";
    }
    for (auto* c : ctxs) spk_destroy(c);
    if (A.states) std::free(A.states);
    spk_host_free(flag_mem);
    return EXIT_SUCCESS;
}
