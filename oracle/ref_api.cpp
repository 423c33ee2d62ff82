// ref_api.cpp -- TEST INFRASTRUCTURE ONLY. A C entry-point layer over the REAL reference code, compiled from the sources where
// they lie under /root/reference by oracle/Makefile.ref into oracle/_ref/libspkref.so (never part of the product; only tests/,
// __graft_entry__.smoke() and bench.py's CPU legs may load it). Nothing of the reference is restated here: every function
// below instantiates and calls the reference's own templates --
//   apegrunt::parse_Alignment / Alignment_filter / cache_sample_weights / calculate_weights   (externals/apegrunt/include/apegrunt)
//   spydrpick::MI_solver::single, get_MI_network, determine_MI_threshold, mutual_information_block_kernel::normalize_and_get_mi_single
//                                                                                               (include/mi.hpp, include/SpydrPick.hpp)
//   apegrunt::Graph::sort, aracne::run_ARACNE, Graph_output_formatter, Outlier_Graph_formatter   (graph/Graph.hpp, include/ARACNE.hpp)
// -- with the option singletons filled by the reference's own option classes from an argv (ref_configure), as its main() does
// (src/SpydrPick.cpp:65-116). Boost / TBB are the stand-ins under oracle/ref_shims (see their headers).
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iomanip>
#include <iostream>
#include <memory>
#include <sstream>
#include <string>
#include <vector>

#include "boost/program_options.hpp"
#include "boost/filesystem/operations.hpp"

#include "apegrunt/Apegrunt.h"
#include "apegrunt/Apegrunt_IO_misc.hpp"
#include "apegrunt/Loci_generators.hpp"
#include "apegrunt/ValueVector_parser.hpp"
#include "graph/Graph_utility.hpp"
#include "misc/Stopwatch.hpp"

#include "SpydrPick.h"
#include "SpydrPick.hpp"
#include "ARACNE.h"

namespace {
using state_t = apegrunt::nucleic_acid_state_t;
using real_t = double;
using alignment_t = apegrunt::Alignment_impl_block_compressed_storage<apegrunt::StateVector_impl_block_compressed_alignment_storage<state_t>>;
using alignment_ptr = apegrunt::Alignment_ptr<state_t>;

struct RefAlignment { alignment_ptr a; };
struct RefNetwork { spydrpick::MI_network<real_t> net; };

std::ostringstream g_log;                      // the reference's verbose chatter lands here (ref_log)
tbb::task_scheduler_init* g_tbb = nullptr;
const char STATE_CHARS[5] = {'A', 'T', 'C', 'G', '-'};   // externals/apegrunt/src/StateVector_state_types.cpp:33-44
}

extern "C" {

/* Fill the reference's option singletons from an argv, exactly as its main() does (src/SpydrPick.cpp:65-116).
 * argv[0] is skipped; an alignment file name must be present (any string: nothing is opened here). Returns 0 / -1. */
int ref_configure(int argc, const char** argv) {
    namespace po = boost::program_options;
    using namespace spydrpick;
    try {
        po::options_description all_options;
        SpydrPick_options spydrpick_options(&g_log, &std::cerr); spydrpick_options.AddOptions(&all_options);
        apegrunt::Apegrunt_options apegrunt_options(&g_log, &std::cerr); apegrunt_options.AddOptions(&all_options);
        aracne::ARACNE_options aracne_options(&g_log, &std::cerr); aracne_options.AddOptions(&all_options);
        po::variables_map options_map;
        po::positional_options_description popt; popt.add("alignmentfile", -1);
        po::store(po::command_line_parser(argc, argv).options(all_options).positional(popt).run(), options_map);
        po::notify(options_map);
        if (!spydrpick_options.CheckOptions(&options_map)) return -1;
        if (!g_tbb) g_tbb = new tbb::task_scheduler_init(tbb::task_scheduler_init::deferred);
        SpydrPick_options::set_threads(SpydrPick_options::threads() > 0 ? SpydrPick_options::threads() : g_tbb->default_num_threads());
        apegrunt_options.set_verbose(SpydrPick_options::verbose());
        if (!apegrunt_options.CheckOptions(&options_map)) return -1;
        apegrunt_options.set_threads(SpydrPick_options::threads());
        aracne_options.set_verbose(SpydrPick_options::verbose());
        if (!aracne_options.CheckOptions(&options_map)) return -1;
        aracne_options.set_threads(SpydrPick_options::threads());
        g_tbb->initialize(SpydrPick_options::threads());
    } catch (std::exception& e) { std::cerr << "ref_configure: " << e.what() << "\n"; return -1; }
    return 0;
}
int ref_threads(void) { return spydrpick::SpydrPick_options::threads(); }

/* the reference's verbose output since the last call (copied into buf, NUL-terminated); returns the full length */
size_t ref_log(char* buf, size_t cap) {
    const std::string s = g_log.str(); g_log.str(std::string());
    if (buf && cap) { const size_t k = std::min(cap - 1, s.size()); std::memcpy(buf, s.data(), k); buf[k] = 0; }
    return s.size();
}

/* dense states (codes a=0,t=1,c=2,g=3,GAP=4, row-major [n][L]) -> FASTA text -> the reference's parse_Alignment(std::istream*)
 * (apegrunt/Alignment_parsers.hpp:49-70) -> block-compressed Alignment. id = alignment id string (the file stem in main()). */
void* ref_alignment_create(const uint8_t* states, uint32_t n, uint64_t L, const char* id) {
    std::string fasta; fasta.reserve((size_t)n * (L + 16));
    for (uint32_t s = 0; s < n; ++s) {
        fasta += ">s" + std::to_string(s) + "\n";
        const uint8_t* row = states + (size_t)s * L;
        for (uint64_t c = 0; c < L; ++c) fasta.push_back(STATE_CHARS[row[c] < 5 ? row[c] : 4]);
        fasta.push_back('\n');
    }
    std::istringstream in(fasta);
    auto a = apegrunt::parse_Alignment<alignment_t>(&in);
    if (!a) return nullptr;
    if (id) a->set_id_string(id);
    return new RefAlignment{a};
}
void* ref_alignment_from_file(const char* path) {          // parse_Alignment(filename), Alignment_parsers.hpp:72-144
    auto a = apegrunt::parse_Alignment<alignment_t>(std::string(path));
    return a ? new RefAlignment{a} : nullptr;
}
void ref_alignment_free(void* h) { delete static_cast<RefAlignment*>(h); }
uint64_t ref_alignment_nloci(void* h) { return static_cast<RefAlignment*>(h)->a->n_loci(); }
uint32_t ref_alignment_size(void* h) { return (uint32_t) static_cast<RefAlignment*>(h)->a->size(); }
uint64_t ref_alignment_n_original_positions(void* h) { return static_cast<RefAlignment*>(h)->a->n_original_positions(); }
size_t ref_alignment_id(void* h, char* buf, size_t cap) {
    const std::string s = static_cast<RefAlignment*>(h)->a->id_string();
    if (buf && cap) { const size_t k = std::min(cap - 1, s.size()); std::memcpy(buf, s.data(), k); buf[k] = 0; }
    return s.size();
}
/* loci translation (original position of every column), as get_alignments() attaches a mapping list (Alignment_utility.hpp:128-160) */
void ref_alignment_set_mapping(void* h, const uint64_t* pos, uint64_t L) {
    std::vector<std::size_t> v(pos, pos + L);
    static_cast<RefAlignment*>(h)->a->set_loci_translation(apegrunt::make_Loci_list(std::move(v), 0));
}
void ref_alignment_get_mapping(void* h, uint64_t* pos) {
    auto tr = static_cast<RefAlignment*>(h)->a->get_loci_translation(); size_t k = 0;
    for (auto p : tr) pos[k++] = p;
}
/* dense states back out of the reference's storage (sequence iteration, Alignment_generators.hpp:45-58) */
void ref_alignment_states(void* h, uint8_t* states) {
    auto a = static_cast<RefAlignment*>(h)->a; const size_t L = a->n_loci(); size_t s = 0;
    for (const auto& sequence : a) { size_t c = 0; for (const auto& state : sequence) { if (c < L) states[s * L + c] = (uint8_t) std::size_t(state); ++c; } ++s; }
}

/* column filter exactly as main() applies it (src/SpydrPick.cpp:234-297): Alignment_filter(AQUIRE_GLOBAL).get_filter_list +
 * alignment->subset(). Returns a NEW alignment handle (or the same columns if nothing is dropped); *n_kept = columns kept. */
void* ref_alignment_filter(void* h, uint64_t* n_kept) {
    auto a = static_cast<RefAlignment*>(h)->a;
    auto alignment_filter = apegrunt::Alignment_filter(apegrunt::Alignment_filter::ParameterPolicy::AQUIRE_GLOBAL);
    const auto filter_list = alignment_filter.get_filter_list(a);
    if (n_kept) *n_kept = filter_list->size();
    if (filter_list->size() == 0) return nullptr;
    if (filter_list->size() != a->n_loci()) a = a->subset(filter_list, nullptr);
    return new RefAlignment{a};
}
/* include-list / exclude-list / sample-list steps of main() (src/SpydrPick.cpp:191-231, 299-318) on a loci list given as indices */
void* ref_alignment_subset(void* h, const uint64_t* loci, uint64_t k) {
    std::vector<std::size_t> v(loci, loci + k);
    auto a = static_cast<RefAlignment*>(h)->a->subset(apegrunt::make_Loci_list(std::move(v), 0));
    return a ? new RefAlignment{a} : nullptr;
}
void* ref_alignment_subsample(void* h, const uint64_t* samples, uint64_t k) {
    std::vector<std::size_t> v(samples, samples + k);
    auto a = static_cast<RefAlignment*>(h)->a->subsample(apegrunt::make_Loci_list(std::move(v), 0), nullptr);
    return a ? new RefAlignment{a} : nullptr;
}

/* apegrunt::cache_sample_weights (Alignment_utility.hpp:186-258; option singletons decide: none / file / computed, rescale),
 * then the per-sequence weights are read back. */
void ref_cache_sample_weights(void* h, double* w) {
    auto a = static_cast<RefAlignment*>(h)->a;
    apegrunt::cache_sample_weights(a);
    size_t s = 0; for (const auto& sequence : a) w[s++] = sequence->weight();
}
/* apegrunt::calculate_weights alone (Alignment_StateVector_weights.hpp:192-353): w = multiplicity / integer weight */
void ref_calculate_weights(void* h, double* w) {
    auto a = static_cast<RefAlignment*>(h)->a;
    const auto v = apegrunt::calculate_weights(a);
    for (size_t s = 0; s < v.size(); ++s) w[s] = v[s];
}
void ref_set_weights(void* h, const double* w) {
    auto a = static_cast<RefAlignment*>(h)->a; size_t s = 0;
    for (auto& sequence : a) sequence->set_weight(w[s++]);
}
double ref_effective_size(void* h) { return static_cast<RefAlignment*>(h)->a->effective_size(); }

/* MI_solver::single over a pair list (include/mi.hpp:367-372, 183-224), serially, as determine_MI_threshold's body does */
void ref_mi_pairs(void* h, const uint64_t* pi, const uint64_t* pj, uint64_t npairs, double* mi) {
    auto a = static_cast<RefAlignment*>(h)->a;
    auto solver = spydrpick::get_MI_solver(std::vector<alignment_ptr>{a}, 0.0, spydrpick::SpydrPick_options::get_mi_pseudocount());
    for (uint64_t k = 0; k < npairs; ++k) mi[k] = solver.single(pi[k], pj[k]);
}
/* the raw weighted 5x5 joint-count table of one column pair as Weighted_crosstable_2Dblock::single builds it
 * (misc/Crosstable.hpp:429-481; table[jstate][istate]) */
void ref_xtable_single(void* h, uint64_t ipos, uint64_t jpos, double* table25) {
    auto a = static_cast<RefAlignment*>(h)->a;
    spydrpick::MI_Parameters<state_t, real_t> params(std::vector<alignment_ptr>{a}, 0.0, spydrpick::SpydrPick_options::get_mi_pseudocount());
    apegrunt::Weighted_crosstable_2Dblock<state_t, real_t> kernel(a, params.get_weights(), 0.0);
    kernel.single(table25, ipos, jpos);
}
/* spydrpick::determine_MI_threshold (include/SpydrPick.hpp:285-343); clock-seeded, so only statistically reproducible */
double ref_determine_mi_threshold(void* h, uint64_t n_values) {
    return spydrpick::determine_MI_threshold<real_t>(static_cast<RefAlignment*>(h)->a, n_values);
}

/* mutual_information_block_kernel::normalize_and_get_mi_single (include/mi.hpp:146-181) on caller-supplied 5x5 tables:
 * tables[k][j][i] (row j = state of column j, as misc/Crosstable.hpp fills it) are MODIFIED in place like the reference does;
 * ipresence / jpresence are the 5-bit state-presence masks; pseudocount is taken from ref_configure's --mi-pseudocount.
 * `h` is any alignment (the kernel object needs one to exist; the tables do not come from it). */
void ref_mi_from_tables(void* h, double* tables, const uint8_t* ipresence, const uint8_t* jpresence, uint64_t ntables, double* mi) {
    auto a = static_cast<RefAlignment*>(h)->a;
    using kernel_t = spydrpick::mutual_information_block_kernel<state_t, real_t>;
    auto weights = std::make_shared<std::vector<real_t>>(a->size(), 1.0);
    kernel_t kernel(a, weights, 0.0, spydrpick::SpydrPick_options::get_mi_pseudocount());
    constexpr std::size_t N = kernel_t::N;
    for (uint64_t k = 0; k < ntables; ++k)
        kernel.normalize_and_get_mi_single(tables + k * N * N, mi + k, typename kernel_t::statepresence_t(ipresence[k]), typename kernel_t::statepresence_t(jpresence[k]));
}

/* spydrpick::get_MI_network (include/SpydrPick.hpp:132-168): the whole scan, threshold strict, options from the singletons */
void* ref_scan(void* h, double mi_threshold) {
    std::vector<alignment_ptr> alignments{static_cast<RefAlignment*>(h)->a};
    auto* r = new RefNetwork{spydrpick::get_MI_network(alignments, mi_threshold)};
    return r;
}
void ref_network_free(void* nh) { delete static_cast<RefNetwork*>(nh); }
uint64_t ref_network_size(void* nh, int wo_gaps) { auto* r = static_cast<RefNetwork*>(nh); return (wo_gaps ? r->net.network_wo_gaps : r->net.network)->size(); }
void ref_network_thresholds(void* nh, double* outlier, double* extreme) { auto* r = static_cast<RefNetwork*>(nh); *outlier = r->net.outlier_threshold; *extreme = r->net.extreme_outlier_threshold; }
/* edges in storage order: node1 < node2, weight, and the flag bit (bool(edge), set on DIRECT edges by ARACNE) */
void ref_network_edges(void* nh, int wo_gaps, uint32_t* n1, uint32_t* n2, double* w, uint8_t* flag) {
    auto* r = static_cast<RefNetwork*>(nh); const auto& g = *(wo_gaps ? r->net.network_wo_gaps : r->net.network); size_t k = 0;
    for (const auto& e : g) { n1[k] = e.node1(); n2[k] = e.node2(); w[k] = e.weight(); if (flag) flag[k] = bool(e) ? 1 : 0; ++k; }
}
void ref_network_sort(void* nh) { static_cast<RefNetwork*>(nh)->net.network->sort(); }                 /* Graph::sort, graph/Graph.hpp:234-242 */
void ref_network_aracne(void* nh) { aracne::run_ARACNE(static_cast<RefNetwork*>(nh)->net.network); }  /* include/ARACNE.hpp:411-556 */

/* a network from a caller-supplied edge list (for ARACNE / sort on arbitrary graphs) */
void* ref_network_from_edges(const uint32_t* n1, const uint32_t* n2, const double* w, uint64_t ne) {
    auto* r = new RefNetwork{};
    r->net.network = apegrunt::make_Graph_ptr<apegrunt::Graph>(); r->net.network_wo_gaps = apegrunt::make_Graph_ptr<apegrunt::Graph>();
    r->net.outlier_threshold = r->net.extreme_outlier_threshold = 0;
    r->net.network->reserve(ne);
    for (uint64_t k = 0; k < ne; ++k) r->net.network->add(n1[k], n2[k], w[k]);
    return r;
}

/* the two listings as main() writes them (src/SpydrPick.cpp:424-480): Graph_output_formatter / Outlier_Graph_formatter */
int ref_write_edges(void* nh, void* h, const char* path) {
    auto* r = static_cast<RefNetwork*>(nh); auto a = static_cast<RefAlignment*>(h)->a;
    std::ofstream os(path, std::ios::binary); if (!os) return -1;
    if (apegrunt::Apegrunt_options::linear_genome()) os << apegrunt::Graph_output_formatter<state_t, apegrunt::LinearDistance>(r->net.network, a);
    else os << apegrunt::Graph_output_formatter<state_t, apegrunt::CircularDistance>(r->net.network, a);
    return os.good() ? 0 : -1;
}
int ref_write_outliers(void* nh, void* h, const char* path) {
    auto* r = static_cast<RefNetwork*>(nh); auto a = static_cast<RefAlignment*>(h)->a;
    std::ofstream os(path, std::ios::binary); if (!os) return -1;
    if (apegrunt::Apegrunt_options::linear_genome()) os << spydrpick::Outlier_Graph_formatter<real_t, apegrunt::LinearDistance>(r->net, a->get_loci_translation(), a->n_original_positions());
    else os << spydrpick::Outlier_Graph_formatter<real_t, apegrunt::CircularDistance>(r->net, a->get_loci_translation(), a->n_original_positions());
    return os.good() ? 0 : -1;
}

const char* ref_version(void) { return "santeripuranen/SpydrPick f6bc110 + apegrunt 5c61810, compiled in place with oracle/ref_shims"; }

} // extern "C"
