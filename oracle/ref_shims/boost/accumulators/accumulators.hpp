// TEST INFRASTRUCTURE ONLY -- the slice of Boost.Accumulators the reference's hot path uses (accumulator_set<T, stats<tag::max>>,
// include/mi.hpp:245-290) plus a plain histogram standing in for apegrunt's distribution accumulators
// (externals/apegrunt/include/accumulators/*.hpp, diagnostics only: --output-state-frequencies), whose own headers are
// written against Boost's accumulator framework and are therefore switched off through their include guards.
#pragma once
#define APEGRUNT_ACCUMULATORS_DISTRIBUTION_HPP
#define APEGRUNT_ACCUMULATORS_DISTRIBUTION_BIN_HPP
#define APEGRUNT_ACCUMULATORS_DISTRIBUTION_BINCOUNT_HPP
#define APEGRUNT_ACCUMULATORS_DISTRIBUTION_COUNT_HPP
#define APEGRUNT_ACCUMULATORS_DISTRIBUTION_CUMULATIVE_HPP
#define APEGRUNT_ACCMULATORS_DISTRIBUTION_EXTRACTOR_HPP
#define APEGRUNT_ACCMULATORS_DISTRIBUTION_GENERATOR_BASE_HPP
#define APEGRUNT_ACCMULATORS_DISTRIBUTION_GENERATOR_CSV_HPP
#define APEGRUNT_ACCUMULATORS_DISTRIBUTION_MEAN_HPP
#define APEGRUNT_ACCUMULATORS_DISTRIBUTION_NORMALIZED_HPP
#define APEGRUNT_ACCUMULATORS_DISTRIBUTION_ORDERED_HPP
#define APEGRUNT_ACCUMULATORS_DISTRIBUTION_STD_HPP
#define APEGRUNT_ACCUMULATORS_DISTRIBUTION_VARIANCE_HPP
#define APEGRUNT_ACCUMULATORS_STD_HPP
#include <cmath>
#include <cstddef>
#include <limits>
#include <map>
#include <ostream>
namespace boost { namespace accumulators {
struct from_distribution {};
namespace tag {
struct max {}; struct count {}; struct mean {}; struct variance {}; struct std {}; struct distribution_bincount {};
struct distribution {
    struct binwidth_arg { double value; };
    struct binwidth_keyword { binwidth_arg operator=(double v) const { return binwidth_arg{v}; } };
    static constexpr binwidth_keyword binwidth{};
};
}
template <class... Features> struct stats {};
template <class Sample, class Stats, class Weight = void> class accumulator_set {
public:
    accumulator_set() = default;
    accumulator_set(tag::distribution::binwidth_arg a) : m_binwidth(a.value) {}
    void operator()(const Sample& v) {
        ++m_count; if (m_max < v) m_max = v;
        if (m_binwidth > 0) ++m_bins[(long long)std::floor(double(v) / m_binwidth)];
    }
    Sample max_value() const { return m_max; }
    std::size_t count_value() const { return m_count; }
    double binwidth() const { return m_binwidth; }
    const std::map<long long, std::size_t>& bins() const { return m_bins; }
private:
    Sample m_max = -(std::numeric_limits<Sample>::max)();      // Boost: numeric::as_min() of a floating-point sample
    std::size_t m_count = 0; double m_binwidth = 0; std::map<long long, std::size_t> m_bins;
};
template <class S, class St, class W> S max(const accumulator_set<S, St, W>& a) { return a.max_value(); }
template <class S, class St, class W> std::size_t count(const accumulator_set<S, St, W>& a) { return a.count_value(); }
template <class Set> struct distribution_view { const Set* set; };
template <class S, class St, class W> distribution_view<accumulator_set<S, St, W>> distribution(const accumulator_set<S, St, W>& a) { return {&a}; }
}}
namespace apegrunt { namespace accumulators {
template <class View> struct csv_generator { View view; };
template <class View> csv_generator<View> csv(View v) { return {v}; }
template <class Set> std::ostream& operator<<(std::ostream& os, const csv_generator<boost::accumulators::distribution_view<Set>>& g) {
    for (const auto& b : g.view.set->bins()) os << double(b.first) * g.view.set->binwidth() << "," << b.second << "\n";
    return os;
}
}}
