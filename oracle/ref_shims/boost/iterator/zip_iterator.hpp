// TEST INFRASTRUCTURE ONLY -- stand-in for boost::zip_iterator over a std::tuple of iterators; dereferences to a tuple of references.
#pragma once
#include <cstddef>
#include <iterator>
#include <tuple>
#include <utility>
#include "../tuple/tuple.hpp"
namespace boost {
template <class Tuple> class zip_iterator;
template <class... Its> class zip_iterator<std::tuple<Its...>> {
public:
    using iterator_category = std::forward_iterator_tag;
    using reference = std::tuple<typename std::iterator_traits<Its>::reference...>;
    using value_type = reference; using difference_type = std::ptrdiff_t; using pointer = void;
    zip_iterator() = default;
    explicit zip_iterator(const std::tuple<Its...>& its) : m_its(its) {}
    reference operator*() const { return deref(std::index_sequence_for<Its...>{}); }
    zip_iterator& operator++() { std::apply([](auto&... it) { (++it, ...); }, m_its); return *this; }
    zip_iterator operator++(int) { auto t = *this; ++*this; return t; }
    friend bool operator==(const zip_iterator& a, const zip_iterator& b) { return std::get<0>(a.m_its) == std::get<0>(b.m_its); }
    friend bool operator!=(const zip_iterator& a, const zip_iterator& b) { return !(a == b); }
    const std::tuple<Its...>& get_iterator_tuple() const { return m_its; }
private:
    template <std::size_t... K> reference deref(std::index_sequence<K...>) const { return reference(*std::get<K>(m_its)...); }
    std::tuple<Its...> m_its;
};
template <class... Its> zip_iterator<std::tuple<Its...>> make_zip_iterator(const std::tuple<Its...>& t) { return zip_iterator<std::tuple<Its...>>(t); }
}
