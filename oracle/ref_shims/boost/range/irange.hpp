// TEST INFRASTRUCTURE ONLY -- stand-in for boost::irange(first, last).
#pragma once
#include <cstddef>
#include <iterator>
#include "iterator_range_core.hpp"
namespace boost {
template <class I> class integer_iterator {
public:
    using iterator_category = std::random_access_iterator_tag; using value_type = I; using difference_type = std::ptrdiff_t;
    using pointer = const I*; using reference = I;
    integer_iterator() = default; explicit integer_iterator(I v) : m_v(v) {}
    I operator*() const { return m_v; }
    I operator[](difference_type k) const { return I(m_v + k); }
    integer_iterator& operator++() { ++m_v; return *this; } integer_iterator operator++(int) { auto t = *this; ++m_v; return t; }
    integer_iterator& operator--() { --m_v; return *this; } integer_iterator operator--(int) { auto t = *this; --m_v; return t; }
    integer_iterator& operator+=(difference_type k) { m_v = I(m_v + k); return *this; }
    integer_iterator& operator-=(difference_type k) { m_v = I(m_v - k); return *this; }
    friend integer_iterator operator+(integer_iterator a, difference_type k) { a += k; return a; }
    friend integer_iterator operator+(difference_type k, integer_iterator a) { a += k; return a; }
    friend integer_iterator operator-(integer_iterator a, difference_type k) { a -= k; return a; }
    friend difference_type operator-(const integer_iterator& a, const integer_iterator& b) { return difference_type(a.m_v) - difference_type(b.m_v); }
    friend bool operator==(const integer_iterator& a, const integer_iterator& b) { return a.m_v == b.m_v; }
    friend bool operator!=(const integer_iterator& a, const integer_iterator& b) { return a.m_v != b.m_v; }
    friend bool operator<(const integer_iterator& a, const integer_iterator& b) { return a.m_v < b.m_v; }
    friend bool operator>(const integer_iterator& a, const integer_iterator& b) { return a.m_v > b.m_v; }
    friend bool operator<=(const integer_iterator& a, const integer_iterator& b) { return a.m_v <= b.m_v; }
    friend bool operator>=(const integer_iterator& a, const integer_iterator& b) { return a.m_v >= b.m_v; }
private:
    I m_v{};
};
template <class I> iterator_range<integer_iterator<I>> irange(I first, I last) { return {integer_iterator<I>(first), integer_iterator<I>(last < first ? first : last)}; }
}
