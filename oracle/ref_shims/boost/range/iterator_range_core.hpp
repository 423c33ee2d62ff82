// TEST INFRASTRUCTURE ONLY -- stand-in for boost::iterator_range / make_iterator_range.
#pragma once
#include <cstddef>
#include <iterator>
namespace boost {
template <class It> class iterator_range {
public:
    using iterator = It; using const_iterator = It;
    using value_type = typename std::iterator_traits<It>::value_type;
    using reference = typename std::iterator_traits<It>::reference;
    using difference_type = typename std::iterator_traits<It>::difference_type;
    using size_type = std::size_t;
    iterator_range() = default;
    iterator_range(It b, It e) : m_b(b), m_e(e) {}
    template <class R> iterator_range(R& r) : m_b(std::begin(r)), m_e(std::end(r)) {}
    It begin() const { return m_b; }
    It end() const { return m_e; }
    bool empty() const { return m_b == m_e; }
    size_type size() const { return size_type(std::distance(m_b, m_e)); }
    reference front() const { return *m_b; }
    reference operator[](difference_type k) const { return m_b[k]; }
    void advance_begin(difference_type k) { std::advance(m_b, k); }
private:
    It m_b{}, m_e{};
};
template <class It> It begin(const iterator_range<It>& r) { return r.begin(); }
template <class It> It end(const iterator_range<It>& r) { return r.end(); }
template <class It> It cbegin(const iterator_range<It>& r) { return r.begin(); }
template <class It> It cend(const iterator_range<It>& r) { return r.end(); }
template <class It> iterator_range<It> make_iterator_range(It b, It e) { return iterator_range<It>(b, e); }
template <class R> auto make_iterator_range(R& r) -> iterator_range<decltype(std::begin(r))> { return {std::begin(r), std::end(r)}; }
}
