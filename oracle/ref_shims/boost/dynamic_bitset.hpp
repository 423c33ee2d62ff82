// TEST INFRASTRUCTURE ONLY -- the part of boost::dynamic_bitset the reference touches (find_first / find_next / npos).
#pragma once
#include <cstddef>
#include <vector>
namespace boost {
template <class Block = unsigned long> class dynamic_bitset {
public:
    using size_type = std::size_t; static const size_type npos = size_type(-1);
    dynamic_bitset() = default; explicit dynamic_bitset(size_type n, bool v = false) : m_bits(n, v) {}
    size_type size() const { return m_bits.size(); }
    void resize(size_type n, bool v = false) { m_bits.resize(n, v); }
    dynamic_bitset& set(size_type k, bool v = true) { m_bits[k] = v; return *this; }
    bool test(size_type k) const { return m_bits[k]; }
    bool operator[](size_type k) const { return m_bits[k]; }
    size_type count() const { size_type c = 0; for (bool b : m_bits) c += b; return c; }
    size_type find_first() const { return find_from(0); }
    size_type find_next(size_type pos) const { return find_from(pos + 1); }
private:
    size_type find_from(size_type k) const { for (; k < m_bits.size(); ++k) if (m_bits[k]) return k; return npos; }
    std::vector<bool> m_bits;
};
template <class Block> const typename dynamic_bitset<Block>::size_type dynamic_bitset<Block>::npos;
}
