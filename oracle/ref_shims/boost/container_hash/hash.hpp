// TEST INFRASTRUCTURE ONLY -- boost::hash / hash_combine / hash_range (Boost's classic 0x9e3779b9 combiner).
#pragma once
#include <cstddef>
#include <functional>
#include <utility>
namespace boost {
template <class T> struct hash;
template <class T> inline void hash_combine(std::size_t& seed, const T& v) { seed ^= hash<T>()(v) + 0x9e3779b9 + (seed << 6) + (seed >> 2); }
template <class T> struct hash { std::size_t operator()(const T& v) const { return std::hash<T>()(v); } };
template <class A, class B> struct hash<std::pair<A, B>> {
    std::size_t operator()(const std::pair<A, B>& p) const { std::size_t s = 0; hash_combine(s, p.first); hash_combine(s, p.second); return s; }
};
template <class It> inline std::size_t hash_range(It first, It last) {
    std::size_t s = 0; for (; first != last; ++first) hash_combine(s, *first); return s;
}
}
