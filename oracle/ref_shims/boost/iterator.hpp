// TEST INFRASTRUCTURE ONLY -- stand-in for <boost/iterator.hpp> (the pre-C++17 std::iterator typedef bundle).
#pragma once
#include <cstddef>
#include <iterator>
namespace boost {
template <class Category, class T, class Distance = std::ptrdiff_t, class Pointer = T*, class Reference = T&>
struct iterator {
    using iterator_category = Category; using value_type = T; using difference_type = Distance; using pointer = Pointer; using reference = Reference;
};
}
