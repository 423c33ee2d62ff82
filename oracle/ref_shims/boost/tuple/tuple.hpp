// TEST INFRASTRUCTURE ONLY -- boost::tuple / boost::get / boost::make_tuple mapped onto std::tuple.
#pragma once
#include <tuple>
#include <utility>
namespace boost {
template <class... T> using tuple = std::tuple<T...>;
using std::get;
using std::make_tuple;
namespace tuples { using std::get; }
}
