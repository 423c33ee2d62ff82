// TEST INFRASTRUCTURE ONLY -- force-included (-include) in front of every reference translation unit by oracle/Makefile.ref:
// brings in the stand-ins whose include guards must be defined before the reference reaches its own Spirit grammars
// (boost/spirit/include/qi.hpp) and Boost.Accumulators-based distribution headers (boost/accumulators/accumulators.hpp).
#pragma once
#include <iostream>
#include <set>
#include <boost/tuple/tuple.hpp>
#include <boost/accumulators/accumulators.hpp>
#include <boost/spirit/include/qi.hpp>
