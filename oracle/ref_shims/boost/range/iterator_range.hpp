#pragma once
#include "iterator_range_core.hpp"
