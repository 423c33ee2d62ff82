#pragma once
#include "qi.hpp"
