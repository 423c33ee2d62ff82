#pragma once
#include "tbb_stddef.h"
