#pragma once
#include "../../container_hash/hash.hpp"
