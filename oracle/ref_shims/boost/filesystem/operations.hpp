// TEST INFRASTRUCTURE ONLY -- boost::filesystem mapped onto std::filesystem.
#pragma once
#include <filesystem>
namespace boost { namespace filesystem = std::filesystem; }
