#pragma once
#include <boost/iostreams/stream.hpp>
