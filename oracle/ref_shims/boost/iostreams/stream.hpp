// TEST INFRASTRUCTURE ONLY -- compressed input (.gz / .xz) needs zlib / liblzma through Boost.Iostreams; this build has neither,
// so the reference's compressed-FASTA branch compiles against these types and fails at run time with a clear message.
#pragma once
#include <cstdio>
#include <cstdlib>
#include <istream>
#include <sstream>
#include "device/mapped_file.hpp"
namespace boost { namespace iostreams {
template <class Device> class stream : public std::istringstream { public: explicit stream(Device&) {} };
struct gzip_decompressor {}; struct lzma_decompressor {};
class filtering_istream : public std::istringstream {
public:
    template <class T> void push(const T&) {}
    void push(const gzip_decompressor&) { unsupported(); } void push(const lzma_decompressor&) { unsupported(); }
private:
    [[noreturn]] static void unsupported() { std::fputs("reference build without Boost.Iostreams: compressed FASTA input is unavailable\n", stderr); std::exit(EXIT_FAILURE); }
};
}}
