// TEST INFRASTRUCTURE ONLY -- boost::iostreams::mapped_file / mapped_file_source on mmap(2).
#pragma once
#include <cstddef>
#include <fcntl.h>
#include <string>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
namespace boost { namespace iostreams {
class mapped_file {
public:
    enum mapmode { readonly = 1, readwrite = 2 };
    mapped_file() = default;
    mapped_file(const std::string& path, mapmode mode = readwrite) { open(path, mode); }
    mapped_file(const char* path, mapmode mode = readwrite) { open(path, mode); }
    ~mapped_file() { close(); }
    mapped_file(const mapped_file&) = delete; mapped_file& operator=(const mapped_file&) = delete;
    void open(const std::string& path, mapmode mode) {
        close();
        const int fd = ::open(path.c_str(), mode == readonly ? O_RDONLY : O_RDWR); if (fd < 0) return;
        struct stat st; if (fstat(fd, &st) != 0) { ::close(fd); return; }
        m_size = std::size_t(st.st_size);
        if (m_size) { void* p = mmap(nullptr, m_size, mode == readonly ? PROT_READ : (PROT_READ | PROT_WRITE), mode == readonly ? MAP_PRIVATE : MAP_SHARED, fd, 0);
                      if (p != MAP_FAILED) m_data = static_cast<char*>(p); else m_size = 0; }
        ::close(fd); m_open = true;
    }
    bool is_open() const { return m_open; }
    void close() { if (m_data) munmap(m_data, m_size); m_data = nullptr; m_size = 0; m_open = false; }
    std::size_t size() const { return m_size; }
    char* data() const { return m_data; }
    const char* const_data() const { return m_data; }
private:
    char* m_data = nullptr; std::size_t m_size = 0; bool m_open = false;
};
class mapped_file_source : public mapped_file {
public:
    mapped_file_source() = default;
    mapped_file_source(const char* path) : mapped_file(path, readonly) {}
    mapped_file_source(const std::string& path) : mapped_file(path, readonly) {}
};
}}
