// TEST INFRASTRUCTURE (oracle/): stand-in for <mitsuba/core/fstream.h>: little-endian binary reads
// and writes, which is all array_from_file()/array_to_file() need
// (reference: include/mitsuba/render/sunsky/sunsky.h:522-556, 576-596).
#pragma once
#include <mitsuba/core/filesystem.h>
#include <mitsuba/core/logger.h>
#include <cstdio>
NAMESPACE_BEGIN(mitsuba)
class FileStream {
public:
    enum EMode { ERead, EReadWrite, ETruncReadWrite };
    FileStream(const fs::path &p, EMode mode = ERead) {
        m_f = fopen(p.string().c_str(), mode == ERead ? "rb" : "wb");
        if (!m_f) Throw("FileStream: cannot open \"%s\"", p.string());
    }
    ~FileStream() { close(); }
    void read(void *dst, size_t n) { if (fread(dst, 1, n, m_f) != n) Throw("FileStream: short read"); }
    template <typename T> void read(T &v) { read(&v, sizeof(T)); }
    template <typename T> void read_array(T *dst, size_t n) { read((void *) dst, n * sizeof(T)); }
    void write(const void *src, size_t n) { if (fwrite(src, 1, n, m_f) != n) Throw("FileStream: short write"); }
    template <typename T> void write(const T &v) { write(&v, sizeof(T)); }
    template <typename T> void write_array(const T *src, size_t n) { write((const void *) src, n * sizeof(T)); }
    void close() { if (m_f) { fclose(m_f); m_f = nullptr; } }
private:
    FILE *m_f = nullptr;
};
NAMESPACE_END(mitsuba)
