// TEST INFRASTRUCTURE (oracle/): stand-in for <mitsuba/core/fresolver.h> + the one Thread accessor
// array_from_file() touches (reference: include/mitsuba/render/sunsky/sunsky.h:517-518).
#pragma once
#include <mitsuba/core/filesystem.h>
#include <mitsuba/core/logger.h>
#include <string>
NAMESPACE_BEGIN(mitsuba)
class FileResolver {
public:
    static std::string &root() { static std::string r; return r; }
    fs::path resolve(const fs::path &p) const {
        if (p.is_absolute() || root().empty()) return p;
        return fs::path(root()) / p;
    }
};
class Thread {
public:
    static Thread *thread() { static Thread t; return &t; }
    FileResolver *file_resolver() { return &m_res; }
private:
    FileResolver m_res;
};
NAMESPACE_END(mitsuba)
