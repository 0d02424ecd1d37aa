// TEST INFRASTRUCTURE (oracle/): stand-in for <mitsuba/core/filesystem.h> on top of std::filesystem.
#pragma once
#include <mitsuba/core/fwd.h>
#include <filesystem>
#include <string>
NAMESPACE_BEGIN(mitsuba)
NAMESPACE_BEGIN(filesystem)
class path : public std::filesystem::path {
public:
    using Base = std::filesystem::path;
    path() { }
    path(const char *s) : Base(s) { }
    path(const std::string &s) : Base(s) { }
    path(const Base &b) : Base(b) { }
    path operator/(const path &o) const { return path((const Base &) *this / (const Base &) o); }
};
inline bool exists(const path &p) { return std::filesystem::exists((const std::filesystem::path &) p); }
NAMESPACE_END(filesystem)
NAMESPACE_END(mitsuba)
