// TEST INFRASTRUCTURE (oracle/): stand-in for <mitsuba/core/logger.h> so that the
// reference's header-only math and src/emitters/sunsky.cpp can be compiled in place
// without libmitsuba. Only the macros the sunsky path uses are provided
// (reference interface: include/mitsuba/core/logger.h:147-182).
#pragma once
#include <mitsuba/core/string.h>
#include <tinyformat.h>
#include <stdexcept>
#include <cstdio>

NAMESPACE_BEGIN(mitsuba)
enum LogLevel : int { Trace = 0, Debug = 100, Info = 200, Warn = 300, Error = 400 };

NAMESPACE_BEGIN(detail)
inline int &oracle_warn_count() { static int c = 0; return c; }
template <typename... Args>
inline void oracle_log(LogLevel level, const char *file, int line, Args &&...args) {
    std::string msg = tfm::format(std::forward<Args>(args)...);
    if (level >= Error)
        throw std::runtime_error(msg);
    if (level >= Warn) {
        oracle_warn_count()++;
        if (getenv("SSK_ORACLE_VERBOSE"))
            fprintf(stderr, "[ref warn] %s:%d %s\n", file, line, msg.c_str());
    }
}
NAMESPACE_END(detail)
NAMESPACE_END(mitsuba)

#define Log(level, ...) ::mitsuba::detail::oracle_log(level, __FILE__, __LINE__, ##__VA_ARGS__)
#define Throw(...) throw std::runtime_error(tfm::format(__VA_ARGS__))
#define Assert(...) ((void) 0)
#define NotImplementedError(Name) Throw("%s(): not implemented!", Name)
