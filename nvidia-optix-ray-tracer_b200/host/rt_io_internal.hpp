// rt_io_internal.hpp — helpers shared by the translation units of librtb200_io.so.
#pragma once
#include <cstdarg>
#include <cstdio>
#include <filesystem>
#include <fstream>
#include <sstream>
#include <stdexcept>
#include <string>

#include "../../include/rtb200_io.h"

namespace fs = std::filesystem;

namespace rtio {

inline std::string &tl_err() {
    static thread_local std::string e;
    return e;
}

inline int fail(int code, const char *fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    tl_err() = buf;
    return code;
}

struct IoError {
    int code;
    std::string msg;
};
[[noreturn]] inline void raise(int code, const std::string &msg) { throw IoError{code, msg}; }

inline bool read_file(const std::string &path, std::string &out) {
    std::ifstream f(path, std::ios::in | std::ios::binary);
    if (!f) return false;
    std::ostringstream ss;
    ss << f.rdbuf();
    out = ss.str();
    return true;
}

// every entry point funnels through this so that no C++ exception crosses the C ABI
template <typename F>
int guarded(F &&f) {
    try {
        f();
        return RT_IO_OK;
    } catch (const IoError &e) {
        return fail(e.code, "%s", e.msg.c_str());
    } catch (const std::exception &e) {
        return fail(RT_IO_ERR_PARSE, "%s", e.what());
    }
}

}  // namespace rtio

using rtio::fail;
using rtio::guarded;
using rtio::raise;
using rtio::read_file;
