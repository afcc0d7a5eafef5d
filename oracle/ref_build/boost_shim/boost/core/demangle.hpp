#pragma once
#include <cstdlib>
#include <cxxabi.h>
#include <string>
namespace boost { namespace core {
inline std::string demangle(const char* name) {
    int status = 0;
    char* d = abi::__cxa_demangle(name, nullptr, nullptr, &status);
    std::string out = (status == 0 && d) ? d : name;
    std::free(d);
    return out;
}
}}
