#pragma once
#include <ostream>
#include <string>
namespace boost { namespace stacktrace {
struct stacktrace { stacktrace() {} stacktrace(size_t, size_t) {} };
inline std::string to_string(const stacktrace&) { return "<no stacktrace in oracle build>\n"; }
inline std::ostream& operator<<(std::ostream& o, const stacktrace& s) { return o << to_string(s); }
}}
