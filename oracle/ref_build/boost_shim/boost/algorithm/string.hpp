// Stand-in for the few Boost.StringAlgo entry points the reference calls.
// Boost is a system dependency of the reference (SConstruct:660) and is absent
// from this image; nothing here does hot-path arithmetic. Test infrastructure only.
#pragma once
#include <algorithm>
#include <cctype>
#include <string>
#include <vector>
namespace boost {
namespace algorithm {
inline bool starts_with(const std::string& s, const std::string& p) {
    return s.size() >= p.size() && s.compare(0, p.size(), p) == 0;
}
inline bool ends_with(const std::string& s, const std::string& p) {
    return s.size() >= p.size() && s.compare(s.size() - p.size(), p.size(), p) == 0;
}
inline std::string to_lower_copy(std::string s) {
    std::transform(s.begin(), s.end(), s.begin(), [](unsigned char c) { return (char) std::tolower(c); });
    return s;
}
inline std::string to_upper_copy(std::string s) {
    std::transform(s.begin(), s.end(), s.begin(), [](unsigned char c) { return (char) std::toupper(c); });
    return s;
}
inline bool iequals(const std::string& a, const std::string& b) {
    return to_lower_copy(a) == to_lower_copy(b);
}
inline std::string trim_copy(const std::string& s) {
    auto issp = [](unsigned char c) { return std::isspace(c) != 0; };
    auto b = std::find_if_not(s.begin(), s.end(), issp);
    auto e = std::find_if_not(s.rbegin(), std::string::const_reverse_iterator(b), issp).base();
    return std::string(b, e);
}
inline std::string replace_all_copy(std::string s, const std::string& from, const std::string& to) {
    if (from.empty()) return s;
    for (size_t pos = 0; (pos = s.find(from, pos)) != std::string::npos; pos += to.size()) {
        s.replace(pos, from.size(), to);
    }
    return s;
}
template <class Seq>
std::string join(const Seq& items, const std::string& sep) {
    std::string out;
    bool first = true;
    for (const auto& it : items) {
        if (!first) out += sep;
        out += it;
        first = false;
    }
    return out;
}
struct is_space {
    bool operator()(char c) const { return std::isspace((unsigned char) c) != 0; }
};
struct is_any_of {
    std::string set;
    is_any_of(const std::string& s) : set(s) {}
    bool operator()(char c) const { return set.find(c) != std::string::npos; }
};
enum token_compress_mode_type { token_compress_off, token_compress_on };
template <class Pred>
std::vector<std::string>& split(std::vector<std::string>& out, const std::string& in, Pred pred,
                                token_compress_mode_type mode = token_compress_off) {
    out.clear();
    std::string cur;
    bool prevDelim = false;
    for (char c : in) {
        if (pred(c)) {
            if (!(mode == token_compress_on && prevDelim)) {
                out.push_back(cur);
                cur.clear();
            }
            prevDelim = true;
        } else {
            cur.push_back(c);
            prevDelim = false;
        }
    }
    out.push_back(cur);
    return out;
}
// ifind_last must give back a range with the constness of its argument and an
// explicit bool conversion (used at src/kinetics/KineticsFactory.cpp:174-178).
template <class It>
struct find_result {
    It first, last;
    bool found;
    explicit operator bool() const { return found; }
    It begin() const { return first; }
    It end() const { return last; }
};
template <class S>
auto ifind_last(S& s, const std::string& pat) -> find_result<decltype(s.begin())> {
    std::string ls = to_lower_copy(std::string(s)), lp = to_lower_copy(pat);
    size_t pos = ls.rfind(lp);
    if (pos == std::string::npos) return {s.end(), s.end(), false};
    return {s.begin() + pos, s.begin() + pos + pat.size(), true};
}
} // namespace algorithm
using algorithm::ends_with; using algorithm::ifind_last; using algorithm::iequals;
using algorithm::is_any_of; using algorithm::is_space; using algorithm::join;
using algorithm::replace_all_copy; using algorithm::split; using algorithm::starts_with;
using algorithm::to_lower_copy; using algorithm::to_upper_copy; using algorithm::token_compress_on;
using algorithm::trim_copy;
} // namespace boost
