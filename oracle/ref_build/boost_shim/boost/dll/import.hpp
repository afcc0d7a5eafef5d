#pragma once
#include <functional>
#include <stdexcept>
#include <string>
namespace boost { namespace dll {
namespace load_mode {
enum type { default_mode = 0, search_system_folders = 1, append_decorations = 2, rtld_global = 4 };
inline type operator|(type a, type b) { return (type) ((int) a | (int) b); }
}
template <class T>
std::function<T> import_alias(const std::string&, const std::string&, load_mode::type = load_mode::default_mode) {
    throw std::runtime_error("extensions are not available in the oracle build");
}
}}
