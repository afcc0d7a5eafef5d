#pragma once
#include <iterator>
namespace boost { namespace adaptors {
template <class C> struct reversed_view { C& c; auto begin() { return std::rbegin(c); } auto end() { return std::rend(c); } };
template <class C> reversed_view<C> reverse(C& c) { return {c}; }
}}
