#pragma once
#include <memory>
namespace boost {
template <class T, class U>
std::shared_ptr<T> polymorphic_pointer_downcast(const std::shared_ptr<U>& p) { return std::dynamic_pointer_cast<T>(p); }
}
