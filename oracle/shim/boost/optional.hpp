// Hand-written minimal stand-in for boost/optional.hpp (test infrastructure only).
// Covers exactly what /root/reference/src/LDPC_DE.cpp:1148-1233 uses:
// optional<T>, optional<T&>, operator bool, operator*, operator->, assignment.
#ifndef LUTLDPC_SHIM_BOOST_OPTIONAL_HPP
#define LUTLDPC_SHIM_BOOST_OPTIONAL_HPP
#include <utility>
namespace boost {
template <class T> class optional {
  bool has_; T val_;
public:
  optional() : has_(false), val_() {}
  optional(const T& v) : has_(true), val_(v) {}
  explicit operator bool() const { return has_; }
  bool operator!() const { return !has_; }
  T& operator*() { return val_; }
  const T& operator*() const { return val_; }
  T* operator->() { return &val_; }
  const T* operator->() const { return &val_; }
  T& get() { return val_; }
  const T& get() const { return val_; }
  bool is_initialized() const { return has_; }
};
template <class T> class optional<T&> {
  T* p_;
public:
  optional() : p_(nullptr) {}
  optional(T& r) : p_(&r) {}
  explicit operator bool() const { return p_ != nullptr; }
  bool operator!() const { return p_ == nullptr; }
  T& operator*() const { return *p_; }
  T* operator->() const { return p_; }
  T& get() const { return *p_; }
  bool is_initialized() const { return p_ != nullptr; }
};
}
#endif
