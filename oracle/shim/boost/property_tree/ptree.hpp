// Hand-written minimal stand-in for boost/property_tree/ptree.hpp (test
// infrastructure only): a two-level section/key tree with the accessors the
// reference sources call (get_child_optional, get_optional<T>, get<T>).
#ifndef LUTLDPC_SHIM_BOOST_PTREE_HPP
#define LUTLDPC_SHIM_BOOST_PTREE_HPP
#include <map>
#include <string>
#include <sstream>
#include <stdexcept>
#include <vector>
#include "../optional.hpp"
namespace boost { namespace property_tree {
class ptree_error : public std::runtime_error {
public: explicit ptree_error(const std::string& s) : std::runtime_error(s) {}
};
class ptree {
  std::string data_;
  std::vector<std::pair<std::string, ptree> > kids_;
  template <class T> static bool conv(const std::string& s, T& out) {
    std::istringstream is(s); is >> std::boolalpha >> out;
    if (is.fail()) { std::istringstream is2(s); is2 >> out; return !is2.fail(); }
    return true;
  }
  static bool conv(const std::string& s, std::string& out) { out = s; return true; }
public:
  ptree() {}
  explicit ptree(const std::string& d) : data_(d) {}
  const std::string& data() const { return data_; }
  ptree* find_child(const std::string& path) {
    std::string::size_type dot = path.find('.');
    std::string head = path.substr(0, dot);
    for (auto& kv : kids_) if (kv.first == head) {
      if (dot == std::string::npos) return &kv.second;
      return kv.second.find_child(path.substr(dot + 1));
    }
    return nullptr;
  }
  const ptree* find_child(const std::string& path) const {
    return const_cast<ptree*>(this)->find_child(path);
  }
  ptree& put_child_local(const std::string& key, const ptree& t) {
    kids_.push_back(std::make_pair(key, t)); return kids_.back().second;
  }
  optional<ptree&> get_child_optional(const std::string& path) {
    ptree* p = find_child(path);
    return p ? optional<ptree&>(*p) : optional<ptree&>();
  }
  template <class T> optional<T> get_optional(const std::string& path) const {
    const ptree* p = find_child(path);
    if (!p) return optional<T>();
    T v; if (!conv(p->data_, v)) return optional<T>();
    return optional<T>(v);
  }
  template <class T> T get(const std::string& path) const {
    const ptree* p = find_child(path);
    if (!p) throw ptree_error("No such node (" + path + ")");
    T v; if (!conv(p->data_, v)) throw ptree_error("conversion failed (" + path + ")");
    return v;
  }
  template <class T> T get(const std::string& path, const T& dflt) const {
    optional<T> o = get_optional<T>(path);
    return o ? *o : dflt;
  }
  std::string get(const std::string& path, const char* dflt) const {
    return get<std::string>(path, std::string(dflt));
  }
};
}}
#endif
