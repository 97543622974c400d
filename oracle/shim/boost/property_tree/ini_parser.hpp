// Hand-written minimal stand-in for boost/property_tree/ini_parser.hpp (test
// infrastructure only): [section], key = value, ';' / '#' comment lines.
#ifndef LUTLDPC_SHIM_BOOST_INI_PARSER_HPP
#define LUTLDPC_SHIM_BOOST_INI_PARSER_HPP
#include <fstream>
#include "ptree.hpp"
namespace boost { namespace property_tree { namespace ini_parser {
inline std::string trim_(const std::string& s) {
  std::string::size_type a = s.find_first_not_of(" \t\r\n");
  if (a == std::string::npos) return "";
  std::string::size_type b = s.find_last_not_of(" \t\r\n");
  return s.substr(a, b - a + 1);
}
inline void read_ini(const std::string& filename, ptree& pt) {
  std::ifstream f(filename.c_str());
  if (!f) throw ptree_error("cannot open " + filename);
  ptree* cur = &pt; std::string line;
  while (std::getline(f, line)) {
    line = trim_(line);
    if (line.empty() || line[0] == ';' || line[0] == '#') continue;
    if (line[0] == '[') {
      std::string::size_type e = line.find(']');
      cur = &pt.put_child_local(trim_(line.substr(1, e - 1)), ptree());
    } else {
      std::string::size_type eq = line.find('=');
      if (eq == std::string::npos) continue;
      cur->put_child_local(trim_(line.substr(0, eq)), ptree(trim_(line.substr(eq + 1))));
    }
  }
}
} using ini_parser::read_ini; }}
#endif
