// Stand-in for <mrpt/containers/yaml.h>: a flat string map with the accessors the element/load parameter loaders use
// (has, operator[], node().as<std::string>(), as<T>()).  The YAML parser itself (parse_file.cpp) is not part of the
// oracle/_ref build.
#pragma once
#include <map>
#include <sstream>
#include <stdexcept>
#include <string>
namespace mrpt::containers {
class yaml {
  std::map<std::string, yaml> map_;
  std::string scalar_;
  bool is_scalar_ = false;

 public:
  using map_t = std::map<std::string, yaml>;
  struct node_t {
    std::string s;
    template <class T> T as() const {
      if constexpr (std::is_same_v<T, std::string>) return s;
      else { std::istringstream is(s); T v{}; is >> v; return v; }
    }
  };
  yaml() = default;
  yaml(const std::string& s) : scalar_(s), is_scalar_(true) {}
  yaml(const char* s) : scalar_(s), is_scalar_(true) {}
  yaml(double v) : is_scalar_(true) { std::ostringstream o; o.precision(17); o << v; scalar_ = o.str(); }
  static yaml Map() { return yaml(); }
  bool has(const std::string& k) const { return map_.count(k) != 0; }
  bool empty() const { return map_.empty() && !is_scalar_; }
  bool isMap() const { return !is_scalar_; }
  bool isScalar() const { return is_scalar_; }
  yaml& operator[](const std::string& k) { return map_[k]; }
  const yaml& operator[](const std::string& k) const {
    auto it = map_.find(k);
    if (it == map_.end()) throw std::out_of_range("yaml: key not found: " + k);
    return it->second;
  }
  node_t node() const { return node_t{scalar_}; }
  template <class T> T as() const { return node().as<T>(); }
  template <class T> T getOrDefault(const std::string& k, const T& d) const { return has(k) ? (*this)[k].template as<T>() : d; }
  const map_t& asMap() const { return map_; }
  map_t& asMap() { return map_; }
  yaml& operator=(const std::string& s) { scalar_ = s; is_scalar_ = true; return *this; }
  yaml& operator=(double v) { return *this = yaml(v); }
  void printAsYAML(std::ostream&) const {}
};
inline std::ostream& operator<<(std::ostream& o, const yaml&) { return o; }
}  // namespace mrpt::containers
