// Declarations-only stand-in for yaml-cpp (not installed in this image), used by tests/test_reference_compiles.py to run
// `g++ -fsyntax-only` over the UNMODIFIED reference translation units against include/.  TEST INFRASTRUCTURE ONLY.
// The real header transitively includes these standard headers and the reference relies on that.
#pragma once
#include <cstddef>
#include <iostream>
#include <map>
#include <sstream>
#include <string>
#include <vector>
namespace YAML {
class Node {
 public:
  Node();
  template <class T> T as() const;
  template <class T, class S> T as(const S &fallback) const;
  template <class K> const Node operator[](const K &key) const;
  template <class K> Node operator[](const K &key);
  explicit operator bool() const;
  bool operator!() const;
  bool IsDefined() const;
  bool IsSequence() const;
  bool IsMap() const;
  std::size_t size() const;
};
Node LoadFile(const std::string &filename);
}  // namespace YAML
