// TEST INFRASTRUCTURE ONLY -- flat "KEY: value  # comment" reader with the three yaml-cpp names
// the reference uses (YAML::LoadFile, Node::operator[], Node::as<T>), enough for the shipped
// config.yml / params.yml.
#pragma once
#include <fstream>
#include <map>
#include <sstream>
#include <stdexcept>
#include <string>

namespace YAML
{
class Node
{
public:
  std::map<std::string, std::string> kv;
  std::string val;
  bool scalar = false;
  Node operator[](const std::string& key) const
  {
    auto it = kv.find(key);
    if (it == kv.end()) throw std::runtime_error("yaml shim: missing key " + key);
    Node n;
    n.val = it->second;
    n.scalar = true;
    return n;
  }
  template <typename T>
  T as() const;
};
template <>
inline std::string Node::as<std::string>() const { return val; }
template <>
inline double Node::as<double>() const { return std::stod(val); }
template <>
inline float Node::as<float>() const { return std::stof(val); }
template <>
inline int Node::as<int>() const { return std::stoi(val); }
template <>
inline unsigned int Node::as<unsigned int>() const { return (unsigned int)std::stoul(val); }
template <>
inline size_t Node::as<size_t>() const { return (size_t)std::stoull(val); }
template <>
inline bool Node::as<bool>() const
{
  return val == "True" || val == "true" || val == "TRUE" || val == "yes" || val == "on" || val == "1";
}
inline Node LoadFile(const std::string& path)
{
  std::ifstream f(path);
  if (!f) throw std::runtime_error("yaml shim: cannot open " + path);
  Node n;
  std::string line;
  while (std::getline(f, line))
  {
    const size_t hash = line.find('#');
    if (hash != std::string::npos) line.erase(hash);
    const size_t colon = line.find(':');
    if (colon == std::string::npos) continue;
    auto trim = [](std::string s) {
      const size_t a = s.find_first_not_of(" \t\r\n");
      const size_t b = s.find_last_not_of(" \t\r\n");
      return a == std::string::npos ? std::string() : s.substr(a, b - a + 1);
    };
    const std::string k = trim(line.substr(0, colon)), v = trim(line.substr(colon + 1));
    if (!k.empty()) n.kv[k] = v;
  }
  return n;
}
}  // namespace YAML
