// yaml-cpp stand-in (TEST INFRASTRUCTURE ONLY, see oracle/ref_shim/README.md).
// The reference reads only flat "key: value   #> comment" maps (configs/3D_Curvix_*.yaml) through
// YAML::LoadFile, Node::operator[] and Node::as<T>() (EdgeSketch_Core.cpp:37-59, edge_mapping.cpp:64-70,
// cmd/main.cpp:101-103).  That is all this header provides.
#pragma once
// yaml-cpp's own headers pull these in; the reference relies on that for std::set (edge_mapping.hpp:250)
#include <algorithm>
#include <fstream>
#include <iostream>
#include <list>
#include <map>
#include <set>
#include <vector>
#include <memory>
#include <sstream>
#include <stdexcept>
#include <string>

namespace YAML {

class Exception : public std::runtime_error {
public:
    explicit Exception(const std::string& m) : std::runtime_error(m) {}
};
class BadFile : public Exception {
public:
    explicit BadFile(const std::string& f) : Exception("bad file: " + f) {}
};

class Node {
public:
    Node() : m_map(std::make_shared<std::map<std::string, std::string>>()), m_defined(true), m_scalar(false) {}
    Node operator[](const std::string& key) const {
        Node n;
        n.m_scalar = true;
        auto it = m_map->find(key);
        n.m_defined = it != m_map->end();
        n.m_key = key;
        if (n.m_defined) n.m_value = it->second;
        return n;
    }
    Node operator[](const char* key) const { return (*this)[std::string(key)]; }
    bool IsDefined() const { return m_defined; }
    explicit operator bool() const { return m_defined; }
    template <class T> T as() const {
        if (!m_defined) throw Exception("bad conversion: key '" + m_key + "' not found");
        return convert<T>(m_value, m_key);
    }
    void set(const std::string& k, const std::string& v) { (*m_map)[k] = v; }
    const std::map<std::string, std::string>& entries() const { return *m_map; }

private:
    template <class T> static T convert(const std::string& s, const std::string& key) {
        if constexpr (std::is_same<T, std::string>::value) {
            return s;
        } else if constexpr (std::is_same<T, bool>::value) {
            if (s == "true" || s == "True" || s == "TRUE" || s == "yes" || s == "on") return true;
            if (s == "false" || s == "False" || s == "FALSE" || s == "no" || s == "off") return false;
            throw Exception("bad conversion to bool for key '" + key + "'");
        } else {
            std::istringstream is(s);
            T v;
            is >> v;
            if (is.fail()) throw Exception("bad conversion for key '" + key + "': " + s);
            return v;
        }
    }
    std::shared_ptr<std::map<std::string, std::string>> m_map;
    bool m_defined, m_scalar;
    std::string m_key, m_value;
};

inline std::string trim_(const std::string& s) {
    size_t b = s.find_first_not_of(" \t\r\n"), e = s.find_last_not_of(" \t\r\n");
    return b == std::string::npos ? std::string() : s.substr(b, e - b + 1);
}

inline Node LoadFile(const std::string& path) {
    std::ifstream f(path);
    if (!f.is_open()) throw BadFile(path);
    Node n;
    std::string line;
    while (std::getline(f, line)) {
        // a comment starts at '#' that opens the line or follows white space
        for (size_t i = 0; i < line.size(); i++)
            if (line[i] == '#' && (i == 0 || line[i - 1] == ' ' || line[i - 1] == '\t')) { line.erase(i); break; }
        if (!line.empty() && line[0] == '%') continue;   // %YAML directive
        const size_t c = line.find(':');
        if (c == std::string::npos) continue;
        std::string k = trim_(line.substr(0, c)), v = trim_(line.substr(c + 1));
        if (v.size() >= 2 && ((v.front() == '"' && v.back() == '"') || (v.front() == '\'' && v.back() == '\''))) v = v.substr(1, v.size() - 2);
        if (!k.empty()) n.set(k, v);
    }
    return n;
}

inline std::ostream& operator<<(std::ostream& s, const Node& n) {
    for (const auto& kv : n.entries()) s << kv.first << ": " << kv.second << "\n";
    return s;
}

}  // namespace YAML
