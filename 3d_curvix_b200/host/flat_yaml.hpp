// flat_yaml.hpp -- reader for the flat `key: value  #> comment` settings files of the reference
// (configs/3D_Curvix_ABC_NEF.yaml, configs/3D_Curvix_DTU.yaml).  The reference parses them with yaml-cpp
// (EdgeSketch_Core.cpp:37-59); yaml-cpp is not available in this image and the files use nothing beyond one scalar
// per line, so a scalar-per-line reader is the whole dependency.
#pragma once
#include <cstdlib>
#include <fstream>
#include <map>
#include <stdexcept>
#include <string>

namespace curvix_host {

class FlatYaml {
public:
    static FlatYaml load(const std::string& path) {
        std::ifstream in(path);
        if (!in) throw std::runtime_error("cannot open settings file: " + path);
        FlatYaml y;
        std::string line;
        while (std::getline(in, line)) {
            const size_t hash = line.find('#');            // comments ("#>", "# %YAML:1.0") run to the end of the line
            if (hash != std::string::npos) line.erase(hash);
            const size_t colon = line.find(':');
            if (colon == std::string::npos) continue;
            std::string key = trim(line.substr(0, colon)), val = trim(line.substr(colon + 1));
            if (key.empty()) continue;
            if (val.size() >= 2 && ((val.front() == '"' && val.back() == '"') || (val.front() == '\'' && val.back() == '\'')))
                val = val.substr(1, val.size() - 2);
            y.kv_[key] = val;
        }
        return y;
    }
    bool has(const std::string& k) const { return kv_.count(k) != 0; }
    const std::string& str(const std::string& k) const {
        auto it = kv_.find(k);
        if (it == kv_.end()) throw std::runtime_error("settings key missing: " + k);   // yaml-cpp throws on .as<T>() of a null node too
        return it->second;
    }
    std::string str(const std::string& k, const std::string& dflt) const { return has(k) ? str(k) : dflt; }
    int as_int(const std::string& k) const { return int(std::strtol(str(k).c_str(), nullptr, 10)); }
    int as_int(const std::string& k, int dflt) const { return has(k) ? as_int(k) : dflt; }
    double as_double(const std::string& k) const { return std::strtod(str(k).c_str(), nullptr); }
    double as_double(const std::string& k, double dflt) const { return has(k) ? as_double(k) : dflt; }
    bool as_bool(const std::string& k) const {
        const std::string& v = str(k);
        return v == "true" || v == "True" || v == "TRUE" || v == "yes" || v == "on" || v == "1";
    }
    bool as_bool(const std::string& k, bool dflt) const { return has(k) ? as_bool(k) : dflt; }
    void set(const std::string& k, const std::string& v) { kv_[k] = v; }

private:
    static std::string trim(const std::string& s) {
        const size_t a = s.find_first_not_of(" \t\r\n");
        if (a == std::string::npos) return "";
        const size_t b = s.find_last_not_of(" \t\r\n");
        return s.substr(a, b - a + 1);
    }
    std::map<std::string, std::string> kv_;
};

}  // namespace curvix_host
