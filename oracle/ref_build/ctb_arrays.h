// TEST INFRASTRUCTURE. Named-array container writer in the CTBM0001 layout
// (include/cantera_b200/MechTables.h): used by the reference-side tools to dump golden vectors.
#pragma once
#include <cstdint>
#include <cstdio>
#include <stdexcept>
#include <string>
#include <map>
#include <vector>

struct ArrayFile {
    struct Entry { std::string name; uint32_t dtype; uint64_t count; std::string bytes; };
    std::vector<Entry> entries;
    void add(const std::string& name, const std::vector<double>& v) {
        entries.push_back({name, 0u, v.size(), std::string((const char*) v.data(), v.size() * 8)});
    }
    void add(const std::string& name, const std::vector<int32_t>& v) {
        entries.push_back({name, 1u, v.size(), std::string((const char*) v.data(), v.size() * 4)});
    }
    void addText(const std::string& name, const std::string& s) {
        entries.push_back({name, 2u, s.size(), s});
    }
    void save(const std::string& path) const {
        FILE* f = std::fopen(path.c_str(), "wb");
        if (!f) throw std::runtime_error("cannot open " + path);
        std::fwrite("CTBM0001", 1, 8, f);
        uint32_t n = (uint32_t) entries.size();
        std::fwrite(&n, 4, 1, f);
        for (auto& e : entries) {
            uint32_t nl = (uint32_t) e.name.size();
            std::fwrite(&nl, 4, 1, f);
            std::fwrite(e.name.data(), 1, nl, f);
            std::fwrite(&e.dtype, 4, 1, f);
            std::fwrite(&e.count, 8, 1, f);
            std::fwrite(e.bytes.data(), 1, e.bytes.size(), f);
        }
        std::fclose(f);
    }
};

//! Reader of the same layout (only the f64 / i32 arrays are kept).
struct ArrayReader {
    std::map<std::string, std::vector<double>> dbl;
    std::map<std::string, std::vector<int32_t>> i32;
    void load(const std::string& path) {
        FILE* f = std::fopen(path.c_str(), "rb");
        if (!f) throw std::runtime_error("cannot open " + path);
        auto rd = [&](void* p, size_t n) {
            if (n && std::fread(p, 1, n, f) != n) {
                std::fclose(f);
                throw std::runtime_error("truncated array file " + path);
            }
        };
        char magic[8];
        rd(magic, 8);
        if (std::string(magic, 8) != "CTBM0001") {
            std::fclose(f);
            throw std::runtime_error(path + " is not a CTBM0001 file");
        }
        uint32_t n;
        rd(&n, 4);
        for (uint32_t a = 0; a < n; a++) {
            uint32_t nl, dtype;
            uint64_t cnt;
            rd(&nl, 4);
            std::string name(nl, '\0');
            rd(&name[0], nl);
            rd(&dtype, 4);
            rd(&cnt, 8);
            if (dtype == 0) { auto& v = dbl[name]; v.resize(cnt); rd(v.data(), cnt * 8); }
            else if (dtype == 1) { auto& v = i32[name]; v.resize(cnt); rd(v.data(), cnt * 4); }
            else { std::string s(cnt, '\0'); rd(&s[0], cnt); }
        }
        std::fclose(f);
    }
    const std::vector<double>& d(const std::string& name) const {
        auto it = dbl.find(name);
        if (it == dbl.end()) throw std::runtime_error("array '" + name + "' missing");
        return it->second;
    }
};
