// TEST INFRASTRUCTURE. Named-array container writer in the CTBM0001 layout
// (include/cantera_b200/MechTables.h): used by the reference-side tools to dump golden vectors.
#pragma once
#include <cstdint>
#include <cstdio>
#include <stdexcept>
#include <string>
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
