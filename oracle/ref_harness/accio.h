/* TEST INFRASTRUCTURE (oracle). Not part of the product path.
 *
 * Reader/writer for the flat "ACCARR1" container: a file is the 8-byte magic followed by records
 *   u32 name_len | name bytes | u8 dtype (0 = int64, 1 = float64) | u64 count | count * 8 bytes.
 * Used for the network snapshot the oracle reads and for the per-step dumps it writes.
 * (The product reads the same container in acclimate_b200/netfile.py.)
 */
#ifndef ORACLE_ACCIO_H
#define ORACLE_ACCIO_H

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

namespace accio {

struct Array {
    int dtype = 0;
    std::vector<std::int64_t> i;
    std::vector<double> f;
};

inline std::map<std::string, Array> read_file(const std::string& path) {
    std::map<std::string, Array> res;
    FILE* fp = std::fopen(path.c_str(), "rb");
    if (fp == nullptr) throw std::runtime_error("cannot open " + path);
    char magic[8];
    if (std::fread(magic, 1, 8, fp) != 8 || std::memcmp(magic, "ACCARR1\0", 8) != 0) throw std::runtime_error("bad magic in " + path);
    for (;;) {
        std::uint32_t nl;
        if (std::fread(&nl, 4, 1, fp) != 1) break;
        std::string name(nl, ' ');
        if (std::fread(&name[0], 1, nl, fp) != nl) throw std::runtime_error("truncated " + path);
        std::uint8_t dt;
        std::uint64_t cnt;
        if (std::fread(&dt, 1, 1, fp) != 1 || std::fread(&cnt, 8, 1, fp) != 1) throw std::runtime_error("truncated " + path);
        Array a;
        a.dtype = dt;
        if (dt == 0) {
            a.i.resize(cnt);
            if (cnt && std::fread(a.i.data(), 8, cnt, fp) != cnt) throw std::runtime_error("truncated " + path);
        } else {
            a.f.resize(cnt);
            if (cnt && std::fread(a.f.data(), 8, cnt, fp) != cnt) throw std::runtime_error("truncated " + path);
        }
        res.emplace(std::move(name), std::move(a));
    }
    std::fclose(fp);
    return res;
}

class Writer {
    FILE* fp;

  public:
    explicit Writer(const std::string& path) : fp(std::fopen(path.c_str(), "wb")) {
        if (fp == nullptr) throw std::runtime_error("cannot write " + path);
        std::fwrite("ACCARR1\0", 1, 8, fp);
    }
    ~Writer() { if (fp) std::fclose(fp); }
    void put(const std::string& name, int dtype, const void* data, std::uint64_t cnt) {
        const std::uint32_t nl = name.size();
        const std::uint8_t dt = dtype;
        std::fwrite(&nl, 4, 1, fp);
        std::fwrite(name.data(), 1, nl, fp);
        std::fwrite(&dt, 1, 1, fp);
        std::fwrite(&cnt, 8, 1, fp);
        if (cnt) std::fwrite(data, 8, cnt, fp);
    }
    void put(const std::string& name, const std::vector<std::int64_t>& v) { put(name, 0, v.data(), v.size()); }
    void put(const std::string& name, const std::vector<double>& v) { put(name, 1, v.data(), v.size()); }
};

}  // namespace accio

#endif
