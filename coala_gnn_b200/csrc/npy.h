// Minimal .npy reader with the acceptance rules of the reference loader
// (M/node_distributor_pybind.cuh:11-109): magic "\x93NUMPY", format version 1.x (2-byte header
// length) or 2.x (4-byte), little-endian 8-byte elements ('<i8' or '<f8'), C order, 1-D or 2-D.
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

struct NpyArray {
    std::vector<char> raw;       // whole file
    size_t data_off = 0;         // offset of the first element
    std::vector<int64_t> shape;
    std::string descr;
    const void* data() const { return raw.data() + data_off; }
    int64_t numel() const { int64_t n = 1; for (auto s : shape) n *= s; return n; }
};

inline NpyArray npy_load(const std::string& path) {
    NpyArray a;
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) throw std::runtime_error("Unable to open file: " + path);
    fseek(f, 0, SEEK_END);
    long sz = ftell(f);
    fseek(f, 0, SEEK_SET);
    a.raw.resize((size_t)sz);
    if (sz > 0 && fread(a.raw.data(), 1, (size_t)sz, f) != (size_t)sz) { fclose(f); throw std::runtime_error("Error reading the file: " + path); }
    fclose(f);
    if (sz < 10 || memcmp(a.raw.data(), "\x93NUMPY", 6) != 0) throw std::runtime_error("Not a valid .npy file.");
    const uint8_t major = (uint8_t)a.raw[6];
    size_t pos = 8;
    uint32_t hlen = 0;
    if (major == 1) { uint16_t h16; memcpy(&h16, a.raw.data() + pos, 2); hlen = h16; pos += 2; }
    else if (major == 2) { memcpy(&hlen, a.raw.data() + pos, 4); pos += 4; }
    else throw std::runtime_error("Unsupported .npy file version: " + std::to_string((int)major));
    if (pos + hlen > (size_t)sz) throw std::runtime_error("Not a valid .npy file.");
    std::string hdr(a.raw.data() + pos, hlen);
    a.data_off = pos + hlen;
    auto key = [&](const char* k) -> size_t { size_t p = hdr.find(k); return p == std::string::npos ? p : p + strlen(k); };
    size_t p = key("'descr':");
    if (p != std::string::npos) {
        size_t q1 = hdr.find('\'', p), q2 = q1 == std::string::npos ? q1 : hdr.find('\'', q1 + 1);
        if (q2 != std::string::npos) a.descr = hdr.substr(q1 + 1, q2 - q1 - 1);
    }
    p = key("'fortran_order':");
    if (p != std::string::npos && hdr.compare(hdr.find_first_not_of(' ', p), 4, "True") == 0)
        throw std::runtime_error("Fortran-ordered .npy files are not supported: " + path);
    p = key("'shape':");
    if (p == std::string::npos) throw std::runtime_error("Not a valid .npy file.");
    size_t o = hdr.find('(', p), c = hdr.find(')', p);
    if (o == std::string::npos || c == std::string::npos) throw std::runtime_error("Not a valid .npy file.");
    int64_t cur = -1;
    for (size_t i = o + 1; i <= c; i++) {
        char ch = hdr[i];
        if (ch >= '0' && ch <= '9') cur = (cur < 0 ? 0 : cur) * 10 + (ch - '0');
        else if (cur >= 0) { a.shape.push_back(cur); cur = -1; }
    }
    if (a.descr != "<i8" && a.descr != "<f8") throw std::runtime_error("Unsupported dtype '" + a.descr + "' in " + path + " (only '<i8' and '<f8')");
    if (a.shape.size() < 1 || a.shape.size() > 2) throw std::runtime_error("Unsupported dimension for .npy file");
    if (a.data_off + (size_t)a.numel() * 8 > (size_t)sz) throw std::runtime_error("Truncated .npy file: " + path);
    return a;
}
