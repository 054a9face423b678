// qtts_gguf.hpp — minimal GGUF (v2/v3) reader for the C++ host adapter: scalar KV metadata and the
// F16 / F32 tensor table.  Inside the reference tree the maintainer would use ggml's gguf_* API
// instead (src/gguf_loader.cpp); this reader keeps the adapter free of any ggml dependency.
// Format reference: ggml/src/gguf.cpp (header, KV section, tensor infos, aligned data section).
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <map>
#include <string>
#include <vector>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

namespace qtts_b200 {

struct GgufTensor {
    std::string name;
    int ggml_type = 0;               // 0 = F32, 1 = F16 (anything else is rejected by the engine)
    std::vector<int64_t> dims;       // ggml order: dims[0] innermost
    uint64_t offset = 0;             // relative to the data section
    size_t n_bytes = 0;
};

class GgufFile {
public:
    ~GgufFile() {
        if (map_) munmap(map_, map_size_);
        if (f_) fclose(f_);
    }
    GgufFile() = default;
    GgufFile(const GgufFile &) = delete;
    GgufFile & operator=(const GgufFile &) = delete;

    bool open(const std::string & path, std::string & err) {
        f_ = fopen(path.c_str(), "rb");
        if (!f_) { err = "Failed to open GGUF file: " + path; return false; }
        path_ = path;
        uint32_t magic = 0, version = 0;
        uint64_t n_tensors = 0, n_kv = 0;
        if (!rd(magic) || magic != 0x46554747u) { err = "not a GGUF file: " + path; return false; }
        if (!rd(version) || version < 2) { err = "unsupported GGUF version"; return false; }
        if (!rd(n_tensors) || !rd(n_kv)) { err = "truncated GGUF header"; return false; }
        uint32_t alignment = 32;
        for (uint64_t i = 0; i < n_kv; ++i) {
            std::string key;
            uint32_t type = 0;
            if (!rd_str(key) || !rd(type)) { err = "truncated GGUF metadata"; return false; }
            double num = 0.0;
            bool is_num = false;
            if (!rd_value(type, num, is_num)) { err = "bad GGUF metadata value for " + key; return false; }
            if (is_num) {
                kv_[key] = num;
                if (key == "general.alignment") alignment = (uint32_t) num;
            }
        }
        tensors_.resize(n_tensors);
        for (auto & t : tensors_) {
            uint32_t n_dims = 0, type = 0;
            if (!rd_str(t.name) || !rd(n_dims)) { err = "truncated GGUF tensor table"; return false; }
            t.dims.resize(n_dims);
            size_t n = 1;
            for (auto & d : t.dims) { uint64_t v; if (!rd(v)) { err = "truncated GGUF tensor table"; return false; } d = (int64_t) v; n *= (size_t) v; }
            if (!rd(type) || !rd(t.offset)) { err = "truncated GGUF tensor table"; return false; }
            t.ggml_type = (int) type;
            t.n_bytes = n * (type == 0 ? 4 : type == 1 ? 2 : 0);
        }
        const long pos = ftell(f_);
        data_start_ = ((uint64_t) pos + alignment - 1) / alignment * alignment;
        return true;
    }

    bool has(const std::string & key) const { return kv_.count(key) != 0; }
    double num(const std::string & key) const { return kv_.at(key); }
    const std::vector<GgufTensor> & tensors() const { return tensors_; }

    // ggml type id -> name, for the messages that reject quantised files (ggml/include/ggml.h: enum ggml_type)
    static const char * type_name(int t) {
        switch (t) {
            case 0: return "F32"; case 1: return "F16"; case 2: return "Q4_0"; case 3: return "Q4_1"; case 6: return "Q5_0";
            case 7: return "Q5_1"; case 8: return "Q8_0"; case 9: return "Q8_1"; case 10: return "Q2_K"; case 11: return "Q3_K";
            case 12: return "Q4_K"; case 13: return "Q5_K"; case 14: return "Q6_K"; case 15: return "Q8_K"; case 30: return "BF16";
            default: return "unknown";
        }
    }

    // Read-only mapping of the whole file: tensor data is then addressed in place (data(t)), so a loader can copy
    // file pages straight into its pinned staging buffers without an intermediate std::vector per tensor.
    bool map(std::string & err) {
        if (map_) return true;
        const int fd = ::open(path_.c_str(), O_RDONLY);
        struct stat st;
        if (fd < 0 || fstat(fd, &st) != 0) { if (fd >= 0) ::close(fd); err = "Failed to open GGUF file: " + path_; return false; }
        void * p = mmap(nullptr, (size_t) st.st_size, PROT_READ, MAP_PRIVATE, fd, 0);
        ::close(fd);
        if (p == MAP_FAILED) { err = "mmap failed for " + path_; return false; }
        map_ = p; map_size_ = (size_t) st.st_size;
        madvise(map_, map_size_, MADV_SEQUENTIAL);
        return true;
    }
    const uint8_t * data(const GgufTensor & t) const {
        if (!map_ || data_start_ + t.offset + t.n_bytes > map_size_) return nullptr;
        return static_cast<const uint8_t *>(map_) + data_start_ + t.offset;
    }

    bool read_tensor(const GgufTensor & t, std::vector<uint8_t> & out, std::string & err) {
        if (t.n_bytes == 0) { err = "tensor " + t.name + ": unsupported GGUF type " + type_name(t.ggml_type) + " (only F32/F16; quantised GGUF files are rejected)"; return false; }
        out.resize(t.n_bytes);
        if (fseek(f_, (long) (data_start_ + t.offset), SEEK_SET) != 0 || fread(out.data(), 1, t.n_bytes, f_) != t.n_bytes) {
            err = "failed to read tensor " + t.name;
            return false;
        }
        return true;
    }

private:
    template <typename T> bool rd(T & v) { return fread(&v, sizeof(T), 1, f_) == 1; }
    bool rd_str(std::string & s) {
        uint64_t n = 0;
        if (!rd(n) || n > (1u << 28)) return false;
        s.resize(n);
        return n == 0 || fread(&s[0], 1, n, f_) == n;
    }
    // GGUF value types: 0 u8, 1 i8, 2 u16, 3 i16, 4 u32, 5 i32, 6 f32, 7 bool, 8 string, 9 array, 10 u64, 11 i64, 12 f64
    bool rd_value(uint32_t type, double & num, bool & is_num) {
        is_num = true;
        switch (type) {
            case 0: { uint8_t v;  if (!rd(v)) return false; num = v; return true; }
            case 1: { int8_t v;   if (!rd(v)) return false; num = v; return true; }
            case 2: { uint16_t v; if (!rd(v)) return false; num = v; return true; }
            case 3: { int16_t v;  if (!rd(v)) return false; num = v; return true; }
            case 4: { uint32_t v; if (!rd(v)) return false; num = v; return true; }
            case 5: { int32_t v;  if (!rd(v)) return false; num = v; return true; }
            case 6: { float v;    if (!rd(v)) return false; num = v; return true; }
            case 7: { uint8_t v;  if (!rd(v)) return false; num = v; return true; }
            case 10: { uint64_t v; if (!rd(v)) return false; num = (double) v; return true; }
            case 11: { int64_t v;  if (!rd(v)) return false; num = (double) v; return true; }
            case 12: { double v;   if (!rd(v)) return false; num = v; return true; }
            case 8: { std::string s; is_num = false; return rd_str(s); }
            case 9: {
                is_num = false;
                uint32_t et = 0; uint64_t n = 0;
                if (!rd(et) || !rd(n)) return false;
                for (uint64_t i = 0; i < n; ++i) { double d; bool b; if (et == 9 || !rd_value(et, d, b)) return false; }
                return true;
            }
            default: return false;
        }
    }

    FILE * f_ = nullptr;
    std::string path_;
    void * map_ = nullptr;
    size_t map_size_ = 0;
    std::map<std::string, double> kv_;
    std::vector<GgufTensor> tensors_;
    uint64_t data_start_ = 0;
};

}  // namespace qtts_b200
