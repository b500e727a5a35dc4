// Test infrastructure (oracle side): tiny self-describing binary dump format shared by
// the reference harness (oracle/ref_harness.cpp), the CPU restatement (oracle/hera_oracle.cpp)
// and the Python tests (tests/heradump.py reads it into numpy arrays).
//
// File = repeated sections:  u32 name_len | name bytes | u32 dtype | u64 count | raw little-endian data
// dtype codes: 0=u8 1=u32 2=i32 3=u64 4=i64 5=f32 6=f64
#ifndef HERA_DUMP_HPP
#define HERA_DUMP_HPP

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

namespace heradump {

enum DType : uint32_t { U8 = 0, U32 = 1, I32 = 2, U64 = 3, I64 = 4, F32 = 5, F64 = 6 };

template <typename T> struct dtype_of;
template <> struct dtype_of<uint8_t> { static constexpr DType v = U8; };
template <> struct dtype_of<uint32_t> { static constexpr DType v = U32; };
template <> struct dtype_of<int32_t> { static constexpr DType v = I32; };
template <> struct dtype_of<uint64_t> { static constexpr DType v = U64; };
template <> struct dtype_of<int64_t> { static constexpr DType v = I64; };
template <> struct dtype_of<float> { static constexpr DType v = F32; };
template <> struct dtype_of<double> { static constexpr DType v = F64; };

class Writer {
public:
    explicit Writer(const char *path) : f_(std::fopen(path, "wb")) {}
    ~Writer() { if (f_) std::fclose(f_); }
    bool ok() const { return f_ != nullptr; }

    template <typename T>
    void put(const std::string &name, const std::vector<T> &v) {
        raw(name, dtype_of<T>::v, v.size(), v.data(), sizeof(T));
    }
    template <typename T>
    void scalar(const std::string &name, T x) {
        raw(name, dtype_of<T>::v, 1, &x, sizeof(T));
    }
    void text(const std::string &name, const std::string &s) {
        raw(name, U8, s.size(), s.data(), 1);
    }

private:
    void raw(const std::string &name, DType dt, uint64_t count, const void *data, size_t esz) {
        uint32_t nl = static_cast<uint32_t>(name.size());
        uint32_t d = dt;
        std::fwrite(&nl, 4, 1, f_);
        std::fwrite(name.data(), 1, nl, f_);
        std::fwrite(&d, 4, 1, f_);
        std::fwrite(&count, 8, 1, f_);
        if (count) std::fwrite(data, esz, count, f_);
    }
    std::FILE *f_;
};

}  // namespace heradump
#endif
