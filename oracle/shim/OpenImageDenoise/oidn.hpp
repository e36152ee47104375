// TEST INFRASTRUCTURE — stub of the Intel OIDN C++ API, just enough for the reference's camera.hpp
// (camera.hpp:10, :581-699) to compile.  OIDN binaries are not in the reference checkout
// (libs/oidn_linux holds only .gitkeep) and the denoiser is off by default (camera.hpp:45), so none
// of this is ever executed by the oracle harness.
#pragma once
#include <cstddef>
namespace oidn {
enum class DeviceType { Default, CPU, SYCL, CUDA, HIP, Metal };
enum class Format { Undefined, Float, Float2, Float3, Float4 };
enum class Error { None, Unknown };
struct BufferRef {
    void write(size_t, size_t, const void*) {}
    void read(size_t, size_t, void*) {}
};
struct FilterRef {
    void setImage(const char*, BufferRef&, Format, size_t, size_t) {}
    template <class T> void set(const char*, T) {}
    void commit() {}
    void execute() {}
};
struct DeviceRef {
    void commit() {}
    BufferRef newBuffer(size_t) { return {}; }
    FilterRef newFilter(const char*) { return {}; }
    template <class T> T get(const char*) { return T{}; }
    Error getError(const char*& msg) { msg = "OIDN stub"; return Error::Unknown; }
    explicit operator bool() const { return false; }
};
inline DeviceRef newDevice(DeviceType = DeviceType::Default) { return {}; }
}  // namespace oidn
