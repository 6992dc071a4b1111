// TEST INFRASTRUCTURE — not product code.
//
// Host-serial stand-in for <sycl/CL/sycl.hpp> so that the reference's kernel-side
// translation units (/root/reference/src/{kernel,BVH,Disney,HDRI,Texture,...}.cpp) compile
// UNMODIFIED with g++ and run on the CPU as the parity oracle (oracle/_ref/).
// Only the subset the reference touches is provided (reference include site: src/sycl.h:8).
//
//  * sycl::{sqrt,sin,cos,abs,pow,floor,log,acos,atan2,atan,isnan} -> glibc float libm
//  * SYCL_EXTERNAL -> nothing
//  * queue/handler/nd_range/nd_item/malloc_device/is_compatible -> serial host equivalents
#pragma once

#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <string>

#define SYCL_EXTERNAL

namespace sycl {

inline float sqrt(float v) { return ::sqrtf(v); }
inline double sqrt(double v) { return ::sqrt(v); }
inline float sin(float v) { return ::sinf(v); }
inline float cos(float v) { return ::cosf(v); }
inline float abs(float v) { return ::fabsf(v); }
inline float fabs(float v) { return ::fabsf(v); }
inline float pow(float b, float e) { return ::powf(b, e); }
inline float floor(float v) { return ::floorf(v); }
inline float log(float v) { return ::logf(v); }
inline float acos(float v) { return ::acosf(v); }
inline float atan2(float y, float x) { return ::atan2f(y, x); }
inline float atan(float v) { return ::atanf(v); }
inline bool isnan(float v) { return std::isnan(v); }

class device {
public:
    device() = default;
};

class range {
public:
    size_t dims[2];
    range(size_t a, size_t b) : dims{a, b} {}
    size_t get(int i) const { return dims[i]; }
};

class nd_range {
public:
    range global, local;
    nd_range(range g, range l) : global(g), local(l) {}
};

template <int D>
class nd_item {
public:
    size_t gid[D];
    size_t get_global_id(int i) const { return gid[i]; }
};

class event {
public:
    void wait() {}
};

class handler {
public:
    template <typename Name = void, typename F>
    void parallel_for(nd_range r, F f) {
        nd_item<2> it;
        for (size_t a = 0; a < r.global.get(0); a++) {
            for (size_t b = 0; b < r.global.get(1); b++) {
                it.gid[0] = a;
                it.gid[1] = b;
                f(it);
            }
        }
    }
};

class queue {
public:
    queue() = default;
    explicit queue(const device&) {}
    template <typename F>
    event submit(F f) {
        handler h;
        f(h);
        return event();
    }
    event memcpy(void* dst, const void* src, size_t n) {
        std::memcpy(dst, src, n);
        return event();
    }
};

template <typename T>
T* malloc_device(size_t n, queue&) {
    // zero-initialised so never-written tails (reference SYCLCopy.cpp:95) are deterministic
    return static_cast<T*>(std::calloc(n ? n : 1, sizeof(T)));
}

template <typename K>
bool is_compatible(const device&) { return true; }

}  // namespace sycl

namespace cl {
namespace sycl = ::sycl;
}
