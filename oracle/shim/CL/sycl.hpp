// TEST INFRASTRUCTURE — see ../sycl/CL/sycl.hpp (reference include site: src/shader_placeholder.cpp:1)
#pragma once
#include "../sycl/CL/sycl.hpp"
