// Host stand-in for <cuda_runtime.h>: lets tests/tools/*.cpp compile the register-level device headers
// (fft_regs.cuh) with g++ and check their arithmetic on the CPU.  Test infrastructure only.
#pragma once
#include <cmath>
#define __device__
#define __host__
#define __forceinline__ inline
struct double2 { double x, y; };
static inline double2 make_double2(double a, double b) { return {a, b}; }
