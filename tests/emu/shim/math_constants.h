// tests/emu/shim/math_constants.h — TEST INFRASTRUCTURE ONLY (see cuda_runtime.h in this directory).
#pragma once
#include <math.h>
#define CUDART_INF ((double)INFINITY)
#define CUDART_INF_F ((float)INFINITY)
