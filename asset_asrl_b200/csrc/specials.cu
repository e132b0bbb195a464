// Transcendental helpers for the specialised segment kernels (jit.cpp).
//
// The kernel specialiser writes PTX directly; PTX has no fp64 sin/cos/exp/log/..., so these wrappers around the CUDA
// math library are compiled to PTX when the library is built (`nvcc -ptx -rdc=true`, see Makefile) and the text is
// embedded in libasset_b200.so.  A specialised kernel whose program uses one of them is assembled together with this
// module, so both execution back-ends (interpreter in kernels.cu, specialised kernels) call the same math functions.
// Derivative formulas per op: /root/reference/src/VectorFunctions/CommonFunctions/CwiseOperators.h:31-860.
#include <math.h>

extern "C" {
__device__ double ab_sp_exp(double a) { return exp(a); }
__device__ double ab_sp_log(double a) { return log(a); }
__device__ double ab_sp_sin(double a) { return sin(a); }
__device__ double ab_sp_cos(double a) { return cos(a); }
__device__ double ab_sp_tan(double a) { return tan(a); }
__device__ double ab_sp_asin(double a) { return asin(a); }
__device__ double ab_sp_acos(double a) { return acos(a); }
__device__ double ab_sp_atan(double a) { return atan(a); }
__device__ double ab_sp_sinh(double a) { return sinh(a); }
__device__ double ab_sp_cosh(double a) { return cosh(a); }
__device__ double ab_sp_tanh(double a) { return tanh(a); }
__device__ double ab_sp_asinh(double a) { return asinh(a); }
__device__ double ab_sp_acosh(double a) { return acosh(a); }
__device__ double ab_sp_atanh(double a) { return atanh(a); }
__device__ double ab_sp_pow(double a, double b) { return pow(a, b); }
__device__ double ab_sp_atan2(double a, double b) { return atan2(a, b); }
}
