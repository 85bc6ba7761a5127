// cr_libm.cpp — TEST INFRASTRUCTURE (oracle/_ref/ref_harness_crlibm only).
//
// Correctly rounded float transcendentals: every libm float function the reference reaches (through glm / <cmath>: sin, cos,
// tan, asin, acos, atan, atan2, pow, exp, log) evaluated in double and rounded once.  glibc's own float versions are < 1 ulp
// functions, not correctly rounded ones, and which of its variants runs depends on the host CPU (the FMA ifunc variants); the
// device evaluates the same functions in double and rounds once (csrc/common.cuh: cr_*).  Linking this file into the oracle
// interposes the float entry points, so that `ref_harness_crlibm` is the UNMODIFIED reference on a libm whose float results are
// the correctly rounded ones.  It exists to answer one question: how much of the per-path disagreement between the device and
// the oracle on a scene is the last bit of the host's libm?  (tests/test_large_scene_gpu.py)
#include <math.h>

extern "C" {
float sinf(float x) { return (float)sin((double)x); }
float cosf(float x) { return (float)cos((double)x); }
float tanf(float x) { return (float)tan((double)x); }
float asinf(float x) { return (float)asin((double)x); }
float acosf(float x) { return (float)acos((double)x); }
float atanf(float x) { return (float)atan((double)x); }
float atan2f(float y, float x) { return (float)atan2((double)y, (double)x); }
float powf(float x, float y) { return (float)pow((double)x, (double)y); }
float expf(float x) { return (float)exp((double)x); }
float logf(float x) { return (float)log((double)x); }
float exp2f(float x) { return (float)exp2((double)x); }
float log2f(float x) { return (float)log2((double)x); }
void sincosf(float x, float* s, float* c) { double ds, dc; sincos((double)x, &ds, &dc); *s = (float)ds; *c = (float)dc; }
}
