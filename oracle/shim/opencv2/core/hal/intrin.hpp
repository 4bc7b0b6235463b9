// oracle/shim/opencv2/core/hal/intrin.hpp — TEST INFRASTRUCTURE: the 128-bit "universal intrinsics"
// the reference's FGS solver uses (fgs_filter.cpp:293-355, 387-428, 516-548, 566-580), on SSE.
#ifndef SHIM_OPENCV2_INTRIN_HPP
#define SHIM_OPENCV2_INTRIN_HPP
#include <xmmintrin.h>
#define CV_SIMD128 1
namespace cv {
struct v_float32x4 {
    __m128 val;
    v_float32x4() : val(_mm_setzero_ps()) {}
    explicit v_float32x4(__m128 v) : val(v) {}
    v_float32x4(float a, float b, float c, float d) : val(_mm_setr_ps(a, b, c, d)) {}
};
static inline v_float32x4 operator+(const v_float32x4& a, const v_float32x4& b) { return v_float32x4(_mm_add_ps(a.val, b.val)); }
static inline v_float32x4 operator-(const v_float32x4& a, const v_float32x4& b) { return v_float32x4(_mm_sub_ps(a.val, b.val)); }
static inline v_float32x4 operator*(const v_float32x4& a, const v_float32x4& b) { return v_float32x4(_mm_mul_ps(a.val, b.val)); }
static inline v_float32x4 operator/(const v_float32x4& a, const v_float32x4& b) { return v_float32x4(_mm_div_ps(a.val, b.val)); }
static inline v_float32x4 v_load(const float* p) { return v_float32x4(_mm_loadu_ps(p)); }
static inline void v_store(float* p, const v_float32x4& a) { _mm_storeu_ps(p, a.val); }
static inline void v_transpose4x4(const v_float32x4& a0, const v_float32x4& a1, const v_float32x4& a2, const v_float32x4& a3,
                                  v_float32x4& b0, v_float32x4& b1, v_float32x4& b2, v_float32x4& b3)
{
    __m128 r0 = a0.val, r1 = a1.val, r2 = a2.val, r3 = a3.val;
    _MM_TRANSPOSE4_PS(r0, r1, r2, r3);
    b0.val = r0; b1.val = r1; b2.val = r2; b3.val = r3;
}
}  // namespace cv
#endif
