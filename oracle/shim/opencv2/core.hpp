// oracle/shim/opencv2/core.hpp — TEST INFRASTRUCTURE.  A minimal stand-in for the parts of
// OpenCV core / imgproc / calib3d that the reference's two hot-path sources
// (modules/ximgproc/src/fgs_filter.cpp, disparity_filters.cpp) use, so that those sources can be
// compiled UNMODIFIED from /root/reference into oracle/_ref/libref_ximgproc.so (oracle/Makefile,
// target `ref`) when no OpenCV C++ development package exists on the machine.
// Nothing here is product code, and nothing here is copied from OpenCV: the classes implement
// the documented behaviour the two files rely on (row-major ref-counted Mat with ROI views,
// saturate_cast = round-half-even + clamp, eager matrix expressions, stripe parallel_for_).
// The imgproc primitives (boxFilter, sqrBoxFilter, resize) follow the rounding sequences
// verified against cv2 in tests/test_oracle_primitives.py.
#ifndef SHIM_OPENCV2_CORE_HPP
#define SHIM_OPENCV2_CORE_HPP

#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

#define CV_EXPORTS
#define CV_EXPORTS_W
#define CV_EXPORTS_W_SIMPLE
#define CV_WRAP
#define CV_OUT
#define CV_IN_OUT
#define CV_PROP_RW

#define CV_8U 0
#define CV_8S 1
#define CV_16U 2
#define CV_16S 3
#define CV_32S 4
#define CV_32F 5
#define CV_64F 6
#define CV_CN_SHIFT 3
#define CV_MAT_DEPTH(t) ((t) & 7)
#define CV_MAT_CN(t) ((((t) >> CV_CN_SHIFT) & 511) + 1)
#define CV_MAKETYPE(d, c) (CV_MAT_DEPTH(d) + (((c) - 1) << CV_CN_SHIFT))
#define CV_8UC1 CV_MAKETYPE(CV_8U, 1)
#define CV_8UC3 CV_MAKETYPE(CV_8U, 3)
#define CV_16SC1 CV_MAKETYPE(CV_16S, 1)
#define CV_32FC1 CV_MAKETYPE(CV_32F, 1)

namespace cv {

using std::abs; using std::exp; using std::log; using std::max; using std::min; using std::pow; using std::sqrt; using std::swap;
typedef std::string String;
typedef unsigned char uchar;

class Exception : public std::runtime_error {
public:
    int code;
    Exception(int c, const std::string& m) : std::runtime_error(m), code(c) {}
};
namespace Error { enum { StsBadArg = -5, StsBadSize = -201, StsAssert = -215 }; }
#define CV_Error(code, msg) throw cv::Exception((code), std::string(msg))
#define CV_Assert(expr) do { if (!(expr)) throw cv::Exception(cv::Error::StsAssert, "Assertion failed: " #expr); } while (0)

// ---- saturate_cast / cvRound: round-half-to-even, NaN and overflow -> INT_MIN (x86 cvtss2si) ----
static inline int cvRound(double v) { if (!(std::fabs(v) < 2147483648.0)) return INT_MIN; return (int)std::nearbyint(v); }
static inline int cvRound(float v) { if (!(std::fabs(v) < 2147483648.0f)) return INT_MIN; return (int)std::nearbyintf(v); }
static inline int cvFloor(double v) { return (int)std::floor(v); }
static inline int cvCeil(double v) { return (int)std::ceil(v); }
template <typename T> static inline T saturate_cast(int v) { return (T)v; }
template <> inline uchar saturate_cast<uchar>(int v) { return (uchar)(v < 0 ? 0 : v > 255 ? 255 : v); }
template <> inline short saturate_cast<short>(int v) { return (short)(v < -32768 ? -32768 : v > 32767 ? 32767 : v); }
template <typename T> static inline T saturate_cast(float v) { return saturate_cast<T>(cvRound(v)); }
template <> inline float saturate_cast<float>(float v) { return v; }
template <typename T> static inline T saturate_cast(double v) { return saturate_cast<T>(cvRound(v)); }
template <> inline float saturate_cast<float>(double v) { return (float)v; }
template <> inline double saturate_cast<double>(double v) { return v; }

// ---- small geometry / value types -----------------------------------------------------------------
struct Point { int x, y; Point(int _x = 0, int _y = 0) : x(_x), y(_y) {} };
struct Size {
    int width, height;
    Size(int w = 0, int h = 0) : width(w), height(h) {}
    bool operator==(const Size& o) const { return width == o.width && height == o.height; }
    bool operator!=(const Size& o) const { return !(*this == o); }
    int area() const { return width * height; }
};
struct Rect {
    int x, y, width, height;
    Rect(int _x = 0, int _y = 0, int w = 0, int h = 0) : x(_x), y(_y), width(w), height(h) {}
    int area() const { return width * height; }
};
struct Range { int start, end; Range(int s = 0, int e = 0) : start(s), end(e) {} };

template <typename T, int N> struct Vec {
    T val[N];
    enum { depth = std::is_same<T, float>::value ? CV_32F : std::is_same<T, uchar>::value ? CV_8U : std::is_same<T, short>::value ? CV_16S : CV_64F,
           channels = N, type = CV_MAKETYPE(depth, N) };
    Vec() { for (int i = 0; i < N; ++i) val[i] = T(); }
    Vec(T a, T b, T c) { val[0] = a; if (N > 1) val[1] = b; if (N > 2) val[2] = c; }
    T& operator[](int i) { return val[i]; }
    const T& operator[](int i) const { return val[i]; }
};
typedef Vec<uchar, 3> Vec3b;

struct Scalar {
    double val[4];
    Scalar(double a = 0, double b = 0, double c = 0, double d = 0) { val[0] = a; val[1] = b; val[2] = c; val[3] = d; }
    Scalar(const Vec3b& v) { val[0] = v[0]; val[1] = v[1]; val[2] = v[2]; val[3] = 0; }
};

template <typename T> class Ptr : public std::shared_ptr<T> {
public:
    Ptr() {}
    Ptr(T* p) : std::shared_ptr<T>(p) {}
    template <typename U> Ptr(const std::shared_ptr<U>& o) : std::shared_ptr<T>(o) {}
    template <typename U> Ptr(const Ptr<U>& o) : std::shared_ptr<T>(std::static_pointer_cast<T>(std::shared_ptr<U>(o))) {}
    template <typename U> Ptr<U> dynamicCast() const { return Ptr<U>(std::dynamic_pointer_cast<U>(std::shared_ptr<T>(*this))); }
    bool empty() const { return !this->get(); }
};
template <typename T, typename... A> static inline Ptr<T> makePtr(A&&... a) { return Ptr<T>(new T(std::forward<A>(a)...)); }

class Algorithm { public: virtual ~Algorithm() {} };

static inline size_t elemSize1(int depth) { static const size_t s[] = { 1, 1, 2, 2, 4, 4, 8 }; return s[depth]; }

}  // namespace cv

#include "opencv2/shim_mat.hpp"
#include "opencv2/shim_ops.hpp"

#endif
