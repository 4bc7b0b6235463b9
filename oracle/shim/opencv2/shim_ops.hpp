// oracle/shim/opencv2/shim_ops.hpp — TEST INFRASTRUCTURE (see core.hpp): element-wise operations,
// eager "matrix expressions", parallel_for_, and the imgproc / calib3d pieces the two files call.
#ifndef SHIM_OPENCV2_OPS_HPP
#define SHIM_OPENCV2_OPS_HPP

namespace cv {

// ---- threads: parallel_for_ splits the range into getNumThreads() contiguous stripes ------------
inline int& shim_num_threads() { static int n = (int)std::max(1u, std::thread::hardware_concurrency()); return n; }
inline int getNumThreads() { return shim_num_threads(); }
inline void setNumThreads(int n) { shim_num_threads() = n <= 0 ? (int)std::max(1u, std::thread::hardware_concurrency()) : n; }
inline int getNumberOfCPUs() { return (int)std::max(1u, std::thread::hardware_concurrency()); }
class ParallelLoopBody { public: virtual ~ParallelLoopBody() {} virtual void operator()(const Range& r) const = 0; };
inline void parallel_for_(const Range& range, const ParallelLoopBody& body, double = -1.)
{
    const int n = range.end - range.start, nt = std::min(getNumThreads(), n);
    if (nt <= 1) { body(range); return; }
    std::vector<std::thread> th;
    for (int t = 0; t < nt; ++t) {
        Range r(range.start + (int)((int64_t)n * t / nt), range.start + (int)((int64_t)n * (t + 1) / nt));
        if (t + 1 < nt) th.emplace_back([&body, r] { body(r); }); else body(r);
    }
    for (auto& x : th) x.join();
}

// ---- fill / convert / multiply ---------------------------------------------------------------------
template <typename T> static inline void shim_fill(Mat& m, const Scalar& s)
{
    const int cn = m.channels();
    for (int i = 0; i < m.rows; ++i) { T* p = m.ptr<T>(i); for (int j = 0; j < m.cols; ++j) for (int c = 0; c < cn; ++c) p[j * cn + c] = saturate_cast<T>(s.val[c]); }
}
inline Mat& Mat::operator=(const Scalar& s)
{
    switch (depth()) { case CV_8U: shim_fill<uchar>(*this, s); break; case CV_16S: shim_fill<short>(*this, s); break;
                       case CV_32F: shim_fill<float>(*this, s); break; default: CV_Error(Error::StsBadArg, "shim: unsupported depth"); }
    return *this;
}
// cv::Mat::convertTo: work type float (OpenCV's cvtScale for 8U/16S/32F), dst = saturate(src*alpha + beta)
template <typename S, typename D> static inline void shim_cvt(const Mat& src, Mat& dst, float a, float b, bool scale)
{
    const int n = src.cols * src.channels();
    for (int i = 0; i < src.rows; ++i) {
        const S* s = src.ptr<S>(i); D* d = dst.ptr<D>(i);
        if (scale) for (int j = 0; j < n; ++j) d[j] = saturate_cast<D>((float)s[j] * a + b);
        else for (int j = 0; j < n; ++j) d[j] = saturate_cast<D>((float)s[j]);
    }
}
inline void Mat::convertTo(Mat& dst, int rtype, double alpha, double beta) const
{
    const int dd = CV_MAT_DEPTH(rtype < 0 ? type() : rtype), dt = CV_MAKETYPE(dd, channels());
    Mat src = *this;                                    // keeps the buffer alive for in-place calls
    Mat out;
    if (dst.data && dst.rows == rows && dst.cols == cols && dst.type() == dt && dst.data != data) out = dst; else out.create(rows, cols, dt);
    const bool sc = !(alpha == 1 && beta == 0);
    const float a = (float)alpha, b = (float)beta;
#define SHIM_CVT(SD, ST, DD, DT) if (src.depth() == SD && dd == DD) { shim_cvt<ST, DT>(src, out, a, b, sc); dst = out; return; }
    SHIM_CVT(CV_8U, uchar, CV_8U, uchar) SHIM_CVT(CV_8U, uchar, CV_16S, short) SHIM_CVT(CV_8U, uchar, CV_32F, float)
    SHIM_CVT(CV_16S, short, CV_8U, uchar) SHIM_CVT(CV_16S, short, CV_16S, short) SHIM_CVT(CV_16S, short, CV_32F, float)
    SHIM_CVT(CV_32F, float, CV_8U, uchar) SHIM_CVT(CV_32F, float, CV_16S, short) SHIM_CVT(CV_32F, float, CV_32F, float)
#undef SHIM_CVT
    CV_Error(Error::StsBadArg, "shim: unsupported conversion");
}
inline Mat Mat::mul(const Mat& o) const
{
    CV_Assert(depth() == CV_32F && o.type() == type() && o.rows == rows && o.cols == cols);
    Mat r(rows, cols, type());
    const int n = cols * channels();
    for (int i = 0; i < rows; ++i) { const float* a = ptr<float>(i); const float* b = o.ptr<float>(i); float* d = r.ptr<float>(i); for (int j = 0; j < n; ++j) d[j] = a[j] * b[j]; }
    return r;
}
// eager stand-ins for the MatExpr forms the reference writes (disparity_filters.cpp:209, 244, 273, 295)
inline Mat operator*(const Mat& m, double s) { Mat r; m.convertTo(r, -1, s, 0); return r; }      // Mat * scalar -> convertTo(alpha)
inline Mat operator*(double s, const Mat& m) { return m * s; }
inline Mat operator+(const Mat& m, double s)                                                      // cv::add(Mat32F, Scalar): float add
{
    CV_Assert(m.depth() == CV_32F);
    Mat r(m.rows, m.cols, m.type());
    const int n = m.cols * m.channels(); const float fs = (float)s;
    for (int i = 0; i < m.rows; ++i) { const float* a = m.ptr<float>(i); float* d = r.ptr<float>(i); for (int j = 0; j < n; ++j) { volatile float t = a[j] + fs; d[j] = t; } }
    return r;
}
inline Mat operator/(double s, const Mat& m)                                                      // cv::divide(scale, Mat32F): scale/x in double, 0 where x == 0
{
    CV_Assert(m.depth() == CV_32F);
    Mat r(m.rows, m.cols, m.type());
    const int n = m.cols * m.channels();
    for (int i = 0; i < m.rows; ++i) { const float* a = m.ptr<float>(i); float* d = r.ptr<float>(i); for (int j = 0; j < n; ++j) d[j] = a[j] != 0 ? (float)(s / (double)a[j]) : 0.f; }
    return r;
}
inline Mat& operator/=(Mat& m, double s) { Mat r; m.convertTo(r, -1, 1.0 / s, 0); m = r; return m; }

inline void split(InputArray src, std::vector<Mat>& mv)
{
    Mat s = src.getMat(); const int cn = s.channels(); const size_t es = elemSize1(s.depth());
    mv.resize(cn);
    for (int c = 0; c < cn; ++c) {
        mv[c].create(s.rows, s.cols, CV_MAKETYPE(s.depth(), 1));
        for (int i = 0; i < s.rows; ++i) for (int j = 0; j < s.cols; ++j) std::memcpy(mv[c].ptr(i) + j * es, s.ptr(i) + ((size_t)j * cn + c) * es, es);
    }
}
inline void merge(const std::vector<Mat>& mv, OutputArray dst)
{
    const int cn = (int)mv.size(); const size_t es = elemSize1(mv[0].depth());
    Mat out(mv[0].rows, mv[0].cols, CV_MAKETYPE(mv[0].depth(), cn));
    for (int c = 0; c < cn; ++c)
        for (int i = 0; i < out.rows; ++i) for (int j = 0; j < out.cols; ++j) std::memcpy(out.ptr(i) + ((size_t)j * cn + c) * es, mv[c].ptr(i) + j * es, es);
    dst.getMatRef() = out;
}

}  // namespace cv
#include "opencv2/shim_imgproc.hpp"
#endif
