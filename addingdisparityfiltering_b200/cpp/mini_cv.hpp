// mini_cv.hpp — the handful of OpenCV core / calib3d types the facade in dwls_ximgproc.hpp needs,
// for machines without an OpenCV development package (neither the build container nor the GPU box
// has one).  With real OpenCV installed this header is not used: dwls_ximgproc.hpp includes
// <opencv2/core.hpp> and <opencv2/calib3d.hpp> instead and the facade compiles against them.
#ifndef DWLS_MINI_CV_HPP
#define DWLS_MINI_CV_HPP

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>

#define CV_8U 0
#define CV_16S 3
#define CV_32F 5
#define CV_MAT_DEPTH(t) ((t) & 7)
#define CV_MAT_CN(t) ((((t) >> 3) & 511) + 1)
#define CV_MAKETYPE(d, c) (((d) & 7) + (((c) - 1) << 3))
#define CV_8UC1 CV_MAKETYPE(CV_8U, 1)
#define CV_8UC3 CV_MAKETYPE(CV_8U, 3)
#define CV_16SC1 CV_MAKETYPE(CV_16S, 1)
#define CV_32FC1 CV_MAKETYPE(CV_32F, 1)

namespace cv {

namespace Error { enum { StsBadArg = -5, StsBadSize = -201, StsAssert = -215, GpuApiCallError = -217, StsNotImplemented = -213 }; }
class Exception : public std::runtime_error {
public:
    int code;
    Exception(int c, const std::string& m) : std::runtime_error(m), code(c) {}
};

struct Size { int width = 0, height = 0; Size() {} Size(int w, int h) : width(w), height(h) {} };
struct Rect {
    int x = 0, y = 0, width = 0, height = 0;
    Rect() {}
    Rect(int _x, int _y, int w, int h) : x(_x), y(_y), width(w), height(h) {}
    int area() const { return width * height; }
};

class Mat {
public:
    int rows = 0, cols = 0, flags = 0;
    size_t step = 0;
    unsigned char* data = nullptr;

    Mat() {}
    Mat(int r, int c, int t) { create(r, c, t); }
    Mat(int r, int c, int t, void* ext, size_t st) : rows(r), cols(c), flags(t), step(st), data((unsigned char*)ext) {}
    Mat(const Mat& m, const Rect& r) : rows(r.height), cols(r.width), flags(m.flags), step(m.step), buf_(m.buf_)
    {
        if (r.x < 0 || r.y < 0 || r.width < 0 || r.height < 0 || r.x + r.width > m.cols || r.y + r.height > m.rows)
            throw Exception(Error::StsAssert, "ROI outside the matrix");
        data = m.data + (size_t)r.y * m.step + (size_t)r.x * m.elemSize();
    }
    void create(int r, int c, int t)
    {
        if (data && r == rows && c == cols && t == flags) return;
        rows = r; cols = c; flags = t; step = (size_t)c * elemSize();
        buf_.reset(new unsigned char[step * (size_t)(r > 0 ? r : 1)], std::default_delete<unsigned char[]>());
        data = buf_.get();
    }
    int type() const { return flags; }
    int depth() const { return CV_MAT_DEPTH(flags); }
    int channels() const { return CV_MAT_CN(flags); }
    size_t elemSize() const { static const size_t s[] = { 1, 1, 2, 2, 4, 4, 8, 2 }; return s[depth()] * channels(); }
    bool empty() const { return !data || rows == 0 || cols == 0; }
    Size size() const { return Size(cols, rows); }
    template <typename T> T* ptr(int i = 0) { return (T*)(data + (size_t)i * step); }
    template <typename T> const T* ptr(int i = 0) const { return (const T*)(data + (size_t)i * step); }
    Mat clone() const
    {
        Mat m(rows, cols, flags);
        for (int i = 0; i < rows; ++i) std::memcpy(m.data + (size_t)i * m.step, data + (size_t)i * step, (size_t)cols * elemSize());
        return m;
    }

private:
    std::shared_ptr<unsigned char> buf_;
};

// imread stand-in for readGT: binary netpbm only (P5 -> CV_8UC1, P6 -> CV_8UC3 in BGR order, like cv::imread(IMREAD_UNCHANGED)
// of the same pixels).  With real OpenCV the facade calls cv::imread and reads whatever imgcodecs decodes.
enum { IMREAD_UNCHANGED = -1 };
inline Mat imread(const std::string& path, int = IMREAD_UNCHANGED)
{
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) return Mat();
    char magic[3] = { 0, 0, 0 };
    int w = 0, h = 0, maxv = 0;
    if (std::fscanf(f, "%2s %d %d %d", magic, &w, &h, &maxv) != 4 || magic[0] != 'P' || (magic[1] != '5' && magic[1] != '6') || maxv != 255 || w <= 0 || h <= 0) {
        std::fclose(f);
        return Mat();
    }
    std::fgetc(f);
    const int cn = magic[1] == '6' ? 3 : 1;
    Mat m(h, w, CV_MAKETYPE(CV_8U, cn));
    for (int i = 0; i < h; ++i) {
        unsigned char* r = m.ptr<unsigned char>(i);
        if (std::fread(r, 1, (size_t)w * cn, f) != (size_t)w * cn) { std::fclose(f); return Mat(); }
        if (cn == 3) for (int j = 0; j < w; ++j) { const unsigned char t = r[3 * j]; r[3 * j] = r[3 * j + 2]; r[3 * j + 2] = t; }
    }
    std::fclose(f);
    return m;
}
typedef std::string String;

typedef const Mat& InputArray;
typedef Mat& OutputArray;
template <typename T> using Ptr = std::shared_ptr<T>;
template <typename T, typename... A> Ptr<T> makePtr(A&&... a) { return std::make_shared<T>(std::forward<A>(a)...); }

class Algorithm { public: virtual ~Algorithm() {} };

// calib3d's matcher interface, as far as createDisparityWLSFilter / createRightMatcher use it
// (reference: src/disparity_filters.cpp:385-449)
class StereoMatcher : public Algorithm {
public:
    virtual int getMinDisparity() const = 0;    virtual void setMinDisparity(int) = 0;
    virtual int getNumDisparities() const = 0;  virtual void setNumDisparities(int) = 0;
    virtual int getBlockSize() const = 0;       virtual void setBlockSize(int) = 0;
    virtual int getSpeckleWindowSize() const = 0; virtual void setSpeckleWindowSize(int) = 0;
    virtual int getDisp12MaxDiff() const = 0;   virtual void setDisp12MaxDiff(int) = 0;
    // cv::StereoMatcher::compute; the stand-in classes below only hold parameters, cv::ximgproc::StereoBMB200 computes
    virtual void compute(InputArray, InputArray, OutputArray) { throw Exception(Error::StsNotImplemented, "this matcher stand-in holds parameters only"); }
};
#define DWLS_PROP(T, Name, field) T get##Name() const override { return field; } void set##Name(T v) override { field = v; }
#define DWLS_PROP_NEW(T, Name, field) virtual T get##Name() const { return field; } virtual void set##Name(T v) { field = v; }
class StereoBM : public StereoMatcher {
public:
    int minDisparity = 0, numDisparities = 64, blockSize = 21, speckleWindowSize = 0, disp12MaxDiff = -1, textureThreshold = 10, uniquenessRatio = 15, preFilterCap = 31;
    DWLS_PROP(int, MinDisparity, minDisparity) DWLS_PROP(int, NumDisparities, numDisparities) DWLS_PROP(int, BlockSize, blockSize)
    DWLS_PROP(int, SpeckleWindowSize, speckleWindowSize) DWLS_PROP(int, Disp12MaxDiff, disp12MaxDiff)
    DWLS_PROP_NEW(int, TextureThreshold, textureThreshold) DWLS_PROP_NEW(int, UniquenessRatio, uniquenessRatio) DWLS_PROP_NEW(int, PreFilterCap, preFilterCap)
    static Ptr<StereoBM> create(int numDisparities = 0, int blockSize = 21)
    {
        Ptr<StereoBM> m = std::make_shared<StereoBM>();
        m->numDisparities = numDisparities ? numDisparities : 64; m->blockSize = blockSize;
        return m;
    }
};
class StereoSGBM : public StereoMatcher {
public:
    int minDisparity = 0, numDisparities = 16, blockSize = 3, speckleWindowSize = 0, disp12MaxDiff = 0, P1 = 0, P2 = 0, preFilterCap = 0, uniquenessRatio = 0, mode = 0;
    DWLS_PROP(int, MinDisparity, minDisparity) DWLS_PROP(int, NumDisparities, numDisparities) DWLS_PROP(int, BlockSize, blockSize)
    DWLS_PROP(int, SpeckleWindowSize, speckleWindowSize) DWLS_PROP(int, Disp12MaxDiff, disp12MaxDiff)
    DWLS_PROP_NEW(int, P1, P1) DWLS_PROP_NEW(int, P2, P2) DWLS_PROP_NEW(int, PreFilterCap, preFilterCap)
    DWLS_PROP_NEW(int, UniquenessRatio, uniquenessRatio) DWLS_PROP_NEW(int, Mode, mode)
    static Ptr<StereoSGBM> create(int minDisparity = 0, int numDisparities = 16, int blockSize = 3)
    {
        Ptr<StereoSGBM> m = std::make_shared<StereoSGBM>();
        m->minDisparity = minDisparity; m->numDisparities = numDisparities; m->blockSize = blockSize;
        return m;
    }
};

}  // namespace cv
#endif
