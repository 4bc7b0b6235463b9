#include <cstdio>
// oracle/shim/opencv2/shim_imgproc.hpp — TEST INFRASTRUCTURE (see core.hpp): the imgproc / imgcodecs /
// calib3d entry points the reference calls.  Rounding sequences as verified against cv2 (non-IPP
// build path) in tests/test_oracle_primitives.py.
#ifndef SHIM_OPENCV2_IMGPROC_HPP
#define SHIM_OPENCV2_IMGPROC_HPP

namespace cv {

enum { BORDER_REFLECT_101 = 4, BORDER_DEFAULT = 4, INTER_LINEAR = 1, IMREAD_UNCHANGED = -1, IMREAD_GRAYSCALE = 0, IMREAD_COLOR = 1 };

inline int borderInterpolate(int p, int len, int = BORDER_REFLECT_101)
{
    if (len == 1) return 0;
    while (p < 0 || p >= len) p = p < 0 ? -p : 2 * (len - 1) - p;
    return p;
}

// boxFilter(16S -> 32F, normalized): exact int window sum, float(sum) * float(1/k^2)
inline void boxFilter(InputArray _src, OutputArray _dst, int ddepth, Size ksize, Point = Point(-1, -1), bool normalize = true, int = BORDER_DEFAULT)
{
    Mat src = _src.getMat();
    CV_Assert(src.type() == CV_16SC1 && ddepth == CV_32F && normalize);
    const int w = src.cols, h = src.rows, rx = ksize.width / 2, ry = ksize.height / 2;
    const float scale = (float)(1.0 / (ksize.width * ksize.height));
    std::vector<int> hs((size_t)w * h);
    for (int i = 0; i < h; ++i) { const short* s = src.ptr<short>(i);
        for (int j = 0; j < w; ++j) { int a = 0; for (int d = -rx; d <= rx; ++d) a += s[borderInterpolate(j + d, w)]; hs[(size_t)i * w + j] = a; } }
    Mat out(h, w, CV_32FC1);
    for (int i = 0; i < h; ++i) { float* o = out.ptr<float>(i);
        for (int j = 0; j < w; ++j) { int a = 0; for (int d = -ry; d <= ry; ++d) a += hs[(size_t)borderInterpolate(i + d, h) * w + j]; o[j] = (float)a * scale; } }
    _dst.getMatRef() = out;
}
// sqrBoxFilter(16S -> 32F, normalized): double sum of squares * double(1/k^2), cast to float
inline void sqrBoxFilter(InputArray _src, OutputArray _dst, int ddepth, Size ksize, Point = Point(-1, -1), bool normalize = true, int = BORDER_DEFAULT)
{
    Mat src = _src.getMat();
    CV_Assert(src.type() == CV_16SC1 && ddepth == CV_32F && normalize);
    const int w = src.cols, h = src.rows, rx = ksize.width / 2, ry = ksize.height / 2;
    const double scale = 1.0 / (ksize.width * ksize.height);
    std::vector<double> hs((size_t)w * h);
    for (int i = 0; i < h; ++i) { const short* s = src.ptr<short>(i);
        for (int j = 0; j < w; ++j) { double a = 0; for (int d = -rx; d <= rx; ++d) { double v = s[borderInterpolate(j + d, w)]; a += v * v; } hs[(size_t)i * w + j] = a; } }
    Mat out(h, w, CV_32FC1);
    for (int i = 0; i < h; ++i) { float* o = out.ptr<float>(i);
        for (int j = 0; j < w; ++j) { double a = 0; for (int d = -ry; d <= ry; ++d) a += hs[(size_t)borderInterpolate(i + d, h) * w + j]; o[j] = (float)(a * scale); } }
    _dst.getMatRef() = out;
}

// resize, INTER_LINEAR, one channel, work type float: horizontal pass (border coefficients zeroed) then vertical pass (rows clipped)
template <typename T> static inline void shim_resize_linear(const Mat& src, Mat& dst)
{
    const int sw = src.cols, sh = src.rows, dw = dst.cols, dh = dst.rows;
    auto coeff = [](int ssize, int dsize, std::vector<int>& ofs, std::vector<float>& fr, bool clampc) {
        const double scale = 1.0 / ((double)dsize / ssize);
        ofs.resize(dsize); fr.resize(dsize);
        for (int d = 0; d < dsize; ++d) {
            float f = (float)((d + 0.5) * scale - 0.5); int s = cvFloor(f); f -= s;
            if (clampc) { if (s < 0) { f = 0; s = 0; } if (s >= ssize - 1) { f = 0; s = ssize - 1; } }
            ofs[d] = s; fr[d] = f;
        }
    };
    std::vector<int> xo, yo; std::vector<float> xf, yf;
    coeff(sw, dw, xo, xf, true); coeff(sh, dh, yo, yf, false);
    std::vector<float> hr((size_t)sh * dw);
    for (int i = 0; i < sh; ++i) { const T* s = src.ptr<T>(i);
        for (int d = 0; d < dw; ++d) { const int x = xo[d]; const float a1 = xf[d], a0 = 1.f - a1;
            hr[(size_t)i * dw + d] = (x >= sw - 1) ? (float)s[sw - 1] * 1.0f : (float)s[x] * a0 + (float)s[x + 1] * a1; } }
    for (int d = 0; d < dh; ++d) {
        const int y0 = std::min(std::max(yo[d], 0), sh - 1), y1 = std::min(std::max(yo[d] + 1, 0), sh - 1);
        const float b1 = yf[d], b0 = 1.f - b1; const float* S0 = &hr[(size_t)y0 * dw]; const float* S1 = &hr[(size_t)y1 * dw];
        T* o = dst.ptr<T>(d);
        for (int x = 0; x < dw; ++x) o[x] = saturate_cast<T>(S0[x] * b0 + S1[x] * b1);
    }
}
inline void resize(InputArray _src, OutputArray _dst, Size dsize, double fx = 0, double fy = 0, int interpolation = INTER_LINEAR)
{
    Mat src = _src.getMat();
    if (dsize.area() == 0) dsize = Size(cvRound(src.cols * fx), cvRound(src.rows * fy));
    CV_Assert(interpolation == INTER_LINEAR && src.channels() == 1 && (src.depth() == CV_16S || src.depth() == CV_32F));
    Mat out(dsize.height, dsize.width, src.type());
    if (src.depth() == CV_16S) shim_resize_linear<short>(src, out); else shim_resize_linear<float>(src, out);
    _dst.getMatRef() = out;
}

// imread stand-in for readGT (disparity_filters.cpp:462): binary netpbm only, P5 -> CV_8UC1, P6 -> CV_8UC3 in BGR order like
// cv::imread(IMREAD_UNCHANGED) of the same pixels; anything else -> empty Mat (no PNG/JPEG codec on this box)
inline Mat imread(const String& path, int = IMREAD_COLOR)
{
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) return Mat();
    char magic[3] = { 0, 0, 0 };
    int w = 0, h = 0, maxv = 0;
    if (fscanf(f, "%2s %d %d %d", magic, &w, &h, &maxv) != 4 || magic[0] != 'P' || (magic[1] != '5' && magic[1] != '6') || maxv != 255 || w <= 0 || h <= 0) { fclose(f); return Mat(); }
    fgetc(f);
    const int cn = magic[1] == '6' ? 3 : 1;
    Mat m(h, w, CV_MAKETYPE(CV_8U, cn));
    for (int i = 0; i < h; ++i) {
        uchar* r = m.ptr<uchar>(i);
        if (fread(r, 1, (size_t)w * cn, f) != (size_t)w * cn) { fclose(f); return Mat(); }
        if (cn == 3) for (int j = 0; j < w; ++j) { const uchar t = r[3 * j]; r[3 * j] = r[3 * j + 2]; r[3 * j + 2] = t; }
    }
    fclose(f);
    return m;
}

// ---- calib3d: the matcher interface createDisparityWLSFilter / createRightMatcher talk to ------
class StereoMatcher : public Algorithm {
public:
    int minDisparity = 0, numDisparities = 64, blockSize = 21, speckleWindowSize = 0, speckleRange = 0, disp12MaxDiff = -1;
    virtual int getMinDisparity() const { return minDisparity; }        virtual void setMinDisparity(int v) { minDisparity = v; }
    virtual int getNumDisparities() const { return numDisparities; }    virtual void setNumDisparities(int v) { numDisparities = v; }
    virtual int getBlockSize() const { return blockSize; }              virtual void setBlockSize(int v) { blockSize = v; }
    virtual int getSpeckleWindowSize() const { return speckleWindowSize; } virtual void setSpeckleWindowSize(int v) { speckleWindowSize = v; }
    virtual int getDisp12MaxDiff() const { return disp12MaxDiff; }      virtual void setDisp12MaxDiff(int v) { disp12MaxDiff = v; }
};
class StereoBM : public StereoMatcher {
public:
    int textureThreshold = 10, uniquenessRatio = 15;
    virtual int getTextureThreshold() const { return textureThreshold; } virtual void setTextureThreshold(int v) { textureThreshold = v; }
    virtual int getUniquenessRatio() const { return uniquenessRatio; }   virtual void setUniquenessRatio(int v) { uniquenessRatio = v; }
    static Ptr<StereoBM> create(int numDisparities = 0, int blockSize = 21) { StereoBM* m = new StereoBM(); m->numDisparities = numDisparities ? numDisparities : 64; m->blockSize = blockSize; return Ptr<StereoBM>(m); }
};
class StereoSGBM : public StereoMatcher {
public:
    int P1 = 0, P2 = 0, preFilterCap = 0, uniquenessRatio = 0, mode = 0;
    virtual int getP1() const { return P1; } virtual void setP1(int v) { P1 = v; }
    virtual int getP2() const { return P2; } virtual void setP2(int v) { P2 = v; }
    virtual int getPreFilterCap() const { return preFilterCap; } virtual void setPreFilterCap(int v) { preFilterCap = v; }
    virtual int getUniquenessRatio() const { return uniquenessRatio; } virtual void setUniquenessRatio(int v) { uniquenessRatio = v; }
    virtual int getMode() const { return mode; } virtual void setMode(int v) { mode = v; }
    static Ptr<StereoSGBM> create(int minDisparity, int numDisparities, int blockSize, int P1 = 0, int P2 = 0, int disp12MaxDiff = 0, int preFilterCap = 0,
                                  int uniquenessRatio = 0, int speckleWindowSize = 0, int speckleRange = 0, int mode = 0)
    {
        StereoSGBM* m = new StereoSGBM();
        m->minDisparity = minDisparity; m->numDisparities = numDisparities; m->blockSize = blockSize; m->P1 = P1; m->P2 = P2; m->disp12MaxDiff = disp12MaxDiff;
        m->preFilterCap = preFilterCap; m->uniquenessRatio = uniquenessRatio; m->speckleWindowSize = speckleWindowSize; m->speckleRange = speckleRange; m->mode = mode;
        return Ptr<StereoSGBM>(m);
    }
};

}  // namespace cv
#endif
