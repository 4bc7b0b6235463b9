// oracle/shim/opencv2/shim_mat.hpp — TEST INFRASTRUCTURE (see core.hpp): cv::Mat stand-in.
#ifndef SHIM_OPENCV2_MAT_HPP
#define SHIM_OPENCV2_MAT_HPP

namespace cv {

class Mat;
struct MatSizeView {                      // what `m.size` (data member) compares as in the reference (disparity_filters.cpp:239)
    const Mat* m;
    bool operator==(const MatSizeView& o) const;
    bool operator!=(const MatSizeView& o) const { return !(*this == o); }
    Size operator()() const;
};

class Mat {
public:
    int rows = 0, cols = 0, flags = 0;
    size_t step = 0;
    uchar* data = nullptr;
    std::shared_ptr<uchar> buf;
    MatSizeView size{ this };

    Mat() {}
    Mat(int r, int c, int t) { create(r, c, t); }
    Mat(Size s, int t) { create(s.height, s.width, t); }
    Mat(int r, int c, int t, void* ext, size_t st) : rows(r), cols(c), flags(t), step(st), data((uchar*)ext) {}      // external (not owned)
    Mat(const Mat& o) : rows(o.rows), cols(o.cols), flags(o.flags), step(o.step), data(o.data), buf(o.buf) {}
    Mat(const Mat& o, const Rect& r) : rows(r.height), cols(r.width), flags(o.flags), step(o.step), buf(o.buf)
    {
        CV_Assert(0 <= r.x && 0 <= r.width && r.x + r.width <= o.cols && 0 <= r.y && 0 <= r.height && r.y + r.height <= o.rows);
        data = o.data + (size_t)r.y * o.step + (size_t)r.x * o.elemSize();
    }
    Mat& operator=(const Mat& o)
    {
        if (this != &o) { rows = o.rows; cols = o.cols; flags = o.flags; step = o.step; data = o.data; buf = o.buf; }
        return *this;
    }
    Mat& operator=(const Scalar& s);       // fill

    int type() const { return flags; }
    int depth() const { return CV_MAT_DEPTH(flags); }
    int channels() const { return CV_MAT_CN(flags); }
    size_t elemSize() const { return elemSize1(depth()) * channels(); }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    size_t total() const { return (size_t)rows * cols; }
    bool isContinuous() const { return step == (size_t)cols * elemSize(); }

    void create(int r, int c, int t)
    {
        if (data && r == rows && c == cols && t == flags) return;          // cv::Mat::create keeps a matching buffer (ROI views included)
        rows = r; cols = c; flags = t; step = (size_t)c * elemSize();
        buf.reset(new uchar[step * (size_t)std::max(r, 1) + 64], std::default_delete<uchar[]>());
        data = buf.get();
    }
    void create(Size s, int t) { create(s.height, s.width, t); }

    uchar* ptr(int i = 0) { return data + (size_t)i * step; }
    const uchar* ptr(int i = 0) const { return data + (size_t)i * step; }
    template <typename T> T* ptr(int i = 0) { return (T*)(data + (size_t)i * step); }
    template <typename T> const T* ptr(int i = 0) const { return (const T*)(data + (size_t)i * step); }
    template <typename T> T& at(int i, int j) { return ((T*)(data + (size_t)i * step))[j]; }
    template <typename T> const T& at(int i, int j) const { return ((const T*)(data + (size_t)i * step))[j]; }

    Mat clone() const { Mat m; copyTo(m); return m; }
    void copyTo(Mat& dst) const
    {
        if (dst.data == data && dst.rows == rows && dst.cols == cols && dst.step == step) return;
        if (!(dst.data && dst.rows == rows && dst.cols == cols && dst.flags == flags)) dst.create(rows, cols, flags);
        for (int i = 0; i < rows; ++i) std::memcpy(dst.ptr(i), ptr(i), (size_t)cols * elemSize());
    }
    void convertTo(Mat& dst, int rtype, double alpha = 1, double beta = 0) const;
    Mat mul(const Mat& o) const;
    static Mat zeros(int r, int c, int t) { Mat m(r, c, t); for (int i = 0; i < r; ++i) std::memset(m.ptr(i), 0, (size_t)c * m.elemSize()); return m; }
};

inline bool MatSizeView::operator==(const MatSizeView& o) const { return m->rows == o.m->rows && m->cols == o.m->cols; }
inline Size MatSizeView::operator()() const { return Size(m->cols, m->rows); }

// InputArray / OutputArray: thin handles on a Mat (the only array kind the two reference files pass)
class _InputArray {
public:
    Mat m;                    // header copy (shares data)
    Mat* ref = nullptr;       // the caller's Mat for output arrays
    _InputArray() {}
    _InputArray(const Mat& a) : m(a) {}
    bool empty() const { return (ref ? *ref : m).empty(); }
    int depth() const { return (ref ? *ref : m).depth(); }
    int channels() const { return (ref ? *ref : m).channels(); }
    int type() const { return (ref ? *ref : m).type(); }
    int rows() const { return (ref ? *ref : m).rows; }
    int cols() const { return (ref ? *ref : m).cols; }
    Size size() const { return Size(cols(), rows()); }
    Mat getMat() const { return ref ? *ref : m; }
};
class _OutputArray : public _InputArray {
public:
    _OutputArray() {}
    _OutputArray(Mat& a) { ref = &a; }
    void create(Size s, int t) const { ref->create(s.height, s.width, t); }
    void create(int r, int c, int t) const { ref->create(r, c, t); }
    Mat& getMatRef() const { return *ref; }
};
typedef const _InputArray& InputArray;
typedef const _OutputArray& OutputArray;
typedef const _OutputArray& InputOutputArray;
inline const _OutputArray& noArray() { static _OutputArray none; return none; }

}  // namespace cv
#endif
