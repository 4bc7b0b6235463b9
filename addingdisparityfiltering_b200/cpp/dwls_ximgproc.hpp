// dwls_ximgproc.hpp — header-only C++ host facade: the reference's public API for this path, same
// names, signatures and error behaviour, implemented on top of the C-ABI in include/dwls_b200.h.
//   reference: include/opencv2/ximgproc/disparity_filter.hpp:52-149  (DisparityFilter, DisparityWLSFilter,
//              createDisparityWLSFilter, createRightMatcher, createDisparityWLSFilterGeneric)
//              include/opencv2/ximgproc/edge_filter.hpp:361-413      (FastGlobalSmootherFilter,
//              createFastGlobalSmootherFilter, fastGlobalSmootherFilter)
// Host code stays C++ on cv::Mat; everything numerical happens in the CUDA library.  Compiles
// against real OpenCV when its headers are found, else against the stand-in in mini_cv.hpp.
// Define DWLS_NAMESPACE to place the classes somewhere other than cv::ximgproc (e.g. to link
// next to an OpenCV build that already has ximgproc).
#ifndef DWLS_XIMGPROC_HPP
#define DWLS_XIMGPROC_HPP

#if defined(DWLS_USE_OPENCV) || (defined(__has_include) && __has_include(<opencv2/core.hpp>) && !defined(DWLS_USE_MINI_CV))
#include <opencv2/calib3d.hpp>
#include <opencv2/core.hpp>
#include <opencv2/imgcodecs.hpp>
#define DWLS_REAL_OPENCV 1
#else
#include "mini_cv.hpp"
#define DWLS_REAL_OPENCV 0
#endif

#include <cmath>
#include <cstdlib>
#include <string>

#include "../../include/dwls_b200.h"

#ifndef DWLS_NAMESPACE
#define DWLS_NAMESPACE ximgproc
#endif

namespace cv {
namespace DWLS_NAMESPACE {

namespace detail {
#if DWLS_REAL_OPENCV
inline Mat mat_of(InputArray a) { return a.getMat(); }
inline void create_out(OutputArray o, int rows, int cols, int type) { o.create(rows, cols, type); }
inline Mat out_mat(OutputArray o) { return o.getMat(); }
inline bool arr_empty(InputArray a) { return a.empty(); }
#else
inline const Mat& mat_of(InputArray a) { return a; }
inline void create_out(OutputArray o, int rows, int cols, int type) { o.create(rows, cols, type); }
inline Mat& out_mat(OutputArray o) { return o; }
inline bool arr_empty(InputArray a) { return a.empty(); }
#endif
inline dwls_image image_of(const Mat& m)
{
    dwls_image im;
    im.data = (void*)m.data; im.step = (int64_t)m.step; im.rows = m.rows; im.cols = m.cols; im.depth = m.depth(); im.channels = m.channels();
    return im;
}
// status -> the exception the reference would have raised (CV_Assert / CV_Error)
inline void check(dwls_ctx* ctx, dwls_status s)
{
    if (s == DWLS_OK) return;
    const std::string msg = dwls_last_error(ctx);
    throw Exception(s == DWLS_BAD_SIZE ? Error::StsBadSize : s == DWLS_BAD_ARG ? Error::StsAssert
                    : s == DWLS_UNSUPPORTED ? Error::StsNotImplemented : Error::GpuApiCallError, msg
#if DWLS_REAL_OPENCV
                    , "dwls_b200", __FILE__, __LINE__
#endif
    );
}
// one context per thread, created on first use (the reference's objects are not re-entrant either)
inline dwls_ctx* thread_ctx()
{
    struct Holder { dwls_ctx* c = nullptr; ~Holder() { dwls_ctx_destroy(c); } };
    static thread_local Holder h;
    if (!h.c) {
        dwls_status s = dwls_ctx_create(0, &h.c);
        if (s != DWLS_OK) check(nullptr, s);
    }
    return h.c;
}
}  // namespace detail

class DisparityFilter : public Algorithm {
public:
    virtual void filter(InputArray disparity_map_left, InputArray left_view, OutputArray filtered_disparity_map,
                        InputArray disparity_map_right = Mat(), Rect ROI = Rect(), InputArray right_view = Mat()) = 0;
};

class DisparityWLSFilter : public DisparityFilter {
public:
    virtual double getLambda() = 0;
    virtual void setLambda(double _lambda) = 0;
    virtual double getSigmaColor() = 0;
    virtual void setSigmaColor(double _sigma_color) = 0;
    virtual int getLRCthresh() = 0;
    virtual void setLRCthresh(int _LRC_thresh) = 0;
    virtual int getDepthDiscontinuityRadius() = 0;
    virtual void setDepthDiscontinuityRadius(int _disc_radius) = 0;
    virtual Mat getConfidenceMap() = 0;
    virtual Rect getROI() = 0;
};

// DisparityWLSFilterImpl (src/disparity_filters.cpp:49-298): parameters on the host, filter() on the GPU
class DisparityWLSFilterB200 : public DisparityWLSFilter {
    dwls_params p_;
    Mat confidence_map_;
    Rect roi_;

public:
    DisparityWLSFilterB200(bool use_confidence, int l_offs = 0, int r_offs = 0, int t_offs = 0, int b_offs = 0)
    {
        dwls_default_params(&p_);                       // init(): lambda 8000, sigma 1.0, LRC 24, radius 5 (:142-159, :215)
        p_.use_confidence = use_confidence ? 1 : 0;
        p_.left_offset = l_offs; p_.right_offset = r_offs; p_.top_offset = t_offs; p_.bottom_offset = b_offs;
    }
    double getLambda() override { return p_.lambda; }
    void setLambda(double v) override { p_.lambda = v; }
    double getSigmaColor() override { return p_.sigma_color; }
    void setSigmaColor(double v) override { p_.sigma_color = v; }
    int getLRCthresh() override { return p_.lrc_thresh; }
    void setLRCthresh(int v) override { p_.lrc_thresh = v; }
    int getDepthDiscontinuityRadius() override { return p_.disc_radius; }
    void setDepthDiscontinuityRadius(int v) override { p_.disc_radius = v; }
    Mat getConfidenceMap() override { return confidence_map_; }
    Rect getROI() override { return roi_; }

    void filter(InputArray disparity_map_left, InputArray left_view, OutputArray filtered_disparity_map, InputArray disparity_map_right = Mat(),
                Rect ROI = Rect(), InputArray = Mat()) override
    {
        dwls_ctx* ctx = detail::thread_ctx();
        const Mat dl = detail::mat_of(disparity_map_left), view = detail::mat_of(left_view);
        if (dl.empty() || dl.depth() != CV_16S || dl.channels() != 1 || view.empty())
            detail::check(ctx, DWLS_BAD_ARG);           // message filled below by the library for the remaining cases
        const bool has_right = !detail::arr_empty(disparity_map_right);
        const Mat dr = has_right ? detail::mat_of(disparity_map_right) : Mat();
        detail::create_out(filtered_disparity_map, view.rows, view.cols, CV_16SC1);      // :252 / :282
        Mat out = detail::out_mat(filtered_disparity_map);
        dwls_image i_dl = detail::image_of(dl), i_dr = detail::image_of(dr), i_view = detail::image_of(view), i_out = detail::image_of(out), i_conf;
        if (p_.use_confidence) { confidence_map_.create(view.rows, view.cols, CV_32FC1); i_conf = detail::image_of(confidence_map_); }
        const int roi_in[4] = { ROI.x, ROI.y, ROI.width, ROI.height };
        int roi_out[4] = { 0, 0, 0, 0 };
        const dwls_status s = dwls_filter_host(ctx, &i_dl, has_right ? &i_dr : nullptr, &i_view, ROI.area() != 0 ? roi_in : nullptr, &p_, &i_out,
                                               p_.use_confidence ? &i_conf : nullptr, roi_out);
        roi_ = Rect(roi_out[0], roi_out[1], roi_out[2], roi_out[3]);
        detail::check(ctx, s);
    }
};

// src/disparity_filters.cpp:451-455
inline Ptr<DisparityWLSFilter> createDisparityWLSFilterGeneric(bool use_confidence)
{
    return Ptr<DisparityWLSFilter>(new DisparityWLSFilterB200(use_confidence));
}

#if !DWLS_REAL_OPENCV
// cv::StereoBM whose compute() runs on the GPU (dwls_stereo_bm_host): the matcher of the sample
// (samples/disparity_filtering.cpp:148-160) with the disparities cv::StereoBM::compute returns, bit for bit.  A matcher made
// by createRightMatcher from one of these is one of these.  (With a real OpenCV build use cv::StereoBM itself or call
// dwls_stereo_bm_host directly: cv::StereoBM has more pure virtuals than this facade needs to restate.)
class StereoBMB200 : public StereoBM {
public:
    static Ptr<StereoBMB200> create(int numDisparities = 0, int blockSize = 21)
    {
        Ptr<StereoBMB200> m = std::make_shared<StereoBMB200>();
        m->numDisparities = numDisparities ? numDisparities : 64; m->blockSize = blockSize;
        return m;
    }
    void compute(InputArray left_, InputArray right_, OutputArray disparity_) override
    {
        dwls_ctx* ctx = detail::thread_ctx();
        const Mat left = detail::mat_of(left_), right = detail::mat_of(right_);
        if (left.empty() || right.empty() || left.rows != right.rows || left.cols != right.cols || left.type() != right.type()) detail::check(nullptr, DWLS_BAD_ARG);
        if (speckleWindowSize > 0 || (disp12MaxDiff >= 0 && disp12MaxDiff < 1000000)) detail::check(nullptr, DWLS_UNSUPPORTED);
        detail::create_out(disparity_, left.rows, left.cols, CV_16SC1);
        Mat disp = detail::out_mat(disparity_);
        dwls_bm_params p;
        p.num_disparities = numDisparities; p.block_size = blockSize; p.min_disparity = minDisparity; p.pre_filter_cap = preFilterCap;
        p.texture_threshold = textureThreshold; p.uniqueness_ratio = uniquenessRatio;
        dwls_image l = detail::image_of(left), r = detail::image_of(right), d = detail::image_of(disp);
        detail::check(ctx, dwls_stereo_bm_host(ctx, &l, &r, &p, &d, nullptr));
    }
};
#endif

// src/disparity_filters.cpp:385-414 — mutates the matcher, derives ROI margins and the discontinuity radius
inline Ptr<DisparityWLSFilter> createDisparityWLSFilter(Ptr<StereoMatcher> matcher_left)
{
    Ptr<DisparityWLSFilter> wls;
    matcher_left->setDisp12MaxDiff(1000000);
    matcher_left->setSpeckleWindowSize(0);
    const int min_disp = matcher_left->getMinDisparity(), num_disp = matcher_left->getNumDisparities(), wsize = matcher_left->getBlockSize();
    const int wsize2 = wsize / 2;
    if (Ptr<StereoBM> bm = std::dynamic_pointer_cast<StereoBM>(matcher_left)) {
        bm->setTextureThreshold(0);
        bm->setUniquenessRatio(0);
        wls = Ptr<DisparityWLSFilter>(new DisparityWLSFilterB200(true, std::max(0, min_disp + num_disp) + wsize2, std::max(0, -min_disp) + wsize2, wsize2, wsize2));
        wls->setDepthDiscontinuityRadius((int)std::ceil(0.33 * wsize));
    } else if (Ptr<StereoSGBM> sgbm = std::dynamic_pointer_cast<StereoSGBM>(matcher_left)) {
        sgbm->setUniquenessRatio(0);
        wls = Ptr<DisparityWLSFilter>(new DisparityWLSFilterB200(true, std::max(0, min_disp + num_disp), std::max(0, -min_disp), 0, 0));
        wls->setDepthDiscontinuityRadius((int)std::ceil(0.5 * wsize));
    } else
        detail::check(nullptr, DWLS_BAD_ARG);           // "DisparityWLSFilter natively supports only StereoBM and StereoSGBM" (:411)
    return wls;
}

// src/disparity_filters.cpp:416-449
inline Ptr<StereoMatcher> createRightMatcher(Ptr<StereoMatcher> matcher_left)
{
    const int min_disp = matcher_left->getMinDisparity(), num_disp = matcher_left->getNumDisparities(), wsize = matcher_left->getBlockSize();
    if (std::dynamic_pointer_cast<StereoBM>(matcher_left)) {
#if !DWLS_REAL_OPENCV
        Ptr<StereoBM> right_bm = std::dynamic_pointer_cast<StereoBMB200>(matcher_left) ? Ptr<StereoBM>(StereoBMB200::create(num_disp, wsize)) : StereoBM::create(num_disp, wsize);
        right_bm->setPreFilterCap(std::dynamic_pointer_cast<StereoBM>(matcher_left)->getPreFilterCap());
#else
        Ptr<StereoBM> right_bm = StereoBM::create(num_disp, wsize);
#endif
        right_bm->setMinDisparity(-(min_disp + num_disp) + 1);
        right_bm->setTextureThreshold(0);
        right_bm->setUniquenessRatio(0);
        right_bm->setDisp12MaxDiff(1000000);
        right_bm->setSpeckleWindowSize(0);
        return right_bm;
    }
    if (Ptr<StereoSGBM> sgbm = std::dynamic_pointer_cast<StereoSGBM>(matcher_left)) {
        Ptr<StereoSGBM> right_sgbm = StereoSGBM::create(-(min_disp + num_disp) + 1, num_disp, wsize);
        right_sgbm->setUniquenessRatio(0);
        right_sgbm->setP1(sgbm->getP1());
        right_sgbm->setP2(sgbm->getP2());
        right_sgbm->setMode(sgbm->getMode());
        right_sgbm->setPreFilterCap(sgbm->getPreFilterCap());
        right_sgbm->setDisp12MaxDiff(1000000);
        right_sgbm->setSpeckleWindowSize(0);
        return right_sgbm;
    }
    detail::check(nullptr, DWLS_BAD_ARG);               // "createRightMatcher supports only StereoBM and StereoSGBM" (:446)
    return Ptr<StereoMatcher>();
}

class FastGlobalSmootherFilter : public Algorithm {
public:
    virtual void filter(InputArray src, OutputArray dst) = 0;
};

// FastGlobalSmootherFilterImpl (src/fgs_filter.cpp:61-233): keeps the guide and the parameters; weights are rebuilt on the
// device per call (the reference rebuilds them per create(), and DisparityWLSFilter creates one per filter(), :292)
class FastGlobalSmootherFilterB200 : public FastGlobalSmootherFilter {
    Mat guide_;
    double lambda_, sigma_, att_;
    int num_iter_;

public:
    FastGlobalSmootherFilterB200(const Mat& guide, double lambda, double sigma_color, double lambda_attenuation, int num_iter)
        : guide_(guide.clone()), lambda_(lambda), sigma_(sigma_color), att_(lambda_attenuation), num_iter_(num_iter)
    {
        if (guide.empty() || !(lambda >= 0) || !(sigma_color >= 0) || num_iter < 1 || guide.depth() != CV_8U || (guide.channels() != 1 && guide.channels() != 3))
            detail::check(nullptr, DWLS_BAD_ARG);        // CV_Assert (:143-144)
    }
    void filter(InputArray src_, OutputArray dst_) override
    {
        dwls_ctx* ctx = detail::thread_ctx();
        const Mat src = detail::mat_of(src_).clone();    // filter(src, src) is legal in the reference
        detail::create_out(dst_, src.rows, src.cols, src.type());
        Mat dst = detail::out_mat(dst_);
        dwls_image g = detail::image_of(guide_), s = detail::image_of(src), d = detail::image_of(dst);
        detail::check(ctx, dwls_fgs_host(ctx, &g, &s, &d, lambda_, sigma_, att_, num_iter_));
    }
};

// include/opencv2/ximgproc/edge_filter.hpp:393, :413
inline Ptr<FastGlobalSmootherFilter> createFastGlobalSmootherFilter(InputArray guide, double lambda, double sigma_color, double lambda_attenuation = 0.25,
                                                                    int num_iter = 3)
{
    return Ptr<FastGlobalSmootherFilter>(new FastGlobalSmootherFilterB200(detail::mat_of(guide), lambda, sigma_color, lambda_attenuation, num_iter));
}
inline void fastGlobalSmootherFilter(InputArray guide, InputArray src, OutputArray dst, double lambda, double sigma_color, double lambda_attenuation = 0.25,
                                     int num_iter = 3)
{
    createFastGlobalSmootherFilter(guide, lambda, sigma_color, lambda_attenuation, num_iter)->filter(src, dst);
}

// ---- ground-truth / quality helpers of the sample (include/opencv2/ximgproc/disparity_filter.hpp:163-204,
// src/disparity_filters.cpp:460-556): host-side like the reference's, a few integer operations per pixel ----
namespace detail {
enum { UNKNOWN_DISPARITY = 16320 };                                         // :460
inline void check_s16_pair(const Mat& GT, const Mat& src, const Rect& ROI)
{
    if (GT.empty() || GT.depth() != CV_16S || GT.channels() != 1 || src.empty() || src.depth() != CV_16S || src.channels() != 1 ||
        src.rows != GT.rows || src.cols != GT.cols)                         // CV_Assert (:499-501, :522-524)
        throw Exception(Error::StsAssert, "GT and src must be non-empty CV_16SC1 images of equal size"
#if DWLS_REAL_OPENCV
                        , "dwls_b200", __FILE__, __LINE__
#endif
        );
    if (ROI.x < 0 || ROI.y < 0 || ROI.width < 0 || ROI.height < 0 || ROI.x + ROI.width > GT.cols || ROI.y + ROI.height > GT.rows)
        throw Exception(Error::StsAssert, "ROI outside the disparity map"   // Mat(m, roi) asserts
#if DWLS_REAL_OPENCV
                        , "dwls_b200", __FILE__, __LINE__
#endif
        );
}
}  // namespace detail

// readGT (:462-495): Middlebury (1 channel: 0 = unknown, else 16 x value) and MPI-Sintel (3 channels: 64 R + G / 4) formats
inline int readGT(String src_path, OutputArray dst_)
{
    const Mat src = imread(src_path, IMREAD_UNCHANGED);
    detail::create_out(dst_, src.rows, src.cols, CV_16SC1);
    Mat dst = detail::out_mat(dst_);
    if (!src.empty() && src.channels() == 3 && src.depth() == CV_8U) {
        for (int i = 0; i < src.rows; ++i) {
            const unsigned char* s = src.ptr<unsigned char>(i);
            short* d = dst.ptr<short>(i);
            for (int j = 0; j < src.cols; ++j) d[j] = (short)(64 * s[3 * j + 2] + s[3 * j + 1] / 4);
        }
        return 0;
    }
    if (!src.empty() && src.channels() == 1 && src.depth() == CV_8U) {
        for (int i = 0; i < src.rows; ++i) {
            const unsigned char* s = src.ptr<unsigned char>(i);
            short* d = dst.ptr<short>(i);
            for (int j = 0; j < src.cols; ++j) d[j] = s[j] == 0 ? (short)detail::UNKNOWN_DISPARITY : (short)(16 * s[j]);
        }
        return 0;
    }
    return 1;
}

// computeMSE (:497-517): mean squared error over the known pixels of the ROI, in pixels^2 (disparities are stored x16)
inline double computeMSE(InputArray GT_, InputArray src_, Rect ROI)
{
    const Mat GT = detail::mat_of(GT_), src = detail::mat_of(src_);
    detail::check_s16_pair(GT, src, ROI);
    double res = 0;
    int cnt = 0;
    for (int i = 0; i < ROI.height; ++i) {
        const short* g = GT.ptr<short>(ROI.y + i) + ROI.x;
        const short* s = src.ptr<short>(ROI.y + i) + ROI.x;
        for (int j = 0; j < ROI.width; ++j)
            if (g[j] != detail::UNKNOWN_DISPARITY) { res += (g[j] - s[j]) * (g[j] - s[j]); cnt++; }
    }
    return res / (cnt * 256);
}

// computeBadPixelPercent (:519-540)
inline double computeBadPixelPercent(InputArray GT_, InputArray src_, Rect ROI, int thresh = 24 /*1.5 pixels*/)
{
    const Mat GT = detail::mat_of(GT_), src = detail::mat_of(src_);
    detail::check_s16_pair(GT, src, ROI);
    int bad = 0, cnt = 0;
    for (int i = 0; i < ROI.height; ++i) {
        const short* g = GT.ptr<short>(ROI.y + i) + ROI.x;
        const short* s = src.ptr<short>(ROI.y + i) + ROI.x;
        for (int j = 0; j < ROI.width; ++j)
            if (g[j] != detail::UNKNOWN_DISPARITY) { if (std::abs(g[j] - s[j]) >= thresh) bad++; cnt++; }
    }
    return (100.0 * bad) / cnt;
}

// getDisparityVis (:542-556): saturate_cast<uchar>(scale * d / 16.0), unknown pixels black
inline void getDisparityVis(InputArray src_, OutputArray dst_, double scale = 1.0)
{
    const Mat src = detail::mat_of(src_);
    if (src.empty() || src.depth() != CV_16S || src.channels() != 1) detail::check(nullptr, DWLS_BAD_ARG);
    detail::create_out(dst_, src.rows, src.cols, CV_8UC1);
    Mat dst = detail::out_mat(dst_);
    for (int i = 0; i < src.rows; ++i) {
        const short* s = src.ptr<short>(i);
        unsigned char* d = dst.ptr<unsigned char>(i);
        for (int j = 0; j < src.cols; ++j) {
            if (s[j] == detail::UNKNOWN_DISPARITY) { d[j] = 0; continue; }
            const long r = std::lrint(scale * s[j] / 16.0);                 // cvRound: round half to even
            d[j] = (unsigned char)(r < 0 ? 0 : r > 255 ? 255 : r);
        }
    }
}

}  // namespace DWLS_NAMESPACE
}  // namespace cv
#endif
