// oracle/shim/ref_capi.cpp — TEST INFRASTRUCTURE.  C entry points over the REFERENCE's own public API
// (cv::ximgproc::fastGlobalSmootherFilter, createDisparityWLSFilterGeneric + filter), for the copy of
// the reference compiled from /root/reference by `make -C oracle ref`.  Used to pin the CPU oracle
// (tests/test_oracle_vs_ref.py, tests/golden/make_golden.py) and as the CPU baseline of bench.py.
#include <opencv2/core.hpp>
#include <opencv2/ximgproc.hpp>

using namespace cv;

static int depth_type(int depth, int cn) { return CV_MAKETYPE(depth, cn); }

extern "C" {

int ref_set_num_threads(int n) { setNumThreads(n); return getNumThreads(); }

// fastGlobalSmootherFilter(guide, src, dst, lambda, sigma_color, lambda_attenuation, num_iter)
int ref_fgs_filter(const unsigned char* guide, long long gstep, int w, int h, int gcn, const void* src, long long sstep, int sdepth, int scn,
                   void* dst, long long dstep, double lambda, double sigma, double att, int num_iter)
{
    try {
        Mat g(h, w, depth_type(CV_8U, gcn), (void*)guide, (size_t)gstep);
        Mat s(h, w, depth_type(sdepth, scn), (void*)src, (size_t)sstep);
        Mat d;
        ximgproc::fastGlobalSmootherFilter(g, s, d, lambda, sigma, att, num_iter);
        Mat out(h, w, d.type(), dst, (size_t)dstep);
        d.copyTo(out);
        return 0;
    } catch (const cv::Exception& e) { return e.code == Error::StsBadSize ? 2 : 1; }
}

// DisparityWLSFilter::filter through createDisparityWLSFilterGeneric; roi = user ROI (must have non-zero area,
// the generic factory has no margins).  conf_out (view size) may be NULL.
int ref_wls_filter(const short* dl, long long dlstep, const short* dr, long long drstep, int dw, int dh, const unsigned char* view, long long vstep,
                   int W, int H, int vcn, int use_conf, double lambda, double sigma, int lrc_thresh, int radius, const int* roi,
                   short* out, long long ostep, float* conf_out, long long cstep, int* roi_out)
{
    try {
        Mat L(dh, dw, CV_16SC1, (void*)dl, (size_t)dlstep), V(H, W, depth_type(CV_8U, vcn), (void*)view, (size_t)vstep);
        Mat R = dr ? Mat(dh, dw, CV_16SC1, (void*)dr, (size_t)drstep) : Mat();
        Ptr<ximgproc::DisparityWLSFilter> f = ximgproc::createDisparityWLSFilterGeneric(use_conf != 0);
        f->setLambda(lambda); f->setSigmaColor(sigma); f->setLRCthresh(lrc_thresh); f->setDepthDiscontinuityRadius(radius);
        Mat res;
        f->filter(L, V, res, R, Rect(roi[0], roi[1], roi[2], roi[3]));
        Mat o(H, W, CV_16SC1, out, (size_t)ostep);
        res.copyTo(o);
        if (conf_out && use_conf) { Mat c = f->getConfidenceMap(); Mat co(c.rows, c.cols, CV_32FC1, conf_out, (size_t)cstep); c.copyTo(co); }
        if (roi_out) { Rect r = f->getROI(); roi_out[0] = r.x; roi_out[1] = r.y; roi_out[2] = r.width; roi_out[3] = r.height; }
        return 0;
    } catch (const cv::Exception&) { return 1; }
}

// createDisparityWLSFilter(matcher) geometry probe: kind 0 = StereoBM, 1 = StereoSGBM.  Filters a frame to expose the derived ROI.
int ref_factory_roi(int kind, int min_disp, int num_disp, int block, int dw, int dh, int* roi_out, int* radius_out, int* right_min_disp)
{
    try {
        Ptr<StereoMatcher> m;
        if (kind == 0) { Ptr<StereoBM> bm = StereoBM::create(num_disp, block); bm->setMinDisparity(min_disp); m = bm; }
        else m = StereoSGBM::create(min_disp, num_disp, block);
        Ptr<ximgproc::DisparityWLSFilter> f = ximgproc::createDisparityWLSFilter(m);
        Ptr<StereoMatcher> r = ximgproc::createRightMatcher(m);
        *radius_out = f->getDepthDiscontinuityRadius();
        *right_min_disp = r->getMinDisparity();
        Mat L = Mat::zeros(dh, dw, CV_16SC1), R = Mat::zeros(dh, dw, CV_16SC1), V = Mat::zeros(dh, dw, CV_8UC1), res;
        f->setLambda(1.0);
        f->filter(L, V, res, R);
        Rect q = f->getROI();
        roi_out[0] = q.x; roi_out[1] = q.y; roi_out[2] = q.width; roi_out[3] = q.height;
        return 0;
    } catch (const cv::Exception&) { return 1; }
}

// quality / IO helpers of the sample (disparity_filters.cpp:460-556), the reference's own code
int ref_read_gt(const char* path, short* dst, int cap, int* rows, int* cols)
{
    try {
        Mat d;
        const int rc = ximgproc::readGT(path, d);
        *rows = d.rows; *cols = d.cols;
        if (rc == 0 && (long long)d.rows * d.cols <= cap) { Mat o(d.rows, d.cols, CV_16SC1, dst, (size_t)d.cols * 2); d.copyTo(o); }
        return rc;
    } catch (const cv::Exception&) { return -1; }
}
int ref_compute_mse(const short* gt, const short* src, int w, int h, const int* roi, double* out)
{
    try {
        Mat G(h, w, CV_16SC1, (void*)gt, (size_t)w * 2), S(h, w, CV_16SC1, (void*)src, (size_t)w * 2);
        *out = ximgproc::computeMSE(G, S, Rect(roi[0], roi[1], roi[2], roi[3]));
        return 0;
    } catch (const cv::Exception&) { return 1; }
}
int ref_bad_pixel_percent(const short* gt, const short* src, int w, int h, const int* roi, int thresh, double* out)
{
    try {
        Mat G(h, w, CV_16SC1, (void*)gt, (size_t)w * 2), S(h, w, CV_16SC1, (void*)src, (size_t)w * 2);
        *out = ximgproc::computeBadPixelPercent(G, S, Rect(roi[0], roi[1], roi[2], roi[3]), thresh);
        return 0;
    } catch (const cv::Exception&) { return 1; }
}
int ref_disparity_vis(const short* src, int w, int h, double scale, unsigned char* dst)
{
    try {
        Mat S(h, w, CV_16SC1, (void*)src, (size_t)w * 2), d;
        ximgproc::getDisparityVis(S, d, scale);
        Mat o(h, w, CV_8UC1, dst, (size_t)w);
        d.copyTo(o);
        return 0;
    } catch (const cv::Exception&) { return 1; }
}

}  // extern "C"
