// Exercises the C++ facade (cv::ximgproc API on cv::Mat) end to end against libdwls_b200.so.
// usage: facade_test <dir> <dw> <dh> <W> <H> <cn> <rx> <ry> <rw> <rh> <radius> <lambda> <sigma>
// reads <dir>/dl.bin dr.bin view.bin src.bin, writes out.bin conf.bin fgs.bin, prints "ROI x y w h" lines.
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include "../../addingdisparityfiltering_b200/cpp/dwls_ximgproc.hpp"

using namespace cv;
using namespace cv::ximgproc;

static void load(const std::string& p, Mat& m)
{
    FILE* f = fopen(p.c_str(), "rb");
    if (!f || fread(m.data, 1, m.step * m.rows, f) != m.step * m.rows) { fprintf(stderr, "cannot read %s\n", p.c_str()); exit(2); }
    fclose(f);
}
static void save(const std::string& p, const Mat& m)
{
    FILE* f = fopen(p.c_str(), "wb");
    for (int i = 0; i < m.rows; ++i) fwrite(m.data + (size_t)i * m.step, 1, (size_t)m.cols * m.elemSize(), f);
    fclose(f);
}

int main(int argc, char** argv)
{
    if (argc < 14) return 2;
    const std::string dir = argv[1];
    const int dw = atoi(argv[2]), dh = atoi(argv[3]), W = atoi(argv[4]), H = atoi(argv[5]), cn = atoi(argv[6]);
    const Rect roi(atoi(argv[7]), atoi(argv[8]), atoi(argv[9]), atoi(argv[10]));
    const int radius = atoi(argv[11]);
    const double lambda = atof(argv[12]), sigma = atof(argv[13]);
    Mat dl(dh, dw, CV_16SC1), dr(dh, dw, CV_16SC1), view(H, W, CV_MAKETYPE(CV_8U, cn)), src(H, W, CV_32FC1);
    load(dir + "/dl.bin", dl); load(dir + "/dr.bin", dr); load(dir + "/view.bin", view); load(dir + "/src.bin", src);

    // 1. generic factory + setters + filter with a user ROI (the reference's ReferenceAccuracy test shape)
    Ptr<DisparityWLSFilter> wls = createDisparityWLSFilterGeneric(true);
    wls->setLambda(lambda); wls->setSigmaColor(sigma); wls->setDepthDiscontinuityRadius(radius);
    Mat out;
    wls->filter(dl, view, out, dr, roi);
    Rect r = wls->getROI();
    printf("ROI %d %d %d %d\n", r.x, r.y, r.width, r.height);
    save(dir + "/out.bin", out);
    save(dir + "/conf.bin", wls->getConfidenceMap());

    // 2. matcher factory: margins and radius derived from the matcher, matcher mutated
    Ptr<StereoBM> bm = StereoBM::create(80, 7);
    Ptr<DisparityWLSFilter> wls2 = createDisparityWLSFilter(bm);
    Ptr<StereoMatcher> right = createRightMatcher(bm);
    printf("FACTORY radius %d right_min %d disp12 %d\n", wls2->getDepthDiscontinuityRadius(), right->getMinDisparity(), bm->getDisp12MaxDiff());
    if (dw == W && dw > 2 * 90) {
        Mat out2;
        wls2->filter(dl, view, out2, dr);
        Rect q = wls2->getROI();
        printf("ROI2 %d %d %d %d\n", q.x, q.y, q.width, q.height);
    }

    // 3. fastGlobalSmootherFilter on a float image
    Mat fgs;
    fastGlobalSmootherFilter(view, src, fgs, lambda, sigma);
    save(dir + "/fgs.bin", fgs);

    // 3b. the device matcher in place of cv::StereoBM: the sample's matching step (samples/disparity_filtering.cpp:148-160)
    //     on <dir>/gl.bin, gr.bin (8UC1, dw x dh) when present -> bm_l.bin, bm_r.bin, bm_out.bin
    {
        FILE* probe = fopen((dir + "/gl.bin").c_str(), "rb");
        if (probe) {
            fclose(probe);
            Mat gl(dh, dw, CV_8UC1), gr(dh, dw, CV_8UC1), bl, br, fout;
            load(dir + "/gl.bin", gl); load(dir + "/gr.bin", gr);
            Ptr<StereoBMB200> lm = StereoBMB200::create(48, 7);
            Ptr<DisparityWLSFilter> w3 = createDisparityWLSFilter(lm);
            Ptr<StereoMatcher> rm = createRightMatcher(lm);
            lm->compute(gl, gr, bl);
            rm->compute(gr, gl, br);
            w3->setLambda(lambda); w3->setSigmaColor(sigma);
            w3->filter(bl, view, fout, br);
            Rect q = w3->getROI();
            printf("ROI3 %d %d %d %d\n", q.x, q.y, q.width, q.height);
            save(dir + "/bm_l.bin", bl); save(dir + "/bm_r.bin", br); save(dir + "/bm_out.bin", fout);
        }
    }

    // 4. error behaviour: exceptions like CV_Assert / CV_Error(StsBadSize)
    int caught = 0;
    try { Mat bad(dh, dw, CV_32FC1), o; wls->filter(bad, view, o, dr, roi); } catch (const cv::Exception&) { caught |= 1; }
    try { Mat small(4, 4, CV_32FC1), o; fastGlobalSmootherFilter(view, small, o, 10.0, 1.0); } catch (const cv::Exception& e) { if (e.code == Error::StsBadSize) caught |= 2; }
    printf("CAUGHT %d\n", caught);
    return 0;
}
