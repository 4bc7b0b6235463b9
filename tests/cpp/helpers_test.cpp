// Exercises the quality / IO helpers of the C++ facade (readGT, computeMSE, computeBadPixelPercent, getDisparityVis;
// reference: disparity_filter.hpp:163-204).  Pure host code: runs without a GPU.
// usage: helpers_test <dir> <w> <h> <rx> <ry> <rw> <rh> <thresh> <scale>
// reads <dir>/gt.bin src.bin (CV_16SC1) and <dir>/g.pgm c.ppm, writes vis.bin gt_pgm.bin gt_ppm.bin, prints the metrics.
#include <cstdio>
#include <cstdlib>
#include <string>

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
    if (argc < 10) return 2;
    const std::string dir = argv[1];
    const int w = atoi(argv[2]), h = atoi(argv[3]);
    const Rect roi(atoi(argv[4]), atoi(argv[5]), atoi(argv[6]), atoi(argv[7]));
    const int thresh = atoi(argv[8]);
    const double scale = atof(argv[9]);
    Mat gt(h, w, CV_16SC1), src(h, w, CV_16SC1);
    load(dir + "/gt.bin", gt); load(dir + "/src.bin", src);
    printf("MSE %.17g\n", computeMSE(gt, src, roi));
    printf("BAD %.17g\n", computeBadPixelPercent(gt, src, roi, thresh));
    printf("BAD_DEFAULT %.17g\n", computeBadPixelPercent(gt, src, roi));
    Mat vis;
    getDisparityVis(src, vis, scale);
    save(dir + "/vis.bin", vis);
    Mat a, b, c;
    printf("READ %d %d %d\n", readGT(dir + "/g.pgm", a), readGT(dir + "/c.ppm", b), readGT(dir + "/missing.pgm", c));
    printf("SIZE %d %d %d %d\n", a.cols, a.rows, b.cols, b.rows);
    save(dir + "/gt_pgm.bin", a); save(dir + "/gt_ppm.bin", b);
    int caught = 0;
    try { Mat f32(h, w, CV_32FC1); computeMSE(f32, src, roi); } catch (const cv::Exception&) { caught |= 1; }
    try { computeBadPixelPercent(gt, src, Rect(0, 0, w + 1, h)); } catch (const cv::Exception&) { caught |= 2; }
    try { Mat f32(h, w, CV_32FC1), o; getDisparityVis(f32, o); } catch (const cv::Exception&) { caught |= 4; }
    printf("CAUGHT %d\n", caught);
    return 0;
}
