// scripts/cpp/bench_operator.cpp — latency of the C++ drop-in call as the reference issues it: one frame per
// ORBextractor::operator() (/root/reference/src/Frame.cc:248-254), host cv::Mat in, std::vector<cv::KeyPoint> + cv::Mat
// out, with and without the mvImagePyramid refresh (needed only by Frame::ComputeStereoMatches' host path).
//   g++ -O2 -std=c++14 -Iinclude -Ioracle/shim scripts/cpp/bench_operator.cpp -Lorbslam2_b200 -lorbx -Wl,-rpath,$PWD/orbslam2_b200
//   ./bench_operator <w> <h> <nfeatures> <raw 8-bit image file> [iterations]
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "dropin/ORBextractor.h"

static double now_ms() {
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

int main(int argc, char** argv) {
    if (argc < 5) return 1;
    const int w = atoi(argv[1]), h = atoi(argv[2]), nf = atoi(argv[3]);
    const int iters = argc > 5 ? atoi(argv[5]) : 300;
    std::vector<unsigned char> img((size_t)w * h);
    FILE* f = fopen(argv[4], "rb");
    if (!f || fread(img.data(), 1, img.size(), f) != img.size()) return 2;
    fclose(f);
    cv::Mat im(h, w, CV_8UC1, img.data(), (size_t)w);
    double med[2] = {0, 0}, p95[2] = {0, 0};
    size_t nk = 0;
    for (int keep = 0; keep < 2; keep++) {
        ORB_SLAM2::ORBextractor ex(nf, 1.2f, 8, 20, 7);
        ex.keepHostPyramid = keep != 0;
        std::vector<cv::KeyPoint> kps;
        cv::Mat desc;
        for (int i = 0; i < 20; i++) ex(im, cv::Mat(), kps, desc);
        std::vector<double> t(iters);
        for (int i = 0; i < iters; i++) {
            const double t0 = now_ms();
            ex(im, cv::Mat(), kps, desc);
            t[i] = now_ms() - t0;
        }
        std::sort(t.begin(), t.end());
        med[keep] = t[iters / 2];
        p95[keep] = t[(int)(iters * 0.95)];
        nk = kps.size();
    }
    printf("{\"call\": \"ORB_SLAM2::ORBextractor::operator() (C++ drop-in, pageable host cv::Mat)\", \"shape\": \"%dx%d\", "
           "\"nfeatures\": %d, \"keypoints\": %zu, \"iterations\": %d, \"ms_median_no_host_pyramid\": %.4f, "
           "\"ms_p95_no_host_pyramid\": %.4f, \"ms_median_with_host_pyramid\": %.4f, \"ms_p95_with_host_pyramid\": %.4f}\n",
           w, h, nf, nk, iters, med[0], p95[0], med[1], p95[1]);
    return 0;
}
