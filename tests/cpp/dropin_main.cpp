// Exercises the C++ drop-in class exactly the way Frame::ExtractORB does (src/Frame.cc:410-417):
//   monoLeft = (*mpORBextractorLeft)(im, cv::Mat(), mvKeys, mDescriptors, vLapping);
// Compiled against orb-slam3-mac_b200/host/ORBextractor.h + liborbx_b200.so with oracle/cvshim
// standing in for OpenCV's headers (tests/test_cpp_dropin.py).
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "ORBextractor.h"

int main(int argc, char** argv) {
    if (argc != 9) return 2;
    const int w = atoi(argv[1]), h = atoi(argv[2]), nf = atoi(argv[3]), lap0 = atoi(argv[4]), lap1 = atoi(argv[5]);
    const int level = atoi(argv[6]);
    std::vector<unsigned char> buf((size_t)w * h);
    FILE* f = fopen(argv[7], "rb");
    if (!f || fread(buf.data(), 1, buf.size(), f) != buf.size()) return 3;
    fclose(f);
    cv::Mat im(h, w, CV_8UC1, buf.data(), (size_t)w);

    ORB_SLAM3::ORBextractor* mpORBextractorLeft = new ORB_SLAM3::ORBextractor(nf, 1.2f, 8, 20, 7);
    std::vector<cv::KeyPoint> mvKeys;
    cv::Mat mDescriptors;
    std::vector<int> vLapping = {lap0, lap1};
    int monoLeft = -7;
    for (int rep = 0; rep < 2; ++rep)  // called once per frame in the reference; twice to check reuse
        monoLeft = (*mpORBextractorLeft)(im, cv::Mat(), mvKeys, mDescriptors, vLapping);
    int empty = (*mpORBextractorLeft)(cv::Mat(), cv::Mat(), mvKeys, mDescriptors, vLapping);
    monoLeft = (*mpORBextractorLeft)(im, cv::Mat(), mvKeys, mDescriptors, vLapping);

    FILE* o = fopen(argv[8], "wb");
    int n = (int)mvKeys.size();
    int hdr[6] = {monoLeft, n, empty, mpORBextractorLeft->GetLevels(), 0, 0};
    const cv::Mat& L = mpORBextractorLeft->mvImagePyramid[level];
    hdr[4] = L.cols;
    hdr[5] = L.rows;
    fwrite(hdr, sizeof(int), 6, o);
    std::vector<float> sf = mpORBextractorLeft->GetScaleFactors(), s2 = mpORBextractorLeft->GetInverseScaleSigmaSquares();
    fwrite(sf.data(), sizeof(float), sf.size(), o);
    fwrite(s2.data(), sizeof(float), s2.size(), o);
    if (n) fwrite(mvKeys.data(), sizeof(cv::KeyPoint), n, o);
    for (int i = 0; i < n; ++i) fwrite(mDescriptors.ptr(i), 1, 32, o);
    // the pyramid level INCLUDING its 19-px border, addressed relative to the ROI like Frame.cc does
    for (int y = -19; y < L.rows + 19; ++y) fwrite(L.data + (ptrdiff_t)y * (ptrdiff_t)L.step - 19, 1, L.cols + 38, o);
    fclose(o);
    delete mpORBextractorLeft;
    return 0;
}
