// Checks of include/vo_stream.h, the streaming wrapper that stands in for the reference's ROS node
// (src/stereo_vo.cpp) and debug overlays (src/vo.cpp:28-39).  `test_vo_stream host` runs the host-only parts
// (pairing, MONO8 conversion, calibration file, overlays); without arguments the node is also driven on the GPU.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "vo_stream.h"

using namespace visual_odometry;

#define REQUIRE(cond)                                                            \
    do {                                                                         \
        if (!(cond)) {                                                           \
            std::printf("FAILED %s:%d: %s\n", __FILE__, __LINE__, #cond);        \
            std::exit(1);                                                        \
        }                                                                        \
    } while (0)

static cv::Mat blank(int rows, int cols) { return cv::Mat::zeros(rows, cols, CV_8UC1); }
static void stamp(cv::Mat& img, int x, int y, int r) {   // main.cpp:89-100
    for (int i = -r; i <= r; ++i)
        for (int j = 0; j <= r; ++j)
            if (std::abs(i) <= r - std::abs(j)) img.at<unsigned char>(i + y, j + x) = 120;
    img.at<unsigned char>(y, x) = 0;
}

static void pairing() {
    std::vector<std::pair<double, double>> got;
    ApproximateTimePairer p([&](const cv::Mat& l, const cv::Mat& r, double tl, double tr) {
        REQUIRE(l.at<unsigned char>(0, 0) == (unsigned char)std::lround(tl * 10) && r.at<unsigned char>(0, 0) == (unsigned char)std::lround(tr * 10));
        got.emplace_back(tl, tr);
    });
    auto img = [](double t) { cv::Mat m = blank(2, 2); m.at<unsigned char>(0, 0) = (unsigned char)std::lround(t * 10); return m; };
    // right lags left by 0.02 s, one right frame is missing (t = 0.3), one left frame arrives twice as fast (0.45)
    const double L[] = {0.0, 0.1, 0.2, 0.3, 0.4, 0.45, 0.5, 0.6};
    const double R[] = {0.02, 0.12, 0.22, 0.42, 0.52, 0.62};
    size_t il = 0, ir = 0;
    while (il < 8 || ir < 6) {   // deliver in time order
        if (ir >= 6 || (il < 8 && L[il] <= R[ir])) { p.addLeft(L[il], img(L[il])); ++il; }
        else { p.addRight(R[ir], img(R[ir])); ++ir; }
    }
    p.flush();
    const double want[6][2] = {{0.0, 0.02}, {0.1, 0.12}, {0.2, 0.22}, {0.4, 0.42}, {0.5, 0.52}, {0.6, 0.62}};
    REQUIRE(got.size() == 6 && p.pairs() == 6);
    for (int i = 0; i < 6; ++i) REQUIRE(got[i].first == want[i][0] && got[i].second == want[i][1]);
    REQUIRE(p.dropped() == 2);   // left 0.3 (no partner) and 0.45 (0.4 is closer to 0.42)
    // exactly synchronised topics pair at once, without waiting for the next message
    got.clear();
    ApproximateTimePairer q([&](const cv::Mat&, const cv::Mat&, double tl, double tr) { got.emplace_back(tl, tr); });
    q.addLeft(1.0, img(0));
    q.addRight(1.0, img(0));
    REQUIRE(got.size() == 1);
}

static void mono8_and_overlays() {
    cv::Mat bgr(2, 2, CV_8UC3);
    const unsigned char px[4][3] = {{255, 0, 0}, {0, 255, 0}, {0, 0, 255}, {12, 200, 77}};
    for (int i = 0; i < 4; ++i) std::memcpy(bgr.ptr<unsigned char>(i / 2) + 3 * (i % 2), px[i], 3);
    const cv::Mat g = to_mono8(bgr);
    const int want[4] = {29, 150, 76, 142};   // cv2.cvtColor(BGR2GRAY) of the four pixels
    for (int i = 0; i < 4; ++i) REQUIRE(g.at<unsigned char>(i / 2, i % 2) == want[i]);
    cv::Mat m = blank(3, 3);
    m.at<unsigned char>(1, 1) = 9;
    const cv::Mat c = to_mono8(m);
    REQUIRE(c.at<unsigned char>(1, 1) == 9 && c.data != m.data);   // a copy, like toCvCopy
    cv::Mat img = blank(12, 12);
    displayPoints(img, {cv::Point2f(5.7f, 6.2f), cv::Point2f(0.f, 0.f)});
    int n = 0;
    for (int y = 0; y < 12; ++y)
        for (int x = 0; x < 12; ++x) n += img.at<unsigned char>(y, x) == 255;
    REQUIRE(n == 8 + 3);   // the 8-pixel ring of cv::circle(radius 2) around (5, 6); 3 of the corner circle's pixels are inside
    REQUIRE(img.at<unsigned char>(6, 7) == 255 && img.at<unsigned char>(4, 5) == 255 && img.at<unsigned char>(7, 4) == 255 && img.at<unsigned char>(7, 3) == 0 &&
            img.at<unsigned char>(6, 5) == 0);
    const cv::Mat t = displayTracking(img, blank(12, 12), {cv::Point2f(2, 2)}, {cv::Point2f(8, 5)});
    REQUIRE(t.at<unsigned char>(2, 2) == 255 && t.at<unsigned char>(5, 8 + 2) == 255 && t.at<unsigned char>(5, 8) == 255);
}

static void calibration(const char* dir) {
    const std::string a = std::string(dir) + "/a.yaml", b = std::string(dir) + "/b.yaml";
    { std::ofstream f(a); f << "%YAML:1.0\nfx: 718.856\nfy: 718.856\ncx: 607.1928\ncy: 185.2157\nbf: -386.1448\n"; }
    { std::ofstream f(b); f << "Camera.fx: 718.856\nCamera.fy: 718.856   # focal\nCamera.cx: 607.1928\nCamera.cy: 185.2157\nCamera.bf: -386.1448\nCamera.width: 1241\n"; }
    cv::Mat_<float> Pl, Pr, Pl2, Pr2;
    REQUIRE(load_calibration(a, Pl, Pr) && load_calibration(b, Pl2, Pr2));
    for (int i = 0; i < 12; ++i) REQUIRE(Pl(i / 4, i % 4) == Pl2(i / 4, i % 4) && Pr(i / 4, i % 4) == Pr2(i / 4, i % 4));
    REQUIRE(Pl(0, 0) == 718.856f && Pl(1, 2) == 185.2157f && Pl(0, 3) == 0.f && Pr(0, 3) == -386.1448f && Pl(2, 2) == 1.f);
    REQUIRE(!load_calibration(std::string(dir) + "/missing.yaml", Pl, Pr));
}

// the node fed through its two callbacks (right topic lagging, one left frame without partner) returns what
// direct stereo_callback calls on the paired frames return, and integrates the pose like main.cpp:396
static void node_on_gpu() {
    cv::Mat imgs[6];
    for (int k = 0; k < 6; ++k) {
        imgs[k] = blank(200, 320);
        for (int y = 40; y < 180; y += 24)
            for (int x = 30 + 2 * k; x < 300; x += 28) stamp(imgs[k], x, y, 6);
    }
    cv::Mat_<float> P(3, 4);
    const float pv[12] = {300.f, 0.f, 160.f, 0.f, 0.f, 300.f, 100.f, 0.f, 0.f, 0.f, 1.f, 0.f};
    for (int i = 0; i < 12; ++i) P(i / 4, i % 4) = pv[i];
    cv::Mat_<float> Pr = P.clone();
    Pr(0, 3) = -30.f;
    VisualOdometry direct;
    direct.initalize_projection_matricies(P, Pr);
    std::vector<std::pair<bool, cv::Mat_<double>>> want;
    const int used[5] = {0, 1, 2, 4, 5};   // frame 3's right image never arrives
    for (int k : used) want.push_back(direct.stereo_callback(imgs[k], imgs[k]));
    std::vector<std::pair<bool, cv::Mat_<double>>> got;
    StereoVONode node(P, Pr, [&](bool ok, const cv::Mat_<double>& T, const cv::Mat_<double>&) { got.emplace_back(ok, T.clone()); });
    for (int k = 0; k < 6; ++k) {
        node.leftImage(0.1 * k, imgs[k]);
        if (k != 3) node.rightImage(0.1 * k + 0.01, imgs[k]);
    }
    node.flush();
    REQUIRE(got.size() == 5 && node.framesProcessed() == 5);
    cv::Mat_<double> pose = cv::Mat_<double>::eye(4, 4);
    for (size_t i = 0; i < 5; ++i) {
        REQUIRE(got[i].first == want[i].first);
        for (int r = 0; r < 4; ++r)
            for (int c = 0; c < 4; ++c) REQUIRE(got[i].second(r, c) == want[i].second(r, c));
        cv::Mat_<double> next(4, 4);
        for (int r = 0; r < 4; ++r)
            for (int c = 0; c < 4; ++c) {
                double s = 0;
                for (int q = 0; q < 4; ++q) s += pose(r, q) * want[i].second(q, c);
                next(r, c) = s;
            }
        pose = next;
    }
    for (int r = 0; r < 4; ++r)
        for (int c = 0; c < 4; ++c) REQUIRE(node.pose()(r, c) == pose(r, c));
}

int main(int argc, char** argv) {
    const bool host_only = argc > 1 && std::strcmp(argv[1], "host") == 0;
    std::printf("TEST PAIRING\n");            pairing();
    std::printf("TEST MONO8 / OVERLAYS\n");   mono8_and_overlays();
    std::printf("TEST CALIBRATION\n");        calibration(argc > 2 ? argv[2] : "/tmp");
    if (!host_only) { std::printf("TEST NODE\n"); node_on_gpu(); }
    std::printf("ALL TESTS PASS\n");
    return 0;
}
