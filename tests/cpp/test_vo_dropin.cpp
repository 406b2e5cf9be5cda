// C++ acceptance checks of the drop-in header include/vo.h (this repo) -- the known answers of the
// reference's own assert-tests (src/main.cpp:50-264 of Alex7Li/stereo_visual_odometry), re-expressed
// against the same class / function signatures.  Needs a CUDA device; run by tests/test_gpu_cpp.py.
#include <cmath>
#include <cstdio>
#include <cstdlib>

#include "vo.h"

using namespace visual_odometry;

#define REQUIRE(cond)                                                            \
    do {                                                                         \
        if (!(cond)) {                                                           \
            std::printf("FAILED %s:%d: %s\n", __FILE__, __LINE__, #cond);        \
            std::exit(1);                                                        \
        }                                                                        \
    } while (0)

static cv::Mat blank(int rows, int cols) { return cv::Mat::zeros(rows, cols, CV_8UC1); }

// filled triangle of value 120 with a one-pixel hole at its apex column (main.cpp:89-100)
static void stamp(cv::Mat& img, int x, int y, int r) {
    for (int i = -r; i <= r; ++i)
        for (int j = 0; j <= r; ++j)
            if (std::abs(i) <= r - std::abs(j)) img.at<unsigned char>(i + y, j + x) = 120;
    img.at<unsigned char>(y, x) = 0;
}

static void bucket_policy() {
    const int ages[5] = {6, 2, 3, 4, 5}, strengths[5] = {60, 20, 30, 40, 50};
    Bucket none(0), three(3);
    for (int i = 0; i < 5; ++i) {
        none.add_feature(cv::Point2f(1, 1), ages[i], strengths[i]);
        three.add_feature(cv::Point2f(1, 1), ages[i], strengths[i]);
    }
    REQUIRE(none.max_size == 0 && none.features.size() == 0);
    REQUIRE(three.features.size() == 3);
    REQUIRE(three.features.ages[0] == 6 && three.features.ages[1] == 4 && three.features.ages[2] == 5);
    REQUIRE(three.features.strengths[0] == 60 && three.features.strengths[1] == 40 && three.features.strengths[2] == 50);
    // the same policy through the GPU bucketing kernel (one bucket of capacity 3)
    FeatureSet fs;
    for (int i = 0; i < 5; ++i) {
        fs.points.push_back(cv::Point2f(1, 1));
        fs.ages.push_back(ages[i]);
        fs.strengths.push_back(strengths[i]);
    }
    cv::Mat img = blank(10, 10);
    fs.filterByBucketLocationInternal(img, 1, 1, 0, 3);
    REQUIRE(fs.size() == 3 && fs.ages[0] == 6 && fs.ages[1] == 4 && fs.ages[2] == 5);
}

static void featureset_counts() {
    FeatureSet fs;
    REQUIRE(fs.size() == 0);
    cv::Mat img = blank(300, 200);
    for (int i = 0; i <= 10; ++i) stamp(img, 20, (i + 1) * 20, 8);
    fs.appendFeaturesFromImage(img, 1);
    for (int a : fs.ages) REQUIRE(a == 0);
    for (int s : fs.strengths) REQUIRE(s <= 128);
    REQUIRE(fs.size() == 11);
    fs.filterByBucketLocationInternal(img, 1, 1, 0, 7);
    REQUIRE(fs.size() == 7);
}

static void featureset_edges() {
    FeatureSet fs;
    cv::Mat img = blank(300, 300);
    for (int i = 0; i < 15; ++i) { fs.points.push_back({299.f, 299.f}); fs.ages.push_back(0); fs.strengths.push_back(40); }
    for (int i = 0; i < 10; ++i) { fs.points.push_back({299.f, 0.f}); fs.ages.push_back(0); fs.strengths.push_back(40); }
    for (int i = 0; i < 5; ++i) { fs.points.push_back({0.f, 150.f}); fs.ages.push_back(0); fs.strengths.push_back(40); }
    fs.filterByBucketLocationInternal(img, 2, 2, 0, 11);
    FeatureSet copy = fs;
    REQUIRE(fs.size() == 26);
    fs.filterByBucketLocationInternal(img, 2, 1, 0, 11);
    REQUIRE(fs.size() == 21);
    copy.filterByBucketLocationInternal(img, 1, 2, 0, 11);
    REQUIRE(copy.size() == 16);
}

static void unmoved_points() {
    std::vector<cv::Point2f> a, b;
    for (int i = 0; i < 35; ++i) {
        a.push_back({float(i), float(i)});
        b.push_back({float(i) + !(i % 5), float(i) + !(i % 7)});
    }
    std::vector<bool> ok = findClosePoints(a, b, .5f);
    for (int i = 0; i < 35; ++i) REQUIRE(ok[i] == ((i % 5) && (i % 7)));
    std::vector<cv::Point2f> kept = a;
    deletePointsWithFailureStatus(kept, ok);
    size_t expect = 0;
    for (bool k : ok) expect += k;
    REQUIRE(kept.size() == expect);
}

static void circular_matching() {
    cv::Mat iL0 = blank(600, 600), iR0 = blank(600, 600), iL1 = blank(600, 600), iR1 = blank(600, 600);
    for (int i = 0; i <= 10; ++i)
        for (int j = 0; j <= 10; ++j) {
            stamp(iL0, (j + 1) * 40, (i + 1) * 40, 8);
            stamp(iL1, (j + 1) * 40 + 1, (i + 1) * 40, 8);
            stamp(iR0, (j + 1) * 40, (i + 1) * 40 + 1, 8);
            stamp(iR1, (j + 1) * 40 + 1, (i + 1) * 40 + 1, 8);
        }
    VisualOdometry vo;
    std::vector<cv::Point2f> pl0, pr0, pl1, pr1;
    vo.stereo_callback(iL0, iR0);
    FeatureSet fs;
    fs.appendFeaturesFromImage(iL0, FAST_THRESHOLD);
    vo.circularMatching(iL1, iR1, pl0, pr0, pl1, pr1, fs);   // empty input must not crash
    pl0.push_back(fs.points[0]);
    FeatureSet one;
    vo.circularMatching(iL1, iR1, pl0, pr0, pl1, pr1, one);  // single element must not crash
    pl0 = fs.points;
    vo.circularMatching(iL1, iR1, pl0, pr0, pl1, pr1, fs);
    const size_t n = fs.points.size();
    REQUIRE(pl0.size() == n && pl1.size() == n && pr0.size() == n && pr1.size() == n);
    REQUIRE(n == 121);
}

// VisualOdometry::matchingFeatures (include/vo.h:354-362, src/vo.cpp:315-366 of the reference) through the header:
// FAST on the t0 left image (+ the threshold / 4 retry below 100 features), bucketing, circular matching, bounds,
// compaction -- all five containers index-aligned.  The scene is the reference's own test_circularMatching grid
// (main.cpp:174-209): 121 triangles, t1 shifted by +1 px in x, the right images by +1 px in y.
static void matching_features() {
    cv::Mat iL0 = blank(600, 600), iR0 = blank(600, 600), iL1 = blank(600, 600), iR1 = blank(600, 600);
    for (int i = 0; i <= 10; ++i)
        for (int j = 0; j <= 10; ++j) {
            stamp(iL0, (j + 1) * 40, (i + 1) * 40, 8);
            stamp(iL1, (j + 1) * 40 + 1, (i + 1) * 40, 8);
            stamp(iR0, (j + 1) * 40, (i + 1) * 40 + 1, 8);
            stamp(iR1, (j + 1) * 40 + 1, (i + 1) * 40 + 1, 8);
        }
    VisualOdometry vo;
    FeatureSet fs;   // empty: everything comes from FAST on iL0
    std::vector<cv::Point2f> pl0, pr0, pl1, pr1;
    vo.matchingFeatures(iL0, iR0, iL1, iR1, fs, pl0, pr0, pl1, pr1);
    const size_t n = fs.points.size();
    REQUIRE(n == 121);
    REQUIRE(pl0.size() == n && pr0.size() == n && pl1.size() == n && pr1.size() == n);
    REQUIRE(fs.ages.size() == n && fs.strengths.size() == n);
    FeatureSet expect;   // what FAST + bucketing alone yield, in the same (bucket row-major) order
    expect.appendFeaturesFromImage(iL0, FAST_THRESHOLD);
    REQUIRE(expect.size() == (int)n);
    for (size_t k = 0; k < n; ++k) {
        REQUIRE(fs.points[k].x == expect.points[k].x && fs.points[k].y == expect.points[k].y);   // t0 coordinates (vo.cpp:360)
        REQUIRE(fs.ages[k] == 0 && fs.strengths[k] == expect.strengths[k]);                       // ages++ is stereo_callback's job
        REQUIRE(pl0[k].x == fs.points[k].x && pl0[k].y == fs.points[k].y);
        REQUIRE(std::fabs(pl1[k].x - pl0[k].x - 1) < 0.05f && std::fabs(pl1[k].y - pl0[k].y) < 0.05f);
        REQUIRE(std::fabs(pr0[k].x - pl0[k].x) < 0.05f && std::fabs(pr0[k].y - pl0[k].y - 1) < 0.05f);
        REQUIRE(std::fabs(pr1[k].x - pl0[k].x - 1) < 0.05f && std::fabs(pr1[k].y - pl0[k].y - 1) < 0.05f);
    }
    // second call with the tracked set (t1 positions, one frame older) as stereo_callback would hand it over:
    // the old features win their buckets (age 1 beats age 0 at equal strength), nothing new is appended beside them
    FeatureSet tracked;
    for (size_t k = 0; k < n; ++k)   // exact pixel positions, so that every tracked feature shares the bucket of its corner
        tracked.points.push_back(cv::Point2f(std::round(pl1[k].x), std::round(pl1[k].y)));
    tracked.ages.assign(n, 1);
    tracked.strengths = fs.strengths;
    vo.matchingFeatures(iL1, iR1, iL0, iR0, tracked, pl0, pr0, pl1, pr1);   // back to the first pair: shift -1 px
    REQUIRE(tracked.points.size() == n && pl0.size() == n);
    for (size_t k = 0; k < n; ++k) {
        REQUIRE(tracked.ages[k] == 1);
        REQUIRE(std::fabs(pl1[k].x - pl0[k].x + 1) < 0.05f && std::fabs(pl1[k].y - pl0[k].y) < 0.05f);
    }
    // no texture at all: the retry fires, nothing is found, every container comes back empty (vo.cpp:179-181)
    cv::Mat z = blank(200, 320);
    FeatureSet none;
    vo.matchingFeatures(z, z, z, z, none, pl0, pr0, pl1, pr1);
    REQUIRE(none.size() == 0 && pl0.empty() && pr0.empty() && pl1.empty() && pr1.empty());
}

static void camera_to_world() {
    cv::Mat_<float> K = cv::Mat_<float>::eye(3, 3);
    const int n = 27;
    cv::Mat world(n, 1, CV_32FC3);
    std::vector<cv::Point2f> cam(n);
    int k = 0;
    for (int i = -1; i <= 1; ++i)
        for (int j = -1; j <= 1; ++j)
            for (int z = 5; z <= 7; ++z) {
                float* w = world.ptr<float>(k);
                w[0] = -j; w[1] = i; w[2] = z;                       // 90 degrees about z, then +1 along z
                cam[k] = cv::Point2f(float(i) / (z + 1), float(j) / (z + 1));
                ++k;
            }
    cv::Mat rotation = cv::Mat::eye(3, 3, CV_64F), translation = cv::Mat::zeros(3, 1, CV_64F);
    cv::Mat_<float> wtyped;
    static_cast<cv::Mat&>(wtyped) = world;
    auto res = cameraToWorld(K, cam, wtyped, rotation, translation);
    REQUIRE(res.second);
    REQUIRE(res.first.size().height == n);
    REQUIRE(std::fabs(translation.at<double>(0)) < 1e-6 && std::fabs(translation.at<double>(1)) < 1e-6);
    REQUIRE(std::fabs(translation.at<double>(2) - 1) < 1e-6);
    const double expect[3][3] = {{0, 1, 0}, {-1, 0, 0}, {0, 0, 1}};
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) REQUIRE(std::fabs(rotation.at<double>(i, j) - expect[i][j]) < 1e-8);
    cv::Mat T = getInverseTransform(rotation, translation);
    REQUIRE(std::fabs(T.at<double>(2, 3) + 1) < 1e-6 && std::fabs(T.at<double>(3, 3) - 1) < 1e-12);
}

static void first_frame_contract() {
    cv::Mat a = blank(200, 320), b = blank(200, 320);
    StereoVO vo;   // upstream alias
    auto r0 = vo.run(a, b);
    REQUIRE(r0.first == false);
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) REQUIRE(r0.second(i, j) == (i == j ? 1.0 : 0.0));
    auto r1 = vo.run(a, b);   // nothing to track: soft failure, last_transform (identity) returned
    REQUIRE(r1.first == false && r1.second(0, 0) == 1.0 && r1.second(0, 3) == 0.0);
}

// streaming calls return what the synchronous ones return, in order
static void streaming_matches_sync() {
    cv::Mat imgs[5];
    for (int k = 0; k < 5; ++k) {
        imgs[k] = blank(200, 320);
        for (int y = 40; y < 180; y += 24)
            for (int x = 30 + 2 * k; x < 300; x += 28) stamp(imgs[k], x, y, 6);
    }
    VisualOdometry a, b;
    cv::Mat_<float> P(3, 4);
    const float pv[12] = {300.f, 0.f, 160.f, 0.f, 0.f, 300.f, 100.f, 0.f, 0.f, 0.f, 1.f, 0.f};
    for (int i = 0; i < 12; ++i) P(i / 4, i % 4) = pv[i];
    cv::Mat_<float> Pr = P.clone();
    Pr(0, 3) = -30.f;
    a.initalize_projection_matricies(P, Pr);
    b.initalize_projection_matricies(P, Pr);
    std::pair<bool, cv::Mat_<double>> ref[5];
    for (int k = 0; k < 5; ++k) ref[k] = a.stereo_callback(imgs[k], imgs[k]);
    for (int k = 0; k < 5; ++k) b.submit_frame(imgs[k], imgs[k]);
    for (int k = 0; k < 5; ++k) {
        auto r = b.collect_frame();
        REQUIRE(r.first == ref[k].first);
        for (int i = 0; i < 4; ++i)
            for (int j = 0; j < 4; ++j) REQUIRE(r.second(i, j) == ref[k].second(i, j));
    }
}

int main() {
    std::printf("TEST BUCKET POLICY\n");        bucket_policy();
    std::printf("TEST FEATURE SET\n");          featureset_counts();
    std::printf("TEST FEATURE SET EDGE\n");     featureset_edges();
    std::printf("TEST FIND UNMOVED POINTS\n");  unmoved_points();
    std::printf("TEST CIRCULAR MATCHING\n");    circular_matching();
    std::printf("TEST MATCHING FEATURES\n");    matching_features();
    std::printf("TEST CAMERA TO WORLD\n");      camera_to_world();
    std::printf("TEST FIRST FRAME\n");          first_frame_contract();
    std::printf("TEST STREAMING\n");            streaming_matches_sync();
    std::printf("ALL TESTS PASS\n");
    return 0;
}
