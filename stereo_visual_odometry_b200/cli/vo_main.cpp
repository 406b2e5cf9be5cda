// vo_main.cpp -- command-line trajectory run with the semantics of the reference's `vo N folder`
// (Alex7Li/stereo_visual_odometry src/main.cpp:315-407), on top of this repo's drop-in include/vo.h:
//
//     vo N folder [--gray] [--calib file.yaml] [--pose0 pitch26|identity] [--out file.csv] [--stream]
//
//   * frames folder/left/frameNNNNNN.png + folder/right/... (6 digits + .png for "run1", 4 digits + .jpg
//     otherwise, SeqID_* names for "cfe_cameras": main.cpp:20-36); a missing left image skips the frame
//   * images go to stereo_callback as cv::imread returns them, 3-channel BGR (main.cpp:37-46) --
//     `--gray` converts to 8UC1 first (the documented contract of the class)
//   * projection matrices: the literals of main.cpp:357-362, or fx/fy/cx/cy/bf of a calibration YAML in either key
//     style (`fx:` as stereo_vo.cpp:39-43 reads them, `Camera.fx:` as calibration/*.yaml spell them)
//   * folder/gt.csv (time,x,y,dx,dy) is passed through to the gtx,gty columns; "rand_feats" and "cfe_cameras"
//     have none (main.cpp:330-344)
//   * folder/result.csv: header x,y,z,gtx,gty, one row per processed frame, default stream formatting
//     (6 significant digits), pose integrated as frame_pose = frame_pose * transform from a 26 degree pitch
//     (main.cpp:368-373, 394-400); `--pose0 identity` reproduces the committed run1/result.csv
//   * `--stream` feeds the frames through submit_frame / collect_frame (several frames in flight); same rows.
//   * N = 0 ran the in-file assert tests upstream (main.cpp:293-312); here they are tests/cpp/test_vo_dropin.cpp.
#include <cmath>
#include <fstream>
#include <iomanip>
#include <iostream>
#include <sstream>
#include <string>
#include <vector>

#include "image_io.h"
#include "vo.h"

using namespace visual_odometry;

namespace {

std::string base_name(std::string p) {
    while (p.size() > 1 && p.back() == '/') p.pop_back();
    const size_t k = p.find_last_of('/');
    return k == std::string::npos ? p : p.substr(k + 1);
}

// main.cpp:20-36
void frame_paths(const std::string& folder, const std::string& name, int i, std::string& left, std::string& right) {
    std::ostringstream l, r;
    if (name == "cfe_cameras") {
        l << folder << "/SeqID_" << i << "-CamId_0-vo-rectified-stereo_app.png";
        r << folder << "/SeqID_" << i << "-CamId_1-vo-rectified-stereo_app.png";
    } else {
        const bool run1 = name == "run1";
        const int zeros = run1 ? 6 : 4;
        const char* ext = run1 ? ".png" : ".jpg";
        l << folder << "/left/frame" << std::setw(zeros) << std::setfill('0') << i << ext;
        r << folder << "/right/frame" << std::setw(zeros) << std::setfill('0') << i << ext;
    }
    left = l.str();
    right = r.str();
}

// fx / fy / cx / cy / bf from a calibration YAML, `key: value` lines, keys with or without the `Camera.` prefix
bool load_calibration(const std::string& path, float P_left[12], float P_right[12]) {
    std::ifstream f(path);
    if (!f.good()) return false;
    double fx = 0, fy = 0, cx = 0, cy = 0, bf = 0;
    int seen = 0;
    std::string line;
    while (std::getline(f, line)) {
        const size_t hash = line.find('#');
        if (hash != std::string::npos) line.erase(hash);
        const size_t colon = line.find(':');
        if (colon == std::string::npos) continue;
        std::string key = line.substr(0, colon);
        key.erase(0, key.find_first_not_of(" \t"));
        key.erase(key.find_last_not_of(" \t") + 1);
        if (key.rfind("Camera.", 0) == 0) key.erase(0, 7);
        char* end = nullptr;
        const std::string val = line.substr(colon + 1);
        const double v = std::strtod(val.c_str(), &end);
        if (end == val.c_str()) continue;
        if (key == "fx") { fx = v; seen |= 1; }
        else if (key == "fy") { fy = v; seen |= 2; }
        else if (key == "cx") { cx = v; seen |= 4; }
        else if (key == "cy") { cy = v; seen |= 8; }
        else if (key == "bf") { bf = v; seen |= 16; }
    }
    if (seen != 31) return false;
    const float l[12] = {(float)fx, 0.f, (float)cx, 0.f, 0.f, (float)fy, (float)cy, 0.f, 0.f, 0.f, 1.f, 0.f};
    for (int i = 0; i < 12; ++i) P_left[i] = P_right[i] = l[i];
    P_right[3] = (float)bf;   // stereo_vo.cpp:46-47
    return true;
}

struct GroundTruth {
    std::vector<std::pair<double, double>> xy;
    size_t next = 0;
    // gt.csv: header line, then time,x,y,dx,dy rows; one row is consumed per processed frame (main.cpp:381-392)
    bool load(const std::string& path) {
        std::ifstream f(path);
        if (!f.good()) return false;
        std::string line;
        std::getline(f, line);
        while (std::getline(f, line)) {
            std::istringstream ls(line);
            std::string t, x, y;
            if (std::getline(ls, t, ',') && std::getline(ls, x, ',') && std::getline(ls, y, ','))
                xy.emplace_back(std::atof(x.c_str()), std::atof(y.c_str()));
        }
        return true;
    }
    std::pair<double, double> take() { return next < xy.size() ? xy[next++] : std::make_pair(0.0, 0.0); }
};

}  // namespace

int main(int argc, char** argv) {
    int n_frames = 2;
    std::string folder = "cfe_cameras", calib, out_path, pose0 = "pitch26";
    bool gray = false, stream = false;
    std::vector<std::string> pos;
    for (int i = 1; i < argc; ++i) {
        const std::string a = argv[i];
        if (a == "--gray") gray = true;
        else if (a == "--stream") stream = true;
        else if (a == "--calib" && i + 1 < argc) calib = argv[++i];
        else if (a == "--pose0" && i + 1 < argc) pose0 = argv[++i];
        else if (a == "--out" && i + 1 < argc) out_path = argv[++i];
        else if (a == "--dump-image" && i + 2 < argc) {   // decoder check: image file -> binary PPM (no GPU needed)
            std::string why;
            const cv::Mat m = svo_cli::read_image(argv[i + 1], &why);
            if (m.empty()) { std::cerr << why << std::endl; return 2; }
            std::ofstream o(argv[i + 2], std::ios::binary);
            o << "P6\n" << m.cols << " " << m.rows << "\n255\n";
            for (int y = 0; y < m.rows; ++y)
                for (int x = 0; x < m.cols; ++x) {
                    const uint8_t* q = m.ptr<uint8_t>(y) + 3 * x;
                    const char rgb[3] = {(char)q[2], (char)q[1], (char)q[0]};
                    o.write(rgb, 3);
                }
            return 0;
        } else if (a == "--dump-gray" && i + 2 < argc) {   // --gray conversion check: image file -> binary PGM
            std::string why;
            const cv::Mat m = svo_cli::read_image(argv[i + 1], &why);
            if (m.empty()) { std::cerr << why << std::endl; return 2; }
            const cv::Mat g = svo_cli::bgr_to_gray(m);
            std::ofstream o(argv[i + 2], std::ios::binary);
            o << "P5\n" << g.cols << " " << g.rows << "\n255\n";
            for (int y = 0; y < g.rows; ++y) o.write((const char*)g.ptr<uint8_t>(y), g.cols);
            return 0;
        } else if (a == "-h" || a == "--help") {
            std::cout << "usage: vo N folder [--gray] [--calib file.yaml] [--pose0 pitch26|identity] [--out file.csv] [--stream]\n";
            return 0;
        } else pos.push_back(a);
    }
    if (pos.empty()) std::cout << "No N_FRAMES given, defaulting to 2" << std::endl;
    else n_frames = std::atoi(pos[0].c_str());
    if (pos.size() >= 2) folder = pos[1];
    if (n_frames == 0) {
        std::cout << "The reference's in-file tests live in tests/cpp/test_vo_dropin.cpp (python -m pytest tests -m gpu)." << std::endl;
        return 0;
    }
    const std::string name = base_name(folder);
    GroundTruth gt;
    const bool has_gt = name != "rand_feats" && name != "cfe_cameras" && gt.load(folder + "/gt.csv");

    float left_P[12] = {322.11376f, 0.f, 327.47336f, 0.f, 0.f, 322.11376f, 176.33722f, 0.f, 0.f, 0.f, 1.f, 0.f};   // main.cpp:357-362
    float right_P[12] = {322.11376f, 0.f, 327.47336f, -22.5428f, 0.f, 322.11376f, 176.33722f, 0.f, 0.f, 0.f, 1.f, 0.f};
    if (!calib.empty() && !load_calibration(calib, left_P, right_P)) {
        std::cerr << "cannot read fx/fy/cx/cy/bf from " << calib << std::endl;
        return 2;
    }
    cv::Mat_<float> Pl(3, 4), Pr(3, 4);
    for (int i = 0; i < 12; ++i) { Pl(i / 4, i % 4) = left_P[i]; Pr(i / 4, i % 4) = right_P[i]; }

    std::ofstream result(out_path.empty() ? folder + "/result.csv" : out_path);
    if (!result.good()) {
        std::cerr << "cannot write the result file" << std::endl;
        return 2;
    }
    result << "x,y,z,gtx,gty" << "\n";

    try {
        VisualOdometry vo;
        vo.initalize_projection_matricies(Pl, Pr);
        cv::Mat frame_pose = cv::Mat::eye(4, 4, CV_64F);
        if (pose0 != "identity") {   // the cameras look down by 26 degrees (main.cpp:368-373)
            const float theta = (26.0 / 360) * 2 * M_PI;
            frame_pose.at<double>(1, 1) = std::cos(theta); frame_pose.at<double>(1, 2) = std::sin(theta);
            frame_pose.at<double>(2, 1) = -std::sin(theta); frame_pose.at<double>(2, 2) = std::cos(theta);
        }
        struct Pending { int frame; double gtx, gty; cv::Mat l, r; };
        std::vector<Pending> pending;
        int n_ok = 0, n_done = 0;
        auto emit = [&](const Pending& p, const std::pair<bool, cv::Mat_<double>>& out) {
            frame_pose = frame_pose * out.second;   // main.cpp:394-396
            result << frame_pose.at<double>(0, 3) << "," << frame_pose.at<double>(1, 3) << "," << frame_pose.at<double>(2, 3)
                   << "," << p.gtx << "," << p.gty << "\n";
            std::cout << "Frame " << p.frame << ": " << (out.first ? "ok" : "no motion estimate") << "\n";
            n_ok += out.first ? 1 : 0;
            ++n_done;
        };
        for (int i = 0; i < n_frames; ++i) {
            std::string lp, rp, why;
            frame_paths(folder, name, i, lp, rp);
            cv::Mat l = svo_cli::read_image(lp, &why);
            if (l.empty()) {
                if (why.find("cannot open") == std::string::npos) std::cerr << why << std::endl;
                continue;   // main.cpp:376-378
            }
            cv::Mat r = svo_cli::read_image(rp, &why);
            if (r.empty()) {   // the reference asserts here (main.cpp:379)
                std::cerr << "right image missing for frame " << i << ": " << why << std::endl;
                return 3;
            }
            if (gray) { l = svo_cli::bgr_to_gray(l); r = svo_cli::bgr_to_gray(r); }
            Pending p{i, 0.0, 0.0, l, r};
            if (has_gt) { const auto g = gt.take(); p.gtx = g.first; p.gty = g.second; }
            if (!stream) {
                emit(p, vo.stereo_callback(l, r));
                continue;
            }
            vo.submit_frame(p.l, p.r);
            pending.push_back(p);
            if (pending.size() > 3) {
                emit(pending.front(), vo.collect_frame());
                pending.erase(pending.begin());
            }
        }
        while (!pending.empty()) {
            emit(pending.front(), vo.collect_frame());
            pending.erase(pending.begin());
        }
        std::cout << n_done << " frames, " << n_ok << " motion estimates" << std::endl;
    } catch (const std::exception& e) {
        std::cerr << "error: " << e.what() << std::endl;
        return 1;
    }
    return 0;
}
