// tests/host_cores.cpp -- TEST HARNESS, not product code and not a fallback.
//
// Compiles the arithmetic cores of stereo_visual_odometry_b200/csrc/svo_core.cuh for the
// host (g++, LANES = 1) so that the exact integer / fixed-point / FP64 arithmetic the CUDA
// kernels execute can be checked against the cv2 oracle on a box without a GPU.  The
// shipped library (libsvo_b200.so) never contains or calls any of this.
#include <stdint.h>
#include <string.h>

#include <vector>

#include "../stereo_visual_odometry_b200/csrc/svo_core.cuh"
#include "../stereo_visual_odometry_b200/csrc/svo_epnp.cuh"
#include "../stereo_visual_odometry_b200/csrc/svo_fast_packed.cuh"

using namespace svo;

struct HostImg {
    const uint8_t* p;
    int pitch;
    int operator()(int x, int y) const { return p[(size_t)y * pitch + x]; }
};

extern "C" {

// dense response map (0 = not a keypoint), same semantics as fast_kernel
int hc_fast_map(const uint8_t* img, int w, int h, int stride, int thr, uint8_t* resp_map) {
    static const int RDX[16] = SVO_FAST_RING_DX;
    static const int RDY[16] = SVO_FAST_RING_DY;
    std::vector<uint8_t> score((size_t)w * h, 0);
    for (int y = 3; y < h - 3; ++y)
        for (int x = 3; x < w - 3; ++x) {
            const int v = img[(size_t)y * stride + x];
            int r[16];
            for (int k = 0; k < 16; ++k) r[k] = img[(size_t)(y + RDY[k]) * stride + x + RDX[k]];
            const int s = fast9_arc_score(v, r);
            score[(size_t)y * w + x] = (uint8_t)(s > thr ? s - 1 : 0);
        }
    int n = 0;
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) {
            const int v = score[(size_t)y * w + x];
            bool keep = v > 0;
            for (int dy = -1; dy <= 1 && keep; ++dy)
                for (int dx = -1; dx <= 1; ++dx) {
                    if (!dx && !dy) continue;
                    const int xx = x + dx, yy = y + dy;
                    const int nv = (xx >= 0 && xx < w && yy >= 0 && yy < h) ? score[(size_t)yy * w + xx] : 0;
                    if (!(v > nv)) { keep = false; break; }
                }
            resp_map[(size_t)y * w + x] = keep ? (uint8_t)v : 0;
            n += keep;
        }
    return n;
}

// the packed pixel-pair FAST arithmetic of fast_kernel (svo_fast_packed.cuh) against the scalar statement, for every
// quad whose 12-byte row windows lie inside the image.  Returns the number of disagreements: a score that differs, a
// compass bit that differs from the scalar compass bound, or a corner (score > thr) the compass test would drop.
int hc_fast_packed_check(const uint8_t* img, int w, int h, int stride, int thr) {
    static const int RDX[16] = SVO_FAST_RING_DX;
    static const int RDY[16] = SVO_FAST_RING_DY;
    const unsigned kt = (unsigned)(0x8000 - thr - 1) * 0x00010001u;
    int bad = 0;
    for (int y = 3; y < h - 3; ++y)
        for (int x0 = 4; x0 + 8 <= w; ++x0) {   // quad = pixels x0 .. x0 + 3, windows = bytes x0 - 4 .. x0 + 7
            unsigned rows[7][3];
            for (int r = 0; r < 7; ++r)
                for (int k = 0; k < 3; ++k) {
                    const uint8_t* p = img + (size_t)(y - 3 + r) * stride + x0 - 4 + 4 * k;
                    rows[r][k] = p[0] | (p[1] << 8) | (p[2] << 16) | ((unsigned)p[3] << 24);
                }
            const unsigned s[2] = {fast_pair_score_ring<4>(rows), fast_pair_score_ring<6>(rows)};
            const unsigned cp[2] = {fast_pair_compass<4>(rows[0], rows[3], rows[6], kt),
                                    fast_pair_compass<6>(rows[0], rows[3], rows[6], kt)};
            const unsigned packed = fast_quad_pack(s[0], s[1], (unsigned)(0x8000 - (thr + 256) - 1) * 0x00010001u);
            for (int e = 0; e < 4; ++e) {
                const int x = x0 + e, v = img[(size_t)y * stride + x];
                int r[16];
                for (int k = 0; k < 16; ++k) r[k] = img[(size_t)(y + RDY[k]) * stride + x + RDX[k]];
                const int ref = fast9_arc_score(v, r);
                const int got = (int)((s[e >> 1] >> (16 * (e & 1))) & 0xFFFFu) - 256;
                const int d0 = v - r[0], d4 = v - r[4], d8 = v - r[8], d12 = v - r[12];
                const int a = imax(imax(imin(d0, d4), imin(d4, d8)), imax(imin(d8, d12), imin(d12, d0)));
                const int b = imin(imin(imax(d0, d4), imax(d4, d8)), imin(imax(d8, d12), imax(d12, d0)));
                const bool ref_pass = imax(a, -b) > thr;
                const bool got_pass = ((cp[e >> 1] >> (16 * (e & 1) + 15)) & 1u) != 0;
                const int want_byte = ref > thr ? ref - 1 : 0;
                if (ref != got || ref_pass != got_pass || (ref > thr && !got_pass) ||
                    (int)((packed >> (8 * e)) & 0xFFu) != want_byte)
                    ++bad;
            }
        }
    return bad;
}

void hc_pyrdown(const uint8_t* src, int sw, int sh, int spitch, uint8_t* dst, int dpitch) {
    const int dw = (sw + 1) / 2, dh = (sh + 1) / 2;
    HostImg g{src, spitch};
    for (int y = 0; y < dh; ++y)
        for (int x = 0; x < dw; ++x) dst[(size_t)y * dpitch + x] = (uint8_t)pyrdown_pixel(g, sw, sh, x, y);
}

struct HostPyr {
    std::vector<std::vector<uint8_t>> lv;
    LkPyramid p;
    int nlevels;
};
static void build_pyr(HostPyr& hp, const uint8_t* img, int w, int h, int stride, int max_level, int win) {
    hp.lv.clear();
    hp.lv.emplace_back((size_t)w * h);
    for (int y = 0; y < h; ++y) memcpy(&hp.lv[0][(size_t)y * w], img + (size_t)y * stride, w);
    int lw = w, lh = h;
    hp.p.lv[0] = LkLevel{hp.lv[0].data(), w, h, w};
    int l = 1;
    for (; l <= max_level && l < SVO_MAX_LEVELS; ++l) {
        const int nw = (lw + 1) / 2, nh = (lh + 1) / 2;
        if (nw <= win || nh <= win) break;
        hp.lv.emplace_back((size_t)nw * nh);
        hc_pyrdown(hp.lv[l - 1].data(), lw, lh, lw, hp.lv[l].data(), nw);
        lw = nw;
        lh = nh;
    }
    hp.nlevels = l;
    for (int k = 0; k < SVO_MAX_LEVELS; ++k) {
        const int q = k < l ? k : l - 1;
        int qw = w, qh = h;
        for (int i = 0; i < q; ++i) { qw = (qw + 1) / 2; qh = (qh + 1) / 2; }
        hp.p.lv[k] = LkLevel{hp.lv[q].data(), qw, qh, qw};
    }
}

#ifdef SVO_LK_COUNT
void hc_lk_counts(long long* out) {
    for (int i = 0; i < 8; ++i) out[i] = g_lk_count[i];
}
#endif
static int* g_iters_out = nullptr;   // optional per-feature iteration counts (analysis scripts)
void hc_lk_set_iters_out(int* p) { g_iters_out = p; }
int hc_lk(const uint8_t* prev, const uint8_t* next, int w, int h, int stride, int max_level, int max_iters,
          double eps, float min_eig, const float* pts, int n, float* out, uint8_t* status) {
    HostPyr I, J;
    build_pyr(I, prev, w, h, stride, max_level, SVO_LK_WIN);
    build_pyr(J, next, w, h, stride, max_level, SVO_LK_WIN);
    LkConfig cfg;
    cfg.nlevels = I.nlevels;
    cfg.max_iters = max_iters;
    cfg.eps_sq = eps * eps;
    cfg.min_eig = min_eig;
    lk_config_finish(cfg);
    static LkScratch patch;
    int total_iters = 0;
    for (int i = 0; i < n; ++i) {
        float ox, oy;
        int it;
        status[i] = (uint8_t)lk_track_feature<1, SerialRed>(I.p, J.p, cfg, pts[2 * i], pts[2 * i + 1], 0, patch, ox,
                                                            oy, it);
        out[2 * i] = ox;
        out[2 * i + 1] = oy;
        total_iters += it;
        if (g_iters_out) g_iters_out[i] = it;
    }
    return total_iters;
}

// ---- interleaved multi-channel images (the CLI-faithful 8UC3 mode)
struct HostImgCn {
    const uint8_t* p;
    int pitch, cn, ch;
    int operator()(int x, int y) const { return p[(size_t)y * pitch + x * cn + ch]; }
};
void hc_pyrdown_cn(const uint8_t* src, int sw, int sh, int spitch, int cn, uint8_t* dst, int dpitch) {
    const int dw = (sw + 1) / 2, dh = (sh + 1) / 2;
    for (int ch = 0; ch < cn; ++ch) {
        HostImgCn g{src, spitch, cn, ch};
        for (int y = 0; y < dh; ++y)
            for (int x = 0; x < dw; ++x) dst[(size_t)y * dpitch + x * cn + ch] = (uint8_t)pyrdown_pixel(g, sw, sh, x, y);
    }
}
static void build_pyr_cn(HostPyr& hp, const uint8_t* img, int w, int h, int stride, int cn, int max_level, int win) {
    hp.lv.clear();
    hp.lv.emplace_back((size_t)w * cn * h);
    for (int y = 0; y < h; ++y) memcpy(&hp.lv[0][(size_t)y * w * cn], img + (size_t)y * stride, (size_t)w * cn);
    int lw = w, lh = h, l = 1;
    std::vector<int> ws{w}, hs{h};
    for (; l <= max_level && l < SVO_MAX_LEVELS; ++l) {
        const int nw = (lw + 1) / 2, nh = (lh + 1) / 2;
        if (nw <= win || nh <= win) break;
        hp.lv.emplace_back((size_t)nw * cn * nh);
        hc_pyrdown_cn(hp.lv[l - 1].data(), lw, lh, lw * cn, cn, hp.lv[l].data(), nw * cn);
        lw = nw;
        lh = nh;
        ws.push_back(nw);
        hs.push_back(nh);
    }
    hp.nlevels = l;
    for (int k = 0; k < SVO_MAX_LEVELS; ++k) {
        const int q = k < l ? k : l - 1;
        hp.p.lv[k] = LkLevel{hp.lv[q].data(), ws[q], hs[q], ws[q] * cn};
    }
}
int hc_lk_cn(const uint8_t* prev, const uint8_t* next, int w, int h, int stride, int cn, int max_level, int max_iters,
             double eps, float min_eig, const float* pts, int n, float* out, uint8_t* status) {
    HostPyr I, J;
    build_pyr_cn(I, prev, w, h, stride, cn, max_level, SVO_LK_WIN);
    build_pyr_cn(J, next, w, h, stride, cn, max_level, SVO_LK_WIN);
    LkConfig cfg;
    cfg.nlevels = I.nlevels;
    cfg.max_iters = max_iters;
    cfg.eps_sq = eps * eps;
    cfg.min_eig = min_eig;
    lk_config_finish(cfg);
    static LkScratchCn patch;
    int total_iters = 0;
    for (int i = 0; i < n; ++i) {
        float ox, oy;
        int it;
        status[i] = (uint8_t)lk_track_feature_cn<1, SerialRed>(I.p, J.p, cfg, cn, pts[2 * i], pts[2 * i + 1], 0, patch,
                                                               ox, oy, it);
        out[2 * i] = ox;
        out[2 * i + 1] = oy;
        total_iters += it;
    }
    return total_iters;
}

void hc_triangulate(const float* Pl, const float* Pr, const float* pL, const float* pR, int n, float* X3) {
    for (int i = 0; i < n; ++i)
        triangulate_dlt(Pl, Pr, pL[2 * i], pL[2 * i + 1], pR[2 * i], pR[2 * i + 1], X3 + 3 * i);
}

struct SetEval {
    const float* X3;
    const float* uv;
    const int* idx;
    int n;
    Intrinsics K;
    void operator()(const Pose& P, double* acc, bool jac) const {
        for (int i = 0; i < SVO_PNP_ACC; ++i) acc[i] = 0.0;
        for (int i = 0; i < n; ++i) {
            const int k = idx ? idx[i] : i;
            pnp_accumulate(P, K, X3 + 3 * k, uv + 2 * k, acc, jac);
        }
    }
};

// serial emulation of select_kernel's sampling + pnp_hyp_kernel + pnp_final_kernel (same cores)
int g_chain = 2;   // 2 = EPnP hypotheses (what the kernels run); 0/1 = LM hypotheses (experiments)
void hc_set_chain(int c) { g_chain = c; }
int hc_pnp_ransac(const float* K9, const float* uv, const float* X3, int n, double* R, double* t, int n_hyp,
                  int max_iters, double confidence, float reproj_thr_sq, int* inliers, int* success) {
    Intrinsics K{K9[0], K9[4], K9[2], K9[5]};
    *success = 0;
    if (n < 5) return 0;
    std::vector<unsigned> raw(SVO_RNG_TABLE);
    uint64_t state = 0xFFFFFFFFFFFFFFFFull;
    for (int i = 0; i < SVO_RNG_TABLE; ++i) raw[i] = cv_rng_next(state);
    std::vector<int> samples((size_t)n_hyp * 5), counts(n_hyp, 0);
    ransac_build_samples([&](int pos) { return (int)(raw[pos] % (unsigned)n); }, n_hyp, samples.data());
    std::vector<Pose> poses(n_hyp);
    for (int hyp = 0; hyp < n_hyp; ++hyp) {
        const int* idx = samples.data() + 5 * hyp;
        if (idx[0] < 0) continue;
        Pose P;
        memcpy(P.R, R, sizeof(double) * 9);
        memcpy(P.t, t, sizeof(double) * 3);
        if (g_chain && hyp > 0) P = poses[hyp - 1];
        if (g_chain == 2) {
            float Xs[15], Us[10];
            for (int i = 0; i < 5; ++i) {
                for (int j = 0; j < 3; ++j) Xs[3 * i + j] = X3[3 * idx[i] + j];
                Us[2 * i] = uv[2 * idx[i]];
                Us[2 * i + 1] = uv[2 * idx[i] + 1];
            }
            static EpnpWork w;
            if (!epnp5_solve<1, 1, SerialRedD>(w, K, Xs, Us, 0, 0, P)) P.t[0] = NAN;
        } else {
            SetEval ev{X3, uv, idx, 5, K};
            pnp_refine_lm(P, ev, 20);
        }
        bool finite = true;
        for (int i = 0; i < 9; ++i) finite &= std::isfinite(P.R[i]);
        for (int i = 0; i < 3; ++i) finite &= std::isfinite(P.t[i]);
        int cnt = 0;
        for (int i = 0; i < n; ++i) cnt += reproj_err_sq(P, K, X3 + 3 * i, uv + 2 * i) <= reproj_thr_sq;
        counts[hyp] = finite ? cnt : 0;
        poses[hyp] = P;
    }
    int best_cnt = 0;
    const int bh = ransac_replay(counts.data(), n_hyp, n, max_iters, confidence, best_cnt);
    if (bh < 0 || best_cnt < 5) return 0;
    Pose best = poses[bh];
    int n_in = 0;
    for (int i = 0; i < n; ++i)
        if (reproj_err_sq(best, K, X3 + 3 * i, uv + 2 * i) <= reproj_thr_sq) inliers[n_in++] = i;
    SetEval ev{X3, uv, inliers, n_in, K};
    pnp_refine_lm(best, ev, 20);
    memcpy(R, best.R, sizeof(double) * 9);
    memcpy(t, best.t, sizeof(double) * 3);
    *success = 1;
    return n_in;
}

// plain LM refinement on all n points (solvePnP ITERATIVE with useExtrinsicGuess)
void hc_pnp_refine(const float* K9, const float* uv, const float* X3, int n, double* R, double* t, int max_iter) {
    Intrinsics K{K9[0], K9[4], K9[2], K9[5]};
    Pose P;
    memcpy(P.R, R, sizeof(double) * 9);
    memcpy(P.t, t, sizeof(double) * 3);
    SetEval ev{X3, uv, nullptr, n, K};
    pnp_refine_lm(P, ev, max_iter);
    memcpy(R, P.R, sizeof(double) * 9);
    memcpy(t, P.t, sizeof(double) * 3);
}

// EPnP on exactly 5 points (the RANSAC hypothesis generator)
int hc_epnp5(const float* K9, const float* uv, const float* X3, double* R, double* t) {
    Intrinsics K{K9[0], K9[4], K9[2], K9[5]};
    static EpnpWork w;
    Pose P;
    const bool ok = epnp5_solve<1, 1, SerialRedD>(w, K, X3, uv, 0, 0, P);
    memcpy(R, P.R, sizeof(double) * 9);
    memcpy(t, P.t, sizeof(double) * 3);
    return ok ? 1 : 0;
}

double hc_rotation_angle(const double* R) { return rotation_angle(R); }

int hc_bucket_score(int age, int strength, int fast_threshold) { return bucket_score(age, strength, fast_threshold); }

// key round trip + ordering property used by bucket_merge_kernel
int hc_bucket_key_roundtrip(int score, unsigned order, int strength) {
    const unsigned long long k = bucket_key(score, order, strength);
    return bucket_key_score(k) == score && bucket_key_order(k) == order && bucket_key_strength(k) == strength;
}
int hc_bucket_key_less(int s1, unsigned o1, int s2, unsigned o2) {
    return bucket_key(s1, o1, 0) < bucket_key(s2, o2, 0);
}

}  // extern "C"
