// svo_core.cuh -- arithmetic cores of the stereo-VO front end, written once as
// host+device inline functions.  The CUDA kernels (svo_kernels.cu) call them
// with LANES = 32 (one warp per feature); tests/host_cores.cpp compiles the same
// header with g++ and LANES = 1 so the arithmetic can be unit-tested on a box
// without a GPU.  Nothing here is a CPU fallback: the shipped library only
// instantiates the device side.
//
// Behavioural spec: SURVEY.md Appendix B (OpenCV calls made by the reference at
// src/feature_set.cpp:61, src/vo.cpp:50-53, 90-94, 200-215, 301-304).
#pragma once
#include <float.h>
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define SVO_HD __host__ __device__ __forceinline__
#else
#define SVO_HD inline
#endif

namespace svo {

// ---------------------------------------------------------------------------
// Border handling (cv::BORDER_REFLECT_101), clamped for degenerate sizes.
SVO_HD int reflect101(int i, int n) {
    if (i < 0) i = -i;
    if (i >= n) i = 2 * n - 2 - i;
    if (i < 0) i = 0;
    if (i >= n) i = n - 1;
    return i;
}

SVO_HD int imin(int a, int b) { return a < b ? a : b; }
SVO_HD int imax(int a, int b) { return a > b ? a : b; }

// ---------------------------------------------------------------------------
// FAST-9/16 (cv::FAST, TYPE_9_16) -- featureDetectionFast, feature_set.cpp:55-68.
// Ring offsets k = 0..15 (dx, dy), SURVEY Appendix B.1.
#define SVO_FAST_RING_DX {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1}
#define SVO_FAST_RING_DY {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3}

// s = the largest t such that 9 contiguous ring pixels are all >= t darker, or
// all >= t brighter, than the centre.  Corner <=> s > threshold; KeyPoint
// response = s - 1 (threshold independent).
SVO_HD int fast9_arc_score(int v, const int* r) {
    int d[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) d[k] = v - r[k];
    int lo2[16], hi2[16], lo4[16], hi4[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        lo2[k] = imin(d[k], d[(k + 1) & 15]);
        hi2[k] = imax(d[k], d[(k + 1) & 15]);
    }
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        lo4[k] = imin(lo2[k], lo2[(k + 2) & 15]);
        hi4[k] = imax(hi2[k], hi2[(k + 2) & 15]);
    }
    int a = -256, b = 256;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        int lo9 = imin(imin(lo4[k], lo4[(k + 4) & 15]), d[(k + 8) & 15]);
        int hi9 = imax(imax(hi4[k], hi4[(k + 4) & 15]), d[(k + 8) & 15]);
        a = imax(a, lo9);
        b = imin(b, hi9);
    }
    return imax(a, -b);
}

// ---------------------------------------------------------------------------
// Bucketing, FeatureSet::filterByBucketLocationInternal, feature_set.cpp:95-147.
// C integer division (truncation toward zero) as in Bucket::compute_score,
// feature_set.cpp:16-18.
SVO_HD int bucket_score(int age, int strength, int fast_threshold) {
    return age + (strength - fast_threshold) / 20;
}

// 64-bit selection key for the one-feature-per-bucket production path:
// max key == highest score, ties -> earliest position in the reference's
// feature order.  [63:48] score + 0x8000 | [47:16] ~order | [15:0] strength.
SVO_HD unsigned long long bucket_key(int score, unsigned order, int strength) {
    return ((unsigned long long)(unsigned)(score + 0x8000) << 48) |
           ((unsigned long long)(0xFFFFFFFFu - order) << 16) | (unsigned long long)(strength & 0xFFFF);
}
SVO_HD int bucket_key_score(unsigned long long k) { return (int)(k >> 48) - 0x8000; }
SVO_HD unsigned bucket_key_order(unsigned long long k) {
    return 0xFFFFFFFFu - (unsigned)((k >> 16) & 0xFFFFFFFFull);
}
SVO_HD int bucket_key_strength(unsigned long long k) { return (int)(k & 0xFFFF); }
// order values >= SVO_FAST_ORDER_BASE denote a FAST pixel (order - base = y*W + x)
#define SVO_FAST_ORDER_BASE (1u << 24)

// ---------------------------------------------------------------------------
// pyrDown (cv::pyrDown u8, 5x5 binomial, BORDER_REFLECT_101): SURVEY B.2.
template <class Img>
SVO_HD int pyrdown_pixel(const Img& src, int sw, int sh, int x, int y) {
    int acc = 0;
#pragma unroll
    for (int j = 0; j < 5; ++j) {
        const int yy = reflect101(2 * y + j - 2, sh);
        const int kj = (j == 0 || j == 4) ? 1 : (j == 2 ? 6 : 4);
        int row = 0;
#pragma unroll
        for (int i = 0; i < 5; ++i) {
            const int xx = reflect101(2 * x + i - 2, sw);
            const int ki = (i == 0 || i == 4) ? 1 : (i == 2 ? 6 : 4);
            row += ki * (int)src(xx, yy);
        }
        acc += kj * row;
    }
    return (acc + 128) >> 8;
}

// ---------------------------------------------------------------------------
// Pyramidal Lucas-Kanade (cv::calcOpticalFlowPyrLK CPU path), SURVEY B.3.
struct LkLevel {
    const uint8_t* img;
    int w, h, pitch;
    SVO_HD int at(int x, int y) const { return img[(size_t)y * pitch + x]; }
    SVO_HD int at_r(int x, int y) const { return img[(size_t)reflect101(y, h) * pitch + reflect101(x, w)]; }
};

#define SVO_MAX_LEVELS 4
struct LkPyramid {
    LkLevel lv[SVO_MAX_LEVELS];
    SVO_HD LkLevel level(int l) const { return lv[l]; }
};

struct LkConfig {
    int nlevels;      // maxLevel + 1 actually available
    int max_iters;    // criteria.maxCount
    double eps_sq;    // criteria.epsilon^2
    float min_eig;    // minEigThreshold
    // float32 brackets of eps_sq (1e-6 relative): the FP64 termination test of OpenCV is only evaluated when the
    // float32 value falls between them.  lk_config_finish() fills them.
    float eps_lo, eps_hi;
};
SVO_HD void lk_config_finish(LkConfig& c) {
    c.eps_lo = (float)(c.eps_sq * (1.0 - 1e-6));
    c.eps_hi = (float)(c.eps_sq * (1.0 + 1e-6));
}

#ifdef SVO_LK_COUNT
static long long g_lk_count[8];
#endif
// Reduction policy for LANES == 1 (host unit tests).
struct SerialRed {
    static SVO_HD int sum(int v) { return v; }
    static SVO_HD long long sum64(long long v) { return v; }
    static SVO_HD long long sum64g(long long v) { return v; }
    // exact integer sums of the two mismatch accumulators, converted to float32 (round to nearest)
    static SVO_HD void sum2f(long long a, long long b, bool, float& fa, float& fb) { fa = (float)a; fb = (float)b; }
    static SVO_HD void sync() {}
    // float32 window sums in OpenCV's order (cv_lk_sum_prod / cv_lk_sum_b below); one lane holds all 100 pixels
    static inline void cv_sum_prod3(const int* ix, const int* iy, int, unsigned*, float& a11, float& a12, float& a22);
    static inline void cv_sum_b(const int* p1, const int* p2, int, unsigned*, float& b1, float& b2);
    // OpenCV folds its five accumulators as tail + ((c0 + c2) + (c1 + c3)), c_k = the chain of pixels k | k + 4 of every
    // row.  While the absolute products of a GROUP -- even columns < 8 (c0, c2), odd columns < 8 (c1, c3), columns 8, 9
    // (tail) -- total <= 2^24, every float32 partial sum inside the group is an exactly representable integer in any
    // order, so the result is fl(T + fl(A + B)) of the three exact integer group sums.  p1 / p2: the products per pixel.
    static SVO_HD int lk_group(int col) { return col >= 8 ? 2 : (col & 1); }
    static inline bool sum_b_grouped(const int* p1, const int* p2, int, long long, long long, long long, long long,
                                     float& b1, float& b2);
    static inline bool sum_prod3_grouped(const int* ix, const int* iy, int, float& a11, float& a12, float& a22);
};

// cvRound(float): round half to even.
SVO_HD int round_half_even(float v) {
#if defined(__CUDA_ARCH__)
    return __float2int_rn(v);
#else
    return (int)lrintf(v);
#endif
}
SVO_HD int floor_int(float v) {
#if defined(__CUDA_ARCH__)
    return __float2int_rd(v);
#else
    return (int)floorf(v);
#endif
}

// Explicitly unfused float ops: OpenCV's x86 build does not contract these and the
// decision quantities (minEig, D, delta) must not depend on FMA contraction.
SVO_HD float fmul(float a, float b) {
#if defined(__CUDA_ARCH__)
    return __fmul_rn(a, b);
#else
    volatile float r = a * b;
    return r;
#endif
}
SVO_HD float fadd(float a, float b) {
#if defined(__CUDA_ARCH__)
    return __fadd_rn(a, b);
#else
    volatile float r = a + b;
    return r;
#endif
}
SVO_HD float fsub(float a, float b) {
#if defined(__CUDA_ARCH__)
    return __fsub_rn(a, b);
#else
    volatile float r = a - b;
    return r;
#endif
}

#define SVO_LK_WIN 10
#define SVO_LK_PATCH (SVO_LK_WIN + 3)     // 13: window + 1 (bilinear) + 2 (Scharr halo)
#define SVO_LK_PATCH_PITCH 16
#define SVO_LK_NPIX (SVO_LK_WIN * SVO_LK_WIN)
#define SVO_LK_DW (SVO_LK_WIN + 1)        // 11 x 11 derivative positions
#define SVO_LK_DP 12                      // derivative patch pitch
#define SVO_LK_JR 18                      // rows of the staged J patch
#define SVO_LK_JC 24                      // pixel pairs per row of the staged J patch

// Per-feature scratch shared by the cooperating lanes (shared memory on the device).
struct LkScratch {
    uint8_t ipatch[SVO_LK_PATCH * SVO_LK_PATCH_PITCH];     // template neighbourhood, u8
    short dxp[SVO_LK_DW * SVO_LK_DP];                      // Scharr dx of the 11x11 window positions
    short dyp[SVO_LK_DW * SVO_LK_DP];                      // Scharr dy
    unsigned short jp[SVO_LK_JR * SVO_LK_JC];              // J patch as horizontal pixel pairs p(x) | p(x+1) << 8
    unsigned stage[3 * SVO_LK_NPIX];                       // window products regrouped into OpenCV's accumulation chains
};

SVO_HD void lk_weights(float a, float b, int& iw00, int& iw01, int& iw10, int& iw11) {
    const float W = 16384.f;   // 1 << W_BITS
    iw00 = round_half_even(fmul(fmul(fsub(1.f, a), fsub(1.f, b)), W));
    iw01 = round_half_even(fmul(fmul(a, fsub(1.f, b)), W));
    iw10 = round_half_even(fmul(fmul(fsub(1.f, a), b), W));
    iw11 = 16384 - iw00 - iw01 - iw10;
}

template <int LANES> struct LkAcc { typedef int type; };
template <> struct LkAcc<1> { typedef long long type; };   // host unit tests: one lane carries all 100 pixels

// Which window pixel (row, col of the 10x10 window) a (lane, slot) pair owns.
//   LANES == 32: lanes 0..29 own column lane % 10 of rows lane / 10 + 3 * slot (slot 3 only exists for the
//   first row group), so the four pixels of a lane sit at CONSTANT row strides (one address register, immediate
//   offsets) and no per-slot index arithmetic is left in the iteration loop.  Slots without a pixel point at
//   a valid pixel and carry zero template values, which makes their contribution exactly zero without predicates.
//   LANES == 1 (host unit tests): one lane walks all 100 pixels.
template <int LANES> struct LkMap;
template <> struct LkMap<32> {
    static constexpr int SLOTS = 4;
    static SVO_HD bool valid(int lane, int s) { return lane < 30 && lane / 10 + 3 * s < SVO_LK_WIN; }
    // slots 0..2 keep the constant stride for every lane (lanes 30, 31 read rows 3, 6, 9 with zero template values)
    static SVO_HD int row(int lane, int s) { return s < 3 ? lane / 10 + 3 * s : (lane < 10 ? 9 : 0); }
    static SVO_HD int col(int lane, int) { return lane % 10; }
};
template <> struct LkMap<1> {
    static constexpr int SLOTS = SVO_LK_NPIX;
    static SVO_HD bool valid(int, int) { return true; }
    static SVO_HD int row(int, int s) { return s / SVO_LK_WIN; }
    static SVO_HD int col(int, int s) { return s % SVO_LK_WIN; }
};

// c + a.lo16 * b.byte0 + a.hi16 * b.byte1 with SIGNED 16-bit weights (iw11 = 2^14 - iw00 - iw01 - iw10
// can be -1 or -2) and UNSIGNED pixels: one IDP.2A (dp2a.lo.s32.u32) on the device
SVO_HD int dp2a_lo_su(unsigned a, unsigned b, int c) {
#if defined(__CUDA_ARCH__)
    int d;
    asm("dp2a.lo.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
#else
    return c + (int)(short)(a & 0xFFFFu) * (int)(b & 0xFFu) + (int)(short)(a >> 16) * (int)((b >> 8) & 0xFFu);
#endif
}
SVO_HD unsigned pack_weights(int lo, int hi) { return ((unsigned)lo & 0xFFFFu) | ((unsigned)hi << 16); }

// __byte_perm(a, b, sel): byte i of the result = byte sel[4i+2:4i] of the 8-byte pool {b, a} (bit 4i+3: its sign bit
// replicated instead)
SVO_HD unsigned byte_perm(unsigned a, unsigned b, unsigned sel) {
#if defined(__CUDA_ARCH__)
    return __byte_perm(a, b, sel);   // NB: masks the selector to three bits per nibble; sign mode: byte_perm_sx
#else
    const uint64_t pool = ((uint64_t)b << 32) | a;
    unsigned r = 0;
    for (int i = 0; i < 4; ++i) {
        const unsigned n = (sel >> (4 * i)) & 15u;
        unsigned byte = (unsigned)((pool >> (8 * (n & 7u))) & 0xFFu);
        if (n & 8u) byte = (byte & 0x80u) ? 0xFFu : 0u;   // selector bit 3: replicate the byte's sign bit
        r |= byte << (8 * i);
    }
    return r;
#endif
}

// prmt.b32 with the full selector: bit 3 of a nibble replicates the selected byte's sign bit over the result byte
// (__byte_perm() cannot express that)
template <unsigned SEL>
SVO_HD unsigned byte_perm_sx(unsigned a, unsigned b) {
#if defined(__CUDA_ARCH__)
    unsigned r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "n"(SEL));
    return r;
#else
    return byte_perm(a, b, SEL);   // the host statement implements the sign mode
#endif
}

// Stage the J patch whose top-left pixel is (jx0, jy0): jp[y][x] = J(jx0+x, jy0+y) | J(jx0+x+1, jy0+y) << 8
// with BORDER_REFLECT_101.  jx0 is a multiple of 4, so interior patches are fetched as aligned words.
template <int LANES>
SVO_HD void lk_stage_j(const LkLevel& J, int jx0, int jy0, int lane, unsigned short* jp) {
    const bool interior = jx0 >= 0 && jy0 >= 0 && jx0 + SVO_LK_JC + 4 <= J.w && jy0 + SVO_LK_JR <= J.h &&
                          (J.pitch & 3) == 0 && (((size_t)J.img) & 3) == 0;
    if (interior) {
        constexpr int WORDS = SVO_LK_JC / 4;   // 6 words -> 24 pairs (the 25th pixel comes from the next word)
        for (int it = lane; it < SVO_LK_JR * WORDS; it += LANES) {
            const int y = it / WORDS, k = it - y * WORDS;
            const uint8_t* src = J.img + (size_t)(jy0 + y) * J.pitch + jx0 + 4 * k;
            unsigned w0, w1;
#if defined(__CUDA_ARCH__)
            w0 = *reinterpret_cast<const unsigned*>(src);
            w1 = *reinterpret_cast<const unsigned*>(src + 4);
#else
            w0 = src[0] | (src[1] << 8) | (src[2] << 16) | ((unsigned)src[3] << 24);
            w1 = src[4] | (src[5] << 8) | (src[6] << 16) | ((unsigned)src[7] << 24);
#endif
            // pairs (b0 b1)(b1 b2) and (b2 b3)(b3 b4): one PRMT each, one 8-byte store
            unsigned* dst = reinterpret_cast<unsigned*>(jp + y * SVO_LK_JC + 4 * k);
#if defined(__CUDA_ARCH__)
            *reinterpret_cast<uint2*>(dst) = make_uint2(byte_perm(w0, w0, 0x2110u), byte_perm(w0, w1, 0x4332u));
#else
            dst[0] = byte_perm(w0, w0, 0x2110u);
            dst[1] = byte_perm(w0, w1, 0x4332u);
#endif
        }
    } else {
        for (int it = lane; it < SVO_LK_JR * SVO_LK_JC; it += LANES) {
            const int y = it / SVO_LK_JC, x = it - y * SVO_LK_JC;
            const uint8_t* row = J.img + (size_t)reflect101(jy0 + y, J.h) * J.pitch;
            jp[it] = (unsigned short)(row[reflect101(jx0 + x, J.w)] | (row[reflect101(jx0 + x + 1, J.w)] << 8));
        }
    }
}

// Stage the 13x13 neighbourhood of the template window whose top-left pixel is (x0, y0) (BORDER_REFLECT_101)
// into sc.ipatch (pitch 16).  Interior windows are fetched as 13 x 4 aligned words; the patch then starts at
// byte column `xo` = x0 & 3 of each row.  Returns xo.
template <int LANES>
SVO_HD int lk_stage_i(const LkLevel& I, int x0, int y0, int lane, uint8_t* ip) {
    const bool interior = x0 >= 0 && y0 >= 0 && x0 + SVO_LK_PATCH <= I.w && y0 + SVO_LK_PATCH <= I.h &&
                          (I.pitch & 3) == 0 && (((size_t)I.img) & 3) == 0 && ((x0 & ~3) + 16 <= I.pitch);
    if (interior) {
        const int xa = x0 & ~3;
        for (int it = lane; it < SVO_LK_PATCH * 4; it += LANES) {
            const int y = it >> 2, k = it & 3;
            const uint8_t* src = I.img + (size_t)(y0 + y) * I.pitch + xa + 4 * k;
            unsigned w0;
#if defined(__CUDA_ARCH__)
            w0 = *reinterpret_cast<const unsigned*>(src);
#else
            w0 = src[0] | (src[1] << 8) | (src[2] << 16) | ((unsigned)src[3] << 24);
#endif
            *reinterpret_cast<unsigned*>(ip + y * SVO_LK_PATCH_PITCH + 4 * k) = w0;
        }
        return x0 & 3;
    }
    for (int idx = lane; idx < SVO_LK_PATCH * SVO_LK_PATCH; idx += LANES) {
        const int py = idx / SVO_LK_PATCH, px = idx - py * SVO_LK_PATCH;
        ip[py * SVO_LK_PATCH_PITCH + px] = (uint8_t)I.at_r(x0 + px, y0 + py);
    }
    return 0;
}

// ---- float32 sums in the order cv::calcOpticalFlowPyrLK's SIMD128 code forms them (lkpyramid.cpp, winSize 10, cn 1):
// per window row the first 8 pixels go through 4-lane float accumulators, pixels 8 and 9 through a scalar float
// accumulator; the lanes are folded at the end.  Inputs: row-major 10 x 10 integer arrays.
SVO_HD float cv_lk_reduce4(const float* q) { return fadd(fadd(q[0], q[2]), fadd(q[1], q[3])); }
// sum a[s] * b[s]: lane l accumulates pixels l and l + 4 of every row (float products), scalar tail pixels 8, 9
SVO_HD float cv_lk_sum_prod(const int* a, const int* b) {
    float q[4] = {0.f, 0.f, 0.f, 0.f}, tail = 0.f;
    for (int y = 0; y < SVO_LK_WIN; ++y) {
        for (int h = 0; h < 2; ++h)
            for (int l = 0; l < 4; ++l) {
                const int sidx = y * SVO_LK_WIN + 4 * h + l;
                q[l] = fadd(q[l], fmul((float)a[sidx], (float)b[sidx]));
            }
        for (int x = 8; x < SVO_LK_WIN; ++x) {
            const int sidx = y * SVO_LK_WIN + x;
            tail = fadd(tail, (float)(a[sidx] * b[sidx]));
        }
    }
    return fadd(tail, cv_lk_reduce4(q));
}
// mismatch vector: b1 = sum diff * Ix, b2 = sum diff * Iy with OpenCV's pairing (the products of pixels x and
// x + 4 of a row are added as int32 before the conversion to float).  p1 / p2: the integer products per pixel.
SVO_HD void cv_lk_sum_b(const int* p1, const int* p2, float& b1, float& b2) {
    float qb0[4] = {0.f, 0.f, 0.f, 0.f}, qb1[4] = {0.f, 0.f, 0.f, 0.f}, t1 = 0.f, t2 = 0.f;
    for (int y = 0; y < SVO_LK_WIN; ++y) {
        const int* gx = p1 + y * SVO_LK_WIN;
        const int* gy = p2 + y * SVO_LK_WIN;
        qb0[0] = fadd(qb0[0], (float)(gx[0] + gx[4]));
        qb0[1] = fadd(qb0[1], (float)(gy[0] + gy[4]));
        qb0[2] = fadd(qb0[2], (float)(gx[1] + gx[5]));
        qb0[3] = fadd(qb0[3], (float)(gy[1] + gy[5]));
        qb1[0] = fadd(qb1[0], (float)(gx[2] + gx[6]));
        qb1[1] = fadd(qb1[1], (float)(gy[2] + gy[6]));
        qb1[2] = fadd(qb1[2], (float)(gx[3] + gx[7]));
        qb1[3] = fadd(qb1[3], (float)(gy[3] + gy[7]));
        for (int x = 8; x < SVO_LK_WIN; ++x) {
            t1 = fadd(t1, (float)gx[x]);
            t2 = fadd(t2, (float)gy[x]);
        }
    }
    float sv[4];
    for (int l = 0; l < 4; ++l) sv[l] = fadd(qb0[l], qb1[l]);
    b1 = fadd(t1, fadd(sv[0], sv[2]));
    b2 = fadd(t2, fadd(sv[1], sv[3]));
}

inline void SerialRed::cv_sum_prod3(const int* ix, const int* iy, int, unsigned*, float& a11, float& a12, float& a22) {
    a11 = cv_lk_sum_prod(ix, ix);
    a12 = cv_lk_sum_prod(ix, iy);
    a22 = cv_lk_sum_prod(iy, iy);
}
inline void SerialRed::cv_sum_b(const int* p1, const int* p2, int, unsigned*, float& b1, float& b2) {
#ifdef SVO_LK_COUNT
    {   // how many of the ten chains would pass a per-chain exactness bound
        int bad = 0;
        for (int k = 0; k < 2; ++k) {
            const int* p = k ? p2 : p1;
            for (int c = 0; c < 5; ++c) {
                long long a = 0;
                for (int y = 0; y < SVO_LK_WIN; ++y) {
                    const int x0 = c < 4 ? c : 8, x1 = c < 4 ? c + 4 : 9;
                    a += llabs((long long)p[y * SVO_LK_WIN + x0]) + llabs((long long)p[y * SVO_LK_WIN + x1]);
                }
                if (a > (1 << 24)) ++bad;
            }
        }
        ++g_lk_count[bad == 0 ? 3 : 4];
        g_lk_count[5] += bad;
    }
#endif
    cv_lk_sum_b(p1, p2, b1, b2);
}

// While the absolute products of an accumulator group total <= 2^24 the group's float32 partial sums are exact
// integers in any order: the result is the fold of the three integer group sums (see SerialRed).
inline bool SerialRed::sum_b_grouped(const int* p1, const int* p2, int, long long, long long, long long, long long,
                                     float& b1, float& b2) {
    long long S[2][3] = {{0, 0, 0}, {0, 0, 0}}, A[2][3] = {{0, 0, 0}, {0, 0, 0}};
    for (int s = 0; s < SVO_LK_NPIX; ++s) {
        const int g = lk_group(s % SVO_LK_WIN);
        S[0][g] += p1[s];
        S[1][g] += p2[s];
        A[0][g] += p1[s] < 0 ? -(long long)p1[s] : p1[s];
        A[1][g] += p2[s] < 0 ? -(long long)p2[s] : p2[s];
    }
    bool ok = true;
    for (int k = 0; k < 2; ++k)
        for (int g = 0; g < 3; ++g) ok = ok && A[k][g] <= (1 << 24);
#ifdef SVO_LK_COUNT
    ++g_lk_count[ok ? 0 : 1];
#endif
    if (!ok) return false;
    b1 = fadd((float)S[0][2], fadd((float)S[0][0], (float)S[0][1]));
    b2 = fadd((float)S[1][2], fadd((float)S[1][0], (float)S[1][1]));
    return true;
}
// the same for the structure tensor (every product of a group is exact when its squares total <= 2^24)
inline bool SerialRed::sum_prod3_grouped(const int* ix, const int* iy, int, float& a11, float& a12, float& a22) {
    long long S[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
    for (int s = 0; s < SVO_LK_NPIX; ++s) {
        const int g = lk_group(s % SVO_LK_WIN);
        S[0][g] += (long long)ix[s] * ix[s];
        S[1][g] += (long long)ix[s] * iy[s];
        S[2][g] += (long long)iy[s] * iy[s];
    }
    bool ok = true;
    for (int g = 0; g < 3; ++g) ok = ok && S[0][g] <= (1 << 24) && S[2][g] <= (1 << 24);
#ifdef SVO_LK_COUNT
    ++g_lk_count[ok ? 6 : 7];
#endif
    if (!ok) return false;
    a11 = fadd((float)S[0][2], fadd((float)S[0][0], (float)S[0][1]));
    a12 = fadd((float)S[1][2], fadd((float)S[1][0], (float)S[1][1]));
    a22 = fadd((float)S[2][2], fadd((float)S[2][0], (float)S[2][1]));
    return true;
}

// Track ONE feature through all pyramid levels.  Executed by LANES cooperating lanes (every lane runs
// the same scalar control flow; window pixels are distributed by LkMap; Red sums across lanes).
// Returns status (1 = tracked) and writes the output position.
// Pyr: anything with `LkLevel level(int) const` (LkPyramid; the kernels pass a view that reads the level geometry from
// the kernel parameters, so that no per-leg pyramid struct has to live in local memory).
template <int LANES, class Red, class Pyr>
SVO_HD int lk_track_feature(const Pyr& Ipyr, const Pyr& Jpyr, const LkConfig& cfg,
                            float ptx, float pty, int lane, LkScratch& sc, float& outx,
                            float& outy, int& iters_out) {
    typedef LkMap<LANES> Map;
    constexpr int SLOTS = Map::SLOTS;
    const float half_win = 4.5f;   // (winSize - 1) * 0.5f
    const float FLT_SCALE = 1.f / (1 << 20);
    // termination test `ddx^2 + ddy^2 <= eps^2` is evaluated in FP64 by OpenCV; a float32 evaluation decides
    // it whenever the value is not within 1e-6 relative of the threshold (exact FP64 otherwise)
    const float eps_lo = cfg.eps_lo, eps_hi = cfg.eps_hi;
    int status = 1;
    float nx = 0.f, ny = 0.f;      // nextPts[ptidx]
    int iters = 0;
    // per-lane window pixel offsets inside the template patch (pitch 16), the derivative patch (pitch 12) and
    // the staged J patch (pitch SVO_LK_JC)
    int offI[SLOTS], offD[SLOTS], offJ[SLOTS];
    bool val[SLOTS];
#pragma unroll
    for (int s = 0; s < SLOTS; ++s) {
        const int wy = Map::row(lane, s), wx = Map::col(lane, s);
        val[s] = Map::valid(lane, s);
        offI[s] = (wy + 1) * SVO_LK_PATCH_PITCH + wx + 1;
        offD[s] = wy * SVO_LK_DP + wx;
        offJ[s] = wy * SVO_LK_JC + wx;
    }

    for (int level = cfg.nlevels - 1; level >= 0; --level) {
        const LkLevel I = Ipyr.level(level);
        const LkLevel J = Jpyr.level(level);
        const float scale = 1.f / (float)(1 << level);
        float prx = fmul(ptx, scale), pry = fmul(pty, scale);
        float nxt_x, nxt_y;
        if (level == cfg.nlevels - 1) {
            nxt_x = prx;
            nxt_y = pry;
        } else {
            nxt_x = fmul(nx, 2.f);
            nxt_y = fmul(ny, 2.f);
        }
        nx = nxt_x;
        ny = nxt_y;

        prx = fsub(prx, half_win);
        pry = fsub(pry, half_win);
        const int ipx = floor_int(prx), ipy = floor_int(pry);
        if (ipx < -SVO_LK_WIN || ipx >= I.w || ipy < -SVO_LK_WIN || ipy >= I.h) {
            if (level == 0) status = 0;
            continue;
        }
        int iw00, iw01, iw10, iw11;
        lk_weights(fsub(prx, (float)ipx), fsub(pry, (float)ipy), iw00, iw01, iw10, iw11);

        // ---- stage the 13x13 neighbourhood of the template window (BORDER_REFLECT_101)
        Red::sync();
        const int xo = lk_stage_i<LANES>(I, ipx - 1, ipy - 1, lane, sc.ipatch);
        // J patch around the start of the search (independent of the template work)
        float cx = fsub(nxt_x, half_win), cy = fsub(nxt_y, half_win);   // nextPt -= halfWin
        int jx0 = (floor_int(cx) - 5) & ~3, jy0 = floor_int(cy) - 3;
        {
            const int fx0 = floor_int(cx), fy0 = floor_int(cy);
            if (!(fx0 < -SVO_LK_WIN || fx0 >= J.w || fy0 < -SVO_LK_WIN || fy0 >= J.h)) lk_stage_j<LANES>(J, jx0, jy0, lane, sc.jp);
        }
        Red::sync();
        // ---- Scharr derivatives on the 11x11 positions of the window (zero outside the level:
        // derivBorder = BORDER_CONSTANT), packed (dx, dy) int16
        const bool dinside = ipx >= 0 && ipx + SVO_LK_DW <= I.w && ipy >= 0 && ipy + SVO_LK_DW <= I.h;
        for (int idx = lane, y = lane / SVO_LK_DW, x = lane - y * SVO_LK_DW; idx < SVO_LK_DW * SVO_LK_DW;
             idx += LANES, x += LANES % SVO_LK_DW, y += LANES / SVO_LK_DW + (x >= SVO_LK_DW ? 1 : 0),
             x -= x >= SVO_LK_DW ? SVO_LK_DW : 0) {
            int dx = 0, dy = 0;
            if (dinside || (ipx + x >= 0 && ipx + x < I.w && ipy + y >= 0 && ipy + y < I.h)) {
                const uint8_t* p = sc.ipatch + y * SVO_LK_PATCH_PITCH + x + xo;   // top-left of the 3x3 around patch (x+1, y+1)
                const int a00 = p[0], a01 = p[1], a02 = p[2];
                const int a10 = p[SVO_LK_PATCH_PITCH], a12 = p[SVO_LK_PATCH_PITCH + 2];
                const int a20 = p[2 * SVO_LK_PATCH_PITCH], a21 = p[2 * SVO_LK_PATCH_PITCH + 1], a22 = p[2 * SVO_LK_PATCH_PITCH + 2];
                dx = 3 * (a02 + a22 - a00 - a20) + 10 * (a12 - a10);
                dy = 3 * (a20 + a22 - a00 - a02) + 10 * (a21 - a01);
            }
            sc.dxp[y * SVO_LK_DP + x] = (short)dx;
            sc.dyp[y * SVO_LK_DP + x] = (short)dy;
        }
        Red::sync();

        // ---- template window: Ival (x32), Ix, Iy (bilinear), structure tensor
        int Iv[SLOTS], Ixv[SLOTS], Iyv[SLOTS];
#pragma unroll
        for (int s = 0; s < SLOTS; ++s) {
            const uint8_t* p = sc.ipatch + offI[s] + xo;
            const int iv = (p[0] * iw00 + p[1] * iw01 + p[SVO_LK_PATCH_PITCH] * iw10 + p[SVO_LK_PATCH_PITCH + 1] * iw11 + (1 << 8)) >> 9;
            const short* qx = sc.dxp + offD[s];
            const short* qy = sc.dyp + offD[s];
            const int ix = ((int)qx[0] * iw00 + (int)qx[1] * iw01 + (int)qx[SVO_LK_DP] * iw10 + (int)qx[SVO_LK_DP + 1] * iw11 + (1 << 13)) >> 14;
            const int iy = ((int)qy[0] * iw00 + (int)qy[1] * iw01 + (int)qy[SVO_LK_DP] * iw10 + (int)qy[SVO_LK_DP + 1] * iw11 + (1 << 13)) >> 14;
            Iv[s] = val[s] ? (int)(short)iv : 0;
            Ixv[s] = val[s] ? (int)(short)ix : 0;
            Iyv[s] = val[s] ? (int)(short)iy : 0;
        }
        // OpenCV accumulates the structure tensor in float32 (4 SIMD lanes + a scalar tail): same order, same roundings
        float A11, A12, A22;
        if (!Red::sum_prod3_grouped(Ixv, Iyv, lane, A11, A12, A22)) Red::cv_sum_prod3(Ixv, Iyv, lane, sc.stage, A11, A12, A22);
        A11 = fmul(A11, FLT_SCALE);
        A12 = fmul(A12, FLT_SCALE);
        A22 = fmul(A22, FLT_SCALE);
        float D = fsub(fmul(A11, A22), fmul(A12, A12));
        const float dA = fsub(A11, A22);
        const float disc = fadd(fmul(dA, dA), fmul(fmul(4.f, A12), A12));
        const float min_eig = fsub(fadd(A22, A11), sqrtf(disc)) / (float)(2 * SVO_LK_NPIX);
        if (min_eig < cfg.min_eig || D < FLT_EPSILON) {
            if (level == 0) status = 0;
            continue;
        }
        D = 1.f / D;
        // iteration-invariant part of the window test: (inx, iny) inside the staged patch AND not fully outside
        // the level  <=>  0 <= inx - lox <= rx, 0 <= iny - loy <= ry   (rx / ry < 0: never)
        int lox = imax(jx0, -SVO_LK_WIN), loy = imax(jy0, -SVO_LK_WIN);
        int rx = imin(jx0 + SVO_LK_JC - SVO_LK_WIN, J.w - 1) - lox, ry = imin(jy0 + SVO_LK_JR - SVO_LK_WIN - 1, J.h - 1) - loy;

        float pdx = 0.f, pdy = 0.f;
        for (int j = 0; j < cfg.max_iters; ++j) {
            const int inx = floor_int(cx), iny = floor_int(cy);
            const int ax = inx - lox, ay = iny - loy;
            if ((ax | (rx - ax) | ay | (ry - ay)) < 0) {   // rare
                if (inx < -SVO_LK_WIN || inx >= J.w || iny < -SVO_LK_WIN || iny >= J.h) {   // window fully outside the level
                    if (level == 0) status = 0;
                    break;
                }
                // the window left the staged patch: re-centre it
                jx0 = (inx - 5) & ~3;
                jy0 = iny - 3;
                Red::sync();
                lk_stage_j<LANES>(J, jx0, jy0, lane, sc.jp);
                Red::sync();
                lox = imax(jx0, -SVO_LK_WIN);
                loy = imax(jy0, -SVO_LK_WIN);
                rx = imin(jx0 + SVO_LK_JC - SVO_LK_WIN, J.w - 1) - lox;
                ry = imin(jy0 + SVO_LK_JR - SVO_LK_WIN - 1, J.h - 1) - loy;
            }
            ++iters;
#ifdef SVO_LK_TRACE
            printf("L%d j%d c=(%.4f,%.4f) in=(%d,%d) j0=(%d,%d)\n", level, j, cx, cy, inx, iny, jx0, jy0);
#endif
            lk_weights(fsub(cx, (float)inx), fsub(cy, (float)iny), iw00, iw01, iw10, iw11);
            const unsigned jt = pack_weights(iw00, iw01), jb = pack_weights(iw10, iw11);
            const unsigned short* jbase = sc.jp + (iny - jy0) * SVO_LK_JC + (inx - jx0);
            // mismatch vector.  OpenCV sums it in float32 in a fixed order; while the absolute products of each of the
            // three accumulator groups stay below 2^24 every partial sum inside a group is an exactly representable
            // integer and the result is the fold of the three integer group sums (Red::sum_b_grouped).  Otherwise the
            // sums are formed term by term in OpenCV's order (Red::cv_sum_b).
            int p1[SLOTS], p2[SLOTS];
            typename LkAcc<LANES>::type sb1 = 0, sb2 = 0, sa1 = 0, sa2 = 0;   // <= 4 pixels per lane fit 32 bits
#pragma unroll
            for (int s = 0; s < SLOTS; ++s) {
                const unsigned short* q = jbase + offJ[s];
                const int v = dp2a_lo_su(jb, q[SVO_LK_JC], dp2a_lo_su(jt, q[0], 1 << 8));
                const int diff = (v >> 9) - Iv[s];
                p1[s] = diff * Ixv[s];
                p2[s] = diff * Iyv[s];
                sb1 += p1[s];
                sb2 += p2[s];
                sa1 += p1[s] < 0 ? -p1[s] : p1[s];
                sa2 += p2[s] < 0 ? -p2[s] : p2[s];
            }
            float b1, b2;
            if (!Red::sum_b_grouped(p1, p2, lane, sb1, sb2, sa1, sa2, b1, b2)) Red::cv_sum_b(p1, p2, lane, sc.stage, b1, b2);
            b1 = fmul(b1, FLT_SCALE);
            b2 = fmul(b2, FLT_SCALE);
            const float ddx = fmul(fsub(fmul(A12, b2), fmul(A22, b1)), D);
            const float ddy = fmul(fsub(fmul(A12, b1), fmul(A11, b2)), D);
#ifdef SVO_LK_TRACE
            printf("   b=(%g,%g) d=(%g,%g) cy=%.9g iny=%d\n", b1, b2, ddx, ddy, cy, iny);
#endif
            cx = fadd(cx, ddx);
            cy = fadd(cy, ddy);
            nx = fadd(cx, half_win);
            ny = fadd(cy, half_win);
            const float d2 = fadd(fmul(ddx, ddx), fmul(ddy, ddy));
            if (d2 < eps_lo) break;
            if (!(d2 > eps_hi) && (double)ddx * (double)ddx + (double)ddy * (double)ddy <= cfg.eps_sq) break;
            // std::abs(float) < 0.01 (a double): 0.01f is the largest float32 below 0.01, so `<= 0.01f` is the same test
            if (j > 0 && fabsf(fadd(ddx, pdx)) <= 0.01f && fabsf(fadd(ddy, pdy)) <= 0.01f) {
                nx = fsub(nx, fmul(ddx, 0.5f));
                ny = fsub(ny, fmul(ddy, 0.5f));
                break;
            }
            pdx = ddx;
            pdy = ddy;
        }
    }
    // err is requested by the reference (vo.cpp:203-215) without
    // OPTFLOW_LK_GET_MIN_EIGENVALS: OpenCV re-checks the final window position.
    if (status) {
        const int fx = floor_int(fsub(nx, half_win)), fy = floor_int(fsub(ny, half_win));
        const LkLevel J0 = Jpyr.level(0);
        if (fx < -SVO_LK_WIN || fx >= J0.w || fy < -SVO_LK_WIN || fy >= J0.h) status = 0;
    }
    outx = nx;
    outy = ny;
    iters_out = iters;
    return status;
}

// ---------------------------------------------------------------------------
// Multi-channel variant (interleaved 8UCn images, n <= 3): what cv::calcOpticalFlowPyrLK does when the
// reference's CLI feeds it the BGR images cv::imread returns (src/main.cpp:37-46).  Same arithmetic
// as lk_track_feature with the window widened to 10 x 10n interleaved samples (horizontal neighbour = +n bytes,
// Scharr per channel, the 2 x 2 system summed over all samples, minEig still divided by 2 * 100).  This path
// serves the CLI-faithful mode only and is written for clarity: samples are distributed round-robin over the
// lanes, the next image is read through the cache instead of a staged patch.
#define SVO_LK_CN_MAX 3
#define SVO_LK_CN_IP (SVO_LK_PATCH * SVO_LK_CN_MAX + 1)   // template neighbourhood pitch (bytes)
#define SVO_LK_CN_DP (SVO_LK_DW * SVO_LK_CN_MAX + 1)      // derivative patch pitch (samples)
#define SVO_LK_CN_N (SVO_LK_NPIX * SVO_LK_CN_MAX)         // samples of the window
struct LkScratchCn {
    uint8_t ipatch[SVO_LK_PATCH * SVO_LK_CN_IP];
    short dxp[SVO_LK_DW * SVO_LK_CN_DP];
    short dyp[SVO_LK_DW * SVO_LK_CN_DP];
    short iv[SVO_LK_CN_N], ixv[SVO_LK_CN_N], iyv[SVO_LK_CN_N];
};

template <int LANES, class Red>
SVO_HD int lk_track_feature_cn(const LkPyramid& Ipyr, const LkPyramid& Jpyr, const LkConfig& cfg, int cn,
                               float ptx, float pty, int lane, LkScratchCn& sc, float& outx, float& outy,
                               int& iters_out) {
    const float half_win = 4.5f;
    const float FLT_SCALE = 1.f / (1 << 20);
    const int wn = SVO_LK_WIN * cn;          // samples per window row
    const int nsamp = SVO_LK_WIN * wn;
    int status = 1;
    float nx = 0.f, ny = 0.f;
    int iters = 0;
    for (int level = cfg.nlevels - 1; level >= 0; --level) {
        const LkLevel I = Ipyr.lv[level];
        const LkLevel J = Jpyr.lv[level];
        const float scale = 1.f / (float)(1 << level);
        float prx = fmul(ptx, scale), pry = fmul(pty, scale);
        if (level == cfg.nlevels - 1) {
            nx = prx;
            ny = pry;
        } else {
            nx = fmul(nx, 2.f);
            ny = fmul(ny, 2.f);
        }
        prx = fsub(prx, half_win);
        pry = fsub(pry, half_win);
        const int ipx = floor_int(prx), ipy = floor_int(pry);
        if (ipx < -SVO_LK_WIN || ipx >= I.w || ipy < -SVO_LK_WIN || ipy >= I.h) {
            if (level == 0) status = 0;
            continue;
        }
        int iw00, iw01, iw10, iw11;
        lk_weights(fsub(prx, (float)ipx), fsub(pry, (float)ipy), iw00, iw01, iw10, iw11);

        // 13 x 13 pixel neighbourhood of the template window, BORDER_REFLECT_101 per pixel
        Red::sync();
        const int pn = SVO_LK_PATCH * cn;
        for (int idx = lane; idx < SVO_LK_PATCH * pn; idx += LANES) {
            const int py = idx / pn, pb = idx - py * pn;
            const int px = pb / cn, ch = pb - px * cn;
            sc.ipatch[py * SVO_LK_CN_IP + pb] =
                I.img[(size_t)reflect101(ipy - 1 + py, I.h) * I.pitch + reflect101(ipx - 1 + px, I.w) * cn + ch];
        }
        Red::sync();
        // Scharr per channel on the 11 x 11 pixel positions (zero outside the level)
        const int dn = SVO_LK_DW * cn;
        for (int idx = lane; idx < SVO_LK_DW * dn; idx += LANES) {
            const int y = idx / dn, xb = idx - y * dn;
            const int px = xb / cn;
            int dx = 0, dy = 0;
            if (ipx + px >= 0 && ipx + px < I.w && ipy + y >= 0 && ipy + y < I.h) {
                const uint8_t* p = sc.ipatch + y * SVO_LK_CN_IP + xb;
                const int a00 = p[0], a01 = p[cn], a02 = p[2 * cn];
                const int a10 = p[SVO_LK_CN_IP], a12 = p[SVO_LK_CN_IP + 2 * cn];
                const int a20 = p[2 * SVO_LK_CN_IP], a21 = p[2 * SVO_LK_CN_IP + cn], a22 = p[2 * SVO_LK_CN_IP + 2 * cn];
                dx = 3 * (a02 + a22 - a00 - a20) + 10 * (a12 - a10);
                dy = 3 * (a20 + a22 - a00 - a02) + 10 * (a21 - a01);
            }
            sc.dxp[y * SVO_LK_CN_DP + xb] = (short)dx;
            sc.dyp[y * SVO_LK_CN_DP + xb] = (short)dy;
        }
        Red::sync();
        long long sA11 = 0, sA12 = 0, sA22 = 0;
        for (int e = lane; e < nsamp; e += LANES) {
            const int wy = e / wn, xb = e - wy * wn;
            const uint8_t* p = sc.ipatch + (wy + 1) * SVO_LK_CN_IP + xb + cn;
            const int iv = (p[0] * iw00 + p[cn] * iw01 + p[SVO_LK_CN_IP] * iw10 + p[SVO_LK_CN_IP + cn] * iw11 + (1 << 8)) >> 9;
            const short* qx = sc.dxp + wy * SVO_LK_CN_DP + xb;
            const short* qy = sc.dyp + wy * SVO_LK_CN_DP + xb;
            const int ix = ((int)qx[0] * iw00 + (int)qx[cn] * iw01 + (int)qx[SVO_LK_CN_DP] * iw10 + (int)qx[SVO_LK_CN_DP + cn] * iw11 + (1 << 13)) >> 14;
            const int iy = ((int)qy[0] * iw00 + (int)qy[cn] * iw01 + (int)qy[SVO_LK_CN_DP] * iw10 + (int)qy[SVO_LK_CN_DP + cn] * iw11 + (1 << 13)) >> 14;
            sc.iv[e] = (short)iv;
            sc.ixv[e] = (short)ix;
            sc.iyv[e] = (short)iy;
            const int sx = (short)ix, sy = (short)iy;
            sA11 += sx * sx; sA12 += sx * sy; sA22 += sy * sy;
        }
        sA11 = Red::sum64g(sA11); sA12 = Red::sum64g(sA12); sA22 = Red::sum64g(sA22);
        Red::sync();
        const float A11 = fmul((float)sA11, FLT_SCALE);
        const float A12 = fmul((float)sA12, FLT_SCALE);
        const float A22 = fmul((float)sA22, FLT_SCALE);
        float D = fsub(fmul(A11, A22), fmul(A12, A12));
        const float dA = fsub(A11, A22);
        const float disc = fadd(fmul(dA, dA), fmul(fmul(4.f, A12), A12));
        const float min_eig = fsub(fadd(A22, A11), sqrtf(disc)) / (float)(2 * SVO_LK_NPIX);
        if (min_eig < cfg.min_eig || D < FLT_EPSILON) {
            if (level == 0) status = 0;
            continue;
        }
        D = 1.f / D;
        float cx = fsub(nx, half_win), cy = fsub(ny, half_win);
        float pdx = 0.f, pdy = 0.f;
        for (int j = 0; j < cfg.max_iters; ++j) {
            const int inx = floor_int(cx), iny = floor_int(cy);
            if (inx < -SVO_LK_WIN || inx >= J.w || iny < -SVO_LK_WIN || iny >= J.h) {
                if (level == 0) status = 0;
                break;
            }
            ++iters;
            lk_weights(fsub(cx, (float)inx), fsub(cy, (float)iny), iw00, iw01, iw10, iw11);
            const bool interior = inx >= 0 && iny >= 0 && inx + SVO_LK_WIN <= J.w - 1 && iny + SVO_LK_WIN <= J.h - 1;
            long long sb1 = 0, sb2 = 0;
            for (int e = lane; e < nsamp; e += LANES) {
                const int wy = e / wn, xb = e - wy * wn;
                int j00, j01, j10, j11;
                if (interior) {
                    const uint8_t* q = J.img + (size_t)(iny + wy) * J.pitch + inx * cn + xb;
                    j00 = q[0]; j01 = q[cn]; j10 = q[J.pitch]; j11 = q[J.pitch + cn];
                } else {
                    const int px = xb / cn, ch = xb - px * cn;
                    const uint8_t* r0 = J.img + (size_t)reflect101(iny + wy, J.h) * J.pitch + ch;
                    const uint8_t* r1 = J.img + (size_t)reflect101(iny + wy + 1, J.h) * J.pitch + ch;
                    const int x0 = reflect101(inx + px, J.w) * cn, x1 = reflect101(inx + px + 1, J.w) * cn;
                    j00 = r0[x0]; j01 = r0[x1]; j10 = r1[x0]; j11 = r1[x1];
                }
                const int v = (j00 * iw00 + j01 * iw01 + j10 * iw10 + j11 * iw11 + (1 << 8)) >> 9;
                const int diff = v - (int)sc.iv[e];
                sb1 += diff * (int)sc.ixv[e];
                sb2 += diff * (int)sc.iyv[e];
            }
            const float b1 = fmul((float)Red::sum64g(sb1), FLT_SCALE);
            const float b2 = fmul((float)Red::sum64g(sb2), FLT_SCALE);
            const float ddx = fmul(fsub(fmul(A12, b2), fmul(A22, b1)), D);
            const float ddy = fmul(fsub(fmul(A12, b1), fmul(A11, b2)), D);
            cx = fadd(cx, ddx);
            cy = fadd(cy, ddy);
            nx = fadd(cx, half_win);
            ny = fadd(cy, half_win);
            if ((double)ddx * (double)ddx + (double)ddy * (double)ddy <= cfg.eps_sq) break;
            if (j > 0 && fabsf(fadd(ddx, pdx)) <= 0.01f && fabsf(fadd(ddy, pdy)) <= 0.01f) {
                nx = fsub(nx, fmul(ddx, 0.5f));
                ny = fsub(ny, fmul(ddy, 0.5f));
                break;
            }
            pdx = ddx;
            pdy = ddy;
        }
    }
    if (status) {
        const int fx = floor_int(fsub(nx, half_win)), fy = floor_int(fsub(ny, half_win));
        const LkLevel J0 = Jpyr.lv[0];
        if (fx < -SVO_LK_WIN || fx >= J0.w || fy < -SVO_LK_WIN || fy >= J0.h) status = 0;
    }
    outx = nx;
    outy = ny;
    iters_out = iters;
    return status;
}

// ---------------------------------------------------------------------------
// findClosePoints, vo.cpp:265-280 (NaN => kept, like `offset > threshold`).
SVO_HD bool close_points(float x1, float y1, float x2, float y2, float thr) {
    const float off = fmaxf(fabsf(x1 - x2), fabsf(y1 - y2));
    return !(off > thr);
}
// in-bounds test of matchingFeatures, vo.cpp:341-359
SVO_HD bool in_bounds(float x, float y, int w, int h) {
    return !((x < 0) || (y < 0) || (y >= (float)h) || (x >= (float)w));
}

// ---------------------------------------------------------------------------
// DLT triangulation (cv::triangulatePoints + convertPointsFromHomogeneous as at
// vo.cpp:89-94), SURVEY B.4.  FP64; the right singular vector of the smallest singular
// value is stored as float32 and dehomogenised in float32.
SVO_HD void triangulate_dlt(const float* Pl, const float* Pr, float xl, float yl, float xr, float yr,
                            float* X3) {
    // Right singular vector of the smallest singular value of the 4x4 DLT system A == eigenvector of
    // the smallest eigenvalue of G = A^T A.  Inverse iteration on G + eps*I (same eigenvectors, safely
    // positive definite) with one LDL^T factorisation: ~10x cheaper than a Jacobi SVD per point and
    // accurate far below float32 resolution for any point with a usable depth.
    double A[4][4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        A[0][k] = (double)xl * (double)Pl[8 + k] - (double)Pl[k];
        A[1][k] = (double)yl * (double)Pl[8 + k] - (double)Pl[4 + k];
        A[2][k] = (double)xr * (double)Pr[8 + k] - (double)Pr[k];
        A[3][k] = (double)yr * (double)Pr[8 + k] - (double)Pr[4 + k];
    }
    double G[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            G[i][j] = 0.0;
            if (j >= i) {
#pragma unroll
                for (int r = 0; r < 4; ++r) G[i][j] += A[r][i] * A[r][j];
            }
        }
    const double tr = G[0][0] + G[1][1] + G[2][2] + G[3][3];
    const double eps = 1e-13 * tr + 1e-300;
#pragma unroll
    for (int i = 0; i < 4; ++i) G[i][i] += eps;
    // LDL^T (upper part of G holds the input): L unit lower in Lm, D in Dv
    double Lm[4][4], Dinv[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        double dj = G[j][j];
#pragma unroll
        for (int k = 0; k < 4; ++k)
            if (k < j) dj -= Lm[j][k] * Lm[j][k] / Dinv[k];
        if (!(dj > 1e-300)) dj = 1e-300;
        Dinv[j] = 1.0 / dj;
#pragma unroll
        for (int i = 0; i < 4; ++i)
            if (i > j) {
                double v = G[j][i];
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    if (k < j) v -= Lm[i][k] * Lm[j][k] / Dinv[k];
                Lm[i][j] = v * Dinv[j];
            }
    }
    double x[4] = {0.3, 0.2, 0.7, 0.6};
    for (int it = 0; it < 12; ++it) {
        double y[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {   // L z = x
            double v = x[i];
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (k < i) v -= Lm[i][k] * y[k];
            y[i] = v;
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) y[i] *= Dinv[i];
#pragma unroll
        for (int i = 3; i >= 0; --i) {   // L^T y = z
            double v = y[i];
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (k > i) v -= Lm[k][i] * y[k];
            y[i] = v;
        }
        const double inv = 1.0 / sqrt(y[0] * y[0] + y[1] * y[1] + y[2] * y[2] + y[3] * y[3]);
        double diff = 0.0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const double v = y[i] * inv;
            diff += fabs(v - x[i]);
            x[i] = v;
        }
        if (diff < 1e-15) break;
    }
    float h[4];
#pragma unroll
    for (int r = 0; r < 4; ++r) h[r] = (float)x[r];
    const float sc = (h[3] != 0.f) ? 1.f / h[3] : 1.f;
    X3[0] = h[0] * sc;
    X3[1] = h[1] * sc;
    X3[2] = h[2] * sc;
}

// ---------------------------------------------------------------------------
// PnP (cv::solvePnPRansac with SOLVEPNP_ITERATIVE as at vo.cpp:287-304),
// SURVEY B.5.  Pose = (R row-major 3x3, t) in FP64.
struct Pose {
    double R[9];
    double t[3];
};
struct Intrinsics {
    double fx, fy, cx, cy;
};

// exp([w]x) * R  (Rodrigues formula)
SVO_HD void rotate_left(const double* w, const double* R, double* out) {
    const double th2 = w[0] * w[0] + w[1] * w[1] + w[2] * w[2];
    const double th = sqrt(th2);
    double a, b;   // exp = I + a [w]x + b [w]x^2
    if (th < 1e-8) {
        a = 1.0 - th2 / 6.0;
        b = 0.5 - th2 / 24.0;
    } else {
        a = sin(th) / th;
        b = (1.0 - cos(th)) / th2;
    }
    double E[9];
    const double wx = w[0], wy = w[1], wz = w[2];
    E[0] = 1.0 - b * (wy * wy + wz * wz); E[1] = -a * wz + b * wx * wy;        E[2] = a * wy + b * wx * wz;
    E[3] = a * wz + b * wx * wy;          E[4] = 1.0 - b * (wx * wx + wz * wz); E[5] = -a * wx + b * wy * wz;
    E[6] = -a * wy + b * wx * wz;         E[7] = a * wx + b * wy * wz;         E[8] = 1.0 - b * (wx * wx + wy * wy);
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j)
            out[i * 3 + j] = E[i * 3] * R[j] + E[i * 3 + 1] * R[3 + j] + E[i * 3 + 2] * R[6 + j];
}

// rotation angle of R = || cv::Rodrigues(R) ||
SVO_HD double rotation_angle(const double* R) {
    double c = (R[0] + R[4] + R[8] - 1.0) * 0.5;
    c = c > 1.0 ? 1.0 : (c < -1.0 ? -1.0 : c);
    const double sx = R[7] - R[5], sy = R[2] - R[6], sz = R[3] - R[1];
    const double s = 0.5 * sqrt(sx * sx + sy * sy + sz * sz);
    return atan2(s, c);
}

// 1/z to ~4e-15 relative: float32 reciprocal refined by one Newton step in FP64 (the projection is
// rounded to float32 afterwards, so this is far inside the noise; a full FP64 division costs ~10x more).
SVO_HD double fast_rcp(double z) {
#if defined(__CUDA_ARCH__)
    if (!(fabs(z) > 1e-30 && fabs(z) < 1e30)) return 1.0 / z;
    const double r = (double)__frcp_rn((float)z);
    const double r1 = r * (2.0 - z * r);
    return r1 * (2.0 - z * r1);
#else
    return 1.0 / z;
#endif
}

// cv::projectPoints (no distortion), FP64.
SVO_HD void project_point(const Pose& P, const Intrinsics& K, double X, double Y, double Z, double& u,
                          double& v) {
    const double x = P.R[0] * X + P.R[1] * Y + P.R[2] * Z + P.t[0];
    const double y = P.R[3] * X + P.R[4] * Y + P.R[5] * Z + P.t[1];
    const double z = P.R[6] * X + P.R[7] * Y + P.R[8] * Z + P.t[2];
    const double iz = z != 0.0 ? fast_rcp(z) : 1.0;
    u = K.fx * (x * iz) + K.cx;
    v = K.fy * (y * iz) + K.cy;
}

// Squared reprojection error exactly as PnPRansacCallback::computeError sees it:
// projection rounded to float32, difference in float32, squares summed.
SVO_HD float reproj_err_sq(const Pose& P, const Intrinsics& K, const float* X3, const float* uv) {
    double u, v;
    project_point(P, K, (double)X3[0], (double)X3[1], (double)X3[2], u, v);
    const float du = uv[0] - (float)u, dv = uv[1] - (float)v;
    return (float)((double)du * (double)du + (double)dv * (double)dv);
}

// Normal-equation accumulator for Gauss-Newton / LM on (dw, dt), left-multiplied
// rotation increment.  acc[0..20] = upper triangle of J^T J (row-major), acc[21..26]
// = J^T e, acc[27] = e^T e.
#define SVO_PNP_ACC 28
SVO_HD void pnp_accumulate(const Pose& P, const Intrinsics& K, const float* X3, const float* uv,
                           double* acc, bool with_jacobian) {
    const double X = X3[0], Y = X3[1], Z = X3[2];
    const double rx = P.R[0] * X + P.R[1] * Y + P.R[2] * Z;
    const double ry = P.R[3] * X + P.R[4] * Y + P.R[5] * Z;
    const double rz = P.R[6] * X + P.R[7] * Y + P.R[8] * Z;
    const double x = rx + P.t[0], y = ry + P.t[1], z = rz + P.t[2];
    const double iz = z != 0.0 ? 1.0 / z : 1.0;
    const double eu = K.fx * (x * iz) + K.cx - (double)uv[0];
    const double ev = K.fy * (y * iz) + K.cy - (double)uv[1];
    acc[27] += eu * eu + ev * ev;
    if (!with_jacobian) return;
    // d(u,v)/d(xc)
    const double a0 = K.fx * iz, a2 = -K.fx * x * iz * iz;
    const double b1 = K.fy * iz, b2 = -K.fy * y * iz * iz;
    // d(xc)/d(w) = -[r]x ; d(xc)/d(t) = I
    double Ju[6], Jv[6];
    Ju[0] = a2 * ry;              // a . (-[r]x col 0) = a . (0, -rz... ) see below
    Ju[1] = a0 * rz - a2 * rx;
    Ju[2] = -a0 * ry;
    Jv[0] = -b1 * rz + b2 * ry;
    Jv[1] = -b2 * rx;
    Jv[2] = b1 * rx;
    Ju[3] = a0; Ju[4] = 0.0; Ju[5] = a2;
    Jv[3] = 0.0; Jv[4] = b1; Jv[5] = b2;
    int k = 0;
#pragma unroll
    for (int i = 0; i < 6; ++i) {
#pragma unroll
        for (int j = i; j < 6; ++j) acc[k++] += Ju[i] * Ju[j] + Jv[i] * Jv[j];
    }
#pragma unroll
    for (int i = 0; i < 6; ++i) acc[21 + i] += Ju[i] * eu + Jv[i] * ev;
}

// Solve (JtJ with diag * (1 + lambda)) d = Jte by Cholesky.  Returns false when
// the damped matrix is not positive definite.
SVO_HD bool pnp_solve_step(const double* acc, double lambda, double* d) {
    double M[6][6];
    int k = 0;
#pragma unroll
    for (int i = 0; i < 6; ++i)
#pragma unroll
        for (int j = i; j < 6; ++j) {
            M[i][j] = acc[k];
            M[j][i] = acc[k];
            ++k;
        }
#pragma unroll
    for (int i = 0; i < 6; ++i) M[i][i] *= (1.0 + lambda);
    // Cholesky M = L L^T (in place, lower)
#pragma unroll
    for (int j = 0; j < 6; ++j) {
        double s = M[j][j];
#pragma unroll
        for (int q = 0; q < 6; ++q)
            if (q < j) s -= M[j][q] * M[j][q];
        if (!(s > 0.0)) return false;
        const double l = sqrt(s);
        M[j][j] = l;
#pragma unroll
        for (int i = 0; i < 6; ++i)
            if (i > j) {
                double v = M[i][j];
#pragma unroll
                for (int q = 0; q < 6; ++q)
                    if (q < j) v -= M[i][q] * M[j][q];
                M[i][j] = v / l;
            }
    }
    double yv[6];
#pragma unroll
    for (int i = 0; i < 6; ++i) {
        double v = acc[21 + i];
#pragma unroll
        for (int q = 0; q < 6; ++q)
            if (q < i) v -= M[i][q] * yv[q];
        yv[i] = v / M[i][i];
    }
#pragma unroll
    for (int i = 5; i >= 0; --i) {
        double v = yv[i];
#pragma unroll
        for (int q = 0; q < 6; ++q)
            if (q > i) v -= M[q][i] * d[q];
        d[i] = v / M[i][i];
    }
    return true;
}

// param = prevParam - delta (CvLevMarq::step): R' = exp(-dw) R, t' = t - dt
SVO_HD void pnp_apply_step(const Pose& P, const double* d, Pose& out) {
    const double w[3] = {-d[0], -d[1], -d[2]};
    rotate_left(w, P.R, out.R);
    out.t[0] = P.t[0] - d[3];
    out.t[1] = P.t[1] - d[4];
    out.t[2] = P.t[2] - d[5];
}

// Levenberg-Marquardt control flow of CvLevMarq (OpenCV calib3d compat_ptsetreg.cpp)
// as used by cvFindExtrinsicCameraParams2: lambda = 10^lg, lg starts at -3, a
// worse step raises lg (up to 16) and retries, a better one lowers it; stops after
// 20 accepted steps or when the relative parameter change drops below FLT_EPSILON.
// `Eval` computes the 28 accumulators for a pose: eval(pose, acc, with_jacobian).
template <class Eval>
SVO_HD void pnp_refine_lm(Pose& P, Eval& eval, int max_iter) {
    double acc[SVO_PNP_ACC], cand_acc[SVO_PNP_ACC];
    eval(P, acc, true);
    double prev_err = acc[27];
    int lg = -3;
    int iters = 0;
    for (int guard = 0; guard < 64; ++guard) {
        double d[6];
        Pose cand;
        bool solved = pnp_solve_step(acc, pow(10.0, (double)lg), d);
        if (solved) {
            pnp_apply_step(P, d, cand);
            eval(cand, cand_acc, true);
        }
        const double ang = rotation_angle(P.R);
        const double pn = sqrt(ang * ang + P.t[0] * P.t[0] + P.t[1] * P.t[1] + P.t[2] * P.t[2]);
        const double dn = solved ? sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2] + d[3] * d[3] + d[4] * d[4] + d[5] * d[5]) : 0.0;
        if (!solved || !(cand_acc[27] <= prev_err)) {
            // CvLevMarq would now raise lambda up to 16 times, re-evaluating the error each time, until rounding
            // noise lets a (still smaller) step pass and the relative-change test ends the loop.  When the rejected
            // step is already below that termination size the outcome is the current pose to ~1e-12: stop here.
            if (solved && dn < (double)FLT_EPSILON * pn) return;
            if (++lg <= 16) continue;   // retry with more damping from the same linearisation
            if (!solved) return;
        }
        lg = lg - 1 < -16 ? -16 : lg - 1;
        // accepted
        P = cand;
        ++iters;
        if (iters >= max_iter || dn < (double)FLT_EPSILON * pn) return;
        prev_err = cand_acc[27];
#pragma unroll
        for (int i = 0; i < SVO_PNP_ACC; ++i) acc[i] = cand_acc[i];
    }
}

// ---- RANSAC control flow of cv::solvePnPRansac (RANSACPointSetRegistrator::run) -------
// OpenCV seeds a local RNG with (uint64)-1 on every call, so the sample sequence depends
// only on the number of points: it can be replayed exactly.  cv::RNG is a multiply-with-
// carry generator; uniform(0, n) = next() % n.
#define SVO_RNG_TABLE 1024
SVO_HD unsigned cv_rng_next(uint64_t& state) {
    state = (uint64_t)(unsigned)state * 4164903690u + (unsigned)(state >> 32);
    return (unsigned)state;
}
// getSubset(): 5 distinct indices per hypothesis, redrawing on duplicates.  mod_at(pos) is
// the pos-th RNG output modulo the point count.  A hypothesis that would run past the table
// is marked invalid (-1); unreachable in practice (needs > 1024 draws for <= 100 samples).
template <class ModAt>
SVO_HD void ransac_build_samples(ModAt mod_at, int n_hyp, int* samples) {
    int pos = 0;
    for (int h = 0; h < n_hyp; ++h) {
        int idx[5] = {-1, -1, -1, -1, -1};
        bool ok = true;
        for (int i = 0; i < 5 && ok; ++i) {
            for (;;) {
                if (pos >= SVO_RNG_TABLE) { ok = false; break; }
                const int v = mod_at(pos++);
                bool dup = false;
                for (int j = 0; j < i; ++j) dup |= (idx[j] == v);
                if (!dup) { idx[i] = v; break; }
            }
        }
        for (int i = 0; i < 5; ++i) samples[h * 5 + i] = ok ? idx[i] : -1;
    }
}
// cv::RANSACUpdateNumIters
SVO_HD int ransac_update_num_iters(double p, double ep, int model_points, int max_iters) {
    p = p < 0. ? 0. : (p > 1. ? 1. : p);
    ep = ep < 0. ? 0. : (ep > 1. ? 1. : ep);
    double num = 1. - p;
    if (num < DBL_MIN) num = DBL_MIN;
    double denom = 1. - pow(1. - ep, (double)model_points);
    if (denom < DBL_MIN) return 0;
    num = log(num);
    denom = log(denom);
    if (denom >= 0 || -num >= max_iters * (-denom)) return max_iters;
#if defined(__CUDA_ARCH__)
    return __double2int_rn(num / denom);
#else
    return (int)lrint(num / denom);
#endif
}
// The iteration budget of the sequential loop after it has looked at the first n_have counts (it stops earlier
// when the budget drops below n_have): > n_have means the loop goes on to hypothesis n_have.
SVO_HD int ransac_replay_budget(const int* counts, int n_have, int n_points, int max_iters, double confidence,
                                int& best_count) {
    int niters = max_iters < 1 ? 1 : max_iters;
    int max_good = 0;
    for (int k = 0; k < niters && k < n_have; ++k) {
        const int c = counts[k];
        if (c > (max_good > 4 ? max_good : 4)) {
            max_good = c;
            niters = ransac_update_num_iters(confidence, (double)(n_points - c) / n_points, 5, niters);
        }
    }
    best_count = max_good;
    return niters;
}
// Replays the sequential loop over per-hypothesis inlier counts: a model replaces the best
// iff count > max(best, 4); the iteration budget shrinks adaptively.  Returns the index of
// the winning hypothesis or -1.
SVO_HD int ransac_replay(const int* counts, int n_hyp, int n_points, int max_iters, double confidence,
                         int& best_count) {
    int niters = max_iters < 1 ? 1 : max_iters;
    int best = -1, max_good = 0;
    for (int k = 0; k < niters && k < n_hyp; ++k) {
        const int c = counts[k];
        if (c > (max_good > 4 ? max_good : 4)) {
            max_good = c;
            best = k;
            niters = ransac_update_num_iters(confidence, (double)(n_points - c) / n_points, 5, niters);
        }
    }
    best_count = max_good;
    return best;
}

}  // namespace svo
