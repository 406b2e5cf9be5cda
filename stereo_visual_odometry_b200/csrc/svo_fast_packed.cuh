// svo_fast_packed.cuh -- FAST-9/16 arithmetic on pixel PAIRS packed as int16x2 (the inner loops of fast_kernel).
//
// Everything works on the raw ring pixels r_k (0..255 per half), not on differences: with d_k = c - r_k
//   max_k min_arc d = c - min_k max_arc r      and      min_k max_arc d = c - max_k min_arc r,
// so the 16 subtractions and the bias that keeps packed differences positive disappear, and the 16 arcs of 9 are
// folded in pairs (arcs k and k + 1 share r[k+1 .. k+8]; OpenCV's cornerScore uses the same pairing):
//   min(max9(arc k), max9(arc k+1)) = max(max8(r[k+1 .. k+8]), min(r[k], r[k+9])).
// 75 packed min / max / add operations per pixel pair instead of 146.  Reference semantics: cv::FAST TYPE_9_16 as
// called by featureDetectionFast (feature_set.cpp:55-68); scalar statement: fast9_arc_score (svo_core.cuh).
#pragma once
#include "svo_core.cuh"

namespace svo {

// ---- packed int16x2 primitives (values 0 .. 0x7FFF per half); the host versions serve tests/host_cores.cpp
SVO_HD unsigned pk_min(unsigned a, unsigned b) {
#if defined(__CUDA_ARCH__)
    return __vmins2(a, b);
#else
    const unsigned al = a & 0xFFFFu, bl = b & 0xFFFFu, ah = a >> 16, bh = b >> 16;
    return (al < bl ? al : bl) | ((ah < bh ? ah : bh) << 16);
#endif
}
SVO_HD unsigned pk_max(unsigned a, unsigned b) {
#if defined(__CUDA_ARCH__)
    return __vmaxs2(a, b);
#else
    const unsigned al = a & 0xFFFFu, bl = b & 0xFFFFu, ah = a >> 16, bh = b >> 16;
    return (al > bl ? al : bl) | ((ah > bh ? ah : bh) << 16);
#endif
}
SVO_HD unsigned pk_min3(unsigned a, unsigned b, unsigned c) {
#if defined(__CUDA_ARCH__)
    return __vimin3_s16x2(a, b, c);
#else
    return pk_min(pk_min(a, b), c);
#endif
}
SVO_HD unsigned pk_max3(unsigned a, unsigned b, unsigned c) {
#if defined(__CUDA_ARCH__)
    return __vimax3_s16x2(a, b, c);
#else
    return pk_max(pk_max(a, b), c);
#endif
}

// bytes i, i + 1 of the 12-byte window w[0..2] as (b_i) | (b_{i+1} << 16)
SVO_HD unsigned pk_pick2(const unsigned* w, int i) {
    const int wi = i >> 2, bi = i & 3;
    if (bi < 3) return byte_perm(w[wi], 0u, (unsigned)(bi | (4 << 4) | ((bi + 1) << 8) | (4 << 12)));
    return byte_perm(w[wi], w[wi + 1], 0x4433u) & 0x00FF00FFu;
}

// ---- compass test of one pixel pair (centre bytes C0, C0 + 1 of the row windows): can a pixel hold a 9-arc whose
// pixels all differ from the centre by more than t?  Any 9-arc contains two ADJACENT compass pixels (k = 0, 4, 8, 12),
// i.e. one of {r0, r8} and one of {r4, r12}:
//   dark  arc possible  <=>  max(min(r0, r8), min(r4, r12)) < c - t
//   bright arc possible <=>  min(max(r0, r8), max(r4, r12)) > c + t
// kt = (0x8000 - t - 1) per half.  Bit 15 of a half of the result <=> that pixel passes.
template <int C0>
SVO_HD unsigned fast_pair_compass(const unsigned* top, const unsigned* mid, const unsigned* bot, unsigned kt) {
    const unsigned c = pk_pick2(mid, C0);
    const unsigned r0 = pk_pick2(bot, C0), r8 = pk_pick2(top, C0), r4 = pk_pick2(mid, C0 + 3), r12 = pk_pick2(mid, C0 - 3);
    const unsigned lo = pk_max(pk_min(r0, r8), pk_min(r4, r12));
    const unsigned hi = pk_min(pk_max(r0, r8), pk_max(r4, r12));
    // 0x8000 + (c - lo) - t - 1 and 0x8000 + (hi - c) - t - 1: both stay inside their half (no borrow, no carry)
    return ((c + kt) - lo) | ((hi + kt) - c);
}

// ---- exact arc scores of the pixel pair whose first pixel is byte C0 of the row windows; rows[r] = the 12-byte
// window of tile row (centre row - 3 + r).  Returns the two scores + 256 packed as int16x2.
template <int C0>
SVO_HD unsigned fast_pair_score_ring(const unsigned (&rows)[7][3]) {
    constexpr int RDX[16] = SVO_FAST_RING_DX;
    constexpr int RDY[16] = SVO_FAST_RING_DY;
    const unsigned c = pk_pick2(rows[3], C0);
    unsigned r[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) r[k] = pk_pick2(rows[3 + RDY[k]], C0 + RDX[k]);
    // windows of 2 and 4 ring pixels starting at the odd positions
    unsigned mx2[8], mn2[8], mx4[8], mn4[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        mx2[i] = pk_max(r[2 * i + 1], r[(2 * i + 2) & 15]);
        mn2[i] = pk_min(r[2 * i + 1], r[(2 * i + 2) & 15]);
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        mx4[i] = pk_max(mx2[i], mx2[(i + 1) & 7]);   // r[2i+1 .. 2i+4]
        mn4[i] = pk_min(mn2[i], mn2[(i + 1) & 7]);
    }
    // arcs k = 2i and 2i + 1 share r[2i+1 .. 2i+8] = mx4[i] u mx4[i+2]; their private pixels are r[2i] and r[2i+9]
    unsigned ta[8], tb[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const unsigned p = r[2 * i], q = r[(2 * i + 9) & 15];
        ta[i] = pk_max3(mx4[i], mx4[(i + 2) & 7], pk_min(p, q));   // min over the two arcs of their maximum
        tb[i] = pk_min3(mn4[i], mn4[(i + 2) & 7], pk_max(p, q));   // max over the two arcs of their minimum
    }
    const unsigned A = pk_min(pk_min3(pk_min3(pk_min3(ta[0], ta[1], ta[2]), ta[3], ta[4]), ta[5], ta[6]), ta[7]);
    const unsigned B = pk_max(pk_max3(pk_max3(pk_max3(tb[0], tb[1], tb[2]), tb[3], tb[4]), tb[5], tb[6]), tb[7]);
    // score = max(c - A, B - c); + 256 per half keeps both halves positive (no borrow across the halves)
    return pk_max((c + 0x01000100u) - A, (B + 0x01000100u) - c);
}

// ---- the four response bytes of a quad from its two packed scores (+ 256): byte e = score - 1 where the score is
// above the threshold, else 0 (cv::FAST keeps corners with score > threshold; KeyPoint response = score - 1).
// kz = (0x8000 - (thr + 256) - 1) per half: bit 15 of (s + kz) <=> score > thr; PRMT's sign-replicate mode turns the
// four bits into a byte mask.
SVO_HD unsigned fast_quad_pack(unsigned sa, unsigned sb, unsigned kz) {
    const unsigned za = sa + kz, zb = sb + kz;
    const unsigned oa = pk_max(sa, 0x01010101u) - 0x01010101u, ob = pk_max(sb, 0x01010101u) - 0x01010101u;   // max(score, 1) - 1
    return byte_perm(oa, ob, 0x6420u) & byte_perm_sx<0xFDB9u>(za, zb);
}

}  // namespace svo
