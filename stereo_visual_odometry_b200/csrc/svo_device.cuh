// svo_device.cuh -- device-side data layout shared by the kernels (svo_kernels.cu)
// and the C-ABI host layer (svo_capi.cu).
//
// HBM layout (one context = n_seq independent sequences, all arrays SoA per sequence):
//   pyramids  [seq][slot 0..2][cam L/R][level 0..3] u8, row pitch 128-B aligned; frame f lives in slot f % 3
//   fastkeys  [frame % 3][A|B][seq][bucket]         u64 per-bucket FAST winners of one left image
//   trackkeys [seq][bucket]                         u64 scratch for the tracked-feature winners
//   feature set / bucketed list / LK legs / matched set / 3-D points: [seq][cap]
//   hypotheses [seq][H] pose (12 f64) + inlier count
//   SeqState [seq], svo_frame_result [seq]
#pragma once
#include <stdint.h>

#include "../../include/svo.h"
#include "svo_core.cuh"

namespace svo {

// Ring of pyramid slots / FAST key buffers: frame f uses slot f % 3, so the ingest of frame f + 1 (its own
// stream) never touches what the tracking chain of frame f reads (slots f % 3 and (f - 1) % 3).
#define SVO_PYR_SLOTS 3

struct PyrGeom {
    int nlevels;
    int cn;                          // interleaved channels per pixel (1; 3 = CLI-faithful 8UC3 mode): w in pixels, pitch in bytes
    int w[SVO_MAX_LEVELS], h[SVO_MAX_LEVELS], pitch[SVO_MAX_LEVELS];
    long long off[SVO_MAX_LEVELS];   // byte offset of each level inside one image pyramid
    long long image_bytes;           // bytes of one image pyramid (all levels)
    int n_slots;                     // pyramid ring slots per sequence (3, or lookahead + 3: svo_submit_frames)
    // pyramid of (seq, slot, cam)
    __host__ __device__ long long base(int seq, int slot, int cam) const {
        return ((long long)(seq * n_slots + slot) * 2 + cam) * image_bytes;
    }
};

struct BucketGeom {
    int n_bh, n_bw;      // buckets along height / width
    int bh, bw;          // bucket size in pixels (ceil division, feature_set.cpp:103-104)
    int start_row;
    int age_threshold;
    int fast_threshold;  // the constant inside Bucket::compute_score
    int nb;              // n_bh * n_bw
};

struct SeqState {
    int spec_bad;        // 1: the speculative KLT results for the NEXT frame must not be used (this frame kept the
                         // old pyramids, vo.cpp:179-181, so they were tracked from the wrong images)
    int n_feat;          // currentVOFeatures.size()
    int n_bucketed;
    int n_matched;
    int n_inliers;
    int retry;
    int pnp_ok;
    int lk_iters;
    int frame_id;
    int fs_from_inliers; // 1: the feature set is the inlier list of the last frame (fs[i] == matched[inl_idx[i]] at its
                         // t1 position), i.e. the tracked-candidate cache is addressable; 0: it holds t0 positions
    double R[9];         // rotation      (vo.h:262)
    double t[3];         // translation   (vo.h:263)
    double last_T[16];   // last_transform (vo.h:264)
};

// What pnp_final_a hands to pnp_final_b (the LM refit runs off the critical chain, while the next frame's
// bucketing / KLT / hypotheses already overwrite the per-frame counters of SeqState).
struct FinalJob {
    int action;          // 0: refit + gates; otherwise an svo_fail_reason to report without a refit
    int n_in;            // inliers of the winning hypothesis
    int enough;          // n_in >= FEATURES_THRESHOLD
    int n_bucketed, n_matched, retry, lk_iters;
    int pad_;
    double R[9], t[3];   // winning hypothesis
};

// All device pointers of a context, passed by value to the kernels.
struct DevArrays {
    uint8_t* pyr;                     // pyramids
    unsigned long long* fastkeys;     // [3][2][n_seq][nb]
    int* fast_nA;                     // [3][n_seq] buckets with a thr_hi FAST winner (gates the thr_lo pass)
    unsigned long long* trackkeys;    // [n_seq][nb]
    int* tile_cnt;                    // [n_seq][16] winners per 1024-bucket tile: without | with the thr/4 FAST set << 16 (bucket_count_kernel)
    SeqState* state;                  // [n_seq]
    svo_frame_result* result;         // [n_seq]
    // tracked feature set (currentVOFeatures)
    float2* fs_xy; int* fs_age; int* fs_str;
    // bucketed list == pointsLeftT0 entering circular matching
    float2* bk_xy; int* bk_age; int* bk_str;
    // LK leg outputs + status bits (bit l = status of leg l)
    float2* lk_L1; float2* lk_R1; float2* lk_R0; float2* lk_C; uint8_t* lk_status;
    // matched (compacted) sets
    float2* m_L0; float2* m_L1; float2* m_R0; float2* m_R1; int* m_age; int* m_str;
    int* m_src;                       // [cap] bucketed-list index of each matched feature
    float* X3;                        // [cap][3] matched (compacted)
    float* X3u;                       // [cap][3] inlier 3-D points, compacted for the LM refit
    float2* in_uv;                    // [cap] inlier image points, compacted for the LM refit
    // RANSAC hypotheses
    double* hyp_pose;                 // [n_seq][H][12]
    int* hyp_count;                   // [n_seq][H]
    int* hyp_samples;                 // [n_seq][H][5] OpenCV-order sample indices (-1 = invalid)
    unsigned* rng_raw;                // [SVO_RNG_TABLE] cv::RNG((uint64)-1) output stream
    int* inl_idx;                     // [n_seq][cap] ascending inlier indices
    int* klt_next;                    // [n_seq] next todo index for the persistent KLT warps
    // speculative KLT cache (see svo_kernels.cu "speculative tracking"): circular-matching results of
    //   F: the FAST winner of every bucket of the previous left image   [frame & 1][n_seq][nb]
    //   T: every matched feature of the previous frame at its t1 position [n_seq][cap], indexed by its position in
    //      the previous frame's BUCKETED list (m_src maps matched -> bucketed)
    // entry = 4 leg outputs + status bits (bit 4 = keep) + LK iteration count
    float2* specF_pts; uint8_t* specF_st; int* specF_it;
    float2* specT_pts; uint8_t* specT_st; int* specT_it;
    int* cntF; int* cntT;             // [n_seq] work counters of the two speculative kernels
    int* fast_work;                   // [slots][2] tile counter + finished-block counter of the persistent FAST launch
    int* predT;                       // [2][n_seq][nb] best predicted tracked score + 0x8000 per bucket of the NEXT frame (0: none)
    int* listF; int* nF;              // [n_seq][nb], [n_seq]: buckets holding a thr-20 FAST winner (any order)
    int* todo; int* n_todo;           // [n_seq][cap], [n_seq]: bucketed features whose result was not cached
    FinalJob* job;                    // [n_seq]
    int nb;                           // buckets per sequence
    int cap;                          // per-sequence feature capacity
    int n_seq;
    int n_hyp;
};

struct PipeConfig {
    int width, height;
    int features_threshold;      // 15
    int pre_matching_threshold;  // 100
    int thr_hi, thr_lo;          // FAST_THRESHOLD, FAST_THRESHOLD / 4
    float circular_threshold;
    float reproj_thr_sq;         // (float)(8*8)
    double max_t_norm, max_r_norm;
    int ransac_iterations;       // RANSAC_ITERATIONS (vo.h) = OpenCV's iteration cap
    double ransac_confidence;    // (double)(float)0.98 (vo.cpp:295)
    int check_bounds;            // in-bounds test of matchingFeatures (vo.cpp:341-359)
    int age_increment;           // ages++ of stereo_callback (vo.cpp:70-72)
    LkConfig lk;
    Intrinsics K;
    float Pl[12], Pr[12];
};

}  // namespace svo
