// svo_kernels.cuh -- launch wrappers of the sm_100a kernels (definitions in svo_kernels.cu).
#pragma once
#include <cuda_runtime.h>

#include "svo_device.cuh"

namespace svo {

// ---- ingest: level-0 copy, FAST-9/16 + NMS (+ per-bucket winners), pyrDown ----
// FAST over `n_img` images (image i at img + i * img_stride).  Any of resp_map /
// keysA / keysB may be null.  resp_map: dense u8 (pitch = map_pitch) holding the
// keypoint response at NMS survivors with response >= thr_lo, 0 elsewhere.
void launch_fast(const uint8_t* img, long long img_stride, int w, int h, int pitch, int n_img,
                 int thr_lo, int thr_hi, uint8_t* resp_map, long long map_stride, int map_pitch,
                 unsigned long long* keysA, unsigned long long* keysB, BucketGeom bg,
                 cudaStream_t st, int* countA = nullptr, const int* gate = nullptr, int gate_min = 0,
                 int* work = nullptr /* 2 zeroed ints: dynamic tile order for the persistent blocks */,
                 long long key_stride = 0 /* distance of the key buffers of consecutive images (0: bg.nb) */);
// countA[i]: incremented once per bucket of image i that receives its first keysA entry.  gate: the blocks of image
// i return at once when gate[i] >= gate_min (the pipeline's low-threshold pass behind the thr_hi pass).
// ordered (row-major) compaction of a dense response map into a keypoint list
void launch_compact_keypoints(const uint8_t* resp_map, int w, int h, int map_pitch, int thr,
                              int* row_count /*[h+1] scratch*/, float2* xy_out, float* resp_out,
                              int cap, int* n_out, cudaStream_t st);
// generic pyrDown: n_img images, src/dst strides in bytes
void launch_pyrdown(const uint8_t* src, long long src_stride, int sw, int sh, int spitch,
                    uint8_t* dst, long long dst_stride, int dw, int dh, int dpitch, int n_img,
                    cudaStream_t st);
// context ingest of one frame into pyramid slot `slot` (ring, frame f -> f % 3): copies the new frame
// (left/right planes, n_seq each) into level 0 -- skipped when left == nullptr, i.e. level 0 was already
// written in place (H2D straight into the slot) --, builds levels 1.., runs FAST on the new LEFT image into
// fastkeys[slot].  Returns the number of kernel launches.
int launch_ingest(const DevArrays& d, const PyrGeom& pg, const BucketGeom& bg, const PipeConfig& pc,
                  const uint8_t* left, const uint8_t* right, int stride, long long seq_stride, int slot,
                  cudaStream_t st, bool pdl = false);

// look-ahead ingest of `n_frames` consecutive frames of ONE sequence into ring slots slot0 .. (n_seq == 1, 8UC1)
int launch_ingest_frames(const DevArrays& d, const PyrGeom& pg, const BucketGeom& bg, const PipeConfig& pc,
                         const uint8_t* left, const uint8_t* right, int stride, long long frame_stride, int slot0,
                         int n_frames, cudaStream_t st);

// ---- per-frame pipeline stages (all sequences of the context per launch) ----
void launch_first_frame(const DevArrays& d, cudaStream_t st);
// keybuf: FAST key buffer of the previous left image; specbuf / use_spec: speculative-cache buffer of this frame
// and whether klt_spec_fast / klt_spec_trk ran for it (0: every bucketed feature is queued for launch_klt)
void launch_bucket_merge(const DevArrays& d, const BucketGeom& bg, const PipeConfig& pc, int keybuf, int specbuf,
                         int use_spec, cudaStream_t st);
// speculative circular matching of the candidates of the coming frame (results looked up by bucket_merge)
void launch_klt_spec_fast(const DevArrays& d, const PyrGeom& pg, const PipeConfig& pc, const BucketGeom& bg,
                          int blocks_per_seq, int slot_prev, int slot_cur, int keybuf, int specbuf, cudaStream_t st);
void launch_klt_spec_trk(const DevArrays& d, const PyrGeom& pg, const PipeConfig& pc, const BucketGeom& bg,
                         int blocks_per_seq, int slot_prev, int slot_cur, int specbuf, cudaStream_t st);
int klt_resident_blocks_per_sm();
int klt_spec_resident_blocks_per_sm();   // same for the speculative kernels   // persistent KLT blocks that fit one SM (occupancy query)
void launch_klt(const DevArrays& d, const PyrGeom& pg, const PipeConfig& pc, int blocks_per_seq, int slot_prev,
                int slot_cur, cudaStream_t st);
// keybuf >= 0: also clears the FAST key buffer `keybuf` and the tracked-key scratch (consumed by bucket_merge)
// predbuf >= 0: also resets that buffer of FAST-pruning predictions (the one the frame after next reads)
void launch_select(const DevArrays& d, const PyrGeom& pg, const PipeConfig& pc, int slot_prev, int slot_cur,
                   int keybuf, int predbuf, cudaStream_t st);
void launch_pnp(const DevArrays& d, const PipeConfig& pc, cudaStream_t st);
void launch_pnp_hyp(const DevArrays& d, const PipeConfig& pc, cudaStream_t st);
void launch_pnp_final(const DevArrays& d, const PipeConfig& pc, cudaStream_t st);     // a then b
void launch_pnp_final_a(const DevArrays& d, const PipeConfig& pc, cudaStream_t st);   // RANSAC replay, inliers, feature set
void launch_pnp_final_b(const DevArrays& d, const PipeConfig& pc, cudaStream_t st);   // LM refit, gates, frame result
void launch_pnp_samples(const DevArrays& d, cudaStream_t st);   // standalone (select_kernel fuses it)

// ---- stage-level helpers used by the stateless API ----
void launch_lk_single(const LkPyramid& I, const LkPyramid& J, LkConfig cfg, const float2* pts, int n,
                      float2* out, uint8_t* status, cudaStream_t st);
// interleaved multi-channel variants (CLI-faithful 8UC3 mode)
void launch_lk_single_cn(const LkPyramid& I, const LkPyramid& J, LkConfig cfg, int cn, const float2* pts, int n,
                         float2* out, uint8_t* status, cudaStream_t st);
void launch_pyrdown_cn(const uint8_t* src, int sw, int sh, int spitch, uint8_t* dst, int dw, int dh, int dpitch,
                       int cn, cudaStream_t st);
void launch_bucket_generic(float2* xy, int* ages, int* strengths, int n, BucketGeom bg,
                           int per_bucket, int* scratch /* see svo_capi.cu */, float2* xy_out,
                           int* age_out, int* str_out, int* n_out, cudaStream_t st);
void launch_triangulate(const float* Pl_dev, const float* Pr_dev, const float2* pL, const float2* pR,
                        int n, float* X3, cudaStream_t st);
void launch_compact_bytes(const uint8_t* src, uint8_t* dst, int elem_bytes, int n, const uint8_t* keep,
                          int* n_out, cudaStream_t st);
void launch_close_points(const float2* a, const float2* b, int n, float thr, uint8_t* ok,
                         cudaStream_t st);

}  // namespace svo
