// svo_kernels.cu -- hand-written sm_100a kernels of the stereo-VO front end.
//
// Stage map (reference call stack):
//   fast_kernel            cv::FAST + NMS (feature_set.cpp:55-68) and the age-0 half of
//                          filterByBucketLocation (feature_set.cpp:95-147): per-bucket winners
//   pyrdown / ctx_* ingest cv::buildOpticalFlowPyramid image levels (vo.cpp:50-53, 200-201)
//   bucket_merge_kernel    appendFeaturesFromImage + bucketing + low-threshold retry (vo.cpp:325-332)
//   klt_kernel             the 4 calcOpticalFlowPyrLK legs L0->L1->R1->R0->L0 (vo.cpp:203-215)
//   select_kernel          findClosePoints + status AND + bounds + compaction (vo.cpp:217-238,
//                          341-364), ages++ (vo.cpp:70-72), triangulation (vo.cpp:89-94)
//   pnp_hyp_kernel / pnp_final_kernel   cameraToWorld (vo.cpp:282-313) + the failure gates of
//                          stereo_callback (vo.cpp:106-136)
#include <cuda.h>
#include <stdio.h>
#include <string.h>

#include "svo_kernels.cuh"
#include "svo_fast_packed.cuh"
#include "svo_epnp.cuh"

namespace svo {

// ============================================================================
// TMA (cp.async.bulk.tensor) tile loads: the image planes are described to the copy engine as 3-D uint8 tensors
// (x, y, image); one elected thread per block arms an mbarrier with the tile's byte count and issues the copy,
// the block waits on the barrier.  Out-of-range coordinates (negative starts, tiles hanging over the right / bottom
// edge) are zero-filled by the hardware, so the loaders need no address clamping.
// Requirements of the engine: base address and row / image strides multiples of 16 bytes, box width a multiple
// of 16 bytes and <= 256; planes that do not qualify take the kernels' LDG loaders.
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
// ---- programmatic dependent launch (the ingest is five short launches in a row): a kernel launched with the attribute
// may be set up while its predecessor in the stream still runs; it must not touch global memory before pdl_wait()
// (which returns when the predecessor has completed and flushed).  pdl_trigger() lets the successor's launch begin
// once every block of this grid has called it.  Both are no-ops in a launch without the attribute.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
// Only launch sets that own their SMs ask for it (batched contexts, serial frames): an early-launched grid waits ON
// the SMs, and on the shared 16-SM chain / ingest partition of the streaming pipeline that starves the pose chain.
static thread_local bool g_pdl_allow = false;
struct PdlScope {
    bool prev;
    explicit PdlScope(bool on) : prev(g_pdl_allow) { g_pdl_allow = on; }
    ~PdlScope() { g_pdl_allow = prev; }
};
static bool pdl_enabled() {
    static int on = -1;
    if (on < 0) {
        const char* e = getenv("SVO_PDL");
        on = e ? (atoi(e) != 0) : 1;
    }
    return on != 0;
}
template <class... KArgs, class... Args>
static void launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, cudaStream_t st, Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = g_pdl_allow && pdl_enabled() ? 1 : 0;
    cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}
__device__ __forceinline__ void mbar_init(unsigned long long* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned phase) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "SVO_MBAR_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra SVO_MBAR_DONE;\n"
        "bra SVO_MBAR_WAIT;\n"
        "SVO_MBAR_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(phase)
        : "memory");
}
__device__ __forceinline__ void tma_load_3d(void* dst, const CUtensorMap* map, int x, int y, int z, unsigned long long* bar) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(
            smem_u32(dst)),
        "l"(map), "r"(x), "r"(y), "r"(z), "r"(smem_u32(bar))
        : "memory");
}
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn tma_encode_fn() {
    static EncodeTiledFn fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qr;
        if (getenv("SVO_NO_TMA") || cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qr) != cudaSuccess ||
            qr != cudaDriverEntryPointSuccess)
            p = nullptr;
        return (EncodeTiledFn)p;
    }();
    return fn;
}
// uint8 planes [n_img][h][w] with row pitch `pitch` and image stride `img_stride` (bytes), box = box_w x box_h x 1
bool make_tile_map(CUtensorMap* map, const void* base, int w, int h, int pitch, long long img_stride, int n_img, int box_w,
                   int box_h) {
    EncodeTiledFn enc = tma_encode_fn();
    if (!enc || ((size_t)base & 15) || (pitch & 15) || box_w > 256 || (box_w & 15) || box_h > 256) return false;
    if (n_img > 1 && (img_stride & 15)) return false;
    const cuuint64_t gdim[3] = {(cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)n_img};
    const cuuint64_t gstr[2] = {(cuuint64_t)pitch, (cuuint64_t)(n_img > 1 ? img_stride : (long long)pitch * h)};
    const cuuint32_t box[3] = {(cuuint32_t)box_w, (cuuint32_t)box_h, 1u};
    const cuuint32_t estr[3] = {1u, 1u, 1u};
    return enc(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<void*>(base), gdim, gstr, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// ============================================================================
// helpers
struct WarpRed {
    static __device__ __forceinline__ int sum(int v) { return __reduce_add_sync(0xffffffffu, v); }
    // |v| < 2^27 per lane (4 window pixels x 8160 x 4080)
    static __device__ __forceinline__ long long sum64(long long v) {
        const unsigned u = (unsigned)((int)v + (1 << 27));
        const unsigned lo = __reduce_add_sync(0xffffffffu, u & 0xFFFFu);
        const unsigned hi = __reduce_add_sync(0xffffffffu, u >> 16);
        return ((long long)hi << 16) + (long long)lo - (32ll << 27);
    }
    // ---- float32 window sums in the order cv::calcOpticalFlowPyrLK forms them (svo_core.cuh cv_lk_sum_prod /
    // cv_lk_sum_b are the serial statements).  Window pixel (row y, column x) lives in lane (y % 3) * 10 + x, slot
    // y / 3 (LkMap<32>).  OpenCV runs five float32 accumulators -- SIMD lanes 0..3 (pixels l and l + 4 of every row)
    // and a scalar tail (pixels 8, 9) -- so the products are regrouped through shared memory into five chains of 20
    // consecutive words (chain c: [row y][first | second pixel]) and lane 5 k + c sums chain c of component k
    // sequentially (one 128-bit load per 4 terms); the fold of the five accumulators is OpenCV's.
    static __device__ __forceinline__ int chain_slot(int lane, int s) {   // word index of the lane's pixel `s` in its component
        const int x = lane % 10, y = lane / 10 + 3 * s;
        return (x < 8 ? (x & 3) : 4) * 20 + y * 2 + (x < 8 ? (x >> 2) : x - 8);
    }
    static __device__ __forceinline__ bool slot_valid(int lane, int s) { return lane < 30 && lane / 10 + 3 * s < 10; }
    // acc (lanes 5 k .. 5 k + 4) -> tail + ((q0 + q2) + (q1 + q3)) in lane 5 k
    static __device__ __forceinline__ float fold5(float acc) {
        const float q1 = __shfl_down_sync(0xffffffffu, acc, 1), q2 = __shfl_down_sync(0xffffffffu, acc, 2),
                    q3 = __shfl_down_sync(0xffffffffu, acc, 3), tail = __shfl_down_sync(0xffffffffu, acc, 4);
        return __fadd_rn(tail, __fadd_rn(__fadd_rn(acc, q2), __fadd_rn(q1, q3)));
    }
    static __device__ __forceinline__ void cv_sum_prod3(const int* ix, const int* iy, int lane, unsigned* stage,
                                                        float& a11, float& a12, float& a22) {
        float* st = reinterpret_cast<float*>(stage);
        __syncwarp();
#pragma unroll
        for (int s = 0; s < 4; ++s)
            if (slot_valid(lane, s)) {
                const int o = chain_slot(lane, s);
                const float fx = (float)ix[s], fy = (float)iy[s];
                st[o] = __fmul_rn(fx, fx);
                st[100 + o] = __fmul_rn(fx, fy);
                st[200 + o] = __fmul_rn(fy, fy);
            }
        __syncwarp();
        float acc = 0.f;
        if (lane < 15) {
            const float4* src = reinterpret_cast<const float4*>(st + (lane / 5) * 100 + (lane % 5) * 20);
#pragma unroll
            for (int i = 0; i < 5; ++i) {
                const float4 v = src[i];
                acc = __fadd_rn(__fadd_rn(__fadd_rn(__fadd_rn(acc, v.x), v.y), v.z), v.w);
            }
        }
        const float r = fold5(acc);
        a11 = __shfl_sync(0xffffffffu, r, 0);
        a12 = __shfl_sync(0xffffffffu, r, 5);
        a22 = __shfl_sync(0xffffffffu, r, 10);
    }
    // p1 / p2: the integer products diff * Ix, diff * Iy of the lane's (up to) four pixels.  Chains 0..3 add the
    // products of pixels l and l + 4 as int32 before the conversion (v_dotprod), the tail converts each product.
    static __device__ __forceinline__ void cv_sum_b(const int* p1, const int* p2, int lane, unsigned* stage, float& b1,
                                                    float& b2) {
        int* st = reinterpret_cast<int*>(stage);
        __syncwarp();
#pragma unroll
        for (int s = 0; s < 4; ++s)
            if (slot_valid(lane, s)) {
                const int o = chain_slot(lane, s);
                st[o] = p1[s];
                st[100 + o] = p2[s];
            }
        __syncwarp();
        float acc = 0.f;
        if (lane < 10) {
            const int4* src = reinterpret_cast<const int4*>(st + (lane / 5) * 100 + (lane % 5) * 20);
            const bool pairs = lane % 5 < 4;   // x + 0.f == x: the tail's second term per pixel pair is a real addend
#pragma unroll
            for (int i = 0; i < 5; ++i) {
                const int4 v = src[i];
                acc = __fadd_rn(__fadd_rn(acc, pairs ? (float)(v.x + v.y) : (float)v.x), pairs ? 0.f : (float)v.y);
                acc = __fadd_rn(__fadd_rn(acc, pairs ? (float)(v.z + v.w) : (float)v.z), pairs ? 0.f : (float)v.w);
            }
        }
        // chains 0..3 = qb0[0], qb0[2], qb1[0], qb1[2] (pixels 0|4, 1|5, 2|6, 3|7): b = tail + ((c0 + c2) + (c1 + c3))
        const float r = fold5(acc);
        b1 = __shfl_sync(0xffffffffu, r, 0);
        b2 = __shfl_sync(0xffffffffu, r, 5);
    }
    // Exact tiers (svo_core.cuh SerialRed::sum_b_grouped states the principle serially).  A lane's pixels share one
    // column, i.e. one accumulator group: A = even columns < 8 (chains c0, c2), B = odd columns < 8 (c1, c3), T = columns
    // 8, 9.  While the absolute products of every group total <= 2^24 the float32 sums inside a group are exact integers
    // in any order; when A + B is exact as well, b = fl(T + (A + B)) is the correctly rounded integer total, i.e.
    // (float)S of the plain warp total S.  Cost model (measured, tools/warp_ops_lat.cu): REDUX issues once per 8
    // cycles per SM sub-partition, SHFL once per 4, a vote like an ALU op -- so the tiers are ordered by REDUX count:
    //   V   every lane's own |sum| <= 2^24 / (lanes of its union): a vote, no REDUX beyond the two totals
    //   G   group bounds as 16-bit fields of (|sum| >> 13) + 1 per lane in three REDUX (field <= 2048 proves the group
    //       total < 2^24; <= 13 lanes x 2049 per field: no carry); A + B exact either because |A| + |B| <= 2^24 or,
    //       with the two tail totals, because |S - T| <= 2^24
    // otherwise the caller forms the sums term by term in OpenCV's order (cv_sum_b).
    static __device__ __forceinline__ bool sum_b_grouped(const int*, const int*, int lane, int s1, int s2, int a1, int a2,
                                                         float& b1, float& b2) {
        const int x = lane % 10;
        const bool gT = x >= 8, gB = !gT && (x & 1);
        const int S1 = __reduce_add_sync(0xffffffffu, s1), S2 = __reduce_add_sync(0xffffffffu, s2);
        // V: 26 lanes hold A u B (24 + the two idle lanes, which hold zeros), 6 lanes hold T
        const int lim = gT ? (1 << 24) / 6 : (1 << 24) / 26;
        if (__all_sync(0xffffffffu, max(a1, a2) <= lim)) {
            b1 = (float)S1;
            b2 = (float)S2;
            return true;
        }
        const unsigned f1 = min((unsigned)a1 >> 13, 2048u) + 1u, f2 = min((unsigned)a2 >> 13, 2048u) + 1u;
        const unsigned sh = gB ? 16u : 0u;
        const unsigned r1 = __reduce_add_sync(0xffffffffu, gT ? 0u : f1 << sh);            // component 1: A | B
        const unsigned r2 = __reduce_add_sync(0xffffffffu, gT ? 0u : f2 << sh);            // component 2: A | B
        const unsigned r3 = __reduce_add_sync(0xffffffffu, gT ? (f1 | (f2 << 16)) : 0u);   // tail: component 1 | 2
        constexpr unsigned K = 0x77FF77FFu;   // field + 0x77FF reaches bit 15 iff field > 2048
        if (((r1 + K) | (r2 + K) | (r3 + K)) & 0x80008000u) return false;
        if (max((r1 & 0xFFFFu) + (r1 >> 16), (r2 & 0xFFFFu) + (r2 >> 16)) > 2048u) {   // |A| + |B| may exceed 2^24: look at A + B
            const int T1 = __reduce_add_sync(0xffffffffu, gT ? s1 : 0), T2 = __reduce_add_sync(0xffffffffu, gT ? s2 : 0);
            if (max(abs(S1 - T1), abs(S2 - T2)) > (1 << 24)) return false;
        }
        b1 = (float)S1;
        b2 = (float)S2;
        return true;
    }
    // structure tensor: the squares bound the mixed products (|ix iy| <= (ix^2 + iy^2) / 2), so the two diagonal sums
    // of a group <= 2^24 prove all three exact.  The terms are squares -- no cancellation, A + B usually needs its own
    // rounding -- so the three group sums are formed separately (nine REDUX); a vote on the lanes' own sums decides
    // first whether that can pay (bucket winners are strong corners: it holds for about a fifth of the windows).
    static __device__ __forceinline__ bool sum_prod3_grouped(const int* ix, const int* iy, int lane, float& a11, float& a12,
                                                             float& a22) {
        const int x = lane % 10;
        const bool gT = x >= 8, gB = !gT && (x & 1), gA = !gT && !(x & 1);
        unsigned u11 = 0, u22 = 0;
        int s12 = 0;
#pragma unroll
        for (int s = 0; s < 4; ++s) {
            u11 += (unsigned)(ix[s] * ix[s]);
            u22 += (unsigned)(iy[s] * iy[s]);
            s12 += ix[s] * iy[s];
        }
        const unsigned lim = gT ? (1u << 24) / 6 : (1u << 24) / 13;   // 12 lanes + one idle lane per group, 6 tail lanes
        if (!__all_sync(0xffffffffu, max(u11, u22) <= lim)) return false;
        const unsigned A11 = __reduce_add_sync(0xffffffffu, gA ? u11 : 0u), B11 = __reduce_add_sync(0xffffffffu, gB ? u11 : 0u),
                       T11 = __reduce_add_sync(0xffffffffu, gT ? u11 : 0u);
        const unsigned A22 = __reduce_add_sync(0xffffffffu, gA ? u22 : 0u), B22 = __reduce_add_sync(0xffffffffu, gB ? u22 : 0u),
                       T22 = __reduce_add_sync(0xffffffffu, gT ? u22 : 0u);
        const int A12 = __reduce_add_sync(0xffffffffu, gA ? s12 : 0), B12 = __reduce_add_sync(0xffffffffu, gB ? s12 : 0),
                  T12 = __reduce_add_sync(0xffffffffu, gT ? s12 : 0);
        a11 = __fadd_rn((float)T11, __fadd_rn((float)A11, (float)B11));
        a12 = __fadd_rn((float)T12, __fadd_rn((float)A12, (float)B12));
        a22 = __fadd_rn((float)T22, __fadd_rn((float)A22, (float)B22));
        return true;
    }
    // any 64-bit per-lane value (multi-channel windows)
    static __device__ __forceinline__ long long sum64g(long long v) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        return v;
    }
    static __device__ __forceinline__ void sync() { __syncwarp(); }
};

// exclusive scan of one int per thread over the block; returns the offset, *total = sum
template <int NT>
__device__ __forceinline__ int block_excl_scan(int v, int* total, int* smem /* NT/32 + 1 ints */) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int n = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += n;
    }
    __syncthreads();   // protect smem reuse across calls
    if (lane == 31) smem[wid] = incl;
    __syncthreads();
    if (wid == 0) {
        int w = (lane < NT / 32) ? smem[lane] : 0;
        int wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int n = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane >= o) wi += n;
        }
        if (lane < NT / 32) smem[lane] = wi - w;
        if (lane == 31) smem[NT / 32] = wi;
    }
    __syncthreads();
    *total = smem[NT / 32];
    return smem[wid] + incl - v;
}

__device__ __forceinline__ unsigned long long ld_key(const unsigned long long* p) {
    return __ldcg(p);   // keys are produced by L2 atomics: bypass L1
}

// ============================================================================
// FAST-9/16 + 3x3 NMS.  One block = 120 x 14 output pixels; scores are computed on the 128 x 16 region around
// them from a 136 x 22 byte tile in shared memory.  A thread scores a QUAD of 4 horizontally adjacent pixels as
// two int16x2 lanes (DPX min / max on packed halves, ring pixels extracted from 12-byte row windows by PRMT;
// arithmetic: svo_fast_packed.cuh, checked on the host against the scalar statement):
//   phase 1  every quad: the compass test (any 9-arc contains two adjacent compass pixels) on raw ring pixels, six
//            packed min / max per pixel pair; quads that cannot hold a corner at thr_lo are dropped (~3/4)
//   phase 2  the surviving quads, compacted through a shared list so no warp idles: the exact arc score (arcs
//            folded in pairs, 75 packed operations per pixel pair)
//   phase 3  strict 3x3 NMS + per-bucket winners (64-bit atomicMax keys) / dense response map
constexpr int FT_OW = 120, FT_OH = 14, FT_SW = 128, FT_SH = 16, FT_TW = 144, FT_TH = 22;   // tile: 136 columns used, pitch 144
constexpr int FT_TASKS = (FT_SW / 4) * FT_SH;   // 512 quads

template <int NT>
__global__ void __launch_bounds__(NT)
fast_kernel(const uint8_t* __restrict__ img, long long img_stride, int w, int h, int pitch, int thr_lo,
            int thr_hi, uint8_t* __restrict__ resp_map, long long map_stride, int map_pitch,
            unsigned long long* __restrict__ keysA, unsigned long long* __restrict__ keysB,
            BucketGeom bg, int* __restrict__ countA, const int* __restrict__ gate, int gate_min, int tiles_x, int tiles_y,
            const __grid_constant__ CUtensorMap tmap, int use_tma, int n_img, int* __restrict__ work, long long key_stride) {
    pdl_wait();      // the pyramid kernels (level 0) / the threshold-20 pass (gate) in front of this launch are complete
    pdl_trigger();
    // low-threshold pass of the pipeline: only needed when the retry of vo.cpp:327-332 can happen at all, i.e. when
    // the thr_hi pass left fewer than gate_min buckets with a winner (every such bucket yields one bucketed feature)
    if (gate && gate[blockIdx.z] >= gate_min) return;
    // Tiles are double buffered: in the ungated launches the blocks are persistent (1-D grid over all (image, tile)
    // pairs) and the copy engine fetches tile k + 1 of a block while the block scores tile k.
    __shared__ __align__(128) uint8_t tile2[2][(FT_TH * FT_TW + 127) & ~127];   // a TMA destination is 128-byte aligned
    __shared__ __align__(8) unsigned long long tma_bar[2];
    __shared__ __align__(16) uint8_t score[FT_SH * FT_SW];
    __shared__ unsigned short list[FT_TASKS];
    __shared__ unsigned short hits[FT_OW * FT_OH];   // score-array positions of the nonzero scores among the outputs
    __shared__ int n_list, n_hits;
    __shared__ int tinfo[2][4];   // per tile buffer: tile id, image, x0, y0 (decoded once, by the thread that fetches the id)
    const int tid = threadIdx.x, lane = tid & 31;
    if (use_tma && tid == 0) {
        mbar_init(&tma_bar[0], 1);
        mbar_init(&tma_bar[1], 1);
    }
    unsigned tma_phase = 0u;   // bit b: the phase buffer b's barrier completes next
    // ungated launches: persistent blocks over all tiles of all images; the gated pass, which almost never has work,
    // runs a few blocks per image (blockIdx.z) that loop over its tiles, so that its early exit costs next to nothing
    const int tpi = tiles_x * tiles_y, n_tiles = gate ? tpi : tpi * n_img;
    // thread 0 only: decode tile `id` into tinfo[buf] and start its bulk tensor copy (the 144 x 22 box at the aligned
    // column at or below x0 - 8)
    auto fetch = [&](int id, int buf) {
        tinfo[buf][0] = id;
        if (id >= n_tiles) return;
        const int zz = gate ? (int)blockIdx.z : id / tpi;
        const int t = gate ? id : id - zz * tpi;
        const int ty = t / tiles_x;
        const int yy = ty * FT_OH, xx = (t - ty * tiles_x) * FT_OW;
        tinfo[buf][1] = zz;
        tinfo[buf][2] = xx;
        tinfo[buf][3] = yy;
        if (use_tma) {
            mbar_expect_tx(&tma_bar[buf], FT_TH * FT_TW);
            tma_load_3d(tile2[buf], &tmap, (xx - 8) & ~15, yy - 4, zz, &tma_bar[buf]);
        }
    };
    // tile order: a global work counter when the launch has one (work[0]; the tiles cost very different amounts, so
    // the persistent blocks take them dynamically; work[1] counts finished blocks and the last one re-arms both),
    // else block b takes tiles b, b + gridDim.x, ...
    __syncthreads();
    if (tid == 0) {
        n_list = 0;
        fetch(work ? atomicAdd(work, 1) : (int)blockIdx.x, 0);
    }
    __syncthreads();
    const unsigned kt = (unsigned)(0x8000 - thr_lo - 1) * 0x00010001u;           // compass test (fast_pair_compass)
    const unsigned kz = (unsigned)(0x8000 - (thr_lo + 256) - 1) * 0x00010001u;   // score > thr_lo (fast_quad_pack)
    const float inv_bh = 1.f / (float)bg.bh, inv_bw = 1.f / (float)bg.bw;
    for (int iter = 0;; ++iter) {
    const int buf = iter & 1;
    const int tile_id = tinfo[buf][0];
    if (tile_id >= n_tiles) break;
    uint8_t* tile = tile2[buf];
    const int z = tinfo[buf][1], x0 = tinfo[buf][2], y0 = tinfo[buf][3];
    const uint8_t* src = img + (long long)z * img_stride;
    if (tid == 0) {
        n_hits = 0;   // (n_list is reset after phase 2: with TMA tiles there is no barrier between here and phase 1)
        fetch(work ? atomicAdd(work, 1) : tile_id + (int)gridDim.x, buf ^ 1);   // the block's next tile: prefetched now
    }
    // ---- tile rows [y0 - 4, y0 + 18) x cols [x0 - 8, x0 + 128), clamped (clamped values only ever reach pixels
    // whose score is forced to 0).  Word loads when the block is inside the image and the plane is word aligned.
    const int tx0 = x0 - 8, ty0 = y0 - 4;
    // Out-of-image bytes only ever feed pixels whose score is forced to 0 (the 3-pixel border), so whole words
    // are fetched with the word column / row clamped into the plane (needs a word-aligned plane whose pitch
    // covers the last word that holds valid pixels).
    const bool fastpath = (pitch & 3) == 0 && (((size_t)src) & 3) == 0 && ((w + 3) & ~3) <= pitch;
    // the copy engine wants the box to start at a multiple of 16 bytes along x: the 144-byte box starts at the aligned
    // column at or below tx0 (tx0 - 0 or tx0 - 8) and the 136 columns that are used sit `toff` bytes into each row
    const int toff = use_tma ? (tx0 & 15) : 0;
    const uint8_t* tl = tile + toff;
    if (use_tma) {   // zero fill outside the image
        mbar_wait(&tma_bar[buf], (tma_phase >> buf) & 1u);
        tma_phase ^= 1u << buf;
    } else if (fastpath) {
        const int wmax = ((w + 3) >> 2) - 1;
        constexpr int NW = FT_TH * (FT_TW / 4), NIT = (NW + NT - 1) / NT;
        unsigned v[NIT];
#pragma unroll
        for (int it = 0; it < NIT; ++it) {   // all loads of a thread in flight before the first store
            const int i = min(tid + it * NT, NW - 1);
            const int r = i / (FT_TW / 4), k = i - r * (FT_TW / 4);
            const int gy = min(max(ty0 + r, 0), h - 1), gk = min(max((tx0 >> 2) + k, 0), wmax);
            v[it] = __ldg(reinterpret_cast<const unsigned*>(src + (long long)gy * pitch) + gk);
        }
#pragma unroll
        for (int it = 0; it < NIT; ++it) {
            const int i = tid + it * NT;
            if (i < NW) reinterpret_cast<unsigned*>(tile)[i] = v[it];
        }
    } else {
        for (int i = tid; i < FT_TH * FT_TW; i += NT) {
            const int r = i / FT_TW, c = i - r * FT_TW;
            const int gy = min(max(ty0 + r, 0), h - 1), gx = min(max(tx0 + c, 0), w - 1);
            tile[i] = __ldg(src + (long long)gy * pitch + gx);
        }
    }
    if (!use_tma) __syncthreads();   // TMA tiles: every thread has waited on the tile's mbarrier itself

    // ---- phase 1: compass test.  Task t: score row t / 32, quad t % 32 (score col 4q <-> tile col 4q + 4,
    // score row r <-> tile row r + 3).  The 12-byte row window of a quad starts at tile col 4q.  A warp's tasks are
    // whole score rows; its survivors are appended to the list with one shared atomic per warp.
    {
        constexpr int TPT = FT_TASKS / NT;
        unsigned m[TPT];
        int total = 0;
#pragma unroll
        for (int s = 0; s < TPT; ++s) {
            const int t = tid + s * NT;
            const int r = t >> 5, q = t & 31;
            const unsigned* top = reinterpret_cast<const unsigned*>(tl + (r + 0) * FT_TW) + q;
            const unsigned* mid = reinterpret_cast<const unsigned*>(tl + (r + 3) * FT_TW) + q;
            const unsigned* bot = reinterpret_cast<const unsigned*>(tl + (r + 6) * FT_TW) + q;
            const unsigned wt[3] = {top[0], top[1], top[2]}, wm[3] = {mid[0], mid[1], mid[2]}, wb[3] = {bot[0], bot[1], bot[2]};
            const unsigned pass = (fast_pair_compass<4>(wt, wm, wb, kt) | fast_pair_compass<6>(wt, wm, wb, kt)) & 0x80008000u;
            reinterpret_cast<unsigned*>(score)[t] = 0u;
            m[s] = __ballot_sync(0xffffffffu, pass != 0u);
            total += __popc(m[s]);
        }
        int base = 0;
        if (lane == 0 && total) base = atomicAdd(&n_list, total);
        base = __shfl_sync(0xffffffffu, base, 0);
        const unsigned lt = (1u << lane) - 1u;
#pragma unroll
        for (int s = 0; s < TPT; ++s) {
            if ((m[s] >> lane) & 1u) list[base + __popc(m[s] & lt)] = (unsigned short)(tid + s * NT);
            base += __popc(m[s]);
        }
    }
    __syncthreads();

    // ---- phase 2: exact scores of the surviving quads
    const int n = n_list;
    const bool edge_tile = x0 - 4 < 3 || x0 + FT_SW - 4 > w - 3 || y0 - 1 < 3 || y0 + FT_SH - 1 > h - 3;
    for (int i = tid; i < n; i += NT) {
        const int t = list[i];
        const int r = t >> 5, q = t & 31;
        unsigned rows[7][3];
#pragma unroll
        for (int k = 0; k < 7; ++k) {
            const unsigned* p = reinterpret_cast<const unsigned*>(tl + (r + k) * FT_TW) + q;
            rows[k][0] = p[0]; rows[k][1] = p[1]; rows[k][2] = p[2];
        }
        unsigned out = fast_quad_pack(fast_pair_score_ring<4>(rows), fast_pair_score_ring<6>(rows), kz);
        if (edge_tile) {   // the 3-pixel image border never holds a keypoint
            const int gy = y0 - 1 + r, gx = x0 - 4 + 4 * q;
            if (gy < 3 || gy >= h - 3) out = 0u;
            if (gx < 3 || gx + 3 >= w - 3) {
#pragma unroll
                for (int e = 0; e < 4; ++e)
                    if (gx + e < 3 || gx + e >= w - 3) out &= ~(0xFFu << (8 * e));
            }
        }
        if (out == 0u) continue;
        reinterpret_cast<unsigned*>(score)[t] = out;
        // scores among the block's outputs (not the 1-pixel score halo: rows 1 .. FT_OH, quads 1 .. FT_OW / 4) are the
        // candidates of phase 3
        if (!resp_map && r >= 1 && r <= FT_OH && q >= 1 && q <= FT_OW / 4) {
            const int cnt = (out & 0xFFu ? 1 : 0) + (out & 0xFF00u ? 1 : 0) + (out & 0xFF0000u ? 1 : 0) + (out >> 24 ? 1 : 0);
            int hb;   // one shared atomic per thread with hits (few per warp: cheaper than a warp-aggregated prefix scan)
            asm volatile("atom.shared.add.u32 %0, [%1], %2;" : "=r"(hb) : "r"(smem_u32(&n_hits)), "r"(cnt) : "memory");
#pragma unroll
            for (int e = 0; e < 4; ++e)
                if ((out >> (8 * e)) & 0xFFu) hits[hb++] = (unsigned short)(r * FT_SW + 4 * q + e);
        }
    }
    __syncthreads();
    if (tid == 0) n_list = 0;   // every thread has read it; the next tile's phase 1 comes after the barrier that ends this tile

    // ---- phase 3: NMS (strict maximum over the 8 neighbours) + outputs
    if (!resp_map) {   // pipeline mode (bucket keys only): one thread per nonzero score instead of the dense sweep
        const int nh = n_hits;
        for (int i = tid; i < nh; i += NT) {
            const int idx = hits[i];
            const uint8_t* sp = score + idx;
            const int v = *sp;
            const bool keep = v > sp[-1] && v > sp[1] && v > sp[-FT_SW - 1] && v > sp[-FT_SW] && v > sp[-FT_SW + 1] &&
                              v > sp[FT_SW - 1] && v > sp[FT_SW] && v > sp[FT_SW + 1];
            if (!keep) continue;
            const int gy = y0 + (idx / FT_SW) - 1, gx = x0 + (idx % FT_SW) - 4;
            const int br = (int)(((float)gy + 0.5f) * inv_bh), bc = (int)(((float)gx + 0.5f) * inv_bw);
            if (br >= bg.start_row && 0 < bg.age_threshold) {
                const unsigned long long key = bucket_key(bucket_score(0, v, bg.fast_threshold),
                                                          SVO_FAST_ORDER_BASE + (unsigned)(gy * w + gx), v);
                const long long bidx = (long long)z * key_stride + br * bg.n_bw + bc;
                if (keysB) atomicMax(keysB + bidx, key);
                if (keysA && v >= thr_hi) {
                    const unsigned long long old = atomicMax(keysA + bidx, key);
                    if (old == 0ull && countA) atomicAdd(countA + z, 1);   // buckets with a thr_hi winner
                }
            }
        }
        __syncthreads();   // every thread is done with this tile's shared arrays
        continue;
    }
    // dense sweep, one output quad per task: most score words are zero and are dismissed with one load
    for (int i = tid; i < FT_OH * (FT_OW / 4); i += NT) {
        const int oy = i / (FT_OW / 4), oq = i - oy * (FT_OW / 4);
        const int gx0 = x0 + 4 * oq, gy = y0 + oy;
        if (gx0 >= w || gy >= h) continue;
        const uint8_t* sp0 = score + (oy + 1) * FT_SW + 4 * (oq + 1);
        const unsigned sw = *reinterpret_cast<const unsigned*>(sp0);
        unsigned outw = 0;
        if (sw != 0u) {
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int v = (sw >> (8 * e)) & 0xFF;
                if (v == 0) continue;
                const uint8_t* sp = sp0 + e;
                const bool keep = v > sp[-1] && v > sp[1] && v > sp[-FT_SW - 1] && v > sp[-FT_SW] && v > sp[-FT_SW + 1] &&
                                  v > sp[FT_SW - 1] && v > sp[FT_SW] && v > sp[FT_SW + 1];
                const int gx = gx0 + e;
                if (!keep || gx >= w) continue;
                outw |= (unsigned)v << (8 * e);
                if (keysA || keysB) {
                    // exact for these small integers: the quotient is never within an ulp of an integer from below
                    const int br = (int)(((float)gy + 0.5f) * inv_bh), bc = (int)(((float)gx + 0.5f) * inv_bw);
                    if (br >= bg.start_row && 0 < bg.age_threshold) {
                        const unsigned long long key = bucket_key(bucket_score(0, v, bg.fast_threshold),
                                                                  SVO_FAST_ORDER_BASE + (unsigned)(gy * w + gx), v);
                        const long long bidx = (long long)z * key_stride + br * bg.n_bw + bc;
                        if (keysB) atomicMax(keysB + bidx, key);
                        if (keysA && v >= thr_hi) {
                            const unsigned long long old = atomicMax(keysA + bidx, key);
                            if (old == 0ull && countA) atomicAdd(countA + z, 1);   // buckets with a thr_hi winner
                        }
                    }
                }
            }
        }
        uint8_t* o = resp_map + (long long)z * map_stride + (long long)gy * map_pitch + gx0;
        if (gx0 + 3 < w && (map_pitch & 3) == 0 && ((size_t)(resp_map + (long long)z * map_stride) & 3) == 0) {
            *reinterpret_cast<unsigned*>(o) = outw;
        } else {
            for (int e = 0; e < 4; ++e)
                if (gx0 + e < w) o[e] = (uint8_t)(outw >> (8 * e));
        }
    }
    __syncthreads();   // every thread is done with this tile's shared arrays
    }   // tile loop
    if (work && tid == 0 && atomicAdd(work + 1, 1) == (int)gridDim.x - 1) {   // last block out: ready for the next launch
        work[0] = 0;
        work[1] = 0;
    }
}

void launch_fast(const uint8_t* img, long long img_stride, int w, int h, int pitch, int n_img,
                 int thr_lo, int thr_hi, uint8_t* resp_map, long long map_stride, int map_pitch,
                 unsigned long long* keysA, unsigned long long* keysB, BucketGeom bg, cudaStream_t st,
                 int* countA, const int* gate, int gate_min, int* work, long long key_stride) {
    const int tiles_x = (w + FT_OW - 1) / FT_OW, tiles_y = (h + FT_OH - 1) / FT_OH;
    // ungated: persistent blocks (as many as stay resident) over all (image, tile) pairs; gated: a few blocks per image
    static int n_sm = 0;
    if (!n_sm) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev);
        if (n_sm <= 0) n_sm = 148;
    }
    static int nt = 0;    // threads per block (SVO_FAST_NT = 128 | 256)
    static int bps = 0;   // persistent blocks per SM (SVO_FAST_BPS; what the register file holds is the useful maximum)
    if (!bps) {
        const char* e = getenv("SVO_FAST_NT");
        nt = e && atoi(e) == 256 ? 256 : (e && atoi(e) == 64 ? 64 : 128);
        e = getenv("SVO_FAST_BPS");
        bps = e ? atoi(e) : 0;
        if (bps <= 0) {
            const cudaError_t rc = nt == 256 ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, fast_kernel<256>, 256, 0)
                                   : nt == 64 ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, fast_kernel<64>, 64, 0)
                                              : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, fast_kernel<128>, 128, 0);
            if (rc != cudaSuccess || bps <= 0) bps = 1024 / nt;
        }
    }
    const long long total = (long long)tiles_x * tiles_y * n_img, cap = (long long)bps * n_sm;
    // (gated pass: 3072 blocks that return at once measured 4.4 us per 64 images, 1024 about half of that)
    const int gated_blocks = 2048 / nt;
    const dim3 grid = gate ? dim3(tiles_x * tiles_y < gated_blocks ? tiles_x * tiles_y : gated_blocks, 1, n_img)
                           : dim3((unsigned)(total < cap ? total : cap), 1, 1);
    CUtensorMap tmap;
    memset(&tmap, 0, sizeof(tmap));
    const int use_tma = make_tile_map(&tmap, img, w, h, pitch, img_stride, n_img, FT_TW, FT_TH) ? 1 : 0;
    const long long ks = key_stride > 0 ? key_stride : (long long)bg.nb;
    if (nt == 256)
        launch_pdl(fast_kernel<256>, grid, dim3(256), st, img, img_stride, w, h, pitch, thr_lo, thr_hi, resp_map, map_stride, map_pitch, keysA, keysB,
                                               bg, countA, gate, gate_min, tiles_x, tiles_y, tmap, use_tma, n_img, gate ? nullptr : work, ks);
    else if (nt == 64)
        launch_pdl(fast_kernel<64>, grid, dim3(64), st, img, img_stride, w, h, pitch, thr_lo, thr_hi, resp_map, map_stride, map_pitch, keysA, keysB,
                                             bg, countA, gate, gate_min, tiles_x, tiles_y, tmap, use_tma, n_img, gate ? nullptr : work, ks);
    else
        launch_pdl(fast_kernel<128>, grid, dim3(128), st, img, img_stride, w, h, pitch, thr_lo, thr_hi, resp_map, map_stride, map_pitch, keysA, keysB,
                                               bg, countA, gate, gate_min, tiles_x, tiles_y, tmap, use_tma, n_img, gate ? nullptr : work, ks);
}

// ---- ordered compaction of the response map (row-major == cv::FAST output order)
__global__ void kp_rowcount_kernel(const uint8_t* __restrict__ map, int w, int h, int pitch, int thr,
                                   int* __restrict__ row_count) {
    const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= h) return;
    int c = 0;
    for (int x = lane; x < w; x += 32) c += (map[(long long)row * pitch + x] >= thr) ? 1 : 0;
    c = __reduce_add_sync(0xffffffffu, c);
    if (lane == 0) row_count[row] = c;
}
__global__ void __launch_bounds__(1024) scan_kernel(int* __restrict__ v, int n, int* __restrict__ total_out) {
    __shared__ int sm[33];
    const int chunk = (n + 1023) / 1024;
    const int b0 = min(n, (int)threadIdx.x * chunk), b1 = min(n, b0 + chunk);
    int s = 0;
    for (int i = b0; i < b1; ++i) s += v[i];
    int total;
    int off = block_excl_scan<1024>(s, &total, sm);
    for (int i = b0; i < b1; ++i) {
        const int c = v[i];
        v[i] = off;
        off += c;
    }
    if (threadIdx.x == 0 && total_out) *total_out = total;
}
__global__ void kp_write_kernel(const uint8_t* __restrict__ map, int w, int h, int pitch, int thr,
                                const int* __restrict__ row_off, float2* __restrict__ xy,
                                float* __restrict__ resp, int cap) {
    const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= h) return;
    int base = row_off[row];
    for (int x0 = 0; x0 < w; x0 += 32) {
        const int x = x0 + lane;
        const int v = (x < w) ? map[(long long)row * pitch + x] : 0;
        const bool k = v >= thr && v > 0;
        const unsigned m = __ballot_sync(0xffffffffu, k);
        if (k) {
            const int o = base + __popc(m & ((1u << lane) - 1));
            if (o < cap) {
                xy[o] = make_float2((float)x, (float)row);
                resp[o] = (float)v;
            }
        }
        base += __popc(m);
    }
}
void launch_compact_keypoints(const uint8_t* resp_map, int w, int h, int map_pitch, int thr, int* row_count,
                              float2* xy_out, float* resp_out, int cap, int* n_out, cudaStream_t st) {
    const int wpb = 8;
    kp_rowcount_kernel<<<(h + wpb - 1) / wpb, wpb * 32, 0, st>>>(resp_map, w, h, map_pitch, thr, row_count);
    scan_kernel<<<1, 1024, 0, st>>>(row_count, h, n_out);
    kp_write_kernel<<<(h + wpb - 1) / wpb, wpb * 32, 0, st>>>(resp_map, w, h, map_pitch, thr, row_count, xy_out,
                                                              resp_out, cap);
}

// ============================================================================
// pyrDown (5x5 binomial, (s + 128) >> 8, BORDER_REFLECT_101): shared-memory tiled, separable, packed 16-bit lanes.
// One block = 128 x 16 output pixels <- a 288 x 35 byte source tile (16 bytes of margin left and right keep the
// tile rows 16-byte aligned with the image: 128-bit loads).  A thread owns one output word (4 pixels) of 4 output
// rows: per source row it splits the 16-byte window around its 8 source columns into even / odd pixels as
// int16x2 pairs (LOP / PRMT), forms the 1 4 6 4 1 row sums of two adjacent outputs per register (<= 4080 per
// half), keeps the 11 row sums of its column strip in registers and runs the vertical 1 4 6 4 1 on the packed
// pairs (<= 65408 per half: no carry between the halves), one PRMT packs the four (v >> 8) bytes.  In copy mode
// the block also writes the 256 x 32 source pixels it covers into the context's level-0 plane, so a new frame is
// read from HBM exactly once for both the level-0 copy and level 1.
// RPT = output rows per thread (block = 128 x 4 * RPT outputs): 4 for batched launches (fewest row sums per output),
// 2 or 1 when one frame's launch would otherwise be a handful of long blocks (latency instead of throughput).
constexpr int PD_OW = 128, PD_TW = 2 * PD_OW + 32, PD_BW = 160, PD_THREADS = 128;
struct PyrDownJob {
    const uint8_t* src[2];    // per camera (n_cam == 1: only [0])
    long long src_seq_stride;
    int sw, sh, spitch;
    uint8_t* dst[2];
    long long dst_seq_stride;
    int dw, dh, dpitch;
    uint8_t* copy[2];         // optional level-0 copy of the source (nullptr: none)
    long long copy_seq_stride;
    int cpitch;
    int n_cam;
    int use_tma;              // tile loads by cp.async.bulk.tensor (the source planes qualify, see make_tile_map)
};
// the 4 bytes at an arbitrary byte address, from the two aligned words around it
__device__ __forceinline__ unsigned ld_word_any(const uint8_t* p) {
    const unsigned sh = (unsigned)((size_t)p & 3);
    const unsigned* a = reinterpret_cast<const unsigned*>(p - sh);
    const unsigned w0 = __ldg(a);
    if (sh == 0) return w0;
    return __funnelshift_r(w0, __ldg(a + 1), 8 * sh);
}
// row sums 1 4 6 4 1 of the four outputs whose source columns are b0, b2, b4, b6 of the window
// w0 = b[-4..-1], w1 = b[0..3], w2 = b[4..7], w3 = b[8..11]; returns (h0, h1) and (h2, h3) as int16x2
__device__ __forceinline__ void pd_row_sums(unsigned w0, unsigned w1, unsigned w2, unsigned w3, unsigned& hA,
                                            unsigned& hB) {
    const unsigned e0 = w0 & 0x00FF00FFu, e1 = w1 & 0x00FF00FFu, e2 = w2 & 0x00FF00FFu, e3 = w3 & 0x00FF00FFu;
    const unsigned o0 = __byte_perm(w0, 0u, 0x4341u), o1 = __byte_perm(w1, 0u, 0x4341u), o2 = __byte_perm(w2, 0u, 0x4341u);
    const unsigned g0 = o0 * 4u + e0, g1 = o1 * 4u + e1, g2 = o2 * 4u + e2;     // even + 4 * odd-to-the-right
    // outputs (0, 1): centres b0, b2: E[-1..1] + 4 * (O[-1] + O[0]) + 5 * E[0] more
    hA = e1 * 5u + g1 + __byte_perm(g0, g1, 0x5432u) + __byte_perm(e1, e2, 0x5432u);
    hB = e2 * 5u + g2 + __byte_perm(g1, g2, 0x5432u) + __byte_perm(e2, e3, 0x5432u);
}
template <int PD_ROWS_PER_THREAD>
__global__ void __launch_bounds__(PD_THREADS)
pyrdown_tile_kernel(PyrDownJob j, const __grid_constant__ CUtensorMap map0, const __grid_constant__ CUtensorMap map1) {
    constexpr int PD_OH = 4 * PD_ROWS_PER_THREAD, PD_TH = 2 * PD_OH + 3;
    // The 288-byte tile rows are kept as two overlapping 160-byte halves (columns 0..159 and 128..287): a TMA box is
    // at most 256 bytes wide, and each half-warp's 16-byte windows then lie inside one half.
    __shared__ __align__(128) uint8_t tileA[PD_TH * PD_BW];
    __shared__ __align__(128) uint8_t tileB[PD_TH * PD_BW];
    __shared__ __align__(8) unsigned long long tma_bar;
    auto st_chunk = [&](int r, int k, const uint4& v) {   // 16-byte chunk k (tile columns 16 k ..) of tile row r
        if (k <= 9) *reinterpret_cast<uint4*>(tileA + r * PD_BW + 16 * k) = v;
        if (k >= 8) *reinterpret_cast<uint4*>(tileB + r * PD_BW + 16 * (k - 8)) = v;
    };
    auto ld_chunk = [&](int r, int k) -> uint4 {
        return k <= 9 ? *reinterpret_cast<const uint4*>(tileA + r * PD_BW + 16 * k)
                      : *reinterpret_cast<const uint4*>(tileB + r * PD_BW + 16 * (k - 8));
    };
    if (j.use_tma && threadIdx.x == 0) mbar_init(&tma_bar, 1);
    pdl_wait();      // the level this launch reads is complete
    pdl_trigger();
    const int cam = j.n_cam == 2 ? (blockIdx.z & 1) : 0, seq = j.n_cam == 2 ? (blockIdx.z >> 1) : blockIdx.z;
    const uint8_t* src = j.src[cam] + (long long)seq * j.src_seq_stride;
    uint8_t* dst = j.dst[cam] + (long long)seq * j.dst_seq_stride;
    const int X0 = blockIdx.x * PD_OW, Y0 = blockIdx.y * PD_OH, tid = threadIdx.x;
    const int sx0 = 2 * X0 - 16, sy0 = 2 * Y0 - 2;   // image position of tile[0][0]
    // ---- tile load, 16 bytes per task.  Chunks that lie inside their image row ("inner") are fetched by every
    // thread unconditionally (addresses of the other chunks are clamped, their data dropped), so all loads of a
    // thread are in flight before the first shared-memory store: one 128-bit load per chunk where the rows are
    // 16-byte aligned, else the five aligned words around the chunk + funnel shifts.  The at most two chunks per
    // side that touch the image's left / right edge are then filled word by word / byte by byte with the reflected
    // column by a second, dense sweep; chunks beyond column sw + 1 only feed outputs >= dw and are not loaded.
    constexpr int NCH = PD_TW / 16, NT = PD_TH * NCH, NIT = (NT + PD_THREADS - 1) / PD_THREADS;
    const bool al16 = (((size_t)src) & 15) == 0 && (j.spitch & 15) == 0;
    const int lo = al16 ? 0 : 4, lim = al16 ? 16 : 20;   // inner chunk: lo <= gx && gx + lim <= sw
    if (j.use_tma) {
        // ---- TMA: the two halves as two bulk tensor copies on one mbarrier (zero fill outside the plane), then
        // BORDER_REFLECT_101 is restored in the tiles that touch an image edge: two columns per side within the
        // rows that exist, then whole rows copied from their mirror rows inside the tile.
        __syncthreads();   // barrier initialised
        if (tid == 0) {
            const CUtensorMap* map = cam ? &map1 : &map0;
            mbar_expect_tx(&tma_bar, 2 * PD_TH * PD_BW);
            tma_load_3d(tileA, map, sx0, sy0, seq, &tma_bar);
            tma_load_3d(tileB, map, sx0 + 128, sy0, seq, &tma_bar);
        }
        mbar_wait(&tma_bar, 0);
        if (sx0 < 0 || sx0 + PD_TW > j.sw || sy0 < 0 || sy0 + PD_TH > j.sh) {
            auto get = [&](int r, int c) -> uint8_t { return c < PD_BW ? tileA[r * PD_BW + c] : tileB[r * PD_BW + c - 128]; };
            auto put = [&](int r, int c, uint8_t v) {
                if (c < PD_BW) tileA[r * PD_BW + c] = v;
                if (c >= 128) tileB[r * PD_BW + c - 128] = v;
            };
            for (int i = tid; i < PD_TH * 4; i += PD_THREADS) {
                const int r = i >> 2, e = i & 3, gy = sy0 + r;
                const int gx = e < 2 ? e - 2 : j.sw + (e - 2), c = gx - sx0;
                if (gy < 0 || gy >= j.sh || c < 0 || c >= PD_TW) continue;
                put(r, c, get(r, reflect101(gx, j.sw) - sx0));
            }
            __syncthreads();
            for (int i = tid; i < 4 * (2 * PD_BW / 16); i += PD_THREADS) {
                const int e = i / (2 * PD_BW / 16), k = i - e * (2 * PD_BW / 16);
                const int gy = e < 2 ? e - 2 : j.sh + (e - 2), r = gy - sy0;
                if (r < 0 || r >= PD_TH) continue;
                const int rs = reflect101(gy, j.sh) - sy0;
                uint8_t* base = k < PD_BW / 16 ? tileA : tileB;
                const int off = 16 * (k < PD_BW / 16 ? k : k - PD_BW / 16);
                *reinterpret_cast<uint4*>(base + r * PD_BW + off) = *reinterpret_cast<const uint4*>(base + rs * PD_BW + off);
            }
        }
    } else {
    if (j.sw >= lo + lim) {
        const int gx_hi = al16 ? ((j.sw - 16) & ~15) : (j.sw - 20);
        if (al16) {
            uint4 v[NIT];
#pragma unroll
            for (int it = 0; it < NIT; ++it) {
                const int i = min(tid + it * PD_THREADS, NT - 1);
                const int r = i / NCH, k = i - r * NCH;
                const int gx = min(max(sx0 + 16 * k, 0), gx_hi);
                v[it] = __ldg(reinterpret_cast<const uint4*>(src + (long long)reflect101(sy0 + r, j.sh) * j.spitch + gx));
            }
#pragma unroll
            for (int it = 0; it < NIT; ++it) {
                const int i = tid + it * PD_THREADS;
                const int gx = sx0 + 16 * (i % NCH);
                if (i < NT && gx >= 0 && gx + 16 <= j.sw) st_chunk(i / NCH, i % NCH, v[it]);
            }
        } else {
            unsigned a[NIT][5];
#pragma unroll
            for (int it = 0; it < NIT; ++it) {
                const int i = min(tid + it * PD_THREADS, NT - 1);
                const int r = i / NCH, k = i - r * NCH;
                const int gx = min(max(sx0 + 16 * k, 4), gx_hi);
                const uint8_t* p = src + (long long)reflect101(sy0 + r, j.sh) * j.spitch + gx;
                const unsigned* w = reinterpret_cast<const unsigned*>(p - ((size_t)p & 3));
#pragma unroll
                for (int q = 0; q < 5; ++q) a[it][q] = __ldg(w + q);
            }
#pragma unroll
            for (int it = 0; it < NIT; ++it) {
                const int i = tid + it * PD_THREADS;
                const int r = min(i, NT - 1) / NCH, gx = sx0 + 16 * (i % NCH);
                // rows differ in alignment when the pitch is odd: the shift is the chunk's own
                const unsigned sh = 8u * (unsigned)((size_t)(src + (long long)reflect101(sy0 + r, j.sh) * j.spitch + gx) & 3);
                if (i < NT && gx >= 4 && gx + 20 <= j.sw)
                    st_chunk(i / NCH, i % NCH,
                             make_uint4(__funnelshift_r(a[it][0], a[it][1], sh), __funnelshift_r(a[it][1], a[it][2], sh),
                                        __funnelshift_r(a[it][2], a[it][3], sh), __funnelshift_r(a[it][3], a[it][4], sh)));
            }
        }
    }
    {
        const int rem = j.sw - lim - sx0;
        const int kr = (j.sw >= lo + lim && rem >= 0) ? (rem >> 4) + 1 : 0;   // first chunk right of the inner ones
        for (int i = tid; i < PD_TH * 4; i += PD_THREADS) {
            const int r = i >> 2, e = i & 3;
            const int k = e < 2 ? e : kr + (e - 2);
            const int gx = sx0 + 16 * k;
            if (k >= NCH || gx > j.sw + 1 || (j.sw >= lo + lim && gx >= lo && gx + lim <= j.sw)) continue;   // none / unused / inner
            const uint8_t* row = src + (long long)reflect101(sy0 + r, j.sh) * j.spitch;
            unsigned wv[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int wx = gx + 4 * q;
                if (wx >= 4 && wx + 8 <= j.sw) {   // the unaligned word load touches the aligned words around it
                    wv[q] = ld_word_any(row + wx);
                } else {
                    unsigned t = 0;
#pragma unroll
                    for (int b = 0; b < 4; ++b) t |= (unsigned)__ldg(row + reflect101(wx + b, j.sw)) << (8 * b);
                    wv[q] = t;
                }
            }
            st_chunk(r, k, make_uint4(wv[0], wv[1], wv[2], wv[3]));
        }
    }
    }   // LDG loader
    __syncthreads();
    if (j.copy[cam]) {   // source rows 2*Y0 .. 2*Y0+31, columns 2*X0 .. 2*X0+255 = tile rows 2.., columns 16..
        uint8_t* cp = j.copy[cam] + (long long)seq * j.copy_seq_stride;
        const bool cal16 = (((size_t)cp) & 15) == 0 && (j.cpitch & 15) == 0;
        for (int i = tid; i < 2 * PD_OH * (2 * PD_OW / 16); i += PD_THREADS) {
            const int r = i >> 4, k = i & 15;
            const int gy = 2 * Y0 + r, gx = 2 * X0 + 16 * k;
            if (gy >= j.sh || gx >= j.sw) continue;
            const uint4 v = ld_chunk(r + 2, k + 1);
            uint8_t* o = cp + (long long)gy * j.cpitch + gx;
            if (cal16 && gx + 16 <= j.cpitch) {   // whole chunks: the copy plane's pitch is a multiple of 128
                *reinterpret_cast<uint4*>(o) = v;
            } else {
                const unsigned wv[4] = {v.x, v.y, v.z, v.w};
                for (int e = 0; e < 16; ++e)
                    if (gx + e < j.sw) o[e] = (uint8_t)(wv[e >> 2] >> (8 * (e & 3)));
            }
        }
    }
    // ---- a thread: output columns X0 + 4c .. + 3, output rows Y0 + 4g .. + 3 <- tile rows 8g .. 8g + 10,
    // window bytes b[-4 .. 11] around image column 2 * (X0 + 4c) = tile column 16 + 8c
    const int c = tid & 31, g = tid >> 5;
    unsigned hA[2 * PD_ROWS_PER_THREAD + 3], hB[2 * PD_ROWS_PER_THREAD + 3];
#pragma unroll
    for (int r = 0; r < 2 * PD_ROWS_PER_THREAD + 3; ++r) {
        const uint8_t* t = (c < 16 ? tileA : tileB) + (2 * PD_ROWS_PER_THREAD * g + r) * PD_BW + 16 + 8 * c - (c < 16 ? 0 : 128);
        const unsigned w0 = *reinterpret_cast<const unsigned*>(t - 4);
        const uint2 w12 = *reinterpret_cast<const uint2*>(t);
        const unsigned w3 = *reinterpret_cast<const unsigned*>(t + 8);
        pd_row_sums(w0, w12.x, w12.y, w3, hA[r], hB[r]);
    }
    const int ox = X0 + 4 * c;
    const bool wordok = ox + 3 < j.dw && (j.dpitch & 3) == 0 && (((size_t)dst) & 3) == 0;
#pragma unroll
    for (int y = 0; y < PD_ROWS_PER_THREAD; ++y) {
        const int oy = Y0 + PD_ROWS_PER_THREAD * g + y;
        if (oy >= j.dh || ox >= j.dw) continue;
        const unsigned vA = (hA[2 * y] + hA[2 * y + 4] + 0x00800080u) + (hA[2 * y + 1] + hA[2 * y + 3]) * 4u + hA[2 * y + 2] * 6u;
        const unsigned vB = (hB[2 * y] + hB[2 * y + 4] + 0x00800080u) + (hB[2 * y + 1] + hB[2 * y + 3]) * 4u + hB[2 * y + 2] * 6u;
        const unsigned out = __byte_perm(vA, vB, 0x7531u);   // (v >> 8) of the four halves
        uint8_t* o = dst + (long long)oy * j.dpitch + ox;
        if (wordok) {
            *reinterpret_cast<unsigned*>(o) = out;
        } else {
            for (int q = 0; q < 4; ++q)
                if (ox + q < j.dw) o[q] = (uint8_t)(out >> (8 * q));
        }
    }
}
static void launch_pyrdown_job(const PyrDownJob& j_in, int n_img, cudaStream_t st) {
    PyrDownJob j = j_in;
    const int gx = (j.dw + PD_OW - 1) / PD_OW;
    const auto blocks = [&](int rpt) { return (long long)gx * ((j.dh + 4 * rpt - 1) / (4 * rpt)) * n_img; };
    const long long enough = 4 * 148;   // a few blocks per SM
    const int rpt = blocks(4) >= enough ? 4 : (blocks(2) >= enough ? 2 : 1);
    // tensor maps of the source planes (x, y, sequence), one per camera; box = half a tile row x the tile height
    CUtensorMap map[2];
    memset(map, 0, sizeof(map));
    const int th = 8 * rpt + 3, n_pl = j.n_cam == 2 ? n_img / 2 : n_img;
    j.use_tma = j.sw >= 32 && j.sh >= 8;
    for (int cam = 0; cam < (j.n_cam == 2 ? 2 : 1) && j.use_tma; ++cam)
        j.use_tma = make_tile_map(&map[cam], j.src[cam], j.sw, j.sh, j.spitch, j.src_seq_stride, n_pl, PD_BW, th) ? 1 : 0;
    if (rpt == 4) launch_pdl(pyrdown_tile_kernel<4>, dim3(gx, (j.dh + 15) / 16, n_img), dim3(PD_THREADS), st, j, map[0], map[1]);
    else if (rpt == 2) launch_pdl(pyrdown_tile_kernel<2>, dim3(gx, (j.dh + 7) / 8, n_img), dim3(PD_THREADS), st, j, map[0], map[1]);
    else launch_pdl(pyrdown_tile_kernel<1>, dim3(gx, (j.dh + 3) / 4, n_img), dim3(PD_THREADS), st, j, map[0], map[1]);
}
void launch_pyrdown(const uint8_t* src, long long src_stride, int sw, int sh, int spitch, uint8_t* dst,
                    long long dst_stride, int dw, int dh, int dpitch, int n_img, cudaStream_t st) {
    PyrDownJob j;
    memset(&j, 0, sizeof(j));
    j.src[0] = src; j.src_seq_stride = src_stride; j.sw = sw; j.sh = sh; j.spitch = spitch;
    j.dst[0] = dst; j.dst_seq_stride = dst_stride; j.dw = dw; j.dh = dh; j.dpitch = dpitch;
    j.n_cam = 1;
    launch_pyrdown_job(j, n_img, st);
}

// ---- interleaved multi-channel images (CLI-faithful 8UC3 mode): pyrDown per channel, one thread per
// output sample.  Small images, off the benchmarked path: written for clarity.
struct ImgCn {
    const uint8_t* p;
    int pitch, cn, ch;
    __device__ int operator()(int x, int y) const { return __ldg(p + (long long)y * pitch + x * cn + ch); }
};
__global__ void __launch_bounds__(256) pyrdown_cn_kernel(PyrDownJob j, int cn) {
    const int cam = blockIdx.z & 1, seq = blockIdx.z >> 1;
    const int xb = blockIdx.x * 256 + threadIdx.x, y = blockIdx.y;
    if (xb >= j.dw * cn) return;
    const int x = xb / cn, ch = xb - x * cn;
    ImgCn g{j.src[cam] + (long long)seq * j.src_seq_stride, j.spitch, cn, ch};
    uint8_t* dst = j.dst[cam] + (long long)seq * j.dst_seq_stride;
    dst[(long long)y * j.dpitch + xb] = (uint8_t)pyrdown_pixel(g, j.sw, j.sh, x, y);
}
__global__ void __launch_bounds__(256) copy_rows_kernel(PyrDownJob j, int row_bytes) {
    const int cam = blockIdx.z & 1, seq = blockIdx.z >> 1;
    const int xb = blockIdx.x * 256 + threadIdx.x, y = blockIdx.y;
    if (xb >= row_bytes) return;
    j.copy[cam][(long long)seq * j.copy_seq_stride + (long long)y * j.cpitch + xb] =
        __ldg(j.src[cam] + (long long)seq * j.src_seq_stride + (long long)y * j.spitch + xb);
}
static int launch_ingest_cn(const DevArrays& d, const PyrGeom& pg, const uint8_t* left, const uint8_t* right,
                            int stride, long long seq_stride, int slot, cudaStream_t st) {
    int launches = 0;
    const long long ctx_seq_stride = pg.base(1, 0, 0) - pg.base(0, 0, 0);
    PyrDownJob j;
    memset(&j, 0, sizeof(j));
    if (left) {
        for (int cam = 0; cam < 2; ++cam) {
            j.src[cam] = cam ? right : left;
            j.copy[cam] = d.pyr + pg.base(0, slot, cam) + pg.off[0];
        }
        j.src_seq_stride = seq_stride;
        j.spitch = stride;
        j.copy_seq_stride = ctx_seq_stride;
        j.cpitch = pg.pitch[0];
        const int rb = pg.w[0] * pg.cn;
        copy_rows_kernel<<<dim3((rb + 255) / 256, pg.h[0], d.n_seq * 2), 256, 0, st>>>(j, rb);
        ++launches;
    }
    for (int l = 1; l < pg.nlevels; ++l) {
        memset(&j, 0, sizeof(j));
        for (int cam = 0; cam < 2; ++cam) {
            uint8_t* img = d.pyr + pg.base(0, slot, cam);
            j.src[cam] = img + pg.off[l - 1];
            j.dst[cam] = img + pg.off[l];
        }
        j.src_seq_stride = j.dst_seq_stride = ctx_seq_stride;
        j.sw = pg.w[l - 1]; j.sh = pg.h[l - 1]; j.spitch = pg.pitch[l - 1];
        j.dw = pg.w[l]; j.dh = pg.h[l]; j.dpitch = pg.pitch[l];
        pyrdown_cn_kernel<<<dim3((j.dw * pg.cn + 255) / 256, j.dh, d.n_seq * 2), 256, 0, st>>>(j, pg.cn);
        ++launches;
    }
    return launches;
}
void launch_pyrdown_cn(const uint8_t* src, int sw, int sh, int spitch, uint8_t* dst, int dw, int dh, int dpitch, int cn,
                       cudaStream_t st) {
    PyrDownJob j;
    memset(&j, 0, sizeof(j));
    j.src[0] = src; j.sw = sw; j.sh = sh; j.spitch = spitch;
    j.dst[0] = dst; j.dw = dw; j.dh = dh; j.dpitch = dpitch;
    pyrdown_cn_kernel<<<dim3((dw * cn + 255) / 256, dh, 1), 256, 0, st>>>(j, cn);
}

// ---- context ingest into pyramid slot `slot` (host-known ring, frame f -> f % 3)
int launch_ingest(const DevArrays& d, const PyrGeom& pg, const BucketGeom& bg, const PipeConfig& pc,
                  const uint8_t* left, const uint8_t* right, int stride, long long seq_stride, int slot,
                  cudaStream_t st, bool pdl) {
    PdlScope pdl_scope(pdl);
    int launches = 0;
    const long long ctx_seq_stride = pg.base(1, 0, 0) - pg.base(0, 0, 0);
    if (pg.cn != 1) launches = launch_ingest_cn(d, pg, left, right, stride, seq_stride, slot, st);
    for (int l = 1; l < pg.nlevels && pg.cn == 1; ++l) {
        PyrDownJob j;
        memset(&j, 0, sizeof(j));
        for (int cam = 0; cam < 2; ++cam) {
            uint8_t* img = d.pyr + pg.base(0, slot, cam);
            j.dst[cam] = img + pg.off[l];
            j.src[cam] = img + pg.off[l - 1];
            if (l == 1 && left) {   // images outside the context: read once, level-0 copy + level 1 from one tile
                j.src[cam] = cam ? right : left;
                j.copy[cam] = img + pg.off[0];
            }
        }
        const bool ext = l == 1 && left;
        j.src_seq_stride = ext ? seq_stride : ctx_seq_stride;
        j.spitch = ext ? stride : pg.pitch[l - 1];
        j.sw = pg.w[l - 1]; j.sh = pg.h[l - 1];
        j.dst_seq_stride = ctx_seq_stride;
        j.dw = pg.w[l]; j.dh = pg.h[l]; j.dpitch = pg.pitch[l];
        j.copy_seq_stride = ctx_seq_stride;
        j.cpitch = pg.pitch[0];
        j.n_cam = 2;
        launch_pyrdown_job(j, d.n_seq * 2, st);
        ++launches;
    }
    if (pg.nlevels == 1 && left && pg.cn == 1) {   // no pyramid level to fuse the copy with: every sequence's planes
        PyrDownJob j;
        memset(&j, 0, sizeof(j));
        for (int cam = 0; cam < 2; ++cam) {
            j.src[cam] = cam ? right : left;
            j.copy[cam] = d.pyr + pg.base(0, slot, cam) + pg.off[0];
        }
        j.src_seq_stride = seq_stride;
        j.spitch = stride;
        j.copy_seq_stride = ctx_seq_stride;
        j.cpitch = pg.pitch[0];
        copy_rows_kernel<<<dim3((pg.w[0] + 255) / 256, pg.h[0], d.n_seq * 2), 256, 0, st>>>(j, pg.w[0]);
        ++launches;
    }
    unsigned long long* keysA = d.fastkeys + ((long long)(slot * 2 + 0) * d.n_seq) * bg.nb;
    unsigned long long* keysB = d.fastkeys + ((long long)(slot * 2 + 1) * d.n_seq) * bg.nb;
    // FAST on the left level 0 of the slot (the pitch-aligned copy): the thr_hi winners of every bucket, then the
    // thr_lo winners -- a pass whose blocks return at once unless fewer than pre_matching_threshold buckets
    // have a thr_hi winner (the only case in which bucket_merge_kernel can take the retry and read keysB)
    int* nA = d.fast_nA + (long long)slot * d.n_seq;
    const uint8_t* l0 = d.pyr + pg.base(0, slot, 0) + pg.off[0];
    launch_fast(l0, ctx_seq_stride, pc.width, pc.height, pg.pitch[0], d.n_seq, pc.thr_hi, pc.thr_hi, nullptr, 0, 0,
                keysA, nullptr, bg, st, nA, nullptr, 0, d.fast_work + 2 * slot);
    launch_fast(l0, ctx_seq_stride, pc.width, pc.height, pg.pitch[0], d.n_seq, pc.thr_lo, pc.thr_hi, nullptr, 0, 0,
                nullptr, keysB, bg, st, nullptr, nA, pc.pre_matching_threshold, nullptr);
    return launches + 2;
}

// ---- look-ahead ingest for ONE trajectory: pyramids + FAST of `n_frames` consecutive frames in one
// launch set, into the consecutive ring slots slot0 .. slot0 + n_frames - 1 (the stages at src/vo.cpp:200-201 and
// src/vo.cpp:325 do not depend on the trajectory state).  The frame index takes the place of the sequence index of
// the batched kernels; frame_stride = distance of consecutive source frames.
int launch_ingest_frames(const DevArrays& d, const PyrGeom& pg, const BucketGeom& bg, const PipeConfig& pc,
                         const uint8_t* left, const uint8_t* right, int stride, long long frame_stride, int slot0,
                         int n_frames, cudaStream_t st) {
    int launches = 0;
    const long long slot_stride = pg.base(0, 1, 0) - pg.base(0, 0, 0);
    for (int l = 1; l < pg.nlevels; ++l) {
        PyrDownJob j;
        memset(&j, 0, sizeof(j));
        for (int cam = 0; cam < 2; ++cam) {
            uint8_t* img = d.pyr + pg.base(0, slot0, cam);
            j.dst[cam] = img + pg.off[l];
            j.src[cam] = img + pg.off[l - 1];
            if (l == 1) {   // frames outside the context: read once, level-0 copy + level 1 from one tile
                j.src[cam] = cam ? right : left;
                j.copy[cam] = img + pg.off[0];
            }
        }
        j.src_seq_stride = l == 1 ? frame_stride : slot_stride;
        j.spitch = l == 1 ? stride : pg.pitch[l - 1];
        j.sw = pg.w[l - 1]; j.sh = pg.h[l - 1];
        j.dst_seq_stride = slot_stride;
        j.dw = pg.w[l]; j.dh = pg.h[l]; j.dpitch = pg.pitch[l];
        j.copy_seq_stride = slot_stride;
        j.cpitch = pg.pitch[0];
        j.n_cam = 2;
        launch_pyrdown_job(j, n_frames * 2, st);
        ++launches;
    }
    if (pg.nlevels == 1) {
        PyrDownJob j;
        memset(&j, 0, sizeof(j));
        for (int cam = 0; cam < 2; ++cam) {
            j.src[cam] = cam ? right : left;
            j.copy[cam] = d.pyr + pg.base(0, slot0, cam) + pg.off[0];
        }
        j.src_seq_stride = frame_stride;
        j.spitch = stride;
        j.copy_seq_stride = slot_stride;
        j.cpitch = pg.pitch[0];
        copy_rows_kernel<<<dim3((pg.w[0] + 255) / 256, pg.h[0], n_frames * 2), 256, 0, st>>>(j, pg.w[0]);
        ++launches;
    }
    // FAST keys of slot s: d.fastkeys + (s * 2 + {A, B}) * nb (n_seq == 1), i.e. a stride of 2 nb between frames
    unsigned long long* keysA = d.fastkeys + (long long)(slot0 * 2 + 0) * bg.nb;
    unsigned long long* keysB = d.fastkeys + (long long)(slot0 * 2 + 1) * bg.nb;
    int* nA = d.fast_nA + slot0;
    const uint8_t* l0 = d.pyr + pg.base(0, slot0, 0) + pg.off[0];
    launch_fast(l0, slot_stride, pc.width, pc.height, pg.pitch[0], n_frames, pc.thr_hi, pc.thr_hi, nullptr, 0, 0, keysA,
                nullptr, bg, st, nA, nullptr, 0, d.fast_work + 2 * slot0, 2ll * bg.nb);
    launch_fast(l0, slot_stride, pc.width, pc.height, pg.pitch[0], n_frames, pc.thr_lo, pc.thr_hi, nullptr, 0, 0, nullptr,
                keysB, bg, st, nullptr, nA, pc.pre_matching_threshold, nullptr, 2ll * bg.nb);
    return launches + 2;
}

// ============================================================================
// first frame (vo.cpp:47-56): nothing but a "fail" result carrying last_transform
__global__ void first_frame_kernel(DevArrays d) {
    const int seq = blockIdx.x * blockDim.x + threadIdx.x;
    if (seq >= d.n_seq) return;
    SeqState* s = d.state + seq;
    s->frame_id = 1;
    svo_frame_result* r = d.result + seq;
    r->success = 0;
    r->reason = SVO_REASON_FIRST_FRAME;
    r->n_bucketed = r->n_matched = r->n_inliers = r->retry = r->pnp_ok = r->lk_iters = 0;
    for (int i = 0; i < 9; ++i) r->R[i] = s->R[i];
    for (int i = 0; i < 3; ++i) r->t[i] = s->t[i];
    for (int i = 0; i < 16; ++i) r->transform[i] = s->last_T[i];
}
void launch_first_frame(const DevArrays& d, cudaStream_t st) {
    first_frame_kernel<<<(d.n_seq + 127) / 128, 128, 0, st>>>(d);
}

// ============================================================================
// bucket merge: tracked features + FAST winners -> bucketed list, incl. the retry.  Two kernels:
//   bucket_keys_kernel   every tracked feature atomicMax-es its key into its bucket (grid over the features)
//   bucket_count_kernel  block t counts the winners of buckets [1024 t, 1024 t + 1024), without and with the
//                        low-threshold FAST set (two __syncthreads_count)
//   bucket_merge_kernel  block t owns the same buckets: the 15 tile counts give it the winners before its tile and in
//                        total -- the total decides the low-threshold retry (vo.cpp:327-332); one block scan orders
//                        its own winners, which are written together with their circular-matching result from the
//                        speculative cache (or queued for klt_kernel).  (One kernel in which every block re-counted
//                        all 14720 x 3 keys took 24 us on 16 SMs and 52 us on the 8-SM chain partition -- on the
//                        loop that bounds the streaming rate.)
// The key buffers are cleared by select_kernel (every block here reads all of them).
__global__ void __launch_bounds__(256) bucket_keys_kernel(DevArrays d, BucketGeom bg) {
    const int seq = blockIdx.y;
    SeqState* s = d.state + seq;
    const long long fo = (long long)seq * d.cap;
    unsigned long long* keyT = d.trackkeys + (long long)seq * bg.nb;
    const int n = s->n_feat;
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i == 0) {   // per-frame accumulators of bucket_merge_kernel / klt_kernel
        s->lk_iters = 0;
        d.n_todo[seq] = 0;
        d.klt_next[seq] = 0;
        d.cntT[seq] = 0;   // klt_spec_trk_kernel of the NEXT frame (this frame's has finished)
    }
    if (i >= n) return;
    const float2 p = d.fs_xy[fo + i];
    const int age = d.fs_age[fo + i], str = d.fs_str[fo + i];
    const int br = (int)__fdiv_rn(p.y, (float)bg.bh), bc = (int)__fdiv_rn(p.x, (float)bg.bw);
    if (br >= bg.start_row && br < bg.n_bh && bc >= 0 && bc < bg.n_bw && age < bg.age_threshold)
        atomicMax(keyT + br * bg.n_bw + bc, bucket_key(bucket_score(age, str, bg.fast_threshold), (unsigned)i, str));
}
__global__ void __launch_bounds__(1024) bucket_count_kernel(DevArrays d, BucketGeom bg, int keybuf) {
    const int seq = blockIdx.y, b = (blockIdx.x << 10) + threadIdx.x;
    const unsigned long long* keyT = d.trackkeys + (long long)seq * bg.nb;
    const unsigned long long* keyA = d.fastkeys + ((long long)(keybuf * 2 + 0) * d.n_seq + seq) * bg.nb;
    const unsigned long long* keyB = d.fastkeys + ((long long)(keybuf * 2 + 1) * d.n_seq + seq) * bg.nb;
    bool w1 = false, w2 = false;
    if (b < bg.nb) {
        w1 = (ld_key(keyT + b) | ld_key(keyA + b)) != 0ull;
        w2 = w1 || ld_key(keyB + b) != 0ull;
    }
    const int c1 = __syncthreads_count(w1), c2 = __syncthreads_count(w2);
    if (threadIdx.x == 0) d.tile_cnt[seq * 16 + blockIdx.x] = c1 | (c2 << 16);
}
__global__ void __launch_bounds__(1024, 2)   // 32 registers: two blocks per SM, the 15 blocks of a frame fit the 8-SM chain partition
bucket_merge_kernel(DevArrays d, BucketGeom bg, PipeConfig pc, int keybuf, int specbuf, int use_spec) {
    __shared__ int sm[33];
    __shared__ int sm_iters;
    const int seq = blockIdx.y, tile = blockIdx.x, tid = threadIdx.x, lane = tid & 31;
    SeqState* s = d.state + seq;
    const long long fo = (long long)seq * d.cap;
    const unsigned long long* keyT = d.trackkeys + (long long)seq * bg.nb;
    const unsigned long long* keyA = d.fastkeys + ((long long)(keybuf * 2 + 0) * d.n_seq + seq) * bg.nb;
    const unsigned long long* keyB = d.fastkeys + ((long long)(keybuf * 2 + 1) * d.n_seq + seq) * bg.nb;
    const bool spec_ok = use_spec != 0 && s->spec_bad == 0;
    const bool trk_ok = s->fs_from_inliers != 0;
    int* pred_out = use_spec ? d.predT + ((long long)(specbuf ^ 1) * d.n_seq + seq) * bg.nb : nullptr;
    if (tid == 0) sm_iters = 0;

    // winners without / with the thr/4 FAST set: in total and before this tile (low 16 bits | high 16 bits), from the
    // per-tile counts of bucket_count_kernel
    const int ntiles = (bg.nb + 1023) >> 10;
    int tot_all = 0, pre_all = 0;   // packed: count(w1 != 0) | count(w1 != 0 || kb != 0) << 16
    for (int j = 0; j < ntiles; ++j) {
        const int c = __ldcg(d.tile_cnt + seq * 16 + j);
        tot_all += c;
        if (j < tile) pre_all += c;
    }
    unsigned long long my_w1 = 0ull, my_kb = 0ull;
    bool my_isA = false;
    {
        const int b = (tile << 10) + tid;
        if (b < bg.nb) {
            const unsigned long long kt = ld_key(keyT + b), ka = ld_key(keyA + b);
            my_kb = ld_key(keyB + b);
            my_w1 = kt > ka ? kt : ka;
            my_isA = ka != 0ull && !(kt > ka);   // the winner is this bucket's thr-20 FAST winner (cached)
        }
    }
    const int n1 = tot_all & 0xFFFF;
    const bool retry = n1 < pc.pre_matching_threshold;   // vo.cpp:327-332
    const int total = retry ? (tot_all >> 16) : n1;
    const int base = retry ? (pre_all >> 16) : (pre_all & 0xFFFF);

    unsigned long long wv = my_w1;
    // second append: the survivors come first in the feature order, so a thr/4 FAST feature only wins with a
    // strictly greater score (a bucket with a thr-20 corner has kb == ka, which never replaces)
    if (retry && my_kb != 0ull && (wv == 0ull || bucket_key_score(my_kb) > bucket_key_score(wv))) {
        my_isA = false;
        wv = my_kb;
    }
    int tile_total;
    const int off = base + block_excl_scan<1024>(wv != 0ull ? 1 : 0, &tile_total, sm);
    if (wv != 0ull && off < d.cap) {
        const unsigned order = bucket_key_order(wv);
        float2 p;
        int age, str;
        if (order >= SVO_FAST_ORDER_BASE) {
            const int pix = (int)(order - SVO_FAST_ORDER_BASE);
            const int y = pix / pc.width;
            p = make_float2((float)(pix - y * pc.width), (float)y);
            age = 0;
            str = bucket_key_strength(wv);
        } else {
            p = d.fs_xy[fo + order];
            age = d.fs_age[fo + order];
            str = d.fs_str[fo + order];
        }
        d.bk_xy[fo + off] = p;
        d.bk_age[fo + off] = age;
        d.bk_str[fo + off] = str;
        // circular-matching result from the speculative cache, else queue the feature for klt_kernel
        long long src = -1;
        const float2* sp = nullptr;
        const uint8_t* sst = nullptr;
        const int* sit = nullptr;
        if (spec_ok) {
            if (order >= SVO_FAST_ORDER_BASE) {
                if (my_isA) {
                    src = ((long long)specbuf * d.n_seq + seq) * bg.nb + (tile << 10) + tid;
                    sp = d.specF_pts; sst = d.specF_st; sit = d.specF_it;
                }
            } else if (trk_ok) {
                src = fo + d.m_src[fo + d.inl_idx[fo + order]];   // feature -> matched -> bucketed index of the last frame
                sp = d.specT_pts; sst = d.specT_st; sit = d.specT_it;
            }
        }
        uint8_t cst = 0;
        if (src >= 0) {
            cst = sst[src];
            if (cst & 0x80) src = -1;   // pruned from the speculative pass
        }
        if (src >= 0) {
            const float4 ab = *reinterpret_cast<const float4*>(sp + src * 4);
            const float4 ce = *reinterpret_cast<const float4*>(sp + src * 4 + 2);
            d.lk_L1[fo + off] = make_float2(ab.x, ab.y);
            d.lk_R1[fo + off] = make_float2(ab.z, ab.w);
            d.lk_R0[fo + off] = make_float2(ce.x, ce.y);
            d.lk_C[fo + off] = make_float2(ce.z, ce.w);
            d.lk_status[fo + off] = cst;
            atomicAdd(&sm_iters, sit[src]);
            // prediction for the next frame's FAST pruning: this feature, if kept, is tracked there one frame older
            if (pred_out && (cst & 16) && age + pc.age_increment < bg.age_threshold) {
                const int br = (int)__fdiv_rn(ab.y, (float)bg.bh), bc = (int)__fdiv_rn(ab.x, (float)bg.bw);
                if (br >= bg.start_row && br < bg.n_bh && bc >= 0 && bc < bg.n_bw)
                    atomicMax(pred_out + br * bg.n_bw + bc,
                              bucket_score(age + pc.age_increment, str, bg.fast_threshold) + 0x8000);
            }
        } else {
            d.todo[fo + atomicAdd(d.n_todo + seq, 1)] = off;
        }
    }
    __syncthreads();
    if (tid == 0) {
        if (sm_iters) atomicAdd(&s->lk_iters, sm_iters);
        if (tile == 0) {
            s->n_bucketed = min(total, d.cap);
            s->retry = retry ? 1 : 0;
            s->frame_id += 1;
        }
    }
    (void)lane;
}
void launch_bucket_merge(const DevArrays& d, const BucketGeom& bg, const PipeConfig& pc, int keybuf, int specbuf,
                         int use_spec, cudaStream_t st) {
    bucket_keys_kernel<<<dim3((d.cap + 255) / 256, d.n_seq), 256, 0, st>>>(d, bg);
    bucket_count_kernel<<<dim3((bg.nb + 1023) / 1024, d.n_seq), 1024, 0, st>>>(d, bg, keybuf);
    bucket_merge_kernel<<<dim3((bg.nb + 1023) / 1024, d.n_seq), 1024, 0, st>>>(d, bg, pc, keybuf, specbuf, use_spec);
}

// ============================================================================
// KLT: one warp per feature, the whole circular chain L0->L1->R1->R0->L0 per work item
#ifndef SVO_KLT_WARPS
#define SVO_KLT_WARPS 8
#endif
constexpr int KLT_WARPS = SVO_KLT_WARPS;
#ifndef SVO_KLT_MINB
#define SVO_KLT_MINB 3       // resident blocks per SM the speculative kernels are compiled for (80 registers: measured best)
#endif
#ifndef SVO_KLT_MINB_SER
#define SVO_KLT_MINB_SER 3   // ... and klt_kernel
#endif
__device__ __forceinline__ LkPyramid make_pyr(const uint8_t* base, const PyrGeom& pg) {
    LkPyramid p;
#pragma unroll
    for (int l = 0; l < SVO_MAX_LEVELS; ++l) {
        p.lv[l].img = base + pg.off[l];
        p.lv[l].w = pg.w[l];
        p.lv[l].h = pg.h[l];
        p.lv[l].pitch = pg.pitch[l];
    }
    return p;
}
// one image pyramid of the context as lk_track_feature sees it: the level geometry is read from the kernel parameter
// (constant bank, indexable by the level) instead of a per-leg LkPyramid copy in local memory
struct GeomPyr {
    const uint8_t* base;
    const PyrGeom& pg;
    __device__ __forceinline__ LkLevel level(int l) const {   // selects over constant indices: operands straight from the parameter bank
        long long off = pg.off[0];
        int w = pg.w[0], h = pg.h[0], pitch = pg.pitch[0];
#pragma unroll
        for (int k = 1; k < SVO_MAX_LEVELS; ++k)
            if (l == k) {
                off = pg.off[k];
                w = pg.w[k];
                h = pg.h[k];
                pitch = pg.pitch[k];
            }
        return LkLevel{base + off, w, h, pitch};
    }
};
struct Circ {
    float2 a, b, c, e;   // pL1, pR1, pR0, pL0'
    unsigned st;         // bit l = status of leg l, bit 4 = keep (findClosePoints + status AND + bounds)
    int iters;
};
// The four legs L0->L1->R1->R0->L0 (vo.cpp:203-215) run in ONE loop body to keep the code I-cache sized.
// pyramid k: 0 = L0, 1 = L1, 2 = R1, 3 = R0 ; leg l tracks from pyramid l to pyramid (l + 1) & 3
__device__ __forceinline__ Circ track_circular(const DevArrays& d, const PyrGeom& pg, const PipeConfig& pc,
                                               const long long* base, float2 p0, int lane, LkScratch& sc) {
    float2 pts[5];
    pts[0] = p0;
    int iters = 0;
    unsigned st = 0;
#pragma unroll 1
    for (int leg = 0; leg < 4; ++leg) {
        const GeomPyr I{d.pyr + base[leg], pg}, J{d.pyr + base[(leg + 1) & 3], pg};
        int it;
        float ox, oy;
        const float ix = leg == 0 ? pts[0].x : (leg == 1 ? pts[1].x : (leg == 2 ? pts[2].x : pts[3].x));
        const float iy = leg == 0 ? pts[0].y : (leg == 1 ? pts[1].y : (leg == 2 ? pts[2].y : pts[3].y));
        const int ok = lk_track_feature<32, WarpRed>(I, J, pc.lk, ix, iy, lane, sc, ox, oy, it);
        st |= ok ? (1u << leg) : 0u;
        iters += it;
        if (leg == 0) pts[1] = make_float2(ox, oy);
        else if (leg == 1) pts[2] = make_float2(ox, oy);
        else if (leg == 2) pts[3] = make_float2(ox, oy);
        else pts[4] = make_float2(ox, oy);
    }
    Circ r;
    r.a = pts[1]; r.b = pts[2]; r.c = pts[3]; r.e = pts[4];
    // findClosePoints + status AND (vo.cpp:217-230) + in-bounds test (vo.cpp:341-359): the keep bit
    bool ok = st == 15u && close_points(p0.x, p0.y, r.e.x, r.e.y, pc.circular_threshold);
    if (pc.check_bounds) {
        const int W = pc.width, H = pc.height;
        ok = ok && in_bounds(p0.x, p0.y, W, H) && in_bounds(r.a.x, r.a.y, W, H) && in_bounds(r.c.x, r.c.y, W, H) &&
             in_bounds(r.b.x, r.b.y, W, H);
    }
    r.st = st | (ok ? 16u : 0u);
    r.iters = iters;
    return r;
}
__device__ __forceinline__ void circ_bases(const PyrGeom& pg, int seq, int slot_prev, int slot_cur, long long* base) {
    base[0] = pg.base(seq, slot_prev, 0);
    base[1] = pg.base(seq, slot_cur, 0);
    base[2] = pg.base(seq, slot_cur, 1);
    base[3] = pg.base(seq, slot_prev, 1);
}

// Persistent warps: every warp pulls the next work item from a per-sequence counter, so warps that finish a
// quickly converging feature immediately start another one (the per-feature cost varies 10x).
//
// klt_kernel tracks the bucketed features listed in d.todo (the ones whose result was not found in the
// speculative cache by bucket_merge_kernel; all of them when speculation is off).
#ifdef SVO_KLT_PROBE   // debug build only (tools/klt_probe.py): per work item start / duration (ns) / LK iterations / SM
__device__ int g_klt_probe[4 * 16384];
extern "C" int svo_debug_klt_probe(int* out, int n) {
    return (int)cudaMemcpyFromSymbol(out, g_klt_probe, sizeof(int) * (size_t)n);
}
__device__ __forceinline__ unsigned long long probe_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
#endif
__global__ void __launch_bounds__(KLT_WARPS * 32, SVO_KLT_MINB_SER)
klt_kernel(DevArrays d, PyrGeom pg, PipeConfig pc, int slot_prev, int slot_cur) {
    __shared__ LkScratch patch[KLT_WARPS];
    const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // A warp works through its own sequence's list first and then helps the other sequences of the batch (the
    // sequences differ in feature count and convergence, and a launch lasts as long as its slowest sequence; which
    // warp tracks a feature never changes the result).  Warps never synchronise with each other.
    for (int k = 0; k < d.n_seq; ++k) {
        int seq = (int)blockIdx.y + k;
        if (seq >= d.n_seq) seq -= d.n_seq;
        const int n = d.n_todo[seq];
        int* counter = d.klt_next + seq;
        if (n <= 0 || (k > 0 && __ldcg(counter) >= n)) continue;   // nothing (left) to do there
        SeqState* s = d.state + seq;
        long long base[4];
        circ_bases(pg, seq, slot_prev, slot_cur, base);
        const int* todo = d.todo + (long long)seq * d.cap;
        for (;;) {
            int t = 0;
            if (lane == 0) t = atomicAdd(counter, 1);
            t = __shfl_sync(0xffffffffu, t, 0);
            if (t >= n) break;
            const int f = todo[t];
            const long long o = (long long)seq * d.cap + f;
#ifdef SVO_KLT_PROBE
            const unsigned long long pt0 = probe_ns();
#endif
            const Circ r = track_circular(d, pg, pc, base, d.bk_xy[o], lane, patch[wid]);
#ifdef SVO_KLT_PROBE
            if (lane == 0 && seq == 0 && t < 16384) {
                unsigned smid;
                asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
                g_klt_probe[4 * t + 0] = (int)(unsigned)pt0;
                g_klt_probe[4 * t + 1] = (int)(probe_ns() - pt0);
                g_klt_probe[4 * t + 2] = r.iters;
                g_klt_probe[4 * t + 3] = (int)smid;
            }
#endif
            if (lane == 0) {
                d.lk_L1[o] = r.a;
                d.lk_R1[o] = r.b;
                d.lk_R0[o] = r.c;
                d.lk_C[o] = r.e;
                d.lk_status[o] = (uint8_t)r.st;
                atomicAdd(&s->lk_iters, r.iters);
            }
        }
    }
}
int klt_resident_blocks_per_sm() {
    int nb = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, klt_kernel, KLT_WARPS * 32, 0) != cudaSuccess || nb < 1) nb = 3;
    return nb;
}
int klt_spec_resident_blocks_per_sm();
// interleaved multi-channel pyramids (CLI-faithful 8UC3 mode): same work distribution, generic LK core
__global__ void __launch_bounds__(KLT_WARPS * 32)
klt_cn_kernel(DevArrays d, PyrGeom pg, PipeConfig pc, int slot_prev, int slot_cur) {
    __shared__ LkScratchCn patch[KLT_WARPS];
    const int seq = blockIdx.y;
    const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
    SeqState* s = d.state + seq;
    const int n = d.n_todo[seq];
    if (n <= 0) return;
    long long base[4];
    circ_bases(pg, seq, slot_prev, slot_cur, base);
    int* counter = d.klt_next + seq;
    const int* todo = d.todo + (long long)seq * d.cap;
    for (;;) {
        int t = 0;
        if (lane == 0) t = atomicAdd(counter, 1);
        t = __shfl_sync(0xffffffffu, t, 0);
        if (t >= n) break;
        const int f = todo[t];
        const long long o = (long long)seq * d.cap + f;
        const float2 p0 = d.bk_xy[o];
        float2 pts[5];
        pts[0] = p0;
        int iters = 0;
        unsigned st = 0;
#pragma unroll 1
        for (int leg = 0; leg < 4; ++leg) {
            const LkPyramid I = make_pyr(d.pyr + base[leg], pg);
            const LkPyramid J = make_pyr(d.pyr + base[(leg + 1) & 3], pg);
            int it;
            float ox, oy;
            const int ok = lk_track_feature_cn<32, WarpRed>(I, J, pc.lk, pg.cn, pts[leg].x, pts[leg].y, lane,
                                                            patch[wid], ox, oy, it);
            st |= ok ? (1u << leg) : 0u;
            iters += it;
            pts[leg + 1] = make_float2(ox, oy);
        }
        bool ok = st == 15u && close_points(p0.x, p0.y, pts[4].x, pts[4].y, pc.circular_threshold);
        if (pc.check_bounds) {
            const int W = pc.width, H = pc.height;
            ok = ok && in_bounds(p0.x, p0.y, W, H) && in_bounds(pts[1].x, pts[1].y, W, H) &&
                 in_bounds(pts[3].x, pts[3].y, W, H) && in_bounds(pts[2].x, pts[2].y, W, H);
        }
        if (lane == 0) {
            d.lk_L1[o] = pts[1];
            d.lk_R1[o] = pts[2];
            d.lk_R0[o] = pts[3];
            d.lk_C[o] = pts[4];
            d.lk_status[o] = (uint8_t)(st | (ok ? 16u : 0u));
            atomicAdd(&s->lk_iters, iters);
        }
    }
}
void launch_klt(const DevArrays& d, const PyrGeom& pg, const PipeConfig& pc, int blocks_per_seq, int slot_prev,
                int slot_cur, cudaStream_t st) {
    dim3 grid(blocks_per_seq, d.n_seq);
    if (pg.cn != 1) klt_cn_kernel<<<grid, KLT_WARPS * 32, 0, st>>>(d, pg, pc, slot_prev, slot_cur);
    else klt_kernel<<<grid, KLT_WARPS * 32, 0, st>>>(d, pg, pc, slot_prev, slot_cur);
}

// ---- speculative tracking -------------------------------------------------------------------------------------
// Which features enter circular matching of frame f is only known after the PnP of frame f - 1 (inliers) --
// but every candidate is known earlier and tracking one feature does not depend on any other:
//   F(f): the FAST winner of each bucket of left image f - 1   (known once image f - 1 and f are ingested)
//   T(f): every feature matched in frame f - 1, at its t1 position, young enough (the keep bits of the bucketed list of f - 1)
// Both are tracked ahead of time on their own SM partition while PnP(f - 1) runs; bucket_merge_kernel(f) then
// looks its winners up in this cache.  Anything not cached (low-threshold retry, failed PnP, ...) goes through
// klt_kernel afterwards, so the cache never changes a result, only when it is computed.
// the buckets that hold a thr-20 FAST winner, as an unordered list (one warp-aggregated atomic per 32 buckets)
// The list has two ends: candidates that the pruning prediction of the PREVIOUS frame would not have skipped go
// to the front (they almost certainly have to be tracked), the others to the back -- the warps reach those last,
// by which time this frame's prediction has arrived and most of them are skipped (klt_spec_fast_kernel).
// d.nF[seq] / d.nF[n_seq + seq] count the two ends.  Order only: which candidates are tracked never changes a result.
__global__ void __launch_bounds__(256) spec_fast_list_kernel(DevArrays d, BucketGeom bg, int keybuf, int specbuf) {
    const int seq = blockIdx.y, lane = threadIdx.x & 31;
    const int b = blockIdx.x * 256 + threadIdx.x;
    const unsigned long long* keyA = d.fastkeys + ((long long)(keybuf * 2 + 0) * d.n_seq + seq) * bg.nb;
    const int* pold = d.predT + ((long long)(specbuf ^ 1) * d.n_seq + seq) * bg.nb;
    const int* pnow = d.predT + ((long long)specbuf * d.n_seq + seq) * bg.nb;
    const unsigned long long ka = b < bg.nb ? ld_key(keyA + b) : 0ull;
    const bool has = ka != 0ull;
    const bool late = has && bucket_key_score(ka) + 0x8000 <= max(__ldcg(pold + b), __ldcg(pnow + b));
    const unsigned mf = __ballot_sync(0xffffffffu, has && !late), ml = __ballot_sync(0xffffffffu, late);
    int bf = 0, bl = 0;
    if (lane == 0) {
        if (mf) bf = atomicAdd(d.nF + seq, __popc(mf));
        if (ml) bl = atomicAdd(d.nF + d.n_seq + seq, __popc(ml));
    }
    bf = __shfl_sync(0xffffffffu, bf, 0);
    bl = __shfl_sync(0xffffffffu, bl, 0);
    int* list = d.listF + (long long)seq * bg.nb;
    const unsigned below = (1u << lane) - 1u;
    if (has && !late) list[bf + __popc(mf & below)] = b;
    if (late) list[bg.nb - 1 - (bl + __popc(ml & below))] = b;
}
__global__ void __launch_bounds__(KLT_WARPS * 32, SVO_KLT_MINB)
klt_spec_fast_kernel(DevArrays d, PyrGeom pg, PipeConfig pc, BucketGeom bg, int slot_prev, int slot_cur, int keybuf,
                     int specbuf) {
    __shared__ LkScratch patch[KLT_WARPS];
    const int seq = blockIdx.y;
    const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
    long long base[4];
    circ_bases(pg, seq, slot_prev, slot_cur, base);
    const unsigned long long* keyA = d.fastkeys + ((long long)(keybuf * 2 + 0) * d.n_seq + seq) * bg.nb;
    const long long so = ((long long)specbuf * d.n_seq + seq) * bg.nb;
    const int* list = d.listF + (long long)seq * bg.nb;
    const int* pred = d.predT + ((long long)specbuf * d.n_seq + seq) * bg.nb;
    const int n_front = d.nF[seq], n = n_front + d.nF[d.n_seq + seq];
    int* counter = d.cntF + seq;
    for (;;) {
        int t = 0;
        if (lane == 0) t = atomicAdd(counter, 1);
        t = __shfl_sync(0xffffffffu, t, 0);
        if (t >= n) break;
        const int b = list[t < n_front ? t : bg.nb - 1 - (t - n_front)];
        const unsigned long long ka = ld_key(keyA + b);
        // A FAST winner only enters circular matching when it beats every tracked feature of its bucket (strictly
        // greater score).  bucket_merge_kernel of the previous frame predicts, per bucket, the best score a tracked
        // feature will have there; once that prediction is in (it arrives while this kernel runs), FAST winners
        // that cannot beat it are skipped and flagged "not computed".  A wrong prediction (the tracked feature
        // turns out a PnP outlier) is only a cache miss: the FAST winner then goes through klt_kernel.
        if (bucket_key_score(ka) + 0x8000 <= __ldcg(pred + b)) {
            if (lane == 0) d.specF_st[so + b] = 0x80;
            continue;
        }
        const unsigned order = bucket_key_order(ka);
        const int pix = (int)(order - SVO_FAST_ORDER_BASE);
        const int y = pix / pc.width;
        const float2 p0 = make_float2((float)(pix - y * pc.width), (float)y);
        const Circ r = track_circular(d, pg, pc, base, p0, lane, patch[wid]);
        if (lane == 0) {
            const long long o = so + b;
            d.specF_pts[o * 4 + 0] = r.a;
            d.specF_pts[o * 4 + 1] = r.b;
            d.specF_pts[o * 4 + 2] = r.c;
            d.specF_pts[o * 4 + 3] = r.e;
            d.specF_st[o] = (uint8_t)r.st;
            d.specF_it[o] = r.iters;
        }
    }
}
__global__ void __launch_bounds__(KLT_WARPS * 32, SVO_KLT_MINB)
klt_spec_trk_kernel(DevArrays d, PyrGeom pg, PipeConfig pc, BucketGeom bg, int slot_prev, int slot_cur, int specbuf) {
    __shared__ LkScratch patch[KLT_WARPS];
    const int seq = blockIdx.y;
    const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
    long long base[4];
    circ_bases(pg, seq, slot_prev, slot_cur, base);
    // work list = the bucketed list of the frame that has just been matched: every entry with the keep bit becomes
    // a matched feature (select_kernel only compacts), so this kernel need not wait for the compaction
    const int n = d.state[seq].n_bucketed;
    const long long fo = (long long)seq * d.cap;
    const int* pred = d.predT + ((long long)specbuf * d.n_seq + seq) * bg.nb;
    int* counter = d.cntT + seq;
    for (;;) {
        int k = 0;
        if (lane == 0) k = atomicAdd(counter, 1);
        k = __shfl_sync(0xffffffffu, k, 0);
        if (k >= n) break;
        if (!(d.lk_status[fo + k] & 16)) continue;                                    // not matched
        const int age1 = d.bk_age[fo + k] + pc.age_increment;
        if (age1 >= bg.age_threshold) continue;                                       // bucketing drops it (feature_set.cpp:24)
        const float2 p0 = d.lk_L1[fo + k];
        {
            // bucketing of the coming frame only keeps the best tracked feature of a bucket: a feature that the
            // prediction (best score per bucket, complete by now) beats is not tracked ahead.  If the predicted
            // winner then turns out a PnP outlier, this one is a cache miss and goes through klt_kernel.
            const int br = (int)__fdiv_rn(p0.y, (float)bg.bh), bc = (int)__fdiv_rn(p0.x, (float)bg.bw);
            const bool in_grid = br >= bg.start_row && br < bg.n_bh && bc >= 0 && bc < bg.n_bw;
            if (!in_grid || bucket_score(age1, d.bk_str[fo + k], bg.fast_threshold) + 0x8000 < __ldcg(pred + br * bg.n_bw + bc)) {
                if (lane == 0) d.specT_st[fo + k] = 0x80;
                continue;
            }
        }
        const Circ r = track_circular(d, pg, pc, base, p0, lane, patch[wid]);
        if (lane == 0) {
            const long long o = fo + k;
            d.specT_pts[o * 4 + 0] = r.a;
            d.specT_pts[o * 4 + 1] = r.b;
            d.specT_pts[o * 4 + 2] = r.c;
            d.specT_pts[o * 4 + 3] = r.e;
            d.specT_st[o] = (uint8_t)r.st;
            d.specT_it[o] = r.iters;
        }
    }
}
int klt_spec_resident_blocks_per_sm() {
    int nb = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, klt_spec_trk_kernel, KLT_WARPS * 32, 0) != cudaSuccess || nb < 1) nb = 3;
    return nb;
}
void launch_klt_spec_fast(const DevArrays& d, const PyrGeom& pg, const PipeConfig& pc, const BucketGeom& bg,
                          int blocks_per_seq, int slot_prev, int slot_cur, int keybuf, int specbuf, cudaStream_t st) {
    cudaMemsetAsync(d.cntF, 0, sizeof(int) * d.n_seq, st);
    cudaMemsetAsync(d.nF, 0, sizeof(int) * 2 * d.n_seq, st);
    spec_fast_list_kernel<<<dim3((bg.nb + 255) / 256, d.n_seq), 256, 0, st>>>(d, bg, keybuf, specbuf);
    dim3 grid(blocks_per_seq, d.n_seq);
    klt_spec_fast_kernel<<<grid, KLT_WARPS * 32, 0, st>>>(d, pg, pc, bg, slot_prev, slot_cur, keybuf, specbuf);
}
void launch_klt_spec_trk(const DevArrays& d, const PyrGeom& pg, const PipeConfig& pc, const BucketGeom& bg,
                         int blocks_per_seq, int slot_prev, int slot_cur, int specbuf, cudaStream_t st) {
    dim3 grid(blocks_per_seq, d.n_seq);   // d.cntT was reset by bucket_keys_kernel
    klt_spec_trk_kernel<<<grid, KLT_WARPS * 32, 0, st>>>(d, pg, pc, bg, slot_prev, slot_cur, specbuf);
}

__global__ void __launch_bounds__(KLT_WARPS * 32)
lk_single_kernel(LkPyramid I, LkPyramid J, LkConfig cfg, const float2* __restrict__ pts, int n,
                 float2* __restrict__ out, uint8_t* __restrict__ status) {
    __shared__ LkScratch patch[KLT_WARPS];
    const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int f = blockIdx.x * KLT_WARPS + wid;
    if (f >= n) return;
    const float2 p = pts[f];
    float2 q;
    int it;
    const int ok = lk_track_feature<32, WarpRed>(I, J, cfg, p.x, p.y, lane, patch[wid], q.x, q.y, it);
    if (lane == 0) {
        out[f] = q;
        status[f] = (uint8_t)ok;
    }
}
__global__ void __launch_bounds__(KLT_WARPS * 32)
lk_single_cn_kernel(LkPyramid I, LkPyramid J, LkConfig cfg, int cn, const float2* __restrict__ pts, int n,
                    float2* __restrict__ out, uint8_t* __restrict__ status) {
    __shared__ LkScratchCn patch[KLT_WARPS];
    const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int f = blockIdx.x * KLT_WARPS + wid;
    if (f >= n) return;
    const float2 p = pts[f];
    float2 q;
    int it;
    const int ok = lk_track_feature_cn<32, WarpRed>(I, J, cfg, cn, p.x, p.y, lane, patch[wid], q.x, q.y, it);
    if (lane == 0) {
        out[f] = q;
        status[f] = (uint8_t)ok;
    }
}
void launch_lk_single_cn(const LkPyramid& I, const LkPyramid& J, LkConfig cfg, int cn, const float2* pts, int n,
                         float2* out, uint8_t* status, cudaStream_t st) {
    if (n <= 0) return;
    lk_single_cn_kernel<<<(n + KLT_WARPS - 1) / KLT_WARPS, KLT_WARPS * 32, 0, st>>>(I, J, cfg, cn, pts, n, out, status);
}
void launch_lk_single(const LkPyramid& I, const LkPyramid& J, LkConfig cfg, const float2* pts, int n, float2* out,
                      uint8_t* status, cudaStream_t st) {
    if (n <= 0) return;
    lk_single_kernel<<<(n + KLT_WARPS - 1) / KLT_WARPS, KLT_WARPS * 32, 0, st>>>(I, J, cfg, pts, n, out, status);
}

// ============================================================================
// select: keep mask -> stable compaction, ages++, triangulation, RANSAC sample indices
//
// OpenCV's getSubset() redraws on duplicates, so hypothesis h starts at RNG position 5h + (redraws so far).
// Lane l replays hypotheses 4l..4l+3 sequentially from a speculated start position; the per-lane draw totals
// are prefix-summed and the replay repeats until the start positions are stable (one redraw event costs one
// extra round; with thousands of points there is usually none).  Result == the sequential replay.
__device__ __forceinline__ int lane_replay_samples(const int* __restrict__ mods, int pos, int h0, int h1,
                                                   int* __restrict__ samples) {
    for (int h = h0; h < h1; ++h) {
        int idx[5] = {-1, -1, -1, -1, -1};
        bool ok = true;
        for (int i = 0; i < 5 && ok; ++i) {
            for (;;) {
                if (pos >= SVO_RNG_TABLE) { ok = false; break; }
                const int v = mods[pos++];
                bool dup = false;
                for (int j = 0; j < i; ++j) dup |= (idx[j] == v);
                if (!dup) { idx[i] = v; break; }
            }
        }
        if (samples)
            for (int i = 0; i < 5; ++i) samples[h * 5 + i] = ok ? idx[i] : -1;
    }
    return pos;
}
__device__ __forceinline__ void warp_build_samples(const unsigned* __restrict__ raw, int n, int n_hyp,
                                                   int* __restrict__ samples, int* mods /* smem [SVO_RNG_TABLE] */) {
    const int lane = threadIdx.x & 31;
    for (int i = lane; i < SVO_RNG_TABLE; i += 32) mods[i] = (int)(raw[i] % (unsigned)n);
    __syncwarp();
    const int per = (n_hyp + 31) / 32;
    const int h0 = min(n_hyp, lane * per), h1 = min(n_hyp, h0 + per);
    int start = 5 * h0;
    for (int round = 0; round <= 32; ++round) {
        const int used = lane_replay_samples(mods, start, h0, h1, nullptr) - start;
        int incl = used;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        const int ns = incl - used;
        const bool changed = ns != start;
        start = ns;
        if (!__any_sync(0xffffffffu, changed)) break;
    }
    lane_replay_samples(mods, start, h0, h1, samples);
}
__global__ void pnp_samples_kernel(DevArrays d) {
    __shared__ int mods[SVO_RNG_TABLE];
    const int seq = blockIdx.x;
    const int n = d.state[seq].n_matched;
    if (n >= 5) warp_build_samples(d.rng_raw, n, d.n_hyp, d.hyp_samples + (long long)seq * d.n_hyp * 5, mods);
}
void launch_pnp_samples(const DevArrays& d, cudaStream_t st) { pnp_samples_kernel<<<d.n_seq, 32, 0, st>>>(d); }

// grid (tiles + 1, n_seq): block b < tiles compacts features [256 b, 256 b + 256) -- its output offset is the
// number of kept features before the tile, counted redundantly from the status bytes (<= 14 KB, L2 hits) so no
// inter-block dependency exists; the last block only counts, builds the RANSAC samples and updates the state.
constexpr int SEL_T = 256;
__global__ void __launch_bounds__(SEL_T) select_kernel(DevArrays d, PyrGeom pg, PipeConfig pc, int slot_prev,
                                                       int slot_cur, int nb, int keybuf, int predbuf) {
    __shared__ int sm[SEL_T / 32 + 1];
    __shared__ int mods[SVO_RNG_TABLE];
    const int seq = blockIdx.y, tid = threadIdx.x;
    SeqState* s = d.state + seq;
    const long long fo = (long long)seq * d.cap;
    const int n = s->n_bucketed;
    const bool tail = blockIdx.x == gridDim.x - 1;
    const int t0 = tail ? n : blockIdx.x * SEL_T;
    if (predbuf >= 0) {   // FAST-pruning predictions of two frames ahead start from "none"
        int* pr = d.predT + ((long long)predbuf * d.n_seq + seq) * nb;
        for (int i = blockIdx.x * SEL_T + tid; i < nb; i += gridDim.x * SEL_T) pr[i] = 0;
    }
    if (keybuf >= 0) {
        // the bucket keys of this frame are consumed (every block of bucket_merge_kernel read all of them):
        // ready for the next frame / the next FAST pass into this ring buffer
        unsigned long long* kT = d.trackkeys + (long long)seq * nb;
        unsigned long long* kA = d.fastkeys + ((long long)(keybuf * 2 + 0) * d.n_seq + seq) * nb;
        unsigned long long* kB = d.fastkeys + ((long long)(keybuf * 2 + 1) * d.n_seq + seq) * nb;
        for (int i = blockIdx.x * SEL_T + tid; i < nb; i += gridDim.x * SEL_T) {
            kT[i] = 0ull;
            kA[i] = 0ull;
            kB[i] = 0ull;
        }
        if (blockIdx.x == 0 && tid == 0) d.fast_nA[(long long)keybuf * d.n_seq + seq] = 0;
    }
    if (n == 0) {
        // circularMatching returns before lastLeft/RightPyramid are replaced (vo.cpp:179-181): keep the old
        // pyramids as "previous" by copying them into this frame's ring slot
        const uint4* src = reinterpret_cast<const uint4*>(d.pyr + pg.base(seq, slot_prev, 0));
        uint4* dst = reinterpret_cast<uint4*>(d.pyr + pg.base(seq, slot_cur, 0));
        const long long nv = pg.image_bytes * 2 / 16;
        if (slot_prev != slot_cur)
            for (long long i = (long long)blockIdx.x * SEL_T + tid; i < nv; i += (long long)gridDim.x * SEL_T) dst[i] = src[i];
    }
    if (!tail && t0 >= n) return;
    // kept features before the tile
    const uint8_t* stt = d.lk_status + fo;
    int c = 0;
    if ((fo & 3) == 0) {
        const unsigned* w = reinterpret_cast<const unsigned*>(stt);
        const int nw = t0 >> 2;
        for (int i = tid; i < nw; i += SEL_T) c += __popc(w[i] & 0x10101010u);
        for (int i = (nw << 2) + tid; i < t0; i += SEL_T) c += (stt[i] >> 4) & 1;
    } else {
        for (int i = tid; i < t0; i += SEL_T) c += (stt[i] >> 4) & 1;
    }
    const int i = t0 + tid;
    const bool ok = !tail && i < n && (stt[i] & 16) != 0;
    // one scan gives both: low 16 bits = own flag, high bits = the count (<= 2^15 features)
    int packed_total;
    const int packed = block_excl_scan<SEL_T>((ok ? 1 : 0) | (c << 16), &packed_total, sm);
    const int base = packed_total >> 16;
    const int off = base + (packed & 0xFFFF);
    if (ok) {
        const float2 p0 = d.bk_xy[fo + i];
        const int age = d.bk_age[fo + i] + pc.age_increment, str = d.bk_str[fo + i];   // ages++ (vo.cpp:70-72)
        d.m_L0[fo + off] = p0;
        d.m_L1[fo + off] = d.lk_L1[fo + i];
        d.m_R1[fo + off] = d.lk_R1[fo + i];
        d.m_age[fo + off] = age;
        d.m_str[fo + off] = str;
        d.m_src[fo + off] = i;
        // currentVOFeatures after matchingFeatures: t0 coordinates (vo.cpp:360)
        d.fs_xy[fo + off] = p0;
        d.fs_age[fo + off] = age;
        d.fs_str[fo + off] = str;
        const float2 r0 = d.lk_R0[fo + i];
        d.m_R0[fo + off] = r0;
        triangulate_dlt(pc.Pl, pc.Pr, p0.x, p0.y, r0.x, r0.y, d.X3 + (fo + off) * 3);   // vo.cpp:89-94
    }
    if (tail) {
        const int total = base;
        // RANSAC sample indices in OpenCV's order (they depend only on the survivor count)
        if (tid < 32 && total >= 5)
            warp_build_samples(d.rng_raw, total, d.n_hyp, d.hyp_samples + (long long)seq * d.n_hyp * 5, mods);
        if (tid == 0) {
            s->n_matched = total;
            s->n_feat = total;
            s->n_inliers = 0;
            s->pnp_ok = 0;
            s->spec_bad = n == 0 ? 1 : 0;   // the old pyramids stay: what was tracked ahead for the next frame is void
            s->fs_from_inliers = 0;         // the feature set now holds the matched set at t0 positions
        }
    }
}
void launch_select(const DevArrays& d, const PyrGeom& pg, const PipeConfig& pc, int slot_prev, int slot_cur,
                   int keybuf, int predbuf, cudaStream_t st) {
    dim3 grid((d.cap + SEL_T - 1) / SEL_T + 1, d.n_seq);
    select_kernel<<<grid, SEL_T, 0, st>>>(d, pg, pc, slot_prev, slot_cur, d.nb, keybuf, predbuf);
}

// ============================================================================
// PnP-RANSAC: parallel 5-point hypotheses (LM from the extrinsic guess, like
// solvePnP(ITERATIVE, useExtrinsicGuess) inside OpenCV's RANSAC loop), inlier counting
// by warp reduction, then an FP64 LM refit on the inliers of the best hypothesis.
// One block of three warps per RANSAC hypothesis: EPnP on the 5-point sample (what cv::solvePnPRansac runs per
// iteration; svo_epnp.cuh restates it operation for operation, so its time is one long dependent FP64 chain --
// the three beta approximations run side by side in the three warps), then the inlier count over all matched
// points by the whole block.  OpenCV's loop is adaptive: with >= 75 % inliers it never looks beyond the first 16
// hypotheses, so the launch is split -- hypotheses [0, HYP_FIRST) always, the rest only when the replay of the
// first counts says the loop goes on (checked on the device by every block of the second launch).
constexpr int HYP_THREADS = 96, HYP_FIRST = 16;

struct WarpRedD {
    static __device__ __forceinline__ void sync() { __syncwarp(); }
    static __device__ __forceinline__ void sync_all() { __syncthreads(); }
    static __device__ __forceinline__ bool any(bool v) { return __any_sync(0xffffffffu, v); }
    static __device__ __forceinline__ double sumd(double v) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        return v;
    }
};

// MINB: resident blocks per SM the kernel is compiled for.  1 (255 registers) is the fastest single hypothesis and what a
// single sequence gets; a batch of sequences launches 16 x n_seq blocks and is wave bound, so it takes the 168-register
// build (4 per SM: 64 sequences 709 -> 421 us per frame, one sequence 146 -> 148).
template <int MINB>
__global__ void __launch_bounds__(HYP_THREADS, MINB) pnp_hyp_kernel(DevArrays d, PipeConfig pc, int hyp0) {
    __shared__ EpnpWork work;
    __shared__ int sm_cnt[HYP_THREADS / 32];
    const int seq = blockIdx.y;
    const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int hyp = hyp0 + blockIdx.x;
    SeqState* s = d.state + seq;
    const int n = s->n_matched;
    const int min_n = max(4, pc.features_threshold);
    if (hyp >= d.n_hyp || n <= min_n) return;   // vo.cpp:82-84
    if (hyp0 > 0) {
        // the sequential loop only reaches this hypothesis if its iteration budget, after the first hyp0 counts,
        // is still larger than hyp0
        int bc;
        if (ransac_replay_budget(d.hyp_count + (long long)seq * d.n_hyp, hyp0, n, pc.ransac_iterations,
                                 pc.ransac_confidence, bc) <= hyp0)
            return;
    }
    const long long fo = (long long)seq * d.cap;
    const float* X3 = d.X3 + fo * 3;
    const float2* uv = d.m_L1 + fo;

    const int* smp = d.hyp_samples + ((long long)seq * d.n_hyp + hyp) * 5;
    float X[15], U[10];
    bool valid = true;
#pragma unroll
    for (int i = 0; i < 5; ++i) {
        int k = smp[i];
        valid = valid && k >= 0 && k < n;
        k = min(max(k, 0), n - 1);
        X[3 * i] = X3[3 * k];
        X[3 * i + 1] = X3[3 * k + 1];
        X[3 * i + 2] = X3[3 * k + 2];
        const float2 q = uv[k];
        U[2 * i] = q.x;
        U[2 * i + 1] = q.y;
    }
    Pose P;
    const bool solved = epnp5_solve<32, 3, WarpRedD>(work, pc.K, X, U, lane, wid, P);
    int cnt = 0;
    if (solved && valid) {
        for (int i = threadIdx.x; i < n; i += HYP_THREADS) {
            const float2 q = uv[i];
            const float uvf[2] = {q.x, q.y};
            const float e = reproj_err_sq(P, pc.K, X3 + 3 * i, uvf);
            cnt += (e <= pc.reproj_thr_sq) ? 1 : 0;
        }
    }
    cnt = __reduce_add_sync(0xffffffffu, cnt);
    if (lane == 0) sm_cnt[wid] = cnt;
    __syncthreads();
    if (threadIdx.x == 0) {
        double* hp = d.hyp_pose + ((long long)seq * d.n_hyp + hyp) * 12;
        for (int i = 0; i < 9; ++i) hp[i] = P.R[i];
        for (int i = 0; i < 3; ++i) hp[9 + i] = P.t[i];
        d.hyp_count[(long long)seq * d.n_hyp + hyp] = (solved && valid) ? sm_cnt[0] + sm_cnt[1] + sm_cnt[2] : 0;
    }
}

constexpr int FIN_THREADS = 512;
struct BlockInlierEval {
    const float* X3;
    const float2* uv;
    const int* idx;
    int n;
    Intrinsics K;
    double* sm;   // [FIN_THREADS/32][SVO_PNP_ACC]
    __device__ void operator()(const Pose& P, double* acc, bool jac) const {
#pragma unroll
        for (int i = 0; i < SVO_PNP_ACC; ++i) acc[i] = 0.0;
#pragma unroll 2
        for (int i = threadIdx.x; i < n; i += FIN_THREADS) {   // compact inlier arrays: coalesced, no indirection
            const float2 q = uv[i];
            const float uvf[2] = {q.x, q.y};
            pnp_accumulate(P, K, X3 + 3 * i, uvf, acc, jac);
        }
        const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
        // warp sums of the 28 accumulators by recursive halving: at every step a lane keeps one half of its vector and
        // trades the other half with its partner (28 -> 14 -> 7 -> 4 -> 2 -> 1 values: 28 exchanges instead of the
        // 28 x 5 of one butterfly per value; SHFL issues once per 4 cycles, it was 60 % of this kernel).  Lane l ends
        // with the warp total of accumulator halving_slot(l) in acc[0].
        static_assert(SVO_PNP_ACC == 28, "the halving schedule below is written for 28 accumulators");
        {
            const bool up4 = lane & 16, up3 = lane & 8, up2 = lane & 4, up1 = lane & 2, up0 = lane & 1;
#pragma unroll
            for (int i = 0; i < 14; ++i) {
                const double send = up4 ? acc[i] : acc[i + 14], keep = up4 ? acc[i + 14] : acc[i];
                acc[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
            }
#pragma unroll
            for (int i = 0; i < 7; ++i) {
                const double send = up3 ? acc[i] : acc[i + 7], keep = up3 ? acc[i + 7] : acc[i];
                acc[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
            }
            acc[7] = 0.0;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const double send = up2 ? acc[i] : acc[i + 4], keep = up2 ? acc[i + 4] : acc[i];
                acc[i] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
            }
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                const double send = up1 ? acc[i] : acc[i + 2], keep = up1 ? acc[i + 2] : acc[i];
                acc[i] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
            }
            {
                const double send = up0 ? acc[0] : acc[1], keep = up0 ? acc[1] : acc[0];
                acc[0] = keep + __shfl_xor_sync(0xffffffffu, send, 1);
            }
        }
        const int sub = ((lane & 4) ? 4 : 0) + ((lane & 2) ? 2 : 0) + (lane & 1);   // 7 = the padding value of a group of 7
        const int slot = ((lane & 16) ? 14 : 0) + ((lane & 8) ? 7 : 0) + sub;
        __syncthreads();
        if (sub < 7) sm[wid * SVO_PNP_ACC + slot] = acc[0];
        __syncthreads();
        if (threadIdx.x < SVO_PNP_ACC) {   // fixed summation order -> every thread sees identical totals
            double v = 0.0;
            for (int w = 0; w < FIN_THREADS / 32; ++w) v += sm[w * SVO_PNP_ACC + threadIdx.x];
            sm[(FIN_THREADS / 32) * SVO_PNP_ACC + threadIdx.x] = v;
        }
        __syncthreads();
#pragma unroll
        for (int i = 0; i < SVO_PNP_ACC; ++i) acc[i] = sm[(FIN_THREADS / 32) * SVO_PNP_ACC + i];
    }
};

__device__ void write_result(const DevArrays& d, int seq, const FinalJob& j, int success, int reason, int pnp_ok,
                             int n_inliers, const double* T) {
    SeqState* s = d.state + seq;
    svo_frame_result* r = d.result + seq;
    r->success = success;
    r->reason = reason;
    r->n_bucketed = j.n_bucketed;
    r->n_matched = j.n_matched;
    r->n_inliers = n_inliers;
    r->retry = j.retry;
    r->pnp_ok = pnp_ok;
    r->lk_iters = j.lk_iters;
    for (int i = 0; i < 9; ++i) r->R[i] = s->R[i];
    for (int i = 0; i < 3; ++i) r->t[i] = s->t[i];
    for (int i = 0; i < 16; ++i) r->transform[i] = T[i];
}

// cameraToWorld is split in two kernels.  pnp_final_a_kernel does what the NEXT frame depends on: the replay of
// OpenCV's sequential RANSAC loop over the hypothesis scores, the inlier list of the winner and the feature-set
// update (vo.cpp:115-121).  pnp_final_b_kernel refits the pose on the inliers (LM, FP64), applies the gates and
// writes the frame result -- it only feeds the pose output, so it runs beside the next frame's chain.
__global__ void __launch_bounds__(FIN_THREADS) pnp_final_a_kernel(DevArrays d, PipeConfig pc) {
    __shared__ int sm_i[33];
    const int seq = blockIdx.x, tid = threadIdx.x;
    SeqState* s = d.state + seq;
    FinalJob* job = d.job + seq;
    const long long fo = (long long)seq * d.cap;
    const int n = s->n_matched;
    const int min_n = max(4, pc.features_threshold);
    if (tid == 0) {
        job->n_bucketed = s->n_bucketed;
        job->n_matched = n;
        job->retry = s->retry;
        job->lk_iters = s->lk_iters;
        job->n_in = 0;
        job->enough = 0;
    }
    if (n <= min_n) {   // vo.cpp:82-84
        if (tid == 0) job->action = SVO_REASON_FEW_FEATURES;
        return;
    }
    // ---- winning hypothesis: replay of OpenCV's sequential adaptive RANSAC loop over the counts
    if (tid == 0) {
        int bc = 0;
        sm_i[0] = ransac_replay(d.hyp_count + (long long)seq * d.n_hyp, d.n_hyp, n, pc.ransac_iterations,
                                pc.ransac_confidence, bc);
        sm_i[1] = bc;
    }
    __syncthreads();
    const int best_h = sm_i[0], best_cnt = sm_i[1];
    __syncthreads();
    if (best_h < 0 || best_cnt < 5) {   // no model: solvePnPRansac returns false, translation restored (vo.cpp:309-311)
        if (tid == 0) {
            s->pnp_ok = 0;
            s->n_inliers = 0;
            job->action = SVO_REASON_PNP_FAILED;
        }
        return;
    }
    Pose P;
    {
        const double* hp = d.hyp_pose + ((long long)seq * d.n_hyp + best_h) * 12;
        for (int i = 0; i < 9; ++i) P.R[i] = hp[i];
        for (int i = 0; i < 3; ++i) P.t[i] = hp[9 + i];
    }
    // ---- inlier mask of the best model -> ascending index list
    const float* X3 = d.X3 + fo * 3;
    const float2* uv = d.m_L1 + fo;
    int* idx = d.inl_idx + fo;
    int n_in = 0;
    for (int t0 = 0; t0 < n; t0 += FIN_THREADS) {   // coalesced tiles, ascending order
        const int i = t0 + tid;
        bool in = false;
        float2 q = make_float2(0.f, 0.f);
        if (i < n) {
            q = uv[i];
            const float uvf[2] = {q.x, q.y};
            in = reproj_err_sq(P, pc.K, X3 + 3 * i, uvf) <= pc.reproj_thr_sq;
        }
        int tile_total;
        const int off = n_in + block_excl_scan<FIN_THREADS>(in ? 1 : 0, &tile_total, sm_i);
        if (in) {
            idx[off] = i;
            d.in_uv[fo + off] = q;
            d.X3u[(fo + off) * 3] = X3[3 * i];
            d.X3u[(fo + off) * 3 + 1] = X3[3 * i + 1];
            d.X3u[(fo + off) * 3 + 2] = X3[3 * i + 2];
        }
        n_in += tile_total;
    }
    __syncthreads();
    const bool enough = n_in >= pc.features_threshold;
    if (enough) {
        // currentVOFeatures.points = pointsLeftT1; keep inliers (vo.cpp:115-121).  fs_* holds the matched set
        // (t0 coordinates); entry i of the new set comes from matched entry idx[i] >= i, read from the m_* copies.
        for (int base = 0; base < n_in; base += FIN_THREADS) {
            const int i = base + tid;
            if (i < n_in) {
                const int k = idx[i];
                d.fs_xy[fo + i] = d.m_L1[fo + k];
                d.fs_age[fo + i] = d.m_age[fo + k];
                d.fs_str[fo + i] = d.m_str[fo + k];
            }
        }
    }
    if (tid == 0) {
        s->pnp_ok = 1;
        s->n_inliers = n_in;
        if (enough) {
            s->n_feat = n_in;
            s->fs_from_inliers = 1;
        }
        job->action = 0;
        job->n_in = n_in;
        job->enough = enough ? 1 : 0;
        for (int i = 0; i < 9; ++i) job->R[i] = P.R[i];
        for (int i = 0; i < 3; ++i) job->t[i] = P.t[i];
    }
}

__global__ void __launch_bounds__(FIN_THREADS) pnp_final_b_kernel(DevArrays d, PipeConfig pc) {
    __shared__ double sm_acc[(FIN_THREADS / 32 + 1) * SVO_PNP_ACC];
    const int seq = blockIdx.x, tid = threadIdx.x;
    SeqState* s = d.state + seq;
    const FinalJob job = d.job[seq];
    const long long fo = (long long)seq * d.cap;
    if (job.action != 0) {
        if (tid == 0) write_result(d, seq, job, 0, job.action, 0, 0, s->last_T);
        return;
    }
    Pose P;
    for (int i = 0; i < 9; ++i) P.R[i] = job.R[i];
    for (int i = 0; i < 3; ++i) P.t[i] = job.t[i];
    const int n_in = job.n_in;
    // ---- LM refit on all inliers (uniform control flow across the block)
    BlockInlierEval ev{d.X3u + fo * 3, d.in_uv + fo, d.inl_idx + fo, n_in, pc.K, sm_acc};
    pnp_refine_lm(P, ev, 20);
    bool finite = true;
    for (int i = 0; i < 9; ++i) finite &= isfinite(P.R[i]);
    for (int i = 0; i < 3; ++i) finite &= isfinite(P.t[i]);
    if (tid != 0) return;
    if (!finite) {   // cannot happen with >= 5 consistent points; reported as a PnP failure, pose state untouched
        s->pnp_ok = 0;
        s->n_inliers = 0;
        write_result(d, seq, job, 0, SVO_REASON_PNP_FAILED, 0, 0, s->last_T);
        return;
    }
    for (int i = 0; i < 9; ++i) s->R[i] = P.R[i];   // rotation / translation overwritten on PnP success
    for (int i = 0; i < 3; ++i) s->t[i] = P.t[i];
    if (!job.enough) {   // vo.cpp:106-113
        write_result(d, seq, job, 0, SVO_REASON_PNP_FAILED, 1, n_in, s->last_T);
        return;
    }
    const double tn = sqrt(P.t[0] * P.t[0] + P.t[1] * P.t[1] + P.t[2] * P.t[2]);
    const double ang = rotation_angle(P.R);
    if (tn > pc.max_t_norm || ang > pc.max_r_norm) {   // vo.cpp:123-132
        write_result(d, seq, job, 0, SVO_REASON_TOO_BIG, 1, n_in, s->last_T);
        return;
    }
    // getInverseTransform (vo.cpp:246-258): inverse of the rigid [R|t]
    double T[16];
    for (int i = 0; i < 3; ++i) {
        for (int j = 0; j < 3; ++j) T[i * 4 + j] = P.R[j * 3 + i];
        T[i * 4 + 3] = -(P.R[0 * 3 + i] * P.t[0] + P.R[1 * 3 + i] * P.t[1] + P.R[2 * 3 + i] * P.t[2]);
    }
    T[12] = T[13] = T[14] = 0.0;
    T[15] = 1.0;
    for (int i = 0; i < 16; ++i) s->last_T[i] = T[i];
    write_result(d, seq, job, 1, SVO_REASON_OK, 1, n_in, T);
}
void launch_pnp_hyp(const DevArrays& d, const PipeConfig& pc, cudaStream_t st) {
    const int first = d.n_hyp < HYP_FIRST ? d.n_hyp : HYP_FIRST;
    if (d.n_seq > 1) {
        pnp_hyp_kernel<4><<<dim3(first, d.n_seq), HYP_THREADS, 0, st>>>(d, pc, 0);
        if (d.n_hyp > first) pnp_hyp_kernel<4><<<dim3(d.n_hyp - first, d.n_seq), HYP_THREADS, 0, st>>>(d, pc, first);
    } else {
        pnp_hyp_kernel<1><<<dim3(first, d.n_seq), HYP_THREADS, 0, st>>>(d, pc, 0);
        if (d.n_hyp > first) pnp_hyp_kernel<1><<<dim3(d.n_hyp - first, d.n_seq), HYP_THREADS, 0, st>>>(d, pc, first);
    }
}
void launch_pnp_final_a(const DevArrays& d, const PipeConfig& pc, cudaStream_t st) {
    pnp_final_a_kernel<<<d.n_seq, FIN_THREADS, 0, st>>>(d, pc);
}
void launch_pnp_final_b(const DevArrays& d, const PipeConfig& pc, cudaStream_t st) {
    pnp_final_b_kernel<<<d.n_seq, FIN_THREADS, 0, st>>>(d, pc);
}
void launch_pnp_final(const DevArrays& d, const PipeConfig& pc, cudaStream_t st) {
    launch_pnp_final_a(d, pc, st);
    launch_pnp_final_b(d, pc, st);
}
void launch_pnp(const DevArrays& d, const PipeConfig& pc, cudaStream_t st) {
    launch_pnp_hyp(d, pc, st);
    launch_pnp_final(d, pc, st);
}

// ============================================================================
// stage-level helpers for the stateless API
__global__ void triangulate_kernel(const float* __restrict__ Pl, const float* __restrict__ Pr,
                                   const float2* __restrict__ pL, const float2* __restrict__ pR, int n,
                                   float* __restrict__ X3) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float l[12], r[12];
    for (int k = 0; k < 12; ++k) {
        l[k] = Pl[k];
        r[k] = Pr[k];
    }
    triangulate_dlt(l, r, pL[i].x, pL[i].y, pR[i].x, pR[i].y, X3 + 3 * (long long)i);
}
void launch_triangulate(const float* Pl_dev, const float* Pr_dev, const float2* pL, const float2* pR, int n,
                        float* X3, cudaStream_t st) {
    if (n <= 0) return;
    triangulate_kernel<<<(n + 127) / 128, 128, 0, st>>>(Pl_dev, Pr_dev, pL, pR, n, X3);
}

__global__ void __launch_bounds__(1024)
compact_bytes_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, int elem, int n,
                     const uint8_t* __restrict__ keep, int* __restrict__ n_out) {
    __shared__ int sm[33];
    const int chunk = (n + 1023) / 1024;
    const int i0 = min(n, (int)threadIdx.x * chunk), i1 = min(n, i0 + chunk);
    int c = 0;
    for (int i = i0; i < i1; ++i) c += keep[i] ? 1 : 0;
    int total;
    int off = block_excl_scan<1024>(c, &total, sm);
    for (int i = i0; i < i1; ++i)
        if (keep[i]) {
            for (int b = 0; b < elem; ++b) dst[(long long)off * elem + b] = src[(long long)i * elem + b];
            ++off;
        }
    if (threadIdx.x == 0) *n_out = total;
}
void launch_compact_bytes(const uint8_t* src, uint8_t* dst, int elem_bytes, int n, const uint8_t* keep, int* n_out,
                          cudaStream_t st) {
    compact_bytes_kernel<<<1, 1024, 0, st>>>(src, dst, elem_bytes, n, keep, n_out);
}

__global__ void close_points_kernel(const float2* __restrict__ a, const float2* __restrict__ b, int n, float thr,
                                    uint8_t* __restrict__ ok) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) ok[i] = close_points(a[i].x, a[i].y, b[i].x, b[i].y, thr) ? 1 : 0;
}
void launch_close_points(const float2* a, const float2* b, int n, float thr, uint8_t* ok, cudaStream_t st) {
    if (n <= 0) return;
    close_points_kernel<<<(n + 255) / 256, 256, 0, st>>>(a, b, n, thr, ok);
}

// ---- generic bucketing (any features_per_bucket): exact Bucket::add_feature semantics.
// scratch layout (ints): bidx[n] | cnt[nb] | kept[nb + 1] | slots[nb * per_bucket]
__global__ void bucket_index_kernel(const float2* __restrict__ xy, int n, BucketGeom bg, int* __restrict__ bidx,
                                    int* __restrict__ cnt) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int br = (int)__fdiv_rn(xy[i].y, (float)bg.bh), bc = (int)__fdiv_rn(xy[i].x, (float)bg.bw);
    int b = -1;
    if (br >= 0 && br < bg.n_bh && bc >= 0 && bc < bg.n_bw) {
        b = br * bg.n_bw + bc;
        atomicAdd(cnt + b, 1);
    }
    bidx[i] = b;
}
__global__ void bucket_fill_kernel(const int* __restrict__ ages, const int* __restrict__ strengths, int n,
                                   BucketGeom bg, int per_bucket, const int* __restrict__ bidx,
                                   const int* __restrict__ cnt, int* __restrict__ kept, int* __restrict__ slots) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= bg.nb) return;
    int size = 0;
    const int max_size = (b / bg.n_bw >= bg.start_row) ? per_bucket : 0;   // feature_set.cpp:112-116
    if (cnt[b] > 0 && max_size > 0) {
        int* sl = slots + (long long)b * per_bucket;
        int seen = 0;
        for (int i = 0; i < n && seen < cnt[b]; ++i) {
            if (bidx[i] != b) continue;
            ++seen;
            const int age = ages[i], str = strengths[i];
            if (age >= bg.age_threshold) continue;
            if (size < max_size) {
                sl[size++] = i;
            } else {
                const int score = bucket_score(age, str, bg.fast_threshold);
                int smin = bucket_score(ages[sl[0]], strengths[sl[0]], bg.fast_threshold), imin_ = 0;
                for (int k = 1; k < size; ++k) {
                    const int sc = bucket_score(ages[sl[k]], strengths[sl[k]], bg.fast_threshold);
                    if (sc < smin) {
                        smin = sc;
                        imin_ = k;
                    }
                }
                if (score > smin) sl[imin_] = i;
            }
        }
    }
    kept[b] = size;
}
__global__ void bucket_emit_kernel(const float2* __restrict__ xy, const int* __restrict__ ages,
                                   const int* __restrict__ strengths, BucketGeom bg, int per_bucket,
                                   const int* __restrict__ kept_off, const int* __restrict__ slots,
                                   float2* __restrict__ xy_out, int* __restrict__ age_out, int* __restrict__ str_out) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= bg.nb) return;
    const int o0 = kept_off[b], o1 = kept_off[b + 1];
    for (int k = 0; k < o1 - o0; ++k) {
        const int i = slots[(long long)b * per_bucket + k];
        xy_out[o0 + k] = xy[i];
        age_out[o0 + k] = ages[i];
        str_out[o0 + k] = strengths[i];
    }
}
void launch_bucket_generic(float2* xy, int* ages, int* strengths, int n, BucketGeom bg, int per_bucket, int* scratch,
                           float2* xy_out, int* age_out, int* str_out, int* n_out, cudaStream_t st) {
    int* bidx = scratch;
    int* cnt = bidx + n;
    int* kept = cnt + bg.nb;
    int* slots = kept + bg.nb + 1;
    cudaMemsetAsync(cnt, 0, sizeof(int) * (2 * (size_t)bg.nb + 1), st);
    if (n > 0) bucket_index_kernel<<<(n + 255) / 256, 256, 0, st>>>(xy, n, bg, bidx, cnt);
    bucket_fill_kernel<<<(bg.nb + 127) / 128, 128, 0, st>>>(ages, strengths, n, bg, per_bucket, bidx, cnt, kept, slots);
    scan_kernel<<<1, 1024, 0, st>>>(kept, bg.nb + 1, nullptr);
    bucket_emit_kernel<<<(bg.nb + 127) / 128, 128, 0, st>>>(xy, ages, strengths, bg, per_bucket, kept, slots, xy_out,
                                                            age_out, str_out);
    cudaMemcpyAsync(n_out, kept + bg.nb, sizeof(int), cudaMemcpyDeviceToDevice, st);
}

}  // namespace svo
